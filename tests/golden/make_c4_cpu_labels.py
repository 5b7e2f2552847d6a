"""Generates tests/golden/c4_scene0_cpu_labels.npz: vertex labels of the FULL-SIZE C4 scene (BASELINE.json configs[3],
bench.py's scene 0: 4480x2880 px, 0.005 m voxels, 16 labels x 10 seeds, FAST arithmetic mode) from the CPU oracle port,
twice: with the float32 sparse LDL^T the reference uses (Eigen::SimplicialCholesky<SparseMatrix<float>>,
impl/random_walker.hpp:194) and with the float64 LDL^T, plus the vertices whose float64 top-2 potential gap is below
1e-3.  One core, ~25 + ~20 minutes, 5-7 GB (this run: 1459 s / 1202 s, profiles/r02_cpu_port_full_c4*.{json,txt}).
Test infrastructure: loads oracle/ like tests/ do.
    python tests/golden/make_c4_cpu_labels.py"""
import os, resource, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
import label_gate
import orc


class A:
    small = False


w, h, xyz, rgba, label, sx, sl = bench.c4_scene(A, 0)
g = orc.Graph.from_cloud(xyz, rgba, bench.C4_RES)          # the class path of orc_segment_cloud, stage by stage
g.normals(mode=orc.MODE_FAST); g.signed_curvatures(); g.edge_weights(mode=orc.MODE_FAST)
colours, K = label_gate.seed_colours(g, sx, sl)
out = {}
for name, solver in (("f32", orc.SOLVE_LDLT_F32), ("f64", orc.SOLVE_LDLT_F64)):
    t = time.time()
    st, lab, pot, res = g.random_walker(colours, K, solver=solver, want_potentials=True)
    out["seconds_" + name] = np.float64(time.time() - t)
    out["peak_rss_gb_" + name] = np.float64(resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6)
    out["labels_" + name] = lab.astype(np.uint8)
    print(name, "status", st, "seconds", float(out["seconds_" + name]), flush=True)
p = np.sort(pot.astype(np.float64), axis=1)
gap = (p[:, -1] - p[:, -2]).astype(np.float32)
tie = np.nonzero(gap < 1e-3)[0].astype(np.uint32)
d = np.nonzero(out["labels_f32"] != out["labels_f64"])[0]
print("float32 vs float64 LDL^T: %d vertices differ, float64 gaps there up to %.3g" % (len(d), gap[d].max() if len(d) else 0.0))
np.savez_compressed(os.path.join(HERE, "c4_scene0_cpu_labels.npz"), tie_idx=tie, tie_gap=gap[tie], V=np.int64(g.V), E=np.int64(g.E), **out)
