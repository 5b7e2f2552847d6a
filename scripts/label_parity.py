"""Label parity survey: GPU labels vs the oracle's float64 direct solve on every parity configuration, with the
oracle's top-2 potential gap at each mismatching vertex.  python scripts/label_parity.py [tolerance ...]
Prints a markdown table (profiles/r02_label_parity.md)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import orc, label_gate
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels

tols = [float(a) for a in sys.argv[1:]] or [1e-6, 1e-8]
ctx = capi.Context(0)
cases = []
for i, name in enumerate(("test1", "test13", "test47", "test56", "test60")):
    d = np.load(os.path.join(ROOT, "tests", "golden", "scene_%s.npz" % name))
    for preset in (0, 1):
        cases.append(("%s preset %s PARITY" % (name, "class" if preset == 0 else "cli"), d["xyz"], d["rgba"], d["label"], i, 0.006, 5, preset, 0))
for s in (0, 1, 2):
    xyz, rgba, label = synth.make_c5(s)
    cases.append(("C5 scene %d FAST" % s, xyz, rgba, label, s, 0.006, 5, 0, 1))
w, h = 4480, 2880
xyz, rgba, label = synth.make_scene(w, h, 15, 20261104)
cw, ch = 1024, 768
r0, c0 = (h - ch) // 2, (w - cw) // 2
X = xyz.reshape(h, w, 3)[r0:r0 + ch, c0:c0 + cw].reshape(-1, 3).copy()
C = rgba.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
Lb = label.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
cases.append(("C4 centre crop 1024x768 FAST", X, C, Lb, 0, 0.005, 10, 0, 1))
print("| configuration | V | K | tolerance | iterations | max rel. residual | mismatching vertices | largest oracle top-2 gap |")
print("|---|---|---|---|---|---|---|---|")
for tag, xyz, rgba, label, sid, res, per, preset, mode in cases:
    sx, sl, _ = seeds_from_labels(xyz, label, sid, per)
    ref = orc.segment_cloud(xyz, rgba, res, sx, sl, preset=preset, mode=mode, solver=orc.SOLVE_LDLT_F64)
    ctx.upload_cloud(xyz, rgba)
    for tol in tols:
        r = ctx.segment_cloud(res, sx, sl, preset=preset, mode=mode, opts=capi.default_opts(tolerance=tol))
        info = r["info"].as_dict()
        diff = np.nonzero(r["vertex_colour"] != ref["vertex_colour"])[0]
        gaps = label_gate.oracle_gaps(ref["graph"], sx, sl, diff) if len(diff) else np.zeros(0)
        print("| %s | %d | %d | %.0e | %d | %.2e | %d | %s |" % (tag, len(ref["vertex_colour"]), ref["K"], tol, info["iterations"],
              max(info["rel_residual"]), len(diff), ("%.3e" % gaps.max()) if len(gaps) else "-"), flush=True)
        r["graph"].close()
