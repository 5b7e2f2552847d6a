"""Multilevel preconditioner vs plain Jacobi-PCG on one scene: iterations, times, label agreement.
python scripts/run_amg.py [W H OBJECTS RES SEEDS]      (default: the 640 x 480 C5 shape)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
a = sys.argv[1:]
w, h, nobj = (int(v) for v in a[:3]) if len(a) >= 3 else (640, 480, 7)
res = float(a[3]) if len(a) > 3 else 0.006
per = int(a[4]) if len(a) > 4 else 5
modes = a[5].split(",") if len(a) > 5 else ["1", "0"]
xyz, rgba, label = synth.make_scene(w, h, nobj, 0 if (w, h) == (640, 480) else 20261104)
sx, sl, _ = seeds_from_labels(xyz, label, 0, per)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
out, pot = {}, {}
for amg in modes:
    os.environ["TCS_AMG"] = amg
    best = None
    for rep in range(3):
        t = time.time()
        r = ctx.segment_cloud(res, sx, sl, preset=0, mode=capi.MODE_FAST, opts=capi.default_opts())
        dt = time.time() - t
        info = r["info"].as_dict()
        if best is None or dt < best[0]:
            best = (dt, info["solve_ms"], info["iterations"], max(info["rel_residual"]), r["graph"].V, info["spmm_kernel"], info["stage_ms"])
        out[amg] = r["vertex_colour"].copy()
        pot[amg] = r["graph"].best()[0]
        r["graph"].close()
    print("TCS_AMG=%s: segment %.2f ms, solve %.2f ms, %d iterations (%.1f us / iteration), max rel residual %.2e, V %d, kernel %d, status %d" %
          (amg, 1e3 * best[0], best[1], best[2], 1e3 * best[1] / max(1, best[2]), best[3], best[4], best[5], r["status"]), flush=True)
    print("   stage_ms", [round(float(v), 3) for v in best[6]], flush=True)
if len(out) == 2:
    d = out["1"] != out["0"]
    print("labels differ at", int(np.count_nonzero(d)), "of", len(d), "vertices")
    if pot["1"] is not None and pot["0"] is not None:
        print("max |potential difference| %.3e" % float(np.abs(pot["1"] - pot["0"]).max()))
