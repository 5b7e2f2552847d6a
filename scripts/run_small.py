"""Small scenes (the reference's 640 x 480 case): persistent chunk kernel vs one launch per kernel.
python scripts/run_small.py [W H OBJECTS]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
w, h, nobj = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (640, 480, 7)
xyz, rgba, label = synth.make_scene(w, h, nobj, 20261104)
sx, sl, _ = seeds_from_labels(xyz, label, 0, 10)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
out = {}
for coop in ("1", "0", "1", "0"):
    os.environ["TCS_COOP"] = coop
    os.environ["TCS_COOP_MAX_ROWS"] = "100000000"
    best = None
    for rep in range(4):
        t = time.time()
        r = ctx.segment_cloud(0.006, sx, sl, preset=0, mode=capi.MODE_FAST,
                              opts=capi.default_opts(max_iterations=int(os.environ.get('MAXIT', '200000'))))
        dt = time.time() - t
        info = r["info"].as_dict()
        if best is None or dt < best[0]:
            best = (dt, info["solve_ms"], info["iterations"], max(info["rel_residual"]), r["graph"].V, info["spmm_kernel"])
        out[coop] = r["vertex_colour"].copy()
        r["graph"].close()
    print("TCS_COOP=%s: segment %.2f ms, solve %.2f ms, %d iterations (%.2f us / iteration), max rel residual %.2e, V %d, kernel %d" %
          (coop, 1e3 * best[0], best[1], best[2], 1e3 * best[1] / max(1, best[2]), best[3], best[4], best[5]), flush=True)
print("labels differ at", int(np.count_nonzero(out["1"] != out["0"])), "vertices")
