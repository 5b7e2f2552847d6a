"""Second preconditioner level (csrc/coarse.cuh) on the C4 scene: iterations, solve time and label agreement with the
Jacobi-only solve for a list of TCS_COARSE_* settings.  python scripts/sweep_coarse.py [W H] 'THETA:BLOCK' ..."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
args = sys.argv[1:]
w, h = 4480, 2880
if len(args) >= 2 and args[0].isdigit():
    w, h = int(args[0]), int(args[1]); args = args[2:]
xyz, rgba, label = synth.make_scene(w, h, 15, 20261104)
sx, sl, _ = seeds_from_labels(xyz, label, 0, 10)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
def run(env):
    for k in ("TCS_COARSE", "TCS_COARSE_THETA", "TCS_COARSE_BLOCK"):
        os.environ.pop(k, None)
    os.environ.update(env)
    best = None
    for rep in range(2):
        t = time.time()
        r = ctx.segment_cloud(0.005, sx, sl, preset=0, mode=capi.MODE_FAST)
        dt = time.time() - t
        info = r["info"].as_dict()
        vc = r["vertex_colour"].copy()
        r["graph"].close()
        if best is None or dt < best[0]:
            best = (dt, info, vc)
    return best
base = run({"TCS_COARSE": "0"})
print("jacobi: %.1f ms wall, solve %.1f ms, %d iterations, converged %s, max rel %.3g" %
      (1e3 * base[0], base[1]["solve_ms"], base[1]["iterations"], base[1]["converged"], max(base[1]["rel_residual"])), flush=True)
for spec in args:
    th, blk = spec.split(":")
    b = run({"TCS_COARSE": "1", "TCS_COARSE_THETA": th, "TCS_COARSE_BLOCK": blk})
    diff = int(np.count_nonzero(b[2] != base[2]))
    print("theta %s block %s: %.1f ms wall, solve %.1f ms, %d iterations, converged %s, max rel %.3g, labels differing from jacobi: %d of %d" %
          (th, blk, 1e3 * b[0], b[1]["solve_ms"], b[1]["iterations"], b[1]["converged"], max(b[1]["rel_residual"]), diff, len(b[2])), flush=True)
