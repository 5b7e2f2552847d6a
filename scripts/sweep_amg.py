"""Solver switches on one scene: iterations and solve time per variant.
python scripts/sweep_amg.py c4|mid|c5 'KEY=VAL,KEY=VAL' 'KEY=VAL' ...   (the empty string '' = defaults)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
kind, _, off = sys.argv[1].partition(":")     # e.g. c4:1 = the C4 shape with generator seed + 1
shape = {"c4": (4480, 2880, 15, 0.005, 10, 20261104), "mid": (2240, 1440, 15, 0.005, 10, 20261104), "c5": (640, 480, 7, 0.006, 5, 0)}[kind]
w, h, nobj, res, per, seed = shape
xyz, rgba, label = synth.make_scene(w, h, nobj, seed + int(off or 0))
sx, sl, _ = seeds_from_labels(xyz, label, int(off or 0), per)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
base = None
keys = set()
for variant in sys.argv[2:] or [""]:
    for k in keys:
        os.environ.pop(k, None)
    kv = dict(p.split("=") for p in variant.split(",") if p)
    keys |= set(kv)
    os.environ.update(kv)
    best = None
    for rep in range(2):
        r = ctx.segment_cloud(res, sx, sl, preset=0, mode=capi.MODE_PARITY if os.environ.get('TCS_SWEEP_MODE') == 'parity' else capi.MODE_FAST)
        info = r["info"].as_dict()
        if best is None or info["solve_ms"] < best["solve_ms"]:
            best = info
        lab = r["vertex_colour"].copy()
        r["graph"].close()
    if base is None:
        base = lab
    print("%-40s iterations %4d  solve %8.2f ms  (%6.1f us / iteration)  max rel %.2e  converged %s  labels differ from first variant at %d" %
          (variant or "(defaults)", best["iterations"], best["solve_ms"], 1e3 * best["solve_ms"] / max(1, best["iterations"]),
           max(best["rel_residual"]), best["converged"], int(np.count_nonzero(lab != base))), flush=True)
