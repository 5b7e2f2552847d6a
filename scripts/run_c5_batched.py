"""C5 shape (640 x 480 scenes) through tcs_segment_batch: throughput per batch size and label agreement with the
single-scene path.  python scripts/run_c5_batched.py [NSCENES] [BATCH ...]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 32
batches = [int(a) for a in sys.argv[2:]] or [1, 4, 8, 16, 32]
scenes = []
for i in range(ns):
    xyz, rgba, label = synth.make_c5(i)
    sx, sl, _ = seeds_from_labels(xyz, label, i, 5)
    scenes.append((xyz, rgba, sx, sl))
ctx = capi.Context(0)
single = []
t = time.time()
for xyz, rgba, sx, sl in scenes:
    ctx.upload_cloud(xyz, rgba)
    r = ctx.segment_cloud(0.006, sx, sl, preset=capi.PRESET_CLASS, mode=capi.MODE_FAST)
    single.append(r["vertex_colour"]); r["graph"].close()
dt = time.time() - t
print("single-scene path: %d scenes in %.3f s = %.1f scenes/s" % (ns, dt, ns / dt), flush=True)
for B in batches:
    for rep in range(2):
        t = time.time()
        outs, infos = [], []
        for b0 in range(0, ns, B):
            o, info, st = ctx.segment_batch(scenes[b0:b0 + B], 0.006, preset=capi.PRESET_CLASS, mode=capi.MODE_FAST)
            assert st == 0
            outs += o; infos.append(info.as_dict())
        dt = time.time() - t
    V = sum(o["V"] for o in outs)
    bad = sum(int(np.count_nonzero(o["vertex_colour"] != s)) for o, s in zip(outs, single))
    i0 = infos[0]
    print("batch %2d: %d scenes in %.3f s = %.1f scenes/s, %.2f M vertices/s; first batch: rows %d, %d iterations, prepare %.1f ms, merge %.2f ms, solve %.1f ms, labels %.2f ms, kernel %d, max rel %.2e; label differences vs single-scene path: %d of %d" %
          (B, ns, dt, ns / dt, V / dt / 1e6, i0["rows"], i0["iterations"], i0["stage_ms"][0], i0["stage_ms"][8], i0["stage_ms"][9], i0["stage_ms"][10], i0["spmm_kernel"], max(i0["rel_residual"]), bad, V), flush=True)
