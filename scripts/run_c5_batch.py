"""C5 (BASELINE configs[4]): many 640x480 scenes on one GPU, T host threads each driving its own context / stream.
python scripts/run_c5_batch.py SCENES THREADS"""
import sys, os, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
ns, nt = int(sys.argv[1]), int(sys.argv[2])
scenes = []
for i in range(min(ns, 16)):     # 16 distinct scenes, cycled
    xyz, rgba, label = synth.make_c5(i)
    sx, sl, _ = seeds_from_labels(xyz, label, i, 5)
    scenes.append((xyz, rgba, sx, sl))
verts = [0] * nt
def work(t):
    ctx = capi.Context(0)
    for i in range(t, ns, nt):
        xyz, rgba, sx, sl = scenes[i % len(scenes)]
        ctx.upload_cloud(xyz, rgba)
        r = ctx.segment_cloud(0.006, sx, sl, preset=0, mode=capi.MODE_FAST)
        assert r["status"] == 0
        verts[t] += r["graph"].V
        r["graph"].close()
    ctx.close()
work_warm = threading.Thread(target=work, args=(0,)); 
t0 = time.time()
th = [threading.Thread(target=work, args=(t,)) for t in range(nt)]
for x in th: x.start()
for x in th: x.join()
dt = time.time() - t0
print("scenes %d threads %d: %.3f s, %.1f scenes/s, %.0f vertices/s" % (ns, nt, dt, ns / dt, sum(verts) / dt))
