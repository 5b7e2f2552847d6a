"""Nearest-neighbours graph builder on a full 640 x 480 scene: python scripts/run_knn.py [K]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
k = int(sys.argv[1]) if len(sys.argv) > 1 else 10
ctx = capi.Context(0)
for (w, h) in ((640, 480), (1280, 960), (4480, 2880)):
    xyz, rgba, label = synth.make_scene(w, h, 7, 20261104)
    ctx.upload_cloud(xyz, rgba)
    for rep in range(3):
        t = time.time()
        g = ctx.build_nn_graph(k)
        dt = time.time() - t
        V, E = g.V, g.E
        g.close()
    print("%dx%d: %d points -> V %d E %d, kNN(k=%d) graph in %.2f ms (%.1f M points/s)" % (w, h, len(xyz), V, E, k, 1e3 * dt, V / dt / 1e6), flush=True)
