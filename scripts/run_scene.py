"""Segment one synthetic scene and time the SpMM alone: python scripts/run_scene.py W H OBJECTS RES MAXIT [CHECK_EVERY]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
w, h, nobj, res, maxit = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5])
ce = int(sys.argv[6]) if len(sys.argv) > 6 else 50
xyz, rgba, label = synth.make_scene(w, h, nobj, 20261104)
sx, sl, _ = seeds_from_labels(xyz, label, 0, 10)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
for rep in range(2):
    t = time.time()
    r = ctx.segment_cloud(res, sx, sl, preset=0, mode=capi.MODE_FAST,
                          opts=capi.default_opts(max_iterations=maxit, time_spmm=1, check_every=ce))
    dt = time.time() - t
    g = r["graph"]; info = r["info"].as_dict()
    print("segment %.3f s V %d E %d status %d K %d iterations %d solve_ms %.1f kernel %d" %
          (dt, g.V, g.E, r["status"], info["ncolours"], info["iterations"], info["solve_ms"], info["spmm_kernel"]), flush=True)
    ms, nb = g.bench_spmm(20)
    print("spmm ms %.4f bytes %d GB/s %.1f" % (ms, nb, nb / ms / 1e6), flush=True)
    g.close()
