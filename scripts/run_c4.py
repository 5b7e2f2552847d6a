import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi, synth
from tcs_b200.seeds import seeds_from_labels
t = time.time(); xyz, rgba, label = synth.make_c4(); print("gen", time.time() - t, flush=True)
sx, sl, _ = seeds_from_labels(xyz, label, 0, 10)
print("seeds", len(sl), flush=True)
ctx = capi.Context(0)
t = time.time(); ctx.upload_cloud(xyz, rgba); print("upload", time.time() - t, flush=True)
for rep in range(2):
    t = time.time()
    r = ctx.segment_cloud(0.005, sx, sl, preset=0, mode=capi.MODE_FAST,
                          opts=capi.default_opts(max_iterations=int(sys.argv[1]) if len(sys.argv) > 1 else 40000, time_spmm=1, check_every=int(sys.argv[2]) if len(sys.argv) > 2 else 50))
    dt = time.time() - t
    g = r["graph"]; info = r["info"].as_dict()
    print("segment", dt, "V", g.V, "E", g.E, "status", r["status"], flush=True)
    print(info, flush=True)
    ms, nb = g.bench_spmm(20)
    print("spmm ms", ms, "bytes", nb, "GB/s", nb / ms / 1e6, flush=True)
    vc = r["vertex_colour"]; print(np.bincount(vc))
    g.close()
