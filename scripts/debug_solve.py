import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tcs_b200 import capi
from tcs_b200.seeds import seeds_from_labels
name = sys.argv[1] if len(sys.argv) > 1 else "test56"
ids = dict(test1=0, test13=1, test47=2, test56=3, test60=4)
d = np.load("tests/golden/scene_%s.npz" % name)
xyz, rgba, label = d["xyz"], d["rgba"], d["label"]
sx, sl, _ = seeds_from_labels(xyz, label, ids[name], 5)
ctx = capi.Context(0)
ctx.upload_cloud(xyz, rgba)
r = ctx.segment_cloud(0.006, sx, sl, preset=int(os.environ.get("PRESET", "0")), mode=0, opts=capi.default_opts(max_iterations=int(sys.argv[2]) if len(sys.argv) > 2 else 6000, check_every=int(sys.argv[3]) if len(sys.argv) > 3 else 50))
print(r["status"], r["info"].as_dict())
