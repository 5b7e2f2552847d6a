"""Turn the reference's five RGB-D scenes (data/test*.pcd: fields x y z rgba label, binary_compressed)
into compact .npz fixtures that travel to the GPU box (where /root/reference does not exist).

Run once in the build container:  python tests/golden/make_scene_fixtures.py
Inputs are read-only data files, not source code; the arrays are stored bit-for-bit (float32 xyz incl. NaN,
uint32 rgba with the alpha byte as stored in the file, uint32 ground-truth label).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from tcs_b200.pcd import read_pcd  # noqa: E402

REF = os.environ.get("TCS_REFERENCE", "/root/reference")
for name in ("test1", "test13", "test47", "test56", "test60"):
    f, h = read_pcd(os.path.join(REF, "data", name + ".pcd"))
    xyz = np.stack([f["x"], f["y"], f["z"]], axis=1).astype(np.float32)
    out = os.path.join(HERE, "scene_%s.npz" % name)
    np.savez_compressed(out, xyz=xyz, rgba=f["rgba"].astype(np.uint32), label=f["label"].astype(np.uint32),
                        width=np.int32(h["WIDTH"][0]), height=np.int32(h["HEIGHT"][0]))
    print(name, xyz.shape, os.path.getsize(out))
