"""e2e of the C++ class on the C4 scene: writes the scene as api_check's input and runs
`api_check class` with TCS_API_REPEAT (RandomWalkerSegmentation::segment from a host cloud to host clusters).
python scripts/class_e2e_c4.py [repeat]"""
import os, struct, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench

class A: small = False
w, h, xyz, rgba, label, sx, sl = bench.c4_scene(A, 0)
d = tempfile.mkdtemp()
inp, out = os.path.join(d, "c4.bin"), os.path.join(d, "out.bin")
with open(inp, "wb") as f:
    f.write(struct.pack("<Q", len(xyz))); f.write(np.ascontiguousarray(xyz, np.float32).tobytes())
    f.write(np.ascontiguousarray(rgba, np.uint32).tobytes())
    f.write(struct.pack("<Q", len(sl))); f.write(np.ascontiguousarray(sx, np.float32).tobytes())
    f.write(np.ascontiguousarray(sl, np.uint32).tobytes())
subprocess.check_call(["make", "-C", os.path.join(ROOT, "tcs_b200", "cpp"), "-s", "api_check"])
env = dict(os.environ, TCS_API_REPEAT=sys.argv[1] if len(sys.argv) > 1 else "3", TCS_API_POTENTIALS="0")
r = subprocess.run([os.path.join(ROOT, "tcs_b200", "cpp", "api_check"), "class", inp, out, str(bench.C4_RES)], env=env,
                   capture_output=True, text=True)
print(r.stdout.strip()); print(r.stderr.strip()[-500:])
os.remove(inp); os.path.exists(out) and os.remove(out)
