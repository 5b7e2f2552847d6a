"""Golden vectors of the reference's own assignment solver (third-party/hungarian-v2.0 driven like
src/hungarian/hungarian.cpp drives it; built by `make -C oracle _ref` into oracle/_ref/libhungarian_ref.so).
Run here (needs /root/reference); writes tests/golden/hungarian_ref.json.  Each case: overlap matrix (rows = ground-
truth labels, columns = segmentation labels, rows <= columns: the reference solver exits with "No free vertex can be
picked" when there are more ground-truth labels than segmentation labels), the reference's matching and its weight."""
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s", "_ref"])
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libhungarian_ref.so"))
    rng = np.random.default_rng(20261105)
    cases = []
    for it in range(60):
        R = int(rng.integers(1, 19))
        Cc = int(rng.integers(R, 20))
        hi = [3, 40, 300000][it % 3]                       # many ties / few ties / point-count sized
        w = rng.integers(0, hi, (R, Cc)).astype(np.int32)
        w[rng.random((R, Cc)) < 0.3] = 0
        out = np.zeros(R, np.int32)
        m = L.ref_find_assignment(w.ctypes.data_as(C.c_void_p), R, Cc, out.ctypes.data_as(C.c_void_p))
        assert m == R and len(set(out.tolist())) == R, (R, Cc, m, out.tolist(), w.tolist())
        cases.append({"weights": w.tolist(), "match": out.tolist(), "weight": int(w[np.arange(R), out].sum())})
    with open(os.path.join(ROOT, "tests", "golden", "hungarian_ref.json"), "w") as f:
        json.dump({"source": "third-party/hungarian-v2.0 via oracle/ref_hungarian_shim.cpp", "cases": cases}, f)
    print("wrote %d cases" % len(cases))


if __name__ == "__main__":
    main()
