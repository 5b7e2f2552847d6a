#!/usr/bin/env python
"""Writes synthetic scenes as the binary file tcs_b200/cpp/tcs_batch_bench reads.
   python scripts/write_scenes.py out.bin [--config c5|c4] [--scenes 32] [--first 0]
Format: u64 count, then per scene u64 n, f32 xyz[3n], u32 rgba[n], u64 ns, f32 seed_xyz[3ns], u32 seed_label[ns]."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def write_scenes(path, scenes):
    with open(path, "wb") as f:
        f.write(np.uint64(len(scenes)).tobytes())
        for xyz, rgba, sx, sl in scenes:
            f.write(np.uint64(len(xyz)).tobytes())
            f.write(np.ascontiguousarray(xyz, np.float32).tobytes())
            f.write(np.ascontiguousarray(rgba, np.uint32).tobytes())
            f.write(np.uint64(len(sl)).tobytes())
            f.write(np.ascontiguousarray(sx, np.float32).tobytes())
            f.write(np.ascontiguousarray(sl, np.uint32).tobytes())


def read_labels(path, count):
    out = []
    with open(path, "rb") as f:
        for _ in range(count):
            v = int(np.frombuffer(f.read(8), np.uint64)[0])
            out.append(np.frombuffer(f.read(4 * v), np.uint32).copy())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--config", default="c5")
    ap.add_argument("--scenes", type=int, default=32)
    ap.add_argument("--first", type=int, default=0)
    a = ap.parse_args()
    import bench
    make = bench.c5_scene if a.config == "c5" else bench.c4_scene
    write_scenes(a.out, [make(a.first + i) for i in range(a.scenes)])
    print(a.out, os.path.getsize(a.out), "bytes")


if __name__ == "__main__":
    main()
