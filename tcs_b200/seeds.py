"""Deterministic seed recipe for the benchmark / parity scenes.

The reference's ``create_seeds`` tool (src/create_seeds.cpp:36-52) needs a directory of per-label PCDs and numbers
labels from 1 in directory order; the shipped scenes only carry a ground-truth ``label`` field, so seeds are drawn
from it: for each ground-truth label l present among the finite points (ascending) draw S distinct finite points
with that label using splitmix64 seeded with 1000*scene_id + l (index = next() mod count, repeats rejected);
seed xyz = that point's xyz, seed label = l + 1.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & _M64

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & _M64
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
        return z ^ (z >> 31)


def seeds_from_labels(xyz: np.ndarray, label: np.ndarray, scene_id: int, per_label: int = 5):
    """Return (seed_xyz float32 (S,3), seed_label uint32 (S,), seed_point_index int64 (S,))."""
    finite = np.isfinite(xyz).all(axis=1) & (xyz[:, 2] > 0)
    sx, sl, si = [], [], []
    for l in np.unique(label[finite]):
        idx = np.nonzero(finite & (label == l))[0]
        rng = SplitMix64(1000 * scene_id + int(l))
        chosen = []
        want = min(per_label, len(idx))
        while len(chosen) < want:
            k = rng.next() % len(idx)
            if k not in chosen:
                chosen.append(k)
        for k in chosen:
            sx.append(xyz[idx[k]])
            sl.append(int(l) + 1)
            si.append(int(idx[k]))
    return np.asarray(sx, np.float32).reshape(-1, 3), np.asarray(sl, np.uint32), np.asarray(si, np.int64)
