"""Synthetic organised RGB-D clouds for the throughput configurations.

A pinhole camera (fx = fy = 525, principal point at the image centre, like the Kinect frames of the reference's
data/*.pcd) looks at a tilted table plane (30 degrees, z about 1.0-1.3 m in front of the camera), a back wall and
`n_objects` convex objects (spheres, boxes, upright cylinders, 8-25 cm) standing on the plane.  Depth gets noise
sigma = 1 mm * z^2; a border and random 8x8 holes are set to NaN (~22 % of the pixels); colour = per-object base
colour + N(0, 8).  Ground-truth label 0 = plane + wall, 1..n_objects = objects  =>  n_objects + 1 labels.

Everything is numpy with a seeded Generator, so the CPU oracle and the GPU path see the same bytes.
"""
from __future__ import annotations

import numpy as np

FX = 525.0


def make_scene(width: int = 640, height: int = 480, n_objects: int = 7, seed: int = 0, nan_fraction: float = 0.22):
    """Return (xyz float32 (N,3) row-major organised, rgba uint32 (N,), label uint32 (N,))."""
    rng = np.random.default_rng(seed)
    cx, cy = width / 2.0, height / 2.0
    u = (np.arange(width, dtype=np.float32) - np.float32(cx)) / np.float32(FX)
    v = (np.arange(height, dtype=np.float32) - np.float32(cy)) / np.float32(FX)
    rx = np.broadcast_to(u[None, :], (height, width))
    ry = np.broadcast_to(v[:, None], (height, width))

    # table plane n.p = d, tilted 30 degrees about the x axis; where the ray misses it: wall at z = 2.6
    th = np.deg2rad(30.0)
    sn, cs = np.float32(np.sin(th)), np.float32(np.cos(th))
    d_plane = np.float32(1.15 * cs)
    den = cs + sn * ry
    with np.errstate(divide="ignore", invalid="ignore"):
        z = np.where(den > 0.3, d_plane / den, np.float32(np.inf)).astype(np.float32)
    z = np.minimum(z, np.float32(2.6))
    label = np.zeros((height, width), np.uint32)

    # objects: centres picked on the plane inside the view, ray-cast and z-buffered
    span_x = 0.85 * (width / 2.0) / FX
    span_y = 0.85 * (height / 2.0) / FX
    min_pixels = max(64, int(2e-4 * width * height))
    for k in range(n_objects):
        kind = k % 3
        for attempt in range(32):  # re-draw until the object is visible (not occluded / outside the view)
            size = np.float32(rng.uniform(0.08, 0.25))
            # a ray inside the view, its hit point on the plane = object base
            ax, ay = np.float32(rng.uniform(-span_x, span_x)), np.float32(rng.uniform(-0.6 * span_y, span_y))
            dn = cs + sn * ay
            if dn < 0.35:
                ay = np.float32((0.35 - cs) / sn + 0.05)
                dn = cs + sn * ay
            zc = d_plane / dn
            base = np.array([ax * zc, ay * zc, zc], np.float32)
            nrm = np.array([0.0, sn, cs], np.float32)          # plane normal (points away from the camera side)
            c = base - nrm * (size * np.float32(0.5))          # centre: half a size above the plane, towards the camera
            if kind == 0:       # sphere |p - c| = r
                r = size * np.float32(0.5)
                b = rx * c[0] + ry * c[1] + c[2]
                a = rx * rx + ry * ry + np.float32(1.0)
                disc = b * b - a * (np.float32(c @ c) - r * r)
                with np.errstate(invalid="ignore"):
                    t = (b - np.sqrt(disc)) / a
                hit = (disc > 0) & (t > 0.3)
            elif kind == 1:     # axis-aligned box (slab test)
                h = size * np.float32(0.5) * np.array([1.0, 0.8, 0.9], np.float32)
                lo, hi = c - h, c + h
                with np.errstate(divide="ignore", invalid="ignore"):
                    tx1, tx2 = lo[0] / rx, hi[0] / rx
                    ty1, ty2 = lo[1] / ry, hi[1] / ry
                tz1, tz2 = lo[2], hi[2]
                tmin = np.maximum(np.maximum(np.minimum(tx1, tx2), np.minimum(ty1, ty2)), min(tz1, tz2))
                tmax = np.minimum(np.minimum(np.maximum(tx1, tx2), np.maximum(ty1, ty2)), max(tz1, tz2))
                t = tmin
                hit = (tmax >= tmin) & (tmin > 0.3)
            else:               # cylinder with axis along y, radius r, half height hh (side surface + caps via slab)
                r = size * np.float32(0.35)
                hh = size * np.float32(0.5)
                a = rx * rx + np.float32(1.0)
                b = rx * c[0] + c[2]
                disc = b * b - a * (c[0] * c[0] + c[2] * c[2] - r * r)
                with np.errstate(invalid="ignore", divide="ignore"):
                    sq = np.sqrt(disc)
                    t0, t1 = (b - sq) / a, (b + sq) / a
                    ty1, ty2 = (c[1] - hh) / ry, (c[1] + hh) / ry
                tmin = np.maximum(t0, np.minimum(ty1, ty2))
                tmax = np.minimum(t1, np.maximum(ty1, ty2))
                t = tmin
                hit = (disc > 0) & (tmax >= tmin) & (tmin > 0.3)
            with np.errstate(invalid="ignore"):
                closer = hit & (t < z)
            if np.count_nonzero(closer) < min_pixels and attempt < 31:
                continue
            z = np.where(closer, t, z).astype(np.float32)
            label[closer] = k + 1
            break

    # depth noise sigma = 1 mm * z^2
    z = (z + rng.standard_normal((height, width), dtype=np.float32) * np.float32(0.001) * z * z).astype(np.float32)

    # invalid pixels: border + random 8x8 holes up to nan_fraction
    invalid = np.zeros((height, width), bool)
    bw, bh = max(1, int(0.025 * width)), max(1, int(0.025 * height))
    invalid[:bh, :] = invalid[-bh:, :] = True
    invalid[:, :bw] = invalid[:, -bw:] = True
    border = invalid.mean()
    hole_p = max(0.0, (nan_fraction - border) / max(1e-9, 1.0 - border))
    holes = rng.random((height + 7) // 8 * ((width + 7) // 8)) < hole_p
    holes = holes.reshape((height + 7) // 8, (width + 7) // 8)
    invalid |= np.kron(holes, np.ones((8, 8), bool))[:height, :width]

    xyz = np.empty((height, width, 3), np.float32)
    xyz[..., 0] = rx * z
    xyz[..., 1] = ry * z
    xyz[..., 2] = z
    xyz[invalid] = np.nan

    # colours: per-label base colour + N(0, 8), alpha 255
    base_col = rng.integers(40, 216, size=(n_objects + 1, 3)).astype(np.float32)
    base_col[0] = (150, 140, 130)
    col = base_col[label] + rng.standard_normal((height, width, 3), dtype=np.float32) * np.float32(8.0)
    col = np.clip(np.rint(col), 0, 255).astype(np.uint32)
    rgba = (np.uint32(255) << np.uint32(24)) | (col[..., 0] << np.uint32(16)) | (col[..., 1] << np.uint32(8)) | col[..., 2]
    return xyz.reshape(-1, 3), rgba.reshape(-1).astype(np.uint32), label.reshape(-1)


def make_c4(seed: int = 20261104):
    """BASELINE config 3: ~10 M finite points, 15 objects + background = 16 labels (voxelise at 0.005)."""
    return make_scene(4480, 2880, 15, seed)


def make_c5(scene_index: int):
    """BASELINE config 4: one of the 1024 Kinect-sized scenes, 7 objects + background = 8 labels (voxelise at 0.006)."""
    return make_scene(640, 480, 7, scene_index)
