"""Shared input generators for the parity tests (numpy only; seeded)."""
import zlib

import numpy as np


def cotangent(name, kind, shape):
    """Same stream as tests/golden/make_golden.py::cotangent."""
    seed = zlib.crc32(f"{name}/{kind}".encode())
    return np.random.default_rng(seed).standard_normal(shape, dtype=np.float32)


def coords_grid(B, h, w):
    ys, xs = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    return np.repeat(np.stack([xs, ys], 0)[None], B, 0).copy()


def make_coords(kind, B, h, w, rng):
    base = coords_grid(B, h, w)
    if kind == "grid":
        return base
    if kind.startswith("sigma"):
        return base + float(kind[5:]) * rng.standard_normal((B, 2, h, w)).astype(np.float32)
    if kind == "oob":
        return base + ((rng.random((B, 2, h, w)) * 2 - 1) * (w + 10)).astype(np.float32)
    if kind == "border":
        c = base.copy()
        c[:, 0] = np.where(rng.random((B, h, w)) < 0.5, -0.37, w - 1 + 0.41)
        c[:, 1] = np.where(rng.random((B, h, w)) < 0.5, 0.25, h - 1 - 0.6)
        return c.astype(np.float32)
    if kind == "huge":
        c = base.copy()
        c[:, 0, ::2] = 3.0e9
        c[:, 1, 1::2] = -7.0e12
        return c
    raise ValueError(kind)


def relmax(a, b):
    """max |a-b| relative to max |b| (values cross zero, so never element-wise relative)."""
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


def normwise(a, b):
    return float(np.linalg.norm((a - b).ravel()) / (np.linalg.norm(b.ravel()) + 1e-30))
