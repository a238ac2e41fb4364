"""numpy front end of the C oracle (oracle/corr_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(raft_ros_b200/) never imports this module.

Each function mirrors one piece of the reference (see corr_oracle.c for file:line).
Parity pinning: checked against outputs of the reference's own Python executed in the
build container (tests/golden/make_golden.py -> tests/golden/*.npz); the reference has
no stored golden vectors of its own.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import List, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libcorr_oracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)


def build(force: bool = False) -> str:
    """Compile corr_oracle.c with gcc (recipe: oracle/Makefile)."""
    src = os.path.join(_HERE, "corr_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libcorr_oracle.so"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_version.restype = ctypes.c_int
        _lib.oracle_num_threads.restype = ctypes.c_int
    return _lib


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_f32p)


def _c(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def level_shapes(h: int, w: int, num_levels: int) -> List[Tuple[int, int]]:
    return [(h >> l, w >> l) for l in range(num_levels)]


def _level_ptrs(levels: Sequence[np.ndarray]):
    arr = (_f32p * len(levels))()
    for i, lv in enumerate(levels):
        arr[i] = _ptr(lv)
    return arr


def corr_build(fmap1, fmap2, num_levels: int = 4) -> List[np.ndarray]:
    """CorrBlock.__init__ (core/corr.py:13-27): list of (B*h*w, 1, h_l, w_l) fp32."""
    f1, f2 = _c(fmap1), _c(fmap2)
    B, D, h, w = f1.shape
    assert f2.shape == f1.shape
    levels = [np.empty((B * h * w, 1, hl, wl), np.float32) for hl, wl in level_shapes(h, w, num_levels)]
    rc = lib().oracle_corr_build(_ptr(f1), _ptr(f2), B, D, h, w, num_levels, _level_ptrs(levels))
    if rc:
        raise RuntimeError(f"oracle_corr_build failed rc={rc}")
    return levels


def corr_lookup(levels: Sequence[np.ndarray], coords, radius: int = 4, roundtrip: bool = False) -> np.ndarray:
    """CorrBlock.__call__ (core/corr.py:29-50): (B, L*(2r+1)^2, h, w) fp32."""
    c = _c(coords)
    B, two, h, w = c.shape
    assert two == 2
    L = len(levels)
    levels = [_c(lv) for lv in levels]
    out = np.empty((B, L * (2 * radius + 1) ** 2, h, w), np.float32)
    rc = lib().oracle_corr_lookup(_level_ptrs(levels), _ptr(c), B, h, w, L, radius, int(roundtrip), _ptr(out))
    if rc:
        raise RuntimeError(f"oracle_corr_lookup failed rc={rc}")
    return out


def corr_lookup_bwd(grad_out, coords, num_levels: int = 4, radius: int = 4) -> List[np.ndarray]:
    """d(lookup)/d(pyramid): list of (B*h*w, 1, h_l, w_l) fp32 gradients."""
    g, c = _c(grad_out), _c(coords)
    B, _, h, w = c.shape
    grads = [np.zeros((B * h * w, 1, hl, wl), np.float32) for hl, wl in level_shapes(h, w, num_levels)]
    rc = lib().oracle_corr_lookup_bwd(_ptr(g), _ptr(c), B, h, w, num_levels, radius, _level_ptrs(grads))
    if rc:
        raise RuntimeError(f"oracle_corr_lookup_bwd failed rc={rc}")
    return grads


def corr_build_bwd(grad_levels: Sequence[np.ndarray], fmap1, fmap2) -> Tuple[np.ndarray, np.ndarray]:
    """autograd of the build: grad pyramid -> (dfmap1, dfmap2), both (B,D,h,w)."""
    f1, f2 = _c(fmap1), _c(fmap2)
    B, D, h, w = f1.shape
    gl = [_c(g) for g in grad_levels]
    df1, df2 = np.empty_like(f1), np.empty_like(f2)
    rc = lib().oracle_corr_build_bwd(_level_ptrs(gl), _ptr(f1), _ptr(f2), B, D, h, w, len(gl), _ptr(df1), _ptr(df2))
    if rc:
        raise RuntimeError(f"oracle_corr_build_bwd failed rc={rc}")
    return df1, df2


def alt_fwd(fmap1_nhwc, fmap2_nhwc, coords, radius: int) -> np.ndarray:
    """alt_cuda_corr.forward (correlation_kernel.cu:18-119): (B,N,(2r+1)^2,H1,W1), unscaled."""
    f1, f2, c = _c(fmap1_nhwc), _c(fmap2_nhwc), _c(coords)
    B, H1, W1, C = f1.shape
    _, H2, W2, _ = f2.shape
    Nc = c.shape[1]
    assert c.shape == (B, Nc, H1, W1, 2)
    rd = 2 * radius + 1
    out = np.empty((B, Nc, rd * rd, H1, W1), np.float32)
    rc = lib().oracle_alt_fwd(_ptr(f1), _ptr(f2), _ptr(c), B, Nc, H1, W1, H2, W2, C, radius, _ptr(out))
    if rc:
        raise RuntimeError(f"oracle_alt_fwd failed rc={rc}")
    return out


def alt_bwd(fmap1_nhwc, fmap2_nhwc, coords, corr_grad, radius: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """alt_cuda_corr.backward (correlation_kernel.cu:122-256): (f1_grad, f2_grad, coords_grad==0)."""
    f1, f2, c, g = _c(fmap1_nhwc), _c(fmap2_nhwc), _c(coords), _c(corr_grad)
    B, H1, W1, C = f1.shape
    _, H2, W2, _ = f2.shape
    Nc = c.shape[1]
    g1, g2 = np.empty_like(f1), np.empty_like(f2)
    rc = lib().oracle_alt_bwd(_ptr(f1), _ptr(f2), _ptr(c), _ptr(g), B, Nc, H1, W1, H2, W2, C, radius, _ptr(g1), _ptr(g2))
    if rc:
        raise RuntimeError(f"oracle_alt_bwd failed rc={rc}")
    return g1, g2, np.zeros_like(c)


def coords_grid(batch: int, ht: int, wd: int) -> np.ndarray:
    """core/utils/utils.py:74-77: (B,2,h,w), channel 0 = x (column), 1 = y (row)."""
    ys, xs = np.meshgrid(np.arange(ht, dtype=np.float32), np.arange(wd, dtype=np.float32), indexing="ij")
    g = np.stack([xs, ys], axis=0)
    return np.repeat(g[None], batch, axis=0).copy()
