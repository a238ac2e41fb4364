"""`alt_cuda_corr`-shaped module: the two entry points of the reference's extension.

Mirrors the pybind module of the reference (alt_cuda_corr/correlation.cpp:51-54):

    forward(fmap1, fmap2, coords, radius)             -> [corr]
    backward(fmap1, fmap2, coords, corr_grad, radius) -> [fmap1_grad, fmap2_grad, coords_grad]

with the same tensor layouts (fmaps NHWC, coords (B,N,H1,W1,2), corr (B,N,(2r+1)^2,H1,W1)),
the same list-return convention (`corr, = alt_cuda_corr.forward(...)`, core/corr.py:86) and
the same CHECK_INPUT errors (correlation.cpp:19-21).  The work is done by rc_alt_corr_fwd /
rc_alt_corr_bwd in libraftcorr.so.  To make `import alt_cuda_corr` itself resolve to this
module, see INTEGRATION.md.
"""
from __future__ import annotations

import os
from typing import List

import torch

from . import _lib


def _check_input(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != torch.float32:
        # the reference instantiates its kernels for float only (correlation_kernel.cu:278-283)
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")


def _dims(fmap1, fmap2, coords):
    if fmap1.dim() != 4 or fmap2.dim() != 4 or coords.dim() != 5 or coords.shape[-1] != 2:
        raise RuntimeError("expected fmap1 (B,H1,W1,C), fmap2 (B,H2,W2,C), coords (B,N,H1,W1,2)")
    B, H1, W1, C = fmap1.shape
    B2, H2, W2, C2 = fmap2.shape
    Bc, N, Hc, Wc, _ = coords.shape
    if B2 != B or Bc != B or C2 != C or Hc != H1 or Wc != W1:
        raise RuntimeError(f"shape mismatch: fmap1 {tuple(fmap1.shape)}, fmap2 {tuple(fmap2.shape)}, coords {tuple(coords.shape)}")
    if not (fmap1.device == fmap2.device == coords.device):
        raise RuntimeError("fmap1, fmap2 and coords must be on the same device")
    return B, N, H1, W1, H2, W2, C


def forward(fmap1: torch.Tensor, fmap2: torch.Tensor, coords: torch.Tensor, radius: int) -> List[torch.Tensor]:
    _check_input(fmap1, "fmap1")
    _check_input(fmap2, "fmap2")
    _check_input(coords, "coords")
    B, N, H1, W1, H2, W2, C = _dims(fmap1, fmap2, coords)
    rd = 2 * int(radius) + 1
    with torch.cuda.device(fmap1.device):
        corr = torch.empty((B, N, rd * rd, H1, W1), dtype=torch.float32, device=fmap1.device)
        rc = _lib.lib().rc_alt_corr_fwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(), corr.data_ptr(),
                                        B, N, H1, W1, H2, W2, C, int(radius),
                                        torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "alt_cuda_corr.forward")
    return [corr]


def backward(fmap1: torch.Tensor, fmap2: torch.Tensor, coords: torch.Tensor, corr_grad: torch.Tensor,
             radius: int) -> List[torch.Tensor]:
    _check_input(fmap1, "fmap1")
    _check_input(fmap2, "fmap2")
    _check_input(coords, "coords")
    _check_input(corr_grad, "corr_grad")
    B, N, H1, W1, H2, W2, C = _dims(fmap1, fmap2, coords)
    rd = 2 * int(radius) + 1
    if tuple(corr_grad.shape) != (B, N, rd * rd, H1, W1):
        raise RuntimeError(f"corr_grad must be {(B, N, rd * rd, H1, W1)}, got {tuple(corr_grad.shape)}")
    with torch.cuda.device(fmap1.device):
        fmap1_grad = torch.empty_like(fmap1)
        coords_grad = torch.zeros_like(coords)  # never written by the reference either (:307)
        L = _lib.lib()
        st = torch.cuda.current_stream().cuda_stream
        if os.environ.get("RAFT_ALT_BWD_ATOMICS") == "1":
            # the reference's scheme: float atomics into a zeroed buffer (run-to-run differences in the low bits)
            fmap2_grad = torch.zeros_like(fmap2)
            rc = L.rc_alt_corr_bwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(), corr_grad.data_ptr(),
                                   fmap1_grad.data_ptr(), fmap2_grad.data_ptr(), B, N, H1, W1, H2, W2, C, int(radius), st)
        else:
            # default: bit-reproducible fmap2_grad (fixed-point accumulation, rc_alt_corr_bwd_det)
            fmap2_grad = torch.empty_like(fmap2)
            nbytes = int(L.rc_alt_corr_bwd_det_workspace_bytes(B, H2, W2, C))
            ws = torch.empty(nbytes, dtype=torch.uint8, device=fmap1.device)
            rc = L.rc_alt_corr_bwd_det(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(), corr_grad.data_ptr(),
                                       fmap1_grad.data_ptr(), fmap2_grad.data_ptr(), B, N, H1, W1, H2, W2, C, int(radius),
                                       ws.data_ptr(), nbytes, st)
            ws.record_stream(torch.cuda.current_stream())
        _lib.check(rc, "alt_cuda_corr.backward")
    return [fmap1_grad, fmap2_grad, coords_grad]
