"""Sampling helpers of the hot path (reference core/utils/utils.py:57-77)."""
from __future__ import annotations

import torch

from . import _lib


class _BilinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, coords, want_mask):
        im = img.detach().float().contiguous()
        co = coords.detach().float().contiguous()
        M, C, H, W = im.shape
        _, Ho, Wo, _ = co.shape
        with torch.cuda.device(im.device):
            out = torch.empty((M, C, Ho, Wo), dtype=torch.float32, device=im.device)
            mask = torch.empty((M, Ho, Wo, 1), dtype=torch.float32, device=im.device) if want_mask else None
            rc = _lib.lib().rc_bilinear_sampler_fwd(im.data_ptr(), co.data_ptr(), out.data_ptr(),
                                                    mask.data_ptr() if want_mask else None, M, C, H, W, Ho, Wo,
                                                    torch.cuda.current_stream().cuda_stream)
            _lib.check(rc, "rc_bilinear_sampler_fwd")
        ctx.save_for_backward(co)
        ctx.img_shape = (M, C, H, W)
        if want_mask:
            ctx.mark_non_differentiable(mask)
            return out, mask
        return out, None

    @staticmethod
    def backward(ctx, gout, _gmask):
        (co,) = ctx.saved_tensors
        M, C, H, W = ctx.img_shape
        _, Ho, Wo, _ = co.shape
        g = gout.detach().float().contiguous()
        with torch.cuda.device(g.device):
            gimg = torch.zeros((M, C, H, W), dtype=torch.float32, device=g.device)
            rc = _lib.lib().rc_bilinear_sampler_bwd(g.data_ptr(), co.data_ptr(), gimg.data_ptr(), M, C, H, W, Ho, Wo,
                                                    torch.cuda.current_stream().cuda_stream)
            _lib.check(rc, "rc_bilinear_sampler_bwd")
        return gimg, None, None


def bilinear_sampler(img, coords, mode='bilinear', mask=False):
    """Pixel-coordinate bilinear sampling, zeros outside (reference core/utils/utils.py:57-71).

    img (M,C,H,W), coords (M,Ho,Wo,2) [x,y] in pixels -> (M,C,Ho,Wo); with mask=True also the
    in-bounds mask (M,Ho,Wo,1) of utils.py:67-69.  `mode` is accepted and ignored, as in the
    reference.  Gradients flow to `img` only: the one caller on the hot path detaches coords
    (core/raft.py:123).
    """
    if not img.is_cuda:
        raise RuntimeError("img must be a CUDA tensor")
    if img.dim() != 4 or coords.dim() != 4 or coords.shape[-1] != 2 or coords.shape[0] != img.shape[0]:
        raise RuntimeError(f"expected img (M,C,H,W) and coords (M,Ho,Wo,2); got {tuple(img.shape)}, {tuple(coords.shape)}")
    if coords.requires_grad and torch.is_grad_enabled():
        raise RuntimeError("bilinear_sampler: gradients w.r.t. coords are not provided (detach them, as core/raft.py:123 does)")
    out, m = _BilinearFn.apply(img, coords, bool(mask))
    if mask:
        return out, m
    return out


def coords_grid(batch, ht, wd, device):
    """(B,2,h,w) pixel grid, channel 0 = x (column), 1 = y (row) (reference utils.py:74-77)."""
    ys = torch.arange(ht, device=device, dtype=torch.float32).view(1, 1, ht, 1).expand(batch, 1, ht, wd)
    xs = torch.arange(wd, device=device, dtype=torch.float32).view(1, 1, 1, wd).expand(batch, 1, ht, wd)
    return torch.cat([xs, ys], dim=1).contiguous()


def upsample_flow(flow, mask):
    """Upsample flow field [H/8, W/8, 2] -> [H, W, 2] using convex combination: `RAFT.upsample_flow` (reference
    core/raft.py:72-83) as one kernel (rc_convex_upsample_fwd).  flow (N,2,H,W), mask (N,576,H,W) -> (N,2,8H,8W) fp32.
    Inference only: tensors that require a gradient are refused."""
    if not flow.is_cuda or not mask.is_cuda:
        raise RuntimeError("flow and mask must be CUDA tensors")
    if flow.dim() != 4 or flow.shape[1] != 2 or mask.dim() != 4 or mask.shape[1] != 576 or mask.shape[0] != flow.shape[0] \
            or mask.shape[2:] != flow.shape[2:]:
        raise RuntimeError(f"expected flow (N,2,H,W) and mask (N,576,H,W); got {tuple(flow.shape)}, {tuple(mask.shape)}")
    if torch.is_grad_enabled() and (flow.requires_grad or mask.requires_grad):
        raise RuntimeError("upsample_flow: inference only (no gradient); use the reference's torch ops for training")
    f = flow.detach().float().contiguous()
    m = mask.detach().float().contiguous()
    N, _, H, W = f.shape
    with torch.cuda.device(f.device):
        out = torch.empty((N, 2, 8 * H, 8 * W), dtype=torch.float32, device=f.device)
        rc = _lib.lib().rc_convex_upsample_fwd(f.data_ptr(), m.data_ptr(), out.data_ptr(), N, H, W,
                                               torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "rc_convex_upsample_fwd")
    return out
