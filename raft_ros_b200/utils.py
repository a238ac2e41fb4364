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


_ACTS = {None: 0, "none": 0, "relu": 1, "sigmoid": 2, "tanh": 3}
_CONV_CACHE = None


_EPIS = {None: 0, "mul_from": 1, "gru": 2}


def _chw(t, name):
    """(B, C, H, W) fp32 CUDA tensor whose items are dense C x H x W blocks (a channel slice of a wider tensor is fine):
    returns its batch stride in floats."""
    if (not t.is_cuda) or t.dim() != 4 or t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be a (B,C,H,W) fp32 CUDA tensor")
    B, C, H, W = t.shape
    if t.stride(3) != 1 or t.stride(2) != W or t.stride(1) != H * W or (B > 1 and t.stride(0) < C * H * W):
        raise RuntimeError(f"{name}: items must be dense C*H*W blocks (got strides {t.stride()})")
    return t.stride(0) if B > 1 else C * H * W


def _prepared_filters(weight, bias, o0, o1, dev, st):
    """fp16 filter blocks (and the padded bias) of output channels [o0, o1), cached per weight / bias tensor object."""
    global _CONV_CACHE
    from .corr import _WeightCache
    if _CONV_CACHE is None:
        _CONV_CACHE = _WeightCache()
    L = _lib.lib()
    _, Cin, KH, KW = weight.shape
    co = (o1 - o0 + 15) // 16 * 16  # the kernel's N granule; padded filters are zero, their outputs dropped by the caller

    def make():
        nbytes = int(L.rc_conv_weight_bytes(co, Cin, KH, KW))
        wp_ = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
        w32 = torch.zeros((co, Cin, KH, KW), dtype=torch.float32, device=dev)
        w32[:o1 - o0] = weight.detach()[o0:o1]
        _lib.check(L.rc_conv_prep_weight(w32.data_ptr(), wp_.data_ptr(), co, Cin, KH, KW, st), "rc_conv_prep_weight")
        return wp_

    def make_bias():
        b32_ = torch.zeros(co, dtype=torch.float32, device=dev)
        b32_[:o1 - o0] = bias.detach()[o0:o1]
        return b32_

    wp = _CONV_CACHE.get(weight, ("conv", o0, o1, dev.index), make)
    b32 = None if bias is None else _CONV_CACHE.get(bias, ("bias", o0, o1, dev.index), make_bias)
    return wp, b32, co


def conv2d_fused(x, weight, bias=None, act=None, out=None, epilogue=None, split=0, aux1=None, aux2=None):
    """rc_conv2d_ex_fwd: `conv2d_act` for at most 256 filters whose inputs, output and epilogue operands may be channel
    slices of wider tensors (no `.contiguous()` copies, the output written in place into `out`), with the GRU's
    element-wise steps fused: epilogue "mul_from" multiplies the output channels from `split` on by `aux1` (z | r*h from
    one convolution, core/update.py:43-45), "gru" returns (1 - aux1) * aux2 + aux1 * act(conv) (:46).  Inference only."""
    xs = x if isinstance(x, (tuple, list)) else (x,)
    if len(xs) not in (1, 2):
        raise RuntimeError("conv2d_fused: x must be one tensor or a pair")
    if torch.is_grad_enabled() and (any(t.requires_grad for t in xs) or weight.requires_grad):
        raise RuntimeError("conv2d_fused: inference only (no gradient); use torch's convolution for training")
    if act not in _ACTS or epilogue not in _EPIS:
        raise ValueError(f"unknown act {act!r} / epilogue {epilogue!r}")
    bs = [_chw(t, "x") for t in xs]
    B, C1, H, W = xs[0].shape
    Cin = C1 + (xs[1].shape[1] if len(xs) == 2 else 0)
    if len(xs) == 2 and (xs[1].shape[0] != B or xs[1].shape[2:] != xs[0].shape[2:]):
        raise RuntimeError("conv2d_fused: the two inputs must agree in batch and spatial size")
    if weight.dim() != 4 or weight.shape[1] != Cin:
        raise RuntimeError(f"conv2d_fused: weight must be (Cout, {Cin}, KH, KW), got {tuple(weight.shape)}")
    Cout, _, KH, KW = weight.shape
    if Cout > 256:
        raise RuntimeError("conv2d_fused: at most 256 filters per call (conv2d_act slices wider layers)")
    dev = xs[0].device
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        wp, b32, co = _prepared_filters(weight, bias, 0, Cout, dev, st)
        store = 0
        if out is None:
            out = torch.empty((B, co, H, W), dtype=torch.float32, device=dev)
        elif out.shape == (B, Cout, H, W):
            store = Cout  # the zero filters padding Cout to a multiple of 16 are computed but not stored
        elif out.shape != (B, co, H, W):
            raise RuntimeError(f"conv2d_fused: out must be {(B, Cout, H, W)}, got {tuple(out.shape)}")
        obs = _chw(out, "out")
        a1 = a2 = None
        a1bs = a2bs = 0
        if epilogue == "mul_from":
            if aux1 is None or aux1.shape != (B, co - split, H, W):
                raise RuntimeError(f"conv2d_fused: 'mul_from' needs aux1 of shape {(B, co - split, H, W)}")
            a1, a1bs = aux1, _chw(aux1, "aux1")
        elif epilogue == "gru":
            if aux1 is None or aux2 is None or aux1.shape != (B, co, H, W) or aux2.shape != (B, co, H, W):
                raise RuntimeError(f"conv2d_fused: 'gru' needs aux1 (z) and aux2 (h) of shape {(B, co, H, W)}")
            a1, a1bs, a2, a2bs = aux1, _chw(aux1, "aux1"), aux2, _chw(aux2, "aux2")
        rc = _lib.lib().rc_conv2d_ex_fwd(xs[0].data_ptr(), C1, bs[0], xs[1].data_ptr() if len(xs) == 2 else None,
                                         bs[1] if len(xs) == 2 else 0, wp.data_ptr(), None if b32 is None else b32.data_ptr(),
                                         out.data_ptr(), obs, store, B, Cin, co, H, W, KH, KW, _ACTS[act], _EPIS[epilogue], split,
                                         None if a1 is None else a1.data_ptr(), a1bs, None if a2 is None else a2.data_ptr(), a2bs, st)
        _lib.check(rc, "rc_conv2d_ex_fwd")
    return out if out.shape[1] == Cout else out[:, :Cout]


def small_conv_relu(x, weight, bias=None, relu=True):
    """`F.relu(F.conv2d(x, weight, bias, padding=K//2))` for a map with 1..4 channels and K x K filters, K odd <= 7: the
    motion encoder's convf1 on the flow field (reference core/update.py:86,92; small :67,73) as one fp32 CUDA-core launch
    (rc_small_conv_relu_fwd) instead of cuDNN's convolution between two layout conversions.  Inference only."""
    if (not x.is_cuda) or x.dim() != 4 or weight.dim() != 4 or weight.shape[1] != x.shape[1] or weight.shape[2] != weight.shape[3]:
        raise RuntimeError("small_conv_relu: x (B,Cin,H,W) CUDA, weight (Cout,Cin,K,K)")
    if torch.is_grad_enabled() and (x.requires_grad or weight.requires_grad):
        raise RuntimeError("small_conv_relu: inference only (no gradient)")
    x = x.detach().float().contiguous()
    w = weight.detach().float().contiguous()
    b = None if bias is None else bias.detach().float().contiguous()
    B, Cin, H, W = x.shape
    Cout, _, K, _ = w.shape
    with torch.cuda.device(x.device):
        out = torch.empty((B, Cout, H, W), dtype=torch.float32, device=x.device)
        rc = _lib.lib().rc_small_conv_relu_fwd(x.data_ptr(), w.data_ptr(), None if b is None else b.data_ptr(), out.data_ptr(), B, Cin,
                                               Cout, H, W, K, 1 if relu else 0, torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "rc_small_conv_relu_fwd")
    return out


def flow_update(coords1, delta, coords0, flow_out, flow_out2=None):
    """coords1 += delta; flow_out (and flow_out2) = coords1 - coords0, in place and in one launch (rc_flow_update): reference
    core/raft.py:123,131-132.  coords1 / coords0 dense (B,2,H,W) fp32; delta and the outputs may be channel slices."""
    B, _, H, W = coords1.shape
    if not (coords1.is_contiguous() and coords0.is_contiguous() and coords1.dtype == torch.float32 and coords0.dtype == torch.float32):
        raise RuntimeError("flow_update: coords1 / coords0 must be dense fp32 (B,2,H,W)")
    for t, name in ((delta, "delta"), (flow_out, "flow_out")) + (((flow_out2, "flow_out2"),) if flow_out2 is not None else ()):
        if tuple(t.shape) != (B, 2, H, W):
            raise RuntimeError(f"flow_update: {name} must be {(B, 2, H, W)}")
    with torch.cuda.device(coords1.device):
        rc = _lib.lib().rc_flow_update(coords1.data_ptr(), delta.data_ptr(), _chw(delta, "delta"), coords0.data_ptr(), flow_out.data_ptr(),
                                       _chw(flow_out, "flow_out"), None if flow_out2 is None else flow_out2.data_ptr(),
                                       0 if flow_out2 is None else _chw(flow_out2, "flow_out2"), B, H, W,
                                       torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "rc_flow_update")
    return coords1


def conv2d_act(x, weight, bias=None, act=None):
    """`act(F.conv2d(x, weight, bias, padding="same"))` for the stride-1 convolutions of the per-iteration update
    (reference core/update.py:6-14,33-60,84-97,122-125) on tensor cores (rc_conv2d_fwd): filter taps re-use one fp16
    conversion of the input tile, bias and activation are fused, and `x` may be a tuple (a, b) standing for
    `torch.cat([a, b], dim=1)` without the copy.  weight (Cout, Cin, KH, KW) with KH x KW in 1x1, 3x3, 1x5, 5x1;
    act in None / "relu" / "sigmoid" / "tanh".  fp16 operands (TF32-class), fp32 accumulate and output.  Inference only.
    A filter count that is not a multiple of 16 (126 in the motion encoder, 2 in the flow head) is padded with zero
    filters inside and the result is then a channel-sliced view."""
    xs = x if isinstance(x, (tuple, list)) else (x,)
    if len(xs) not in (1, 2) or any((not t.is_cuda) or t.dim() != 4 for t in xs) or not weight.is_cuda:
        raise RuntimeError("conv2d_act: x must be one or two (B,C,H,W) CUDA tensors and weight a CUDA tensor")
    if torch.is_grad_enabled() and (any(t.requires_grad for t in xs) or weight.requires_grad):
        raise RuntimeError("conv2d_act: inference only (no gradient); use torch's convolution for training")
    if act not in _ACTS:
        raise ValueError(f"act must be one of {sorted(k for k in _ACTS if k)}, got {act!r}")
    xs = [t.detach().float().contiguous() for t in xs]
    B, C1, H, W = xs[0].shape
    Cin = C1 + (xs[1].shape[1] if len(xs) == 2 else 0)
    if len(xs) == 2 and (xs[1].shape[0] != B or xs[1].shape[2:] != xs[0].shape[2:]):
        raise RuntimeError("conv2d_act: the two inputs must agree in batch and spatial size")
    if weight.dim() != 4 or weight.shape[1] != Cin:
        raise RuntimeError(f"conv2d_act: weight must be (Cout, {Cin}, KH, KW), got {tuple(weight.shape)}")
    Cout, _, KH, KW = weight.shape
    L = _lib.lib()
    dev = xs[0].device
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        outs = []
        for o0 in range(0, Cout, 256):  # wider layers (the mask head's 576 channels) go slice by slice
            o1 = min(Cout, o0 + 256)
            wp, b32, co = _prepared_filters(weight, bias, o0, o1, dev, st)
            out = torch.empty((B, co, H, W), dtype=torch.float32, device=dev)
            rc = L.rc_conv2d_fwd(xs[0].data_ptr(), C1, xs[1].data_ptr() if len(xs) == 2 else None, wp.data_ptr(),
                                 None if bias is None else b32.data_ptr(), out.data_ptr(), B, Cin, co, H, W, KH, KW, _ACTS[act], st)
            _lib.check(rc, "rc_conv2d_fwd")
            outs.append(out if co == o1 - o0 else out[:, :o1 - o0])
    return outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)
