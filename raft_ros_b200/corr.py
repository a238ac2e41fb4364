"""Drop-in replacements for the reference's correlation operators.

Same names, constructor arguments, attributes and call convention as the reference's
`core/corr.py` (CorrBlock :12-60, AlternateCorrBlock :63-91), so `core/raft.py:104-107,124`
works unchanged with `from raft_ros_b200.corr import CorrBlock, AlternateCorrBlock`.

Host code here is PyTorch only for device memory, streams and autograd plumbing; all the
arithmetic runs in hand-written sm_100a kernels behind the C ABI of libraftcorr.so
(include/raftcorr.h).  There is no CPU or eager fallback: CPU tensors raise.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import torch

from . import _lib

_PRECISIONS = {"fp32": _lib.RC_PREC_FP32, "bf16x3": _lib.RC_PREC_BF16X3, "bf16": _lib.RC_PREC_BF16,
               "bf16_vol16": _lib.RC_PREC_BF16_VOL16, "fp16_fp8": _lib.RC_PREC_FP16_FP8}

#: arithmetic of the all-pairs contraction when the caller does not say: "bf16x3" is the
#: fp32-class tensor-core path (error-compensated bf16 operands, ~1e-5 normwise), "bf16" the
#: single-pass fast path, "fp32" plain fp32 FMA on CUDA cores, "bf16_vol16" = "bf16" with the
#: pyramid stored as fp16 (inference with radius 3 or 4 and <= 4 levels only; half the HBM traffic).
DEFAULT_PRECISION = os.environ.get("RAFT_CORR_PRECISION", "bf16x3")


def _precision_code(precision: Optional[str]) -> int:
    name = DEFAULT_PRECISION if precision is None else precision
    if name not in _PRECISIONS:
        raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {name!r}")
    return _PRECISIONS[name]


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(t: torch.Tensor, name: str) -> None:
    # wording of the reference's CHECK_CUDA (alt_cuda_corr/correlation.cpp:19)
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().float().contiguous()


def level_shapes(h: int, w: int, num_levels: int):
    shapes = [(h >> l, w >> l) for l in range(num_levels)]
    if shapes[-1][0] < 1 or shapes[-1][1] < 1:
        raise RuntimeError(f"level {num_levels - 1} of a {h}x{w} feature grid is empty")
    return shapes


# --------------------------------------------------------------------------- raw ops
def build_pyramid_raw(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int, precision: Optional[str] = None,
                      scale: float = 0.0) -> List[torch.Tensor]:
    """rc_corr_build_pyramid: (B,D,h,w) x2 -> list of (B*h*w, 1, h_l, w_l) fp32, one launch."""
    _require_cuda(fmap1, "fmap1")
    _require_cuda(fmap2, "fmap2")
    if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
        raise RuntimeError(f"fmap1/fmap2 must both be (B,D,h,w); got {tuple(fmap1.shape)} and {tuple(fmap2.shape)}")
    if fmap1.device != fmap2.device:
        raise RuntimeError("fmap1 and fmap2 must be on the same device")
    f1, f2 = _f32c(fmap1), _f32c(fmap2)
    B, D, h, w = f1.shape
    shapes = level_shapes(h, w, num_levels)
    prec = _precision_code(precision)
    L = _lib.lib()
    with torch.cuda.device(f1.device):
        # one allocation for the whole pyramid; levels are views into it
        sizes = [B * h * w * hl * wl for hl, wl in shapes]
        # keep every level 256-byte aligned (TMA stores want 16 B; be generous)
        offs, total = [], 0
        for s in sizes:
            offs.append(total)
            total += (s + 63) // 64 * 64
        flat = torch.empty(total, dtype=torch.float32, device=f1.device)
        levels = [flat[o:o + s].view(B * h * w, 1, hl, wl) for o, s, (hl, wl) in zip(offs, sizes, shapes)]
        ws_bytes = int(L.rc_corr_build_workspace_bytes(B, D, h, w, prec))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=f1.device) if ws_bytes else None
        rc = L.rc_corr_build_pyramid(f1.data_ptr(), f2.data_ptr(), B, D, h, w,
                                     _lib.ptr_array([lv.data_ptr() for lv in levels]), num_levels, prec,
                                     float(scale), ws.data_ptr() if ws is not None else None, ws_bytes, _stream())
        _lib.check(rc, "rc_corr_build_pyramid")
        if ws is not None:
            ws.record_stream(torch.cuda.current_stream())
    return levels


_TILE_GROUP = None


def _tile_group() -> int:
    global _TILE_GROUP
    if _TILE_GROUP is None:
        _TILE_GROUP = int(_lib.lib().rc_corr_tile_group())
    return _TILE_GROUP


def tiled_level_elems(B: int, N: int, hl: int, wl: int, vol16: bool = False) -> int:
    """elements of one level in the tiled layout (include/raftcorr.h: rc_corr_tiled_level_elems /
    rc_corr_tiled16_level_elems): [B][ceil(N/G)][hp/2][wp/cw][G pixels][4 chunks of 16 B], G = rc_corr_tile_group()
    = 32, cw = 8 floats or 16 halves per tile row; chunk k = 2*(row & 1) + (column half) of pixel u sits at
    position k ^ ((u >> 1) & 3).  Only the size is computed here; the addressing is in the header."""
    G = _tile_group()
    cw = 16 if vol16 else 8
    return B * ((N + G - 1) // G) * G * ((hl + 1) & ~1) * ((wl + cw - 1) & ~(cw - 1))


def build_pyramid_tiled_raw(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int,
                            precision: Optional[str] = None, scale: float = 0.0) -> List[torch.Tensor]:
    """rc_corr_build_pyramid_tiled: (B,D,h,w) x2 -> list of flat levels in the TILED layout (fp32, or
    fp16 for precision "bf16_vol16")."""
    _require_cuda(fmap1, "fmap1")
    _require_cuda(fmap2, "fmap2")
    if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
        raise RuntimeError(f"fmap1/fmap2 must both be (B,D,h,w); got {tuple(fmap1.shape)} and {tuple(fmap2.shape)}")
    if fmap1.device != fmap2.device:
        raise RuntimeError("fmap1 and fmap2 must be on the same device")
    f1, f2 = _f32c(fmap1), _f32c(fmap2)
    B, D, h, w = f1.shape
    shapes = level_shapes(h, w, num_levels)
    prec = _precision_code(precision)
    L = _lib.lib()
    with torch.cuda.device(f1.device):
        vol16 = prec == _lib.RC_PREC_BF16_VOL16
        sizes = [tiled_level_elems(B, h * w, hl, wl, vol16) for hl, wl in shapes]  # multiples of 64 bytes
        flat = torch.empty(sum(sizes), dtype=torch.float16 if vol16 else torch.float32, device=f1.device)
        levels, o = [], 0
        for s_ in sizes:
            levels.append(flat[o:o + s_])
            o += s_
        ws_bytes = int(L.rc_corr_build_workspace_bytes(B, D, h, w, prec))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=f1.device)
        rc = L.rc_corr_build_pyramid_tiled(f1.data_ptr(), f2.data_ptr(), B, D, h, w,
                                           _lib.ptr_array([lv.data_ptr() for lv in levels]), num_levels, prec,
                                           float(scale), ws.data_ptr(), ws_bytes, _stream())
        _lib.check(rc, "rc_corr_build_pyramid_tiled")
        ws.record_stream(torch.cuda.current_stream())
    return levels


def relayout_raw(src: torch.Tensor, B: int, N: int, hl: int, wl: int, to_tiled: bool, vol16: bool = False) -> torch.Tensor:
    """rc_corr_relayout_level / rc_corr_relayout16_level: one level, row layout (B*N, 1, hl, wl) fp32 <->
    tiled layout (flat; fp16 when vol16)."""
    _require_cuda(src, "level")
    if to_tiled or not vol16:
        s = src if (src.dtype == torch.float32 and src.is_contiguous()) else _f32c(src)
    else:
        s = src if (src.dtype == torch.float16 and src.is_contiguous()) else src.detach().half().contiguous()
    with torch.cuda.device(s.device):
        if to_tiled:
            dst = torch.empty(tiled_level_elems(B, N, hl, wl, vol16), dtype=torch.float16 if vol16 else torch.float32,
                              device=s.device)
        else:
            dst = torch.empty((B * N, 1, hl, wl), dtype=torch.float32, device=s.device)
        fn = _lib.lib().rc_corr_relayout16_level if vol16 else _lib.lib().rc_corr_relayout_level
        rc = fn(s.data_ptr(), dst.data_ptr(), B, N, hl, wl, 1 if to_tiled else 0, _stream())
        _lib.check(rc, "rc_corr_relayout16_level" if vol16 else "rc_corr_relayout_level")
    return dst


def lookup_tiled_raw(levels: Sequence[torch.Tensor], coords: torch.Tensor, radius: int) -> torch.Tensor:
    """rc_corr_lookup_tiled_fwd: lookup_raw on levels kept in the tiled layout."""
    _require_cuda(coords, "coords")
    if coords.dim() != 4 or coords.shape[1] != 2:
        raise RuntimeError(f"coords must be (B,2,h,w); got {tuple(coords.shape)}")
    c = _f32c(coords)
    B, _, h, w = c.shape
    nl = len(levels)
    shapes = level_shapes(h, w, nl)
    vol16 = levels[0].dtype == torch.float16
    for l, (t, (hl, wl)) in enumerate(zip(levels, shapes)):
        if (t.numel() != tiled_level_elems(B, h * w, hl, wl, vol16) or t.dtype != levels[0].dtype
                or t.dtype not in (torch.float32, torch.float16) or not t.is_contiguous()):
            raise RuntimeError(f"tiled level {l} does not match a {B}x{h}x{w} problem")
    rd = 2 * radius + 1
    with torch.cuda.device(c.device):
        out = torch.empty((B, nl * rd * rd, h, w), dtype=torch.float32, device=c.device)
        fn = _lib.lib().rc_corr_lookup_tiled16_fwd if vol16 else _lib.lib().rc_corr_lookup_tiled_fwd
        rc = fn(_lib.ptr_array([t.data_ptr() for t in levels]), c.data_ptr(), out.data_ptr(), B, h, w, nl, radius, _stream())
        _lib.check(rc, "rc_corr_lookup_tiled16_fwd" if vol16 else "rc_corr_lookup_tiled_fwd")
    return out


class _TiledLookup:
    """The per-iteration call of an inference CorrBlock with everything that does not change between
    iterations resolved once (level pointer table, entry point, geometry): the host side of a lookup
    is then one shape check, one allocation and one C call.  At batch 1 the kernel takes ~8 us, so
    this is what the call costs."""

    def __init__(self, levels, B, h, w, radius):
        self.levels = levels  # keeps the storage alive
        self.B, self.h, self.w, self.nl, self.radius = B, h, w, len(levels), radius
        self.shape = (B, 2, h, w)
        self.out_shape = (B, len(levels) * (2 * radius + 1) ** 2, h, w)
        self.device = levels[0].device
        vol16 = levels[0].dtype == torch.float16
        lib = _lib.lib()
        self.fn = lib.rc_corr_lookup_tiled16_fwd if vol16 else lib.rc_corr_lookup_tiled_fwd
        self.name = "rc_corr_lookup_tiled16_fwd" if vol16 else "rc_corr_lookup_tiled_fwd"
        self.ptrs = _lib.ptr_array([t.data_ptr() for t in levels])

    def __call__(self, coords):
        if (coords.dtype is not torch.float32 or coords.device != self.device or not coords.is_contiguous()
                or coords.requires_grad or tuple(coords.shape) != self.shape
                or torch.cuda.current_device() != self.device.index):
            return lookup_tiled_raw(self.levels, coords, self.radius)  # the checked, general path
        out = torch.empty(self.out_shape, dtype=torch.float32, device=self.device)
        rc = self.fn(self.ptrs, coords.data_ptr(), out.data_ptr(), self.B, self.h, self.w, self.nl, self.radius,
                     torch.cuda.current_stream().cuda_stream)
        if rc:
            _lib.check(rc, self.name)
        return out


def lookup_raw(levels: Sequence[torch.Tensor], coords: torch.Tensor, radius: int) -> torch.Tensor:
    """rc_corr_lookup_fwd: (B,2,h,w) coords -> (B, L*(2r+1)^2, h, w) fp32, one launch."""
    _require_cuda(coords, "coords")
    if coords.dim() != 4 or coords.shape[1] != 2:
        raise RuntimeError(f"coords must be (B,2,h,w); got {tuple(coords.shape)}")
    c = _f32c(coords)
    B, _, h, w = c.shape
    nl = len(levels)
    shapes = level_shapes(h, w, nl)
    lv = []
    for l, (t, (hl, wl)) in enumerate(zip(levels, shapes)):
        _require_cuda(t, f"corr_pyramid[{l}]")
        if t.numel() != B * h * w * hl * wl:
            raise RuntimeError(f"corr_pyramid[{l}] has {t.numel()} elements, expected {B * h * w * hl * wl}")
        lv.append(t if (t.dtype == torch.float32 and t.is_contiguous()) else _f32c(t))
    rd = 2 * radius + 1
    with torch.cuda.device(c.device):
        out = torch.empty((B, nl * rd * rd, h, w), dtype=torch.float32, device=c.device)
        rc = _lib.lib().rc_corr_lookup_fwd(_lib.ptr_array([t.data_ptr() for t in lv]), c.data_ptr(), out.data_ptr(),
                                           B, h, w, nl, radius, _stream())
        _lib.check(rc, "rc_corr_lookup_fwd")
    return out


def lookup_bwd_raw(grad_out: torch.Tensor, coords: torch.Tensor, grad_levels: Sequence[torch.Tensor], radius: int) -> None:
    """rc_corr_lookup_bwd: accumulate d(out)/d(pyramid) into grad_levels (deterministic)."""
    g, c = _f32c(grad_out), _f32c(coords)
    B, _, h, w = c.shape
    with torch.cuda.device(c.device):
        rc = _lib.lib().rc_corr_lookup_bwd(g.data_ptr(), c.data_ptr(), _lib.ptr_array([t.data_ptr() for t in grad_levels]),
                                           B, h, w, len(grad_levels), radius, _stream())
        _lib.check(rc, "rc_corr_lookup_bwd")


def build_bwd_raw(grad_levels: Sequence[torch.Tensor], fmap1: torch.Tensor, fmap2: torch.Tensor, scale: float = 0.0,
                  precision: Optional[str] = None):
    """rc_corr_build_bwd: grad pyramid -> (dfmap1, dfmap2), both (B,D,h,w)."""
    f1, f2 = _f32c(fmap1), _f32c(fmap2)
    B, D, h, w = f1.shape
    prec = _precision_code(precision)
    L = _lib.lib()
    with torch.cuda.device(f1.device):
        df1, df2 = torch.empty_like(f1), torch.empty_like(f2)
        ws_bytes = int(L.rc_corr_build_bwd_workspace_bytes(B, D, h, w, prec))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=f1.device)
        rc = L.rc_corr_build_bwd(_lib.ptr_array([t.data_ptr() for t in grad_levels]), f1.data_ptr(), f2.data_ptr(),
                                 df1.data_ptr(), df2.data_ptr(), B, D, h, w, len(grad_levels), prec, float(scale),
                                 ws.data_ptr(), ws_bytes, _stream())
        _lib.check(rc, "rc_corr_build_bwd")
        ws.record_stream(torch.cuda.current_stream())
    return df1, df2


class PreparedFeatures:
    """Feature maps in the layout the tensor-core build reads: K-major bf16 `hi` (+ `lo` residual, x ~= hi + lo),
    shaped (Bt, h, w, C) with C a multiple of 64.  Produced by `encoder_head`; consumed by `CorrBlock.from_prepared`."""

    def __init__(self, hi, lo, h, w):
        self.hi, self.lo, self.h, self.w = hi, lo, h, w

    def __getitem__(self, sl):
        return PreparedFeatures(self.hi[sl], None if self.lo is None else self.lo[sl], self.h, self.w)

    @property
    def batch(self):
        return self.hi.shape[0]

    @property
    def channels(self):
        return self.hi.shape[3]

    def float(self):
        """(Bt, C, h, w) fp32: hi + lo back in the reference's layout (tests, debugging)"""
        v = self.hi.float() if self.lo is None else self.hi.float() + self.lo.float()
        return v.permute(0, 3, 1, 2).contiguous()


class _WeightCache:
    """Converted weights per weight TENSOR OBJECT (weak: the entry dies with the tensor) and variant; an in-place
    update of the tensor (its `_version` moves) re-converts.  Never keyed on data_ptr: a freed tensor's address is
    handed out again."""

    def __init__(self):
        self._d = {}  # id(tensor) -> (weak reference to the tensor, {variant: (version, converted)})

    def get(self, weight, variant, make):
        import weakref
        key = id(weight)
        slot = self._d.get(key)
        if slot is None or slot[0]() is not weight:  # (an id can be reused once its tensor is gone)
            d = self._d
            slot = d[key] = (weakref.ref(weight, lambda _r, k=key: d.pop(k, None)), {})
        per = slot[1]
        ent = per.get(variant)
        if ent is None or ent[0] != weight._version:
            ent = per[variant] = (weight._version, make())
        return ent[1]


_HEAD_CACHE = _WeightCache()
_CONVC1_CACHE = _WeightCache()


def encoder_head(x, weight, bias, with_lo=True):
    """The feature encoder's output layer -- `conv2`, a 1x1 convolution (reference core/extractor.py:144,184; small
    :217,256) -- emitting the operand maps of the build directly (rc_fnet_conv1x1_prep): the fp32 (B, 256, h, w)
    feature maps of core/raft.py:99-103 are never written and the build's layout prepass never runs.

    x: (Bt, Cin, h, w) fp32, the input of conv2 (Bt = 2B when both images go through the encoder as one batch,
    core/raft.py:100); weight: (Cout, Cin[, 1, 1]), bias: (Cout,).  Cin a multiple of 16 (<= 128), Cout a multiple of
    64 (<= 256).  Error-compensated bf16 products with fp32 accumulation: fp32-class, like precision "bf16x3".
    The converted weights are cached per weight tensor object (and its version)."""
    _require_cuda(x, "x")
    _require_cuda(weight, "weight")
    _require_cuda(bias, "bias")
    weight_obj = weight
    if weight.dim() == 4 and weight.shape[2:] == (1, 1):
        weight = weight[:, :, 0, 0]
    if x.dim() != 4 or weight.dim() != 2 or weight.shape[1] != x.shape[1] or bias.shape != (weight.shape[0],):
        raise RuntimeError(f"expected x (Bt,Cin,h,w), weight (Cout,Cin[,1,1]), bias (Cout,); got {tuple(x.shape)}, "
                           f"{tuple(weight.shape)}, {tuple(bias.shape)}")
    xs = _f32c(x)
    Bt, Cin, h, w = xs.shape
    Cout = weight.shape[0]
    L = _lib.lib()
    with torch.cuda.device(xs.device):
        def make():
            nbytes = int(L.rc_fnet_conv1x1_weight_bytes(Cout, Cin))
            wp = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=xs.device)
            w32 = _f32c(weight)
            _lib.check(L.rc_fnet_conv1x1_prep_weight(w32.data_ptr(), wp.data_ptr(), Cout, Cin, _stream()), "rc_fnet_conv1x1_prep_weight")
            return wp

        ent = _HEAD_CACHE.get(weight_obj, ("head", xs.device.index), make)
        b32 = _f32c(bias)
        hi = torch.empty((Bt, h, w, Cout), dtype=torch.bfloat16, device=xs.device)
        lo = torch.empty_like(hi) if with_lo else None
        rc = L.rc_fnet_conv1x1_prep(xs.data_ptr(), ent.data_ptr(), b32.data_ptr(), hi.data_ptr(), lo.data_ptr() if with_lo else None,
                                    Bt, Cin, Cout, h, w, _stream())
        _lib.check(rc, "rc_fnet_conv1x1_prep")
    return PreparedFeatures(hi, lo, h, w)


# --------------------------------------------------------------------------- autograd
class _GradState:
    """Per-CorrBlock bookkeeping of one backward pass.

    The reference lets autograd add up the K per-iteration gradients of the pyramid: every
    grid_sample backward allocates a zeros_like(level) and the engine sums them (K zero-fills and
    K-1 adds over a volume-sized tensor).  Here all K lookups of a CorrBlock accumulate straight
    into ONE gradient pyramid (rc_corr_lookup_bwd is `+=` and deterministic), and the build's
    backward consumes it once.  The autograd edge between the lookups and the build is a
    one-element `token` tensor, so the engine still orders build.backward after every lookup.
    """

    def __init__(self):
        self.levels = None        # detached pyramid in the row layout, read by the lookups ...
        self.tiled = None         # ... or in the tiled layout (then `levels` is only made on demand)
        self.shapes = None        # [(B*N, 1, h_l, w_l)] of the row layout
        self.geom = None          # (B, h, w)
        self.device = None
        self.grad_levels = None   # allocated lazily by the first lookup backward of a pass


def _alloc_like_levels(levels):
    return _alloc_levels([tuple(lv.shape) for lv in levels], levels[0].device)


def _alloc_levels(shapes, device):
    """zeroed gradient pyramid (row layout), one allocation"""
    sizes = [int(torch.Size(sh).numel()) for sh in shapes]
    offs, total = [], 0
    for s_ in sizes:
        offs.append(total)
        total += (s_ + 63) // 64 * 64
    flat = torch.zeros(total, dtype=torch.float32, device=device)
    return [flat[o:o + n].view(sh) for o, n, sh in zip(offs, sizes, shapes)]


def _use_tiled(num_levels, radius, precision) -> bool:
    return (radius in (3, 4) and 1 <= num_levels <= 4 and _precision_code(precision) != _lib.RC_PREC_FP32
            and os.environ.get("RAFT_CORR_LAYOUT", "tiled") != "rows")


class _BuildPyramidFn(torch.autograd.Function):
    """outputs: (token, level_0, ..., level_{L-1}) in the row layout, or (token,) alone when the
    pyramid is kept in the tiled layout (`tiled=True`).  Gradients may arrive through the levels
    (generic autograd use of `corr_pyramid`) and/or through `state.grad_levels` (the lookups and,
    in the tiled case, `_UntileFn`)."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, num_levels, precision, state, tiled):
        ctx.save_for_backward(fmap1, fmap2)
        ctx.state = state
        ctx.precision = precision
        B, _, h, w = fmap1.shape
        state.geom, state.device = (B, h, w), fmap1.device
        state.shapes = [(B * h * w, 1, hl, wl) for hl, wl in level_shapes(h, w, num_levels)]
        token = torch.zeros(1, dtype=torch.float32, device=fmap1.device)
        if tiled:
            state.tiled = build_pyramid_tiled_raw(fmap1, fmap2, num_levels, precision)
            return (token,)
        levels = build_pyramid_raw(fmap1, fmap2, num_levels, precision)
        state.levels = [lv.detach() for lv in levels]
        return (token,) + tuple(levels)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_token, *grad_levels):
        fmap1, fmap2 = ctx.saved_tensors
        state = ctx.state
        gl = state.grad_levels
        state.grad_levels = None  # a later backward pass (retain_graph) starts from zero again
        for l, g in enumerate(grad_levels):
            if g is None:
                continue
            if gl is None:
                gl = _alloc_levels(state.shapes, state.device)
            gl[l].add_(g.view(gl[l].shape))
        if gl is None:
            return torch.zeros_like(fmap1), torch.zeros_like(fmap2), None, None, None, None
        df1, df2 = build_bwd_raw(gl, fmap1, fmap2, precision=ctx.precision)
        return df1.to(fmap1.dtype), df2.to(fmap2.dtype), None, None, None, None


class _UntileFn(torch.autograd.Function):
    """`corr_pyramid` of a block that keeps its pyramid tiled: row-layout copies of the levels whose
    gradients flow into the block's single gradient pyramid (and from there into the build)."""

    @staticmethod
    def forward(ctx, token, state):
        ctx.state = state
        B, h, w = state.geom
        return tuple(relayout_raw(t, B, h * w, sh[2], sh[3], to_tiled=False) for t, sh in zip(state.tiled, state.shapes))

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, *grad_levels):
        state = ctx.state
        for l, g in enumerate(grad_levels):
            if g is None:
                continue
            if state.grad_levels is None:
                state.grad_levels = _alloc_levels(state.shapes, state.device)
            state.grad_levels[l].add_(g.view(state.grad_levels[l].shape))
        return torch.zeros(1, dtype=torch.float32, device=state.device), None


class _LookupFn(torch.autograd.Function):
    """Generic path: the pyramid levels themselves are the differentiable inputs."""

    @staticmethod
    def forward(ctx, coords, radius, *levels):
        ctx.radius = radius
        ctx.save_for_backward(coords)
        ctx.level_shapes = [tuple(lv.shape) for lv in levels]
        return lookup_raw(levels, coords, radius)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_out):
        (coords,) = ctx.saved_tensors
        grads = [torch.zeros(s, dtype=torch.float32, device=grad_out.device) for s in ctx.level_shapes]
        lookup_bwd_raw(grad_out, coords, grads, ctx.radius)
        # coords carry no gradient (they are detached at core/raft.py:123)
        return (None, None) + tuple(grads)


class _LookupTokenFn(torch.autograd.Function):
    """CorrBlock's own path: accumulate into the block's single gradient pyramid."""

    @staticmethod
    def forward(ctx, coords, token, state, radius):
        ctx.radius = radius
        ctx.state = state
        ctx.save_for_backward(coords)
        if state.tiled is not None:
            return lookup_tiled_raw(state.tiled, coords, radius)
        return lookup_raw(state.levels, coords, radius)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_out):
        (coords,) = ctx.saved_tensors
        state = ctx.state
        if state.grad_levels is None:
            state.grad_levels = _alloc_levels(state.shapes, state.device)
        lookup_bwd_raw(grad_out, coords, state.grad_levels, ctx.radius)
        return None, torch.zeros(1, dtype=torch.float32, device=grad_out.device), None, None


# --------------------------------------------------------------------------- public classes
class CorrBlock:
    """All-pairs correlation volume + pyramid + windowed lookup (reference core/corr.py:12-60).

    `corr_pyramid` is a list of `num_levels` fp32 tensors shaped (B*h*w, 1, h>>l, w>>l), as in
    the reference; it is produced by ONE kernel launch (contraction, 1/sqrt(D) scale and every
    pooling level fused, each level written once).  `precision` is an extension: see
    DEFAULT_PRECISION.

    With radius 3 or 4, <= 4 levels and a tensor-core precision the block keeps the
    pyramid in the TILED layout (include/raftcorr.h) that the lookup kernel reads with 35 % less
    DRAM traffic; `corr_pyramid` then converts to the reference's row layout on first access (a
    copy: writing into it does not change what the lookups read -- assign `corr_pyramid` to
    replace the pyramid).  RAFT_CORR_LAYOUT=rows keeps the row layout throughout.
    """

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, precision=None):
        self.num_levels = num_levels
        self.radius = radius
        self.precision = precision
        self._state = None
        self._token = None
        self._tiled = None      # tiled levels (inference fast path) or None
        self._rows = None       # row-layout levels (reference attribute), lazily made from _tiled
        self._geom = None
        tiled = _use_tiled(num_levels, radius, precision)
        if _precision_code(precision) == _lib.RC_PREC_BF16_VOL16:
            # the fp16 volume exists only where the fused tiled kernels do
            if not tiled or (torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad)):
                raise RuntimeError("precision 'bf16_vol16' is an inference mode for radius 3 or 4 and at most 4 levels "
                                   "(tiled layout); use 'bf16' or 'bf16x3' here")
        if _precision_code(precision) == _lib.RC_PREC_FP16_FP8 and torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
            raise RuntimeError("precision 'fp16_fp8' is a forward-only mode; train with 'bf16x3' or 'bf16'")
        if torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
            self._state = _GradState()
            out = _BuildPyramidFn.apply(fmap1, fmap2, num_levels, precision, self._state, tiled)
            self._token = out[0]
            self._rows = None if tiled else list(out[1:])
        elif tiled:
            self._tiled = build_pyramid_tiled_raw(fmap1, fmap2, num_levels, precision)
            self._geom = (fmap1.shape[0], fmap1.shape[2], fmap1.shape[3])
            self._fast = _TiledLookup(self._tiled, *self._geom, radius)
        else:
            self._rows = build_pyramid_raw(fmap1, fmap2, num_levels, precision)

    @property
    def corr_pyramid(self):
        if self._rows is None and self._token is not None:
            # training block with a tiled pyramid: differentiable row-layout copies
            self._rows = list(_UntileFn.apply(self._token, self._state))
        if self._rows is None:
            B, h, w = self._geom
            self._rows = [relayout_raw(t, B, h * w, hl, wl, to_tiled=False, vol16=t.dtype == torch.float16)
                          for t, (hl, wl) in zip(self._tiled, level_shapes(h, w, self.num_levels))]
        return self._rows

    @corr_pyramid.setter
    def corr_pyramid(self, levels):
        self._rows = list(levels)
        self._tiled = None  # the lookups follow the assigned pyramid
        self._token = None

    def __call__(self, coords):
        if coords.requires_grad and torch.is_grad_enabled():
            # the reference's grid_sample would propagate to coords; its one caller detaches them first
            raise RuntimeError("CorrBlock: gradients w.r.t. coords are not provided (detach them, as core/raft.py:123 does)")
        if self._token is not None:
            if torch.is_grad_enabled():
                return _LookupTokenFn.apply(coords, self._token, self._state, self.radius)
            st = self._state  # a block built for training, looked up under no_grad
            if st.tiled is not None:
                return lookup_tiled_raw(st.tiled, coords, self.radius)
            return lookup_raw(st.levels, coords, self.radius)
        if self._tiled is not None:
            return self._fast(coords)
        if torch.is_grad_enabled() and any(lv.requires_grad for lv in self._rows):
            # an assigned pyramid that is part of someone else's graph: generic differentiable path
            return _LookupFn.apply(coords, self.radius, *self._rows)
        return lookup_raw(self._rows, coords, self.radius)

    @classmethod
    def from_prepared(cls, prep1, prep2=None, num_levels=4, radius=4, precision=None):
        """An inference CorrBlock from `encoder_head` output (rc_corr_build_pyramid_tiled_prepared): `prep1`, `prep2`
        are the two images' PreparedFeatures; with `prep2` omitted, `prep1` holds both (first half = image 1, as the
        encoder returns them for `fnet([image1, image2])`, core/raft.py:100).  Same object as CorrBlock(fmap1, fmap2)
        otherwise; precision "bf16x3" needs the `lo` parts."""
        if prep2 is None:
            if prep1.batch % 2:
                raise RuntimeError("from_prepared: one PreparedFeatures must hold an even number of images (2B)")
            prep1, prep2 = prep1[:prep1.batch // 2], prep1[prep1.batch // 2:]
        if prep1.hi.shape != prep2.hi.shape or prep1.hi.device != prep2.hi.device:
            raise RuntimeError("from_prepared: the two images' maps must have one shape and device")
        prec = _precision_code(precision)
        if not _use_tiled(num_levels, radius, precision):
            raise RuntimeError("from_prepared builds the tiled inference pyramid: radius 3 or 4, at most 4 levels, a tensor-core precision")
        if prec == _lib.RC_PREC_FP16_FP8:
            raise RuntimeError("from_prepared: the prepared maps are bf16 hi/lo; precision 'fp16_fp8' converts the fp32 maps itself (CorrBlock(fmap1, fmap2, precision='fp16_fp8'))")
        if prec == _lib.RC_PREC_BF16X3 and (prep1.lo is None or prep2.lo is None):
            raise RuntimeError("precision 'bf16x3' needs the lo parts (encoder_head(..., with_lo=True))")
        B, h, w, D = prep1.hi.shape
        shapes = level_shapes(h, w, num_levels)
        self = cls.__new__(cls)
        self.num_levels, self.radius, self.precision = num_levels, radius, precision
        self._state = self._token = self._rows = None
        dev = prep1.hi.device
        with torch.cuda.device(dev):
            vol16 = prec == _lib.RC_PREC_BF16_VOL16
            sizes = [tiled_level_elems(B, h * w, hl, wl, vol16) for hl, wl in shapes]
            flat = torch.empty(sum(sizes), dtype=torch.float16 if vol16 else torch.float32, device=dev)
            levels, o = [], 0
            for s_ in sizes:
                levels.append(flat[o:o + s_])
                o += s_
            ops = [prep1.hi, prep1.lo, prep2.hi, prep2.lo]
            ops = [None if t is None else (t if t.is_contiguous() else t.contiguous()) for t in ops]
            rc = _lib.lib().rc_corr_build_pyramid_tiled_prepared(
                ops[0].data_ptr(), None if ops[1] is None else ops[1].data_ptr(), ops[2].data_ptr(),
                None if ops[3] is None else ops[3].data_ptr(), B, D, h, w, _lib.ptr_array([lv.data_ptr() for lv in levels]),
                num_levels, prec, 0.0, _stream())
            _lib.check(rc, "rc_corr_build_pyramid_tiled_prepared")
            for t in ops:
                if t is not None:
                    t.record_stream(torch.cuda.current_stream())
        self._tiled = levels
        self._geom = (B, h, w)
        self._fast = _TiledLookup(self._tiled, B, h, w, radius)
        return self

    def lookup_convc1(self, coords, weight, bias, operands="fp16"):
        """`F.relu(F.conv2d(self(coords), weight, bias))` in ONE launch (rc_corr_lookup_convc1_fwd): the lookup
        fused into its only consumer, the motion encoder's 1x1 convolution convc1 + ReLU (reference
        core/update.py:83,90; raft-small :66,72), so the (B, L*(2r+1)^2, h, w) lookup result is never
        written to global memory.  Extension: the reference has no such call; `self(coords)` stays.

        weight: (Cout, L*(2r+1)^2) or (Cout, L*(2r+1)^2, 1, 1) fp32, bias: (Cout,) fp32; Cout a multiple of 16, <= 256.
        operands: "fp16" (taps and weights rounded to fp16, fp32 accumulate -- what the reference's
        --mixed_precision autocast does to this convolution) or "bf16".  Returns (B, Cout, h, w) fp32.
        Inference blocks on the tiled layout only (radius 3 or 4, <= 4 levels, a tensor-core precision).
        The converted weights are cached per weight tensor object (and its version) and operand type."""
        if self._tiled is None:
            raise RuntimeError("lookup_convc1 needs an inference CorrBlock on the tiled layout "
                               "(radius 3 or 4, at most 4 levels, precision bf16x3 / bf16 / bf16_vol16, no gradient)")
        if operands not in ("fp16", "bf16"):
            raise ValueError(f"operands must be 'fp16' or 'bf16', got {operands!r}")
        _require_cuda(coords, "coords")
        _require_cuda(weight, "weight")
        _require_cuda(bias, "bias")
        B, h, w = self._geom
        nl, r = self.num_levels, self.radius
        K = nl * (2 * r + 1) ** 2
        weight_obj = weight
        if weight.dim() == 4 and weight.shape[2:] == (1, 1):
            weight = weight[:, :, 0, 0]
        if weight.dim() != 2 or weight.shape[1] != K or bias.shape != (weight.shape[0],):
            raise RuntimeError(f"expected weight (Cout, {K}[, 1, 1]) and bias (Cout,); got {tuple(weight.shape)}, {tuple(bias.shape)}")
        Cout = weight.shape[0]
        if tuple(coords.shape) != (B, 2, h, w):
            raise RuntimeError(f"coords must be {(B, 2, h, w)}, got {tuple(coords.shape)}")
        c = _f32c(coords)
        L = _lib.lib()
        bf = 1 if operands == "bf16" else 0
        with torch.cuda.device(c.device):
            def make():
                nbytes = int(L.rc_corr_convc1_weight_bytes(Cout, nl, r))
                if nbytes == 0:
                    raise RuntimeError(f"lookup_convc1: unsupported Cout={Cout}, levels={nl}, radius={r}")
                wp_ = torch.empty(nbytes, dtype=torch.uint8, device=c.device)
                w32 = _f32c(weight)
                _lib.check(L.rc_corr_convc1_prep_weight(w32.data_ptr(), wp_.data_ptr(), Cout, nl, r, bf, _stream()),
                           "rc_corr_convc1_prep_weight")
                return wp_

            wp = _CONVC1_CACHE.get(weight_obj, (bf, nl, r, c.device.index), make)
            b32 = _f32c(bias)
            out = torch.empty((B, Cout, h, w), dtype=torch.float32, device=c.device)
            vol16 = 1 if self._tiled[0].dtype == torch.float16 else 0
            rc = L.rc_corr_lookup_convc1_fwd(self._fast.ptrs, vol16, c.data_ptr(), wp.data_ptr(), b32.data_ptr(), out.data_ptr(),
                                             B, h, w, nl, r, Cout, bf, _stream())
            _lib.check(rc, "rc_corr_lookup_convc1_fwd")
        return out

    @staticmethod
    def corr(fmap1, fmap2, precision=None):
        """(B,D,h,w) x2 -> (B,h,w,1,h,w): fmap1^T fmap2 / sqrt(D) (reference core/corr.py:52-60)."""
        B, _, h, w = fmap1.shape
        if torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
            _, vol = _BuildPyramidFn.apply(fmap1, fmap2, 1, precision, _GradState(), False)
        else:
            (vol,) = build_pyramid_raw(fmap1, fmap2, 1, precision)
        return vol.view(B, h, w, 1, h, w)


class AlternateCorrBlock:
    """On-the-fly windowed correlation, no materialised volume (reference core/corr.py:63-91).

    `pyramid` keeps the reference's attribute (num_levels+1 pairs of average-pooled feature maps) but is
    only computed if somebody reads it: the kernels need neither the pooled fmap1 nor the fp32 pooled
    fmap2 that core/corr.py:68-72 builds.  With precision "bf16x3" (default, fp32-class) or "bf16", radius
    3 or 4 and <= 4 levels every call is ONE tensor-core launch for all levels (rc_alt_corr_pyramid_run):
    the feature maps are pooled, converted and described to the TMA unit once here -- not re-permuted on
    every call as in core/corr.py:82-83 -- so a call is one shape check, one allocation and one C call.
    precision="fp32" or other radii use the per-level CUDA-core kernel behind `alt_cuda_corr.forward`.
    fmap1 may be a band of query rows, (B, C, H1, W1) against the full (B, C, H2, W2) fmap2 (coords in
    fmap2's frame): how one large pair is split over several GPUs (raft_ros_b200.shard).  Forward only,
    as in the reference.  A block is used from one stream at a time (its plan scratch is reused).
    """

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, precision=None):
        self.num_levels = num_levels
        self.radius = radius
        _require_cuda(fmap1, "fmap1")
        _require_cuda(fmap2, "fmap2")
        self._fmaps = (fmap1, fmap2)
        self._pyramid = None
        self._prec = _precision_code(precision)
        if self._prec == _lib.RC_PREC_BF16_VOL16:
            raise ValueError("precision 'bf16_vol16' applies to CorrBlock (there is no volume here); use 'bf16'")
        if fmap1.dim() != 4 or fmap2.dim() != 4 or fmap1.shape[:2] != fmap2.shape[:2] or fmap1.device != fmap2.device:
            raise RuntimeError(f"fmap1/fmap2 must be (B,C,H,W) with the same B, C and device; got {tuple(fmap1.shape)} and "
                               f"{tuple(fmap2.shape)}")
        self._tc = self._prec != _lib.RC_PREC_FP32 and radius in (3, 4) and 1 <= num_levels <= 4
        if self._tc:
            self._setup_tc()
        else:
            from torch.nn import functional as F
            self._f1_nhwc = _f32c(fmap1.permute(0, 2, 3, 1))
            self._f2_nhwc, f2 = [], fmap2
            for _ in range(self.num_levels):
                self._f2_nhwc.append(_f32c(f2.permute(0, 2, 3, 1)))
                f2 = F.avg_pool2d(f2, 2, stride=2)

    @property
    def pyramid(self):
        """[(fmap1, fmap2)] + num_levels x avg_pool2d of both, as core/corr.py:67-72 (computed on first access)."""
        if self._pyramid is None:
            from torch.nn import functional as F
            fmap1, fmap2 = self._fmaps
            self._pyramid = [(fmap1, fmap2)]
            for _ in range(self.num_levels):
                fmap1 = F.avg_pool2d(fmap1, 2, stride=2)
                fmap2 = F.avg_pool2d(fmap2, 2, stride=2)
                self._pyramid.append((fmap1, fmap2))
        return self._pyramid

    def _setup_tc(self):
        import ctypes
        L = _lib.lib()
        f1, f2 = _f32c(self._fmaps[0]), _f32c(self._fmaps[1])
        B, C, H1, W1 = f1.shape
        _, _, H2, W2 = f2.shape
        level_shapes(H2, W2, self.num_levels)  # raises if a level would be empty
        f8 = self._prec == _lib.RC_PREC_FP16_FP8
        x3 = self._prec == _lib.RC_PREC_BF16X3 or f8  # two maps per feature map
        Cp = (C + 63) // 64 * 64
        dev = f1.device
        with torch.cuda.device(dev):
            if f8:  # power-of-two normalisation per batch item from the largest magnitude of each level-0 map
                self._amax = torch.zeros((2, B), dtype=torch.int32, device=dev)
                for i, (f, H, W) in enumerate(((f1, H1, W1), (f2, H2, W2))):
                    _lib.check(L.rc_alt_absmax(f.data_ptr(), self._amax[i].data_ptr(), B, C, H, W, _stream()), "rc_alt_absmax")

            def prep(f, H, W, level, side=0):
                hi = torch.empty((B, H >> level, W >> level, Cp), dtype=torch.float16 if f8 else torch.bfloat16, device=dev)
                lo = torch.empty_like(hi) if x3 else None
                if f8:
                    rc = L.rc_alt_prep_fmap_level_f8(f.data_ptr(), hi.data_ptr(), lo.data_ptr(), self._amax[side].data_ptr(), side,
                                                     B, C, H, W, level, _stream())
                    _lib.check(rc, "rc_alt_prep_fmap_level_f8")
                    return hi, lo
                rc = L.rc_alt_prep_fmap_level(f.data_ptr(), hi.data_ptr(), lo.data_ptr() if x3 else None, B, C, H, W, level,
                                              _stream())
                _lib.check(rc, "rc_alt_prep_fmap_level")
                return hi, lo

            self._f1_tc = prep(f1, H1, W1, 0)
            self._f2_tc = [prep(f2, H2, W2, l, 1) for l in range(self.num_levels)]
            self._setup = ctypes.create_string_buffer(int(L.rc_alt_corr_pyramid_setup_bytes()))
            f2_hi = _lib.ptr_array([t[0].data_ptr() for t in self._f2_tc])
            f2_lo = _lib.ptr_array([t[1].data_ptr() if x3 else None for t in self._f2_tc])
            rc = L.rc_alt_corr_pyramid_setup(self._setup, self._f1_tc[0].data_ptr(), self._f1_tc[1].data_ptr() if x3 else None,
                                             f2_hi, f2_lo, B, H1, W1, H2, W2, C, self.num_levels, self.radius, self._prec)
            _lib.check(rc, "rc_alt_corr_pyramid_setup")
            if f8:
                _lib.check(L.rc_alt_corr_pyramid_set_norm(self._setup, self._amax[0].data_ptr(), self._amax[1].data_ptr()),
                           "rc_alt_corr_pyramid_set_norm")
            self._plan_bytes = int(L.rc_alt_corr_pyramid_plan_bytes(B, H1, W1, self.num_levels))
            self._plan = torch.empty(self._plan_bytes, dtype=torch.uint8, device=dev)
        rd = 2 * self.radius + 1
        self._shape, self._out_shape, self._device = (B, 2, H1, W1), (B, self.num_levels * rd * rd, H1, W1), dev
        # the reference divides by torch.sqrt(torch.tensor(dim).float()): an fp32 square root (core/corr.py:91)
        self._scale = 1.0 / float(torch.sqrt(torch.tensor(float(C))))
        self._run = L.rc_alt_corr_pyramid_run

    def __call__(self, coords):
        if self._tc:
            return self._call_tc(coords)
        from . import alt_cuda_corr

        coords = coords.permute(0, 2, 3, 1)
        B, H, W, _ = coords.shape
        dim = self._fmaps[0].shape[1]
        corr_list = []
        for i in range(self.num_levels):
            coords_i = (coords / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
            (corr,) = alt_cuda_corr.forward(self._f1_nhwc, self._f2_nhwc[i], coords_i, self.radius)
            corr_list.append(corr.squeeze(1))
        corr = torch.stack(corr_list, dim=1).reshape(B, -1, H, W)
        # the reference divides by torch.sqrt(torch.tensor(dim).float()): an fp32 square root
        return corr / float(torch.sqrt(torch.tensor(float(dim))))

    def _call_tc(self, coords):
        _require_cuda(coords, "coords")
        if tuple(coords.shape) != self._shape:
            raise RuntimeError(f"coords must be {self._shape}, got {tuple(coords.shape)}")
        c = coords if (coords.dtype is torch.float32 and coords.is_contiguous() and not coords.requires_grad) else _f32c(coords)
        if c.device != self._device:
            raise RuntimeError("coords must be on the feature maps' device")
        with torch.cuda.device(self._device):
            out = torch.empty(self._out_shape, dtype=torch.float32, device=self._device)
            rc = self._run(self._setup, c.data_ptr(), out.data_ptr(), self._scale, self._plan.data_ptr(), self._plan_bytes,
                           torch.cuda.current_stream().cuda_stream)
            if rc:
                _lib.check(rc, "rc_alt_corr_pyramid_run")
        return out
