"""rc_conv2d_fwd (-m gpu): the stride-1 "same" convolutions of the per-iteration update (reference core/update.py) on
tensor cores, against torch's convolution in fp64 on the same inputs.  Operands are rounded to fp16, so the bound is the
fused convc1's: relmax <= 2e-3, normwise <= 6e-4; against the same convolution on fp16-rounded inputs (fp64 accumulate)
what is left is fp32 summation order: relmax <= 2e-5."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

CASES = [
    # B, C1, C2, Cout, H, W, KH, KW, act                         the layer of core/update.py it stands for
    (1, 256, 0, 192, 55, 128, 3, 3, "relu"),      # convc2 (:85,91)
    (2, 128, 0, 64, 23, 39, 3, 3, "relu"),        # convf2 (:87,93), ragged width
    (1, 192, 64, 126, 24, 40, 3, 3, "relu"),      # conv on cat([cor, flo]) (:88,95-96): 126 filters, padded to 128 inside
    (2, 256, 0, 2, 12, 20, 3, 3, None),           # FlowHead conv2 (:10): two filters, padded to 16 inside
    (1, 128, 256, 128, 30, 130, 1, 5, "sigmoid"), # SepConvGRU convz1 / convr1 on cat([h, x]) (:36-37,42-43), two row tiles
    (1, 128, 256, 128, 21, 64, 5, 1, "tanh"),     # SepConvGRU convq2 (:42,50)
    (1, 128, 0, 256, 16, 33, 3, 3, "relu"),       # FlowHead conv1 / mask head conv (:9,123)
    (1, 256, 0, 576, 9, 17, 1, 1, None),          # mask head 1x1, 576 channels = three slices (:125)
    (3, 40, 0, 16, 5, 7, 3, 3, None),             # channel count that is not a multiple of 64, tiny image
    (2, 16, 0, 16, 2, 8, 3, 3, None),             # image lower and narrower than the staged box (box pinned to the origin)
    (1, 24, 40, 32, 2, 12, 5, 1, "relu"),         # the same for the column filter's 8 x 16 tiles, two inputs
    (1, 4, 0, 16, 8, 8, 3, 3, None),              # fewer channels than a staged box holds: direct-load path
    (1, 128, 256, 256, 55, 128, 1, 5, "sigmoid"), # z and r gates as one 256-filter convolution at the real size
    (8, 256, 0, 192, 55, 128, 3, 3, "relu"),      # batch 8: 440 tiles, three per CTA (the persistent loop and its ring phases)
    (3, 64, 64, 64, 55, 128, 1, 5, "tanh"),       # 165 tiles: some CTAs do two, some one
    (8, 128, 256, 128, 21, 64, 5, 1, "sigmoid"),  # many 8 x 16 tiles of the column filters
    (5, 48, 0, 126, 31, 40, 3, 3, None),          # several tiles per CTA with a padded filter count
    (5, 48, 0, 32, 31, 39, 3, 3, None),           # the same on the direct-load path (odd width)
    (1, 64, 0, 32, 47, 156, 3, 3, "relu"),        # KITTI's 47 x 156 grid: two tiles per row, staged boxes that start left of x0
    (1, 96, 146, 96, 47, 156, 3, 3, "tanh"),      # raft-small's GRU candidate convolution there (two inputs, 146 channels)
    (2, 128, 256, 128, 9, 300, 1, 5, "sigmoid"),  # three tiles per row, 1x5
    (1, 64, 0, 48, 11, 260, 3, 5, None),          # 3x5 filters, a last tile of 4 columns
]


def _relmax(a, b):
    return float((a - b).abs().max() / b.abs().max())


@pytest.mark.parametrize("case", CASES)
def test_conv2d_act_matches_torch(case):
    import raft_ros_b200 as rr
    from raft_ros_b200 import _lib
    _lib.build()
    B, C1, C2, Cout, H, W, KH, KW, act = case
    g = torch.Generator().manual_seed(C1 * 100 + H * W + KH)
    a = torch.randn(B, C1, H, W, generator=g).to(DEV)
    b = torch.randn(B, C2, H, W, generator=g).to(DEV) if C2 else None
    Cin = C1 + C2
    wgt = (torch.randn(Cout, Cin, KH, KW, generator=g) / (Cin * KH * KW) ** 0.5).to(DEV)
    bias = (0.2 * torch.randn(Cout, generator=g)).to(DEV)
    x = a if b is None else torch.cat([a, b], dim=1)
    fn = {None: lambda t: t, "relu": torch.relu, "sigmoid": torch.sigmoid, "tanh": torch.tanh}[act]
    pad = (KH // 2, KW // 2)
    want = fn(F.conv2d(x.double(), wgt.double(), bias.double(), padding=pad)).float()
    got = rr.conv2d_act(a if b is None else (a, b), wgt, bias, act)
    assert got.shape == want.shape and got.dtype == torch.float32
    assert torch.isfinite(got).all()
    scale = float(F.conv2d(x.double(), wgt.double(), bias.double(), padding=pad).abs().max())  # pre-activation magnitude
    assert float((got - want).abs().max()) <= 2e-3 * max(scale, 1.0), (case, float((got - want).abs().max()), scale)
    want16 = fn(F.conv2d(x.half().double(), wgt.half().double(), bias.double(), padding=pad)).float()
    assert float((got - want16).abs().max()) <= 2e-5 * max(scale, 1.0), (case, float((got - want16).abs().max()))
    assert torch.equal(rr.conv2d_act(a if b is None else (a, b), wgt, bias, act), got)


def test_conv2d_act_rejects():
    import raft_ros_b200 as rr
    x = torch.zeros(1, 64, 8, 8, device=DEV)
    with pytest.raises(RuntimeError, match="unsupported"):
        rr.conv2d_act(x, torch.zeros(64, 64, 7, 7, device=DEV))
    from raft_ros_b200 import _lib
    L = _lib.lib()
    assert L.rc_conv2d_fwd(x.data_ptr(), 64, None, x.data_ptr(), None, x.data_ptr(), 1, 64, 30, 8, 8, 3, 3, 0, None) != 0  # C ABI: N granule 16
    with pytest.raises(RuntimeError, match="inference"):
        rr.conv2d_act(x.clone().requires_grad_(True), torch.zeros(64, 64, 3, 3, device=DEV))


def test_conv2d_act_sees_a_new_bias_and_an_updated_weight():
    import raft_ros_b200 as rr
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, 32, 6, 9, generator=g).to(DEV)
    w = torch.randn(16, 32, 3, 3, generator=g).to(DEV)
    b1, b2 = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    y1 = rr.conv2d_act(x, w, b1)
    y2 = rr.conv2d_act(x, w, b2)
    assert torch.allclose(y2 - y1, torch.ones_like(y1), atol=1e-5)
    b2.mul_(3.0)
    assert torch.allclose(rr.conv2d_act(x, w, b2) - y1, 3 * torch.ones_like(y1), atol=1e-5)
    w.mul_(2.0)
    assert torch.allclose(rr.conv2d_act(x, w, b1), 2 * y1, atol=1e-4)
