"""rc_fnet_conv1x1_prep / rc_corr_build_pyramid_tiled_prepared (-m gpu): the encoder's output 1x1 convolution
(`conv2`, reference core/extractor.py:144,184; raft-small :217,256) emitting the build's operand maps, against the
reference's own steps: `F.conv2d` in fp32 (TF32 off), `.float()`, then CorrBlock on the fp32 maps.

Bounds: the operand maps hold hi + lo = 16 mantissa bits of the convolution result: |hi + lo - conv| <= 2e-5 relative
to the largest magnitude (products are error-compensated bf16, accumulation fp32).  A CorrBlock built from them meets
the same bound as CorrBlock(fmap1, fmap2) at the same precision: bf16x3 relmax <= 1e-4, normwise <= 3e-5 (+ the
lookup's 2e-5 coordinate-rounding allowance)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import corr_oracle as O
from tests.helpers import make_coords, normwise, relmax

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def rr():
    import raft_ros_b200
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return raft_ros_b200


def n(x):
    return x.detach().float().cpu().numpy()


CASES = [
    # B, Cin, Cout, h, w, L, r
    (1, 128, 256, 24, 40, 4, 4),
    (2, 128, 256, 23, 39, 4, 4),   # ragged: 897 pixels per image
    (1, 96, 128, 16, 18, 4, 3),    # raft-small: conv2 96 -> 128, radius 3
    (1, 128, 256, 55, 128, 4, 4),  # BASELINE config 1 grid
    (3, 64, 64, 9, 17, 2, 4),      # fewer pixels than a tile, small channel counts
]


@pytest.mark.parametrize("case", CASES)
def test_encoder_head_and_build_from_prepared(rr, case):
    B, Cin, Cout, h, w, L, r = case
    rng = np.random.default_rng(Cin * 1000 + h * w)
    x = torch.from_numpy(rng.standard_normal((2 * B, Cin, h, w)).astype(np.float32)).to(DEV)
    W = torch.from_numpy((rng.standard_normal((Cout, Cin, 1, 1)) / np.sqrt(Cin)).astype(np.float32)).to(DEV)
    bias = torch.from_numpy(rng.standard_normal(Cout).astype(np.float32) * 0.3).to(DEV)
    prep = rr.encoder_head(x, W, bias)
    assert prep.hi.shape == (2 * B, h, w, Cout) and prep.hi.dtype == torch.bfloat16 and prep.lo.shape == prep.hi.shape
    ref = F.conv2d(x.double(), W.double(), bias.double()).float()      # what conv2 + .float() gives (core/raft.py:99-103)
    got = prep.float()
    assert torch.isfinite(got).all()
    assert relmax(n(got), n(ref)) <= 2e-5, relmax(n(got), n(ref))
    # hi is the correctly rounded bf16 of (something within 2e-5 of) the result, lo the remainder
    assert relmax(n(prep.hi.float().permute(0, 3, 1, 2)), n(ref)) <= 2 ** -8
    if h * w > 1200:
        from oracle import torch_port as TP
        with torch.no_grad():
            pyr = TP.build_pyramid(ref[:B], ref[B:], L)
            want_pyr = [n(p) for p in pyr]
    else:
        want_pyr = O.corr_build(n(ref[:B]), n(ref[B:]), L)
    for precision, tol in (("bf16x3", dict(rel=1e-4, norm=3e-5)), ("bf16", dict(rel=3e-2, norm=6e-3))):
        blk = rr.CorrBlock.from_prepared(prep, num_levels=L, radius=r, precision=precision)
        assert blk._tiled is not None and len(blk.corr_pyramid) == L
        for l in range(L):
            a = n(blk.corr_pyramid[l])
            assert a.shape == want_pyr[l].shape
            assert relmax(a, want_pyr[l]) <= tol["rel"], (case, precision, l, relmax(a, want_pyr[l]))
            assert normwise(a, want_pyr[l]) <= tol["norm"], (case, precision, l)
        if precision == "bf16x3" and h * w <= 1200:
            coords = make_coords("sigma4", B, h, w, rng)
            want = O.corr_lookup(want_pyr, coords, r)
            got_l = n(blk(torch.from_numpy(coords).to(DEV)))
            assert relmax(got_l, want) <= 1e-4 + 2e-5
    # against the two-step path on the reconstructed fp32 maps at the single-pass precision: the same volume up to the
    # few operands (~2^-8 of them) whose hi + lo lands across a bf16 rounding boundary from the accumulator's value
    a = rr.CorrBlock.from_prepared(prep, num_levels=L, radius=r, precision="bf16")
    b_ = rr.CorrBlock(got[:B], got[B:], num_levels=L, radius=r, precision="bf16")
    for x_, y_ in zip(a.corr_pyramid, b_.corr_pyramid):
        assert normwise(n(x_), n(y_)) <= 1e-3


def test_encoder_head_rejects(rr):
    x = torch.zeros(2, 100, 8, 8, device=DEV)
    with pytest.raises(RuntimeError, match="multiple of 16"):
        rr.encoder_head(x, torch.zeros(256, 100, device=DEV), torch.zeros(256, device=DEV))
    x = torch.zeros(2, 128, 8, 8, device=DEV)
    with pytest.raises(RuntimeError, match="multiple of 64"):
        rr.encoder_head(x, torch.zeros(100, 128, device=DEV), torch.zeros(100, device=DEV))
    prep = rr.encoder_head(x, torch.zeros(256, 128, device=DEV), torch.zeros(256, device=DEV), with_lo=False)
    with pytest.raises(RuntimeError, match="lo"):
        rr.CorrBlock.from_prepared(prep, precision="bf16x3")
    assert rr.CorrBlock.from_prepared(prep, precision="bf16", num_levels=2)(rr.coords_grid(1, 8, 8, DEV)).shape == (1, 162, 8, 8)
