"""rc_corr_lookup_convc1_fwd (-m gpu): the lookup fused into convc1 + ReLU (reference core/update.py:83,90;
raft-small :66,72) against the reference's two steps, `F.relu(F.conv2d(lookup, W, b))`, with the lookup from the
checker (the C oracle on small grids, the reference's torch ops at 55x128).

Bounds (relative to the largest reference magnitude / normwise; the convolution sums 324 products):
  fp16 operands: taps and weights rounded once to 11-bit mantissas -> relmax <= 2e-3, normwise <= 6e-4.  The same
                 rounding is what the reference's --mixed_precision autocast applies to this convolution; against that
                 computation (fp16 inputs, fp32 accumulate) the kernel agrees to relmax <= 3e-4: fp32 summation
                 order, plus the few taps whose fp32 value sits within the lookup's 2e-5 of an fp16 rounding boundary and
                 so round the other way (one 16-bit ulp of one product out of 324).
  bf16 operands: 8-bit mantissas -> relmax <= 1.5e-2, normwise <= 5e-3.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import corr_oracle as O
from tests.helpers import make_coords

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
BOUND = {"fp16": dict(rel=2e-3, norm=6e-4), "bf16": dict(rel=1.5e-2, norm=5e-3)}


@pytest.fixture(scope="module")
def rr():
    import raft_ros_b200
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return raft_ros_b200


def _relmax(a, b):
    return float((a - b).abs().max() / b.abs().max())


def _normwise(a, b):
    return float((a - b).double().norm() / b.double().norm())


CASES = [
    # B, D, h, w, L, r, Cout
    (1, 64, 24, 40, 4, 4, 256),
    (2, 96, 23, 39, 4, 4, 256),     # ragged: 897 pixels = 7 tiles + 1, two pairs
    (1, 32, 16, 18, 4, 3, 96),      # raft-small: radius 3, 96 output channels
    (3, 32, 9, 17, 1, 4, 48),       # one level, fewer pixels than a tile, Cout with a 16-column tail
    (2, 256, 55, 128, 4, 4, 256),   # BASELINE config 1/2 grid
]


@pytest.mark.parametrize("operands", ["fp16", "bf16"])
@pytest.mark.parametrize("precision", ["bf16x3", "bf16_vol16"])
@pytest.mark.parametrize("case", CASES)
def test_fused_lookup_convc1(rr, case, precision, operands):
    B, D, h, w, L, r, Cout = case
    rng = np.random.default_rng(B * 1000 + h * w + Cout)
    f1 = torch.from_numpy(rng.standard_normal((B, D, h, w)).astype(np.float32)).to(DEV)
    f2 = torch.from_numpy(rng.standard_normal((B, D, h, w)).astype(np.float32)).to(DEV)
    K = L * (2 * r + 1) ** 2
    W = torch.from_numpy((rng.standard_normal((Cout, K, 1, 1)) / np.sqrt(K)).astype(np.float32)).to(DEV)
    bias = torch.from_numpy(rng.standard_normal(Cout).astype(np.float32) * 0.1).to(DEV)
    blk = rr.CorrBlock(f1, f2, num_levels=L, radius=r, precision=precision)
    for kind in ("sigma4", "border", "oob"):
        cn = make_coords(kind, B, h, w, rng)
        coords = torch.from_numpy(cn).to(DEV)
        out = torch.full((1,), float("nan"), device=DEV)
        out = blk.lookup_convc1(coords, W, bias, operands=operands)
        assert out.shape == (B, Cout, h, w) and out.dtype == torch.float32 and out.is_contiguous()
        assert torch.isfinite(out).all() and float(out.min()) >= 0.0
        # the checker's lookup on the pyramid this block holds (the build has its own tests)
        pyr = [lv.cpu().numpy() for lv in blk.corr_pyramid]
        if h * w <= 1200:
            look = torch.from_numpy(O.corr_lookup(pyr, cn, r)).to(DEV)
        else:
            from oracle import torch_port as TP
            with torch.no_grad():
                look = TP.lookup([lv.clone() for lv in blk.corr_pyramid], coords, r)
        ref = F.relu(F.conv2d(look, W, bias))
        tol = BOUND[operands]
        assert _relmax(out, ref) <= tol["rel"], (case, precision, operands, kind, _relmax(out, ref))
        assert _normwise(out, ref) <= tol["norm"], (case, precision, operands, kind, _normwise(out, ref))
        # the same operand rounding applied to the checker's lookup: what is left is summation order and rounding flips
        cast = (lambda t: t.half().float()) if operands == "fp16" else (lambda t: t.bfloat16().float())
        ref16 = F.relu(F.conv2d(cast(look.clamp(-65504, 65504)).double(), cast(W).double(), bias.double())).float()
        assert _relmax(out, ref16) <= (3e-4 if operands == "fp16" else 2.5e-3), (case, precision, operands, kind, _relmax(out, ref16))
        assert torch.equal(blk.lookup_convc1(coords, W, bias, operands=operands), out)  # deterministic


def test_fused_rejects(rr):
    f = torch.randn(1, 32, 16, 16, device=DEV)
    blk = rr.CorrBlock(f, f, precision="bf16")
    c = rr.coords_grid(1, 16, 16, DEV)
    with pytest.raises(RuntimeError, match="weight"):
        blk.lookup_convc1(c, torch.zeros(256, 100, device=DEV), torch.zeros(256, device=DEV))
    with pytest.raises(RuntimeError, match="multiple of 16"):
        blk.lookup_convc1(c, torch.zeros(100, 324, device=DEV), torch.zeros(100, device=DEV))
    with pytest.raises(RuntimeError, match="tiled"):
        rr.CorrBlock(f, f, precision="fp32").lookup_convc1(c, torch.zeros(256, 324, device=DEV), torch.zeros(256, device=DEV))
