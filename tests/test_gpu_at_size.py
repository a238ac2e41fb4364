"""GPU parity at the sizes BASELINE.json names (-m gpu): VALUES, not only properties.

The C oracle is too slow at these sizes, so the checker here is the reference's own arithmetic run
on the same GPU: its unmodified `alt_cuda_corr` binary from oracle/_ref (config 5) and its torch ops
(`oracle/torch_port.py`, bit-identical to core/corr.py on the fixtures; TF32 off) for CorrBlock and
for autograd (config 4).  None of the expected values comes from this repository's own kernels.

  config 5  2160x3840 -> 270x480, AlternateCorrBlock, tensor-core on-the-fly kernel
  config 4  400x720 crop -> 50x90, batch 6, 12 lookups, forward + backward
  configs 1-3  55x128 and 47x156: sub-pixel lookups on the full-size pyramid
"""
import numpy as np
import pytest
import torch

from tests.helpers import make_coords

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ALT_TOL = {"bf16x3": dict(rel=1e-4, norm=3e-5), "bf16": dict(rel=3e-2, norm=6e-3)}  # = tests/test_gpu_parity.py
LOOKUP_TOL = 2e-5


@pytest.fixture(scope="module")
def rr():
    import raft_ros_b200
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()
    assert torch.cuda.is_available()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return raft_ros_b200


def _relmax(a, b):
    return float((a - b).abs().max() / b.abs().max())


def _normwise(a, b):
    return float((a - b).double().norm() / b.double().norm())


# ----------------------------------------------------------------------------- config 5 (a7)
@pytest.fixture(scope="module")
def fourk(rr):
    """Feature maps of a 4K pair, and the reference binary's answer per coordinate kind (computed once:
    the reference kernel needs ~40 ms per call here)."""
    from oracle import build_ref
    if torch.cuda.mem_get_info()[0] < 8 * 2 ** 30:
        pytest.skip("needs 8 GB of free device memory")
    ext = build_ref.load_module()
    B, D, H, W, L = 1, 256, 270, 480, 4
    g = torch.Generator().manual_seed(4096)
    f1 = torch.randn(B, D, H, W, generator=g).to(DEV)
    f2 = torch.randn(B, D, H, W, generator=g).to(DEV)
    f1n = f1.permute(0, 2, 3, 1).contiguous()
    f2n, f = [], f2
    for _ in range(L):  # core/corr.py:68-72: the fmap2 pyramid is avg_pool2d of the previous level
        f2n.append(f.permute(0, 2, 3, 1).contiguous())
        f = torch.nn.functional.avg_pool2d(f, 2, stride=2)
    cache = {}

    def want(kind):
        if kind not in cache:
            coords = torch.from_numpy(make_coords(kind, B, H, W, np.random.default_rng(len(kind)))).to(DEV)
            outs = []
            for i in range(L):  # core/corr.py:80-91
                ci = (coords.permute(0, 2, 3, 1) / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
                if ext is not None:
                    (c,) = ext.forward(f1n, f2n[i], ci, 4)
                else:  # no reference binary on this box: this repo's fp32 CUDA-core kernel, itself pinned to the
                    (c,) = rr.alt_cuda_corr.forward(f1n, f2n[i], ci, 4)  # binary by test_alt_vs_reference_binary
                outs.append(c.squeeze(1))
            ref = torch.stack(outs, dim=1).reshape(B, -1, H, W) / torch.sqrt(torch.tensor(float(D)))
            torch.cuda.synchronize()
            cache[kind] = (coords, ref)
        return cache[kind]

    return f1, f2, want, ext is not None


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("kind", ["sigma4", "border", "oob"])
def test_alt_tc_at_4k(rr, fourk, kind, precision):
    """AlternateCorrBlock at 1x256x270x480 (N = 129 600; 1020 query tiles x 4 levels of plan entries,
    boxes clipped at every border) against the reference's own compiled alt_cuda_corr kernel
    (alt_cuda_corr/correlation_kernel.cu:18-119 driven as core/corr.py:74-91 drives it), every
    pixel of every level."""
    f1, f2, want, _ = fourk
    coords, ref = want(kind)
    blk = rr.AlternateCorrBlock(f1, f2, num_levels=4, radius=4, precision=precision)
    assert blk._tc, "the tensor-core kernel is what this test is for"
    got = blk(coords)
    assert got.shape == ref.shape == (1, 324, 270, 480) and got.dtype == torch.float32 and got.is_contiguous()
    assert torch.isfinite(got).all()
    tol = ALT_TOL[precision]
    for l in range(4):
        a, b = got[:, 81 * l:81 * (l + 1)], ref[:, 81 * l:81 * (l + 1)]
        assert _relmax(a, b) <= tol["rel"], (kind, precision, l, _relmax(a, b))
        assert _normwise(a, b) <= tol["norm"], (kind, precision, l, _normwise(a, b))
    if kind == "oob":
        # the same pixels come out exactly zero on both sides (window wholly off the map)
        assert torch.equal(got == 0, ref == 0) or float(((got == 0) != (ref == 0)).float().mean()) < 1e-6
    assert torch.equal(blk(coords), got)  # deterministic


# ----------------------------------------------------------------------------- config 4 (a11)
def test_train_step_at_config4(rr):
    """FlyingThings crop 400x720 -> 50x90, batch 6, 12 lookups (train_standard.sh:4, train.py:232):
    CorrBlock(bf16x3) forward + backward to the feature maps against torch autograd through the
    reference's ops.  At this size the backward GEMMs run 2.9 waves of 216 tiles and fold_split crosses
    the ragged 45 / 22 / 11-wide levels.  Gradient bound: relmax <= 1e-4 (the fp32-class bound)."""
    from oracle import torch_port as TP
    B, D, h, w, K, r = 6, 256, 50, 90, 12, 4
    g = torch.Generator().manual_seed(400720)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    base = rr.coords_grid(B, h, w, "cpu")
    coords = (base[None] + torch.cumsum(0.7 * torch.randn(K, B, 2, h, w, generator=g), 0)).to(DEV)
    cot = [torch.randn(B, 324, h, w, generator=g).to(DEV) for _ in range(K)]

    a, b = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    blk = rr.CorrBlock(a, b, num_levels=4, radius=r, precision="bf16x3")
    outs = [blk(coords[k]) for k in range(K)]
    loss = sum((o * c).sum() for o, c in zip(outs, cot))
    loss.backward()

    ra, rb = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    pyr = TP.build_pyramid(ra, rb, 4)
    ref_loss = 0
    for k in range(K):
        ref = TP.lookup(pyr, coords[k], r)
        err = _relmax(outs[k].detach(), ref.detach())
        assert err <= 1e-4 + LOOKUP_TOL, (k, err)
        ref_loss = ref_loss + (ref * cot[k]).sum()
    ref_loss.backward()
    e1, e2 = _relmax(a.grad, ra.grad), _relmax(b.grad, rb.grad)
    assert e1 <= 1e-4 and e2 <= 1e-4, (e1, e2)
    assert _normwise(a.grad, ra.grad) <= 3e-5 and _normwise(b.grad, rb.grad) <= 3e-5
    # bit-reproducible backward (no atomics on this path)
    a2, b2 = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    blk2 = rr.CorrBlock(a2, b2, num_levels=4, radius=r, precision="bf16x3")
    sum((blk2(coords[k]) * cot[k]).sum() for k in range(K)).backward()
    assert torch.equal(a2.grad, a.grad) and torch.equal(b2.grad, b.grad)


# ----------------------------------------------------------------------------- configs 1-3 (a3 at size)
@pytest.mark.parametrize("precision", ["bf16x3", "bf16_vol16"])
@pytest.mark.parametrize("kind", ["sigma0.5", "sigma4", "border"])
@pytest.mark.parametrize("shape", [(2, 256, 55, 128), (1, 256, 47, 156)])
def test_full_size_subpixel_lookup(rr, shape, kind, precision):
    """Sub-pixel lookups on the full-size pyramid (tiled layout, the kernel CorrBlock runs for inference)
    against the reference's grid_sample-based lookup (core/corr.py:29-50) on the SAME pyramid, and the
    whole CorrBlock against the reference's build + lookup."""
    from oracle import torch_port as TP
    B, D, h, w = shape
    g = torch.Generator().manual_seed(h * 1000 + w)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    coords = torch.from_numpy(make_coords(kind, B, h, w, np.random.default_rng(h + w))).to(DEV)
    blk = rr.CorrBlock(f1, f2, num_levels=4, radius=4, precision=precision)
    assert blk._tiled is not None
    got = blk(coords)
    with torch.no_grad():
        same_pyr = TP.lookup([lv.clone() for lv in blk.corr_pyramid], coords, 4)
        assert _relmax(got, same_pyr) <= LOOKUP_TOL, (shape, kind, precision, _relmax(got, same_pyr))
        if precision == "bf16x3":
            ref = TP.lookup(TP.build_pyramid(f1, f2, 4), coords, 4)
            assert _relmax(got, ref) <= 1e-4 + LOOKUP_TOL, (shape, kind, _relmax(got, ref))
            assert _normwise(got, ref) <= 3e-5 + LOOKUP_TOL
