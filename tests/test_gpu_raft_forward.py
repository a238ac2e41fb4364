"""The reference's own `RAFT.forward` (core/raft.py:86-144) run through the drop-ins (-m gpu).

The model, encoders, update block and the iteration loop are the reference's unmodified files
(ferried to the box as oracle/_ref/core/ by oracle/build_ref_core.py).  Each case runs the forward
once as the reference ships it (its own CorrBlock; for `alternate_corr` its own AlternateCorrBlock
on its own compiled alt_cuda_corr binary from oracle/_ref) and once with ONLY the two names at the
call sites core/raft.py:104-107 rebound to this repository's classes -- the swap INTEGRATION.md
describes.  Random-init weights, seed 1234 (train.py:241; checkpoints are offline), 12 iterations,
test_mode=True.  Bound: mean end-point error between the two final flows <= 0.005 px at 1/8
resolution for the fp32-class precision (bf16x3); measured values are printed.
"""
import argparse

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
EPE_BOUND = {"bf16x3": 0.005, "bf16": 0.02}


@pytest.fixture(scope="module")
def ref_core():
    from oracle import build_ref_core
    core = build_ref_core.load()
    if core is None:
        pytest.skip("oracle/_ref/core not present (built in the build container by __graft_entry__.build())")
    import core.raft  # noqa: F401
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()
    return core


def _model(core, small, alternate):
    torch.manual_seed(1234)
    args = argparse.Namespace(small=small, mixed_precision=False, alternate_corr=alternate)
    return core.raft.RAFT(args).to(DEV).eval()


def _images(H, W):
    # smooth random textures, second frame = first shifted by a few pixels plus noise: a flow field exists
    g = torch.Generator().manual_seed(77)
    low = torch.rand(1, 3, H // 8 + 2, W // 8 + 2, generator=g)
    big = torch.nn.functional.interpolate(low, size=(H + 16, W + 16), mode="bicubic", align_corners=False).clamp(0, 1)
    tex = (big + 0.15 * torch.rand(1, 3, H + 16, W + 16, generator=g)).clamp(0, 1) * 255.0
    return tex[:, :, 8:8 + H, 8:8 + W].contiguous().to(DEV), tex[:, :, 5:5 + H, 11:11 + W].contiguous().to(DEV)


@pytest.mark.parametrize("small", [False, True], ids=["things", "small"])
@pytest.mark.parametrize("alternate", [False, True], ids=["corrblock", "alternate"])
def test_reference_forward_through_dropins(ref_core, monkeypatch, small, alternate):
    import raft_ros_b200 as rr
    from oracle import build_ref
    core = ref_core
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    if alternate:
        ext = build_ref.load_module()
        if ext is None:
            pytest.skip("oracle/_ref/alt_cuda_corr_ref.so not present")
        monkeypatch.setattr(core.corr, "alt_cuda_corr", ext, raising=False)  # what `import alt_cuda_corr` would bind
    H, W = 440, 1024  # config 1: a Sintel pair after InputPadder
    im1, im2 = _images(H, W)
    model = _model(core, small, alternate)
    with torch.no_grad():
        flow_ref, up_ref = model(im1, im2, iters=12, test_mode=True)
        again, _ = model(im1, im2, iters=12, test_mode=True)
    noise = float((again - flow_ref).norm(dim=1).mean())  # the reference against itself (cuDNN / atomics)

    made = []

    def spy(cls):
        def make(*a, **k):
            made.append(cls.__name__)
            return cls(*a, **k)
        return make

    monkeypatch.setattr(core.raft, "CorrBlock", spy(rr.CorrBlock))
    monkeypatch.setattr(core.raft, "AlternateCorrBlock", spy(rr.AlternateCorrBlock))
    with torch.no_grad():
        flow_our, up_our = model(im1, im2, iters=12, test_mode=True)
    assert made == ["AlternateCorrBlock" if alternate else "CorrBlock"]
    assert flow_our.shape == flow_ref.shape == (1, 2, H // 8, W // 8) and up_our.shape == (1, 2, H, W)
    assert torch.isfinite(flow_our).all()
    epe = float((flow_our - flow_ref).norm(dim=1).mean())
    mag = float(flow_ref.norm(dim=1).mean())

    # sensitivity: the same forward with the correlation features off by 1 % must move the flow by far more
    # than the bound, otherwise the bound would say nothing about the operator
    class Off(rr.CorrBlock):
        def __call__(self, coords):
            return super().__call__(coords) * 1.01

    monkeypatch.setattr(core.raft, "CorrBlock", Off)
    monkeypatch.setattr(core.raft, "AlternateCorrBlock", Off)
    with torch.no_grad():
        flow_off, _ = model(im1, im2, iters=12, test_mode=True)
    sens = float((flow_off - flow_ref).norm(dim=1).mean())
    print(f"[raft.forward small={small} alternate={alternate}] mean |flow| {mag:.3f} px, EPE vs reference {epe:.3e} px, "
          f"reference run-to-run {noise:.3e} px, EPE with corr features off by 1 % {sens:.3e} px")
    assert epe <= EPE_BOUND["bf16x3"], (epe, noise, mag)
    assert sens > 10 * max(epe, 1e-7), (sens, epe)
