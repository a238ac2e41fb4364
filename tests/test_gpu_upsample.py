"""rc_convex_upsample_fwd (-m gpu) against the reference's own RAFT.upsample_flow (core/raft.py:72-83), restated with
the same torch ops.  Bound: relmax <= 2e-6 (fp32; the softmax's exp and the order of the 9-term sums differ)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def reference_upsample_flow(flow, mask):
    N, _, H, W = flow.shape
    mask = mask.view(N, 1, 9, 8, 8, H, W)
    mask = torch.softmax(mask, dim=2)
    up_flow = F.unfold(8 * flow, [3, 3], padding=1)
    up_flow = up_flow.view(N, 2, 9, 1, 1, H, W)
    up_flow = torch.sum(mask * up_flow, dim=2)
    up_flow = up_flow.permute(0, 1, 4, 2, 5, 3)
    return up_flow.reshape(N, 2, 8 * H, 8 * W)


@pytest.mark.parametrize("shape", [(1, 55, 128), (2, 47, 156), (3, 9, 17), (1, 1, 1), (1, 50, 90)])
def test_convex_upsample_matches_reference(shape):
    import raft_ros_b200 as rr
    from raft_ros_b200 import _lib
    _lib.build()
    N, H, W = shape
    g = torch.Generator().manual_seed(H * W)
    flow = (5 * torch.randn(N, 2, H, W, generator=g)).to(DEV)
    mask = (2 * torch.randn(N, 576, H, W, generator=g)).to(DEV)
    mask[0, :, 0, 0] = 40.0 * torch.randn(576, generator=g).to(DEV)  # a peaked softmax
    want = reference_upsample_flow(flow, mask)
    got = rr.upsample_flow(flow, mask)
    assert got.shape == want.shape == (N, 2, 8 * H, 8 * W) and got.dtype == torch.float32 and got.is_contiguous()
    assert torch.isfinite(got).all()
    err = float((got - want).abs().max() / want.abs().max())
    assert err <= 2e-6, err
    with pytest.raises(RuntimeError, match="inference"):
        rr.upsample_flow(flow.clone().requires_grad_(True), mask)
