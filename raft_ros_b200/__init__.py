"""raft_ros_b200: RAFT's correlation hot path, B200-native (sm_100a), as a drop-in.

    from raft_ros_b200 import CorrBlock, AlternateCorrBlock, bilinear_sampler, coords_grid
    from raft_ros_b200 import alt_cuda_corr          # forward / backward, reference signatures

Only the correlation path of the reference (core/corr.py, core/utils/utils.py:57-77,
alt_cuda_corr/) lives here, plus the two fusions into its neighbours:
`encoder_head` (the encoder's output 1x1 convolution emitting the build's operand maps) and
`CorrBlock.lookup_convc1` (the lookup fused into the motion encoder's first convolution);
`conv2d_act` / `conv2d_fused` / `upsample_flow` for the rest of the per-iteration update, and the host mirrors `UpdateBlock`
and `FusedRAFT` that run a user's reference-built model through all of them; the encoders' residual stages stay torch.
"""
from . import alt_cuda_corr  # noqa: F401
from .corr import AlternateCorrBlock, CorrBlock, DEFAULT_PRECISION, PreparedFeatures, encoder_head  # noqa: F401
from .utils import (bilinear_sampler, conv2d_act, conv2d_fused, coords_grid, flow_update, small_conv_relu,  # noqa: F401
                    upsample_flow)
from .update import FusedRAFT, UpdateBlock  # noqa: F401

__all__ = ["CorrBlock", "AlternateCorrBlock", "bilinear_sampler", "coords_grid", "alt_cuda_corr", "encoder_head",
           "PreparedFeatures", "upsample_flow", "conv2d_act", "conv2d_fused", "small_conv_relu", "flow_update", "UpdateBlock",
           "FusedRAFT"]
