"""Recipe that freezes the REFERENCE's own RAFT sub-networks into oracle/_ref/ as TorchScript.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  The feature encoder, context encoder and update block
are out of this repository's scope: they stay the reference's
torch code.  To report the END-TO-END flow drift of swapping the correlation operator, the
GPU box needs them, and the reference checkout does not exist there; this script imports the
reference (read-only) in the build container, instantiates raft-things with random weights
(seed 1234, the reference's own seed, train.py:241; checkpoints are offline) and traces the
three sub-modules.  Only the traced artifacts land in oracle/_ref/ (git-ignored, shipped with
the snapshot); no reference source is copied.  tools/epe_drift.py drives them.
"""
import argparse
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
FILES = ["raft_fnet.pt", "raft_cnet.pt", "raft_update.pt"]


def built():
    return all(os.path.exists(os.path.join(OUT, f)) for f in FILES)


def build(reference="/root/reference", H=440, W=1024):
    if built():
        return OUT
    warnings.filterwarnings("ignore")
    sys.path.insert(0, reference)
    from core.raft import RAFT  # the reference, unmodified

    torch.manual_seed(1234)
    model = RAFT(argparse.Namespace(small=False, mixed_precision=False, alternate_corr=False)).eval()
    os.makedirs(OUT, exist_ok=True)
    img = torch.rand(1, 3, H, W) * 2 - 1
    h, w = H // 8, W // 8
    with torch.no_grad():
        fnet = torch.jit.trace(model.fnet, img)
        cnet = torch.jit.trace(model.cnet, img)
        upd = torch.jit.trace(model.update_block,
                              (torch.randn(1, 128, h, w), torch.randn(1, 128, h, w), torch.randn(1, 324, h, w),
                               torch.randn(1, 2, h, w)))
    fnet.save(os.path.join(OUT, FILES[0]))
    cnet.save(os.path.join(OUT, FILES[1]))
    upd.save(os.path.join(OUT, FILES[2]))
    return OUT


if __name__ == "__main__":
    print(build())
