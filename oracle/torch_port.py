"""Torch-CPU port of the reference CorrBlock, op for op, used ONLY as the timed CPU baseline.

TEST / MEASUREMENT INFRASTRUCTURE ONLY (bench.py `cpu_baseline` and `--impl reference`,
tests/).  The reference's CPU path *is* PyTorch's CPU kernels (MKL sgemm, ATen avg_pool2d
and grid_sampler): the reference checkout is not present on the GPU box, so this module
issues the same vendor calls in the same order so that the timing is representative:

    build   : core/corr.py:52-60 (view, matmul, /sqrt(D)) then :21-27 (3x avg_pool2d)
    lookup  : core/corr.py:29-50 -> core/utils/utils.py:57-65 (normalise, grid_sample
              bilinear / zeros / align_corners=True), cat, permute, contiguous

It is checked against the real reference in tests/golden/make_golden.py (bit-identical
on the fixture inputs) and against the C oracle in tests/test_oracle.py.
"""
from __future__ import annotations

import math
from typing import List

import torch
import torch.nn.functional as F


def build_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4) -> List[torch.Tensor]:
    b, d, h, w = fmap1.shape
    n = h * w
    vol = torch.matmul(fmap1.reshape(b, d, n).transpose(1, 2), fmap2.reshape(b, d, n))
    vol = vol / torch.sqrt(torch.tensor(d).float())
    level = vol.reshape(b * n, 1, h, w)
    pyramid = [level]
    for _ in range(1, num_levels):
        level = F.avg_pool2d(level, 2, stride=2)
        pyramid.append(level)
    return pyramid


def _window_offsets(radius: int, device) -> torch.Tensor:
    # offsets[i, j] = (i - r, j - r): first window index moves x, second moves y
    k = torch.linspace(-radius, radius, 2 * radius + 1, device=device)
    a, bb = torch.meshgrid(k, k, indexing="ij")
    return torch.stack((a, bb), dim=-1)


def _sample(img: torch.Tensor, xy: torch.Tensor) -> torch.Tensor:
    hh, ww = img.shape[-2:]
    gx = 2 * xy[..., 0:1] / (ww - 1) - 1
    gy = 2 * xy[..., 1:2] / (hh - 1) - 1
    return F.grid_sample(img, torch.cat((gx, gy), dim=-1), align_corners=True)


def lookup(pyramid: List[torch.Tensor], coords: torch.Tensor, radius: int = 4) -> torch.Tensor:
    b, _, h, w = coords.shape
    centre = coords.permute(0, 2, 3, 1).reshape(b * h * w, 1, 1, 2)
    win = _window_offsets(radius, coords.device).view(1, 2 * radius + 1, 2 * radius + 1, 2)
    feats = []
    for lvl, vol in enumerate(pyramid):
        taps = _sample(vol, centre / 2 ** lvl + win)
        feats.append(taps.view(b, h, w, -1))
    return torch.cat(feats, dim=-1).permute(0, 3, 1, 2).contiguous().float()


def scale_of(dim: int) -> float:
    return 1.0 / math.sqrt(dim)
