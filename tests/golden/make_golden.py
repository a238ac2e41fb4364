"""Generates tests/golden/*.npz by EXECUTING THE REFERENCE's own Python.

Run in the build container only (the reference checkout is not on the GPU box):

    python tests/golden/make_golden.py [/root/reference]

The reference ships no tests and no golden vectors; these fixtures are the outputs of its
own `core/corr.py::CorrBlock` (and torch.autograd through it) on seeded inputs, on CPU,
fp32, torch 2.11.0.  They pin oracle/corr_oracle.c and oracle/torch_port.py, which in
turn check the CUDA kernels.  Nothing at test time reads the reference checkout.
"""
import os
import sys
import warnings

import numpy as np
import torch

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.filterwarnings("ignore")

from core.corr import CorrBlock  # noqa: E402  (the reference, unmodified)
from core.utils.utils import coords_grid  # noqa: E402

from oracle import corr_oracle as O  # noqa: E402
from oracle import torch_port as TP  # noqa: E402

torch.set_num_threads(4)


def make_coords(kind, B, h, w, gen):
    base = coords_grid(B, h, w, "cpu")
    if kind == "grid":
        return base
    if kind == "sigma0.5":
        return base + 0.5 * torch.randn(B, 2, h, w, generator=gen)
    if kind == "sigma4":
        return base + 4.0 * torch.randn(B, 2, h, w, generator=gen)
    if kind == "oob":
        return base + (torch.rand(B, 2, h, w, generator=gen) * 2 - 1) * (w + 10)
    if kind == "border":
        c = base.clone()
        c[:, 0] = torch.where(torch.rand(B, h, w, generator=gen) < 0.5, -0.37, w - 1 + 0.41)
        c[:, 1] = torch.where(torch.rand(B, h, w, generator=gen) < 0.5, 0.25, h - 1 - 0.6)
        return c
    raise ValueError(kind)


def cotangent(name, kind, shape):
    """Seeded cotangent for the backward fixtures; regenerated (not stored) by the tests.
    numpy's PCG64 stream is stable across versions."""
    import zlib
    seed = zlib.crc32(f"{name}/{kind}".encode())
    return np.random.default_rng(seed).standard_normal(shape, dtype=np.float32)


def case(name, B, D, h, w, L, r, kinds, seed, feat_scale=1.0):
    gen = torch.Generator().manual_seed(seed)
    f1 = (feat_scale * torch.randn(B, D, h, w, generator=gen)).requires_grad_(True)
    f2 = (feat_scale * torch.randn(B, D, h, w, generator=gen)).requires_grad_(True)
    blk = CorrBlock(f1, f2, num_levels=L, radius=r)
    out = {"fmap1": f1.detach().numpy(), "fmap2": f2.detach().numpy(), "num_levels": L, "radius": r}
    for l, lv in enumerate(blk.corr_pyramid):
        out[f"level{l}"] = lv.detach().numpy()
    # oracle / port pinned on the build
    o_levels = O.corr_build(out["fmap1"], out["fmap2"], L)
    p_levels = TP.build_pyramid(f1.detach(), f2.detach(), L)
    for l in range(L):
        d_or = np.abs(o_levels[l] - out[f"level{l}"]).max()
        d_pt = (p_levels[l] - blk.corr_pyramid[l].detach()).abs().max().item()
        print(f"[{name}] level{l} {tuple(out[f'level{l}'].shape)} |oracle-ref|max={d_or:.3e} |port-ref|max={d_pt:.3e}")
        assert d_pt == 0.0
    grad_total = None
    for kind in kinds:
        coords = make_coords(kind, B, h, w, gen)
        res = blk(coords)
        out[f"coords_{kind}"] = coords.numpy()
        out[f"out_{kind}"] = res.detach().numpy()
        ref_levels = [out[f"level{l}"] for l in range(L)]
        o_rt = O.corr_lookup(ref_levels, coords.numpy(), r, roundtrip=True)
        o_px = O.corr_lookup(ref_levels, coords.numpy(), r, roundtrip=False)
        p_out = TP.lookup([lv.detach() for lv in blk.corr_pyramid], coords, r)
        scale = np.abs(out[f"out_{kind}"]).max() + 1e-30
        print(f"[{name}] lookup {kind:9s} |oracle(roundtrip)-ref|max={np.abs(o_rt - out[f'out_{kind}']).max():.3e} "
              f"|oracle(pixel)-ref|max={np.abs(o_px - out[f'out_{kind}']).max():.3e} (out max {scale:.3f}) "
              f"|port-ref|max={(p_out - res.detach()).abs().max().item():.3e}")
        assert (p_out - res.detach()).abs().max().item() == 0.0
        # backward golden: one seeded cotangent per coords set, summed like the 12 iterations do
        g = torch.from_numpy(cotangent(name, kind, tuple(res.shape)))
        grad_total = (res * g).sum() if grad_total is None else grad_total + (res * g).sum()
    grad_total.backward()
    out["dfmap1"] = f1.grad.numpy()
    out["dfmap2"] = f2.grad.numpy()
    # oracle backward vs autograd of the reference
    gl = None
    for kind in kinds:
        gk = O.corr_lookup_bwd(cotangent(name, kind, out[f"out_{kind}"].shape), out[f"coords_{kind}"], L, r)
        gl = gk if gl is None else [a + b for a, b in zip(gl, gk)]
    d1, d2 = O.corr_build_bwd(gl, out["fmap1"], out["fmap2"])
    print(f"[{name}] backward |oracle-ref|max df1={np.abs(d1 - out['dfmap1']).max():.3e} df2={np.abs(d2 - out['dfmap2']).max():.3e} "
          f"(|df1|max {np.abs(out['dfmap1']).max():.3f})")
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"[{name}] wrote {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    # odd sizes on every level (17x20 -> 8x10 -> 4x5 -> 2x2), full 4 levels r=4 like core/raft.py:43
    case("corrblock_17x20_r4", B=1, D=16, h=17, w=20, L=4, r=4,
         kinds=["grid", "sigma4", "border"], seed=1234)
    # raft-small geometry: radius 3 (core/raft.py:37), 2 levels, larger feature magnitudes
    case("corrblock_16x18_r3", B=2, D=32, h=16, w=18, L=2, r=3,
         kinds=["grid", "sigma0.5", "sigma4", "oob", "border"], seed=4321, feat_scale=3.0)
