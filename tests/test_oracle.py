"""CPU suite: the oracle against the fixtures produced by the reference's own Python
(tests/golden/make_golden.py), and the oracle's internal identities."""
import os

import numpy as np
import pytest

from oracle import corr_oracle as O
from tests.helpers import cotangent, make_coords, relmax

CASES = ["corrblock_17x20_r4", "corrblock_16x18_r3"]


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def _kinds(z):
    return sorted(k[len("coords_"):] for k in z.files if k.startswith("coords_"))


@pytest.mark.parametrize("name", CASES)
def test_build_matches_reference(golden_dir, name):
    z = _load(golden_dir, name)
    L = int(z["num_levels"])
    levels = O.corr_build(z["fmap1"], z["fmap2"], L)
    for l in range(L):
        ref = z[f"level{l}"]
        assert levels[l].shape == ref.shape
        # fp32 SGEMM vs double accumulation: a few ulps of the largest entry
        assert relmax(levels[l], ref) < 2e-6, (name, l)


@pytest.mark.parametrize("name", CASES)
def test_lookup_matches_reference(golden_dir, name):
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    ref_levels = [z[f"level{l}"] for l in range(L)]
    for kind in _kinds(z):
        ref = z[f"out_{kind}"]
        rt = O.corr_lookup(ref_levels, z[f"coords_{kind}"], r, roundtrip=True)
        px = O.corr_lookup(ref_levels, z[f"coords_{kind}"], r, roundtrip=False)
        assert rt.shape == ref.shape and rt.dtype == np.float32
        # with grid_sample's normalise/un-normalise round trip restated: 1-2 ulp
        assert relmax(rt, ref) < 5e-7, (name, kind)
        # pixel-space sampling (what the fused kernel does): the round trip moves coordinates
        # by O(1e-6 * W) and the result is continuous in them
        assert relmax(px, ref) < 5e-6, (name, kind)


@pytest.mark.parametrize("name", CASES)
def test_window_channel_order(golden_dir, name):
    """channel = l*rd^2 + i*rd + j with i the x offset, j the y offset (core/corr.py:37-43)."""
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    rd = 2 * r + 1
    lv0 = z["level0"]
    coords = z["coords_grid"]
    B, _, h, w = coords.shape
    out = z["out_grid"]
    # integer coords at level 0: tap (i,j) of pixel (y,x) is exactly level0[p, y+j-r, x+i-r]
    for (b, y, x, i, j) in [(0, 8, 9, 0, 0), (0, 8, 9, rd - 1, 0), (0, 5, 7, 2, rd - 1), (B - 1, 10, 12, 1, 3)]:
        p = (b * h + y) * w + x
        yy, xx = y + j - r, x + i - r
        want = lv0[p, 0, yy, xx] if (0 <= yy < h and 0 <= xx < w) else 0.0
        got = out[b, i * rd + j, y, x]
        assert abs(got - want) < 1e-5 * max(1.0, abs(want)), (b, y, x, i, j)


@pytest.mark.parametrize("name", CASES)
def test_backward_matches_reference_autograd(golden_dir, name):
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    gl = None
    for kind in _kinds(z):
        g = cotangent(name, kind, z[f"out_{kind}"].shape)
        gk = O.corr_lookup_bwd(g, z[f"coords_{kind}"], L, r)
        gl = gk if gl is None else [a + b for a, b in zip(gl, gk)]
    d1, d2 = O.corr_build_bwd(gl, z["fmap1"], z["fmap2"])
    assert relmax(d1, z["dfmap1"]) < 5e-6
    assert relmax(d2, z["dfmap2"]) < 5e-6


def test_alt_equals_pyramid_lookup():
    """alt_cuda_corr.forward on pooled fmap2 == CorrBlock lookup * sqrt(D) (pooling is linear,
    reference README.md:109-114 / core/corr.py:80-91)."""
    rng = np.random.default_rng(7)
    B, D, h, w, L, r = 2, 32, 16, 20, 3, 3
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    coords = make_coords("sigma4", B, h, w, rng)
    levels = O.corr_build(f1, f2, L)
    want = O.corr_lookup(levels, coords, r) * np.sqrt(np.float32(D))
    rd = 2 * r + 1
    f2l = f2
    for l in range(L):
        if l > 0:
            hh, ww = f2l.shape[2] // 2, f2l.shape[3] // 2
            f2l = f2l[:, :, :2 * hh, :2 * ww].reshape(B, D, hh, 2, ww, 2).mean(axis=(3, 5)).astype(np.float32)
        c = (coords.transpose(0, 2, 3, 1) / 2 ** l).reshape(B, 1, h, w, 2)
        got = O.alt_fwd(f1.transpose(0, 2, 3, 1), f2l.transpose(0, 2, 3, 1), c, r)
        assert got.shape == (B, 1, rd * rd, h, w)
        blk = want[:, l * rd * rd:(l + 1) * rd * rd]
        assert relmax(got[:, 0], blk) < 2e-5, l


def test_alt_bwd_is_adjoint_of_fwd():
    """<alt_fwd(f1,f2), g> differentiated numerically agrees with alt_bwd (both bilinear in f1,f2)."""
    rng = np.random.default_rng(11)
    B, H, W, C, r, N = 1, 6, 7, 8, 2, 2
    f1 = rng.standard_normal((B, H, W, C)).astype(np.float32)
    f2 = rng.standard_normal((B, H, W, C)).astype(np.float32)
    coords = (make_coords("sigma4", B, H, W, rng).transpose(0, 2, 3, 1)[:, None] +
              rng.standard_normal((B, N, H, W, 2)).astype(np.float32))
    g = rng.standard_normal((B, N, (2 * r + 1) ** 2, H, W)).astype(np.float32)
    g1, g2, gc = O.alt_bwd(f1, f2, coords, g, r)
    assert not gc.any()
    # bilinearity: <fwd(d1, f2), g> == <d1, g1> and <fwd(f1, d2), g> == <d2, g2>
    d1 = rng.standard_normal(f1.shape).astype(np.float32)
    d2 = rng.standard_normal(f2.shape).astype(np.float32)
    lhs1 = float((O.alt_fwd(d1, f2, coords, r).astype(np.float64) * g).sum())
    lhs2 = float((O.alt_fwd(f1, d2, coords, r).astype(np.float64) * g).sum())
    assert abs(lhs1 - float((d1.astype(np.float64) * g1).sum())) < 1e-3 * max(1.0, abs(lhs1))
    assert abs(lhs2 - float((d2.astype(np.float64) * g2).sum())) < 1e-3 * max(1.0, abs(lhs2))


def test_lookup_bwd_is_adjoint_of_lookup():
    rng = np.random.default_rng(3)
    B, h, w, L, r = 1, 16, 17, 3, 4
    shapes = O.level_shapes(h, w, L)
    levels = [rng.standard_normal((B * h * w, 1, hl, wl)).astype(np.float32) for hl, wl in shapes]
    coords = make_coords("sigma4", B, h, w, rng)
    g = rng.standard_normal((B, L * (2 * r + 1) ** 2, h, w)).astype(np.float32)
    lhs = float((O.corr_lookup(levels, coords, r).astype(np.float64) * g).sum())
    grads = O.corr_lookup_bwd(g, coords, L, r)
    rhs = sum(float((a.astype(np.float64) * b).sum()) for a, b in zip(levels, grads))
    assert abs(lhs - rhs) < 1e-4 * max(1.0, abs(lhs))


def test_torch_port_matches_oracle():
    torch = pytest.importorskip("torch")
    from oracle import torch_port as TP
    rng = np.random.default_rng(5)
    B, D, h, w, L, r = 1, 24, 18, 21, 4, 4
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    coords = make_coords("sigma4", B, h, w, rng)
    pyr = TP.build_pyramid(torch.from_numpy(f1), torch.from_numpy(f2), L)
    out = TP.lookup(pyr, torch.from_numpy(coords), r).numpy()
    levels = O.corr_build(f1, f2, L)
    for a, b in zip(levels, pyr):
        assert relmax(a, b.numpy()) < 2e-6
    assert relmax(O.corr_lookup(levels, coords, r, roundtrip=True), out) < 2e-6
