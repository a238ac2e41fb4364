"""GPU parity suite (-m gpu): the CUDA path, called through the Python host layer and the
C ABI of libraftcorr.so, against (a) the fixtures produced by the reference's own Python,
(b) the CPU oracle on seeded inputs, (c) size-independent properties at full size.

Tolerances (values cross zero, so always relative to the largest reference magnitude or
normwise, never element-wise relative):
  fp32 build          relmax <= 3e-6   (fp32 FMA vs the reference's SGEMM / double oracle)
  bf16x3 build        relmax <= 1e-4, normwise <= 3e-5   (the "<=1e-4" fp32-class bound)
  bf16 build          normwise <= 6e-3 (single-pass bf16 operands, 8-bit mantissa)
  lookup (same pyramid) relmax <= 2e-5: fp32 coordinate rounding.  The reference (and the oracle)
                      form each tap's coordinate as fl(x + k) and grid_sample round-trips it through
                      [-1,1], which moves it by up to ~2^-23 * |coord| (1.5e-5 px at |coord| = 128);
                      the kernel uses one exact fraction for the whole window.  The result is
                      continuous in the coordinate, so the effect is that shift times the local slope.
"""
import os

import numpy as np
import pytest
import torch

from oracle import corr_oracle as O
import zlib

from tests.helpers import cotangent, make_coords, normwise, relmax

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
BUILD_TOL = {"fp32": dict(rel=3e-6, norm=2e-6), "bf16x3": dict(rel=1e-4, norm=3e-5), "bf16": dict(rel=3e-2, norm=6e-3),
             "fp16_fp8": dict(rel=1e-4, norm=3e-5)}
PRECISIONS = ["fp32", "bf16x3", "bf16"]
LOOKUP_TOL = 2e-5


@pytest.fixture(scope="module")
def rr():
    import raft_ros_b200
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()  # raises if the CUDA library is missing: no fallback
    assert torch.cuda.is_available()
    return raft_ros_b200


def seed_of(*args):
    return zlib.crc32(repr(args).encode())


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def n(x):
    return x.detach().float().cpu().numpy()


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def _kinds(z):
    return sorted(k[len("coords_"):] for k in z.files if k.startswith("coords_"))


GOLDEN = ["corrblock_17x20_r4", "corrblock_16x18_r3"]


# ----------------------------------------------------------------------------- golden fixtures
@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", GOLDEN)
def test_golden_build(rr, golden_dir, name, precision):
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    blk = rr.CorrBlock(t(z["fmap1"]), t(z["fmap2"]), num_levels=L, radius=r, precision=precision)
    assert blk.num_levels == L and blk.radius == r and len(blk.corr_pyramid) == L
    tol = BUILD_TOL[precision]
    for l in range(L):
        ref = z[f"level{l}"]
        got = blk.corr_pyramid[l]
        assert tuple(got.shape) == ref.shape and got.dtype == torch.float32 and got.is_contiguous()
        assert relmax(n(got), ref) <= tol["rel"], (name, l, precision)
        assert normwise(n(got), ref) <= tol["norm"], (name, l, precision)


@pytest.mark.parametrize("name", GOLDEN)
def test_golden_lookup_on_reference_pyramid(rr, golden_dir, name):
    """Lookup alone: feed the reference's own pyramid, compare with the reference's output."""
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    from raft_ros_b200.corr import lookup_raw
    levels = [t(z[f"level{l}"]) for l in range(L)]
    for kind in _kinds(z):
        out = lookup_raw(levels, t(z[f"coords_{kind}"]), r)
        ref = z[f"out_{kind}"]
        assert tuple(out.shape) == ref.shape and out.dtype == torch.float32 and out.is_contiguous()
        assert relmax(n(out), ref) <= LOOKUP_TOL, (name, kind)
        # and against the oracle in the same (pixel-space) arithmetic
        orc = O.corr_lookup([z[f"level{l}"] for l in range(L)], z[f"coords_{kind}"], r, roundtrip=False)
        assert relmax(n(out), orc) <= LOOKUP_TOL, (name, kind)


@pytest.mark.parametrize("name", GOLDEN)
def test_golden_end_to_end_and_backward(rr, golden_dir, name):
    """CorrBlock(fmaps)(coords) and autograd through it vs the reference's autograd."""
    z = _load(golden_dir, name)
    L, r = int(z["num_levels"]), int(z["radius"])
    f1 = t(z["fmap1"]).requires_grad_(True)
    f2 = t(z["fmap2"]).requires_grad_(True)
    blk = rr.CorrBlock(f1, f2, num_levels=L, radius=r, precision="fp32")
    total = None
    for kind in _kinds(z):
        out = blk(t(z[f"coords_{kind}"]))
        assert relmax(n(out), z[f"out_{kind}"]) <= LOOKUP_TOL, (name, kind)
        g = t(cotangent(name, kind, z[f"out_{kind}"].shape))
        total = (out * g).sum() if total is None else total + (out * g).sum()
    total.backward()
    assert relmax(n(f1.grad), z["dfmap1"]) <= 1e-5
    assert relmax(n(f2.grad), z["dfmap2"]) <= 1e-5


# ----------------------------------------------------------------------------- seeded vs oracle
SHAPES = [
    # (B, D, h, w, L, r)  -- ragged against every tile size: 8x16 / 8x32 patches, 32-pixel tiles
    (2, 64, 27, 40, 4, 4),
    (1, 256, 23, 39, 3, 4),     # odd everywhere; D of the real network
    (3, 32, 16, 16, 4, 3),      # exact multiples, raft-small radius
    (1, 128, 33, 67, 4, 4),     # one past multiples
    (1, 48, 9, 130, 2, 2),      # wide and flat, D not a multiple of 64
]


@pytest.mark.parametrize("precision", PRECISIONS + ["fp16_fp8"])
@pytest.mark.parametrize("shape", SHAPES)
def test_build_vs_oracle(rr, shape, precision):
    B, D, h, w, L, r = shape
    rng = np.random.default_rng(seed_of(shape))
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    want = O.corr_build(f1, f2, L)
    blk = rr.CorrBlock(t(f1), t(f2), num_levels=L, radius=r, precision=precision)
    tol = BUILD_TOL[precision]
    for l in range(L):
        got = n(blk.corr_pyramid[l])
        assert got.shape == want[l].shape
        assert np.isfinite(got).all()
        assert relmax(got, want[l]) <= tol["rel"], (shape, l, relmax(got, want[l]))
        assert normwise(got, want[l]) <= tol["norm"], (shape, l, normwise(got, want[l]))


def test_fp16_fp8_build_is_scale_free_per_item_and_map(rr):
    """RC_PREC_FP16_FP8 normalises every (feature map, batch item) by a power of two taken from its largest magnitude:
    the error must not depend on the magnitudes (1e-6 ... 1e+5 here, far outside fp16's and e4m3's own ranges, different
    for the two maps and the items of one batch), an all-zero map gives zeros, and an infinity stays non-finite."""
    B, D, h, w, L = 3, 96, 12, 21, 3
    rng = np.random.default_rng(77)
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    m1 = np.array([1e-6, 1.0, 3e4], np.float32).reshape(B, 1, 1, 1)
    m2 = np.array([1e5, 7e-3, 1.0], np.float32).reshape(B, 1, 1, 1)
    f1, f2 = f1 * m1, f2 * m2
    want = O.corr_build(f1, f2, L)
    blk = rr.CorrBlock(t(f1), t(f2), num_levels=L, radius=4, precision="fp16_fp8")
    N = h * w
    for l in range(L):
        got = n(blk.corr_pyramid[l]).reshape(B, N, -1)
        ref = want[l].reshape(B, N, -1)
        for b in range(B):  # per item: a relative bound against the whole batch would hide the small items
            assert relmax(got[b], ref[b]) <= 1e-4, (l, b, relmax(got[b], ref[b]))
            assert normwise(got[b], ref[b]) <= 3e-5, (l, b, normwise(got[b], ref[b]))
    z = rr.CorrBlock(t(np.zeros_like(f1)), t(f2), num_levels=L, radius=4, precision="fp16_fp8")
    assert all(bool((lv == 0).all()) for lv in z.corr_pyramid)
    f1[1, 5, 3, 4] = np.inf
    bad = n(rr.CorrBlock(t(f1), t(f2), num_levels=1, radius=4, precision="fp16_fp8").corr_pyramid[0]).reshape(B, N, -1)
    assert not np.isfinite(bad[1]).all() and np.isfinite(bad[0]).all() and np.isfinite(bad[2]).all()
    with pytest.raises(RuntimeError, match="forward-only"):
        rr.CorrBlock(t(f2).requires_grad_(True), t(f2), num_levels=L, radius=4, precision="fp16_fp8")


def _fuzz_geometries(count=12, seed=20261106):
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < count:
        L = int(rng.integers(1, 5))
        h, w = int(rng.integers(8, 31)), int(rng.integers(8, 51))
        if (h >> (L - 1)) < 1 or (w >> (L - 1)) < 1:
            continue
        out.append((int(rng.integers(1, 4)), int(rng.integers(8, 301)), h, w, L, 4 if len(out) % 3 else 3))
    return out


@pytest.mark.parametrize("geom", _fuzz_geometries())
def test_corrblock_random_geometries(rr, geom):
    """CorrBlock as a user gets it (default precision, tiled pyramid, one build + one lookup) on random
    small geometries -- odd sizes, channel counts that are not multiples of 64, 1 to 4 levels -- against
    the oracle end to end."""
    B, D, h, w, L, radius = geom
    rng = np.random.default_rng(seed_of(geom))
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    coords = make_coords("sigma4", B, h, w, rng)
    want_pyr = O.corr_build(f1, f2, L)
    want = O.corr_lookup(want_pyr, coords, radius)
    blk = rr.CorrBlock(t(f1), t(f2), num_levels=L, radius=radius)
    assert blk._tiled is not None
    got = n(blk(t(coords)))
    assert got.shape == want.shape and np.isfinite(got).all()
    assert relmax(got, want) <= 1e-4, (geom, relmax(got, want))
    for l in range(L):
        assert relmax(n(blk.corr_pyramid[l]), want_pyr[l]) <= BUILD_TOL["bf16x3"]["rel"], (geom, l)


@pytest.mark.parametrize("kind", ["grid", "sigma0.5", "sigma4", "sigma32", "oob", "border", "huge"])
@pytest.mark.parametrize("shape", SHAPES[:4])
def test_lookup_vs_oracle(rr, shape, kind):
    B, D, h, w, L, r = shape
    from raft_ros_b200.corr import lookup_raw
    rng = np.random.default_rng(seed_of(shape, kind))
    shapes = O.level_shapes(h, w, L)
    levels = [rng.standard_normal((B * h * w, 1, hl, wl)).astype(np.float32) for hl, wl in shapes]
    coords = make_coords(kind, B, h, w, rng)
    want = O.corr_lookup(levels, coords, r)
    got = n(lookup_raw([t(lv) for lv in levels], t(coords), r))
    assert got.shape == want.shape
    assert relmax(got, want) <= LOOKUP_TOL, (shape, kind, relmax(got, want))
    if kind == "huge":
        assert np.isfinite(got).all()


@pytest.mark.parametrize("shape", SHAPES[:3])
def test_lookup_bwd_vs_oracle_and_determinism(rr, shape):
    B, D, h, w, L, r = shape
    from raft_ros_b200.corr import lookup_bwd_raw
    rng = np.random.default_rng(99)
    coords = make_coords("sigma4", B, h, w, rng)
    g = rng.standard_normal((B, L * (2 * r + 1) ** 2, h, w)).astype(np.float32)
    want = O.corr_lookup_bwd(g, coords, L, r)
    runs = []
    for _ in range(2):
        grads = [torch.zeros(B * h * w, 1, hl, wl, device=DEV) for hl, wl in O.level_shapes(h, w, L)]
        lookup_bwd_raw(t(g), t(coords), grads, r)
        runs.append([n(x) for x in grads])
    for l in range(L):
        assert relmax(runs[0][l], want[l]) <= 2e-6, (shape, l)
        assert np.array_equal(runs[0][l], runs[1][l]), "lookup backward must be bit-reproducible"
    # accumulation semantics: a second call adds on top
    grads = [torch.zeros(B * h * w, 1, hl, wl, device=DEV) for hl, wl in O.level_shapes(h, w, L)]
    lookup_bwd_raw(t(g), t(coords), grads, r)
    lookup_bwd_raw(t(g), t(coords), grads, r)
    assert relmax(n(grads[0]), 2 * want[0]) <= 2e-6


@pytest.mark.parametrize("precision,tol", [("fp32", 5e-6), ("bf16x3", 5e-5)])
@pytest.mark.parametrize("dims", [(2, 48, 18, 21, 4), (1, 256, 23, 39, 3), (1, 96, 33, 67, 4)])
def test_build_bwd_vs_oracle(rr, dims, precision, tol):
    from raft_ros_b200.corr import build_bwd_raw
    B, D, h, w, L = dims
    rng = np.random.default_rng(5)
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    gl = [rng.standard_normal((B * h * w, 1, hl, wl)).astype(np.float32) for hl, wl in O.level_shapes(h, w, L)]
    w1, w2 = O.corr_build_bwd(gl, f1, f2)
    d1, d2 = build_bwd_raw([t(g) for g in gl], t(f1), t(f2), precision=precision)
    assert relmax(n(d1), w1) <= tol and relmax(n(d2), w2) <= tol, (relmax(n(d1), w1), relmax(n(d2), w2))


@pytest.mark.parametrize("prec", ["fp32", "bf16x3"])
def test_autograd_paths(rr, prec):
    """(a) CorrBlock's own lookups accumulate into one gradient pyramid; (b) user code may also
    differentiate through `corr_pyramid` tensors directly; (c) a second backward pass over a
    retained graph starts from zero again; all against torch autograd of the same math.  fp32 keeps
    the pyramid in the row layout, bf16x3 in the tiled one (`corr_pyramid` = differentiable copies)."""
    from oracle import torch_port as TP
    gtol = 2e-5 if prec == "fp32" else 1e-4
    B, D, h, w, r = 2, 64, 20, 24, 4
    g = torch.Generator(device="cpu").manual_seed(7)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    coords = [rr.coords_grid(B, h, w, DEV) + 2.0 * torch.randn(B, 2, h, w, generator=g).to(DEV) for _ in range(3)]
    cot = [torch.randn(B, 324, h, w, generator=g).to(DEV) for _ in range(3)]

    def loss_of(build, look, a, b):
        pyr = build(a, b)
        tot = (pyr[1] ** 2).sum() * 1e-3 + pyr[3].sum()     # direct use of the pyramid tensors
        for c, ct in zip(coords, cot):
            tot = tot + (look(pyr, c) * ct).sum()
        return tot

    a, b = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    blk = {}

    def our_build(x, y):
        blk["b"] = rr.CorrBlock(x, y, radius=r, precision=prec)
        assert (blk["b"]._state.tiled is not None) == (prec != "fp32")
        return blk["b"].corr_pyramid

    ours = loss_of(our_build, lambda pyr, c: blk["b"](c), a, b)
    ours.backward(retain_graph=True)
    ga1, gb1 = a.grad.clone(), b.grad.clone()
    a.grad = None
    b.grad = None
    ours.backward()
    assert torch.equal(a.grad, ga1) and torch.equal(b.grad, gb1)  # deterministic, re-entrant
    ra, rb = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    ref = loss_of(lambda x, y: TP.build_pyramid(x, y, 4), lambda pyr, c: TP.lookup(pyr, c, r), ra, rb)
    ref.backward()
    assert abs(float(ours) - float(ref)) <= 1e-4 * abs(float(ref))
    assert relmax(n(ga1), n(ra.grad)) <= gtol and relmax(n(gb1), n(rb.grad)) <= gtol
    # a block built for training can also be looked up under no_grad
    with torch.no_grad():
        again = blk["b"](coords[0])
    assert not again.requires_grad and relmax(n(again), n(TP.lookup(TP.build_pyramid(f1, f2, 4), coords[0], r))) <= 1e-4
    # CorrBlock.corr
    a2, b2 = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    v = rr.CorrBlock.corr(a2, b2, precision="bf16x3")
    assert v.shape == (B, h, w, 1, h, w)
    wgt = torch.randn(v.shape, generator=g).to(DEV)
    (v * wgt).sum().backward()
    a3, b3 = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    vr = (torch.matmul(a3.view(B, D, h * w).transpose(1, 2), b3.view(B, D, h * w)) / 8.0).view(B, h, w, 1, h, w)
    (vr * wgt).sum().backward()
    assert relmax(n(a2.grad), n(a3.grad)) <= 5e-5 and relmax(n(b2.grad), n(b3.grad)) <= 5e-5


# ----------------------------------------------------------------------------- alt (on the fly)
ALT_SHAPES = [(2, 1, 13, 18, 13, 18, 64, 4), (1, 2, 9, 11, 4, 5, 32, 3), (1, 1, 16, 16, 8, 8, 256, 4), (1, 1, 7, 33, 7, 33, 36, 2)]


@pytest.mark.parametrize("shape", ALT_SHAPES)
def test_alt_fwd_bwd_vs_oracle(rr, shape):
    B, N, H1, W1, H2, W2, C, r = shape
    rng = np.random.default_rng(seed_of(shape))
    f1 = rng.standard_normal((B, H1, W1, C)).astype(np.float32)
    f2 = rng.standard_normal((B, H2, W2, C)).astype(np.float32)
    base = make_coords("sigma4", B, H1, W1, rng).transpose(0, 2, 3, 1)[:, None] * (W2 / W1)
    coords = (base + rng.standard_normal((B, N, H1, W1, 2))).astype(np.float32)
    coords[0, 0, 0, 0] = (-3.5, 2.25)          # window straddling the border
    coords[0, 0, -1, -1] = (W2 + 40.0, -77.0)  # far outside -> zeros
    (corr,) = rr.alt_cuda_corr.forward(t(f1), t(f2), t(coords), r)
    want = O.alt_fwd(f1, f2, coords, r)
    assert tuple(corr.shape) == want.shape
    assert relmax(n(corr), want) <= 2e-6
    g = rng.standard_normal(want.shape).astype(np.float32)
    g1, g2, gc = rr.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(g), r)
    w1, w2, _ = O.alt_bwd(f1, f2, coords, g, r)
    assert relmax(n(g1), w1) <= 3e-6 and relmax(n(g2), w2) <= 1e-5
    assert tuple(gc.shape) == coords.shape and not bool(gc.any())


def test_alt_backward_fmap2_grad_is_bit_reproducible(rr, monkeypatch):
    """alt_cuda_corr.backward: thousands of query pixels whose windows all land on the same few fmap2 pixels -- the
    case in which a float atomicAdd scatter (the reference's, correlation_kernel.cu:229-238) gives different low bits
    from run to run.  The default path (rc_alt_corr_bwd_det, fixed-point accumulation) must give the same bits every
    time and agree with the oracle; the float-atomics path stays available and agrees to rounding."""
    B, N, H1, W1, H2, W2, C, r = 2, 1, 48, 64, 20, 24, 64, 4
    rng = np.random.default_rng(2025)
    f1 = rng.standard_normal((B, H1, W1, C)).astype(np.float32)
    f2 = rng.standard_normal((B, H2, W2, C)).astype(np.float32)
    coords = np.empty((B, N, H1, W1, 2), np.float32)
    coords[..., 0] = 9.3 + 0.5 * rng.standard_normal((B, N, H1, W1))
    coords[..., 1] = 7.7 + 0.5 * rng.standard_normal((B, N, H1, W1))
    g = rng.standard_normal((B, N, 81, H1, W1)).astype(np.float32)
    w1, w2, _ = O.alt_bwd(f1, f2, coords, g, r)
    runs = [rr.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(g), r) for _ in range(3)]
    torch.cuda.synchronize()
    for g1, g2, gc in runs:
        assert relmax(n(g1), w1) <= 3e-6 and relmax(n(g2), w2) <= 3e-6, (relmax(n(g1), w1), relmax(n(g2), w2))
    assert all(torch.equal(runs[0][0], x[0]) and torch.equal(runs[0][1], x[1]) for x in runs[1:])
    monkeypatch.setenv("RAFT_ALT_BWD_ATOMICS", "1")
    a1, a2, _ = rr.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(g), r)
    assert torch.equal(a1, runs[0][0]) and relmax(n(a2), w2) <= 1e-5
    # all-zero gradients (scale undefined) and a workspace that is too small
    z = rr.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(np.zeros_like(g)), r)
    assert not bool(z[1].any())
    from raft_ros_b200 import _lib
    L = _lib.lib()
    x = torch.zeros(1024, device=DEV)
    rc = L.rc_alt_corr_bwd_det(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), B, N, H1, W1,
                               H2, W2, C, r, x.data_ptr(), 1024, None)
    assert rc == 4 and b"workspace" in L.rc_last_error()


ALT_TOL = {"fp32": dict(rel=5e-6, norm=3e-6), "bf16x3": dict(rel=1e-4, norm=3e-5), "bf16": dict(rel=3e-2, norm=6e-3),
           "fp16_fp8": dict(rel=1e-4, norm=3e-5)}


_ALT_WANT = {}


def _alt_want(geom, kind):
    """What the lookup must give, from the checker and never from this repository's own kernels:
    the C oracle (build + lookup, core/corr.py:13-50) where it finishes in seconds, the reference's
    torch ops (oracle/torch_port.py, TF32 off) on the GPU at 55x128."""
    key = (geom, kind)
    if key not in _ALT_WANT:
        B, D, h, w, L = geom
        rng = np.random.default_rng(seed_of(geom, kind))
        f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
        f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
        coords = make_coords(kind, B, h, w, rng)
        if h * w <= 1200:
            want = O.corr_lookup(O.corr_build(f1, f2, L), coords, 4)
        else:
            from oracle import torch_port as TP
            torch.backends.cuda.matmul.allow_tf32 = False
            with torch.no_grad():
                want = n(TP.lookup(TP.build_pyramid(t(f1), t(f2), L), t(coords), 4))
        _ALT_WANT[key] = (f1, f2, coords, want)
    return _ALT_WANT[key]


@pytest.mark.parametrize("precision", ["fp32", "bf16x3", "bf16", "fp16_fp8"])
@pytest.mark.parametrize("geom", [(2, 64, 24, 32, 4), (1, 256, 27, 40, 4), (1, 96, 23, 39, 3), (1, 256, 55, 128, 4)])
@pytest.mark.parametrize("kind", ["sigma4", "sigma32", "grid", "oob", "huge"])
def test_alternate_block_vs_oracle(rr, geom, kind, precision):
    """AlternateCorrBlock(coords) against the oracle's CorrBlock restatement (the two are equal by
    reference README.md:109-114), for the per-level CUDA-core kernel (fp32) and the one-launch
    tensor-core kernel (bf16x3 / bf16), coherent and incoherent flow, ragged tiles, windows off
    the map."""
    B, D, h, w, L = geom
    if kind in ("grid", "oob", "huge") and geom != (1, 256, 27, 40, 4):
        pytest.skip("edge-case coordinates are exercised on one geometry")
    f1n, f2n, cn, want = _alt_want(geom, kind)
    f1, f2, coords = t(f1n), t(f2n), t(cn)
    blk = rr.AlternateCorrBlock(f1, f2, num_levels=L, radius=4, precision=precision)
    assert len(blk.pyramid) == L + 1 and blk.pyramid[1][1].shape == (B, D, h // 2, w // 2)
    b = blk(coords)
    assert tuple(b.shape) == want.shape == (B, L * 81, h, w) and b.dtype == torch.float32 and b.is_contiguous()
    assert torch.isfinite(b).all()
    tol = ALT_TOL[precision]
    # + the lookup's own coordinate-rounding allowance (LOOKUP_TOL) on top of the contraction's bound
    assert relmax(n(b), want) <= tol["rel"] + LOOKUP_TOL, (geom, kind, precision, relmax(n(b), want))
    assert normwise(n(b), want) <= tol["norm"] + LOOKUP_TOL, (geom, kind, precision, normwise(n(b), want))
    assert torch.equal(blk(coords), b)  # deterministic


@pytest.mark.parametrize("precision", ["bf16x3", "bf16", "fp16_fp8"])
@pytest.mark.parametrize("geom", [(2, 128, 24, 32, 4), (1, 128, 23, 39, 3), (1, 96, 46, 62, 4)])
def test_alternate_block_radius3(rr, geom, precision):
    """raft-small (core/raft.py:33: corr_radius = 3, 128 feature channels) on the tensor-core on-the-fly kernel,
    against the oracle's CorrBlock restatement."""
    B, D, h, w, L = geom
    rng = np.random.default_rng(seed_of(geom, "r3"))
    f1 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2 = rng.standard_normal((B, D, h, w)).astype(np.float32)
    blk = rr.AlternateCorrBlock(t(f1), t(f2), num_levels=L, radius=3, precision=precision)
    assert blk._tc
    pyr = O.corr_build(f1, f2, L)
    tol = ALT_TOL[precision]
    for kind in ("sigma4", "border", "oob", "grid"):
        coords = make_coords(kind, B, h, w, rng)
        want = O.corr_lookup(pyr, coords, 3)
        got = n(blk(t(coords)))
        assert got.shape == want.shape == (B, L * 49, h, w) and np.isfinite(got).all()
        assert relmax(got, want) <= tol["rel"] + LOOKUP_TOL, (geom, kind, precision, relmax(got, want))
        assert normwise(got, want) <= tol["norm"] + LOOKUP_TOL, (geom, kind, precision)


@pytest.mark.parametrize("shape", [(2, 96, 27, 40), (1, 256, 55, 128), (1, 40, 8, 8)])
def test_alt_prep_pools_like_avg_pool2d(rr, shape):
    """rc_alt_prep_fmap_level forms pyramid level l of a feature map in registers; its hi / lo outputs are
    bit-identical to converting the reference's avg_pool2d chain (core/corr.py:68-72)."""
    from raft_ros_b200 import _lib
    lib = _lib.lib()
    B, C, H, W = shape
    g = torch.Generator().manual_seed(H * W)
    f = (torch.randn(B, C, H, W, generator=g) * 3).to(DEV)
    Cp = (C + 63) // 64 * 64
    st = torch.cuda.current_stream().cuda_stream
    pooled = f
    for level in range(4):
        if level:
            pooled = torch.nn.functional.avg_pool2d(pooled, 2, stride=2)
        Hl, Wl = H >> level, W >> level
        assert pooled.shape == (B, C, Hl, Wl)
        outs = []
        for src, fn, extra in ((f, lib.rc_alt_prep_fmap_level, (B, C, H, W, level)),
                               (pooled.contiguous(), lib.rc_alt_prep_fmap, (B, C, Hl, Wl))):
            hi = torch.full((B, Hl, Wl, Cp), float("nan"), dtype=torch.bfloat16, device=DEV)
            lo = torch.full_like(hi, float("nan"))
            assert fn(src.data_ptr(), hi.data_ptr(), lo.data_ptr(), *extra, st) == 0, lib.rc_last_error()
            outs.append((hi, lo))
        torch.cuda.synchronize()
        assert torch.isfinite(outs[0][0].float()).all()
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]), (shape, level)


@pytest.mark.parametrize("precision", ["bf16x3", "bf16", "fp16_fp8"])
def test_alternate_block_query_row_band(rr, precision):
    """A band of query rows against the full fmap2 (how one large pair is split over GPUs):
    bit-identical to the same rows of the whole-frame block when the band starts on a (16-row) tile row, and within
    rounding of it otherwise (the box of a query tile, and with it the summation split, depends on the tiling)."""
    B, D, H, W, L = 1, 128, 43, 70, 4
    rng = np.random.default_rng(43)
    f1 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    coords = t(make_coords("sigma4", B, H, W, rng))
    full = rr.AlternateCorrBlock(f1, f2, num_levels=L, radius=4, precision=precision)(coords)
    for r0, r1 in ((0, 16), (16, 43), (32, 43), (11, 30), (8, 24)):
        band = rr.AlternateCorrBlock(f1[:, :, r0:r1].contiguous(), f2, num_levels=L, radius=4, precision=precision)
        got = band(coords[:, :, r0:r1].contiguous())
        want = full[:, :, r0:r1]
        assert got.shape == want.shape
        if r0 % 16 == 0:  # the query tile is 16 rows (a CTA pair)
            assert torch.equal(got, want), (r0, r1)
        else:
            assert relmax(n(got), n(want)) <= 2e-6, (r0, r1, relmax(n(got), n(want)))


# ----------------------------------------------------------------------------- bilinear_sampler
def test_bilinear_sampler_vs_grid_sample(rr):
    g = torch.Generator(device="cpu").manual_seed(3)
    img = torch.randn(5, 3, 11, 17, generator=g).to(DEV).requires_grad_(True)
    coords = torch.rand(5, 9, 9, 2, generator=g) * torch.tensor([20.0, 14.0]) - 2.0
    # samples exactly on, and one ulp either side of, the four borders
    edge = torch.tensor([0.0, 16.0, 10.0, 1e-45, -1e-45, 15.999998, 16.000002, 9.999999, 10.000001])
    coords[0, 0, :, 0] = edge
    coords[0, 1, :, 1] = edge
    coords = coords.to(DEV)
    out, mask = rr.bilinear_sampler(img, coords, mask=True)
    H, W = 11, 17
    grid = torch.stack([2 * coords[..., 0] / (W - 1) - 1, 2 * coords[..., 1] / (H - 1) - 1], -1)
    ref = torch.nn.functional.grid_sample(img.detach(), grid, align_corners=True)
    assert relmax(n(out), n(ref)) <= 5e-6
    # mask (utils.py:67-69): the kernel evaluates the reference's expression on the normalised grid with
    # IEEE fp32 operations, i.e. exactly what the reference computes on the CPU ...
    cc = coords.cpu()
    gx, gy = 2 * cc[..., :1] / (W - 1) - 1, 2 * cc[..., 1:] / (H - 1) - 1
    cpumask = ((gx > -1) & (gy > -1) & (gx < 1) & (gy < 1)).float()
    assert mask.shape == cpumask.shape and torch.equal(mask.cpu(), cpumask)
    # ... and what it computes on the GPU wherever that is not within an ulp of a border (ATen's CUDA
    # division by a scalar multiplies by the reciprocal, which may round the other way there)
    refmask = ((grid[..., :1] > -1) & (grid[..., 1:] > -1) & (grid[..., :1] < 1) & (grid[..., 1:] < 1)).float()
    away = ((grid.abs() - 1).abs() > 1e-6).all(-1, keepdim=True)
    assert torch.equal(mask[away], refmask[away])
    go = torch.randn(out.shape, generator=g).to(DEV)
    out.backward(go)
    img2 = img.detach().clone().requires_grad_(True)
    torch.nn.functional.grid_sample(img2, grid, align_corners=True).backward(go)
    assert relmax(n(img.grad), n(img2.grad)) <= 1e-5


# ----------------------------------------------------------------------------- full-size properties
FULL = [(1, 256, 55, 128), (1, 256, 47, 156), (1, 256, 50, 90)]


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("shape", FULL)
def test_full_size_properties(rr, shape, precision):
    """BASELINE.json shapes: volume vs an fp32 torch contraction (TF32 off), pooling consistency,
    transpose symmetry, lookup identities.  The oracle is too slow here; these are its
    size-independent consequences."""
    B, D, h, w = shape
    N = h * w
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cpu").manual_seed(1234)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    blk = rr.CorrBlock(f1, f2, num_levels=4, radius=4, precision=precision)
    ref0 = (torch.matmul(f1.view(B, D, N).double().transpose(1, 2), f2.view(B, D, N).double()) / 16.0).float()
    tol = BUILD_TOL[precision]
    l0 = blk.corr_pyramid[0].view(B, N, N)
    err = float((l0 - ref0).abs().max() / ref0.abs().max())
    nerr = float((l0 - ref0).norm() / ref0.norm())
    assert err <= tol["rel"] and nerr <= tol["norm"], (err, nerr)
    # pooling consistency: level l+1 == floor-mode 2x2 mean of level l (core/corr.py:26)
    for l in range(3):
        pooled = torch.nn.functional.avg_pool2d(blk.corr_pyramid[l], 2, stride=2)
        assert tuple(pooled.shape) == tuple(blk.corr_pyramid[l + 1].shape)
        assert float((pooled - blk.corr_pyramid[l + 1]).abs().max()) <= 2e-6 * float(ref0.abs().max())
    # transpose symmetry: corr(f1,f2)[p,q] == corr(f2,f1)[q,p]
    blk_t = rr.CorrBlock(f2, f1, num_levels=1, radius=4, precision=precision)
    assert torch.equal(blk_t.corr_pyramid[0].view(B, N, N).transpose(1, 2), l0) or \
        float((blk_t.corr_pyramid[0].view(B, N, N).transpose(1, 2) - l0).abs().max()) <= 2e-6 * float(ref0.abs().max())
    # lookup at integer coords is a pure gather of level 0 (channel i*9+j <-> (x+i-4, y+j-4))
    coords = rr.coords_grid(B, h, w, DEV)
    out = blk(coords)
    assert out.shape == (B, 324, h, w) and out.dtype == torch.float32 and out.is_contiguous()
    vol = blk.corr_pyramid[0].view(B, h, w, h, w)
    for (i, j) in [(4, 4), (0, 0), (8, 8), (2, 7)]:
        ys = torch.arange(h, device=DEV).view(h, 1).expand(h, w)
        xs = torch.arange(w, device=DEV).view(1, w).expand(h, w)
        yy, xx = ys + (j - 4), xs + (i - 4)
        ok = (yy >= 0) & (yy < h) & (xx >= 0) & (xx < w)
        want = torch.where(ok, vol[0, ys, xs, yy.clamp(0, h - 1), xx.clamp(0, w - 1)], torch.zeros((), device=DEV))
        assert torch.equal(out[0, i * 9 + j], want), (i, j)
    # far out-of-bounds coords give exact zeros; idempotence of repeated lookups
    far = coords + 10000.0
    assert not bool(blk(far).any())
    assert torch.equal(blk(coords), out)


# ----------------------------------------------------------------------------- host-side errors
def test_error_conventions(rr):
    f = torch.zeros(1, 8, 8, 32, device=DEV)
    c = torch.zeros(1, 1, 8, 8, 2, device=DEV)
    with pytest.raises(RuntimeError, match="fmap2 must be contiguous"):
        rr.alt_cuda_corr.forward(f, f.permute(0, 2, 1, 3), c, 4)
    with pytest.raises(RuntimeError, match="coords must be a CUDA tensor"):
        rr.alt_cuda_corr.forward(f, f, c.cpu(), 4)
    with pytest.raises(RuntimeError, match="empty"):
        rr.CorrBlock(torch.zeros(1, 32, 4, 64, device=DEV), torch.zeros(1, 32, 4, 64, device=DEV), num_levels=4)
    blk = rr.CorrBlock(torch.zeros(1, 32, 16, 16, device=DEV), torch.zeros(1, 32, 16, 16, device=DEV), precision="fp32")
    with pytest.raises(RuntimeError):
        blk(torch.zeros(1, 2, 8, 8, device=DEV))  # coords grid does not match the pyramid
    with pytest.raises(RuntimeError, match="radius"):
        rr.CorrBlock(torch.zeros(1, 32, 16, 16, device=DEV), torch.zeros(1, 32, 16, 16, device=DEV), radius=9, precision="fp32")(
            torch.zeros(1, 2, 16, 16, device=DEV))


# ----------------------------------------------------------------------------- the reference's own binary
def _ref_ext():
    from oracle import build_ref
    try:
        return build_ref.load_module()
    except Exception:  # built for another torch ABI, etc.
        return None


@pytest.mark.parametrize("shape", ALT_SHAPES[:3])
def test_alt_vs_reference_binary(rr, shape):
    """rc_alt_corr_fwd/bwd against the reference's UNMODIFIED alt_cuda_corr extension
    (oracle/_ref/, compiled from the reference's sources by oracle/build_ref.py).  C must be a
    multiple of 32 for the reference kernel; H1/W1 multiples of the reference's 4x8 block avoid
    its out-of-bounds coordinate read in the backward (correlation_kernel.cu:178-179)."""
    ext = _ref_ext()
    if ext is None:
        pytest.skip("oracle/_ref/alt_cuda_corr_ref.so not present")
    B, N, H1, W1, H2, W2, C, r = shape
    H1, W1 = (H1 + 3) // 4 * 4, (W1 + 7) // 8 * 8
    g = torch.Generator(device="cpu").manual_seed(seed_of(shape))
    f1 = torch.randn(B, H1, W1, C, generator=g).to(DEV)
    f2 = torch.randn(B, H2, W2, C, generator=g).to(DEV)
    coords = (torch.rand(B, N, H1, W1, 2, generator=g) * torch.tensor([W2 + 6.0, H2 + 6.0]) - 3.0).to(DEV)
    (want,) = ext.forward(f1, f2, coords, r)
    (got,) = rr.alt_cuda_corr.forward(f1, f2, coords, r)
    torch.cuda.synchronize()
    assert got.shape == want.shape
    assert relmax(n(got), n(want)) <= 5e-6
    cg = torch.randn(want.shape, generator=g).to(DEV)
    w1, w2, wc = ext.backward(f1, f2, coords, cg, r)
    g1, g2, gc = rr.alt_cuda_corr.backward(f1, f2, coords, cg, r)
    torch.cuda.synchronize()
    assert relmax(n(g1), n(w1)) <= 1e-5 and relmax(n(g2), n(w2)) <= 1e-5
    assert not bool(wc.any()) and not bool(gc.any())


# ----------------------------------------------------------------------------- guard bands
@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("geom", [(1, 96, 23, 39, 4), (2, 64, 17, 20, 4), (1, 256, 9, 130, 2)])
def test_no_writes_outside_outputs(rr, geom, precision):
    """compute-sanitizer is closed on this GPU pool, so out-of-bounds WRITES are checked with
    guard bands: every output lives between two sentinel regions that must stay untouched, and
    every output element must be overwritten (outputs are pre-filled with NaN)."""
    from raft_ros_b200 import _lib
    L_ = _lib.lib()
    B, D, h, w, nl = geom
    N, G = h * w, 4096
    rng = np.random.default_rng(seed_of(geom))
    f1 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    sizes = [B * N * (h >> l) * (w >> l) for l in range(nl)]
    padded = [(s + 63) // 64 * 64 for s in sizes]
    SENT = 1234.5
    buf = torch.full((G * (nl + 1) + sum(padded),), SENT, device=DEV)
    levels, off = [], G
    for s, pz in zip(sizes, padded):
        lv = buf[off:off + s]
        lv.fill_(float("nan"))
        levels.append(lv)
        off += pz + G
    prec = {"fp32": 0, "bf16x3": 1, "bf16": 2}[precision]
    nbytes = int(L_.rc_corr_build_workspace_bytes(B, D, h, w, prec))
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=DEV)
    rc = L_.rc_corr_build_pyramid(f1.data_ptr(), f2.data_ptr(), B, D, h, w, _lib.ptr_array([x.data_ptr() for x in levels]),
                                  nl, prec, 0.0, ws.data_ptr() if nbytes else None, nbytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L_.rc_last_error()
    torch.cuda.synchronize()
    for lv in levels:
        assert torch.isfinite(lv).all(), "an output element was not written"
    mask = torch.ones_like(buf, dtype=torch.bool)
    off = G
    for s, pz in zip(sizes, padded):
        mask[off:off + s] = False
        off += pz + G
    assert bool((buf[mask] == SENT).all()), "write outside the pyramid levels"
    # lookup output and on-the-fly output, same scheme
    rd2 = 81
    obuf = torch.full((2 * G + B * nl * rd2 * N,), SENT, device=DEV)
    out = obuf[G:G + B * nl * rd2 * N]
    out.fill_(float("nan"))
    coords = t(make_coords("sigma4", B, h, w, rng))
    rc = L_.rc_corr_lookup_fwd(_lib.ptr_array([x.data_ptr() for x in levels]), coords.data_ptr(), out.data_ptr(), B, h, w, nl, 4,
                               torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L_.rc_last_error()
    torch.cuda.synchronize()
    assert torch.isfinite(out).all() and bool((obuf[:G] == SENT).all()) and bool((obuf[G + out.numel():] == SENT).all())


# ----------------------------------------------------------------------------- tiled volume layout
def _tile_levels(levels, B):
    """(B*N,1,hl,wl) rows -> [B][ceil(N/G)][ceil2(hl)/2][ceil8(wl)/8][G][4 swizzled 16 B chunks] zero-padded,
    pixel-interleaved tiles (test-side conversion in torch; layout documented in include/raftcorr.h)."""
    from raft_ros_b200 import _lib
    G = int(_lib.lib().rc_corr_tile_group())
    out = []
    for lv in levels:
        P, _, hl, wl = lv.shape
        N = P // B
        hp, wp, ng = (hl + 1) // 2 * 2, (wl + 7) // 8 * 8, (N + G - 1) // G
        x = torch.nn.functional.pad(lv.view(B, N, hl, wl), (0, wp - wl, 0, hp - hl, 0, ng * G - N))
        x = x.view(B, ng, G, hp // 2, 2, wp // 8, 2, 4).permute(0, 1, 3, 5, 2, 4, 6, 7)  # b, g, ty, tx, u, r, half, c
        x = x.reshape(B, ng, hp // 2, wp // 8, G, 4, 4)  # chunk k = 2 r + half
        out.append(x.gather(5, _chunk_perm(G).view(1, 1, 1, 1, G, 4, 1).expand_as(x)).contiguous().view(-1))
    return out


def _chunk_perm(G):
    """position -> chunk stored there, per pixel u of a group: k = pos ^ ((u >> 1) & 3) (an involution)."""
    u = torch.arange(G, device=DEV).view(G, 1)
    return torch.arange(4, device=DEV).view(1, 4) ^ ((u >> 1) & 3)


def _untile(flat, B, N, hl, wl):
    from raft_ros_b200 import _lib
    G = int(_lib.lib().rc_corr_tile_group())
    hp, wp, ng = (hl + 1) // 2 * 2, (wl + 7) // 8 * 8, (N + G - 1) // G
    x = flat.view(B, ng, hp // 2, wp // 8, G, 4, 4)
    x = x.gather(5, _chunk_perm(G).view(1, 1, 1, 1, G, 4, 1).expand_as(x))  # undo the chunk swizzle
    x = x.view(B, ng, hp // 2, wp // 8, G, 2, 8).permute(0, 1, 4, 2, 5, 3, 6).reshape(B, ng * G, hp, wp)
    return x  # padded planes, query pixel major


@pytest.mark.parametrize("shape", [(1, 7, 5, 9), (2, 16, 16, 32), (1, 5, 1, 1), (3, 11, 23, 39)])
def test_relayout_level_round_trip(rr, shape):
    """rc_corr_relayout_level: rows -> tiled equals the torch conversion (padding zeroed), and back is exact."""
    from raft_ros_b200.corr import relayout_raw
    B, N, hl, wl = shape
    rng = np.random.default_rng(seed_of(shape))
    rows = t(rng.standard_normal((B * N, 1, hl, wl)).astype(np.float32))
    tiled = relayout_raw(rows, B, N, hl, wl, to_tiled=True)
    assert torch.equal(tiled.view(-1), _tile_levels([rows], B)[0])
    assert torch.equal(relayout_raw(tiled, B, N, hl, wl, to_tiled=False), rows)


TILED_GEOMS = [(1, 64, 8, 32, 4), (2, 256, 27, 40, 4), (1, 96, 23, 39, 3), (1, 256, 47, 156, 4), (1, 128, 17, 20, 4),
               (2, 256, 16, 18, 2), (1, 64, 9, 65, 1), (1, 256, 55, 128, 4)]


@pytest.mark.parametrize("prec", ["bf16x3", "bf16", "fp16_fp8"])
@pytest.mark.parametrize("geom", TILED_GEOMS)
def test_tiled_build_equals_row_build(rr, geom, prec):
    """rc_corr_build_pyramid_tiled writes, tile by tile, exactly the numbers rc_corr_build_pyramid writes
    row by row (same accumulators, same pooling order), zeros in the padding a lookup can reach, and the
    tiled lookup on it equals the row lookup -- with every buffer pre-filled with NaN, so anything the
    kernels read but never wrote would show."""
    from raft_ros_b200 import _lib
    from raft_ros_b200.corr import build_pyramid_raw, lookup_raw, relayout_raw, tiled_level_elems, _PRECISIONS
    B, D, h, w, L = geom
    rng = np.random.default_rng(seed_of(geom, prec))
    f1 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    rows = build_pyramid_raw(f1, f2, L, prec)
    lib = _lib.lib()
    shapes = O.level_shapes(h, w, L)
    N = h * w
    tiled = [torch.full((tiled_level_elems(B, N, hl, wl),), float("nan"), device=DEV) for hl, wl in shapes]
    code = _PRECISIONS[prec]
    ws_bytes = int(lib.rc_corr_build_workspace_bytes(B, D, h, w, code))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    rc = lib.rc_corr_build_pyramid_tiled(f1.data_ptr(), f2.data_ptr(), B, D, h, w,
                                         _lib.ptr_array([x.data_ptr() for x in tiled]), L, code, 0.0, ws.data_ptr(),
                                         ws_bytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.rc_last_error()
    torch.cuda.synchronize()
    for l, (hl, wl) in enumerate(shapes):
        back = relayout_raw(tiled[l], B, N, hl, wl, to_tiled=False)
        assert torch.equal(back, rows[l]), (geom, prec, l)
        grid = _untile(tiled[l], B, N, hl, wl)
        wb = (wl + 3) // 4 * 4
        assert bool((grid[:, :N, :hl, wl:wb] == 0).all()), (geom, prec, l, "padding columns a lookup fetches must be zero")
    for kind in ("sigma4", "border", "oob"):
        coords = t(make_coords(kind, B, h, w, rng))
        out = torch.full((B, L * 81, h, w), float("nan"), device=DEV)
        rc = lib.rc_corr_lookup_tiled_fwd(_lib.ptr_array([x.data_ptr() for x in tiled]), coords.data_ptr(),
                                          out.data_ptr(), B, h, w, L, 4, torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.rc_last_error()
        want = lookup_raw(rows, coords, 4)
        assert torch.isfinite(out).all(), (geom, prec, kind)
        assert relmax(n(out), n(want)) <= 2e-6, (geom, prec, kind, relmax(n(out), n(want)))


@pytest.mark.parametrize("geom", [(1, 64, 8, 32, 4), (2, 256, 27, 40, 4), (1, 96, 23, 39, 3), (1, 256, 47, 156, 4),
                                  (1, 128, 17, 20, 4), (2, 200, 16, 18, 2), (1, 64, 9, 65, 1), (1, 256, 55, 128, 4),
                                  (1, 256, 50, 90, 4), (1, 32, 40, 160, 4)])
def test_vol16_build_and_lookup(rr, geom):
    """precision 'bf16_vol16': the pyramid stored as fp16 in the tiled layout.  (a) Untiled, it equals
    the 'bf16' build rounded to fp16 (same accumulators, same pooling order, one rounding at the store);
    (b) the fp16 lookup equals the fp32-layout lookup on exactly those rounded values; (c) against the
    oracle the error stays within the bf16 bound plus the fp16 storage rounding; (d) buffers are
    pre-filled with NaN so that anything read but never written would show."""
    from raft_ros_b200 import _lib
    from raft_ros_b200.corr import build_pyramid_raw, lookup_raw, relayout_raw, tiled_level_elems
    B, D, h, w, L = geom
    rng = np.random.default_rng(seed_of(geom, "vol16"))
    f1n = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f2n = rng.standard_normal((B, D, h, w)).astype(np.float32)
    f1, f2 = t(f1n), t(f2n)
    rows = build_pyramid_raw(f1, f2, L, "bf16")
    lib = _lib.lib()
    shapes = O.level_shapes(h, w, L)
    N = h * w
    tiled = [torch.full((tiled_level_elems(B, N, hl, wl, True),), float("nan"), dtype=torch.float16, device=DEV)
             for hl, wl in shapes]
    ws_bytes = int(lib.rc_corr_build_workspace_bytes(B, D, h, w, 3))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    rc = lib.rc_corr_build_pyramid_tiled(f1.data_ptr(), f2.data_ptr(), B, D, h, w,
                                         _lib.ptr_array([x.data_ptr() for x in tiled]), L, 3, 0.0, ws.data_ptr(),
                                         ws_bytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.rc_last_error()
    torch.cuda.synchronize()
    rounded = []
    for l, (hl, wl) in enumerate(shapes):
        back = relayout_raw(tiled[l], B, N, hl, wl, to_tiled=False, vol16=True)
        want = rows[l].half().float()
        assert torch.equal(back, want), (geom, l, float((back - want).abs().max()))
        rounded.append(want)
        # and the converter's own rows -> tiled direction agrees with what the build wrote (where written)
        again = relayout_raw(rows[l], B, N, hl, wl, to_tiled=True, vol16=True)
        assert torch.equal(relayout_raw(again, B, N, hl, wl, to_tiled=False, vol16=True), want)
    for kind in ("sigma4", "border", "oob", "grid"):
        coords = t(make_coords(kind, B, h, w, rng))
        out = torch.full((B, L * 81, h, w), float("nan"), device=DEV)
        rc = lib.rc_corr_lookup_tiled16_fwd(_lib.ptr_array([x.data_ptr() for x in tiled]), coords.data_ptr(),
                                            out.data_ptr(), B, h, w, L, 4, torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.rc_last_error()
        want = lookup_raw(rounded, coords, 4)
        assert torch.isfinite(out).all(), (geom, kind)
        assert relmax(n(out), n(want)) <= 2e-6, (geom, kind, relmax(n(out), n(want)))
    want_pyr = O.corr_build(f1n, f2n, L)
    for l in range(L):
        assert normwise(n(rounded[l]), want_pyr[l]) <= BUILD_TOL["bf16"]["norm"] + 5e-4, (geom, l)


def test_vol16_corrblock_and_limits(rr):
    """CorrBlock(precision='bf16_vol16'): fp16 tiled pyramid inside, fp32 `corr_pyramid` outside; training,
    other radii and AlternateCorrBlock refuse it; values beyond the fp16 range saturate instead of overflowing."""
    rng = np.random.default_rng(5)
    B, D, h, w = 1, 64, 24, 40
    f1 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    coords = t(make_coords("sigma4", B, h, w, rng))
    blk = rr.CorrBlock(f1, f2, precision="bf16_vol16")
    assert blk._tiled[0].dtype == torch.float16
    ref = rr.CorrBlock(f1, f2, precision="bf16")
    assert relmax(n(blk(coords)), n(ref(coords))) <= 1e-3
    assert blk.corr_pyramid[0].dtype == torch.float32 and blk.corr_pyramid[0].shape == ref.corr_pyramid[0].shape
    with pytest.raises(RuntimeError, match="bf16_vol16"):
        rr.CorrBlock(f1, f2, radius=2, precision="bf16_vol16")
    with pytest.raises(RuntimeError, match="bf16_vol16"):
        rr.CorrBlock(f1.clone().requires_grad_(True), f2, precision="bf16_vol16")
    with pytest.raises(ValueError, match="bf16_vol16"):
        rr.AlternateCorrBlock(f1, f2, precision="bf16_vol16")
    big = rr.CorrBlock(f1 * 3e3, f2 * 3e3, precision="bf16_vol16")  # |corr| up to ~1e7 > 65504
    lv0 = big.corr_pyramid[0]
    assert torch.isfinite(lv0).all() and float(lv0.abs().max()) == 65504.0
    assert torch.isfinite(big(coords)).all()


def test_tiled_equals_rows_at_1080p(rr):
    """A 1080p pair (feature grid 135x240, 32 400 query pixels, 5.6 GB pyramid, odd height): the tiled
    pipeline and the row-layout pipeline agree -- pyramid bit for bit, lookups to fp32 rounding -- so the
    64-bit offsets of both layouts hold beyond 2^31 bytes."""
    from raft_ros_b200.corr import build_pyramid_raw, build_pyramid_tiled_raw, lookup_raw, lookup_tiled_raw, relayout_raw
    if torch.cuda.mem_get_info()[0] < 40 * 2 ** 30:
        pytest.skip("needs 40 GB of free device memory")
    B, D, h, w, L = 1, 256, 135, 240, 4
    g = torch.Generator().manual_seed(11)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    coords = t(make_coords("sigma4", B, h, w, np.random.default_rng(11)))
    rows = build_pyramid_raw(f1, f2, L, "bf16")
    tiled = build_pyramid_tiled_raw(f1, f2, L, "bf16")
    a, b = lookup_tiled_raw(tiled, coords, 4), lookup_raw(rows, coords, 4)
    assert torch.isfinite(a).all() and relmax(n(a), n(b)) <= 2e-6
    for l, (hl, wl) in enumerate(O.level_shapes(h, w, L)):
        back = relayout_raw(tiled[l], B, h * w, hl, wl, to_tiled=False)
        assert torch.equal(back, rows[l]), l
        del back
    # the last query pixels live beyond 2^31 bytes of level 0: spot-check them against a direct product
    p = h * w - 5
    want = (f1.view(D, -1)[:, p].double() @ f2.view(D, -1).double() / 16.0).float()
    got = rows[0][p].view(-1)
    assert float((got - want).abs().max()) <= 6e-3 * float(want.abs().max())


def test_tiled_build_rejects(rr):
    from raft_ros_b200 import _lib
    lib = _lib.lib()
    f = torch.zeros((1, 64, 32, 32), device=DEV)
    lv = [torch.zeros(1 << 22, device=DEV) for _ in range(5)]
    ws = torch.empty(1 << 22, dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.rc_corr_build_pyramid_tiled(f.data_ptr(), f.data_ptr(), 1, 64, 32, 32, _lib.ptr_array([x.data_ptr() for x in lv[:4]]),
                                         4, 0, 0.0, ws.data_ptr(), ws.numel(), st)
    assert rc == 3 and b"tensor-core" in lib.rc_last_error()
    rc = lib.rc_corr_build_pyramid_tiled(f.data_ptr(), f.data_ptr(), 1, 64, 32, 32, _lib.ptr_array([x.data_ptr() for x in lv]),
                                         5, 1, 0.0, ws.data_ptr(), ws.numel(), st)
    assert rc == 3 and b"4 levels" in lib.rc_last_error()
    rc = lib.rc_corr_build_pyramid_tiled(f.data_ptr(), f.data_ptr(), 1, 64, 32, 32,
                                         _lib.ptr_array([lv[0].data_ptr() + 16] + [x.data_ptr() for x in lv[1:4]]),
                                         4, 1, 0.0, ws.data_ptr(), ws.numel(), st)
    assert rc == 1 and b"aligned" in lib.rc_last_error()


def test_corrblock_layouts_agree(rr, monkeypatch):
    """CorrBlock's inference path (tiled) and RAFT_CORR_LAYOUT=rows give the same lookups and the same
    `corr_pyramid`; assigning `corr_pyramid` redirects the lookups, as with the reference's attribute."""
    from raft_ros_b200.corr import CorrBlock
    rng = np.random.default_rng(77)
    B, D, h, w = 2, 256, 27, 41
    f1 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, h, w)).astype(np.float32))
    coords = t(make_coords("sigma4", B, h, w, rng))
    blk = CorrBlock(f1, f2)
    assert blk._tiled is not None
    monkeypatch.setenv("RAFT_CORR_LAYOUT", "rows")
    ref = CorrBlock(f1, f2)
    assert ref._tiled is None
    a, b = blk(coords), ref(coords)
    assert relmax(n(a), n(b)) <= 2e-6
    for x, y in zip(blk.corr_pyramid, ref.corr_pyramid):
        assert x.shape == y.shape and torch.equal(x, y)
    blk.corr_pyramid = [torch.zeros_like(x) for x in blk.corr_pyramid]
    assert float(blk(coords).abs().max()) == 0.0


@pytest.mark.parametrize("radius", [4, 3])
@pytest.mark.parametrize("kind", ["grid", "sigma4", "sigma32", "oob", "border", "huge"])
@pytest.mark.parametrize("geom", [(2, 27, 40, 4), (1, 23, 39, 3), (1, 55, 128, 4), (3, 16, 16, 4)])
def test_lookup_on_tiled_layout(rr, geom, kind, radius):
    """rc_corr_lookup_tiled_fwd (2x8-tiled volume rows, cp.async-fed persistent kernel) gives the
    row-layout lookup's result on every geometry, including widths / heights that need padding."""
    from raft_ros_b200 import _lib
    from raft_ros_b200.corr import lookup_raw
    B, h, w, L = geom
    rng = np.random.default_rng(seed_of(geom, kind))
    levels = [t(rng.standard_normal((B * h * w, 1, hl, wl)).astype(np.float32)) for hl, wl in O.level_shapes(h, w, L)]
    coords = t(make_coords(kind, B, h, w, rng))
    want = lookup_raw(levels, coords, radius)
    tiles = _tile_levels(levels, B)
    out = torch.full((B, L * (2 * radius + 1) ** 2, h, w), float("nan"), device=DEV)
    lib = _lib.lib()
    rc = lib.rc_corr_lookup_tiled_fwd(_lib.ptr_array([x.data_ptr() for x in tiles]), coords.data_ptr(), out.data_ptr(),
                                      B, h, w, L, radius, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.rc_last_error()
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    assert relmax(n(out), n(want)) <= 2e-6, (geom, kind, radius, relmax(n(out), n(want)))
    # the fp16 flavour on the same (rounded) values
    from raft_ros_b200.corr import relayout_raw, lookup_tiled_raw
    shapes = O.level_shapes(h, w, L)
    t16 = [relayout_raw(lv, B, h * w, hl, wl, to_tiled=True, vol16=True) for lv, (hl, wl) in zip(levels, shapes)]
    got16 = lookup_tiled_raw(t16, coords, radius)
    want16 = lookup_raw([lv.half().float() for lv in levels], coords, radius)
    assert relmax(n(got16), n(want16)) <= 2e-6, (geom, kind, radius, "fp16 volume")
    rc = lib.rc_corr_lookup_tiled_fwd(_lib.ptr_array([x.data_ptr() for x in tiles]), coords.data_ptr(), out.data_ptr(),
                                      B, h, w, L, 5, torch.cuda.current_stream().cuda_stream)
    assert rc == 3 and b"radius" in lib.rc_last_error()


# ----------------------------------------------------------------------------- documented build-time / run-time knobs
def test_single_cta_build_variant_in_a_subprocess(rr):
    """RAFT_CORR_TC_CTAS=1 selects the single-CTA (cta_group::1) instantiations of the tensor-core build: other
    TMA boxes, barrier counts and commit paths than the default CTA pairs.  The variable is read once per process,
    so the golden, ragged tiled and fp16-volume build / lookup cases are re-run in a child process with it set."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RAFT_CORR_TC_CTAS="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-x", "-q", "-m", "gpu",
                        "-k", "golden_build or test_build_vs_oracle or tiled_build_equals_row_build or vol16_build_and_lookup",
                        "-p", "no:cacheprovider"], capture_output=True, text=True, env=env, cwd=root, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout


def test_pdl_off_gives_the_same_results(rr):
    """RAFT_CORR_PDL=0 launches every kernel without programmatic dependent launch; a child process with it set
    re-runs a build + lookup parity slice (the variable is read once per process)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RAFT_CORR_PDL="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-x", "-q", "-m", "gpu",
                        "-k", "corrblock_random_geometries or lookup_on_tiled_layout", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, env=env, cwd=root, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_corrblock_refuses_coords_that_want_a_gradient(rr):
    f = torch.randn(1, 32, 16, 16, device=DEV)
    c = rr.coords_grid(1, 16, 16, DEV).requires_grad_(True)
    for blk in (rr.CorrBlock(f, f), rr.CorrBlock(f.clone().requires_grad_(True), f)):
        with pytest.raises(RuntimeError, match="coords"):
            blk(c)
        with torch.no_grad():
            assert blk(c).shape == (1, 324, 16, 16)
        assert blk(c.detach()).shape == (1, 324, 16, 16)


def test_cta_pair_on_the_fly_variant_in_a_subprocess(rr):
    """RAFT_CORR_ALT_CTAS=2 selects the CTA-pair on-the-fly kernel (16x16 query tiles, cta_group::2) instead of the default
    single CTA per 8x16 tile; the variable is read once per process, so the oracle comparisons re-run in a child process
    with it set."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RAFT_CORR_ALT_CTAS="2")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-x", "-q", "-m", "gpu",
                        "-k", "alternate_block_vs_oracle or alternate_block_radius3", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, env=env, cwd=root, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_alt_pyramid_one_call_entry_point(rr):
    """rc_alt_corr_pyramid_fwd (maps described and kernel launched in one call) gives what the setup-once / run pair behind
    AlternateCorrBlock gives, through the C ABI with raw pointers."""
    from raft_ros_b200 import _lib
    L_ = _lib.lib()
    B, D, H, W, nl = 1, 128, 27, 40, 4
    rng = np.random.default_rng(27)
    f1 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    coords = t(make_coords("sigma4", B, H, W, rng))
    blk = rr.AlternateCorrBlock(f1, f2, num_levels=nl, radius=4, precision="bf16x3")
    want = blk(coords)
    out = torch.full_like(want, float("nan"))
    nbytes = int(L_.rc_alt_corr_pyramid_plan_bytes(B, H, W, nl))
    plan = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    rc = L_.rc_alt_corr_pyramid_fwd(blk._f1_tc[0].data_ptr(), blk._f1_tc[1].data_ptr(),
                                    _lib.ptr_array([x[0].data_ptr() for x in blk._f2_tc]),
                                    _lib.ptr_array([x[1].data_ptr() for x in blk._f2_tc]), coords.data_ptr(), out.data_ptr(),
                                    B, H, W, D, nl, 4, 1, 0.0, plan.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L_.rc_last_error()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
