"""The per-iteration update on this library's kernels (-m gpu): conv2d_fused's strided operands and GRU epilogues against
torch in fp64, `UpdateBlock` against the reference's own BasicUpdateBlock / SmallUpdateBlock, and `FusedRAFT` against the
reference's own RAFT.forward (unmodified files ferried as oracle/_ref/core; random-init weights, seed 1234).  Operands
are fp16 (TF32-class): element bounds 3e-3 of the layer's magnitude; flow drift <= 0.005 px mean EPE, as for the drop-in."""
import argparse

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def rr():
    import raft_ros_b200
    from raft_ros_b200 import _lib
    _lib.build()
    _lib.lib()
    return raft_ros_b200


@pytest.fixture(scope="module")
def ref_core(rr):
    from oracle import build_ref_core
    core = build_ref_core.load()
    if core is None:
        pytest.skip("oracle/_ref/core not present (built in the build container by __graft_entry__.build())")
    import core.raft  # noqa: F401
    return core


def _rand(g, *shape):
    return torch.randn(*shape, generator=g).to(DEV)


@pytest.mark.parametrize("geom", [(2, 128, 256, 24, 40, 1, 5), (1, 96, 146, 13, 22, 3, 3), (2, 128, 256, 17, 16, 5, 1),
                                  (7, 128, 256, 24, 40, 1, 5)])  # the last one: 168 tiles, more than one per CTA, with the GRU epilogues
def test_gru_half_in_two_launches(rr, geom):
    """z | r*h from one convolution (epilogue mul_from), then (1-z)h + z tanh(conv([r*h, x])) (epilogue gru), with the
    operands of the second launch being channel slices of the first one's output (core/update.py:16-31, 33-60)."""
    B, hd, xd, H, W, KH, KW = geom
    g = torch.Generator().manual_seed(hd + H)
    h, x = torch.tanh(_rand(g, B, hd, H, W)), _rand(g, B, xd, H, W)
    s = ((hd + xd) * KH * KW) ** -0.5
    wz, wr, wq = (s * _rand(g, hd, hd + xd, KH, KW) for _ in range(3))
    bz, br, bq = (0.1 * _rand(g, hd) for _ in range(3))
    pad = (KH // 2, KW // 2)
    hx = torch.cat([h, x], 1).double()
    z = torch.sigmoid(F.conv2d(hx, wz.double(), bz.double(), padding=pad))
    r = torch.sigmoid(F.conv2d(hx, wr.double(), br.double(), padding=pad))
    q = torch.tanh(F.conv2d(torch.cat([r * h.double(), x.double()], 1), wq.double(), bq.double(), padding=pad))
    want = ((1 - z) * h.double() + z * q).float()
    zr = rr.conv2d_fused((h, x), torch.cat([wz, wr]), torch.cat([bz, br]), "sigmoid", epilogue="mul_from", split=hd, aux1=h)
    assert float((zr[:, :hd] - z.float()).abs().max()) <= 3e-3
    assert float((zr[:, hd:] - (r * h.double()).float()).abs().max()) <= 3e-3
    got = rr.conv2d_fused((zr[:, hd:], x), wq, bq, "tanh", epilogue="gru", aux1=zr[:, :hd], aux2=h)
    assert got.shape == want.shape and torch.isfinite(got).all()
    assert float((got - want).abs().max()) <= 3e-3, float((got - want).abs().max())


def test_conv_writes_into_a_channel_slice(rr):
    g = torch.Generator().manual_seed(3)
    B, H, W = 2, 12, 24
    a, b = _rand(g, B, 64, H, W), _rand(g, B, 32, H, W)
    wgt, bias = 0.05 * _rand(g, 126, 96, 3, 3), 0.1 * _rand(g, 126)
    x = torch.full((B, 256, H, W), float("nan"), device=DEV)
    x[:, :128] = 7.0
    rr.conv2d_fused((a, b), wgt, bias, "relu", out=x[:, 128:254])  # 126 filters: the two that pad them to 128 store nothing
    want = F.relu(F.conv2d(torch.cat([a, b], 1).double(), wgt.double(), bias.double(), padding=1)).float()
    assert float((x[:, 128:254] - want).abs().max()) <= 3e-3 * float(want.abs().max())
    assert bool(torch.isnan(x[:, 254:]).all()) and bool((x[:, :128] == 7.0).all())  # nothing else touched
    rr.conv2d_fused((a, b), wgt, bias, "relu", out=x[:, 128:256])  # a slice of the padded count takes the zero filters' zeros
    assert bool((x[:, 254:] == 0).all())
    with pytest.raises(RuntimeError, match="out must be"):
        rr.conv2d_fused((a, b), wgt, bias, "relu", out=x[:, 128:250])
    with pytest.raises(RuntimeError, match="dense"):
        rr.conv2d_fused(a[:, :, :, ::2], wgt[:, :64], bias)


def test_small_conv_relu_and_flow_update(rr):
    g = torch.Generator().manual_seed(9)
    for (B, Cin, Cout, H, W, K) in [(2, 2, 128, 21, 37, 7), (1, 2, 64, 55, 128, 7), (1, 3, 20, 9, 11, 3)]:
        x, w, b = _rand(g, B, Cin, H, W), 0.1 * _rand(g, Cout, Cin, K, K), 0.1 * _rand(g, Cout)
        want = F.relu(F.conv2d(x.double(), w.double(), b.double(), padding=K // 2)).float()
        got = rr.small_conv_relu(x, w, b)
        assert got.shape == want.shape and float((got - want).abs().max()) <= 2e-5 * max(1.0, float(want.abs().max()))
    B, H, W = 2, 13, 22
    c0 = rr.coords_grid(B, H, W, DEV)
    c1 = c0 + _rand(g, B, 2, H, W)
    wide = _rand(g, B, 16, H, W)
    xbuf = torch.zeros(B, 40, H, W, device=DEV)
    flow = torch.empty(B, 2, H, W, device=DEV)
    want_c1 = c1 + wide[:, :2]
    rr.flow_update(c1, wide[:, :2], c0, flow, xbuf[:, 30:32])
    assert torch.equal(c1, want_c1) and torch.equal(flow, want_c1 - c0) and torch.equal(xbuf[:, 30:32], flow)
    assert bool((xbuf[:, :30] == 0).all()) and bool((xbuf[:, 32:] == 0).all())


def _update_inputs(core, small, B, H, W, g):
    hd, cd = (96, 64) if small else (128, 128)
    r = 3 if small else 4
    net = torch.tanh(_rand(g, B, hd, H, W))
    inp = torch.relu(_rand(g, B, cd, H, W))
    corr = _rand(g, B, 4 * (2 * r + 1) ** 2, H, W)
    flow = 3.0 * _rand(g, B, 2, H, W)
    return net, inp, corr, flow


@pytest.mark.parametrize("small", [False, True], ids=["basic", "small"])
def test_update_block_matches_the_reference_module(rr, ref_core, small):
    import core.update as ref_update
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(1234)
    args = argparse.Namespace(corr_levels=4, corr_radius=3 if small else 4)
    ub = (ref_update.SmallUpdateBlock(args, hidden_dim=96) if small else ref_update.BasicUpdateBlock(args, hidden_dim=128)).to(DEV).eval()
    g = torch.Generator().manual_seed(11)
    net, inp, corr, flow = _update_inputs(ref_core, small, 2, 20, 36, g)
    with torch.no_grad():
        n0, m0, d0 = ub(net, inp, corr, flow)
    n1, m1, d1 = rr.UpdateBlock(ub)(net, inp, corr, flow)
    assert float((n1 - n0).abs().max()) <= 3e-3, float((n1 - n0).abs().max())
    assert float((d1 - d0).abs().max()) <= 3e-3 * max(1.0, float(d0.abs().max())), float((d1 - d0).abs().max())
    if small:
        assert m0 is None and m1 is None
    else:
        assert float((m1 - m0).abs().max()) <= 3e-3 * max(1.0, float(m0.abs().max())), float((m1 - m0).abs().max())
    # a weight updated in place (load_state_dict, an optimiser step) is seen: the stacked gate filters are rebuilt
    with torch.no_grad():
        (ub.gru.convz if small else ub.gru.convz1).weight.mul_(0.5)
        n0b, _, _ = ub(net, inp, corr, flow)
    blk = rr.UpdateBlock(ub)
    n1b, _, _ = blk(net, inp, corr, flow)
    assert float((n1b - n0b).abs().max()) <= 3e-3
    with torch.no_grad():
        (ub.gru.convz if small else ub.gru.convz1).weight.mul_(2.0)
    n1c, _, _ = blk(net, inp, corr, flow)
    assert float((n1c - n0).abs().max()) <= 3e-3


def _images(H, W):
    g = torch.Generator().manual_seed(77)
    low = torch.rand(1, 3, H // 8 + 2, W // 8 + 2, generator=g)
    big = F.interpolate(low, size=(H + 16, W + 16), mode="bicubic", align_corners=False).clamp(0, 1)
    tex = (big + 0.15 * torch.rand(1, 3, H + 16, W + 16, generator=g)).clamp(0, 1) * 255.0
    return tex[:, :, 8:8 + H, 8:8 + W].contiguous().to(DEV), tex[:, :, 5:5 + H, 11:11 + W].contiguous().to(DEV)


def test_fused_raft_on_a_kitti_frame(rr, ref_core):
    """376 x 1248 (the ROS launch file's KITTI input, raft-small and raft-things): a 47 x 156 feature grid, i.e. more than
    one convolution tile per row and ragged everything."""
    core = ref_core
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    im1, im2 = _images(376, 1248)
    for small in (True, False):
        torch.manual_seed(1234)
        model = core.raft.RAFT(argparse.Namespace(small=small, mixed_precision=False, alternate_corr=False)).to(DEV).eval()
        with torch.no_grad():
            flow_ref, up_ref = model(im1, im2, iters=8, test_mode=True)
        flow, up = rr.FusedRAFT(model, graph=True)(im1, im2, iters=8)
        epe = float((flow - flow_ref).norm(dim=1).mean())
        print(f"FusedRAFT 376x1248 small={small}: EPE {epe:.2e} px")
        assert epe <= 0.005 and float((up - up_ref).norm(dim=1).mean()) <= 0.04, (small, epe)


def test_fused_raft_with_a_mixed_precision_model(rr, ref_core):
    """`--mixed_precision` (core/raft.py:98-115,127: encoders and update block under fp16 autocast): FusedRAFT keeps the
    encoders under autocast and runs the update block on its fp16-operand kernels.  Both sides are fp16-class here, so the
    bound is the reference's own distance between its fp16 and fp32 runs (measured and printed), with 0.05 px as a cap."""
    core = ref_core
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(1234)
    im1, im2 = _images(184, 320)
    model = core.raft.RAFT(argparse.Namespace(small=False, mixed_precision=False, alternate_corr=False)).to(DEV).eval()
    with torch.no_grad():
        f32, _ = model(im1, im2, iters=12, test_mode=True)
        model.args.mixed_precision = True
        f16, _ = model(im1, im2, iters=12, test_mode=True)
    ours, up = rr.FusedRAFT(model)(im1, im2, iters=12)
    d_ref = float((f16 - f32).norm(dim=1).mean())
    d_ours = float((ours - f32).norm(dim=1).mean())
    print(f"mixed precision: reference fp16 vs fp32 {d_ref:.2e} px, FusedRAFT vs fp32 {d_ours:.2e} px")
    assert torch.isfinite(up).all() and d_ours <= max(2.0 * d_ref, 0.005) and d_ours <= 0.05, (d_ref, d_ours)


@pytest.mark.parametrize("small,alternate", [(False, False), (True, False), (False, True)], ids=["things", "small", "things-alternate"])
def test_fused_raft_matches_the_reference_forward(rr, ref_core, monkeypatch, small, alternate):
    core = ref_core
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(1234)
    H, W = (440, 1024) if not alternate else (184, 320)
    model = core.raft.RAFT(argparse.Namespace(small=small, mixed_precision=False, alternate_corr=False)).to(DEV).eval()
    im1, im2 = _images(H, W)
    with torch.no_grad():
        flow_ref, up_ref = model(im1, im2, iters=12, test_mode=True)
    model.args.alternate_corr = alternate  # the reference's alternate path needs its compiled extension; ours does not
    fused = rr.FusedRAFT(model)
    flow, up = fused(im1, im2, iters=12)
    assert flow.shape == flow_ref.shape and up.shape == up_ref.shape
    epe = float((flow - flow_ref).norm(dim=1).mean())
    epe_up = float((up - up_ref).norm(dim=1).mean())
    print(f"FusedRAFT small={small} alternate={alternate}: EPE {epe:.2e} px (1/8 res), {epe_up:.2e} px (full res), mean |flow| {float(flow_ref.norm(dim=1).mean()):.2f}")
    assert epe <= 0.005 and epe_up <= 0.04, (epe, epe_up)
    # warm start (core/raft.py:118-119) and the captured graph give the same numbers as the eager call
    init = 0.5 * torch.ones_like(flow_ref)
    with torch.no_grad():
        ref_i, _ = model.__class__.forward(model, im1, im2, iters=4, flow_init=init, test_mode=True) if not alternate else (None, None)
    got_i, _ = fused(im1, im2, iters=4, flow_init=init)
    if ref_i is not None:
        assert float((got_i - ref_i).norm(dim=1).mean()) <= 0.005
    graphed = rr.FusedRAFT(model, graph=True)
    for _ in range(2):
        gflow, gup = graphed(im1, im2, iters=12)
    assert float((gflow - flow).abs().max()) <= 1e-4 and float((gup - up).abs().max()) <= 1e-3
    with pytest.raises(RuntimeError, match="test-mode"):
        fused(im1, im2, iters=2, test_mode=False)
    if not small and not alternate:
        # weights changed in place after the capture (load_state_dict): the graph is rebuilt, not replayed stale
        with torch.no_grad():
            model.update_block.flow_head.conv2.weight.mul_(0.5)
            ref2, _ = model(im1, im2, iters=12, test_mode=True)
        g2, _ = graphed(im1, im2, iters=12)
        assert float((g2 - ref2).norm(dim=1).mean()) <= 0.005 and float((g2 - gflow).abs().max()) > 1e-3
