"""The reference's whole `RAFT.forward` (core/raft.py:86-144) on one B200 with the correlation path swapped in, step
by step -- what a user of the ROS node gets (ros/scripts/main.py:117-149: one pair, iters = 20, test_mode):

  reference   the reference as shipped (its torch-op CorrBlock), fp32, TF32 as torch defaults it
  drop-in     only core.raft.CorrBlock rebound to raft_ros_b200.CorrBlock
  fused       + encoder_head for the encoder's last layer, CorrBlock.lookup_convc1 for lookup -> convc1 -> ReLU,
              upsample_flow for the convex upsampling; everything else is still the reference's own modules
  graph       the fused forward captured once in a CUDA graph and replayed
  FusedRAFT   the package's host object (raft_ros_b200/update.py): + conv2d_fused (tcgen05 implicit GEMM; cat, bias,
              activation and the GRU's element-wise steps fused) for the update block's convolutions, a GRU half in two
              launches, mask head + upsampling once after the loop; eager and as a CUDA graph

Model, encoders and update block come from the reference's unmodified files (oracle/_ref/core, ferried by
oracle/build_ref_core.py: measurement infrastructure, not product).  The fused variants re-type the loop of
core/raft.py:122-139 around those modules, so their flow is checked against the reference's (mean EPE printed).
Random-init raft-things weights (seed 1234), synthetic smooth frames, 440x1024.
"""
import argparse
import json
import os
import sys
import warnings

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raft_ros_b200 as rr  # noqa: E402
from oracle import build_ref_core  # noqa: E402

DEV = "cuda:0"


def images(H, W):
    g = torch.Generator().manual_seed(77)
    low = torch.rand(1, 3, H // 8 + 2, W // 8 + 2, generator=g)
    big = F.interpolate(low, size=(H + 16, W + 16), mode="bicubic", align_corners=False).clamp(0, 1)
    tex = (big + 0.15 * torch.rand(1, 3, H + 16, W + 16, generator=g)).clamp(0, 1) * 255.0
    return tex[:, :, 8:8 + H, 8:8 + W].contiguous().to(DEV), tex[:, :, 5:5 + H, 11:11 + W].contiguous().to(DEV)


def fused_forward(model, image1, image2, iters, precision="bf16x3"):
    """core/raft.py:86-144 in test_mode with the three fusions; every module called is the reference's own."""
    image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
    image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
    hdim, cdim = model.hidden_dim, model.context_dim
    enc = model.fnet
    x = torch.cat([image1, image2], dim=0)                      # core/extractor.py:168-192 up to conv2's input
    x = enc.layer3(enc.layer2(enc.layer1(enc.relu1(enc.norm1(enc.conv1(x))))))
    prep = rr.encoder_head(x, enc.conv2.weight, enc.conv2.bias)  # conv2 + .float() + layout prepass
    corr_fn = rr.CorrBlock.from_prepared(prep, radius=model.args.corr_radius, precision=precision)
    cnet = model.cnet(image1)
    net, inp = torch.split(cnet, [hdim, cdim], dim=1)
    net, inp = torch.tanh(net), torch.relu(inp)
    coords0, coords1 = model.initialize_flow(image1)
    ub = model.update_block
    me = ub.encoder
    flow_up = None
    for _ in range(iters):
        coords1 = coords1.detach()
        flow = coords1 - coords0
        cor = corr_fn.lookup_convc1(coords1, me.convc1.weight, me.convc1.bias)  # lookup -> convc1 -> ReLU (core/update.py:90)
        cor = F.relu(me.convc2(cor))                                            # core/update.py:91-97, unchanged
        flo = F.relu(me.convf2(F.relu(me.convf1(flow))))
        out = F.relu(me.conv(torch.cat([cor, flo], dim=1)))
        motion = torch.cat([out, flow], dim=1)
        net = ub.gru(net, torch.cat([inp, motion], dim=1))                      # core/update.py:127-136
        delta_flow = ub.flow_head(net)
        up_mask = 0.25 * ub.mask(net)
        coords1 = coords1 + delta_flow
        flow_up = rr.upsample_flow(coords1 - coords0, up_mask)                  # core/raft.py:72-83
    return coords1 - coords0, flow_up


def timed(fn, warm=2, reps=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--height", type=int, default=440)
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=1, help="frame pairs per forward (the pair is repeated with small noise)")
    ap.add_argument("--small", action="store_true", help="raft-small (the ROS launch file's default model)")
    a = ap.parse_args()
    core = build_ref_core.load()
    if core is None:
        print(json.dumps({"unavailable": "oracle/_ref/core not present"}))
        return
    import core.raft  # noqa: F401
    warnings.filterwarnings("ignore")
    torch.manual_seed(1234)
    model = core.raft.RAFT(argparse.Namespace(small=a.small, mixed_precision=False, alternate_corr=False)).to(DEV).eval()
    im1, im2 = images(a.height, a.width)
    if a.batch > 1:
        g = torch.Generator(device=DEV).manual_seed(3)
        im1 = (im1.repeat(a.batch, 1, 1, 1) + 2.0 * torch.rand(a.batch, *im1.shape[1:], device=DEV, generator=g)).clamp(0, 255)
        im2 = (im2.repeat(a.batch, 1, 1, 1) + 2.0 * torch.rand(a.batch, *im2.shape[1:], device=DEV, generator=g)).clamp(0, 255)
    res = {"case": f"RAFT.forward {'raft-small' if a.small else 'raft-things'} {a.height}x{a.width}, batch {a.batch}, iters {a.iters}, test_mode", "tf32": "torch defaults"}
    with torch.no_grad():
        ms_ref, (flow_ref, up_ref) = timed(lambda: model(im1, im2, iters=a.iters, test_mode=True))
        res["reference_ms"] = ms_ref
        saved = core.raft.CorrBlock
        core.raft.CorrBlock = rr.CorrBlock
        try:
            ms_drop, (flow_drop, _) = timed(lambda: model(im1, im2, iters=a.iters, test_mode=True))
        finally:
            core.raft.CorrBlock = saved
        res["dropin_ms"] = ms_drop
        res["dropin_epe_px"] = float((flow_drop - flow_ref).norm(dim=1).mean())
        ms_fused, (flow_fused, _) = timed(lambda: fused_forward(model, im1, im2, a.iters)) if not a.small else (float("nan"), (flow_ref, None))
        res["fused_ms"] = ms_fused
        res["fused_epe_px"] = float((flow_fused - flow_ref).norm(dim=1).mean())
        # CUDA graph of the fused forward
        s1, s2 = im1.clone(), im2.clone()
        if not a.small:  # (the round-2 mid-way variant: raft-things only)
            graph = torch.cuda.CUDAGraph()
            fused_forward(model, s1, s2, a.iters)
            torch.cuda.synchronize()
            with torch.cuda.graph(graph):
                g_flow, g_up = fused_forward(model, s1, s2, a.iters)
            ms_graph, _ = timed(graph.replay)
            res["fused_graph_ms"] = ms_graph
            res["fused_graph_epe_px"] = float((g_flow - flow_ref).norm(dim=1).mean())
        fr = rr.FusedRAFT(model)
        ms_c, (flow_c, up_c) = timed(lambda: fr(im1, im2, iters=a.iters))
        res["fusedraft_ms"] = ms_c
        res["fusedraft_epe_px"] = float((flow_c - flow_ref).norm(dim=1).mean())
        res["fusedraft_epe_fullres_px"] = float((up_c - up_ref).norm(dim=1).mean())
        frg = rr.FusedRAFT(model, graph=True)
        ms_cg, (flow_g, _) = timed(lambda: frg(im1, im2, iters=a.iters))
        res["fusedraft_graph_ms"] = ms_cg
        res["fusedraft_graph_epe_px"] = float((flow_g - flow_ref).norm(dim=1).mean())
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        flow32, _ = model(im1, im2, iters=a.iters, test_mode=True)  # the reference in true fp32: how far TF32 itself moves the flow
        res["reference_tf32_vs_fp32_epe_px"] = float((flow_ref - flow32).norm(dim=1).mean())
        res["fusedraft_vs_fp32_epe_px"] = float((flow_c - flow32).norm(dim=1).mean())
        res["mean_flow_px"] = float(flow_ref.norm(dim=1).mean())
        if a.batch == 1:
            # the ROS node's callback (ros/scripts/main.py:117-149) by wall clock: two uint8 HxWx3 host images in, the
            # upsampled flow as a host HxWx2 array out
            import time
            import numpy as np
            h1 = np.ascontiguousarray(im1[0].permute(1, 2, 0).byte().cpu().numpy())
            h2 = np.ascontiguousarray(im2[0].permute(1, 2, 0).byte().cpu().numpy())

            def callback(fwd):
                i1 = torch.from_numpy(h1).permute(2, 0, 1).float()[None].to(DEV)
                i2 = torch.from_numpy(h2).permute(2, 0, 1).float()[None].to(DEV)
                _, up = fwd(i1, i2)
                return up[0].permute(1, 2, 0).cpu().numpy()

            def wall(fwd, reps=20):
                for _ in range(3):
                    callback(fwd)
                t0 = time.perf_counter()
                for _ in range(reps):
                    callback(fwd)
                return (time.perf_counter() - t0) / reps * 1e3

            torch.backends.cudnn.allow_tf32 = True
            torch.backends.cuda.matmul.allow_tf32 = True
            res["ros_callback_reference_ms"] = wall(lambda i1, i2: model(i1, i2, iters=a.iters, test_mode=True))
            res["ros_callback_fusedraft_graph_ms"] = wall(lambda i1, i2: frg(i1, i2, iters=a.iters))
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
