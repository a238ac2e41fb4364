"""End-to-end flow drift of swapping the correlation operator inside RAFT (BASELINE.json:
"with end-to-end flow EPE drift reported").

The encoders and the update block are the REFERENCE's own networks, frozen to TorchScript by
oracle/build_ref_model.py (raft-things architecture, random init seed 1234: checkpoints are
offline).  The refinement loop below only re-states the data flow around the corr call sites
(core/raft.py:99-131): features -> corr_fn -> 12 x [lookup, update, coords += delta].  The
baseline arm uses the reference's torch ops for the correlation (oracle/torch_port.py on the
GPU, TF32 off); the other arms use this repository's CorrBlock / AlternateCorrBlock.  EPE drift
= mean / max end-point distance between the two final flow fields, in full-resolution pixels.
"""
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import raft_ros_b200 as rr  # noqa: E402
from oracle import torch_port as TP  # noqa: E402

DEV = "cuda:0"
REF = os.path.join(ROOT, "oracle", "_ref")


def synthetic_pair(H, W, seed=1234):
    """Smooth random texture and a copy warped by a smooth flow (a few pixels)."""
    g = torch.Generator().manual_seed(seed)
    tex = sum(F.interpolate(torch.rand(1, 3, H // s, W // s, generator=g), size=(H, W), mode="bicubic", align_corners=False) / k
              for s, k in ((32, 1.0), (8, 2.0), (2, 4.0)))
    img1 = (tex - tex.min()) / (tex.max() - tex.min()) * 255.0
    flow = F.interpolate(torch.randn(1, 2, H // 64, W // 64, generator=g) * 6.0, size=(H, W), mode="bicubic", align_corners=False)
    ys, xs = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    gx = 2 * (xs + flow[0, 0]) / (W - 1) - 1
    gy = 2 * (ys + flow[0, 1]) / (H - 1) - 1
    img2 = F.grid_sample(img1, torch.stack([gx, gy], -1)[None], align_corners=True, padding_mode="border")
    return img1.clamp(0, 255), img2.clamp(0, 255)


def run(fnet, cnet, upd, img1, img2, make_corr, iters=12):
    with torch.no_grad():
        a = 2 * (img1 / 255.0) - 1.0
        b = 2 * (img2 / 255.0) - 1.0
        f1, f2 = fnet(a).float(), fnet(b).float()
        corr_fn = make_corr(f1, f2)
        c = cnet(a)
        net, inp = torch.tanh(c[:, :128]), torch.relu(c[:, 128:])
        B, _, h, w = f1.shape
        coords0 = rr.coords_grid(B, h, w, f1.device)
        coords1 = coords0.clone()
        for _ in range(iters):
            corr = corr_fn(coords1)
            net, _, delta = upd(net, inp, corr, coords1 - coords0)
            coords1 = coords1 + delta
        return (coords1 - coords0) * 8.0


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    fnet = torch.jit.load(os.path.join(REF, "raft_fnet.pt"), map_location=DEV).eval()
    cnet = torch.jit.load(os.path.join(REF, "raft_cnet.pt"), map_location=DEV).eval()
    upd = torch.jit.load(os.path.join(REF, "raft_update.pt"), map_location=DEV).eval()
    img1, img2 = [x.to(DEV) for x in synthetic_pair(440, 1024)]

    def ref_corr(f1, f2):
        pyr = TP.build_pyramid(f1, f2, 4)
        return lambda coords: TP.lookup(pyr, coords, 4)

    base = run(fnet, cnet, upd, img1, img2, ref_corr)
    again = run(fnet, cnet, upd, img1, img2, ref_corr)
    out = {"what": "EPE drift vs the reference's torch CorrBlock inside RAFT (raft-things arch, random init, 12 iters, 440x1024)",
           "flow_magnitude_px_mean": float(base.norm(dim=1).mean()), "flow_magnitude_px_max": float(base.norm(dim=1).max()),
           "reference_rerun_epe_mean": float((again - base).norm(dim=1).mean())}
    arms = {
        "CorrBlock_fp32": lambda f1, f2: rr.CorrBlock(f1, f2, radius=4, precision="fp32"),
        "CorrBlock_bf16x3": lambda f1, f2: rr.CorrBlock(f1, f2, radius=4, precision="bf16x3"),
        "CorrBlock_bf16": lambda f1, f2: rr.CorrBlock(f1, f2, radius=4, precision="bf16"),
        "CorrBlock_bf16_vol16": lambda f1, f2: rr.CorrBlock(f1, f2, radius=4, precision="bf16_vol16"),
        "AlternateCorrBlock_fp32": lambda f1, f2: rr.AlternateCorrBlock(f1, f2, radius=4, precision="fp32"),
        "AlternateCorrBlock_bf16x3": lambda f1, f2: rr.AlternateCorrBlock(f1, f2, radius=4, precision="bf16x3"),
        "AlternateCorrBlock_bf16": lambda f1, f2: rr.AlternateCorrBlock(f1, f2, radius=4, precision="bf16"),
    }
    for name, mk in arms.items():
        flow = run(fnet, cnet, upd, img1, img2, mk)
        d = (flow - base).norm(dim=1)
        out[name] = {"epe_mean_px": float(d.mean()), "epe_max_px": float(d.max())}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
