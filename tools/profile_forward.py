"""Where the time of one FusedRAFT forward (440x1024, iters 20, batch 1) goes: device time per kernel name (torch.profiler)."""
import argparse, json, os, sys, warnings
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_ros_b200 as rr
from oracle import build_ref_core
from tools.raft_e2e import images

warnings.filterwarnings("ignore")
core = build_ref_core.load()
import core.raft  # noqa
torch.manual_seed(1234)
model = core.raft.RAFT(argparse.Namespace(small=False, mixed_precision=False, alternate_corr=False)).cuda().eval()
im1, im2 = images(440, 1024)
BATCH = int(sys.argv[1]) if len(sys.argv) > 1 else 1
ITERS = int(sys.argv[2]) if len(sys.argv) > 2 else 20
if BATCH > 1:
    im1, im2 = im1.repeat(BATCH, 1, 1, 1), im2.repeat(BATCH, 1, 1, 1)
fr = rr.FusedRAFT(model)
for _ in range(3):
    fr(im1, im2, iters=ITERS)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    fr(im1, im2, iters=ITERS)
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total if hasattr(e, "device_time_total") else e.cuda_time_total, e.count) for e in prof.key_averages()]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print("total device us", round(tot))
for k, us, n in rows[:40]:
    print(f"{us:9.0f} us  {n:4d} x  {us / n:7.1f}  {k[:110]}")
