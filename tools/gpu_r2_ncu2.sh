#!/bin/bash
# ncu captures of the kernels added late in round 2: the convolution (one pair and batch 8), the fp16_fp8 build, the fp16_fp8 on-the-fly kernel
set -u
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
bash tools/gpu_prof_any.sh conv_tc_kernel r02_conv_tc_b1 4 python tools/run_conv_one.py 1 256 0 192 3 3
bash tools/gpu_prof_any.sh conv_tc_kernel r02_conv_tc_b8 4 python tools/run_conv_one.py 8 256 0 192 3 3
bash tools/gpu_prof_any.sh conv_tc_kernel r02_conv_tc_gru_b8 4 python tools/run_conv_one.py 8 128 256 128 1 5 gru
cat > /tmp/run_alt.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
import raft_ros_b200 as rr
g = torch.Generator().manual_seed(5)
a1 = torch.randn(1, 256, 270, 480, generator=g).cuda(); a2 = torch.randn(1, 256, 270, 480, generator=g).cuda()
c4 = (rr.coords_grid(1, 270, 480, "cpu") + 2.0 * torch.randn(1, 2, 270, 480, generator=g)).cuda()
blk = rr.AlternateCorrBlock(a1, a2, num_levels=4, radius=4, precision="fp16_fp8")
for _ in range(4): o = blk(c4)
torch.cuda.synchronize(); print("ok", float(o.abs().mean()))
PY
bash tools/gpu_prof_any.sh alt_tc_kernel r02_alt_tc_fp16_fp8 2 python /tmp/run_alt.py
bash tools/gpu_prof_any.sh build_tc_kernel r02_build_tc_fp16_fp8 3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference --precision fp16_fp8
