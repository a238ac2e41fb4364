#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 gpurun_out/pytest_noalloc_alone.log
cat > /tmp/fused_once.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
import raft_ros_b200 as rr
prec = sys.argv[1]
g = torch.Generator().manual_seed(1)
B, D, h, w = 8, 256, 55, 128
f1 = torch.randn(B, D, h, w, generator=g).cuda(); f2 = torch.randn(B, D, h, w, generator=g).cuda()
W = (torch.randn(256, 324, generator=g) / 18).cuda(); b = torch.zeros(256).cuda()
coords = (rr.coords_grid(B, h, w, "cpu") + 1.5 * torch.randn(B, 2, h, w, generator=g)).cuda()
blk = rr.CorrBlock(f1, f2, precision=prec)
for _ in range(5): o = blk.lookup_convc1(coords, W, b)
torch.cuda.synchronize(); print("ok", float(o.mean()))
PY
for prec in bf16x3 bf16_vol16; do
python /tmp/fused_once.py $prec && timeout 600 ncu --set full --clock-control none --import-source on -k regex:lookup_convc1_kernel -s 2 -c 1 -o gpurun_out/r02_lookup_convc1_$prec -f python /tmp/fused_once.py $prec > gpurun_out/ncu_fused_$prec.log 2>&1; echo "ncu fused $prec rc=$?"
done
python bench.py --steps 30 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference --l2-fetch 32 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('l2fetch32', round(d['value'],1), {k:round(v['ms'],4) for k,v in d['kernels'].items()})"
