#!/usr/bin/env python
"""Benchmark of the correlation hot path: frame-pairs/sec for corr build + K lookups.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]      # the reference's CPU path

Workload (BASELINE.json metric, config[1] geometry): Sintel 436x1024 padded to 440x1024
(feature grid 55x128, D=256), batch 8 frame pairs per GPU, CorrBlock with 4 levels and r=4,
one build + 12 lookups per step.  Synthetic data (seeded randn feature maps; lookup
coordinates = pixel grid + a random-walk flow field), since datasets/checkpoints are offline.

One JSON line on stdout (rank 0); everything else goes to stderr.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frame-pairs/sec, corr build+12 lookups @440x1024"
UNIT = "frame-pairs/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="frame pairs per GPU per step")
    ap.add_argument("--height", type=int, default=440)
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--iters", type=int, default=12, help="lookups per build")
    ap.add_argument("--levels", type=int, default=4)
    ap.add_argument("--radius", type=int, default=4)
    ap.add_argument("--precision", default=os.environ.get("RAFT_CORR_PRECISION", "fp16_fp8"),
                    choices=["fp32", "bf16x3", "bf16", "bf16_vol16", "fp16_fp8"])
    ap.add_argument("--l2-fetch", type=int, default=int(os.environ.get("RAFT_CORR_L2_FETCH", "0")),
                    help="set cudaLimitMaxL2FetchGranularity (32/64/128); 0 = leave the driver default")
    ap.add_argument("--layout", default=os.environ.get("RAFT_CORR_LAYOUT", "tiled"), choices=["tiled", "rows"],
                    help="HBM layout of the pyramid: 'tiled' is what CorrBlock keeps for inference, "
                         "'rows' the reference's (B*h*w,1,h_l,w_l) arrays")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fast-paths", action="store_true",
                    help="skip the other precisions' measurements appended to the default line")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true",
                    help="skip the comparator leg: the reference's own torch ops / alt_cuda_corr binary on this GPU")
    ap.add_argument("--cpu-budget-s", type=float, default=150.0)
    return ap.parse_args()


def algorithmic_bytes(B, N, D, h, w, L, r, K, s_vol=4):
    """read both fmaps once, write every pyramid element once, and per
    lookup read the (2r+2)^2 integer patch per level, write L*(2r+1)^2 fp32, read 2 coords.
    s_vol = bytes per volume element (4, or 2 for precision bf16_vol16)."""
    pyr = sum(N * (h >> l) * (w >> l) for l in range(L)) * s_vol
    build = B * (2 * N * D * 4 + pyr)
    lookup = B * N * (L * (2 * r + 2) ** 2 * s_vol + L * (2 * r + 1) ** 2 * 4 + 8)
    return build, lookup, build + K * lookup


def measured_traffic(kernel, args, layout="rows"):
    """Per-launch memory traffic of `kernel` from this round's committed `ncu --set full` capture
    (profiles/r02_traffic.json, written by tools/ncu_traffic.py from the .ncu-rep; it records the commit
    it was captured at).  Only valid for the configuration it was captured on, else None.
    Returns (dram_read + dram_write, detail dict)."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    default = (args.batch, args.height, args.width, args.dim, args.levels, args.radius)
    if default != (8, 440, 1024, 256, 4, 4) or args.precision not in ("bf16x3", "fp16_fp8") or layout != "tiled" or not os.path.exists(p):
        return None, None
    try:
        d = json.load(open(p))
        k = d["kernels"][kernel]
        if kernel == "build" and args.precision == "fp16_fp8":  # (the lookup is the same kernel for every build precision)
            k = d["kernels"]["build_fp16_fp8"]
        detail = {"dram_read": k["dram_read"], "dram_write": k["dram_write"], "sm_write": k["sm_write"],
                  "sm_read": k.get("sm_read"), "kernel_us_under_ncu": k.get("us"), "commit": d.get("commit"),
                  "note": "dram_write counts only what reached DRAM before the kernel ended; sm_write is what the SMs "
                          "stored (l1tex__m_l1tex2xbar_write_bytes): in the bench loop the rest drains during the next launch"}
        return float(k["dram_read"]) + float(k["dram_write"]), detail
    except Exception:
        return None, None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    """Samples SM clock + throttle reasons DURING the timed region (NVML, 10 ms period)."""

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            log(f"[bench] NVML unavailable: {e}")
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.01)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ----------------------------------------------------------------------------- CPU (reference) arm
def cpu_reference_run(args, steps, warmup, budget_s):
    """The reference's CPU CorrBlock path on the box's host cores.  When the reference's own `core/`
    package was ferried to this machine (oracle/_ref/core/, oracle/build_ref_core.py) the timed code IS the
    reference's unmodified `core/corr.py::CorrBlock` (kind "reference"); otherwise the op-for-op port
    oracle/torch_port.py (kind "port").  One step = ONE frame pair of the workload: build + K lookups."""
    import torch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    h, w = args.height // 8, args.width // 8
    g = torch.Generator().manual_seed(1234)
    f1 = torch.randn(1, args.dim, h, w, generator=g)
    f2 = torch.randn(1, args.dim, h, w, generator=g)
    ys, xs = torch.meshgrid(torch.arange(h).float(), torch.arange(w).float(), indexing="ij")
    base = torch.stack([xs, ys], 0)[None]
    flows = torch.cumsum(0.5 * torch.randn(args.iters, 1, 2, h, w, generator=g), 0)

    ref_core = None
    try:
        from oracle import build_ref_core
        ref_core = build_ref_core.load()
    except Exception as e:
        log(f"[bench] reference core/ package unavailable ({e}); timing the port")
    if ref_core is not None:
        kind = "reference"
        RefCorrBlock = ref_core.corr.CorrBlock

        def step():
            blk = RefCorrBlock(f1, f2, num_levels=args.levels, radius=args.radius)
            s = 0.0
            for k in range(args.iters):
                s += float(blk(base + flows[k])[0, 0, 0, 0])
            return s
    else:
        kind = "port"
        from oracle import torch_port as TP

        def step():
            pyr = TP.build_pyramid(f1, f2, args.levels)
            s = 0.0
            for k in range(args.iters):
                s += float(TP.lookup(pyr, base + flows[k], args.radius)[0, 0, 0, 0])
            return s

    t_start = time.perf_counter()
    with torch.no_grad():
        for _ in range(max(1, warmup)):
            step()
            if time.perf_counter() - t_start > budget_s * 0.4:
                break
        times = []
        for _ in range(steps):
            t0 = time.perf_counter()
            step()
            times.append(time.perf_counter() - t0)
            if time.perf_counter() - t_start > budget_s:
                break
    mean = sum(times) / len(times)
    return {"pairs_per_s": 1.0 / mean, "ms_per_step": mean * 1e3, "steps_timed": len(times), "cores": cores,
            "threads": torch.get_num_threads(), "kind": kind,
            "sample": f"{'core/corr.py CorrBlock (unmodified, oracle/_ref/core)' if kind == 'reference' else 'oracle/torch_port.py'}: "
                      f"1 frame pair per step (of the {args.batch}-pair batch): build + {args.iters} lookups, fp32, "
                      f"{len(times)} timed steps, mean"}


def raft_forward_run(args):
    """Its own key, never the headline: the reference's whole test-mode RAFT.forward (core/raft.py:86-144; its own files
    from oracle/_ref/core, random-init raft-things weights, one 440x1024 pair, args.iters iterations, TF32 as torch
    defaults it) next to raft_ros_b200.FusedRAFT on the same model -- the call the ROS node makes per image pair
    (ros/scripts/main.py:117-149).  None when the reference's files are not on the box."""
    import argparse as ap
    import warnings

    import torch
    import torch.nn.functional as F

    import raft_ros_b200 as rr
    from oracle import build_ref_core

    core = build_ref_core.load()
    if core is None:
        return None
    import core.raft  # noqa: F401
    warnings.filterwarnings("ignore")
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1234)
    model = core.raft.RAFT(ap.Namespace(small=False, mixed_precision=False, alternate_corr=False)).to(dev).eval()
    H, W = args.height, args.width
    g = torch.Generator().manual_seed(77)
    low = torch.rand(1, 3, H // 8 + 2, W // 8 + 2, generator=g)
    big = F.interpolate(low, size=(H + 16, W + 16), mode="bicubic", align_corners=False).clamp(0, 1)
    tex = (big + 0.15 * torch.rand(1, 3, H + 16, W + 16, generator=g)).clamp(0, 1) * 255.0
    im1, im2 = tex[:, :, 8:8 + H, 8:8 + W].contiguous().to(dev), tex[:, :, 5:5 + H, 11:11 + W].contiguous().to(dev)

    def timeit(fn, warm=3, reps=10):
        for _ in range(warm):
            out = fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps, out

    K = args.iters
    with torch.no_grad():
        ms_ref, (flow_ref, _) = timeit(lambda: model(im1, im2, iters=K, test_mode=True))
        ms_eager, _ = timeit(lambda fr=rr.FusedRAFT(model): fr(im1, im2, iters=K))
        ms_graph, (flow_g, _) = timeit(lambda fr=rr.FusedRAFT(model, graph=True): fr(im1, im2, iters=K))
    return {"what": f"RAFT.forward, raft-things architecture, random init, one {H}x{W} pair, {K} iterations, test_mode",
            "reference_ms": ms_ref, "reference": "the reference's own files on the same GPU, TF32 as torch defaults it",
            "fused_raft_ms": ms_eager, "fused_raft_graph_ms": ms_graph, "speedup_graph": ms_ref / ms_graph,
            "epe_px_vs_reference": float((flow_g - flow_ref).norm(dim=1).mean()), "mean_flow_px": float(flow_ref.norm(dim=1).mean())}


def fused_convc1_run(args, f1, f2, coords):
    """Its own key, never the headline: the step with every lookup fused into the motion encoder's convc1 + ReLU
    (CorrBlock.lookup_convc1 = rc_corr_lookup_convc1_fwd, fp16 operands) against the unfused pair it replaces on
    the same GPU (this repository's lookup, then cuDNN's 1x1 convolution + ReLU under fp16 autocast, as the
    reference's --mixed_precision update block runs it)."""
    import torch
    import torch.nn.functional as F

    import raft_ros_b200 as rr

    K, L, r = args.iters, args.levels, args.radius
    B, _, h, w = f1.shape
    cout = 256 if r == 4 else 96
    g = torch.Generator().manual_seed(7)
    cin = L * (2 * r + 1) ** 2
    W = (torch.randn(cout, cin, 1, 1, generator=g) / cin ** 0.5).to(f1.device)
    bias = (0.1 * torch.randn(cout, generator=g)).to(f1.device)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def timeit(fn, warm=3, reps=20):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out = {"what": f"1 build + {K} x (lookup -> convc1 {cin}->{cout} 1x1 + ReLU), batch {B}", "operands": "fp16"}
    N = h * w
    for prec in (args.precision, "bf16_vol16"):
        def fused_step():
            blk = rr.CorrBlock(f1, f2, num_levels=L, radius=r, precision=prec)
            for k in range(K):
                blk.lookup_convc1(coords[k], W, bias)

        def unfused_step():
            blk = rr.CorrBlock(f1, f2, num_levels=L, radius=r, precision=prec)
            for k in range(K):
                with torch.autocast("cuda", dtype=torch.float16):
                    F.relu(F.conv2d(blk(coords[k]), W, bias))

        blk = rr.CorrBlock(f1, f2, num_levels=L, radius=r, precision=prec)
        with torch.no_grad():
            ms_f, ms_u = timeit(fused_step), timeit(unfused_step)
            us_call = timeit(lambda: blk.lookup_convc1(coords[0], W, bias), reps=50) * 1e3
        s_vol = 2 if prec == "bf16_vol16" else 4
        alg = B * N * (L * (2 * r + 2) ** 2 * s_vol + cout * 4 + 8)  # read the integer patches, write Cout fp32, read coords
        out[prec] = {"value": B / (ms_f * 1e-3), "unit": UNIT, "ms_per_step": ms_f, "unfused_ms_per_step": ms_u,
                     "fused_call_us": us_call, "fused_call_alg_bytes": alg, "fused_call_gbs": alg / (us_call * 1e-6) / 1e9}
    return out


def gpu_reference_run(args, f1, f2, coords, our_value):
    """Comparator leg, outside every timed region of this repository's own numbers: the reference's own code
    on the SAME GPU.  (1) Its torch-op CorrBlock (core/corr.py unmodified from oracle/_ref/core when ferried,
    else the op-for-op port), same batch, fp32 with TF32 off: build + K lookups.  (2) Its compiled
    alt_cuda_corr binary (oracle/_ref) against AlternateCorrBlock at config 5 (2160x3840 -> 270x480), one
    iteration = 4 levels."""
    import torch

    import raft_ros_b200 as rr

    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    K, L, r = args.iters, args.levels, args.radius
    B = f1.shape[0]
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def timeit(fn, warm, reps):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out = {}
    ref_core = None
    try:
        from oracle import build_ref_core
        ref_core = build_ref_core.load()
    except Exception:
        pass
    if ref_core is not None:
        RefCorrBlock = ref_core.corr.CorrBlock
        kind = "reference (core/corr.py unmodified, torch ops on this GPU)"

        def ref_step():
            blk = RefCorrBlock(f1, f2, num_levels=L, radius=r)
            for k in range(K):
                blk(coords[k])
    else:
        from oracle import torch_port as TP
        kind = "port (oracle/torch_port.py, torch ops on this GPU)"

        def ref_step():
            pyr = TP.build_pyramid(f1, f2, L)
            for k in range(K):
                TP.lookup(pyr, coords[k], r)

    import warnings
    with torch.no_grad(), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ms = timeit(ref_step, 2, 5)
    out["corrblock"] = {"value": B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "kind": kind,
                        "ours_over_reference": our_value / (B / (ms * 1e-3))}
    torch.cuda.empty_cache()
    # on-the-fly path at 4K
    try:
        from oracle import build_ref
        ext = build_ref.load_module()
    except Exception:
        ext = None
    if r == 4 and torch.cuda.mem_get_info()[0] > 8 * 2 ** 30:
        H, W, D = 270, 480, 256
        g = torch.Generator().manual_seed(5)
        a1 = torch.randn(1, D, H, W, generator=g).to(f1.device)
        a2 = torch.randn(1, D, H, W, generator=g).to(f1.device)
        c4 = (rr.coords_grid(1, H, W, "cpu") + 2.0 * torch.randn(1, 2, H, W, generator=g)).to(f1.device)
        alt = {"workload": "2160x3840 -> 270x480, D=256, 4 levels r=4, one iteration"}
        for prec in ("bf16x3", "fp16_fp8", "bf16"):
            blk = rr.AlternateCorrBlock(a1, a2, num_levels=4, radius=4, precision=prec)
            alt[f"ours_{prec}_ms"] = timeit(lambda: blk(c4), 2, 10)
        if ext is not None:
            blk = rr.AlternateCorrBlock(a1, a2, num_levels=4, radius=4)
            f1n = a1.permute(0, 2, 3, 1).contiguous()
            f2n = [x[1].permute(0, 2, 3, 1).contiguous() for x in blk.pyramid[:4]]
            cs = [(c4.permute(0, 2, 3, 1) / 2 ** i).reshape(1, 1, H, W, 2).contiguous() for i in range(4)]

            def ref_alt():
                for i in range(4):
                    ext.forward(f1n, f2n[i], cs[i], 4)

            alt["reference_binary_ms"] = timeit(ref_alt, 1, 3)
            alt["kind"] = "reference (alt_cuda_corr compiled unmodified, oracle/_ref)"
        out["alt_corr_4k"] = alt
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU arm runs once, on rank 0
    r = cpu_reference_run(args, args.steps, min(args.warmup, 2), args.cpu_budget_s)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["pairs_per_s"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "steps_timed": r["steps_timed"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args), "where": "cpu",
        "cpu_baseline": {"value": r["pairs_per_s"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": r["pairs_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args):
    h, w = args.height // 8, args.width // 8
    return {
        "workload": f"sintel 436x1024 padded to {args.height}x{args.width}: feature grid {h}x{w}, D={args.dim}, "
                    f"batch {args.batch} frame pairs per GPU, CorrBlock {args.levels} levels r={args.radius}, "
                    f"1 build + {args.iters} lookups per step",
        "batch_per_gpu": args.batch, "iters": args.iters, "levels": args.levels, "radius": args.radius,
        "feature_grid": [h, w], "dim": args.dim,
        "l2": "pyramid per step (2.09 GB fp32 at batch 8) far exceeds the 126 MB L2; no explicit flush",
    }


# ----------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    import raft_ros_b200 as rr
    from raft_ros_b200 import _lib
    from raft_ros_b200.corr import build_pyramid_raw, build_pyramid_tiled_raw, lookup_raw, lookup_tiled_raw

    tiled = args.layout == "tiled" and args.precision != "fp32" and args.radius in (3, 4) and args.levels <= 4
    os.environ["RAFT_CORR_LAYOUT"] = "tiled" if tiled else "rows"  # the e2e arm's CorrBlock follows
    build_fn = build_pyramid_tiled_raw if tiled else build_pyramid_raw
    look_fn = lookup_tiled_raw if tiled else lookup_raw

    _lib.build()
    _lib.lib()  # fail loudly if the CUDA library is missing

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    # host placement of the end-to-end arm: this rank's threads and the first touch of its pinned buffers go to
    # the NUMA node its GPU hangs off (no effect on the device-timed numbers)
    import contextlib

    from raft_ros_b200 import hostbind
    bind = hostbind.near_gpu(local, rank, world) if os.environ.get("RAFT_BENCH_NUMA_BIND", "1") != "0" else None

    def pinned_alloc():
        # `with pinned_alloc(): x = t.pin_memory()`: pages land on the GPU's NUMA node, then the thread is released
        return bind if bind is not None else contextlib.nullcontext()

    if args.l2_fetch:
        _lib.check(_lib.lib().rc_set_l2_fetch_granularity(args.l2_fetch), "rc_set_l2_fetch_granularity")
    import ctypes as _ct
    _g = _ct.c_int(0)
    _lib.lib().rc_get_l2_fetch_granularity(_ct.byref(_g))
    l2_fetch = int(_g.value)

    B, D, L, r, K = args.batch, args.dim, args.levels, args.radius, args.iters
    h, w = args.height // 8, args.width // 8
    N = h * w
    # every rank owns its own shard of frame pairs: no data-path collective
    g = torch.Generator().manual_seed(1234 + rank)
    base = rr.coords_grid(B, h, w, "cpu")
    with pinned_alloc():
        f1_host = torch.randn(B, D, h, w, generator=g).pin_memory()
        f2_host = torch.randn(B, D, h, w, generator=g).pin_memory()
        flows = torch.cumsum(0.5 * torch.randn(K, B, 2, h, w, generator=g), 0)
        coords_host = (base[None] + flows).contiguous().pin_memory()  # (K,B,2,h,w)
    f1, f2 = f1_host.to(dev), f2_host.to(dev)
    coords = coords_host.to(dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    build_ev, look_ev = [], []

    def step(timed):
        if timed:
            e0, e1, e2 = ev(), ev(), ev()
            e0.record()
        levels = build_fn(f1, f2, L, args.precision)
        if timed:
            e1.record()
        out = None
        for k in range(K):
            out = look_fn(levels, coords[k], r)
        if timed:
            e2.record()
            build_ev.append((e0, e1))
            look_ev.append((e1, e2))
        return out

    for _ in range(max(args.warmup, 3)):
        step(False)
    torch.cuda.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    t0, t1 = ev(), ev()
    t0.record()
    for _ in range(args.steps):
        step(True)
    t1.record()
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.stop()
    ms_total = t0.elapsed_time(t1)
    ms_step = ms_total / args.steps
    build_ms = sum(a.elapsed_time(b) for a, b in build_ev) / len(build_ev)
    look_ms = sum(a.elapsed_time(b) for a, b in look_ev) / len(look_ev) / K  # per launch
    if world > 1:
        tt = torch.tensor([ms_step], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_step = float(tt.item())
    value = world * B / (ms_step * 1e-3)

    # ---- end to end through the public API with HOST buffers (H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        # Host side: pinned inputs, one pinned result buffer per lookup of a step.  Three streams so that
        # the PCIe directions and the kernels overlap: feature maps of the next pair go up on `s_up`
        # while the previous pair's lookups are still being read back on `s_down`.
        with pinned_alloc():
            out_host = [torch.empty((B, L * (2 * r + 1) ** 2, h, w), dtype=torch.float32).pin_memory() for _ in range(K)]
            for o_ in out_host:
                o_.zero_()  # first touch
        s_main = torch.cuda.current_stream()
        s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
        ring = [[torch.empty_like(f1), torch.empty_like(f2), None] for _ in range(2)]  # device inputs, double-buffered

        def e2e_step(i):
            a, b, busy = ring[i & 1]
            with torch.cuda.stream(s_up):
                if busy is not None:
                    s_up.wait_event(busy)  # the build that read this slot two steps ago is done
                a.copy_(f1_host, non_blocking=True)
                b.copy_(f2_host, non_blocking=True)
                up = torch.cuda.Event()
                up.record(s_up)
            s_main.wait_event(up)
            blk = rr.CorrBlock(a, b, num_levels=L, radius=r, precision=args.precision)
            done = torch.cuda.Event()
            done.record(s_main)
            ring[i & 1][2] = done
            for k in range(K):
                c = coords_host[k].to(dev, non_blocking=True)
                o = blk(c)
                ready = torch.cuda.Event()
                ready.record(s_main)
                with torch.cuda.stream(s_down):
                    s_down.wait_event(ready)
                    out_host[k].copy_(o, non_blocking=True)
                o.record_stream(s_down)

        n_e2e = max(3, min(args.steps, 10))
        for i in range(2):
            e2e_step(i)
        torch.cuda.synchronize()
        barrier()
        s0, s1 = ev(), ev()
        s0.record()
        for i in range(n_e2e):
            e2e_step(i)
        s_main.wait_stream(s_down)  # the step is over when its last result is in host memory
        s_main.wait_stream(s_up)
        s1.record()
        torch.cuda.synchronize()
        barrier()
        e_ms = s0.elapsed_time(s1) / n_e2e
        if world > 1:
            tt = torch.tensor([e_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e_ms = float(tt.item())
        d2h_bytes = int(K * out_host[0].numel() * 4)
        e2e = {"value": world * B / (e_ms * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(2 * f1_host.numel() * 4 + K * coords_host[0].numel() * 4),
               "d2h_bytes_per_step": d2h_bytes, "ms_per_step": e_ms, "steps": n_e2e,
               "overlap": "H2D, kernels and D2H on three streams; device inputs double-buffered",
               "d2h_gbs": world * d2h_bytes / (e_ms * 1e-3) / 1e9, "host_binding": bind.info if bind is not None else {"bound": False}}
        # ceiling measured on this box type by tools/probe_d2h.py (aggregate pinned D2H rate at this GPU count)
        try:
            probe = json.load(open(os.path.join(ROOT, "profiles", "r02_d2h_probe.json")))
            ceil = probe.get(str(world), {}).get("bound_d2h", {}).get("aggregate_gbs")
            if ceil:
                e2e["probe_d2h_gbs"] = ceil
                e2e["frac_of_probe"] = e2e["d2h_gbs"] / ceil
        except Exception:
            pass

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    b_build, b_look, b_total = algorithmic_bytes(B, N, D, h, w, L, r, K, 2 if args.precision == "bf16_vol16" else 4)
    kern = {
        "build": {"launches_per_step": 2, "ms": build_ms, "alg_bytes": b_build, "gbs": b_build / (build_ms * 1e-3) / 1e9,
                  "share_of_step": build_ms / ms_step, "flops": 2.0 * B * N * N * D * {"bf16x3": 3, "fp16_fp8": 2}.get(args.precision, 1),
                  "tflops": 2.0 * B * N * N * D * {"bf16x3": 3, "fp16_fp8": 2}.get(args.precision, 1) / (build_ms * 1e-3) / 1e12,
                  "flops_note": "bf16-equivalent tensor work: bf16x3 = 3 passes; fp16_fp8 = 1 fp16 pass + 1 e4m3 pass of twice the K at twice the rate"},
        "lookup": {"launches_per_step": K, "ms": look_ms, "alg_bytes": b_look, "gbs": b_look / (look_ms * 1e-3) / 1e9,
                   "share_of_step": look_ms * K / ms_step},
    }
    dom = "build" if build_ms >= look_ms * K else "lookup"
    traffic, traffic_ncu = measured_traffic(dom, args, "tiled" if tiled else "rows")
    roofline = {"bound": "hbm", "kernel": "build_tc_kernel (prep + fused volume/pyramid)" if dom == "build"
                else ("lookup_tiled_fwd_kernel" if tiled else "lookup_fwd_kernel"),
                "achieved": kern[dom]["gbs"], "peak": peak, "unit": "GB/s", "frac": kern[dom]["gbs"] / peak,
                "traffic": traffic, "traffic_ncu": traffic_ncu, "peak_source": peak_src,
                "whole_step": {"alg_bytes": b_total, "gbs": b_total / (ms_step * 1e-3) / 1e9,
                               "frac": b_total / (ms_step * 1e-3) / 1e9 / peak}}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"fp32": "f32", "bf16x3": "bf16x3 (hi/lo split, f32 accumulate) + f32 lookup", "bf16": "bf16 (f32 accumulate) + f32 lookup",
                  "bf16_vol16": "bf16 (f32 accumulate), fp16 volume + f32 lookup",
                  "fp16_fp8": "fp16 products + e4m3 cross terms (f32 accumulate) + f32 lookup"}[args.precision],
        "data": "synthetic", "config": workload_config(args), "where": "gpu", "precision": args.precision,
        "roofline": roofline, "kernels": kern, "clocks": clocks, "gpu_launches": int(args.steps * (2 + K)),
        "l2_fetch_granularity": l2_fetch, "layout": "tiled" if tiled else "rows",
    }
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and tiled and not args.no_fast_paths and args.precision in ("bf16x3", "fp16_fp8"):
        # the same step with the other precisions, for the record (not the headline value): the other fp32-class one and the
        # stated-bounds fast paths
        extra = {}
        for prec in (("bf16x3" if args.precision == "fp16_fp8" else "fp16_fp8"), "bf16", "bf16_vol16"):
            def fstep():
                lv = build_pyramid_tiled_raw(f1, f2, L, prec)
                for k in range(K):
                    lookup_tiled_raw(lv, coords[k], r)
            for _ in range(3):
                fstep()
            torch.cuda.synchronize()
            a0, a1 = ev(), ev()
            a0.record()
            nrep = max(10, min(args.steps, 30))
            for _ in range(nrep):
                fstep()
            a1.record()
            torch.cuda.synchronize()
            ms = a0.elapsed_time(a1) / nrep
            extra[prec] = {"value": B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": nrep}
        line["fast_paths"] = extra
    if world == 1 and tiled and not args.no_fast_paths and args.radius in (3, 4) and args.levels <= 4:
        try:
            line["fused"] = fused_convc1_run(args, f1, f2, coords)
        except Exception as e:
            log(f"[bench] fused lookup+convc1 leg failed: {type(e).__name__}: {e}")
    if world == 1 and tiled and not args.no_fast_paths and args.radius == 4 and args.levels == 4:
        try:
            rf = raft_forward_run(args)
            if rf is not None:
                line["raft_forward"] = rf
        except Exception as e:
            log(f"[bench] RAFT.forward leg failed: {type(e).__name__}: {e}")
    if world == 1 and not args.no_gpu_reference:
        try:
            line["gpu_reference"] = gpu_reference_run(args, f1, f2, coords, value)
        except Exception as e:
            log(f"[bench] gpu reference leg failed: {type(e).__name__}: {e}")
    if world == 1 and not args.no_cpu_baseline:
        try:
            c = cpu_reference_run(args, 2, 1, min(args.cpu_budget_s, 60.0))
            line["cpu_baseline"] = {"value": c["pairs_per_s"], "unit": UNIT, "cores": c["cores"], "kind": c["kind"], "sample": c["sample"]}
        except Exception as e:
            log(f"[bench] cpu baseline failed: {e}")
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
