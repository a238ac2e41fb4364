"""ctypes binding of libraftcorr.so (the C ABI declared in include/raftcorr.h).

The library is built in-tree (raft_ros_b200/libraftcorr.so) by `build()` below, which is what
`__graft_entry__.build()` calls.  There is NO fallback: if the shared library is missing or a
kernel fails, the call raises.  Nothing under oracle/ is ever imported from here.
"""
from __future__ import annotations

import ctypes
import glob
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("RAFTCORR_LIB") or os.path.join(_HERE, "libraftcorr.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "raftcorr.h")

RC_OK = 0
RC_PREC_FP32 = 0
RC_PREC_BF16X3 = 1
RC_PREC_BF16 = 2
RC_PREC_BF16_VOL16 = 3
RC_PREC_FP16_FP8 = 4
RC_MAX_LEVELS = 8
RC_MAX_RADIUS = 8

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]

_lock = threading.Lock()
_lib = None


def sources():
    return sorted(glob.glob(os.path.join(_CSRC, "*.cu")))


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + glob.glob(os.path.join(_CSRC, "*.cuh")) + [HEADER_PATH]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo ... -shared -> libraftcorr.so

    Safe to call from several processes at once (one rank per GPU under torchrun): the compile
    runs under an exclusive file lock, into a temporary file that is renamed into place."""
    if not force and not _stale():
        return LIB_PATH
    import fcntl

    lock_path = os.path.join(_HERE, ".build.lock")
    with open(lock_path, "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():  # another process built it while we waited
                return LIB_PATH
            nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
            if not os.path.exists(nvcc):
                nvcc = "nvcc"
            # the image exports CC/CXX=/opt/gcc/...; nvcc picks its host compiler with -ccbin
            ccbin = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
            tmp = LIB_PATH + f".tmp{os.getpid()}"
            cmd = [nvcc, "-ccbin", ccbin] + NVCC_FLAGS + ["-shared", "-o", tmp] + sources()
            res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.remove(tmp)
                raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout)
            os.replace(tmp, LIB_PATH)
            if verbose:
                print(res.stdout)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH


_vp = ctypes.c_void_p
_i = ctypes.c_int
_f = ctypes.c_float
_sz = ctypes.c_size_t
_ll = ctypes.c_longlong

# name -> (restype, argtypes); mirrors include/raftcorr.h one to one
SIGNATURES = {
    "rc_version": (_i, []),
    "rc_last_error": (ctypes.c_char_p, []),
    "rc_set_l2_fetch_granularity": (_i, [_i]),
    "rc_get_l2_fetch_granularity": (_i, [ctypes.POINTER(_i)]),
    "rc_level_dims": (_i, [_i, _i, _i, ctypes.POINTER(_i), ctypes.POINTER(_i)]),
    "rc_corr_build_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "rc_corr_build_pyramid": (_i, [_vp, _vp, _i, _i, _i, _i, ctypes.POINTER(_vp), _i, _i, _f, _vp, _sz, _vp]),
    "rc_corr_build_pyramid_noalloc": (_i, [_vp, _vp, _i, _i, _i, _i, ctypes.POINTER(_vp), _i, _i, _f, _vp]),
    "rc_corr_lookup_fwd": (_i, [ctypes.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_tile_group": (_i, []),
    "rc_corr_tiled_level_elems": (_i, [_i, _i, _i, _i, ctypes.POINTER(_sz)]),
    "rc_corr_build_pyramid_tiled": (_i, [_vp, _vp, _i, _i, _i, _i, ctypes.POINTER(_vp), _i, _i, _f, _vp, _sz, _vp]),
    "rc_corr_build_pyramid_tiled_prepared": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, ctypes.POINTER(_vp), _i, _i, _f, _vp]),
    "rc_fnet_conv1x1_weight_bytes": (_sz, [_i, _i]),
    "rc_fnet_conv1x1_prep_weight": (_i, [_vp, _vp, _i, _i, _vp]),
    "rc_fnet_conv1x1_prep": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_relayout_level": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_tiled16_level_elems": (_i, [_i, _i, _i, _i, ctypes.POINTER(_sz)]),
    "rc_corr_lookup_tiled16_fwd": (_i, [ctypes.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_relayout16_level": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_lookup_tiled_fwd": (_i, [ctypes.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_convc1_weight_bytes": (_sz, [_i, _i, _i]),
    "rc_corr_convc1_prep_weight": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "rc_corr_lookup_convc1_fwd": (_i, [ctypes.POINTER(_vp), _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_corr_lookup_bwd": (_i, [_vp, _vp, ctypes.POINTER(_vp), _i, _i, _i, _i, _i, _vp]),
    "rc_corr_build_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "rc_corr_build_bwd": (_i, [ctypes.POINTER(_vp), _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _vp, _sz, _vp]),
    "rc_alt_corr_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_alt_prep_fmap": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "rc_alt_prep_fmap_level": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rc_alt_corr_pyramid_setup_bytes": (_sz, []),
    "rc_alt_corr_pyramid_setup": (_i, [_vp, _vp, _vp, ctypes.POINTER(_vp), ctypes.POINTER(_vp), _i, _i, _i, _i, _i, _i, _i, _i, _i]),
    "rc_alt_corr_pyramid_run": (_i, [_vp, _vp, _vp, _f, _vp, _sz, _vp]),
    "rc_alt_corr_pyramid_plan_bytes": (_sz, [_i, _i, _i, _i]),
    "rc_alt_corr_pyramid_fwd": (_i, [_vp, _vp, ctypes.POINTER(_vp), ctypes.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _i,
                                     _i, _f, _vp, _sz, _vp]),
    "rc_alt_corr_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_conv_weight_bytes": (_sz, [_i, _i, _i, _i]),
    "rc_conv_prep_weight": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "rc_conv2d_fwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_alt_absmax": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "rc_alt_prep_fmap_level_f8": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_alt_corr_pyramid_set_norm": (_i, [_vp, _vp, _vp]),
    "rc_conv2d_ex_fwd": (_i, [_vp, _i, _ll, _vp, _ll, _vp, _vp, _vp, _ll, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _ll, _vp, _ll, _vp]),
    "rc_small_conv_relu_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_flow_update": (_i, [_vp, _vp, _ll, _vp, _vp, _ll, _vp, _ll, _i, _i, _i, _vp]),
    "rc_convex_upsample_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "rc_alt_corr_bwd_det_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "rc_alt_corr_bwd_det": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "rc_bilinear_sampler_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "rc_bilinear_sampler_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
}


def lib():
    """Load libraftcorr.so; raise loudly if it is not there (no CPU / eager fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(raft_ros_b200 has no fallback path)")
            handle = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(handle, name)  # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != RC_OK:
        msg = lib().rc_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what}: {msg} (rc={rc})" if what else f"{msg} (rc={rc})")


def ptr_array(ptrs):
    arr = (_vp * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr
