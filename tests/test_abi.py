"""CPU suite: the C-ABI library builds for sm_100a, loads without a GPU and exports every
symbol include/raftcorr.h declares; host-side argument checking mirrors the reference."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "raftcorr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(rc_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_expected_surface():
    names = declared_functions()
    for must in ["rc_corr_build_pyramid", "rc_corr_lookup_fwd", "rc_corr_lookup_bwd", "rc_corr_build_bwd",
                 "rc_alt_corr_fwd", "rc_alt_corr_bwd", "rc_bilinear_sampler_fwd", "rc_last_error", "rc_version"]:
        assert must in names


def test_library_exports_every_declared_symbol(built_lib):
    from raft_ros_b200 import _lib
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_functions():
        assert hasattr(handle, name), f"{name} declared in include/raftcorr.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in raft_ros_b200/_lib.py"
    assert built_lib.rc_version() >= 100


def test_header_is_plain_c():
    gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    code = '#include "raftcorr.h"\nint (*fp)(void) = rc_version;\nint main(void){return fp != 0;}\n'
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-x", "c", "-", "-c", "-o", os.devnull],
                       input=code, text=True, capture_output=True)
    assert r.returncode == 0, r.stderr


def test_sass_is_sm100a(built_lib):
    from raft_ros_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_argument_validation_without_gpu(built_lib):
    i, p = ctypes.c_int, ctypes.POINTER(ctypes.c_int)
    hl, wl = i(), i()
    assert built_lib.rc_level_dims(55, 128, 3, ctypes.byref(hl), ctypes.byref(wl)) == 0
    assert (hl.value, wl.value) == (6, 16)  # floor mode: 55>>3, 128>>3
    assert built_lib.rc_level_dims(4, 128, 3, ctypes.byref(hl), ctypes.byref(wl)) != 0
    assert b"empty" in built_lib.rc_last_error()
    # null pointers are rejected before any CUDA call
    rc = built_lib.rc_corr_lookup_fwd(None, None, None, 1, 16, 16, 4, 4, None)
    assert rc == 1 and b"null" in built_lib.rc_last_error()
    rc = built_lib.rc_alt_corr_fwd(None, None, None, None, 1, 1, 8, 8, 8, 8, 30, 4, None)
    assert rc == 3 and b"multiple of 4" in built_lib.rc_last_error()


def test_cpu_tensors_raise_like_the_reference(built_lib):
    """alt_cuda_corr/correlation.cpp:19-21: 'fmap1 must be a CUDA tensor' / 'must be contiguous'."""
    torch = pytest.importorskip("torch")
    import raft_ros_b200 as rr
    f = torch.zeros(1, 8, 8, 32)
    c = torch.zeros(1, 1, 8, 8, 2)
    with pytest.raises(RuntimeError, match="fmap1 must be a CUDA tensor"):
        rr.alt_cuda_corr.forward(f, f, c, 4)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        rr.CorrBlock(torch.zeros(1, 32, 16, 16), torch.zeros(1, 32, 16, 16))
    g = rr.coords_grid(2, 3, 5, "cpu")
    assert g.shape == (2, 2, 3, 5) and g[0, 0, 1, 4] == 4 and g[0, 1, 2, 0] == 2


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "raft_ros_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, fn)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "corr_oracle" not in txt, fn
