"""Recipe that compiles the REFERENCE's own alt_cuda_corr extension into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Sources are compiled where they lie under the reference checkout
(never copied into this repository); only the resulting binary lands in oracle/_ref/ (git-ignored,
but it travels to the GPU box with the snapshot).  It is the second oracle and the kernel-to-beat
for the on-the-fly path (alt_cuda_corr/correlation.cpp + correlation_kernel.cu, unmodified).

Runs only in the build container (the reference checkout does not exist on the GPU box).
"""
import glob
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "alt_cuda_corr_ref"


def built_path():
    hits = glob.glob(os.path.join(OUT, NAME + "*.so"))
    return hits[0] if hits else None


def build(reference="/root/reference", verbose=False):
    src = os.path.join(reference, "alt_cuda_corr")
    sources = [os.path.join(src, "correlation.cpp"), os.path.join(src, "correlation_kernel.cu")]
    for s in sources:
        if not os.path.exists(s):
            raise FileNotFoundError(s)
    if built_path():
        return built_path()
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    for k in ("CC", "CXX"):  # the image's /opt/gcc lacks pieces; let torch pick the system g++
        if os.environ.get(k, "").startswith("/opt/gcc"):
            os.environ[k] = "/usr/bin/gcc" if k == "CC" else "/usr/bin/g++"
    from torch.utils.cpp_extension import load

    load(name=NAME, sources=sources, build_directory=OUT, extra_cflags=["-O3"],
         extra_cuda_cflags=["-O3", "-gencode", "arch=compute_100a,code=sm_100a"], verbose=verbose, is_python_module=True)
    return built_path()


def load_module():
    """Import the prebuilt reference extension (needs torch imported first); None if absent."""
    p = built_path()
    if p is None:
        return None
    import importlib.util

    import torch  # noqa: F401  (libtorch must be loaded before the extension)

    spec = importlib.util.spec_from_file_location(NAME, p)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
