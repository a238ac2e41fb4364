"""Recipe that places the REFERENCE's own `core/` Python package under oracle/_ref/core/.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  The reference's CorrBlock path is ~170 lines of pure Python
over torch (core/corr.py, core/utils/utils.py) and its model shell, encoders and update block are
more of the same (core/raft.py, core/extractor.py, core/update.py).  The GPU box has no reference
checkout, so -- exactly like the compiled alt_cuda_corr binary (build_ref.py) and the TorchScript
sub-networks (build_ref_model.py) -- the files are ferried through oracle/_ref/, which is git-ignored
(nothing of the reference enters this repository's history) but travels with the snapshot.  They are
used, unmodified, as

  * the CPU baseline of bench.py (`cpu_baseline.kind = "reference"`, `--impl reference`): the
    reference's own `CorrBlock` timed on the box's host cores;
  * the harness of tests/test_gpu_raft_forward.py: the reference's own `RAFT.forward` run once as is
    and once with this repository's CorrBlock / AlternateCorrBlock swapped in at core/raft.py:104-107.

Runs only in the build container.  `load()` imports the copy as the top-level package `core`.
"""
import importlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
FILES = ["__init__.py", "raft.py", "corr.py", "update.py", "extractor.py", os.path.join("utils", "__init__.py"),
         os.path.join("utils", "utils.py")]


def built():
    return all(os.path.exists(os.path.join(OUT, "core", f)) for f in FILES)


def build(reference="/root/reference"):
    src = os.path.join(reference, "core")
    if not os.path.isdir(src):
        raise FileNotFoundError(src)
    for f in FILES:
        dst = os.path.join(OUT, "core", f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(os.path.join(src, f)):
            shutil.copyfile(os.path.join(src, f), dst)
    return os.path.join(OUT, "core")


def load():
    """The reference's `core` package (core.raft, core.corr, core.utils.utils), or None if it was not ferried."""
    if not built():
        return None
    if "core" in sys.modules and not getattr(sys.modules["core"], "__file__", "").startswith(OUT):
        raise RuntimeError("another top-level package named 'core' is already imported")
    if OUT not in sys.path:
        sys.path.insert(0, OUT)
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        core = importlib.import_module("core")
        importlib.import_module("core.corr")
        importlib.import_module("core.utils.utils")
    return core


if __name__ == "__main__":
    print(build())
