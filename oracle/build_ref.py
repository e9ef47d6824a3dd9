"""Build the UNMODIFIED reference extension for sm_100 into oracle/_ref/ (TEST INFRASTRUCTURE ONLY).

The reference's CUDA kernels have no CPU path, so the runnable oracle for
iter_proj / refine_matches / gauss_newton_* is the reference's own extension
(mast3r_slam/backend/src/{gn.cpp,gn_kernels.cu,matching_kernels.cu}) compiled
from the sources where they lie under /root/reference, with
  * oracle/ref_compat.h force-included (two arithmetic-neutral API shims for
    torch 2.11, see that header), and
  * the same optimisation flags as the reference's setup.py:20-37 (-O3, default
    FMA contraction, no fast-math), retargeted at sm_100 (the reference ships
    sm_60..sm_86 SASS only).
The reference's own setup.py cannot be used here: it gates the CUDA sources on
torch.cuda.is_available() (setup.py:9,24,46-50) and this container has no GPU.

Output: oracle/_ref/mast3r_slam_backends_ref.so (git-ignored, travels to the
GPU box with gpurun).  Nothing under oracle/ is imported by the product.
"""
import os
import sys
import time

REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "mast3r_slam_backends_ref"


def so_path():
    return os.path.join(OUT, NAME + ".so")


def build(verbose=False, force=False):
    if os.path.exists(so_path()) and not force:
        return so_path()
    src = os.path.join(REF, "mast3r_slam", "backend", "src")
    if not os.path.isdir(src):
        raise FileNotFoundError(f"reference sources not found under {REF}")
    os.makedirs(OUT, exist_ok=True)
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils.cpp_extension import load

    compat = os.path.join(HERE, "ref_compat.h")
    t0 = time.time()
    load(
        name=NAME,
        sources=[os.path.join(src, f) for f in ("gn.cpp", "gn_kernels.cu", "matching_kernels.cu")],
        extra_include_paths=[
            os.path.join(REF, "mast3r_slam", "backend", "include"),
            os.path.join(REF, "thirdparty", "eigen"),
        ],
        extra_cflags=["-O3", "-include", compat],
        extra_cuda_cflags=["-O3", "-include", compat],
        build_directory=OUT,
        verbose=verbose,
        is_python_module=True,
    )
    print(f"[oracle/_ref] built {so_path()} in {time.time() - t0:.0f}s")
    return so_path()


def load_ref():
    """Import the prebuilt reference extension (GPU box: prebuilt file only)."""
    import importlib.util

    import torch  # noqa: F401  (libtorch must be loaded first)

    p = so_path()
    if not os.path.exists(p):
        return None
    spec = importlib.util.spec_from_file_location(NAME, p)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    build(verbose="-v" in sys.argv, force="--force" in sys.argv)
