"""Build the REFERENCE's own CUDA extensions (gsplat.csrc and clm_kernels._C) for sm_100 from the
sources where they lie under /root/reference, into oracle/_ref/ (git-ignored, travels to the GPU
box).  Test infrastructure only: used by tests/test_gpu_vs_reference_cuda.py and tests/ref_gpu_compare.py on the GPU box to compare our
kernels against the reference's kernels on identical inputs and to time them.

No reference source is copied; this script only points torch.utils.cpp_extension at the files.
Runs only where /root/reference exists (the build container).
"""
import glob
import os
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")


def build():
    if not os.path.isdir(REF):
        print("no /root/reference here: nothing to build")
        return
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "6")
    from torch.utils.cpp_extension import load

    g = os.path.join(REF, "submodules/gsplat/gsplat/cuda/csrc")
    srcs = sorted(glob.glob(os.path.join(g, "*.cu"))) + [os.path.join(g, "ext.cpp")]
    load(name="gsplat_ref_csrc", sources=srcs, extra_include_paths=[os.path.join(g, "third_party", "glm")],
         extra_cflags=["-O3"], extra_cuda_cflags=["-O3", "--use_fast_math"], build_directory=_mk("gsplat"), verbose=True,
         is_python_module=False)
    c = os.path.join(REF, "submodules/clm_kernels")
    srcs = sorted(glob.glob(os.path.join(c, "*.cu"))) + [os.path.join(c, "ext.cpp")]
    load(name="clm_kernels_ref_C", sources=srcs, extra_include_paths=[os.path.join(c, "third_party", "glm")],
         extra_cflags=["-O3"], extra_cuda_cflags=["-O3"], build_directory=_mk("clm_kernels"), verbose=True,
         is_python_module=False)


def _mk(name):
    d = os.path.join(OUT, name)
    os.makedirs(d, exist_ok=True)
    return d


if __name__ == "__main__":
    build()
