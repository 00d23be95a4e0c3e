"""Build the UNMODIFIED reference CUDA extension into oracle/_ref/ (test infrastructure only).

The reference (leejaeyong7/nerfacc v0.5.3) keeps its whole hot path in four CUDA
sources + one pybind file under /root/reference/nerfacc/cuda/csrc.  They compile
as-is for sm_100a with torch 2.11 / nvcc 12.9, so this script compiles them *where
they lie* (no source is copied into this repo) and drops `nerfacc_cuda.so` into
`oracle/_ref/` (git-ignored; it still travels to the GPU box with gpurun).

The result cannot run in the authoring container (no GPU).  It is used on the B200
box, by tests/ and tests/golden/make_golden_gpu.py only, as the strongest oracle:
the reference's own kernels on the same inputs.  Nothing in the product path
(nerfacc_b200/) ever imports it.

Usage:  python oracle/build_ref.py          (≈5 min, needs /root/reference)
"""
import glob
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_CSRC = os.environ.get("NERFACC_REF_CSRC", "/root/reference/nerfacc/cuda/csrc")
OUT = os.path.join(HERE, "_ref")


def build(verbose: bool = False) -> str:
    so = os.path.join(OUT, "nerfacc_cuda.so")
    if not os.path.isdir(REF_CSRC):
        if os.path.exists(so):
            return so
        raise FileNotFoundError(f"reference sources not found at {REF_CSRC}")
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    from torch.utils.cpp_extension import load

    sources = sorted(glob.glob(os.path.join(REF_CSRC, "*.cu"))) + sorted(
        glob.glob(os.path.join(REF_CSRC, "*.cpp"))
    )
    # Same flags as the reference's own JIT path (nerfacc/cuda/_backend.py:44-46).
    load(
        name="nerfacc_cuda",
        sources=sources,
        extra_cflags=["-O3"],
        extra_cuda_cflags=["-O3"],
        build_directory=OUT,
        verbose=verbose,
        is_python_module=False,
    )
    assert os.path.exists(so), so
    return so


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
