"""TEST INFRASTRUCTURE ONLY.  Compiles the reference's own CUDA sources, where they lie under /root/reference, into
oracle/_ref/ (git-ignored; travels to the GPU box with the snapshot):

  libref_rasterizer.so  R/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu + oracle/ref_shim.cu
  libref_knn.so         K/simple_knn.cu + oracle/ref_knn_shim.cu

No reference source is copied or edited: the two `-include` flags only supply the <cstdint>/<cfloat> declarations the
sources rely on transitively with older toolchains.  The reference's own build system (setup.py /
CMake + torch extension) is not used.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("GSR_REFERENCE_ROOT", "/root/reference")
R = os.path.join(REF, "submodules_fsgs", "diff-gaussian-rasterization-confidence")
K = os.path.join(REF, "submodules_fsgs", "simple-knn")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared", "-w", "-lineinfo"]


def available() -> bool:
    return os.path.isdir(os.path.join(R, "cuda_rasterizer"))


def build(force: bool = False) -> bool:
    """Returns True when both libraries exist afterwards."""
    ras = os.path.join(OUT, "libref_rasterizer.so")
    knn = os.path.join(OUT, "libref_knn.so")
    if not available():
        return os.path.exists(ras) and os.path.exists(knn)
    os.makedirs(OUT, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    if force or not os.path.exists(ras):
        srcs = [os.path.join(R, "cuda_rasterizer", f) for f in ("forward.cu", "backward.cu", "rasterizer_impl.cu")]
        cmd = [nvcc, *ARCH, *COMMON, "-include", "cstdint", "-I", R, "-I", os.path.join(R, "third_party", "glm"),
               *srcs, os.path.join(HERE, "ref_shim.cu"), "-o", ras]
        subprocess.check_call(cmd)
    if force or not os.path.exists(knn):
        cmd = [nvcc, *ARCH, *COMMON, "-include", "cfloat", "-I", K, os.path.join(K, "simple_knn.cu"),
               os.path.join(HERE, "ref_knn_shim.cu"), "-o", knn]
        subprocess.check_call(cmd)
    return True


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
