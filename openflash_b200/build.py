"""Build libmeemb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m openflash_b200.build [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libmeemb200.so")
SOURCES = ["meem_kernels.cu"]
DEPS = ["meem_kernels.cu", "meem_lu.cuh", "meem_forms.cuh", "meem_math.cuh", "meem_dmma.cuh", "meem_chol.cuh",
        "meem_geom_coop.cuh", "meem_geom_band.cuh", "bessel_coeffs.h",
        os.path.join("..", "..", "include", "meemb200.h")]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + SOURCES
    subprocess.check_call(cmd, cwd=CSRC)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
