"""Builds libo3dpp_b200.so (the C ABI of include/o3dpp.h) in-tree with nvcc for sm_100a.

    python open3d-py-extension_b200/build.py [--force]

The library is one translation unit (csrc/o3dpp.cu).  -fmad=false / -ffp-contract=off are part
of the contract, not tuning: the FP64 expressions must round like the CPU path they replace.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libo3dpp_b200.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "o3dpp.h")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden",
    "-shared",
]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))) + [HEADER]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources())


def nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def build(force=False, verbose=False, out=None, defines=()):
    """out / defines: tuning builds (tools/variants.py), e.g. defines=("KL_WARPS=2",) into another file"""
    if out is None and not force and not stale():
        return LIB
    out = out or LIB
    cmd = [nvcc()] + NVCC_FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-o", out, os.path.join(CSRC, "o3dpp.cu")]
    subprocess.check_call(cmd, cwd=CSRC)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
