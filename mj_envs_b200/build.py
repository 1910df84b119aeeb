"""Builds the CUDA library in-tree: mj_envs_b200/libadroit_b200.so (sm_100a only)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "adroit_b200.cu")
LIB = os.path.join(HERE, "libadroit_b200.so")
import glob

DEPS = sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")) + glob.glob(os.path.join(HERE, "csrc", "*.cuh")))   # every source the library includes
DEPS.append(os.path.join(HERE, "..", "include", "adroit_b200.h"))


# fp32 division / square root as MUFU + one Newton step (2 ulp) instead of the IEEE-rounded slow-path sequences: the fp32 path's error
# against the fp64 oracle is unchanged to three digits (profiles/R2f_fastdiv.txt), the step is 3 % faster.  fp64 arithmetic (the
# ADROIT_F64 build, the MPR portal refinement) is not affected by these flags.
FP32_FLAGS = ["-prec-div=false", "-prec-sqrt=false"]


def nvcc():
    for p in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if p and (os.path.exists(p) or p == "nvcc"):
            return p
    return "nvcc"


def build(force=False, verbose=False, defines=(), out=None):
    """`defines`/`out` build an experiment variant (tools/ab.sh); the product library takes neither."""
    if out or defines:
        if not out:
            raise ValueError("an experiment variant (defines=...) needs its own output path (out=...)")
        cmd = [nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--expt-relaxed-constexpr", *FP32_FLAGS, "-shared",
               "-Xcompiler", "-fPIC", "-o", out, SRC] + [f"-D{d}" for d in defines] + os.environ.get("ADROIT_NVCC_EXTRA", "").split()
        subprocess.check_call(cmd)
        return out
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in DEPS):
        return LIB
    cmd = [nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--expt-relaxed-constexpr", *FP32_FLAGS, "-shared",
           "-Xcompiler", "-fPIC", "-o", LIB, SRC]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a[2:] for a in sys.argv[1:] if a.startswith("-o")]
    print(build(force="-f" in sys.argv, verbose="-v" in sys.argv, defines=defs, out=outs[0] if outs else None))
