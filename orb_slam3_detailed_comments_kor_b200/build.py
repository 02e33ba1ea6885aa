"""Build liborbfe.so (CUDA, sm_100a) and libsynth.so in-tree.  nvcc cross-compiles without a GPU."""
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(_HERE, "liborbfe.so")
SOURCES = ["orbfe_host.cu", "orbfe_extract.cu", "orbfe_match.cu", "orbfe_condition.cu", "orbfe_project.cu", "orbfe_bow.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--fmad=false",
              "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function,-Wno-unknown-pragmas", "-shared", "-cudart", "static", "-ldl"] + os.environ.get("ORBFE_EXTRA_NVCC", "").split()


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, "orbfe_internal.h"), os.path.join(_HERE, "..", "include", "orbfe.h"),
                   os.path.join(_HERE, "..", "include", "orbfe_pattern31.inc"), os.path.abspath(__file__)]
    if force or _stale(LIB, deps):
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + srcs
        subprocess.check_call(cmd)
    from . import synth
    synth.build(force)
    return LIB


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(_HERE))
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
