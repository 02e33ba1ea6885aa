"""Build liborbfe.so (CUDA, sm_100a) and libsynth.so in-tree.  nvcc cross-compiles without a GPU."""
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(_HERE, "liborbfe.so")
SOURCES = ["orbfe_host.cu", "orbfe_extract.cu", "orbfe_match.cu", "orbfe_condition.cu", "orbfe_project.cu", "orbfe_bow.cu", "orbfe_init.cu"]
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
    build_shim_probe(force)
    return LIB


SHIM_PROBES = {"shim_probe": [], "shim_probe_lazy": ["-DORBFE_SHIM_LAZY_PYRAMID"]}


def build_shim_probe(force=False):
    """tests/shim_smoke.cpp compiled against the C++ drop-in headers (include/orbfe_shim) and liborbfe.so: bench.py runs it with
    --latency to time ORBextractor::operator() as a C++ caller sees it, with the mvImagePyramid mirrors and without."""
    root = os.path.dirname(_HERE)
    src = os.path.join(root, "tests", "shim_smoke.cpp")
    deps = [src, os.path.join(root, "include", "orbfe_shim", "ORBextractor.h"), os.path.join(root, "include", "orbfe_shim", "orbfe_matching.h"),
            os.path.join(root, "include", "orbfe_shim", "opencv_min.hpp"), os.path.join(root, "include", "orbfe.h"), LIB]
    for name, flags in SHIM_PROBES.items():
        exe = os.path.join(_HERE, name + ".bin")
        if force or _stale(exe, deps):
            subprocess.check_call(["g++", "-std=c++17", "-O2"] + flags + ["-I" + os.path.join(root, "include"), src, "-o", exe, "-L" + _HERE, "-lorbfe",
                                   "-L" + os.path.join(_HERE, "synth"), "-lsynth", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/synth"])


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(_HERE))
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
