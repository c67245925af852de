"""Builds csrc/libhlsmap.so (sm_100a only) with nvcc, in-tree."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
# HLSMAP_TRACE_BUILD=1 in the environment selects a separate diagnostics library whose search
# kernel records per-tile phase timestamps (scripts/trace_run.py); the product library has none.
TRACE = os.environ.get("HLSMAP_TRACE_BUILD") == "1"
# Experiments: HLSMAP_VARIANT=name (+ HLSMAP_VARIANT_FLAGS="-DX=1 ...") builds and loads
# csrc/libhlsmap_<name>.so beside the product library, so one GPU session can compare builds.
VARIANT = os.environ.get("HLSMAP_VARIANT", "")
VARIANT_FLAGS = os.environ.get("HLSMAP_VARIANT_FLAGS", "").split()
LIB = os.path.join(CSRC, "libhlsmap_trace.so" if TRACE else
                   ("libhlsmap_%s.so" % VARIANT if VARIANT else "libhlsmap.so"))
SOURCES = ["resolve.cu", "views.cu", "ingest.cu", "json.cu", "api.cu"]
HEADERS = ["common.cuh", "scan.cuh", "decimal.cuh", "dtoa.cuh", "swar.cuh", "pow10_table.inc"]

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.sep not in c or os.path.exists(c)):
            return c
    return "nvcc"


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(INCLUDE, "hlsmap.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every CUDA source into one shared library; returns its path."""
    if not force and not needs_build():
        return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + (["-DHLSMAP_TRACE_BUILD=1"] if TRACE else []) + VARIANT_FLAGS + \
        ["-I", INCLUDE, "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
