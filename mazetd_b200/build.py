"""Builds libmazetd_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmazetd_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # no FMA contraction: float32 results must match the reference's unfused arithmetic
    "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "--extended-lambda",
    "-shared", "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + [os.path.join(ROOT, "include", "mazetd_b200.h"), os.path.abspath(__file__)]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("MTD_NVCC_EXTRA", "").split()  # developer probes only (e.g. -DMTD_TIMING)
    tmp = LIB + ".tmp"  # the library appears atomically: a snapshot taken for the GPU box meanwhile ships the old or the new one
    cmd = [nvcc] + NVCC_FLAGS + extra + ["-o", tmp, os.path.join(CSRC, "capi.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(HERE, "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libmazetd_b200.so")
    os.replace(tmp, LIB)
    if verbose:
        print(r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    build(force=True, verbose=True)
