"""Builds liblbm_b200.so in-tree with nvcc for sm_100a.

Flags that matter for parity: ``-fmad=false`` (one IEEE operation
per source operation, no FMA contraction), default IEEE division/sqrt
(``-prec-div=true -prec-sqrt=true``, never ``--use_fast_math``), and the host
compiler is told ``-ffp-contract=off`` as well because the MRT tables and the
cos() force profile are evaluated on the host.
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "liblbm_b200.so")
# same ABI, D3Q19 only, populations stored as float (csrc/lbm_pop.h); not the parity build
LIB_F32 = os.path.join(PKG, "liblbm_b200_f32.so")
SOURCES = ["lbm_api.cu", "d3q19_binary.cu", "d2q9_models.cu", "diagnostics.cu", "noalloc.cu"]
SOURCES_F32 = ["lbm_api.cu", "d3q19_binary.cu", "diagnostics.cu", "noalloc.cu"]
HOST_CXX = "/usr/bin/g++"   # the image's /opt/gcc toolchain links libstdc++ statically; do not use it


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build(lib=None):
    lib = lib or LIB
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, f32=False):
    lib, sources = (LIB_F32, SOURCES_F32) if f32 else (LIB, SOURCES)
    if not force and not needs_build(lib):
        return lib
    cmd = [nvcc_path(), "-std=c++17", "-O3", "-lineinfo"] + (["-DLBM_POP_FLOAT"] if f32 else []) + [
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-fmad=false", "-prec-div=true", "-prec-sqrt=true",
           "-ccbin", HOST_CXX,
           "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math",
           "-shared", "-I", os.path.join(ROOT, "include"), "-I", CSRC,
           "-o", lib] + [os.path.join(CSRC, s) for s in sources] + ["-ldl"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building " + os.path.basename(lib))
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, f32=True))
