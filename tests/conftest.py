import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("%s failed:\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


@pytest.fixture(scope="session", autouse=True)
def _checkers_built():
    """Builds the CPU checkers if they are missing (the product library is built by __graft_entry__.build())."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "liblbm_oracle.so")):
        _run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    yield


@pytest.fixture(scope="session")
def emu_libs():
    """CPU emulation of the plain CUDA kernels (tests/emu): the per-thread functions run from host loops."""
    emu = os.path.join(ROOT, "tests", "emu")
    csrc = os.path.join(ROOT, "latticeboltzmannmethod_b200", "csrc")
    out = {}
    for name in ("d3q19", "d2q9"):
        so = os.path.join(emu, "libemu_%s.so" % name)
        src = os.path.join(emu, "emu_%s.cpp" % name)
        deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".h")]
        if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
            _run(["/usr/bin/g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-I", csrc,
                  "-I", os.path.join(ROOT, "include"), "-o", so, src])
        out[name] = so
    return out


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
