"""The C-ABI library loads and exports every entry point include/lbm_b200.h declares; configuration
defaults mirror the reference constants; without a GPU every compute call fails loudly (no fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, has_gpu

HEADER = os.path.join(ROOT, "include", "lbm_b200.h")


def _declared():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(lbm_[a-z0-9_]+)\s*\(", txt)))


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_library_exports_every_declared_symbol(precision):
    """Both builds (fp64, and the fp32-storage D3Q19 build) export the whole ABI."""
    from latticeboltzmannmethod_b200 import api
    L = C.CDLL(api.library_path(precision))
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "%s does not export %s" % (os.path.basename(api.library_path(precision)), n)
    assert set(api.EXPORTS) <= set(names)


def test_product_does_not_touch_the_oracle():
    """The product package must not import, link or dlopen anything under oracle/."""
    pkg = os.path.join(ROOT, "latticeboltzmannmethod_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.lower().replace("no oracle", ""), "%s mentions the oracle" % f
    import subprocess
    out = subprocess.run(["ldd", os.path.join(pkg, "liblbm_b200.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out


def test_config_defaults_match_reference_constants():
    from latticeboltzmannmethod_b200 import api
    c = api.default_config("binary_d3q19")   # QuarterChannel/main.cpp:18-62
    assert (c.nx, c.ny, c.nz) == (27, 27, 52)
    assert (c.aconst, c.kconst, c.gammaconst, c.rho_wall) == (0.04, 0.04, 1.0, 0.5)
    assert (c.tau_phi, c.tau_liq, c.tau_gas) == (2.0, 2.5, 0.7)
    assert list(c.force) == [0.0, 0.0, 0.00003]
    c = api.microchannel_config("full")      # FullChannel/main.cpp:19-70
    assert (c.nx, c.ny, c.nz, c.tau_phi) == (92, 52, 1500, 1.0)
    c = api.default_config("shanchen_d2q9")  # shanchen.cpp:11-27
    assert (c.nx, c.ny, c.tau, c.g) == (128, 128, 1.0, -5.0)
    c = api.default_config("binary_d2q9")    # binary.cpp:12-41
    assert (c.wall_gradient, c.tau_gas, c.tau_liq) == (-0.15, 1.0, 1.0)
    c = api.default_config("mrt_hermite_d2q9")
    assert (c.omega, c.omega_bulk) == (0.9, 0.1)


@pytest.mark.skipif(has_gpu(), reason="only meaningful on a box without a GPU")
def test_no_gpu_fails_loudly():
    from latticeboltzmannmethod_b200 import api
    with pytest.raises(api.LbmError) as e:
        api.Context(api.default_config("binary_d3q19"))
    assert "LBM_ECUDA" in str(e.value)
