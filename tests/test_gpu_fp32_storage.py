"""liblbm_b200_f32.so: the D3Q19 model with the populations STORED as float (arithmetic, phase field and the
macroscopic fields stay double).  This is not the parity build; the test states how far it is from the fp64
build and checks that it is a consistent model of its own: conservation, slab invariance, ABI behaviour."""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
FIELDS = ("rho", "phase", "ux", "uy", "uz")


def _api():
    from latticeboltzmannmethod_b200 import api
    return api


def _run(api, precision, nx, ny, nz, width, steps, params=None):
    cfg = api.microchannel_config("quarter", nx, ny, nz)
    if params:
        (cfg.aconst, cfg.kconst, cfg.gammaconst, cfg.wall_gradient, _pw, cfg.rho_wall, cfg.tau_phi, cfg.tau_liq, cfg.tau_gas) = params[:9]
        cfg.force[0], cfg.force[1], cfg.force[2] = params[9:12]
    ctx = api.Context(cfg, precision)
    for name, a in zip(FIELDS, api.microchannel_initial_fields(nx, ny, nz, width)):
        ctx.set_field(name, a)
    ctx.init()
    out = []
    done = 0
    for s in steps:
        ctx.step(s - done)
        done = s
        out.append(({k: ctx.get_field(k) for k in FIELDS}, ctx.diagnostics()))
    ctx.close()
    return out


# Stated tolerance of the fp32-storage mode against the fp64 build, QuarterChannel 27x27x52 (reference parameters),
# field-normalised max norm ||a-b||_inf / ||b||_inf: float rounding of the stored state is 6e-8 per value and
# step; it accumulates like a random walk, not linearly.
TOL_AFTER = {1: 2e-6, 100: 2e-5, 1001: 2e-4}


def test_fp32_storage_against_fp64_build():
    api = _api()
    steps = sorted(TOL_AFTER)
    a = _run(api, "f32", 27, 27, 52, 6, steps)
    b = _run(api, "f64", 27, 27, 52, 6, steps)
    for s, (fa, da), (fb, db) in zip(steps, a, b):
        assert not da["has_nan"]
        umax = max(np.max(np.abs(fb[k])) for k in ("ux", "uy", "uz"))
        for k in FIELDS:
            scale = umax if k.startswith("u") else np.max(np.abs(fb[k]))
            err = np.max(np.abs(fa[k] - fb[k])) / scale
            assert err <= TOL_AFTER[s], (s, k, err)
        # mass and order parameter are conserved by both builds; float storage keeps them to ~1e-7 relative
        assert abs(da["mass"] - db["mass"]) <= 2e-6 * abs(db["mass"])
        assert abs(da["phase_mass"] - db["phase_mass"]) <= 2e-5 * abs(db["mass"])


def test_fp32_storage_roundtrip_and_limits():
    api = _api()
    cfg = api.microchannel_config("quarter", 12, 10, 8)
    ctx = api.Context(cfg, "f32")
    for name, a in zip(FIELDS, api.microchannel_initial_fields(12, 10, 8, 3)):
        ctx.set_field(name, a)
    ctx.init()
    ctx.step(3)
    f, g = ctx.get_populations(0), ctx.get_populations(1)
    assert np.array_equal(f, f.astype(np.float32).astype(np.float64))       # what comes back is what float holds
    other = api.Context(cfg, "f32")
    for name, a in zip(FIELDS, api.microchannel_initial_fields(12, 10, 8, 3)):
        other.set_field(name, a)
    other.init()
    other.set_populations(0, f)
    other.set_populations(1, g)
    ctx.step(4)
    other.step(4)
    assert np.array_equal(ctx.get_populations(0), other.get_populations(0))  # exact restart within the mode
    assert np.array_equal(ctx.get_field("phase"), other.get_field("phase"))
    ctx.close()
    other.close()
    cfg.kernel_variant = 91                                                   # no such kernel
    with pytest.raises(api.LbmError):
        api.Context(cfg, "f32")
    with pytest.raises(api.LbmError):                                         # D2Q9 models live in the fp64 library
        api.Context(api.shanchen_config(), "f32")


@pytest.mark.parametrize("variant", [2, 6, 5, 0])
def test_fp32_storage_kernel_variants_agree_bit_for_bit(variant):
    """Within the fp32-storage build every kernel variant rounds the same values at the same stores, so the in-place
    forms on one population array (2: one pass, 6: two passes; 4-byte asynchronous copies there) and the two-array
    forms (5, 0) must leave exactly the state of the plain two-array kernel (1)."""
    api = _api()
    nx, ny, nz = 70, 37, 12
    runs = []
    for v in (variant, 1):
        cfg = api.microchannel_config("quarter", nx, ny, nz)
        cfg.kernel_variant = v
        ctx = api.Context(cfg, "f32")
        for name, a in zip(FIELDS, api.microchannel_initial_fields(nx, ny, nz, 5)):
            ctx.set_field(name, a)
        ctx.init()
        out = []
        for s in (1, 2, 6):
            ctx.step(s)
            out.append([ctx.get_field(k) for k in FIELDS] + [ctx.get_populations(0), ctx.get_populations(1)])
        ctx.close()
        runs.append(out)
    for a, b in zip(*runs):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
