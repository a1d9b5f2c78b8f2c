"""GPU parity of the TRT diffusion model (DiffusionCylinder/cylinder.cpp) through the C ABI:
against the golden vectors taken from the unmodified program, against the oracle on random
geometries with a velocity field, plus the geometry rules and the conserved-mass property."""
import numpy as np
import pytest

import golden_lib as gl
import oracle_lib as ol
from oracle_lib import ndiff

pytestmark = pytest.mark.gpu


def _api():
    from latticeboltzmannmethod_b200 import api
    return api


def _ctx(api, nx, ny, mask, rho, ux, uy, **kw):
    ctx = api.Context(api.trt_config(nx, ny, **kw))
    ctx.set_geometry(mask)
    ctx.set_field("rho", rho)
    ctx.set_field("ux", ux)
    ctx.set_field("uy", uy)
    ctx.init()
    return ctx


def test_trt_cylinder_golden():
    """The program as shipped, through its complete run of 64001 iterations (cylinder.cpp:21,353)."""
    api = _api()
    mask, rho, ux, uy = api.trt_cylinder_setup()
    assert ndiff(mask.astype(np.float64), gl.samples("trt_geometry", "mask")[0]) == 0
    m0, r0, _, _ = ol.trt_reference_setup()
    assert ndiff(mask, m0) == 0 and ndiff(rho, r0) == 0
    ctx = _ctx(api, 129, 129, mask, rho, ux, uy)
    done = 0
    for steps in (1, 2, 201, 4001, 64001):
        ctx.step(steps - done)
        done = steps
        tag = "trt_t%d" % steps
        nd, sha = gl.check(tag, "rho", ctx.get_field("rho"))
        assert nd == 0 and sha, tag
        nd, sha = gl.check(tag, "f", ctx.get_populations(0))
        assert nd == 0 and sha, tag
    ctx.close()


@pytest.mark.parametrize("steps", [1, 500])
def test_trt_cylinder_advected_golden(steps):
    api = _api()
    tag = "trt_adv_t%d" % steps
    m = gl.meta(tag)
    mask, _, _, _ = api.trt_cylinder_setup()
    rho, ux, uy = ol.trt_advected_fields(mask, m["conc_wall"])
    ctx = _ctx(api, 129, 129, mask, rho, ux, uy, omega=m["omega"], ce=m["ce"], conc_wall=m["conc_wall"])
    ctx.step(steps)
    assert gl.check(tag, "rho", ctx.get_field("rho")) == (0, True)
    assert gl.check(tag, "f", ctx.get_populations(0)) == (0, True)
    ctx.close()


@pytest.mark.parametrize("nx,ny", [(40, 33), (97, 61), (515, 259)])
def test_trt_random_geometry_vs_oracle(nx, ny):
    api = _api()
    mask, rho, ux, uy = ol.trt_random_case(nx, ny, seed=nx + ny)
    omega, ce, cw = 0.9, 0.3, 1.5
    ctx = _ctx(api, nx, ny, mask, rho, ux, uy, omega=omega, ce=ce, conc_wall=cw)
    o = ol.OracleTRT(nx, ny, mask, omega, ce, cw)
    o.init(rho, ux, uy)
    for s in (0, 1, 2, 60):
        ctx.step(s)
        o.step(s)
        assert ndiff(o.rho(), ctx.get_field("rho")) == 0, s
        assert ndiff(o.pops(), ctx.get_populations(0)) == 0, s
    # diagnostics: fluid nodes only
    d = ctx.diagnostics()
    f = o.pops()[mask == 1]
    assert not d["has_nan"]
    assert abs(d["mass"] - f.sum()) <= 1e-12 * abs(f.sum())
    cx = np.array([0, 1, 0, -1, 0, 1, -1, -1, 1.0])
    assert abs(d["momentum"][0] - (f * cx).sum()) <= 1e-10
    ctx.close()
    o.close()


def test_trt_population_roundtrip():
    api = _api()
    nx, ny = 64, 48
    mask, rho, ux, uy = ol.trt_random_case(nx, ny, seed=9)
    ctx = _ctx(api, nx, ny, mask, rho, ux, uy)
    ctx.step(7)
    f = ctx.get_populations(0)
    ctx2 = _ctx(api, nx, ny, mask, rho, ux, uy)
    ctx2.set_populations(0, f)
    ctx.step(5)
    ctx2.step(5)
    assert ndiff(ctx.get_populations(0), ctx2.get_populations(0)) == 0
    ctx.close()
    ctx2.close()


def test_trt_uniform_state_is_stationary_at_scale():
    """Size-independent property at a size the oracle does not run: with zero velocity and the wall held
    at the initial concentration c, f = w c is the stationary solution (the anti-bounce-back link returns
    -w c + 2 w c), so 50 steps may move the populations by rounding only and the mass stays c * #fluid."""
    api = _api()
    nx = ny = 2048
    mask = np.full((ny, nx), -1, np.int32)
    mask[1:-1, 1:-1] = 0                   # ring of boundary nodes between the solid edge and the fluid
    mask[2:-2, 2:-2] = 1
    c = 0.75
    rho = np.where(mask == 1, c, np.where(mask == 0, c, -1.0)).astype(np.float64)
    z = np.zeros(nx * ny)
    ctx = _ctx(api, nx, ny, mask.ravel(), rho.ravel(), z, z, conc_wall=c)
    f0 = ctx.get_populations(0)
    ctx.step(50)
    f1 = ctx.get_populations(0)
    fl = (mask == 1).ravel()
    assert np.max(np.abs(f1.reshape(-1, 9)[fl] - f0.reshape(-1, 9)[fl])) <= 1e-13
    d = ctx.diagnostics()
    assert not d["has_nan"] and abs(d["mass"] - c * fl.sum()) <= 1e-9 * c * fl.sum()
    ctx.close()


def test_trt_geometry_rules():
    api = _api()
    nx = ny = 16
    ctx = api.Context(api.trt_config(nx, ny))
    with pytest.raises(api.LbmError):      # init before the geometry
        ctx.init()
    m = np.full((ny, nx), -1, np.int32)
    m[0, 5] = 1                            # fluid on the edge of the array
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m)
    m[:] = -1
    m[5, 5] = 1                            # fluid touching solid
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m)
    m[4:7, 4:7] = 0
    m[5, 5] = 1
    ctx.set_geometry(m)                    # one fluid node inside a ring of boundary nodes
    m[8, 8] = 3
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m)
    ctx.close()
    other = api.Context(api.mrt_config("gs"))
    with pytest.raises(api.LbmError):
        other.set_geometry(np.zeros(other.local_nodes, np.int32))
    other.close()
