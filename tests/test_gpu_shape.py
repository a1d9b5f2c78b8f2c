"""GPU parity of the curved-solid binary-liquid model (CurvedSolid/binary_shape.cpp) through the C ABI:
golden vectors from the unmodified program (complete 20001-iteration run included), the oracle on random
geometries, the geometry rules, and conservation properties at a size the oracle does not run."""
import numpy as np
import pytest

import golden_lib as gl
import oracle_lib as ol
from oracle_lib import ndiff

pytestmark = pytest.mark.gpu
FIELDS = ("phase", "rho", "ux", "uy")


def _api():
    from latticeboltzmannmethod_b200 import api
    return api


def _ctx(api, nx, ny, mask, rho, phi, ux, uy, params=ol.PARAMS_SHAPE, variant=0):
    cfg = api.shape_config(nx, ny, wall_gradient=params[3], aconst=params[0], kconst=params[1], gammaconst=params[2],
                           tau_gas=params[4], tau_liq=params[5], tau_phi=1.0 / params[6])
    cfg.force[0], cfg.force[1] = params[7], params[8]
    cfg.kernel_variant = variant
    ctx = api.Context(cfg)
    ctx.set_geometry(mask)
    for name, a in (("rho", rho), ("phase", phi), ("ux", ux), ("uy", uy)):
        ctx.set_field(name, a)
    ctx.init()
    return ctx


def _check_golden(ctx, tag):
    for k in FIELDS:
        assert gl.check(tag, k, ctx.get_field(k))[0] == 0, (tag, k)
    assert gl.check(tag, "f", ctx.get_populations(0))[0] == 0, tag
    assert gl.check(tag, "g", ctx.get_populations(1))[0] == 0, tag


def test_shape_golden_complete_run():
    api = _api()
    mask, rho, phi, ux, uy = api.shape_setup()
    assert ndiff(mask.astype(np.float64), gl.samples("shape_geometry", "mask")[0]) == 0
    ctx = _ctx(api, 301, 301, mask, rho, phi, ux, uy)
    done = 0
    for steps in (1, 2, 201, 2001, 20001):
        ctx.step(steps - done)
        done = steps
        tag = "shape_t%d" % steps
        _check_golden(ctx, tag)
        # the "Mass is" line of binary_shape.cpp:639-646, summed in the program's order
        mass = 0.0
        for v in ctx.get_field("phase")[mask == 1]:
            mass += v
        assert mass == gl.meta(tag)["mass"]
    ctx.close()


@pytest.mark.parametrize("steps", [1, 300])
def test_shape_perturbed_golden(steps):
    api = _api()
    tag = "shape_pert_t%d" % steps
    m = gl.meta(tag)
    mask, rho, phi, _, _ = api.shape_setup()
    rho2, phi2, ux2, uy2 = ol.shape_perturbed_fields(mask, rho, phi)
    ctx = _ctx(api, 301, 301, mask, rho2, phi2, ux2, uy2, m["params"])
    ctx.step(steps)
    _check_golden(ctx, tag)
    ctx.close()


@pytest.mark.parametrize("nx,ny", [(64, 48), (131, 77), (517, 263)])
def test_shape_random_geometry_vs_oracle(nx, ny):
    api = _api()
    mask, rho, phi, ux, uy = ol.shape_random_case(nx, ny, seed=nx)
    params = [0.04, 0.04, 1.0, -0.15, 0.7, 2.5, 0.5, 0.0, 3e-5]
    ctx = _ctx(api, nx, ny, mask, rho, phi, ux, uy, params)
    o = ol.OracleShape(nx, ny, mask, params)
    o.init(rho, phi, ux, uy)
    for s in (0, 1, 2, 60):
        ctx.step(s)
        o.step(s)
        a = o.fields()
        for k in FIELDS:
            assert ndiff(a[k], ctx.get_field(k)) == 0, (k, s)
        fo, go = o.pops()
        assert ndiff(fo, ctx.get_populations(0)) == 0 and ndiff(go, ctx.get_populations(1)) == 0, s
    d = ctx.diagnostics()
    fl = mask == 1
    assert not d["has_nan"]
    assert abs(d["mass"] - fo[fl].sum()) <= 1e-12 * fo[fl].sum()
    assert abs(d["phase_mass"] - go[fl].sum()) <= 1e-10
    ctx.close()
    o.close()


def test_shape_conservation_at_scale():
    """Size-independent properties on 2048x2048 (4.2 M nodes, 16 discs): half-way bounce-back and periodic
    streaming lose nothing, so total mass and total order parameter over the fluid nodes are conserved
    to rounding over 200 steps, and no NaN appears."""
    api = _api()
    nx = ny = 2048
    iy, ix = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    solid = np.zeros((ny, nx), bool)
    for cy in range(256, ny, 512):
        for cx in range(256, nx, 512):
            solid |= (ix - cx) ** 2 + (iy - cy) ** 2 < 90 * 90
    mask = np.where(solid, -1, 1).astype(np.int32)
    near = np.zeros(mask.shape, bool)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dx or dy:
                near |= np.roll(np.roll(mask, -dy, 0), -dx, 1) == 1
    mask[(mask == -1) & near] = 0
    fl = mask == 1
    phi = np.where(fl, np.where((ix // 64 + iy // 64) % 2 == 0, 1.0, -1.0), 1.5).astype(np.float64)
    rho = np.where(fl, 1.0, 0.5)
    z = np.zeros(nx * ny)
    ctx = _ctx(api, nx, ny, mask.ravel(), rho.ravel(), phi.ravel(), z, z)
    d0 = ctx.diagnostics()
    ctx.step(200)
    d1 = ctx.diagnostics()
    assert not d1["has_nan"]
    assert abs(d0["mass"] - fl.sum()) <= 1e-6
    assert abs(d1["mass"] - d0["mass"]) <= 1e-9 * d0["mass"]
    assert abs(d1["phase_mass"] - d0["phase_mass"]) <= 1e-9 * fl.sum()
    ctx.close()


def test_shape_geometry_rules():
    api = _api()
    nx = ny = 24
    ctx = api.Context(api.shape_config(nx, ny))
    with pytest.raises(api.LbmError):                  # init before the geometry
        ctx.init()
    m = np.ones((ny, nx), np.int32)
    m[10:13, 10:13] = -1                               # fluid touching solid
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m)
    m[10:13, 10:13] = 0
    m[11, 11] = -1
    ctx.set_geometry(m)                                # 3x3 obstacle: ring of bounce-back nodes around one solid node
    m2 = np.ones((ny, nx), np.int32)
    m2[0, 5] = 0                                       # bounce-back node on the edge of the array
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m2)
    m3 = np.ones((ny, nx), np.int32)
    m3[5, 5] = 0                                       # fluid on all sides: the axis directions sum to zero, no wall normal
    with pytest.raises(api.LbmError):
        ctx.set_geometry(m3)
    ctx.close()
