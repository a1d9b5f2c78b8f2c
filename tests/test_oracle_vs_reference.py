"""The C restatement against the reference ITSELF (oracle/_ref, compiled from /root/reference by
`make -C oracle ref`) on inputs the golden file does not hold: random fields, odd sizes, non-zero wall
gradient and all three force components.  Skipped where the reference build is absent."""
import numpy as np
import pytest

import oracle_lib as ol

need_ref3d = pytest.mark.skipif(not ol.ref_available("quarter"), reason="oracle/_ref not built")


def _random_fields(nx, ny, nz, seed):
    rng = np.random.default_rng(seed)
    n = nx * ny * nz
    _, ph, _, _, _ = ol.bubble_fields(nx, ny, nz, max(1, nx // 4))
    return (1 + 0.05 * rng.standard_normal(n), np.clip(ph + 0.3 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n))


@need_ref3d
@pytest.mark.parametrize("which,shape,params", [
    ("quarter", (9, 7, 5), ol.PARAMS_QUARTER),
    ("full", (12, 10, 6), [0.04, 0.04, 1.0, 0.05, 0.0, 0.5, 1.0, 2.5, 0.7, 1e-5, -2e-5, 3e-5]),
    ("quarter", (4, 4, 2), [0.03, 0.05, 0.9, -0.1, 0.0, 0.6, 1.5, 2.0, 0.8, 0.0, 1e-5, 0.0]),
])
def test_d3q19_random(which, shape, params):
    nx, ny, nz = shape
    flds = _random_fields(nx, ny, nz, 7)
    r = ol.Ref3D(nx, ny, nz, params, which)
    r.set_fields(*flds)
    r.init()
    o = ol.Oracle3D(nx, ny, nz, params)
    o.set_fields(*flds)
    o.init_single()
    nw = ol.nonwall_mask_3d(nx, ny, nz)
    for n in (0, 1, 2, 15):
        r.step(n)
        o.step_single(n)
        a, b = r.fields(), o.fields()
        for k in a:
            assert ol.ndiff(a[k], b[k]) == 0, (k, n)
        (fr, gr), (fo, go) = r.pops(), o.pops()
        assert ol.ndiff(fr[nw], fo[nw]) == 0 and ol.ndiff(gr[nw], go[nw]) == 0
    r.close()
    o.close()


@need_ref3d
def test_d3q19_quarter_reference_case():
    nx, ny, nz = 27, 27, 52
    r = ol.Ref3D(nx, ny, nz, ol.PARAMS_QUARTER)
    r.set_bubble(6)
    r.init()
    o = ol.Oracle3D(nx, ny, nz, ol.PARAMS_QUARTER)
    o.set_fields(*ol.bubble_fields(nx, ny, nz, 6))
    o.init_single()
    r.step(30)
    o.step_single(30)
    a, b = r.fields(), o.fields()
    for k in a:
        assert ol.ndiff(a[k], b[k]) == 0, k
    r.close()
    o.close()


@pytest.mark.skipif(not ol.ref_available("binary"), reason="oracle/_ref not built")
def test_binary2d():
    r = ol.ref_binary_run(57)
    o = ol.OracleBinary2D(r["nx"], r["ny"])
    o.init_reference()
    o.step(58)
    fl = o.fields()
    for k in fl:
        assert ol.ndiff(r[k], fl[k]) == 0, k
    fo, go = o.pops()
    assert ol.ndiff(r["f"], fo) == 0 and ol.ndiff(r["g"], go) == 0
    o.close()


@pytest.mark.skipif(not ol.ref_available("shanchen"), reason="oracle/_ref not built")
@pytest.mark.parametrize("shared", [True, False])
def test_shanchen(shared):
    r = ol.ref_shanchen_run(120, shared_exp=shared)
    o = ol.OracleShanChen(r["nx"], r["ny"], use_libm_exp=not shared)
    o.init_droplet()
    o.step(121)
    fl = o.fields()
    for k in fl:
        assert ol.ndiff(r[k], fl[k]) == 0, k
    assert ol.ndiff(r["f"], o.pops()) == 0
    o.close()


@pytest.mark.skipif(not ol.ref_available("cylinder"), reason="oracle/_ref not built")
@pytest.mark.parametrize("omega,ce,cw", [(0.0, 0.0, 0.0), (0.8, 0.3, 2.0)])
def test_trt_cylinder(omega, ce, cw):
    mask, rho0, _, _ = ol.trt_reference_setup(conc_wall=cw if omega > 0 else 1.0)
    rho, ux, uy = ol.trt_advected_fields(mask, cw if omega > 0 else 1.0, speed=0.08)
    args = (omega, ce, cw) if omega > 0 else ()
    for steps in (0, 1, 3, 120):
        o = ol.OracleTRT(129, 129, mask, *args)
        o.init(rho, ux, uy)
        o.step(steps)
        r = ol.ref_cylinder_run(steps, omega, ce, cw, rho, ux, uy)
        assert ol.ndiff(r["geometry"], mask) == 0
        assert ol.ndiff(r["rho"], o.rho()) == 0 and ol.ndiff(r["f"], o.pops()) == 0, steps
        o.close()


@pytest.mark.skipif(not ol.ref_available("shape"), reason="oracle/_ref not built")
@pytest.mark.parametrize("params,rd", [(None, 0), ([0.05, 0.03, 1.2, -0.2, 0.6, 2.0, 0.8, 0.0, 0.0], 14)])
def test_shape_curved_solid(params, rd):
    mask, rho, phi, ux, uy = ol.shape_reference_setup(radius_droplet=rd or 10)
    rho, phi, ux, uy = ol.shape_perturbed_fields(mask, rho, phi, amp=0.08, speed=0.03)
    for steps in (0, 1, 3, 80):
        o = ol.OracleShape(301, 301, mask, params or ol.PARAMS_SHAPE)
        o.init(rho, phi, ux, uy)
        o.step(steps)
        r = ol.ref_shape_run(steps, params, rd, phi, rho, ux, uy)
        assert ol.ndiff(r["geometry"], mask) == 0
        a = o.fields()
        for k in a:
            assert ol.ndiff(r[k], a[k]) == 0, (k, steps)
        f, g = o.pops()
        assert ol.ndiff(r["f"], f) == 0 and ol.ndiff(r["g"], g) == 0, steps
        o.close()
