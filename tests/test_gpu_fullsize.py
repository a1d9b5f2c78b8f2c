"""BASELINE.json's bench size (QuarterChannel scaled to 512 x 512 x 512, 134 M nodes, 82 GB of populations), where no
oracle run is affordable: size-independent properties instead.
  * z is periodic and every node runs the same arithmetic, so moving the initial bubble by s planes must move every
    field by s planes -- bit for bit;
  * the order parameter and the mass are conserved (bounce-back and mirror planes lose nothing; the body force adds
    momentum, not mass);
  * the in-place step on one population array must leave exactly the fields of the default two-array step."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
N, WIDTH, STEPS, SHIFT = 512, 114, 6, 37


def _run(api, shift, variant=0):
    cfg = api.microchannel_config("quarter", N, N, N)
    cfg.kernel_variant = variant
    ctx = api.Context(cfg)
    x = np.arange(N)[None, :] <= N - WIDTH - 1
    y = np.arange(N)[:, None] <= N - WIDTH - 1
    plane_gas = np.where(x & y, -1.0, 1.0)
    phase = np.ones((N, N, N))
    zlo, zhi = (N - 1) // 3, 2 * (N - 1) // 3
    for z in range(zlo, zhi + 1):
        phase[(z + shift) % N] = plane_gas
    ctx.set_field("phase", phase.ravel())
    del phase
    ctx.set_field("rho", np.ones(N ** 3))
    zero = np.zeros(N ** 3)
    for k in ("ux", "uy", "uz"):
        ctx.set_field(k, zero)
    del zero
    ctx.init()
    d0 = ctx.diagnostics()
    ctx.step(STEPS)
    d1 = ctx.diagnostics()
    out = {k: ctx.get_field(k).reshape(N, N * N) for k in ("phase", "uz", "rho")}
    ctx.close()
    return out, d0, d1


def test_z_translation_invariance_and_conservation_at_bench_size():
    import torch
    from latticeboltzmannmethod_b200 import api
    if torch.cuda.get_device_properties(0).total_memory < 120e9:
        pytest.skip("needs the 180 GB of a B200")
    a, d0, d1 = _run(api, 0)
    assert not d1["has_nan"]
    assert abs(d1["mass"] - d0["mass"]) <= 1e-9 * d0["mass"]
    assert abs(d1["phase_mass"] - d0["phase_mass"]) <= 1e-9 * d0["mass"]
    # keep a few planes around both interfaces and in the bulk, drop the rest before the second run
    zs = [0, 1, 169, 170, 171, 255, 340, 341, 342, 511]
    keep = {k: {z: v[z].copy() for z in zs} for k, v in a.items()}
    del a
    b, _, e1 = _run(api, SHIFT)
    assert e1["mass"] == pytest.approx(d1["mass"], rel=1e-12)
    for k in keep:
        for z in zs:
            assert np.array_equal(keep[k][z], b[k][(z + SHIFT) % N]), (k, z)
    # the same run in place on ONE population array (kernel_variant 2, one pass): every field of every node, bit for bit
    c, _, f1 = _run(api, SHIFT, variant=2)
    assert f1["mass"] == e1["mass"] and f1["phase_mass"] == e1["phase_mass"] and not f1["has_nan"]
    for k in b:
        assert np.array_equal(b[k], c[k]), ("in place", k)


@pytest.mark.parametrize("variant,omega,omega_bulk", [("gs", 1.2, 0.8), ("hermite", 0.9, 0.1)])
def test_mrt_8192_squared_equals_the_y_independent_oracle_run(variant, omega, omega_bulk):
    """BASELINE.json configs[2]: the shear wave of mrt.cpp on 8192 x 8192.  The problem does not depend on y, so every
    column of the big run must equal the oracle's 8192 x 3 run -- bit for bit."""
    import oracle_lib as ol
    from latticeboltzmannmethod_b200 import api
    n, steps = 8192, 25
    v = 0 if variant == "gs" else 1
    o = ol.OracleMRT(v, n, 3, omega, omega_bulk, 0.002 if v == 0 else 0.0)
    o.init_wave(0.001)
    o.step(steps)
    want = {k: a.reshape(n, 3)[:, 1] for k, a in o.fields().items()}
    o.close()
    ctx = api.Context(api.mrt_config(variant, n, n, omega, omega_bulk))
    rho, ux, uy = api.mrt_initial_fields(variant, n, 1)
    for name, a in (("rho", rho), ("ux", ux), ("uy", uy)):
        ctx.set_field(name, np.repeat(a, n))
    ctx.init()
    ctx.step(steps)
    for k in ("rho", "ux", "uy"):
        got = ctx.get_field(k).reshape(n, n)
        assert np.array_equal(got, np.broadcast_to(got[:, :1], got.shape)), k      # every row of y equal
        assert np.array_equal(got[:, 0], want[k]), k
    ctx.close()


def test_shanchen_8192_squared_equals_the_tiled_128_squared_oracle_run():
    """The droplet of shanchen.cpp tiled with its own period: each 128 x 128 tile of the big periodic run must equal
    the oracle's 128 x 128 run -- bit for bit."""
    import oracle_lib as ol
    from latticeboltzmannmethod_b200 import api
    n, t, steps = 8192, 128, 40
    o = ol.OracleShanChen(t, t)
    o.init_droplet()
    rho0 = o.fields()["rho"].copy()
    o.step(steps)
    want = o.fields()["rho"].reshape(t, t)
    o.close()
    ctx = api.Context(api.shanchen_config(n, n))
    ctx.set_field("rho", np.tile(rho0.reshape(t, t), (n // t, n // t)).ravel())
    ctx.init()
    ctx.step(steps)
    got = ctx.get_field("rho").reshape(n // t, t, n // t, t)
    assert np.array_equal(got, np.broadcast_to(want[None, :, None, :], got.shape))
    ctx.close()


def test_binary2d_tiled_in_x_equals_the_128_squared_oracle_run():
    """binary.cpp's box on a wetting wall, repeated 256 times along the periodic x direction (32768 x 128): each
    128-column tile must equal the oracle's 128 x 128 run -- bit for bit."""
    import oracle_lib as ol
    from latticeboltzmannmethod_b200 import api
    t, reps, steps = 128, 256, 60
    o = ol.OracleBinary2D(t, t)
    o.init_reference()
    o.step(steps)
    want = {k: a.reshape(t, t) for k, a in o.fields().items()}        # [iX][iY]
    o.close()
    ctx = api.Context(api.binary2d_config(t * reps, t))
    ctx.set_field("phase", np.tile(api.binary2d_initial_phase().reshape(t, t), (reps, 1)).ravel())
    ctx.init()
    ctx.step(steps)
    for k in ("phase", "rho", "ux", "uy"):
        got = ctx.get_field(k).reshape(reps, t, t)
        assert np.array_equal(got, np.broadcast_to(want[k][None], got.shape)), k
    ctx.close()
