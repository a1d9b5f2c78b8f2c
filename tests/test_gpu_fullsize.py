"""BASELINE.json's bench size (QuarterChannel scaled to 512 x 512 x 512, 134 M nodes, 82 GB of populations), where no
oracle run is affordable: size-independent properties instead.
  * z is periodic and every node runs the same arithmetic, so moving the initial bubble by s planes must move every
    field by s planes -- bit for bit;
  * the order parameter and the mass are conserved (bounce-back and mirror planes lose nothing; the body force adds
    momentum, not mass)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
N, WIDTH, STEPS, SHIFT = 512, 114, 6, 37


def _run(api, shift):
    cfg = api.microchannel_config("quarter", N, N, N)
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
