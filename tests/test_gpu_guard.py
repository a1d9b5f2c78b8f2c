"""Out-of-bounds writes, checked without compute-sanitizer (closed on this GPU pool): with LBM_GUARD=1 at lbm_create every
device buffer of a context lies between two 1 MiB guard bands holding a byte pattern, and lbm_guard_check counts the
guard bytes a kernel has overwritten.  Small grids on purpose: a push that leaves its buffer by a row or a plane lands
inside the band.  Every model, every D3Q19 kernel variant, the z-slab phases included."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def guarded(monkeypatch):
    monkeypatch.setenv("LBM_GUARD", "1")
    from latticeboltzmannmethod_b200 import api
    return api


@pytest.mark.parametrize("variant", [0, 5, 1, 2, 6])
@pytest.mark.parametrize("shape", [(35, 18, 9), (4, 4, 2), (140, 37, 21)])
@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_d3q19_kernels_stay_inside_their_buffers(guarded, shape, variant, precision):
    api = guarded
    nx, ny, nz = shape
    cfg = api.microchannel_config("quarter", nx, ny, nz)
    cfg.kernel_variant = variant
    cfg.wall_gradient = 0.02
    cfg.force[0], cfg.force[1] = 1e-5, -2e-5
    ctx = api.Context(cfg, precision)
    for name, a in zip(("rho", "phase", "ux", "uy", "uz"), api.microchannel_initial_fields(nx, ny, nz, max(1, nx // 5))):
        ctx.set_field(name, a)
    ctx.init()
    ctx.step(7)
    ctx.step(1)
    f = ctx.get_populations(0)
    ctx.set_populations(0, f)
    ctx.step(2)
    ctx.diagnostics()
    assert ctx.guard_check() == 0
    assert not np.isnan(ctx.get_field("phase")).any()
    ctx.close()


def test_guard_check_sees_a_stray_write(guarded):
    """The check itself: a deliberate write one element before a halo message's buffer must be counted."""
    import torch
    from latticeboltzmannmethod_b200.slab import cuda_tensor_of
    api = guarded
    ctx = api.Context(api.microchannel_config("quarter", 12, 10, 6))
    assert ctx.guard_check() == 0
    ptr, count, eb = ctx.slab_halo("R_F_FROM_BELOW")          # starts 1 plane into the f back buffer
    plane = 19 * 128                                            # 19 directions x plane pitch (12*10 rounded up to 128)
    cuda_tensor_of(ptr - (plane + 1) * 8, 1, 8).fill_(1.0)     # one double before the start of the allocation
    torch.cuda.synchronize()
    assert ctx.guard_check() == 8
    ctx.close()


@pytest.mark.parametrize("model", ["shanchen", "binary", "mrt_gs", "mrt_hermite", "trt", "shape"])
def test_d2q9_kernels_stay_inside_their_buffers(guarded, model):
    api = guarded
    n = 96
    mask = None
    if model == "shanchen":
        cfg, fields = api.shanchen_config(n, n), dict(rho=np.tile(api.shanchen_initial_rho(32, 32, radius=6).reshape(32, 32), (3, 3)).ravel())
    elif model == "binary":
        cfg, fields = api.binary2d_config(n, n), dict(phase=api.binary2d_initial_phase(n, n).ravel())
    elif model.startswith("mrt"):
        v = model.split("_")[1]
        cfg = api.mrt_config(v, n, n, 1.2, 0.8)
        fields = dict(rho=np.ones(n * n), ux=np.repeat(1e-3 * np.cos(np.arange(n) / n * 2 * np.pi), n), uy=np.zeros(n * n))
    elif model == "trt":
        cfg = api.trt_config(n, n)
        mask, rho, ux, uy = api.trt_cylinder_setup(n, n, radius=20)
        fields = dict(rho=rho, ux=ux, uy=uy)
    else:
        cfg = api.shape_config(n, n)
        mask, rho, phi, ux, uy = api.shape_setup(n, n, radius=16, radius_droplet=5)
        fields = dict(rho=rho, phase=phi, ux=ux, uy=uy)
    for variant in ((0, 1) if model in ("shanchen", "binary", "mrt_gs", "mrt_hermite") else (0,)):
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        if mask is not None:
            ctx.set_geometry(mask)
        for k, a in fields.items():
            ctx.set_field(k, a)
        ctx.init()
        ctx.step(5)
        ctx.diagnostics()
        assert ctx.guard_check() == 0, (model, variant)
        ctx.close()
