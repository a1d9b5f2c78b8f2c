"""The z-slab protocol of the C ABI (lbm_slab_halo / lbm_slab_collide / lbm_slab_finish ...) on real kernels with
ONE GPU: P contexts on device 0, each owning slab r of P, the four halo messages per step moved between them by
device-to-device copies issued from the test (the "caller's own transport" of INTEGRATION.md).  The ring of slabs
must equal the single-slab run bit for bit (the reference is 1-vs-N bit-identical), for even and
uneven splits (solver.hpp:25-30 gives the remainder to the last slab)."""
import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import ndiff

pytestmark = pytest.mark.gpu
FIELDS = ("rho", "phase", "ux", "uy", "uz")


def _ring(api, torch, nx, ny, nz, world, params, flds):
    from latticeboltzmannmethod_b200.slab import PHI, STATE, cuda_tensor_of
    ctxs = []
    for r in range(world):
        cfg = api.microchannel_config("quarter", nx, ny, nz, slab_rank=r, slab_count=world)
        cfg.aconst, cfg.kconst, cfg.gammaconst, cfg.wall_gradient = params[0], params[1], params[2], params[3]
        cfg.rho_wall, cfg.tau_phi, cfg.tau_liq, cfg.tau_gas = params[5], params[6], params[7], params[8]
        cfg.force[0], cfg.force[1], cfg.force[2] = params[9], params[10], params[11]
        c = api.Context(cfg)
        z0, nzl = c.slab_extent()
        for name, a in zip(FIELDS, flds):
            c.set_field(name, a.reshape(nz, ny * nx)[z0:z0 + nzl].ravel())
        ctxs.append((c, z0, nzl))

    def exchange(pairs):
        for c, _, _ in ctxs:
            c.sync()
        half = len(pairs) // 2
        for r, (c, _, _) in enumerate(ctxs):
            up, down = ctxs[(r + 1) % world][0], ctxs[(r - 1 + world) % world][0]
            for i, (s, rcv) in enumerate(pairs):
                dst = up if i < half else down        # sends [up..., down...] land in "from below" / "from above"
                cuda_tensor_of(*dst.slab_halo(rcv)).copy_(cuda_tensor_of(*c.slab_halo(s)))
        torch.cuda.synchronize()

    def init():
        for c, _, _ in ctxs:
            c.slab_init_fields()
        exchange(PHI)
        for c, _, _ in ctxs:
            c.slab_init_pops()
        exchange(STATE)
        for c, _, _ in ctxs:
            c.slab_finish()
            c.slab_after_init()

    def step(n):
        for t in range(n):
            for c, _, _ in ctxs:
                c.slab_collide(t == n - 1)
            exchange(STATE)
            for c, _, _ in ctxs:
                c.slab_finish()

    def fields():
        return {k: np.concatenate([c.get_field(k) for c, _, _ in ctxs]) for k in FIELDS}

    def close():
        for c, _, _ in ctxs:
            c.close()

    return init, step, fields, close


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("shape,params", [((27, 27, 52), ol.PARAMS_QUARTER),
                                          ((12, 10, 23), [0.04, 0.04, 1.0, 0.05, 0.0, 0.5, 1.0, 2.5, 0.7, 1e-5, -2e-5, 3e-5])])
def test_slab_ring_on_one_gpu_equals_single_slab(world, shape, params):
    import torch
    from latticeboltzmannmethod_b200 import api
    nx, ny, nz = shape
    rng = np.random.default_rng(3)
    base = ol.bubble_fields(nx, ny, nz, max(1, nx // 4))
    flds = [base[0] * (1 + 0.02 * rng.standard_normal(base[0].size)), np.clip(base[1] + 0.2 * rng.standard_normal(base[1].size), -1.2, 1.2),
            0.01 * rng.standard_normal(base[0].size), 0.01 * rng.standard_normal(base[0].size), 0.01 * rng.standard_normal(base[0].size)]
    init, step, fields, close = _ring(api, torch, nx, ny, nz, world, params, flds)
    o = ol.Oracle3D(nx, ny, nz, params)
    o.set_fields(*flds)
    o.init_single()
    init()
    for n in (0, 1, 2, 25):
        step(n)
        o.step_single(n)
        a, b = o.fields(), fields()
        for k in FIELDS:
            assert ndiff(a[k], b[k]) == 0, (k, n, world)
    close()
    o.close()
