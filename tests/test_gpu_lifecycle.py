"""Life-cycle behaviour of a context through the C ABI: re-initialising after steps equals a fresh context,
an external stream can be installed, profiling can be toggled, odd/even step splits give the same state."""
import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import ndiff

pytestmark = pytest.mark.gpu


def _api():
    from latticeboltzmannmethod_b200 import api
    return api


def _models(api):
    """(name, make_context, set_inputs) for every model; inputs are the reference initial conditions."""
    def sc():
        c = api.Context(api.shanchen_config(64, 48))
        # init reads rho AND the velocity fields (zero in a fresh context, the last step's output afterwards)
        return c, dict(rho=api.shanchen_initial_rho(64, 48, radius=10), ux=np.zeros(64 * 48), uy=np.zeros(64 * 48)), None

    def b2():
        c = api.Context(api.binary2d_config(48, 40))
        return c, dict(phase=np.random.default_rng(0).uniform(-1, 1, 48 * 40)), None

    def mrt():
        c = api.Context(api.mrt_config("gs", 40, 12, 1.2, 0.8))
        rho, ux, uy = api.mrt_initial_fields("gs", 40, 12)
        return c, dict(rho=rho, ux=ux, uy=uy), None

    def trt():
        mask, rho, ux, uy = ol.trt_random_case(60, 44, 5)
        return api.Context(api.trt_config(60, 44)), dict(rho=rho, ux=ux, uy=uy), mask

    def shape():
        mask, rho, phi, ux, uy = api.shape_setup(90, 80, 14, 4)
        return api.Context(api.shape_config(90, 80)), dict(rho=rho, phase=phi, ux=ux, uy=uy), mask

    def d3():
        c = api.Context(api.microchannel_config("quarter", 14, 11, 12))
        return c, dict(zip(("rho", "phase", "ux", "uy", "uz"), api.microchannel_initial_fields(14, 11, 12, 3))), None

    return dict(shanchen=sc, binary2d=b2, mrt=mrt, trt=trt, shape=shape, d3q19=d3)


def _start(ctx, fields, mask):
    if mask is not None:
        ctx.set_geometry(mask)
    for k, a in fields.items():
        ctx.set_field(k, a)
    ctx.init()


def _state(ctx):
    return [ctx.get_populations(s) for s in range(2 if ctx.q == 19 or ctx.cfg.model in (1, 6) else 1)]


@pytest.mark.parametrize("name", ["shanchen", "binary2d", "mrt", "trt", "shape", "d3q19"])
def test_reinit_after_steps_equals_fresh_context(name):
    api = _api()
    make = _models(api)[name]
    a, fields, mask = make()
    _start(a, fields, mask)
    a.step(7)                      # odd: leaves the other buffer current
    _start(a, fields, mask)        # same inputs again
    a.step(5)
    b, fields, mask = make()
    _start(b, fields, mask)
    b.step(5)
    for x, y in zip(_state(a), _state(b)):
        assert ndiff(x, y) == 0
    # 5 steps as 2 + 3 on a third context
    c, fields, mask = make()
    _start(c, fields, mask)
    c.step(2)
    c.step(3)
    for x, y in zip(_state(c), _state(b)):
        assert ndiff(x, y) == 0
    for k in fields:
        assert ndiff(c.get_field(k), b.get_field(k)) == 0, k
    for ctx in (a, b, c):
        ctx.close()


def test_external_stream_and_profile_toggle():
    import torch
    api = _api()
    a, fields, mask = _models(api)["d3q19"]()
    b, _, _ = _models(api)["d3q19"]()
    s = torch.cuda.Stream()
    a.set_stream(s.cuda_stream)
    assert a.stream == s.cuda_stream
    _start(a, fields, mask)
    _start(b, fields, mask)
    a.profile(True)
    a.step(4)
    rep = a.profile_report()
    assert rep and all(v["launches"] > 0 and v["ms"] > 0 for v in rep.values())
    a.profile(False)
    n0 = a.launch_count
    a.step(2)
    assert a.launch_count > n0 and not a.profile_report()
    b.step(6)
    s.synchronize()
    for x, y in zip(_state(a), _state(b)):
        assert ndiff(x, y) == 0
    a.close()
    b.close()
