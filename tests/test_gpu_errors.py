"""Error behaviour of the C ABI on a GPU box: every misuse returns a status and a message (lbm_last_error),
nothing aborts, and the context stays usable afterwards."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EINVAL, ESTATE = 1, 6


def _api():
    from latticeboltzmannmethod_b200 import api
    return api


def _raises(api, status, fn, *a):
    with pytest.raises(api.LbmError) as e:
        fn(*a)
    assert e.value.status == status, str(e.value)
    assert len(str(e.value)) > len("LBM_EINVAL: ")        # there is a message


def test_create_rejects_bad_configs():
    api = _api()
    cfg = api.shanchen_config()
    cfg.model = 99
    _raises(api, EINVAL, api.Context, cfg)
    cfg = api.shanchen_config(2, 128)
    _raises(api, EINVAL, api.Context, cfg)
    cfg = api.shanchen_config()
    cfg.device = 4096
    _raises(api, EINVAL, api.Context, cfg)
    cfg = api.microchannel_config("quarter", 3, 27, 52)
    _raises(api, EINVAL, api.Context, cfg)
    cfg = api.microchannel_config("quarter", 27, 27, 52, slab_rank=2, slab_count=2)
    _raises(api, EINVAL, api.Context, cfg)
    cfg = api.microchannel_config("quarter", 27, 27, 7, slab_rank=0, slab_count=7)     # one plane per slab
    _raises(api, EINVAL, api.Context, cfg)


def test_calls_in_the_wrong_order_or_with_wrong_sizes():
    api = _api()
    ctx = api.Context(api.mrt_config("gs"))
    L, h, n = ctx.L, ctx.h, ctx.local_nodes
    _raises(api, ESTATE, ctx.step, 1)                                                   # before init
    buf = np.zeros(n + 1)
    assert L.lbm_set_field(h, 0, buf.ctypes.data, n + 1) == EINVAL
    assert b"lbm_local_nodes" in L.lbm_last_error(h)
    assert L.lbm_set_field(h, 0, None, n) == EINVAL
    assert L.lbm_get_field(h, 0, buf.ctypes.data, n - 1) == EINVAL
    assert L.lbm_set_field(h, 1, buf.ctypes.data, n) == EINVAL                          # MRT has no phase field
    assert L.lbm_get_populations(h, 1, buf.ctypes.data, n * 9) == EINVAL                # ... and no g set
    assert L.lbm_get_populations(h, 0, buf.ctypes.data, n) == EINVAL
    assert L.lbm_set_geometry(h, buf.ctypes.data, n) == EINVAL                          # not a masked model
    assert L.lbm_comm_attach(h, None) == EINVAL
    ptr, cnt = C.c_void_p(), C.c_size_t()
    assert L.lbm_slab_halo(h, 0, C.byref(ptr), C.byref(cnt)) == EINVAL                  # 2-D models have no slabs
    assert L.lbm_slab_collide(h, 0) == EINVAL
    # the context still works
    rho, ux, uy = api.mrt_initial_fields("gs")
    for name, a in (("rho", rho), ("ux", ux), ("uy", uy)):
        ctx.set_field(name, a)
    ctx.init()
    _raises(api, EINVAL, ctx.step, -1)
    ctx.step(3)
    assert not ctx.diagnostics()["has_nan"]
    ctx.close()


def test_null_context_is_an_error_not_a_crash():
    api = _api()
    L = api.load_library()
    assert L.lbm_step(None, 1) == EINVAL
    assert L.lbm_init(None) == EINVAL
    assert L.lbm_local_nodes(None) == 0
    assert L.lbm_get_stream(None) is None
    L.lbm_destroy(None)


def test_slab_step_without_a_communicator():
    api = _api()
    cfg = api.microchannel_config("quarter", 12, 10, 16, slab_rank=0, slab_count=2)
    ctx = api.Context(cfg)
    for name, a in zip(("rho", "phase", "ux", "uy", "uz"), api.microchannel_initial_fields(12, 10, 16, 3, 0, 8)):
        ctx.set_field(name, a)
    _raises(api, ESTATE, ctx.init)          # slab_count > 1 and neither lbm_comm_attach nor the lbm_slab_* calls
    ctx.close()
