"""Two-relaxation-time collision for the D3Q19 binary liquid: an EXTENSION with no counterpart in the reference
(the north star names TRT on D3Q19; the reference has BGK only), so parity is unpinned by construction.
What can be checked is checked: with constant tau and Lambda = (tau - 1/2)^2 both rates coincide and the scheme
must reproduce the BGK run up to rounding; the CUDA kernels must equal the oracle's restatement bit for bit;
mass and order parameter must stay conserved."""
import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import ndiff

F3 = ("rho", "phase", "ux", "uy", "uz")
# aconst, kconst, gamma, phase_gradient, phase_wall, rho_wall, tau_phi, tau_liq, tau_gas, fx, fy, fz
CONST_TAU = [0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, 1.3, 1.3, 0.0, 0.0, 3e-5]
TWO_TAU = [0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, 2.5, 0.7, 1e-5, -2e-5, 3e-5]


def _fields(nx, ny, nz, seed=2):
    rng = np.random.default_rng(seed)
    n = nx * ny * nz
    ph = ol.bubble_fields(nx, ny, nz, max(1, nx // 4))[1]
    return [1 + 0.03 * rng.standard_normal(n), np.clip(ph + 0.2 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]


def _oracle(shape, params, flds, magic, steps):
    o = ol.Oracle3D(*shape, params)
    o.set_trt(magic)
    o.set_fields(*flds)
    o.init_single()
    o.step_single(steps)
    out = o.fields(), o.pops()
    o.close()
    return out


def test_oracle_trt_reduces_to_bgk_when_both_rates_coincide():
    shape = (12, 10, 8)
    flds = _fields(*shape)
    tau = CONST_TAU[7]
    bgk, (fb, gb) = _oracle(shape, CONST_TAU, flds, 0.0, 40)
    trt, (ft, gt) = _oracle(shape, CONST_TAU, flds, (tau - 0.5) ** 2, 40)
    for k in F3:
        assert np.max(np.abs(bgk[k] - trt[k])) <= 1e-13 * max(1.0, np.max(np.abs(bgk[k]))), k
    assert np.max(np.abs(fb - ft)) <= 1e-13
    other, _ = _oracle(shape, CONST_TAU, flds, 3.0 / 16.0, 40)          # a different Lambda is a different scheme
    assert np.max(np.abs(other["uz"] - bgk["uz"])) > 1e-9


@pytest.mark.parametrize("magic", [0.0, 3.0 / 16.0, 0.25, 0.64, 2.0])
@pytest.mark.parametrize("tau_liq,tau_gas", [(1.3, 1.3), (2.5, 0.7)])
def test_trt_collision_adds_exactly_the_force_to_the_momentum(emu_libs, magic, tau_liq, tau_gas):
    """The per-node collision of csrc/d3q19_node.h (the very code the kernels compile, run here on the CPU): whatever
    Lambda is, one collision must change the node's momentum by exactly the body force F -- the antisymmetric half of
    the Guo force populations has to carry (1 - omega_minus/2), not the BGK prefactor -- and must conserve the mass
    and the order parameter."""
    import ctypes as C
    L = C.CDLL(emu_libs["d3q19"])
    dp = C.POINTER(C.c_double)
    L.emu3_collide_node.argtypes = [dp, C.c_double, dp, dp, dp, dp, dp]
    L.emu3_collide_node.restype = None
    no_mrt = np.zeros(5)
    cx = np.array([0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1], dtype=np.float64)
    cy = np.array([0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0], dtype=np.float64)
    cz = np.array([0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1], dtype=np.float64)
    w = np.array([1 / 3.0] + [1 / 18.0] * 6 + [1 / 36.0] * 12)
    force = np.array([1e-3, -2e-3, 3e-3])
    params = np.array([0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, tau_liq, tau_gas, force[0], force[1], force[2]])
    rng = np.random.default_rng(7)
    for trial in range(20):
        F = w * (1.0 + 0.1 * rng.standard_normal(19))
        ph0 = rng.uniform(-1, 1)
        PH = ph0 + 0.05 * rng.standard_normal(19)
        PH[0] = ph0
        G = w * ph0 * (1.0 + 0.1 * rng.standard_normal(19))
        G[0] += ph0 - G.sum()                                   # g sums to the phase value the stencil sees
        F0, G0 = F.copy(), G.copy()
        out4 = np.zeros(4)
        L.emu3_collide_node(params.ctypes.data_as(dp), magic, no_mrt.ctypes.data_as(dp), F.ctypes.data_as(dp), G.ctypes.data_as(dp),
                            PH.ctypes.data_as(dp), out4.ctypes.data_as(dp))
        for c, fc in zip((cx, cy, cz), force):
            assert abs((c * F).sum() - (c * F0).sum() - fc) <= 1e-15 * 19, (trial, magic)
        assert abs(F.sum() - F0.sum()) <= 1e-15 * 19
        assert abs(G.sum() - G0.sum()) <= 1e-15 * 19


@pytest.mark.gpu
@pytest.mark.parametrize("shape,params,magic", [((12, 10, 8), CONST_TAU, 0.64), ((27, 27, 20), TWO_TAU, 3.0 / 16.0),
                                                ((9, 7, 5), TWO_TAU, 0.25)])
@pytest.mark.parametrize("variant", [0, 2])
def test_gpu_trt_equals_oracle_restatement(shape, params, magic, variant):
    from latticeboltzmannmethod_b200 import api
    nx, ny, nz = shape
    flds = _fields(nx, ny, nz)
    cfg = api.microchannel_config("quarter", nx, ny, nz)
    (cfg.aconst, cfg.kconst, cfg.gammaconst, cfg.wall_gradient, _pw, cfg.rho_wall, cfg.tau_phi, cfg.tau_liq, cfg.tau_gas) = params[:9]
    cfg.force[0], cfg.force[1], cfg.force[2] = params[9:12]
    cfg.trt_magic = magic
    cfg.kernel_variant = variant        # 0: two-pass with the register kernel, 2: in place on one population array
    ctx = api.Context(cfg)
    for name, a in zip(F3, flds):
        ctx.set_field(name, a)
    ctx.init()
    d0 = ctx.diagnostics()
    o = ol.Oracle3D(nx, ny, nz, params)
    o.set_trt(magic)
    o.set_fields(*flds)
    o.init_single()
    for s in (1, 2, 30):
        ctx.step(s)
        o.step_single(s)
        want = o.fields()
        for k in F3:
            assert ndiff(want[k], ctx.get_field(k)) == 0, (k, s)
        fo, go = o.pops()
        assert ndiff(fo, ctx.get_populations(0)) == 0 and ndiff(go, ctx.get_populations(1)) == 0, s
    d1 = ctx.diagnostics()
    assert not d1["has_nan"]
    assert abs(d1["phase_mass"] - d0["phase_mass"]) <= 1e-9 * abs(d0["mass"])
    ctx.close()
    o.close()


@pytest.mark.gpu
def test_gpu_trt_limits():
    from latticeboltzmannmethod_b200 import api
    cfg = api.microchannel_config("quarter", 12, 10, 8)
    cfg.trt_magic = -1.0
    with pytest.raises(api.LbmError):
        api.Context(cfg)
    cfg.trt_magic = 0.25
    cfg.kernel_variant = 1                                   # the plain cross-check kernel has no TRT form
    with pytest.raises(api.LbmError):
        api.Context(cfg)
