"""Multiple-relaxation-time collision for the D3Q19 binary liquid: an EXTENSION with no counterpart in the reference
(the north star names MRT on D3Q19; the reference has BGK only), so parity is unpinned by construction.  What can be
checked is checked: the moment basis is orthogonal; with every rate equal to 1/tau the scheme must reproduce the BGK
run up to rounding; whatever the rates, one collision must add exactly the body force to the momentum and conserve
mass and order parameter; the CUDA kernels must equal the oracle's restatement bit for bit."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import ndiff

F3 = ("rho", "phase", "ux", "uy", "uz")
CX = np.array([0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1])
CY = np.array([0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0])
CZ = np.array([0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1])
# aconst, kconst, gamma, phase_gradient, phase_wall, rho_wall, tau_phi, tau_liq, tau_gas, fx, fy, fz
CONST_TAU = [0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, 1.3, 1.3, 0.0, 0.0, 3e-5]
TWO_TAU = [0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, 2.5, 0.7, 1e-5, -2e-5, 3e-5]
DHUMIERES = [1.19, 1.4, 1.2, 1.4, 1.98]


def _basis():
    c2 = CX * CX + CY * CY + CZ * CZ
    rows = [np.ones(19, dtype=int), 19 * c2 - 30, (21 * c2 * c2 - 53 * c2 + 24) // 2, CX, (5 * c2 - 9) * CX, CY, (5 * c2 - 9) * CY,
            CZ, (5 * c2 - 9) * CZ, 3 * CX * CX - c2, (3 * c2 - 5) * (3 * CX * CX - c2), CY * CY - CZ * CZ,
            (3 * c2 - 5) * (CY * CY - CZ * CZ), CX * CY, CY * CZ, CX * CZ, (CY * CY - CZ * CZ) * CX, (CZ * CZ - CX * CX) * CY,
            (CX * CX - CY * CY) * CZ]
    return np.array(rows)


def test_moment_basis_is_the_published_one_and_orthogonal():
    M = _basis()
    gram = M @ M.T
    assert np.array_equal(gram, np.diag(np.diag(gram)))
    assert np.diag(gram).tolist() == [19, 2394, 252, 10, 40, 10, 40, 10, 40, 36, 72, 12, 24, 4, 4, 4, 8, 8, 8]   # d'Humieres et al. 2002
    assert M[1].tolist()[:7] == [-30, -11, -11, -11, -11, -11, -11] and M[2].tolist()[0] == 12 and M[2].tolist()[7] == 1


def _fields(nx, ny, nz, seed=2):
    rng = np.random.default_rng(seed)
    n = nx * ny * nz
    ph = ol.bubble_fields(nx, ny, nz, max(1, nx // 4))[1]
    return [1 + 0.03 * rng.standard_normal(n), np.clip(ph + 0.2 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]


def _oracle(shape, params, flds, rates, steps):
    o = ol.Oracle3D(*shape, params)
    o.set_mrt(rates)
    o.set_fields(*flds)
    o.init_single()
    o.step_single(steps)
    out = o.fields(), o.pops()
    o.close()
    return out


def test_oracle_mrt_reduces_to_bgk_when_all_rates_coincide():
    shape = (12, 10, 8)
    flds = _fields(*shape)
    w = 1.0 / CONST_TAU[7]
    bgk, (fb, gb) = _oracle(shape, CONST_TAU, flds, [0.0] * 5, 40)
    mrt, (fm, gm) = _oracle(shape, CONST_TAU, flds, [w] * 5, 40)
    for k in F3:
        assert np.max(np.abs(bgk[k] - mrt[k])) <= 1e-13 * max(1.0, np.max(np.abs(bgk[k]))), k
    assert np.max(np.abs(fb - fm)) <= 1e-13
    same, (fs, _) = _oracle(shape, CONST_TAU, flds, [-1.0] * 5, 40)      # "<= 0 means 1/tau": the same limit
    assert np.max(np.abs(fb - fs)) <= 1e-13
    other, _ = _oracle(shape, CONST_TAU, flds, DHUMIERES, 40)             # other rates: a different scheme
    assert np.max(np.abs(other["uz"] - bgk["uz"])) > 1e-9


@pytest.mark.parametrize("rates", [DHUMIERES, [1.0, 1.0, 1.0, 1.0, 1.0], [-1.0, 1.7, -1.0, 0.3, 1.9]])
@pytest.mark.parametrize("tau_liq,tau_gas", [(1.3, 1.3), (2.5, 0.7)])
def test_mrt_collision_adds_exactly_the_force_to_the_momentum(emu_libs, rates, tau_liq, tau_gas):
    """The per-node collision of csrc/d3q19_node.h (the code the kernels compile, run here on the CPU)."""
    L = C.CDLL(emu_libs["d3q19"])
    dp = C.POINTER(C.c_double)
    L.emu3_collide_node.argtypes = [dp, C.c_double, dp, dp, dp, dp, dp]
    L.emu3_collide_node.restype = None
    w = np.array([1 / 3.0] + [1 / 18.0] * 6 + [1 / 36.0] * 12)
    force = np.array([1e-3, -2e-3, 3e-3])
    params = np.array([0.04, 0.04, 1.0, 0.02, 0.0, 0.5, 2.0, tau_liq, tau_gas, force[0], force[1], force[2]])
    r5 = np.array(rates, dtype=np.float64)
    rng = np.random.default_rng(9)
    for trial in range(20):
        F = w * (1.0 + 0.1 * rng.standard_normal(19))
        ph0 = rng.uniform(-1, 1)
        PH = ph0 + 0.05 * rng.standard_normal(19)
        PH[0] = ph0
        G = w * ph0 * (1.0 + 0.1 * rng.standard_normal(19))
        G[0] += ph0 - G.sum()
        F0, G0 = F.copy(), G.copy()
        out4 = np.zeros(4)
        L.emu3_collide_node(params.ctypes.data_as(dp), 0.0, r5.ctypes.data_as(dp), F.ctypes.data_as(dp), G.ctypes.data_as(dp),
                            PH.ctypes.data_as(dp), out4.ctypes.data_as(dp))
        for c, fc in zip((CX, CY, CZ), force):
            assert abs((c * F).sum() - (c * F0).sum() - fc) <= 2e-15 * 19, (trial, rates)
        assert abs(F.sum() - F0.sum()) <= 2e-15 * 19
        assert abs(G.sum() - G0.sum()) <= 2e-15 * 19


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [0, 5, 2])
@pytest.mark.parametrize("shape,params,rates", [((12, 10, 8), CONST_TAU, DHUMIERES), ((27, 27, 20), TWO_TAU, DHUMIERES),
                                                ((9, 7, 5), TWO_TAU, [-1.0, 1.7, -1.0, 0.3, 1.9])])
def test_gpu_mrt_equals_oracle_restatement(shape, params, rates, variant):
    from latticeboltzmannmethod_b200 import api
    nx, ny, nz = shape
    flds = _fields(nx, ny, nz)
    cfg = api.microchannel_config("quarter", nx, ny, nz)
    (cfg.aconst, cfg.kconst, cfg.gammaconst, cfg.wall_gradient, _pw, cfg.rho_wall, cfg.tau_phi, cfg.tau_liq, cfg.tau_gas) = params[:9]
    cfg.force[0], cfg.force[1], cfg.force[2] = params[9:12]
    cfg.kernel_variant = variant
    for i in range(5):
        cfg.mrt_rates[i] = rates[i]
    ctx = api.Context(cfg)
    for name, a in zip(F3, flds):
        ctx.set_field(name, a)
    ctx.init()
    o = ol.Oracle3D(nx, ny, nz, params)
    o.set_mrt(rates)
    o.set_fields(*flds)
    o.init_single()
    for s in (1, 2, 14):
        ctx.step(s)
        o.step_single(s)
        want = o.fields()
        for k in F3:
            assert ndiff(want[k], ctx.get_field(k)) == 0, (s, k)
        fo, go = o.pops()
        assert ndiff(fo, ctx.get_populations(0)) == 0 and ndiff(go, ctx.get_populations(1)) == 0, s
        if s == 1:
            d0 = ctx.diagnostics()
    d1 = ctx.diagnostics()
    assert abs(d1["phase_mass"] - d0["phase_mass"]) <= 1e-9 * abs(d0["mass"])     # order parameter conserved
    ctx.close()
    o.close()


@pytest.mark.gpu
def test_gpu_mrt_limits():
    from latticeboltzmannmethod_b200 import api
    cfg = api.microchannel_config("quarter", 12, 10, 8)
    cfg.mrt_rates[0] = 2.5                                     # outside (0, 2)
    with pytest.raises(api.LbmError):
        api.Context(cfg)
    cfg.mrt_rates[0] = 1.2
    cfg.trt_magic = 0.25                                       # the two extensions exclude each other
    with pytest.raises(api.LbmError):
        api.Context(cfg)
    cfg.trt_magic = 0.0
    cfg.kernel_variant = 1                                     # the plain cross-check kernel has no MRT form
    with pytest.raises(api.LbmError):
        api.Context(cfg)
