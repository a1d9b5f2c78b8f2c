"""The per-thread functions the CUDA kernels are made of (csrc/d3q19_threads.h, d2q9_threads.h) run on the
CPU from host loops (tests/emu) and compared with the oracle.  This checks the kernels' arithmetic, boundary
folding, storage layout and ghost-plane protocol without a GPU; the GPU tests then only have to show that
nvcc (-fmad=false) and g++ (-ffp-contract=off) produce the same bits from the same source."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import _p, dp, ndiff


def _emu3(path):
    L = C.CDLL(path)
    L.emu3_create.restype = C.c_void_p
    L.emu3_create.argtypes = [C.c_int] * 3 + [dp]
    L.emu3_set_fields.argtypes = [C.c_void_p] + [dp] * 5
    L.emu3_get_fields.argtypes = [C.c_void_p] + [dp] * 5
    L.emu3_init.argtypes = [C.c_void_p]
    L.emu3_step.argtypes = [C.c_void_p, C.c_int]
    L.emu3_get_pops.argtypes = [C.c_void_p, dp, dp]
    L.emu3_destroy.argtypes = [C.c_void_p]
    L.emu3_use_streamed.argtypes = [C.c_void_p, C.c_int]
    return L


def _fields(nx, ny, nz, width, rand):
    flds = list(ol.bubble_fields(nx, ny, nz, width))
    if rand:
        rng = np.random.default_rng(5)
        n = nx * ny * nz
        flds = [1 + 0.05 * rng.standard_normal(n), np.clip(flds[1] + 0.3 * rng.standard_normal(n), -1.2, 1.2),
                0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]
    return [np.ascontiguousarray(a, dtype=np.float64) for a in flds]


GENERAL = [0.04, 0.04, 1.0, 0.05, 0.0, 0.5, 1.0, 2.5, 0.7, 1e-5, -2e-5, 3e-5]


@pytest.mark.parametrize("shape,params,width,steps,rand", [
    ((27, 27, 52), ol.PARAMS_QUARTER, 6, (1, 2, 9), False),
    ((9, 7, 5), ol.PARAMS_QUARTER, 3, (1, 3, 10), True),
    ((12, 10, 6), GENERAL, 3, (1, 3, 10), True),
    ((4, 4, 2), GENERAL, 1, (1, 3), True),
])
@pytest.mark.parametrize("streamed", [0, 1, 3, 4])
def test_d3q19_threads_match_oracle(emu_libs, shape, params, width, steps, rand, streamed):
    nx, ny, nz = shape
    n = nx * ny * nz
    L = _emu3(emu_libs["d3q19"])
    flds = _fields(nx, ny, nz, width, rand)
    o = ol.Oracle3D(nx, ny, nz, params)
    o.set_fields(*flds)
    o.init_single()
    p = np.asarray(params, dtype=np.float64)
    e = L.emu3_create(nx, ny, nz, _p(p))
    L.emu3_set_fields(e, *[_p(a) for a in flds])
    L.emu3_use_streamed(e, streamed)   # 0 collide(), 1 collide_streamed(), 3 interior (PART 3) + rim rows, 4 in-place (AA) on one array
    L.emu3_init(e)

    def compare(tag):
        a = o.fields()
        out = [np.empty(n) for _ in range(5)]
        L.emu3_get_fields(e, *[_p(x) for x in out])
        for k, b in zip(("rho", "phase", "ux", "uy", "uz"), out):
            assert ndiff(a[k], b) == 0, (tag, k)
        f, g = np.empty(n * 19), np.empty(n * 19)
        L.emu3_get_pops(e, _p(f), _p(g))
        fo, go = o.pops()
        assert ndiff(fo.ravel(), f) == 0 and ndiff(go.ravel(), g) == 0, tag

    compare("init")
    for s in steps:
        o.step_single(s)
        L.emu3_step(e, s)
        compare("+%d" % s)
    L.emu3_destroy(e)
    o.close()


@pytest.mark.parametrize("nslabs", [1, 2, 3])
@pytest.mark.parametrize("steps", [1, 2, 3, 6])
def test_d3q19_in_place_slab_protocol_matches_one_slab(emu_libs, nslabs, steps):
    """The halo protocol of the in-place step between slabs (Slab3D::aa_exchange + k_aa_merge: five directions per face,
    ghost rows after even steps, masked merge into the owned planes after odd ones, ghost-aware swap for reading in the
    odd half), emulated with memcpy as the transport, against the same steps on one periodic slab."""
    L = _emu3(emu_libs["d3q19"])
    L.emu3_aa_slabs_vs_whole.restype = C.c_longlong
    L.emu3_aa_slabs_vs_whole.argtypes = [C.c_void_p, C.c_int, C.c_int]
    nx, ny, nz = 12, 10, 9
    e = L.emu3_create(nx, ny, nz, _p(np.asarray(GENERAL, dtype=np.float64)))
    L.emu3_set_fields(e, *[_p(a) for a in _fields(nx, ny, nz, 3, True)])
    L.emu3_init(e)
    assert L.emu3_aa_slabs_vs_whole(e, nslabs, steps) == 0
    L.emu3_destroy(e)


@pytest.mark.parametrize("zwrap", [1, 0])
def test_d3q19_in_place_one_pass_schedule(emu_libs, zwrap):
    """aa_schedule (the plane-step arithmetic of k_aa_one): over a whole launch every owned plane is summed once and
    collided once, a plane's phase part comes before its collision, and every plane of a collision's phase stencil was
    summed at an earlier plane-step that is among the dependencies the block waits for -- one periodic slab and a slab
    of several, slabs from 2 planes up, every lag the library can choose."""
    L = C.CDLL(emu_libs["d3q19"])
    L.emu3_aa_schedule_violations.argtypes = [C.c_int] * 3
    for nz in list(range(2, 12)) + [52, 512, 1024]:
        for lag in (3, 4, 5, 17, 40):
            assert L.emu3_aa_schedule_violations(zwrap, nz, lag) == 0, (nz, lag)
    # a lag that is too short is caught (the check has teeth): the periodic order needs 3 plane-steps, a slab's order 2
    assert L.emu3_aa_schedule_violations(zwrap, 8, 2 if zwrap else 1) > 0


def _emu2(path):
    L = C.CDLL(path)
    L.emusc_create.restype = C.c_void_p
    L.emusc_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double]
    L.emusc_init.argtypes = [C.c_void_p, dp]
    L.emusc_step.argtypes = [C.c_void_p, C.c_int]
    L.emusc_get.argtypes = [C.c_void_p] + [dp] * 4
    L.emub2_create.restype = C.c_void_p
    L.emub2_create.argtypes = [C.c_int, C.c_int, dp]
    L.emub2_init.argtypes = [C.c_void_p, dp]
    L.emub2_step.argtypes = [C.c_void_p, C.c_int]
    L.emub2_get.argtypes = [C.c_void_p] + [dp] * 6
    L.emutrt_create.restype = C.c_void_p
    L.emutrt_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p]
    L.emutrt_init.argtypes = [C.c_void_p] + [dp] * 3
    L.emutrt_step.argtypes = [C.c_void_p, C.c_int]
    L.emutrt_get.argtypes = [C.c_void_p] + [dp] * 2
    L.emubs_create.restype = C.c_void_p
    L.emubs_create.argtypes = [C.c_int, C.c_int, dp, C.c_void_p]
    L.emubs_init.argtypes = [C.c_void_p] + [dp] * 4
    L.emubs_step.argtypes = [C.c_void_p, C.c_int]
    L.emubs_get.argtypes = [C.c_void_p] + [dp] * 6
    L.emumrt_create.restype = C.c_void_p
    L.emumrt_create.argtypes = [C.c_int] * 3 + [C.c_double] * 3
    L.emumrt_init.argtypes = [C.c_void_p] + [dp] * 3
    L.emumrt_step.argtypes = [C.c_void_p, C.c_int]
    L.emumrt_get.argtypes = [C.c_void_p] + [dp] * 4
    L.emumrt_wave.argtypes = [C.c_int] * 3 + [C.c_double] * 2 + [dp] * 3
    return L


def test_shanchen_threads_match_oracle(emu_libs):
    L = _emu2(emu_libs["d2q9"])
    nx = ny = 64
    n = nx * ny
    o = ol.OracleShanChen(nx, ny)
    o.init_droplet(radius=12)
    rho0 = o.fields()["rho"].copy()
    e = L.emusc_create(nx, ny, 1.0, -5.0)
    L.emusc_init(e, _p(rho0))
    for s in (1, 10, 60):
        o.step(s)
        L.emusc_step(e, s)
        r, u1, u2, f = np.empty(n), np.empty(n), np.empty(n), np.empty(9 * n)
        L.emusc_get(e, _p(r), _p(u1), _p(u2), _p(f))
        a = o.fields()
        assert ndiff(a["rho"], r) == 0 and ndiff(a["u1"], u1) == 0 and ndiff(a["u2"], u2) == 0
        assert ndiff(o.pops().ravel(), f) == 0
    o.close()


@pytest.mark.parametrize("nx,ny", [(128, 128), (17, 11)])
def test_binary2d_threads_match_oracle(emu_libs, nx, ny):
    L = _emu2(emu_libs["d2q9"])
    n = nx * ny
    o = ol.OracleBinary2D(nx, ny)
    if (nx, ny) == (128, 128):
        o.init_reference()
        ix, iy = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
        ph0 = np.where((abs(ix - nx // 2) <= 20) & (iy < 40) & (iy >= 1), 1.0, -1.0).ravel()
    else:
        ph0 = np.random.default_rng(1).uniform(-1, 1, n)
        o.init_phase(ph0)
    p = np.asarray(ol.PARAMS_BINARY2D)
    e = L.emub2_create(nx, ny, _p(p))
    L.emub2_init(e, _p(np.ascontiguousarray(ph0)))
    for s in (0, 1, 10, 40):
        o.step(s)
        L.emub2_step(e, s)
        out = [np.empty(n) for _ in range(4)] + [np.empty(9 * n), np.empty(9 * n)]
        L.emub2_get(e, *[_p(x) for x in out])
        a = o.fields()
        fo, go = o.pops()
        for i, k in enumerate(("phase", "rho", "ux", "uy")):
            assert ndiff(a[k], out[i]) == 0, (k, s)
        assert ndiff(fo.ravel(), out[4]) == 0 and ndiff(go.ravel(), out[5]) == 0
    o.close()


@pytest.mark.parametrize("aa", [0, 1])
@pytest.mark.parametrize("variant,om,ob,fa,shape", [(0, 1.0, 1.0, 0.002, (40, 3)), (0, 1.2, 0.8, 0.002, (33, 7)),
                                                    (1, 0.9, 0.1, 0.0, (40, 3))])
def test_mrt_threads_match_oracle(emu_libs, variant, om, ob, fa, shape, aa):
    """aa = 1: the in-place AA-pattern form (one population array, even and odd steps alternating); the streamed
    populations gathered out of it must equal the oracle's after an odd as well as after an even number of steps."""
    L = _emu2(emu_libs["d2q9"])
    L.emumrt_use_aa.argtypes = [C.c_void_p, C.c_int]
    nx, ny = shape
    n = nx * ny
    o = ol.OracleMRT(variant, nx, ny, om, ob, fa)
    o.init_wave(0.001)
    rho, ux, uy = np.empty(n), np.empty(n), np.empty(n)
    L.emumrt_wave(variant, nx, ny, 0.001, fa, _p(rho), _p(ux), _p(uy))
    e = L.emumrt_create(variant, nx, ny, om, ob, fa)
    L.emumrt_use_aa(e, aa)
    L.emumrt_init(e, _p(rho), _p(ux), _p(uy))
    for s in (0, 1, 50, 200, 3):
        o.step(s)
        L.emumrt_step(e, s)
        out = [np.empty(n) for _ in range(3)] + [np.empty(9 * n)]
        L.emumrt_get(e, *[_p(x) for x in out])
        a = o.fields()
        for i, k in enumerate(("rho", "ux", "uy")):
            assert ndiff(a[k], out[i]) == 0, (k, s)
        assert ndiff(o.pops().ravel(), out[3]) == 0
    o.close()


@pytest.mark.parametrize("case", ["cylinder", "cylinder_advected", "blobs_40x33", "blobs_97x61"])
def test_trt_threads_match_oracle(emu_libs, case):
    """k_trt_init / k_trt_step thread functions (push with folded anti-bounce-back, link bytes) against the oracle."""
    L = _emu2(emu_libs["d2q9"])
    omega, ce, cw = ol.TRT_OMEGA, 1.0 / 3.0, 1.0
    if case.startswith("cylinder"):
        nx = ny = 129
        mask, rho, ux, uy = ol.trt_reference_setup()
        if case == "cylinder_advected":
            omega, ce, cw = 1.3, 0.25, 0.7
            rho, ux, uy = ol.trt_advected_fields(mask, cw)
    else:
        nx, ny = [int(v) for v in case.split("_")[1].split("x")]
        mask, rho, ux, uy = ol.trt_random_case(nx, ny, seed=nx)
        omega, ce, cw = 0.9, 0.3, 1.5
    n = nx * ny
    o = ol.OracleTRT(nx, ny, mask, omega, ce, cw)
    o.init(rho, ux, uy)
    e = L.emutrt_create(nx, ny, omega, ce, cw, mask.ctypes.data)
    L.emutrt_init(e, _p(rho), _p(ux), _p(uy))
    for s in (0, 1, 2, 150):
        o.step(s)
        L.emutrt_step(e, s)
        r, f = np.empty(n), np.empty(9 * n)
        L.emutrt_get(e, _p(r), _p(f))
        assert ndiff(o.rho(), r) == 0, s
        assert ndiff(o.pops().ravel(), f) == 0, s
    L.emutrt_destroy(e)
    o.close()


@pytest.mark.parametrize("case", ["disc_301", "disc_perturbed_160x120", "discs_64x48", "discs_131x77"])
def test_shape_threads_match_oracle(emu_libs, case):
    """k_bs_init / k_bs_phase / k_bs_collide thread functions (periodic push with folded bounce-back, wall-normal
    phase extrapolation recomputed in the phase pass) against the oracle."""
    L = _emu2(emu_libs["d2q9"])
    params = list(ol.PARAMS_SHAPE)
    if case == "disc_301":
        nx = ny = 301
        mask, rho, phi, ux, uy = ol.shape_reference_setup()
        steps = (0, 1, 2, 40)
    elif case.startswith("disc_perturbed"):
        nx, ny = 160, 120
        mask, rho, phi, ux, uy = ol.shape_reference_setup(nx, ny, 30, 7)
        rho, phi, ux, uy = ol.shape_perturbed_fields(mask, rho, phi)
        params = [0.03, 0.05, 0.9, 0.1, 0.8, 1.4, 1.0 / 1.3, 1e-5, -2e-5]
        steps = (0, 1, 2, 120)
    else:
        nx, ny = [int(v) for v in case.split("_")[1].split("x")]
        mask, rho, phi, ux, uy = ol.shape_random_case(nx, ny, seed=nx)
        params = [0.04, 0.04, 1.0, -0.15, 0.7, 2.5, 0.5, 0.0, 3e-5]
        steps = (0, 1, 2, 150)
    n = nx * ny
    p = np.array(params)
    o = ol.OracleShape(nx, ny, mask, params)
    o.init(rho, phi, ux, uy)
    e = L.emubs_create(nx, ny, _p(p), mask.ctypes.data)
    L.emubs_init(e, _p(rho), _p(phi), _p(ux), _p(uy))
    for s in steps:
        o.step(s)
        L.emubs_step(e, s)
        out = [np.empty(n) for _ in range(4)] + [np.empty(9 * n), np.empty(9 * n)]
        L.emubs_get(e, *[_p(x) for x in out])
        a = o.fields()
        for i, k in enumerate(("phase", "rho", "ux", "uy")):
            assert ndiff(a[k], out[i]) == 0, (k, s)
        fo, go = o.pops()
        assert ndiff(fo.ravel(), out[4]) == 0 and ndiff(go.ravel(), out[5]) == 0, s
    L.emubs_destroy(e)
    o.close()
