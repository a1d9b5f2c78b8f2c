"""ctypes bindings for the CPU checkers (test infrastructure only).

* ``oracle/liblbm_oracle.so`` -- this repo's C restatement (oracle/lbm_oracle.c).
* ``oracle/_ref/libref_*.so`` -- the unmodified reference compiled in place by
  ``make -C oracle ref`` (only where /root/reference exists; the prebuilt files
  travel to the GPU box).

Nothing in the product package imports this module.
"""
import ctypes as C
import os
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liblbm_oracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

dp = C.POINTER(C.c_double)


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(dp)


def ref_available(name):
    return os.path.exists(os.path.join(REF_DIR, "libref_%s.so" % name))


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            raise RuntimeError("oracle library missing: run `make -C oracle oracle` (or __graft_entry__.build())")
        L = C.CDLL(ORACLE_SO)
        L.lbo3_create.restype = C.c_void_p
        L.lbo3_create.argtypes = [C.c_int, C.c_int, C.c_int, dp]
        L.lbo3_halo.restype = dp
        L.lbo3_halo.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
        for fn in ("lbo3_destroy", "lbo3_init_fields", "lbo3_init_pops", "lbo3_after_init", "lbo3_finish",
                   "lbo3_self_exchange_phi", "lbo3_self_exchange_step", "lbo3_init_single"):
            getattr(L, fn).argtypes = [C.c_void_p]
            getattr(L, fn).restype = None
        L.lbo3_collide.argtypes = [C.c_void_p, C.c_int]
        L.lbo3_step_single.argtypes = [C.c_void_p, C.c_int]
        L.lbo3_set_fields.argtypes = [C.c_void_p] + [dp] * 5
        L.lbo3_get_fields.argtypes = [C.c_void_p] + [dp] * 5
        L.lbo3_get_pops.argtypes = [C.c_void_p, dp, dp]
        L.lbo3_set_pops.argtypes = [C.c_void_p, dp, dp]
        L.lbo2b_create.restype = C.c_void_p
        L.lbo2b_create.argtypes = [C.c_int, C.c_int, dp]
        for fn in ("lbo2b_destroy", "lbo2b_init_reference_ic", "lbo2b_init"):
            getattr(L, fn).argtypes = [C.c_void_p]
        L.lbo2b_set_phase.argtypes = [C.c_void_p, dp]
        L.lbo2b_step.argtypes = [C.c_void_p, C.c_int]
        L.lbo2b_get_fields.argtypes = [C.c_void_p] + [dp] * 4
        L.lbo2b_get_pops.argtypes = [C.c_void_p, dp, dp]
        L.lbo2b_set_pops.argtypes = [C.c_void_p, dp, dp]
        L.lbosc_create.restype = C.c_void_p
        L.lbosc_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_int]
        L.lbosc_destroy.argtypes = [C.c_void_p]
        L.lbosc_init_droplet.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
        L.lbosc_set_pops.argtypes = [C.c_void_p, dp]
        L.lbosc_step.argtypes = [C.c_void_p, C.c_int]
        L.lbosc_get_fields.argtypes = [C.c_void_p, dp, dp, dp]
        L.lbosc_get_pops.argtypes = [C.c_void_p, dp]
        L.lbomrt_create.restype = C.c_void_p
        L.lbomrt_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
        L.lbomrt_destroy.argtypes = [C.c_void_p]
        L.lbomrt_init_wave.argtypes = [C.c_void_p, C.c_double]
        L.lbomrt_set_pops.argtypes = [C.c_void_p, dp]
        L.lbomrt_step.argtypes = [C.c_void_p, C.c_int]
        L.lbomrt_get_fields.argtypes = [C.c_void_p, dp, dp, dp]
        L.lbomrt_get_pops.argtypes = [C.c_void_p, dp]
        L.lbotrt_create.restype = C.c_void_p
        L.lbotrt_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.lbotrt_destroy.argtypes = [C.c_void_p]
        L.lbotrt_init.argtypes = [C.c_void_p, dp, dp, dp]
        L.lbotrt_step.argtypes = [C.c_void_p, C.c_int]
        L.lbotrt_get_rho.argtypes = [C.c_void_p, dp]
        L.lbotrt_get_pops.argtypes = [C.c_void_p, dp]
        L.lbotrt_reference_setup.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p, dp, dp, dp]
        L.lbobs_create.restype = C.c_void_p
        L.lbobs_create.argtypes = [C.c_int, C.c_int, dp, C.c_void_p]
        L.lbobs_destroy.argtypes = [C.c_void_p]
        L.lbobs_init.argtypes = [C.c_void_p] + [dp] * 4
        L.lbobs_step.argtypes = [C.c_void_p, C.c_int]
        L.lbobs_get_fields.argtypes = [C.c_void_p] + [dp] * 4
        L.lbobs_get_pops.argtypes = [C.c_void_p, dp, dp]
        L.lbobs_reference_setup.argtypes = [C.c_int] * 4 + [C.c_void_p] + [dp] * 4
        _oracle = L
    return _oracle


# QuarterChannel/main.cpp:48-62 and FullChannel/main.cpp:49-70, in the LBO3_* order
PARAMS_QUARTER = [0.04, 0.04, 1.0, 0.0, 0.0, 0.5, 2.0, 2.5, 0.7, 0.0, 0.0, 0.00003]
PARAMS_FULL = [0.04, 0.04, 1.0, 0.0, 0.0, 0.5, 1.0, 2.5, 0.7, 0.0, 0.0, 0.00003]
HALO = dict(S_F_UP=0, S_G_UP=1, S_F_DOWN=2, S_G_DOWN=3,
            R_F_FROM_BELOW=4, R_G_FROM_BELOW=5, R_F_FROM_ABOVE=6, R_G_FROM_ABOVE=7,
            S_PHI_UP=8, S_PHI_DOWN=9, R_PHI_FROM_BELOW=10, R_PHI_FROM_ABOVE=11)


def bubble_fields(nx, ny, nz, width):
    """Initial condition of Microchannel3D main.cpp:83-100 (rho, phase, ux, uy, uz), node order [z][y][x]."""
    z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    gas = (z >= (nz - 1) // 3) & (z <= 2 * (nz - 1) // 3) & (x <= nx - width - 1) & (y <= ny - width - 1)
    phase = np.where(gas, -1.0, 1.0).astype(np.float64).ravel()
    n = nx * ny * nz
    return np.ones(n), phase, np.zeros(n), np.zeros(n), np.zeros(n)


class Oracle3D:
    """One z-slab of the D3Q19 binary-liquid oracle."""

    def __init__(self, nx, ny, nz, params):
        self.L = oracle()
        self.nx, self.ny, self.nz = nx, ny, nz
        self.n = nx * ny * nz
        p = np.asarray(params, dtype=np.float64)
        self.h = self.L.lbo3_create(nx, ny, nz, _p(p))
        assert self.h

    def set_mrt(self, rates5):
        """Extension without a reference counterpart: multiple-relaxation-time collision of f (all 0 = BGK)."""
        r = (C.c_double * 5)(*rates5)
        self.L.lbo3_set_mrt.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        self.L.lbo3_set_mrt(self.h, r)

    def set_trt(self, magic):
        """Extension without a reference counterpart: two-relaxation-time collision of f (0 = the reference's BGK)."""
        self.L.lbo3_set_trt.argtypes = [C.c_void_p, C.c_double]
        self.L.lbo3_set_trt(self.h, magic)

    def close(self):
        if self.h:
            self.L.lbo3_destroy(self.h)
            self.h = None

    def set_fields(self, rho, phase, ux, uy, uz):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (rho, phase, ux, uy, uz)]
        self.L.lbo3_set_fields(self.h, *[_p(a) for a in arrs])

    def init_single(self):
        self.L.lbo3_init_single(self.h)

    def step_single(self, n):
        self.L.lbo3_step_single(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(5)]
        self.L.lbo3_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("rho", "phase", "ux", "uy", "uz"), out))

    def pops(self):
        f, g = np.empty(self.n * 19), np.empty(self.n * 19)
        self.L.lbo3_get_pops(self.h, _p(f), _p(g))
        return f.reshape(-1, 19), g.reshape(-1, 19)

    def set_pops(self, f, g):
        f = np.ascontiguousarray(f, dtype=np.float64).ravel()
        g = np.ascontiguousarray(g, dtype=np.float64).ravel()
        self.L.lbo3_set_pops(self.h, _p(f), _p(g))

    def halo(self, name):
        cnt = C.c_size_t()
        ptr = self.L.lbo3_halo(self.h, HALO[name], C.byref(cnt))
        return np.ctypeslib.as_array(ptr, shape=(cnt.value,))


class Ref3D:
    """The unmodified reference Lattice/Solver (oracle/_ref/libref_{quarter,full}.so), one rank."""

    def __init__(self, nx, ny, nz, params, which="quarter"):
        L = C.CDLL(os.path.join(REF_DIR, "libref_%s.so" % which))
        L.ref3d_create.restype = C.c_void_p
        L.ref3d_create.argtypes = [C.c_int, C.c_int, C.c_int, dp]
        L.ref3d_set_fields.argtypes = [C.c_void_p] + [dp] * 5
        L.ref3d_get_fields.argtypes = [C.c_void_p] + [dp] * 5
        L.ref3d_set_bubble.argtypes = [C.c_void_p, C.c_int]
        L.ref3d_init.argtypes = [C.c_void_p]
        L.ref3d_step.argtypes = [C.c_void_p, C.c_int]
        L.ref3d_get_pops.argtypes = [C.c_void_p, dp, dp]
        L.ref3d_set_pops.argtypes = [C.c_void_p, dp, dp]
        L.ref3d_destroy.argtypes = [C.c_void_p]
        self.L = L
        self.nx, self.ny, self.nz = nx, ny, nz
        self.n = nx * ny * nz
        p = np.asarray(params, dtype=np.float64)
        self.h = L.ref3d_create(nx, ny, nz, _p(p))

    def close(self):
        if self.h:
            self.L.ref3d_destroy(self.h)
            self.h = None

    def set_fields(self, rho, phase, ux, uy, uz):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (rho, phase, ux, uy, uz)]
        self.L.ref3d_set_fields(self.h, *[_p(a) for a in arrs])

    def set_bubble(self, width):
        self.L.ref3d_set_bubble(self.h, width)

    def init(self):
        self.L.ref3d_init(self.h)

    def step(self, n):
        self.L.ref3d_step(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(5)]
        self.L.ref3d_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("rho", "phase", "ux", "uy", "uz"), out))

    def pops(self):
        f, g = np.empty(self.n * 19), np.empty(self.n * 19)
        self.L.ref3d_get_pops(self.h, _p(f), _p(g))
        return f.reshape(-1, 19), g.reshape(-1, 19)


def bulk_mask_3d(nx, ny, nz):
    z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    return ((x >= 1) & (x <= nx - 2) & (y >= 1) & (y <= ny - 2)).ravel()


def nonwall_mask_3d(nx, ny, nz):
    z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    return ((x <= nx - 2) & (y <= ny - 2)).ravel()


# ---------------- 2-D programs ----------------
# FreeEnergy/binary.cpp:29-41: aconst,kconst,gammaconst,wall_gradient,tau_gas,tau_liq,omega_phi
PARAMS_BINARY2D = [0.04, 0.04, 1.0, -0.15, 1.0, 1.0, 1.0]


class OracleBinary2D:
    def __init__(self, nx, ny, params=PARAMS_BINARY2D):
        self.L = oracle()
        self.nx, self.ny, self.n = nx, ny, nx * ny
        p = np.asarray(params, dtype=np.float64)
        self.h = self.L.lbo2b_create(nx, ny, _p(p))

    def close(self):
        if self.h:
            self.L.lbo2b_destroy(self.h)
            self.h = None

    def init_reference(self):
        self.L.lbo2b_init_reference_ic(self.h)
        self.L.lbo2b_init(self.h)

    def init_phase(self, phase):
        a = np.ascontiguousarray(phase, dtype=np.float64).ravel()
        self.L.lbo2b_set_phase(self.h, _p(a))
        self.L.lbo2b_init(self.h)

    def step(self, n):
        self.L.lbo2b_step(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(4)]
        self.L.lbo2b_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("phase", "rho", "ux", "uy"), out))

    def pops(self):
        f, g = np.empty(self.n * 9), np.empty(self.n * 9)
        self.L.lbo2b_get_pops(self.h, _p(f), _p(g))
        return f.reshape(-1, 9), g.reshape(-1, 9)

    def set_pops(self, f, g):
        f = np.ascontiguousarray(f, dtype=np.float64).ravel()
        g = np.ascontiguousarray(g, dtype=np.float64).ravel()
        self.L.lbo2b_set_pops(self.h, _p(f), _p(g))


class OracleShanChen:
    def __init__(self, nx, ny, tau=1.0, g=-5.0, use_libm_exp=False):
        self.L = oracle()
        self.nx, self.ny, self.n = nx, ny, nx * ny
        self.h = self.L.lbosc_create(nx, ny, tau, g, int(use_libm_exp))

    def close(self):
        if self.h:
            self.L.lbosc_destroy(self.h)
            self.h = None

    def init_droplet(self, rhol=1.95, rhog=0.15, radius=20):
        self.L.lbosc_init_droplet(self.h, rhol, rhog, radius)

    def set_pops(self, f):
        f = np.ascontiguousarray(f, dtype=np.float64).ravel()
        self.L.lbosc_set_pops(self.h, _p(f))

    def step(self, n):
        self.L.lbosc_step(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(3)]
        self.L.lbosc_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("rho", "u1", "u2"), out))

    def pops(self):
        f = np.empty(self.n * 9)
        self.L.lbosc_get_pops(self.h, _p(f))
        return f.reshape(-1, 9)


class OracleMRT:
    def __init__(self, variant, nx, ny, omega, omega_bulk, force_ampl=0.002):
        self.L = oracle()
        self.nx, self.ny, self.n = nx, ny, nx * ny
        self.h = self.L.lbomrt_create(variant, nx, ny, omega, omega_bulk, force_ampl)

    def close(self):
        if self.h:
            self.L.lbomrt_destroy(self.h)
            self.h = None

    def init_wave(self, ampl=0.001):
        self.L.lbomrt_init_wave(self.h, ampl)

    def set_pops(self, f):
        f = np.ascontiguousarray(f, dtype=np.float64).ravel()
        self.L.lbomrt_set_pops(self.h, _p(f))

    def step(self, n):
        self.L.lbomrt_step(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(3)]
        self.L.lbomrt_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("rho", "ux", "uy"), out))

    def pops(self):
        f = np.empty(self.n * 9)
        self.L.lbomrt_get_pops(self.h, _p(f))
        return f.reshape(-1, 9)


TRT_OMEGA = 1.0 / (0.5 + 0.5 / 32.0)   # cylinder.cpp:38


class OracleTRT:
    """DiffusionCylinder/cylinder.cpp restated (oracle/lbm_oracle.c, lbotrt_*); node i = iY*nx+iX."""

    def __init__(self, nx, ny, mask, omega=TRT_OMEGA, ce=1.0 / 3.0, conc_wall=1.0):
        self.L = oracle()
        self.nx, self.ny, self.n = nx, ny, nx * ny
        self.mask = np.ascontiguousarray(mask, dtype=np.int32).ravel()
        self.h = self.L.lbotrt_create(nx, ny, omega, ce, conc_wall, self.mask.ctypes.data)

    def close(self):
        if self.h:
            self.L.lbotrt_destroy(self.h)
            self.h = None

    def init(self, rho, ux, uy):
        a = [np.ascontiguousarray(x, dtype=np.float64).ravel() for x in (rho, ux, uy)]
        self.L.lbotrt_init(self.h, *[_p(x) for x in a])

    def step(self, n):
        self.L.lbotrt_step(self.h, n)

    def rho(self):
        r = np.empty(self.n)
        self.L.lbotrt_get_rho(self.h, _p(r))
        return r

    def pops(self):
        f = np.empty(self.n * 9)
        self.L.lbotrt_get_pops(self.h, _p(f))
        return f.reshape(-1, 9)


def trt_reference_setup(nx=129, ny=129, radius=40, conc_wall=1.0):
    """cylinder.cpp:224-296 through the oracle: mask, rho, ux, uy."""
    n = nx * ny
    mask, rho, ux, uy = np.zeros(n, np.int32), np.empty(n), np.empty(n), np.empty(n)
    oracle().lbotrt_reference_setup(nx, ny, radius, conc_wall, mask.ctypes.data, _p(rho), _p(ux), _p(uy))
    return mask, rho, ux, uy


def trt_random_case(nx, ny, seed, speed=0.05):
    """A random blob geometry that satisfies lbm_set_geometry's rules, with random rho and a random
    velocity field: fluid = a thresholded smooth random field kept off the array edge."""
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((ny, nx))
    for _ in range(3):
        a = (a + np.roll(a, 1, 0) + np.roll(a, -1, 0) + np.roll(a, 1, 1) + np.roll(a, -1, 1)) / 5.0
    fluid = a > 0.0
    fluid[:2, :] = fluid[-2:, :] = False
    fluid[:, :2] = fluid[:, -2:] = False
    pad = np.pad(fluid, 1)
    near = np.zeros_like(fluid)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            near |= pad[1 + dy:1 + dy + ny, 1 + dx:1 + dx + nx]
    mask = np.where(fluid, 1, np.where(near, 0, -1)).astype(np.int32)
    rho = rng.uniform(0.0, 1.0, (ny, nx))
    ux, uy = rng.uniform(-speed, speed, (ny, nx)), rng.uniform(-speed, speed, (ny, nx))
    return mask.ravel(), rho.ravel(), ux.ravel(), uy.ravel()


def hashed_uniform(n, salt):
    """n doubles in [0,1) from integer arithmetic only (identical on every machine and numpy version)."""
    i = np.arange(n, dtype=np.uint64)
    return (((i * np.uint64(2654435761) + np.uint64(salt) * np.uint64(40503)) % np.uint64(2 ** 32)).astype(np.float64)) / 2.0 ** 32


def trt_advected_fields(mask, conc_wall, speed=0.05):
    """A non-trivial initial state on the reference geometry: hashed rho in the fluid, hashed velocity everywhere."""
    n = mask.size
    rho = np.where(mask == 1, hashed_uniform(n, 1), np.where(mask == 0, conc_wall, -1.0))
    ux = (hashed_uniform(n, 2) - 0.5) * (2 * speed)
    uy = (hashed_uniform(n, 3) - 0.5) * (2 * speed)
    return rho, ux, uy


def ref_cylinder_run(steps, omega=0.0, ce=0.0, conc_wall=0.0, rho=None, ux=None, uy=None):
    """The unmodified cylinder.cpp (129x129, radius 40) for `steps` loop iterations; omega <= 0 keeps its constants."""
    L = C.CDLL(os.path.join(REF_DIR, "libref_cylinder.so"))
    L.refcyl_run.argtypes = [C.c_int, C.c_char_p, C.c_double, C.c_double, C.c_double, dp, dp, dp, dp, dp, C.c_void_p]
    n = 129 * 129
    keep = [None if x is None else np.ascontiguousarray(x, dtype=np.float64).ravel() for x in (rho, ux, uy)]
    r, f, g = np.empty(n), np.empty(n * 9), np.zeros(n, np.int32)
    cwd = os.getcwd()
    try:
        rc = L.refcyl_run(steps, _scratch().encode(), omega, ce, conc_wall, *[None if x is None else _p(x) for x in keep],
                          _p(r), _p(f), g.ctypes.data)
    finally:
        os.chdir(cwd)
    assert rc == 0, rc
    return dict(nx=129, ny=129, rho=r, f=f.reshape(-1, 9), geometry=g)


# binary_shape.cpp:31,37-48: aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi, force_x, force_y
PARAMS_SHAPE = [0.04, 0.04, 1.0, -0.05, 1.0, 1.0, 1.0, 0.0, 0.0]


class OracleShape:
    """CurvedSolid/binary_shape.cpp restated (oracle/lbm_oracle.c, lbobs_*); node i = iY*nx+iX."""

    def __init__(self, nx, ny, mask, params=PARAMS_SHAPE):
        self.L = oracle()
        self.nx, self.ny, self.n = nx, ny, nx * ny
        self.mask = np.ascontiguousarray(mask, dtype=np.int32).ravel()
        p = np.array(params, dtype=np.float64)
        self.h = self.L.lbobs_create(nx, ny, _p(p), self.mask.ctypes.data)
        if not self.h:
            raise ValueError("geometry rejected (a bounce-back node without a wall normal towards fluid)")

    def close(self):
        if self.h:
            self.L.lbobs_destroy(self.h)
            self.h = None

    def init(self, rho, phi, ux, uy):
        a = [np.ascontiguousarray(x, dtype=np.float64).ravel() for x in (rho, phi, ux, uy)]
        self.L.lbobs_init(self.h, *[_p(x) for x in a])

    def step(self, n):
        self.L.lbobs_step(self.h, n)

    def fields(self):
        out = [np.empty(self.n) for _ in range(4)]
        self.L.lbobs_get_fields(self.h, *[_p(a) for a in out])
        return dict(zip(("phase", "rho", "ux", "uy"), out))

    def pops(self):
        f, g = np.empty(self.n * 9), np.empty(self.n * 9)
        self.L.lbobs_get_pops(self.h, _p(f), _p(g))
        return f.reshape(-1, 9), g.reshape(-1, 9)


def shape_reference_setup(nx=301, ny=301, radius=50, radius_droplet=10):
    """binary_shape.cpp:439-520 through the oracle: mask, rho, phi, ux, uy."""
    n = nx * ny
    mask = np.zeros(n, np.int32)
    rho, phi, ux, uy = (np.empty(n) for _ in range(4))
    oracle().lbobs_reference_setup(nx, ny, radius, radius_droplet, mask.ctypes.data, _p(rho), _p(phi), _p(ux), _p(uy))
    return mask, rho, phi, ux, uy


def shape_perturbed_fields(mask, rho, phi, amp=0.05, speed=0.02):
    """Hashed perturbation of the reference initial state on the fluid nodes plus a hashed velocity everywhere."""
    n = mask.size
    fl = mask == 1
    rho2 = np.where(fl, rho * (1.0 + amp * (hashed_uniform(n, 4) - 0.5)), rho)
    phi2 = np.where(fl, phi * (1.0 - amp * hashed_uniform(n, 5)), phi)
    ux = (hashed_uniform(n, 6) - 0.5) * (2 * speed)
    uy = (hashed_uniform(n, 7) - 0.5) * (2 * speed)
    return rho2, phi2, ux, uy


def shape_random_case(nx, ny, seed):
    """Several solid discs (one straddling the periodic seam in x is not allowed: bounce-back nodes must stay off
    the array edge) with the reference's rule for bounce-back nodes, a perturbed two-phase state and a velocity field."""
    rng = np.random.default_rng(seed)
    iy, ix = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    solid = np.zeros((ny, nx), bool)
    for _ in range(3):
        r = int(rng.integers(4, max(5, min(nx, ny) // 6)))
        cx, cy = int(rng.integers(r + 3, nx - r - 3)), int(rng.integers(r + 3, ny - r - 3))
        solid |= (ix - cx) ** 2 + (iy - cy) ** 2 < r * r
    mask = np.where(solid, -1, 1).astype(np.int32)
    near = np.zeros(mask.shape, bool)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dx or dy:
                near |= np.roll(np.roll(mask, -dy, 0), -dx, 1) == 1
    mask[(mask == -1) & near] = 0
    fl = mask == 1
    phi = np.where(fl, np.where(rng.uniform(size=(ny, nx)) < 0.4, 1.0, -1.0) * rng.uniform(0.8, 1.0, (ny, nx)), 1.5)
    rho = np.where(fl, rng.uniform(0.95, 1.05, (ny, nx)), 0.5)
    ux, uy = rng.uniform(-0.02, 0.02, (ny, nx)), rng.uniform(-0.02, 0.02, (ny, nx))
    return mask.ravel(), rho.ravel(), phi.ravel(), ux.ravel(), uy.ravel()


def ref_shape_run(steps, params=None, radius_droplet=0, phi=None, rho=None, ux=None, uy=None, outdir=None):
    """The unmodified binary_shape.cpp (301x301, radius 50) for `steps` loop iterations."""
    L = C.CDLL(os.path.join(REF_DIR, "libref_shape.so"))
    L.refshape_run.argtypes = [C.c_int, C.c_char_p, dp, C.c_int] + [dp] * 4 + [dp] * 6 + [C.c_void_p]
    n = 301 * 301
    keep = [None if x is None else np.ascontiguousarray(x, dtype=np.float64).ravel() for x in (phi, rho, ux, uy)]
    p = None if params is None else np.array(params[:7], dtype=np.float64)
    o = [np.empty(n) for _ in range(4)] + [np.empty(n * 9), np.empty(n * 9)]
    geom = np.zeros(n, np.int32)
    cwd = os.getcwd()
    try:
        rc = L.refshape_run(steps, (outdir or _scratch()).encode(), None if p is None else _p(p), radius_droplet,
                            *[None if x is None else _p(x) for x in keep], *[_p(x) for x in o], geom.ctypes.data)
    finally:
        os.chdir(cwd)
    assert rc == 0, rc
    return dict(nx=301, ny=301, phase=o[0], rho=o[1], ux=o[2], uy=o[3], f=o[4].reshape(-1, 9), g=o[5].reshape(-1, 9), geometry=geom)


def _scratch():
    return tempfile.mkdtemp(prefix="lbmref_")


def ref_shanchen_run(last_step, shared_exp=True, outdir=None):
    """Run the unmodified shanchen.cpp through step `last_step`; returns rho,u1,u2,f(next)."""
    L = C.CDLL(os.path.join(REF_DIR, "libref_shanchen%s.so" % ("_sharedexp" if shared_exp else "")))
    L.ref_shanchen_run.argtypes = [C.c_int, C.c_char_p, dp, dp, dp, dp]
    nx, ny = L.ref_shanchen_nx(), L.ref_shanchen_ny()
    n = nx * ny
    rho, u1, u2, f = np.empty(n), np.empty(n), np.empty(n), np.empty(n * 9)
    rc = L.ref_shanchen_run(last_step, (outdir or _scratch()).encode(), _p(rho), _p(u1), _p(u2), _p(f))
    assert rc in (0, 1), rc
    return dict(nx=nx, ny=ny, rho=rho, u1=u1, u2=u2, f=f.reshape(-1, 9))


def ref_binary_run(last_step, noutput=0, outdir=None):
    L = C.CDLL(os.path.join(REF_DIR, "libref_binary.so"))
    L.ref_binary_run.argtypes = [C.c_int, C.c_int, C.c_char_p] + [dp] * 6
    nx, ny = L.ref_binary_nx(), L.ref_binary_ny()
    n = nx * ny
    ph, rho, ux, uy = (np.empty(n) for _ in range(4))
    f, g = np.empty(n * 9), np.empty(n * 9)
    rc = L.ref_binary_run(last_step, noutput, (outdir or _scratch()).encode(), _p(ph), _p(rho), _p(ux), _p(uy), _p(f), _p(g))
    assert rc == 0, rc
    return dict(nx=nx, ny=ny, phase=ph, rho=rho, ux=ux, uy=uy, f=f.reshape(-1, 9), g=g.reshape(-1, 9))


def ref_mrt_run(variant, arg1=10, arg2=10, outdir=None):
    name = "mrt" if variant == 0 else "another_mrt"
    L = C.CDLL(os.path.join(REF_DIR, "libref_%s.so" % name))
    run = getattr(L, "ref_%s_run" % name)
    run.argtypes = [C.c_int, C.c_int, C.c_char_p, dp, dp, dp, dp]
    nx, ny = getattr(L, "ref_%s_nx" % name)(), getattr(L, "ref_%s_ny" % name)()
    nsteps = getattr(L, "ref_%s_nsteps" % name)()
    n = nx * ny
    rho, ux, uy, f = np.empty(n), np.empty(n), np.empty(n), np.empty(n * 9)
    rc = run(arg1, arg2, (outdir or _scratch()).encode(), _p(rho), _p(ux), _p(uy), _p(f))
    assert rc == 0, rc
    return dict(nx=nx, ny=ny, nsteps=nsteps, rho=rho, ux=ux, uy=uy, f=f.reshape(-1, 9))


def ndiff(a, b):
    """number of entries whose values differ (NaN == NaN; -0.0 == 0.0)"""
    a, b = np.asarray(a), np.asarray(b)
    return int(np.count_nonzero(~((a == b) | (np.isnan(a) & np.isnan(b)))))
