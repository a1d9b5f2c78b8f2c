"""ctypes binding of include/lbm_b200.h plus the reference programs' parameter sets.

The class and method names follow the C ABI one to one (``Context.step`` is
``lbm_step`` and so on); see the header for which reference routine each call
replaces.  Host arrays are numpy float64 in the reference's own layouts.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_PKG, "liblbm_b200.so")

MODELS = dict(shanchen_d2q9=0, binary_d2q9=1, mrt_gs_d2q9=2, mrt_hermite_d2q9=3, binary_d3q19=4,
              trt_diffusion_d2q9=5, binary_shape_d2q9=6)
FIELDS = dict(rho=0, phase=1, ux=2, uy=3, uz=4)
HALO = dict(S_F_UP=0, S_G_UP=1, S_F_DOWN=2, S_G_DOWN=3, R_F_FROM_BELOW=4, R_G_FROM_BELOW=5, R_F_FROM_ABOVE=6,
            R_G_FROM_ABOVE=7, S_PHI_UP=8, S_PHI_DOWN=9, R_PHI_FROM_BELOW=10, R_PHI_FROM_ABOVE=11)
COMM_ID_BYTES = 128
_STATUS = {0: "LBM_OK", 1: "LBM_EINVAL", 2: "LBM_ENOMEM", 3: "LBM_ECUDA", 4: "LBM_ENCCL", 5: "LBM_ENAN", 6: "LBM_ESTATE"}


class LbmError(RuntimeError):
    def __init__(self, status, text):
        super().__init__("%s: %s" % (_STATUS.get(status, status), text))
        self.status = status


class Config(C.Structure):
    """lbm_config (include/lbm_b200.h)."""
    _fields_ = [("model", C.c_int), ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int),
                ("tau", C.c_double), ("g", C.c_double),
                ("aconst", C.c_double), ("kconst", C.c_double), ("gammaconst", C.c_double),
                ("tau_gas", C.c_double), ("tau_liq", C.c_double), ("tau_phi", C.c_double),
                ("wall_gradient", C.c_double), ("rho_wall", C.c_double), ("force", C.c_double * 3),
                ("omega", C.c_double), ("omega_bulk", C.c_double), ("force_ampl", C.c_double),
                ("ce", C.c_double), ("conc_wall", C.c_double),
                ("device", C.c_int), ("slab_rank", C.c_int), ("slab_count", C.c_int), ("kernel_variant", C.c_int),
                ("trt_magic", C.c_double), ("mrt_rates", C.c_double * 5)]


_LIB_F32 = os.path.join(_PKG, "liblbm_b200_f32.so")
_libs = {}
dp = C.POINTER(C.c_double)


def library_path(precision="f64"):
    return _LIB_F32 if precision == "f32" else _LIB


def load_library(precision="f64"):
    """Loads liblbm_b200.so (precision "f64", the parity build) or liblbm_b200_f32.so ("f32": same ABI, D3Q19 model
    only, populations stored as float); raises (never falls back) when it has not been built."""
    if precision in _libs:
        return _libs[precision]
    path = library_path(precision)
    if not os.path.exists(path):
        raise ImportError("%s is missing: build it with `python -m latticeboltzmannmethod_b200.build` "
                          "(or __graft_entry__.build()); there is no CPU fallback" % os.path.basename(path))
    L = C.CDLL(path)
    vp, sz, ci = C.c_void_p, C.c_size_t, C.c_int
    sig = {
        "lbm_config_default": (ci, [C.POINTER(Config), ci]),
        "lbm_create": (ci, [C.POINTER(Config), C.POINTER(vp)]),
        "lbm_destroy": (None, [vp]),
        "lbm_last_error": (C.c_char_p, [vp]),
        "lbm_slab_extent": (ci, [vp, C.POINTER(ci), C.POINTER(ci)]),
        "lbm_local_nodes": (sz, [vp]),
        "lbm_set_field": (ci, [vp, ci, vp, sz]),
        "lbm_get_field": (ci, [vp, ci, vp, sz]),
        "lbm_get_field_async": (ci, [vp, ci, vp, sz]),
        "lbm_set_geometry": (ci, [vp, vp, sz]),
        "lbm_init": (ci, [vp]),
        "lbm_set_populations": (ci, [vp, ci, vp, sz]),
        "lbm_get_populations": (ci, [vp, ci, vp, sz]),
        "lbm_step": (ci, [vp, ci]),
        "lbm_sync": (ci, [vp]),
        "lbm_diagnostics": (ci, [vp, dp, dp, dp, C.POINTER(ci)]),
        "lbm_get_stream": (vp, [vp]),
        "lbm_set_stream": (ci, [vp, vp]),
        "lbm_launch_count": (C.c_ulonglong, [vp]),
        "lbm_population_bytes": (ci, [vp]),
        "lbm_device_bytes": (C.c_ulonglong, [vp]),
        "lbm_phase_fallbacks": (C.c_ulonglong, [vp]),
        "lbm_guard_check": (C.c_ulonglong, [vp]),
        "lbm_profile": (ci, [vp, ci]),
        "lbm_profile_report": (ci, [vp, C.c_char_p, sz]),
        "lbm_host_alloc": (vp, [sz]),
        "lbm_host_free": (None, [vp]),
        "lbm_comm_unique_id": (ci, [vp]),
        "lbm_comm_attach": (ci, [vp, vp]),
        "lbm_slab_halo": (ci, [vp, ci, C.POINTER(vp), C.POINTER(sz)]),
        "lbm_slab_init_fields": (ci, [vp]),
        "lbm_slab_init_pops": (ci, [vp]),
        "lbm_slab_after_init": (ci, [vp]),
        "lbm_slab_collide": (ci, [vp, ci]),
        "lbm_slab_finish": (ci, [vp]),
        "lbm_population_sums_noalloc": (ci, [vp, vp, ci, sz, sz, vp, vp]),
        "lbm_noalloc_last_error": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError if the library does not export it
        fn.restype, fn.argtypes = res, args
    _libs[precision] = L
    return L


EXPORTS = ["lbm_config_default", "lbm_create", "lbm_destroy", "lbm_last_error", "lbm_slab_extent", "lbm_local_nodes",
           "lbm_set_field", "lbm_get_field", "lbm_get_field_async", "lbm_population_bytes", "lbm_device_bytes", "lbm_phase_fallbacks", "lbm_guard_check", "lbm_set_geometry", "lbm_init", "lbm_set_populations", "lbm_get_populations", "lbm_step",
           "lbm_sync", "lbm_diagnostics", "lbm_get_stream", "lbm_set_stream", "lbm_launch_count", "lbm_profile",
           "lbm_profile_report", "lbm_host_alloc",
           "lbm_host_free", "lbm_comm_unique_id", "lbm_comm_attach", "lbm_slab_halo", "lbm_slab_init_fields",
           "lbm_slab_init_pops", "lbm_slab_after_init", "lbm_slab_collide", "lbm_slab_finish",
           "lbm_population_sums_noalloc", "lbm_noalloc_last_error"]


def default_config(model):
    L = load_library()
    cfg = Config()
    st = L.lbm_config_default(C.byref(cfg), MODELS[model] if isinstance(model, str) else model)
    if st:
        raise LbmError(st, "lbm_config_default")
    return cfg


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if isinstance(a, np.ndarray) else C.c_void_p(a)


class Context:
    """One lbm_ctx: one lattice model on one GPU (for the 3-D model: one z-slab)."""

    def __init__(self, cfg, precision="f64"):
        self.L = load_library(precision)
        self.cfg = cfg
        h = C.c_void_p()
        st = self.L.lbm_create(C.byref(cfg), C.byref(h))
        if st:
            raise LbmError(st, (self.L.lbm_last_error(None) or b"").decode())
        self.h = h
        self.q = 19 if cfg.model == MODELS["binary_d3q19"] else 9

    def _ck(self, st):
        if st:
            raise LbmError(st, (self.L.lbm_last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.lbm_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def local_nodes(self):
        return int(self.L.lbm_local_nodes(self.h))

    def slab_extent(self):
        a, b = C.c_int(), C.c_int()
        self._ck(self.L.lbm_slab_extent(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_field(self, field, host):
        """host: numpy float64 (copied if not contiguous) or a raw host address (e.g. pinned memory)."""
        if isinstance(host, np.ndarray):
            host = np.ascontiguousarray(host, dtype=np.float64).ravel()
            assert host.size == self.local_nodes, (host.size, self.local_nodes)
        self._ck(self.L.lbm_set_field(self.h, FIELDS[field] if isinstance(field, str) else field, _ptr(host), self.local_nodes))

    def set_geometry(self, mask):
        """mask: int array, 1 fluid / 0 boundary / -1 solid, node i = iY*nx+iX (TRT diffusion model only)."""
        mask = np.ascontiguousarray(mask, dtype=np.int32).ravel()
        self._keep = mask
        self._ck(self.L.lbm_set_geometry(self.h, mask.ctypes.data, mask.size))

    def get_field(self, field, out=None):
        if out is None:
            out = np.empty(self.local_nodes, dtype=np.float64)
        self._ck(self.L.lbm_get_field(self.h, FIELDS[field] if isinstance(field, str) else field, _ptr(out), self.local_nodes))
        return out

    def get_field_async(self, field, host_addr):
        """Queues the copy into page-locked host memory at `host_addr`; valid after sync()."""
        self._ck(self.L.lbm_get_field_async(self.h, FIELDS[field] if isinstance(field, str) else field, _ptr(host_addr), self.local_nodes))

    def init(self):
        self._ck(self.L.lbm_init(self.h))

    def step(self, n=1):
        self._ck(self.L.lbm_step(self.h, int(n)))

    def sync(self):
        self._ck(self.L.lbm_sync(self.h))

    def set_populations(self, which, host):
        host = np.ascontiguousarray(host, dtype=np.float64).ravel()
        self._ck(self.L.lbm_set_populations(self.h, which, _ptr(host), host.size))

    def get_populations(self, which):
        out = np.empty(self.local_nodes * self.q, dtype=np.float64)
        self._ck(self.L.lbm_get_populations(self.h, which, _ptr(out), out.size))
        return out.reshape(-1, self.q)

    def diagnostics(self):
        m, pm = C.c_double(), C.c_double()
        mom = (C.c_double * 3)()
        nan = C.c_int()
        self._ck(self.L.lbm_diagnostics(self.h, C.byref(m), mom, C.byref(pm), C.byref(nan)))
        return dict(mass=m.value, momentum=[mom[0], mom[1], mom[2]], phase_mass=pm.value, has_nan=bool(nan.value))

    @property
    def stream(self):
        return self.L.lbm_get_stream(self.h)

    def set_stream(self, cuda_stream):
        self._ck(self.L.lbm_set_stream(self.h, C.c_void_p(cuda_stream)))

    @property
    def launch_count(self):
        return int(self.L.lbm_launch_count(self.h))

    @property
    def population_bytes(self):
        return int(self.L.lbm_population_bytes(self.h))

    @property
    def device_bytes(self):
        """Device memory the context's model holds (populations, fields, scratch)."""
        return int(self.L.lbm_device_bytes(self.h))

    def guard_check(self):
        """Guard bytes overwritten so far (contexts created with LBM_GUARD=1 in the environment; else 0)."""
        return int(self.L.lbm_guard_check(self.h))

    @property
    def phase_fallbacks(self):
        return int(self.L.lbm_phase_fallbacks(self.h))

    def profile(self, enable=True):
        self._ck(self.L.lbm_profile(self.h, int(bool(enable))))

    def profile_report(self):
        import json
        buf = C.create_string_buffer(1 << 16)
        self._ck(self.L.lbm_profile_report(self.h, buf, len(buf)))
        return json.loads(buf.value.decode())

    # ---- z-slab phases ----
    def comm_attach(self, unique_id_bytes):
        buf = C.create_string_buffer(bytes(unique_id_bytes), COMM_ID_BYTES)
        self._ck(self.L.lbm_comm_attach(self.h, buf))

    def slab_halo(self, name):
        """(device address, element count, bytes per element) of one halo message: F/G messages are made of
        population values (8 bytes; 4 in the fp32-storage build), PHI messages of doubles."""
        p, n = C.c_void_p(), C.c_size_t()
        hid = HALO[name] if isinstance(name, str) else name
        self._ck(self.L.lbm_slab_halo(self.h, hid, C.byref(p), C.byref(n)))
        return p.value, n.value, (8 if hid >= HALO["S_PHI_UP"] else self.population_bytes)

    def slab_init_fields(self):
        self._ck(self.L.lbm_slab_init_fields(self.h))

    def slab_init_pops(self):
        self._ck(self.L.lbm_slab_init_pops(self.h))

    def slab_after_init(self):
        self._ck(self.L.lbm_slab_after_init(self.h))

    def slab_collide(self, write_fields):
        self._ck(self.L.lbm_slab_collide(self.h, int(bool(write_fields))))

    def slab_finish(self):
        self._ck(self.L.lbm_slab_finish(self.h))


def population_sums_noalloc(f_dev, g_dev, q, n_nodes, stride_k, cuda_stream, sums6_dev):
    """lbm_population_sums_noalloc: raw device addresses in, six sums written to device memory; allocates nothing."""
    L = load_library()
    st = L.lbm_population_sums_noalloc(C.c_void_p(f_dev), C.c_void_p(g_dev) if g_dev else None, int(q), int(n_nodes),
                                       int(stride_k), C.c_void_p(cuda_stream) if cuda_stream else None, C.c_void_p(sums6_dev))
    if st:
        raise LbmError(st, (L.lbm_noalloc_last_error() or b"").decode())


def comm_unique_id():
    L = load_library()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    st = L.lbm_comm_unique_id(buf)
    if st:
        raise LbmError(st, (L.lbm_last_error(None) or b"").decode())
    return buf.raw


# ---- the reference programs' configurations and initial conditions ----
def shanchen_config(nx=128, ny=128, device=0):
    """ShanChen/shanchen.cpp:11-27."""
    c = default_config("shanchen_d2q9")
    c.nx, c.ny, c.device = nx, ny, device
    return c


def shanchen_initial_rho(nx=128, ny=128, rhol=1.95, rhog=0.15, radius=20):
    """shanchen.cpp:88-95, node i = iY*nx+iX."""
    i = np.arange(nx * ny)
    inside = (i // nx - ny / 2.0) ** 2 + (i % nx - nx / 2.0) ** 2 <= radius * radius
    return np.where(inside, rhol, rhog).astype(np.float64)


def binary2d_config(nx=128, ny=128, device=0):
    """FreeEnergy/binary.cpp:12-41."""
    c = default_config("binary_d2q9")
    c.nx, c.ny, c.device = nx, ny, device
    return c


def binary2d_initial_phase(nx=128, ny=128):
    """binary.cpp:126-132, layout [iX][iY]."""
    ix, iy = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    return np.where((np.abs(ix - nx // 2) <= 20) & (iy < 40) & (iy >= 1), 1.0, -1.0).astype(np.float64).ravel()


def mrt_config(variant="gs", nx=40, ny=3, omega=None, omega_bulk=None, device=0):
    """MRT/mrt.cpp (variant 'gs', argv omega*10, omega_bulk*10) or MRT/another_mrt.cpp ('hermite')."""
    c = default_config("mrt_gs_d2q9" if variant == "gs" else "mrt_hermite_d2q9")
    c.nx, c.ny, c.device = nx, ny, device
    if omega is not None:
        c.omega = omega
    if omega_bulk is not None:
        c.omega_bulk = omega_bulk
    return c


def mrt_initial_fields(variant="gs", nx=40, ny=3, ampl=0.001, force_ampl=0.002):
    """mrt.cpp:100-105 / another_mrt.cpp:92-96, layout [iX][iY].  Uses libm cos with the reference's
    argument association (differs from the time loop's)."""
    import math
    pi = 3.141592653589793238462643383279502884197
    rho, ux, uy = np.ones(nx * ny), np.empty(nx * ny), np.zeros(nx * ny)
    for ix in range(nx):
        if variant == "gs":
            forcex = force_ampl * math.cos(float(ix) / float(nx) * 2.0 * pi)
            u = ampl * math.cos(float(ix) / float(nx) * 2.0 * pi) - forcex / 2.0
        else:
            u = ampl * math.cos(float(ix) / float(nx) * 2 * pi)
        ux[ix * ny:(ix + 1) * ny] = u
    return rho, ux, uy


def trt_config(nx=129, ny=129, omega=None, ce=None, conc_wall=None, device=0):
    """DiffusionCylinder/cylinder.cpp:226-227,33-43."""
    c = default_config("trt_diffusion_d2q9")
    c.nx, c.ny, c.device = nx, ny, device
    if omega is not None:
        c.omega = omega
    if ce is not None:
        c.ce = ce
    if conc_wall is not None:
        c.conc_wall = conc_wall
    return c


def trt_cylinder_setup(nx=129, ny=129, radius=40, conc_wall=1.0):
    """cylinder.cpp:224-296: the disc, its ring of boundary nodes, and the initial concentration
    (0 in the fluid, conc_wall on the ring, -1 in the solid).  Returns mask, rho, ux, uy flat [iY*nx+iX]."""
    iy, ix = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    inside = (ix - (nx - 1) // 2) ** 2 + (iy - (ny - 1) // 2) ** 2 <= radius * radius
    mask = np.where(inside, 1, -1).astype(np.int32)

    def touches(m, value):
        pad = np.pad(m, 1, constant_values=2)
        hit = np.zeros(m.shape, bool)
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                hit |= pad[1 + dy:1 + dy + ny, 1 + dx:1 + dx + nx] == value
        return hit

    mask[(mask == 1) & touches(mask, -1)] = 0          # fluid next to solid becomes a boundary node
    mask[(mask == 0) & ~touches(mask, 1)] = -1         # boundary nodes with no fluid neighbour are solid
    rho = np.where(mask == 0, conc_wall, np.where(mask == -1, -1.0, 0.0)).astype(np.float64)
    return mask.ravel(), rho.ravel(), np.zeros(nx * ny), np.zeros(nx * ny)


def shape_config(nx=301, ny=301, wall_gradient=None, device=0, **kw):
    """CurvedSolid/binary_shape.cpp:17-48; kw: aconst, kconst, gammaconst, tau_gas, tau_liq, tau_phi."""
    c = default_config("binary_shape_d2q9")
    c.nx, c.ny, c.device = nx, ny, device
    if wall_gradient is not None:
        c.wall_gradient = wall_gradient
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def shape_setup(nx=301, ny=301, radius=50, radius_droplet=10, num_droplets=3):
    """binary_shape.cpp:439-520: a solid disc in the centre of a periodic box, the solid nodes that touch
    fluid turned into bounce-back nodes, `num_droplets` droplets (phi = +1) sitting on the disc in a phi = -1
    bath; rho = 1 in the fluid, rho_wall = 0.5 and phi_wall = 1.5 elsewhere.
    Returns mask, rho, phi, ux, uy flat [iY*nx+iX]."""
    import math
    iy, ix = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    solid = (ix - (nx - 1) // 2) ** 2 + (iy - (ny - 1) // 2) ** 2 < radius * radius
    mask = np.where(solid, -1, 1).astype(np.int32)
    near_fluid = np.zeros(mask.shape, bool)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dx or dy:
                near_fluid |= np.roll(np.roll(mask, -dy, 0), -dx, 1) == 1
    mask[(mask == -1) & near_fluid] = 0
    rho = np.where(mask == 1, 1.0, 0.5)
    pi = 4.0 * math.atan(1.0)
    angle = 360 // num_droplets
    drop = np.zeros(mask.shape, bool)
    for d in range(num_droplets):
        cx = (nx - 1) // 2 - math.sin(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius)
        cy = (ny - 1) // 2 + math.cos(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius)
        drop |= ((ix - cx) * (ix - cx) + (iy - cy) * (iy - cy) <= radius_droplet * radius_droplet) & (mask == 1)
    phi = np.where(drop, 1.0, np.where(mask == 1, -1.0, 1.5))
    return mask.ravel(), rho.ravel().astype(np.float64), phi.ravel().astype(np.float64), np.zeros(nx * ny), np.zeros(nx * ny)


def microchannel_config(which="quarter", nx=None, ny=None, nz=None, device=0, slab_rank=0, slab_count=1):
    """Microchannel3D/QuarterChannel/main.cpp:18-62 or FullChannel/main.cpp:19-70."""
    c = default_config("binary_d3q19")
    if which == "full":
        c.nx, c.ny, c.nz, c.tau_phi = 92, 52, 1500, 1.0
    if nx:
        c.nx = nx
    if ny:
        c.ny = ny
    if nz:
        c.nz = nz
    c.device, c.slab_rank, c.slab_count = device, slab_rank, slab_count
    return c


def microchannel_initial_fields(nx, ny, nz, width, z_begin=0, nz_local=None):
    """The bubble of main.cpp:83-100 for global planes [z_begin, z_begin+nz_local): rho, phase, ux, uy, uz."""
    nz_local = nz if nz_local is None else nz_local
    z, y, x = np.meshgrid(np.arange(z_begin, z_begin + nz_local), np.arange(ny), np.arange(nx), indexing="ij")
    gas = (z >= (nz - 1) // 3) & (z <= 2 * (nz - 1) // 3) & (x <= nx - width - 1) & (y <= ny - width - 1)
    n = nx * ny * nz_local
    return np.ones(n), np.where(gas, -1.0, 1.0).astype(np.float64).ravel(), np.zeros(n), np.zeros(n), np.zeros(n)
