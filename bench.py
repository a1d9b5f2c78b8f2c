#!/usr/bin/env python
"""bench.py -- MLUPS of the D3Q19 binary-liquid collide-stream on B200, with roofline, CPU baseline,
end-to-end figure and clocks, one JSON line per run (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload quarter|full] [--impl ours|reference]

Workload (default `quarter`): the QuarterChannel program (Microchannel3D/QuarterChannel/main.cpp: parameters
:48-62, bubble initial condition :83-100) scaled to 512 x 512 x (512 per GPU) nodes, z-slab per GPU --
BASELINE.json configs[4], the weak-scaling case; `full` is FullChannel at its own 92x52x1500 (configs[3]).
A "step" is one lattice update of the whole grid.  Under torchrun every rank owns one z-slab.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_ALG = 608.0  # algorithmic bytes per lattice update: 2 sets x 19 populations x 8 B, read once + written once


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="quarter", choices=["quarter", "full"])
    ap.add_argument("--nx", type=int, default=512)
    ap.add_argument("--ny", type=int, default=512)
    ap.add_argument("--nz-per-gpu", type=int, default=512)
    ap.add_argument("--variant", type=int, default=0, help="kernel_variant (0 default: one-pass step; 5 two-pass with interior + rim kernels; 1 two-pass, plain kernel; 2 / 6 in-place streaming on one population array, one pass / two passes)")
    ap.add_argument("--e2e-steps", type=int, default=100, help="lattice steps per end-to-end segment (QuarterChannel NUMOUTPUT)")
    ap.add_argument("--precision", default="f64", choices=["f64", "f32"], help="f32: liblbm_b200_f32.so, populations stored as float (not the parity build; the headline is f64)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the 2-D models (N = 1) / the FullChannel strong-scaling point (N > 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the bit-equality check of the multi-GPU path (N > 1 only)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------ reference on the host cores
def ref_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "ref_quarter")
    return p if os.path.exists(p) else None


def run_reference_cpu(nx, ny, nz, warm, steps, ranks, width=6):
    """The unmodified reference Lattice/Solver (oracle/_ref/ref_quarter) on `ranks` forked mini-MPI ranks."""
    env = dict(os.environ, LBM_MINIMPI_RANKS=str(ranks))
    out = subprocess.run([ref_binary(), str(nx), str(ny), str(nz), str(warm), str(steps), str(width)], env=env,
                         capture_output=True, text=True, timeout=1200)
    if out.returncode != 0:
        raise RuntimeError("reference run failed: " + out.stderr[-400:])
    return json.loads(out.stdout.strip().splitlines()[-1])


def host_ranks():
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    return max(1, min(n, 32))   # the reference's own job script asks for 32 ranks (FullChannel/binary.pbs:6)


def cpu_sample_grid(ranks):
    # a bounded sample of the workload: same program, same parameters, 128x128 cross-section, >= 8 planes per rank
    return 128, 128, max(64, 8 * ranks)


def cpu_baseline(seconds):
    """Reference (or, without oracle/_ref, the C port) timed on the host for about `seconds`."""
    if ref_binary():
        ranks = host_ranks()
        nx, ny, nz = cpu_sample_grid(ranks)
        probe = run_reference_cpu(nx, ny, nz, 0, 2, ranks)
        steps = max(3, int(seconds / max(probe["seconds"] / 2, 1e-3)))
        r = run_reference_cpu(nx, ny, nz, 1, steps, ranks)
        # the serial build as well (one rank, a few seconds): the north star asks for both
        serial = None
        try:
            p1 = run_reference_cpu(nx, ny, 32, 0, 1, 1)
            s1 = run_reference_cpu(nx, ny, 32, 1, max(2, int(4.0 / max(p1["seconds"], 1e-3))), 1)
            serial = {"value": s1["mlups"], "cores": 1, "sample": "%dx%dx32, %.1f s" % (nx, ny, s1["seconds"])}
        except Exception:
            pass
        return {"value": r["mlups"], "unit": "MLUPS", "cores": ranks, "kind": "reference",
                "sample": "unmodified Microchannel3D/QuarterChannel Lattice/Solver on %d mini-MPI z-slab ranks, %dx%dx%d, %d steps, %.1f s"
                          % (ranks, nx, ny, nz, steps, r["seconds"]), "serial": serial}
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    nx, ny, nz = 64, 64, 64
    o = ol.Oracle3D(nx, ny, nz, ol.PARAMS_QUARTER)
    o.set_fields(*ol.bubble_fields(nx, ny, nz, 6))
    o.init_single()
    o.step_single(1)
    t0 = time.time()
    steps = 0
    while time.time() - t0 < seconds:
        o.step_single(2)
        steps += 2
    dt = time.time() - t0
    return {"value": nx * ny * nz * steps / dt / 1e6, "unit": "MLUPS", "cores": 1, "kind": "port",
            "sample": "oracle/lbm_oracle.c, %dx%dx%d, %d steps, %.1f s" % (nx, ny, nz, steps, dt)}


def reference_arm(args, rank):
    if rank != 0:
        return
    ranks = host_ranks()
    nx, ny, nz = cpu_sample_grid(ranks)
    # `config` is the grid the reference's CPU code actually runs here: a bounded sample of the job our arm runs
    # for this --gpus (same program, same parameters, smaller cross-section and fewer planes); the job is named in `sample_of`
    world = max(1, args.gpus)
    if args.workload == "full":
        job = "Microchannel3D/FullChannel 92x52x1500 (BASELINE.json configs[3]), z-slabs over %d GPU(s)" % world
    else:
        job = "Microchannel3D/QuarterChannel scaled to %dx%dx%d (%d planes per GPU; BASELINE.json configs[4])" % (
            args.nx, args.ny, args.nz_per_gpu * world, args.nz_per_gpu)
    cfg = {"workload": "Microchannel3D/QuarterChannel Lattice/Solver (unmodified reference) on %dx%dx%d, %d host ranks (mini-MPI z-slabs)" % (nx, ny, nz, ranks),
           "nx": nx, "ny": ny, "nz": nz, "nodes": nx * ny * nz, "precision": "fp64", "sample_of": job, "host_ranks": ranks}
    if ref_binary():
        r = run_reference_cpu(nx, ny, nz, args.warmup, args.steps, ranks)
        kind, val, sec = "reference", r["mlups"], r["seconds"]
    else:
        cb = cpu_baseline(max(5.0, args.steps * 0.2))
        kind, val, sec, ranks = cb["kind"], cb["value"], None, cb["cores"]
    line = {"impl": "reference", "metric": "MLUPS (aggregate), D3Q19 binary-liquid collide-stream, fp64", "value": val, "unit": "MLUPS",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": (sec * 1e3 / args.steps) if sec else None, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": val, "unit": "MLUPS", "cores": ranks, "kind": kind,
                             "sample": "%dx%dx%d, %d steps" % (nx, ny, nz, args.steps)},
            "e2e": {"value": val, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks
class ClockSampler:
    def __init__(self, device_index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr:
            self._stop.set()
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------ our arm
def slab_initial_fields(api, nx, ny, nz_global, z_begin, nz_local, width, np):
    """main.cpp:83-100 for one slab, built plane-wise to stay light on host memory."""
    x = np.arange(nx)[None, :] <= nx - width - 1
    y = np.arange(ny)[:, None] <= ny - width - 1
    inxy = (x & y)
    phase = np.ones((nz_local, ny, nx))
    zlo, zhi = (nz_global - 1) // 3, 2 * (nz_global - 1) // 3
    for zl in range(nz_local):
        z = z_begin + zl
        if zlo <= z <= zhi:
            phase[zl][inxy] = -1.0
    return phase.reshape(-1)


def bind_to_gpu_numa_node(device_index):
    """Pins this process to the CPU cores NVML reports as local to its GPU, so that the pinned host buffers of the
    end-to-end leg are first-touched on the GPU's own socket (torchrun does not bind ranks)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass      # placement is an optimisation of the e2e figure only


class quiet_stdout:
    """Sends file descriptor 1 to stderr while communicators come up: NCCL prints its version banner on stdout at the
    first communicator, and stdout has to stay the one JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def multi_gpu_parity(api, dist, torch, np, args, rank, world, local_rank):
    """Bit-equality check of the path the timed region runs at N > 1 -- lbm_comm_attach, then lbm_init / lbm_step with
    the NCCL halo exchange overlapped with the collision (Slab3D::step_overlapped) -- against ONE slab holding the whole
    grid on rank 0's GPU (the reference's own property: 1 rank == N ranks bit for bit, solver.hpp:893-928,
    lattice.hpp:809-840).  Small grid, 12 planes per rank so that the overlapped branch is the one taken."""
    nx, ny, nzl, steps, width = 35, 18, 12, 9, 4
    nz = nzl * world
    params = (0.04, 0.04, 1.0, 0.03, 0.5, 2.0, 2.5, 0.7)     # a, k, gamma, wall gradient, rho_wall, tau_phi, tau_liq, tau_gas
    force = (1e-5, 0.0, 3e-5)

    def make(slab_rank, slab_count):
        c = api.microchannel_config("quarter", nx, ny, nz, device=local_rank, slab_rank=slab_rank, slab_count=slab_count)
        c.aconst, c.kconst, c.gammaconst, c.wall_gradient, c.rho_wall, c.tau_phi, c.tau_liq, c.tau_gas = params
        c.force[0], c.force[1], c.force[2] = force
        c.kernel_variant = args.variant
        return api.Context(c, args.precision)

    rng = np.random.default_rng(11)
    n = nx * ny * nz
    x = np.arange(nx)[None, None, :] <= nx - width - 1
    y = np.arange(ny)[None, :, None] <= ny - width - 1
    z = np.arange(nz)[:, None, None]
    ph = np.where(x & y & (z >= (nz - 1) // 3) & (z <= 2 * (nz - 1) // 3), -1.0, 1.0).reshape(-1)
    flds = [1 + 0.05 * rng.standard_normal(n), np.clip(ph + 0.3 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]
    names = ("rho", "phase", "ux", "uy", "uz")

    ctx = make(rank, world)
    with quiet_stdout():
        uid = [api.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_attach(uid[0])
    z0, nl = ctx.slab_extent()
    pl = nx * ny
    for name, a in zip(names, flds):
        ctx.set_field(name, a[z0 * pl:(z0 + nl) * pl])
    ctx.init()
    ctx.step(steps)
    mine = [ctx.get_field(k) for k in names] + [ctx.get_populations(0).ravel(), ctx.get_populations(1).ravel()]
    fb = ctx.phase_fallbacks
    ctx.close()
    parts = [None] * world if rank == 0 else None
    dist.gather_object(mine, parts, dst=0)
    res = None
    if rank == 0:
        one = make(0, 1)
        for name, a in zip(names, flds):
            one.set_field(name, a)
        one.init()
        one.step(steps)
        want = [one.get_field(k) for k in names] + [one.get_populations(0).ravel(), one.get_populations(1).ravel()]
        one.close()
        ndiff = 0
        for i, w in enumerate(want):
            got = np.concatenate([p[i] for p in parts])
            ndiff += int(np.count_nonzero(~((got == w) | (np.isnan(got) & np.isnan(w)))))
        res = {"grid": "%dx%dx%d" % (nx, ny, nz), "slabs": world, "steps": steps, "path": "lbm_comm_attach + lbm_init + lbm_step (NCCL%s)" % (", five directions per face after every in-place sub-step" if args.variant in (2, 6) else ", overlapped"),
               "compared": "rho, phase, ux, uy, uz, f, g against one slab on one GPU", "values": int(sum(w.size for w in want)),
               "ndiff": ndiff, "phase_fallbacks": fb}
    flag = torch.tensor([0 if res is None else res["ndiff"]], dtype=torch.int64, device="cuda")
    dist.broadcast(flag, src=0)
    if int(flag.item()) != 0:
        if rank == 0:
            sys.stderr.write("bench.py: multi-GPU parity check FAILED: %s\n" % json.dumps(res))
        sys.exit(3)
    return res


def strong_scaling_point(api, dist, torch, args, rank, world, local_rank, peak):
    """FullChannel 92x52x1500 (BASELINE.json configs[3]) over the `world` GPUs of this run: the fixed-size job of the
    reference's own 32-rank script (binary.pbs:58), timed like the headline (CUDA events, max over ranks)."""
    nx, ny, nz, steps, warm = 92, 52, 1500, 200, 20
    cfg = api.microchannel_config("full", nx, ny, nz, device=local_rank, slab_rank=rank, slab_count=world)
    cfg.kernel_variant = args.variant
    ctx = api.Context(cfg, args.precision)
    with quiet_stdout():
        uid = [api.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_attach(uid[0])
    z0, nzl = ctx.slab_extent()
    import numpy as np
    flds = {"rho": np.ones(nx * ny * nzl), "phase": slab_initial_fields(api, nx, ny, nz, z0, nzl, 8, np)}
    for k in ("ux", "uy", "uz"):
        flds[k] = np.zeros(nx * ny * nzl)
    for k, a in flds.items():
        ctx.set_field(k, a)
    ctx.init()
    ctx.step(warm)
    stream = torch.cuda.ExternalStream(ctx.stream)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.sync(); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev0.record(stream)
    ctx.step(steps)
    ev1.record(stream)
    ctx.sync(); torch.cuda.synchronize(); dist.barrier()
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    fb = ctx.phase_fallbacks
    ctx.close()
    mlups = nx * ny * nz / ms / 1e3
    return {"workload": "Microchannel3D/FullChannel 92x52x1500 (BASELINE.json configs[3]), z-slabs over %d GPUs" % world,
            "scaling": "strong", "steps": steps, "warmup": warm, "ms_per_step": ms, "value": mlups, "unit": "MLUPS",
            "frac_of_measured_hbm_per_gpu": mlups * 1e6 * B_ALG / world / 1e9 / peak, "phase_fallbacks": fb}


def ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    from latticeboltzmannmethod_b200 import api

    torch.cuda.set_device(local_rank)
    if world > 1:      # at N=1 the whole host stays available to the CPU baseline timed in the same run
        bind_to_gpu_numa_node(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        with quiet_stdout():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()

    parity = None
    if world > 1 and not args.no_parity:
        parity = multi_gpu_parity(api, dist, torch, np, args, rank, world, local_rank)

    if args.workload == "full":
        nx, ny, nz, width, which = 92, 52, 1500, 8, "full"
        wl = "Microchannel3D/FullChannel 92x52x1500 (BASELINE.json configs[3]), z-slabs over %d GPU(s)" % world
        scaling = "strong"
    else:
        nx, ny, nz, which = args.nx, args.ny, args.nz_per_gpu * world, "quarter"
        width = max(1, round(nx * 6 / 27))
        wl = "Microchannel3D/QuarterChannel scaled to %dx%dx%d (%d planes per GPU; BASELINE.json configs[4])" % (nx, ny, nz, args.nz_per_gpu)
        scaling = "weak"
    cfg = api.microchannel_config(which, nx, ny, nz, device=local_rank, slab_rank=rank, slab_count=world)
    cfg.kernel_variant = args.variant
    ctx = api.Context(cfg, args.precision)
    b_alg = B_ALG if args.precision == "f64" else B_ALG // 2     # populations stored as float: half the bytes per update
    z_begin, nz_local = ctx.slab_extent()
    nloc = ctx.local_nodes
    if world > 1:
        with quiet_stdout():
            uid = [api.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            ctx.comm_attach(uid[0])

    # pinned host buffers: initial fields in, result fields out (the e2e leg copies through them every segment)
    L = api.load_library(args.precision)
    import ctypes as C
    host = {}
    for name in ("rho", "phase", "ux", "uy", "uz"):
        ptr = L.lbm_host_alloc(nloc * 8)
        if not ptr:
            raise RuntimeError("pinned host allocation failed")
        host[name] = (ptr, np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(nloc,)))
    host["rho"][1][:] = 1.0
    host["phase"][1][:] = slab_initial_fields(api, nx, ny, nz, z_begin, nz_local, width, np)
    for k in ("ux", "uy", "uz"):
        host[k][1][:] = 0.0

    def upload_and_init():
        for name in ("rho", "phase", "ux", "uy", "uz"):
            ctx.set_field(name, host[name][0])
        ctx.init()

    def barrier():
        ctx.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    upload_and_init()
    stream = torch.cuda.ExternalStream(ctx.stream)
    ctx.step(args.warmup)
    barrier()

    # ---- device-resident timing: K steps, CUDA events on the stream the kernels run on ----
    # the sampler thread (NVML start-up takes tens of ms, different on every rank) is running before the barrier, so
    # that all ranks record their first event and queue their first step together: a slab cannot run ahead of its
    # neighbours, and any start skew would be billed to the early ranks as waiting time
    clocks = ClockSampler(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ctx.launch_count
    clocks.start()
    barrier()
    ev0.record(stream)
    ctx.step(args.steps)
    ev1.record(stream)
    barrier()
    clk = clocks.stop()
    ms = ev0.elapsed_time(ev1)
    launches = ctx.launch_count - launches0
    fallbacks = ctx.phase_fallbacks
    device_bytes = ctx.device_bytes
    # per-kernel split: a second, untimed-for-the-headline pass with one CUDA event pair around every launch
    psteps = min(args.steps, 20)
    ctx.profile(True)
    ctx.step(psteps)
    barrier()
    prof = ctx.profile_report()
    ctx.profile(False)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    total_nodes = nx * ny * nz
    value = total_nodes * args.steps / (ms_max * 1e-3) / 1e6

    # roofline: the whole step first (what the MLUPS figure means in bytes), then the dominant kernel
    peak, which_peak = peaks()
    step_achieved = value * 1e6 * b_alg / world / 1e9          # GB/s per GPU on the algorithmic-bytes basis
    roof = {"bound": "hbm", "peak": peak, "unit": "GB/s", "peak_source": which_peak,
            "step_achieved": step_achieved, "step_frac": step_achieved / peak,
            "algorithmic_bytes_per_lattice_update": b_alg}
    if prof:
        name = max(prof, key=lambda k: prof[k]["ms"])
        launches_per_step = prof[name]["launches"] / psteps
        per_launch_ms = prof[name]["ms"] / prof[name]["launches"]
        # nodes one launch of the dominant kernel updates: the collide kernels take every bulk node except the rim
        # rows iY = 1, NY-2 (those go to k_collide_rim) of nz_local + 2 planes (the two redundantly collided ghost
        # planes are work the slab design adds, not lattice updates: they are NOT counted)
        launch_nodes = (nx - 2) * (ny - 4) * nz_local / launches_per_step if name in ("k_collide_lag", "k_collide_interior", "k_step_one", "k_collide_core") else nloc / launches_per_step
        achieved = b_alg * launch_nodes / (per_launch_ms * 1e-3) / 1e9
        traffic = step_traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp) and args.precision == "f64" and (nx, ny, nz_local) == (512, 512, 512):
            tj = json.load(open(tp)).get("kernels", {})
            if name in tj:      # only a capture of this very kernel counts
                traffic = tj[name]["dram_bytes_per_launch"]
            if all(k in tj for k in prof):
                step_traffic = sum(tj[k]["dram_bytes_per_launch"] * prof[k]["launches"] / psteps for k in prof)
        roof.update({"kernel": name, "achieved": achieved, "frac": achieved / peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": b_alg * launch_nodes, "launches_per_step": launches_per_step,
                     "avg_launch_ms": per_launch_ms, "share_of_step": prof[name]["ms"] / sum(v["ms"] for v in prof.values()),
                     "step_traffic": step_traffic, "step_traffic_per_lattice_update": step_traffic / nloc if step_traffic else None,
                     "kernels": prof, "profiled_steps": psteps})

    # every rank's own view (the step time above is the max over ranks; a slab waits for its neighbours, so
    # one slow GPU sets the pace): kernel time per step, wall per step, median SM clock
    per_rank = None
    if dist is not None:
        mine = torch.tensor([sum(v["ms"] for v in prof.values()) / psteps if prof else 0.0, ms / args.steps,
                             float(clk.get("sm_mhz") or 0)], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [{"kernel_ms_per_step": round(float(a[0]), 4), "ms_per_step": round(float(a[1]), 4), "sm_mhz": int(a[2])} for a in allr]

    # ---- end to end through the C ABI with host buffers: H2D initial fields, init, S steps, D2H result fields ----
    e2e = None
    if not args.no_e2e:
        S = args.e2e_steps
        barrier()
        t0 = time.perf_counter()
        upload_and_init()
        ctx.step(S)
        for name in ("rho", "phase", "ux", "uy", "uz"):
            ctx.get_field_async(name, host[name][0])
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": total_nodes * S / dt / 1e6, "unit": "MLUPS", "h2d_bytes_per_step": 5 * 8 * nloc / S,
               "d2h_bytes_per_step": 5 * 8 * nloc / S, "segment": "lbm_set_field x5 (pinned host) + lbm_init + lbm_step(%d) + lbm_get_field_async x5 + lbm_sync" % S,
               "seconds": dt}

    # ---- the other configurations of BASELINE.json, outside the headline's timed region ----
    secondary = None
    if rank == 0 and world == 1 and not args.no_secondary:
        ctx.close()            # give the 87 GB of the headline run back first
        ctx = None
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench2d
            sc = ClockSampler(local_rank)
            sc.start()
            models = bench2d.measure(8192, 20, alternatives=False, peak=peak)
            secondary = {"what": "the D2Q9 programs on a synthetic 8192x8192 grid (BASELINE.json configs[0..2] + the two widened models), "
                                 "20 steps each after 5 warm-up steps, CUDA events on the library's stream, fp64",
                         "models": models, "clocks": sc.stop()}
        except Exception as e:
            secondary = {"error": str(e)[:300]}
    strong = None
    if world > 1 and args.workload == "quarter" and not args.no_secondary:
        strong = strong_scaling_point(api, dist, torch, args, rank, world, local_rank, peak)

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cb = cpu_baseline(args.cpu_seconds)
        except Exception as e:  # the baseline is reported, never the thing measured
            cb = {"value": None, "unit": "MLUPS", "cores": 0, "kind": "unavailable", "sample": str(e)[:200]}

    if rank == 0:
        line = {"metric": "MLUPS (aggregate), D3Q19 binary-liquid collide-stream, %s" % ("fp64" if args.precision == "f64" else "fp64 arithmetic on fp32-stored populations"),
                "value": value, "unit": "MLUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
                "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64" if args.precision == "f64" else "f64 (populations stored as f32)", "data": "synthetic",
                "config": {"workload": wl, "nx": nx, "ny": ny, "nz": nz, "nodes": total_nodes,
                           "precision": "fp64" if args.precision == "f64" else "fp32 storage of the populations, fp64 arithmetic and fields (not the parity build)",
                           "kernel_variant": args.variant, "l2": "inputs (%.1f GB of populations per GPU) exceed the 126 MB L2" % (nloc * b_alg / 2 / 1e9),
                           "parallelism": "z-slab x%d" % world},
                "mlups_per_gpu": value / world, "roofline": roof, "cpu_baseline": cb, "e2e": e2e, "clocks": clk,
                "gpu_launches": launches, "phase_fallbacks": fallbacks,
                "device_bytes_per_gpu": device_bytes}
        if parity is not None:
            line["parity"] = parity
        if secondary is not None:
            line["secondary"] = secondary
        if strong is not None:
            line["strong"] = strong
        if per_rank:
            line["per_rank"] = per_rank
        print(json.dumps(line), flush=True)
    if ctx is not None:
        ctx.close()
    for ptr, _ in host.values():
        L.lbm_host_free(ptr)
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        sys.stderr.write("bench.py: --gpus %d needs torchrun (one rank per GPU); running on 1 GPU\n" % args.gpus)
    ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
