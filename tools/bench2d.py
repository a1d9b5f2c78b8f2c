#!/usr/bin/env python
"""MLUPS of the D2Q9 programs on large synthetic grids (BASELINE.json configs[0..2] scaled to 8192x8192,
SURVEY.md section 8d "synthetic inputs").  Not the headline bench (that is bench.py); prints one JSON line
per model with the HBM roofline fraction (144 B per update for one population set, 288 B for two)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmannmethod_b200 import api  # noqa: E402


def timed(ctx, steps, warm=5):
    import torch
    ctx.step(warm)
    ctx.sync()
    s = torch.cuda.ExternalStream(ctx.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    ctx.step(steps)
    e1.record(s)
    ctx.sync()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


# static count of double-precision instructions (DADD, DMUL, DFMA, DSETP) in the SASS of k_mrt_step_aa<variant> (the default, in-place form), one
# node per thread, straight-line code: `cuobjdump -sass liblbm_b200.so` (re-count after touching the kernel)
MRT_DP_INSTR = {"gs": 340, "hermite": 368}


def fp64_peak():
    """tools/fp64bench: sustained rate of unfused DADD/DMUL on this GPU (G thread-instructions/s), or None."""
    import subprocess
    exe = os.path.join(ROOT, "tools", "fp64bench")
    if not os.path.exists(exe):
        return None
    try:
        r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        return None


def measure(n=8192, steps=20, alternatives=True, peak=None):
    """Every D2Q9 model on an n x n synthetic grid; returns one dict per model (MLUPS, bytes per update, fraction of
    the measured HBM peak; the MRT models additionally against the measured unfused-FP64 instruction rate)."""
    if peak is None:
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
    res = []

    def add(model, ms, nodes, bytes_per_update, extra_bytes=0.0, **more):
        gbs = (nodes * bytes_per_update + extra_bytes) / ms / 1e6
        d = {"model": model, "grid": [n, n], "ms_per_step": ms, "mlups": nodes / ms / 1e3, "bytes_per_update": bytes_per_update,
             "achieved_gbs": gbs, "frac_of_measured_hbm": gbs / peak}
        d.update(more)
        res.append(d)

    # Shan-Chen: the reference droplet tiled with period 128
    rho = np.tile(api.shanchen_initial_rho().reshape(128, 128), (n // 128, n // 128)).ravel()
    for variant, label in ((0, "shanchen_d2q9"), (1, "shanchen_d2q9 (two kernels)")):
        if variant and not alternatives:
            continue
        cfg = api.shanchen_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("rho", rho)
        ctx.init()
        add(label, timed(ctx, steps), n * n, 144)
        ctx.close()
    del rho
    ph = np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel()
    for variant, label in ((0, "binary_d2q9"), (1, "binary_d2q9 (phase pass + collide)")):
        if variant and not alternatives:
            continue
        cfg = api.binary2d_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("phase", ph)
        ctx.init()
        add(label, timed(ctx, steps), n * n, 288)
        ctx.close()
    del ph
    f64 = fp64_peak()
    for variant, om, ob, kv in (("gs", 1.2, 0.8, 0), ("hermite", 0.9, 0.1, 0), ("gs", 1.2, 0.8, 1)):
        if kv and not alternatives:
            continue
        cfg = api.mrt_config(variant, n, n, om, ob)
        cfg.kernel_variant = kv
        ctx = api.Context(cfg)
        x = np.arange(n)
        fa = 0.002 if variant == "gs" else 0.0
        ux = np.repeat(1e-3 * np.cos(x / n * 2 * np.pi) - fa / 2 * np.cos(x / n * 2 * np.pi), n)
        ctx.set_field("rho", np.ones(n * n))
        ctx.set_field("ux", ux)
        ctx.set_field("uy", np.zeros(n * n))
        ctx.init()
        ms = timed(ctx, steps)
        more = {}
        if f64 and f64.get("fp64_unfused_ginstr_per_s"):
            rate = n * n / ms / 1e6 * MRT_DP_INSTR[variant]          # G double-precision thread-instructions per second
            more = {"dp_instr_per_node": MRT_DP_INSTR[variant], "dp_ginstr_per_s": rate,
                    "fp64_peak_ginstr_per_s": f64["fp64_unfused_ginstr_per_s"], "frac_of_measured_fp64": rate / f64["fp64_unfused_ginstr_per_s"]}
        add("mrt_%s_d2q9" % variant + (" (two population arrays, push streaming)" if kv else ""), ms, n * n, 144, **more)
        ctx.close()
    # TRT diffusion: the reference disc scaled to the grid (radius 40/129 of the side); updates = fluid nodes;
    # 9 reads + 9 writes + rho write + ux, uy reads = 168 B per fluid node, + 2 B link word per node of the array
    mask, rho, ux, uy = api.trt_cylinder_setup(n, n, radius=int(n * 40 / 129))
    ctx = api.Context(api.trt_config(n, n))
    ctx.set_geometry(mask)
    ctx.set_field("rho", rho)
    ctx.set_field("ux", ux)
    ctx.set_field("uy", uy)
    ctx.init()
    fluid = int((mask == 1).sum())
    add("trt_diffusion_d2q9", timed(ctx, steps), fluid, 168, extra_bytes=n * n * 2.0, fluid_nodes=fluid)
    ctx.close()
    # curved solid: the reference disc and droplets scaled to the grid
    mask, rho, phi, ux, uy = api.shape_setup(n, n, radius=int(n * 50 / 301), radius_droplet=int(n * 10 / 301))
    fluid = int((mask == 1).sum())
    ctx = api.Context(api.shape_config(n, n))
    ctx.set_geometry(mask)
    for name, a in (("rho", rho), ("phase", phi), ("ux", ux), ("uy", uy)):
        ctx.set_field(name, a)
    ctx.init()
    add("binary_shape_d2q9", timed(ctx, steps), fluid, 288, fluid_nodes=fluid)
    ctx.close()
    return res


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    for d in measure(n):
        print(json.dumps(d))


if __name__ == "__main__":
    main()
