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


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    out = []
    # Shan-Chen: the reference droplet tiled with period 128
    rho = np.tile(api.shanchen_initial_rho().reshape(128, 128), (n // 128, n // 128)).ravel()
    for variant, label in ((0, "shanchen_d2q9"), (1, "shanchen_d2q9 (two kernels)")):
        cfg = api.shanchen_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("rho", rho)
        ctx.init()
        ms = timed(ctx, 20)
        out.append((label, ms, 144))
        ctx.close()
    ph = np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel()
    for variant, label in ((0, "binary_d2q9"), (3, "binary_d2q9 (fused tile 4x32, 96 regs)")):
        cfg = api.binary2d_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("phase", ph)
        ctx.init()
        ms = timed(ctx, 20)
        out.append((label, ms, 288))
        ctx.close()
    for variant, om, ob in (("gs", 1.2, 0.8), ("hermite", 0.9, 0.1)):
        ctx = api.Context(api.mrt_config(variant, n, n, om, ob))
        x = np.arange(n)
        fa = 0.002 if variant == "gs" else 0.0
        ux = np.repeat(1e-3 * np.cos(x / n * 2 * np.pi) - fa / 2 * np.cos(x / n * 2 * np.pi), n)
        ctx.set_field("rho", np.ones(n * n))
        ctx.set_field("ux", ux)
        ctx.set_field("uy", np.zeros(n * n))
        ctx.init()
        ms = timed(ctx, 20)
        out.append(("mrt_%s_d2q9" % variant, ms, 144))
        ctx.close()
    # TRT diffusion: the reference disc scaled to the grid (radius 40/129 of the side); updates = fluid nodes
    mask, rho, ux, uy = api.trt_cylinder_setup(n, n, radius=int(n * 40 / 129))
    ctx = api.Context(api.trt_config(n, n))
    ctx.set_geometry(mask)
    ctx.set_field("rho", rho)
    ctx.set_field("ux", ux)
    ctx.set_field("uy", uy)
    ctx.init()
    ms = timed(ctx, 20)
    fluid = int((mask == 1).sum())
    ctx.close()
    # 9 reads + 9 writes + rho write + ux, uy reads = 168 B per fluid node, + 2 B link word per node of the array
    print(json.dumps({"model": "trt_diffusion_d2q9", "grid": [n, n], "fluid_nodes": fluid, "ms_per_step": ms,
                      "mlups": fluid / ms / 1e3, "bytes_per_update": 168,
                      "achieved_gbs": (fluid * 168 + n * n * 2) / ms / 1e6,
                      "frac_of_measured_hbm": (fluid * 168 + n * n * 2) / ms / 1e6 / peak}))
    # curved solid: the reference disc and droplets scaled to the grid
    mask, rho, phi, ux, uy = api.shape_setup(n, n, radius=int(n * 50 / 301), radius_droplet=int(n * 10 / 301))
    fluid = int((mask == 1).sum())
    for variant in (0,):
        cfg = api.shape_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_geometry(mask)
        for name, a in (("rho", rho), ("phase", phi), ("ux", ux), ("uy", uy)):
            ctx.set_field(name, a)
        ctx.init()
        ms = timed(ctx, 20)
        ctx.close()
        print(json.dumps({"model": "binary_shape_d2q9" if variant == 0 else "binary_shape_d2q9 (variant %d)" % variant, "grid": [n, n], "fluid_nodes": fluid, "ms_per_step": ms,
                          "mlups": fluid / ms / 1e3, "bytes_per_update": 288, "achieved_gbs": fluid * 288 / ms / 1e6,
                          "frac_of_measured_hbm": fluid * 288 / ms / 1e6 / peak}))
    for name, ms, b in out:
        mlups = n * n / ms / 1e3
        print(json.dumps({"model": name, "grid": [n, n], "ms_per_step": ms, "mlups": mlups, "bytes_per_update": b,
                          "achieved_gbs": mlups * b / 1e3, "frac_of_measured_hbm": mlups * b / 1e3 / peak}))


if __name__ == "__main__":
    main()
