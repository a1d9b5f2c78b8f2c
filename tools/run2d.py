#!/usr/bin/env python
"""Runs a few steps of one D2Q9 model on an n x n grid (for ncu captures): run2d.py MODEL [n] [steps] [variant]."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmannmethod_b200 import api  # noqa: E402


def main():
    if sys.argv[1] == "all":
        for m in ("shanchen", "binary", "mrt_gs", "mrt_hermite", "trt", "shape"):
            run(m, *[int(a) for a in sys.argv[2:]])
        return
    run(sys.argv[1], *[int(a) for a in sys.argv[2:]])


def run(model, n=8192, steps=4, variant=0):
    if model == "shanchen":
        cfg = api.shanchen_config(n, n)
        fields = dict(rho=np.tile(api.shanchen_initial_rho().reshape(128, 128), (n // 128, n // 128)).ravel())
        mask = None
    elif model == "binary":
        cfg = api.binary2d_config(n, n)
        fields = dict(phase=np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel())
        mask = None
    elif model in ("mrt_gs", "mrt_hermite"):
        v = model.split("_")[1]
        cfg = api.mrt_config(v, n, n, 1.2 if v == "gs" else 0.9, 0.8 if v == "gs" else 0.1)
        x = np.arange(n)
        fields = dict(rho=np.ones(n * n), ux=np.repeat(1e-3 * np.cos(x / n * 2 * np.pi), n), uy=np.zeros(n * n))
        mask = None
    elif model == "trt":
        cfg = api.trt_config(n, n)
        mask, rho, ux, uy = api.trt_cylinder_setup(n, n, radius=int(n * 40 / 129))
        fields = dict(rho=rho, ux=ux, uy=uy)
    elif model == "shape":
        cfg = api.shape_config(n, n)
        mask, rho, phi, ux, uy = api.shape_setup(n, n, radius=int(n * 50 / 301), radius_droplet=int(n * 10 / 301))
        fields = dict(rho=rho, phase=phi, ux=ux, uy=uy)
    else:
        raise SystemExit("unknown model " + model)
    cfg.kernel_variant = variant
    ctx = api.Context(cfg)
    if mask is not None:
        ctx.set_geometry(mask)
    for k, a in fields.items():
        ctx.set_field(k, a)
    ctx.init()
    ctx.step(steps)
    ctx.sync()
    print(model, n, steps, "ok", ctx.diagnostics()["mass"])
    ctx.close()


if __name__ == "__main__":
    main()
