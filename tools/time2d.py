"""Times one D2Q9 model at n x n for a list of kernel variants (measurement helper for the optimisation log).
usage: python tools/time2d.py binary|shanchen|shape n variant[:ENV=VALUE[,ENV=VALUE]] ..."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from latticeboltzmannmethod_b200 import api   # noqa: E402
from bench2d import timed                       # noqa: E402


def make(model, n, variant):
    if model == "binary":
        cfg = api.binary2d_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("phase", np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel())
        return ctx, 288
    if model == "shanchen":
        cfg = api.shanchen_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_field("rho", np.tile(api.shanchen_initial_rho().reshape(128, 128), (n // 128, n // 128)).ravel())
        return ctx, 144
    if model == "shape":
        mask, rho, phi, ux, uy = api.shape_setup(n, n, radius=int(n * 50 / 301), radius_droplet=int(n * 10 / 301))
        cfg = api.shape_config(n, n)
        cfg.kernel_variant = variant
        ctx = api.Context(cfg)
        ctx.set_geometry(mask)
        for name, a in (("rho", rho), ("phase", phi), ("ux", ux), ("uy", uy)):
            ctx.set_field(name, a)
        make.fluid = int((mask == 1).sum())
        return ctx, 288
    raise SystemExit("unknown model " + model)


def main():
    model, n = sys.argv[1], int(sys.argv[2])
    peak = 6547.2
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = json.load(open(pk))["hbm_gbs"]
    for spec in sys.argv[3:]:
        v, _, envs = spec.partition(":")
        sets = dict(e.split("=") for e in envs.split(",") if e)
        os.environ.update(sets)
        ctx, b = make(model, n, int(v))
        ctx.init()
        ms = timed(ctx, 20)
        ctx.close()
        for k in sets:
            del os.environ[k]
        nodes = getattr(make, "fluid", n * n)      # the curved-solid model counts fluid nodes
        print(json.dumps({"model": model, "n": n, "variant": int(v), "env": sets, "ms_per_step": round(ms, 4),
                          "glups": round(nodes / ms / 1e6, 2), "frac_of_measured_hbm": round(nodes * b / ms / 1e6 / peak, 4)}), flush=True)


if __name__ == "__main__":
    main()
