#!/usr/bin/env python
"""Runs a few steps of the D3Q19 model on a small channel (for compute-sanitizer / ncu): run3d.py NX NY NZ STEPS VARIANT
With WORLD_SIZE > 1 in the environment (torchrun) every rank takes one z-slab and the halos go over NCCL."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmannmethod_b200 import api  # noqa: E402


def main():
    nx, ny, nz, steps, variant = [int(a) for a in sys.argv[1:6]]
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = api.microchannel_config("quarter", nx, ny, nz, device=local, slab_rank=rank, slab_count=world)
    cfg.kernel_variant = variant
    cfg.wall_gradient = 0.02
    ctx = api.Context(cfg)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = [api.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_attach(uid[0])
    z0, nzl = ctx.slab_extent()
    pl = nx * ny
    for name, a in zip(("rho", "phase", "ux", "uy", "uz"), api.microchannel_initial_fields(nx, ny, nz, max(1, nx // 5))):
        ctx.set_field(name, np.ascontiguousarray(a[z0 * pl:(z0 + nzl) * pl]))
    ctx.init()
    ctx.step(steps)
    ctx.step(2)
    ctx.sync()
    d = ctx.diagnostics()
    print("rank", rank, "d3q19", nx, ny, nz, "variant", variant, "ok mass", d["mass"], "nan", d["has_nan"], "fallbacks", ctx.phase_fallbacks)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
