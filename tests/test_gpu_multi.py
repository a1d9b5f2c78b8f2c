"""z-slab decomposition on real GPUs (needs >= 2; skipped otherwise): the slab ring must reproduce the
one-GPU run bit for bit, with the library's own NCCL exchange and with the
torch.distributed transport of latticeboltzmannmethod_b200/slab.py."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NX, NY, NZ, WIDTH, STEPS = 35, 18, 23, 4, 9
PARAMS = [0.04, 0.04, 1.0, 0.03, 0.0, 0.5, 2.0, 2.5, 0.7, 1e-5, 0.0, 3e-5]
F3 = ("rho", "phase", "ux", "uy", "uz")


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _fields():
    import oracle_lib as ol
    rng = np.random.default_rng(11)
    n = NX * NY * NZ
    ph = ol.bubble_fields(NX, NY, NZ, WIDTH)[1]
    return [1 + 0.05 * rng.standard_normal(n), np.clip(ph + 0.3 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]


def _cfg(api, rank, world, variant=0):
    cfg = api.microchannel_config("quarter", NX, NY, NZ, device=rank, slab_rank=rank, slab_count=world)
    cfg.kernel_variant = variant
    (cfg.aconst, cfg.kconst, cfg.gammaconst, cfg.wall_gradient, _pw, cfg.rho_wall, cfg.tau_phi, cfg.tau_liq, cfg.tau_gas) = PARAMS[:9]
    cfg.force[0], cfg.force[1], cfg.force[2] = PARAMS[9:12]
    return cfg


def _worker(rank, world, port, out_dir, transport, precision="f64", variant=0, more_steps=0):
    import torch
    import torch.distributed as dist
    from latticeboltzmannmethod_b200 import api
    from latticeboltzmannmethod_b200.slab import SlabDriver
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ctx = api.Context(_cfg(api, rank, world, variant), precision)
    z0, nzl = ctx.slab_extent()
    pl = NX * NY
    for name, a in zip(F3, _fields()):
        ctx.set_field(name, a[z0 * pl:(z0 + nzl) * pl])
    if transport == "nccl":
        uid = [api.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_attach(uid[0])
        ctx.init()
        ctx.step(STEPS)
    else:
        ext = torch.cuda.ExternalStream(ctx.stream)
        with torch.cuda.stream(ext):
            drv = SlabDriver(ctx, rank, world)
            drv.init()
            drv.step(STEPS)
        ext.synchronize()
    ctx.sync()
    out = {k: ctx.get_field(k) for k in F3}
    out["f"], out["g"] = ctx.get_populations(0), ctx.get_populations(1)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **out)
    if more_steps:      # the reads above must have left the run undisturbed
        ctx.step(more_steps)
        out = {k: ctx.get_field(k) for k in F3}
        out["f"], out["g"] = ctx.get_populations(0), ctx.get_populations(1)
        np.savez(os.path.join(out_dir, "more_rank%d.npz" % rank), **out)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("transport,precision", [("nccl", "f64"), ("torch", "f64"), ("nccl", "f32"), ("torch", "f32")])
def test_slab_ring_on_gpus_equals_one_gpu(transport, precision, tmp_path):
    """f32: the fp32-storage build sends float halo messages; its slab ring must equal its own one-GPU run."""
    import torch.multiprocessing as mp
    import oracle_lib as ol
    from latticeboltzmannmethod_b200 import api
    world = min(_ngpu(), 4)
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), transport, precision), nprocs=world, join=True)
    ctx = api.Context(_cfg(api, 0, 1), precision)
    for name, a in zip(F3, _fields()):
        ctx.set_field(name, a)
    ctx.init()
    ctx.step(STEPS)
    parts = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    for k in F3:
        assert ol.ndiff(ctx.get_field(k), np.concatenate([p[k] for p in parts])) == 0, k
    assert ol.ndiff(ctx.get_populations(0), np.concatenate([p["f"] for p in parts])) == 0
    assert ol.ndiff(ctx.get_populations(1), np.concatenate([p["g"] for p in parts])) == 0
    ctx.close()


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("variant,precision", [(2, "f64"), (6, "f64"), (2, "f32")])
def test_in_place_slab_ring_on_gpus_equals_one_gpu(variant, precision, tmp_path):
    """kernel_variant 2 / 6 on several GPUs: one population array per slab, only the five z-crossing directions travel
    (after every sub-step, in alternating directions).  Bit-equal to the one-GPU two-array run; populations read in the
    odd half of the cycle (9 steps) are the streamed ones and reading them does not disturb the run (4 more steps)."""
    import torch.multiprocessing as mp
    import oracle_lib as ol
    from latticeboltzmannmethod_b200 import api
    world = min(_ngpu(), 4)
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), "nccl", precision, variant, 4), nprocs=world, join=True)
    ctx = api.Context(_cfg(api, 0, 1, 1), precision)      # the plain two-array kernel on one GPU
    for name, a in zip(F3, _fields()):
        ctx.set_field(name, a)
    ctx.init()
    for steps, prefix in ((STEPS, "rank"), (4, "more_rank")):
        ctx.step(steps)
        parts = [np.load(os.path.join(str(tmp_path), "%s%d.npz" % (prefix, r))) for r in range(world)]
        for k in F3:
            assert ol.ndiff(ctx.get_field(k), np.concatenate([p[k] for p in parts])) == 0, (prefix, k)
        assert ol.ndiff(ctx.get_populations(0), np.concatenate([p["f"] for p in parts])) == 0, prefix
        assert ol.ndiff(ctx.get_populations(1), np.concatenate([p["g"] for p in parts])) == 0, prefix
    ctx.close()
