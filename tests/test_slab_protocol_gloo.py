"""The z-slab ring protocol (what is sent where, in which order, around which phase) on two and three CPU
ranks over gloo: the driver of latticeboltzmannmethod_b200/slab.py moves the halo messages, the oracle plays
the slab engine.  The gathered result must equal the single-slab run bit for bit (the reference's own
decomposition is arithmetic-neutral)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as ol
from latticeboltzmannmethod_b200.slab import SlabDriver, slab_bounds

NX, NY, NZ, WIDTH, STEPS = 11, 9, 13, 3, 7
PARAMS = [0.04, 0.04, 1.0, 0.03, 0.0, 0.5, 2.0, 2.5, 0.7, 1e-5, 0.0, 3e-5]


class OracleSlabEngine:
    """Adapts one oracle slab to the phase interface of api.Context."""

    def __init__(self, nx, ny, nz_local, params):
        self.o = ol.Oracle3D(nx, ny, nz_local, params)
        L, h = self.o.L, self.o.h
        self.slab_init_fields = lambda: L.lbo3_init_fields(h)
        self.slab_init_pops = lambda: L.lbo3_init_pops(h)
        self.slab_after_init = lambda: L.lbo3_after_init(h)
        self.slab_collide = lambda w: L.lbo3_collide(h, int(bool(w)))
        self.slab_finish = lambda: L.lbo3_finish(h)

    def slab_halo(self, name):
        a = self.o.halo(name)
        return a, a.size


def _fields():
    rng = np.random.default_rng(11)
    n = NX * NY * NZ
    ph = ol.bubble_fields(NX, NY, NZ, WIDTH)[1]
    return [1 + 0.05 * rng.standard_normal(n), np.clip(ph + 0.3 * rng.standard_normal(n), -1.2, 1.2),
            0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n), 0.01 * rng.standard_normal(n)]


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    z0, nzl = slab_bounds(NZ, rank, world)
    pl = NX * NY
    eng = OracleSlabEngine(NX, NY, nzl, PARAMS)
    eng.o.set_fields(*[f[z0 * pl:(z0 + nzl) * pl] for f in _fields()])
    drv = SlabDriver(eng, rank, world, tensor_of=lambda a, n, eb=8: torch.from_numpy(a))
    drv.init()
    snaps = {"init": eng.o.fields()}
    drv.step(1)
    snaps["t1"] = eng.o.fields()
    drv.step(STEPS - 1)
    snaps["tN"] = eng.o.fields()
    f, g = eng.o.pops()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), f=f, g=g, **{"%s_%s" % (s, k): v for s, d in snaps.items() for k, v in d.items()})
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_slab_ring_equals_single_slab(world, tmp_path):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    o = ol.Oracle3D(NX, NY, NZ, PARAMS)
    o.set_fields(*_fields())
    o.init_single()
    want = {"init": o.fields()}
    o.step_single(1)
    want["t1"] = o.fields()
    o.step_single(STEPS - 1)
    want["tN"] = o.fields()
    fo, go = o.pops()
    parts = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    for snap, d in want.items():
        for k, v in d.items():
            got = np.concatenate([p["%s_%s" % (snap, k)] for p in parts])
            assert ol.ndiff(v, got) == 0, (snap, k)
    assert ol.ndiff(fo, np.concatenate([p["f"] for p in parts])) == 0
    assert ol.ndiff(go, np.concatenate([p["g"] for p in parts])) == 0


def test_slab_bounds_follow_the_reference():
    # solver.hpp:25-30 on the reference's own split: 52 planes over 3 ranks -> 17/17/18
    assert [slab_bounds(52, r, 3) for r in range(3)] == [(0, 17), (17, 17), (34, 18)]
    assert slab_bounds(1500, 7, 8) == (1309, 191)
