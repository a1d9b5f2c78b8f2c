"""z-slab ring driven from Python with torch.distributed as the transport.

The library can move its halos itself over NCCL (``Context.comm_attach`` then ``init``/``step``); this
module is the other option the C ABI offers -- the caller moves the four contiguous halo messages per
step with whatever it has.  Here that is ``torch.distributed`` point-to-point (NCCL between GPUs, gloo in
the CPU tests).  It replaces Solver::updatePhase and Solver::propagate (Microchannel3D solver.hpp:821-868,
893-928) and follows the same call order as the library's own ``lbm_init`` / ``lbm_step``.

`engine` is anything with the slab phase methods of :class:`latticeboltzmannmethod_b200.api.Context`;
`tensor_of(ptr, count, elem_bytes)` turns one halo message into a 1-D tensor (float64, or float32 for the
population messages of the fp32-storage build) that aliases it.
"""
import torch
import torch.distributed as dist

STATE = (("S_F_UP", "R_F_FROM_BELOW"), ("S_G_UP", "R_G_FROM_BELOW"), ("S_F_DOWN", "R_F_FROM_ABOVE"), ("S_G_DOWN", "R_G_FROM_ABOVE"))
PHI = (("S_PHI_UP", "R_PHI_FROM_BELOW"), ("S_PHI_DOWN", "R_PHI_FROM_ABOVE"))


def slab_bounds(nz, rank, world):
    """solver.hpp:25-30: nz/world planes each, the last rank takes the remainder."""
    zstep = nz // world
    z0 = rank * zstep
    z1 = nz - 1 if rank == world - 1 else z0 + zstep - 1
    return z0, z1 - z0 + 1


def cuda_tensor_of(ptr, count, elem_bytes=8):
    """A CUDA tensor aliasing `count` values of `elem_bytes` bytes (8: float64, 4: float32) of device memory at `ptr`."""
    class _Mem:
        __cuda_array_interface__ = {"shape": (count,), "typestr": "<f%d" % elem_bytes, "data": (ptr, False), "version": 2}
    return torch.as_tensor(_Mem(), device="cuda")


class SlabDriver:
    def __init__(self, engine, rank, world, tensor_of=cuda_tensor_of, group=None):
        self.e, self.rank, self.world, self.tensor_of, self.group = engine, rank, world, tensor_of, group
        self.up, self.down = (rank + 1) % world, (rank - 1 + world) % world

    def exchange(self, pairs):
        # sends ordered [up..., down...], receives [from below..., from above...]: with two slabs both
        # neighbours are the same peer and messages between a pair are matched in posting order
        half = len(pairs) // 2
        ops = []
        for i, (s, _) in enumerate(pairs):
            ops.append(dist.P2POp(dist.isend, self.tensor_of(*self.e.slab_halo(s)), self.up if i < half else self.down, self.group))
        for i, (_, r) in enumerate(pairs):
            ops.append(dist.P2POp(dist.irecv, self.tensor_of(*self.e.slab_halo(r)), self.down if i < half else self.up, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def init(self):
        self.e.slab_init_fields()
        self.exchange(PHI)
        self.e.slab_init_pops()
        self.exchange(STATE)
        self.e.slab_finish()
        self.e.slab_after_init()

    def step(self, n=1):
        for t in range(n):
            self.e.slab_collide(t == n - 1)
            self.exchange(STATE)
            self.e.slab_finish()
