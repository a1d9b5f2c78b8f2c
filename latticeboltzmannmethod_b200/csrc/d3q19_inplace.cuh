// d3q19_inplace.cuh -- the D3Q19 binary-liquid step on ONE population array (in-place, AA-pattern streaming),
// kernel_variant 2 (one pass) and 6 (two passes).  Replaces the second population array and exchangeMatrices
// (lattice.hpp:842-855) on top of Lattice::collide_stream / updateMacro (lattice.hpp:4-108, 377-688).
//
// The per-thread address rules and the reasons why they are race-free are in d3q19_threads.h (aa_loc and friends);
// the CPU unit test runs those.  Here: the kernels.  A slab that holds the whole grid takes z periodic by addressing
// (Geom::zwrap): blocks address the planes below and above explicitly, populations need no ghost planes.  A slab of
// several uses its two ghost planes as the places beyond its z faces; Slab3D::aa_exchange moves them (five directions
// per face) after every sub-step.
//
//   k_aa_phase / k_aa_core (k_aa_collide for the TRT and MRT extensions)   two passes: phase field, then collide in place
//   k_aa_one + k_aa_repair                                                 one pass, see below
//   k_aa_swap                                                              odd half of the cycle <-> streamed populations
//   k_aa_merge                                                             several slabs: second half of the exchange after an odd step
#ifndef LBM_D3Q19_INPLACE_CUH
#define LBM_D3Q19_INPLACE_CUH

#include "d3q19_step.cuh"
#include "d3q19_threads.h"

namespace lbm3 {

// one slab holding the whole grid: z periodic by addressing; a slab of several: the ghost planes 1 and nz+2, which the
// halo exchange of the in-place step fills and empties (Slab3D::aa_exchange)
__device__ __forceinline__ AaPlanes aa_planes(const Geom &g, int zs)
{
    const long long pp = 19 * g.pl;
    const int zm = g.zwrap && zs == 2 ? g.nz + 1 : zs - 1, zp = g.zwrap && zs == g.nz + 1 ? 2 : zs + 1;
    return AaPlanes{zm * pp, zs * pp, zp * pp};
}

// planes zs0, zs0 + zstep, ... (blockIdx.y): all owned planes (2, 1), or just the first and the last one (2, nz - 1)
template <bool ODD>
__global__ void __launch_bounds__(256) k_aa_phase(Geom g, Params P, const pop_t *gq, double *phi, int zs0, int zstep)
{
    const unsigned idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= (unsigned)(g.nx * g.ny)) return;
    const int zs = zs0 + (int)blockIdx.y * zstep;
    aa_phase_thread<ODD>(g, P, gq, phi, zs, aa_planes(g, zs), (int)(idx % (unsigned)g.nx), (int)(idx / (unsigned)g.nx));
}

// array form of the collision (any operator), one thread per node of an owned plane
template <bool FXY, bool WRITE, int COLL, bool ODD>
__global__ void __launch_bounds__(128, 3) k_aa_collide(Geom g, Params P, pop_t *f, pop_t *gq, const double *__restrict__ phi,
                                                        FieldsOut o)
{
    const unsigned idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= (unsigned)(g.nx * g.ny)) return;
    const int zs = 2 + blockIdx.y;
    aa_collide_thread<FXY, WRITE, COLL, ODD>(g, P, f, gq, phi, o, zs, aa_planes(g, zs), (int)(idx % (unsigned)g.nx),
                                             (int)(idx / (unsigned)g.nx));
}

__global__ void __launch_bounds__(256) k_aa_swap(Geom g, pop_t *f, pop_t *gq)
{
    const unsigned idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= (unsigned)(g.nx * g.ny)) return;
    const int zs = 2 + blockIdx.y;
    aa_swap_thread(g, f, gq, aa_planes(g, zs), (int)(idx % (unsigned)g.nx), (int)(idx / (unsigned)g.nx), !g.zwrap && zs == 2,
                   !g.zwrap && zs == g.nz + 1);
}

// Second half of the halo exchange after an odd step (Slab3D::aa_exchange): what the neighbours pushed across the two z
// faces has arrived in the unused rows of the ghost planes (c_z = +1 rows of plane 1, c_z = -1 rows of plane nz+2) and
// goes into the first / last owned plane -- except where the slot holds the node's own bounced-back population.
__global__ void __launch_bounds__(256) k_aa_merge(Geom g, pop_t *f, pop_t *gq)
{
    const unsigned idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= (unsigned)(g.nx * g.ny)) return;
    const int ix = (int)(idx % (unsigned)g.nx), iy = (int)(idx / (unsigned)g.nx);
    if (ix < 1 || ix > g.nx - 2 || iy < 1 || iy > g.ny - 2) return;
    constexpr int UP[5] = {5, 11, 12, 15, 16}, DOWN[5] = {6, 13, 14, 17, 18};
    const size_t pl = (size_t)g.pl;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        if (aa_crossed_face(g, UP[i], ix, iy)) {          // from below into the first owned plane
            const size_t src = ((size_t)1 * 19 + UP[i]) * pl + idx, dst = ((size_t)2 * 19 + UP[i]) * pl + idx;
            f[dst] = f[src];
            gq[dst] = gq[src];
        }
        if (aa_crossed_face(g, DOWN[i], ix, iy)) {        // from above into the last owned plane
            const size_t src = ((size_t)(g.nz + 2) * 19 + DOWN[i]) * pl + idx, dst = ((size_t)(g.nz + 1) * 19 + DOWN[i]) * pl + idx;
            f[dst] = f[src];
            gq[dst] = gq[src];
        }
    }
}

// ---- the BGK step in place with the collision in its streamed form (collide_streamed, as in k_step_one) ----
// Every thread parks its own 19 g values in its own column of shared memory with asynchronous copies issued before
// anything else -- they are needed only at the end of the collision -- and every post-collision pair leaves for its
// place as soon as it exists: 128 registers, 16 warps per SM.  No thread reads another thread's column: no barrier.
// A direction group of collide_streamed is closed under reversal and fetches its g values before it emits, and all
// f values are in registers before the first store, so no place is overwritten before it has been read.
__device__ __forceinline__ void aa_cp_async(pop_t *smem, const pop_t *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    if (sizeof(pop_t) == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem) : "memory");
}
template <bool ODD>
struct AaEmit {
    const Geom &g;
    const AaPlanes &z;
    pop_t *f, *gq;
    int ix, iy;
    template <int K>
    __device__ __forceinline__ void operator()(double fv, double gv) const
    {
        const long long t = aa_slot<ODD, K>(g, z, ix, iy);
        f[t] = (pop_t)fv;
        gq[t] = (pop_t)gv;
    }
};
// first half of a node: request the g values (into the column gs[k * 128]) and load the f values
template <bool ODD>
__device__ __forceinline__ void aa_node_fetch(const Geom &g, const AaPlanes &z, const pop_t *f, const pop_t *gq, int ix, int iy,
                                              pop_t *gs, double *F)
{
    aa_each([&](auto m) {
        constexpr int M = decltype(m)::value;
        constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
        aa_cp_async(gs + OPP[M] * 128, gq + aa_slot<ODD, M>(g, z, ix, iy));
    });
    aa_load<ODD>(g, z, f, ix, iy, F);
}
// second half: phase stencil, collision, stores (the caller has waited for the asynchronous copies)
template <bool FXY, bool WRITE, bool ODD>
__device__ __forceinline__ void aa_node_finish(const Geom &g, const Params &P, const AaPlanes &z, pop_t *f, pop_t *gq,
                                               const double *__restrict__ phi, const FieldsOut &o, int zs, int idx, int ix,
                                               int iy, const pop_t *gs, double *F)
{
    const int NX = g.nx;
    double PH[19];
    {
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        const double *pp = phi + (long long)zs * g.pl + idx;
        const int pl32 = (int)g.pl;
#pragma unroll
        for (int k = 0; k < 19; k++) PH[k] = __ldg(pp + (CZ[k] * pl32 + CY[k] * NX + CX[k]));
    }
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    if (WRITE) {
        if (iy == 1) mirror_fields<2>(F, P, o, (long long)(zs - 2) * g.pl + ix);
        if (ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * g.pl + (long long)iy * NX);
        if (ix == 1 && iy == 1) mirror_fields<3>(F, P, o, (long long)(zs - 2) * g.pl);
    }
    auto getG = [gs](int k) { return (double)gs[k * 128]; };
    // the stores go to the places the loads came from; with the pitch hidden from the optimiser the 38 store
    // addresses are formed when they are needed instead of being kept in registers from the loads on (spills)
    Geom gst = g;
    {
        int pitch = (int)g.pl;
        asm volatile("" : "+r"(pitch));
        gst.pl = pitch;
    }
    double rho, ux, uy, uz;
    collide_streamed<FXY>(F, getG, AaEmit<ODD>{gst, z, f, gq, ix, iy}, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
    if (WRITE) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
    }
}
// wall node fields, lattice.hpp:75-106
__device__ __forceinline__ void aa_wall_fields(const Geom &g, const Params &P, const FieldsOut &o, int zs, int idx, int ix, int iy)
{
    if (ix == g.nx - 1 || iy == g.ny - 1) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
    }
}

template <bool FXY, bool WRITE, bool ODD>
__global__ void __launch_bounds__(128, 4) k_aa_core(Geom g, Params P, pop_t *f, pop_t *gq, const double *__restrict__ phi,
                                                     FieldsOut o)
{
    __shared__ pop_t gs_all[19 * 128];
    const int NX = g.nx, NY = g.ny;
    const unsigned idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= (unsigned)(NX * NY)) return;
    const int zs = 2 + blockIdx.y;
    const int ix = (int)(idx % (unsigned)NX), iy = (int)(idx / (unsigned)NX);
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE) aa_wall_fields(g, P, o, zs, (int)idx, ix, iy);
        return;
    }
    const AaPlanes z = aa_planes(g, zs);
    pop_t *gs = gs_all + threadIdx.x;
    double F[19];
    aa_node_fetch<ODD>(g, z, f, gq, ix, iy, gs, F);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    aa_node_finish<FXY, WRITE, ODD>(g, P, z, f, gq, phi, o, zs, (int)idx, ix, iy, gs, F);
}

// ---- one pass: the phase field of THIS step runs ahead of the collision inside the same launch ----
// The collision at x needs phi = sum_k g_k at x and its 18 neighbours, so phi of the neighbourhood must exist before
// x is collided; the two-pass form sums phi in a kernel of its own that reads all of g from HBM a second time.  In
// place, the g values a node's phase sum needs are exactly the node's own 19 places (in either half of the cycle),
// untouched until the node itself collides.  So the grid walks the slab in narrow x-y strips, plane after plane
// (the geometry of k_step_one), and block (strip s, plane-step t, segment i)
//   * sums the phase field of plane-sequence element t, segment i -- no dependency on anything in this launch --
//     which also pulls those g values into L2, and
//   * collides plane-sequence element t - L, segment i - D, whose phase stencil is complete once the phase parts of
//     plane-steps t-L .. t-L+2 are (strip s and, for the first segments, strip s-1: counters as in k_step_one, one
//     release-add per block, relaxed polls), and whose g values are mostly still in L2 (674 B of DRAM traffic per
//     lattice update measured at 512^3, 624 if none had left it; the two-pass form moves 780).
// Plane sequence, one slab holding the whole grid: nz-1, 0, 1, .., nz-2 for the phase parts (the plane below plane 0
// first: z is periodic) and 0 .. nz-1 for the collisions; the phase part of plane 0 / nz-1 also writes the ghost plane
// above / below.  A slab of several: 0 .. nz-1 for both; the phase field of the two ghost planes is there before the
// launch (the neighbours' first / last plane, summed by k_aa_phase and exchanged).
// A warp whose counters are not full after a bounded wait does NOT collide: it records itself in `skipped`
// (one word per 128-node segment of an owned plane, one bit per warp) and raises flags[0]; k_aa_repair then collides
// those segments from the finished phase field.  In place that is all a repair takes -- a node's collision touches
// only its own places -- so an unexpected dispatch order costs time, never a hang and never a different result.
struct AaPlan {
    int sb, pz, nstrips;   // grid (sb, pz, nstrips): blocks per strip and plane-step, plane-steps (nz + L), strips
    int L;                 // the collision trails the phase part by L plane-steps (>= 3)
    int dshift;            // ... and by dshift segments
    int nblk;              // 128-node segments per plane
    unsigned full;         // counter value of a finished group (= sb)
    unsigned spin_limit;
};

// No L2 residency hints: every one tried (phase-part loads or phase stores asked to stay, last-use loads and stores
// asked to go first, an evict_last L2 prefetch) cut the DRAM traffic and cost more time than it saved.
template <bool FXY, bool WRITE, bool ODD>
__global__ void __launch_bounds__(128, 4)
k_aa_one(Geom g, Params P, pop_t *f, pop_t *gq, double *phi, FieldsOut o, unsigned *done, unsigned *flags, unsigned *skipped, AaPlan A)
{
    __shared__ pop_t gs_all[19 * 128];   // g of the nodes this block collides, one column per thread
    __shared__ pop_t ps_all[19 * 128];   // g of the nodes whose phase field it sums
    const int NX = g.nx, NY = g.ny, nz = g.nz, nplane = NX * NY;
    const int t = (int)blockIdx.y, strip = (int)blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const AaSchedule sched = aa_schedule(g.zwrap, nz, A.L, t);
    const int c = sched.collide_plane;                      // plane this block collides (-1: none)
    const bool c_step = c >= 0;
    const int bbase = (strip * A.sb + (int)blockIdx.x) * 128;   // first node of the phase segment
    const int cbase = bbase - A.dshift * 128;                   // first node of the collide segment

    // counters the collision depends on: phase groups (c, c+1, c+2) mod nz of this strip and -- for the segments whose
    // stencil reaches back into it -- of the strip before.  Requested first, looked at after the loads are on their way.
    bool need = false;
    unsigned seen = A.full;
    const unsigned *cnt = done;
    if (c_step && lane < 6) {
        const int j = lane % 3, q = j == 0 ? sched.dep[0] : (j == 1 ? sched.dep[1] : sched.dep[2]), sdep = strip - lane / 3;
        need = sdep >= 0 && q >= 0 && (lane < 3 || (int)blockIdx.x < 2 * A.dshift);
        cnt = done + (size_t)(sdep * nz + (need ? q : 0)) * LAG_PAD;
        if (need) seen = lag_ld_relaxed(cnt);
    }

    // ---- collide part, first half: everything that does not need the phase field
    const int cidx = cbase + (int)threadIdx.x;
    const int zs = c + 2;
    int ix = 0, iy = 0;
    bool do_c = c_step && cbase >= 0 && cidx < nplane;
    if (do_c) {
        ix = (int)((unsigned)cidx % (unsigned)NX);
        iy = (int)((unsigned)cidx / (unsigned)NX);
        if (WRITE) aa_wall_fields(g, P, o, zs, cidx, ix, iy);
        do_c = ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    }
    AaPlanes zc{0, 0, 0};
    pop_t *gs = gs_all + threadIdx.x;
    double F[19];
    if (do_c) {
        zc = aa_planes(g, zs);
        aa_node_fetch<ODD>(g, zc, f, gq, ix, iy, gs, F);
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");

    // ---- phase part, requests: the g values of plane-sequence element t (nothing in this launch has touched these
    // places yet) into this thread's second column; summed at the very end of the block, when they have long arrived
    const int pidx = bbase + (int)threadIdx.x;
    const int pz = sched.phase_plane;
    int px = 0, py = 0;
    bool do_p = pz >= 0 && pidx < nplane;
    if (do_p) {
        px = (int)((unsigned)pidx % (unsigned)NX);
        py = (int)((unsigned)pidx / (unsigned)NX);
        do_p = px >= 1 && px <= NX - 2 && py >= 1 && py <= NY - 2;
    }
    pop_t *ps = ps_all + threadIdx.x;
    if (do_p) {
        const AaPlanes zp = aa_planes(g, pz + 2);
        aa_each([&](auto m) {
            constexpr int M = decltype(m)::value;
            constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
            aa_cp_async(ps + OPP[M] * 128, gq + aa_slot<ODD, M>(g, zp, px, py));
        });
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");

    // ---- wait for the phase field of the collide plane's neighbourhood (normally long there)
    bool ok = true;
    if (c_step && cbase >= 0 && cbase + (warp << 5) < nplane) {          // warp-uniform
        unsigned spins = 0;
        for (;;) {
            ok = __all_sync(0xffffffffu, seen >= A.full);
            if (ok || ++spins > A.spin_limit) break;
            if (__any_sync(0xffffffffu, lane == 6 && lag_ld_relaxed(flags) != 0u)) break;
            __nanosleep(200);
            if (need) seen = lag_ld_relaxed(cnt);
        }
        if (!ok && lane == 0) {
            atomicOr(skipped + (size_t)c * A.nblk + (cbase >> 7), 1u << warp);
            atomicOr(flags, 1u);
        }
    }
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");              // the g values of the collision have landed
    if (do_c && ok) aa_node_finish<FXY, WRITE, ODD>(g, P, zc, f, gq, phi, o, zs, cidx, ix, iy, gs, F);

    // ---- phase part, sums
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    if (do_p) {
        double G2[19];
#pragma unroll
        for (int k = 0; k < 19; k++) G2[k] = (double)ps[k * 128];
        phase_node(g, P, G2, phi, pz + 2, px, py);
        if (g.zwrap && pz == 0) phase_node(g, P, G2, phi, nz + 2, px, py);       // ghost plane above the slab
        if (g.zwrap && pz == nz - 1) phase_node(g, P, G2, phi, 1, px, py);       // ghost plane below
    }

    // publish the end of this block's phase part: one release-add per block, after a block barrier
    __syncthreads();
    if (t < nz && threadIdx.x == 0) lag_add_release(done + (size_t)(strip * nz + t) * LAG_PAD, 1u);
}

// collides what k_aa_one skipped (array form of the collision: same bits); a few blocks that look at the flag and leave
template <bool FXY, bool WRITE, bool ODD>
__global__ void __launch_bounds__(128, 3) k_aa_repair(Geom g, Params P, pop_t *f, pop_t *gq, const double *__restrict__ phi,
                                                       FieldsOut o, const unsigned *flags, const unsigned *skipped,
                                                       unsigned *sticky, int nblk)
{
    if (*reinterpret_cast<const volatile unsigned *>(flags) == 0u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sticky, 1u);
    const int total = nblk * g.nz, warp = threadIdx.x >> 5;
    for (int b = blockIdx.x; b < total; b += gridDim.x) {
        if (!((skipped[b] >> warp) & 1u)) continue;
        const int seg = b % nblk, zs = 2 + b / nblk;
        const unsigned idx = (unsigned)seg * 128 + threadIdx.x;
        if (idx >= (unsigned)(g.nx * g.ny)) continue;
        const int ix = (int)(idx % (unsigned)g.nx), iy = (int)(idx / (unsigned)g.nx);
        if (ix < 1 || ix > g.nx - 2 || iy < 1 || iy > g.ny - 2) continue;   // wall-node fields were written by k_aa_one
        aa_collide_thread<FXY, WRITE, 0, ODD>(g, P, f, gq, phi, o, zs, aa_planes(g, zs), ix, iy);
    }
}

}  // namespace lbm3
#endif
