// d3q19_step.cuh -- the D3Q19 binary-liquid step as ONE pass over HBM (k_step_one): collide-push of plane z and, in
// the same launch, the phase field of the NEXT step for a plane a few planes behind, read back while the freshly
// pushed g populations are still in L2.
//
// Reference work covered per launch: Lattice::collide_stream (lattice.hpp:377-688) for step t and the phase sum
// of Lattice::updateMacro (lattice.hpp:19-22, :75-106, :645-686) for step t+1.
//
// Why one pass: the collision at x needs phi = sum_k g_k at the 18 neighbours of x, so phi of the whole
// neighbourhood must exist before x is collided.  The two-pass path forms phi in a kernel of its own that re-reads
// all of g from HBM (161 B per lattice update on top of the 608 the collision moves).  Here the grid walks the slab
// in narrow x-y strips, plane after plane: block b collides (strip s, plane z, segment i) and also sums phi_{t+1}
// for (strip s, plane z - LAG, segment i - D) from the g* values that the blocks before it pushed into the back
// buffer.  Between the push and the read-back only a few tens of MB pass through the 126 MB L2, so the read-back
// is served from it and HBM sees the populations once in and once out (661 B per lattice update measured).
//
// Why it is not slower than the plain collide kernel (every earlier fused form was): nothing is added to the
// collision's register budget and no warp ever waits for another one inside the hot loop.
//   * The g populations of the collision itself and the g* rows of the phase part travel through shared memory
//     with 16-byte L2-only asynchronous copies issued before the arithmetic starts; g is needed only at the very
//     end of the collision, so the streamed form (collide_streamed) runs in 128 registers: 16 warps per SM.
//   * Ordering between blocks: blocks are dispatched in ascending order (x fastest); a block's phase part depends
//     on collide parts (LAG-1) plane-steps or more before it.  A block publishes the end of its collide part with
//     ONE release-add (after a block barrier) on the counter of its group (= the SB blocks of one strip and
//     plane-step; one counter per 128-byte line).  The phase part reads the six counters it depends on with
//     relaxed loads requested before anything else; if they are full (the normal case) the g* copies are issued at
//     once and land during the collision, if not the block collides first and looks again afterwards.  No acquire
//     loads, no __threadfence: both empty L1 for every block on the SM, which the phase stencil of the collision
//     cannot afford.
//   * The second look is a bounded wait: a warp that gives up raises flags[0] and k_phase_finish then recomputes
//     the phase field the two-pass way, so an unexpected dispatch order can cost time but can neither hang the
//     GPU nor change a result.
#ifndef LBM_D3Q19_STEP_CUH
#define LBM_D3Q19_STEP_CUH

#include "d3q19_threads.h"

namespace lbm3 {

constexpr int LAG_PAD = 32;   // k_step_one: one group counter per 128-byte line (they are hammered: keep them on different L2 slices)

struct LagPlan {
    int sb;          // blocks (of 128 in-plane nodes) per strip
    int pz;          // plane-steps per strip
    int nstrips;
    int z0;          // collide plane of plane-step 0 (storage plane index)
    int zjump_at, zjump_by;   // plane-steps from zjump_at on are zjump_by planes further up (both slab ends in one launch; PHI = false only)
    int c0, c1;      // collide planes [c0, c1)
    int p0, p1;      // phi planes [p0, p1)
    int lag;         // phi plane = collide plane - lag
    int dshift;      // phi segment = collide segment - dshift
    unsigned full;   // counter value of a finished group (= blocks per group)
    unsigned spin_limit;
};

// 16 bytes global -> shared, L2 only (never through the non-coherent L1: the source was written by other SMs in
// this very launch)
__device__ __forceinline__ void lag_cp_async16(void *smem, const void *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
// The counter protocol.  Writer: all threads store, __syncthreads, thread 0 adds with RELEASE semantics at GPU scope
// (the block barrier makes the other threads' stores part of what the release publishes).  Reader: RELAXED loads of the
// counters until they are full, then the data is fetched with L2-only copies issued behind the branch that depends
// on the counter values.  Neither side invalidates or goes through L1, which the collision needs for its phase
// stencil (an acquire load or a __threadfence would empty it for every block on the SM, several times per microsecond).
__device__ __forceinline__ unsigned lag_ld_relaxed(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void lag_add_release(unsigned *p, unsigned v)
{
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

// L2 residency hints (POL = 1): the g* values are read back by the phi part a few plane-steps later, so their
// stores ask L2 to keep them (evict_last); everything that streams through once -- the f and g loads, the f*
// stores -- asks to be evicted first.
__device__ __forceinline__ unsigned long long lag_policy_last()
{
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;\n" : "=l"(p));
    return p;
}
__device__ __forceinline__ unsigned long long lag_policy_first()
{
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;\n" : "=l"(p));
    return p;
}
__device__ __forceinline__ double lag_ld_hint(const double *p, unsigned long long pol)
{
    double v;
    asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;\n" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double lag_ld_hint(const float *p, unsigned long long pol)
{
    float v;
    asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;\n" : "=f"(v) : "l"(p), "l"(pol));
    return (double)v;
}
__device__ __forceinline__ void lag_st_hint(double *p, double v, unsigned long long pol)
{
    asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;\n" ::"l"(p), "d"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ void lag_st_hint(float *p, double v, unsigned long long pol)
{
    asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;\n" ::"l"(p), "f"((float)v), "l"(pol) : "memory");
}
struct HintedStore {
    unsigned long long first, last;
    __device__ __forceinline__ void f(pop_t *p, double v) const { lag_st_hint(p, v, first); }
    __device__ __forceinline__ void g(pop_t *p, double v) const { lag_st_hint(p, v, last); }
};

// Interior push with the x mirror line / x wall folded in as an address select (PART 3 of collide_push_node) for
// collide_streamed(): stores one post-collision pair as soon as it exists.
template <class ST>
struct PushXRim {
    pop_t *fo, *go;      // fout/gout + pop_index(zs, 0, idx)
    int pl, nx;          // plane pitch and row length: offsets from fo/go stay below 2^31 elements (checked at set-up), so
                         // an address is ONE 32x32+64 multiply-add from a 32-bit offset, and the select is on 32 bits
    bool xlo, xhi;
    ST st;
    template <int K>
    __device__ __forceinline__ void operator()(double fv, double gv) const
    {
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
        constexpr int SYMX[19] = {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17};
        int off = (CZ[K] * 19 + K) * pl + (CY[K] * nx + CX[K]);
        if (CX[K] == -1) {   // into the x mirror line: reflected, stays in column 1 (lattice.hpp:534-557)
            const int om = (CZ[K] * 19 + SYMX[K]) * pl + CY[K] * nx;
            off = xlo ? om : off;
        }
        if (CX[K] == 1) {    // into the x wall: back to its own node reversed (lattice.hpp:609-626)
            const int ob = OPP[K] * pl;
            off = xhi ? ob : off;
        }
        st.f(fo + off, fv);
        st.g(go + off, gv);
    }
};

// 16-byte chunks of one warp's 32 nodes, all 19 directions, global -> shared through L2 only
__device__ __forceinline__ void lag_copy_warp_rows(pop_t *dst_warp, const pop_t *src_warp, long long pl, long long room, int lane)
{
    constexpr int EPC = 16 / (int)sizeof(pop_t), CPW = 32 / EPC;
    const int k0 = lane / CPW, e = (lane % CPW) * EPC;
    if (e < room) {
        const pop_t *sp = src_warp + k0 * pl + e;
        pop_t *dp = dst_warp + k0 * 128 + e;
#pragma unroll
        for (int j = 0; j < (19 * CPW + 31) / 32; j++) {
            if (k0 + j * (32 / CPW) < 19) lag_cp_async16(dp, sp);
            sp += (32 / CPW) * pl;
            dp += (32 / CPW) * 128;
        }
    }
}

// The one-pass step with the g populations of the collision staged through shared memory as well: they are needed
// only at the very end of the collision (the relaxation), so asynchronous copies issued up front replace 19 loads
// whose 38 registers would be held for the whole collision.  The collision then runs in its streamed form
// (collide_streamed: post-collision pairs leave as soon as they exist) within 128 registers, four blocks = 16 warps
// per SM instead of 12.  BGK only (the TRT extension keeps the array form and the two-pass path).
//   PHI    false: collide only (the two-pass path and the slab ends of the multi-GPU step): no counters, no phi part
template <bool FXY, bool WRITE, int POL, bool PHI>
__global__ void __launch_bounds__(128, 4)
k_step_one(Geom g, Params P, const pop_t *__restrict__ fin, const pop_t *__restrict__ gin, const double *__restrict__ phi,
           pop_t *__restrict__ fout, pop_t *gout, FieldsOut o, double *phi_next, unsigned *done, unsigned *flags, LagPlan L)
{
    extern __shared__ __align__(16) unsigned char lag_smem[];
    pop_t *gs = reinterpret_cast<pop_t *>(lag_smem);   // [19][128]: g of the nodes this block collides
    pop_t *ps = gs + 19 * 128;                          // [19][128]: g* of the nodes whose phase field it sums (PHI)
    const int NX = g.nx, NY = g.ny;
    // grid (sb, pz, nstrips): x runs fastest in the dispatch order, so blocks arrive segment by segment,
    // plane-step by plane-step, strip by strip -- the order the counters assume
    const int zc = (int)blockIdx.y, strip = (int)blockIdx.z;
    const unsigned grp = (unsigned)(strip * L.pz + zc);
    const int nplane = NX * NY;          // in-plane indices fit 32 bits (checked at set-up)
    const long long pl = g.pl;
    const int lane = threadIdx.x & 31, wbase = threadIdx.x & ~31;
    const int zs = L.z0 + zc + (zc >= L.zjump_at ? L.zjump_by : 0);
    const int zp = zs - L.lag;
    const bool c_plane = zs >= L.c0 && zs < L.c1;
    const bool p_plane = PHI && zp >= L.p0 && zp < L.p1;

    // counters the phase part depends on -- the g* of plane zp come from the collide parts of planes zp-1 .. zp+1,
    // segments (this - 2*dshift) .. this: groups (strip, zc-lag-1 .. zc-lag+1) and, for the first segments of a
    // strip, the same of strip-1.  Requested first, looked at after the loads have been issued.
    const unsigned *cnt = done;
    bool need = false;
    unsigned seen = 0;
    if (PHI) {
        const int dz = lane % 3 - 1, ds = lane / 3;          // lanes 0..5
        const int zdep = zp + dz, sdep = strip - ds;
        need = p_plane && lane < 6 && zdep >= L.c0 && zdep < L.c1 && sdep >= 0;
        cnt = done + (size_t)(sdep * L.pz + (zdep - L.z0)) * LAG_PAD;
        seen = need ? lag_ld_relaxed(cnt) : L.full;
    }
    unsigned long long pol_first = 0, pol_last = 0;
    if (POL == 1) { pol_first = lag_policy_first(); pol_last = lag_policy_last(); }

    {
        const int bbase = (strip * L.sb + (int)blockIdx.x) * 128;
        const int jbase = bbase - L.dshift * 128;
        bool p_warp = p_plane && jbase >= 0 && jbase + wbase < nplane;   // warp-uniform

        // ---- collide part: loads ----
        const int idx = bbase + (int)threadIdx.x;
        const int ix = (int)((unsigned)idx % (unsigned)NX), iy = (int)((unsigned)idx / (unsigned)NX);
        const bool do_c = c_plane && idx < nplane && iy >= 2 && iy <= NY - 3 && ix >= 1 && ix <= NX - 2;
        double F[19], PH[19];
        if (do_c) {
            const pop_t *pf = fin + pop_index(g, zs, 0, idx);
            const double *pp = phi + (long long)zs * pl + idx;
            const int pl32 = (int)pl;            // 19 pl < 2^31 (set-up): 32-bit offsets from one 64-bit base each
            constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
            for (int k = 0; k < 19; k++) {
                F[k] = POL == 1 ? lag_ld_hint(pf + k * pl32, pol_first) : (double)__ldg(pf + k * pl32);
                PH[k] = __ldg(pp + (CZ[k] * pl32 + CY[k] * NX + CX[k]));
            }
        }
        if (c_plane && bbase + wbase < nplane)
            lag_copy_warp_rows(gs + wbase, gin + pop_index(g, zs, 0, bbase + wbase), pl, pl - (bbase + wbase), lane);
        asm volatile("cp.async.commit_group;\n" ::: "memory");

        // ---- phi part, first chance: if the producers have finished (the normal case) request the g* rows now, so
        // that they arrive during the collision; otherwise do not wait here -- collide first, look again afterwards
        bool early = false;
        if (PHI) {
            early = p_warp && __all_sync(0xffffffffu, seen >= L.full);
            if (early) lag_copy_warp_rows(ps + wbase, gout + pop_index(g, zp, 0, jbase + wbase), pl, pl - (jbase + wbase), lane);
            asm volatile("cp.async.commit_group;\n" ::: "memory");
        }

        // ---- collide part: arithmetic and push ----
        if (PHI) asm volatile("cp.async.wait_group 1;\n" ::: "memory");   // the g rows (the first group) have landed ...
        else asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncwarp();                                                     // ... those the other lanes copied as well
        if (do_c) {
            double gx, gy, gz, lap;
            stencils(PH, gx, gy, gz, lap);
            const bool own = zs >= 2 && zs < g.nz + 2;
            if (WRITE && own && ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * pl + (long long)iy * NX);
            const pop_t *gcol = gs + threadIdx.x;
            auto getG = [gcol](int k) { return (double)gcol[k * 128]; };
            const long long base = pop_index(g, zs, 0, idx);
            double rho, ux, uy, uz;
            if (POL == 1)
                collide_streamed<FXY>(F, getG, PushXRim<HintedStore>{fout + base, gout + base, (int)pl, NX, ix == 1, ix == NX - 2, HintedStore{pol_first, pol_last}},
                                      PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
            else
                collide_streamed<FXY>(F, getG, PushXRim<PlainStore>{fout + base, gout + base, (int)pl, NX, ix == 1, ix == NX - 2, PlainStore{}},
                                      PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
            if (WRITE && own) {
                const long long at = (long long)(zs - 2) * pl + idx;
                o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
            }
        }

        if (PHI) {
            // ---- phi part, second chance: a whole collision later the producers are done unless something is badly
            // out of order; the wait is bounded, and a warp that gives up leaves the phase field to the repair kernel
            if (p_warp && !early) {
                unsigned spins = 0;
                bool ok;
                for (;;) {
                    if (need) seen = lag_ld_relaxed(cnt);
                    ok = __all_sync(0xffffffffu, seen >= L.full);
                    if (ok || ++spins > L.spin_limit) break;
                    if (__any_sync(0xffffffffu, lane == 6 && lag_ld_relaxed(flags) != 0u)) break;
                    __nanosleep(200);
                }
                if (ok) {
                    lag_copy_warp_rows(ps + wbase, gout + pop_index(g, zp, 0, jbase + wbase), pl, pl - (jbase + wbase), lane);
                    asm volatile("cp.async.commit_group;\n" ::: "memory");
                } else {
                    if (lane == 0) atomicOr(flags, 1u);
                    p_warp = false;
                }
            }
            // ---- phi part: the sums ----
            asm volatile("cp.async.wait_group 0;\n" ::: "memory");
            __syncwarp();
            const int jdx = jbase + (int)threadIdx.x;
            if (p_warp && jdx < nplane) {
                const int jx = (int)((unsigned)jdx % (unsigned)NX), jy = (int)((unsigned)jdx / (unsigned)NX);
                if (jx >= 1 && jx <= NX - 2 && jy >= 1 && jy <= NY - 2) {
                    double G2[19];
#pragma unroll
                    for (int k = 0; k < 19; k++) G2[k] = (double)ps[k * 128 + threadIdx.x];
                    phase_node(g, P, G2, phi_next, zp, jx, jy);
                }
            }
        }
    }
    // publish the end of this block's collide part: one release-add per block, after a block barrier (the release
    // of thread 0 then covers the stores of all four warps)
    if (PHI) {
        __syncthreads();
        if (threadIdx.x == 0) lag_add_release(done + (size_t)grp * LAG_PAD, 1u);
    }
}

}  // namespace lbm3
#endif
