// d3q19_lag.cuh -- the D3Q19 binary-liquid step as ONE pass over HBM: collide-push of plane z and, in the same
// launch, the phase field of the NEXT step for a plane a few planes behind, read back while the freshly pushed
// g populations are still in L2.
//
// Reference work covered per launch: Lattice::collide_stream (lattice.hpp:377-688) for step t and the phase sum
// of Lattice::updateMacro (lattice.hpp:19-22, :75-106, :645-686) for step t+1.
//
// Why: the collision at x needs phi = sum_k g_k at the 18 neighbours of x, so phi of the whole neighbourhood must
// exist before x is collided.  The two-pass path forms phi in a kernel of its own that re-reads all of g from HBM
// (161 B per lattice update on top of the 608 the collision moves).  Here the grid walks the slab in narrow x-y
// strips, plane after plane: block b collides (strip s, plane z, segment i) and then sums phi_{t+1} for
// (strip s, plane z - LAG, segment i - D) from the g* values that the blocks before it pushed into the back buffer.
// Between the push and the read-back only ~2 x (resident blocks) x 128 nodes x 152 B = ~17 MB of g* are written,
// so the read-back is served by the 126 MB L2 and HBM sees the populations once in and once out.
//
// Ordering: blocks are dispatched in ascending blockIdx; a block's phi part depends only on collide parts of
// blocks (LAG-1) plane-steps or more before it.  Every warp publishes the end of its collide part with
// one release-add on its group's counter (group = the SB blocks of one strip and plane-step), and the
// phi part polls the six counters it depends on before it issues its loads.  The poll is bounded: a warp
// that gives up raises `flags[0]`, and the host-side step then recomputes phi with the plain kernel, so an
// unexpected dispatch order can cost time but cannot hang the GPU or change a result.
#ifndef LBM_D3Q19_LAG_CUH
#define LBM_D3Q19_LAG_CUH

#include "d3q19_threads.h"

namespace lbm3 {

struct LagPlan {
    int sb;          // blocks (of 128 in-plane nodes) per strip
    int pz;          // plane-steps per strip
    int nstrips;
    int z0;          // collide plane of plane-step 0 (storage plane index)
    int c0, c1;      // collide planes [c0, c1)
    int p0, p1;      // phi planes [p0, p1)
    int lag;         // phi plane = collide plane - lag
    int dshift;      // phi segment = collide segment - dshift
    unsigned warps_per_group;   // counter value of a finished group
    unsigned spin_limit;
    int ablate;      // measurement only: 1 no publish, 2 no wait, 4 no phi part (results are then wrong)
};

// 16 bytes global -> shared, L2 only (never through the non-coherent L1: the source was written by other SMs in
// this very launch)
__device__ __forceinline__ void lag_cp_async16(void *smem, const void *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
// The counter protocol.  Writer: all lanes store, __syncwarp, lane 0 adds with RELEASE semantics at GPU scope (the
// warp barrier makes the other lanes' stores part of what the release publishes).  Reader: RELAXED loads of the
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

template <bool FXY, bool WRITE, bool TRT, int POL>
__global__ void __launch_bounds__(128, 3)
k_collide_lag(Geom g, Params P, const pop_t *__restrict__ fin, const pop_t *__restrict__ gin, const double *__restrict__ phi,
              pop_t *__restrict__ fout, pop_t *gout, FieldsOut o, double *phi_next, unsigned *done, unsigned *flags, LagPlan L)
{
    extern __shared__ __align__(16) unsigned char lag_smem[];
    pop_t *slots = reinterpret_cast<pop_t *>(lag_smem);   // [19][128], every thread uses its own column
    const int NX = g.nx, NY = g.ny;
    const unsigned b = blockIdx.x;
    const int seg = (int)(b % (unsigned)L.sb);
    const unsigned grp = b / (unsigned)L.sb;
    const int zc = (int)(grp % (unsigned)L.pz), strip = (int)(grp / (unsigned)L.pz);
    const long long nplane = (long long)NX * NY;
    const long long pl = g.pl;

    // ---- collide part: node idx of plane zs ----
    const int zs = L.z0 + zc;
    const long long idx = ((long long)strip * L.sb + seg) * 128 + threadIdx.x;
    const int ix = (int)(idx % NX), iy = (int)(idx / NX);
    const bool do_c = zs >= L.c0 && zs < L.c1 && idx < nplane && iy >= 2 && iy <= NY - 3 && ix >= 1 && ix <= NX - 2;
    double F[19], G[19], PH[19];
    unsigned long long pol_first = 0, pol_last = 0;
    if (POL == 1) { pol_first = lag_policy_first(); pol_last = lag_policy_last(); }
    if (do_c) {
        const pop_t *pf = fin + pop_index(g, zs, 0, idx), *pg = gin + pop_index(g, zs, 0, idx);
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
        for (int k = 0; k < 19; k++) {
            if (POL == 1) {
                F[k] = lag_ld_hint(pf + k * pl, pol_first);
                G[k] = lag_ld_hint(pg + k * pl, pol_first);
            } else {
                F[k] = __ldg(pf + k * pl);
                G[k] = __ldg(pg + k * pl);
            }
            PH[k] = __ldg(phi + (long long)(zs + CZ[k]) * pl + idx + CY[k] * NX + CX[k]);
        }
    }

    // ---- phi part: the 128 nodes starting at jbase of plane zp; every warp fetches and sums its own 32 ----
    constexpr int EPC = 16 / (int)sizeof(pop_t);   // population values per 16-byte chunk
    constexpr int CPW = 32 / EPC;                   // chunks per direction for one warp's 32 nodes
    const int zp = zs - L.lag;
    const long long jbase = ((long long)strip * L.sb + seg - L.dshift) * 128;
    const long long jdx = jbase + threadIdx.x;
    const int lane = threadIdx.x & 31, wbase = threadIdx.x & ~31;
    bool p_warp = zp >= L.p0 && zp < L.p1 && jbase >= 0 && jbase + wbase < nplane && !(L.ablate & 4);   // warp-uniform
    const int jx = (int)(jdx % NX), jy = (int)(jdx / NX);
    if (p_warp && !(L.ablate & 2)) {
        // the g* of plane zp come from the collide parts of planes zp-1 .. zp+1, segments (this - 2*dshift) .. this:
        // groups (strip, zc-lag-1 .. zc-lag+1) and, for the first segments of a strip, the same of strip-1
        const int dz = lane % 3 - 1, ds = lane / 3;          // lanes 0..5
        const int zdep = zp + dz, sdep = strip - ds;
        const bool need = lane < 6 && zdep >= L.c0 && zdep < L.c1 && sdep >= 0;
        const unsigned *cnt = done + ((long long)sdep * L.pz + (zdep - L.z0));
        unsigned spins = 0;
        bool ok;
        for (;;) {
            const bool mine = !need || lag_ld_relaxed(cnt) >= L.warps_per_group;
            ok = __all_sync(0xffffffffu, mine);
            if (ok || ++spins > L.spin_limit) break;
            // somebody already gave up in this step: the phase field gets recomputed anyway, do not wait again
            if (__any_sync(0xffffffffu, lane == 6 && lag_ld_relaxed(flags) != 0u)) break;
            __nanosleep(200);
        }
        if (!ok) {
            if (lane == 0) atomicOr(flags, 1u);
            p_warp = false;
        }
    }
    if (p_warp) {
        const pop_t *src = gout + pop_index(g, zp, 0, jbase + wbase);
        pop_t *dst = slots + wbase;
        for (int c = lane; c < 19 * CPW; c += 32) {
            const int k = c / CPW, e = (c % CPW) * EPC;
            if (jbase + wbase + e < pl) lag_cp_async16(dst + k * 128 + e, src + k * pl + e);
        }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    const bool do_p = p_warp && jdx < nplane && jx >= 1 && jx <= NX - 2 && jy >= 1 && jy <= NY - 2;

    if (do_c) {
        if (POL == 1)
            collide_push_node<FXY, WRITE, 3, TRT>(g, P, F, G, PH, fout, gout, o, zs, ix, iy, HintedStore{pol_first, pol_last});
        else
            collide_push_node<FXY, WRITE, 3, TRT>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
    }

    // phi part, second half: the sums (the copies had the whole collision to arrive)
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
    if (do_p) {
        double G2[19];
#pragma unroll
        for (int k = 0; k < 19; k++) G2[k] = (double)slots[k * 128 + threadIdx.x];
        phase_node(g, P, G2, phi_next, zp, jx, jy);
    }

    // publish the end of this warp's collide part (every warp counts, with or without nodes)
    if (!(L.ablate & 1)) {
        __syncwarp();
        if (lane == 0) lag_add_release(done + grp, 1u);
    }
}

}  // namespace lbm3
#endif
