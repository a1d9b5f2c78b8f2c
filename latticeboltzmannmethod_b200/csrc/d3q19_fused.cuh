// d3q19_fused.cuh -- the D3Q19 binary-liquid step as ONE kernel.
//
// Reference work covered per launch: Lattice::updateMacro (lattice.hpp:4-108),
// the phase-field stencils, Lattice::collide_stream (lattice.hpp:377-688) --
// the same per-node arithmetic as the plain path (d3q19_threads.h), so the
// results are bit-identical to it.
//
// Why one kernel: the collision at x needs the phase field phi = sum_k g_k at
// the 18 neighbours of x.  The plain path computes phi for every node in a first
// kernel (reads g: 152 B per lattice update) and reads g again in the second.
// Here a block owns an x-y tile of 32 x TY nodes and marches along z.  For each
// plane it brings the g populations of the tile AND of its one-node halo ring
// into shared memory, sums phi for tile+halo there (4-plane ring buffer), and
// keeps the tile's own g in shared memory until that plane is collided one
// iteration later.  g is read from HBM once (plus the halo ring, which mostly
// hits L2 because neighbouring tiles read it at about the same time).
//
// Why cp.async: with 19+19 populations per node the kernel needs ~240
// registers, i.e. 8 resident warps per SM; with synchronous loads those warps
// all wait on HBM at the same time (ncu on the plain kernel: long-scoreboard is
// the top stall, 11 % warps active, 30 % of the issue slots used).  All global
// loads are therefore asynchronous copies into per-thread shared-memory slots
// issued one plane ahead (LDGSTS, 8 bytes each, any alignment), so arithmetic on
// plane z overlaps the loads of planes z+1 / z+2 without holding them in
// registers.  Every thread sums/consumes only slots it filled itself; the only
// cross-thread data is the phi ring, guarded by one __syncthreads per plane.
//
// Shared memory per block (TY = 8, 256 threads): f slot 38 KB, g slots 3 x 38 KB,
// halo/extra positions 2 x 19 KB, phi ring 11 KB  ->  ~204 KB, one block per SM.
#ifndef LBM_D3Q19_FUSED_CUH
#define LBM_D3Q19_FUSED_CUH

#include "d3q19_threads.h"

namespace lbm3 {

constexpr int FT_TX = 32;          // tile width = one warp per row
constexpr int FT_NE = 128;         // capacity of the extra-position list (halo ring + non-bulk tile nodes; <= 123)

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

template <int TY>
struct FusedSmem {
    static constexpr int NT = FT_TX * TY;                 // threads per block
    static constexpr int RW = FT_TX + 2, RH = TY + 2;     // phi ring plane: tile + halo
    static constexpr int RP = RW * RH;
    static constexpr size_t doubles = (size_t)19 * NT        // f slot
                                      + (size_t)3 * 19 * NT  // g slots
                                      + (size_t)2 * 19 * FT_NE  // extra positions, two stages
                                      + (size_t)4 * RP;      // phi ring
    static constexpr size_t bytes = doubles * sizeof(double) + (size_t)3 * FT_NE * sizeof(int) + 16;
};

// permuted 19-term sums read straight from a shared-memory column (stride `st` doubles between directions)
template <int MODE>
__device__ __forceinline__ double smem_sum19(const double *c, int st)
{
    // MODE 0: lattice.hpp:19-22 order; 1,2,3: the same sum over the x / y / xy mirrored copy (lattice.h:58-81)
    constexpr int ORD[4][19] = {{0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18},
                                {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17},
                                {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18},
                                {0, 2, 1, 4, 3, 5, 6, 10, 9, 8, 7, 12, 11, 14, 13, 16, 15, 18, 17}};
    double s = c[ORD[MODE][0] * st] + c[ORD[MODE][1] * st];
#pragma unroll
    for (int k = 2; k < 19; k++) s += c[ORD[MODE][k] * st];
    return s;
}

template <int TY, bool FXY, bool WRITE>
__global__ void __launch_bounds__(FT_TX *TY, 1)
    k_fused_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                 double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = FusedSmem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Fs = reinterpret_cast<double *>(smem_raw);      // [19][NT]
    double *Gs = Fs + 19 * NT;                               // [3][19][NT]
    double *Es = Gs + 3 * 19 * NT;                           // [2][19][FT_NE]
    double *ring = Es + 2 * 19 * FT_NE;                      // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);   // [FT_NE] in-plane index of the node whose g is summed
    int *e_ring = e_owner + FT_NE;                           // [FT_NE] offset in a ring plane
    int *e_flags = e_ring + FT_NE;                           // [FT_NE] bits 0-1 mirror mode, bit 2 wall, bit 3 inside tile
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const long long idx = (long long)iy * NX + ix;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    // ---- list of positions of the tile+halo plane whose phi does not come from a bulk thread of this tile ----
    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) | (inside ? 8 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;

    double *myF = Fs + tid;                     // direction stride NT
    auto myG = [&](int zs) { return Gs + (size_t)(zs % 3) * 19 * NT + tid; };
    auto myE = [&](int zs) { return Es + (size_t)(zs & 1) * 19 * FT_NE + tid; };  // direction stride FT_NE

    auto load_g_plane = [&](int zs) {
        if (bulk) {
            const double *src = gin + pop_index(g, zs, 0, idx);
            double *dst = myG(zs);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(dst + k * NT, src + (long long)k * g.pl);
        }
        if (has_extra) {
            const double *src = gin + pop_index(g, zs, 0, my_owner);
            double *dst = myE(zs);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(dst + k * FT_NE, src + (long long)k * g.pl);
        }
    };
    auto load_f_plane = [&](int zs) {
        if (bulk) {
            const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(myF + k * NT, src + (long long)k * g.pl);
        }
    };
    // phi of storage plane zs into the ring (and, on output steps, into the global phase field for owned planes)
    auto phase_plane = [&](int zs) {
        double *rp = ring + (size_t)(zs & 3) * RP;
        const bool store = WRITE && zs >= 2 && zs < g.nz + 2 && zs >= zs0 && zs < zs1;
        if (bulk) {
            const double v = smem_sum19<0>(myG(zs), NT);
            rp[(ty + 1) * RW + tx + 1] = v;
            if (store) phi_out[(long long)zs * g.pl + idx] = v;
        }
        if (has_extra) {
            const double *c = myE(zs);
            double v;
            switch (my_flags & 3) {
            case 0: v = smem_sum19<0>(c, FT_NE); break;
            case 1: v = smem_sum19<1>(c, FT_NE); break;
            case 2: v = smem_sum19<2>(c, FT_NE); break;
            default: v = smem_sum19<3>(c, FT_NE); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            rp[my_ring] = v;
            if (store && (my_flags & 8)) {
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * g.pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    // ---- prologue: phi(zs0-1), phi(zs0); g(zs0+1) and f(zs0) in flight ----
    load_g_plane(zs0 - 1);
    cp_async_commit();
    load_g_plane(zs0);
    cp_async_commit();
    cp_async_wait<1>();
    phase_plane(zs0 - 1);
    cp_async_wait<0>();
    phase_plane(zs0);
    load_g_plane(zs0 + 1);   // reuses the extra stage of plane zs0-1, which this thread has just consumed
    cp_async_commit();
    load_f_plane(zs0);
    cp_async_commit();

    for (int zs = zs0; zs < zs1; zs++) {
        // g(zs+1) has landed (f(zs) may still be in flight)
        cp_async_wait<1>();
        phase_plane(zs + 1);
        const bool more = zs + 1 < zs1;
        if (more) load_g_plane(zs + 2);
        cp_async_commit();
        cp_async_wait<1>();   // f(zs)
        __syncthreads();       // phi ring planes zs-1, zs, zs+1 complete

        if (bulk) {
            double F[19], G[19], PH[19];
#pragma unroll
            for (int k = 0; k < 19; k++) F[k] = myF[k * NT];
            {
                const double *gs = myG(zs);
#pragma unroll
                for (int k = 0; k < 19; k++) G[k] = gs[k * NT];
            }
            {
                constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
                for (int k = 0; k < 19; k++)
                    PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            }
            double gx, gy, gz, lap;
            stencils(PH, gx, gy, gz, lap);
            const bool own = zs >= 2 && zs < g.nz + 2;
            if (WRITE && own) {
                if (iy == 1) mirror_fields<2>(F, P, o, (long long)(zs - 2) * g.pl + ix);
                if (ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * g.pl + (long long)iy * NX);
                if (ix == 1 && iy == 1) mirror_fields<3>(F, P, o, (long long)(zs - 2) * g.pl);
            }
            double rho, ux, uy, uz;
            // the f slot is free once its 19 values sit in registers: start fetching the next plane into it
            if (more) load_f_plane(zs + 1);
            cp_async_commit();
            collide<FXY>(F, G, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
            if (WRITE && own) {
                const long long at = (long long)(zs - 2) * g.pl + idx;
                o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
            }
            const bool border = ix == 1 || iy == 1 || ix == NX - 2 || iy == NY - 2;
            if (!border) {
                constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
                for (int k = 0; k < 19; k++) {
                    const long long t = pop_index(g, zs + CZ[k], k, idx + CY[k] * NX + CX[k]);
                    fout[t] = F[k];
                    gout[t] = G[k];
                }
            } else {
                for (int k = 0; k < 19; k++) push_border(g, fout, gout, zs, ix, iy, k, F[k], G[k]);
            }
        } else {
            cp_async_commit();   // keep the group count identical across threads
            if (WRITE && in_dom && zs >= 2 && zs < g.nz + 2 && (ix == NX - 1 || iy == NY - 1)) {
                const long long at = (long long)(zs - 2) * g.pl + idx;   // wall node fields, lattice.hpp:75-106
                o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
            }
        }
    }
    cp_async_wait<0>();
}

// ---------------------------------------------------------------------------
// Second fused kernel: occupancy instead of asynchronous staging.
//
// ncu on k_fused_step above (profiles/r01_*): with 8 warps per SM the arithmetic
// itself is latency-bound -- "wait" (fixed-latency dependency) and "barrier" are
// the top stalls and only a quarter of the issue slots are used -- so hiding the
// HBM latency with cp.async did not help.  The fp64 dependency chains need more
// warps per scheduler.  This kernel therefore
//   * collides through collide_streamed() (d3q19_node.h): only the 19 f values
//     live in registers, g is read from shared memory one value at a time and
//     every result is stored the moment it exists -> <= 128 registers;
//   * keeps just two shared-memory slots of g per thread (the plane being
//     collided and the plane whose phi was just summed) plus the phi ring, so
//     four 128-thread blocks (TY=4) or two 256-thread blocks (TY=8) fit on an
//     SM: 16 resident warps;
//   * loads with plain LDG: the f loads of a plane are issued before the phi
//     phase and the barrier, the other warps cover the rest.
template <int TY>
struct Fused2Smem {
    static constexpr int NT = FT_TX * TY;
    static constexpr int RW = FT_TX + 2, RH = TY + 2, RP = RW * RH;
    static constexpr size_t doubles = (size_t)2 * 19 * NT + (size_t)4 * RP;
    static constexpr size_t bytes = doubles * sizeof(double) + (size_t)3 * FT_NE * sizeof(int) + 16;
};

template <int TY, bool FXY, bool WRITE, int MINB, bool FEARLY>
__global__ void __launch_bounds__(FT_TX *TY, MINB)
    k_fused2_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                  double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = Fused2Smem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gs = reinterpret_cast<double *>(smem_raw);       // [2][19][NT]
    double *ring = Gs + 2 * 19 * NT;                          // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);
    int *e_ring = e_owner + FT_NE;
    int *e_flags = e_ring + FT_NE;
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const long long idx = (long long)iy * NX + ix;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) | (inside ? 8 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;
    const long long pl = g.pl;

    // phi of storage plane zs: own node (g kept in the shared slot when `keep`) and this thread's extra position
    auto phase_plane = [&](int zs, bool keep) {
        double *rp = ring + (size_t)(zs & 3) * RP;
        const bool store = WRITE && zs >= 2 && zs < g.nz + 2 && zs >= zs0 && zs < zs1;
        double G[19], E[19];
        if (bulk) {
            const double *src = gin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) G[k] = __ldg(src + k * pl);
        }
        if (has_extra) {
            const double *src = gin + pop_index(g, zs, 0, my_owner);
#pragma unroll
            for (int k = 0; k < 19; k++) E[k] = __ldg(src + k * pl);
        }
        if (bulk) {
            const double v = sum19(G);
            rp[(ty + 1) * RW + tx + 1] = v;
            if (keep) {
                double *dst = Gs + (size_t)(zs & 1) * 19 * NT + tid;
#pragma unroll
                for (int k = 0; k < 19; k++) dst[k * NT] = G[k];
            }
            if (store) phi_out[(long long)zs * pl + idx] = v;
        }
        if (has_extra) {
            double v;
            switch (my_flags & 3) {
            case 0: v = sum19(E); break;
            case 1: v = sum19_mirror<1>(E); break;
            case 2: v = sum19_mirror<2>(E); break;
            default: v = sum19_mirror<3>(E); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            rp[my_ring] = v;
            if (store && (my_flags & 8)) {
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    phase_plane(zs0 - 1, false);
    phase_plane(zs0, true);

    for (int zs = zs0; zs < zs1; zs++) {
        double F[19];
        if (FEARLY && bulk) {   // issued early: in flight during the phi phase and the barrier
            const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) F[k] = __ldg(src + k * pl);
        }
        phase_plane(zs + 1, true);
        __syncthreads();   // phi ring planes zs-1, zs, zs+1 complete
        if (bulk) {
            if (!FEARLY) {
                const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
                for (int k = 0; k < 19; k++) F[k] = __ldg(src + k * pl);
            }
            double PH[19];
            {
                constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
                for (int k = 0; k < 19; k++)
                    PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            }
            const double *gs = Gs + (size_t)(zs & 1) * 19 * NT + tid;
            collide_push_node_streamed<FXY, WRITE>(g, P, F, [gs](int k) { return gs[k * NT]; }, PH, fout, gout, o, zs, ix, iy);
        } else if (WRITE && in_dom && zs >= 2 && zs < g.nz + 2 && (ix == NX - 1 || iy == NY - 1)) {
            const long long at = (long long)(zs - 2) * pl + idx;   // wall node fields, lattice.hpp:75-106
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
    }
}

// ---------------------------------------------------------------------------
// Third fused kernel = k_fused2_step restricted to the interior nodes (2..NX-3 x 2..NY-3): no boundary
// code in the collision, array form of the collision with all loads up front, 168 registers so that three
// 128-thread blocks (12 warps) share an SM.  The one-node rim is collided afterwards by k_collide_rim
// (d3q19_binary.cu), which reads the phase field from global memory; this kernel therefore records phi
// for the three-node band next to the mirror lines and walls on every step.
template <int TY, bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(FT_TX *TY, MINB)
    k_fused3_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                  double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = Fused2Smem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gs = reinterpret_cast<double *>(smem_raw);       // [2][19][NT]
    double *ring = Gs + 2 * 19 * NT;                          // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);
    int *e_ring = e_owner + FT_NE;
    int *e_flags = e_ring + FT_NE;
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const long long idx = (long long)iy * NX + ix;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) | (inside ? 8 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;
    const long long pl = g.pl;
    const bool band = ix <= 2 || ix >= NX - 3 || iy <= 2 || iy >= NY - 3;
    const bool interior = in_dom && ix >= 2 && ix <= NX - 3 && iy >= 2 && iy <= NY - 3;

    // phi of storage plane zs: own node (g kept in the shared slot when `keep`) and this thread's extra position
    auto phase_plane = [&](int zs, bool keep) {
        double *rp = ring + (size_t)(zs & 3) * RP;
        // which block records phi(zs) in global memory: the one whose chunk holds the plane (the first /
        // last chunk also take the outermost ghost planes).  Always for the three-node band along the
        // mirror lines and walls (the rim kernel's stencils), everywhere on output steps.
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        const bool store_all = WRITE && mine && zs >= 2 && zs < g.nz + 2;
        double G[19], E[19];
        if (bulk) {
            const double *src = gin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) G[k] = __ldg(src + k * pl);
        }
        if (has_extra) {
            const double *src = gin + pop_index(g, zs, 0, my_owner);
#pragma unroll
            for (int k = 0; k < 19; k++) E[k] = __ldg(src + k * pl);
        }
        if (bulk) {
            const double v = sum19(G);
            rp[(ty + 1) * RW + tx + 1] = v;
            if (keep) {
                double *dst = Gs + (size_t)(zs & 1) * 19 * NT + tid;
#pragma unroll
                for (int k = 0; k < 19; k++) dst[k * NT] = G[k];
            }
            if (store_all || (mine && band)) phi_out[(long long)zs * pl + idx] = v;
        }
        if (has_extra) {
            double v;
            switch (my_flags & 3) {
            case 0: v = sum19(E); break;
            case 1: v = sum19_mirror<1>(E); break;
            case 2: v = sum19_mirror<2>(E); break;
            default: v = sum19_mirror<3>(E); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            rp[my_ring] = v;
            if (mine && (my_flags & 8)) {   // non-bulk nodes inside the tile are all within the band
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    phase_plane(zs0 - 1, false);
    phase_plane(zs0, true);

    for (int zs = zs0; zs < zs1; zs++) {
        double F[19];
        if (interior) {   // issued early: in flight during the phi phase and the barrier
            const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) F[k] = __ldg(src + k * pl);
        }
        phase_plane(zs + 1, true);
        __syncthreads();   // phi ring planes zs-1, zs, zs+1 complete
        if (interior) {
            double G[19], PH[19];
            {
                constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
                for (int k = 0; k < 19; k++)
                    PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            }
            const double *gs = Gs + (size_t)(zs & 1) * 19 * NT + tid;
#pragma unroll
            for (int k = 0; k < 19; k++) G[k] = gs[k * NT];
            collide_push_node<FXY, WRITE, 1>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
        }
    }
}

// ---------------------------------------------------------------------------
// Fourth fused kernel: interior nodes only (rim by k_collide_rim), g and the extra (halo / non-bulk)
// positions prefetched with cp.async one plane ahead into shared memory -- no registers held by loads in
// flight --, f loaded with plain LDG at the top of the iteration, array form of the collision.
// Shared memory per 128-thread block (TY = 4): g 2 x 19.5 KB, extras 19.5 KB, phi ring 6.5 KB = 65 KB,
// so three blocks (12 warps) fit on an SM at <= 168 registers and one block's barrier is covered by the
// other two.
template <int TY>
struct Fused4Smem {
    static constexpr int NT = FT_TX * TY;
    static constexpr int RW = FT_TX + 2, RH = TY + 2, RP = RW * RH;
    static constexpr size_t doubles = (size_t)2 * 19 * NT + (size_t)19 * FT_NE + (size_t)4 * RP;
    static constexpr size_t bytes = doubles * sizeof(double) + (size_t)3 * FT_NE * sizeof(int) + 16;
};

template <int TY, bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(FT_TX *TY, MINB)
    k_fused4_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                  double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = Fused4Smem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gs = reinterpret_cast<double *>(smem_raw);       // [2][19][NT]
    double *Es = Gs + 2 * 19 * NT;                            // [19][FT_NE]
    double *ring = Es + 19 * FT_NE;                           // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);
    int *e_ring = e_owner + FT_NE;
    int *e_flags = e_ring + FT_NE;
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const bool interior = in_dom && ix >= 2 && ix <= NX - 3 && iy >= 2 && iy <= NY - 3;
    const bool band = ix <= 2 || ix >= NX - 3 || iy <= 2 || iy >= NY - 3;
    const long long idx = (long long)iy * NX + ix;
    const long long pl = g.pl;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) | (inside ? 8 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;

    auto myG = [&](int zs) { return Gs + (size_t)(zs & 1) * 19 * NT + tid; };
    double *myE = Es + tid;
    auto load_g = [&](int zs) {
        if (bulk) {
            const double *src = gin + pop_index(g, zs, 0, idx);
            double *dst = myG(zs);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(dst + k * NT, src + k * pl);
        }
    };
    auto load_e = [&](int zs) {
        if (has_extra) {
            const double *src = gin + pop_index(g, zs, 0, my_owner);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(myE + k * FT_NE, src + k * pl);
        }
    };
    auto phase_own = [&](int zs) {
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        if (bulk) {
            const double v = smem_sum19<0>(myG(zs), NT);
            ring[(size_t)(zs & 3) * RP + (ty + 1) * RW + tx + 1] = v;
            if (mine && (band || (WRITE && zs >= 2 && zs < g.nz + 2))) phi_out[(long long)zs * pl + idx] = v;
        }
    };
    auto phase_extra = [&](int zs) {
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        if (has_extra) {
            double v;
            switch (my_flags & 3) {
            case 0: v = smem_sum19<0>(myE, FT_NE); break;
            case 1: v = smem_sum19<1>(myE, FT_NE); break;
            case 2: v = smem_sum19<2>(myE, FT_NE); break;
            default: v = smem_sum19<3>(myE, FT_NE); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            ring[(size_t)(zs & 3) * RP + my_ring] = v;
            if (mine && (my_flags & 8)) {
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    // prologue: phi(zs0-1), phi(zs0); g(zs0) kept; g(zs0+1) and extras(zs0+1) in flight
    load_e(zs0 - 1);
    load_g(zs0 - 1);
    load_g(zs0);
    cp_async_commit();
    cp_async_wait<0>();
    phase_own(zs0 - 1);
    phase_extra(zs0 - 1);
    phase_own(zs0);
    load_e(zs0);
    cp_async_commit();
    cp_async_wait<0>();
    phase_extra(zs0);
    load_e(zs0 + 1);
    load_g(zs0 + 1);     // slot of plane zs0-1, already summed
    cp_async_commit();

    for (int zs = zs0; zs < zs1; zs++) {
        double F[19];
        if (interior) {
            const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) F[k] = __ldg(src + k * pl);
        }
        cp_async_wait<0>();          // g(zs+1), extras(zs+1)
        phase_own(zs + 1);
        phase_extra(zs + 1);
        const bool more = zs + 1 < zs1;
        if (more) load_e(zs + 2);    // the extras stage is free again
        __syncthreads();             // phi ring planes zs-1, zs, zs+1 complete
        double G[19];
        if (interior) {
            const double *gs = myG(zs);
#pragma unroll
            for (int k = 0; k < 19; k++) G[k] = gs[k * NT];
        }
        if (more) load_g(zs + 2);    // into the slot of plane zs, whose values now sit in registers
        cp_async_commit();
        if (interior) {
            double PH[19];
            constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
            for (int k = 0; k < 19; k++)
                PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            collide_push_node<FXY, WRITE, 1>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
        }
    }
    cp_async_wait<0>();
}

// ---------------------------------------------------------------------------
// Fifth fused kernel: as k_fused4_step, but the collision goes through collide_streamed() with g read
// from shared memory at the point of use (cheap there, unlike from global memory), so nothing but the 19 f
// values is held in registers; g gets a third slot so that plane zs+2 can be prefetched while plane zs is
// still being read; the extra positions are loaded synchronously (a minority of threads).
template <int TY>
struct Fused5Smem {
    static constexpr int NT = FT_TX * TY;
    static constexpr int RW = FT_TX + 2, RH = TY + 2, RP = RW * RH;
    static constexpr size_t doubles = (size_t)3 * 19 * NT + (size_t)4 * RP;
    static constexpr size_t bytes = doubles * sizeof(double) + (size_t)3 * FT_NE * sizeof(int) + 16;
};

template <int TY, bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(FT_TX *TY, MINB)
    k_fused5_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                  double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = Fused5Smem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gs = reinterpret_cast<double *>(smem_raw);       // [3][19][NT]
    double *ring = Gs + 3 * 19 * NT;                          // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);
    int *e_ring = e_owner + FT_NE;
    int *e_flags = e_ring + FT_NE;
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const bool interior = in_dom && ix >= 2 && ix <= NX - 3 && iy >= 2 && iy <= NY - 3;
    const bool band = ix <= 2 || ix >= NX - 3 || iy <= 2 || iy >= NY - 3;
    const long long idx = (long long)iy * NX + ix;
    const long long pl = g.pl;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) | (inside ? 8 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;

    auto myG = [&](int zs) { return Gs + (size_t)(zs % 3) * 19 * NT + tid; };
    auto load_g = [&](int zs) {
        if (bulk) {
            const double *src = gin + pop_index(g, zs, 0, idx);
            double *dst = myG(zs);
#pragma unroll
            for (int k = 0; k < 19; k++) cp_async8(dst + k * NT, src + k * pl);
        }
    };
    auto phase_own = [&](int zs) {
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        if (bulk) {
            const double v = smem_sum19<0>(myG(zs), NT);
            ring[(size_t)(zs & 3) * RP + (ty + 1) * RW + tx + 1] = v;
            if (mine && (band || (WRITE && zs >= 2 && zs < g.nz + 2))) phi_out[(long long)zs * pl + idx] = v;
        }
    };
    auto phase_extra = [&](int zs) {   // synchronous: only some threads have an extra position
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        if (has_extra) {
            double E[19];
            const double *src = gin + pop_index(g, zs, 0, my_owner);
#pragma unroll
            for (int k = 0; k < 19; k++) E[k] = __ldcg(src + k * pl);
            double v;
            switch (my_flags & 3) {
            case 0: v = sum19(E); break;
            case 1: v = sum19_mirror<1>(E); break;
            case 2: v = sum19_mirror<2>(E); break;
            default: v = sum19_mirror<3>(E); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            ring[(size_t)(zs & 3) * RP + my_ring] = v;
            if (mine && (my_flags & 8)) {
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    // prologue: phi(zs0-1), phi(zs0); g(zs0) kept, g(zs0+1) in flight
    load_g(zs0 - 1);
    load_g(zs0);
    cp_async_commit();
    load_g(zs0 + 1);
    cp_async_commit();
    phase_extra(zs0 - 1);
    phase_extra(zs0);
    cp_async_wait<1>();
    phase_own(zs0 - 1);
    phase_own(zs0);

    for (int zs = zs0; zs < zs1; zs++) {
        double F[19];
        if (interior) {
            const double *src = fin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) F[k] = __ldcg(src + k * pl);
        }
        const bool more = zs + 1 < zs1;
        if (more) load_g(zs + 2);    // third slot: plane zs-1 was collided in the previous iteration
        cp_async_commit();
        phase_extra(zs + 1);
        cp_async_wait<1>();          // g(zs+1)
        phase_own(zs + 1);
        __syncthreads();             // phi ring planes zs-1, zs, zs+1 complete
        if (interior) {
            double PH[19];
            constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
            for (int k = 0; k < 19; k++)
                PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            const double *gs = myG(zs);
            collide_push_node_streamed<FXY, WRITE, 1>(g, P, F, [gs](int k) { return gs[k * NT]; }, PH, fout, gout, o, zs, ix, iy);
        }
    }
    cp_async_wait<0>();
}

// ---------------------------------------------------------------------------
// Sixth fused kernel: the plain interior kernel (k_collide_interior, PART 3: x columns folded in, 168
// registers, three 128-thread blocks per SM) with the phase pass pulled in.  Only the phi ring lives in
// shared memory (6.5 KB per block), so nearly the whole 228 KB stay L1: the g populations of plane z+1 are
// loaded once for the phi sums (tile + halo positions) and again, one iteration later, for the collision --
// by then they sit in L1/L2, HBM sees them once.  Rows iY = 1 and NY-2 are left to k_collide_rim, which
// reads phi from global memory; this kernel records phi for the rows iY <= 2 and iY >= NY-3 on every step.
template <int TY>
struct Fused6Smem {
    static constexpr int NT = FT_TX * TY;
    static constexpr int RW = FT_TX + 2, RH = TY + 2, RP = RW * RH;
    static constexpr size_t bytes = (size_t)4 * RP * sizeof(double) + (size_t)3 * FT_NE * sizeof(int) + 16;
};

template <int TY, bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(FT_TX *TY, MINB)
    k_fused6_step(Geom g, Params P, const double *__restrict__ fin, const double *__restrict__ gin, double *__restrict__ fout,
                  double *__restrict__ gout, FieldsOut o, double *__restrict__ phi_out, int zs_begin, int zs_end, int chunk)
{
    using S = Fused6Smem<TY>;
    constexpr int NT = S::NT, RW = S::RW, RP = S::RP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *ring = reinterpret_cast<double *>(smem_raw);      // [4][RP]
    int *e_owner = reinterpret_cast<int *>(ring + 4 * RP);
    int *e_ring = e_owner + FT_NE;
    int *e_flags = e_ring + FT_NE;
    int *n_extra_p = e_flags + FT_NE;

    const int NX = g.nx, NY = g.ny;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int ix0 = blockIdx.x * FT_TX, iy0 = blockIdx.y * TY;
    const int ix = ix0 + tx, iy = iy0 + ty;
    const bool in_dom = ix < NX && iy < NY;
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    const bool mine_to_collide = bulk && iy >= 2 && iy <= NY - 3;     // rows 1 and NY-2: rim kernel
    const bool band = iy <= 2 || iy >= NY - 3;
    const long long idx = (long long)iy * NX + ix;
    const long long pl = g.pl;
    const int zs0 = zs_begin + blockIdx.z * chunk;
    const int zs1 = min(zs0 + chunk, zs_end);
    if (zs0 >= zs1) return;

    if (tid == 0) *n_extra_p = 0;
    __syncthreads();
    for (int p = tid; p < RP; p += NT) {
        const int px = p % RW, py = p / RW;
        const int ex = ix0 - 1 + px, ey = iy0 - 1 + py;
        if (ex < 0 || ex > NX - 1 || ey < 0 || ey > NY - 1) continue;
        const bool inside = px >= 1 && px <= FT_TX && py >= 1 && py <= TY;
        const bool ebulk = ex >= 1 && ex <= NX - 2 && ey >= 1 && ey <= NY - 2;
        if (inside && ebulk) continue;
        const int ox = min(max(ex, 1), NX - 2), oy = min(max(ey, 1), NY - 2);
        const int slot = atomicAdd(n_extra_p, 1);
        e_owner[slot] = oy * NX + ox;
        e_ring[slot] = p;
        e_flags[slot] = (ex == 0 ? 1 : 0) | (ey == 0 ? 2 : 0) | ((ex == NX - 1 || ey == NY - 1) ? 4 : 0) |
                        ((inside && (ey <= 2 || ey >= NY - 3)) ? 8 : 0) | (inside ? 16 : 0);
    }
    __syncthreads();
    const int n_extra = *n_extra_p;
    const bool has_extra = tid < n_extra;
    const int my_owner = has_extra ? e_owner[tid] : 0;
    const int my_ring = has_extra ? e_ring[tid] : 0;
    const int my_flags = has_extra ? e_flags[tid] : 0;

    auto phase_plane = [&](int zs) {
        double *rp = ring + (size_t)(zs & 3) * RP;
        const bool mine = (zs >= zs0 && zs < zs1) || (zs == zs0 - 1 && zs0 == zs_begin) || (zs == zs1 && zs1 == zs_end);
        const bool all = WRITE && mine && zs >= 2 && zs < g.nz + 2;
        if (bulk) {
            double G[19];
            const double *src = gin + pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) G[k] = src[k * pl];
            const double v = sum19(G);
            rp[(ty + 1) * RW + tx + 1] = v;
            if (all || (mine && band)) phi_out[(long long)zs * pl + idx] = v;
        }
        if (has_extra) {
            double E[19];
            const double *src = gin + pop_index(g, zs, 0, my_owner);
#pragma unroll
            for (int k = 0; k < 19; k++) E[k] = src[k * pl];
            double v;
            switch (my_flags & 3) {
            case 0: v = sum19(E); break;
            case 1: v = sum19_mirror<1>(E); break;
            case 2: v = sum19_mirror<2>(E); break;
            default: v = sum19_mirror<3>(E); break;
            }
            if (my_flags & 4) v = v - P.phase_gradient;
            rp[my_ring] = v;
            if ((mine && (my_flags & 8)) || (all && (my_flags & 16))) {
                const int px = my_ring % RW, py = my_ring / RW;
                phi_out[(long long)zs * pl + (long long)(iy0 - 1 + py) * NX + (ix0 - 1 + px)] = v;
            }
        }
    };

    phase_plane(zs0 - 1);
    phase_plane(zs0);
    for (int zs = zs0; zs < zs1; zs++) {
        phase_plane(zs + 1);
        __syncthreads();   // phi ring planes zs-1, zs, zs+1 complete
        if (mine_to_collide) {
            double F[19], G[19], PH[19];
            const double *pf = fin + pop_index(g, zs, 0, idx), *pg = gin + pop_index(g, zs, 0, idx);
            constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
            for (int k = 0; k < 19; k++) {
                F[k] = __ldg(pf + k * pl);
                G[k] = pg[k * pl];      // read for the phi sums one iteration ago: L1 / L2
                PH[k] = ring[(size_t)((zs + CZ[k]) & 3) * RP + (ty + 1 + CY[k]) * RW + tx + 1 + CX[k]];
            }
            collide_push_node<FXY, WRITE, 3>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
        }
    }
}

}  // namespace lbm3
#endif
