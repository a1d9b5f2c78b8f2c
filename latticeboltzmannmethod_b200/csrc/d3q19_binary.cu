// d3q19_binary.cu -- D3Q19 binary-liquid z-slab on one GPU.
//
// Reference: Microchannel3D/{FullChannel,QuarterChannel}: Lattice<D> (lattice.h,
// lattice.hpp) and the per-step part of Solver<D> (solver.hpp:821-928).
// Per-thread arithmetic: d3q19_threads.h / d3q19_node.h.
#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <vector>

#ifndef LBM_POP_FLOAT
#include "d3q19_fused.cuh"   // the alternative kernels exist in the fp64 build only
#endif
#include "d3q19_threads.h"
#include "lbm_internal.h"

namespace lbm {
namespace {

using lbm3::FieldsOut;
using lbm3::Geom;
using lbm3::Params;

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) k_phase(Geom g, Params P, const pop_t *gpop, double *phi, int zs0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::phase_thread(g, P, gpop, phi, zs0 + blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

#ifndef LBM_POP_FLOAT   // alternative kernels: fp64 build only
template <bool FXY, bool WRITE>
__global__ void __launch_bounds__(kThreads) k_collide_push(Geom g, Params P, const pop_t *fin, const pop_t *gin,
                                                            const double *phi, pop_t *fout, pop_t *gout, FieldsOut o,
                                                            int zs0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::collide_push_thread<FXY, WRITE>(g, P, fin, gin, phi, fout, gout, o, zs0 + blockIdx.y, (int)(idx % g.nx),
                                          (int)(idx / g.nx));
}

template <bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(128, MINB) k_collide_push_occ(Geom g, Params P, const pop_t *fin, const pop_t *gin,
                                                                const double *phi, pop_t *fout, pop_t *gout, FieldsOut o,
                                                                int zs0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::collide_push_thread<FXY, WRITE>(g, P, fin, gin, phi, fout, gout, o, zs0 + blockIdx.y, (int)(idx % g.nx),
                                          (int)(idx / g.nx));
}

#endif  // LBM_POP_FLOAT
// Interior / rim split of the plain collide-push.  The interior kernel contains no boundary code at all
// (nodes 2..NX-3 x 2..NY-3: every push is a plain shifted store), which removes the control-flow merges
// and register moves the folded boundary path costs every thread; the rim kernel runs the one-node rim
// (and, on output steps, the wall-node fields) -- 2(NX+NY) nodes per plane.
template <bool FXY, bool WRITE, int MINB, bool XRIM = false, bool TILE2D = false, int NT = 128, bool TRT = false>
__global__ void __launch_bounds__(NT, MINB) k_collide_interior(Geom g, Params P, const pop_t *__restrict__ fin,
                                                                const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                                pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                                int zs0)
{
    const int NX = g.nx, NY = g.ny;
    long long idx;
    int ix, iy;
    if (TILE2D) {
        // 32 x 4 nodes per block instead of 128 x 1: the same coalescing (a warp is 32 consecutive iX), but the
        // block's phase stencils overlap in y as well, 612 distinct values instead of 1170
        const int tiles_x = (NX + 31) / 32;
        ix = (int)(blockIdx.x % tiles_x) * 32 + (threadIdx.x & 31);
        iy = (int)(blockIdx.x / tiles_x) * 4 + (threadIdx.x >> 5);
        if (ix >= NX || iy >= NY) return;
        idx = (long long)iy * NX + ix;
    } else {
        idx = (long long)blockIdx.x * NT + threadIdx.x;
        if (idx >= (long long)NX * NY) return;
        ix = (int)(idx % NX);
        iy = (int)(idx / NX);
    }
    const int zs = zs0 + blockIdx.y;
    // XRIM: the columns iX = 1 and iX = NX-2 are done here too (PART 3), only the rows iY = 1, NY-2 are left
    // to the rim kernel -- column accesses from the rim kernel cost a 32-byte sector per 8-byte value
    if (iy < 2 || iy > NY - 3 || ix < (XRIM ? 1 : 2) || ix > NX - (XRIM ? 2 : 3)) return;
    const long long pl = g.pl;
    double F[19], G[19], PH[19];
    const pop_t *pf = fin + lbm3::pop_index(g, zs, 0, idx), *pg = gin + lbm3::pop_index(g, zs, 0, idx);
    constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
    for (int k = 0; k < 19; k++) {
        F[k] = __ldg(pf + k * pl);
        G[k] = __ldg(pg + k * pl);
        PH[k] = __ldg(phi + (long long)(zs + CZ[k]) * pl + idx + CY[k] * NX + CX[k]);
    }
    lbm3::collide_push_node<FXY, WRITE, XRIM ? 3 : 1, TRT>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
}

template <bool FXY, bool WRITE, bool ROWS_ONLY = false, bool TRT = false>
__global__ void __launch_bounds__(128) k_collide_rim(Geom g, Params P, const pop_t *__restrict__ fin,
                                                     const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                     pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o, int zs0)
{
    const int NX = g.nx, NY = g.ny, zs = zs0 + blockIdx.y;
    const int rows = 2 * (NX - 2), cols = ROWS_ONLY ? 0 : 2 * (NY - 4 > 0 ? NY - 4 : 0);
    int t = blockIdx.x * 128 + threadIdx.x;
    int ix, iy;
    if (t < rows) {
        iy = t < NX - 2 ? 1 : NY - 2;
        ix = 1 + (t < NX - 2 ? t : t - (NX - 2));
        if (NY - 2 == 1 && t >= NX - 2) return;
    } else if (t < rows + cols) {
        t -= rows;
        const int h = NY - 4;
        ix = t < h ? 1 : NX - 2;
        iy = 2 + (t < h ? t : t - h);
        if (NX - 2 == 1 && t >= h) return;
    } else {
        if (!WRITE) return;
        t -= rows + cols;   // wall nodes, lattice.hpp:75-106: row NY-1 then column NX-1
        if (t < NX) { ix = t; iy = NY - 1; }
        else if (t < NX + NY - 1) { ix = NX - 1; iy = t - NX; }
        else return;
        if (zs >= 2 && zs < g.nz + 2) {
            const long long at = (long long)(zs - 2) * g.pl + (long long)iy * NX + ix;
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
        return;
    }
    const long long idx = (long long)iy * NX + ix, pl = g.pl;
    double F[19], G[19], PH[19];
    constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
    for (int k = 0; k < 19; k++) {
        F[k] = fin[lbm3::pop_index(g, zs, k, idx)];
        G[k] = gin[lbm3::pop_index(g, zs, k, idx)];
        PH[k] = phi[(long long)(zs + CZ[k]) * pl + idx + CY[k] * NX + CX[k]];
    }
    if (TRT) {   // the two-rate extension exists in the array form of the collision only
        lbm3::collide_push_node<FXY, WRITE, 0, true>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
        return;
    }
    // compile-time direction indices in the folded push (no table look-ups in local memory)
    const double *Gp = G;
    lbm3::collide_push_node_streamed<FXY, WRITE>(g, P, F, [Gp](int k) { return Gp[k]; }, PH, fout, gout, o, zs, ix, iy);
}

#ifndef LBM_POP_FLOAT   // alternative kernels: fp64 build only
// The interior kernel with g fetched where the collision consumes it (collide_streamed): f and the phase stencil
// are loaded up front, each g_k only when its post-collision value is formed and stored, so the 19 g registers
// are never live together.  Nodes 2..NX-3 x 2..NY-3 (PART 1); the one-node rim incl. the columns goes to k_collide_rim.
template <bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(128, MINB) k_collide_interior_sg(Geom g, Params P, const pop_t *__restrict__ fin,
                                                                   const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                                   pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                                   int zs0)
{
    const long long idx = (long long)blockIdx.x * 128 + threadIdx.x;
    const int NX = g.nx, NY = g.ny;
    if (idx >= (long long)NX * NY) return;
    const int ix = (int)(idx % NX), iy = (int)(idx / NX), zs = zs0 + blockIdx.y;
    if (iy < 2 || iy > NY - 3 || ix < 2 || ix > NX - 3) return;
    const long long pl = g.pl;
    double F[19], PH[19];
    const pop_t *pf = fin + lbm3::pop_index(g, zs, 0, idx), *pg = gin + lbm3::pop_index(g, zs, 0, idx);
    constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
    for (int k = 0; k < 19; k++) {
        F[k] = __ldg(pf + k * pl);
        PH[k] = __ldg(phi + (long long)(zs + CZ[k]) * pl + idx + CY[k] * NX + CX[k]);
    }
    lbm3::collide_push_node_streamed<FXY, WRITE, 1>(g, P, F, [pg, pl](int k) { return __ldg(pg + k * pl); }, PH, fout, gout, o, zs, ix, iy);
}

// Collide-push with every global load asynchronous and no block-wide barrier.
//
// ncu on k_collide_push (profiles/r01_collide_plain.*): 8 resident warps per SM (238 registers), top stall
// long-scoreboard, 33 % of the issue slots used, 4.7 TB/s.  The loads of a node are only in flight while its
// thread waits for them; during the ~1500 instructions of arithmetic nothing is.  Here a thread keeps one
// (x,y) and marches along z: as soon as the 57 values of plane z (19 f, 19 g, 19 neighbour phi) have been
// moved from its private shared-memory slots into registers, it issues the 57 cp.async (LDGSTS, 8 B) for
// plane z+1 into the same slots and then does the arithmetic of plane z.  Loads of the next node overlap the
// whole collision of the current one without holding registers; every thread only ever touches its own
// slots, so there is no __syncthreads and warps drift freely.
template <bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(128, MINB) k_collide_async(Geom g, Params P, const pop_t *__restrict__ fin,
                                                             const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                             pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                             int zs_begin, int zs_end, int chunk)
{
    extern __shared__ __align__(16) double slots[];   // [57][128]
    const int tid = threadIdx.x;
    const long long idx = (long long)blockIdx.x * 128 + tid;
    const int NX = g.nx, NY = g.ny;
    if (idx >= (long long)NX * NY) return;
    const int ix = (int)(idx % NX), iy = (int)(idx / NX);
    const int zs0 = zs_begin + blockIdx.y * chunk, zs1 = min(zs0 + chunk, zs_end);
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE && (ix == NX - 1 || iy == NY - 1))   // wall node fields, lattice.hpp:75-106
            for (int zs = max(zs0, 2); zs < min(zs1, g.nz + 2); zs++) {
                const long long at = (long long)(zs - 2) * g.pl + idx;
                o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
            }
        return;
    }
    double *slot = slots + tid;
    const long long pl = g.pl;
    auto issue = [&](int zs) {
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        const double *pf = fin + lbm3::pop_index(g, zs, 0, idx);
        const double *pg = gin + lbm3::pop_index(g, zs, 0, idx);
#pragma unroll
        for (int k = 0; k < 19; k++) {
            lbm3::cp_async8(slot + k * 128, pf + k * pl);
            lbm3::cp_async8(slot + (19 + k) * 128, pg + k * pl);
            lbm3::cp_async8(slot + (38 + k) * 128, phi + (long long)(zs + CZ[k]) * pl + idx + CY[k] * NX + CX[k]);
        }
        lbm3::cp_async_commit();
    };
    issue(zs0);
    for (int zs = zs0; zs < zs1; zs++) {
        double F[19], G[19], PH[19];
        lbm3::cp_async_wait<0>();
#pragma unroll
        for (int k = 0; k < 19; k++) {
            F[k] = slot[k * 128];
            G[k] = slot[(19 + k) * 128];
            PH[k] = slot[(38 + k) * 128];
        }
        if (zs + 1 < zs1) issue(zs + 1);
        lbm3::collide_push_node<FXY, WRITE>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
    }
}

// Plain collide-push through collide_streamed(): fewer live registers, more resident warps.
template <bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(128, MINB) k_collide_streamed(Geom g, Params P, const pop_t *__restrict__ fin,
                                                                const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                                pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                                int zs0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::collide_push_thread_streamed<FXY, WRITE>(g, P, fin, gin, phi, fout, gout, o, zs0 + blockIdx.y, (int)(idx % g.nx),
                                                   (int)(idx / g.nx));
}

// Plain collide-push with the f and the g populations of a node on two different threads: a block
// of 256 threads updates 128 nodes, warps 0-3 hold f, warps 4-7 hold g (d3q19_threads.h, "split").
template <bool FXY, bool WRITE, int MINB>
__global__ void __launch_bounds__(256, MINB) k_collide_split(Geom g, Params P, const pop_t *__restrict__ fin,
                                                             const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                             pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                             int zs0)
{
    __shared__ double xch[8][128];
    const int ln = threadIdx.x & 127;
    const bool is_g = threadIdx.x >= 128;
    const long long idx = (long long)blockIdx.x * 128 + ln;
    const int zs = zs0 + blockIdx.y;
    const int NX = g.nx, NY = g.ny;
    const bool in_dom = idx < (long long)NX * NY;
    const int ix = (int)(idx % NX), iy = (int)(idx / NX);
    const bool bulk = in_dom && ix >= 1 && ix <= NX - 2 && iy >= 1 && iy <= NY - 2;
    double V[19];
    double rho = 0, ux = 0, uy = 0, uz = 0;
    if (bulk) {
        if (!is_g) {
            const double *src = fin + lbm3::pop_index(g, zs, 0, idx);
#pragma unroll
            for (int k = 0; k < 19; k++) V[k] = __ldg(src + (long long)k * g.pl);
            lbm3::moments_shifted<FXY>(V, P, rho, ux, uy, uz);
            xch[0][ln] = ux; xch[1][ln] = uy; xch[2][ln] = uz;
        } else {
            const double *src = gin + lbm3::pop_index(g, zs, 0, idx);
            double PH[19];
            constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
            for (int k = 0; k < 19; k++) {
                V[k] = __ldg(src + (long long)k * g.pl);
                PH[k] = __ldg(phi + (long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]);
            }
            double gx, gy, gz, lap;
            lbm3::stencils(PH, gx, gy, gz, lap);
            xch[3][ln] = PH[0]; xch[4][ln] = gx; xch[5][ln] = gy; xch[6][ln] = gz; xch[7][ln] = lap;
        }
    }
    __syncthreads();
    if (bulk) {
        if (!is_g)
            lbm3::f_thread_finish<FXY, WRITE>(g, P, V, xch[3][ln], xch[4][ln], xch[5][ln], xch[6][ln], xch[7][ln], rho, ux, uy, uz,
                                              fout, o, zs, ix, iy);
        else
            lbm3::g_thread_finish(g, P, V, xch[3][ln], xch[7][ln], xch[0][ln], xch[1][ln], xch[2][ln], gout, zs, ix, iy);
    } else if (WRITE && !is_g && in_dom && zs >= 2 && zs < g.nz + 2 && (ix == NX - 1 || iy == NY - 1)) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
    }
}

#endif  // LBM_POP_FLOAT
__global__ void k_init_fields(Geom g, Params P, double *phi, double *rho)
{
    const int z = blockIdx.x * blockDim.x + threadIdx.x;
    if (z < g.nz) lbm3::init_fields_plane(g, P, phi, rho, z);
}

__global__ void __launch_bounds__(kThreads) k_init_pops(Geom g, Params P, const double *phi, const double *rho,
                                                         const double *ux, const double *uy, const double *uz,
                                                         pop_t *fout, pop_t *gout)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::init_pops_thread(g, P, phi, rho, ux, uy, uz, fout, gout, blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

__global__ void __launch_bounds__(kThreads) k_after_init(Geom g, Params P, const pop_t *fin, FieldsOut o)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::after_init_thread(g, P, fin, o, blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

__global__ void __launch_bounds__(kThreads) k_gather_aos(Geom g, const pop_t *pop, double *aos)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::gather_aos_thread(g, pop, aos, blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

__global__ void __launch_bounds__(kThreads) k_scatter_aos(Geom g, const double *aos, pop_t *pop)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::scatter_aos_thread(g, aos, pop, blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

// ---- NCCL, loaded on demand so the library has no hard dependency on it ----
struct Id128 { char b[128]; };   // ncclUniqueId
static_assert(sizeof(Id128) == LBM_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(void **, int, Id128, int) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
constexpr int kNcclFloat64 = 8;  // ncclDouble
constexpr int kNcclFloat32 = 7;  // ncclFloat

NcclApi *nccl_api(std::string *why)
{
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (api.lib) {
            api.GetUniqueId = (int (*)(void *))dlsym(api.lib, "ncclGetUniqueId");
            api.CommInitRank = (int (*)(void **, int, Id128, int))dlsym(api.lib, "ncclCommInitRank");
            api.CommDestroy = (int (*)(void *))dlsym(api.lib, "ncclCommDestroy");
            api.GroupStart = (int (*)())dlsym(api.lib, "ncclGroupStart");
            api.GroupEnd = (int (*)())dlsym(api.lib, "ncclGroupEnd");
            api.Send = (int (*)(const void *, size_t, int, int, void *, cudaStream_t))dlsym(api.lib, "ncclSend");
            api.Recv = (int (*)(void *, size_t, int, int, void *, cudaStream_t))dlsym(api.lib, "ncclRecv");
            api.GetErrorString = (const char *(*)(int))dlsym(api.lib, "ncclGetErrorString");
        }
    }
    if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd) {
        if (why) *why = "NCCL (libnccl.so.2) could not be loaded";
        return nullptr;
    }
    return &api;
}

class Slab3D : public Model {
public:
    Slab3D(Runtime &rt, const lbm_config &cfg) : rt_(rt), cfg_(cfg) {}

    lbm_status setup()
    {
        if (cfg_.nx < 4 || cfg_.ny < 4) return rt_.fail(LBM_EINVAL, "D3Q19: nx and ny must be at least 4");
        if (cfg_.slab_count < 1 || cfg_.slab_rank < 0 || cfg_.slab_rank >= cfg_.slab_count)
            return rt_.fail(LBM_EINVAL, "bad slab_rank/slab_count");
        // solver.hpp:25-30
        const int zstep = cfg_.nz / cfg_.slab_count;
        z_begin_ = cfg_.slab_rank * zstep;
        int z_end = z_begin_ + zstep - 1;
        if (cfg_.slab_rank == cfg_.slab_count - 1) z_end = cfg_.nz - 1;
        g_.nx = cfg_.nx; g_.ny = cfg_.ny; g_.nz = z_end - z_begin_ + 1;
        if (g_.nz < 2) return rt_.fail(LBM_EINVAL, "D3Q19: every slab needs at least 2 z planes");
        g_.pl = ((long long)g_.nx * g_.ny + 15) / 16 * 16;
        P_.aconst = cfg_.aconst; P_.kconst = cfg_.kconst; P_.gammaconst = cfg_.gammaconst;
        P_.tau_gas = cfg_.tau_gas; P_.dtau = cfg_.tau_liq - cfg_.tau_gas;
        P_.omega_phi = 1.0 / cfg_.tau_phi;
        P_.fx = cfg_.force[0]; P_.fy = cfg_.force[1]; P_.fz = cfg_.force[2];
        P_.phase_gradient = cfg_.wall_gradient; P_.rho_wall = cfg_.rho_wall;
        P_.trt_magic = cfg_.trt_magic;
        if (cfg_.trt_magic < 0.0) return rt_.fail(LBM_EINVAL, "trt_magic must be >= 0 (0 = BGK, the reference's collision)");
        if (cfg_.trt_magic > 0.0 && cfg_.kernel_variant != 0)
            return rt_.fail(LBM_EINVAL, "the TRT extension exists for kernel_variant 0 only");
        fxy_ = P_.fx != 0.0 || P_.fy != 0.0;

        const size_t pop_bytes = (size_t)(g_.nz + 4) * 19 * g_.pl * sizeof(pop_t);
        for (int b = 0; b < 2; b++) {
            LBM_CUDA(rt_, cudaMalloc(&f_[b], pop_bytes));
            LBM_CUDA(rt_, cudaMalloc(&gq_[b], pop_bytes));
            LBM_CUDA(rt_, cudaMemsetAsync(f_[b], 0, pop_bytes, rt_.stream));
            LBM_CUDA(rt_, cudaMemsetAsync(gq_[b], 0, pop_bytes, rt_.stream));
        }
        const size_t phi_bytes = (size_t)(g_.nz + 4) * g_.pl * sizeof(double);
        LBM_CUDA(rt_, cudaMalloc(&phi_, phi_bytes));
        LBM_CUDA(rt_, cudaMemsetAsync(phi_, 0, phi_bytes, rt_.stream));
        const size_t fld_bytes = (size_t)g_.nz * g_.pl * sizeof(double);
        double **fl[4] = {&rho_, &ux_, &uy_, &uz_};
        for (auto p : fl) {
            LBM_CUDA(rt_, cudaMalloc(p, fld_bytes));
            LBM_CUDA(rt_, cudaMemsetAsync(*p, 0, fld_bytes, rt_.stream));
        }
        LBM_CUDA(rt_, cudaMalloc(&diag_, (size_t)(diagnostics_scratch_blocks() + 1) * 6 * sizeof(double)));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }

    ~Slab3D() override
    {
        cudaSetDevice(rt_.device);
        if (comm_) {
            NcclApi *api = nccl_api(nullptr);
            if (api && api->CommDestroy) api->CommDestroy(comm_);
        }
        if (side_) { cudaStreamDestroy(side_); cudaEventDestroy(ev_fork_); cudaEventDestroy(ev_join_); }
        for (int b = 0; b < 2; b++) { cudaFree(f_[b]); cudaFree(gq_[b]); }
        cudaFree(phi_); cudaFree(rho_); cudaFree(ux_); cudaFree(uy_); cudaFree(uz_); cudaFree(diag_);
    }

    size_t local_nodes() const override { return (size_t)g_.nx * g_.ny * g_.nz; }
    int q() const override { return 19; }
    int nsets() const override { return 2; }
    lbm_status slab_extent(int *zb, int *nzl) const override { *zb = z_begin_; *nzl = g_.nz; return LBM_OK; }

    double *field_ptr(int field)
    {
        switch (field) {
        case LBM_RHO: return rho_;
        case LBM_PHASE: return phi_ + 2 * g_.pl;
        case LBM_UX: return ux_;
        case LBM_UY: return uy_;
        case LBM_UZ: return uz_;
        }
        return nullptr;
    }
    lbm_status set_field(int field, const double *host) override
    {
        double *d = field_ptr(field);
        if (!d) return rt_.fail(LBM_EINVAL, "unknown field");
        const size_t row = (size_t)g_.nx * g_.ny * sizeof(double);
        LBM_CUDA(rt_, cudaMemcpy2DAsync(d, g_.pl * sizeof(double), host, row, row, g_.nz, cudaMemcpyHostToDevice, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }
    lbm_status get_field(int field, double *host) override
    {
        double *d = field_ptr(field);
        if (!d) return rt_.fail(LBM_EINVAL, "unknown field");
        const size_t row = (size_t)g_.nx * g_.ny * sizeof(double);
        LBM_CUDA(rt_, cudaMemcpy2DAsync(host, row, d, g_.pl * sizeof(double), row, g_.nz, cudaMemcpyDeviceToHost, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }

    dim3 plane_grid(int planes) const { return dim3((unsigned)(((long long)g_.nx * g_.ny + kThreads - 1) / kThreads), planes); }
    FieldsOut fields_out() const { return FieldsOut{rho_, ux_, uy_, uz_}; }

    // ---- slab phases ----
    lbm_status slab_init_fields() override
    {
        k_init_fields<<<(g_.nz + 63) / 64, 64, 0, rt_.stream>>>(g_, P_, phi_, rho_);
        return rt_.launched("k_init_fields");
    }
    lbm_status slab_init_pops() override
    {
        const int b = 1 - cur_;
        k_init_pops<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, P_, phi_, rho_, ux_, uy_, uz_, f_[b], gq_[b]);
        return rt_.launched("k_init_pops");
    }
    lbm_status slab_after_init() override
    {
        k_phase<<<plane_grid(g_.nz + 4), kThreads, 0, rt_.stream>>>(g_, P_, gq_[cur_], phi_, 0);
        LBM_TRY(rt_.launched("k_phase"));
        k_after_init<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, P_, f_[cur_], fields_out());
        return rt_.launched("k_after_init");
    }
#ifndef LBM_POP_FLOAT   // alternative kernels: fp64 build only
    // kernel_variant 0: the whole step as one kernel (d3q19_fused.cuh)
    template <int TY>
    lbm_status launch_fused(int write_fields)
    {
        using S = lbm3::FusedSmem<TY>;
        const int a = cur_, b = 1 - cur_;
        const int planes = g_.nz + 2;
        const int gx = (g_.nx + lbm3::FT_TX - 1) / lbm3::FT_TX, gy = (g_.ny + TY - 1) / TY;
        int nch = std::max((planes + 63) / 64, (148 * 8 + gx * gy - 1) / (gx * gy));
        nch = std::max(1, std::min(nch, planes / 4));
        const int chunk = (planes + nch - 1) / nch;
        nch = (planes + chunk - 1) / chunk;
        const dim3 grid(gx, gy, nch);
        const FieldsOut o = fields_out();
#define LBM_LAUNCH_FUSED(FXY, WR)                                                                                        \
    do {                                                                                                                 \
        auto kern = lbm3::k_fused_step<TY, FXY, WR>;                                                                     \
        LBM_CUDA(rt_, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));          \
        rt_.about_to_launch();                                                                                           \
        kern<<<grid, S::NT, S::bytes, rt_.stream>>>(g_, P_, f_[a], gq_[a], f_[b], gq_[b], o, phi_, 1, g_.nz + 3, chunk); \
    } while (0)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_FUSED(true, true); else LBM_LAUNCH_FUSED(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_FUSED(false, true); else LBM_LAUNCH_FUSED(false, false);
        }
#undef LBM_LAUNCH_FUSED
        return rt_.launched("k_fused_step");
    }

    // kernel_variant 0/2: fused, occupancy-oriented (k_fused2_step, TY = 8 / 4)
    template <int TY, int MINB, bool FEARLY>
    lbm_status launch_fused2(int write_fields)
    {
        using S = lbm3::Fused2Smem<TY>;
        const int a = cur_, b = 1 - cur_;
        const int planes = g_.nz + 2;
        const int gx = (g_.nx + lbm3::FT_TX - 1) / lbm3::FT_TX, gy = (g_.ny + TY - 1) / TY;
        int nch = std::max((planes + 63) / 64, (148 * 16 + gx * gy - 1) / (gx * gy));
        nch = std::max(1, std::min(nch, planes / 4));
        const int chunk = (planes + nch - 1) / nch;
        nch = (planes + chunk - 1) / chunk;
        const dim3 grid(gx, gy, nch);
        const FieldsOut o = fields_out();
#define LBM_LAUNCH_FUSED2(FXY, WR)                                                                                       \
    do {                                                                                                                 \
        auto kern = lbm3::k_fused2_step<TY, FXY, WR, MINB, FEARLY>;                                                                   \
        LBM_CUDA(rt_, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));          \
        rt_.about_to_launch();                                                                                           \
        kern<<<grid, S::NT, S::bytes, rt_.stream>>>(g_, P_, f_[a], gq_[a], f_[b], gq_[b], o, phi_, 1, g_.nz + 3, chunk); \
    } while (0)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_FUSED2(true, true); else LBM_LAUNCH_FUSED2(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_FUSED2(false, true); else LBM_LAUNCH_FUSED2(false, false);
        }
#undef LBM_LAUNCH_FUSED2
        return rt_.launched("k_fused2_step");
    }

    // kernel_variant 0: fused interior kernel + rim kernel
    template <int TY, int MINB, int GEN>
    lbm_status launch_fused3(int write_fields)
    {
        const size_t smem_bytes = GEN == 3 ? lbm3::Fused2Smem<TY>::bytes : (GEN == 4 ? lbm3::Fused4Smem<TY>::bytes : lbm3::Fused5Smem<TY>::bytes);
        constexpr int NT = lbm3::FT_TX * TY;
        const int a = cur_, b = 1 - cur_;
        const int planes = g_.nz + 2;
        const int gx = (g_.nx + lbm3::FT_TX - 1) / lbm3::FT_TX, gy = (g_.ny + TY - 1) / TY;
        int nch = std::max((planes + 63) / 64, (148 * 16 + gx * gy - 1) / (gx * gy));
        nch = std::max(1, std::min(nch, planes / 4));
        const int chunk = (planes + nch - 1) / nch;
        nch = (planes + chunk - 1) / chunk;
        const dim3 grid(gx, gy, nch);
        const FieldsOut o = fields_out();
        const int rim = 2 * (g_.nx - 2) + 2 * std::max(0, g_.ny - 4) + g_.nx + g_.ny - 1;
        const dim3 gr((rim + 127) / 128, g_.nz + 2);
#define LBM_LAUNCH_FUSED3(FXY, WR)                                                                                       \
    do {                                                                                                                 \
        auto kern = GEN == 3 ? lbm3::k_fused3_step<TY, FXY, WR, MINB>                                                     \
                             : (GEN == 4 ? lbm3::k_fused4_step<TY, FXY, WR, MINB> : lbm3::k_fused5_step<TY, FXY, WR, MINB>); \
        LBM_CUDA(rt_, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));        \
        rt_.about_to_launch();                                                                                           \
        kern<<<grid, NT, smem_bytes, rt_.stream>>>(g_, P_, f_[a], gq_[a], f_[b], gq_[b], o, phi_, 1, g_.nz + 3, chunk);  \
        LBM_TRY(rt_.launched(GEN == 3 ? "k_fused3_step" : (GEN == 4 ? "k_fused4_step" : "k_fused5_step")));                                                                          \
        rt_.about_to_launch();                                                                                           \
        k_collide_rim<FXY, WR><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);             \
    } while (0)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_FUSED3(true, true); else LBM_LAUNCH_FUSED3(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_FUSED3(false, true); else LBM_LAUNCH_FUSED3(false, false);
        }
#undef LBM_LAUNCH_FUSED3
        return rt_.launched("k_collide_rim");
    }

#endif  // LBM_POP_FLOAT
    // default collide: interior incl. the x columns (168 registers, 12 warps/SM) + rim rows, for `count`
    // storage planes starting at zs0; the phase field of planes zs0-1 .. zs0+count must be up to date
    lbm_status collide_default(int zs0, int count, int write_fields)
    {
        if (count <= 0) return LBM_OK;
        const int a = cur_, b = 1 - cur_;
        const FieldsOut o = fields_out();
        const dim3 gi((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), count);
        const dim3 gt((unsigned)(((g_.nx + 31) / 32) * ((g_.ny + 3) / 4)), count);
        const dim3 g64((unsigned)(((long long)g_.nx * g_.ny + 63) / 64), count);
        const int rim = 2 * (g_.nx - 2) + g_.nx + g_.ny - 1;
        const dim3 gr((rim + 127) / 128, count);
#define LBM_LAUNCH_DEFAULT(FXY, WR)                                                                                            \
    do {                                                                                                                       \
        rt_.about_to_launch();                                                                                                 \
        if (P_.trt_magic > 0.0) {                                                                                              \
            k_collide_interior<FXY, WR, 3, true, false, 128, true><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
            LBM_TRY(rt_.launched("k_collide_interior"));                                                                       \
            rt_.about_to_launch();                                                                                             \
            k_collide_rim<FXY, WR, true, true><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
            break;                                                                                                             \
        }                                                                                                                      \
        if (cfg_.kernel_variant == 65)                                                                                         \
            k_collide_interior<FXY, WR, 2, true><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
        else if (cfg_.kernel_variant == 67)                                                                                    \
            k_collide_interior<FXY, WR, 6, true, false, 64><<<g64, 64, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
        else if (cfg_.kernel_variant == 66)                                                                                    \
            k_collide_interior<FXY, WR, 3, true, true><<<gt, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
        else                                                                                                                   \
            k_collide_interior<FXY, WR, 3, true><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0); \
        LBM_TRY(rt_.launched("k_collide_interior"));                                                                           \
        rt_.about_to_launch();                                                                                                 \
        k_collide_rim<FXY, WR, true><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, zs0);           \
    } while (0)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_DEFAULT(true, true); else LBM_LAUNCH_DEFAULT(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_DEFAULT(false, true); else LBM_LAUNCH_DEFAULT(false, false);
        }
#undef LBM_LAUNCH_DEFAULT
        return rt_.launched("k_collide_rim");
    }

#ifndef LBM_POP_FLOAT   // alternative kernels: fp64 build only
    // fused interior (phase pulled in, x columns folded) + rim rows
    template <int TY, int MINB>
    lbm_status launch_fused6(int write_fields)
    {
        using S = lbm3::Fused6Smem<TY>;
        const int a = cur_, b = 1 - cur_;
        const int planes = g_.nz + 2;
        const int gx = (g_.nx + lbm3::FT_TX - 1) / lbm3::FT_TX, gy = (g_.ny + TY - 1) / TY;
        int nch = std::max((planes + 63) / 64, (148 * 24 + gx * gy - 1) / (gx * gy));
        nch = std::max(1, std::min(nch, planes / 4));
        const int chunk = (planes + nch - 1) / nch;
        nch = (planes + chunk - 1) / chunk;
        const dim3 grid(gx, gy, nch);
        const FieldsOut o = fields_out();
        const int rim = 2 * (g_.nx - 2) + g_.nx + g_.ny - 1;
        const dim3 gr((rim + 127) / 128, g_.nz + 2);
#define LBM_LAUNCH_FUSED6(FXY, WR)                                                                                        \
    do {                                                                                                                  \
        auto kern = lbm3::k_fused6_step<TY, FXY, WR, MINB>;                                                               \
        LBM_CUDA(rt_, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));           \
        rt_.about_to_launch();                                                                                            \
        kern<<<grid, S::NT, S::bytes, rt_.stream>>>(g_, P_, f_[a], gq_[a], f_[b], gq_[b], o, phi_, 1, g_.nz + 3, chunk);  \
        LBM_TRY(rt_.launched("k_fused6_step"));                                                                           \
        rt_.about_to_launch();                                                                                            \
        k_collide_rim<FXY, WR, true><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);        \
    } while (0)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_FUSED6(true, true); else LBM_LAUNCH_FUSED6(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_FUSED6(false, true); else LBM_LAUNCH_FUSED6(false, false);
        }
#undef LBM_LAUNCH_FUSED6
        return rt_.launched("k_collide_rim");
    }

#endif  // LBM_POP_FLOAT
    lbm_status slab_collide(int write_fields) override
    {
#ifdef LBM_POP_FLOAT
        // fp32-storage build: the default path and its block-shape variants only
        if (cfg_.kernel_variant != 0 && cfg_.kernel_variant != 65 && cfg_.kernel_variant != 66 && cfg_.kernel_variant != 67)
            return rt_.fail(LBM_EINVAL, "this kernel_variant exists in the fp64 build only");
        rt_.about_to_launch();
        k_phase<<<plane_grid(g_.nz + 4), kThreads, 0, rt_.stream>>>(g_, P_, gq_[cur_], phi_, 0);
        LBM_TRY(rt_.launched("k_phase"));
        return collide_default(1, g_.nz + 2, write_fields);
#else
        switch (cfg_.kernel_variant) {
        case 70: return launch_fused3<4, 3, 3>(write_fields);
        case 71: return launch_fused3<4, 2, 3>(write_fields);
        case 72: return launch_fused3<8, 1, 3>(write_fields);
        case 100: return launch_fused6<4, 3>(write_fields);
        case 101: return launch_fused6<4, 2>(write_fields);
        case 102: return launch_fused6<8, 1>(write_fields);
        case 103: return launch_fused6<2, 6>(write_fields);
        case 90: return launch_fused3<4, 3, 5>(write_fields);
        case 91: return launch_fused3<4, 2, 5>(write_fields);
        case 92: return launch_fused3<8, 1, 5>(write_fields);
        case 93: return launch_fused3<4, 4, 5>(write_fields);
        case 80: return launch_fused3<4, 3, 4>(write_fields);
        case 81: return launch_fused3<4, 2, 4>(write_fields);
        case 82: return launch_fused3<8, 1, 4>(write_fields);
        case 3: return launch_fused2<4, 3, true>(write_fields);
        case 2: return launch_fused2<8, 1, true>(write_fields);
        case 20: return launch_fused2<4, 3, false>(write_fields);
        case 21: return launch_fused2<4, 4, true>(write_fields);
        case 22: return launch_fused2<4, 4, false>(write_fields);
        case 23: return launch_fused2<4, 2, true>(write_fields);
        case 24: return launch_fused2<8, 2, true>(write_fields);
        case 25: return launch_fused2<8, 2, false>(write_fields);
        default: break;
        }
        if (cfg_.kernel_variant == 10) return launch_fused<8>(write_fields);
        if (cfg_.kernel_variant == 12) return launch_fused<4>(write_fields);
        const int a = cur_, b = 1 - cur_;
        rt_.about_to_launch();
        k_phase<<<plane_grid(g_.nz + 4), kThreads, 0, rt_.stream>>>(g_, P_, gq_[a], phi_, 0);
        LBM_TRY(rt_.launched("k_phase"));
        const dim3 grid = plane_grid(g_.nz + 2);
        const FieldsOut o = fields_out();
        rt_.about_to_launch();
        if (cfg_.kernel_variant == 63 || cfg_.kernel_variant == 64) {
            // interior on the main stream, rim concurrently on a high-priority side stream: both read the
            // front buffers and the phase field, and every (node, direction) slot of the back buffers has
            // exactly one writer, so the two kernels are independent
            LBM_TRY(ensure_side_stream());
            const dim3 gi((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
            const int rim = 2 * (g_.nx - 2) + 2 * std::max(0, g_.ny - 4) + g_.nx + g_.ny - 1;
            const dim3 gr((rim + 127) / 128, g_.nz + 2);
            LBM_CUDA(rt_, cudaEventRecord(ev_fork_, rt_.stream));
            LBM_CUDA(rt_, cudaStreamWaitEvent(side_, ev_fork_, 0));
#define LBM_LAUNCH_SPLIT3(FXY, WR)                                                                                           \
    do {                                                                                                                     \
        k_collide_rim<FXY, WR><<<gr, 128, 0, side_>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);                      \
        LBM_TRY(rt_.launched("k_collide_rim"));                                                                              \
        rt_.about_to_launch();                                                                                               \
        if (cfg_.kernel_variant == 63)                                                                                       \
            k_collide_interior<FXY, WR, 3><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        else                                                                                                                 \
            k_collide_interior<FXY, WR, 2><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_SPLIT3(true, true); else LBM_LAUNCH_SPLIT3(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_SPLIT3(false, true); else LBM_LAUNCH_SPLIT3(false, false);
            }
#undef LBM_LAUNCH_SPLIT3
            LBM_TRY(rt_.launched("k_collide_interior"));
            LBM_CUDA(rt_, cudaEventRecord(ev_join_, side_));
            LBM_CUDA(rt_, cudaStreamWaitEvent(rt_.stream, ev_join_, 0));
            return LBM_OK;
        }
        if (cfg_.kernel_variant == 0 || cfg_.kernel_variant == 65 || cfg_.kernel_variant == 66 || cfg_.kernel_variant == 67) return collide_default(1, g_.nz + 2, write_fields);
        if (cfg_.kernel_variant == 68 || cfg_.kernel_variant == 69) {
            const dim3 gi((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
            const int rim = 2 * (g_.nx - 2) + 2 * std::max(0, g_.ny - 4) + g_.nx + g_.ny - 1;
            const dim3 gr((rim + 127) / 128, g_.nz + 2);
#define LBM_LAUNCH_SG(FXY, WR)                                                                                                  \
    do {                                                                                                                        \
        if (cfg_.kernel_variant == 68)                                                                                          \
            k_collide_interior_sg<FXY, WR, 4><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        else                                                                                                                    \
            k_collide_interior_sg<FXY, WR, 3><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        LBM_TRY(rt_.launched("k_collide_interior_sg"));                                                                         \
        rt_.about_to_launch();                                                                                                  \
        k_collide_rim<FXY, WR><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);                    \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_SG(true, true); else LBM_LAUNCH_SG(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_SG(false, true); else LBM_LAUNCH_SG(false, false);
            }
#undef LBM_LAUNCH_SG
            return rt_.launched("k_collide_rim");
        }
        if (cfg_.kernel_variant >= 60 && cfg_.kernel_variant <= 62) {
            const dim3 gi((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
            const int rim = 2 * (g_.nx - 2) + 2 * std::max(0, g_.ny - 4) + g_.nx + g_.ny - 1;
            const dim3 gr((rim + 127) / 128, g_.nz + 2);
#define LBM_LAUNCH_SPLIT2(FXY, WR)                                                                                           \
    do {                                                                                                                     \
        if (cfg_.kernel_variant == 60)                                                                                       \
            k_collide_interior<FXY, WR, 2><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        else if (cfg_.kernel_variant == 61)                                                                                  \
            k_collide_interior<FXY, WR, 3><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        else                                                                                                                 \
            k_collide_interior<FXY, WR, 4><<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);     \
        LBM_TRY(rt_.launched("k_collide_interior"));                                                                         \
        rt_.about_to_launch();                                                                                               \
        k_collide_rim<FXY, WR><<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);                 \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_SPLIT2(true, true); else LBM_LAUNCH_SPLIT2(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_SPLIT2(false, true); else LBM_LAUNCH_SPLIT2(false, false);
            }
#undef LBM_LAUNCH_SPLIT2
            return rt_.launched("k_collide_rim");
        }
        if (cfg_.kernel_variant >= 50 && cfg_.kernel_variant <= 52) {
            const int planes = g_.nz + 2;
            const int gxb = (int)(((long long)g_.nx * g_.ny + 127) / 128);
            int nch = std::max((planes + 31) / 32, (148 * 16 + gxb - 1) / gxb);
            nch = std::max(1, std::min(nch, planes / 4));
            const int chunk = (planes + nch - 1) / nch;
            nch = (planes + chunk - 1) / chunk;
            const dim3 grid2(gxb, nch);
            const size_t smem = 57 * 128 * sizeof(double);
#define LBM_LAUNCH_ASYNC1(FXY, WR, MB)                                                                                \
    do {                                                                                                              \
        auto kern = k_collide_async<FXY, WR, MB>;                                                                     \
        LBM_CUDA(rt_, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
        kern<<<grid2, 128, smem, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1, g_.nz + 3, chunk);   \
    } while (0)
#define LBM_LAUNCH_ASYNC(FXY, WR)                                                                                     \
    do {                                                                                                              \
        if (cfg_.kernel_variant == 50) LBM_LAUNCH_ASYNC1(FXY, WR, 2);                                                 \
        else if (cfg_.kernel_variant == 51) LBM_LAUNCH_ASYNC1(FXY, WR, 3);                                            \
        else LBM_LAUNCH_ASYNC1(FXY, WR, 4);                                                                           \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_ASYNC(true, true); else LBM_LAUNCH_ASYNC(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_ASYNC(false, true); else LBM_LAUNCH_ASYNC(false, false);
            }
#undef LBM_LAUNCH_ASYNC
#undef LBM_LAUNCH_ASYNC1
            return rt_.launched("k_collide_async");
        }
        if (cfg_.kernel_variant >= 40 && cfg_.kernel_variant <= 42) {
            const dim3 grid2((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
#define LBM_LAUNCH_STREAMED(FXY, WR)                                                                                         \
    do {                                                                                                                     \
        if (cfg_.kernel_variant == 40)                                                                                       \
            k_collide_streamed<FXY, WR, 3><<<grid2, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);  \
        else if (cfg_.kernel_variant == 41)                                                                                  \
            k_collide_streamed<FXY, WR, 4><<<grid2, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);  \
        else                                                                                                                 \
            k_collide_streamed<FXY, WR, 2><<<grid2, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);  \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_STREAMED(true, true); else LBM_LAUNCH_STREAMED(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_STREAMED(false, true); else LBM_LAUNCH_STREAMED(false, false);
            }
#undef LBM_LAUNCH_STREAMED
            return rt_.launched("k_collide_streamed");
        }
        if (cfg_.kernel_variant >= 30 && cfg_.kernel_variant <= 32) {
            const dim3 grid2((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
#define LBM_LAUNCH_SPLIT(FXY, WR)                                                                                          \
    do {                                                                                                                   \
        if (cfg_.kernel_variant == 30)                                                                                     \
            k_collide_split<FXY, WR, 1><<<grid2, 256, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);   \
        else if (cfg_.kernel_variant == 31)                                                                                \
            k_collide_split<FXY, WR, 2><<<grid2, 256, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);   \
        else                                                                                                               \
            k_collide_split<FXY, WR, 3><<<grid2, 256, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);   \
    } while (0)
            if (fxy_) {
                if (write_fields) LBM_LAUNCH_SPLIT(true, true); else LBM_LAUNCH_SPLIT(true, false);
            } else {
                if (write_fields) LBM_LAUNCH_SPLIT(false, true); else LBM_LAUNCH_SPLIT(false, false);
            }
#undef LBM_LAUNCH_SPLIT
            return rt_.launched("k_collide_split");
        }
        if ((cfg_.kernel_variant == 13 || cfg_.kernel_variant == 14) && !fxy_ && !write_fields) {
            const dim3 grid2((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), g_.nz + 2);
            if (cfg_.kernel_variant == 13)
                k_collide_push_occ<false, false, 3><<<grid2, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);
            else
                k_collide_push_occ<false, false, 4><<<grid2, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1);
            return rt_.launched("k_collide_push_occ");
        }
#define LBM_LAUNCH_COLLIDE(FXY, WR) \
    k_collide_push<FXY, WR><<<grid, kThreads, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phi_, f_[b], gq_[b], o, 1)
        if (fxy_) {
            if (write_fields) LBM_LAUNCH_COLLIDE(true, true); else LBM_LAUNCH_COLLIDE(true, false);
        } else {
            if (write_fields) LBM_LAUNCH_COLLIDE(false, true); else LBM_LAUNCH_COLLIDE(false, false);
        }
#undef LBM_LAUNCH_COLLIDE
        return rt_.launched("k_collide_push");
#endif
    }
    lbm_status slab_finish() override
    {
        cur_ = 1 - cur_;
        return LBM_OK;
    }
    // halo messages always refer to the back buffers (what collide / init_pops just wrote)
    lbm_status slab_halo(int id, void **ptr, size_t *count) override
    {
        const int b = 1 - cur_;
        const size_t pp = 19 * (size_t)g_.pl;
        auto at = [&](pop_t *arr, int z) { return arr + (size_t)(z + 2) * pp; };
        const int nz = g_.nz;
        void *p = nullptr;   // population messages are pop_t (double; float in the fp32-storage build), phase messages double
        size_t n = 0;
        switch (id) {
        case LBM_HALO_S_F_UP: p = at(f_[b], nz - 1); n = pp; break;
        case LBM_HALO_S_G_UP: p = at(gq_[b], nz - 2); n = 2 * pp; break;
        case LBM_HALO_S_F_DOWN: p = at(f_[b], 0); n = pp; break;
        case LBM_HALO_S_G_DOWN: p = at(gq_[b], 0); n = 2 * pp; break;
        case LBM_HALO_R_F_FROM_BELOW: p = at(f_[b], -1); n = pp; break;
        case LBM_HALO_R_G_FROM_BELOW: p = at(gq_[b], -2); n = 2 * pp; break;
        case LBM_HALO_R_F_FROM_ABOVE: p = at(f_[b], nz); n = pp; break;
        case LBM_HALO_R_G_FROM_ABOVE: p = at(gq_[b], nz); n = 2 * pp; break;
        case LBM_HALO_S_PHI_UP: p = phi_ + (size_t)(nz - 1 + 2) * g_.pl; n = g_.pl; break;
        case LBM_HALO_S_PHI_DOWN: p = phi_ + (size_t)2 * g_.pl; n = g_.pl; break;
        case LBM_HALO_R_PHI_FROM_BELOW: p = phi_ + (size_t)1 * g_.pl; n = g_.pl; break;
        case LBM_HALO_R_PHI_FROM_ABOVE: p = phi_ + (size_t)(nz + 2) * g_.pl; n = g_.pl; break;
        default: return rt_.fail(LBM_EINVAL, "unknown halo id");
        }
        *ptr = p;
        *count = n;
        return LBM_OK;
    }

    // ---- exchange: with itself (one slab, z periodic) or with the ring neighbours over NCCL ----
    lbm_status exchange(bool phi_only, cudaStream_t st = nullptr)
    {
        if (!st) st = rt_.stream;
        static const int kState[4][2] = {{LBM_HALO_S_F_UP, LBM_HALO_R_F_FROM_BELOW}, {LBM_HALO_S_G_UP, LBM_HALO_R_G_FROM_BELOW},
                                         {LBM_HALO_S_F_DOWN, LBM_HALO_R_F_FROM_ABOVE}, {LBM_HALO_S_G_DOWN, LBM_HALO_R_G_FROM_ABOVE}};
        static const int kPhi[2][2] = {{LBM_HALO_S_PHI_UP, LBM_HALO_R_PHI_FROM_BELOW}, {LBM_HALO_S_PHI_DOWN, LBM_HALO_R_PHI_FROM_ABOVE}};
        const int(*pairs)[2] = phi_only ? kPhi : kState;
        const int npairs = phi_only ? 2 : 4;
        if (cfg_.slab_count == 1) {
            for (int i = 0; i < npairs; i++) {
                void *s, *r;
                size_t ns, nr;
                LBM_TRY(slab_halo(pairs[i][0], &s, &ns));
                LBM_TRY(slab_halo(pairs[i][1], &r, &nr));
                LBM_CUDA(rt_, cudaMemcpyAsync(r, s, ns * (phi_only ? sizeof(double) : sizeof(pop_t)), cudaMemcpyDeviceToDevice, st));
            }
            return LBM_OK;
        }
        if (!comm_) return rt_.fail(LBM_ESTATE, "slab_count > 1: call lbm_comm_attach first, or drive lbm_slab_* yourself");
        NcclApi *api = nccl_api(&rt_.err);
        if (!api) return LBM_ENCCL;
        const int up = (cfg_.slab_rank + 1) % cfg_.slab_count, down = (cfg_.slab_rank - 1 + cfg_.slab_count) % cfg_.slab_count;
        const int dtype = (phi_only || sizeof(pop_t) == sizeof(double)) ? kNcclFloat64 : kNcclFloat32;
        int rc = api->GroupStart();
        // sends in the order [up..., down...], receives [from below..., from above...]: with two
        // slabs both neighbours are the same peer and NCCL matches a pair's messages in order
        for (int i = 0; i < npairs && rc == 0; i++) {
            void *s;
            size_t ns;
            LBM_TRY(slab_halo(pairs[i][0], &s, &ns));
            const bool to_up = i < npairs / 2;
            rc = api->Send(s, ns, dtype, to_up ? up : down, comm_, st);
        }
        for (int i = 0; i < npairs && rc == 0; i++) {
            void *r;
            size_t nr;
            LBM_TRY(slab_halo(pairs[i][1], &r, &nr));
            const bool from_below = i < npairs / 2;
            rc = api->Recv(r, nr, dtype, from_below ? down : up, comm_, st);
        }
        const int rc2 = api->GroupEnd();
        if (rc == 0) rc = rc2;
        if (rc != 0) return rt_.fail(LBM_ENCCL, std::string("NCCL halo exchange: ") + (api->GetErrorString ? api->GetErrorString(rc) : "error"));
        return LBM_OK;
    }

    lbm_status comm_attach(const void *id) override
    {
        if (cfg_.slab_count == 1) return LBM_OK;
        NcclApi *api = nccl_api(&rt_.err);
        if (!api) return LBM_ENCCL;
        Id128 uid;
        memcpy(&uid, id, sizeof uid);
        const int rc = api->CommInitRank(&comm_, cfg_.slab_count, uid, cfg_.slab_rank);
        if (rc != 0) {
            comm_ = nullptr;
            return rt_.fail(LBM_ENCCL, std::string("ncclCommInitRank: ") + (api->GetErrorString ? api->GetErrorString(rc) : "error"));
        }
        return LBM_OK;
    }

    lbm_status init() override
    {
        LBM_TRY(slab_init_fields());
        LBM_TRY(exchange(true));
        LBM_TRY(slab_init_pops());
        LBM_TRY(exchange(false));
        LBM_TRY(slab_finish());
        return slab_after_init();
    }
    // One step with the halo exchange hidden behind the bulk of the collision (replaces the blocking
    // MPI_Sendrecv chain of solver.hpp:893-928): phase field, then the four planes at either end of the
    // slab -- everything the outgoing messages depend on --, then the exchange on a side stream while the
    // main stream collides the planes in between.
    lbm_status step_overlapped(int write_fields)
    {
        LBM_TRY(ensure_side_stream());
        const int nz = g_.nz;
        rt_.about_to_launch();
        k_phase<<<plane_grid(nz + 4), kThreads, 0, rt_.stream>>>(g_, P_, gq_[cur_], phi_, 0);
        LBM_TRY(rt_.launched("k_phase"));
        LBM_TRY(collide_default(1, 4, write_fields));            // z = -1 .. 2
        LBM_TRY(collide_default(nz - 1, 4, write_fields));       // z = nz-3 .. nz
        LBM_CUDA(rt_, cudaEventRecord(ev_fork_, rt_.stream));
        LBM_CUDA(rt_, cudaStreamWaitEvent(side_, ev_fork_, 0));
        LBM_TRY(exchange(false, side_));
        LBM_CUDA(rt_, cudaEventRecord(ev_join_, side_));
        LBM_TRY(collide_default(5, nz - 6, write_fields));       // z = 3 .. nz-4
        LBM_CUDA(rt_, cudaStreamWaitEvent(rt_.stream, ev_join_, 0));
        return slab_finish();
    }

    lbm_status step(int nsteps) override
    {
        const bool overlap = comm_ && cfg_.slab_count > 1 && (cfg_.kernel_variant == 0 || cfg_.kernel_variant == 65 || cfg_.kernel_variant == 66 || cfg_.kernel_variant == 67) && g_.nz >= 8;
        for (int t = 0; t < nsteps; t++) {
            if (overlap) {
                LBM_TRY(step_overlapped(t == nsteps - 1));
                continue;
            }
            LBM_TRY(slab_collide(t == nsteps - 1));
            LBM_TRY(exchange(false));
            LBM_TRY(slab_finish());
        }
        return LBM_OK;
    }

    lbm_status set_populations(int set, const double *host) override
    {
        const size_t n = local_nodes() * 19;
        double *stage = nullptr;
        LBM_CUDA(rt_, cudaMalloc(&stage, n * sizeof(double)));
        lbm_status st = rt_.cuda(cudaMemcpyAsync(stage, host, n * sizeof(double), cudaMemcpyHostToDevice, rt_.stream), "H2D populations");
        if (st == LBM_OK) {
            pop_t *dst = set == 0 ? f_[cur_] : gq_[cur_];
            k_scatter_aos<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, stage, dst);
            st = rt_.launched("k_scatter_aos");
        }
        if (st == LBM_OK) st = rt_.cuda(cudaStreamSynchronize(rt_.stream), "sync");
        cudaFree(stage);
        LBM_TRY(st);
        // refresh the ghost planes of the front buffers: slab_halo() addresses the back buffers, so
        // flip, exchange, flip back
        cur_ = 1 - cur_;
        st = exchange(false);
        cur_ = 1 - cur_;
        return st;
    }
    lbm_status get_populations(int set, double *host) override
    {
        const size_t n = local_nodes() * 19;
        double *stage = nullptr;
        LBM_CUDA(rt_, cudaMalloc(&stage, n * sizeof(double)));
        k_gather_aos<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, set == 0 ? f_[cur_] : gq_[cur_], stage);
        lbm_status st = rt_.launched("k_gather_aos");
        if (st == LBM_OK) st = rt_.cuda(cudaMemcpyAsync(host, stage, n * sizeof(double), cudaMemcpyDeviceToHost, rt_.stream), "D2H populations");
        if (st == LBM_OK) st = rt_.cuda(cudaStreamSynchronize(rt_.stream), "sync");
        cudaFree(stage);
        return st;
    }

    lbm_status diagnostics(double *mass, double *mom, double *phase_mass, int *has_nan) override
    {
        DiagLayout L;
        const size_t pp = 19 * (size_t)g_.pl;
        L.f = f_[cur_] + 2 * pp;
        L.g = gq_[cur_] + 2 * pp;
        L.stride_k = g_.pl;
        L.plane_stride = (long long)pp;
        L.planes = g_.nz; L.nx = g_.nx; L.ny = g_.ny; L.q = 19; L.skip_border = 1; L.node_major_x = 1;
        double out[6];
        LBM_TRY(run_diagnostics(rt_, L, diag_, diagnostics_scratch_blocks(), out));
        if (mass) *mass = out[0];
        if (mom) { mom[0] = out[1]; mom[1] = out[2]; mom[2] = out[3]; }
        if (phase_mass) *phase_mass = out[4];
        if (has_nan) *has_nan = out[5] != 0.0;
        return LBM_OK;
    }

private:
    Runtime &rt_;
    lbm_config cfg_;
    Geom g_{};
    Params P_{};
    bool fxy_ = false;
    int z_begin_ = 0, cur_ = 0;
    pop_t *f_[2] = {nullptr, nullptr}, *gq_[2] = {nullptr, nullptr};
    double *phi_ = nullptr, *rho_ = nullptr, *ux_ = nullptr, *uy_ = nullptr, *uz_ = nullptr, *diag_ = nullptr;
    void *comm_ = nullptr;
    cudaStream_t side_ = nullptr;
    cudaEvent_t ev_fork_ = nullptr, ev_join_ = nullptr;

    lbm_status ensure_side_stream()
    {
        if (side_) return LBM_OK;
        int lo = 0, hi = 0;
        LBM_CUDA(rt_, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        LBM_CUDA(rt_, cudaStreamCreateWithPriority(&side_, cudaStreamNonBlocking, hi));
        LBM_CUDA(rt_, cudaEventCreateWithFlags(&ev_fork_, cudaEventDisableTiming));
        LBM_CUDA(rt_, cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming));
        return LBM_OK;
    }
};

}  // namespace

Model *make_d3q19_binary(Runtime &rt, const lbm_config &cfg, lbm_status *st)
{
    Slab3D *m = new Slab3D(rt, cfg);
    *st = m->setup();
    if (*st != LBM_OK) {
        delete m;
        return nullptr;
    }
    return m;
}

// used by lbm_comm_unique_id (lbm_api.cu)
lbm_status nccl_unique_id(void *out, std::string *why)
{
    NcclApi *api = nccl_api(why);
    if (!api) return LBM_ENCCL;
    const int rc = api->GetUniqueId(out);
    if (rc != 0) {
        if (why) *why = "ncclGetUniqueId failed";
        return LBM_ENCCL;
    }
    return LBM_OK;
}

}  // namespace lbm
