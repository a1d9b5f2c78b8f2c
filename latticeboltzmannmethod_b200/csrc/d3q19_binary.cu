// d3q19_binary.cu -- D3Q19 binary-liquid z-slab on one GPU.
//
// Reference: Microchannel3D/{FullChannel,QuarterChannel}: Lattice<D> (lattice.h,
// lattice.hpp) and the per-step part of Solver<D> (solver.hpp:821-928).
// Per-thread arithmetic: d3q19_threads.h / d3q19_node.h.
#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <vector>

#include <mutex>

#include "d3q19_inplace.cuh"
#include "d3q19_step.cuh"
#include "d3q19_threads.h"
#include "lbm_internal.h"

namespace lbm {
namespace {

using lbm3::FieldsOut;
using lbm3::k_aa_collide;
using lbm3::k_aa_core;
using lbm3::k_aa_merge;
using lbm3::k_aa_one;
using lbm3::k_aa_phase;
using lbm3::k_aa_repair;
using lbm3::k_aa_swap;
using lbm3::Geom;
using lbm3::Params;

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) k_phase(Geom g, Params P, const pop_t *gpop, double *phi, int zs0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::phase_thread(g, P, gpop, phi, zs0 + blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

// Plain collide-push, one thread per node of a whole plane incl. the boundary folding (kernel_variant 1): the
// straightforward form, kept as the cross-check of the default path.
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

// Interior / rim split of the collide-push.  The interior kernel takes every bulk node with 2 <= iY <= NY-3: no
// y-boundary code is compiled in, and the x mirror column (iX = 1) and the x wall column (iX = NX-2) are an address
// select on the ten directions with cx != 0 (PART 3 of collide_push_node), so they stay in the coalesced warps.
// The rim kernel runs the rows iY = 1 and NY-2 (and, on output steps, the wall-node fields).
// zsel: planes zs0 .. zs0+n0-1 come first in blockIdx.y, then planes zs1 .. (both ends of a slab in one launch).
template <bool FXY, bool WRITE, int COLL = 0>
__global__ void __launch_bounds__(128, 3) k_collide_interior(Geom g, Params P, const pop_t *__restrict__ fin,
                                                              const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                              pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                              int zs0, int n0, int zs1)
{
    const int NX = g.nx, NY = g.ny;
    const long long idx = (long long)blockIdx.x * 128 + threadIdx.x;
    if (idx >= (long long)NX * NY) return;
    const int ix = (int)(idx % NX), iy = (int)(idx / NX);
    const int zs = (int)blockIdx.y < n0 ? zs0 + (int)blockIdx.y : zs1 + ((int)blockIdx.y - n0);
    if (iy < 2 || iy > NY - 3 || ix < 1 || ix > NX - 2) return;
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
    lbm3::collide_push_node<FXY, WRITE, 3, COLL>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
}

template <bool FXY, bool WRITE, int COLL = 0>
__global__ void __launch_bounds__(128) k_collide_rim(Geom g, Params P, const pop_t *__restrict__ fin,
                                                     const pop_t *__restrict__ gin, const double *__restrict__ phi,
                                                     pop_t *__restrict__ fout, pop_t *__restrict__ gout, FieldsOut o,
                                                     int zs0, int n0, int zs1)
{
    const int NX = g.nx, NY = g.ny;
    const int zs = (int)blockIdx.y < n0 ? zs0 + (int)blockIdx.y : zs1 + ((int)blockIdx.y - n0);
    const int rows = 2 * (NX - 2);
    int t = blockIdx.x * 128 + threadIdx.x;
    int ix, iy;
    if (t < rows) {
        iy = t < NX - 2 ? 1 : NY - 2;
        ix = 1 + (t < NX - 2 ? t : t - (NX - 2));
        if (NY - 2 == 1 && t >= NX - 2) return;
    } else {
        if (!WRITE) return;
        t -= rows;   // wall nodes, lattice.hpp:75-106: row NY-1 then column NX-1
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
    if (COLL != 0) {   // the TRT and MRT extensions exist in the array form of the collision only
        lbm3::collide_push_node<FXY, WRITE, 0, COLL>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
        return;
    }
    // compile-time direction indices in the folded push (no table look-ups in local memory)
    const double *Gp = G;
    lbm3::collide_push_node_streamed<FXY, WRITE>(g, P, F, [Gp](int k) { return Gp[k]; }, PH, fout, gout, o, zs, ix, iy);
}

// End of a one-pass step: the phase field of the four ghost planes (0, 1, nz+2, nz+3), whose g populations the halo
// exchange has just delivered, and -- only when the one-pass kernel reported (ctrl[0] != 0) that a warp gave up
// waiting for the blocks it depends on -- the phase field of the owned planes [zs0, zs1) all over again.
__global__ void __launch_bounds__(kThreads) k_phase_finish(Geom g, Params P, const pop_t *gpop, double *phi, const unsigned *ctrl,
                                                            unsigned *sticky, int zs0, int zs1)
{
    const long long per = (long long)g.nx * g.ny;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int ghost = (int)blockIdx.y < 2 ? (int)blockIdx.y : g.nz + (int)blockIdx.y;
    if (idx < per) lbm3::phase_thread(g, P, gpop, phi, ghost, (int)(idx % g.nx), (int)(idx / g.nx));
    if (*reinterpret_cast<const volatile unsigned *>(ctrl) == 0u) return;
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) atomicAdd(sticky, 1u);
    const long long total = per * (zs1 - zs0), nthreads = (long long)gridDim.x * gridDim.y * blockDim.x;
    for (long long t = ((long long)blockIdx.y * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += nthreads) {
        const long long i = t % per;
        lbm3::phase_thread(g, P, gpop, phi, zs0 + (int)(t / per), (int)(i % g.nx), (int)(i / g.nx));
    }
}

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

// plane z0 + blockIdx.y of the slab <-> plane blockIdx.y of the AoS staging chunk
__global__ void __launch_bounds__(kThreads) k_gather_aos(Geom g, const pop_t *pop, double *aos, int z0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::gather_aos_thread(g, pop, aos - (long long)z0 * g.nx * g.ny * 19, z0 + blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
}

__global__ void __launch_bounds__(kThreads) k_scatter_aos(Geom g, const double *aos, pop_t *pop, int z0)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)g.nx * g.ny) return;
    lbm3::scatter_aos_thread(g, aos - (long long)z0 * g.nx * g.ny * 19, pop, z0 + blockIdx.y, (int)(idx % g.nx), (int)(idx / g.nx));
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
    static std::once_flag once;   // contexts of several host threads may attach at the same time
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) return;
        api.GetUniqueId = (int (*)(void *))dlsym(api.lib, "ncclGetUniqueId");
        api.CommInitRank = (int (*)(void **, int, Id128, int))dlsym(api.lib, "ncclCommInitRank");
        api.CommDestroy = (int (*)(void *))dlsym(api.lib, "ncclCommDestroy");
        api.GroupStart = (int (*)())dlsym(api.lib, "ncclGroupStart");
        api.GroupEnd = (int (*)())dlsym(api.lib, "ncclGroupEnd");
        api.Send = (int (*)(const void *, size_t, int, int, void *, cudaStream_t))dlsym(api.lib, "ncclSend");
        api.Recv = (int (*)(void *, size_t, int, int, void *, cudaStream_t))dlsym(api.lib, "ncclRecv");
        api.GetErrorString = (const char *(*)(int))dlsym(api.lib, "ncclGetErrorString");
    });
    if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd) {
        if (why) *why = "NCCL (libnccl.so.2) could not be loaded";
        return nullptr;
    }
    return &api;
}

class Slab3D : public Model {
public:
    Slab3D(Runtime &rt, const lbm_config &cfg) : rt_(rt), cfg_(cfg) {}

    // kernel_variant: 0 one-pass step (collide + next step's phase field in one launch, d3q19_step.cuh); 5 two-pass
    // (k_phase, then the collide kernels); 1 two-pass with the plain whole-plane collide kernel (the cross-check).
    // The TRT extension exists in the array form of the collision only and always takes the two-pass path.
    // 2 and 6: in-place (AA-pattern) streaming on ONE population array (d3q19_inplace.cuh) -- half the population
    // memory; any collision operator; on several slabs only five directions per z face travel (aa_exchange).  2: one pass (the phase field of
    // the step runs ahead of the collision inside the same launch; BGK -- the extensions take the two-pass form);
    // 6: two passes (phase field, then collide in place).
    bool aa() const { return cfg_.kernel_variant == 2 || cfg_.kernel_variant == 6; }
    bool aa_one() const { return cfg_.kernel_variant == 2 && coll() == 0; }
    int coll() const { return mrt_ ? 2 : (P_.trt_magic > 0.0 ? 1 : 0); }
    bool one_pass() const { return cfg_.kernel_variant == 0 && coll() == 0; }

    lbm_status setup()
    {
        if (cfg_.nx < 4 || cfg_.ny < 4) return rt_.fail(LBM_EINVAL, "D3Q19: nx and ny must be at least 4");
        if (cfg_.slab_count < 1 || cfg_.slab_rank < 0 || cfg_.slab_rank >= cfg_.slab_count)
            return rt_.fail(LBM_EINVAL, "bad slab_rank/slab_count");
        if (cfg_.kernel_variant != 0 && cfg_.kernel_variant != 1 && cfg_.kernel_variant != 5 && !aa())
            return rt_.fail(LBM_EINVAL, "D3Q19: kernel_variant must be 0 (one-pass), 5 (two-pass, interior + rim), 1 (two-pass, plain), 2 (in place on one array, one pass) or 6 (in place, two passes)");
        // solver.hpp:25-30
        const int zstep = cfg_.nz / cfg_.slab_count;
        z_begin_ = cfg_.slab_rank * zstep;
        int z_end = z_begin_ + zstep - 1;
        if (cfg_.slab_rank == cfg_.slab_count - 1) z_end = cfg_.nz - 1;
        g_.nx = cfg_.nx; g_.ny = cfg_.ny; g_.nz = z_end - z_begin_ + 1;
        g_.zwrap = cfg_.slab_count == 1;   // in-place kernels: z periodic by addressing / ghost planes filled by aa_exchange
        if (g_.nz < 2) return rt_.fail(LBM_EINVAL, "D3Q19: every slab needs at least 2 z planes");
        g_.pl = ((long long)g_.nx * g_.ny + 15) / 16 * 16;
        P_.aconst = cfg_.aconst; P_.kconst = cfg_.kconst; P_.gammaconst = cfg_.gammaconst;
        P_.tau_gas = cfg_.tau_gas; P_.dtau = cfg_.tau_liq - cfg_.tau_gas;
        P_.omega_phi = 1.0 / cfg_.tau_phi;
        P_.fx = cfg_.force[0]; P_.fy = cfg_.force[1]; P_.fz = cfg_.force[2];
        P_.phase_gradient = cfg_.wall_gradient; P_.rho_wall = cfg_.rho_wall;
        P_.trt_magic = cfg_.trt_magic;
        if (cfg_.trt_magic < 0.0) return rt_.fail(LBM_EINVAL, "trt_magic must be >= 0 (0 = BGK, the reference's collision)");
        for (int i = 0; i < 5; i++) {
            P_.mrt[i] = cfg_.mrt_rates[i];
            mrt_ = mrt_ || cfg_.mrt_rates[i] != 0.0;
            if (cfg_.mrt_rates[i] >= 2.0) return rt_.fail(LBM_EINVAL, "mrt_rates must be below 2 (<= 0: use 1/tau)");
        }
        if (mrt_ && cfg_.trt_magic > 0.0) return rt_.fail(LBM_EINVAL, "trt_magic and mrt_rates exclude each other");
        if (coll() != 0 && cfg_.kernel_variant == 1)
            return rt_.fail(LBM_EINVAL, "the TRT and MRT extensions exist for kernel_variant 0 and 5 only");
        fxy_ = P_.fx != 0.0 || P_.fy != 0.0;

        const size_t pop_bytes = (size_t)(g_.nz + 4) * 19 * g_.pl * sizeof(pop_t);
        if (aa() && 20 * g_.pl >= (1ll << 31))
            return rt_.fail(LBM_EINVAL, "D3Q19 in-place streaming: a plane must hold fewer than 107 million nodes");
        for (int b = 0; b < (aa() ? 1 : 2); b++) {
            LBM_CUDA(rt_, rt_.dev_alloc(&f_[b], pop_bytes));
            LBM_CUDA(rt_, rt_.dev_alloc(&gq_[b], pop_bytes));
            LBM_CUDA(rt_, cudaMemsetAsync(f_[b], 0, pop_bytes, rt_.stream));
            LBM_CUDA(rt_, cudaMemsetAsync(gq_[b], 0, pop_bytes, rt_.stream));
        }
        if (aa()) { f_[1] = f_[0]; gq_[1] = gq_[0]; }   // one array: "front" and "back" are the same memory
        // two phase-field buffers: the one-pass kernel writes the field of step t+1 while step t reads its own
        const size_t phi_bytes = (size_t)(g_.nz + 4) * g_.pl * sizeof(double);
        for (int b = 0; b < (cfg_.kernel_variant == 0 ? 2 : 1); b++) {
            LBM_CUDA(rt_, rt_.dev_alloc(&phib_[b], phi_bytes));
            LBM_CUDA(rt_, cudaMemsetAsync(phib_[b], 0, phi_bytes, rt_.stream));
        }
        const size_t fld_bytes = (size_t)g_.nz * g_.pl * sizeof(double);
        double **fl[4] = {&rho_, &ux_, &uy_, &uz_};
        for (auto p : fl) {
            LBM_CUDA(rt_, rt_.dev_alloc(p, fld_bytes));
            LBM_CUDA(rt_, cudaMemsetAsync(*p, 0, fld_bytes, rt_.stream));
        }
        LBM_CUDA(rt_, rt_.dev_alloc(&diag_, (size_t)(diagnostics_scratch_blocks() + 1) * 6 * sizeof(double)));
        if (one_pass() || aa_one()) {
            // strip / lag geometry of the one-pass kernels (d3q19_step.cuh, d3q19_inplace.cuh).  A strip is `sb_` blocks
            // of 128 in-plane nodes; one part of a block trails the other by `lag_` plane-steps = kLagWindows times the
            // number of blocks the GPU holds at once (measured at 512^3 with k_step_one: 1.0 -> 16.4, 1.25 -> 16.2,
            // 1.5 -> 15.3, 1.75 -> 15.4, 2 -> 15.5, 3 -> 16.3 ms: shorter and blocks wait for their producers, longer
            // and L2 loses the g rows).
            if (g_.pl * 40 >= (1ll << 31) || g_.nz + 4 + 64 > 65535)
                return rt_.fail(LBM_EINVAL, "D3Q19 one-pass step: a plane must hold fewer than 53 million nodes and a slab fewer than 65 000 planes (use kernel_variant 1)");
            int sms = 148;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, rt_.device);
            sms_ = sms;
            double lag_windows = 1.5;
            if (const char *e = getenv("LBM_TUNE_LAGK")) lag_windows = atof(e);     // measurement knobs, not part of the ABI
            if (const char *e = getenv("LBM_TUNE_POL")) pol_ = atoi(e);
            if (const char *e = getenv("LBM_TUNE_SPIN")) spin_limit_ = (unsigned)atoi(e);   // 0 with LAGK 0: every late block gives up (tests)
            nblk_ = (int)(((long long)g_.nx * g_.ny + 127) / 128);
            dshift_ = 1 + g_.nx / 128;
            sb_ = std::max(64, 2 * dshift_);
            if (nblk_ + dshift_ <= sb_) sb_ = nblk_ + dshift_;   // one strip
            nstrips_ = (nblk_ + dshift_ + sb_ - 1) / sb_;
            lag_ = 2 + (int)(lag_windows * (sms * 4) / sb_);
            // control words: [0] give-up flag, [32 + 32 i] counter of group i, then (in place) one word per segment of
            // an owned plane for the warps that did not collide, [last] number of steps that fell back
            ctrl_words_ = ((size_t)nstrips_ * (g_.nz + 4 + lag_) + 2) * lbm3::LAG_PAD + 64;
            if (aa_one()) ctrl_words_ += (size_t)nblk_ * g_.nz;
            LBM_CUDA(rt_, rt_.dev_alloc(&ctrl_, ctrl_words_ * sizeof(unsigned)));
            LBM_CUDA(rt_, cudaMemsetAsync(ctrl_, 0, ctrl_words_ * sizeof(unsigned), rt_.stream));
        }
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
        for (auto &ge : graph_exec_) if (ge) cudaGraphExecDestroy(ge);
        if (side_) { cudaStreamDestroy(side_); cudaEventDestroy(ev_fork_); cudaEventDestroy(ev_join_); }
        for (int b = 0; b < 2; b++) {
            if (b == 0 || f_[1] != f_[0]) { rt_.dev_free(f_[b]); rt_.dev_free(gq_[b]); }
            rt_.dev_free(phib_[b]);
        }
        rt_.dev_free(rho_); rt_.dev_free(ux_); rt_.dev_free(uy_); rt_.dev_free(uz_); rt_.dev_free(diag_); rt_.dev_free(ctrl_); rt_.dev_free(stage_);
    }

    size_t local_nodes() const override { return (size_t)g_.nx * g_.ny * g_.nz; }
    int q() const override { return 19; }
    int nsets() const override { return 2; }
    lbm_status slab_extent(int *zb, int *nzl) const override { *zb = z_begin_; *nzl = g_.nz; return LBM_OK; }
    int pop_bytes() const override { return (int)sizeof(pop_t); }
    unsigned long long fallbacks() override
    {
        unsigned v = 0;
        if (ctrl_ && cudaMemcpyAsync(&v, ctrl_ + ctrl_words_ - 1, sizeof v, cudaMemcpyDeviceToHost, rt_.stream) == cudaSuccess)
            cudaStreamSynchronize(rt_.stream);
        return v;
    }

    // the public phase field: what the last step read (the reference's output timing)
    double *phi_pub() const { return phib_[po_]; }
    double *field_ptr(int field)
    {
        switch (field) {
        case LBM_RHO: return rho_;
        case LBM_PHASE: return phi_pub() + 2 * g_.pl;
        case LBM_UX: return ux_;
        case LBM_UY: return uy_;
        case LBM_UZ: return uz_;
        }
        return nullptr;
    }
    // queued on the context's stream; a pageable host buffer is staged by the runtime before the call returns, a
    // page-locked one (lbm_host_alloc) must stay untouched until the next call that waits (lbm_sync, lbm_get_field ...)
    lbm_status set_field(int field, const double *host) override
    {
        double *d = field_ptr(field);
        if (!d) return rt_.fail(LBM_EINVAL, "unknown field");
        const size_t row = (size_t)g_.nx * g_.ny * sizeof(double);
        LBM_CUDA(rt_, cudaMemcpy2DAsync(d, g_.pl * sizeof(double), host, row, row, g_.nz, cudaMemcpyHostToDevice, rt_.stream));
        return LBM_OK;
    }
    lbm_status get_field_async(int field, double *host) override
    {
        double *d = field_ptr(field);
        if (!d) return rt_.fail(LBM_EINVAL, "unknown field");
        const size_t row = (size_t)g_.nx * g_.ny * sizeof(double);
        LBM_CUDA(rt_, cudaMemcpy2DAsync(host, row, d, g_.pl * sizeof(double), row, g_.nz, cudaMemcpyDeviceToHost, rt_.stream));
        return LBM_OK;
    }
    lbm_status get_field(int field, double *host) override
    {
        LBM_TRY(get_field_async(field, host));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }

    dim3 plane_grid(int planes) const { return dim3((unsigned)(((long long)g_.nx * g_.ny + kThreads - 1) / kThreads), planes); }
    FieldsOut fields_out() const { return FieldsOut{rho_, ux_, uy_, uz_}; }

    // compile-time flavour of a collide kernel from the run-time switches: fn(FXY, WRITE, COLL) with integral_constant tags
    template <class Fn>
    void with_flavour(int write_fields, Fn &&fn)
    {
        const int c = coll();
        auto l3 = [&](auto fxy, auto wr) {
            if (c == 2) fn(fxy, wr, std::integral_constant<int, 2>{});
            else if (c == 1) fn(fxy, wr, std::integral_constant<int, 1>{});
            else fn(fxy, wr, std::integral_constant<int, 0>{});
        };
        auto l2 = [&](auto fxy) {
            if (write_fields) l3(fxy, std::true_type{});
            else l3(fxy, std::false_type{});
        };
        if (fxy_) l2(std::true_type{});
        else l2(std::false_type{});
    }

    lbm_status phase_planes(const pop_t *gpop, double *phi, int zs0, int count)
    {
        if (count <= 0) return LBM_OK;
        rt_.about_to_launch();
        k_phase<<<plane_grid(count), kThreads, 0, rt_.stream>>>(g_, P_, gpop, phi, zs0);
        return rt_.launched("k_phase");
    }

    // ---- in-place streaming (kernel_variant 2) ----
    dim3 owned_grid(int threads) const { return dim3((unsigned)(((long long)g_.nx * g_.ny + threads - 1) / threads), g_.nz); }
    // phase field out of either half of the cycle -- of all owned planes, or of the first and the last one only --
    // then its two ghost planes (copies inside one slab, the neighbours' planes between slabs)
    lbm_status aa_phase(bool ends_only = false)
    {
        const dim3 grid = ends_only ? dim3(owned_grid(kThreads).x, 2) : owned_grid(kThreads);
        const int zstep = ends_only ? g_.nz - 1 : 1;
        rt_.about_to_launch();
        if (aa_odd_) k_aa_phase<true><<<grid, kThreads, 0, rt_.stream>>>(g_, P_, gq_[0], phib_[0], 2, zstep);
        else k_aa_phase<false><<<grid, kThreads, 0, rt_.stream>>>(g_, P_, gq_[0], phib_[0], 2, zstep);
        LBM_TRY(rt_.launched("k_aa_phase"));
        return exchange(true);
    }
    // Halo exchange of the in-place step between slabs: only the five directions that cross a z face travel, a quarter
    // of what the two-array step sends.  After an even step (the array holds post-collision values in the opposite
    // slots) the slots with c_z = -1 of the top owned plane go to the upper neighbour's ghost plane below its slab and
    // the slots with c_z = +1 of the bottom owned plane to the lower neighbour's ghost plane above: the places its odd
    // step loads from.  After an odd step the same places hold what this slab pushed across the face and travel on to
    // the neighbour's first / last owned plane -- through the unused rows of its ghost planes, from where k_aa_merge
    // moves them into the owned plane except where a slot holds the node's own bounced-back population (a wall next
    // to the z face: that place is written by the node itself, not from across the face).  Slot numbers and in-plane
    // positions are the same on both sides (the folded push never changes c_z).  Directions with c_z = -1 are 6, 13, 14,
    // 17, 18; c_z = +1: 5, 11, 12, 15, 16.
    lbm_status aa_exchange(bool after_even)
    {
        if (cfg_.slab_count == 1) return LBM_OK;
        LBM_TRY(need_comm());
        NcclApi *api = nccl_api(&rt_.err);
        if (!api) return LBM_ENCCL;
        static const int kDown[3][2] = {{6, 1}, {13, 2}, {17, 2}}, kUp[3][2] = {{5, 1}, {11, 2}, {15, 2}};   // first slot, count
        const int up = (cfg_.slab_rank + 1) % cfg_.slab_count, down = (cfg_.slab_rank - 1 + cfg_.slab_count) % cfg_.slab_count;
        const int dtype = sizeof(pop_t) == sizeof(double) ? kNcclFloat64 : kNcclFloat32;
        const size_t pl = (size_t)g_.pl;
        const int nz = g_.nz;
        auto at = [&](pop_t *arr, int zs, int k) { return arr + ((size_t)zs * 19 + k) * pl; };
        // planes: what goes to the upper / lower neighbour and where its counterpart from below / above arrives
        const int s_up = after_even ? nz + 1 : nz + 2, s_down = after_even ? 2 : 1;
        const int r_below = 1, r_above = nz + 2;      // after an odd step: the rows of the ghost planes no place of this slab uses
        // to the upper neighbour: after an even step my c_z = -1 slots (it loads them as its plane below), after an odd
        // step the c_z = +1 slots of my ghost plane above (what I pushed up); the lower neighbour mirrors that
        const int(*to_up)[2] = after_even ? kDown : kUp, (*to_down)[2] = after_even ? kUp : kDown;
        int rc = api->GroupStart();
        pop_t *arrs[2] = {f_[0], gq_[0]};
        for (pop_t *arr : arrs)
            for (int i = 0; i < 3 && rc == 0; i++) rc = api->Send(at(arr, s_up, to_up[i][0]), to_up[i][1] * pl, dtype, up, comm_, rt_.stream);
        for (pop_t *arr : arrs)
            for (int i = 0; i < 3 && rc == 0; i++) rc = api->Send(at(arr, s_down, to_down[i][0]), to_down[i][1] * pl, dtype, down, comm_, rt_.stream);
        // receives in the order the peer sends: from below comes what the lower neighbour sent up, from above what the
        // upper one sent down (with two slabs both are the same peer: NCCL matches a pair's messages in order)
        for (pop_t *arr : arrs)
            for (int i = 0; i < 3 && rc == 0; i++) rc = api->Recv(at(arr, r_below, to_up[i][0]), to_up[i][1] * pl, dtype, down, comm_, rt_.stream);
        for (pop_t *arr : arrs)
            for (int i = 0; i < 3 && rc == 0; i++) rc = api->Recv(at(arr, r_above, to_down[i][0]), to_down[i][1] * pl, dtype, up, comm_, rt_.stream);
        const int rc2 = api->GroupEnd();
        if (rc == 0) rc = rc2;
        if (rc != 0) return rt_.fail(LBM_ENCCL, std::string("NCCL halo exchange (in-place step): ") + (api->GetErrorString ? api->GetErrorString(rc) : "error"));
        if (!after_even) {
            rt_.about_to_launch();
            k_aa_merge<<<owned_grid(kThreads).x, kThreads, 0, rt_.stream>>>(g_, f_[0], gq_[0]);
            LBM_TRY(rt_.launched("k_aa_merge"));
        }
        return LBM_OK;
    }
    // one pass (d3q19_inplace.cuh): phase parts run L plane-steps ahead of the collisions of the same launch
    lbm_status aa_step_one(int write_fields)
    {
        // between slabs the collision of the first and the last plane needs the neighbours' phase field: those two
        // planes are summed ahead of the launch and exchanged (the launch sums them again: same values)
        if (cfg_.slab_count > 1) LBM_TRY(aa_phase(true));
        lbm3::AaPlan A{};
        A.sb = sb_; A.nstrips = nstrips_; A.L = lag_ + 1; A.dshift = dshift_; A.nblk = nblk_;
        A.pz = g_.nz + A.L;
        A.full = (unsigned)sb_;
        A.spin_limit = spin_limit_;
        const size_t counters = (size_t)nstrips_ * g_.nz * lbm3::LAG_PAD, masks = (size_t)nblk_ * g_.nz;
        if (32 + counters + masks + 1 > ctrl_words_) return rt_.fail(LBM_ESTATE, "in-place one-pass kernel: control words too small");
        LBM_CUDA(rt_, cudaMemsetAsync(ctrl_, 0, (32 + counters + masks) * sizeof(unsigned), rt_.stream));
        unsigned *done = ctrl_ + 32, *skipped = ctrl_ + 32 + counters, *sticky = ctrl_ + ctrl_words_ - 1;
        const FieldsOut o = fields_out();
        const dim3 grid(sb_, A.pz, nstrips_);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto) {
            constexpr bool FXY = decltype(fxy)::value, WR = decltype(wr)::value;
            if (aa_odd_) k_aa_one<FXY, WR, true><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o, done, ctrl_, skipped, A);
            else k_aa_one<FXY, WR, false><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o, done, ctrl_, skipped, A);
        });
        LBM_TRY(rt_.launched("k_aa_one"));
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto) {
            constexpr bool FXY = decltype(fxy)::value, WR = decltype(wr)::value;
            if (aa_odd_) k_aa_repair<FXY, WR, true><<<sms_ * 3, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o, ctrl_, skipped, sticky, nblk_);
            else k_aa_repair<FXY, WR, false><<<sms_ * 3, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o, ctrl_, skipped, sticky, nblk_);
        });
        LBM_TRY(rt_.launched("k_aa_repair"));
        LBM_TRY(aa_exchange(!aa_odd_));
        aa_odd_ = !aa_odd_;
        return LBM_OK;
    }
    lbm_status aa_step(int write_fields)
    {
        if (aa_one()) return aa_step_one(write_fields);
        LBM_TRY(aa_phase());
        const FieldsOut o = fields_out();
        const dim3 grid = owned_grid(128);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto c) {
            constexpr bool FXY = decltype(fxy)::value, WR = decltype(wr)::value;
            constexpr int COLL = decltype(c)::value;
            if (COLL == 0) {   // BGK: the streamed form; the TRT and MRT extensions exist in the array form only
                if (aa_odd_) k_aa_core<FXY, WR, true><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o);
                else k_aa_core<FXY, WR, false><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o);
            } else {
                if (aa_odd_) k_aa_collide<FXY, WR, COLL == 0 ? 1 : COLL, true><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o);
                else k_aa_collide<FXY, WR, COLL == 0 ? 1 : COLL, false><<<grid, 128, 0, rt_.stream>>>(g_, P_, f_[0], gq_[0], phib_[0], o);
            }
        });
        LBM_TRY(rt_.launched(coll() == 0 ? "k_aa_core" : "k_aa_collide"));
        LBM_TRY(aa_exchange(!aa_odd_));
        aa_odd_ = !aa_odd_;
        return LBM_OK;
    }
    // everything that reads or writes populations from outside a step wants the streamed populations: after an odd
    // number of steps the pairs of places swap contents once (in place; the next step is then an even one again)
    // Between slabs the swap is local as well (the partner places in the ghost planes hold the neighbours' values since
    // the exchange after the even step), but the slabs have to stay in the same half of the cycle: readers swap, read and
    // swap back (aa_denormalize), and populations can only be loaded in the even half.
    lbm_status aa_swap()
    {
        rt_.about_to_launch();
        k_aa_swap<<<owned_grid(kThreads), kThreads, 0, rt_.stream>>>(g_, f_[0], gq_[0]);
        return rt_.launched("k_aa_swap");
    }
    lbm_status aa_normalize()
    {
        if (!aa_odd_) return LBM_OK;
        LBM_TRY(aa_swap());
        if (cfg_.slab_count == 1) aa_odd_ = false;
        else aa_swapped_for_read_ = true;
        return LBM_OK;
    }
    lbm_status aa_denormalize()
    {
        if (!aa_swapped_for_read_) return LBM_OK;
        aa_swapped_for_read_ = false;
        return aa_swap();
    }

    // ---- slab phases ----
    lbm_status slab_init_fields() override
    {
        k_init_fields<<<(g_.nz + 63) / 64, 64, 0, rt_.stream>>>(g_, P_, phi_pub(), rho_);
        return rt_.launched("k_init_fields");
    }
    lbm_status slab_init_pops() override
    {
        const int b = 1 - cur_;
        k_init_pops<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, P_, phi_pub(), rho_, ux_, uy_, uz_, f_[b], gq_[b]);
        return rt_.launched("k_init_pops");
    }
    lbm_status slab_after_init() override
    {
        if (aa()) return rt_.fail(LBM_EINVAL, "kernel_variant 2 / 6 (in-place streaming) is driven by lbm_init / lbm_step, not by the lbm_slab_* protocol");
        pf_ = po_;
        LBM_TRY(phase_planes(gq_[cur_], phib_[pf_], 0, g_.nz + 4));
        phi_valid_ = true;
        k_after_init<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, P_, f_[cur_], fields_out());
        return rt_.launched("k_after_init");
    }

    // rows iY = 1 and NY-2 (+ wall-node fields on output steps) of planes [zs0, zs0+n0) and [zs1, zs1+n1)
    lbm_status collide_rim(int zs0, int n0, int zs1, int n1, int write_fields)
    {
        if (n0 + n1 <= 0) return LBM_OK;
        const int a = cur_, b = 1 - cur_;
        const FieldsOut o = fields_out();
        const int rim = 2 * (g_.nx - 2) + g_.nx + g_.ny - 1;
        const dim3 gr((rim + 127) / 128, n0 + n1);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto trt) {
            k_collide_rim<decltype(fxy)::value, decltype(wr)::value, decltype(trt)::value>
                <<<gr, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phib_[pf_], f_[b], gq_[b], o, zs0, n0, zs1);
        });
        return rt_.launched("k_collide_rim");
    }
    // every other bulk node of the same planes
    lbm_status collide_interior(int zs0, int n0, int zs1, int n1, int write_fields)
    {
        if (n0 + n1 <= 0) return LBM_OK;
        const int a = cur_, b = 1 - cur_;
        const FieldsOut o = fields_out();
        const dim3 gi((unsigned)(((long long)g_.nx * g_.ny + 127) / 128), n0 + n1);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto trt) {
            k_collide_interior<decltype(fxy)::value, decltype(wr)::value, decltype(trt)::value>
                <<<gi, 128, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phib_[pf_], f_[b], gq_[b], o, zs0, n0, zs1);
        });
        return rt_.launched("k_collide_interior");
    }

    // the same planes through k_step_one<PHI = false>: the streamed collision with g staged through shared memory
    // (128 registers, 16 warps per SM), collide only
    lbm_status collide_core(int zs0, int n0, int zs1, int n1, int write_fields)
    {
        if (n0 + n1 <= 0) return LBM_OK;
        const int a = cur_, b = 1 - cur_;
        lbm3::LagPlan L{};
        const int nb = (int)(((long long)g_.nx * g_.ny + 127) / 128);
        L.sb = nb; L.nstrips = 1; L.lag = 0; L.dshift = 0;
        L.c0 = -(1 << 30); L.c1 = 1 << 30; L.p0 = 0; L.p1 = 0;
        L.z0 = zs0; L.pz = n0 + n1;
        L.zjump_at = n0; L.zjump_by = zs1 - (zs0 + n0);
        const FieldsOut o = fields_out();
        const dim3 grid(L.sb, L.pz, 1);
        const size_t smem = 19 * 128 * sizeof(pop_t);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto) {
            lbm3::k_step_one<decltype(fxy)::value, decltype(wr)::value, 0, false>
                <<<grid, 128, smem, rt_.stream>>>(g_, P_, f_[a], gq_[a], phib_[pf_], f_[b], gq_[b], o, nullptr, nullptr, nullptr, L);
        });
        return rt_.launched("k_collide_core");
    }

    // every bulk node except the rim rows, collide only: the staged kernel, or (the TRT extension, which exists in
    // the array form of the collision only) the register kernel
    lbm_status collide_bulk(int zs0, int n0, int zs1, int n1, int write_fields)
    {
        if (coll() != 0) return collide_interior(zs0, n0, zs1, n1, write_fields);
        return collide_core(zs0, n0, zs1, n1, write_fields);
    }

    // One-pass kernel: collide planes [c0, c1) (all bulk nodes except the rim rows) and the next step's phase field
    // of planes [p0, p1) into the other phase buffer.  Every collide that pushes into planes [p0, p1) must either
    // be part of this launch or have been launched before it.
    lbm_status step_one(int c0, int c1, int p0, int p1, int write_fields)
    {
        const int a = cur_, b = 1 - cur_;
        lbm3::LagPlan L{};
        L.sb = sb_; L.nstrips = nstrips_; L.lag = lag_; L.dshift = dshift_;
        L.c0 = c0; L.c1 = c1; L.p0 = p0; L.p1 = p1;
        L.z0 = std::min(c0, p0 + lag_);
        L.zjump_at = 1 << 30; L.zjump_by = 0;
        L.pz = std::max(c1, p1 + lag_) - L.z0;
        L.full = (unsigned)sb_;
        L.spin_limit = spin_limit_;
        const size_t groups = (size_t)nstrips_ * L.pz;
        if (groups * lbm3::LAG_PAD + 34 > ctrl_words_) return rt_.fail(LBM_ESTATE, "one-pass kernel: group counters too small");
        // [0] give-up flag, [32 + 32 i] counter of group i, [last] number of steps that fell back
        LBM_CUDA(rt_, cudaMemsetAsync(ctrl_, 0, (groups * lbm3::LAG_PAD + 33) * sizeof(unsigned), rt_.stream));
        const FieldsOut o = fields_out();
        const dim3 grid(sb_, L.pz, nstrips_);
        const size_t smem = 2 * 19 * 128 * sizeof(pop_t);
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto) {
            auto launch = [&](auto kern) {
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                kern<<<grid, 128, smem, rt_.stream>>>(g_, P_, f_[a], gq_[a], phib_[pf_], f_[b], gq_[b], o, phib_[1 - pf_], ctrl_ + 32, ctrl_, L);
            };
            constexpr bool FXY = decltype(fxy)::value, WR = decltype(wr)::value;
            if (pol_ == 1) launch(lbm3::k_step_one<FXY, WR, 1, true>);
            else launch(lbm3::k_step_one<FXY, WR, 0, true>);
        });
        LBM_TRY(rt_.launched("k_step_one"));
        pend_p0_ = p0; pend_p1_ = p1;
        lag_pending_ = true;
        return LBM_OK;
    }
    lbm_status ensure_phi()
    {
        if (phi_valid_) return LBM_OK;
        LBM_TRY(phase_planes(gq_[cur_], phib_[pf_], 0, g_.nz + 4));
        phi_valid_ = true;
        return LBM_OK;
    }

    lbm_status slab_collide(int write_fields) override
    {
        if (aa()) return rt_.fail(LBM_EINVAL, "kernel_variant 2 / 6 (in-place streaming) is driven by lbm_step, not by the lbm_slab_* protocol");
        const int nz = g_.nz;
        if (one_pass()) {
            LBM_TRY(ensure_phi());
            LBM_TRY(collide_rim(1, nz + 2, 0, 0, write_fields));
            return step_one(1, nz + 3, 2, nz + 2, write_fields);
        }
        LBM_TRY(phase_planes(gq_[cur_], phib_[pf_], 0, nz + 4));
        po_ = pf_;
        if (cfg_.kernel_variant != 1) {
            LBM_TRY(collide_bulk(1, nz + 2, 0, 0, write_fields));
            return collide_rim(1, nz + 2, 0, 0, write_fields);
        }
        const int a = cur_, b = 1 - cur_;
        const FieldsOut o = fields_out();
        rt_.about_to_launch();
        with_flavour(write_fields, [&](auto fxy, auto wr, auto) {
            k_collide_push<decltype(fxy)::value, decltype(wr)::value>
                <<<plane_grid(nz + 2), kThreads, 0, rt_.stream>>>(g_, P_, f_[a], gq_[a], phib_[pf_], f_[b], gq_[b], o, 1);
        });
        return rt_.launched("k_collide_push");
    }
    // swap (lattice.hpp:842-855 exchangeMatrices); after a one-pass collide also the phase field of the four ghost
    // planes, whose g populations the exchange has just delivered
    lbm_status slab_finish() override
    {
        if (aa()) return rt_.fail(LBM_EINVAL, "kernel_variant 2 / 6 (in-place streaming) is driven by lbm_step, not by the lbm_slab_* protocol");
        cur_ = 1 - cur_;
        if (lag_pending_) {
            lag_pending_ = false;
            rt_.about_to_launch();
            k_phase_finish<<<plane_grid(4), kThreads, 0, rt_.stream>>>(g_, P_, gq_[cur_], phib_[1 - pf_], ctrl_, ctrl_ + ctrl_words_ - 1, pend_p0_, pend_p1_);
            LBM_TRY(rt_.launched("k_phase_finish"));
            po_ = pf_;
            pf_ = 1 - pf_;
            phi_valid_ = true;
        } else {
            phi_valid_ = false;
        }
        return LBM_OK;
    }
    // halo messages always refer to the back buffers (what collide / init_pops just wrote)
    lbm_status slab_halo(int id, void **ptr, size_t *count) override
    {
        const int b = 1 - cur_;
        const size_t pp = 19 * (size_t)g_.pl;
        auto at = [&](pop_t *arr, int z) { return arr + (size_t)(z + 2) * pp; };
        const int nz = g_.nz;
        double *phi = phi_pub();
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
        case LBM_HALO_S_PHI_UP: p = phi + (size_t)(nz - 1 + 2) * g_.pl; n = g_.pl; break;
        case LBM_HALO_S_PHI_DOWN: p = phi + (size_t)2 * g_.pl; n = g_.pl; break;
        case LBM_HALO_R_PHI_FROM_BELOW: p = phi + (size_t)1 * g_.pl; n = g_.pl; break;
        case LBM_HALO_R_PHI_FROM_ABOVE: p = phi + (size_t)(nz + 2) * g_.pl; n = g_.pl; break;
        default: return rt_.fail(LBM_EINVAL, "unknown halo id");
        }
        *ptr = p;
        *count = n;
        return LBM_OK;
    }

    // ---- exchange: with itself (one slab, z periodic) or with the ring neighbours over NCCL ----
    lbm_status need_comm()
    {
        if (cfg_.slab_count > 1 && !comm_)
            return rt_.fail(LBM_ESTATE, "slab_count > 1: call lbm_comm_attach first, or drive lbm_slab_* yourself");
        return LBM_OK;
    }
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
        LBM_TRY(need_comm());
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
        LBM_TRY(need_comm());
        if (aa()) {   // one array, no ghost planes of populations
            LBM_TRY(slab_init_fields());
            LBM_TRY(exchange(true));
            LBM_TRY(slab_init_pops());
            aa_odd_ = false;
            LBM_TRY(aa_phase());
            k_after_init<<<plane_grid(g_.nz), kThreads, 0, rt_.stream>>>(g_, P_, f_[0], fields_out());
            return rt_.launched("k_after_init");
        }
        LBM_TRY(slab_init_fields());
        LBM_TRY(exchange(true));
        LBM_TRY(slab_init_pops());
        LBM_TRY(exchange(false));
        LBM_TRY(slab_finish());
        return slab_after_init();
    }
    // One step with the halo exchange hidden behind the bulk of the collision (replaces the blocking
    // MPI_Sendrecv chain of solver.hpp:893-928): the four planes at either end of the slab -- everything
    // the outgoing messages depend on -- first, then the exchange on a side stream while the main stream
    // collides the planes in between.
    lbm_status step_overlapped(int write_fields)
    {
        LBM_TRY(ensure_side_stream());
        const int nz = g_.nz;
        if (one_pass()) {
            LBM_TRY(ensure_phi());
            LBM_TRY(collide_rim(1, nz + 2, 0, 0, write_fields));
        } else {
            LBM_TRY(phase_planes(gq_[cur_], phib_[pf_], 0, nz + 4));
            po_ = pf_;
            LBM_TRY(collide_rim(1, 4, nz - 1, 4, write_fields));
        }
        LBM_TRY(collide_bulk(1, 4, nz - 1, 4, write_fields));   // z = -1 .. 2 and nz-3 .. nz in one launch
        LBM_CUDA(rt_, cudaEventRecord(ev_fork_, rt_.stream));
        LBM_CUDA(rt_, cudaStreamWaitEvent(side_, ev_fork_, 0));
        LBM_TRY(exchange(false, side_));
        LBM_CUDA(rt_, cudaEventRecord(ev_join_, side_));
        if (one_pass()) {
            LBM_TRY(step_one(5, nz - 1, 2, nz + 2, write_fields));   // z = 3 .. nz-4, phase field of all owned planes
        } else {
            LBM_TRY(collide_bulk(5, nz - 6, 0, 0, write_fields));
            LBM_TRY(collide_rim(5, nz - 6, 0, 0, write_fields));
        }
        LBM_CUDA(rt_, cudaStreamWaitEvent(rt_.stream, ev_join_, 0));
        return slab_finish();
    }

    lbm_status one_step(bool overlap, int write_fields)
    {
        if (overlap) return step_overlapped(write_fields);
        LBM_TRY(slab_collide(write_fields));
        LBM_TRY(exchange(false));
        return slab_finish();
    }
    // A pair of steps (no field output) as a CUDA graph: after two steps the buffer indices are back where they were,
    // so one captured pair can be replayed.  Worth it on small slabs, where the five launches, the memset and the
    // NCCL group of a step take as long to issue as to run (FullChannel on 8 GPUs: 0.15 ms per step); the side stream
    // of the overlapped exchange is part of the capture.  Not used while the profiler is on (its event records would
    // be captured) or on a caller's stream (which the caller may be capturing itself).
    lbm_status step_pairs_graph(bool overlap, int pairs)
    {
        const int key = cur_ * 4 + pf_ * 2 + po_;
        if (!graph_exec_[key]) {
            if (overlap) LBM_TRY(ensure_side_stream());
            const unsigned long long launches_before = rt_.launches;
            // the capture records two steps without running them: the host-side indices they advance are put back
            const int s_cur = cur_, s_pf = pf_, s_po = po_, s_p0 = pend_p0_, s_p1 = pend_p1_;
            const bool s_valid = phi_valid_, s_pending = lag_pending_;
            cudaGraph_t graph = nullptr;
            LBM_CUDA(rt_, cudaStreamBeginCapture(rt_.stream, cudaStreamCaptureModeThreadLocal));
            lbm_status st = one_step(overlap, 0);
            if (st == LBM_OK) st = one_step(overlap, 0);
            const cudaError_t e = cudaStreamEndCapture(rt_.stream, &graph);
            cur_ = s_cur; pf_ = s_pf; po_ = s_po; pend_p0_ = s_p0; pend_p1_ = s_p1;
            phi_valid_ = s_valid; lag_pending_ = s_pending;
            if (st != LBM_OK || e != cudaSuccess || !graph) {
                rt_.launches = launches_before;
                if (graph) cudaGraphDestroy(graph);
                graphs_off_ = true;               // e.g. a transport that cannot be captured: plain launches from now on
                (void)cudaGetLastError();
                return st != LBM_OK ? st : rt_.fail(LBM_ECUDA, std::string("graph capture of a step pair: ") + cudaGetErrorString(e));
            }
            const cudaError_t ei = cudaGraphInstantiate(&graph_exec_[key], graph, 0);
            cudaGraphDestroy(graph);
            if (ei != cudaSuccess) {
                graph_exec_[key] = nullptr;
                graphs_off_ = true;
                return rt_.fail(LBM_ECUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei));
            }
            graph_launches_[key] = rt_.launches - launches_before;
            rt_.launches = launches_before;      // nothing has run yet: the capture only recorded the pair
        }
        for (int i = 0; i < pairs; i++) {
            LBM_CUDA(rt_, cudaGraphLaunch(graph_exec_[key], rt_.stream));
            rt_.launches += graph_launches_[key];
        }
        return LBM_OK;
    }

    lbm_status step(int nsteps) override
    {
        LBM_TRY(need_comm());
        if (aa()) {
            for (int t = 0; t < nsteps; t++) LBM_TRY(aa_step(t == nsteps - 1));
            return LBM_OK;
        }
        const bool overlap = comm_ && cfg_.slab_count > 1 && cfg_.kernel_variant != 1 && g_.nz >= 8;
        int t = 0;
        // small slabs: everything but the first and the last steps in replayed pairs
        const bool small = (long long)g_.nx * g_.ny * g_.nz <= kGraphMaxNodes;
        if (small && !graphs_off_ && rt_.own_stream && !rt_.profile && nsteps >= 8 && cfg_.slab_count == 1) {
            LBM_TRY(one_step(overlap, 0));        // brings the phase field of the one-pass path into its steady state
            t = 1;
            const int pairs = (nsteps - 1 - t) / 2;
            const lbm_status st = step_pairs_graph(overlap, pairs);
            if (st == LBM_OK) t += 2 * pairs;
            else if (!graphs_off_) return st;     // a real error; otherwise fall through to plain launches
        }
        for (; t < nsteps; t++) LBM_TRY(one_step(overlap, t == nsteps - 1));
        return LBM_OK;
    }

    // AoS <-> SoA through a bounded staging buffer (allocated on first use, kept): a chunk of planes at a time
    lbm_status ensure_stage()
    {
        if (stage_) return LBM_OK;
        const size_t per_plane = (size_t)g_.nx * g_.ny * 19 * sizeof(double);
        stage_planes_ = (int)std::max<size_t>(1, std::min<size_t>((size_t)g_.nz, ((size_t)256 << 20) / per_plane));
        LBM_CUDA(rt_, rt_.dev_alloc(&stage_, per_plane * stage_planes_));
        return LBM_OK;
    }
    lbm_status set_populations(int set, const double *host) override
    {
        LBM_TRY(need_comm());
        if (aa() && aa_odd_ && cfg_.slab_count > 1)
            return rt_.fail(LBM_ESTATE, "in-place streaming on several slabs: populations can be loaded after lbm_init or an even number of steps only");
        LBM_TRY(ensure_stage());
        LBM_TRY(aa_normalize());
        const size_t per_plane = (size_t)g_.nx * g_.ny * 19;
        pop_t *dst = set == 0 ? f_[cur_] : gq_[cur_];
        for (int z0 = 0; z0 < g_.nz; z0 += stage_planes_) {
            const int n = std::min(stage_planes_, g_.nz - z0);
            LBM_CUDA(rt_, cudaMemcpyAsync(stage_, host + per_plane * z0, per_plane * n * sizeof(double), cudaMemcpyHostToDevice, rt_.stream));
            k_scatter_aos<<<plane_grid(n), kThreads, 0, rt_.stream>>>(g_, stage_, dst, z0);
            LBM_TRY(rt_.launched("k_scatter_aos"));
        }
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        if (aa()) return LBM_OK;   // no ghost planes of populations
        // refresh the ghost planes of the front buffers: slab_halo() addresses the back buffers, so
        // flip, exchange, flip back
        cur_ = 1 - cur_;
        const lbm_status st = exchange(false);
        cur_ = 1 - cur_;
        phi_valid_ = false;   // the front g populations changed
        return st;
    }
    lbm_status get_populations(int set, double *host) override
    {
        LBM_TRY(ensure_stage());
        LBM_TRY(aa_normalize());
        const size_t per_plane = (size_t)g_.nx * g_.ny * 19;
        for (int z0 = 0; z0 < g_.nz; z0 += stage_planes_) {
            const int n = std::min(stage_planes_, g_.nz - z0);
            k_gather_aos<<<plane_grid(n), kThreads, 0, rt_.stream>>>(g_, set == 0 ? f_[cur_] : gq_[cur_], stage_, z0);
            LBM_TRY(rt_.launched("k_gather_aos"));
            LBM_CUDA(rt_, cudaMemcpyAsync(host + per_plane * z0, stage_, per_plane * n * sizeof(double), cudaMemcpyDeviceToHost, rt_.stream));
        }
        LBM_TRY(aa_denormalize());
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }

    lbm_status diagnostics(double *mass, double *mom, double *phase_mass, int *has_nan) override
    {
        LBM_TRY(aa_normalize());
        DiagLayout L;
        const size_t pp = 19 * (size_t)g_.pl;
        L.f = f_[cur_] + 2 * pp;
        L.g = gq_[cur_] + 2 * pp;
        L.stride_k = g_.pl;
        L.plane_stride = (long long)pp;
        L.planes = g_.nz; L.nx = g_.nx; L.ny = g_.ny; L.q = 19; L.skip_border = 1; L.node_major_x = 1;
        double out[6];
        LBM_TRY(run_diagnostics(rt_, L, diag_, diagnostics_scratch_blocks(), out));
        LBM_TRY(aa_denormalize());
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
    bool fxy_ = false, mrt_ = false;
    int z_begin_ = 0, cur_ = 0;
    pop_t *f_[2] = {nullptr, nullptr}, *gq_[2] = {nullptr, nullptr};
    // phase field: phib_[pf_] belongs to the front populations (valid when phi_valid_), phib_[po_] is the public one
    double *phib_[2] = {nullptr, nullptr};
    int pf_ = 0, po_ = 0;
    bool phi_valid_ = false, lag_pending_ = false;
    bool aa_odd_ = false;   // kernel_variant 2 / 6: the array holds post-collision values in the opposite slots (odd half of the cycle)
    bool aa_swapped_for_read_ = false;   // ... and is swapped into streamed order for the duration of a read (several slabs)
    cudaGraphExec_t graph_exec_[8] = {};   // one replayable pair of steps per (cur_, pf_, po_)
    unsigned long long graph_launches_[8] = {};
    bool graphs_off_ = false;
    static constexpr long long kGraphMaxNodes = 16ll << 20;   // slabs up to 16 M nodes (~2 ms per step) use the graph
    int pend_p0_ = 0, pend_p1_ = 0;      // owned planes whose phase field the pending one-pass step has written
    double *rho_ = nullptr, *ux_ = nullptr, *uy_ = nullptr, *uz_ = nullptr, *diag_ = nullptr;
    // one-pass kernel: strip geometry and its control words ([0] give-up flag, [1..] group counters, last: fallback count)
    int nblk_ = 0, dshift_ = 0, sb_ = 0, nstrips_ = 0, lag_ = 0, sms_ = 148;
    int pol_ = 1;     // L2 residency hints on (15.28 vs 15.63 ms at 512^3)
    unsigned spin_limit_ = 4000;   // polls of 200 ns before a warp of the one-pass kernel gives up (~1 ms)
    unsigned *ctrl_ = nullptr;
    size_t ctrl_words_ = 0;
    double *stage_ = nullptr;
    int stage_planes_ = 0;
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
