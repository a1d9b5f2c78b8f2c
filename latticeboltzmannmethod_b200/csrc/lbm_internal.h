// lbm_internal.h -- shared between the translation units of liblbm_b200.so.
#ifndef LBM_INTERNAL_H
#define LBM_INTERNAL_H

#include <cuda_runtime.h>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "lbm_b200.h"
#include "lbm_pop.h"

namespace lbm {

// What lbm_ctx owns besides the model: device, stream, error text, launch counter.
struct Runtime {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    unsigned long long launches = 0;
    // optional per-kernel timing: one event pair around every launch made while `profile` is on
    bool profile = false;
    struct Timed { const char *name; cudaEvent_t t0, t1; };
    std::vector<Timed> timed;
    cudaEvent_t pending_start = nullptr;

    // Device allocations of the models go through here.  With LBM_GUARD=1 in the environment when the context is
    // created every buffer sits between two 1 MiB guard bands filled with a byte pattern; guard_check() counts the
    // guard bytes that no longer hold it -- the out-of-bounds-write check that can run where compute-sanitizer cannot.
    bool guard = false;
    static constexpr size_t kGuardBytes = 1u << 20;
    struct Alloc { char *base; size_t bytes; };
    std::map<void *, Alloc> allocs;
    cudaError_t dev_alloc(void **p, size_t bytes)
    {
        const size_t g = guard ? kGuardBytes : 0;
        char *base = nullptr;
        cudaError_t e = cudaMalloc((void **)&base, bytes + 2 * g);
        if (e != cudaSuccess) return e;
        if (g) {
            cudaMemset(base, 0xA5, g);
            cudaMemset(base + g + bytes, 0xA5, g);
        }
        *p = base + g;
        allocs[*p] = Alloc{base, bytes};
        return cudaSuccess;
    }
    template <class T>
    cudaError_t dev_alloc(T **p, size_t bytes) { return dev_alloc((void **)p, bytes); }
    void dev_free(void *p)
    {
        if (!p) return;
        auto it = allocs.find(p);
        if (it == allocs.end()) { cudaFree(p); return; }
        cudaFree(it->second.base);
        allocs.erase(it);
    }
    // number of guard bytes overwritten so far (0 without LBM_GUARD); waits for the stream
    unsigned long long guard_check()
    {
        if (!guard) return 0;
        cudaStreamSynchronize(stream);
        std::vector<unsigned char> h(kGuardBytes);
        unsigned long long bad = 0;
        for (auto &kv : allocs)
            for (int side = 0; side < 2; side++) {
                const char *src = side == 0 ? kv.second.base : kv.second.base + kGuardBytes + kv.second.bytes;
                if (cudaMemcpy(h.data(), src, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) return ~0ull;
                for (unsigned char c : h) bad += c != 0xA5;
            }
        return bad;
    }

    lbm_status fail(lbm_status st, const std::string &what)
    {
        err = what;
        return st;
    }
    lbm_status cuda(cudaError_t e, const char *what)
    {
        if (e == cudaSuccess) return LBM_OK;
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return e == cudaErrorMemoryAllocation ? LBM_ENOMEM : LBM_ECUDA;
    }
    // call right before / right after a kernel launch
    std::vector<cudaEvent_t> spare;   // events of collected records, reused: no cudaEventCreate in steady state
    cudaEvent_t take_event()
    {
        cudaEvent_t e = nullptr;
        if (!spare.empty()) { e = spare.back(); spare.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    void about_to_launch()
    {
        if (!profile) return;
        if (!pending_start) pending_start = take_event();
        cudaEventRecord(pending_start, stream);
    }
    lbm_status launched(const char *kernel)
    {
        ++launches;
        if (profile && pending_start) {
            cudaEvent_t t1 = take_event();
            cudaEventRecord(t1, stream);
            timed.push_back(Timed{kernel, pending_start, t1});
            pending_start = nullptr;
        }
        return cuda(cudaGetLastError(), kernel);
    }
    // drops every recorded and spare event (profiling switched off, context destroyed)
    void profile_release()
    {
        for (auto &t : timed) { spare.push_back(t.t0); spare.push_back(t.t1); }
        timed.clear();
        if (pending_start) { spare.push_back(pending_start); pending_start = nullptr; }
        for (auto e : spare) cudaEventDestroy(e);
        spare.clear();
    }
    // synchronises, sums the recorded durations per kernel name, releases the events
    std::map<std::string, std::pair<int, double>> profile_collect()
    {
        std::map<std::string, std::pair<int, double>> out;
        cudaStreamSynchronize(stream);
        for (auto &t : timed) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, t.t0, t.t1) == cudaSuccess) {
                auto &e = out[t.name];
                e.first += 1;
                e.second += ms;
            }
            spare.push_back(t.t0);
            spare.push_back(t.t1);
        }
        timed.clear();
        return out;
    }
};

#define LBM_TRY(expr)                          \
    do {                                       \
        lbm_status _st = (expr);               \
        if (_st != LBM_OK) return _st;         \
    } while (0)
#define LBM_CUDA(rt, call) LBM_TRY((rt).cuda((call), #call))

// One lattice model on one GPU.
class Model {
public:
    virtual ~Model() {}
    virtual size_t local_nodes() const = 0;
    virtual int q() const = 0;
    virtual int nsets() const = 0;  // 1 = f only, 2 = f and g
    virtual lbm_status set_field(int field, const double *host) = 0;
    virtual lbm_status set_geometry(const int *) { return LBM_EINVAL; }  // masked models only
    virtual lbm_status get_field(int field, double *host) = 0;
    // queues the copy only (valid after the next lbm_sync); models without an asynchronous form copy and wait
    virtual lbm_status get_field_async(int field, double *host) { return get_field(field, host); }
    virtual int pop_bytes() const { return (int)sizeof(double); }      // storage size of one population value
    virtual unsigned long long fallbacks() { return 0; }               // steps whose phase field had to be recomputed
    virtual lbm_status init() = 0;
    virtual lbm_status step(int nsteps) = 0;
    virtual lbm_status set_populations(int set, const double *host_aos) = 0;
    virtual lbm_status get_populations(int set, double *host_aos) = 0;
    virtual lbm_status diagnostics(double *mass, double *mom, double *phase_mass, int *has_nan) = 0;
    // z-slab interface; only the 3-D model implements it
    virtual lbm_status slab_extent(int *z_begin, int *nz_local) const { *z_begin = 0; *nz_local = 1; return LBM_OK; }
    virtual lbm_status comm_attach(const void *) { return LBM_EINVAL; }
    virtual lbm_status slab_halo(int, void **, size_t *) { return LBM_EINVAL; }
    virtual lbm_status slab_init_fields() { return LBM_EINVAL; }
    virtual lbm_status slab_init_pops() { return LBM_EINVAL; }
    virtual lbm_status slab_after_init() { return LBM_EINVAL; }
    virtual lbm_status slab_collide(int) { return LBM_EINVAL; }
    virtual lbm_status slab_finish() { return LBM_EINVAL; }
};

Model *make_d3q19_binary(Runtime &rt, const lbm_config &cfg, lbm_status *st);
Model *make_d2q9_shanchen(Runtime &rt, const lbm_config &cfg, lbm_status *st);
Model *make_d2q9_binary(Runtime &rt, const lbm_config &cfg, lbm_status *st);
Model *make_d2q9_mrt(Runtime &rt, const lbm_config &cfg, lbm_status *st);
Model *make_d2q9_trt(Runtime &rt, const lbm_config &cfg, lbm_status *st);
Model *make_d2q9_binary_shape(Runtime &rt, const lbm_config &cfg, lbm_status *st);

// diagnostics: block-level partial sums into scratch, then one block folds them.
// sums[0..4] = sum f, sum f*cx, sum f*cy, sum f*cz, sum g ; sums[5] = NaN count
// Populations are SoA with `stride_k` doubles between directions; the node list is
// `count` consecutive in-plane indices starting at `first`, repeated for `planes`
// planes `plane_stride` apart (2-D models: planes = 1); `nx,ny` + `skip_border`
// exclude mirror/wall nodes of the 3-D model.
struct DiagLayout {
    const pop_t *f, *g;       // g may be null (pop_t: double; float only for D3Q19 in the fp32-storage build)
    long long stride_k;       // doubles between consecutive directions
    long long plane_stride;   // doubles between consecutive planes (of direction 0)
    int planes, nx, ny, q;
    int skip_border;          // 1: only 1<=ix<=nx-2, 1<=iy<=ny-2 count (3-D); 2: only 1<=fast<=ny-2 (binary 2-D)
    int node_major_x;         // in-plane index = iy*nx+ix (1) or ix*ny+iy (0)
    const unsigned short *fluid = nullptr;  // optional per-node link word, bit 0 = counts (masked 2-D model)
};
lbm_status run_diagnostics(Runtime &rt, const DiagLayout &L, double *scratch, int scratch_blocks, double *host_out6);
int diagnostics_scratch_blocks();

}  // namespace lbm

struct lbm_ctx {
    lbm_config cfg;
    lbm::Runtime rt;
    lbm::Model *model = nullptr;
    bool initialised = false;
    bool bound = false;   // cudaSetDevice(rt.device) has succeeded at least once
};

#endif
