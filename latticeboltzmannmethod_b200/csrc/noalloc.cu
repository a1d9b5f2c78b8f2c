// noalloc.cu -- lbm_population_sums_noalloc(): the mass / momentum / phase-mass / NaN sums of
// diagnostics.cu for callers that must not allocate.
//
// Who that is.  A solver that has handed every byte of device memory to its own pool, a coupling layer that
// runs inside somebody else's time step, a monitoring hook called between two kernels of a tight loop: for
// all of them cudaMalloc/cudaFree are off limits, either because nothing is left to allocate from or
// because the driver's allocator waits for the whole device and would stall work they do not own.  The
// ordinary lbm_diagnostics() keeps its block partials in a buffer allocated with the context; this entry
// point needs no context and keeps them in an array that is part of the library image.
//
// The price.  The array is file-scope device memory: every GPU that loads the library carries it
// (LBM_NOALLOC_MAX_BLOCKS * 6 doubles = 768 KiB) until the process exits, all calls share it, and it keeps
// whatever the previous call left there.  Every call overwrites exactly the entries it later reads, so
// nothing has to be cleared between calls -- but two calls whose kernels overlap on one GPU (different
// streams, different threads) write the same entries and race.  Serialise them.
//
// Host code cannot use the array's name as a pointer: on the host that name is only a handle.  The device
// address is looked up with cudaGetSymbolAddress, once per device, with that device current, on the first
// call made on that device.  That first call therefore does a little set-up work; make it before any
// stream capture begins.
#include <mutex>

#include "lbm_internal.h"

namespace {

constexpr int kNodesPerBlock = 8192;
constexpr int kThreads = 256;
constexpr int kMaxBlocks = LBM_NOALLOC_MAX_BLOCKS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_noalloc_partials[(size_t)kMaxBlocks * 6];

double *g_partials_addr[kMaxDevices];   // device address of g_noalloc_partials, per device ordinal
bool g_partials_ready[kMaxDevices];
std::mutex g_setup_mutex;
thread_local std::string g_noalloc_error;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void block_sum6(double (&v)[6], double (*sh)[6])
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 6; c++) v[c] = warp_sum(v[c]);
    if (lane == 0)
        for (int c = 0; c < 6; c++) sh[warp][c] = v[c];
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
#pragma unroll
        for (int c = 0; c < 6; c++) {
            double x = lane < nw ? sh[lane][c] : 0.0;
            v[c] = warp_sum(x);
        }
    }
}

// block b sums nodes [b*kNodesPerBlock, (b+1)*kNodesPerBlock) and writes its six partials to partials[b]
template <int Q>
__global__ void __launch_bounds__(kThreads) k_noalloc_partial(const double *__restrict__ f, const double *__restrict__ g,
                                                              long long n_nodes, long long stride_k, double *partials)
{
    constexpr int CX19[19] = {0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1};
    constexpr int CY19[19] = {0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0};
    constexpr int CZ19[19] = {0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1};
    constexpr int CX9[9] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
    constexpr int CY9[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1};
    __shared__ double sh[kThreads / 32][6];
    double acc[6] = {0, 0, 0, 0, 0, 0};
    const long long first = (long long)blockIdx.x * kNodesPerBlock;
    const long long last = min(first + kNodesPerBlock, n_nodes);
    for (long long i = first + threadIdx.x; i < last; i += kThreads) {
#pragma unroll
        for (int k = 0; k < Q; k++) {
            const double fv = f[i + k * stride_k];
            const int cx = Q == 19 ? CX19[k] : CX9[k], cy = Q == 19 ? CY19[k] : CY9[k], cz = Q == 19 ? CZ19[k] : 0;
            acc[0] += fv;
            if (cx) acc[1] += cx > 0 ? fv : -fv;
            if (cy) acc[2] += cy > 0 ? fv : -fv;
            if (cz) acc[3] += cz > 0 ? fv : -fv;
            if (fv != fv) acc[5] += 1.0;
            if (g) {
                const double gv = g[i + k * stride_k];
                acc[4] += gv;
                if (gv != gv) acc[5] += 1.0;
            }
        }
    }
    block_sum6(acc, sh);
    if (threadIdx.x == 0)
        for (int c = 0; c < 6; c++) partials[(size_t)blockIdx.x * 6 + c] = acc[c];
}

__global__ void __launch_bounds__(kThreads) k_noalloc_final(const double *partials, int nblocks, double *out6)
{
    __shared__ double sh[kThreads / 32][6];
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x)
        for (int c = 0; c < 6; c++) acc[c] += partials[(size_t)b * 6 + c];
    block_sum6(acc, sh);
    if (threadIdx.x == 0)
        for (int c = 0; c < 6; c++) out6[c] = acc[c];
}

lbm_status fail(lbm_status st, const std::string &what)
{
    g_noalloc_error = what;
    return st;
}

}  // namespace

extern "C" {

const char *lbm_noalloc_last_error(void) { return g_noalloc_error.c_str(); }

lbm_status lbm_population_sums_noalloc(const double *f_dev, const double *g_dev, int q, size_t n_nodes, size_t stride_k,
                                       void *cuda_stream, double *sums6_dev)
{
    if (!f_dev || !sums6_dev) return fail(LBM_EINVAL, "lbm_population_sums_noalloc: f_dev and sums6_dev must not be null");
    if (q != 9 && q != 19) return fail(LBM_EINVAL, "lbm_population_sums_noalloc: q must be 9 or 19");
    if (n_nodes == 0 || stride_k < n_nodes) return fail(LBM_EINVAL, "lbm_population_sums_noalloc: need 0 < n_nodes <= stride_k");
    const size_t max_nodes = (size_t)kMaxBlocks * kNodesPerBlock;
    if (n_nodes > max_nodes) {
        char msg[200];
        snprintf(msg, sizeof msg, "lbm_population_sums_noalloc: %zu nodes exceed the built-in scratch limit of %zu nodes "
                                  "(LBM_NOALLOC_MAX_BLOCKS x %d); split the range or use lbm_diagnostics",
                 n_nodes, max_nodes, kNodesPerBlock);
        return fail(LBM_EINVAL, msg);
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(LBM_ECUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail(LBM_EINVAL, "lbm_population_sums_noalloc: device ordinal out of range");
    {
        // set-up, once per device, with that device current: turn the symbol into a device address
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_partials_ready[dev]) {
            double *ptr = nullptr;
            e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_partials);
            if (e != cudaSuccess || !ptr)
                return fail(LBM_ECUDA, std::string("cudaGetSymbolAddress(g_noalloc_partials): ") + cudaGetErrorString(e));
            g_partials_addr[dev] = ptr;
            g_partials_ready[dev] = true;
        }
    }
    double *partials = g_partials_addr[dev];
    const int nblocks = (int)((n_nodes + kNodesPerBlock - 1) / kNodesPerBlock);
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    if (q == 19)
        k_noalloc_partial<19><<<nblocks, kThreads, 0, stream>>>(f_dev, g_dev, (long long)n_nodes, (long long)stride_k, partials);
    else
        k_noalloc_partial<9><<<nblocks, kThreads, 0, stream>>>(f_dev, g_dev, (long long)n_nodes, (long long)stride_k, partials);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(LBM_ECUDA, std::string("k_noalloc_partial: ") + cudaGetErrorString(e));
    k_noalloc_final<<<1, kThreads, 0, stream>>>(partials, nblocks, sums6_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(LBM_ECUDA, std::string("k_noalloc_final: ") + cudaGetErrorString(e));
    return LBM_OK;
}

}  // extern "C"
