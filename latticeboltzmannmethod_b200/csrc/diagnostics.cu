// diagnostics.cu -- mass, momentum, phase mass and NaN count of the current
// populations (reference: the sums of shanchen.cpp:49-70, the mass print-out of
// CurvedSolid/binary_shape.cpp:654-661 and the NaN probes binary.cpp:359,
// solver.hpp:49-55; the reference computes them on the host, serially).
//
// Two launches: every block reduces its grid-stride share with warp shuffles and
// writes six partial sums to a scratch array in global memory; one block then
// folds the partials.  Block count and visiting order are fixed, so the result is
// reproducible run to run.  HBM-bound: reads every population once.
#include "lbm_internal.h"

namespace lbm {

namespace {
constexpr int kDiagBlocks = 148 * 4;
constexpr int kDiagThreads = 256;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// folds six per-thread values over the block; valid in thread 0
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

template <int Q>
__global__ void __launch_bounds__(kDiagThreads) k_diag_partial(DiagLayout L, double *partials)
{
    // descriptor.h:43-45 / shanchen.cpp:18-19
    constexpr int CX19[19] = {0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1};
    constexpr int CY19[19] = {0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0};
    constexpr int CZ19[19] = {0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1};
    constexpr int CX9[9] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
    constexpr int CY9[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1};
    __shared__ double sh[kDiagThreads / 32][6];
    double acc[6] = {0, 0, 0, 0, 0, 0};
    const long long per_plane = (long long)L.nx * L.ny;
    const long long total = per_plane * L.planes;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int plane = (int)(t / per_plane);
        const long long idx = t - (long long)plane * per_plane;
        if (L.skip_border == 1) {
            const int ix = (int)(idx % L.nx), iy = (int)(idx / L.nx);
            if (ix < 1 || ix > L.nx - 2 || iy < 1 || iy > L.ny - 2) continue;
        } else if (L.skip_border == 2) {
            const int iy = (int)(idx % L.ny);
            if (iy < 1 || iy > L.ny - 2) continue;
        }
        if (L.fluid && !(L.fluid[idx] & 1)) continue;
        const long long base = (long long)plane * L.plane_stride + idx;
#pragma unroll
        for (int k = 0; k < Q; k++) {
            const double fv = L.f[base + k * L.stride_k];
            const int cx = Q == 19 ? CX19[k] : CX9[k], cy = Q == 19 ? CY19[k] : CY9[k], cz = Q == 19 ? CZ19[k] : 0;
            acc[0] += fv;
            if (cx) acc[1] += cx > 0 ? fv : -fv;
            if (cy) acc[2] += cy > 0 ? fv : -fv;
            if (cz) acc[3] += cz > 0 ? fv : -fv;
            if (fv != fv) acc[5] += 1.0;
            if (L.g) {
                const double gv = L.g[base + k * L.stride_k];
                acc[4] += gv;
                if (gv != gv) acc[5] += 1.0;
            }
        }
    }
    block_sum6(acc, sh);
    if (threadIdx.x == 0)
        for (int c = 0; c < 6; c++) partials[blockIdx.x * 6 + c] = acc[c];
}

__global__ void __launch_bounds__(kDiagThreads) k_diag_final(const double *partials, int nblocks, double *out6)
{
    __shared__ double sh[kDiagThreads / 32][6];
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x)
        for (int c = 0; c < 6; c++) acc[c] += partials[b * 6 + c];
    block_sum6(acc, sh);
    if (threadIdx.x == 0)
        for (int c = 0; c < 6; c++) out6[c] = acc[c];
}
}  // namespace

int diagnostics_scratch_blocks() { return kDiagBlocks; }

// scratch: (scratch_blocks + 1) * 6 doubles of device memory owned by the caller
lbm_status run_diagnostics(Runtime &rt, const DiagLayout &L, double *scratch, int scratch_blocks, double *host_out6)
{
    double *out = scratch + (size_t)scratch_blocks * 6;
    if (L.q == 19)
        k_diag_partial<19><<<scratch_blocks, kDiagThreads, 0, rt.stream>>>(L, scratch);
    else
        k_diag_partial<9><<<scratch_blocks, kDiagThreads, 0, rt.stream>>>(L, scratch);
    LBM_TRY(rt.launched("k_diag_partial"));
    k_diag_final<<<1, kDiagThreads, 0, rt.stream>>>(scratch, scratch_blocks, out);
    LBM_TRY(rt.launched("k_diag_final"));
    LBM_CUDA(rt, cudaMemcpyAsync(host_out6, out, 6 * sizeof(double), cudaMemcpyDeviceToHost, rt.stream));
    LBM_CUDA(rt, cudaStreamSynchronize(rt.stream));
    return LBM_OK;
}

}  // namespace lbm
