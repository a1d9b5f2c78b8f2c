// fp64bench.cu -- what rate of UNFUSED double-precision instructions (DADD, DMUL; the library is built -fmad=false,
// so its kernels contain no DFMA from source arithmetic) does this GPU sustain?  The denominator of the FP64-pipe
// roofline of the MRT kernels.  Every thread runs 16 independent dependency chains, half
// additions and half multiplications; the grid fills every SM with 2048 threads.
// Build: nvcc -O3 -fmad=false -gencode arch=compute_100a,code=sm_100a -o tools/fp64bench tools/fp64bench.cu
// Prints one JSON line: thread-level instructions per second (divide by 32 for warp instructions).
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s: %s\"}\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void __launch_bounds__(256) k_fp64(double *out, int iters, double a, double m)
{
    double x[8], y[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = threadIdx.x + i; y[i] = 1.0 + 1e-9 * (threadIdx.x + i); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) { x[i] = x[i] + a; y[i] = y[i] * m; }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i] + y[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4000;
    double *out;
    CK(cudaMalloc(&out, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    k_fp64<<<blocks, threads>>>(out, iters, 1e-3, 1.0000001);   // warm-up
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        CK(cudaEventRecord(e0));
        k_fp64<<<blocks, threads>>>(out, iters, 1e-3, 1.0000001);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    const double instr = (double)blocks * threads * iters * 8 * 16;   // 8 DADD + 8 DMUL per inner round, 8 rounds
    printf("{\"fp64_unfused_ginstr_per_s\": %.1f, \"ms\": %.3f, \"sms\": %d, \"per_sm_per_clock_at_max\": %.2f, \"how\": "
           "\"16 independent DADD/DMUL chains per thread, 2048 threads per SM, best of 5\"}\n",
           instr / (best * 1e-3) / 1e9, best, prop.multiProcessorCount,
           instr / (best * 1e-3) / prop.multiProcessorCount / (prop.clockRate * 1e3));
    return 0;
}
