// dispatch_order.cu -- in which order does the hardware start the blocks of a 3-D grid?  Every block takes a ticket
// when it starts and burns a few microseconds; the host then compares the ticket of block (x, y, z) with its linear
// index x + gx*(y + gy*z).  The one-pass D3Q19 step (csrc/d3q19_step.cuh) assumes "x fastest, then y, then z" and
// only loses time, not correctness, if that is wrong.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/dispatch_order tools/dispatch_order.cu
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(128, 4) k_order(unsigned *ticket, unsigned *out, int spin)
{
    __shared__ unsigned t;
    if (threadIdx.x == 0) {
        t = atomicAdd(ticket, 1u);
        out[blockIdx.x + gridDim.x * (blockIdx.y + (size_t)gridDim.y * blockIdx.z)] = t;
    }
    __syncthreads();
    long long t0 = clock64();
    while (clock64() - t0 < spin + (t % 7) * (spin / 8)) {}     // blocks of slightly different length
}
int main(int argc, char **argv)
{
    const int gx = argc > 1 ? atoi(argv[1]) : 64, gy = argc > 2 ? atoi(argv[2]) : 525, gz = argc > 3 ? atoi(argv[3]) : 33;
    const size_t n = (size_t)gx * gy * gz;
    unsigned *ticket, *out;
    cudaMalloc(&ticket, 4);
    cudaMalloc(&out, n * 4);
    cudaMemset(ticket, 0, 4);
    k_order<<<dim3(gx, gy, gz), 128, 38912>>>(ticket, out, 8000);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("{\"error\": \"launch failed\"}\n"); return 1; }
    std::vector<unsigned> h(n);
    cudaMemcpy(h.data(), out, n * 4, cudaMemcpyDeviceToHost);
    long long maxdev = 0;
    double sumabs = 0;
    std::vector<long long> dev(n);
    for (size_t i = 0; i < n; i++) {
        dev[i] = (long long)h[i] - (long long)i;
        maxdev = std::max(maxdev, std::llabs(dev[i]));
        sumabs += (double)std::llabs(dev[i]);
    }
    std::vector<long long> a(dev);
    for (auto &v : a) v = std::llabs(v);
    std::sort(a.begin(), a.end());
    printf("{\"grid\": [%d, %d, %d], \"blocks\": %zu, \"ticket_minus_linear_index\": {\"mean_abs\": %.1f, \"p50\": %lld, \"p99\": %lld, \"max\": %lld}, "
           "\"first_tickets_of_block_0_1_gx_gxgy\": [%u, %u, %u, %u]}\n",
           gx, gy, gz, n, sumabs / n, a[n / 2], a[(size_t)(n * 0.99)], maxdev, h[0], h[1], h[gx], h[(size_t)gx * gy]);
    return 0;
}
