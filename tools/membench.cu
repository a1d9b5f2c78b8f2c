// membench.cu -- what HBM bandwidth does the D3Q19 push access pattern get with NO arithmetic?
// 38 read streams + 38 shifted write streams per node, for two layouts:
//   SoA   [plane][k][node]                         (the library's layout)
//   AoSoA [plane][group of GS nodes][k][GS nodes]
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/membench tools/membench.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
__constant__ int CX[19] = {0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1};
__constant__ int CY[19] = {0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0};
__constant__ int CZ[19] = {0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1};

template <int GS>   // GS = 0: SoA
__device__ __forceinline__ long long at(long long pl, int zs, int k, long long idx)
{
    if (GS == 0) return ((long long)zs * 19 + k) * pl + idx;
    return (((long long)zs * (pl / GS) + idx / GS) * 19 + k) * GS + idx % GS;
}
template <int GS, int NT, int SETS>
__global__ void __launch_bounds__(NT) k_push(int nx, int ny, long long pl, const double *fin, const double *gin, double *fout, double *gout)
{
    const long long idx = (long long)blockIdx.x * NT + threadIdx.x;
    const int zs = 1 + blockIdx.y;
    const int ix = idx % nx, iy = idx / nx;
    if (ix < 1 || ix > nx - 2 || iy < 1 || iy > ny - 2) return;
    double F[19], G[19];
#pragma unroll
    for (int k = 0; k < 19; k++) { F[k] = fin[at<GS>(pl, zs, k, idx)]; if (SETS == 2) G[k] = gin[at<GS>(pl, zs, k, idx)]; }
#pragma unroll
    for (int k = 0; k < 19; k++) {
        const long long t = at<GS>(pl, zs + CZ[k], k, idx + CY[k] * nx + CX[k]);
        fout[t] = F[k] * 1.0000001;
        if (SETS == 2) gout[t] = G[k] * 1.0000001;
    }
}
template <int GS, int NT>
float run(int nx, int ny, int nz, long long pl, double *f0, double *g0, double *f1, double *g1, int reps)
{
    dim3 grid((unsigned)(((long long)nx * ny + NT - 1) / NT), nz);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) k_push<GS, NT, 2><<<grid, NT>>>(nx, ny, pl, f0, g0, f1, g1);
    cudaEventRecord(a);
    for (int i = 0; i < reps; i++) k_push<GS, NT, 2><<<grid, NT>>>(nx, ny, pl, i & 1 ? f1 : f0, i & 1 ? g1 : g0, i & 1 ? f0 : f1, i & 1 ? g0 : g1);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}
int main()
{
    const int nx = 512, ny = 512, nz = 256;
    const long long pl = (long long)nx * ny;
    const size_t n = (size_t)(nz + 2) * 19 * pl;
    double *f0, *g0, *f1, *g1;
    CK(cudaMalloc(&f0, n * 8)); CK(cudaMalloc(&g0, n * 8)); CK(cudaMalloc(&f1, n * 8)); CK(cudaMalloc(&g1, n * 8));
    CK(cudaMemset(f0, 0, n * 8)); CK(cudaMemset(g0, 0, n * 8)); CK(cudaMemset(f1, 0, n * 8)); CK(cudaMemset(g1, 0, n * 8));
    const double bytes = 608.0 * nx * ny * nz;
    float ms;
    ms = run<0, 256>(nx, ny, nz, pl, f0, g0, f1, g1, 10);  printf("SoA        NT=256: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<0, 128>(nx, ny, nz, pl, f0, g0, f1, g1, 10);  printf("SoA        NT=128: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<0, 512>(nx, ny, nz, pl, f0, g0, f1, g1, 10);  printf("SoA        NT=512: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<32, 256>(nx, ny, nz, pl, f0, g0, f1, g1, 10); printf("AoSoA GS=32 NT=256: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<64, 256>(nx, ny, nz, pl, f0, g0, f1, g1, 10); printf("AoSoA GS=64 NT=256: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<128, 256>(nx, ny, nz, pl, f0, g0, f1, g1, 10); printf("AoSoA GS=128 NT=256: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = run<512, 256>(nx, ny, nz, pl, f0, g0, f1, g1, 10); printf("AoSoA GS=512 NT=256: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    CK(cudaGetLastError());
    return 0;
}
