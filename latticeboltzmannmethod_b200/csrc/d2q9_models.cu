// d2q9_models.cu -- the D2Q9 programs on one GPU: Shan-Chen
// (ShanChen/shanchen.cpp), binary liquid (FreeEnergy/binary.cpp), MRT
// (MRT/mrt.cpp, MRT/another_mrt.cpp), TRT diffusion (DiffusionCylinder/cylinder.cpp) and the
// binary liquid around a curved solid (CurvedSolid/binary_shape.cpp).  Per-thread arithmetic: d2q9_threads.h.
#include <vector>

#include "d2q9_host.h"
#include "d2q9_threads.h"
#include "lbm_internal.h"

namespace lbm {
namespace {

constexpr int kThreads = 128;   // measured at 8192^2: 128-thread blocks beat 256 for every D2Q9 kernel (TRT 0.68 -> 0.64 ms, MRT-GS 1.95 -> 1.81 ms)
inline unsigned blocks_for(long long n) { return (unsigned)((n + kThreads - 1) / kThreads); }

// ---- Shan-Chen kernels ----
__global__ void __launch_bounds__(kThreads) k_sc_init(lbm2::ScParams P, const double *rho, const double *u1, const double *u2, double *f)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::sc_init_thread(P, rho, u1, u2, f, i);
}
__global__ void __launch_bounds__(kThreads) k_sc_density(lbm2::ScParams P, const double *f, double *rho, double *psi)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::sc_density_thread(P, f, rho, psi, i);
}
// WRITE (all step kernels below): leave the macroscopic fields behind; only the last step of an lbm_step(n) batch does
template <bool WRITE>
__global__ void __launch_bounds__(kThreads) k_sc_collide(lbm2::ScParams P, const double *f, const double *rho, const double *psi,
                                                          double *f2, double *u1, double *u2)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::sc_collide_thread(P, f, rho, psi, f2, u1, u2, (int)(i % P.nx), (int)(i / P.nx), WRITE);
}
// One launch per Shan-Chen step: a TX x TY tile of nodes per block.  Every thread loads the nine populations of
// its node, sums them and puts psi = 1 - exp(-rho) into shared memory; the first threads of the block do the
// same for the ring of nodes around the tile (periodic), whose populations are re-read through L2; after one
// barrier each thread has the nine psi values of its force stencil and collides out of registers.  The rho and
// psi arrays of the two-kernel form and their 2 x 16 B per node of traffic disappear.
template <int TX, int TY, int MINB, bool WRITE>
__global__ void __launch_bounds__(TX *TY, MINB) k_sc_fused(lbm2::ScParams P, const double *__restrict__ f, double *__restrict__ f2,
                                                     double *__restrict__ rho, double *__restrict__ u1, double *__restrict__ u2)
{
    __shared__ double psi_s[TY + 2][TX + 2];
    const int nx = P.nx, ny = P.ny;
    const long long n = (long long)nx * ny;
    const int lx = threadIdx.x % TX, ly = threadIdx.x / TX;
    const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
    // the node of this thread; past the edge of the grid it wraps, which is exactly the periodic neighbour
    // the last column (row) needs, and the thread then only contributes psi
    const bool active = x0 + lx < nx && y0 + ly < ny;
    // periodic wrap by conditional subtraction: the operands are below twice the period plus a tile (an integer
    // remainder costs ~20 instructions, and the ring needs two per position)
    auto wrap = [](int v, int period) { while (v >= period) v -= period; return v; };
    const int gx = wrap(x0 + lx, nx), gy = wrap(y0 + ly, ny);
    const long long i = (long long)gy * nx + gx;
    double F[9], dense = 0;
#pragma unroll
    for (int k = 0; k < 9; k++) F[k] = f[k * n + i];
#pragma unroll
    for (int k = 0; k < 9; k++) dense += F[k];
    psi_s[ly + 1][lx + 1] = 1.0 - lbm_exp(-dense);
    constexpr int kRing = 2 * TX + 2 * TY + 4;
    for (int r = threadIdx.x; r < kRing; r += TX * TY) {
        int rx, ry;   // position in the (TX+2) x (TY+2) frame
        if (r < TX + 2) { rx = r; ry = 0; }
        else if (r < 2 * (TX + 2)) { rx = r - (TX + 2); ry = TY + 1; }
        else if (r < 2 * (TX + 2) + TY) { rx = 0; ry = r - 2 * (TX + 2) + 1; }
        else { rx = TX + 1; ry = r - 2 * (TX + 2) - TY + 1; }
        const int hx = wrap(x0 + rx - 1 + nx, nx), hy = wrap(y0 + ry - 1 + ny, ny);
        const long long h = (long long)hy * nx + hx;
        double d = 0;
#pragma unroll
        for (int k = 0; k < 9; k++) d += f[k * n + h];
        psi_s[ry][rx] = 1.0 - lbm_exp(-d);
    }
    __syncthreads();
    if (!active) return;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    double ps[9];
#pragma unroll
    for (int k = 0; k < 9; k++) ps[k] = psi_s[ly + 1 + CY[k]][lx + 1 + CX[k]];
    double v1, v2;
    lbm2::sc_collide_core(P, F, dense, ps[0], ps, v1, v2);
    if (WRITE) { rho[i] = dense; u1[i] = v1; u2[i] = v2; }
    const int xm = gx == 0 ? nx - 1 : gx - 1, xp = gx == nx - 1 ? 0 : gx + 1;
    const int ym = gy == 0 ? ny - 1 : gy - 1, yp = gy == ny - 1 ? 0 : gy + 1;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int x2 = CX[k] > 0 ? xp : (CX[k] < 0 ? xm : gx), y2 = CY[k] > 0 ? yp : (CY[k] < 0 ? ym : gy);
        f2[k * n + (long long)y2 * nx + x2] = F[k];
    }
}
// ---- binary kernels (node index i = iX*NY+iY, iY fastest) ----
__global__ void __launch_bounds__(kThreads) k_b2_wallphase(lbm2::B2Params P, double *phase)
{
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;   // binary.cpp:136-140
    if (ix >= P.nx) return;
    phase[(long long)ix * P.ny] = phase[(long long)ix * P.ny + 1] - P.wall_gradient;
    phase[(long long)ix * P.ny + P.ny - 1] = phase[(long long)ix * P.ny + P.ny - 2] - P.wall_gradient;
}
__global__ void __launch_bounds__(kThreads) k_b2_init(lbm2::B2Params P, const double *phase, double *fa, double *ga, double *fb,
                                                       double *gb, double *rho, double *ux, double *uy)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::b2_init_thread(P, phase, fa, ga, fb, gb, rho, ux, uy, (int)(i / P.ny), (int)(i % P.ny));
}
__global__ void __launch_bounds__(kThreads) k_b2_phase(lbm2::B2Params P, const double *g, double *phase)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::b2_phase_thread(P, g, phase, (int)((unsigned)i / (unsigned)P.ny), (int)((unsigned)i % (unsigned)P.ny));
}
template <bool WRITE, int NT = kThreads>
__global__ void __launch_bounds__(NT) k_b2_collide(lbm2::B2Params P, const double *f, const double *g, const double *phase,
                                                          double *f2, double *g2, double *rho, double *ux, double *uy)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::b2_collide_thread(P, f, g, phase, f2, g2, rho, ux, uy, (int)((unsigned)i / (unsigned)P.ny), (int)((unsigned)i % (unsigned)P.ny), WRITE);
}
// kernel_variant 0 (the default): one pass per binary-liquid step by MARCHING.  A block owns 128 consecutive rows (iY, the fast
// index) and walks over `xc` consecutive columns (iX).  The phase field lives only in shared memory, as a ring of four
// columns of 128 + 2 values; before column x is collided the block has summed the phase values of column x + 1 from
// the g populations it is about to collide anyway, so HBM sees every population once in and once out and the phase
// pass (88 B per update on top of 288) is gone -- without any ordering between blocks.  What it costs instead: the g
// values of the two columns either side of a block's range and of the two rows either side of its 128 are read by the
// neighbouring blocks as well (2/xc + 2/128 of g).  All populations travel through shared memory in a three-column
// pipeline of 8-byte asynchronous copies (every thread its own column: only the ring needs the one barrier per column).
__device__ __forceinline__ void b2_cp_async8(double *smem, const double *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
constexpr int kB2MarchSmem = ((3 * 9 + 3 * 9) * 128 + 3 * 9 * 2 + 4 * 130) * (int)sizeof(double);
template <bool WRITE>
__global__ void __launch_bounds__(128, 3) k_b2_march(lbm2::B2Params P, const double *f, const double *g, double *f2, double *g2,
                                                     double *phase, double *rho, double *uxo, double *uyo, int xc)
{
    extern __shared__ __align__(16) double b2m_smem[];
    double *fb = b2m_smem;                 // [3][9][128] f of the columns in flight
    double *gb = fb + 3 * 9 * 128;         // [3][9][128] g
    double *hb = gb + 3 * 9 * 128;         // [3][9][2]   g of the rows just outside the block's 128 (threads 0 and 127)
    double *ring = hb + 3 * 9 * 2;         // [4][130]    phase field of columns x-1 .. x+1 (+ one being written)
    const int NX = P.nx, NY = P.ny;
    const long long n = (long long)NX * NY;
    const int t = threadIdx.x;
    const int y0 = blockIdx.x * 128, x0 = blockIdx.y * xc;
    const int ncol = min(xc, NX - x0);                       // columns this block collides: x0 .. x0 + ncol - 1
    const int iy = y0 + t;
    const bool in_y = iy < NY, active = iy >= 1 && iy <= NY - 2;
    const int sy = min(max(iy, 1), NY - 2);                  // wall rows take their neighbour row's sum (binary.cpp:245-249)
    int hy = -1;
    if (t == 0 && y0 > 0) hy = y0 - 1;
    if (t == 127 && y0 + 128 < NY) hy = y0 + 128;
    const int hsy = min(max(hy, 1), NY - 2), hcol = t == 0 ? 0 : 1;

    // column r (relative to x0; -1 and ncol are the neighbours, of which only g is needed)
    auto request = [&](int r) {
        if (r <= ncol) {
            int x = x0 + r;
            x = x < 0 ? x + NX : (x >= NX ? x - NX : x);
            const int bi = (r + 1) % 3;
            const long long base = (long long)x * NY;
            if (in_y) {
#pragma unroll
                for (int k = 0; k < 9; k++) b2_cp_async8(gb + (bi * 9 + k) * 128 + t, g + k * n + base + sy);
                if (active && r >= 0 && r < ncol) {
#pragma unroll
                    for (int k = 0; k < 9; k++) b2_cp_async8(fb + (bi * 9 + k) * 128 + t, f + k * n + base + iy);
                }
            }
            if (hy >= 0) {
#pragma unroll
                for (int k = 0; k < 9; k++) b2_cp_async8(hb + (bi * 9 + k) * 2 + hcol, g + k * n + base + hsy);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    // this thread's phase values of column r into the ring (its copies have landed); binary.cpp:239-250, ascending k
    auto phase_of = [&](int r) {
        const int bi = (r + 1) % 3, ri = (r + 1) & 3;
        if (in_y) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < 9; k++) s += gb[(bi * 9 + k) * 128 + t];
            ring[ri * 130 + t + 1] = sy == iy ? s : s - P.wall_gradient;
        }
        if (hy >= 0) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < 9; k++) s += hb[(bi * 9 + k) * 2 + hcol];
            ring[ri * 130 + (t == 0 ? 0 : 129)] = hsy == hy ? s : s - P.wall_gradient;
        }
    };

    request(-1);
    request(0);
    request(1);
    asm volatile("cp.async.wait_group 2;\n" ::: "memory");
    phase_of(-1);
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    phase_of(0);
    for (int r = 0; r < ncol; r++) {
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");   // column r + 1 has landed
        phase_of(r + 1);
        __syncthreads();
        const int bi = (r + 1) % 3, ix = x0 + r;
        double F[9], G[9], ph9[9];
        if (active) {
#pragma unroll
            for (int k = 0; k < 9; k++) {
                F[k] = fb[(bi * 9 + k) * 128 + t];
                G[k] = gb[(bi * 9 + k) * 128 + t];
            }
        }
        request(r + 2);                                          // into the buffers of column r - 1
        const long long i = (long long)ix * NY + iy;
        if (WRITE && in_y) phase[i] = ring[((r + 1) & 3) * 130 + t + 1];   // wall rows included
        if (active) {
            const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
#pragma unroll
            for (int k = 0; k < 9; k++) ph9[k] = ring[((r + 1 + CX[k]) & 3) * 130 + t + 1 + CY[k]];
            double d, vx, vy;
            lbm2::b2_collide_core(P, F, G, ph9, d, vx, vy);
            if (WRITE) { rho[i] = d; uxo[i] = vx; uyo[i] = vy; }
            lbm2::b2_push(P, F, G, f2, g2, ix, iy);
        }
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}
// ---- MRT kernels ----
__global__ void __launch_bounds__(kThreads) k_mrt_init(lbm2::MrtParams P, const double *rho, const double *ux, const double *uy, double *f)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::mrt_init_thread(P, rho, ux, uy, f, i);
}
template <int VARIANT, bool WRITE>
__global__ void __launch_bounds__(kThreads) k_mrt_step(lbm2::MrtParams P, const double *__restrict__ f, const double *__restrict__ force,
                                                        double *__restrict__ f2, double *__restrict__ rho, double *__restrict__ ux,
                                                        double *__restrict__ uy)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny)
        lbm2::mrt_thread_v<VARIANT>(P, f, force, f2, rho, ux, uy, (int)((unsigned)i / (unsigned)P.ny), (int)((unsigned)i % (unsigned)P.ny), WRITE);
}
// in-place AA-pattern form of the MRT step: ONE population array (the default; kernel_variant 1 = two arrays).
// Measured at 8192^2: 1.80 / 1.68 ms (GS / Hermite) against 1.82 / 1.85 ms for the two-array push, 4.8 GB less memory.
template <int VARIANT, bool ODD, bool WRITE>
__global__ void __launch_bounds__(kThreads) k_mrt_step_aa(lbm2::MrtParams P, double *f, const double *__restrict__ force,
                                                           double *__restrict__ rho, double *__restrict__ ux, double *__restrict__ uy)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;          // nx*ny < 2^31 (checked at create)
    if (i < (unsigned)(P.nx * P.ny))
        lbm2::mrt_thread_aa<VARIANT, ODD>(P, f, force, rho, ux, uy, (int)(i / (unsigned)P.ny), (int)(i % (unsigned)P.ny), WRITE);
}
// the streamed populations in the reference's AoS order out of the AA array, whichever half of the cycle comes next
__global__ void __launch_bounds__(kThreads) k_mrt_aa_to_aos9(lbm2::MrtParams P, const double *f, int odd_next, double *aos)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x, n = (long long)P.nx * P.ny;
    if (t >= n * 9) return;
    const long long i = t / 9;
    aos[t] = lbm2::mrt_aa_streamed(P, f, odd_next, (int)(t % 9), (int)(i / P.ny), (int)(i % P.ny));
}
// ---- TRT diffusion kernels (node index i = iY*nx+iX) ----
__global__ void __launch_bounds__(kThreads) k_trt_init(lbm2::TrtParams P, const double *rho, const double *ux, const double *uy,
                                                        double *fa, double *fb)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::trt_init_thread(P, rho, ux, uy, fa, fb, i);
}
template <bool WRITE>
__global__ void __launch_bounds__(kThreads) k_trt_step(lbm2::TrtParams P, const double *__restrict__ f,
                                                        const unsigned short *__restrict__ links, const double *__restrict__ ux,
                                                        const double *__restrict__ uy, double *__restrict__ f2, double *__restrict__ rho)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::trt_thread(P, f, links, ux, uy, f2, rho, i, WRITE);
}
// ---- binary liquid around a curved solid (node index i = iY*nx+iX) ----
__global__ void __launch_bounds__(kThreads) k_bs_init(lbm2::BsParams P, const double *rho, const double *phi, const double *ux,
                                                       const double *uy, double *fa, double *ga, double *fb, double *gb)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::bs_init_thread(P, rho, phi, ux, uy, fa, ga, fb, gb, i);
}
__global__ void __launch_bounds__(kThreads) k_bs_phase(lbm2::BsParams P, const double *__restrict__ g,
                                                        const unsigned short *__restrict__ links, double *__restrict__ phi)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::bs_phase_thread(P, g, links, phi, i);
}
template <bool WRITE, int NT = kThreads, int MINB = 2>
__global__ void __launch_bounds__(NT, MINB) k_bs_collide(lbm2::BsParams P, const double *__restrict__ f, const double *__restrict__ g,
                                                          const double *__restrict__ phi, const unsigned short *__restrict__ links,
                                                          double *__restrict__ f2, double *__restrict__ g2, double *__restrict__ rho,
                                                          double *__restrict__ ux, double *__restrict__ uy)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.nx * P.ny) lbm2::bs_collide_thread(P, f, g, phi, links, f2, g2, rho, ux, uy, i, WRITE);
}
// ---- AoS <-> SoA at the boundary ----
__global__ void __launch_bounds__(kThreads) k_aos_to_soa9(long long n, const double *aos, double *soa)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n * 9) soa[(t % 9) * n + t / 9] = aos[t];
}
__global__ void __launch_bounds__(kThreads) k_soa_to_aos9(long long n, const double *soa, double *aos)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n * 9) aos[t] = soa[(t % 9) * n + t / 9];
}

// common storage of the 2-D models
class Model2D : public Model {
public:
    Model2D(Runtime &rt, const lbm_config &cfg, int nsets) : rt_(rt), cfg_(cfg), nsets_(nsets) {}
    ~Model2D() override
    {
        cudaSetDevice(rt_.device);
        for (int b = 0; b < 2; b++) { rt_.dev_free(f_[b]); rt_.dev_free(g_[b]); }
        for (double *p : fields_) rt_.dev_free(p);
        rt_.dev_free(aux_); rt_.dev_free(diag_);
    }
    lbm_status alloc_common()
    {
        if (cfg_.nx < 3 || cfg_.ny < 3 || cfg_.nz != 1) return rt_.fail(LBM_EINVAL, "2-D model: need nx,ny >= 3 and nz == 1");
        n_ = (long long)cfg_.nx * cfg_.ny;
        if (n_ >= (1ll << 31)) return rt_.fail(LBM_EINVAL, "2-D model: nx*ny must stay below 2^31");
        for (int b = 0; b < 2; b++) {
            LBM_CUDA(rt_, rt_.dev_alloc(&f_[b], 9 * n_ * sizeof(double)));
            LBM_CUDA(rt_, cudaMemsetAsync(f_[b], 0, 9 * n_ * sizeof(double), rt_.stream));
            if (nsets_ == 2) {
                LBM_CUDA(rt_, rt_.dev_alloc(&g_[b], 9 * n_ * sizeof(double)));
                LBM_CUDA(rt_, cudaMemsetAsync(g_[b], 0, 9 * n_ * sizeof(double), rt_.stream));
            }
        }
        for (double *&p : fields_) {
            LBM_CUDA(rt_, rt_.dev_alloc(&p, n_ * sizeof(double)));
            LBM_CUDA(rt_, cudaMemsetAsync(p, 0, n_ * sizeof(double), rt_.stream));
        }
        LBM_CUDA(rt_, rt_.dev_alloc(&diag_, (size_t)(diagnostics_scratch_blocks() + 1) * 6 * sizeof(double)));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }
    size_t local_nodes() const override { return (size_t)n_; }
    int q() const override { return 9; }
    int nsets() const override { return nsets_; }
    virtual bool has_field(int field) const = 0;
    lbm_status set_field(int field, const double *host) override
    {
        if (!has_field(field)) return rt_.fail(LBM_EINVAL, "this model has no such field");
        LBM_CUDA(rt_, cudaMemcpyAsync(fields_[field], host, n_ * sizeof(double), cudaMemcpyHostToDevice, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }
    lbm_status get_field(int field, double *host) override
    {
        if (!has_field(field)) return rt_.fail(LBM_EINVAL, "this model has no such field");
        LBM_CUDA(rt_, cudaMemcpyAsync(host, fields_[field], n_ * sizeof(double), cudaMemcpyDeviceToHost, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        return LBM_OK;
    }
    lbm_status set_populations(int set, const double *host) override
    {
        if (set < 0 || set >= nsets_) return rt_.fail(LBM_EINVAL, "no such population set");
        double *stage = nullptr;
        LBM_CUDA(rt_, rt_.dev_alloc(&stage, 9 * n_ * sizeof(double)));
        lbm_status st = rt_.cuda(cudaMemcpyAsync(stage, host, 9 * n_ * sizeof(double), cudaMemcpyHostToDevice, rt_.stream), "H2D populations");
        if (st == LBM_OK) {
            k_aos_to_soa9<<<blocks_for(9 * n_), kThreads, 0, rt_.stream>>>(n_, stage, set == 0 ? f_[cur_] : g_[cur_]);
            st = rt_.launched("k_aos_to_soa9");
        }
        if (st == LBM_OK) st = after_set_populations(set);
        if (st == LBM_OK) st = rt_.cuda(cudaStreamSynchronize(rt_.stream), "sync");
        rt_.dev_free(stage);
        return st;
    }
    virtual lbm_status after_set_populations(int) { return LBM_OK; }
    lbm_status get_populations(int set, double *host) override
    {
        if (set < 0 || set >= nsets_) return rt_.fail(LBM_EINVAL, "no such population set");
        double *stage = nullptr;
        LBM_CUDA(rt_, rt_.dev_alloc(&stage, 9 * n_ * sizeof(double)));
        k_soa_to_aos9<<<blocks_for(9 * n_), kThreads, 0, rt_.stream>>>(n_, set == 0 ? f_[cur_] : g_[cur_], stage);
        lbm_status st = rt_.launched("k_soa_to_aos9");
        if (st == LBM_OK) st = rt_.cuda(cudaMemcpyAsync(host, stage, 9 * n_ * sizeof(double), cudaMemcpyDeviceToHost, rt_.stream), "D2H populations");
        if (st == LBM_OK) st = rt_.cuda(cudaStreamSynchronize(rt_.stream), "sync");
        rt_.dev_free(stage);
        return st;
    }
    lbm_status diag_common(int skip_border, double *mass, double *mom, double *phase_mass, int *has_nan)
    {
        DiagLayout L;
        L.f = f_[cur_]; L.g = nsets_ == 2 ? g_[cur_] : nullptr;
        L.stride_k = n_; L.plane_stride = 0; L.planes = 1; L.nx = cfg_.nx; L.ny = cfg_.ny; L.q = 9;
        L.skip_border = skip_border; L.node_major_x = 0;
        double out[6];
        LBM_TRY(run_diagnostics(rt_, L, diag_, diagnostics_scratch_blocks(), out));
        if (mass) *mass = out[0];
        if (mom) { mom[0] = out[1]; mom[1] = out[2]; mom[2] = 0.0; }
        if (phase_mass) *phase_mass = out[4];
        if (has_nan) *has_nan = out[5] != 0.0;
        return LBM_OK;
    }

protected:
    Runtime &rt_;
    lbm_config cfg_;
    int nsets_, cur_ = 0;
    long long n_ = 0;
    double *f_[2] = {nullptr, nullptr}, *g_[2] = {nullptr, nullptr};
    double *fields_[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // indexed by lbm_field
    double *aux_ = nullptr, *diag_ = nullptr;
};

class ShanChen : public Model2D {
public:
    ShanChen(Runtime &rt, const lbm_config &cfg) : Model2D(rt, cfg, 1) {}
    lbm_status setup()
    {
        if (cfg_.kernel_variant != 0 && cfg_.kernel_variant != 1)
            return rt_.fail(LBM_EINVAL, "Shan-Chen: kernel_variant must be 0 (fused tile kernel) or 1 (density pass + collide)");
        LBM_TRY(alloc_common());
        P_ = lbm2::sc_params(cfg_.nx, cfg_.ny, cfg_.tau, cfg_.g);
        LBM_CUDA(rt_, rt_.dev_alloc(&aux_, n_ * sizeof(double)));   // psi
        return LBM_OK;
    }
    bool has_field(int f) const override { return f == LBM_RHO || f == LBM_UX || f == LBM_UY; }
    lbm_status init() override
    {
        k_sc_init<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY], f_[cur_]);
        return rt_.launched("k_sc_init");
    }
    lbm_status step(int nsteps) override
    {
        // kernel_variant 0 (default): the fused tile kernel, 32x4 tile capped at 64 registers, 8 blocks/SM (1.98 ms at
        // 8192^2); 1: the two-kernel form (density pass + collide), the cross-check.  Measured and removed:
        // tiles of 32x16 and 32x8 (2.27 - 2.58 ms); a marching kernel (128 columns per block walking over 64 rows, psi
        // in a shared-memory ring -- the scheme that took the binary model from 3.94 to 3.42 ms): bit-exact, 2.03 ms.
        if (cfg_.kernel_variant != 1) {
            for (int t = 0; t < nsteps; t++) {
                rt_.about_to_launch();
                launch_fused<32, 4, 8>(t == nsteps - 1);
                LBM_TRY(rt_.launched("k_sc_fused"));
                cur_ = 1 - cur_;
            }
            return LBM_OK;
        }
        for (int t = 0; t < nsteps; t++) {
            rt_.about_to_launch();
            k_sc_density<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, f_[cur_], fields_[LBM_RHO], aux_);
            LBM_TRY(rt_.launched("k_sc_density"));
            rt_.about_to_launch();
            auto kern = t == nsteps - 1 ? k_sc_collide<true> : k_sc_collide<false>;
            kern<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, f_[cur_], fields_[LBM_RHO], aux_, f_[1 - cur_], fields_[LBM_UX],
                                                              fields_[LBM_UY]);
            LBM_TRY(rt_.launched("k_sc_collide"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    lbm_status diagnostics(double *m, double *mom, double *pm, int *nan) override { return diag_common(0, m, mom, pm, nan); }

private:
    template <int TX, int TY, int MINB>
    void launch_fused(bool write)
    {
        const dim3 grid((cfg_.nx + TX - 1) / TX, (cfg_.ny + TY - 1) / TY);
        auto kern = write ? k_sc_fused<TX, TY, MINB, true> : k_sc_fused<TX, TY, MINB, false>;
        kern<<<grid, TX * TY, 0, rt_.stream>>>(P_, f_[cur_], f_[1 - cur_], fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY]);
    }
    lbm2::ScParams P_{};
};

class Binary2D : public Model2D {
public:
    Binary2D(Runtime &rt, const lbm_config &cfg) : Model2D(rt, cfg, 2) {}
    lbm_status setup()
    {
        if (cfg_.ny < 4) return rt_.fail(LBM_EINVAL, "binary D2Q9: ny must be at least 4");
        if (cfg_.kernel_variant != 0 && cfg_.kernel_variant != 1)
            return rt_.fail(LBM_EINVAL, "binary D2Q9: kernel_variant must be 0 (marching kernel, one pass) or 1 (phase pass + collide)");
        LBM_TRY(alloc_common());
        P_ = lbm2::B2Params{cfg_.nx, cfg_.ny, cfg_.aconst, cfg_.kconst, cfg_.gammaconst, cfg_.wall_gradient,
                            cfg_.tau_gas, cfg_.tau_liq - cfg_.tau_gas, 1.0 / cfg_.tau_phi};
        return LBM_OK;
    }
    bool has_field(int f) const override { return f == LBM_RHO || f == LBM_PHASE || f == LBM_UX || f == LBM_UY; }
    lbm_status init() override
    {
        k_b2_wallphase<<<blocks_for(cfg_.nx), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_PHASE]);
        LBM_TRY(rt_.launched("k_b2_wallphase"));
        k_b2_init<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_PHASE], f_[0], g_[0], f_[1], g_[1],
                                                               fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY]);
        return rt_.launched("k_b2_init");
    }
    lbm_status step_march(int nsteps)
    {
        int xc = 64;
        if (const char *e = getenv("LBM_TUNE_XC")) xc = std::max(1, atoi(e));   // measurement knob
        const dim3 grid((cfg_.ny + 127) / 128, (cfg_.nx + xc - 1) / xc);
        if (!march_ready_) {    // 60 KB of dynamic shared memory: opt in once
            LBM_CUDA(rt_, cudaFuncSetAttribute(k_b2_march<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kB2MarchSmem));
            LBM_CUDA(rt_, cudaFuncSetAttribute(k_b2_march<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kB2MarchSmem));
            march_ready_ = true;
        }
        for (int t = 0; t < nsteps; t++) {
            const bool w = t == nsteps - 1;
            auto kern = w ? k_b2_march<true> : k_b2_march<false>;
            rt_.about_to_launch();
            kern<<<grid, 128, kB2MarchSmem, rt_.stream>>>(P_, f_[cur_], g_[cur_], f_[1 - cur_], g_[1 - cur_], fields_[LBM_PHASE],
                                                          fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY], xc);
            LBM_TRY(rt_.launched("k_b2_march"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    lbm_status step(int nsteps) override
    {
        // kernel_variant 0 (default): the marching kernel, one pass with the phase field in shared memory: 3.42 ms at
        // 8192^2 = 0.86 of the measured HBM peak (columns per block 32 / 64 / 128 / 256: 3.41 / 3.42 / 3.45 / 3.47 ms)
        if (cfg_.kernel_variant == 0) return step_march(nsteps);
        // kernel_variant 1: phase pass + collide with 128-thread blocks, round 1's default and the cross-check of the
        // marching kernel (3.94 ms; 256-thread blocks: 4.23 ms).  What else was built, measured and removed (DESIGN.md 4
        // has the numbers, the history the code): a lagging one-pass kernel in the manner of the D3Q19 k_step_one
        // (4.68 ms), a staged, streamed form of the collide kernel (4.19 - 4.79 ms), shared-memory tile kernels
        // (4.52 - 5.37 ms).
        for (int t = 0; t < nsteps; t++) {
            rt_.about_to_launch();
            k_b2_phase<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, g_[cur_], fields_[LBM_PHASE]);
            LBM_TRY(rt_.launched("k_b2_phase"));
            rt_.about_to_launch();
            auto kern = t == nsteps - 1 ? k_b2_collide<true, 128> : k_b2_collide<false, 128>;
            kern<<<(unsigned)((n_ + 127) / 128), 128, 0, rt_.stream>>>(P_, f_[cur_], g_[cur_], fields_[LBM_PHASE], f_[1 - cur_], g_[1 - cur_],
                                                                      fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY]);
            LBM_TRY(rt_.launched("k_b2_collide"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    // wall rows are never streamed into, so a load must reach both buffers
    lbm_status after_set_populations(int set) override
    {
        double *src = set == 0 ? f_[cur_] : g_[cur_], *dst = set == 0 ? f_[1 - cur_] : g_[1 - cur_];
        return rt_.cuda(cudaMemcpyAsync(dst, src, 9 * n_ * sizeof(double), cudaMemcpyDeviceToDevice, rt_.stream), "copy populations");
    }
    lbm_status diagnostics(double *m, double *mom, double *pm, int *nan) override { return diag_common(2, m, mom, pm, nan); }

private:
    lbm2::B2Params P_{};
    bool march_ready_ = false;
};

class Mrt : public Model2D {
public:
    Mrt(Runtime &rt, const lbm_config &cfg) : Model2D(rt, cfg, 1) {}
    lbm_status setup()
    {
        if (cfg_.kernel_variant != 0 && cfg_.kernel_variant != 1)
            return rt_.fail(LBM_EINVAL, "MRT: kernel_variant must be 0 (in place, one population array) or 1 (two arrays)");
        LBM_TRY(alloc_common());
        aa_ = cfg_.kernel_variant != 1;      // default: in-place AA streaming; 1: two arrays, push streaming
        if (aa_) { rt_.dev_free(f_[1]); f_[1] = nullptr; }      // the point of the in-place form: one population array
        std::vector<double> force(cfg_.nx);
        lbm2::mrt_build_tables(&P_, cfg_.model == LBM_MRT_GS_D2Q9 ? 0 : 1, cfg_.nx, cfg_.ny, cfg_.omega, cfg_.omega_bulk,
                               cfg_.force_ampl, force.data());
        LBM_CUDA(rt_, rt_.dev_alloc(&aux_, cfg_.nx * sizeof(double)));
        LBM_CUDA(rt_, cudaMemcpy(aux_, force.data(), cfg_.nx * sizeof(double), cudaMemcpyHostToDevice));
        return LBM_OK;
    }
    bool has_field(int f) const override { return f == LBM_RHO || f == LBM_UX || f == LBM_UY; }
    lbm_status init() override
    {
        odd_next_ = false;
        k_mrt_init<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY], f_[cur_]);
        return rt_.launched("k_mrt_init");
    }
    lbm_status step(int nsteps) override
    {
        for (int t = 0; t < nsteps; t++) {
            rt_.about_to_launch();
            const bool w = t == nsteps - 1;
            if (aa_) {
                // even and odd half of the AA cycle alternate on the one array
                auto kern = P_.variant == 0
                    ? (odd_next_ ? (w ? k_mrt_step_aa<0, true, true> : k_mrt_step_aa<0, true, false>) : (w ? k_mrt_step_aa<0, false, true> : k_mrt_step_aa<0, false, false>))
                    : (odd_next_ ? (w ? k_mrt_step_aa<1, true, true> : k_mrt_step_aa<1, true, false>) : (w ? k_mrt_step_aa<1, false, true> : k_mrt_step_aa<1, false, false>));
                kern<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, f_[0], aux_, fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY]);
                LBM_TRY(rt_.launched("k_mrt_step_aa"));
                odd_next_ = !odd_next_;
                continue;
            }
            auto kern = P_.variant == 0 ? (w ? k_mrt_step<0, true> : k_mrt_step<0, false>) : (w ? k_mrt_step<1, true> : k_mrt_step<1, false>);
            kern<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, f_[cur_], aux_, f_[1 - cur_], fields_[LBM_RHO], fields_[LBM_UX],
                                                              fields_[LBM_UY]);
            LBM_TRY(rt_.launched("k_mrt_step"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    // AA mode: after an even step the array holds post-collision values in the opposite slots; the streamed populations
    // are gathered out of it (no second array exists to un-swap into)
    lbm_status get_populations(int set, double *host) override
    {
        if (!aa_ || !odd_next_) return Model2D::get_populations(set, host);
        if (set != 0) return rt_.fail(LBM_EINVAL, "no such population set");
        double *stage = nullptr;
        LBM_CUDA(rt_, rt_.dev_alloc(&stage, 9 * n_ * sizeof(double)));
        k_mrt_aa_to_aos9<<<blocks_for(9 * n_), kThreads, 0, rt_.stream>>>(P_, f_[0], 1, stage);
        lbm_status st = rt_.launched("k_mrt_aa_to_aos9");
        if (st == LBM_OK) st = rt_.cuda(cudaMemcpyAsync(host, stage, 9 * n_ * sizeof(double), cudaMemcpyDeviceToHost, rt_.stream), "D2H populations");
        if (st == LBM_OK) st = rt_.cuda(cudaStreamSynchronize(rt_.stream), "sync");
        rt_.dev_free(stage);
        return st;
    }
    lbm_status after_set_populations(int) override { odd_next_ = false; return LBM_OK; }   // loaded in natural order
    lbm_status diagnostics(double *m, double *mom, double *pm, int *nan) override
    {
        LBM_TRY(diag_common(0, m, mom, pm, nan));
        // in the swapped half of the cycle slot k holds direction opp(k): the sums over all (periodic) nodes are those of
        // the streamed populations with the momentum's sign flipped
        if (aa_ && odd_next_ && mom) { mom[0] = -mom[0]; mom[1] = -mom[1]; }
        return LBM_OK;
    }

private:
    lbm2::MrtParams P_{};
    bool aa_ = false, odd_next_ = false;   // aa_: in-place AA streaming on f_[0]; odd_next_: the next step is an odd one
};

// DiffusionCylinder/cylinder.cpp: one launch per step; the geometry mask becomes one 16-bit link word per node
class Trt : public Model2D {
public:
    Trt(Runtime &rt, const lbm_config &cfg) : Model2D(rt, cfg, 1) {}
    ~Trt() override { rt_.dev_free(links_); }
    lbm_status setup()
    {
        if (cfg_.kernel_variant != 0) return rt_.fail(LBM_EINVAL, "TRT diffusion: kernel_variant must be 0");
        LBM_TRY(alloc_common());
        P_ = lbm2::TrtParams{cfg_.nx, cfg_.ny, 2.0 - cfg_.omega, cfg_.omega, cfg_.ce, cfg_.conc_wall};   // cylinder.cpp:38-40
        LBM_CUDA(rt_, rt_.dev_alloc(&links_, n_ * sizeof(unsigned short)));
        LBM_CUDA(rt_, cudaMemset(links_, 0, n_ * sizeof(unsigned short)));
        return LBM_OK;
    }
    bool has_field(int f) const override { return f == LBM_RHO || f == LBM_UX || f == LBM_UY; }
    lbm_status set_geometry(const int *mask) override
    {
        const int nx = cfg_.nx, ny = cfg_.ny;
        const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
        std::vector<unsigned short> links((size_t)n_);
        for (long long i = 0; i < n_; i++) {
            if (mask[i] < -1 || mask[i] > 1) return rt_.fail(LBM_EINVAL, "lbm_set_geometry: mask values are 1 (fluid), 0 (boundary), -1 (solid)");
            if (mask[i] != 1) continue;
            const int ix = (int)(i % nx), iy = (int)(i / nx);
            if (ix == 0 || iy == 0 || ix == nx - 1 || iy == ny - 1)
                return rt_.fail(LBM_EINVAL, "lbm_set_geometry: a fluid node lies on the edge of the array");
            for (int k = 1; k < 9; k++)
                if (mask[i + CY[k] * nx + CX[k]] == -1)
                    return rt_.fail(LBM_EINVAL, "lbm_set_geometry: a fluid node touches a solid node (no boundary node between them)");
            links[(size_t)i] = lbm2::trt_links_thread(P_, mask, i);
        }
        LBM_CUDA(rt_, cudaMemcpyAsync(links_, links.data(), (size_t)n_ * sizeof(unsigned short), cudaMemcpyHostToDevice, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        have_geometry_ = true;
        return LBM_OK;
    }
    lbm_status init() override
    {
        if (!have_geometry_) return rt_.fail(LBM_ESTATE, "TRT diffusion: call lbm_set_geometry before lbm_init");
        k_trt_init<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY], f_[0], f_[1]);
        return rt_.launched("k_trt_init");
    }
    lbm_status step(int nsteps) override
    {
        for (int t = 0; t < nsteps; t++) {
            rt_.about_to_launch();
            auto kern = t == nsteps - 1 ? k_trt_step<true> : k_trt_step<false>;
            kern<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, f_[cur_], links_, fields_[LBM_UX], fields_[LBM_UY], f_[1 - cur_],
                                                              fields_[LBM_RHO]);
            LBM_TRY(rt_.launched("k_trt_step"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    // boundary and solid nodes are never streamed into, so a load must reach both buffers
    lbm_status after_set_populations(int) override
    {
        return rt_.cuda(cudaMemcpyAsync(f_[1 - cur_], f_[cur_], 9 * n_ * sizeof(double), cudaMemcpyDeviceToDevice, rt_.stream), "copy populations");
    }
    lbm_status diagnostics(double *mass, double *mom, double *pm, int *has_nan) override
    {
        DiagLayout L;
        L.f = f_[cur_]; L.g = nullptr;
        L.stride_k = n_; L.plane_stride = 0; L.planes = 1; L.nx = cfg_.nx; L.ny = cfg_.ny; L.q = 9;
        L.skip_border = 0; L.node_major_x = 1; L.fluid = links_;
        double out[6];
        LBM_TRY(run_diagnostics(rt_, L, diag_, diagnostics_scratch_blocks(), out));
        if (mass) *mass = out[0];
        if (mom) { mom[0] = out[1]; mom[1] = out[2]; mom[2] = 0.0; }
        if (pm) *pm = 0.0;
        if (has_nan) *has_nan = out[5] != 0.0;
        return LBM_OK;
    }

private:
    lbm2::TrtParams P_{};
    unsigned short *links_ = nullptr;
    bool have_geometry_ = false;
};

// CurvedSolid/binary_shape.cpp: phase pass + collide/push pass per step on a masked periodic domain
class BinaryShape : public Model2D {
public:
    BinaryShape(Runtime &rt, const lbm_config &cfg) : Model2D(rt, cfg, 2) {}
    ~BinaryShape() override { rt_.dev_free(links_); }
    lbm_status setup()
    {
        if (cfg_.kernel_variant != 0) return rt_.fail(LBM_EINVAL, "curved-solid binary liquid: kernel_variant must be 0");
        LBM_TRY(alloc_common());
        P_ = lbm2::BsParams{cfg_.nx, cfg_.ny, cfg_.aconst, cfg_.kconst, cfg_.gammaconst, cfg_.wall_gradient, cfg_.tau_gas,
                            cfg_.tau_liq - cfg_.tau_gas, 1.0 / cfg_.tau_phi, cfg_.force[0], cfg_.force[1]};
        LBM_CUDA(rt_, rt_.dev_alloc(&links_, n_ * sizeof(unsigned short)));
        LBM_CUDA(rt_, cudaMemset(links_, 0, n_ * sizeof(unsigned short)));
        return LBM_OK;
    }
    bool has_field(int f) const override { return f == LBM_RHO || f == LBM_PHASE || f == LBM_UX || f == LBM_UY; }
    lbm_status set_geometry(const int *mask) override
    {
        const int nx = cfg_.nx, ny = cfg_.ny;
        const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
        std::vector<unsigned short> links((size_t)n_);
        for (long long i = 0; i < n_; i++) {
            if (mask[i] < -1 || mask[i] > 1) return rt_.fail(LBM_EINVAL, "lbm_set_geometry: mask values are 1 (fluid), 0 (bounce-back), -1 (solid)");
            const int ix = (int)(i % nx), iy = (int)(i / nx);
            if (mask[i] == 0) {
                // binary_shape.cpp:364,430 index the neighbours of a bounce-back node without the periodic wrap
                if (ix == 0 || iy == 0 || ix == nx - 1 || iy == ny - 1)
                    return rt_.fail(LBM_EINVAL, "lbm_set_geometry: a bounce-back node lies on the edge of the array");
                const int dir = lbm2::bs_main_dir(mask, nx, i);
                if (dir == 0 || mask[i + CY[dir] * nx + CX[dir]] != 1)
                    return rt_.fail(LBM_EINVAL, "lbm_set_geometry: a bounce-back node has no wall normal pointing at a fluid node "
                                                "(binary_shape.cpp:556-588 would misalign main_dir)");
            } else if (mask[i] == 1) {
                for (int k = 1; k < 9; k++)
                    if (mask[(long long)((iy + CY[k] + ny) % ny) * nx + (ix + CX[k] + nx) % nx] == -1)
                        return rt_.fail(LBM_EINVAL, "lbm_set_geometry: a fluid node touches a solid node (no bounce-back node between them)");
            }
            links[(size_t)i] = lbm2::bs_links_thread(P_, mask, i);
        }
        LBM_CUDA(rt_, cudaMemcpyAsync(links_, links.data(), (size_t)n_ * sizeof(unsigned short), cudaMemcpyHostToDevice, rt_.stream));
        LBM_CUDA(rt_, cudaStreamSynchronize(rt_.stream));
        have_geometry_ = true;
        return LBM_OK;
    }
    lbm_status init() override
    {
        if (!have_geometry_) return rt_.fail(LBM_ESTATE, "curved-solid binary liquid: call lbm_set_geometry before lbm_init");
        k_bs_init<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, fields_[LBM_RHO], fields_[LBM_PHASE], fields_[LBM_UX], fields_[LBM_UY],
                                                               f_[0], g_[0], f_[1], g_[1]);
        return rt_.launched("k_bs_init");
    }
    lbm_status step(int nsteps) override
    {
        for (int t = 0; t < nsteps; t++) {
            rt_.about_to_launch();
            k_bs_phase<<<blocks_for(n_), kThreads, 0, rt_.stream>>>(P_, g_[cur_], links_, fields_[LBM_PHASE]);
            LBM_TRY(rt_.launched("k_bs_phase"));
            rt_.about_to_launch();
            // 128 threads x 4 blocks/SM (128 registers): 3.96 ms at 8192^2.  Measured and removed: 128 x 3 (168 registers,
            // 4.75 ms at the time), 256 x 2 (4.45 ms), and a marching kernel in the manner of k_b2_march -- sums of g in a
            // shared-memory ring of rows y-2 .. y+2, the phase value of a bounce-back neighbour looked up through its
            // link word -- at 4.16 ms: the per-neighbour look-ups and the solid nodes it walks over cost more than the
            // phase pass.
            const bool w = t == nsteps - 1;
            auto kern = w ? k_bs_collide<true, 128, 4> : k_bs_collide<false, 128, 4>;
            kern<<<(unsigned)((n_ + 127) / 128), 128, 0, rt_.stream>>>(P_, f_[cur_], g_[cur_], fields_[LBM_PHASE], links_, f_[1 - cur_],
                                                                      g_[1 - cur_], fields_[LBM_RHO], fields_[LBM_UX], fields_[LBM_UY]);
            LBM_TRY(rt_.launched("k_bs_collide"));
            cur_ = 1 - cur_;
        }
        return LBM_OK;
    }
    // bounce-back and solid nodes are never streamed into, so a load must reach both buffers
    lbm_status after_set_populations(int set) override
    {
        double *src = set == 0 ? f_[cur_] : g_[cur_], *dst = set == 0 ? f_[1 - cur_] : g_[1 - cur_];
        return rt_.cuda(cudaMemcpyAsync(dst, src, 9 * n_ * sizeof(double), cudaMemcpyDeviceToDevice, rt_.stream), "copy populations");
    }
    lbm_status diagnostics(double *mass, double *mom, double *pm, int *has_nan) override
    {
        DiagLayout L;
        L.f = f_[cur_]; L.g = g_[cur_];
        L.stride_k = n_; L.plane_stride = 0; L.planes = 1; L.nx = cfg_.nx; L.ny = cfg_.ny; L.q = 9;
        L.skip_border = 0; L.node_major_x = 1; L.fluid = links_;
        double out[6];
        LBM_TRY(run_diagnostics(rt_, L, diag_, diagnostics_scratch_blocks(), out));
        if (mass) *mass = out[0];
        if (mom) { mom[0] = out[1]; mom[1] = out[2]; mom[2] = 0.0; }
        if (pm) *pm = out[4];
        if (has_nan) *has_nan = out[5] != 0.0;
        return LBM_OK;
    }

private:
    lbm2::BsParams P_{};
    unsigned short *links_ = nullptr;
    bool have_geometry_ = false;
};

template <class T>
Model *make(Runtime &rt, const lbm_config &cfg, lbm_status *st)
{
    T *m = new T(rt, cfg);
    *st = m->setup();
    if (*st != LBM_OK) {
        delete m;
        return nullptr;
    }
    return m;
}

}  // namespace

Model *make_d2q9_shanchen(Runtime &rt, const lbm_config &cfg, lbm_status *st) { return make<ShanChen>(rt, cfg, st); }
Model *make_d2q9_binary(Runtime &rt, const lbm_config &cfg, lbm_status *st) { return make<Binary2D>(rt, cfg, st); }
Model *make_d2q9_mrt(Runtime &rt, const lbm_config &cfg, lbm_status *st) { return make<Mrt>(rt, cfg, st); }
Model *make_d2q9_trt(Runtime &rt, const lbm_config &cfg, lbm_status *st) { return make<Trt>(rt, cfg, st); }
Model *make_d2q9_binary_shape(Runtime &rt, const lbm_config &cfg, lbm_status *st) { return make<BinaryShape>(rt, cfg, st); }

}  // namespace lbm
