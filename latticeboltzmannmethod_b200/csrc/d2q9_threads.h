// d2q9_threads.h -- what one thread of each D2Q9 kernel does (Shan-Chen, binary
// liquid, MRT).  Plain C++ so the CPU unit test (tests/emu/) can run the same
// code; the __global__ wrappers live in d2q9_models.cu.
//
// Storage: populations SoA [k][n]; the flat node index is the reference's own
// (Shan-Chen: i = iY*nx+iX, shanchen.cpp:140; binary/MRT: i = iX*NY+iY,
// binary.cpp:24, mrt.cpp:14), so fields need no transposition at the boundary.
// Arrays hold the streamed populations, i.e. what the reference's f holds at the
// start of a step.  Compile with -fmad=false / -ffp-contract=off.
#ifndef LBM_D2Q9_THREADS_H
#define LBM_D2Q9_THREADS_H

#include "d3q19_node.h"  // LBM_HD
#include "lbm_exp.h"

namespace lbm2 {

// shanchen.cpp:18-20, binary.cpp:44-47, mrt.cpp:63-67
#define LBM2_CX {0, 1, 0, -1, 0, 1, -1, -1, 1}
#define LBM2_CY {0, 0, 1, 0, -1, 1, 1, -1, -1}
#define LBM2_W {4.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0, 1.0 / 36.0, 1.0 / 36.0, 1.0 / 36.0, 1.0 / 36.0}
#define LBM2_OPP {0, 3, 4, 1, 2, 7, 8, 5, 6}

// ======================= Shan-Chen (ShanChen/shanchen.cpp) =======================
struct ScParams {
    int nx, ny;
    double tau, g;
    // the two per-run constants of the node update, divided once on the host (IEEE division: the same bits as
    // dividing in every thread): -1.0/tau of shanchen.cpp:182 and (1 - 0.5/tau) of :163
    double neg_inv_tau, force_factor;
};
inline ScParams sc_params(int nx, int ny, double tau, double g) { return ScParams{nx, ny, tau, g, -1.0 / tau, 1 - 0.5 / tau}; }

// shanchen.cpp:124-132 plus psi = 1 - exp(-rho), which the reference re-evaluates
// for every neighbour (:152-157); same function, same argument, computed once.
LBM_HD void sc_density_thread(const ScParams &P, const double *f, double *rho, double *psi, long long i)
{
    const long long n = (long long)P.nx * P.ny;
    double r = 0;
#pragma unroll
    for (int k = 0; k < 9; k++) r += f[k * n + i];
    rho[i] = r;
    psi[i] = 1.0 - lbm_exp(-r);
}

// shanchen.cpp:136-186 for one node whose populations F, density and the pseudopotential psi of the node (ps_c)
// and of its nine stencil nodes (ps[k], k = 0 is the node itself) are in registers: force (float accumulators),
// shifted velocity, force populations, closed-form equilibria, in-place BGK.  F holds the post-collision values on return.
LBM_HD void sc_collide_core(const ScParams &P, double (&F)[9], double dense, double ps_c, const double (&ps)[9], double &v1, double &v2)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    float fx = 0.0, fy = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        fx += W[k] * CX[k] * ps[k];
        fy += W[k] * CY[k] * ps[k];
    }
    fx = -P.g * ps_c * fx;
    fy = -P.g * ps_c * fy;
    v1 = (F[1] - F[3] + F[5] - F[6] - F[7] + F[8]) / dense + fx / (2.0 * dense);
    v2 = (F[2] - F[4] + F[5] + F[6] - F[7] - F[8]) / dense + fy / (2.0 * dense);
    double fpop[9];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        fpop[k] = W[k] * P.force_factor *
                  ((3 * (cx - v1) + 9 * cx * (cx * v1 + cy * v2)) * fx + (3 * (cy - v2) + 9 * cy * (cx * v1 + cy * v2)) * fy);
    }
    const float usq = v1 * v1 + v2 * v2;
    double feq[9];
    feq[0] = 4.0 / 9.0 * dense * (1.0 - 1.5 * usq);
    feq[1] = 1.0 / 9.0 * dense * (1.0 + 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[2] = 1.0 / 9.0 * dense * (1.0 + 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[3] = 1.0 / 9.0 * dense * (1.0 - 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[4] = 1.0 / 9.0 * dense * (1.0 - 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[5] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 + v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[6] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 + v2) + 4.5 * (-v1 + v2) * (-v1 + v2) - 1.5 * usq);
    feq[7] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 - v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[8] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 - v2) + 4.5 * (v1 - v2) * (v1 - v2) - 1.5 * usq);
#pragma unroll
    for (int k = 0; k < 9; k++) F[k] += P.neg_inv_tau * (F[k] - feq[k]) + fpop[k];
}

// the two-kernel form: rho and psi come from sc_density_thread's arrays; periodic push
LBM_HD void sc_collide_thread(const ScParams &P, const double *f, const double *rho, const double *psi, double *f2,
                              double *u1, double *u2, int ix, int iy, bool write_fields = true)
{
    const int nx = P.nx, ny = P.ny;
    const long long n = (long long)nx * ny, i = (long long)iy * nx + ix;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    long long nb[9];
    double ps[9], F[9];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int x2 = (ix + CX[k] + nx) % nx, y2 = (iy + CY[k] + ny) % ny;
        nb[k] = (long long)nx * y2 + x2;
        ps[k] = psi[nb[k]];
        F[k] = f[k * n + i];
    }
    double v1, v2;
    sc_collide_core(P, F, rho[i], psi[i], ps, v1, v2);
    if (write_fields) {      // only the last step of a batch has to leave the fields behind (shanchen.cpp:190-207 reads them)
        u1[i] = v1;
        u2[i] = v2;
    }
#pragma unroll
    for (int k = 0; k < 9; k++) f2[k * n + nb[k]] = F[k];
}

// shanchen.cpp:95-113 generalised to a given (rho, u): f = feq(rho, u)
LBM_HD void sc_init_thread(const ScParams &P, const double *rho, const double *u1, const double *u2, double *f, long long i)
{
    const long long n = (long long)P.nx * P.ny;
    const double dense = rho[i], v1 = u1[i], v2 = u2[i];
    const float usq = v1 * v1 + v2 * v2;
    double feq[9];
    feq[0] = 4.0 / 9.0 * dense * (1.0 - 1.5 * usq);
    feq[1] = 1.0 / 9.0 * dense * (1.0 + 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[2] = 1.0 / 9.0 * dense * (1.0 + 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[3] = 1.0 / 9.0 * dense * (1.0 - 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[4] = 1.0 / 9.0 * dense * (1.0 - 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[5] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 + v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[6] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 + v2) + 4.5 * (-v1 + v2) * (-v1 + v2) - 1.5 * usq);
    feq[7] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 - v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[8] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 - v2) + 4.5 * (v1 - v2) * (v1 - v2) - 1.5 * usq);
    for (int k = 0; k < 9; k++) f[k * n + i] = feq[k];
}

// ======================= binary liquid (FreeEnergy/binary.cpp) =======================
struct B2Params {
    int nx, ny;
    double aconst, kconst, gammaconst, wall_gradient, tau_gas, dtau, omega_phi;
};

// the float-typed tables of binary.cpp:48-59, promoted exactly as the reference promotes them
#define LBM2_B_WXX {0.0f, (float)(1.0 / 3.0), (float)(-1.0 / 6.0), (float)(1.0 / 3.0), (float)(-1.0 / 6.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0)}
#define LBM2_B_WYY {0.0f, (float)(-1.0 / 6.0), (float)(1.0 / 3.0), (float)(-1.0 / 6.0), (float)(1.0 / 3.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0), (float)(-1.0 / 24.0)}
#define LBM2_B_WXY {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, (float)(1.0 / 4.0), (float)(-1.0 / 4.0), (float)(1.0 / 4.0), (float)(-1.0 / 4.0)}
#define LBM2_B_GX {0.0f, (float)(4.0 / 12.0), 0.0f, (float)(-4.0 / 12.0), 0.0f, (float)(1.0 / 12.0), (float)(-1.0 / 12.0), (float)(-1.0 / 12.0), (float)(1.0 / 12.0)}
#define LBM2_B_GY {0.0f, 0.0f, (float)(4.0 / 12.0), 0.0f, (float)(-4.0 / 12.0), (float)(1.0 / 12.0), (float)(1.0 / 12.0), (float)(-1.0 / 12.0), (float)(-1.0 / 12.0)}
#define LBM2_B_LAP {(float)(-20.0 / 6.0), (float)(4.0 / 6.0), (float)(4.0 / 6.0), (float)(4.0 / 6.0), (float)(4.0 / 6.0), (float)(1.0 / 6.0), (float)(1.0 / 6.0), (float)(1.0 / 6.0), (float)(1.0 / 6.0)}

// binary.cpp:239-250: phase = sum of g on interior rows, wall rows from their neighbour row
LBM_HD void b2_phase_thread(const B2Params &P, const double *g, double *phase, int ix, int iy)
{
    const int NY = P.ny;
    if (iy < 1 || iy > NY - 2) return;
    const long long n = (long long)P.nx * NY, i = (long long)ix * NY + iy;
    double ph = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) ph += g[k * n + i];
    phase[i] = ph;
    if (iy == 1) phase[i - 1] = ph - P.wall_gradient;
    if (iy == NY - 2) phase[i + 1] = ph - P.wall_gradient;
}

// binary.cpp:265-278: the three stencil sums over the nine phase values ph[k] (k = 0 the node itself), ascending k
LBM_HD void b2_stencil_sums(const double (&ph)[9], double &gx, double &gy, double &lap)
{
    const float GX[9] = LBM2_B_GX, GY[9] = LBM2_B_GY, LAP[9] = LBM2_B_LAP;
    double al = 0.0, ax = 0.0, ay = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        al += LAP[k] * ph[k];
        ax += GX[k] * ph[k];
        ay += GY[k] * ph[k];
    }
    gx = ax; gy = ay; lap = al;
}
LBM_HD void b2_stencils(const B2Params &P, const double *phase, int ix, int iy, double &gx, double &gy, double &lap)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    double ph[9];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int x2 = (ix + CX[k] + P.nx) % P.nx, y2 = (iy + CY[k] + P.ny) % P.ny;
        ph[k] = phase[(long long)x2 * P.ny + y2];
    }
    b2_stencil_sums(ph, gx, gy, lap);
}

// binary.cpp:291-313 (and :169-195)
LBM_HD void b2_equilibria(const B2Params &P, double d, double ph, double ux, double uy, double gx, double gy, double lap,
                          double *feq, double *geq)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    const float WXX[9] = LBM2_B_WXX, WYY[9] = LBM2_B_WYY, WXY[9] = LBM2_B_WXY;
    double sum = 0.0, sum_phase = 0.0;
    const double ps = ph * ph;
    const double pb = d / 3.0 + P.aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - P.kconst * ph * lap;
    const double mu = P.gammaconst * (P.aconst * (-ph + ph * ph * ph) - P.kconst * lap);
#pragma unroll
    for (int k = 1; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        feq[k] = W[k] * (3.0 * pb + 3.0 * d * (cx * ux + cy * uy) +
                         4.5 * d * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy + 2.0 * ux * uy * cx * cy)) +
                 P.kconst * (WXX[k] * gx * gx + WYY[k] * gy * gy + WXY[k] * gx * gy);
        geq[k] = W[k] * (3.0 * mu + 3.0 * ph * (cx * ux + cy * uy) +
                         4.5 * ph * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy + 2.0 * ux * uy * cx * cy));
        sum += feq[k];
        sum_phase += geq[k];
    }
    feq[0] = d - sum;
    geq[0] = ph - sum_phase;
}

// binary.cpp:142-196 for one node (all nodes, wall rows included); writes both buffers so the
// wall rows -- which the reference never streams into -- persist across the buffer swap
LBM_HD void b2_init_thread(const B2Params &P, const double *phase, double *fa, double *ga, double *fb, double *gb,
                           double *rho, double *ux, double *uy, int ix, int iy)
{
    const long long n = (long long)P.nx * P.ny, i = (long long)ix * P.ny + iy;
    double gx, gy, lap, feq[9], geq[9];
    b2_stencils(P, phase, ix, iy, gx, gy, lap);
    rho[i] = 1.0; ux[i] = 0.0; uy[i] = 0.0;
    b2_equilibria(P, 1.0, phase[i], 0.0, 0.0, gx, gy, lap, feq, geq);
    for (int k = 0; k < 9; k++) {
        fa[k * n + i] = feq[k]; ga[k * n + i] = geq[k];
        fb[k * n + i] = feq[k]; gb[k * n + i] = geq[k];
    }
}

// binary.cpp:252-327 for one node whose populations F, G and the nine phase values of its stencil are in
// registers; F and G hold the post-collision values on return
LBM_HD void b2_collide_core(const B2Params &P, double (&F)[9], double (&G)[9], const double (&ph9)[9], double &d, double &ux, double &uy)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    double jx = 0.0, jy = 0.0;
    d = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        d += F[k];
        jx += F[k] * CX[k];
        jy += F[k] * CY[k];
    }
    ux = jx / d;
    uy = jy / d;
    double gx, gy, lap, feq[9], geq[9];
    b2_stencil_sums(ph9, gx, gy, lap);
    const double ph = ph9[0];
    b2_equilibria(P, d, ph, ux, uy, gx, gy, lap, feq, geq);
    const double tau_rho = P.tau_gas + (ph + 1.0) / 2.0 * P.dtau;
    const double omega_rho = 1.0 / tau_rho;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        F[k] = F[k] * (1.0 - omega_rho) + omega_rho * feq[k];
        G[k] = G[k] * (1.0 - P.omega_phi) + P.omega_phi * geq[k];
    }
}

// :201-233 (bounce-back rows) and :341-350 (pull streaming) as one push: a population that would enter a
// wall row comes back to its own node reversed
LBM_HD void b2_push(const B2Params &P, const double (&F)[9], const double (&G)[9], double *f2, double *g2, int ix, int iy)
{
    const int NX = P.nx, NY = P.ny;
    const long long n = (long long)NX * NY, i = (long long)ix * NY + iy;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY, OPP[9] = LBM2_OPP;
    const int xs[3] = {ix == 0 ? NX - 1 : ix - 1, ix, ix == NX - 1 ? 0 : ix + 1};   // periodic in x, by comparison
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const int ty = iy + CY[k];
        if (ty == 0 || ty == NY - 1) {
            f2[OPP[k] * n + i] = F[k];
            g2[OPP[k] * n + i] = G[k];
        } else {
            const long long t = (long long)xs[1 + CX[k]] * NY + ty;
            f2[k * n + t] = F[k];
            g2[k * n + t] = G[k];
        }
    }
}

// the two-kernel form: the phase field comes from b2_phase_thread's array
LBM_HD void b2_collide_thread(const B2Params &P, const double *f, const double *g, const double *phase, double *f2,
                              double *g2, double *rho, double *uxo, double *uyo, int ix, int iy, bool write_fields = true)
{
    const int NX = P.nx, NY = P.ny;
    if (iy < 1 || iy > NY - 2) return;
    const long long n = (long long)NX * NY, i = (long long)ix * NY + iy;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    double F[9], G[9], ph9[9];
    const int xs[3] = {ix == 0 ? NX - 1 : ix - 1, ix, ix == NX - 1 ? 0 : ix + 1};   // periodic in x; 1 <= iy <= NY-2 needs no wrap
#pragma unroll
    for (int k = 0; k < 9; k++) {
        F[k] = f[k * n + i];
        G[k] = g[k * n + i];
        ph9[k] = phase[(long long)xs[1 + CX[k]] * NY + (iy + CY[k])];
    }
    double d, ux, uy;
    b2_collide_core(P, F, G, ph9, d, ux, uy);
    if (write_fields) { rho[i] = d; uxo[i] = ux; uyo[i] = uy; }
    b2_push(P, F, G, f2, g2, ix, iy);
}

// ======================= MRT (MRT/mrt.cpp, MRT/another_mrt.cpp) =======================
struct MrtParams {
    int nx, ny, variant;       // 0 = Gram-Schmidt + force (mrt.cpp), 1 = Hermite (another_mrt.cpp)
    double omegamat[9];
    double M[9][9];
    double B[9][9];            // GS: invM (mrt.cpp:82-91); Hermite: omegamat[m]*M[m][k] stored as B[k][m]
    double s3, c32, c3s15;     // sqrt(3), 3/sqrt(2), 3*sqrt(1.5) (another_mrt.cpp:144-152)
};

// one node: moments, (GS) half-force shift, equilibrium moments, relaxation in moment space,
// back-transform, periodic push.  mrt.cpp:146-251 / another_mrt.cpp:114-214.
// forcex_of_ix: the per-column force mrt.cpp:165 evaluates with cos(); built on the host with the
// same expression so the device needs no cos().
//
// The transforms are written against the sparsity of M, which is a property of the lattice, not of the run:
// M[m][k] is structurally zero where the row's polynomial in (cx, cy) vanishes (GS: 22 of 81 entries, Hermite 17),
// and the Gram-Schmidt rows are small integers.  Skipping a structurally zero term and replacing a product with
// +-1 by the operand are exact: the accumulators start at +0.0 and can never become -0.0, and x*1.0 == x.
// (What is NOT done: no re-association, no contraction, every remaining product and sum in the reference order.)
LBM_HD bool mrt_structural_nonzero(int variant, int m, int k)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const bool x = CX[k] != 0, y = CY[k] != 0;
    if (variant == 0) {   // mrt.cpp:71-80
        switch (m) {
        case 3: case 4: return x;
        case 5: case 6: return y;
        case 7: return x != y;      // cx^2 - cy^2
        case 8: return x && y;      // cx cy
        default: return true;
        }
    }
    switch (m) {          // another_mrt.cpp:75-84
    case 1: case 7: return x;
    case 2: case 8: return y;
    case 4: return x && y;
    default: return true;
    }
}
// mrt.cpp:71-80 evaluated: the Gram-Schmidt matrix (exact small integers; the unit test pins it to mrt_build_tables)
#define LBM2_MRT_GS_M                                                                                                       \
    {{1, 1, 1, 1, 1, 1, 1, 1, 1}, {-4, -1, -1, -1, -1, 2, 2, 2, 2}, {4, -2, -2, -2, -2, 1, 1, 1, 1},                         \
     {0, 1, 0, -1, 0, 1, -1, -1, 1}, {0, -2, 0, 2, 0, 1, -1, -1, 1}, {0, 0, 1, 0, -1, 1, 1, -1, -1},                          \
     {0, 0, -2, 0, 2, 1, 1, -1, -1}, {0, 1, -1, 1, -1, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 1, -1, 1, -1}}

// the collision of one node in place on its nine (streamed) populations fl; returns rho and the velocity the program
// writes out (GS: incl. the half-force shift).  mrt.cpp:146-240 / another_mrt.cpp:114-203.
template <int VARIANT>
LBM_HD void mrt_collide_v(const MrtParams &P, double (&fl)[9], const double *forcex_of_ix, int ix, double &rho, double &ux, double &uy)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    double eq[9], add[9];
    rho = 0.0; ux = 0.0; uy = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        rho += fl[k];
        ux += fl[k] * CX[k];
        uy += fl[k] * CY[k];
    }
    const double dense = rho;
    ux /= dense;
    uy /= dense;
    const double *om = P.omegamat;
    double momforce[9];
    if (VARIANT == 0) {
        const double forcex = forcex_of_ix[ix];
        const double forcey = 0;
        ux += forcex / (2.0 * dense);
        ux += forcey / (2.0 * dense);
        const double uxeq = ux, uyeq = uy;
        eq[0] = dense;
        eq[1] = -2.0 * dense + 3.0 * dense * (uxeq * uxeq + uyeq * uyeq);
        eq[2] = dense - 3.0 * dense * (uxeq * uxeq + uyeq * uyeq);
        eq[3] = dense * uxeq;
        eq[4] = -dense * uxeq;
        eq[5] = dense * uyeq;
        eq[6] = -dense * uyeq;
        eq[7] = dense * (uxeq * uxeq - uyeq * uyeq);
        eq[8] = dense * uxeq * uyeq;
        momforce[0] = 0.0;
        momforce[1] = (1.0 - om[1] / 2.0) * 6.0 * (uxeq * forcex + uyeq * forcey);
        momforce[2] = 0.0;
        momforce[3] = (1.0 - om[3] / 2.0) * forcex;
        momforce[4] = 0.0;
        momforce[5] = (1.0 - om[5] / 2.0) * forcey;
        momforce[6] = 0.0;
        momforce[7] = (1.0 - om[7] / 2.0) * 2.0 * (uxeq * forcex - uyeq * forcey);
        momforce[8] = (1.0 - om[8] / 2.0) * (uxeq * forcey + uyeq * forcex);
    } else {
        const double uxeq = ux, uyeq = uy;
        eq[0] = dense;
        eq[1] = P.s3 * dense * uxeq;
        eq[2] = P.s3 * dense * uyeq;
        eq[3] = P.c32 * dense * uxeq * uxeq;
        eq[4] = 3.0 * dense * uxeq * uyeq;
        eq[5] = P.c32 * dense * uyeq * uyeq;
        eq[6] = 4.5 * dense * uxeq * uxeq * uyeq * uyeq;
        eq[7] = P.c3s15 * dense * uxeq * uyeq * uyeq;
        eq[8] = P.c3s15 * dense * uxeq * uxeq * uyeq;
    }
    // forward transform of -f, plus equilibrium moments (and, GS, relaxation and forcing in moment space)
    const int MGS[9][9] = LBM2_MRT_GS_M;
#pragma unroll
    for (int m = 0; m < 9; m++) {
        double a = 0.0;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            if (!mrt_structural_nonzero(VARIANT, m, k)) continue;
            if (VARIANT == 0) {
                const int c = MGS[m][k];
                if (c == 1) a = a + (-fl[k]);
                else if (c == -1) a = a + fl[k];
                else a = a + (double)c * (-fl[k]);
            } else {
                a = a + P.M[m][k] * (-fl[k]);
            }
        }
        a += eq[m];
        if (VARIANT == 0) {
            a *= om[m];
            a += momforce[m];
        }
        add[m] = a;
    }
    // back transform
#pragma unroll
    for (int k = 0; k < 9; k++) {
        double addit = 0.0;
#pragma unroll
        for (int m = 0; m < 9; m++)
            if (mrt_structural_nonzero(VARIANT, m, k)) addit = addit + P.B[k][m] * add[m];
        fl[k] = VARIANT == 0 ? fl[k] + addit : fl[k] + W[k] * addit;
    }
}

template <int VARIANT>
LBM_HD void mrt_thread_v(const MrtParams &P, const double *f, const double *forcex_of_ix, double *f2, double *rho_o,
                         double *ux_o, double *uy_o, int ix, int iy, bool write_fields)
{
    const int NX = P.nx, NY = P.ny;
    const long long n = (long long)NX * NY, i = (long long)ix * NY + iy;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    double fl[9];
#pragma unroll
    for (int k = 0; k < 9; k++) fl[k] = f[k * n + i];
    double rho, ux, uy;
    mrt_collide_v<VARIANT>(P, fl, forcex_of_ix, ix, rho, ux, uy);
    if (write_fields) { rho_o[i] = rho; ux_o[i] = ux; uy_o[i] = uy; }
    const int xs[3] = {ix == 0 ? NX - 1 : ix - 1, ix, ix == NX - 1 ? 0 : ix + 1};
    const int ys[3] = {iy == 0 ? NY - 1 : iy - 1, iy, iy == NY - 1 ? 0 : iy + 1};
#pragma unroll
    for (int k = 0; k < 9; k++) f2[k * n + (xs[1 + CX[k]] * NY + ys[1 + CY[k]])] = fl[k];
}

// The same step with IN-PLACE "AA-pattern" streaming (Bailey et al. 2009) on ONE population array: an even step reads
// the node's own slots and writes the post-collision value of direction k into the OPPOSITE slot of the same node; an
// odd step reads direction k from the opposite slot of the upstream neighbour (which is where the even step left
// what streams into this node) and writes it into slot k of the downstream neighbour (which is where the next even
// step expects it).  Every step touches, per thread, one set of 9 slots for both its loads and its stores, so no second
// buffer exists and no thread reads what another writes.  Arithmetic: mrt_collide_v, unchanged.
template <int VARIANT, bool ODD>
LBM_HD void mrt_thread_aa(const MrtParams &P, double *f, const double *forcex_of_ix, double *rho_o, double *ux_o, double *uy_o,
                          int ix, int iy, bool write_fields)
{
    const int NX = P.nx, NY = P.ny;
    const long long n = (long long)NX * NY, i = (long long)ix * NY + iy;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY, OPP[9] = LBM2_OPP;
    // periodic neighbours by comparison (an integer remainder per direction would cost more than the collision's
    // address arithmetic); nx*ny < 2^31 is checked at create
    const int xs[3] = {ix == 0 ? NX - 1 : ix - 1, ix, ix == NX - 1 ? 0 : ix + 1};
    const int ys[3] = {iy == 0 ? NY - 1 : iy - 1, iy, iy == NY - 1 ? 0 : iy + 1};
    int up[9], down[9];      // neighbour against / along direction k
#pragma unroll
    for (int k = 0; k < 9; k++) {
        up[k] = xs[1 - CX[k]] * NY + ys[1 - CY[k]];
        down[k] = xs[1 + CX[k]] * NY + ys[1 + CY[k]];
    }
    double fl[9];
#pragma unroll
    for (int k = 0; k < 9; k++) fl[k] = ODD ? f[OPP[k] * n + up[k]] : f[k * n + i];
    double rho, ux, uy;
    mrt_collide_v<VARIANT>(P, fl, forcex_of_ix, ix, rho, ux, uy);
    if (write_fields) { rho_o[i] = rho; ux_o[i] = ux; uy_o[i] = uy; }
#pragma unroll
    for (int k = 0; k < 9; k++) {
        if (ODD) f[k * n + down[k]] = fl[k];
        else f[OPP[k] * n + i] = fl[k];
    }
}
// natural (streamed) population k of node (ix, iy) from the array in either phase of the AA cycle
LBM_HD double mrt_aa_streamed(const MrtParams &P, const double *f, int odd_next, int k, int ix, int iy)
{
    const int NX = P.nx, NY = P.ny;
    const long long n = (long long)NX * NY;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY, OPP[9] = LBM2_OPP;
    if (!odd_next) return f[k * n + (long long)ix * NY + iy];
    return f[OPP[k] * n + (long long)((ix - CX[k] + NX) % NX) * NY + (iy - CY[k] + NY) % NY];
}

LBM_HD void mrt_thread(const MrtParams &P, const double *f, const double *forcex_of_ix, double *f2, double *rho_o,
                       double *ux_o, double *uy_o, int ix, int iy, bool write_fields = true)
{
    if (P.variant == 0) mrt_thread_v<0>(P, f, forcex_of_ix, f2, rho_o, ux_o, uy_o, ix, iy, write_fields);
    else mrt_thread_v<1>(P, f, forcex_of_ix, f2, rho_o, ux_o, uy_o, ix, iy, write_fields);
}

// mrt.cpp:107-119 / another_mrt.cpp:97-109: second-order equilibrium from given rho, ux, uy
LBM_HD void mrt_init_thread(const MrtParams &P, const double *rho_i, const double *ux_i, const double *uy_i, double *f,
                            long long i)
{
    const long long n = (long long)P.nx * P.ny;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    const double rho = rho_i[i], ux = ux_i[i], uy = uy_i[i];
    const double fluxx = 3.0 * ux, fluxy = 3.0 * uy;
    const double qxx = 4.5 * ux * ux, qxy = 9.0 * ux * uy, qyy = 4.5 * uy * uy;
    for (int k = 0; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        f[k * n + i] = rho * W[k] * (1.0 + fluxx * cx + fluxy * cy + qxx * (cx * cx - 1.0 / 3.0) + qxy * cx * cy + qyy * (cy * cy - 1.0 / 3.0));
    }
}

// ======================= TRT diffusion with anti-bounce-back (DiffusionCylinder/cylinder.cpp) =======================
// Node index i = iY*nx+iX (cylinder.cpp:127).  `links` is built once from the geometry mask (1 fluid, 0
// boundary, -1 solid; cylinder.cpp:224-283): bit 0 = this node is fluid, bit k (1..8) = the neighbour in
// direction k is a boundary node.  The reference collides every fluid node (cylinder.cpp:126-205), then
// writes -f2[fluid][opp] + 2 w C into the boundary node (:208-221) and pulls (:337-347); pushed from the
// fluid node, the population sent towards a boundary node comes straight back into its own opposite slot.
struct TrtParams {
    int nx, ny;
    double omega_plus, omega_minus, ce, conc_wall;   // cylinder.cpp:33,38-43
};

// bit k of the link word (16 bits: fluid flag + 8 directions) from the mask; the host validates that no fluid node touches the array edge or a solid node
LBM_HD unsigned short trt_links_thread(const TrtParams &P, const int *mask, long long i)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    if (mask[i] != 1) return 0;
    unsigned short b = 1;
    for (int k = 1; k < 9; k++)
        if (mask[i + CY[k] * P.nx + CX[k]] == 0) b |= (unsigned short)(1u << k);
    // bit 0 doubles as "fluid"; k = 0 never leaves the node
    return b;
}

// cylinder.cpp:97-124: second-order equilibrium on every node, into both buffers (non-fluid nodes keep it)
LBM_HD void trt_init_thread(const TrtParams &P, const double *rho_i, const double *ux_i, const double *uy_i, double *fa, double *fb,
                            long long i)
{
    const long long n = (long long)P.nx * P.ny;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    const double dense_temp = rho_i[i], ux_temp = ux_i[i], uy_temp = uy_i[i];
    for (int k = 0; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        const double feq = W[k] * (dense_temp + 3.0 * dense_temp * (cx * ux_temp + cy * uy_temp) +
                                   4.5 * dense_temp * ((cx * cx - 1.0 / 3.0) * ux_temp * ux_temp + (cy * cy - 1.0 / 3.0) * uy_temp * uy_temp +
                                                       2.0 * ux_temp * uy_temp * cx * cy));
        fa[k * n + i] = feq;
        fb[k * n + i] = feq;
    }
}

// the even/odd pair (k, opposite k) of cylinder.cpp:147-200; wt = weights_trt[k], cu = cx[k]*ux+cy[k]*uy
struct TrtPair {
    double plus, minus, eq_plus, eq_minus;
};
LBM_HD TrtPair trt_pair(double fk, double fo, double wt, double dense, double cu, double u_sq, double ce)
{
    TrtPair r;
    r.plus = 0.5 * (fk + fo);
    r.minus = 0.5 * (fk - fo);
    r.eq_minus = wt * dense * cu;
    r.eq_plus = wt * dense * (ce + 0.5 * (3.0 * cu * cu - u_sq));
    return r;
}

LBM_HD void trt_thread(const TrtParams &P, const double *f, const unsigned short *links, const double *ux_i, const double *uy_i,
                       double *f2, double *rho_o, long long i, bool write_fields = true)
{
    const unsigned link = links[i];
    if (!(link & 1u)) return;
    const long long n = (long long)P.nx * P.ny;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY, OPP[9] = LBM2_OPP;
    const double W[9] = LBM2_W;
    double fl[9];
#pragma unroll
    for (int k = 0; k < 9; k++) fl[k] = f[k * n + i];
    double sum = 0;
#pragma unroll
    for (int k = 0; k < 9; k++) sum += fl[k];
    if (write_fields) rho_o[i] = sum;
    const double dense_temp = sum, ux_temp = ux_i[i], uy_temp = uy_i[i];
    const double u_sq = ux_temp * ux_temp + uy_temp * uy_temp;
    // pairs (1,3), (2,4), (5,7), (6,8); weights_trt = 1/3 on the axes, 1/12 on the diagonals (cylinder.cpp:50)
    const TrtPair a = trt_pair(fl[1], fl[3], 1.0 / 3.0, dense_temp, 1 * ux_temp + 0 * uy_temp, u_sq, P.ce);
    const TrtPair b = trt_pair(fl[2], fl[4], 1.0 / 3.0, dense_temp, 0 * ux_temp + 1 * uy_temp, u_sq, P.ce);
    const TrtPair c = trt_pair(fl[5], fl[7], 1.0 / 12.0, dense_temp, 1 * ux_temp + 1 * uy_temp, u_sq, P.ce);
    const TrtPair d = trt_pair(fl[6], fl[8], 1.0 / 12.0, dense_temp, -1 * ux_temp + 1 * uy_temp, u_sq, P.ce);
    const double eq_plus0 = dense_temp - 2.0 * (a.eq_plus + b.eq_plus + c.eq_plus + d.eq_plus);   // cylinder.cpp:196
    double out[9];
    out[0] = fl[0] - P.omega_plus * (fl[0] - eq_plus0) - P.omega_minus * (0.0 - 0.0);
    out[1] = fl[1] - P.omega_plus * (a.plus - a.eq_plus) - P.omega_minus * (a.minus - a.eq_minus);
    out[3] = fl[3] - P.omega_plus * (a.plus - a.eq_plus) - P.omega_minus * (-a.minus - -a.eq_minus);
    out[2] = fl[2] - P.omega_plus * (b.plus - b.eq_plus) - P.omega_minus * (b.minus - b.eq_minus);
    out[4] = fl[4] - P.omega_plus * (b.plus - b.eq_plus) - P.omega_minus * (-b.minus - -b.eq_minus);
    out[5] = fl[5] - P.omega_plus * (c.plus - c.eq_plus) - P.omega_minus * (c.minus - c.eq_minus);
    out[7] = fl[7] - P.omega_plus * (c.plus - c.eq_plus) - P.omega_minus * (-c.minus - -c.eq_minus);
    out[6] = fl[6] - P.omega_plus * (d.plus - d.eq_plus) - P.omega_minus * (d.minus - d.eq_minus);
    out[8] = fl[8] - P.omega_plus * (d.plus - d.eq_plus) - P.omega_minus * (-d.minus - -d.eq_minus);
    f2[i] = out[0];
#pragma unroll
    for (int k = 1; k < 9; k++) {
        if (link & (1u << k))
            f2[OPP[k] * n + i] = -out[k] + 2 * W[OPP[k]] * P.conc_wall;   // cylinder.cpp:217
        else
            f2[k * n + i + CY[k] * P.nx + CX[k]] = out[k];
    }
}

// ======================= binary liquid around a curved solid (CurvedSolid/binary_shape.cpp) =======================
// Node index i = iY*nx+iX, periodic in both directions (binary_shape.cpp:287-288,628-629).  Geometry mask as for
// the TRT model (1 fluid, 0 bounce-back node, -1 solid; :449-480).  One 16-bit link word per node, built once on
// the host: fluid nodes carry bit 0 and bit k (1..8) = "the neighbour in direction k is a bounce-back node";
// bounce-back nodes carry their wall normal main_dir (:556-588) in bits 9..12.  Half-way bounce-back
// (:358-372) + pull streaming (:617-637) become a push in which a population sent towards a bounce-back node
// comes straight back into the node's own opposite slot.
struct BsParams {
    int nx, ny;
    double aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq_minus_gas, omega_phi, force_x, force_y;
};
#define LBM_BS_WXX {0.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0}
#define LBM_BS_WYY {0.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0}
#define LBM_BS_WXY {0.0, 0.0, 0.0, 0.0, 0.0, 1.0 / 4.0, -1.0 / 4.0, 1.0 / 4.0, -1.0 / 4.0}
#define LBM_BS_GX {0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 0.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0, 1.0 / 12.0}
#define LBM_BS_GY {0.0, 0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 1.0 / 12.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0}
#define LBM_BS_LAP {-20.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0}

// the wall normal of a bounce-back node: the lattice direction equal to the sum of the axis directions that
// lead to fluid, or, when there is none, of the diagonal ones (:556-588); 0 = no such direction
LBM_HD int bs_main_dir(const int *mask, int nx, long long c)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    int sx = 0, sy = 0;
    bool any = false;
    for (int k = 1; k < 5; k++)
        if (mask[c + CY[k] * nx + CX[k]] == 1) { any = true; sx += CX[k]; sy += CY[k]; }
    if (!any)
        for (int k = 5; k < 9; k++)
            if (mask[c + CY[k] * nx + CX[k]] == 1) { sx += CX[k]; sy += CY[k]; }
    for (int k = 1; k < 9; k++)
        if (sx == CX[k] && sy == CY[k]) return k;
    return 0;
}

// link word of node i (the host has already checked the geometry rules)
LBM_HD unsigned short bs_links_thread(const BsParams &P, const int *mask, long long i)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    if (mask[i] == 0) return (unsigned short)(bs_main_dir(mask, P.nx, i) << 9);
    if (mask[i] != 1) return 0;
    const int ix = (int)(i % P.nx), iy = (int)(i / P.nx);
    unsigned short b = 1;
    for (int k = 1; k < 9; k++) {
        const long long t = (long long)((iy + CY[k] + P.ny) % P.ny) * P.nx + (ix + CX[k] + P.nx) % P.nx;
        if (mask[t] == 0) b |= (unsigned short)(1u << k);
    }
    return b;
}

// :222-258: equilibrium of every node from rho, phi, ux, uy, into both buffers
LBM_HD void bs_init_thread(const BsParams &P, const double *rho_i, const double *phi_i, const double *ux_i, const double *uy_i,
                           double *fa, double *ga, double *fb, double *gb, long long i)
{
    const long long n = (long long)P.nx * P.ny;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    const double dense_temp = rho_i[i], phase_temp = phi_i[i], ux_temp = ux_i[i], uy_temp = uy_i[i];
    for (int k = 0; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        const double feq = W[k] * (dense_temp + 3.0 * dense_temp * (cx * ux_temp + cy * uy_temp) +
                                   4.5 * dense_temp * ((cx * cx - 1.0 / 3.0) * ux_temp * ux_temp + (cy * cy - 1.0 / 3.0) * uy_temp * uy_temp +
                                                       2.0 * ux_temp * uy_temp * cx * cy));
        const double geq = W[k] * (phase_temp + 3.0 * phase_temp * (cx * ux_temp + cy * uy_temp) +
                                   4.5 * phase_temp * ((cx * cx - 1.0 / 3.0) * ux_temp * ux_temp + (cy * cy - 1.0 / 3.0) * uy_temp * uy_temp +
                                                       2.0 * ux_temp * uy_temp * cx * cy));
        fa[k * n + i] = feq; fb[k * n + i] = feq;
        ga[k * n + i] = geq; gb[k * n + i] = geq;
    }
}

// update_phase (:413-435): phi = sum of g on fluid nodes; a bounce-back node takes the phi of the fluid node
// its wall normal points at, minus wall_gradient -- recomputed here from that node's g (same sum, same order)
// so that the kernel needs no second pass
LBM_HD void bs_phase_thread(const BsParams &P, const double *g, const unsigned short *links, double *phi, long long i)
{
    const unsigned link = links[i];
    const int dir = (int)(link >> 9);
    if (!(link & 1u) && dir == 0) return;
    const long long n = (long long)P.nx * P.ny;
    // c_x+1 and c_y+1 of the nine directions as 2-bit fields (a table indexed by `dir` would live in local memory)
    const int cxd = (int)((0x20919u >> (2 * dir)) & 3u) - 1, cyd = (int)((0x2865u >> (2 * dir)) & 3u) - 1;
    const long long src = i + cyd * P.nx + cxd;
    double phase_temp = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) phase_temp += g[k * n + src];
    phi[i] = dir ? phase_temp - P.wall_gradient : phase_temp;
}

// collide_bulk (:260-356) of one fluid node whose populations fl, gl and the nine phase values of its stencil (ph9[0]
// the node's own) are in registers; fl and gl hold the post-collision values on return
LBM_HD void bs_collide_node(const BsParams &P, double (&fl)[9], double (&gl)[9], const double (&ph9)[9], double &dense_temp,
                            double &ux_temp, double &uy_temp)
{
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const double W[9] = LBM2_W;
    const double WXX[9] = LBM_BS_WXX, WYY[9] = LBM_BS_WYY, WXY[9] = LBM_BS_WXY;
    const double GX[9] = LBM_BS_GX, GY[9] = LBM_BS_GY, LAP[9] = LBM_BS_LAP;
    dense_temp = 0.0; ux_temp = 0.0; uy_temp = 0.0;
    const double phase_temp = ph9[0];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        dense_temp += fl[k];
        ux_temp += fl[k] * CX[k];
        uy_temp += fl[k] * CY[k];
    }
    ux_temp = (ux_temp + 0.5 * P.force_x) / dense_temp;
    uy_temp = (uy_temp + 0.5 * P.force_y) / dense_temp;

    double laplace_temp = 0.0, gradx_temp = 0.0, grady_temp = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const double ph = ph9[k];
        laplace_temp += LAP[k] * ph;
        gradx_temp += GX[k] * ph;
        grady_temp += GY[k] * ph;
    }
    double sum = 0.0, sum_phase = 0.0;
    const double phase_square = phase_temp * phase_temp;
    const double pressure_bulk = dense_temp / 3.0 + P.aconst * (-0.5 * phase_square + 3.0 / 4.0 * phase_square * phase_square) -
                                 P.kconst * phase_temp * laplace_temp;
    const double chemical_pot = P.gammaconst * (P.aconst * (-phase_temp + phase_temp * phase_temp * phase_temp) - P.kconst * laplace_temp);
    double feqeq[9], geqeq[9];
#pragma unroll
    for (int k = 1; k < 9; k++) {
        const int cx = CX[k], cy = CY[k];
        feqeq[k] = W[k] * (3.0 * pressure_bulk + 3.0 * dense_temp * (cx * ux_temp + cy * uy_temp) +
                           4.5 * dense_temp * ((cx * cx - 1.0 / 3.0) * ux_temp * ux_temp + (cy * cy - 1.0 / 3.0) * uy_temp * uy_temp +
                                               2.0 * ux_temp * uy_temp * cx * cy)) +
                   P.kconst * (WXX[k] * gradx_temp * gradx_temp + WYY[k] * grady_temp * grady_temp + WXY[k] * gradx_temp * grady_temp);
        geqeq[k] = W[k] * (3.0 * chemical_pot + 3.0 * phase_temp * (cx * ux_temp + cy * uy_temp) +
                           4.5 * phase_temp * ((cx * cx - 1.0 / 3.0) * ux_temp * ux_temp + (cy * cy - 1.0 / 3.0) * uy_temp * uy_temp +
                                               2.0 * ux_temp * uy_temp * cx * cy));
        sum += feqeq[k];
        sum_phase += geqeq[k];
    }
    feqeq[0] = dense_temp - sum;
    geqeq[0] = phase_temp - sum_phase;
    const double tau_rho = P.tau_gas + (phase_temp + 1.0) / 2.0 * P.tau_liq_minus_gas;
    const double omega_rho = 1.0 / tau_rho;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        fl[k] = fl[k] * (1.0 - omega_rho) + omega_rho * feqeq[k];
        gl[k] = gl[k] * (1.0 - P.omega_phi) + P.omega_phi * geqeq[k];
    }
}
// update_bounce_back + stream folded into the push: a population sent towards a bounce-back node comes straight back
// into the node's own opposite slot (:358-372, :617-637); nb[k] = index of the periodic neighbour in direction k
LBM_HD void bs_push(const BsParams &P, unsigned link, const double (&fl)[9], const double (&gl)[9], const int (&nb)[9], double *f2,
                    double *g2, long long i)
{
    const long long n = (long long)P.nx * P.ny;
    const int OPP[9] = LBM2_OPP;
#pragma unroll
    for (int k = 0; k < 9; k++) {
        if (link & (1u << k)) {      // bit 0 (the fluid flag) sends the rest population to slot OPP[0] = 0 of the node itself
            f2[OPP[k] * n + i] = fl[k];
            g2[OPP[k] * n + i] = gl[k];
        } else {
            f2[k * n + nb[k]] = fl[k];
            g2[k * n + nb[k]] = gl[k];
        }
    }
}

// the two-kernel form: the phase field comes from bs_phase_thread's array
LBM_HD void bs_collide_thread(const BsParams &P, const double *f, const double *g, const double *phi, const unsigned short *links,
                              double *f2, double *g2, double *rho_o, double *ux_o, double *uy_o, long long i,
                              bool write_fields = true)
{
    const unsigned link = links[i];
    if (!(link & 1u)) return;
    const long long n = (long long)P.nx * P.ny;
    const int CX[9] = LBM2_CX, CY[9] = LBM2_CY;
    const int ix = (int)((unsigned)i % (unsigned)P.nx), iy = (int)((unsigned)i / (unsigned)P.nx);   // nx*ny < 2^31 (checked at create)
    // periodic neighbours: (iX+cx+NX)%NX, (iY+cy+NY)%NY
    const int xs[3] = {ix == 0 ? P.nx - 1 : ix - 1, ix, ix == P.nx - 1 ? 0 : ix + 1};
    const int ys[3] = {iy == 0 ? P.ny - 1 : iy - 1, iy, iy == P.ny - 1 ? 0 : iy + 1};
    int nb[9];      // nx*ny < 2^31 (checked at create)
    double fl[9], gl[9], ph9[9];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        nb[k] = ys[CY[k] + 1] * P.nx + xs[CX[k] + 1];
        fl[k] = f[k * n + i];
        gl[k] = g[k * n + i];
        ph9[k] = phi[nb[k]];
    }
    double dense_temp, ux_temp, uy_temp;
    bs_collide_node(P, fl, gl, ph9, dense_temp, ux_temp, uy_temp);
    if (write_fields) { rho_o[i] = dense_temp; ux_o[i] = ux_temp; uy_o[i] = uy_temp; }
    bs_push(P, link, fl, gl, nb, f2, g2, i);
}

}  // namespace lbm2
#endif
