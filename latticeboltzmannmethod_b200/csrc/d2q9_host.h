// d2q9_host.h -- host-side tables of the MRT programs, built with the reference's
// own expressions so that libm's cos()/sqrt() are evaluated on the host exactly
// as the reference evaluates them.  Plain C++ (shared with the CPU unit test).
#ifndef LBM_D2Q9_HOST_H
#define LBM_D2Q9_HOST_H
#include <cmath>
#include "d2q9_threads.h"

namespace lbm2 {

static const double kPiRef = 3.141592653589793238462643383279502884197;  // mrt.cpp:12

// M, invM / omega-scaled M^T, relaxation rates and the per-column force of the
// time loop.  Reference: mrt.cpp:57-92,165 and another_mrt.cpp:52-84.
inline void mrt_build_tables(MrtParams *P, int variant, int nx, int ny, double omega, double omega_bulk,
                             double force_ampl, double *forcex_of_ix)
{
    const int cxs[9] = LBM2_CX, cys[9] = LBM2_CY;
    P->nx = nx; P->ny = ny; P->variant = variant;
    P->s3 = std::sqrt(3.0);
    P->c32 = 3.0 / std::sqrt(2.0);
    P->c3s15 = 3.0 * std::sqrt(1.5);
    if (variant == 0) {
        const double om[9] = {1.0, omega_bulk, omega, 1.0, omega, 1.0, omega, omega, omega};
        const double wm[9] = {1.0 / 9.0, 1.0 / 36.0, 1.0 / 36.0, 1.0 / 6.0, 1.0 / 12.0, 1.0 / 6.0, 1.0 / 12.0, 1.0 / 4.0, 1.0 / 4.0};
        for (int m = 0; m < 9; m++) P->omegamat[m] = om[m];
        for (int i = 0; i < 9; i++) {
            const int cx = cxs[i], cy = cys[i];
            P->M[0][i] = 1.0;
            P->M[1][i] = -4.0 + 3.0 * (cx * cx + cy * cy);
            P->M[2][i] = 4.0 - 21.0 / 2.0 * (cx * cx + cy * cy) + 9.0 / 2.0 * (cx * cx + cy * cy) * (cx * cx + cy * cy);
            P->M[3][i] = cx;
            P->M[4][i] = (-5.0 + 3.0 * (cx * cx + cy * cy)) * cx;
            P->M[5][i] = cy;
            P->M[6][i] = (-5.0 + 3.0 * (cx * cx + cy * cy)) * cy;
            P->M[7][i] = cx * cx - cy * cy;
            P->M[8][i] = cx * cy;
            P->B[i][0] = wm[0] * 1.0;
            P->B[i][1] = wm[1] * (-4.0 + 3.0 * (cx * cx + cy * cy));
            P->B[i][2] = wm[2] * (4.0 - 21.0 / 2.0 * (cx * cx + cy * cy) + 9.0 / 2.0 * (cx * cx + cy * cy) * (cx * cx + cy * cy));
            P->B[i][3] = wm[3] * cx;
            P->B[i][4] = wm[4] * (-5.0 + 3.0 * (cx * cx + cy * cy)) * cx;
            P->B[i][5] = wm[5] * cy;
            P->B[i][6] = wm[6] * (-5.0 + 3.0 * (cx * cx + cy * cy)) * cy;
            P->B[i][7] = wm[7] * (cx * cx - cy * cy);
            P->B[i][8] = wm[8] * cx * cy;
        }
        for (int ix = 0; ix < nx; ix++) forcex_of_ix[ix] = force_ampl * std::cos(2.0 * kPiRef * double(ix) / double(nx));
    } else {
        const double om[9] = {1.0, 1.0, 1.0, omega, omega_bulk, omega_bulk, omega_bulk, omega_bulk, omega_bulk};
        const double gg[9] = {1.0, -2.0, -2.0, -2.0, -2.0, 4.0, 4.0, 4.0, 4.0};
        for (int m = 0; m < 9; m++) P->omegamat[m] = om[m];
        for (int i = 0; i < 9; i++) {
            const int cx = cxs[i], cy = cys[i];
            P->M[0][i] = 1.0;
            P->M[1][i] = cx * std::sqrt(3.0);
            P->M[2][i] = cy * std::sqrt(3.0);
            P->M[3][i] = (cx * cx - 1.0 / 3.0) * 3.0 / std::sqrt(2.0);
            P->M[4][i] = cx * cy * 3.0;
            P->M[5][i] = (cy * cy - 1.0 / 3.0) * 3.0 / std::sqrt(2.0);
            P->M[6][i] = gg[i] / 2.0;
            P->M[7][i] = gg[i] * cx * std::sqrt(1.5) / 2.0;
            P->M[8][i] = gg[i] * cy * std::sqrt(1.5) / 2.0;
        }
        for (int k = 0; k < 9; k++)
            for (int m = 0; m < 9; m++) P->B[k][m] = om[m] * P->M[m][k];
        for (int ix = 0; ix < nx; ix++) forcex_of_ix[ix] = 0.0;
    }
}

// The shear-wave initial fields of mrt.cpp:100-105 / another_mrt.cpp:92-96 (note
// the cos argument is associated differently from the time loop's, mrt.cpp:165).
inline void mrt_reference_wave(int variant, int nx, int ny, double ampl, double force_ampl, double *rho, double *ux, double *uy)
{
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            const long long i = (long long)ix * ny + iy;
            if (variant == 0) {
                const double forcex = force_ampl * std::cos(double(ix) / double(nx) * 2.0 * kPiRef);
                rho[i] = 1.0;
                ux[i] = ampl * std::cos(double(ix) / double(nx) * 2.0 * kPiRef) - forcex / 2.0;
            } else {
                const double dense = 0.0;
                rho[i] = 1.0 + dense;
                ux[i] = ampl * std::cos(double(ix) / double(nx) * 2 * kPiRef);
            }
            uy[i] = 0.0;
        }
}

}  // namespace lbm2
#endif
