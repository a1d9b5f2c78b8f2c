// Drop-in for MRT/another_mrt.cpp: omega = 0.9, omega_bulk = 0.1 (:52-53), initial wave :89-96, 501 steps,
// output precision 10 (:17-46).
#include "mrt_common.h"

int main()
{
    double omega = 0.9, omega_bulk = 0.1;
    const int N = 500;
    double ampl = 0.001;
    std::vector<double> rho(NX * NY), ux(NX * NY), uy(NX * NY);
    for (int iX = 0; iX < NX; iX++)
        for (int iY = 0; iY < NY; iY++) {
            double dense = 0.0;
            rho[iX * NY + iY] = 1.0 + dense;
            ux[iX * NY + iY] = ampl * cos(double(iX) / double(NX) * 2 * pi);
            uy[iX * NY + iY] = 0.0;
        }
    return run_mrt(LBM_MRT_HERMITE_D2Q9, omega, omega_bulk, N, 10, rho, ux, uy);
}
