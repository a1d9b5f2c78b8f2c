// Drop-in for MRT/mrt.cpp: argv omega*10, omega_bulk*10 (:53-54, both required), shear-wave initial fields
// (:97-106), proba/phase%06d.dat in ./tmp every step with precision 15 (:17-46,254-270).
#include "mrt_common.h"

int main(int argc, char *argv[])
{
    if (argc < 3) { std::fprintf(stderr, "usage: %s omega*10 omega_bulk*10\n", argv[0]); return 2; }
    double omega = double(atoi(argv[1])) / 10.0;
    double omega_bulk = double(atoi(argv[2])) / 10.0;
    std::cout << "Omega=" << omega << "\n";
    std::cout << "Omega_bulk=" << omega_bulk << "\n";
    const int N = 800;
    double ampl = 0.001, force_ampl = 0.002;
    std::vector<double> rho(NX * NY), ux(NX * NY), uy(NX * NY);
    for (int iX = 0; iX < NX; iX++)
        for (int iY = 0; iY < NY; iY++) {
            double forcex = force_ampl * cos(double(iX) / double(NX) * 2.0 * pi);
            rho[iX * NY + iY] = 1.0;
            ux[iX * NY + iY] = ampl * cos(double(iX) / double(NX) * 2.0 * pi) - forcex / 2.0;
            uy[iX * NY + iY] = 0.0;
        }
    return run_mrt(LBM_MRT_GS_D2Q9, omega, omega_bulk, N, 15, rho, ux, uy);
}
