// Shared by hosts/mrt.cpp and hosts/another_mrt.cpp (MRT/mrt.cpp, MRT/another_mrt.cpp).
#include <cmath>
#include "host_common.h"

const int NY = 3, NX = 40;
const double pi = 3.141592653589793238462643383279502884197;

static void writefield(const std::string &fName, const double *a, int precision)
{
    std::string fullName = "./tmp/" + fName + ".dat";   // silently nothing if ./tmp is missing, as in the reference
    std::ofstream fout(fullName.c_str());
    fout.precision(precision);
    for (int iY = NY - 1; iY >= 0; --iY) {
        for (int iX = 0; iX < NX; ++iX) fout << a[iX * NY + iY] << " ";
        fout << "\n";
    }
}

static int run_mrt(int model, double omega, double omega_bulk, int N, int precision, const std::vector<double> &rho0,
                   const std::vector<double> &ux0, const std::vector<double> &uy0)
{
    lbm_config cfg;
    lbm_config_default(&cfg, model);
    cfg.omega = omega; cfg.omega_bulk = omega_bulk; cfg.device = device_from_env();
    lbm_ctx *ctx = 0;
    LBM_CHECK(0, lbm_create(&cfg, &ctx));
    std::vector<double> rho(rho0), ux(ux0), uy(uy0);
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_RHO, rho.data(), rho.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UX, ux.data(), ux.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UY, uy.data(), uy.size()));
    LBM_CHECK(ctx, lbm_init(ctx));
    for (int counter = 0; counter <= N; counter++) {
        LBM_CHECK(ctx, lbm_step(ctx, 1));
        std::cout << counter << "\n";
        LBM_CHECK(ctx, lbm_get_field(ctx, LBM_RHO, rho.data(), rho.size()));
        LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UX, ux.data(), ux.size()));
        writefield(padded("proba", 6, counter), rho.data(), precision);
        writefield(padded("phase", 6, counter), ux.data(), precision);
    }
    lbm_destroy(ctx);
    return 0;
}
