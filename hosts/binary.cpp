// Drop-in for FreeEnergy/binary.cpp: constants :12-41, box initial condition :126-132, NaN probe :359-363,
// phase/velocityx/velocityy%06d.dat every NOUTPUT steps with the reference writers (:76-120: precision 10,
// rows iY = NY-1 .. 0).  init() and the loop body (:336-350) run on the GPU.
#include <cmath>
#include "host_common.h"

const int NY = 128, NX = 128;
int N = 20000, NOUTPUT = 100, NSIGNAL = 100;

static void writefield(const std::string &fName, const double *a)
{
    std::string fullName = fName + ".dat";
    std::ofstream fout(fullName.c_str());
    fout.precision(10);
    for (int iY = NY - 1; iY >= 0; --iY) {
        for (int iX = 0; iX < NX; ++iX) fout << a[iX * NY + iY] << " ";
        fout << "\n";
    }
}

int main(int argc, char *argv[])
{
    if (argc > 1) N = std::atoi(argv[1]);   // optional: last step index
    lbm_config cfg;
    lbm_config_default(&cfg, LBM_BINARY_D2Q9);
    cfg.device = device_from_env();
    lbm_ctx *ctx = 0;
    LBM_CHECK(0, lbm_create(&cfg, &ctx));
    std::vector<double> phase(NX * NY), ux(NX * NY), uy(NX * NY);
    for (int iX = 0; iX < NX; iX++)
        for (int iY = 0; iY < NY; iY++)
            phase[iX * NY + iY] = ((std::abs(iX - NX / 2) <= 20) && (iY < 40) && (iY >= 1)) ? 1.0 : -1.0;
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_PHASE, phase.data(), phase.size()));
    LBM_CHECK(ctx, lbm_init(ctx));
    for (int counter = 0; counter <= N; counter++) {
        // steps up to the next step that prints or writes
        int next = counter;
        while (next < N && next % NSIGNAL != 0 && next % NOUTPUT != 0) next++;
        LBM_CHECK(ctx, lbm_step(ctx, next - counter + 1));
        counter = next;
        if (counter % NSIGNAL == 0) std::cout << "Time is " << counter << "\n";
        if (counter % NOUTPUT == 0) {
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_PHASE, phase.data(), phase.size()));
            if (std::isnan(phase[(NX / 2) * NY + NY / 2])) {
                std::cout << "There were some NaN values.\n";
                return 0;
            }
            std::cout << counter << "\n";
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UX, ux.data(), ux.size()));
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UY, uy.data(), uy.size()));
            writefield(padded("phase", 6, counter), phase.data());
            writefield(padded("velocityx", 6, counter), ux.data());
            writefield(padded("velocityy", 6, counter), uy.data());
        }
    }
    lbm_destroy(ctx);
    return 0;
}
