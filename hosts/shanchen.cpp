// Drop-in for ShanChen/shanchen.cpp: same constants (:11-27), droplet (:87-95), output every 1000 steps
// into density/velx/vely%05d.dat with the reference's writer (:36-47), same stdout lines.  The time loop
// body (:120-188) runs on the GPU.  Built against include/lbm_exp.h: bit-identical to the reference
// compiled with `#define exp lbm_exp`.
#include <ctime>
#include "host_common.h"

const int nx = 128, ny = 128, nsteps = 30000, noutput = 1000;
double rhol = 1.95, rhog = 0.15;
int radius = 20;

static void writeFile(const std::string &name, const double *f)
{
    std::ofstream fout(name.c_str());
    for (int counterX = 0; counterX < nx; counterX++) {
        for (int counterY = 0; counterY < ny; counterY++) fout << f[counterX * ny + counterY] << " ";
        fout << "\n";
    }
    fout << std::endl;
}

int main(int argc, char **argv)
{
    lbm_config cfg;
    lbm_config_default(&cfg, LBM_SHANCHEN_D2Q9);
    cfg.device = device_from_env();
    int steps = nsteps;
    if (argc == 2) steps = std::atoi(argv[1]);   // optional: last step index (the reference's own argv handling is broken, :76-80)
    lbm_ctx *ctx = 0;
    LBM_CHECK(0, lbm_create(&cfg, &ctx));
    std::vector<double> rho(nx * ny), u1(nx * ny), u2(nx * ny);
    for (int i = 0; i < nx * ny; i++)
        rho[i] = ((i / nx - ny / 2.0) * (i / nx - ny / 2.0) + (i % nx - nx / 2.0) * (i % nx - nx / 2.0) <= radius * radius) ? rhol : rhog;
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_RHO, rho.data(), rho.size()));
    LBM_CHECK(ctx, lbm_init(ctx));
    time_t start = time(NULL);
    for (int timecounter = 0; timecounter <= steps;) {
        // run up to and including the next output step
        int next_out = (timecounter % noutput == 0) ? timecounter : (timecounter / noutput + 1) * noutput;
        if (next_out > steps) next_out = steps;
        LBM_CHECK(ctx, lbm_step(ctx, next_out - timecounter + 1));
        timecounter = next_out;
        if (timecounter % noutput == 0) {
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_RHO, rho.data(), rho.size()));
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UX, u1.data(), u1.size()));
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UY, u2.data(), u2.size()));
            writeFile(padded("density", 5, timecounter) + ".dat", rho.data());
            writeFile(padded("velx", 5, timecounter) + ".dat", u1.data());
            writeFile(padded("vely", 5, timecounter) + ".dat", u2.data());
            std::cout << "Iteration is " << timecounter << "\n";
        }
        timecounter++;
    }
    time_t finish = time(NULL);
    std::cout << "Overall time is " << finish - start << " sec" << "\n";
    lbm_destroy(ctx);
    return 0;
}
