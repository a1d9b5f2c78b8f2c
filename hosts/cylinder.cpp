// hosts/cylinder.cpp -- DiffusionCylinder/cylinder.cpp after the cut-over: the program's own constants,
// geometry construction, initial concentration, stdout lines and text writer; the time loop body
// (collide, update_bounce_back, stream: cylinder.cpp:355-358) is lbm_step.  Optional argv[1] = number of
// steps N (the program's constant is 64000), argv[2] = NOUTPUT (200).
#include "host_common.h"

static int NY, NX, NUM, radius;
static int N = 64000, NOUTPUT = 200;
static std::vector<double> rho, ux, uy;
static std::vector<int> geometry;
static const double conc_wall = 1.0;
static const double omega = 1.0 / (0.5 + 0.5 / 32.0);
static const double ce = 1.0 / 3.0;
static const int cx[] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
static const int cy[] = {0, 0, 1, 0, -1, 1, 1, -1, -1};

static void writedensity(std::string const &fname)
{
    std::string filename = "antibb/" + fname + ".dat";   // silently nothing if ./antibb is missing, as in the reference
    std::ofstream fout(filename.c_str());
    fout.precision(10);
    for (int iX = 0; iX < NX; ++iX) {
        for (int iY = 0; iY < NY; iY++) fout << rho[iY * NX + iX] << " ";
        fout << "\n";
    }
}

// the disc, then the two clean-up passes that leave a one-node ring of boundary nodes around the fluid
static void initialize_geometry()
{
    NY = 129; NX = 129; NUM = NX * NY; radius = 40;
    geometry.assign(NUM, 0); rho.assign(NUM, 0.0); ux.assign(NUM, 0.0); uy.assign(NUM, 0.0);
    const int mx = (NX - 1) / 2, my = (NY - 1) / 2;
    for (int iX = 0; iX < NX; iX++)
        for (int iY = 0; iY < NY; iY++)
            geometry[iX * NY + iY] = (iX - mx) * (iX - mx) + (iY - my) * (iY - my) > radius * radius ? -1 : 1;
    for (int pass = 0; pass < 2; pass++) {
        const int from = pass == 0 ? 1 : 0, look_for = pass == 0 ? -1 : 1, to = pass == 0 ? 0 : -1;
        for (int iX = 0; iX < NX; iX++)
            for (int iY = 0; iY < NY; iY++) {
                const int counter = iX * NY + iY;
                if (geometry[counter] != from) continue;
                bool found = false;
                for (int k = 0; k < 9; k++)
                    if (geometry[(iX + cx[k]) * NY + iY + cy[k]] == look_for) found = true;
                if (found == (pass == 0)) geometry[counter] = to;
            }
    }
    for (int counter = 0; counter < NUM; counter++)
        rho[counter] = geometry[counter] == 0 ? conc_wall : (geometry[counter] == -1 ? -1.0 : 0.0);
}

int main(int argc, char *argv[])
{
    if (argc > 1) N = std::atoi(argv[1]);
    if (argc > 2) NOUTPUT = std::atoi(argv[2]);
    initialize_geometry();
    lbm_config cfg;
    lbm_config_default(&cfg, LBM_TRT_DIFFUSION_D2Q9);
    cfg.nx = NX; cfg.ny = NY; cfg.omega = omega; cfg.ce = ce; cfg.conc_wall = conc_wall; cfg.device = device_from_env();
    lbm_ctx *ctx = 0;
    LBM_CHECK(0, lbm_create(&cfg, &ctx));
    LBM_CHECK(ctx, lbm_set_geometry(ctx, geometry.data(), geometry.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_RHO, rho.data(), rho.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UX, ux.data(), ux.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UY, uy.data(), uy.size()));
    LBM_CHECK(ctx, lbm_init(ctx));
    for (int counter = 0; counter <= N; counter += NOUTPUT) {
        LBM_CHECK(ctx, lbm_step(ctx, counter == 0 ? 1 : NOUTPUT));
        LBM_CHECK(ctx, lbm_get_field(ctx, LBM_RHO, rho.data(), rho.size()));
        std::cout << "Counter=" << counter << "\n";
        writedensity(padded("density", 7, counter));
    }
    lbm_destroy(ctx);
    return 0;
}
