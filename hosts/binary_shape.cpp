// hosts/binary_shape.cpp -- CurvedSolid/binary_shape.cpp after the cut-over: the program's own constants,
// argv, geometry and droplet construction, stdout lines and writers; the time loop body (update_phase,
// collide_bulk, update_bounce_back, stream: binary_shape.cpp:666-669) is lbm_step.  argv as the reference:
// "radius_droplet wall_gradient" (both or none); an optional third argument shortens the run to N steps.
#include <cmath>
#include "host_common.h"

const int NY = 301, NX = 301, NUM = NX * NY;
static int N = 20000, NOUTPUT = 200, NSIGNAL = 100;
const int radius = 50;
static int radius_droplet = 10;
static double wall_gradient = -0.05;
const int NUM_DROPLETS = 3;
const int angle = 360 / NUM_DROPLETS;
const double pi = 4.0 * std::atan(1.0);
const double phi_wall = 1.5, rho_wall = 0.5;
static std::vector<double> phi, rho, ux, uy;
static std::vector<int> geometry;
static const int cx[] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
static const int cy[] = {0, 0, 1, 0, -1, 1, 1, -1, -1};

template <class T>
static void writefield(std::string const &fname, const std::vector<T> &a)
{
    std::string filename = fname + ".dat";
    std::ofstream fout(filename.c_str());
    fout.precision(10);
    for (int iY = 0; iY < NY; iY++) {
        for (int iX = 0; iX < NX; ++iX) fout << a[iY * NX + iX] << " ";
        fout << "\n";
    }
}

static void writevtk(std::string const &fname)
{
    std::string filename = fname + ".vtk";
    std::ofstream fout(filename.c_str());
    fout << "# vtk DataFile Version 3.0\n";
    fout << "Binary liquid phase field vtk representation\n";
    fout << "ASCII\n\n";
    fout << "DATASET STRUCTURED_GRID\n";
    fout << "DIMENSIONS " << NX << " " << NY << " " << 1 << "\n";
    fout << "POINTS " << NX * NY * 1 << " double\n";
    for (int counter = 0; counter < NUM; counter++) fout << counter % NX << " " << counter / NX << " " << 0 << "\n";
    fout << "\n";
    fout << "POINT_DATA " << NX * NY * 1 << "\n";
    fout << "SCALARS phase double\n";
    fout << "LOOKUP_TABLE phase_table\n";
    for (int counter = 0; counter < NUM; counter++) fout << phi[counter] << "\n";
    fout << "SCALARS density double\n";
    fout << "LOOKUP_TABLE density_table\n";
    for (int counter = 0; counter < NUM; counter++) fout << rho[counter] << "\n";
    fout << "SCALARS geometry double\n";
    fout << "LOOKUP_TABLE geometry_table\n";
    for (int counter = 0; counter < NUM; counter++) fout << geometry[counter] << "\n";
    fout << "VECTORS velocity double\n";
    for (int counter = 0; counter < NUM; counter++) fout << ux[counter] << " " << uy[counter] << " 0\n";
}

// solid disc, its skin of bounce-back nodes, densities, the droplets on the disc (binary_shape.cpp:439-520)
static int initialize_geometry()
{
    geometry.assign(NUM, 1); rho.assign(NUM, 1.0); ux.assign(NUM, 0.0); uy.assign(NUM, 0.0); phi.assign(NUM, -1.0);
    const int mx = (NX - 1) / 2, my = (NY - 1) / 2;
    for (int c = 0; c < NUM; c++)
        if ((c % NX - mx) * (c % NX - mx) + (c / NX - my) * (c / NX - my) < radius * radius) geometry[c] = -1;
    int bb = 0;
    for (int c = 0; c < NUM; c++) {
        if (geometry[c] != -1) continue;
        for (int k = 1; k < 9; k++)
            if (geometry[((c / NX + cy[k] + NY) % NY) * NX + (c % NX + cx[k] + NX) % NX] == 1) { geometry[c] = 0; break; }
    }
    for (int c = 0; c < NUM; c++) {
        if (geometry[c] != 1) rho[c] = rho_wall;
        if (geometry[c] == 0) bb++;
    }
    for (int c = 0; c < NUM; c++) {
        const int iX = c % NX, iY = c / NX;
        bool flag_droplet = false;
        for (int d = 0; d < NUM_DROPLETS; d++) {
            double center_x = (NX - 1) / 2 - std::sin(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius);
            double center_y = (NY - 1) / 2 + std::cos(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius);
            if (((iX - center_x) * (iX - center_x) + (iY - center_y) * (iY - center_y) <= radius_droplet * radius_droplet) && geometry[c] == 1)
                flag_droplet = true;
        }
        phi[c] = flag_droplet ? 1.0 : (geometry[c] == 1 ? -1.0 : phi_wall);
    }
    return bb;
}

int main(int argc, char *argv[])
{
    if (argc >= 3) {
        radius_droplet = atoi(argv[1]);
        wall_gradient = atof(argv[2]);
    }
    if (argc >= 4) N = atoi(argv[3]);
    std::cout << "Radius of the droplet is " << radius_droplet << "\n";
    std::cout << "Wall gradient is " << wall_gradient << "\n";
    const int bb = initialize_geometry();
    std::cout << "BB size=" << bb << "\n";
    std::cout << "Main size=" << bb << "\n";      // equal by construction: lbm_set_geometry rejects a node without a wall normal
    writefield("phase", phi);
    writefield("geometry", geometry);

    lbm_config cfg;
    lbm_config_default(&cfg, LBM_BINARY_SHAPE_D2Q9);
    cfg.nx = NX; cfg.ny = NY; cfg.wall_gradient = wall_gradient; cfg.device = device_from_env();
    lbm_ctx *ctx = 0;
    LBM_CHECK(0, lbm_create(&cfg, &ctx));
    LBM_CHECK(ctx, lbm_set_geometry(ctx, geometry.data(), geometry.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_RHO, rho.data(), rho.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_PHASE, phi.data(), phi.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UX, ux.data(), ux.size()));
    LBM_CHECK(ctx, lbm_set_field(ctx, LBM_UY, uy.data(), uy.size()));
    LBM_CHECK(ctx, lbm_init(ctx));
    for (int counter = 0; counter <= N; counter += NSIGNAL) {
        LBM_CHECK(ctx, lbm_step(ctx, counter == 0 ? 1 : NSIGNAL));
        LBM_CHECK(ctx, lbm_get_field(ctx, LBM_PHASE, phi.data(), phi.size()));
        std::cout << "Counter=" << counter << "\n";
        double mass = 0.0;                          // calculate_mass(): same order, so the same digits
        for (int c = 0; c < NUM; c++)
            if (geometry[c] == 1) mass += phi[c];
        std::cout << "Mass is " << mass << "\n";
        if (counter % NOUTPUT == 0) {
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_RHO, rho.data(), rho.size()));
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UX, ux.data(), ux.size()));
            LBM_CHECK(ctx, lbm_get_field(ctx, LBM_UY, uy.data(), uy.size()));
            // the reference's three .dat names are empty strings (their stems are commented out, :688-690),
            // so all three writers hit the one file ".dat" and the last one, velocity y, is what remains
            writefield("", rho);
            writefield("", ux);
            writefield("", uy);
            writevtk(padded("vtk", 7, counter));
        }
    }
    lbm_destroy(ctx);
    return 0;
}
