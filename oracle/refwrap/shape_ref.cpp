// Test infrastructure: the UNMODIFIED CurvedSolid/binary_shape.cpp compiled in place, its main renamed.
// As with cylinder.cpp the program's main is only the loop "update_phase(); collide_bulk();
// update_bounce_back(); stream();" plus writers (binary_shape.cpp:664-712) around external-linkage globals
// (:17-66), so the wrapper calls the reference's own functions in that order and copies the arrays out
// before finish_simulation() frees them.  NX = NY = 301 and radius = 50 are compile-time constants (:17-19,29).
#include <iostream>
#include <fstream>
#include <string>
#include <sstream>
#include <stdio.h>
#include <stdlib.h>
#include <cmath>
#include <vector>
#include <cstring>
#include <unistd.h>

#define main ref_main
#include "CurvedSolid/binary_shape.cpp"
#undef main

extern "C" {
void refshape_shape(int *nx, int *ny) { *nx = NX; *ny = NY; }
// steps = loop iterations (the program's own run is N+1 = 20001).  workdir: initialize_geometry() writes
// phase.dat and geometry.dat into the current directory (:596-597).  p (may be NULL = compiled-in constants):
// aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi; radius_droplet_in <= 0 keeps 10.
// phi_in/rho_in/ux_in/uy_in (may be NULL) overwrite the fields initialize_geometry() filled, before init().
int refshape_run(int steps, const char *workdir, const double *p, int radius_droplet_in, const double *phi_in, const double *rho_in,
                 const double *ux_in, const double *uy_in, double *phi_out, double *rho_out, double *ux_out, double *uy_out,
                 double *f_out, double *g_out, int *geom_out)
{
    if (workdir && chdir(workdir) != 0) return 2;
    if (p) {
        aconst = p[0]; kconst = p[1]; gammaconst = p[2]; wall_gradient = p[3]; tau_gas = p[4]; tau_liq = p[5]; omega_phi = p[6];
    }
    if (radius_droplet_in > 0) radius_droplet = radius_droplet_in;
    bb_nodes.clear();
    main_dir.clear();
    std::cout.setstate(std::ios_base::failbit);      // "BB size=", "Main size=" lines
    initialize_geometry();
    std::cout.clear();
    if (bb_nodes.size() != main_dir.size()) return 3;
    if (phi_in) memcpy(phi, phi_in, NUM * sizeof(double));
    if (rho_in) memcpy(rho, rho_in, NUM * sizeof(double));
    if (ux_in) memcpy(ux, ux_in, NUM * sizeof(double));
    if (uy_in) memcpy(uy, uy_in, NUM * sizeof(double));
    init();
    for (int counter = 0; counter < steps; counter++) {
        update_phase();
        collide_bulk();
        update_bounce_back();
        stream();
    }
    if (phi_out) memcpy(phi_out, phi, NUM * sizeof(double));
    if (rho_out) memcpy(rho_out, rho, NUM * sizeof(double));
    if (ux_out) memcpy(ux_out, ux, NUM * sizeof(double));
    if (uy_out) memcpy(uy_out, uy, NUM * sizeof(double));
    if (f_out) memcpy(f_out, f, (size_t)NUM * NPOP * sizeof(double));
    if (g_out) memcpy(g_out, g, (size_t)NUM * NPOP * sizeof(double));
    if (geom_out) memcpy(geom_out, geometry, NUM * sizeof(int));
    finish_simulation();
    return 0;
}
}
