// Test infrastructure: the UNMODIFIED DiffusionCylinder/cylinder.cpp compiled in place, its main renamed.
// The program keeps its state in external-linkage globals (cylinder.cpp:11-36) and its main is only the
// loop "collide(); update_bounce_back(); stream();" plus the text writer (cylinder.cpp:350-377), so the
// wrapper calls the reference's own functions in that order and copies the arrays out before
// finish_simulation() frees them.  initialize_geometry() fixes NX = NY = 129, radius = 40 (cylinder.cpp:226-234).
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
#include "DiffusionCylinder/cylinder.cpp"
#undef main

extern "C" {
// sizes the reference hard-codes, for the caller's buffers
void refcyl_shape(int *nx, int *ny, int *npop) { *nx = 129; *ny = 129; *npop = NPOP; }
// steps = the number of loop iterations (the reference's own run is N+1 = 64001); every output is optional.
// rho_out/geom_out: [iY*NX+iX]; f_out: [(iY*NX+iX)*9+k], the streamed populations after the last iteration
// The constants and the velocity field are the program's own globals (cylinder.cpp:27-28,33,38-43): a
// non-positive omega_in keeps the compiled-in constants; ux_in/uy_in/rho_in (may be NULL) overwrite the
// fields initialize_geometry() filled, before init() builds the populations from them.
// workdir: initialize_geometry() writes antibbgeometry_before.dat / antibbgeometry_after.dat into the current
// directory (cylinder.cpp:243,284).
int refcyl_run(int steps, const char *workdir, double omega_in, double ce_in, double conc_wall_in, const double *rho_in,
               const double *ux_in, const double *uy_in, double *rho_out, double *f_out, int *geom_out)
{
    if (workdir && chdir(workdir) != 0) return 2;
    if (omega_in > 0.0) {
        omega = omega_in; omega_plus = 2.0 - omega; omega_minus = omega;
        ce = ce_in; conc_wall = conc_wall_in;
    }
    bb_nodes.clear();
    initialize_geometry();
    if (rho_in) memcpy(rho, rho_in, NUM * sizeof(double));
    if (ux_in) memcpy(ux, ux_in, NUM * sizeof(double));
    if (uy_in) memcpy(uy, uy_in, NUM * sizeof(double));
    init();
    for (int counter = 0; counter < steps; counter++) {
        collide();
        update_bounce_back();
        stream();
    }
    if (rho_out) memcpy(rho_out, rho, NUM * sizeof(double));
    if (f_out) memcpy(f_out, f, (size_t)NUM * NPOP * sizeof(double));
    if (geom_out) memcpy(geom_out, geometry, NUM * sizeof(int));
    finish_simulation();
    return 0;
}
}
