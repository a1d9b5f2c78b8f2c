// Wrapper around the UNMODIFIED FreeEnergy/binary.cpp (test infrastructure).
// N, NOUTPUT, NSIGNAL are plain globals there (binary.cpp:17-19) and the state
// arrays have external linkage (:24-25), so renaming main is all it takes.
#include <iostream>
#include <fstream>
#include <string>
#include <sstream>
#include <stdio.h>
#include <stdlib.h>
#include <cmath>
#include <vector>
#include <limits>
#include <cstring>
#include <unistd.h>

#define main ref_main
#include "FreeEnergy/binary.cpp"
#undef main

extern "C" {
int ref_binary_nx(void) { return NX; }
int ref_binary_ny(void) { return NY; }

// Runs init() and steps 0..last_step.  Returns the fields of the last step and
// the populations the next step would start from; all in the reference's own
// [iX][iY] (x-major) layout.  noutput<=0 switches the .dat files off.
int ref_binary_run(int last_step, int noutput, const char *outdir, double *phase_out, double *rho_out,
                   double *ux_out, double *uy_out, double *f_out, double *g_out)
{
    char cwd[4096];
    if (!getcwd(cwd, sizeof cwd)) return -1;
    if (outdir && chdir(outdir) != 0) return -2;
    N = last_step;
    NOUTPUT = noutput > 0 ? noutput : 2000000000;
    NSIGNAL = 2000000000;
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    char arg0[] = "binary";
    char *argv[] = {arg0, 0};
    ref_main(1, argv);
    std::cout.rdbuf(old);
    if (chdir(cwd) != 0) return -3;
    if (phase_out) memcpy(phase_out, phase, sizeof phase);
    if (rho_out) memcpy(rho_out, rho, sizeof rho);
    if (ux_out) memcpy(ux_out, ux, sizeof ux);
    if (uy_out) memcpy(uy_out, uy, sizeof uy);
    if (f_out) memcpy(f_out, f, sizeof f);
    if (g_out) memcpy(g_out, g, sizeof g);
    return 0;
}
}
