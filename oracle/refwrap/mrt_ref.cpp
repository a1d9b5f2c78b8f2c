// Wrapper around the UNMODIFIED MRT/mrt.cpp (test infrastructure): main is
// renamed and called; the result arrays are external-linkage globals
// (mrt.cpp:14-15).  The program writes two files per step into ./tmp/ and
// silently skips them when that directory is missing, which is what happens
// when outdir has no tmp/ in it.
#include <iostream>
#include <fstream>
#include <string>
#include <sstream>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <unistd.h>

#define main ref_main
#include "MRT/mrt.cpp"
#undef main

extern "C" {
int ref_mrt_nx(void) { return NX; }
int ref_mrt_ny(void) { return NY; }
int ref_mrt_nsteps(void) { return N; }

// arg1/arg2 are the two integers mrt.cpp reads from argv (omega*10, omega_bulk*10);
// another_mrt.cpp ignores them.  Returns the state after the reference's full run.
int ref_mrt_run(int arg1, int arg2, const char *outdir, double *rho_out, double *ux_out,
                 double *uy_out, double *f_out)
{
    char cwd[4096];
    if (!getcwd(cwd, sizeof cwd)) return -1;
    if (outdir && chdir(outdir) != 0) return -2;
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    char a0[] = "mrt", a1[32], a2[32];
    snprintf(a1, sizeof a1, "%d", arg1);
    snprintf(a2, sizeof a2, "%d", arg2);
    char *argv[] = {a0, a1, a2, 0};
    ref_main(3, argv);
    std::cout.rdbuf(old);
    if (chdir(cwd) != 0) return -3;
    if (rho_out) memcpy(rho_out, rho, sizeof rho);
    if (ux_out) memcpy(ux_out, ux, sizeof ux);
    if (uy_out) memcpy(uy_out, uy, sizeof uy);
    if (f_out) memcpy(f_out, f, sizeof f);
    return 0;
}
}
