// Wrapper around the UNMODIFIED ShanChen/shanchen.cpp (test infrastructure).
//
// The reference is a single main() with the step count compiled in
// (shanchen.cpp:14) and its state in external-linkage globals (:29-33).  This
// translation unit includes the standard headers first, then renames three
// tokens with the preprocessor and includes the source where it lies:
//   main  -> ref_main        so it can be called;
//   swap  -> lbm_ref_step_hook   std::swap(f,f2) (:188) is the last statement of
//                                every time step, so a function of that name in
//                                namespace std sees each step end, can snapshot
//                                the state and can stop the run early;
//   exp   -> lbm_exp         only with -DREF_SHARED_EXP: the bit-portable exp of
//                                include/lbm_exp.h.
// Without -DREF_SHARED_EXP the program keeps glibc's exp.
#include <iostream>
#include <cmath>
#include <time.h>
#include <fstream>
#include <sstream>
#include <string>
#include <cstdlib>
#include <algorithm>
#include <utility>
#include <cstring>
#include <unistd.h>

#ifdef REF_SHARED_EXP
#include "lbm_exp.h"
#endif

namespace lbmref {
struct Stop {};
int step = 0;            // index of the step that just finished
int stop_after = -1;     // finish the run once this step is done
int snap_step = -1;      // copy the state when this step is done
double *snap_rho = 0, *snap_u1 = 0, *snap_u2 = 0, *snap_f = 0;
void on_step(double *f_next);
}
namespace std {
inline void lbm_ref_step_hook(double *&a, double *&b)
{
    double *t = a; a = b; b = t;
    lbmref::on_step(a);
}
}

#define main ref_main
#define swap lbm_ref_step_hook
#ifdef REF_SHARED_EXP
#define exp lbm_exp
#endif
#include "ShanChen/shanchen.cpp"
#undef main
#undef swap
#ifdef REF_SHARED_EXP
#undef exp
#endif

namespace lbmref {
void on_step(double *f_next)
{
    if (step == snap_step) {
        if (snap_rho) memcpy(snap_rho, rho, sizeof rho);
        if (snap_u1) memcpy(snap_u1, u1, sizeof u1);
        if (snap_u2) memcpy(snap_u2, u2, sizeof u2);
        if (snap_f) memcpy(snap_f, f_next, sizeof f_mem);
    }
    if (step == stop_after) throw Stop();
    ++step;
}
}

extern "C" {
int ref_shanchen_nx(void) { return nx; }
int ref_shanchen_ny(void) { return ny; }
int ref_shanchen_nsteps(void) { return nsteps; }

// Runs the reference from its own initial condition through step `last_step`
// (0-based, as the reference counts) and returns rho,u1,u2 of that step (the
// values the reference would write, shanchen.cpp:190-207) and the populations
// it would start the next step with.  Output files go to `outdir` (may be "."
// or a scratch directory; the reference writes density/velx/vely%05d.dat every
// 1000 steps into the working directory).
int ref_shanchen_run(int last_step, const char *outdir, double *rho_out, double *u1_out,
                     double *u2_out, double *f_out)
{
    char cwd[4096];
    if (!getcwd(cwd, sizeof cwd)) return -1;
    if (outdir && chdir(outdir) != 0) return -2;
    lbmref::step = 0;
    // the hook fires before the step's file output (shanchen.cpp:190-207); let the final step run to the
    // program's own end so that its .dat files are written too
    lbmref::stop_after = last_step >= nsteps ? -1 : last_step;
    lbmref::snap_step = last_step;
    lbmref::snap_rho = rho_out; lbmref::snap_u1 = u1_out; lbmref::snap_u2 = u2_out; lbmref::snap_f = f_out;
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    char arg0[] = "shanchen";
    char *argv[] = {arg0, 0};
    int rc = 0;
    try { ref_main(1, argv); } catch (lbmref::Stop &) { rc = 1; }
    std::cout.rdbuf(old);
    if (chdir(cwd) != 0) return -3;
    return rc;  // 1 = stopped at last_step, 0 = ran to the reference's own end
}
}
