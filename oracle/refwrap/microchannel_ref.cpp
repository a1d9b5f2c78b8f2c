// Driver around the UNMODIFIED reference Microchannel3D sources (test infrastructure).
//
// The reference's Lattice<D>/Solver<D> templates (lattice.hpp, solver.hpp,
// descriptor.h, params_list.h) are compiled from where they lie under
// /root/reference (-I on the command line, see oracle/Makefile); only the thin
// main.cpp is replaced, because it hard-codes the grid size
// (QuarterChannel/main.cpp:18-26).  What main.cpp does is restated here with the
// sizes as run-time arguments: the parameter list (main.cpp:47-62), the
// initial condition (main.cpp:72-122: a gas bubble phi=-1 in the middle third
// of the channel, liquid phi=+1 elsewhere, rho=1, u=0), solver.init() and the
// time loop of solver.collide_stream() calls.
//
// Two front ends:
//  * a C API (ref3d_*) loaded by the tests through ctypes -- one rank;
//  * with -DREF3D_STANDALONE a main() that times the loop and can run on P
//    forked ranks through oracle/shim/minimpi.cpp (LBM_MINIMPI_RANKS=P): the
//    reference's real z-slab decomposition on the host cores, used by
//    `bench.py --impl reference`.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <iostream>
#include <sstream>
#include <string>

#include <mpi.h>
#include "mpi_singleton.h"
#include "lattice.hpp"
#include "solver.hpp"
#include "params_list.h"
#include "descriptor.h"

namespace {

struct Quiet {  // the Lattice ctor prints 57 table lines (lattice.h:83-88)
    std::streambuf *old;
    std::ostringstream sink;
    Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~Quiet() { std::cout.rdbuf(old); }
};

// order of the 12 doubles in `p`
enum { P_ACONST, P_KCONST, P_GAMMA, P_PHASE_GRADIENT, P_PHASE_WALL, P_RHO_WALL,
       P_TAU_PHI, P_TAU_LIQ, P_TAU_GAS, P_FORCE_X, P_FORCE_Y, P_FORCE_Z, P_COUNT };

ParamsList make_params(int nx, int ny, int nz, const double *p)
{
    ParamsList params;
    params.add("omega", 1.0);
    params.add("aconst", p[P_ACONST]);
    params.add("kconst", p[P_KCONST]);
    params.add("gammaconst", p[P_GAMMA]);
    params.add("phase_gradient", p[P_PHASE_GRADIENT]);
    params.add("phase_wall", p[P_PHASE_WALL]);
    params.add("rho_wall", p[P_RHO_WALL]);
    params.add("tau_phi", p[P_TAU_PHI]);
    params.add("tau_liq", p[P_TAU_LIQ]);
    params.add("tau_gas", p[P_TAU_GAS]);
    params.add("force_x", p[P_FORCE_X]);
    params.add("force_y", p[P_FORCE_Y]);
    params.add("force_z", p[P_FORCE_Z]);
    params.add("NX", nx);
    params.add("NY", ny);
    params.add("NZ", nz);
    return params;
}

MPISingleton *g_mpi = 0;

struct Ref3D {
    int nx, ny, nz;
    Solver<Descriptor> *solver;
};

}  // namespace

extern "C" {

void *ref3d_create(int nx, int ny, int nz, const double *p)
{
    Quiet q;
    if (!g_mpi) g_mpi = new MPISingleton;
    Ref3D *r = new Ref3D;
    r->nx = nx; r->ny = ny; r->nz = nz;
    r->solver = new Solver<Descriptor>(make_params(nx, ny, nz, p));
    return r;
}

// fields in the reference's node order iZ*NX*NY+iY*NX+iX (lattice.h:136); one rank
void ref3d_set_fields(void *h, const double *rho, const double *phase,
                      const double *ux, const double *uy, const double *uz)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    Lattice<Descriptor> *l = r->solver->getLattice();
    size_t n = (size_t)r->nx * r->ny * r->nz;
    memcpy(l->rho, rho, n * sizeof(double));
    memcpy(l->phase, phase, n * sizeof(double));
    memcpy(l->ux, ux, n * sizeof(double));
    memcpy(l->uy, uy, n * sizeof(double));
    memcpy(l->uz, uz, n * sizeof(double));
}

// the bubble of main.cpp:83-100, through the same put* calls
void ref3d_set_bubble(void *h, int width)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    const int NX = r->nx, NY = r->ny, NZ = r->nz;
    double u0[3] = {0.0, 0.0, 0.0};
    for (int iZ = 0; iZ < NZ; iZ++)
        for (int iY = 0; iY < NY; iY++)
            for (int iX = 0; iX < NX; iX++) {
                bool gas = (iZ >= (NZ - 1) / 3) && (iZ <= 2 * (NZ - 1) / 3) &&
                           (iX <= NX - width - 1) && (iY <= NY - width - 1);
                r->solver->putDensity(iX, iY, iZ, iX, iY, iZ, 1.0);
                r->solver->putPhase(iX, iY, iZ, iX, iY, iZ, gas ? -1.0 : 1.0);
                r->solver->putVelocity(iX, iY, iZ, iX, iY, iZ, u0);
            }
}

void ref3d_init(void *h) { static_cast<Ref3D *>(h)->solver->init(); }

void ref3d_step(void *h, int n)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    for (int t = 0; t < n; t++) r->solver->collide_stream();
}

// rho, phase, u as the lattice holds them now: after n steps these are the
// moments computed at the start of the last step (what main.cpp:141-163 writes)
void ref3d_get_fields(void *h, double *rho, double *phase, double *ux, double *uy, double *uz)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    Lattice<Descriptor> *l = r->solver->getLattice();
    size_t n = (size_t)r->nx * r->ny * r->nz;
    memcpy(rho, l->rho, n * sizeof(double));
    memcpy(phase, l->phase, n * sizeof(double));
    memcpy(ux, l->ux, n * sizeof(double));
    memcpy(uy, l->uy, n * sizeof(double));
    memcpy(uz, l->uz, n * sizeof(double));
}

// current populations, reference layout [node*19+k]
void ref3d_get_pops(void *h, double *f, double *g)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    Lattice<Descriptor> *l = r->solver->getLattice();
    size_t n = (size_t)r->nx * r->ny * r->nz * 19;
    memcpy(f, l->f, n * sizeof(double));
    memcpy(g, l->g, n * sizeof(double));
}

void ref3d_set_pops(void *h, const double *f, const double *g)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    Lattice<Descriptor> *l = r->solver->getLattice();
    size_t n = (size_t)r->nx * r->ny * r->nz * 19;
    memcpy(l->f, f, n * sizeof(double));
    memcpy(l->g, g, n * sizeof(double));
}

// every text writer of the reference Solver (solver.hpp:578-787), into `dir`, under the names hosts/solver_selftest.cpp uses
int ref3d_write_text(void *h, const char *dir)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    if (chdir(dir) != 0) return 1;
    Solver<Descriptor> &s = *r->solver;
    s.writeTextXZVelocity("vel_xz"); s.writeTextXYVelocity("vel_xy"); s.writeTextYZVelocity("vel_yz");
    s.writeTextWholeVelocity("vel_whole"); s.writeTextWholePhase("phase_whole"); s.writeTextWholeDensity("density_whole");
    s.writeTextXYPhase("phase_xy"); s.writeTextXZPhase("phase_xz"); s.writeTextYZPhase("phase_yz");
    s.writeTextXYDensity("density_xy"); s.writeTextXZDensity("density_xz"); s.writeTextYZDensity("density_yz");
    s.writeVTKWholePhase("phase_legacy");
    return 0;
}

void ref3d_destroy(void *h)
{
    Ref3D *r = static_cast<Ref3D *>(h);
    delete r->solver;
    delete r;
}

}  // extern "C"

#ifdef REF3D_STANDALONE
// usage: ref_microchannel NX NY NZ WARMUP STEPS [width] ; ranks from LBM_MINIMPI_RANKS
int main(int argc, char **argv)
{
    if (argc < 6) { fprintf(stderr, "usage: %s NX NY NZ WARMUP STEPS [width]\n", argv[0]); return 2; }
    int nx = atoi(argv[1]), ny = atoi(argv[2]), nz = atoi(argv[3]);
    int warm = atoi(argv[4]), steps = atoi(argv[5]);
    int width = argc > 6 ? atoi(argv[6]) : 6;
    // QuarterChannel/main.cpp:48-62
    double p[P_COUNT] = {0.04, 0.04, 1.0, 0.0, 0.0, 0.5, 2.0, 2.5, 0.7, 0.0, 0.0, 0.00003};
    Ref3D *r;
    {
        Quiet q;
        g_mpi = new MPISingleton;  // forks the ranks
        r = static_cast<Ref3D *>(ref3d_create(nx, ny, nz, p));
    }
    ref3d_set_bubble(r, width);
    ref3d_init(r);
    ref3d_step(r, warm);
    MPI_Barrier(MPI_COMM_WORLD);
    auto t0 = std::chrono::steady_clock::now();
    ref3d_step(r, steps);
    MPI_Barrier(MPI_COMM_WORLD);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // checksum over rank 0's slab so the work cannot be optimised away
    Lattice<Descriptor> *l = r->solver->getLattice();
    double sum = 0.0;
    for (int i = 0; i < l->getNUMLOCAL(); i++) sum += l->phase[i];
    if (minimpi_rank() == 0)
        printf("{\"nx\": %d, \"ny\": %d, \"nz\": %d, \"ranks\": %d, \"steps\": %d, \"seconds\": %.6f, "
               "\"mlups\": %.6f, \"phase_sum_rank0\": %.17g}\n",
               nx, ny, nz, minimpi_size(), steps, sec, (double)nx * ny * nz * steps / sec / 1e6, sum);
    fflush(stdout);
    delete g_mpi;  // MPI_Finalize: children exit, rank 0 reaps them
    return 0;
}
#endif
