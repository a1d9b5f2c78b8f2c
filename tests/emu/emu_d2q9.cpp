// CPU emulation of the D2Q9 CUDA kernels (runs d2q9_threads.h from host loops, same
// layout and call order as d2q9_models.cu).  TEST INFRASTRUCTURE only.
#include <cmath>
#include <cstring>
#include <vector>
#include "d2q9_threads.h"
#include "d2q9_host.h"
using namespace lbm2;

struct EmuSc { ScParams P; std::vector<double> f[2], rho, psi, u1, u2; int cur = 0; };
struct EmuB2 { B2Params P; std::vector<double> f[2], g[2], phase, rho, ux, uy; int cur = 0; };
struct EmuTrt { TrtParams P; std::vector<double> f[2], rho, ux, uy; std::vector<unsigned short> links; int cur = 0; };
struct EmuBs { BsParams P; std::vector<double> f[2], g[2], phi, rho, ux, uy; std::vector<unsigned short> links; int cur = 0; };
struct EmuMrt { MrtParams P; std::vector<double> f[2], force, rho, ux, uy; int cur = 0; int aa = 0, odd_next = 0; };

extern "C" {
// ---- Shan-Chen ----
EmuSc *emusc_create(int nx, int ny, double tau, double g)
{
    EmuSc *e = new EmuSc; e->P = sc_params(nx, ny, tau, g);
    size_t n = (size_t)nx * ny;
    for (int i = 0; i < 2; i++) e->f[i].assign(9 * n, 0.0);
    e->rho.assign(n, 0.0); e->psi.assign(n, 0.0); e->u1.assign(n, 0.0); e->u2.assign(n, 0.0);
    return e;
}
void emusc_destroy(EmuSc *e) { delete e; }
void emusc_init(EmuSc *e, const double *rho)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(e->rho.data(), rho, n * sizeof(double));
    for (size_t i = 0; i < n; i++) sc_init_thread(e->P, e->rho.data(), e->u1.data(), e->u2.data(), e->f[e->cur].data(), i);
}
void emusc_step(EmuSc *e, int steps)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    for (int t = 0; t < steps; t++) {
        int a = e->cur, b = 1 - a;
        for (size_t i = 0; i < n; i++) sc_density_thread(e->P, e->f[a].data(), e->rho.data(), e->psi.data(), i);
        for (int iy = 0; iy < e->P.ny; iy++)
            for (int ix = 0; ix < e->P.nx; ix++)
                sc_collide_thread(e->P, e->f[a].data(), e->rho.data(), e->psi.data(), e->f[b].data(), e->u1.data(), e->u2.data(), ix, iy);
        e->cur = b;
    }
}
void emusc_get(EmuSc *e, double *rho, double *u1, double *u2, double *f_aos)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(rho, e->rho.data(), n * 8); memcpy(u1, e->u1.data(), n * 8); memcpy(u2, e->u2.data(), n * 8);
    for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++) f_aos[i * 9 + k] = e->f[e->cur][k * n + i];
}
// ---- binary 2-D ----
EmuB2 *emub2_create(int nx, int ny, const double *p)
{
    EmuB2 *e = new EmuB2;
    e->P = B2Params{nx, ny, p[0], p[1], p[2], p[3], p[4], p[5] - p[4], p[6]};
    size_t n = (size_t)nx * ny;
    for (int i = 0; i < 2; i++) { e->f[i].assign(9 * n, 0.0); e->g[i].assign(9 * n, 0.0); }
    e->phase.assign(n, 0.0); e->rho.assign(n, 0.0); e->ux.assign(n, 0.0); e->uy.assign(n, 0.0);
    return e;
}
void emub2_destroy(EmuB2 *e) { delete e; }
void emub2_init(EmuB2 *e, const double *phase)
{
    const int NX = e->P.nx, NY = e->P.ny;
    memcpy(e->phase.data(), phase, (size_t)NX * NY * 8);
    for (int ix = 0; ix < NX; ix++) {   // binary.cpp:136-140
        e->phase[(size_t)ix * NY] = e->phase[(size_t)ix * NY + 1] - e->P.wall_gradient;
        e->phase[(size_t)ix * NY + NY - 1] = e->phase[(size_t)ix * NY + NY - 2] - e->P.wall_gradient;
    }
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 0; iy < NY; iy++)
            b2_init_thread(e->P, e->phase.data(), e->f[0].data(), e->g[0].data(), e->f[1].data(), e->g[1].data(),
                           e->rho.data(), e->ux.data(), e->uy.data(), ix, iy);
}
void emub2_step(EmuB2 *e, int steps)
{
    const int NX = e->P.nx, NY = e->P.ny;
    for (int t = 0; t < steps; t++) {
        int a = e->cur, b = 1 - a;
        for (int ix = 0; ix < NX; ix++) for (int iy = 0; iy < NY; iy++) b2_phase_thread(e->P, e->g[a].data(), e->phase.data(), ix, iy);
        for (int ix = 0; ix < NX; ix++) for (int iy = 0; iy < NY; iy++)
            b2_collide_thread(e->P, e->f[a].data(), e->g[a].data(), e->phase.data(), e->f[b].data(), e->g[b].data(),
                              e->rho.data(), e->ux.data(), e->uy.data(), ix, iy);
        e->cur = b;
    }
}
void emub2_get(EmuB2 *e, double *phase, double *rho, double *ux, double *uy, double *f, double *g)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(phase, e->phase.data(), n * 8); memcpy(rho, e->rho.data(), n * 8);
    memcpy(ux, e->ux.data(), n * 8); memcpy(uy, e->uy.data(), n * 8);
    for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++) { f[i * 9 + k] = e->f[e->cur][k * n + i]; g[i * 9 + k] = e->g[e->cur][k * n + i]; }
}
// ---- MRT ----
EmuMrt *emumrt_create(int variant, int nx, int ny, double omega, double omega_bulk, double force_ampl)
{
    EmuMrt *e = new EmuMrt;
    e->force.assign(nx, 0.0);
    mrt_build_tables(&e->P, variant, nx, ny, omega, omega_bulk, force_ampl, e->force.data());
    size_t n = (size_t)nx * ny;
    for (int i = 0; i < 2; i++) e->f[i].assign(9 * n, 0.0);
    e->rho.assign(n, 0.0); e->ux.assign(n, 0.0); e->uy.assign(n, 0.0);
    return e;
}
void emumrt_destroy(EmuMrt *e) { delete e; }
void emumrt_init(EmuMrt *e, const double *rho, const double *ux, const double *uy)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(e->rho.data(), rho, n * 8); memcpy(e->ux.data(), ux, n * 8); memcpy(e->uy.data(), uy, n * 8);
    for (size_t i = 0; i < n; i++) mrt_init_thread(e->P, e->rho.data(), e->ux.data(), e->uy.data(), e->f[e->cur].data(), i);
}
// in-place AA-pattern mode: everything on f[0], even and odd steps alternating (mrt_thread_aa)
void emumrt_use_aa(EmuMrt *e, int on) { e->aa = on; e->odd_next = 0; }
static void aa_step(EmuMrt *e)
{
    for (int ix = 0; ix < e->P.nx; ix++) for (int iy = 0; iy < e->P.ny; iy++) {
        double *f = e->f[0].data();
        if (e->P.variant == 0) {
            if (e->odd_next) mrt_thread_aa<0, true>(e->P, f, e->force.data(), e->rho.data(), e->ux.data(), e->uy.data(), ix, iy, true);
            else mrt_thread_aa<0, false>(e->P, f, e->force.data(), e->rho.data(), e->ux.data(), e->uy.data(), ix, iy, true);
        } else {
            if (e->odd_next) mrt_thread_aa<1, true>(e->P, f, e->force.data(), e->rho.data(), e->ux.data(), e->uy.data(), ix, iy, true);
            else mrt_thread_aa<1, false>(e->P, f, e->force.data(), e->rho.data(), e->ux.data(), e->uy.data(), ix, iy, true);
        }
    }
    e->odd_next = !e->odd_next;
}
void emumrt_step(EmuMrt *e, int steps)
{
    if (e->aa) {
        for (int t = 0; t < steps; t++) aa_step(e);
        return;
    }
    for (int t = 0; t < steps; t++) {
        int a = e->cur, b = 1 - a;
        for (int ix = 0; ix < e->P.nx; ix++) for (int iy = 0; iy < e->P.ny; iy++)
            mrt_thread(e->P, e->f[a].data(), e->force.data(), e->f[b].data(), e->rho.data(), e->ux.data(), e->uy.data(), ix, iy);
        e->cur = b;
    }
}
void emumrt_get(EmuMrt *e, double *rho, double *ux, double *uy, double *f)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(rho, e->rho.data(), n * 8); memcpy(ux, e->ux.data(), n * 8); memcpy(uy, e->uy.data(), n * 8);
    if (e->aa) {
        for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++)
            f[i * 9 + k] = mrt_aa_streamed(e->P, e->f[0].data(), e->odd_next, k, (int)(i / e->P.ny), (int)(i % e->P.ny));
        return;
    }
    for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++) f[i * 9 + k] = e->f[e->cur][k * n + i];
}
void emumrt_wave(int variant, int nx, int ny, double ampl, double force_ampl, double *rho, double *ux, double *uy)
{
    mrt_reference_wave(variant, nx, ny, ampl, force_ampl, rho, ux, uy);
}
// ---- TRT diffusion ----
EmuTrt *emutrt_create(int nx, int ny, double omega, double ce, double conc_wall, const int *mask)
{
    EmuTrt *e = new EmuTrt;
    e->P = TrtParams{nx, ny, 2.0 - omega, omega, ce, conc_wall};
    size_t n = (size_t)nx * ny;
    for (int i = 0; i < 2; i++) e->f[i].assign(9 * n, 0.0);
    e->rho.assign(n, 0.0); e->ux.assign(n, 0.0); e->uy.assign(n, 0.0);
    e->links.assign(n, 0);
    for (size_t i = 0; i < n; i++) e->links[i] = trt_links_thread(e->P, mask, (long long)i);
    return e;
}
void emutrt_destroy(EmuTrt *e) { delete e; }
void emutrt_init(EmuTrt *e, const double *rho, const double *ux, const double *uy)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(e->rho.data(), rho, n * 8); memcpy(e->ux.data(), ux, n * 8); memcpy(e->uy.data(), uy, n * 8);
    for (size_t i = 0; i < n; i++) trt_init_thread(e->P, e->rho.data(), e->ux.data(), e->uy.data(), e->f[0].data(), e->f[1].data(), i);
}
void emutrt_step(EmuTrt *e, int steps)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    for (int t = 0; t < steps; t++) {
        int a = e->cur, b = 1 - a;
        for (size_t i = 0; i < n; i++)
            trt_thread(e->P, e->f[a].data(), e->links.data(), e->ux.data(), e->uy.data(), e->f[b].data(), e->rho.data(), (long long)i);
        e->cur = b;
    }
}
void emutrt_get(EmuTrt *e, double *rho, double *f)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(rho, e->rho.data(), n * 8);
    for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++) f[i * 9 + k] = e->f[e->cur][k * n + i];
}
// ---- binary liquid around a curved solid ----
// p[9] = aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi, force_x, force_y
EmuBs *emubs_create(int nx, int ny, const double *p, const int *mask)
{
    EmuBs *e = new EmuBs;
    e->P = BsParams{nx, ny, p[0], p[1], p[2], p[3], p[4], p[5] - p[4], p[6], p[7], p[8]};
    size_t n = (size_t)nx * ny;
    for (int i = 0; i < 2; i++) { e->f[i].assign(9 * n, 0.0); e->g[i].assign(9 * n, 0.0); }
    e->phi.assign(n, 0.0); e->rho.assign(n, 0.0); e->ux.assign(n, 0.0); e->uy.assign(n, 0.0);
    e->links.assign(n, 0);
    for (size_t i = 0; i < n; i++) e->links[i] = bs_links_thread(e->P, mask, (long long)i);
    return e;
}
void emubs_destroy(EmuBs *e) { delete e; }
void emubs_init(EmuBs *e, const double *rho, const double *phi, const double *ux, const double *uy)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(e->rho.data(), rho, n * 8); memcpy(e->phi.data(), phi, n * 8); memcpy(e->ux.data(), ux, n * 8); memcpy(e->uy.data(), uy, n * 8);
    for (size_t i = 0; i < n; i++)
        bs_init_thread(e->P, e->rho.data(), e->phi.data(), e->ux.data(), e->uy.data(), e->f[0].data(), e->g[0].data(), e->f[1].data(),
                       e->g[1].data(), (long long)i);
}
void emubs_step(EmuBs *e, int steps)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    for (int t = 0; t < steps; t++) {
        int a = e->cur, b = 1 - a;
        for (size_t i = 0; i < n; i++) bs_phase_thread(e->P, e->g[a].data(), e->links.data(), e->phi.data(), (long long)i);
        for (size_t i = 0; i < n; i++)
            bs_collide_thread(e->P, e->f[a].data(), e->g[a].data(), e->phi.data(), e->links.data(), e->f[b].data(), e->g[b].data(),
                              e->rho.data(), e->ux.data(), e->uy.data(), (long long)i);
        e->cur = b;
    }
}
void emubs_get(EmuBs *e, double *phi, double *rho, double *ux, double *uy, double *f, double *g)
{
    size_t n = (size_t)e->P.nx * e->P.ny;
    memcpy(phi, e->phi.data(), n * 8); memcpy(rho, e->rho.data(), n * 8); memcpy(ux, e->ux.data(), n * 8); memcpy(uy, e->uy.data(), n * 8);
    for (size_t i = 0; i < n; i++) for (int k = 0; k < 9; k++) { f[i * 9 + k] = e->f[e->cur][k * n + i]; g[i * 9 + k] = e->g[e->cur][k * n + i]; }
}
}
