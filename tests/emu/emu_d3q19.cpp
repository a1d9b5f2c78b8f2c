// CPU emulation of the plain D3Q19 CUDA kernels: runs the per-thread functions of
// latticeboltzmannmethod_b200/csrc/d3q19_threads.h from host loops, with the same
// storage layout and the same call order as the CUDA library.  TEST INFRASTRUCTURE
// (lets the kernel logic be checked against the oracle without a GPU); never part
// of the product library.  Build: g++ -O2 -ffp-contract=off (tests/emu/build.sh).
#include <cstdlib>
#include <cstring>
#include <vector>
#include "d3q19_threads.h"

using namespace lbm3;

struct Emu {
    Geom g;
    Params P;
    bool fxy;
    std::vector<double> f[2], gq[2], phi, rho, ux, uy, uz;
    int cur = 0;
    int streamed = 0;        // 0: collide(), 1: collide_streamed(), 3: interior(PART 3)+rim rows, 4: in-place (AA) on f[cur] alone
    bool odd = false;        // mode 4: the array holds post-collision values in the opposite slots
};

extern "C" {
Emu *emu3_create(int nx, int ny, int nz, const double *p)
{
    Emu *e = new Emu;
    e->g.nx = nx; e->g.ny = ny; e->g.nz = nz; e->g.zwrap = 1;
    e->g.pl = ((long long)nx * ny + 15) / 16 * 16;
    e->P.aconst = p[0]; e->P.kconst = p[1]; e->P.gammaconst = p[2]; e->P.phase_gradient = p[3];
    e->P.rho_wall = p[5]; e->P.omega_phi = 1.0 / p[6]; e->P.tau_gas = p[8]; e->P.dtau = p[7] - p[8];
    e->P.fx = p[9]; e->P.fy = p[10]; e->P.fz = p[11];
    e->fxy = p[9] != 0.0 || p[10] != 0.0;
    size_t npop = (size_t)(nz + 4) * 19 * e->g.pl;
    for (int i = 0; i < 2; i++) { e->f[i].assign(npop, 0.0); e->gq[i].assign(npop, 0.0); }
    e->phi.assign((size_t)(nz + 4) * e->g.pl, 0.0);
    size_t nf = (size_t)nz * e->g.pl;
    e->rho.assign(nf, 0.0); e->ux.assign(nf, 0.0); e->uy.assign(nf, 0.0); e->uz.assign(nf, 0.0);
    return e;
}
void emu3_destroy(Emu *e) { delete e; }
// one node through collide<>() of d3q19_node.h: F, G in/out; PH = the 19 neighbour phase values; out4 = rho, ux, uy, uz;
// mrt5: the five free rates of the MRT extension (all 0: off)
void emu3_collide_node(const double *p, double trt_magic, const double *mrt5, double *F, double *G, const double *PH, double *out4)
{
    Params P{};
    P.aconst = p[0]; P.kconst = p[1]; P.gammaconst = p[2]; P.phase_gradient = p[3];
    P.rho_wall = p[5]; P.omega_phi = 1.0 / p[6]; P.tau_gas = p[8]; P.dtau = p[7] - p[8];
    P.fx = p[9]; P.fy = p[10]; P.fz = p[11];
    P.trt_magic = trt_magic;
    bool mrt = false;
    for (int i = 0; i < 5; i++) { P.mrt[i] = mrt5[i]; mrt = mrt || mrt5[i] != 0.0; }
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    if (mrt) collide<true, 2>(F, G, PH[0], gx, gy, gz, lap, P, out4[0], out4[1], out4[2], out4[3]);
    else if (trt_magic > 0.0) collide<true, 1>(F, G, PH[0], gx, gy, gz, lap, P, out4[0], out4[1], out4[2], out4[3]);
    else collide<true, 0>(F, G, PH[0], gx, gy, gz, lap, P, out4[0], out4[1], out4[2], out4[3]);
}
void emu3_use_streamed(Emu *e, int on) { e->streamed = on; }

static void put(Emu *e, double *dst, const double *src)
{
    size_t n = (size_t)e->g.nx * e->g.ny;
    for (int z = 0; z < e->g.nz; z++) memcpy(dst + (size_t)z * e->g.pl, src + (size_t)z * n, n * sizeof(double));
}
static void get(Emu *e, double *dst, const double *src)
{
    size_t n = (size_t)e->g.nx * e->g.ny;
    for (int z = 0; z < e->g.nz; z++) memcpy(dst + (size_t)z * n, src + (size_t)z * e->g.pl, n * sizeof(double));
}
void emu3_set_fields(Emu *e, const double *rho, const double *phase, const double *ux, const double *uy, const double *uz)
{
    put(e, e->rho.data(), rho); put(e, e->phi.data() + 2 * e->g.pl, phase);
    put(e, e->ux.data(), ux); put(e, e->uy.data(), uy); put(e, e->uz.data(), uz);
}
void emu3_get_fields(Emu *e, double *rho, double *phase, double *ux, double *uy, double *uz)
{
    get(e, rho, e->rho.data()); get(e, phase, e->phi.data() + 2 * e->g.pl);
    get(e, ux, e->ux.data()); get(e, uy, e->uy.data()); get(e, uz, e->uz.data());
}
static void self_exchange(Emu *e, int b)
{
    const size_t pp = 19 * (size_t)e->g.pl;
    const int nz = e->g.nz;
    double *f = e->f[b].data(), *g = e->gq[b].data();
    memcpy(f + (size_t)(-1 + 2) * pp, f + (size_t)(nz - 1 + 2) * pp, pp * sizeof(double));
    memcpy(g + (size_t)(-2 + 2) * pp, g + (size_t)(nz - 2 + 2) * pp, 2 * pp * sizeof(double));
    memcpy(f + (size_t)(nz + 2) * pp, f + (size_t)(0 + 2) * pp, pp * sizeof(double));
    memcpy(g + (size_t)(nz + 2) * pp, g + (size_t)(0 + 2) * pp, 2 * pp * sizeof(double));
}
void emu3_init(Emu *e)
{
    const Geom &g = e->g;
    for (int z = 0; z < g.nz; z++) init_fields_plane(g, e->P, e->phi.data(), e->rho.data(), z);
    memcpy(e->phi.data() + 1 * g.pl, e->phi.data() + (size_t)(g.nz - 1 + 2) * g.pl, g.pl * sizeof(double));
    memcpy(e->phi.data() + (size_t)(g.nz + 2) * g.pl, e->phi.data() + 2 * g.pl, g.pl * sizeof(double));
    int b = 1 - e->cur;
    for (int z = 0; z < g.nz; z++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++)
                init_pops_thread(g, e->P, e->phi.data(), e->rho.data(), e->ux.data(), e->uy.data(), e->uz.data(),
                                 e->f[b].data(), e->gq[b].data(), z, ix, iy);
    self_exchange(e, b);
    e->cur = b;
    FieldsOut o{e->rho.data(), e->ux.data(), e->uy.data(), e->uz.data()};
    for (int zs = 0; zs < g.nz + 4; zs++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++) phase_thread(g, e->P, e->gq[e->cur].data(), e->phi.data(), zs, ix, iy);
    for (int z = 0; z < g.nz; z++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++) after_init_thread(g, e->P, e->f[e->cur].data(), o, z, ix, iy);
}
}  // extern "C"
static AaPlanes aa_planes(const Geom &g, int zs)
{
    const long long pp = 19 * g.pl;
    const int zm = zs == 2 ? g.nz + 1 : zs - 1, zp = zs == g.nz + 1 ? 2 : zs + 1;
    return AaPlanes{zm * pp, zs * pp, zp * pp};
}
template <bool ODD>
static void aa_step(Emu *e, bool w)
{
    const Geom &g = e->g;
    FieldsOut o{e->rho.data(), e->ux.data(), e->uy.data(), e->uz.data()};
    double *f = e->f[e->cur].data(), *gq = e->gq[e->cur].data(), *phi = e->phi.data();
    for (int zs = 2; zs < g.nz + 2; zs++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++) aa_phase_thread<ODD>(g, e->P, gq, phi, zs, aa_planes(g, zs), ix, iy);
    memcpy(phi + 1 * g.pl, phi + (size_t)(g.nz + 1) * g.pl, g.pl * sizeof(double));
    memcpy(phi + (size_t)(g.nz + 2) * g.pl, phi + 2 * g.pl, g.pl * sizeof(double));
    for (int zs = 2; zs < g.nz + 2; zs++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++) {
                const AaPlanes z = aa_planes(g, zs);
                if (e->fxy) {
                    if (w) aa_collide_thread<true, true, 0, ODD>(g, e->P, f, gq, phi, o, zs, z, ix, iy);
                    else aa_collide_thread<true, false, 0, ODD>(g, e->P, f, gq, phi, o, zs, z, ix, iy);
                } else {
                    if (w) aa_collide_thread<false, true, 0, ODD>(g, e->P, f, gq, phi, o, zs, z, ix, iy);
                    else aa_collide_thread<false, false, 0, ODD>(g, e->P, f, gq, phi, o, zs, z, ix, iy);
                }
            }
}
static void aa_normalize(Emu *e)
{
    if (!e->odd) return;
    const Geom &g = e->g;
    for (int zs = 2; zs < g.nz + 2; zs++)
        for (int iy = 0; iy < g.ny; iy++)
            for (int ix = 0; ix < g.nx; ix++) aa_swap_thread(g, e->f[e->cur].data(), e->gq[e->cur].data(), aa_planes(g, zs), ix, iy);
    e->odd = false;
}
extern "C" {
void emu3_step(Emu *e, int n)
{
    const Geom &g = e->g;
    FieldsOut o{e->rho.data(), e->ux.data(), e->uy.data(), e->uz.data()};
    if (e->streamed == 4) {
        for (int t = 0; t < n; t++) {
            if (e->odd) aa_step<true>(e, t == n - 1);
            else aa_step<false>(e, t == n - 1);
            e->odd = !e->odd;
        }
        return;
    }
    for (int t = 0; t < n; t++) {
        const int a = e->cur, b = 1 - a;
        const bool w = t == n - 1;
        for (int zs = 0; zs < g.nz + 4; zs++)
            for (int iy = 0; iy < g.ny; iy++)
                for (int ix = 0; ix < g.nx; ix++) phase_thread(g, e->P, e->gq[a].data(), e->phi.data(), zs, ix, iy);
        for (int zs = 1; zs <= g.nz + 2; zs++)
            for (int iy = 0; iy < g.ny; iy++)
                for (int ix = 0; ix < g.nx; ix++) {
                    if (e->streamed == 3) {
                        if (e->fxy) {
                            if (w) collide_push_thread_rows<true, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                            else collide_push_thread_rows<true, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        } else {
                            if (w) collide_push_thread_rows<false, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                            else collide_push_thread_rows<false, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        }
                        continue;
                    }
                    if (e->streamed) {
                        if (e->fxy) {
                            if (w) collide_push_thread_streamed<true, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                            else collide_push_thread_streamed<true, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        } else {
                            if (w) collide_push_thread_streamed<false, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                            else collide_push_thread_streamed<false, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        }
                        continue;
                    }
                    if (e->fxy) {
                        if (w) collide_push_thread<true, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        else collide_push_thread<true, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                    } else {
                        if (w) collide_push_thread<false, true>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                        else collide_push_thread<false, false>(g, e->P, e->f[a].data(), e->gq[a].data(), e->phi.data(), e->f[b].data(), e->gq[b].data(), o, zs, ix, iy);
                    }
                }
        self_exchange(e, b);
        e->cur = b;
    }
}
void emu3_get_pops(Emu *e, double *f, double *g)
{
    aa_normalize(e);
    for (int z = 0; z < e->g.nz; z++)
        for (int iy = 0; iy < e->g.ny; iy++)
            for (int ix = 0; ix < e->g.nx; ix++) {
                gather_aos_thread(e->g, e->f[e->cur].data(), f, z, ix, iy);
                gather_aos_thread(e->g, e->gq[e->cur].data(), g, z, ix, iy);
            }
}
void emu3_set_pops(Emu *e, const double *f, const double *g)
{
    for (int z = 0; z < e->g.nz; z++)
        for (int iy = 0; iy < e->g.ny; iy++)
            for (int ix = 0; ix < e->g.nx; ix++) {
                scatter_aos_thread(e->g, f, e->f[e->cur].data(), z, ix, iy);
                scatter_aos_thread(e->g, g, e->gq[e->cur].data(), z, ix, iy);
            }
    self_exchange(e, e->cur);
}
}

// ---- the in-place step on several slabs, emulated: the halo protocol of Slab3D::aa_exchange with memcpy as transport ----
// Splits the (initialised, even-half) state of `whole` into `nslabs` slabs, runs `steps` in-place steps on both, and
// returns the number of population values (streamed order) that differ between the slabs and the whole grid.
namespace {
struct SlabEmu {
    Geom g;
    std::vector<double> f, gq, phi;
    int z0;
};
void copy_rows(std::vector<double> &dst, int zd, const std::vector<double> &src, int zs, const Geom &g, bool down_set, bool merge = false)
{
    static const int kDown[5] = {6, 13, 14, 17, 18}, kUp[5] = {5, 11, 12, 15, 16};
    const int *ks = down_set ? kDown : kUp;
    for (int i = 0; i < 5; i++) {
        if (!merge) {
            memcpy(&dst[((size_t)zd * 19 + ks[i]) * g.pl], &src[((size_t)zs * 19 + ks[i]) * g.pl], g.pl * sizeof(double));
            continue;
        }
        for (int iy = 1; iy < g.ny - 1; iy++)     // what the CUDA library does with k_aa_merge
            for (int ix = 1; ix < g.nx - 1; ix++)
                if (aa_crossed_face(g, ks[i], ix, iy))
                    dst[((size_t)zd * 19 + ks[i]) * g.pl + (size_t)iy * g.nx + ix] = src[((size_t)zs * 19 + ks[i]) * g.pl + (size_t)iy * g.nx + ix];
    }
}
template <bool ODD>
void slab_step(std::vector<SlabEmu> &S, const Params &P, bool fxy)
{
    const int n = (int)S.size();
    for (auto &s : S)
        for (int zs = 2; zs < s.g.nz + 2; zs++)
            for (int iy = 0; iy < s.g.ny; iy++)
                for (int ix = 0; ix < s.g.nx; ix++) {
                    const long long pp = 19 * s.g.pl;
                    aa_phase_thread<ODD>(s.g, P, s.gq.data(), s.phi.data(), zs, AaPlanes{(zs - 1) * pp, zs * pp, (zs + 1) * pp}, ix, iy);
                }
    for (int r = 0; r < n; r++) {      // phase ghost planes from the ring neighbours
        SlabEmu &s = S[r], &up = S[(r + 1) % n], &dn = S[(r + n - 1) % n];
        memcpy(&s.phi[1 * s.g.pl], &dn.phi[(size_t)(dn.g.nz + 1) * s.g.pl], s.g.pl * sizeof(double));
        memcpy(&s.phi[(size_t)(s.g.nz + 2) * s.g.pl], &up.phi[2 * s.g.pl], s.g.pl * sizeof(double));
    }
    FieldsOut o{nullptr, nullptr, nullptr, nullptr};
    for (auto &s : S)
        for (int zs = 2; zs < s.g.nz + 2; zs++)
            for (int iy = 0; iy < s.g.ny; iy++)
                for (int ix = 0; ix < s.g.nx; ix++) {
                    const long long pp = 19 * s.g.pl;
                    const AaPlanes z{(zs - 1) * pp, zs * pp, (zs + 1) * pp};
                    if (fxy) aa_collide_thread<true, false, 0, ODD>(s.g, P, s.f.data(), s.gq.data(), s.phi.data(), o, zs, z, ix, iy);
                    else aa_collide_thread<false, false, 0, ODD>(s.g, P, s.f.data(), s.gq.data(), s.phi.data(), o, zs, z, ix, iy);
                }
    // population halos.  Snapshot semantics: every slab's outgoing rows are read before anybody's incoming rows land
    std::vector<SlabEmu> T = S;
    for (int r = 0; r < n; r++) {
        SlabEmu &s = S[r];
        const SlabEmu &up = T[(r + 1) % n], &dn = T[(r + n - 1) % n];
        const int nz = s.g.nz;
        for (int a = 0; a < 2; a++) {
            std::vector<double> &mine = a ? s.gq : s.f;
            const std::vector<double> &fu = a ? up.gq : up.f, &fd = a ? dn.gq : dn.f;
            if (!ODD) {   // after an even step: neighbours' boundary planes into my ghost planes
                copy_rows(mine, 1, fd, dn.g.nz + 1, s.g, true);        // lower neighbour's top plane, c_z = -1 slots
                copy_rows(mine, nz + 2, fu, 2, s.g, false);            // upper neighbour's bottom plane, c_z = +1 slots
            } else {      // after an odd step: neighbours' ghost planes into my boundary planes
                copy_rows(mine, 2, fd, dn.g.nz + 2, s.g, false, true);       // what the lower neighbour pushed up
                copy_rows(mine, nz + 1, fu, 1, s.g, true, true);             // what the upper neighbour pushed down
            }
        }
    }
}
}  // namespace
extern "C" long long emu3_aa_slabs_vs_whole(Emu *whole, int nslabs, int steps)
{
    const Geom &G = whole->g;
    const size_t pp = 19 * (size_t)G.pl;
    std::vector<SlabEmu> S(nslabs);
    const int zstep = G.nz / nslabs;
    for (int r = 0; r < nslabs; r++) {
        SlabEmu &s = S[r];
        s.z0 = r * zstep;
        s.g = G;
        s.g.nz = r == nslabs - 1 ? G.nz - s.z0 : zstep;
        s.g.zwrap = 0;
        s.f.assign((size_t)(s.g.nz + 4) * pp, 0.0);
        s.gq.assign((size_t)(s.g.nz + 4) * pp, 0.0);
        s.phi.assign((size_t)(s.g.nz + 4) * G.pl, 0.0);
        memcpy(&s.f[2 * pp], &whole->f[whole->cur][(size_t)(s.z0 + 2) * pp], (size_t)s.g.nz * pp * sizeof(double));
        memcpy(&s.gq[2 * pp], &whole->gq[whole->cur][(size_t)(s.z0 + 2) * pp], (size_t)s.g.nz * pp * sizeof(double));
    }
    whole->streamed = 4;
    bool odd = false;
    for (int t = 0; t < steps; t++) {
        if (odd) slab_step<true>(S, whole->P, whole->fxy);
        else slab_step<false>(S, whole->P, whole->fxy);
        odd = !odd;
    }
    emu3_step(whole, steps);
    // streamed order on both sides
    aa_normalize(whole);
    if (odd)
        for (auto &s : S)
            for (int zs = 2; zs < s.g.nz + 2; zs++)
                for (int iy = 0; iy < G.ny; iy++)
                    for (int ix = 0; ix < G.nx; ix++) {
                        const long long q = 19 * s.g.pl;
                        aa_swap_thread(s.g, s.f.data(), s.gq.data(), AaPlanes{(zs - 1) * q, zs * q, (zs + 1) * q}, ix, iy, zs == 2, zs == s.g.nz + 1);
                    }
    long long bad = 0;
    for (auto &s : S)
        for (int z = 0; z < s.g.nz; z++)
            for (int k = 0; k < 19; k++)
                for (int iy = 1; iy < G.ny - 1; iy++)
                    for (int ix = 1; ix < G.nx - 1; ix++) {
                        const size_t a = ((size_t)(z + 2) * 19 + k) * G.pl + (size_t)iy * G.nx + ix;
                        const size_t b = ((size_t)(s.z0 + z + 2) * 19 + k) * G.pl + (size_t)iy * G.nx + ix;
                        bad += s.f[a] != whole->f[whole->cur][b];
                        bad += s.gq[a] != whole->gq[whole->cur][b];
                    }
    return bad;
}

// ---- the plane-step arithmetic of k_aa_one (aa_schedule): number of rule violations over a whole launch ----
extern "C" int emu3_aa_schedule_violations(int zwrap, int nz, int L)
{
    std::vector<int> phase_step(nz, -1), collide_step(nz, -1);
    int bad = 0;
    const int steps = nz + L;
    for (int t = 0; t < steps; t++) {
        const AaSchedule s = aa_schedule(zwrap, nz, L, t);
        if (s.phase_plane >= 0) { bad += s.phase_plane >= nz || phase_step[s.phase_plane] != -1; phase_step[s.phase_plane] = t; }
        if (s.collide_plane >= 0) { bad += s.collide_plane >= nz || collide_step[s.collide_plane] != -1; collide_step[s.collide_plane] = t; }
    }
    for (int p = 0; p < nz; p++) {
        bad += phase_step[p] < 0 || collide_step[p] < 0;          // every plane summed once and collided once
        bad += !(phase_step[p] < collide_step[p]);                // its own phase part reads its pre-collision values
    }
    for (int t = 0; t < steps; t++) {
        const AaSchedule s = aa_schedule(zwrap, nz, L, t);
        if (s.collide_plane < 0) continue;
        for (int d = -1; d <= 1; d++) {                            // the three planes of the phase stencil
            int p = s.collide_plane + d;
            if (zwrap) p = (p + nz) % nz;
            if (p < 0 || p >= nz) continue;                        // a ghost plane of a slab: filled before the launch
            bool found = false;
            for (int j = 0; j < 3; j++) found = found || (s.dep[j] >= 0 && s.dep[j] == phase_step[p]);
            bad += !found;
            bad += !(phase_step[p] < t);                           // summed at an earlier plane-step
        }
        for (int j = 0; j < 3; j++) bad += s.dep[j] >= nz || (s.dep[j] >= 0 && !(s.dep[j] < t));
    }
    return bad;
}
