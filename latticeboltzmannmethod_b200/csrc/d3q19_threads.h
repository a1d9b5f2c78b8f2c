// d3q19_threads.h -- what ONE thread of each plain D3Q19 kernel does.
//
// The __global__ wrappers in d3q19_binary.cu only turn (blockIdx, threadIdx)
// into (storage plane, in-plane index) and call these functions; the CPU unit
// test (tests/emu/) calls the same functions from loops.  Everything here is
// plain C++ for that reason.
//
// Storage (this repo's layout, not the reference's AoS):
//   populations  [plane zs][direction k][in-plane index], plane pitch PL doubles,
//                in-plane index = iY*NX+iX, zs = z+2 for z in [-2, nz+1]
//   phase field  [plane zs][in-plane index]   (extended: bulk, mirror, wall nodes)
//   out fields   [z][in-plane index], z in [0, nz)
// A population array holds the STREAMED populations, i.e. exactly what the
// reference's f/g hold at the start of a step (lattice.h:35-38), for bulk nodes
// 1<=iX<=NX-2, 1<=iY<=NY-2.  Mirror nodes (iX=0 or iY=0) and wall nodes
// (iX=NX-1 or iY=NY-1) are not stored: the reference's mirror nodes are permuted
// copies of their inner neighbour (lattice.hpp:645-686) and its wall-node
// populations are never read.
#ifndef LBM_D3Q19_THREADS_H
#define LBM_D3Q19_THREADS_H

#include <type_traits>

#include "d3q19_node.h"
#include "lbm_pop.h"

namespace lbm3 {

struct Geom {
    int nx, ny, nz;        // local slab: nz owned planes
    int zwrap;             // in-place kernels only: 1 = the slab is the whole grid, z is periodic by addressing; 0 = the
                           // planes below / above the slab are ghost planes filled by the halo exchange
    long long pl;          // plane pitch in doubles (>= nx*ny, multiple of 16)
};

LBM_HD long long pop_index(const Geom &g, int zs, int k, long long idx) { return ((long long)zs * 19 + k) * g.pl + idx; }

// ---- phase field of one storage plane from its g populations -------------
// Reference: lattice.hpp:19-22 (bulk sum), :645-686 + :19-22 (mirror nodes sum
// the permuted copy), :75-106 (wall nodes: inner neighbour - phase_gradient).
// The thread of bulk node (ix,iy) also writes the mirror/wall nodes whose inner
// neighbour it is, so no second pass is needed.
// phase field of bulk node (ix,iy) of plane zs from its 19 g populations already in registers (G), incl. the
// mirror / wall nodes this node is the inner neighbour of
LBM_HD void phase_node(const Geom &g, const Params &P, const double *G, double *phi, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    const long long idx = (long long)iy * NX + ix;
    double *ph = phi + (long long)zs * g.pl;
    const double own = sum19(G);
    ph[idx] = own;
    const bool xlo = ix == 1, ylo = iy == 1, xhi = ix == NX - 2, yhi = iy == NY - 2;
    if (yhi) ph[idx + NX] = own - P.phase_gradient;
    if (xhi) ph[idx + 1] = own - P.phase_gradient;
    if (xhi && yhi) ph[idx + NX + 1] = own - P.phase_gradient;
    if (ylo) {
        const double m = sum19_mirror<2>(G);
        ph[idx - NX] = m;
        if (xhi) ph[idx - NX + 1] = m - P.phase_gradient;
    }
    if (xlo) {
        const double m = sum19_mirror<1>(G);
        ph[idx - 1] = m;
        if (yhi) ph[idx + NX - 1] = m - P.phase_gradient;
    }
    if (xlo && ylo) ph[idx - NX - 1] = sum19_mirror<3>(G);
}

LBM_HD void phase_thread(const Geom &g, const Params &P, const pop_t *gpop, double *phi, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) return;
    const long long idx = (long long)iy * NX + ix;
    double G[19];
#pragma unroll
    for (int k = 0; k < 19; k++) G[k] = gpop[pop_index(g, zs, k, idx)];
    phase_node(g, P, G, phi, zs, ix, iy);
}

// ---- push of one post-collision population -------------------------------
// Reference: lattice.hpp:491-499 (push), :503-585 (mirror nodes copy the inner
// node's post-collision populations permuted and push them), :588-643 (half-way
// bounce-back).  Folded: a population heading into a wall returns to its own
// node reversed; one heading into a mirror line is reflected (the mirrored
// velocity component flips, the node coordinate stays on the inner line).
// (Inlined on purpose: as a __noinline__ call it forces the population arrays of every thread into
// local memory -- measured 2x slower.)
LBM_HD void push_border(const Geom &g, pop_t *fout, pop_t *gout, int zs, int ix, int iy, int k, double fv, double gv)
{
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    const int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
    const int SYMX[19] = {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17};
    const int SYMY[19] = {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18};
    int px = ix + CX[k], py = iy + CY[k], pz = zs + CZ[k], kk = k;
    if (px == g.nx - 1 || py == g.ny - 1) {
        px = ix; py = iy; pz = zs; kk = OPP[k];
    } else {
        if (px == 0) { px = 1; kk = SYMX[kk]; }
        if (py == 0) { py = 1; kk = SYMY[kk]; }
    }
    const long long t = pop_index(g, pz, kk, (long long)py * g.nx + px);
    fout[t] = fv;
    gout[t] = gv;
}

struct FieldsOut {
    double *rho, *ux, *uy, *uz;   // [z][in-plane], may be null when WRITE is false
};

// mirror-node moments for the output fields: updateMacro (lattice.hpp:15-38) on
// the permuted populations, then the force shift of lattice.hpp:512-514 etc.
template <int MODE>
LBM_HD void mirror_fields(const double *F, const Params &P, const FieldsOut &o, long long at)
{
    const int SYM[3][19] = {{0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17},
                            {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18},
                            {0, 2, 1, 4, 3, 5, 6, 10, 9, 8, 7, 12, 11, 14, 13, 16, 15, 18, 17}};
    double Pm[19];
#pragma unroll
    for (int k = 0; k < 19; k++) Pm[k] = F[SYM[MODE - 1][k]];
    double d, ux, uy, uz;
    moments(Pm, d, ux, uy, uz);
    o.rho[at] = d;
    o.ux[at] = ux + P.fx / (2 * d);
    o.uy[at] = uy + P.fy / (2 * d);
    o.uz[at] = uz + P.fz / (2 * d);
}

// ---- one bulk node whose 19+19 populations and 19 neighbour phase values are already in registers ----
// PART 0: any bulk node; 1: a node at least two away from every mirror line / wall (no boundary code is
// even compiled in: straight-line kernel for the interior); 2: a node of the one-node rim (always the
// folded boundary push); 3: a node with 2 <= iY <= NY-3 and any bulk iX -- the x mirror line (iX = 1) and
// the x wall (iX = NX-2) are folded in as an address select on the ten directions with cx != 0, so that
// these two columns are stored by the same coalesced warps as their row neighbours.
struct PlainStore {   // how the interior push writes a value: plain stores (kernels may pass cache-hinted ones)
    LBM_HD void f(pop_t *p, double v) const { *p = (pop_t)v; }
    LBM_HD void g(pop_t *p, double v) const { *p = (pop_t)v; }
};
template <bool FXY, bool WRITE, int PART = 0, int COLL = 0, class ST = PlainStore>
LBM_HD void collide_push_node(const Geom &g, const Params &P, double *F, double *G, const double *PH, pop_t *fout,
                              pop_t *gout, const FieldsOut &o, int zs, int ix, int iy, const ST &st = ST())
{
    const int NX = g.nx, NY = g.ny;
    const long long idx = (long long)iy * NX + ix;
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    const bool own = zs >= 2 && zs < g.nz + 2;
    if (WRITE && own && PART != 1) {
        if (PART != 3 && iy == 1) mirror_fields<2>(F, P, o, (long long)(zs - 2) * g.pl + ix);
        if (ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * g.pl + (long long)iy * NX);
        if (PART != 3 && ix == 1 && iy == 1) mirror_fields<3>(F, P, o, (long long)(zs - 2) * g.pl);
    }
    double rho, ux, uy, uz;
    collide<FXY, COLL>(F, G, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
    if (WRITE && own) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
    }
    const bool border = PART == 2 || (PART == 0 && (ix == 1 || iy == 1 || ix == NX - 2 || iy == NY - 2));
    if (!border) {
        const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        const int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
        const int SYMX[19] = {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17};
        const bool xlo = PART == 3 && ix == 1, xhi = PART == 3 && ix == NX - 2;
#pragma unroll
        for (int k = 0; k < 19; k++) {
            long long t = pop_index(g, zs + CZ[k], k, idx + CY[k] * NX + CX[k]);
            if (PART == 3 && CX[k] == -1) {   // into the x mirror line: reflected, stays in column 1 (lattice.hpp:534-557)
                const long long tm = pop_index(g, zs + CZ[k], SYMX[k], idx + CY[k] * NX);
                t = xlo ? tm : t;
            }
            if (PART == 3 && CX[k] == 1) {    // into the x wall: back to its own node reversed (lattice.hpp:609-626)
                const long long tb = pop_index(g, zs, OPP[k], idx);
                t = xhi ? tb : t;
            }
            st.f(fout + t, F[k]);
            st.g(gout + t, G[k]);
        }
    } else {
        for (int k = 0; k < 19; k++) push_border(g, fout, gout, zs, ix, iy, k, F[k], G[k]);
    }
}

// ---- one bulk node: moments, stencils, collision, push -------------------
// Reference: Lattice::updateMacro + Lattice::collide_stream (lattice.hpp:4-108,
// 377-688) for this node.  `phi` must already hold the extended phase field of
// planes zs-1..zs+1.
template <bool FXY, bool WRITE>
LBM_HD void collide_push_thread(const Geom &g, const Params &P, const pop_t *fin, const pop_t *gin,
                                const double *phi, pop_t *fout, pop_t *gout, const FieldsOut &o, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE && zs >= 2 && zs < g.nz + 2 && ix < NX && iy < NY && (ix == NX - 1 || iy == NY - 1)) {
            // wall node fields, lattice.hpp:75-106 (phase is taken from the extended field by the caller)
            const long long at = (long long)(zs - 2) * g.pl + (long long)iy * NX + ix;
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
        return;
    }
    const long long idx = (long long)iy * NX + ix;
    double F[19], G[19], PH[19];
#pragma unroll
    for (int k = 0; k < 19; k++) {
        F[k] = fin[pop_index(g, zs, k, idx)];
        G[k] = gin[pop_index(g, zs, k, idx)];
    }
    {
        const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
        for (int k = 0; k < 19; k++) PH[k] = phi[(long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]];
    }
    collide_push_node<FXY, WRITE>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
}

// the default kernel pair as one per-node function (CPU unit test): rows iY = 1 and NY-2 take the rim
// path, every other bulk node the PART 3 path of k_collide_interior
template <bool FXY, bool WRITE>
LBM_HD void collide_push_thread_rows(const Geom &g, const Params &P, const pop_t *fin, const pop_t *gin,
                                     const double *phi, pop_t *fout, pop_t *gout, const FieldsOut &o, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE && zs >= 2 && zs < g.nz + 2 && ix < NX && iy < NY && (ix == NX - 1 || iy == NY - 1)) {
            const long long at = (long long)(zs - 2) * g.pl + (long long)iy * NX + ix;
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
        return;
    }
    const long long idx = (long long)iy * NX + ix;
    double F[19], G[19], PH[19];
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    for (int k = 0; k < 19; k++) {
        F[k] = fin[pop_index(g, zs, k, idx)];
        G[k] = gin[pop_index(g, zs, k, idx)];
        PH[k] = phi[(long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]];
    }
    if (iy == 1 || iy == NY - 2) collide_push_node<FXY, WRITE, 2>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
    else collide_push_node<FXY, WRITE, 3>(g, P, F, G, PH, fout, gout, o, zs, ix, iy);
}

// ---- the same node update through collide_streamed() ----------------------
// Push functor: stores one post-collision pair as soon as it exists.
struct PushInterior {
    const Geom &g;
    pop_t *fo, *go;     // fout/gout + pop_index(zs, 0, idx)
    template <int K>
    LBM_HD void operator()(double fv, double gv) const
    {
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        const long long off = (long long)(CZ[K] * 19 + K) * g.pl + (CY[K] * g.nx + CX[K]);
#if defined(__CUDA_ARCH__)
        __stcg(fo + off, fv);   // streaming stores: keep them out of L1
        __stcg(go + off, gv);
#else
        fo[off] = fv;
        go[off] = gv;
#endif
    }
};
struct PushBorder {
    const Geom &g;
    pop_t *fout, *gout;
    int zs, ix, iy;
    template <int K>
    LBM_HD void operator()(double fv, double gv) const { push_border(g, fout, gout, zs, ix, iy, K, fv, gv); }
};

// One bulk node whose f populations are in registers (F), whose g populations come from `getG`
// and whose 19 neighbour phase values are in PH.  Same reference lines as collide_push_thread.
template <bool FXY, bool WRITE, int PART = 0, class GetG>
LBM_HD void collide_push_node_streamed(const Geom &g, const Params &P, double *F, GetG getG, const double *PH, pop_t *fout,
                                       pop_t *gout, const FieldsOut &o, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    const long long idx = (long long)iy * NX + ix;
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    const bool own = zs >= 2 && zs < g.nz + 2;
    if (WRITE && own && PART != 1) {
        if (iy == 1) mirror_fields<2>(F, P, o, (long long)(zs - 2) * g.pl + ix);
        if (ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * g.pl + (long long)iy * NX);
        if (ix == 1 && iy == 1) mirror_fields<3>(F, P, o, (long long)(zs - 2) * g.pl);
    }
    double rho, ux, uy, uz;
    const bool border = PART == 2 || (PART == 0 && (ix == 1 || iy == 1 || ix == NX - 2 || iy == NY - 2));
    if (!border) {
        const long long base = pop_index(g, zs, 0, idx);
        collide_streamed<FXY>(F, getG, PushInterior{g, fout + base, gout + base}, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
    } else {
        collide_streamed<FXY>(F, getG, PushBorder{g, fout, gout, zs, ix, iy}, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
    }
    if (WRITE && own) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
    }
}

// plain-kernel flavour of the above (used by the CPU unit test to pin collide_streamed against collide)
template <bool FXY, bool WRITE>
LBM_HD void collide_push_thread_streamed(const Geom &g, const Params &P, const pop_t *fin, const pop_t *gin,
                                         const double *phi, pop_t *fout, pop_t *gout, const FieldsOut &o, int zs, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE && zs >= 2 && zs < g.nz + 2 && ix < NX && iy < NY && (ix == NX - 1 || iy == NY - 1)) {
            const long long at = (long long)(zs - 2) * g.pl + (long long)iy * NX + ix;
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
        return;
    }
    const long long idx = (long long)iy * NX + ix;
    double F[19], PH[19];
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    for (int k = 0; k < 19; k++) {
        F[k] = fin[pop_index(g, zs, k, idx)];
        PH[k] = phi[(long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]];
    }
    const pop_t *gp = gin + pop_index(g, zs, 0, idx);
    const long long pl = g.pl;
    collide_push_node_streamed<FXY, WRITE>(g, P, F, [gp, pl](int k) { return gp[(long long)k * pl]; }, PH, fout, gout, o, zs, ix, iy);
}

// ---- in-place (AA-pattern) streaming on ONE population array (kernel_variant 2) --------------
// Replaces the second population array and exchangeMatrices (lattice.hpp:842-855).
//
// The folded push above is a bijection on (bulk node, direction) with one more property: a value pushed back from
// where it landed, in the direction opposite to the slot it landed in, arrives at the node it came from -- in the
// opposite slot (plain push, bounce-back and both mirror lines alike).  Hence, with loc(x, m) the place the push of
// direction m from node x writes:
//   even step (array holds the streamed populations, slot k of node x = f_k(x)):
//       F[k] = A[x, k];            collide;   A[x, opp(m)] = F*[m]          -- nothing leaves the node
//   odd step  (array holds post-collision values in the opposite slot of their own node):
//       F[opp(m)] = A[loc(x, m)];  collide;   A[loc(x, m)] = F*[m]          -- lands where the two-array push puts it
// Every thread loads and stores the same 19 places, no two threads share one: no second array, no race.  With
// slot(x, m) = (x, opp(m)) in an even and loc(x, m) in an odd step both read F[opp(m)] = A[slot(x, m)] and write
// A[slot(x, m)] = F*[m].  The planes below and above come as explicit offsets: inside a slab that holds the whole grid
// z is periodic by addressing and no ghost planes of populations are needed; a slab of several uses its ghost planes
// as the places beyond its z faces (the halo exchange moves them, see aa_crossed_face below).
struct AaPlanes {
    long long zm, z0, zp;   // element offsets of plane z-1, z, z+1 (19 * pitch each) from the start of the array
};

// offset of the place the folded push of direction K from bulk node (ix, iy) writes (same rules as push_border)
template <int K>
LBM_HD long long aa_loc(const Geom &g, const AaPlanes &z, int ix, int iy)
{
    constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
    constexpr int SYMX[19] = {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17};
    constexpr int SYMY[19] = {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18};
    constexpr int cx = CX[K], cy = CY[K], cz = CZ[K];
    constexpr int KX = SYMX[K], KY = SYMY[K], KXY = SYMY[SYMX[K]], KO = OPP[K];
    const int idx = iy * g.nx + ix;
    const int pl = (int)g.pl;
    const bool wall = (cx == 1 && ix == g.nx - 2) || (cy == 1 && iy == g.ny - 2);
    const bool mx = cx == -1 && ix == 1, my = cy == -1 && iy == 1;
    const int kk = mx ? (my ? KXY : KX) : (my ? KY : K);
    const int d = (my ? 0 : cy * g.nx) + (mx ? 0 : cx);
    const long long zb = cz > 0 ? z.zp : (cz < 0 ? z.zm : z.z0);
    return wall ? z.z0 + (KO * pl + idx) : zb + (kk * pl + idx + d);
}
template <bool ODD, int K>
LBM_HD long long aa_slot(const Geom &g, const AaPlanes &z, int ix, int iy)
{
    constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
    if (ODD) return aa_loc<K>(g, z, ix, iy);
    return z.z0 + (OPP[K] * (int)g.pl + iy * g.nx + ix);
}
// fn(integral_constant<int, K>) for K = 0 .. 18, unrolled at compile time
template <int K = 0, class Fn>
LBM_HD void aa_each(Fn &&fn)
{
    fn(std::integral_constant<int, K>{});
    if constexpr (K < 18) aa_each<K + 1>(fn);
}
// the streamed populations of bulk node (ix, iy) in direction order, out of either half of the cycle
template <bool ODD>
LBM_HD void aa_load(const Geom &g, const AaPlanes &z, const pop_t *a, int ix, int iy, double *V)
{
    constexpr int OPP[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
    aa_each([&](auto m) {
        constexpr int M = decltype(m)::value;
        V[OPP[M]] = a[aa_slot<ODD, M>(g, z, ix, iy)];
    });
}

// phase field of one bulk node (+ the mirror / wall nodes it is the inner neighbour of) from the g array in either half
template <bool ODD>
LBM_HD void aa_phase_thread(const Geom &g, const Params &P, const pop_t *gq, double *phi, int zs, const AaPlanes &z, int ix, int iy)
{
    if (ix < 1 || ix > g.nx - 2 || iy < 1 || iy > g.ny - 2) return;
    double G[19];
    aa_load<ODD>(g, z, gq, ix, iy, G);
    phase_node(g, P, G, phi, zs, ix, iy);
}

// one node of one step, in place (same reference lines as collide_push_thread)
template <bool FXY, bool WRITE, int COLL, bool ODD>
LBM_HD void aa_collide_thread(const Geom &g, const Params &P, pop_t *f, pop_t *gq, const double *phi, const FieldsOut &o,
                              int zs, const AaPlanes &z, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) {
        if (WRITE && ix < NX && iy < NY && (ix == NX - 1 || iy == NY - 1)) {   // wall node fields, lattice.hpp:75-106
            const long long at = (long long)(zs - 2) * g.pl + (long long)iy * NX + ix;
            o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        }
        return;
    }
    const long long idx = (long long)iy * NX + ix;
    double F[19], G[19], PH[19];
    aa_load<ODD>(g, z, f, ix, iy, F);
    aa_load<ODD>(g, z, gq, ix, iy, G);
    {
        constexpr int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
#pragma unroll
        for (int k = 0; k < 19; k++) PH[k] = phi[(long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]];
    }
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    if (WRITE) {
        if (iy == 1) mirror_fields<2>(F, P, o, (long long)(zs - 2) * g.pl + ix);
        if (ix == 1) mirror_fields<1>(F, P, o, (long long)(zs - 2) * g.pl + (long long)iy * NX);
        if (ix == 1 && iy == 1) mirror_fields<3>(F, P, o, (long long)(zs - 2) * g.pl);
    }
    double rho, ux, uy, uz;
    collide<FXY, COLL>(F, G, PH[0], gx, gy, gz, lap, P, rho, ux, uy, uz);
    if (WRITE) {
        const long long at = (long long)(zs - 2) * g.pl + idx;
        o.rho[at] = rho; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
    }
    aa_each([&](auto m) {
        constexpr int M = decltype(m)::value;
        const long long t = aa_slot<ODD, M>(g, z, ix, iy);
        f[t] = (pop_t)F[M];
        gq[t] = (pop_t)G[M];
    });
}

// Between slabs: after an odd step the value in slot k (c_z != 0) of bulk node (ix, iy) of a slab's first or last plane
// has come across the z face -- pushed there by the neighbouring slab, straight or off a mirror line -- unless the
// in-plane position it would have come from is a wall node: then it is the node's own population, bounced back, and
// the halo exchange must leave it alone.
LBM_HD bool aa_crossed_face(const Geom &g, int k, int ix, int iy)
{
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY;
    return !((CX[k] == -1 && ix == g.nx - 2) || (CY[k] == -1 && iy == g.ny - 2));
}

// Plane-step arithmetic of the one-pass in-place kernel (k_aa_one, d3q19_inplace.cuh), host/device so that the CPU test
// can check it: at plane-step t a block sums the phase field of owned plane `phase_plane` and collides owned plane
// `collide_plane` (-1: none), whose phase stencil needs the phase parts of plane-steps dep[0..2] (-1: not summed in this
// launch -- a ghost plane of a slab, which the exchange has filled).  One slab holding the whole grid (zwrap): phase
// planes in the order nz-1, 0, 1, .., nz-2 (the plane below plane 0 first); a slab of several: 0 .. nz-1.  The
// collision trails by L >= 3 plane-steps, so every dependency is an earlier plane-step and a plane's own phase part
// (which reads the plane's pre-collision values) comes before its collision.
struct AaSchedule {
    int phase_plane, collide_plane, dep[3];
};
LBM_HD AaSchedule aa_schedule(int zwrap, int nz, int L, int t)
{
    AaSchedule s;
    s.phase_plane = t < nz ? (zwrap ? (t == 0 ? nz - 1 : t - 1) : t) : -1;
    const int c = t - L;
    s.collide_plane = c >= 0 && c < nz ? c : -1;
    for (int j = 0; j < 3; j++) {
        const int q = zwrap ? (c + j) % nz : c - 1 + j;     // plane-step at which plane c - 1 + j was summed
        s.dep[j] = s.collide_plane >= 0 && q >= 0 && q < nz ? q : -1;
    }
    return s;
}

// Turns an array of the odd half (post-collision values in the opposite slots) into the streamed populations and
// back, in place: the streamed value of (x, k) sits at loc(x, opp(k)), and loc(., opp(.)) is an involution on the
// places, so the two places of every pair swap contents.  The thread of the lower address does the swap.
// `ghost_below` / `ghost_above`: the plane below / above is a ghost plane of a slab (its places have no thread of their
// own in this launch), so this thread swaps with them whatever the addresses.
LBM_HD void aa_swap_thread(const Geom &g, pop_t *f, pop_t *gq, const AaPlanes &z, int ix, int iy, bool ghost_below = false,
                           bool ghost_above = false)
{
    if (ix < 1 || ix > g.nx - 2 || iy < 1 || iy > g.ny - 2) return;
    const long long pp = 19 * g.pl;
    aa_each([&](auto m) {
        constexpr int M = decltype(m)::value;
        const long long own = aa_slot<false, M>(g, z, ix, iy);   // (x, opp(M))
        const long long other = aa_slot<true, M>(g, z, ix, iy);  // loc(x, M): holds the streamed value of (x, opp(M))
        const bool in_ghost = (ghost_below && other >= z.zm && other < z.zm + pp) || (ghost_above && other >= z.zp && other < z.zp + pp);
        if (own < other || in_ghost) {
            const pop_t a = f[own], b = f[other];
            f[own] = b; f[other] = a;
            const pop_t c = gq[own], d = gq[other];
            gq[own] = d; gq[other] = c;
        }
    });
}

// ---- initialisation -------------------------------------------------------
// updateSymmetric + updateWall on the user's fields, one plane per call site
// thread-row; reference lattice.hpp:705-760 including the loops that stop one
// node short.  Sequential dependencies (a wall node reads
// a mirror node) are resolved by doing the work of one z plane in one thread.
LBM_HD void init_fields_plane(const Geom &g, const Params &P, double *phi, double *rho, int z)
{
    const int NX = g.nx, NY = g.ny;
    double *ph = phi + (long long)(z + 2) * g.pl;
    double *rh = rho + (long long)z * g.pl;
    for (int ix = 1; ix < NX - 2; ix++) { ph[ix] = ph[NX + ix]; rh[ix] = rh[NX + ix]; }
    for (int iy = 1; iy < NY - 2; iy++) { ph[(long long)iy * NX] = ph[(long long)iy * NX + 1]; rh[(long long)iy * NX] = rh[(long long)iy * NX + 1]; }
    ph[0] = ph[NX + 1];
    rh[0] = rh[NX + 1];
    for (int ix = 0; ix < NX; ix++) rh[(long long)(NY - 1) * NX + ix] = P.rho_wall;
    for (int iy = 0; iy < NY; iy++) rh[(long long)iy * NX + NX - 1] = P.rho_wall;
    for (int ix = 0; ix < NX - 2; ix++) ph[(long long)(NY - 1) * NX + ix] = ph[(long long)(NY - 2) * NX + ix] - P.phase_gradient;
    for (int iy = 0; iy < NY - 2; iy++) ph[(long long)iy * NX + NX - 1] = ph[(long long)iy * NX + NX - 2] - P.phase_gradient;
    ph[(long long)(NY - 1) * NX + NX - 1] = ph[(long long)(NY - 2) * NX + NX - 2] - P.phase_gradient;
}

// Lattice::init for one bulk node (lattice.hpp:216-276); fields [z][in-plane]
LBM_HD void init_pops_thread(const Geom &g, const Params &P, const double *phi, const double *rho, const double *ux,
                             const double *uy, const double *uz, pop_t *fout, pop_t *gout, int z, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) return;
    const int zs = z + 2;
    const long long idx = (long long)iy * NX + ix, at = (long long)z * g.pl + idx;
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    double PH[19], F[19], G[19];
    for (int k = 0; k < 19; k++) PH[k] = phi[(long long)(zs + CZ[k]) * g.pl + idx + CY[k] * NX + CX[k]];
    double gx, gy, gz, lap;
    stencils(PH, gx, gy, gz, lap);
    init_equilibrium(rho[at], PH[0], ux[at], uy[at], uz[at], gx, gy, gz, lap, P, F, G);
    for (int k = 0; k < 19; k++) {
        fout[pop_index(g, zs, k, idx)] = F[k];
        gout[pop_index(g, zs, k, idx)] = G[k];
    }
}

// updateMacro right after init (solver.hpp:877): moments without the force shift
LBM_HD void after_init_thread(const Geom &g, const Params &P, const pop_t *fin, const FieldsOut &o, int z, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix >= NX || iy >= NY) return;
    const long long base = (long long)z * g.pl;
    if (ix == NX - 1 || iy == NY - 1) {
        const long long at = base + (long long)iy * NX + ix;
        o.rho[at] = P.rho_wall; o.ux[at] = 0.0; o.uy[at] = 0.0; o.uz[at] = 0.0;
        return;
    }
    if (ix < 1 || iy < 1) return;
    const long long idx = (long long)iy * NX + ix;
    double F[19];
    for (int k = 0; k < 19; k++) F[k] = fin[pop_index(g, z + 2, k, idx)];
    double d, ux, uy, uz;
    moments(F, d, ux, uy, uz);
    o.rho[base + idx] = d; o.ux[base + idx] = ux; o.uy[base + idx] = uy; o.uz[base + idx] = uz;
    const int SYM[3][19] = {{0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17},
                            {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18},
                            {0, 2, 1, 4, 3, 5, 6, 10, 9, 8, 7, 12, 11, 14, 13, 16, 15, 18, 17}};
    for (int m = 0; m < 3; m++) {
        long long at;
        if (m == 0) { if (ix != 1) continue; at = base + (long long)iy * NX; }
        else if (m == 1) { if (iy != 1) continue; at = base + ix; }
        else { if (ix != 1 || iy != 1) continue; at = base; }
        double Pm[19];
        for (int k = 0; k < 19; k++) Pm[k] = F[SYM[m][k]];
        moments(Pm, d, ux, uy, uz);
        o.rho[at] = d; o.ux[at] = ux; o.uy[at] = uy; o.uz[at] = uz;
    }
}

// populations in the reference's AoS order for one node of plane z (any node):
// mirror nodes get the permuted copy, wall nodes zero
LBM_HD void gather_aos_thread(const Geom &g, const pop_t *pop, double *aos, int z, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix >= NX || iy >= NY) return;
    double *out = aos + (((long long)z * NY + iy) * NX + ix) * 19;
    if (ix == NX - 1 || iy == NY - 1) {
        for (int k = 0; k < 19; k++) out[k] = 0.0;
        return;
    }
    const int SYM[4][19] = {{0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18},
                            {0, 2, 1, 3, 4, 5, 6, 8, 7, 10, 9, 11, 12, 13, 14, 16, 15, 18, 17},
                            {0, 1, 2, 4, 3, 5, 6, 9, 10, 7, 8, 12, 11, 14, 13, 15, 16, 17, 18},
                            {0, 2, 1, 4, 3, 5, 6, 10, 9, 8, 7, 12, 11, 14, 13, 16, 15, 18, 17}};
    const int m = (ix == 0 ? 1 : 0) + (iy == 0 ? 2 : 0);
    const int sx = ix == 0 ? 1 : ix, sy = iy == 0 ? 1 : iy;
    const long long idx = (long long)sy * NX + sx;
    for (int k = 0; k < 19; k++) out[k] = pop[pop_index(g, z + 2, SYM[m][k], idx)];
}

LBM_HD void scatter_aos_thread(const Geom &g, const double *aos, pop_t *pop, int z, int ix, int iy)
{
    const int NX = g.nx, NY = g.ny;
    if (ix < 1 || ix > NX - 2 || iy < 1 || iy > NY - 2) return;
    const double *in = aos + (((long long)z * NY + iy) * NX + ix) * 19;
    const long long idx = (long long)iy * NX + ix;
    for (int k = 0; k < 19; k++) pop[pop_index(g, z + 2, k, idx)] = in[k];
}

}  // namespace lbm3
#endif
