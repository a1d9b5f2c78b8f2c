/* lbm_oracle.c -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE
 * (see lbm_oracle.h for who may load it and how it is pinned).
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (oracle/Makefile).  No FMA
 * contraction, no fast-math: every * and + below is one IEEE-754 binary64
 * operation in the order the reference's C++ expressions evaluate them.
 */
#include "lbm_oracle.h"
#include "lbm_exp.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ====================================================================== */
/* D3Q19 binary liquid: Microchannel3D/{Quarter,Full}Channel              */
/* ====================================================================== */

/* velocity set, descriptor.h:43-45 */
static const int C3X[19] = {0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1};
static const int C3Y[19] = {0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0};
static const int C3Z[19] = {0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1};
/* opposite direction, descriptor.h:98 */
static const int OPP3[19] = {0, 2, 1, 4, 3, 6, 5, 10, 9, 8, 7, 14, 13, 12, 11, 18, 17, 16, 15};
/* moment basis of d'Humieres et al., Phil. Trans. R. Soc. A 360 (2002), for D3Q19: row i as a polynomial of the
 * direction; used only by the MRT extension (no reference counterpart) */
static int mrt19_entry(int i, int cx, int cy, int cz)
{
    int c2 = cx * cx + cy * cy + cz * cz;
    switch (i) {
    case 0: return 1;
    case 1: return 19 * c2 - 30;
    case 2: return (21 * c2 * c2 - 53 * c2 + 24) / 2;
    case 3: return cx;
    case 4: return (5 * c2 - 9) * cx;
    case 5: return cy;
    case 6: return (5 * c2 - 9) * cy;
    case 7: return cz;
    case 8: return (5 * c2 - 9) * cz;
    case 9: return 3 * cx * cx - c2;
    case 10: return (3 * c2 - 5) * (3 * cx * cx - c2);
    case 11: return cy * cy - cz * cz;
    case 12: return (3 * c2 - 5) * (cy * cy - cz * cz);
    case 13: return cx * cy;
    case 14: return cy * cz;
    case 15: return cx * cz;
    case 16: return (cy * cy - cz * cz) * cx;
    case 17: return (cz * cz - cx * cx) * cy;
    default: return (cx * cx - cy * cy) * cz;
    }
}

static double W3[19];                                    /* descriptor.h:82-84 (3*w, rest 0) */
static double QXX[19], QYY[19], QZZ[19], QXY[19], QXZ[19], QYZ[19];  /* descriptor.h:47-61 */
static double WXX[19], WYY[19], WZZ[19], WXY[19], WYZ[19], WZX[19];  /* descriptor.h:63-81 */
static double GXS[19], GYS[19], GZS[19], LAPS[19];       /* descriptor.h:86-96 */
static int SYMX[19], SYMY[19], SYMXY[19];                /* lattice.h:58-81 */
static int tables3_ready = 0;

static int find_dir(int cx, int cy, int cz)
{
    for (int l = 1; l < 19; l++)
        if (C3X[l] == cx && C3Y[l] == cy && C3Z[l] == cz) return l;
    return 0;
}

static void build_tables3(void)
{
    if (tables3_ready) return;
    for (int k = 0; k < 19; k++) {
        int cx = C3X[k], cy = C3Y[k], cz = C3Z[k];
        int n = cx * cx + cy * cy + cz * cz;
        W3[k] = n == 0 ? 0.0 : (n == 1 ? 1.0 / 6.0 : 1.0 / 12.0);
        QXX[k] = cx ? 2.0 / 3.0 : -1.0 / 3.0;
        QYY[k] = cy ? 2.0 / 3.0 : -1.0 / 3.0;
        QZZ[k] = cz ? 2.0 / 3.0 : -1.0 / 3.0;
        QXY[k] = (double)(cx * cy);
        QXZ[k] = (double)(cx * cz);
        QYZ[k] = (double)(cy * cz);
        /* gradient-tensor weights: axis directions 5/12 along, -1/3 across;
         * diagonals -1/24 in-plane, 1/12 out-of-plane, +-1/4 cross */
        if (n == 0) {
            WXX[k] = WYY[k] = WZZ[k] = 0.0;
        } else if (n == 1) {
            WXX[k] = cx ? 5.0 / 12.0 : -1.0 / 3.0;
            WYY[k] = cy ? 5.0 / 12.0 : -1.0 / 3.0;
            WZZ[k] = cz ? 5.0 / 12.0 : -1.0 / 3.0;
        } else {
            WXX[k] = cx ? -1.0 / 24.0 : 1.0 / 12.0;
            WYY[k] = cy ? -1.0 / 24.0 : 1.0 / 12.0;
            WZZ[k] = cz ? -1.0 / 24.0 : 1.0 / 12.0;
        }
        WXY[k] = (cx * cy) * (1.0 / 4.0);
        WYZ[k] = (cy * cz) * (1.0 / 4.0);
        WZX[k] = (cz * cx) * (1.0 / 4.0);
        double gw = n == 1 ? 1.0 / 6.0 : (n == 2 ? 1.0 / 12.0 : 0.0);
        GXS[k] = cx * gw;
        GYS[k] = cy * gw;
        GZS[k] = cz * gw;
        LAPS[k] = n == 0 ? -4.0 : (n == 1 ? 1.0 / 3.0 : 1.0 / 6.0);
        SYMX[k] = find_dir(-cx, cy, cz);
        SYMY[k] = find_dir(cx, -cy, cz);
        SYMXY[k] = find_dir(-cx, -cy, cz);
    }
    tables3_ready = 1;
}

struct lbo3 {
    int nx, ny, nz;
    size_t pl;                 /* nodes per plane */
    /* A slab owns planes 0..nz-1 and keeps two ghost planes on each side
     * (z = -2,-1 and nz,nz+1; storage plane zs = z+2).  Ghost planes -1 and nz
     * hold the neighbour slab's complete boundary-plane state and are collided
     * redundantly, so that everything landing in planes 0..nz-1 is produced
     * locally and nothing received has to be merged; ghost planes -2 and nz+1
     * only feed the phase-field stencil of that redundant collision. */
    double *f, *g, *f2, *g2;   /* [(nz+4)][19][pl] */
    double *phi;               /* [(nz+4)][pl] extended phase field (bulk, mirror, wall, ghost planes) */
    double *rho, *ux, *uy, *uz;/* [nz][pl] */
    double aconst, kconst, gammaconst, phase_gradient, rho_wall, tau_phi, tau_liq, tau_gas, fx, fy, fz;
    double trt_magic;          /* 0: the reference's BGK; > 0: the two-relaxation-time EXTENSION (no reference counterpart) */
    double mrt[5];             /* all 0: off; else the multiple-relaxation-time EXTENSION (no reference counterpart) */
    int mrt_on;
};

static inline size_t pidx(const lbo3 *s, int zs, int k, size_t node)
{
    return ((size_t)zs * 19 + (size_t)k) * s->pl + node;
}

lbo3 *lbo3_create(int nx, int ny, int nz, const double *p)
{
    build_tables3();
    if (nx < 4 || ny < 4 || nz < 2) return NULL;
    lbo3 *s = (lbo3 *)calloc(1, sizeof *s);
    s->nx = nx; s->ny = ny; s->nz = nz;
    s->pl = (size_t)nx * ny;
    size_t npop = (size_t)(nz + 4) * 19 * s->pl;
    s->f = (double *)calloc(npop, sizeof(double));
    s->g = (double *)calloc(npop, sizeof(double));
    s->f2 = (double *)calloc(npop, sizeof(double));
    s->g2 = (double *)calloc(npop, sizeof(double));
    s->phi = (double *)calloc((size_t)(nz + 4) * s->pl, sizeof(double));
    s->rho = (double *)calloc((size_t)nz * s->pl, sizeof(double));
    s->ux = (double *)calloc((size_t)nz * s->pl, sizeof(double));
    s->uy = (double *)calloc((size_t)nz * s->pl, sizeof(double));
    s->uz = (double *)calloc((size_t)nz * s->pl, sizeof(double));
    s->aconst = p[LBO3_ACONST]; s->kconst = p[LBO3_KCONST]; s->gammaconst = p[LBO3_GAMMA];
    s->phase_gradient = p[LBO3_PHASE_GRADIENT]; s->rho_wall = p[LBO3_RHO_WALL];
    s->tau_phi = p[LBO3_TAU_PHI]; s->tau_liq = p[LBO3_TAU_LIQ]; s->tau_gas = p[LBO3_TAU_GAS];
    s->fx = p[LBO3_FORCE_X]; s->fy = p[LBO3_FORCE_Y]; s->fz = p[LBO3_FORCE_Z];
    return s;
}

void lbo3_destroy(lbo3 *s)
{
    if (!s) return;
    free(s->f); free(s->g); free(s->f2); free(s->g2); free(s->phi);
    free(s->rho); free(s->ux); free(s->uy); free(s->uz);
    free(s);
}

void lbo3_set_fields(lbo3 *s, const double *rho, const double *phase, const double *ux, const double *uy,
                     const double *uz)
{
    size_t n = (size_t)s->nz * s->pl;
    memcpy(s->rho, rho, n * sizeof(double));
    memcpy(s->phi + 2 * s->pl, phase, n * sizeof(double));
    memcpy(s->ux, ux, n * sizeof(double));
    memcpy(s->uy, uy, n * sizeof(double));
    memcpy(s->uz, uz, n * sizeof(double));
}

/* updateSymmetric (lattice.hpp:733-760) then updateWall (lattice.hpp:705-731),
 * including their loops that stop one node short. */
void lbo3_init_fields(lbo3 *s)
{
    const int NX = s->nx, NY = s->ny;
    for (int z = 0; z < s->nz; z++) {
        double *ph = s->phi + (size_t)(z + 2) * s->pl;
        double *rh = s->rho + (size_t)z * s->pl;
        for (int ix = 1; ix < NX - 2; ix++) { ph[ix] = ph[NX + ix]; rh[ix] = rh[NX + ix]; }
        for (int iy = 1; iy < NY - 2; iy++) { ph[iy * NX] = ph[iy * NX + 1]; rh[iy * NX] = rh[iy * NX + 1]; }
        ph[0] = ph[NX + 1]; rh[0] = rh[NX + 1];
        for (int ix = 0; ix < NX; ix++) rh[(NY - 1) * NX + ix] = s->rho_wall;
        for (int iy = 0; iy < NY; iy++) rh[iy * NX + NX - 1] = s->rho_wall;
        for (int ix = 0; ix < NX - 2; ix++) ph[(NY - 1) * NX + ix] = ph[(NY - 2) * NX + ix] - s->phase_gradient;
        for (int iy = 0; iy < NY - 2; iy++) ph[iy * NX + NX - 1] = ph[iy * NX + NX - 2] - s->phase_gradient;
        ph[(NY - 1) * NX + NX - 1] = ph[(NY - 2) * NX + NX - 2] - s->phase_gradient;
    }
}

/* the four 19-point stencils of lattice.hpp:392-409 / :226-243 at a bulk node */
static inline void stencils3(const lbo3 *s, int zs, int ix, int iy, double *gx, double *gy, double *gz, double *lap)
{
    double ax = 0.0, ay = 0.0, az = 0.0, al = 0.0;
    for (int k = 0; k < 19; k++) {
        double ph = s->phi[(size_t)(zs + C3Z[k]) * s->pl + (size_t)(iy + C3Y[k]) * s->nx + (ix + C3X[k])];
        ax += GXS[k] * ph;
        ay += GYS[k] * ph;
        az += GZS[k] * ph;
        al += LAPS[k] * ph;
    }
    *gx = ax; *gy = ay; *gz = az; *lap = al;
}

/* updateMacro (lattice.hpp:15-38) on 19 populations given in reference order */
static inline void moments3(const double *F, double *rho, double *ux, double *uy, double *uz)
{
    double d = F[0] + F[1] + F[2] + F[3] + F[4] + F[5] + F[6] + F[7] + F[8] + F[9] + F[10] + F[11] + F[12] +
               F[13] + F[14] + F[15] + F[16] + F[17] + F[18];
    double jx = F[1] - F[2] + F[7] - F[8] + F[9] - F[10] + F[15] - F[16] + F[17] - F[18];
    double jy = F[3] - F[4] + F[7] + F[8] - F[9] - F[10] + F[11] - F[12] + F[13] - F[14];
    double jz = F[5] - F[6] + F[11] + F[12] - F[13] - F[14] + F[15] + F[16] - F[17] - F[18];
    *rho = d; *ux = jx / d; *uy = jy / d; *uz = jz / d;
}
static inline double sum19(const double *G)
{
    return G[0] + G[1] + G[2] + G[3] + G[4] + G[5] + G[6] + G[7] + G[8] + G[9] + G[10] + G[11] + G[12] + G[13] +
           G[14] + G[15] + G[16] + G[17] + G[18];
}
static inline void gather19(const lbo3 *s, const double *arr, int zs, size_t node, double *out)
{
    for (int k = 0; k < 19; k++) out[k] = arr[pidx(s, zs, k, node)];
}

/* Lattice::init (lattice.hpp:212-276): equilibrium at every bulk node from the
 * prepared fields, written to the back buffers (exchange, then lbo3_finish).
 * Mirror and wall nodes are not stored here; what the reference writes into
 * them (lattice.hpp:278-372) is a permuted copy / junk that this restatement
 * derives on demand. */
void lbo3_init_pops(lbo3 *s)
{
    const int NX = s->nx, NY = s->ny;
    for (int z = 0; z < s->nz; z++)
        for (int iy = 1; iy < NY - 1; iy++)
            for (int ix = 1; ix < NX - 1; ix++) {
                size_t node = (size_t)iy * NX + ix;
                size_t fi = (size_t)z * s->pl + node;
                double gx, gy, gz, lap;
                stencils3(s, z + 2, ix, iy, &gx, &gy, &gz, &lap);
                double ph = s->phi[(size_t)(z + 2) * s->pl + node];
                double d = s->rho[fi], ux = s->ux[fi], uy = s->uy[fi], uz = s->uz[fi];
                double sum = 0.0, sum_phase = 0.0;
                double ps = ph * ph;
                double pb = d / 3.0 + s->aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - s->kconst * ph * lap;
                double mu = s->gammaconst * (s->aconst * (-ph + ph * ph * ph) - s->kconst * lap);
                for (int k = 1; k < 19; k++) {
                    int cx = C3X[k], cy = C3Y[k], cz = C3Z[k];
                    double feq = W3[k] * (pb + d * (cx * ux + cy * uy + cz * uz) +
                                          1.5 * d * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy +
                                                     (cz * cz - 1.0 / 3.0) * uz * uz +
                                                     2.0 * (ux * uy * cx * cy + ux * uz * cx * cz + uy * uz * cy * cz))) +
                                 s->kconst * (WXX[k] * gx * gx + WYY[k] * gy * gy + WZZ[k] * gz * gz + WXY[k] * gx * gy +
                                              WYZ[k] * gy * gz + WZX[k] * gx * gz);
                    double geq = W3[k] * (mu + ph * (cx * ux + cy * uy + cz * uz) +
                                          1.5 * ph * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy +
                                                      (cz * cz - 1.0 / 3.0) * uz * uz +
                                                      2.0 * (ux * uy * cx * cy + ux * uz * cx * cz + uy * uz * cy * cz)));
                    sum += feq;
                    sum_phase += geq;
                    s->f2[pidx(s, z + 2, k, node)] = feq;
                    s->g2[pidx(s, z + 2, k, node)] = geq;
                }
                s->f2[pidx(s, z + 2, 0, node)] = d - sum;
                s->g2[pidx(s, z + 2, 0, node)] = ph - sum_phase;
            }
}

/* phase field on every node of storage plane zs from the populations g of that
 * plane: bulk nodes sum their 19 populations (lattice.hpp:19-22); a mirror node
 * holds its inner neighbour's populations permuted (lattice.hpp:645-686), so its
 * sum runs over the permuted order; a wall node takes its inner neighbour's
 * value minus phase_gradient (lattice.hpp:75-106). */
static void phase_plane(lbo3 *s, int zs)
{
    const int NX = s->nx, NY = s->ny;
    double *ph = s->phi + (size_t)zs * s->pl;
    double G[19], P[19];
    for (int iy = 1; iy < NY - 1; iy++)
        for (int ix = 1; ix < NX - 1; ix++) {
            size_t node = (size_t)iy * NX + ix;
            gather19(s, s->g, zs, node, G);
            ph[node] = sum19(G);
            if (iy == 1) { for (int k = 0; k < 19; k++) P[k] = G[SYMY[k]]; ph[ix] = sum19(P); }
            if (ix == 1) { for (int k = 0; k < 19; k++) P[k] = G[SYMX[k]]; ph[iy * NX] = sum19(P); }
            if (ix == 1 && iy == 1) { for (int k = 0; k < 19; k++) P[k] = G[SYMXY[k]]; ph[0] = sum19(P); }
        }
    for (int ix = 0; ix < NX - 1; ix++) ph[(NY - 1) * NX + ix] = ph[(NY - 2) * NX + ix] - s->phase_gradient;
    for (int iy = 0; iy < NY - 1; iy++) ph[iy * NX + NX - 1] = ph[iy * NX + NX - 2] - s->phase_gradient;
    ph[(NY - 1) * NX + NX - 1] = ph[(NY - 2) * NX + NX - 2] - s->phase_gradient;
}

/* where a post-collision population of bulk node (ix,iy) in direction k lands:
 * the reference's push (lattice.hpp:491-499) with its mirror copy (:503-585)
 * and half-way bounce-back (:588-643) folded in.  Returns the direction index
 * to store under and the target node; *dz is the plane offset. */
static inline int push_target(const lbo3 *s, int ix, int iy, int k, int *tx, int *ty, int *dz)
{
    int px = ix + C3X[k], py = iy + C3Y[k];
    if (px == s->nx - 1 || py == s->ny - 1) { *tx = ix; *ty = iy; *dz = 0; return OPP3[k]; }
    int kk = k;
    if (px == 0) { px = 1; kk = SYMX[kk]; }
    if (py == 0) { py = 1; kk = SYMY[kk]; }
    *tx = px; *ty = py; *dz = C3Z[k];
    return kk;
}

void lbo3_collide(lbo3 *s, int write_fields)
{
    const int NX = s->nx, NY = s->ny;
    for (int zs = 0; zs < s->nz + 4; zs++) phase_plane(s, zs);

    const double omega_phi = 1.0 / s->tau_phi;
    double F[19], G[19], feq[19], geq[19], fp[19], P[19];
    for (int z = -1; z <= s->nz; z++) {
        int zs = z + 2;
        const int own = z >= 0 && z < s->nz;   /* ghost planes are collided too, but report no fields */
        for (int iy = 1; iy < NY - 1; iy++)
            for (int ix = 1; ix < NX - 1; ix++) {
                size_t node = (size_t)iy * NX + ix;
                gather19(s, s->f, zs, node, F);
                gather19(s, s->g, zs, node, G);
                double d, ux, uy, uz;
                moments3(F, &d, &ux, &uy, &uz);
                double ph = s->phi[(size_t)zs * s->pl + node];
                double gx, gy, gz, lap;
                stencils3(s, zs, ix, iy, &gx, &gy, &gz, &lap);
                /* lattice.hpp:412-414 */
                ux += s->fx / (2.0 * d);
                uy += s->fy / (2.0 * d);
                uz += s->fz / (2.0 * d);
                if (write_fields && own) {
                    size_t fi = (size_t)z * s->pl + node;
                    s->rho[fi] = d; s->ux[fi] = ux; s->uy[fi] = uy; s->uz[fi] = uz;
                    /* mirror nodes: moments of the permuted populations, then the same force shift
                     * (lattice.hpp:15-38 on lattice.hpp:645-686 data, then :512-514,:540-542,:564-566) */
                    for (int m = 0; m < 3; m++) {
                        const int *sym; size_t mnode;
                        if (m == 0) { if (iy != 1) continue; sym = SYMY; mnode = ix; }
                        else if (m == 1) { if (ix != 1) continue; sym = SYMX; mnode = (size_t)iy * NX; }
                        else { if (ix != 1 || iy != 1) continue; sym = SYMXY; mnode = 0; }
                        for (int k = 0; k < 19; k++) P[k] = F[sym[k]];
                        double md, mux, muy, muz;
                        moments3(P, &md, &mux, &muy, &muz);
                        size_t mi = (size_t)z * s->pl + mnode;
                        s->rho[mi] = md;
                        s->ux[mi] = mux + s->fx / (2 * md);
                        s->uy[mi] = muy + s->fy / (2 * md);
                        s->uz[mi] = muz + s->fz / (2 * md);
                    }
                }
                /* lattice.hpp:425-462 */
                double ps = ph * ph;
                double pb = d / 3.0 + s->aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - s->kconst * ph * lap;
                double mu = s->gammaconst * (s->aconst * (-ph + ph * ph * ph) - s->kconst * lap);
                double sum = 0.0, sum_phase = 0.0;
                double uxs = ux * ux, uys = uy * uy, uzs = uz * uz;
                double uxuy = ux * uy, uxuz = ux * uz, uyuz = uy * uz;
                double gxs = gx * gx, gys = gy * gy, gzs = gz * gz;
                double gxy = gx * gy, gxz = gx * gz, gyz = gy * gz;
                for (int k = 1; k < 19; k++) {
                    int cx = C3X[k], cy = C3Y[k], cz = C3Z[k];
                    feq[k] = W3[k] * (pb + d * (cx * ux + cy * uy + cz * uz) +
                                      1.5 * d * (QXX[k] * uxs + QYY[k] * uys + QZZ[k] * uzs +
                                                 2.0 * (uxuy * QXY[k] + uxuz * QXZ[k] + uyuz * QYZ[k]))) +
                             s->kconst * (WXX[k] * gxs + WYY[k] * gys + WZZ[k] * gzs + WXY[k] * gxy + WYZ[k] * gyz +
                                          WZX[k] * gxz);
                    geq[k] = W3[k] * (mu + ph * (cx * ux + cy * uy + cz * uz) +
                                      1.5 * ph * (QXX[k] * uxs + QYY[k] * uys + QZZ[k] * uzs +
                                                  2.0 * (uxuy * QXY[k] + uxuz * QXZ[k] + uyuz * QYZ[k])));
                    sum += feq[k];
                    sum_phase += geq[k];
                }
                feq[0] = d - sum;
                geq[0] = ph - sum_phase;
                /* lattice.hpp:465-480 */
                double tau = s->tau_gas + (ph + 1.0) / 2.0 * (s->tau_liq - s->tau_gas);
                double omega_rho = 1.0 / tau;
                double sum_force = 0.0;
                for (int k = 1; k < 19; k++) {
                    int cx = C3X[k], cy = C3Y[k], cz = C3Z[k];
                    fp[k] = (s->mrt_on ? 1.0 : 1.0 - 0.5 * omega_rho) * W3[k] *
                            (s->fx * ((cx - ux) + 3.0 * cx * (cx * ux + cy * uy + cz * uz)) +
                             s->fy * ((cy - uy) + 3.0 * cy * (cx * ux + cy * uy + cz * uz)) +
                             s->fz * ((cz - uz) + 3.0 * cz * (cx * ux + cy * uy + cz * uz)));
                    sum_force += fp[k];
                }
                fp[0] = -sum_force;
                /* extension (no reference counterpart): multiple relaxation times for f in the moment basis of
                 * d'Humieres et al. (2002), stress moments at 1/tau, Guo force populations with (1 - s_i/2) per moment */
                if (s->mrt_on) {
                    double se = s->mrt[0] > 0.0 ? s->mrt[0] : omega_rho, seps = s->mrt[1] > 0.0 ? s->mrt[1] : omega_rho;
                    double sq = s->mrt[2] > 0.0 ? s->mrt[2] : omega_rho, spi = s->mrt[3] > 0.0 ? s->mrt[3] : omega_rho;
                    double sm = s->mrt[4] > 0.0 ? s->mrt[4] : omega_rho;
                    double S[19] = {omega_rho, se, seps, omega_rho, sq, omega_rho, sq, omega_rho, sq, omega_rho, spi, omega_rho, spi,
                                    omega_rho, omega_rho, omega_rho, sm, sm, sm};
                    double fn[19], t[19];
                    for (int k = 0; k < 19; k++) fn[k] = F[k] - feq[k];
                    for (int i = 0; i < 19; i++) {
                        double m = 0.0, mf = 0.0;
                        int norm = 0;
                        for (int k = 0; k < 19; k++) {
                            int e = mrt19_entry(i, C3X[k], C3Y[k], C3Z[k]);
                            norm += e * e;
                            if (e != 0) { m += (double)e * fn[k]; mf += (double)e * fp[k]; }
                        }
                        double inv = 1.0 / (double)norm;
                        t[i] = mf * ((1.0 - 0.5 * S[i]) * inv) - m * (S[i] * inv);
                    }
                    for (int k = 0; k < 19; k++) {
                        double a = 0.0;
                        for (int i = 0; i < 19; i++) {
                            int e = mrt19_entry(i, C3X[k], C3Y[k], C3Z[k]);
                            if (e != 0) a += (double)e * t[i];
                        }
                        F[k] += a;
                    }
                }
                /* extension (no reference counterpart): two relaxation rates for f -- the symmetric half of every pair
                 * (k, opposite k) with omega+ = 1/tau, the antisymmetric half with omega- from the magic parameter */
                if (!s->mrt_on && s->trt_magic > 0.0) {
                    const double omega_minus = 1.0 / (s->trt_magic / (tau - 0.5) + 0.5);
                    const double fr = (1.0 - 0.5 * omega_minus) / (1.0 - 0.5 * omega_rho);
                    F[0] = (F[0] - omega_rho * (F[0] - feq[0])) + fp[0];
                    for (int k = 1; k < 19; k++) {
                        int kb = OPP3[k];
                        if (kb < k) continue;
                        double fpl = 0.5 * (F[k] + F[kb]), fmi = 0.5 * (F[k] - F[kb]);
                        double epl = 0.5 * (feq[k] + feq[kb]), emi = 0.5 * (feq[k] - feq[kb]);
                        double a = omega_rho * (fpl - epl), b = omega_minus * (fmi - emi);
                        /* force populations: symmetric half with (1 - omega/2) (already in fp), antisymmetric half
                         * with (1 - omega_minus/2): the first moment of a step's change is then exactly F */
                        double fsy = 0.5 * (fp[k] + fp[kb]), fan = fr * (0.5 * (fp[k] - fp[kb]));
                        F[k] = ((F[k] - a) - b) + (fsy + fan);
                        F[kb] = ((F[kb] - a) + b) + (fsy - fan);
                    }
                }
                /* lattice.hpp:483-488 then the push */
                for (int k = 0; k < 19; k++) {
                    if (s->trt_magic <= 0.0 && !s->mrt_on) F[k] += -omega_rho * (F[k] - feq[k]) + fp[k];
                    G[k] += -omega_phi * (G[k] - geq[k]);
                    int tx, ty, dz;
                    int kk = push_target(s, ix, iy, k, &tx, &ty, &dz);
                    size_t t = pidx(s, zs + dz, kk, (size_t)ty * NX + tx);
                    s->f2[t] = F[k];
                    s->g2[t] = G[k];
                }
            }
        if (write_fields && own) {
            /* wall nodes (lattice.hpp:75-106) */
            for (int iy = 0; iy < NY; iy++)
                for (int ix = 0; ix < NX; ix++)
                    if (ix == NX - 1 || iy == NY - 1) {
                        size_t fi = (size_t)z * s->pl + (size_t)iy * NX + ix;
                        s->rho[fi] = s->rho_wall; s->ux[fi] = 0.0; s->uy[fi] = 0.0; s->uz[fi] = 0.0;
                    }
        }
    }
}

void lbo3_finish(lbo3 *s)
{
    double *t = s->f; s->f = s->f2; s->f2 = t;
    t = s->g; s->g = s->g2; s->g2 = t;
}

double *lbo3_halo(lbo3 *s, int id, size_t *count)
{
    const size_t pl = s->pl;
    const int nz = s->nz;
#define AT(arr, z) ((arr) + (size_t)((z) + 2) * 19 * pl)
    switch (id) {
    case LBO3_S_F_UP:           *count = 19 * pl; return AT(s->f2, nz - 1);
    case LBO3_S_G_UP:           *count = 38 * pl; return AT(s->g2, nz - 2);
    case LBO3_S_F_DOWN:         *count = 19 * pl; return AT(s->f2, 0);
    case LBO3_S_G_DOWN:         *count = 38 * pl; return AT(s->g2, 0);
    case LBO3_R_F_FROM_BELOW:   *count = 19 * pl; return AT(s->f2, -1);
    case LBO3_R_G_FROM_BELOW:   *count = 38 * pl; return AT(s->g2, -2);
    case LBO3_R_F_FROM_ABOVE:   *count = 19 * pl; return AT(s->f2, nz);
    case LBO3_R_G_FROM_ABOVE:   *count = 38 * pl; return AT(s->g2, nz);
    case LBO3_S_PHI_UP:         *count = pl; return s->phi + (size_t)(nz - 1 + 2) * pl;
    case LBO3_S_PHI_DOWN:       *count = pl; return s->phi + (size_t)2 * pl;
    case LBO3_R_PHI_FROM_BELOW: *count = pl; return s->phi + (size_t)1 * pl;
    case LBO3_R_PHI_FROM_ABOVE: *count = pl; return s->phi + (size_t)(nz + 2) * pl;
    }
#undef AT
    *count = 0;
    return NULL;
}

static void self_copy(lbo3 *s, int send_id, int recv_id)
{
    size_t n, m;
    double *src = lbo3_halo(s, send_id, &n);
    double *dst = lbo3_halo(s, recv_id, &m);
    memmove(dst, src, n * sizeof(double));
}
void lbo3_self_exchange_phi(lbo3 *s)
{
    self_copy(s, LBO3_S_PHI_UP, LBO3_R_PHI_FROM_BELOW);
    self_copy(s, LBO3_S_PHI_DOWN, LBO3_R_PHI_FROM_ABOVE);
}
void lbo3_self_exchange_step(lbo3 *s)
{
    self_copy(s, LBO3_S_F_UP, LBO3_R_F_FROM_BELOW);
    self_copy(s, LBO3_S_G_UP, LBO3_R_G_FROM_BELOW);
    self_copy(s, LBO3_S_F_DOWN, LBO3_R_F_FROM_ABOVE);
    self_copy(s, LBO3_S_G_DOWN, LBO3_R_G_FROM_ABOVE);
}

/* fields as updateMacro leaves them right after init (solver.hpp:877): no force shift */
void lbo3_after_init(lbo3 *s)
{
    const int NX = s->nx, NY = s->ny;
    double F[19], P[19];
    for (int zs = 0; zs < s->nz + 4; zs++) phase_plane(s, zs);
    for (int z = 0; z < s->nz; z++) {
        for (int iy = 1; iy < NY - 1; iy++)
            for (int ix = 1; ix < NX - 1; ix++) {
                size_t node = (size_t)iy * NX + ix, fi = (size_t)z * s->pl + node;
                gather19(s, s->f, z + 2, node, F);
                moments3(F, &s->rho[fi], &s->ux[fi], &s->uy[fi], &s->uz[fi]);
                for (int m = 0; m < 3; m++) {
                    const int *sym; size_t mnode;
                    if (m == 0) { if (iy != 1) continue; sym = SYMY; mnode = ix; }
                    else if (m == 1) { if (ix != 1) continue; sym = SYMX; mnode = (size_t)iy * NX; }
                    else { if (ix != 1 || iy != 1) continue; sym = SYMXY; mnode = 0; }
                    for (int k = 0; k < 19; k++) P[k] = F[sym[k]];
                    size_t mi = (size_t)z * s->pl + mnode;
                    moments3(P, &s->rho[mi], &s->ux[mi], &s->uy[mi], &s->uz[mi]);
                }
            }
        for (int iy = 0; iy < NY; iy++)
            for (int ix = 0; ix < NX; ix++)
                if (ix == NX - 1 || iy == NY - 1) {
                    size_t fi = (size_t)z * s->pl + (size_t)iy * NX + ix;
                    s->rho[fi] = s->rho_wall; s->ux[fi] = 0.0; s->uy[fi] = 0.0; s->uz[fi] = 0.0;
                }
    }
}

void lbo3_init_single(lbo3 *s)
{
    lbo3_init_fields(s);
    lbo3_self_exchange_phi(s);
    lbo3_init_pops(s);
    lbo3_self_exchange_step(s);
    lbo3_finish(s);
    lbo3_after_init(s);
}

void lbo3_step_single(lbo3 *s, int n)
{
    for (int t = 0; t < n; t++) {
        lbo3_collide(s, t == n - 1);
        lbo3_self_exchange_step(s);
        lbo3_finish(s);
    }
}

void lbo3_get_fields(lbo3 *s, double *rho, double *phase, double *ux, double *uy, double *uz)
{
    size_t n = (size_t)s->nz * s->pl;
    memcpy(rho, s->rho, n * sizeof(double));
    memcpy(phase, s->phi + 2 * s->pl, n * sizeof(double));
    memcpy(ux, s->ux, n * sizeof(double));
    memcpy(uy, s->uy, n * sizeof(double));
    memcpy(uz, s->uz, n * sizeof(double));
}

void lbo3_get_pops(lbo3 *s, double *f, double *g)
{
    const int NX = s->nx, NY = s->ny;
    for (int z = 0; z < s->nz; z++)
        for (int iy = 0; iy < NY; iy++)
            for (int ix = 0; ix < NX; ix++) {
                size_t out = ((size_t)z * s->pl + (size_t)iy * NX + ix) * 19;
                if (ix == NX - 1 || iy == NY - 1) {
                    for (int k = 0; k < 19; k++) f[out + k] = g[out + k] = 0.0;
                    continue;
                }
                int sx = ix == 0 ? 1 : ix, sy = iy == 0 ? 1 : iy;
                const int *sym = (ix == 0 && iy == 0) ? SYMXY : (ix == 0 ? SYMX : (iy == 0 ? SYMY : NULL));
                size_t node = (size_t)sy * NX + sx;
                for (int k = 0; k < 19; k++) {
                    int kk = sym ? sym[k] : k;
                    f[out + k] = s->f[pidx(s, z + 2, kk, node)];
                    g[out + k] = s->g[pidx(s, z + 2, kk, node)];
                }
            }
}

void lbo3_set_trt(lbo3 *s, double magic) { s->trt_magic = magic; }
void lbo3_set_mrt(lbo3 *s, const double *rates5)
{
    s->mrt_on = 0;
    for (int i = 0; i < 5; i++) { s->mrt[i] = rates5[i]; if (rates5[i] != 0.0) s->mrt_on = 1; }
}

void lbo3_set_pops(lbo3 *s, const double *f, const double *g)
{
    const int NX = s->nx, NY = s->ny;
    for (int z = 0; z < s->nz; z++)
        for (int iy = 1; iy < NY - 1; iy++)
            for (int ix = 1; ix < NX - 1; ix++) {
                size_t node = (size_t)iy * NX + ix, in = ((size_t)z * s->pl + node) * 19;
                for (int k = 0; k < 19; k++) {
                    s->f[pidx(s, z + 2, k, node)] = f[in + k];
                    s->g[pidx(s, z + 2, k, node)] = g[in + k];
                }
            }
}

/* ====================================================================== */
/* D2Q9 tables shared by the three 2-D programs                            */
/* (shanchen.cpp:18-20, binary.cpp:44-47, mrt.cpp:63-67)                   */
/* ====================================================================== */
static const int C2X[9] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
static const int C2Y[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1};
static const int OPP2[9] = {0, 3, 4, 1, 2, 7, 8, 5, 6};
static const double W2[9] = {4.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0, 1.0 / 9.0,
                             1.0 / 36.0, 1.0 / 36.0, 1.0 / 36.0, 1.0 / 36.0};

/* ====================================================================== */
/* D2Q9 binary liquid: FreeEnergy/binary.cpp                               */
/* ====================================================================== */
/* float-typed tables, binary.cpp:48-59 */
static const float B_WXX[9] = {0.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0};
static const float B_WYY[9] = {0.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0};
static const float B_WXY[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 1.0 / 4.0, -1.0 / 4.0, 1.0 / 4.0, -1.0 / 4.0};
static const float B_GX[9] = {0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 0.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0, 1.0 / 12.0};
static const float B_GY[9] = {0.0, 0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 1.0 / 12.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0};
static const float B_LAP[9] = {-20.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0};

struct lbo2b {
    int nx, ny;
    size_t n;
    double *f, *g, *f2, *g2;         /* SoA [k][iX*NY+iY]; f,g streamed, f2,g2 scratch */
    double *phase, *rho, *ux, *uy;   /* [iX*NY+iY] */
    double aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi;
};

lbo2b *lbo2b_create(int nx, int ny, const double *p)
{
    if (nx < 3 || ny < 4) return NULL;
    lbo2b *s = (lbo2b *)calloc(1, sizeof *s);
    s->nx = nx; s->ny = ny; s->n = (size_t)nx * ny;
    s->f = (double *)calloc(9 * s->n, sizeof(double));
    s->g = (double *)calloc(9 * s->n, sizeof(double));
    s->f2 = (double *)calloc(9 * s->n, sizeof(double));
    s->g2 = (double *)calloc(9 * s->n, sizeof(double));
    s->phase = (double *)calloc(s->n, sizeof(double));
    s->rho = (double *)calloc(s->n, sizeof(double));
    s->ux = (double *)calloc(s->n, sizeof(double));
    s->uy = (double *)calloc(s->n, sizeof(double));
    s->aconst = p[0]; s->kconst = p[1]; s->gammaconst = p[2]; s->wall_gradient = p[3];
    s->tau_gas = p[4]; s->tau_liq = p[5]; s->omega_phi = p[6];
    return s;
}
void lbo2b_destroy(lbo2b *s)
{
    if (!s) return;
    free(s->f); free(s->g); free(s->f2); free(s->g2); free(s->phase); free(s->rho); free(s->ux); free(s->uy);
    free(s);
}
void lbo2b_set_phase(lbo2b *s, const double *ph) { memcpy(s->phase, ph, s->n * sizeof(double)); }

/* binary.cpp:126-132 */
void lbo2b_init_reference_ic(lbo2b *s)
{
    for (int ix = 0; ix < s->nx; ix++)
        for (int iy = 0; iy < s->ny; iy++)
            s->phase[(size_t)ix * s->ny + iy] = (abs(ix - s->nx / 2) <= 20 && iy < 40 && iy >= 1) ? 1.0 : -1.0;
}

static inline void stencils2(const lbo2b *s, int ix, int iy, double *gx, double *gy, double *lap)
{
    double al = 0.0, ax = 0.0, ay = 0.0;
    for (int k = 0; k < 9; k++) {
        int x2 = (ix + C2X[k] + s->nx) % s->nx, y2 = (iy + C2Y[k] + s->ny) % s->ny;
        double ph = s->phase[(size_t)x2 * s->ny + y2];
        al += B_LAP[k] * ph;
        ax += B_GX[k] * ph;
        ay += B_GY[k] * ph;
    }
    *gx = ax; *gy = ay; *lap = al;
}

/* equilibria of binary.cpp:298-313 (and :176-195, the same expression) */
static inline void equilibria2(const lbo2b *s, double d, double ph, double ux, double uy, double gx, double gy,
                               double lap, double *feq, double *geq)
{
    double sum = 0.0, sum_phase = 0.0;
    double ps = ph * ph;
    double pb = d / 3.0 + s->aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - s->kconst * ph * lap;
    double mu = s->gammaconst * (s->aconst * (-ph + ph * ph * ph) - s->kconst * lap);
    for (int k = 1; k < 9; k++) {
        int cx = C2X[k], cy = C2Y[k];
        feq[k] = W2[k] * (3.0 * pb + 3.0 * d * (cx * ux + cy * uy) +
                          4.5 * d * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy +
                                     2.0 * ux * uy * cx * cy)) +
                 s->kconst * (B_WXX[k] * gx * gx + B_WYY[k] * gy * gy + B_WXY[k] * gx * gy);
        geq[k] = W2[k] * (3.0 * mu + 3.0 * ph * (cx * ux + cy * uy) +
                          4.5 * ph * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy +
                                      2.0 * ux * uy * cx * cy));
        sum += feq[k];
        sum_phase += geq[k];
    }
    feq[0] = d - sum;
    geq[0] = ph - sum_phase;
}

/* binary.cpp:136-196: wall phase rows, then equilibrium on EVERY node (walls included) */
void lbo2b_init(lbo2b *s)
{
    const int NX = s->nx, NY = s->ny;
    for (int ix = 0; ix < NX; ix++) {
        s->phase[(size_t)ix * NY] = s->phase[(size_t)ix * NY + 1] - s->wall_gradient;
        s->phase[(size_t)ix * NY + NY - 1] = s->phase[(size_t)ix * NY + NY - 2] - s->wall_gradient;
    }
    double feq[9], geq[9];
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 0; iy < NY; iy++) {
            size_t i = (size_t)ix * NY + iy;
            double gx, gy, lap;
            stencils2(s, ix, iy, &gx, &gy, &lap);
            s->rho[i] = 1.0; s->ux[i] = 0.0; s->uy[i] = 0.0;
            equilibria2(s, 1.0, s->phase[i], 0.0, 0.0, gx, gy, lap, feq, geq);
            for (int k = 0; k < 9; k++) { s->f[k * s->n + i] = feq[k]; s->g[k * s->n + i] = geq[k]; }
        }
}

static void step2b(lbo2b *s)
{
    const int NX = s->nx, NY = s->ny;
    const size_t n = s->n;
    /* binary.cpp:239-250 */
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 1; iy < NY - 1; iy++) {
            size_t i = (size_t)ix * NY + iy;
            double ph = 0.0;
            for (int k = 0; k < 9; k++) ph += s->g[k * n + i];
            s->phase[i] = ph;
        }
    for (int ix = 0; ix < NX; ix++) {
        s->phase[(size_t)ix * NY] = s->phase[(size_t)ix * NY + 1] - s->wall_gradient;
        s->phase[(size_t)ix * NY + NY - 1] = s->phase[(size_t)ix * NY + NY - 2] - s->wall_gradient;
    }
    /* binary.cpp:252-327 collide, then :201-233 bounce-back and :341-350 streaming as one push:
     * a population that would enter a wall row returns to its own node reversed */
    double feq[9], geq[9];
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 1; iy < NY - 1; iy++) {
            size_t i = (size_t)ix * NY + iy;
            double d = 0.0, jx = 0.0, jy = 0.0;
            for (int k = 0; k < 9; k++) {
                double fk = s->f[k * n + i];
                d += fk;
                jx += fk * C2X[k];
                jy += fk * C2Y[k];
            }
            double ux = jx / d, uy = jy / d;
            s->rho[i] = d; s->ux[i] = ux; s->uy[i] = uy;
            double gx, gy, lap;
            stencils2(s, ix, iy, &gx, &gy, &lap);
            double ph = s->phase[i];
            equilibria2(s, d, ph, ux, uy, gx, gy, lap, feq, geq);
            double tau_rho = s->tau_gas + (ph + 1.0) / 2.0 * (s->tau_liq - s->tau_gas);
            double omega_rho = 1.0 / tau_rho;
            for (int k = 0; k < 9; k++) {
                double fo = s->f[k * n + i] * (1.0 - omega_rho) + omega_rho * feq[k];
                double go = s->g[k * n + i] * (1.0 - s->omega_phi) + s->omega_phi * geq[k];
                int ty = iy + C2Y[k];
                if (ty == 0 || ty == NY - 1) {
                    s->f2[OPP2[k] * n + i] = fo;
                    s->g2[OPP2[k] * n + i] = go;
                } else {
                    size_t t = (size_t)((ix + C2X[k] + NX) % NX) * NY + ty;
                    s->f2[k * n + t] = fo;
                    s->g2[k * n + t] = go;
                }
            }
        }
    /* wall rows: populations stay what init() left there (the reference never streams into
     * f[][0] / f[][NY-1]); rho=1, u=0 (binary.cpp:227-230) */
    for (int ix = 0; ix < NX; ix++) {
        size_t b = (size_t)ix * NY, t = b + NY - 1;
        for (int k = 0; k < 9; k++) {
            s->f2[k * n + b] = s->f[k * n + b]; s->g2[k * n + b] = s->g[k * n + b];
            s->f2[k * n + t] = s->f[k * n + t]; s->g2[k * n + t] = s->g[k * n + t];
        }
        s->rho[b] = 1.0; s->rho[t] = 1.0; s->ux[b] = 0.0; s->ux[t] = 0.0; s->uy[b] = 0.0; s->uy[t] = 0.0;
    }
    double *tmp = s->f; s->f = s->f2; s->f2 = tmp;
    tmp = s->g; s->g = s->g2; s->g2 = tmp;
}
void lbo2b_step(lbo2b *s, int nsteps) { for (int t = 0; t < nsteps; t++) step2b(s); }
void lbo2b_get_fields(lbo2b *s, double *phase, double *rho, double *ux, double *uy)
{
    memcpy(phase, s->phase, s->n * sizeof(double));
    memcpy(rho, s->rho, s->n * sizeof(double));
    memcpy(ux, s->ux, s->n * sizeof(double));
    memcpy(uy, s->uy, s->n * sizeof(double));
}
void lbo2b_get_pops(lbo2b *s, double *f, double *g)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) { f[i * 9 + k] = s->f[k * s->n + i]; g[i * 9 + k] = s->g[k * s->n + i]; }
}
void lbo2b_set_pops(lbo2b *s, const double *f, const double *g)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) { s->f[k * s->n + i] = f[i * 9 + k]; s->g[k * s->n + i] = g[i * 9 + k]; }
}

/* ====================================================================== */
/* D2Q9 Shan-Chen: ShanChen/shanchen.cpp                                   */
/* ====================================================================== */
struct lbosc {
    int nx, ny;
    size_t n;
    double tau, g;
    int use_libm;
    double *f, *f2, *rho, *psi, *u1, *u2;   /* f SoA [k][iY*nx+iX] */
};
lbosc *lbosc_create(int nx, int ny, double tau, double g, int use_libm)
{
    if (nx < 3 || ny < 3) return NULL;
    lbosc *s = (lbosc *)calloc(1, sizeof *s);
    s->nx = nx; s->ny = ny; s->n = (size_t)nx * ny; s->tau = tau; s->g = g; s->use_libm = use_libm;
    s->f = (double *)calloc(9 * s->n, sizeof(double));
    s->f2 = (double *)calloc(9 * s->n, sizeof(double));
    s->rho = (double *)calloc(s->n, sizeof(double));
    s->psi = (double *)calloc(s->n, sizeof(double));
    s->u1 = (double *)calloc(s->n, sizeof(double));
    s->u2 = (double *)calloc(s->n, sizeof(double));
    return s;
}
void lbosc_destroy(lbosc *s)
{
    if (!s) return;
    free(s->f); free(s->f2); free(s->rho); free(s->psi); free(s->u1); free(s->u2); free(s);
}
/* closed-form equilibria, shanchen.cpp:101-109 / :169-177 */
static inline void sc_feq(double dense, double v1, double v2, double *feq)
{
    float usq = v1 * v1 + v2 * v2;
    feq[0] = 4.0 / 9.0 * dense * (1.0 - 1.5 * usq);
    feq[1] = 1.0 / 9.0 * dense * (1.0 + 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[2] = 1.0 / 9.0 * dense * (1.0 + 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[3] = 1.0 / 9.0 * dense * (1.0 - 3 * v1 + 4.5 * v1 * v1 - 1.5 * usq);
    feq[4] = 1.0 / 9.0 * dense * (1.0 - 3 * v2 + 4.5 * v2 * v2 - 1.5 * usq);
    feq[5] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 + v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[6] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 + v2) + 4.5 * (-v1 + v2) * (-v1 + v2) - 1.5 * usq);
    feq[7] = 1.0 / 36.0 * dense * (1.0 + 3 * (-v1 - v2) + 4.5 * (v1 + v2) * (v1 + v2) - 1.5 * usq);
    feq[8] = 1.0 / 36.0 * dense * (1.0 + 3 * (v1 - v2) + 4.5 * (v1 - v2) * (v1 - v2) - 1.5 * usq);
}
void lbosc_init_droplet(lbosc *s, double rhol, double rhog, int radius)
{
    const int nx = s->nx, ny = s->ny;
    double feq[9];
    for (int i = 0; i < nx * ny; i++) {
        double d;
        if ((i / nx - ny / 2.0) * (i / nx - ny / 2.0) + (i % nx - nx / 2.0) * (i % nx - nx / 2.0) <= radius * radius)
            d = rhol;
        else
            d = rhog;
        s->rho[i] = d; s->u1[i] = 0.0; s->u2[i] = 0.0;
        sc_feq(d, 0.0, 0.0, feq);
        for (int k = 0; k < 9; k++) s->f[k * s->n + i] = feq[k];
    }
}
void lbosc_set_pops(lbosc *s, const double *f)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) s->f[k * s->n + i] = f[i * 9 + k];
}
static void step_sc(lbosc *s)
{
    const int nx = s->nx, ny = s->ny;
    const size_t n = s->n;
    /* shanchen.cpp:124-132, plus psi = 1-exp(-rho) once per node (the reference recomputes the same
     * function of the same rho for every neighbour, :152-157) */
    for (size_t i = 0; i < n; i++) {
        double r = 0;
        for (int k = 0; k < 9; k++) r += s->f[k * n + i];
        s->rho[i] = r;
        s->psi[i] = 1.0 - (s->use_libm ? exp(-r) : lbm_exp(-r));
    }
    double feq[9], fpop[9];
    for (int iy = 0; iy < ny; iy++)
        for (int ix = 0; ix < nx; ix++) {
            size_t i = (size_t)iy * nx + ix;
            double dense = s->rho[i], v1, v2;
            float fx = 0.0, fy = 0.0;
            for (int k = 0; k < 9; k++) {
                int x2 = (ix + C2X[k] + nx) % nx, y2 = (iy + C2Y[k] + ny) % ny;
                fx += W2[k] * C2X[k] * s->psi[(size_t)nx * y2 + x2];
                fy += W2[k] * C2Y[k] * s->psi[(size_t)nx * y2 + x2];
            }
            fx = -s->g * s->psi[i] * fx;
            fy = -s->g * s->psi[i] * fy;
            const double *F = s->f;
            v1 = (F[1 * n + i] - F[3 * n + i] + F[5 * n + i] - F[6 * n + i] - F[7 * n + i] + F[8 * n + i]) / dense + fx / (2.0 * dense);
            v2 = (F[2 * n + i] - F[4 * n + i] + F[5 * n + i] + F[6 * n + i] - F[7 * n + i] - F[8 * n + i]) / dense + fy / (2.0 * dense);
            s->u1[i] = v1; s->u2[i] = v2;
            for (int k = 0; k < 9; k++) {
                int cx = C2X[k], cy = C2Y[k];
                fpop[k] = W2[k] * (1 - 0.5 / s->tau) * ((3 * (cx - v1) + 9 * cx * (cx * v1 + cy * v2)) * fx +
                                                         (3 * (cy - v2) + 9 * cy * (cx * v1 + cy * v2)) * fy);
            }
            sc_feq(dense, v1, v2, feq);
            for (int k = 0; k < 9; k++) {
                int x2 = (ix + C2X[k] + nx) % nx, y2 = (iy + C2Y[k] + ny) % ny;
                double fk = s->f[k * n + i];
                fk += -1.0 / s->tau * (fk - feq[k]) + fpop[k];
                s->f2[k * n + (size_t)nx * y2 + x2] = fk;
            }
        }
    double *t = s->f; s->f = s->f2; s->f2 = t;
}
void lbosc_step(lbosc *s, int nsteps) { for (int t = 0; t < nsteps; t++) step_sc(s); }
void lbosc_get_fields(lbosc *s, double *rho, double *u1, double *u2)
{
    memcpy(rho, s->rho, s->n * sizeof(double));
    memcpy(u1, s->u1, s->n * sizeof(double));
    memcpy(u2, s->u2, s->n * sizeof(double));
}
void lbosc_get_pops(lbosc *s, double *f)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) f[i * 9 + k] = s->f[k * s->n + i];
}

/* ====================================================================== */
/* D2Q9 MRT: MRT/mrt.cpp (Gram-Schmidt, forced) and MRT/another_mrt.cpp    */
/* ====================================================================== */
static const double PI_REF = 3.141592653589793238462643383279502884197;  /* mrt.cpp:12 */

struct lbomrt {
    int variant, nx, ny;
    size_t n;
    double omega, omega_bulk, force_ampl;
    double omegamat[9], M[9][9], invM[9][9];
    double *f, *f2, *rho, *ux, *uy;   /* SoA [k][iX*NY+iY] */
};

lbomrt *lbomrt_create(int variant, int nx, int ny, double omega, double omega_bulk, double force_ampl)
{
    if (nx < 3 || ny < 3) return NULL;
    lbomrt *s = (lbomrt *)calloc(1, sizeof *s);
    s->variant = variant; s->nx = nx; s->ny = ny; s->n = (size_t)nx * ny;
    s->omega = omega; s->omega_bulk = omega_bulk; s->force_ampl = force_ampl;
    s->f = (double *)calloc(9 * s->n, sizeof(double));
    s->f2 = (double *)calloc(9 * s->n, sizeof(double));
    s->rho = (double *)calloc(s->n, sizeof(double));
    s->ux = (double *)calloc(s->n, sizeof(double));
    s->uy = (double *)calloc(s->n, sizeof(double));
    if (variant == 0) {
        /* mrt.cpp:60-92 */
        double om[9] = {1.0, omega_bulk, omega, 1.0, omega, 1.0, omega, omega, omega};
        double wm[9] = {1.0 / 9.0, 1.0 / 36.0, 1.0 / 36.0, 1.0 / 6.0, 1.0 / 12.0, 1.0 / 6.0, 1.0 / 12.0, 1.0 / 4.0, 1.0 / 4.0};
        memcpy(s->omegamat, om, sizeof om);
        for (int i = 0; i < 9; i++) {
            int cx = C2X[i], cy = C2Y[i];
            s->M[0][i] = 1.0;
            s->M[1][i] = -4.0 + 3.0 * (cx * cx + cy * cy);
            s->M[2][i] = 4.0 - 21.0 / 2.0 * (cx * cx + cy * cy) + 9.0 / 2.0 * (cx * cx + cy * cy) * (cx * cx + cy * cy);
            s->M[3][i] = cx;
            s->M[4][i] = (-5.0 + 3.0 * (cx * cx + cy * cy)) * cx;
            s->M[5][i] = cy;
            s->M[6][i] = (-5.0 + 3.0 * (cx * cx + cy * cy)) * cy;
            s->M[7][i] = cx * cx - cy * cy;
            s->M[8][i] = cx * cy;
            s->invM[i][0] = wm[0] * 1.0;
            s->invM[i][1] = wm[1] * (-4.0 + 3.0 * (cx * cx + cy * cy));
            s->invM[i][2] = wm[2] * (4.0 - 21.0 / 2.0 * (cx * cx + cy * cy) + 9.0 / 2.0 * (cx * cx + cy * cy) * (cx * cx + cy * cy));
            s->invM[i][3] = wm[3] * cx;
            s->invM[i][4] = wm[4] * (-5.0 + 3.0 * (cx * cx + cy * cy)) * cx;
            s->invM[i][5] = wm[5] * cy;
            s->invM[i][6] = wm[6] * (-5.0 + 3.0 * (cx * cx + cy * cy)) * cy;
            s->invM[i][7] = wm[7] * (cx * cx - cy * cy);
            s->invM[i][8] = wm[8] * cx * cy;
        }
    } else {
        /* another_mrt.cpp:66-84 */
        double om[9] = {1.0, 1.0, 1.0, omega, omega_bulk, omega_bulk, omega_bulk, omega_bulk, omega_bulk};
        double gg[9] = {1.0, -2.0, -2.0, -2.0, -2.0, 4.0, 4.0, 4.0, 4.0};
        memcpy(s->omegamat, om, sizeof om);
        for (int i = 0; i < 9; i++) {
            int cx = C2X[i], cy = C2Y[i];
            s->M[0][i] = 1.0;
            s->M[1][i] = cx * sqrt(3.0);
            s->M[2][i] = cy * sqrt(3.0);
            s->M[3][i] = (cx * cx - 1.0 / 3.0) * 3.0 / sqrt(2.0);
            s->M[4][i] = cx * cy * 3.0;
            s->M[5][i] = (cy * cy - 1.0 / 3.0) * 3.0 / sqrt(2.0);
            s->M[6][i] = gg[i] / 2.0;
            s->M[7][i] = gg[i] * cx * sqrt(1.5) / 2.0;
            s->M[8][i] = gg[i] * cy * sqrt(1.5) / 2.0;
        }
    }
    return s;
}
void lbomrt_destroy(lbomrt *s)
{
    if (!s) return;
    free(s->f); free(s->f2); free(s->rho); free(s->ux); free(s->uy); free(s);
}
/* mrt.cpp:97-120 / another_mrt.cpp:89-110 */
void lbomrt_init_wave(lbomrt *s, double ampl)
{
    const int NX = s->nx, NY = s->ny;
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 0; iy < NY; iy++) {
            size_t i = (size_t)ix * NY + iy;
            double rho, ux, uy = 0.0;
            if (s->variant == 0) {
                double forcex = s->force_ampl * cos((double)ix / (double)NX * 2.0 * PI_REF);
                rho = 1.0;
                ux = ampl * cos((double)ix / (double)NX * 2.0 * PI_REF) - forcex / 2.0;
            } else {
                double dense = 0.0;
                rho = 1.0 + dense;
                ux = ampl * cos((double)ix / (double)NX * 2 * PI_REF);
            }
            s->rho[i] = rho; s->ux[i] = ux; s->uy[i] = uy;
            double fluxx = 3.0 * ux, fluxy = 3.0 * uy;
            double qxx = 4.5 * ux * ux, qxy = 9.0 * ux * uy, qyy = 4.5 * uy * uy;
            for (int k = 0; k < 9; k++) {
                int cx = C2X[k], cy = C2Y[k];
                s->f[k * s->n + i] = rho * W2[k] * (1.0 + fluxx * cx + fluxy * cy + qxx * (cx * cx - 1.0 / 3.0) +
                                                    qxy * cx * cy + qyy * (cy * cy - 1.0 / 3.0));
            }
        }
}
void lbomrt_set_pops(lbomrt *s, const double *f)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) s->f[k * s->n + i] = f[i * 9 + k];
}
static void step_mrt(lbomrt *s)
{
    const int NX = s->nx, NY = s->ny;
    const size_t n = s->n;
    double fl[9], eq[9], add[9], momforce[9];
    for (int ix = 0; ix < NX; ix++)
        for (int iy = 0; iy < NY; iy++) {
            size_t i = (size_t)ix * NY + iy;
            double rho = 0.0, ux = 0.0, uy = 0.0;
            for (int k = 0; k < 9; k++) {
                fl[k] = s->f[k * n + i];
                rho += fl[k];
                ux += fl[k] * C2X[k];
                uy += fl[k] * C2Y[k];
            }
            double dense = rho;
            ux /= dense;
            uy /= dense;
            if (s->variant == 0) {
                /* mrt.cpp:165-212 */
                double forcex = s->force_ampl * cos(2.0 * PI_REF * (double)ix / (double)NX);
                double forcey = 0;
                ux += forcex / (2.0 * dense);
                ux += forcey / (2.0 * dense);
                double uxeq = ux, uyeq = uy;
                eq[0] = dense;
                eq[1] = -2.0 * dense + 3.0 * dense * (uxeq * uxeq + uyeq * uyeq);
                eq[2] = dense - 3.0 * dense * (uxeq * uxeq + uyeq * uyeq);
                eq[3] = dense * uxeq;
                eq[4] = -dense * uxeq;
                eq[5] = dense * uyeq;
                eq[6] = -dense * uyeq;
                eq[7] = dense * (uxeq * uxeq - uyeq * uyeq);
                eq[8] = dense * uxeq * uyeq;
                const double *om = s->omegamat;
                momforce[0] = 0.0;
                momforce[1] = (1.0 - om[1] / 2.0) * 6.0 * (uxeq * forcex + uyeq * forcey);
                momforce[2] = 0.0;
                momforce[3] = (1.0 - om[3] / 2.0) * forcex;
                momforce[4] = 0.0;
                momforce[5] = (1.0 - om[5] / 2.0) * forcey;
                momforce[6] = 0.0;
                momforce[7] = (1.0 - om[7] / 2.0) * 2.0 * (uxeq * forcex - uyeq * forcey);
                momforce[8] = (1.0 - om[8] / 2.0) * (uxeq * forcey + uyeq * forcex);
                for (int m = 0; m < 9; m++) {
                    add[m] = 0.0;
                    for (int k = 0; k < 9; k++) add[m] = add[m] + s->M[m][k] * (-fl[k]);
                    add[m] += eq[m];
                    add[m] *= om[m];
                    add[m] += momforce[m];
                }
                for (int k = 0; k < 9; k++) {
                    double addit = 0.0;
                    for (int m = 0; m < 9; m++) addit = addit + s->invM[k][m] * add[m];
                    fl[k] = fl[k] + addit;
                }
            } else {
                /* another_mrt.cpp:140-198 */
                double uxeq = ux, uyeq = uy;
                eq[0] = dense;
                eq[1] = sqrt(3.0) * dense * uxeq;
                eq[2] = sqrt(3.0) * dense * uyeq;
                eq[3] = 3.0 / sqrt(2.0) * dense * uxeq * uxeq;
                eq[4] = 3.0 * dense * uxeq * uyeq;
                eq[5] = 3.0 / sqrt(2.0) * dense * uyeq * uyeq;
                eq[6] = 4.5 * dense * uxeq * uxeq * uyeq * uyeq;
                eq[7] = 3.0 * sqrt(1.5) * dense * uxeq * uyeq * uyeq;
                eq[8] = 3.0 * sqrt(1.5) * dense * uxeq * uxeq * uyeq;
                for (int m = 0; m < 9; m++) {
                    add[m] = 0.0;
                    for (int k = 0; k < 9; k++) add[m] = add[m] + s->M[m][k] * (-fl[k]);
                    add[m] += eq[m];
                }
                for (int k = 0; k < 9; k++) {
                    double addit = 0.0;
                    for (int m = 0; m < 9; m++) addit = addit + s->omegamat[m] * s->M[m][k] * add[m];
                    fl[k] = fl[k] + W2[k] * addit;
                }
            }
            s->rho[i] = rho; s->ux[i] = ux; s->uy[i] = uy;
            /* pull streaming of mrt.cpp:244-251 written as a push, fully periodic */
            for (int k = 0; k < 9; k++) {
                size_t t = (size_t)((ix + C2X[k] + NX) % NX) * NY + (iy + C2Y[k] + NY) % NY;
                s->f2[k * n + t] = fl[k];
            }
        }
    double *t = s->f; s->f = s->f2; s->f2 = t;
}
void lbomrt_step(lbomrt *s, int nsteps) { for (int t = 0; t < nsteps; t++) step_mrt(s); }
void lbomrt_get_fields(lbomrt *s, double *rho, double *ux, double *uy)
{
    memcpy(rho, s->rho, s->n * sizeof(double));
    memcpy(ux, s->ux, s->n * sizeof(double));
    memcpy(uy, s->uy, s->n * sizeof(double));
}
void lbomrt_get_pops(lbomrt *s, double *f)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) f[i * 9 + k] = s->f[k * s->n + i];
}

/* ====================================================================== */
/* D2Q9 TRT diffusion with anti-bounce-back: DiffusionCylinder/cylinder.cpp */
/* ====================================================================== */
static const double WTRT[9] = {0.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 12.0, 1.0 / 12.0, 1.0 / 12.0, 1.0 / 12.0};

struct lbotrt {
    int nx, ny;
    size_t n;
    double omega_plus, omega_minus, ce, conc_wall;
    int *mask;
    double *f, *f2, *rho, *ux, *uy;   /* f SoA [k][iY*nx+iX] */
};

lbotrt *lbotrt_create(int nx, int ny, double omega, double ce, double conc_wall, const int *mask)
{
    lbotrt *s = (lbotrt *)calloc(1, sizeof *s);
    s->nx = nx; s->ny = ny; s->n = (size_t)nx * ny;
    s->omega_plus = 2.0 - omega;      /* cylinder.cpp:39-40 */
    s->omega_minus = omega;
    s->ce = ce; s->conc_wall = conc_wall;
    s->mask = (int *)malloc(s->n * sizeof(int));
    memcpy(s->mask, mask, s->n * sizeof(int));
    s->f = (double *)calloc(9 * s->n, sizeof(double));
    s->f2 = (double *)calloc(9 * s->n, sizeof(double));
    s->rho = (double *)calloc(s->n, sizeof(double));
    s->ux = (double *)calloc(s->n, sizeof(double));
    s->uy = (double *)calloc(s->n, sizeof(double));
    return s;
}
void lbotrt_destroy(lbotrt *s)
{
    if (!s) return;
    free(s->mask); free(s->f); free(s->f2); free(s->rho); free(s->ux); free(s->uy); free(s);
}
/* cylinder.cpp:224-296: circle, then fluid nodes touching solid become boundary nodes, boundary nodes
 * touching no fluid become solid; initial concentration 0 in the fluid, conc_wall on the boundary, -1 in
 * the solid.  The reference indexes iX*NY+iY here and iY*NX+iX elsewhere; NX == NY and the circle is
 * symmetric, so the flat array is the same. */
void lbotrt_reference_setup(int nx, int ny, int radius, double conc_wall, int *mask, double *rho, double *ux, double *uy)
{
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            int c = ix * ny + iy;
            mask[c] = ((ix - (nx - 1) / 2) * (ix - (nx - 1) / 2) + (iy - (ny - 1) / 2) * (iy - (ny - 1) / 2) > radius * radius) ? -1 : 1;
        }
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            int c = ix * ny + iy;
            if (mask[c] == 1) {
                int flag = 0;
                for (int k = 0; k < 9; k++)
                    if (mask[(ix + C2X[k]) * ny + iy + C2Y[k]] == -1) flag = 1;
                if (flag) mask[c] = 0;
            }
        }
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            int c = ix * ny + iy;
            if (mask[c] == 0) {
                int flag = 0;
                for (int k = 0; k < 9; k++)
                    if (mask[(ix + C2X[k]) * ny + iy + C2Y[k]] == 1) flag = 1;
                if (!flag) mask[c] = -1;
            }
        }
    for (int c = 0; c < nx * ny; c++) {
        rho[c] = mask[c] == 0 ? conc_wall : (mask[c] == -1 ? -1.0 : 0.0);
        ux[c] = 0.0; uy[c] = 0.0;
    }
}
void lbotrt_init(lbotrt *s, const double *rho, const double *ux, const double *uy)
{
    memcpy(s->rho, rho, s->n * sizeof(double));
    memcpy(s->ux, ux, s->n * sizeof(double));
    memcpy(s->uy, uy, s->n * sizeof(double));
    for (size_t c = 0; c < s->n; c++) {
        double d = s->rho[c], u = s->ux[c], v = s->uy[c];
        for (int k = 0; k < 9; k++) {
            int cx = C2X[k], cy = C2Y[k];
            s->f[k * s->n + c] = W2[k] * (d + 3.0 * d * (cx * u + cy * v) +
                                          4.5 * d * ((cx * cx - 1.0 / 3.0) * u * u + (cy * cy - 1.0 / 3.0) * v * v + 2.0 * u * v * cx * cy));
        }
    }
    /* boundary and solid nodes are never streamed into (cylinder.cpp:340): they keep these values in f for good */
    memcpy(s->f2, s->f, 9 * s->n * sizeof(double));
}
static void step_trt(lbotrt *s)
{
    const size_t n = s->n;
    const int NX = s->nx;
    for (size_t c = 0; c < n; c++) {
        if (s->mask[c] != 1) continue;
        double F[9], fo[9];
        double sum = 0;
        for (int k = 0; k < 9; k++) { F[k] = s->f[k * n + c]; sum += F[k]; }
        s->rho[c] = sum;
        double d = sum, u = s->ux[c], v = s->uy[c];
        double f_plus[9], f_minus[9], feq_plus[9], feq_minus[9];
        f_plus[0] = F[0];
        f_plus[1] = 0.5 * (F[1] + F[3]); f_plus[2] = 0.5 * (F[2] + F[4]); f_plus[3] = f_plus[1]; f_plus[4] = f_plus[2];
        f_plus[5] = 0.5 * (F[5] + F[7]); f_plus[6] = 0.5 * (F[6] + F[8]); f_plus[7] = f_plus[5]; f_plus[8] = f_plus[6];
        f_minus[0] = 0.0;
        f_minus[1] = 0.5 * (F[1] - F[3]); f_minus[2] = 0.5 * (F[2] - F[4]); f_minus[3] = -f_minus[1]; f_minus[4] = -f_minus[2];
        f_minus[5] = 0.5 * (F[5] - F[7]); f_minus[6] = 0.5 * (F[6] - F[8]); f_minus[7] = -f_minus[5]; f_minus[8] = -f_minus[6];
        feq_minus[0] = 0.0;
        feq_minus[1] = WTRT[1] * d * (C2X[1] * u + C2Y[1] * v);
        feq_minus[2] = WTRT[2] * d * (C2X[2] * u + C2Y[2] * v);
        feq_minus[3] = -feq_minus[1]; feq_minus[4] = -feq_minus[2];
        feq_minus[5] = WTRT[5] * d * (C2X[5] * u + C2Y[5] * v);
        feq_minus[6] = WTRT[6] * d * (C2X[6] * u + C2Y[6] * v);
        feq_minus[7] = -feq_minus[5]; feq_minus[8] = -feq_minus[6];
        double u_sq = u * u + v * v;
        feq_plus[1] = WTRT[1] * d * (s->ce + 0.5 * (3.0 * (C2X[1] * u + C2Y[1] * v) * (C2X[1] * u + C2Y[1] * v) - u_sq));
        feq_plus[2] = WTRT[2] * d * (s->ce + 0.5 * (3.0 * (C2X[2] * u + C2Y[2] * v) * (C2X[2] * u + C2Y[2] * v) - u_sq));
        feq_plus[3] = feq_plus[1]; feq_plus[4] = feq_plus[2];
        feq_plus[5] = WTRT[5] * d * (s->ce + 0.5 * (3.0 * (C2X[5] * u + C2Y[5] * v) * (C2X[5] * u + C2Y[5] * v) - u_sq));
        feq_plus[6] = WTRT[6] * d * (s->ce + 0.5 * (3.0 * (C2X[6] * u + C2Y[6] * v) * (C2X[6] * u + C2Y[6] * v) - u_sq));
        feq_plus[7] = feq_plus[5]; feq_plus[8] = feq_plus[6];
        feq_plus[0] = d - 2.0 * (feq_plus[1] + feq_plus[2] + feq_plus[5] + feq_plus[6]);
        for (int k = 0; k < 9; k++)
            fo[k] = F[k] - s->omega_plus * (f_plus[k] - feq_plus[k]) - s->omega_minus * (f_minus[k] - feq_minus[k]);
        /* cylinder.cpp:208-221 (anti-bounce-back on the boundary nodes) and :337-347 (pull) as one push:
         * a population heading into a boundary node comes back reversed, negated, plus 2 w C */
        for (int k = 0; k < 9; k++) {
            size_t t = c + C2Y[k] * NX + C2X[k];
            if (s->mask[t] == 0)
                s->f2[OPP2[k] * n + c] = -fo[k] + 2 * W2[OPP2[k]] * s->conc_wall;
            else
                s->f2[k * n + t] = fo[k];
        }
    }
    double *t = s->f; s->f = s->f2; s->f2 = t;
}
void lbotrt_step(lbotrt *s, int nsteps) { for (int t = 0; t < nsteps; t++) step_trt(s); }
void lbotrt_get_rho(lbotrt *s, double *rho) { memcpy(rho, s->rho, s->n * sizeof(double)); }
void lbotrt_get_pops(lbotrt *s, double *f)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) f[i * 9 + k] = s->f[k * s->n + i];
}

/* ====================================================================== */
/* D2Q9 binary liquid around a curved solid: CurvedSolid/binary_shape.cpp  */
/* ====================================================================== */
/* binary_shape.cpp:74-87 -- double tables here (binary.cpp's are float) */
static const double BS_WXX[9] = {0.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0};
static const double BS_WYY[9] = {0.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 6.0, 1.0 / 3.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0, -1.0 / 24.0};
static const double BS_WXY[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 1.0 / 4.0, -1.0 / 4.0, 1.0 / 4.0, -1.0 / 4.0};
static const double BS_GX[9] = {0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 0.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0, 1.0 / 12.0};
static const double BS_GY[9] = {0.0, 0.0, 4.0 / 12.0, 0.0, -4.0 / 12.0, 1.0 / 12.0, 1.0 / 12.0, -1.0 / 12.0, -1.0 / 12.0};
static const double BS_LAP[9] = {-20.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 4.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0};

struct lbobs {
    int nx, ny;
    size_t n;
    double aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi, force_x, force_y;
    int *mask;
    int nbb, *bb, *bb_src;       /* bounce-back nodes in ascending index order and the node their phi is copied from */
    double *f, *g, *f2, *g2, *phi, *rho, *ux, *uy;   /* populations SoA [k][i] */
};

/* binary_shape.cpp:556-588: sum of the axis directions that lead to fluid, else of the diagonal ones; the
 * wall normal is the lattice direction equal to that sum */
static int bs_main_dir(const int *mask, int nx, int c)
{
    int sx = 0, sy = 0, flag = 0;
    for (int k = 1; k < 5; k++)
        if (mask[c + C2Y[k] * nx + C2X[k]] == 1) { flag = 1; sx += C2X[k]; sy += C2Y[k]; }
    if (!flag)
        for (int k = 5; k < 9; k++)
            if (mask[c + C2Y[k] * nx + C2X[k]] == 1) { sx += C2X[k]; sy += C2Y[k]; }
    for (int k = 1; k < 9; k++)
        if (sx == C2X[k] && sy == C2Y[k]) return k;
    return 0;
}

lbobs *lbobs_create(int nx, int ny, const double *p, const int *mask)
{
    lbobs *s = (lbobs *)calloc(1, sizeof *s);
    s->nx = nx; s->ny = ny; s->n = (size_t)nx * ny;
    s->aconst = p[0]; s->kconst = p[1]; s->gammaconst = p[2]; s->wall_gradient = p[3];
    s->tau_gas = p[4]; s->tau_liq = p[5]; s->omega_phi = p[6]; s->force_x = p[7]; s->force_y = p[8];
    s->mask = (int *)malloc(s->n * sizeof(int));
    memcpy(s->mask, mask, s->n * sizeof(int));
    s->bb = (int *)malloc(s->n * sizeof(int));
    s->bb_src = (int *)malloc(s->n * sizeof(int));
    for (size_t c = 0; c < s->n; c++) {
        if (mask[c] != 0) continue;
        int ix = (int)(c % nx), iy = (int)(c / nx);
        int k = (ix == 0 || iy == 0 || ix == nx - 1 || iy == ny - 1) ? 0 : bs_main_dir(mask, nx, (int)c);
        if (k == 0 || mask[c + C2Y[k] * nx + C2X[k]] != 1) { lbobs_destroy(s); return NULL; }
        s->bb[s->nbb] = (int)c;
        s->bb_src[s->nbb++] = (int)c + C2Y[k] * nx + C2X[k];
    }
    double **arr[] = {&s->f, &s->g, &s->f2, &s->g2};
    for (int a = 0; a < 4; a++) *arr[a] = (double *)calloc(9 * s->n, sizeof(double));
    double **fld[] = {&s->phi, &s->rho, &s->ux, &s->uy};
    for (int a = 0; a < 4; a++) *fld[a] = (double *)calloc(s->n, sizeof(double));
    return s;
}
void lbobs_destroy(lbobs *s)
{
    if (!s) return;
    free(s->mask); free(s->bb); free(s->bb_src);
    free(s->f); free(s->g); free(s->f2); free(s->g2); free(s->phi); free(s->rho); free(s->ux); free(s->uy);
    free(s);
}

/* binary_shape.cpp:439-520 with NX, NY, radius, radius_droplet as arguments (rho_wall 0.5, phi_wall 1.5,
 * three droplets 120 degrees apart) */
void lbobs_reference_setup(int nx, int ny, int radius, int radius_droplet, int *mask, double *rho, double *phi, double *ux, double *uy)
{
    const double pi = 4.0 * atan(1.0), phi_wall = 1.5, rho_wall = 0.5;
    const int angle = 360 / 3, n = nx * ny;
    for (int c = 0; c < n; c++) {
        int ix = c % nx, iy = c / nx;
        mask[c] = ((ix - (nx - 1) / 2) * (ix - (nx - 1) / 2) + (iy - (ny - 1) / 2) * (iy - (ny - 1) / 2) < radius * radius) ? -1 : 1;
    }
    for (int c = 0; c < n; c++) {
        int flag = 0;
        if (mask[c] == -1) {
            int ix = c % nx, iy = c / nx;
            for (int k = 1; k < 9; k++)
                if (mask[((iy + C2Y[k] + ny) % ny) * nx + (ix + C2X[k] + nx) % nx] == 1) flag = 1;
        }
        if (flag) mask[c] = 0;
    }
    for (int c = 0; c < n; c++) {
        ux[c] = 0.0; uy[c] = 0.0;
        rho[c] = mask[c] == 1 ? 1.0 : rho_wall;
    }
    for (int c = 0; c < n; c++) {
        int ix = c % nx, iy = c / nx, in_droplet = 0;
        for (int d = 0; d < 3; d++) {
            double center_x = (nx - 1) / 2 - sin(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius);
            double center_y = (ny - 1) / 2 + cos(angle * (d - 1) * pi / 180.0) * (radius_droplet + radius);
            if (((ix - center_x) * (ix - center_x) + (iy - center_y) * (iy - center_y) <= radius_droplet * radius_droplet) && mask[c] == 1) {
                phi[c] = 1.0;
                in_droplet = 1;
            }
        }
        if (!in_droplet) phi[c] = mask[c] == 1 ? -1.0 : phi_wall;
    }
}

void lbobs_init(lbobs *s, const double *rho, const double *phi, const double *ux, const double *uy)
{
    const size_t n = s->n;
    memcpy(s->rho, rho, n * sizeof(double)); memcpy(s->phi, phi, n * sizeof(double));
    memcpy(s->ux, ux, n * sizeof(double)); memcpy(s->uy, uy, n * sizeof(double));
    for (size_t c = 0; c < n; c++) {
        double d = rho[c], u = ux[c], v = uy[c], ph = phi[c];
        for (int k = 0; k < 9; k++) {
            int cx = C2X[k], cy = C2Y[k];
            s->f[k * n + c] = W2[k] * (d + 3.0 * d * (cx * u + cy * v) +
                                       4.5 * d * ((cx * cx - 1.0 / 3.0) * u * u + (cy * cy - 1.0 / 3.0) * v * v + 2.0 * u * v * cx * cy));
            s->g[k * n + c] = W2[k] * (ph + 3.0 * ph * (cx * u + cy * v) +
                                       4.5 * ph * ((cx * cx - 1.0 / 3.0) * u * u + (cy * cy - 1.0 / 3.0) * v * v + 2.0 * u * v * cx * cy));
        }
    }
    /* bounce-back and solid nodes are never streamed into: they keep these values in f and g for good */
    memcpy(s->f2, s->f, 9 * n * sizeof(double));
    memcpy(s->g2, s->g, 9 * n * sizeof(double));
}

static void step_bs(lbobs *s)
{
    const size_t n = s->n;
    const int NX = s->nx, NY = s->ny;
    /* update_phase */
    for (size_t c = 0; c < n; c++)
        if (s->mask[c] == 1) {
            double ph = 0.0;
            for (int k = 0; k < 9; k++) ph += s->g[k * n + c];
            s->phi[c] = ph;
        }
    for (int b = 0; b < s->nbb; b++) s->phi[s->bb[b]] = s->phi[s->bb_src[b]] - s->wall_gradient;
    /* collide_bulk, then bounce-back and pull streaming as one push */
    for (size_t c = 0; c < n; c++) {
        if (s->mask[c] != 1) continue;
        const int ix = (int)(c % NX), iy = (int)(c / NX);
        double F[9], G[9];
        double d = 0.0, ph = s->phi[c], u = 0.0, v = 0.0;
        for (int k = 0; k < 9; k++) {
            F[k] = s->f[k * n + c]; G[k] = s->g[k * n + c];
            d += F[k]; u += F[k] * C2X[k]; v += F[k] * C2Y[k];
        }
        u = (u + 0.5 * s->force_x) / d;
        v = (v + 0.5 * s->force_y) / d;
        s->rho[c] = d; s->ux[c] = u; s->uy[c] = v;
        double lap = 0.0, gx = 0.0, gy = 0.0;
        size_t nb[9];
        for (int k = 0; k < 9; k++) {
            nb[k] = (size_t)((iy + C2Y[k] + NY) % NY) * NX + (ix + C2X[k] + NX) % NX;
            lap += BS_LAP[k] * s->phi[nb[k]];
            gx += BS_GX[k] * s->phi[nb[k]];
            gy += BS_GY[k] * s->phi[nb[k]];
        }
        double sum = 0.0, sum_phase = 0.0, sq = ph * ph;
        double pb = d / 3.0 + s->aconst * (-0.5 * sq + 3.0 / 4.0 * sq * sq) - s->kconst * ph * lap;
        double mu = s->gammaconst * (s->aconst * (-ph + ph * ph * ph) - s->kconst * lap);
        double feq[9], geq[9];
        for (int k = 1; k < 9; k++) {
            int cx = C2X[k], cy = C2Y[k];
            feq[k] = W2[k] * (3.0 * pb + 3.0 * d * (cx * u + cy * v) +
                              4.5 * d * ((cx * cx - 1.0 / 3.0) * u * u + (cy * cy - 1.0 / 3.0) * v * v + 2.0 * u * v * cx * cy)) +
                     s->kconst * (BS_WXX[k] * gx * gx + BS_WYY[k] * gy * gy + BS_WXY[k] * gx * gy);
            geq[k] = W2[k] * (3.0 * mu + 3.0 * ph * (cx * u + cy * v) +
                              4.5 * ph * ((cx * cx - 1.0 / 3.0) * u * u + (cy * cy - 1.0 / 3.0) * v * v + 2.0 * u * v * cx * cy));
            sum += feq[k];
            sum_phase += geq[k];
        }
        feq[0] = d - sum;
        geq[0] = ph - sum_phase;
        double tau_rho = s->tau_gas + (ph + 1.0) / 2.0 * (s->tau_liq - s->tau_gas);
        double omega_rho = 1.0 / tau_rho;
        for (int k = 0; k < 9; k++) {
            double fo = F[k] * (1.0 - omega_rho) + omega_rho * feq[k];
            double go = G[k] * (1.0 - s->omega_phi) + s->omega_phi * geq[k];
            if (s->mask[nb[k]] == 0) {      /* :358-372 + :617-637: comes back reversed from the bounce-back node */
                s->f2[OPP2[k] * n + c] = fo;
                s->g2[OPP2[k] * n + c] = go;
            } else {
                s->f2[k * n + nb[k]] = fo;
                s->g2[k * n + nb[k]] = go;
            }
        }
    }
    double *t = s->f; s->f = s->f2; s->f2 = t;
    t = s->g; s->g = s->g2; s->g2 = t;
}
void lbobs_step(lbobs *s, int nsteps) { for (int t = 0; t < nsteps; t++) step_bs(s); }
void lbobs_get_fields(lbobs *s, double *phi, double *rho, double *ux, double *uy)
{
    memcpy(phi, s->phi, s->n * sizeof(double)); memcpy(rho, s->rho, s->n * sizeof(double));
    memcpy(ux, s->ux, s->n * sizeof(double)); memcpy(uy, s->uy, s->n * sizeof(double));
}
void lbobs_get_pops(lbobs *s, double *f, double *g)
{
    for (size_t i = 0; i < s->n; i++)
        for (int k = 0; k < 9; k++) { f[i * 9 + k] = s->f[k * s->n + i]; g[i * 9 + k] = s->g[k * s->n + i]; }
}
