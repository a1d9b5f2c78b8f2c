// d3q19_node.h -- per-node arithmetic of the D3Q19 binary-liquid update.
//
// Reference: Microchannel3D/*/lattice.hpp:15-38 (moments), :412-414 (force
// shift), :425-462 (equilibria), :465-480 (relaxation rates, Guo force
// populations), :483-488 (BGK relaxation).
//
// The reference evaluates everything per direction k with 19-entry look-up
// tables.  Here the tables are folded into the code: terms multiplied by a
// table zero are dropped, multiplications by +-1 become sign changes, and
// sub-expressions that are the same IEEE operation on the same operands for
// several k are computed once.  None of that changes a single bit (x*1 = x,
// x+0 = x for x != -0, (-a)+(-b) = -(a+b), a*(-b) = -(a*b) are exact in
// binary64), but no sum is re-associated and nothing is contracted into an
// FMA: the file must be compiled with -fmad=false (nvcc) / -ffp-contract=off
// (g++), see build.py.  The header is plain C++ so the CPU unit test can
// compile the very same code.
#ifndef LBM_D3Q19_NODE_H
#define LBM_D3Q19_NODE_H

#if defined(__CUDACC__)
#define LBM_HD __host__ __device__ __forceinline__
#define LBM_HD_NOINLINE __host__ __device__ __noinline__
#else
#define LBM_HD inline
#define LBM_HD_NOINLINE inline
#endif

namespace lbm3 {

struct Params {
    double aconst, kconst, gammaconst;
    double tau_gas, dtau;      // dtau = tau_liq - tau_gas (lattice.hpp:465)
    double omega_phi;          // 1.0 / tau_phi (lattice.hpp:467)
    double fx, fy, fz;
    double phase_gradient, rho_wall;
    double trt_magic;          // 0: BGK as in the reference; > 0: two-relaxation-time collision of f (extension, no reference)
    double mrt[5];             // all 0: off; else multiple-relaxation-time collision of f (extension): rates of the energy,
                               // energy-square, heat-flux, fourth-order and third-order moments; an entry <= 0 means 1/tau
};

// Moment basis of d'Humieres et al. (Phil. Trans. R. Soc. A 360, 2002) for D3Q19, row i as a polynomial of the
// direction (cx, cy, cz): density, energy e, energy square eps, j_x, q_x, j_y, q_y, j_z, q_z, 3p_xx, 3pi_xx, p_ww, pi_ww,
// p_xy, p_yz, p_xz, m_x, m_y, m_z.  The rows are mutually orthogonal, so M^-1 = M^T diag(1 / |row|^2).
LBM_HD constexpr int mrt19_row(int i, int cx, int cy, int cz)
{
    const int c2 = cx * cx + cy * cy + cz * cz;
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

// direction table of descriptor.h:43-45 for the code that is not unrolled
#define LBM3_CX {0, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1, 0, 0, 0, 0, 1, -1, 1, -1}
#define LBM3_CY {0, 0, 0, 1, -1, 0, 0, 1, 1, -1, -1, 1, -1, 1, -1, 0, 0, 0, 0}
#define LBM3_CZ {0, 0, 0, 0, 0, 1, -1, 0, 0, 0, 0, 1, 1, -1, -1, 1, 1, -1, -1}

// rho and momentum sums, lattice.hpp:15-26 (left to right as written there)
LBM_HD double sum19(const double *G)
{
    return G[0] + G[1] + G[2] + G[3] + G[4] + G[5] + G[6] + G[7] + G[8] + G[9] + G[10] + G[11] + G[12] + G[13] +
           G[14] + G[15] + G[16] + G[17] + G[18];
}

// Sum of the 19 populations of a mirror node: the mirror node stores its inner
// neighbour's populations permuted (lattice.hpp:645-686), so the reference's
// left-to-right sum (lattice.hpp:19-22) visits them in permuted order.
// MODE 1: x mirror (symmetricx), 2: y mirror (symmetricy), 3: both (symmetricxy); lattice.h:58-81
template <int MODE>
LBM_HD double sum19_mirror(const double *G)
{
    if (MODE == 1)  // {0,2,1,3,4,5,6,8,7,10,9,11,12,13,14,16,15,18,17}
        return G[0] + G[2] + G[1] + G[3] + G[4] + G[5] + G[6] + G[8] + G[7] + G[10] + G[9] + G[11] + G[12] + G[13] +
               G[14] + G[16] + G[15] + G[18] + G[17];
    if (MODE == 2)  // {0,1,2,4,3,5,6,9,10,7,8,12,11,14,13,15,16,17,18}
        return G[0] + G[1] + G[2] + G[4] + G[3] + G[5] + G[6] + G[9] + G[10] + G[7] + G[8] + G[12] + G[11] + G[14] +
               G[13] + G[15] + G[16] + G[17] + G[18];
    // {0,2,1,4,3,5,6,10,9,8,7,12,11,14,13,16,15,18,17}
    return G[0] + G[2] + G[1] + G[4] + G[3] + G[5] + G[6] + G[10] + G[9] + G[8] + G[7] + G[12] + G[11] + G[14] +
           G[13] + G[16] + G[15] + G[18] + G[17];
}

LBM_HD void moments(const double *F, double &rho, double &ux, double &uy, double &uz)
{
    rho = sum19(F);
    double jx = F[1] - F[2] + F[7] - F[8] + F[9] - F[10] + F[15] - F[16] + F[17] - F[18];
    double jy = F[3] - F[4] + F[7] + F[8] - F[9] - F[10] + F[11] - F[12] + F[13] - F[14];
    double jz = F[5] - F[6] + F[11] + F[12] - F[13] - F[14] + F[15] + F[16] - F[17] - F[18];
    ux = jx / rho;
    uy = jy / rho;
    uz = jz / rho;
}

// Collides the 19+19 populations of one bulk node in place.
//   F,G        populations (in: streamed, out: post-collision)
//   ph         phase field at the node, gx,gy,gz,lap its stencils (lattice.hpp:392-409)
//   rho,ux..   out: density and force-shifted velocity (what the reference stores, :412-414)
// FXY = false is the fast path for force_x == force_y == 0 (both shipped
// configurations): the x and y force terms are then exact zeros in the reference.
//
// COLL selects the collision operator of f: 0 the reference's BGK; 1 (TRT) and 2 (MRT) are EXTENSIONS.
// COLL = 2, multiple relaxation times: the non-equilibrium part of f is taken to the moment space of mrt19_row, every
// moment relaxed with its own rate -- the stress moments (rows 9, 11, 13-15) with 1/tau, which keeps the reference's
// viscosity, density and momentum rows with 1/tau as well, the other five groups with P.mrt[] -- and brought back; the
// Guo force populations go the same way with (1 - s_i/2) per moment, so that a step adds exactly F to the momentum.
// Equilibria and the g collision are the reference's.  With every rate equal to 1/tau the scheme is BGK up to rounding.
// COLL = 1 is an EXTENSION with no counterpart in the reference (the north star names TRT on D3Q19, the reference
// has none): the f populations relax with two rates -- the symmetric part (f_k + f_kbar)/2 with omega+ = 1/tau, which
// keeps the reference's viscosity, the antisymmetric part with omega- = 1/(Lambda/(tau - 1/2) + 1/2), Lambda =
// P.trt_magic -- while equilibria, Guo force populations and the g collision stay exactly the reference's.  With
// Lambda = (tau - 1/2)^2 both rates coincide and the scheme is BGK up to rounding; that limit is the validation.
template <bool FXY, int COLL = 0>
LBM_HD void collide(double *F, double *G, double ph, double gx, double gy, double gz, double lap, const Params &P,
                    double &rho, double &ux, double &uy, double &uz)
{
    moments(F, rho, ux, uy, uz);
    const double two_rho = 2.0 * rho;
    if (FXY) {
        ux += P.fx / two_rho;
        uy += P.fy / two_rho;
    }
    uz += P.fz / two_rho;

    // lattice.hpp:425-427
    const double ps = ph * ph;
    const double pb = rho / 3.0 + P.aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - P.kconst * ph * lap;
    const double mu = P.gammaconst * (P.aconst * (-ph + ps * ph) - P.kconst * lap);

    const double uxs = ux * ux, uys = uy * uy, uzs = uz * uz;
    const double uxuy = ux * uy, uxuz = ux * uz, uyuz = uy * uz;
    const double gxs = gx * gx, gys = gy * gy, gzs = gz * gz;
    const double gxy = gx * gy, gxz = gx * gz, gyz = gy * gz;

    // Q_k = qxx*ux^2 + qyy*uy^2 + qzz*uz^2 + 2*(qxy*uxuy + qxz*uxuz + qyz*uyuz): nine distinct values
    const double q2 = 2.0 / 3.0, q1 = -1.0 / 3.0;
    const double a2x = q2 * uxs, a1x = q1 * uxs, a2y = q2 * uys, a1y = q1 * uys, a2z = q2 * uzs, a1z = q1 * uzs;
    const double Qx = (a2x + a1y) + a1z, Qy = (a1x + a2y) + a1z, Qz = (a1x + a1y) + a2z;
    const double Bxy = (a2x + a2y) + a1z, Byz = (a1x + a2y) + a2z, Bxz = (a2x + a1y) + a2z;
    const double cxy = 2.0 * uxuy, cyz = 2.0 * uyuz, cxz = 2.0 * uxuz;
    const double Qxyp = Bxy + cxy, Qxym = Bxy - cxy;
    const double Qyzp = Byz + cyz, Qyzm = Byz - cyz;
    const double Qxzp = Bxz + cxz, Qxzm = Bxz - cxz;

    // kconst * (wxx*gx^2 + wyy*gy^2 + wzz*gz^2 + wxy*gxgy + wyz*gygz + wzx*gxgz): nine distinct values
    const double wa = 5.0 / 12.0, wb = -1.0 / 3.0, wc = -1.0 / 24.0, wd = 1.0 / 12.0, we = 1.0 / 4.0;
    const double K = P.kconst;
    const double Kx = K * ((wa * gxs + wb * gys) + wb * gzs);
    const double Ky = K * ((wb * gxs + wa * gys) + wb * gzs);
    const double Kz = K * ((wb * gxs + wb * gys) + wa * gzs);
    const double Gxy = (wc * gxs + wc * gys) + wd * gzs, exy = we * gxy;
    const double Gyz = (wd * gxs + wc * gys) + wc * gzs, eyz = we * gyz;
    const double Gxz = (wc * gxs + wd * gys) + wc * gzs, exz = we * gxz;
    const double Kxyp = K * (Gxy + exy), Kxym = K * (Gxy - exy);
    const double Kyzp = K * (Gyz + eyz), Kyzm = K * (Gyz - eyz);
    const double Kxzp = K * (Gxz + exz), Kxzm = K * (Gxz - exz);

    // c_k . u for the nine direction pairs
    const double c1 = ux, c3 = uy, c5 = uz;
    const double c7 = ux + uy, c8 = uy - ux, c11 = uy + uz, c12 = uz - uy, c15 = ux + uz, c16 = uz - ux;

    const double w6 = 1.0 / 6.0, w12 = 1.0 / 12.0;
    const double h = 1.5 * rho, hp = 1.5 * ph;
    double feq[19], geq[19];
#define LBM3_EQ(k, kopp, w, c, Q, KG)                       \
    {                                                       \
        const double rc = rho * (c), pc = ph * (c);         \
        const double hq = h * (Q), hpq = hp * (Q);          \
        feq[k] = (w) * ((pb + rc) + hq) + (KG);             \
        feq[kopp] = (w) * ((pb - rc) + hq) + (KG);          \
        geq[k] = (w) * ((mu + pc) + hpq);                   \
        geq[kopp] = (w) * ((mu - pc) + hpq);                \
    }
    LBM3_EQ(1, 2, w6, c1, Qx, Kx)
    LBM3_EQ(3, 4, w6, c3, Qy, Ky)
    LBM3_EQ(5, 6, w6, c5, Qz, Kz)
    LBM3_EQ(7, 10, w12, c7, Qxyp, Kxyp)
    LBM3_EQ(8, 9, w12, c8, Qxym, Kxym)
    LBM3_EQ(11, 14, w12, c11, Qyzp, Kyzp)
    LBM3_EQ(12, 13, w12, c12, Qyzm, Kyzm)
    LBM3_EQ(15, 18, w12, c15, Qxzp, Kxzp)
    LBM3_EQ(16, 17, w12, c16, Qxzm, Kxzm)
#undef LBM3_EQ
    double sum = 0.0, sum_phase = 0.0;
#pragma unroll
    for (int k = 1; k < 19; k++) {
        sum += feq[k];
        sum_phase += geq[k];
    }
    feq[0] = rho - sum;
    geq[0] = ph - sum_phase;

    // lattice.hpp:465-467
    const double tau = P.tau_gas + (ph + 1.0) / 2.0 * P.dtau;
    const double omega = 1.0 / tau;
    const double pref = COLL == 2 ? 1.0 : 1.0 - 0.5 * omega;   // MRT applies (1 - s_i/2) per moment instead
    const double pw6 = pref * w6, pw12 = pref * w12;

    // Guo force populations, lattice.hpp:472-480
    double fp[19];
    if (FXY) {
        const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
        const double cu[19] = {0.0, c1, -c1, c3, -c3, c5, -c5, c7, c8, -c8, -c7, c11, c12, -c12, -c11, c15, c16, -c16, -c15};
#pragma unroll
        for (int k = 1; k < 19; k++) {
            const double pw = k < 7 ? pw6 : pw12;
            fp[k] = pw * (P.fx * ((CX[k] - ux) + 3.0 * CX[k] * cu[k]) + P.fy * ((CY[k] - uy) + 3.0 * CY[k] * cu[k]) +
                          P.fz * ((CZ[k] - uz) + 3.0 * CZ[k] * cu[k]));
        }
    } else {
        // only the z term survives; (cz - uz) + 3*cz*(c.u) with cz in {0,+1,-1}
        const double t0 = P.fz * (-uz);
        const double mp = 1.0 - uz, mm = -1.0 - uz;
        const double e5 = 3.0 * c5, e11 = 3.0 * c11, e12 = 3.0 * c12, e15 = 3.0 * c15, e16 = 3.0 * c16;
        const double a6 = pw6 * t0, a12 = pw12 * t0;
        fp[1] = a6; fp[2] = a6; fp[3] = a6; fp[4] = a6;
        fp[5] = pw6 * (P.fz * (mp + e5));
        fp[6] = pw6 * (P.fz * (mm + e5));
        fp[7] = a12; fp[8] = a12; fp[9] = a12; fp[10] = a12;
        fp[11] = pw12 * (P.fz * (mp + e11));
        fp[12] = pw12 * (P.fz * (mp + e12));
        fp[13] = pw12 * (P.fz * (mm + e12));
        fp[14] = pw12 * (P.fz * (mm + e11));
        fp[15] = pw12 * (P.fz * (mp + e15));
        fp[16] = pw12 * (P.fz * (mp + e16));
        fp[17] = pw12 * (P.fz * (mm + e16));
        fp[18] = pw12 * (P.fz * (mm + e15));
    }
    double sum_force = 0.0;
#pragma unroll
    for (int k = 1; k < 19; k++) sum_force += fp[k];
    fp[0] = -sum_force;

    // lattice.hpp:483-488
    const double mo = -omega, mop = -P.omega_phi;
    if (COLL == 2) {
        constexpr int MX[19] = LBM3_CX, MY[19] = LBM3_CY, MZ[19] = LBM3_CZ;
        const double se = P.mrt[0] > 0.0 ? P.mrt[0] : omega, seps = P.mrt[1] > 0.0 ? P.mrt[1] : omega;
        const double sq = P.mrt[2] > 0.0 ? P.mrt[2] : omega, spi = P.mrt[3] > 0.0 ? P.mrt[3] : omega;
        const double sm = P.mrt[4] > 0.0 ? P.mrt[4] : omega;
        const double S[19] = {omega, se, seps, omega, sq, omega, sq, omega, sq, omega, spi, omega, spi, omega, omega, omega, sm, sm, sm};
        double fn[19], t[19];
#pragma unroll
        for (int k = 0; k < 19; k++) fn[k] = F[k] - feq[k];
#pragma unroll
        for (int i = 0; i < 19; i++) {
            double m = 0.0, mf = 0.0;
            int norm = 0;
#pragma unroll
            for (int k = 0; k < 19; k++) {
                const int e = mrt19_row(i, MX[k], MY[k], MZ[k]);
                norm += e * e;
                if (e != 0) {
                    m += (double)e * fn[k];
                    mf += (double)e * fp[k];
                }
            }
            const double inv = 1.0 / (double)norm;
            t[i] = mf * ((1.0 - 0.5 * S[i]) * inv) - m * (S[i] * inv);
        }
#pragma unroll
        for (int k = 0; k < 19; k++) {
            double a = 0.0;
#pragma unroll
            for (int i = 0; i < 19; i++) {
                const int e = mrt19_row(i, MX[k], MY[k], MZ[k]);
                if (e != 0) a += (double)e * t[i];
            }
            F[k] += a;
            G[k] += mop * (G[k] - geq[k]);
        }
        return;
    }
    if (COLL == 1) {
        const double omega_minus = 1.0 / (P.trt_magic / (tau - 0.5) + 0.5);
        // the force populations split like the populations themselves: the symmetric half keeps Guo's prefactor
        // (1 - omega/2) that fp[] already carries, the antisymmetric half -- the one that injects momentum -- takes
        // (1 - omega_minus/2), so that a step adds exactly F to the momentum whatever Lambda is
        const double fr = (1.0 - 0.5 * omega_minus) / pref;
        F[0] = (F[0] - omega * (F[0] - feq[0])) + fp[0];
#define LBM3_TRT(k, kb)                                                     \
    {                                                                       \
        const double fpl = 0.5 * (F[k] + F[kb]), fmi = 0.5 * (F[k] - F[kb]); \
        const double epl = 0.5 * (feq[k] + feq[kb]), emi = 0.5 * (feq[k] - feq[kb]); \
        const double a = omega * (fpl - epl), b = omega_minus * (fmi - emi); \
        const double fsy = 0.5 * (fp[k] + fp[kb]), fan = fr * (0.5 * (fp[k] - fp[kb])); \
        F[k] = ((F[k] - a) - b) + (fsy + fan);                              \
        F[kb] = ((F[kb] - a) + b) + (fsy - fan);                            \
    }
        LBM3_TRT(1, 2) LBM3_TRT(3, 4) LBM3_TRT(5, 6) LBM3_TRT(7, 10) LBM3_TRT(8, 9)
        LBM3_TRT(11, 14) LBM3_TRT(12, 13) LBM3_TRT(15, 18) LBM3_TRT(16, 17)
#undef LBM3_TRT
#pragma unroll
        for (int k = 0; k < 19; k++) G[k] += mop * (G[k] - geq[k]);
        return;
    }
#pragma unroll
    for (int k = 0; k < 19; k++) {
        F[k] += mo * (F[k] - feq[k]) + fp[k];
        G[k] += mop * (G[k] - geq[k]);
    }
}

// The same collision as collide<>(), restructured so that only the 19 f
// populations live in registers: the g populations are fetched one at a time
// through `getG(K)`, every post-collision pair is handed to `emit<K>(f, g)` as
// soon as it exists, and the per-direction equilibria are formed group by group
// instead of in 19-entry arrays.  Every arithmetic expression is the one in
// collide<>() (hence in lattice.hpp:412-488); only the order in which
// independent values are produced differs, and the three running sums still
// add their terms in ascending direction order.  About 100 fewer live registers,
// which is what lets four 128-thread blocks share an SM.
//   getG : double operator()(int k)
//   emit : template <int K> void operator()(double f_new, double g_new)
template <bool FXY, class GetG, class Emit>
LBM_HD void collide_streamed(double *F, GetG getG, Emit emit, double ph, double gx, double gy, double gz, double lap,
                             const Params &P, double &rho, double &ux, double &uy, double &uz)
{
    moments(F, rho, ux, uy, uz);
    const double two_rho = 2.0 * rho;
    if (FXY) {
        ux += P.fx / two_rho;
        uy += P.fy / two_rho;
    }
    uz += P.fz / two_rho;

    const double ps = ph * ph;
    const double pb = rho / 3.0 + P.aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - P.kconst * ph * lap;
    const double mu = P.gammaconst * (P.aconst * (-ph + ps * ph) - P.kconst * lap);

    const double q2 = 2.0 / 3.0, q1 = -1.0 / 3.0;
    const double uxs = ux * ux, uys = uy * uy, uzs = uz * uz;
    const double a2x = q2 * uxs, a1x = q1 * uxs, a2y = q2 * uys, a1y = q1 * uys, a2z = q2 * uzs, a1z = q1 * uzs;
    const double gxs = gx * gx, gys = gy * gy, gzs = gz * gz;
    const double wa = 5.0 / 12.0, wb = -1.0 / 3.0, wc = -1.0 / 24.0, wd = 1.0 / 12.0, we = 1.0 / 4.0;
    const double K = P.kconst;
    const double w6 = 1.0 / 6.0, w12 = 1.0 / 12.0;
    const double h = 1.5 * rho, hp = 1.5 * ph;

    const double tau = P.tau_gas + (ph + 1.0) / 2.0 * P.dtau;
    const double omega = 1.0 / tau;
    const double pref = 1.0 - 0.5 * omega;
    const double pw6 = pref * w6, pw12 = pref * w12;
    const double mo = -omega, mop = -P.omega_phi;
    // force: only fz survives on the fast path; (cz - uz) + 3*cz*(c.u)
    const double t0 = P.fz * (-uz), mp = 1.0 - uz, mm = -1.0 - uz;

    double sum = 0.0, sum_phase = 0.0, sum_force = 0.0;

    // general force population for direction (cx,cy,cz) with c.u = cu  (lattice.hpp:474-476)
#define LBM3_FP(pw, cx, cy, cz, cu)                                                                                   \
    ((pw) * (P.fx * (((cx) - ux) + 3.0 * (cx) * (cu)) + P.fy * (((cy) - uy) + 3.0 * (cy) * (cu)) +                     \
             P.fz * (((cz) - uz) + 3.0 * (cz) * (cu))))
    // one direction pair (k, kopp = -k): equilibria, then relax and emit in ascending order
#define LBM3_PAIR(k, kopp, w, c, Q, KG, fpk, fpo)                                                                    \
    {                                                                                                                 \
        const double rc = rho * (c), pc = ph * (c);                                                                   \
        const double hq = h * (Q), hpq = hp * (Q);                                                                    \
        const double fe_k = (w) * ((pb + rc) + hq) + (KG), fe_o = (w) * ((pb - rc) + hq) + (KG);                      \
        const double ge_k = (w) * ((mu + pc) + hpq), ge_o = (w) * ((mu - pc) + hpq);                                  \
        const double fp_k = (fpk), fp_o = (fpo);                                                                      \
        sum += fe_k; sum += fe_o;                                                                                     \
        sum_phase += ge_k; sum_phase += ge_o;                                                                         \
        sum_force += fp_k; sum_force += fp_o;                                                                         \
        const double g_k = getG(k), g_o = getG(kopp);                                                                 \
        emit.template operator()<k>(F[k] + (mo * (F[k] - fe_k) + fp_k), g_k + mop * (g_k - ge_k));                    \
        emit.template operator()<kopp>(F[kopp] + (mo * (F[kopp] - fe_o) + fp_o), g_o + mop * (g_o - ge_o));           \
    }
    // a diagonal plane: directions kp (+,+), km (-,+), kmo (+,-) = -km, kpo (-,-) = -kp, numbered kp < km < kmo < kpo
#define LBM3_QUAD(kp, km, kmo, kpo, cp, cm, B, cross, Gd, e, fp_p, fp_m, fp_mo, fp_po)                                \
    {                                                                                                                 \
        const double Qp = (B) + (cross), Qm = (B) - (cross);                                                          \
        const double Kp = K * ((Gd) + (e)), Km = K * ((Gd) - (e));                                                    \
        const double rcp = rho * (cp), pcp = ph * (cp), rcm = rho * (cm), pcm = ph * (cm);                            \
        const double hqp = h * Qp, hpqp = hp * Qp, hqm = h * Qm, hpqm = hp * Qm;                                      \
        const double fe_p = w12 * ((pb + rcp) + hqp) + Kp, fe_po = w12 * ((pb - rcp) + hqp) + Kp;                     \
        const double fe_m = w12 * ((pb + rcm) + hqm) + Km, fe_mo = w12 * ((pb - rcm) + hqm) + Km;                     \
        const double ge_p = w12 * ((mu + pcp) + hpqp), ge_po = w12 * ((mu - pcp) + hpqp);                             \
        const double ge_m = w12 * ((mu + pcm) + hpqm), ge_mo = w12 * ((mu - pcm) + hpqm);                             \
        const double f_p = (fp_p), f_m = (fp_m), f_mo = (fp_mo), f_po = (fp_po);                                      \
        sum += fe_p; sum += fe_m; sum += fe_mo; sum += fe_po;                                                         \
        sum_phase += ge_p; sum_phase += ge_m; sum_phase += ge_mo; sum_phase += ge_po;                                 \
        sum_force += f_p; sum_force += f_m; sum_force += f_mo; sum_force += f_po;                                     \
        const double g_p = getG(kp), g_m = getG(km), g_mo = getG(kmo), g_po = getG(kpo);                              \
        emit.template operator()<kp>(F[kp] + (mo * (F[kp] - fe_p) + f_p), g_p + mop * (g_p - ge_p));                  \
        emit.template operator()<km>(F[km] + (mo * (F[km] - fe_m) + f_m), g_m + mop * (g_m - ge_m));                  \
        emit.template operator()<kmo>(F[kmo] + (mo * (F[kmo] - fe_mo) + f_mo), g_mo + mop * (g_mo - ge_mo));          \
        emit.template operator()<kpo>(F[kpo] + (mo * (F[kpo] - fe_po) + f_po), g_po + mop * (g_po - ge_po));          \
    }

    {   // axis directions 1..6
        const double Qx = (a2x + a1y) + a1z, Qy = (a1x + a2y) + a1z, Qz = (a1x + a1y) + a2z;
        const double Kx = K * ((wa * gxs + wb * gys) + wb * gzs);
        const double Ky = K * ((wb * gxs + wa * gys) + wb * gzs);
        const double Kz = K * ((wb * gxs + wb * gys) + wa * gzs);
        if (FXY) {
            LBM3_PAIR(1, 2, w6, ux, Qx, Kx, LBM3_FP(pw6, 1, 0, 0, ux), LBM3_FP(pw6, -1, 0, 0, -ux))
            LBM3_PAIR(3, 4, w6, uy, Qy, Ky, LBM3_FP(pw6, 0, 1, 0, uy), LBM3_FP(pw6, 0, -1, 0, -uy))
            LBM3_PAIR(5, 6, w6, uz, Qz, Kz, LBM3_FP(pw6, 0, 0, 1, uz), LBM3_FP(pw6, 0, 0, -1, -uz))
        } else {
            const double a6 = pw6 * t0, e5 = 3.0 * uz;
            LBM3_PAIR(1, 2, w6, ux, Qx, Kx, a6, a6)
            LBM3_PAIR(3, 4, w6, uy, Qy, Ky, a6, a6)
            LBM3_PAIR(5, 6, w6, uz, Qz, Kz, pw6 * (P.fz * (mp + e5)), pw6 * (P.fz * (mm + e5)))
        }
    }
    const double a12 = pw12 * t0;
    {   // x-y diagonals 7 (+,+), 8 (-,+), 9 (+,-), 10 (-,-)
        const double c7 = ux + uy, c8 = uy - ux;
        const double Bxy = (a2x + a2y) + a1z, cxy = 2.0 * (ux * uy);
        const double Gxy = (wc * gxs + wc * gys) + wd * gzs, exy = we * (gx * gy);
        if (FXY) {
            LBM3_QUAD(7, 8, 9, 10, c7, c8, Bxy, cxy, Gxy, exy, LBM3_FP(pw12, 1, 1, 0, c7), LBM3_FP(pw12, -1, 1, 0, c8),
                      LBM3_FP(pw12, 1, -1, 0, -c8), LBM3_FP(pw12, -1, -1, 0, -c7))
        } else {
            LBM3_QUAD(7, 8, 9, 10, c7, c8, Bxy, cxy, Gxy, exy, a12, a12, a12, a12)
        }
    }
    {   // y-z diagonals 11 (y+,z+), 12 (y-,z+), 13 (y+,z-), 14 (y-,z-)
        const double c11 = uy + uz, c12 = uz - uy;
        const double Byz = (a1x + a2y) + a2z, cyz = 2.0 * (uy * uz);
        const double Gyz = (wd * gxs + wc * gys) + wc * gzs, eyz = we * (gy * gz);
        if (FXY) {
            LBM3_QUAD(11, 12, 13, 14, c11, c12, Byz, cyz, Gyz, eyz, LBM3_FP(pw12, 0, 1, 1, c11), LBM3_FP(pw12, 0, -1, 1, c12),
                      LBM3_FP(pw12, 0, 1, -1, -c12), LBM3_FP(pw12, 0, -1, -1, -c11))
        } else {
            const double e11 = 3.0 * c11, e12 = 3.0 * c12;
            LBM3_QUAD(11, 12, 13, 14, c11, c12, Byz, cyz, Gyz, eyz, pw12 * (P.fz * (mp + e11)), pw12 * (P.fz * (mp + e12)),
                      pw12 * (P.fz * (mm + e12)), pw12 * (P.fz * (mm + e11)))
        }
    }
    {   // x-z diagonals 15 (x+,z+), 16 (x-,z+), 17 (x+,z-), 18 (x-,z-)
        const double c15 = ux + uz, c16 = uz - ux;
        const double Bxz = (a2x + a1y) + a2z, cxz = 2.0 * (ux * uz);
        const double Gxz = (wc * gxs + wd * gys) + wc * gzs, exz = we * (gx * gz);
        if (FXY) {
            LBM3_QUAD(15, 16, 17, 18, c15, c16, Bxz, cxz, Gxz, exz, LBM3_FP(pw12, 1, 0, 1, c15), LBM3_FP(pw12, -1, 0, 1, c16),
                      LBM3_FP(pw12, 1, 0, -1, -c16), LBM3_FP(pw12, -1, 0, -1, -c15))
        } else {
            const double e15 = 3.0 * c15, e16 = 3.0 * c16;
            LBM3_QUAD(15, 16, 17, 18, c15, c16, Bxz, cxz, Gxz, exz, pw12 * (P.fz * (mp + e15)), pw12 * (P.fz * (mp + e16)),
                      pw12 * (P.fz * (mm + e16)), pw12 * (P.fz * (mm + e15)))
        }
    }
#undef LBM3_FP
#undef LBM3_PAIR
#undef LBM3_QUAD
    {   // rest direction: complements (lattice.hpp:461-462, :480)
        const double fe0 = rho - sum, ge0 = ph - sum_phase, fp0 = -sum_force;
        const double g0 = getG(0);
        emit.template operator()<0>(F[0] + (mo * (F[0] - fe0) + fp0), g0 + mop * (g0 - ge0));
    }
}

// Equilibrium used by Lattice::init (lattice.hpp:259-275).  Not the same
// expression as in collide(): init squares through (c*c - 1/3)*u*u, whose
// constants and association differ in the last bit.  Runs once; kept literal.
LBM_HD void init_equilibrium(double rho, double ph, double ux, double uy, double uz, double gx, double gy, double gz,
                             double lap, const Params &P, double *F, double *G)
{
    const int CX[19] = LBM3_CX, CY[19] = LBM3_CY, CZ[19] = LBM3_CZ;
    const double ps = ph * ph;
    const double pb = rho / 3.0 + P.aconst * (-0.5 * ps + 3.0 / 4.0 * ps * ps) - P.kconst * ph * lap;
    const double mu = P.gammaconst * (P.aconst * (-ph + ph * ph * ph) - P.kconst * lap);
    double sum = 0.0, sum_phase = 0.0;
    for (int k = 1; k < 19; k++) {
        const int cx = CX[k], cy = CY[k], cz = CZ[k];
        const int n = cx * cx + cy * cy + cz * cz;
        const double w = n == 1 ? 1.0 / 6.0 : 1.0 / 12.0;
        double wxx, wyy, wzz;
        if (n == 1) {
            wxx = cx ? 5.0 / 12.0 : -1.0 / 3.0;
            wyy = cy ? 5.0 / 12.0 : -1.0 / 3.0;
            wzz = cz ? 5.0 / 12.0 : -1.0 / 3.0;
        } else {
            wxx = cx ? -1.0 / 24.0 : 1.0 / 12.0;
            wyy = cy ? -1.0 / 24.0 : 1.0 / 12.0;
            wzz = cz ? -1.0 / 24.0 : 1.0 / 12.0;
        }
        const double wxy = (cx * cy) * (1.0 / 4.0), wyz = (cy * cz) * (1.0 / 4.0), wzx = (cz * cx) * (1.0 / 4.0);
        const double feq =
            w * (pb + rho * (cx * ux + cy * uy + cz * uz) +
                 1.5 * rho * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy + (cz * cz - 1.0 / 3.0) * uz * uz +
                              2.0 * (ux * uy * cx * cy + ux * uz * cx * cz + uy * uz * cy * cz))) +
            P.kconst * (wxx * gx * gx + wyy * gy * gy + wzz * gz * gz + wxy * gx * gy + wyz * gy * gz + wzx * gx * gz);
        const double geq =
            w * (mu + ph * (cx * ux + cy * uy + cz * uz) +
                 1.5 * ph * ((cx * cx - 1.0 / 3.0) * ux * ux + (cy * cy - 1.0 / 3.0) * uy * uy + (cz * cz - 1.0 / 3.0) * uz * uz +
                             2.0 * (ux * uy * cx * cy + ux * uz * cx * cz + uy * uz * cy * cz)));
        sum += feq;
        sum_phase += geq;
        F[k] = feq;
        G[k] = geq;
    }
    F[0] = rho - sum;
    G[0] = ph - sum_phase;
}

// The four 19-point stencils of lattice.hpp:392-409 from the 19 neighbour values
// PH[k] = phase(x + c_k), accumulated in direction order with the zero-weight
// terms dropped (descriptor.h:86-96).
LBM_HD void stencils(const double *PH, double &gx, double &gy, double &gz, double &lap)
{
    const double s6 = 1.0 / 6.0, s12 = 1.0 / 12.0, l3 = 1.0 / 3.0, l6 = 1.0 / 6.0;
    gx = s6 * PH[1] - s6 * PH[2] + s12 * PH[7] - s12 * PH[8] + s12 * PH[9] - s12 * PH[10] + s12 * PH[15] - s12 * PH[16] +
         s12 * PH[17] - s12 * PH[18];
    gy = s6 * PH[3] - s6 * PH[4] + s12 * PH[7] + s12 * PH[8] - s12 * PH[9] - s12 * PH[10] + s12 * PH[11] - s12 * PH[12] +
         s12 * PH[13] - s12 * PH[14];
    gz = s6 * PH[5] - s6 * PH[6] + s12 * PH[11] + s12 * PH[12] - s12 * PH[13] - s12 * PH[14] + s12 * PH[15] + s12 * PH[16] -
         s12 * PH[17] - s12 * PH[18];
    lap = -4.0 * PH[0] + l3 * PH[1] + l3 * PH[2] + l3 * PH[3] + l3 * PH[4] + l3 * PH[5] + l3 * PH[6] + l6 * PH[7] +
          l6 * PH[8] + l6 * PH[9] + l6 * PH[10] + l6 * PH[11] + l6 * PH[12] + l6 * PH[13] + l6 * PH[14] + l6 * PH[15] +
          l6 * PH[16] + l6 * PH[17] + l6 * PH[18];
}

}  // namespace lbm3
#endif
