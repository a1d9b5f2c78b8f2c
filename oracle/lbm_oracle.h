/* lbm_oracle.h -- CPU restatement of the reference's collide-stream hot path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; the product
 * (latticeboltzmannmethod_b200/, include/lbm_b200.h) never does.
 *
 * Each function restates one reference routine and cites it.  The restatement
 * is deliberately NOT a transcription: populations are SoA, streaming is a
 * push with the boundary rules folded in (pull-form
 * table), mirror and wall nodes are not stored, and the 3-D path is written
 * per z-slab with ghost planes so the multi-rank protocol can be exercised on
 * the CPU.  What it keeps is the reference's per-node expression order
 * (SURVEY.md Appendix D), which is what makes the results bit-identical.
 *
 * Pinned (see tests/test_oracle_vs_reference.py and tests/golden/): against the
 * unmodified reference compiled into oracle/_ref for every model, zero differing
 * nodes.
 */
#ifndef LBM_ORACLE_H
#define LBM_ORACLE_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- D3Q19 binary liquid (Microchannel3D) ---------------- */
typedef struct lbo3 lbo3;

/* parameter order shared with oracle/refwrap/microchannel_ref.cpp */
enum { LBO3_ACONST, LBO3_KCONST, LBO3_GAMMA, LBO3_PHASE_GRADIENT, LBO3_PHASE_WALL, LBO3_RHO_WALL,
       LBO3_TAU_PHI, LBO3_TAU_LIQ, LBO3_TAU_GAS, LBO3_FORCE_X, LBO3_FORCE_Y, LBO3_FORCE_Z, LBO3_NPARAM };

/* halo buffer ids for lbo3_halo(): what one slab sends (S_*) and where what it
 * receives lands (R_*).  UP = towards larger z. */
enum { LBO3_S_F_UP, LBO3_S_G_UP, LBO3_S_F_DOWN, LBO3_S_G_DOWN,
       LBO3_R_F_FROM_BELOW, LBO3_R_G_FROM_BELOW, LBO3_R_F_FROM_ABOVE, LBO3_R_G_FROM_ABOVE,
       LBO3_S_PHI_UP, LBO3_S_PHI_DOWN, LBO3_R_PHI_FROM_BELOW, LBO3_R_PHI_FROM_ABOVE };

lbo3 *lbo3_create(int nx, int ny, int nz_local, const double *params);
void lbo3_destroy(lbo3 *);
/* fields of the local slab in the reference's node order iZ*NX*NY+iY*NX+iX */
void lbo3_set_fields(lbo3 *, const double *rho, const double *phase, const double *ux, const double *uy,
                     const double *uz);
/* Slab protocol.  A slab owns planes 0..nz-1 plus ghost planes -2,-1,nz,nz+1.
 * After every state change it sends its complete boundary state (f of the
 * outermost plane, g of the two outermost planes) to each ring neighbour and
 * receives theirs into its ghost planes -- four contiguous messages, S_* ->
 * the neighbour's R_*; "UP" is towards larger z.
 *
 * Solver::init (solver.hpp:871-879):
 *   init_fields : updateSymmetric + updateWall           (lattice.hpp:705-760)
 *   [exchange PHI]                                        (solver.hpp:821-868)
 *   init_pops   : Lattice::init, into the back buffers    (lattice.hpp:212-375)
 *   [exchange F,G]  finish  after_init (updateMacro, solver.hpp:877)
 * One time step (Solver::collide_stream, solver.hpp:881-891):
 *   collide : updateMacro + Lattice::collide_stream on planes -1..nz (the two
 *             ghost planes redundantly), into the back buffers
 *   [exchange F,G]
 *   finish  : exchangeMatrices (swap) */
void lbo3_init_fields(lbo3 *);
void lbo3_init_pops(lbo3 *);
void lbo3_after_init(lbo3 *);
void lbo3_collide(lbo3 *, int write_fields);
void lbo3_finish(lbo3 *);
/* pointer + length (doubles) of a halo buffer; contiguous */
double *lbo3_halo(lbo3 *, int id, size_t *count);
/* single-slab conveniences: exchange with itself (z periodic) */
void lbo3_self_exchange_phi(lbo3 *);
void lbo3_self_exchange_step(lbo3 *);
void lbo3_init_single(lbo3 *);           /* the whole init sequence with self exchanges */
void lbo3_step_single(lbo3 *, int n);    /* n x (collide + self exchange + finish), fields of the last step kept */
/* rho,phase,u of the last collide(write_fields=1), all nodes incl. mirror and
 * wall nodes, as the reference's updateMacro/collide_stream leave them */
void lbo3_get_fields(lbo3 *, double *rho, double *phase, double *ux, double *uy, double *uz);
/* populations in the reference's AoS layout [node*19+k]; mirror nodes hold the
 * permuted copy of their inner neighbour, wall nodes 0 (the reference's are junk) */
void lbo3_get_pops(lbo3 *, double *f, double *g);
void lbo3_set_pops(lbo3 *, const double *f, const double *g);
/* EXTENSION, no counterpart in the reference: magic > 0 switches the f collision to two relaxation rates
 * (omega+ = 1/tau, omega- = 1/(magic/(tau-1/2)+1/2)); 0 restores the reference's BGK. */
void lbo3_set_trt(lbo3 *, double magic);

/* ---------------- D2Q9 binary liquid (FreeEnergy/binary.cpp) ---------------- */
typedef struct lbo2b lbo2b;
/* params: aconst,kconst,gammaconst,wall_gradient,tau_gas,tau_liq,omega_phi */
lbo2b *lbo2b_create(int nx, int ny, const double *params7);
void lbo2b_destroy(lbo2b *);
void lbo2b_set_phase(lbo2b *, const double *phase_xmajor);   /* [iX][iY] */
void lbo2b_init_reference_ic(lbo2b *);                       /* binary.cpp:126-132 box */
void lbo2b_init(lbo2b *);                                    /* binary.cpp:136-196 */
void lbo2b_step(lbo2b *, int n);
void lbo2b_get_fields(lbo2b *, double *phase, double *rho, double *ux, double *uy);
void lbo2b_get_pops(lbo2b *, double *f, double *g);          /* [iX][iY][9] */
void lbo2b_set_pops(lbo2b *, const double *f, const double *g);

/* ---------------- D2Q9 Shan-Chen (ShanChen/shanchen.cpp) ---------------- */
typedef struct lbosc lbosc;
/* use_libm_exp: 0 = lbm_exp (bit-portable), 1 = glibc exp (to pin against the stock reference) */
lbosc *lbosc_create(int nx, int ny, double tau, double g, int use_libm_exp);
void lbosc_destroy(lbosc *);
void lbosc_init_droplet(lbosc *, double rhol, double rhog, int radius);  /* shanchen.cpp:87-114 */
void lbosc_set_pops(lbosc *, const double *f);                            /* [i*9+k], i=iY*nx+iX */
void lbosc_step(lbosc *, int n);
void lbosc_get_fields(lbosc *, double *rho, double *u1, double *u2);
void lbosc_get_pops(lbosc *, double *f);

/* ---------------- D2Q9 MRT (MRT/mrt.cpp, MRT/another_mrt.cpp) ---------------- */
typedef struct lbomrt lbomrt;
/* variant 0 = Gram-Schmidt with force (mrt.cpp), 1 = Hermite (another_mrt.cpp) */
lbomrt *lbomrt_create(int variant, int nx, int ny, double omega, double omega_bulk, double force_ampl);
void lbomrt_destroy(lbomrt *);
void lbomrt_init_wave(lbomrt *, double ampl);                 /* mrt.cpp:97-120 / another_mrt.cpp:89-110 */
void lbomrt_set_pops(lbomrt *, const double *f);              /* [iX][iY][9] */
void lbomrt_step(lbomrt *, int n);
void lbomrt_get_fields(lbomrt *, double *rho, double *ux, double *uy);
void lbomrt_get_pops(lbomrt *, double *f);

/* ---------------- D2Q9 TRT diffusion in a masked domain (DiffusionCylinder/cylinder.cpp) ---------------- */
typedef struct lbotrt lbotrt;
/* mask: 1 fluid, 0 anti-bounce-back boundary node, -1 solid; node i = iY*nx+iX (cylinder.cpp:127,135) */
lbotrt *lbotrt_create(int nx, int ny, double omega, double ce, double conc_wall, const int *mask);
void lbotrt_destroy(lbotrt *);
void lbotrt_init(lbotrt *, const double *rho, const double *ux, const double *uy);   /* cylinder.cpp:97-124 */
void lbotrt_step(lbotrt *, int n);                                                    /* :128-205, :208-221, :337-347 */
void lbotrt_get_rho(lbotrt *, double *rho);
void lbotrt_get_pops(lbotrt *, double *f);                                            /* [i*9+k] */
/* the reference's geometry and initial fields (cylinder.cpp:224-312): fills mask[nx*ny], rho, ux, uy */
void lbotrt_reference_setup(int nx, int ny, int radius, double conc_wall, int *mask, double *rho, double *ux, double *uy);

/* ---------------- D2Q9 binary liquid around a curved solid (CurvedSolid/binary_shape.cpp) ---------------- */
typedef struct lbobs lbobs;
/* p[9] = aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi, force_x, force_y
 * (binary_shape.cpp:31,37-48); mask: 1 fluid, 0 bounce-back node, -1 solid, node i = iY*nx+iX (:449-480).
 * Returns NULL when a bounce-back node has no wall normal that points at a fluid node (the reference
 * would then misalign main_dir, :583-587) or lies on the edge of the array. */
lbobs *lbobs_create(int nx, int ny, const double *p, const int *mask);
void lbobs_destroy(lbobs *);
void lbobs_init(lbobs *, const double *rho, const double *phi, const double *ux, const double *uy);  /* :222-258 */
void lbobs_step(lbobs *, int n);                 /* update_phase :413-435, collide_bulk :260-356, update_bounce_back :358-372, stream :617-637 */
void lbobs_get_fields(lbobs *, double *phi, double *rho, double *ux, double *uy);
void lbobs_get_pops(lbobs *, double *f, double *g);   /* [i*9+k] */
/* the reference's geometry and initial fields (:439-520) for any size: disc of `radius` in the centre, three
 * droplets of radius_droplet on it; fills mask, rho, phi, ux, uy */
void lbobs_reference_setup(int nx, int ny, int radius, int radius_droplet, int *mask, double *rho, double *phi, double *ux, double *uy);

#ifdef __cplusplus
}
#endif
#endif
