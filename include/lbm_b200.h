/* lbm_b200.h -- C ABI of liblbm_b200.so, the B200 (sm_100a) replacement for the
 * collide-stream hot path of shurikkuzmin/LatticeBoltzmannMethod.
 *
 * The reference has no library interface: every program is one main() that
 * owns global arrays and runs  moments -> stencil -> equilibrium -> relax ->
 * boundary fix-up -> stream -> swap  in a time loop.  This header is the seam a
 * reference program is cut at: it keeps its parameters, initial condition and
 * output code, and replaces the body of its time loop by lbm_step().  Each
 * entry point names the reference code it stands in for.
 *
 * Conventions
 *  - Plain C, no exceptions cross the boundary, every call returns lbm_status;
 *    lbm_last_error() gives the text for the last failure on that context.
 *  - Host buffers are caller-owned and only touched during the call.  All device
 *    memory, streams and communicators belong to the context.
 *  - One host thread drives a context; a context drives one GPU.  Several
 *    contexts (one per GPU, in one or several processes) form a z-slab ring for
 *    the 3-D model, see "z-slab decomposition" below.
 *  - Host layouts are the reference's own:
 *      LBM_SHANCHEN_D2Q9     node i = iY*nx+iX                 (shanchen.cpp:140)
 *      LBM_BINARY_D2Q9, MRT  node i = iX*ny+iY   ([iX][iY])    (binary.cpp:24-25, mrt.cpp:14-15)
 *      LBM_BINARY_D3Q19      node i = iZ*nx*ny+iY*nx+iX        (lattice.h:136)
 *    populations are array-of-structures [i*Q+k] in the reference's direction order.
 *  - Field semantics: after lbm_step(ctx,n) has finished reference step t, the
 *    fields returned by lbm_get_field() are the ones the reference would write at
 *    step t: the moments taken at the START of step t (3-D and Shan-Chen: velocity
 *    including the F/2rho shift; wall nodes rho_wall / 0; mirror nodes as the
 *    reference derives them).  Populations returned by lbm_get_populations() are
 *    the ones the next step would start from.
 *  - fp64 arithmetic in the reference's operation order, compiled without FMA
 *    contraction: results are bit-identical to the reference C++ programs built
 *    with -ffp-contract=off (Shan-Chen: to the reference built against
 *    include/lbm_exp.h, see that header).
 */
#ifndef LBM_B200_H
#define LBM_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lbm_ctx lbm_ctx;

typedef enum {
    LBM_OK = 0,
    LBM_EINVAL = 1,  /* bad argument or configuration */
    LBM_ENOMEM = 2,  /* device or host allocation failed */
    LBM_ECUDA = 3,   /* a CUDA call or kernel failed (no GPU, wrong architecture, ...) */
    LBM_ENCCL = 4,   /* NCCL missing or failed */
    LBM_ENAN = 5,    /* reserved: NaN is reported through lbm_diagnostics, never aborts */
    LBM_ESTATE = 6   /* call made in the wrong order (e.g. step before init) */
} lbm_status;

/* Two builds export this ABI.  liblbm_b200.so is the fp64 library every parity statement refers to.
 * liblbm_b200_f32.so holds LBM_BINARY_D3Q19 only and stores its populations as float in device memory
 * (arithmetic and all fields stay double; host buffers stay double at every call): half the memory and
 * traffic, results within the tolerance DESIGN.md section 3 states.  No reference counterpart. */
typedef enum {
    LBM_SHANCHEN_D2Q9 = 0,     /* ShanChen/shanchen.cpp */
    LBM_BINARY_D2Q9 = 1,       /* FreeEnergy/binary.cpp */
    LBM_MRT_GS_D2Q9 = 2,       /* MRT/mrt.cpp */
    LBM_MRT_HERMITE_D2Q9 = 3,  /* MRT/another_mrt.cpp */
    LBM_BINARY_D3Q19 = 4,      /* Microchannel3D/{FullChannel,QuarterChannel} */
    LBM_TRT_DIFFUSION_D2Q9 = 5, /* DiffusionCylinder/cylinder.cpp: TRT advection-diffusion, anti-bounce-back on a masked domain */
    LBM_BINARY_SHAPE_D2Q9 = 6  /* CurvedSolid/binary_shape.cpp: binary liquid around a curved solid, wetting through the wall-normal phase gradient */
} lbm_model;

typedef enum { LBM_RHO = 0, LBM_PHASE = 1, LBM_UX = 2, LBM_UY = 3, LBM_UZ = 4 } lbm_field;

/* One-to-one with the reference's constants.  Start from lbm_config_default(). */
typedef struct {
    int model;                 /* lbm_model */
    int nx, ny, nz;            /* GLOBAL grid; nz = 1 for the 2-D models */
    double tau, g;             /* Shan-Chen: shanchen.cpp:21,27 */
    double aconst, kconst, gammaconst;      /* binary.cpp:29-31; main.cpp:49-51 */
    double tau_gas, tau_liq, tau_phi;       /* binary.cpp:36-41 (omega_phi = 1/tau_phi); main.cpp:55-57 */
    double wall_gradient;      /* binary.cpp:32; binary_shape.cpp:31; 3-D: phase_gradient, main.cpp:52 */
    double rho_wall;           /* main.cpp:54 */
    double force[3];           /* main.cpp:58-60 (FullChannel: :50-53); binary_shape.cpp:39-40 (x, y) */
    double omega, omega_bulk;  /* mrt.cpp:53-54; another_mrt.cpp:52-53 */
    double force_ampl;         /* mrt.cpp:98 */
    double ce, conc_wall;      /* cylinder.cpp:43,33; the TRT rates are omega_minus = omega, omega_plus = 2 - omega (:38-40) */
    int device;                /* CUDA device ordinal */
    int slab_rank, slab_count; /* z-slab decomposition (3-D only): this context owns slab `rank` of `count` */
    int kernel_variant;        /* 0 = default (fastest measured).  Other values select alternative kernels of the same
                                  model, kept for cross-checks and held to the same parity tests.  D3Q19: 0 one-pass
                                  step, 5 two-pass (phase kernel + collide kernels), 1 two-pass with the plain
                                  whole-plane kernel, 2 and 6 in-place (AA-pattern) streaming on ONE population array --
                                  half the population memory (replaces the second array and exchangeMatrices,
                                  lattice.hpp:842-855); 2 one pass (the phase field runs ahead of the collision inside
                                  one launch), 6 two passes; on several slabs only the five z-crossing directions travel
                                  (NCCL inside lbm_step after lbm_comm_attach; the lbm_slab_* protocol is for the
                                  two-array forms), and populations can be loaded after lbm_init or an even number of
                                  steps only; Shan-Chen 0 = fused tile kernel, 1 = density pass + collide; binary D2Q9 0 = marching
                                  kernel (one pass, the phase field in a shared-memory ring), 1 = phase pass + collide;
                                  MRT 0 = in-place AA-pattern streaming on ONE population array, 1 = two arrays with
                                  push streaming; TRT diffusion and curved solid: 0 only.  Any other value: LBM_EINVAL.
                                  Results never depend on it. */
    double trt_magic;          /* LBM_BINARY_D3Q19 only, EXTENSION with no counterpart in the reference: 0 (default) keeps
                                  the reference's BGK collision; Lambda > 0 relaxes the antisymmetric half of f with
                                  omega- = 1/(Lambda/(tau-1/2) + 1/2) (two-relaxation-time), everything else unchanged.
                                  Parity unpinned: validated against the BGK limit Lambda = (tau-1/2)^2. */
    double mrt_rates[5];       /* LBM_BINARY_D3Q19 only, EXTENSION with no counterpart in the reference: all 0 (default)
                                  keeps BGK; otherwise f collides in the moment space of d'Humieres et al. (2002) with
                                  the stress moments at 1/tau (the reference's viscosity) and these rates for the energy,
                                  energy-square, heat-flux, fourth-order and third-order moments (an entry <= 0 means
                                  1/tau; e.g. {1.19, 1.4, 1.2, 1.4, 1.98}).  Excludes trt_magic.  Parity unpinned:
                                  validated against the BGK limit (all rates 1/tau) and by momentum/mass balances. */
} lbm_config;

/* Fills *cfg with the constants of the reference program for `model`
 * (QuarterChannel/main.cpp for the 3-D model) on device 0, one slab. */
lbm_status lbm_config_default(lbm_config *cfg, int model);

/* Replaces: the global arrays / `new double[]` of every program
 * (shanchen.cpp:29-33, binary.cpp:24-25, mrt.cpp:14-15, lattice.h:35-55) and the
 * Solver constructor's slab bounds (solver.hpp:25-30). */
lbm_status lbm_create(const lbm_config *cfg, lbm_ctx **out);
void lbm_destroy(lbm_ctx *ctx);
const char *lbm_last_error(const lbm_ctx *ctx); /* ctx may be NULL: error of the last failed lbm_create */

/* Slab geometry of this context: first owned global z plane and plane count
 * (solver.hpp:25-30: nz/count planes each, the last slab takes the remainder). */
lbm_status lbm_slab_extent(const lbm_ctx *ctx, int *z_begin, int *nz_local);
size_t lbm_local_nodes(const lbm_ctx *ctx); /* nx*ny*nz_local */

/* Replaces: Solver::putDensity/putPhase/putVelocity (solver.hpp:58-118) and the
 * 2-D programs' direct writes to rho/phase/u.  n = lbm_local_nodes(). */
lbm_status lbm_set_field(lbm_ctx *ctx, int field, const double *host, size_t n);

/* The two masked models.  Replaces: the geometry array the program builds in initialize_geometry()
 * (cylinder.cpp:224-283; binary_shape.cpp:439-480) together with bb_nodes/dirs/main_dir (cylinder.cpp:264-283,
 * :303-312; binary_shape.cpp:485-588), which the library derives from it.  mask[iY*nx+iX]: 1 fluid,
 * 0 boundary (anti-bounce-back / bounce-back) node, -1 solid; n = lbm_local_nodes().  Call before lbm_init.
 * LBM_EINVAL where the reference itself would read outside its arrays or from nodes it never updates:
 *   both: a fluid node touches a solid node;
 *   LBM_TRT_DIFFUSION_D2Q9 (not periodic): a fluid node lies on the edge of the array;
 *   LBM_BINARY_SHAPE_D2Q9 (periodic): a boundary node lies on the edge of the array, or its wall normal
 *   (binary_shape.cpp:556-588) does not exist or does not point at a fluid node. */
lbm_status lbm_set_geometry(lbm_ctx *ctx, const int *mask, size_t n);

/* Replaces: init() of binary.cpp:122-198; the init loops of shanchen.cpp:87-114
 * (from the rho field and the velocity fields, which are zero in a new context as the program's globals are;
 * after lbm_step they hold the last output, so set them again before re-initialising), mrt.cpp:97-120 / another_mrt.cpp:89-110 and cylinder.cpp:97-124 (from rho,ux,uy),
 * binary_shape.cpp:222-258 (from rho,phase,ux,uy);
 * Solver::init (solver.hpp:871-879) incl. its phase halo exchange. */
lbm_status lbm_init(lbm_ctx *ctx);

/* Test / restart hooks with no reference counterpart other than the arrays being
 * public: load or read back raw populations. set: 0 = f, 1 = g.  n = nodes*Q.
 * With slab_count > 1, lbm_set_populations refreshes the ghost planes through the attached communicator
 * (LBM_ESTATE without lbm_comm_attach: a caller that moves the halos itself restarts through lbm_set_field + lbm_slab_*). */
lbm_status lbm_set_populations(lbm_ctx *ctx, int set, const double *host_aos, size_t n);
lbm_status lbm_get_populations(lbm_ctx *ctx, int set, double *host_aos, size_t n);

/* Replaces the body of the time loop: shanchen.cpp:120-188, binary.cpp:336-350
 * (collide_bulk, update_bounce_back, streaming), mrt.cpp:144-251, cylinder.cpp:355-358
 * (collide, update_bounce_back, stream), binary_shape.cpp:666-669 (update_phase, collide_bulk,
 * update_bounce_back, stream),
 * another_mrt.cpp:112-214, Solver::collide_stream (solver.hpp:881-891: updateMacro,
 * updatePhase, Lattice::collide_stream, propagate, exchangeMatrices).
 * Asynchronous: returns once the work is queued; any call that reads results waits. */
lbm_status lbm_step(lbm_ctx *ctx, int nsteps);
lbm_status lbm_sync(lbm_ctx *ctx);

/* Replaces: reading rho/phase/u before the writers (shanchen.cpp:190-207,
 * binary.cpp:357-384, mrt.cpp:254-270) and Solver::getWhole* for one slab
 * (solver.hpp:121-277).  n = lbm_local_nodes(). */
lbm_status lbm_get_field(lbm_ctx *ctx, int field, double *host, size_t n);
/* Same copy, queued only: `host` (page-locked memory from lbm_host_alloc, or the call degenerates to the blocking
 * form) holds the field after the next lbm_sync.  Lets a host program queue all five getWhole* copies of
 * solver.hpp:121-277 and wait once.  lbm_set_field is queued the same way: a pageable source is staged before the
 * call returns, a page-locked one must stay unchanged until the next call that waits. */
lbm_status lbm_get_field_async(lbm_ctx *ctx, int field, double *host, size_t n);

/* Replaces: calculateDensities' sums (shanchen.cpp:49-70), the mass print-out of
 * CurvedSolid (binary_shape.cpp:654-661) and the NaN probes (binary.cpp:359,
 * solver.hpp:49-55).  Sums run over the fluid nodes of this slab (no mirror or
 * wall nodes) of the populations the next step starts from. Any pointer may be NULL. */
lbm_status lbm_diagnostics(lbm_ctx *ctx, double *mass, double momentum[3], double *phase_mass, int *has_nan);

/* Plumbing for callers that time or order work themselves. */
void *lbm_get_stream(lbm_ctx *ctx);                /* cudaStream_t the step runs on */
lbm_status lbm_set_stream(lbm_ctx *ctx, void *cuda_stream);
unsigned long long lbm_launch_count(const lbm_ctx *ctx); /* kernels launched by this context so far */
unsigned long long lbm_device_bytes(const lbm_ctx *ctx); /* device memory the context's model holds right now (the sum of
                                                      its allocations: populations, fields, scratch) */
int lbm_population_bytes(const lbm_ctx *ctx);      /* storage size of one population value: 8, or 4 in liblbm_b200_f32.so;
                                                      the element size of the F/G halo messages of lbm_slab_halo */
/* Debugging aid with no reference counterpart: with LBM_GUARD=1 in the environment at lbm_create, every device buffer
 * of the context lies between two 1 MiB guard bands holding a byte pattern; this returns how many guard bytes have
 * been overwritten so far (0 = no kernel wrote outside its buffers by less than 1 MiB; always 0 without LBM_GUARD).
 * Waits for the stream. */
unsigned long long lbm_guard_check(lbm_ctx *ctx);
/* D3Q19 one-pass step (kernel_variant 0): number of steps so far in which a block gave up waiting for the blocks its
 * phase-field part depends on and the phase field was recomputed by the separate kernel instead (results are the
 * same either way; a non-zero value only says the step was slower than it should be).  Waits for the stream. */
unsigned long long lbm_phase_fallbacks(lbm_ctx *ctx);
/* Per-kernel device timing (CUDA events around every kernel launched while enabled); the report is a JSON
 * object {"kernel": {"launches": n, "ms": total}} and clears the record.  Replaces the reference's only
 * instrumentation, time(NULL) around the loop (shanchen.cpp:117-118,212-214). */
lbm_status lbm_profile(lbm_ctx *ctx, int enable);
lbm_status lbm_profile_report(lbm_ctx *ctx, char *buf, size_t buf_bytes);
void *lbm_host_alloc(size_t bytes);                /* page-locked host memory for fast set/get */
void lbm_host_free(void *p);

/* ---- z-slab decomposition (replaces solver.hpp:821-868 updatePhase and
 * :893-928 propagate with lattice.hpp:690-703, 809-840) ----
 * Every slab keeps two ghost planes per side and collides the inner ghost plane
 * redundantly, so each exchange is four contiguous messages per slab and step.
 * Either let the library move them with NCCL (attach once, then lbm_init /
 * lbm_step do everything), or drive the phases yourself with any transport. */
#define LBM_COMM_ID_BYTES 128
lbm_status lbm_comm_unique_id(void *id_out);                   /* ncclGetUniqueId; rank 0 then broadcasts it */
lbm_status lbm_comm_attach(lbm_ctx *ctx, const void *id);      /* ncclCommInitRank(slab_count, id, slab_rank) */

typedef enum {
    LBM_HALO_S_F_UP = 0, LBM_HALO_S_G_UP, LBM_HALO_S_F_DOWN, LBM_HALO_S_G_DOWN,
    LBM_HALO_R_F_FROM_BELOW, LBM_HALO_R_G_FROM_BELOW, LBM_HALO_R_F_FROM_ABOVE, LBM_HALO_R_G_FROM_ABOVE,
    LBM_HALO_S_PHI_UP, LBM_HALO_S_PHI_DOWN, LBM_HALO_R_PHI_FROM_BELOW, LBM_HALO_R_PHI_FROM_ABOVE
} lbm_halo_id;
/* device pointer and length in ELEMENTS of one halo message -- F/G messages: elements of lbm_population_bytes()
 * bytes (double; float in liblbm_b200_f32.so), PHI messages: doubles; S_x_UP pairs with the upper
 * neighbour's R_x_FROM_BELOW, S_x_DOWN with the lower neighbour's R_x_FROM_ABOVE */
lbm_status lbm_slab_halo(lbm_ctx *ctx, int halo_id, void **device_ptr, size_t *count);
lbm_status lbm_slab_init_fields(lbm_ctx *ctx);  /* then exchange PHI */
lbm_status lbm_slab_init_pops(lbm_ctx *ctx);    /* then exchange F,G, then lbm_slab_finish + lbm_slab_after_init */
lbm_status lbm_slab_after_init(lbm_ctx *ctx);
lbm_status lbm_slab_collide(lbm_ctx *ctx, int write_fields); /* then exchange F,G, then lbm_slab_finish */
lbm_status lbm_slab_finish(lbm_ctx *ctx);

/* ---- allocation-free population sums ----
 * No counterpart in the reference (its sums run on the host over arrays the program already owns).
 * Same six sums as lbm_diagnostics -- sum f, sum f*cx, sum f*cy, sum f*cz, sum g, number of NaNs --
 * over n_nodes consecutive nodes of structure-of-arrays populations (direction k of node i at
 * f_dev[i + k*stride_k]; q = 9 or 19; g_dev may be NULL), written to sums6_dev (device memory, 6 doubles),
 * queued on cuda_stream.  For callers that cannot allocate -- all device memory already committed to
 * their own pool, or running where the implicit device-wide wait of cudaMalloc/cudaFree is not
 * acceptable: it needs no context, and calls no cudaMalloc, cudaFree or any allocating routine.  The block
 * partials live in a file-scope device array of LBM_NOALLOC_MAX_BLOCKS*6 doubles (768 KiB) that every
 * GPU loading the library holds until exit; one block covers 8192 nodes, so n_nodes is limited to
 * LBM_NOALLOC_MAX_BLOCKS*8192 = 134 217 728 (LBM_EINVAL beyond that).  All calls share that array: calls
 * whose kernels overlap on one GPU race -- serialise them.  Nothing needs clearing between calls (each
 * call overwrites every entry it reads).  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress); make that first call before any stream capture.
 * liblbm_b200_f32.so exports the same function with the same signature: it sums `double` arrays there too, so it
 * cannot be pointed at that build's float-stored populations (use lbm_diagnostics on the context for those). */
#define LBM_NOALLOC_MAX_BLOCKS 16384
lbm_status lbm_population_sums_noalloc(const double *f_dev, const double *g_dev, int q, size_t n_nodes, size_t stride_k,
                                       void *cuda_stream, double *sums6_dev);
const char *lbm_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LBM_B200_H */
