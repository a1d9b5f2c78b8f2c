// lbm_pop.h -- the storage type of the D3Q19 populations in device memory.
// Default build (liblbm_b200.so): double -- every parity statement in this repository is about this build.
// -DLBM_POP_FLOAT (liblbm_b200_f32.so): populations are STORED as float and widened on load; all arithmetic,
// the phase field and the macroscopic fields stay double.  Halves the bytes per lattice update (304 instead of
// 608); results then differ from the fp64 build by float rounding of the stored state (tolerance: DESIGN.md).
#ifndef LBM_POP_H
#define LBM_POP_H
#ifdef LBM_POP_FLOAT
typedef float pop_t;
#else
typedef double pop_t;
#endif
#endif
