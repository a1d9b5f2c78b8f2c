/* lbm_exp.h -- one exp() with the same bits on the host and on the GPU.
 *
 * Why this exists: ShanChen/shanchen.cpp:152-157 calls exp() twenty times per
 * node per step and rounds the result into float accumulators
 * (shanchen.cpp:145-146).  That makes the program sensitive to the last bit of
 * exp(): glibc's exp and CUDA's exp differ by 1 ulp on a small fraction of
 * arguments, and that alone moves rho by ~1e-5 after a few thousand steps
 * (SURVEY.md Appendix C).  Bit parity between the C++ program and the CUDA
 * kernel therefore needs a single exp() that both sides evaluate with exactly
 * the same IEEE-754 operations.
 *
 * lbm_exp() uses only binary64 add, multiply, correctly-rounded fma(), one
 * double->int conversion and an exact power-of-two scale.  It contains no
 * expression a compiler may contract or re-associate, so the result does not
 * depend on -ffp-contract / -fmad settings.  Accuracy is about 1 ulp; it is not
 * correctly rounded and does not need to be -- it only needs to be identical
 * everywhere.
 *
 * To reproduce a CUDA Shan-Chen run with the reference program, include the
 * standard headers, then `#define exp lbm_exp`, then include shanchen.cpp
 * (see INTEGRATION.md); the reference source is not modified.
 */
#ifndef LBM_EXP_H
#define LBM_EXP_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define LBM_EXP_HD __host__ __device__ __forceinline__
#else
#define LBM_EXP_HD static inline
#endif

LBM_EXP_HD double lbm_exp(double x)
{
    /* Out-of-range and NaN handling (never reached by the Shan-Chen path,
     * whose argument is -rho in about [-3, 0]). */
    if (!(x < 709.0)) {
        if (x != x) return x;
        return HUGE_VAL;
    }
    if (x < -708.0) return 0.0;

    /* k = nearest integer to x/ln2, by the 1.5*2^52 shift trick. */
    const double shift = 6755399441055744.0;
    double t = fma(x, 1.4426950408889634074, shift);
    double kd = t - shift;
    int k = (int)kd;

    /* r = x - k*ln2 in two pieces (fdlibm split; hi has 32 trailing zero bits,
     * so k*hi is exact and the first fma is exact). */
    double r = fma(-kd, 6.93147180369123816490e-01, x);
    r = fma(-kd, 1.90821492927058770002e-10, r);

    /* exp(r), |r| <= ln2/2: Taylor to degree 13 (truncation 4e-18), Horner. */
    double p = 1.6059043836821614599e-10;           /* 1/13! */
    p = fma(p, r, 2.0876756987868098979e-09);       /* 1/12! */
    p = fma(p, r, 2.5052108385441718775e-08);       /* 1/11! */
    p = fma(p, r, 2.7557319223985890653e-07);       /* 1/10! */
    p = fma(p, r, 2.7557319223985892511e-06);       /* 1/9!  */
    p = fma(p, r, 2.4801587301587301566e-05);       /* 1/8!  */
    p = fma(p, r, 1.9841269841269841253e-04);       /* 1/7!  */
    p = fma(p, r, 1.3888888888888889419e-03);       /* 1/6!  */
    p = fma(p, r, 8.3333333333333332177e-03);       /* 1/5!  */
    p = fma(p, r, 4.1666666666666664354e-02);       /* 1/4!  */
    p = fma(p, r, 1.6666666666666665741e-01);       /* 1/3!  */
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);

    /* Multiply by 2^k exactly.  k is in [-1022, 1023] here. */
    uint64_t bits = (uint64_t)(int64_t)(k + 1023) << 52;
    double scale;
#if defined(__CUDA_ARCH__)
    scale = __longlong_as_double((long long)bits);
#else
    memcpy(&scale, &bits, sizeof scale);
#endif
    return p * scale;
}

#endif /* LBM_EXP_H */
