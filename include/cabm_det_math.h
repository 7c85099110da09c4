/* include/cabm_det_math.h -- bit-reproducible exp / log for the transmission cycle.
 *
 * Exposure (`U < 1-exp(-hazard)`, Community.cpp:846,868,903,923,937,965) and waning immunity
 * (`remainingEfficacy`, Parameters.h:615-625) go through transcendental functions. Host libm and
 * CUDA libdevice differ in the last ulp, which would turn "integer state transitions bit-exact"
 * into "almost always". These two functions use only IEEE-754 add/mul/div (never contracted into
 * FMA: explicit _rn intrinsics on the device, -ffp-contract=off on the host) and integer bit
 * manipulation, so device and host produce identical bits. Accuracy is ~2 ulp, i.e. within 1e-15
 * relative of libm, far inside the 1e-6 force-of-infection tolerance.
 *
 * This header defines part of the numerical contract of the C ABI (how T is obtained from the
 * hazard); it is shared by the CUDA kernels and by the CPU oracle for that reason only.
 */
#ifndef CABM_DET_MATH_H
#define CABM_DET_MATH_H
#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define CABM_HD __host__ __device__ __forceinline__
#define CABM_MUL(a, b) __dmul_rn((a), (b))
#define CABM_ADD(a, b) __dadd_rn((a), (b))
#define CABM_SUB(a, b) __dsub_rn((a), (b))
#define CABM_DIV(a, b) __ddiv_rn((a), (b))
#elif defined(__CUDACC__)
#define CABM_HD __host__ __device__ __forceinline__
#define CABM_MUL(a, b) ((a) * (b))
#define CABM_ADD(a, b) ((a) + (b))
#define CABM_SUB(a, b) ((a) - (b))
#define CABM_DIV(a, b) ((a) / (b))
#else
#define CABM_HD static inline
#define CABM_MUL(a, b) ((a) * (b))
#define CABM_ADD(a, b) ((a) + (b))
#define CABM_SUB(a, b) ((a) - (b))
#define CABM_DIV(a, b) ((a) / (b))
#endif

CABM_HD uint64_t cabm_d2u(double x) { uint64_t u; memcpy(&u, &x, 8); return u; }
CABM_HD double   cabm_u2d(uint64_t u) { double x; memcpy(&x, &u, 8); return x; }

/* exp(x): x = k ln2 + r, |r| <= ln2/2, degree-14 Taylor in Horner form, scaled by 2^k */
CABM_HD double cabm_exp(double x) {
    if (x != x) return x;
    if (x > 709.0) return cabm_u2d(0x7FF0000000000000ull);
    if (x < -708.0) return 0.0;                       /* subnormal results are not needed by the model */
    const double inv_ln2 = 1.4426950408889634;
    const double ln2_hi = 6.93147180369123816490e-01; /* high part: low 21 bits zero, k*ln2_hi is exact */
    const double ln2_lo = 1.90821492927058770002e-10;
    const double t = CABM_MUL(x, inv_ln2);
    const double kd = (t >= 0.0) ? (double)(int64_t)CABM_ADD(t, 0.5) : (double)(int64_t)CABM_SUB(t, 0.5);
    const int64_t k = (int64_t)kd;
    double r = CABM_SUB(x, CABM_MUL(kd, ln2_hi));
    r = CABM_SUB(r, CABM_MUL(kd, ln2_lo));
    double p = 1.0 / 87178291200.0;                   /* 1/14! */
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 6227020800.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 479001600.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 39916800.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 3628800.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 362880.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 40320.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 5040.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 720.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 120.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 24.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0 / 6.0);
    p = CABM_ADD(CABM_MUL(p, r), 0.5);
    p = CABM_ADD(CABM_MUL(p, r), 1.0);
    p = CABM_ADD(CABM_MUL(p, r), 1.0);
    /* multiply by 2^k; |k| <= 1023 here and p in [0.7, 1.42], result stays normal */
    const double scale = cabm_u2d((uint64_t)(k + 1023) << 52);
    return CABM_MUL(p, scale);
}

/* log(x): x = m 2^e with m in [sqrt(1/2), sqrt(2)), s = (m-1)/(m+1), log m = 2 atanh(s) */
CABM_HD double cabm_log(double x) {
    if (x != x || x < 0.0) return cabm_u2d(0x7FF8000000000000ull);
    if (x == 0.0) return cabm_u2d(0xFFF0000000000000ull);
    uint64_t u = cabm_d2u(x);
    if (u == 0x7FF0000000000000ull) return x;
    int64_t e = 0;
    if ((u >> 52) == 0) { x = CABM_MUL(x, 18014398509481984.0); u = cabm_d2u(x); e = -54; }  /* subnormal */
    e += (int64_t)(u >> 52) - 1023;
    u = (u & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
    double m = cabm_u2d(u);                           /* [1, 2) */
    if (m > 1.4142135623730951) { m = CABM_MUL(m, 0.5); e += 1; }
    const double s = CABM_DIV(CABM_SUB(m, 1.0), CABM_ADD(m, 1.0));
    const double z = CABM_MUL(s, s);
    double p = 1.0 / 27.0;
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 25.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 23.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 21.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 19.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 17.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 15.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 13.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 11.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 9.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 7.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 5.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0 / 3.0);
    p = CABM_ADD(CABM_MUL(p, z), 1.0);
    const double lm = CABM_MUL(CABM_MUL(2.0, s), p);
    const double ed = (double)e;
    const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
    return CABM_ADD(CABM_ADD(CABM_MUL(ed, ln2_hi), lm), CABM_MUL(ed, ln2_lo));
}

/* 1/(1+exp(-l))  (covid::util::logistic, Utility.h:205) */
CABM_HD double cabm_logistic(double l) { return CABM_DIV(1.0, CABM_ADD(1.0, cabm_exp(-l))); }

#endif
