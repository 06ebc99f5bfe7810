/* oracle/det_math.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * Deterministic elementary functions of the PDA draw contract,
 * written in plain C with every rounding step spelled out so that a CPU build
 * (gcc -ffp-contract=off) and the CUDA product (its own copy in
 * gsmfretda_b200/csrc/pda_math.cuh, compiled -fmad=false) produce the same bits.
 * Only + - * / sqrt and explicit fma()/fmaf() are used; no libm transcendental.
 *
 * They stand in for the libdevice calls of the reference:
 *   det_logf        <- logf      in drawTau            (/root/reference/src/gen_rand.cuh:56)
 *   det_norm_icdf   <- curand_normal on a Sobol64 state = sqrt(2)*erfcinvf(2p)
 *                                 in drawE             (/root/reference/src/gen_rand.cuh:65,
 *                                 /usr/local/cuda/include/curand_normal.h:240-257)
 *   det_log/det_exp <- log/exp   in rk_binomial_*      (/root/reference/src/binom.cuh:81,86,109,120-122,145)
 *                      exp       in cuRAND's Poisson   (curand_poisson.h:436-463)
 * The algorithms are the classic Sun fdlibm argument reductions (log, logf, exp)
 * and M. Giles' single-precision erfinv polynomial ("Approximating the erfinv
 * function", GPU Computing Gems 2011).
 */
#ifndef ORACLE_DET_MATH_H
#define ORACLE_DET_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline uint32_t om_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float om_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint64_t om_d2u(double f) { uint64_t u; memcpy(&u, &f, 8); return u; }
static inline double om_u2d(uint64_t u) { double f; memcpy(&f, &u, 8); return f; }

/* natural log of a positive, normal, finite float. */
static inline float det_logf(float x)
{
    const float ln2_hi = 6.9313812256e-01f; /* 0x3f317180 */
    const float ln2_lo = 9.0580006145e-06f; /* 0x3717f7d1 */
    const float Lg1 = 0.66666662693f;       /* 0xaaaaaa.0p-24 */
    const float Lg2 = 0.40000972152f;       /* 0xccce13.0p-25 */
    const float Lg3 = 0.28498786688f;       /* 0x91e9ee.0p-25 */
    const float Lg4 = 0.24279078841f;       /* 0xf89e26.0p-26 */
    uint32_t ix = om_f2u(x);
    ix += 0x3f800000u - 0x3f3504f3u;
    int k = (int)(ix >> 23) - 0x7f;
    ix = (ix & 0x007fffffu) + 0x3f3504f3u;
    float m = om_u2f(ix);
    float f = m - 1.0f;
    float s = f / (2.0f + f);
    float z = s * s;
    float w = z * z;
    float t1 = w * (Lg2 + w * Lg4);
    float t2 = z * (Lg1 + w * Lg3);
    float R = t2 + t1;
    float hfsq = 0.5f * f * f;
    float dk = (float)k;
    return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
}

/* natural log of a positive, normal, finite double. */
static inline double det_log(double x)
{
    const double ln2_hi = 6.93147180369123816490e-01;
    const double ln2_lo = 1.90821492927058770002e-10;
    const double Lg1 = 6.666666666666735130e-01;
    const double Lg2 = 3.999999999940941908e-01;
    const double Lg3 = 2.857142874366239149e-01;
    const double Lg4 = 2.222219843214978396e-01;
    const double Lg5 = 1.818357216161805012e-01;
    const double Lg6 = 1.531383769920937332e-01;
    const double Lg7 = 1.479819860511658591e-01;
    uint64_t ux = om_d2u(x);
    uint32_t hx = (uint32_t)(ux >> 32);
    hx += 0x3ff00000u - 0x3fe6a09eu;
    int k = (int)(hx >> 20) - 0x3ff;
    hx = (hx & 0x000fffffu) + 0x3fe6a09eu;
    ux = ((uint64_t)hx << 32) | (ux & 0xffffffffull);
    double m = om_u2d(ux);
    double f = m - 1.0;
    double hfsq = 0.5 * f * f;
    double s = f / (2.0 + f);
    double z = s * s;
    double w = z * z;
    double t1 = w * (Lg2 + w * (Lg4 + w * Lg6));
    double t2 = z * (Lg1 + w * (Lg3 + w * (Lg5 + w * Lg7)));
    double R = t2 + t1;
    double dk = (double)k;
    return s * (hfsq + R) + dk * ln2_lo - hfsq + f + dk * ln2_hi;
}

/* exp of a finite double in [-700, 700]. */
static inline double det_exp(double x)
{
    const double ln2HI = 6.93147180369123816490e-01;
    const double ln2LO = 1.90821492927058770002e-10;
    const double invln2 = 1.44269504088896338700e+00;
    const double P1 = 1.66666666666666019037e-01;
    const double P2 = -2.77777777770155933842e-03;
    const double P3 = 6.61375632143793436117e-05;
    const double P4 = -1.65339022054652515390e-06;
    const double P5 = 4.13813679705723846039e-08;
    double half = x < 0.0 ? -0.5 : 0.5;
    int k = (int)(invln2 * x + half); /* truncation toward zero */
    double dk = (double)k;
    double hi = x - dk * ln2HI;
    double lo = dk * ln2LO;
    double r = hi - lo;
    double rr = r * r;
    double c = r - rr * (P1 + rr * (P2 + rr * (P3 + rr * (P4 + rr * P5))));
    double y = 1.0 + (r * c / (2.0 - c) - lo + hi);
    /* scale by 2^k, k in [-1010, 1010] so a single exponent add is exact */
    uint64_t uy = om_d2u(y);
    uy += (uint64_t)((int64_t)k << 52);
    return om_u2d(uy);
}

/* Standard normal deviate from one 32-bit word, by inverse CDF, with cuRAND's
 * mirroring (curand_normal.h:240-257): z = s*sqrt(2)*erfcinv(2p),
 * p = t*2^-32 + 2^-33 in (0, 0.5].  erfcinv(2p) = erfinv(1-2p) evaluated with
 * Giles' polynomial on w = -log(4p(1-p)). */
static inline float det_norm_icdf(uint32_t t)
{
    float sgn = -1.41421356237309504880f;   /* CURAND_SQRT2 is NEGATIVE (curand_globals.h:64): z = -sqrt(2) erfcinv(2p) for t <= 2^31 */
    if (t > 0x80000000u) { t = 0xffffffffu - t; sgn = -sgn; }
    float p = (float)t * 2.3283064365386963e-10f + 1.1641532182693481e-10f;
    float x = 1.0f - 2.0f * p;                 /* in [0, 1) */
    float w = -det_logf(4.0f * p * (1.0f - p));
    float r;
    if (w < 5.0f) {
        w = w - 2.5f;
        r = 2.81022636e-08f;
        r = fmaf(r, w, 3.43273939e-07f);
        r = fmaf(r, w, -3.5233877e-06f);
        r = fmaf(r, w, -4.39150654e-06f);
        r = fmaf(r, w, 0.00021858087f);
        r = fmaf(r, w, -0.00125372503f);
        r = fmaf(r, w, -0.00417768164f);
        r = fmaf(r, w, 0.246640727f);
        r = fmaf(r, w, 1.50140941f);
    } else {
        w = sqrtf(w) - 3.0f;
        r = -0.000200214257f;
        r = fmaf(r, w, 0.000100950558f);
        r = fmaf(r, w, 0.00134934322f);
        r = fmaf(r, w, -0.00367342844f);
        r = fmaf(r, w, 0.00573950773f);
        r = fmaf(r, w, -0.0076224613f);
        r = fmaf(r, w, 0.00943887047f);
        r = fmaf(r, w, 1.00167406f);
        r = fmaf(r, w, 2.83297682f);
    }
    return sgn * (r * x);
}

/* x^6 of the E draw (powf(rd/r0, 6) in /root/reference/src/gen_rand.cuh:67):
 * three roundings, fixed order. */
static inline float det_pow6f(float x)
{
    float x2 = x * x;
    float x4 = x2 * x2;
    return x4 * x2;
}

/* x^(1/6) of the E draw (powf(1/e-1, 1.0/6) in /root/reference/src/gen_rand.cuh:65)
 * = exp(log(x)/6) through the double routines, rounded once to float.
 * x <= 0 (E >= 1, outside the model's domain) gives 0. */
static inline float det_root6f(float x)
{
    if (!(x > 0.0f)) return 0.0f;
    double xd = (double)x;
    if (xd < 2.2250738585072014e-308) return 0.0f;
    return (float)det_exp(det_log(xd) / 6.0);
}

#endif
