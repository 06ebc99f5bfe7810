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
 * The algorithms: polynomial logf on [2/3, 4/3); table-driven log (Tang's reduction with
 * Gal's accurate table); Taylor exp after a two-step Cody-Waite reduction; and M. Giles'
 * single-precision erfinv polynomial ("Approximating the erfinv
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

/* ---- draw contract v2: division-free, fma-based, table-driven -------------------------
 * The three functions below are DEFINITIONS (the contract's logf / log / exp), chosen so that
 * the sm_100a kernel evaluates them in a handful of fused multiply-adds: no division, no
 * hardware approximation (MUFU) whose bits a CPU could not restate.  Each operation is one
 * IEEE-754 rounding: +, *, fma.  Measured against extended precision
 * (tests/test_oracle_primitives.py): logf <= 0.94 ulp, log <= 0.80 ulp, exp <= 1.02 ulp.
 * Constants: oracle/det_tables.h (tools/gen_det_tables.py). */
#include "det_tables.h"

/* natural log of a positive, normal, finite float:
 * x = 2^e m, m in [2/3, 4/3); log1p(f) = f - f^2/2 + f^3 Q(f), f = m - 1 */
static inline float det_logf(float x)
{
    uint32_t ix = om_f2u(x);
    uint32_t e = (ix - 0x3f2aaaabu) & 0xff800000u;
    float m = om_u2f(ix - e);
    float fi = (float)((int32_t)e >> 23);
    float f = m - 1.0f;
    float q = PDA_LOGF_Q7;
    q = fmaf(q, f, PDA_LOGF_Q6);
    q = fmaf(q, f, PDA_LOGF_Q5);
    q = fmaf(q, f, PDA_LOGF_Q4);
    q = fmaf(q, f, PDA_LOGF_Q3);
    q = fmaf(q, f, PDA_LOGF_Q2);
    q = fmaf(q, f, PDA_LOGF_Q1);
    q = fmaf(q, f, PDA_LOGF_Q0);
    float r = fmaf(q, f, -0.5f);
    r = r * f;
    r = fmaf(r, f, f);
    return fmaf(fi, 0.693147182f, r);
}

static const double det_log_table[128][2] = PDA_LOG_TABLE_INIT;

/* natural log of a positive, normal, finite double:
 * x = 2^e m, m in [1, 2), j = top 7 mantissa bits, r = fma(m, c_j, -1), |r| < 2^-7,
 * log x = (e + [j >= KSPLIT]) ln2 + L_j + log1p(r), log1p by its Taylor series to r^8 */
static inline double det_log(double x)
{
    const double ln2_hi = 6.93147180369123816490e-01; /* 0x3fe62e42fee00000 */
    const double ln2_lo = 1.90821492927058770002e-10; /* 0x3dea39ef35793c76 */
    uint64_t ux = om_d2u(x);
    uint32_t hx = (uint32_t)(ux >> 32);
    int j = (int)((hx >> 13) & 0x7fu);
    int kk = (int)(hx >> 20) - 1023 + (j >= PDA_LOG_KSPLIT ? 1 : 0);
    double m = om_u2d((ux & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
    double r = fma(m, det_log_table[j][0], -1.0);
    double q = -0.125;
    q = fma(q, r, 1.0 / 7.0);
    q = fma(q, r, -1.0 / 6.0);
    q = fma(q, r, 0.2);
    q = fma(q, r, -0.25);
    q = fma(q, r, 1.0 / 3.0);
    double s = fma(q, r, -0.5);
    s = s * (r * r);
    double dk = (double)kk;
    double lo = fma(dk, ln2_lo, s);
    double t = det_log_table[j][1] + (r + lo);
    return fma(dk, ln2_hi, t);   /* dk ln2_hi is exact (33-bit ln2_hi): one rounding for the large terms */
}

/* exp of a finite double in [-700, 700]: k = nearest integer to x / ln2 (magic-number
 * rounding), r = x - k ln2 in two fused steps, Taylor series to r^13, exponent add */
static inline double det_exp(double x)
{
    const double ln2_hi = 6.93147180369123816490e-01;
    const double ln2_lo = 1.90821492927058770002e-10;
    const double invln2 = 1.44269504088896338700e+00;
    const double magic = 6755399441055744.0; /* 1.5 * 2^52 */
    double t = fma(x, invln2, magic);
    int32_t k = (int32_t)(uint32_t)om_d2u(t);
    double dk = t - magic;
    double r = fma(dk, -ln2_hi, x);
    r = fma(dk, -ln2_lo, r);
    double q = 1.0 / 6227020800.0;      /* 1/13! */
    q = fma(q, r, 1.0 / 479001600.0);   /* 1/12! */
    q = fma(q, r, 1.0 / 39916800.0);
    q = fma(q, r, 1.0 / 3628800.0);
    q = fma(q, r, 1.0 / 362880.0);
    q = fma(q, r, 1.0 / 40320.0);
    q = fma(q, r, 1.0 / 5040.0);
    q = fma(q, r, 1.0 / 720.0);
    q = fma(q, r, 1.0 / 120.0);
    q = fma(q, r, 1.0 / 24.0);
    q = fma(q, r, 1.0 / 6.0);
    q = fma(q, r, 0.5);
    double s = fma(q * r, r, r);
    double y = 1.0 + s;
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
