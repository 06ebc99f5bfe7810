/* oracle/pda_oracle.c — TEST INFRASTRUCTURE ONLY (see pda_oracle.h, oracle/README.md).
 *
 * Plain-C restatement of the reference's Monte-Carlo PDA evaluation on the draw
 * contract of DESIGN.md §4 (Philox4x32-10 counter draws for the walk, the reference's
 * xor128 stream for the binomial, explicit-rounding elementary functions).
 * Build: gcc -O2 -ffp-contract=off -mfma -fopenmp (oracle/Makefile). No fast-math.
 *
 * Each function cites the reference lines it follows (paths under /root/reference).
 */
#include "pda_oracle.h"
#include "det_math.h"
#include "philox.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* physics constants: default arguments of mc_kernel, never overridden by the launch
 * (src/mc.cu:89-91 vs src/mc.cu:571-576) */
#define PDA_GAMMA 0.34f
#define PDA_DEXDIRAEM 0.08f
#define PDA_DCH2ACH 0.07f
#define PDA_R0 52.0f

/* ---- uniform in (0,1] from one 32-bit word -------------------------------------
 * cuRAND's float mapping of the high word of a Sobol64 draw
 * (curand_uniform.h: x*2^-32 + 2^-33), reached from gen_rand.cuh:37,54,101 */
static inline float u01(uint32_t w)
{
    return (float)w * 2.3283064365386963e-10f + 1.1641532182693481e-10f;
}

/* ---- drawDisIdx, src/gen_rand.cuh:35-50 ---------------------------------------- */
static inline int draw_dis_idx(int n, const float *p, float pv)
{
    float a = 0.0f;
    for (int i = 0; i < n; ++i) {
        a += p[i];
        if (a > pv) return i;
    }
    return n - 1;
}

/* ---- drawTau with randN = 0 (src/gen_rand.cuh:52-62, call site src/mc.cu:122):
 * the jitter branch needs randN > 0 and is dead. */
static inline float draw_tau(float k, float pv)
{
    return det_logf(pv) / (-k);
}

/* ---- drawE, src/gen_rand.cuh:63-68 --------------------------------------------- */
static inline float draw_e(float e, float r0, float v, uint32_t wn)
{
    float rd = det_norm_icdf(wn) * v + r0 * det_root6f(1.0f / e - 1.0f);
    return 1.0f / (1.0f + det_pow6f(rd / r0));
}

/* ---- drawJ_Si2Sj, src/gen_rand.cuh:78-89: column i of P_i2j without its diagonal */
static inline int draw_j(const float *P_i2j, int n, int i, float pv)
{
    float col[PDA_MAX_STATES];
    for (int ii = 0; ii < n - 1; ++ii)
        col[ii] = (ii >= i) ? P_i2j[i * n + ii + 1] : P_i2j[i * n + ii];
    int j = draw_dis_idx(n - 1, col, pv);
    if (j >= i) ++j;
    return j;
}

/* ---- Poisson by inversion: cuRAND's _curand_poisson_ITR_double
 * (/usr/local/cuda/include/curand_poisson.h:436-463, Fishman's algorithm), which is
 * what curand_poisson(Sobol64 state, lambda < 1000) resolves to (ibid. :546-551,759).
 * lambda >= 1000 (normal approximation) is outside the contract; the product
 * rejects such inputs at init_data. */
static uint32_t poisson_itr(uint32_t hi, uint32_t lo, double lambda)
{
    uint64_t x = ((uint64_t)hi << 32) | lo;
    double u = (double)(x >> 11) * 1.1102230246251565e-16 + 2.7755575615628914e-17;
    double p = 1.0, q = 1.0, L;
    uint32_t k = 0;
    int pw = 0;
    do {
        if (lambda > (double)pw + 500.0)
            L = det_exp(-500.0);
        else
            L = det_exp((double)pw - lambda);
        p *= L;
        q *= L;
        pw += 500;
        while (u > q) {
            k++;
            p *= lambda / (double)k;
            q += p;
        }
    } while ((double)pw < lambda);
    return k;
}

/* ---- xor128 of rk_state, src/binom.cuh:7-19,190-197 ---------------------------- */
static inline uint32_t rk_random(uint32_t s[4])
{
    uint32_t t = s[0] ^ (s[0] << 11);
    s[0] = s[1];
    s[1] = s[2];
    s[2] = s[3];
    s[3] = s[3] ^ (s[3] >> 19) ^ t ^ (t >> 8);
    return s[3];
}

static inline double rk_double(uint32_t s[4])
{
    uint32_t a = rk_random(s) >> 5;
    uint32_t b = rk_random(s) >> 6;
    return ((double)a * 67108864.0 + (double)b) / 9007199254740992.0;
}

static inline void rk_seed(uint64_t s, uint32_t st[4])
{
    for (uint64_t i = 1; i <= 4; ++i) {
        s = 1812433253ull * (s ^ (s >> 30)) + i;
        st[i - 1] = (uint32_t)s;
    }
}

/* ---- rk_binomial_inversion, src/binom.cuh:135-169.  The rk_state cache only saves
 * recomputation of (q, qn, np, bound) for a repeated (n, p); it never changes a result. */
static int64_t binomial_inversion(uint32_t st[4], int n, double p)
{
    double q = 1.0 - p;
    double qn = det_exp((double)n * det_log(q));
    double np = (double)n * p;
    double bd = np + 10.0 * sqrt(np * q + 1.0);
    if (!(bd <= (double)n)) bd = (double)n; /* min(); a NaN candidate (n < 0) yields n */
    int64_t bound = (int64_t)bd;
    int64_t X = 0;
    double px = qn;
    double U = rk_double(st);
    while (U > px) {
        X++;
        if (X > bound) {
            X = 0;
            px = qn;
            U = rk_double(st);
        } else {
            U -= px;
            px = ((double)((int64_t)n - X + 1) * p * px) / ((double)X * q);
        }
    }
    return X;
}

/* Stirling-type correction of the BTPE squeeze, src/binom.cuh:122-125 */
static inline double btpe_corr(double x, double x2)
{
    return (13680. - (462. - (132. - (99. - 140. / x2) / x2) / x2) / x2) / x / 166320.;
}

/* ---- rk_binomial_btpe, src/binom.cuh:21-134 (Kachitvichyanukul & Schmeiser 1988).
 * Called with p <= 0.5 only (rk_binomial mirrors first), so r = p and the final
 * "p > 0.5" flip at Step60 never fires; it is kept for fidelity. */
static int64_t binomial_btpe(uint32_t st[4], int64_t n, double p)
{
    double r = p < 1.0 - p ? p : 1.0 - p;
    double q = 1.0 - r;
    double fm = (double)n * r + r;
    int64_t m = (int64_t)floor(fm);
    double p1 = floor(2.195 * sqrt((double)n * r * q) - 4.6 * q) + 0.5;
    double xm = (double)m + 0.5;
    double xl = xm - p1;
    double xr = xm + p1;
    double c = 0.134 + 20.5 / (15.3 + (double)m);
    double a = (fm - xl) / (fm - xl * r);
    double laml = a * (1.0 + a / 2.0);
    a = (xr - fm) / (xr * q);
    double lamr = a * (1.0 + a / 2.0);
    double p2 = p1 * (1.0 + 2.0 * c);
    double p3 = p2 + c / laml;
    double p4 = p3 + c / lamr;
    double nrq = (double)n * r * q;
    int64_t y;
    for (;;) {
        double u = rk_double(st) * p4;
        double v = rk_double(st);
        if (u <= p1) { /* triangle: accept at once */
            y = (int64_t)floor(xm - p1 * v + u);
            break;
        }
        if (u <= p2) { /* parallelogram */
            double x = xl + (u - p1) / c;
            v = v * c + 1.0 - fabs((double)m - x + 0.5) / p1;
            if (v > 1.0) continue;
            y = (int64_t)floor(x);
        } else if (u <= p3) { /* left tail */
            y = (int64_t)floor(xl + det_log(v) / laml);
            if (y < 0) continue;
            v = v * (u - p2) * laml;
        } else { /* right tail */
            y = (int64_t)floor(xr - det_log(v) / lamr);
            if (y > n) continue;
            v = v * (u - p3) * lamr;
        }
        /* Step50: acceptance */
        int64_t k = y > m ? y - m : m - y;
        if (!((k > 20) && ((double)k < nrq / 2.0 - 1.0))) {
            double s = r / q;
            double aa = s * (double)(n + 1);
            double F = 1.0;
            if (m < y) {
                for (int64_t i = m + 1; i <= y; ++i) F *= (aa / (double)i - s);
            } else if (m > y) {
                for (int64_t i = y + 1; i <= m; ++i) F /= (aa / (double)i - s);
            }
            if (v > F) continue;
            break;
        }
        /* Step52: squeeze */
        double kd = (double)k;
        double rho = (kd / nrq) * ((kd * (kd / 3.0 + 0.625) + 0.16666666666666666) / nrq + 0.5);
        double t = -kd * kd / (2.0 * nrq);
        double A = det_log(v);
        if (A < t - rho) break;
        if (A > t + rho) continue;
        double x1 = (double)(y + 1);
        double f1 = (double)(m + 1);
        double z = (double)(n + 1 - m);
        double w = (double)(n - y + 1);
        double x2 = x1 * x1, f2 = f1 * f1, z2 = z * z, w2 = w * w;
        double lim = xm * det_log(f1 / x1) + ((double)(n - m) + 0.5) * det_log(z / w) +
                     (double)(y - m) * det_log(w * r / (x1 * q)) + btpe_corr(f1, f2) +
                     btpe_corr(z, z2) + btpe_corr(x1, x2) + btpe_corr(w, w2);
        if (A > lim) continue;
        break;
    }
    if (p > 0.5) y = n - y;
    return y;
}

/* ---- rk_binomial, src/binom.cuh:172-188 ---------------------------------------- */
static int64_t rk_binomial(uint32_t st[4], int n, double p)
{
    if (p <= 0.5) {
        if (p * (double)n <= 30.0) return binomial_inversion(st, n, p);
        return binomial_btpe(st, n, p);
    }
    double q = 1.0 - p;
    if (q * (double)n <= 30.0) return (int64_t)n - binomial_inversion(st, n, q);
    return (int64_t)n - binomial_btpe(st, n, q);
}

/* ---- csd_int_multinomial, src/binom.cuh:234-263 (used by the reference's tests only) */
void pda_oracle_multinomial(uint32_t st[4], int K, int N, const float *p, int *out)
{
    float norm = 0.0f, sum_p = 0.0f;
    int sum_n = 0;
    for (int k = 0; k < K; ++k) norm += p[k];
    for (int k = 0; k < K; ++k) {
        if (p[k] > 0.0f)
            out[k] = (int)rk_binomial(st, N - sum_n, (double)(p[k] / (norm - sum_p)));
        else
            out[k] = 0;
        sum_p += p[k];
        sum_n += out[k];
    }
}

/* ---- mc::set_params + genMatK/genP_i2j/genMatP ---------------------------------
 * src/mc.cu:790-803 (split of args), src/eigenhelper.cpp:4-54 */
int pda_oracle_params(int n, const float *args, int n_args, pda_params_t *out)
{
    if (n < 1 || n > PDA_MAX_STATES) return 0;
    if (n_args != n * n + n) return 0;
    if (n * n - n < 1) return 0; /* genMatK: args.size() < 1 */
    memset(out, 0, sizeof(*out));
    out->n = n;
    const float *e = args, *k = args + n, *v = args + n * n;
    for (int i = 0; i < n; ++i) {
        out->E[i] = e[i];
        out->v[i] = v[i];
    }
    float *K = out->K;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            if (i < j)
                K[i + j * n] = k[i * (n - 1) + j - 1];
            else if (i > j)
                K[i + j * n] = k[i * (n - 1) + j];
            else
                K[i + j * n] = 0.0f;
        }
    for (int j = 0; j < n; ++j)
        for (int ii = 0; ii < n; ++ii)
            if (ii != j) K[j + j * n] -= K[ii + j * n];
    for (int j = 0; j < n; ++j) {
        float sum = 0.0f;
        for (int ii = 0; ii < n; ++ii)
            if (ii != j) sum += K[ii + j * n];
        for (int ii = 0; ii < n; ++ii)
            out->P_i2j[ii + j * n] = (ii != j) ? K[ii + j * n] / sum : 0.0f;
    }
    for (int i = 0; i < n; ++i) {
        float a = 0.0f;
        for (int j = 0; j < n; ++j) a += K[i + j * n];
        a -= K[i + i * n];
        out->p[i] = a; /* normalisation is commented out in the reference (:51-52) */
    }
    return 1;
}

/* ---- second draw source: the reference's own generators -------------------------
 * One cuRAND scrambled Sobol64 sequence per thread (state i/x + 64 direction vectors, gray-code
 * update: curand_kernel.h:1582-1591) consumed in the reference's call order, and the thread's
 * xor128 stream; both persist from one evaluation to the next (src/mc.cu:366-455, quirk Q8). */
typedef struct {
    uint64_t i, x;
    const uint64_t *dv;
    uint32_t xs[4];
} sobol_thread_t;

static inline uint64_t sobol_next(sobol_thread_t *s)
{
    uint64_t r = s->x;
    s->x ^= s->dv[__builtin_ctzll(~s->i)];
    s->i++;
    return r;
}

/* ---- one thread of mc_kernel, src/mc.cu:92-196 ---------------------------------
 * S == NULL: the draw contract (Philox words keyed by burst, repeat, step, slot);
 * S != NULL: the reference's Sobol64 + xor128 generators, drawn where the reference draws. */
static void eval_sample(const pda_bursts_t *bd, const pda_params_t *prm, int64_t idx,
                        uint32_t g, uint32_t rep, uint32_t seed, uint32_t epoch,
                        float *o_mcE, int64_t *o_sum, float *o_F, int32_t *o_steps,
                        sobol_thread_t *S)
{
    const int n = prm->n;
    const int ph = bd->ph_count[idx];
    const int64_t *ad = bd->burst_ad + bd->istart[idx];
    const int64_t *dd = bd->burst_dd + bd->istart[idx];
    const float T = bd->T[idx];
    const float clk_p = bd->clk_p;
    const int with_bg = !((double)bd->bg_dd_rate < 1e-4); /* mc.cu:153 */

    /* draws of step 0; the xor128 stream is seeded like setup_kernel seeds rk_state
     * from one 64-bit draw (gen_rand.cuh:19-21) */
    uint32_t xs[4];
    philox4_t w = {{0u, 0u, 0u, 0u}};
    if (S) {
        memcpy(xs, S->xs, sizeof(xs));
        w.v[3] = (uint32_t)(sobol_next(S) >> 32); /* drawDisIdx before the loop, mc.cu:110 */
    } else {
        w = philox4x32_10(g, rep, 0u, PDA_SLOT_STEP, seed, epoch);
        rk_seed((uint64_t)w.v[3], xs);
        w.v[3] = pda_init_word(&w);
    }

    float F = 0.0f, acc = 0.0f, spent = 0.0f;
    int64_t sum_ai = 0;
    int si = draw_dis_idx(n, prm->p, u01(w.v[3])); /* mc.cu:110 */
    int64_t bin0 = bd->start[idx], bin1 = bd->start[idx];
    int first = 0;
    uint32_t step = 0;
    while (T > spent) { /* mc.cu:117 */
        if (S) {
            w.v[0] = (uint32_t)(sobol_next(S) >> 32); /* drawJ_Si2Sj */
            w.v[1] = (uint32_t)(sobol_next(S) >> 32); /* drawTau */
        } else if (step) {
            w = philox4x32_10(g, rep, step, PDA_SLOT_STEP, seed, epoch);
        }
        int sj = draw_j(prm->P_i2j, n, si, u01(w.v[0]));      /* mc.cu:118 */
        float st = draw_tau(prm->K[sj + si * n], u01(w.v[1])); /* mc.cu:122 */
        spent = spent + st;
        if (spent >= T)
            bin1 = bd->stop[idx];
        else
            bin1 = (int64_t)((float)bin0 + spent / clk_p); /* mc.cu:132: float arithmetic */
        int sdd = 0, sad = 0, f_id = 0, f_ia = 0;
        int64_t ai = 0;
        for (int i = first; i < ph; ++i) { /* mc.cu:140-181 */
            int64_t ddx = dd[i], adx = ad[i];
            if (ddx >= bin1 || adx >= bin1 || i == ph - 1) {
                first = i;
                float f_if = (PDA_GAMMA - PDA_DCH2ACH) * (float)f_id +
                             (1.0f - PDA_DEXDIRAEM) * (float)f_ia;
                float f_i;
                F += f_if;
                if (!with_bg) {
                    f_i = floorf(f_if + 0.5f);
                } else {
                    float t_diff = (float)(bin1 - bin0) * clk_p;
                    float rf_ia = (1.0f - PDA_DEXDIRAEM) * (float)f_ia;
                    /* draw_P_B_Tr: one word per Poisson draw, a fall-back uniform only when the
                     * draw exceeds the photons there are (gen_rand.cuh:92-107) */
                    /* contract draws: one block for both channels, (Poisson word, fallback) each */
                    philox4_t bw = {{0u, 0u, 0u, 0u}};
                    uint32_t p_hi, p_lo = 0u;
                    if (S) { uint64_t x = sobol_next(S); p_hi = (uint32_t)(x >> 32); p_lo = (uint32_t)x; }
                    else { bw = philox4x32_10(g, rep, step, PDA_SLOT_BG, seed, epoch); p_hi = bw.v[0]; }
                    uint32_t sv = poisson_itr(p_hi, p_lo, (double)(t_diff * bd->bg_ad_rate));
                    float bg_a;
                    if ((float)(int32_t)sv <= rf_ia) bg_a = (float)(int32_t)sv;
                    else bg_a = floorf(rf_ia * u01(S ? (uint32_t)(sobol_next(S) >> 32) : bw.v[1]));
                    float rf_id = (PDA_GAMMA - PDA_DCH2ACH) * (float)f_id;
                    if (S) { uint64_t x = sobol_next(S); p_hi = (uint32_t)(x >> 32); p_lo = (uint32_t)x; }
                    else p_hi = bw.v[2];
                    sv = poisson_itr(p_hi, p_lo, (double)(t_diff * bd->bg_dd_rate));
                    float bg_d;
                    if ((float)(int32_t)sv <= rf_id) bg_d = (float)(int32_t)sv;
                    else bg_d = floorf(rf_id * u01(S ? (uint32_t)(sobol_next(S) >> 32) : bw.v[3]));
                    f_i = floorf(f_if - bg_d - bg_a + 0.5f);
                }
                if (S) w.v[2] = (uint32_t)(sobol_next(S) >> 32);             /* curand_normal in drawE */
                float de = draw_e(prm->E[si], PDA_R0, prm->v[si], w.v[2]); /* mc.cu:167 */
                ai = rk_binomial(xs, (int)f_i, (double)de);                /* mc.cu:168 */
                break;
            }
            if (ddx > 0) sdd = 1;
            if (adx > 0) sad = 1;
            if (sad && sdd) continue;
            if (sad) f_ia++;
            if (sdd) f_id++;
        }
        acc += (float)ai; /* mcE[tidx] += ai */
        sum_ai += ai;
        bin0 = bin1;
        si = sj;
        step++;
    }
    if (o_mcE) *o_mcE = (F > 0.0f) ? acc / F : 0.0f; /* mc.cu:192-195 */
    if (o_sum) *o_sum = sum_ai;
    if (o_F) *o_F = F;
    if (o_steps) *o_steps = (int32_t)step;
    if (S) memcpy(S->xs, xs, sizeof(xs));
}

/* ---- boost::histogram regular axis index, 3rdparty/src/boost/libs/histogram/include/
 * boost/histogram/axis/regular.hpp:292-308 with min 0, delta 1 (streamWorker.cpp:33-38) */
static inline int hist_index(float x, int nb)
{
    double z = ((double)x - 0.0) / 1.0;
    if (z < 1.0) {
        if (z >= 0.0) return (int)(z * (double)nb);
        return nb + 1; /* underflow (-1 in boost) */
    }
    return nb; /* overflow, also NaN */
}

void pda_oracle_hist(const float *x, int64_t n, int nb, uint64_t *hist)
{
    for (int64_t i = 0; i < n; ++i) hist[hist_index(x[i], nb)]++;
}

/* ---- reduced chi-square, src/streamWorker.cpp:197-206.  vOEHist/vMcHist are float
 * vectors; pow(float, int) promotes to double; chisqr is a float accumulator. */
float pda_oracle_chi2(const float *ohist, const uint64_t *mhist, int nb, int R,
                      int left_pad, int n_states)
{
    int effN = nb - left_pad;
    float chisqr = 0.0f;
    for (int i = left_pad; i < nb; ++i) {
        float M = (float)mhist[i];
        if (M > 0.0f) {
            float d = ohist[i] - M / (float)R;
            double num = (double)d * (double)d;
            double den = (double)(M / (float)R);
            chisqr = (float)((double)chisqr + num / den);
        } else {
            effN--;
        }
    }
    return chisqr / (float)(effN - n_states * (n_states + 1));
}

void pda_oracle_eval(const pda_bursts_t *bursts, const pda_params_t *prm, int R,
                     uint32_t seed, uint32_t epoch, int64_t b0, int64_t b1,
                     int64_t g_offset, const pda_samples_t *out, int nthreads)
{
    int64_t total = (b1 - b0) * (int64_t)R;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
#endif
    for (int64_t t = 0; t < total; ++t) {
        int64_t b = b0 + t / R;
        uint32_t rep = (uint32_t)(t % R);
        eval_sample(bursts, prm, b, (uint32_t)(b + g_offset), rep, seed, epoch,
                    out->mcE ? out->mcE + t : NULL, out->sum_ai ? out->sum_ai + t : NULL,
                    out->F ? out->F + t : NULL, out->n_steps ? out->n_steps + t : NULL, NULL);
    }
}

/* The reference's generators: thread t = burst * R + repeat takes dimension t % 20000 of
 * CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6 at offset 1234 and seeds its xor128 stream from the
 * first word (setup_kernel, gen_rand.cuh:7-25; mc::init_randstate, mc.cu:366-455); `evals`
 * evaluations run on the persistent states.  mcE: [evals][N*R].  Tables from libcurand's host API. */
void pda_oracle_eval_sobol(const pda_bursts_t *bursts, const pda_params_t *prm, int R, int64_t N,
                           const uint64_t *dirvec20000x64, const uint64_t *scramble20000,
                           int evals, float *mcE, int nthreads)
{
    int64_t total = N * (int64_t)R;
    sobol_thread_t *st = (sobol_thread_t *)malloc((size_t)(total > 0 ? total : 1) * sizeof(sobol_thread_t));
    for (int64_t t = 0; t < total; ++t) {
        sobol_thread_t *s = &st[t];
        s->dv = dirvec20000x64 + 64 * (t % 20000);
        s->i = 1234;
        s->x = scramble20000[t % 20000];
        uint64_t gray = s->i ^ (s->i >> 1);
        for (int k = 0; k < 64; ++k)
            if (gray & (1ull << k)) s->x ^= s->dv[k];
        rk_seed(sobol_next(s), s->xs);
    }
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
    for (int ev = 0; ev < evals; ++ev) {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
#endif
        for (int64_t t = 0; t < total; ++t)
            eval_sample(bursts, prm, t / R, 0u, 0u, 0u, 0u, mcE + (size_t)ev * (size_t)total + t, NULL, NULL,
                        NULL, &st[t]);
    }
    free(st);
}

float pda_oracle_eval_chi2(const pda_bursts_t *bursts, const pda_params_t *prm, int R,
                           uint32_t seed, uint32_t epoch, int64_t b0, int64_t b1,
                           int64_t g_offset, const float *ohist, int nb, int left_pad,
                           uint64_t *hist, int nthreads)
{
    int64_t total = (b1 - b0) * (int64_t)R;
    memset(hist, 0, sizeof(uint64_t) * (size_t)(nb + 2));
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
    {
        uint64_t *local = (uint64_t *)calloc((size_t)nb + 2, sizeof(uint64_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (int64_t t = 0; t < total; ++t) {
            int64_t b = b0 + t / R;
            float e;
            eval_sample(bursts, prm, b, (uint32_t)(b + g_offset), (uint32_t)(t % R), seed,
                        epoch, &e, NULL, NULL, NULL, NULL);
            local[hist_index(e, nb)]++;
        }
#ifdef _OPENMP
#pragma omp critical
#endif
        for (int i = 0; i < nb + 2; ++i) hist[i] += local[i];
        free(local);
    }
    return pda_oracle_chi2(ohist, hist, nb, R, left_pad, prm->n);
}

/* ---- primitive hooks for tests -------------------------------------------------- */
void pda_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    philox4_t r = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    memcpy(out, r.v, sizeof(r.v));
}
float pda_oracle_logf(float x) { return det_logf(x); }
double pda_oracle_log(double x) { return det_log(x); }
double pda_oracle_exp(double x) { return det_exp(x); }
float pda_oracle_norm_icdf(uint32_t w) { return det_norm_icdf(w); }
float pda_oracle_root6f(float x) { return det_root6f(x); }
float pda_oracle_u01(uint32_t w) { return u01(w); }
int pda_oracle_draw_dis_idx(int n, const float *p, uint32_t w) { return draw_dis_idx(n, p, u01(w)); }
float pda_oracle_draw_tau(float k, uint32_t w) { return draw_tau(k, u01(w)); }
float pda_oracle_draw_e(float e, float r0, float v, uint32_t w) { return draw_e(e, r0, v, w); }
int pda_oracle_draw_j(const float *P_i2j, int n, int i, uint32_t w)
{
    return draw_j(P_i2j, n, i, u01(w));
}
uint32_t pda_oracle_poisson(uint32_t hi, uint32_t lo, double lambda)
{
    return poisson_itr(hi, lo, lambda);
}
void pda_oracle_rk_seed(uint64_t s, uint32_t st[4]) { rk_seed(s, st); }
double pda_oracle_rk_double(uint32_t st[4]) { return rk_double(st); }
int64_t pda_oracle_binomial(uint32_t st[4], int n, double p) { return rk_binomial(st, n, p); }
int pda_oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
