/* oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * Host build of the REFERENCE'S OWN hot-path sources, #included from where they lie
 * under /root/reference/src (nothing is copied into this repository):
 *     binom.cuh      xor128 + binomial (inversion / BTPE) + multinomial
 *     gen_rand.cuh   setup_kernel, drawDisIdx, drawTau, drawE, drawJ_Si2Sj, draw_P_B_Tr
 *     mc.cu:83-197   mc_kernel, cut out by oracle/Makefile into a scratch dir under /tmp
 *     mc.hpp:19-30   rk_state,  likewise (scratch removed after the compile)
 *     eigenhelper.cpp (compiled as its own translation unit)
 * What this file supplies is only what the CUDA toolkit supplied to the reference:
 *   - the __global__/__device__ qualifiers and blockIdx/blockDim/threadIdx,
 *   - a stand-in for cuRAND's curandStateScrambledSobol64 whose curand(),
 *     curand_uniform(), curand_normal() and curand_poisson() return the draws of the
 *     contract in the order the reference asks for them,
 *   - logf/powf/log/exp bound to the explicit-rounding routines of det_math.h
 *     (the reference gets them from libdevice).
 * So the reference code itself, unmodified, replays the contract draws and is the
 * bit-exact pin of the restatement in pda_oracle.c.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "eigenhelper.hpp" /* reference header: Eigen typedefs (matXfMapper) + genMat* prototypes */
#ifdef _OPENMP
#include <omp.h>
#endif
#include "det_math.h"
#include "philox.h"

typedef float retype; /* src/loadHdf5.hpp:8 */

/* ---- CUDA execution-model stand-ins ------------------------------------------- */
#define __global__
#define __device__
struct shim_dim3 { unsigned x, y, z; };
static thread_local shim_dim3 blockIdx = {0, 1, 1}, blockDim = {1, 1, 1}, threadIdx = {0, 0, 0};

/* ---- cuRAND stand-in: the draw contract ---------------------------------------- */
struct curandStateScrambledSobol64 {
    uint32_t g, rep, seed, epoch, step;
    int phase;    /* 0 before the initial-state draw, 1 next-state, 2 dwell, 3 inside window */
    int bg_calls; /* Poisson draws made in this step */
    int seeded;   /* curand() for rk_seed already served */
    /* second personality (ref_mc_kernel_sobol): cuRAND's own scrambled Sobol64 generator,
     * restated from the toolkit's curand_kernel.h (state i/x/c + 64 direction vectors, gray-code
     * update, skipahead), so the reference code draws what it draws on the GPU */
    int sobol;
    unsigned long long i, x, c;
    const unsigned long long *dv;
};

static inline unsigned long long sobol_next(curandStateScrambledSobol64 *s)
{
    /* curand(curandStateScrambledSobol64_t*), curand_kernel.h:1582-1591 */
    unsigned long long res = s->x;
    s->x ^= s->dv[__builtin_ctzll(~s->i)];   /* ~i == 0 cannot happen below 2^64 draws */
    s->i++;
    return res;
}

static inline float shim_u01(uint32_t w)
{
    return (float)w * 2.3283064365386963e-10f + 1.1641532182693481e-10f;
}

static void curand_init(unsigned long long *dirvec, unsigned long long scramble, unsigned long long offset,
                        curandStateScrambledSobol64 *s)
{
    /* contract personality: direction vectors are not used, the state was pre-loaded with
     * (g, rep, seed, epoch) */
    if (!s->sobol) return;
    /* curand_init(curandDirectionVectors64_t, scramble_c, offset, state) + skipahead,
     * curand_kernel.h:1491-1503,1340-1357 */
    s->dv = dirvec;
    s->c = scramble;
    s->i = offset;
    s->x = s->c;
    const unsigned long long gray = s->i ^ (s->i >> 1);
    for (unsigned k = 0; k < 64; k++)
        if (gray & (1ULL << k)) s->x ^= s->dv[k];
}

/* _curand_poisson_ITR_double (curand_poisson.h:436-462) on a 64-bit word */
static unsigned int shim_poisson_itr(uint64_t x, double lambda);

static unsigned long long curand(curandStateScrambledSobol64 *s)
{
    if (s->sobol) return sobol_next(s);
    if (s->seeded) { fprintf(stderr, "ref_shim: unexpected curand()\n"); abort(); }
    s->seeded = 1;
    philox4_t w = philox4x32_10(s->g, s->rep, 0u, PDA_SLOT_STEP, s->seed, s->epoch);
    return (unsigned long long)w.v[3];
}

static float curand_uniform(curandStateScrambledSobol64 *s)
{
    if (s->sobol) return shim_u01((uint32_t)(sobol_next(s) >> 32));   /* _curand_uniform(ull), curand_uniform.h:84-89 */
    if (s->phase == 3 && s->bg_calls > 0) { /* fallback uniform of draw_P_B_Tr */
        philox4_t w = philox4x32_10(s->g, s->rep, s->step, PDA_SLOT_BG, s->seed, s->epoch);
        return shim_u01(s->bg_calls == 1 ? w.v[1] : w.v[3]);
    }
    if (s->phase == 3) { /* a window without photons drew no E: next step starts here */
        s->step++;
        s->phase = 1;
    }
    philox4_t w = philox4x32_10(s->g, s->rep, s->step, PDA_SLOT_STEP, s->seed, s->epoch);
    if (s->phase == 0) { s->phase = 1; return shim_u01(pda_init_word(&w)); }
    if (s->phase == 1) { s->phase = 2; return shim_u01(w.v[0]); }
    s->phase = 3;
    s->bg_calls = 0;
    return shim_u01(w.v[1]);
}

static float curand_normal(curandStateScrambledSobol64 *s)
{
    if (s->sobol) return det_norm_icdf((uint32_t)(sobol_next(s) >> 32));   /* _curand_normal_icdf(ull) */
    if (s->phase != 3) { fprintf(stderr, "ref_shim: curand_normal out of order\n"); abort(); }
    philox4_t w = philox4x32_10(s->g, s->rep, s->step, PDA_SLOT_STEP, s->seed, s->epoch);
    s->step++;
    s->phase = 1;
    s->bg_calls = 0;
    return det_norm_icdf(w.v[2]);
}

/* cuRAND: curand_poisson(Sobol64, lambda) = _curand_poisson_ITR_double for lambda < 1000
 * (/usr/local/cuda/include/curand_poisson.h:436-463,546-551,759); that header is NVIDIA's,
 * restated here on det_exp so it replays bit-exactly. */
static unsigned int shim_poisson_itr(uint64_t x, double lambda)
{
    double u = (double)(x >> 11) * 1.1102230246251565e-16 + 2.7755575615628914e-17;
    double L, p = 1.0, q = 1.0;
    unsigned int k = 0;
    int pw = 0;
    do {
        L = (lambda > (double)(pw + 500.0)) ? det_exp(-500.0) : det_exp((double)(pw - lambda));
        p *= L;
        q *= L;
        pw += 500;
        while (u > q) {
            k++;
            p *= (lambda / (double)k);
            q += p;
        }
    } while ((double)pw < lambda);
    return k;
}

static unsigned int curand_poisson(curandStateScrambledSobol64 *s, double lambda)
{
    if (s->sobol) {
        if (!(lambda < 1000.0)) { fprintf(stderr, "ref_shim: lambda >= 1000 outside contract\n"); abort(); }
        return shim_poisson_itr(sobol_next(s), lambda);
    }
    if (s->phase != 3 || s->bg_calls > 1) { fprintf(stderr, "ref_shim: poisson out of order\n"); abort(); }
    if (!(lambda < 1000.0)) { fprintf(stderr, "ref_shim: lambda >= 1000 outside contract\n"); abort(); }
    s->bg_calls++;
    philox4_t w = philox4x32_10(s->g, s->rep, s->step, PDA_SLOT_BG, s->seed, s->epoch);
    return shim_poisson_itr((uint64_t)(s->bg_calls == 1 ? w.v[0] : w.v[2]) << 32, lambda);
}

/* ---- libdevice stand-ins -------------------------------------------------------- */
static inline float shim_powf(float x, float y)
{
    if (y == 6.0f) return det_pow6f(x);
    if (y == (float)(1.0 / 6)) return det_root6f(x);
    fprintf(stderr, "ref_shim: powf exponent %g outside contract\n", (double)y);
    abort();
}
using std::min;
#define logf(x) det_logf(x)
#define powf(x, y) shim_powf((float)(x), (float)(y))
#define log(x) det_log(x)
#define exp(x) det_exp(x)

/* ---- the reference sources ------------------------------------------------------ */
#include "rk_state.inc"   /* typedef struct {...} rk_state;   from src/mc.hpp */
#include "binom.cuh"      /* /root/reference/src/binom.cuh */
#include "gen_rand.cuh"   /* /root/reference/src/gen_rand.cuh */
#include "mc_kernel.inc"  /* mc_kernel from /root/reference/src/mc.cu */

#undef logf
#undef powf
#undef log
#undef exp

/* ---- C entry points for the tests ------------------------------------------------ */
extern "C" {

/* mc::set_params' split of args (src/mc.cu:792-799) + the reference's eigenhelper.cpp */
int ref_params(int n, float *args, float *E, float *v, float *K, float *P_i2j, float *p)
{
    vecFloatMapper evargs(args, n * n + n);
    RowVectorXf eargs = evargs.block(0, 0, 1, n);
    RowVectorXf kargs = evargs.block(0, n, 1, n * n - n);
    RowVectorXf vargs = evargs.block(0, n * n, 1, n);
    for (int i = 0; i < n; ++i) { E[i] = eargs[i]; v[i] = vargs[i]; }
    bool r = genMatK(K, n, kargs);
    if (!r) return 0;
    genP_i2j(K, P_i2j, n);
    r = r && genMatP(p, K, n);
    return r ? 1 : 0;
}

/* genMatK alone with an arbitrary argument count (its false-return cases) */
int ref_genMatK(int n, float *kargs, int n_kargs, float *K)
{
    RowVectorXf a(n_kargs);
    for (int i = 0; i < n_kargs; ++i) a[i] = kargs[i];
    return genMatK(K, n, a) ? 1 : 0;
}

/* Run the reference's setup_kernel + mc_kernel over N bursts x R repeats (R a power of
 * two, as the reference requires: src/mc.cu:101).  Burst data always start at burst 0
 * (quirk Q7).  g_offset shifts the contract's global burst id only. */
void ref_mc_kernel(int64_t *start, int64_t *stop, int64_t *burst_ad, int64_t *burst_dd,
                   int64_t *istart, int *phCount, float *T, float clk_p, float bg_ad_rate,
                   float bg_dd_rate, float *gpe, float *gpv, float *gpk, float *gpp_i2j,
                   float *gpp, int N, int s_n, int R, uint32_t seed, uint32_t epoch,
                   int64_t g_offset, float *mcE, int nthreads)
{
    int64_t NN = (int64_t)N * R;
    std::vector<curandStateScrambledSobol64> q((size_t)NN);
    std::vector<rk_state> st((size_t)NN);
    std::vector<float> scratch((size_t)NN * (size_t)s_n);
    std::vector<unsigned long long> scramble((size_t)NN, 0ull); /* read, then ignored */
    static unsigned long long dirvec[64];                       /* address only, never read */
    int shift = (int)log2((double)R);
    for (int64_t t = 0; t < NN; ++t) {
        memset(&q[t], 0, sizeof(q[t]));
        q[t].g = (uint32_t)((t >> shift) + g_offset);
        q[t].rep = (uint32_t)(t & (R - 1));
        q[t].seed = seed;
        q[t].epoch = epoch;
        memset(&st[t], 0, sizeof(rk_state));
    }
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
#endif
    for (int64_t t = 0; t < NN; ++t) {
        blockDim.x = 1;
        threadIdx.x = 0;
        blockIdx.x = (unsigned)t;
        setup_kernel(st.data(), 0ul, (int)NN, dirvec - 64 * t, scramble.data(), q.data());
        mc_kernel(start, stop, burst_ad, burst_dd, istart, phCount, T, clk_p, bg_ad_rate,
                  bg_dd_rate, gpe, gpv, gpk, gpp_i2j, gpp, scratch.data(), N, s_n, q.data(),
                  st.data(), mcE, R);
    }
}

/* The same with cuRAND's own generator: what the reference's class mc does on the GPU.
 * mc::init_randstate (src/mc.cu:366-455) hands thread t the direction vectors and scramble
 * constant of dimension t % 20000 of CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6 (tables from
 * libcurand's host API, passed in), setup_kernel starts every sequence at offset 1234 and seeds
 * the xor128 stream from its first draw; the states persist over the `evals` evaluations
 * (quirk Q8).  mcE: [evals][N*R].  Differences left against a GPU run: libdevice's
 * logf/powf/erfcinvf/log/exp are a few ulp away from det_math.h. */
void ref_mc_kernel_sobol(int64_t *start, int64_t *stop, int64_t *burst_ad, int64_t *burst_dd,
                         int64_t *istart, int *phCount, float *T, float clk_p, float bg_ad_rate,
                         float bg_dd_rate, float *gpe, float *gpv, float *gpk, float *gpp_i2j,
                         float *gpp, int N, int s_n, int R, unsigned long long *dirvec20000x64,
                         unsigned long long *scramble20000, int evals, float *mcE, int nthreads)
{
    int64_t NN = (int64_t)N * R;
    std::vector<curandStateScrambledSobol64> q((size_t)NN);
    std::vector<rk_state> st((size_t)NN);
    std::vector<float> scratch((size_t)NN * (size_t)s_n);
    std::vector<unsigned long long> scr((size_t)NN);
    for (int64_t t = 0; t < NN; ++t) {
        memset(&q[t], 0, sizeof(q[t]));
        q[t].sobol = 1;
        memset(&st[t], 0, sizeof(rk_state));
        scr[t] = scramble20000[t % 20000];
    }
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
    for (int ev = -1; ev < evals; ++ev) {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
#endif
        for (int64_t t = 0; t < NN; ++t) {
            blockDim.x = 1;
            threadIdx.x = 0;
            blockIdx.x = (unsigned)t;
            if (ev < 0)   /* setup_kernel reads sobolDirectionVectors + 64 * idx */
                setup_kernel(st.data(), 0ul, (int)NN, dirvec20000x64 + 64 * (t % 20000) - 64 * t, scr.data(), q.data());
            else
                mc_kernel(start, stop, burst_ad, burst_dd, istart, phCount, T, clk_p, bg_ad_rate,
                          bg_dd_rate, gpe, gpv, gpk, gpp_i2j, gpp, scratch.data(), N, s_n, q.data(),
                          st.data(), mcE + (size_t)ev * (size_t)NN, R);
        }
    }
}

/* primitives of the reference, one draw each */
int ref_drawDisIdx(int n, float *p, uint32_t g, uint32_t rep, uint32_t seed, uint32_t epoch)
{
    curandStateScrambledSobol64 s = {g, rep, seed, epoch, 0, 1, 0, 1}; /* phase 1: word x0 */
    return drawDisIdx(n, p, &s);
}
float ref_drawTau(float k, uint32_t g, uint32_t rep, uint32_t seed, uint32_t epoch)
{
    curandStateScrambledSobol64 s = {g, rep, seed, epoch, 0, 2, 0, 1}; /* phase 2: word x1 */
    return drawTau(k, &s, 0);
}
float ref_drawE(float e, float r0, float v, uint32_t g, uint32_t rep, uint32_t seed, uint32_t epoch)
{
    curandStateScrambledSobol64 s = {g, rep, seed, epoch, 0, 3, 0, 1}; /* word x2 */
    return drawE(e, r0, v, &s);
}
int ref_drawJ(float *matP_i2j, int n, int i, uint32_t g, uint32_t rep, uint32_t seed, uint32_t epoch)
{
    float scratch[64];
    curandStateScrambledSobol64 s = {g, rep, seed, epoch, 0, 1, 0, 1};
    return drawJ_Si2Sj(scratch, matP_i2j, n, i, &s);
}
void ref_rk_seed(uint64_t seed64, uint32_t out[4])
{
    rk_state s;
    memset(&s, 0, sizeof(s));
    rk_seed(seed64, &s);
    memcpy(out, s.xor128, 16);
}
/* binomial on an xor128 state carried by the caller (cache cold each call, as after a
 * change of (n, p)) */
long ref_rk_binomial(uint32_t st[4], int n, double p)
{
    rk_state s;
    memset(&s, 0, sizeof(s));
    memcpy(s.xor128, st, 16);
    long r = rk_binomial(&s, n, p);
    memcpy(st, s.xor128, 16);
    return r;
}
double ref_rk_double(uint32_t st[4])
{
    rk_state s;
    memset(&s, 0, sizeof(s));
    memcpy(s.xor128, st, 16);
    double r = rk_double(&s);
    memcpy(st, s.xor128, 16);
    return r;
}
void ref_int_multinomial(uint32_t st[4], int K, int N, float *p, int *out)
{
    rk_state s;
    memset(&s, 0, sizeof(s));
    memcpy(s.xor128, st, 16);
    csd_int_multinomial(&s, K, N, p, out);
    memcpy(st, s.xor128, 16);
}

} /* extern "C" */
