/* oracle/pda_oracle.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * CPU restatement of the Monte-Carlo PDA evaluation of liu-kan/gSMFRETda, used as
 * the checker of the CUDA product.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it.  Nothing under
 * gsmfretda_b200/ includes, links or calls it.
 *
 * Parity status: PINNED against the reference's own source compiled for the host
 * (oracle/_ref/libref_cpu.so, recipe oracle/Makefile + oracle/ref_shim.cpp) — per
 * sample bit-exact on the same draws — and against tests/golden/ fixtures generated
 * from that build.  The reference's tests hold no golden vectors for this path
 * (SURVEY.md §8c); its algebraic and statistical pins are re-hosted in tests/.
 */
#ifndef PDA_ORACLE_H
#define PDA_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDA_MAX_STATES 8

/* burst arrays in the reference's layout (mc::init_data_gpu, /root/reference/src/mc.cu:456-521) */
typedef struct {
    int64_t n_bursts;
    const int64_t *start, *stop, *istart; /* [n_bursts] */
    const int32_t *ph_count;              /* [n_bursts] */
    const float *T;                       /* [n_bursts] burst duration, s */
    const int64_t *burst_ad, *burst_dd;   /* [n_photons] timestamp or 0 */
    float clk_p, bg_ad_rate, bg_dd_rate;
} pda_bursts_t;

/* one parameter vector expanded as mc::set_params does (/root/reference/src/mc.cu:790-829) */
typedef struct {
    int n;
    float E[PDA_MAX_STATES];
    float v[PDA_MAX_STATES];
    float K[PDA_MAX_STATES * PDA_MAX_STATES];     /* column-major n x n, K(i,j) = rate j->i */
    float P_i2j[PDA_MAX_STATES * PDA_MAX_STATES]; /* column-major n x n */
    float p[PDA_MAX_STATES];                      /* unnormalised inflow rates */
} pda_params_t;

/* args = [E_0..E_{n-1} | k (n*n-n, row-major without the diagonal) | v_0..v_{n-1}].
 * returns 1 on success, 0 when n or n_args is wrong (genMatK's false). */
int pda_oracle_params(int n, const float *args, int n_args, pda_params_t *out);

/* per-sample outputs; any pointer may be NULL.  sample index = burst * R + repeat. */
typedef struct {
    float *mcE;        /* the reference's result, sum(ai)/F or 0 */
    int64_t *sum_ai;   /* integer acceptor photons, before the float accumulation */
    float *F;          /* the float accumulator F */
    int32_t *n_steps;  /* dwell windows walked */
} pda_samples_t;

/* Evaluate bursts [b0, b1) x R repeats.  g_offset is added to the local burst index to
 * form the global burst id of the draw contract (burst-sharded evaluation). Outputs are
 * indexed from sample 0 = (b0, repeat 0). nthreads <= 0: all OpenMP threads. */
void pda_oracle_eval(const pda_bursts_t *bursts, const pda_params_t *prm, int R,
                     uint32_t seed, uint32_t epoch, int64_t b0, int64_t b1,
                     int64_t g_offset, const pda_samples_t *out, int nthreads);

/* The same evaluation on the reference's own generators (cuRAND scrambled Sobol64 per thread,
 * dimension t % 20000 at offset 1234, + its xor128 stream; src/mc.cu:366-455, gen_rand.cuh:7-25):
 * what the reference's class mc computes on the GPU.  `evals` evaluations on persistent states;
 * mcE is [evals][N*R].  dirvec/scramble: the 20000 x 64 / 20000 tables of
 * CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6 from libcurand's host API. */
void pda_oracle_eval_sobol(const pda_bursts_t *bursts, const pda_params_t *prm, int R, int64_t N,
                           const uint64_t *dirvec20000x64, const uint64_t *scramble20000,
                           int evals, float *mcE, int nthreads);

/* boost::histogram regular axis [0,1) with nb bins; hist has nb+2 entries:
 * [0,nb) inner, [nb] overflow (x>=1 or NaN), [nb+1] underflow (x<0). counts are ADDED. */
void pda_oracle_hist(const float *x, int64_t n, int nb, uint64_t *hist);

/* reduced chi-square of streamWorker.cpp:197-206. ohist = observed counts as float. */
float pda_oracle_chi2(const float *ohist, const uint64_t *mhist, int nb, int R,
                      int left_pad, int n_states);

/* convenience: eval + hist + chi2 in one call, nothing per-sample kept (cpu baseline) */
float pda_oracle_eval_chi2(const pda_bursts_t *bursts, const pda_params_t *prm, int R,
                           uint32_t seed, uint32_t epoch, int64_t b0, int64_t b1,
                           int64_t g_offset, const float *ohist, int nb, int left_pad,
                           uint64_t *hist, int nthreads);

/* --- primitives exposed for the unit tests -------------------------------- */
void pda_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
float pda_oracle_logf(float x);
double pda_oracle_log(double x);
double pda_oracle_exp(double x);
float pda_oracle_norm_icdf(uint32_t w);
float pda_oracle_root6f(float x);
float pda_oracle_u01(uint32_t w);
int pda_oracle_draw_dis_idx(int n, const float *p, uint32_t w);
float pda_oracle_draw_tau(float k, uint32_t w);
float pda_oracle_draw_e(float e, float r0, float v, uint32_t w);
int pda_oracle_draw_j(const float *P_i2j, int n, int i, uint32_t w);
uint32_t pda_oracle_poisson(uint32_t hi, uint32_t lo, double lambda);
/* xor128 state seeded as rk_seed does; binomial consumes it */
void pda_oracle_rk_seed(uint64_t s, uint32_t st[4]);
double pda_oracle_rk_double(uint32_t st[4]);
int64_t pda_oracle_binomial(uint32_t st[4], int n, double p);
void pda_oracle_multinomial(uint32_t st[4], int K, int N, const float *p, int *out);
int pda_oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
