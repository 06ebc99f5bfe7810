/* gsmfret_c.h — C ABI of libgsmfret.so: the drop-in boundary of the B200-native
 * Monte-Carlo PDA evaluation.
 *
 * The reference (liu-kan/gSMFRETda) has no C ABI; its boundary is the C++ class `mc`
 * (/root/reference/src/mc.hpp:78-179) driven by gpuWorker::run
 * (/root/reference/src/gpuWorker.cpp:71-80) and read by streamWorker::run
 * (/root/reference/src/streamWorker.cpp:186-206).  Each entry point below replaces one
 * member of that class (cited per function); gsmfretda_b200/host/mc.hpp re-creates the
 * C++ class on top of this header, INTEGRATION.md shows the reference-side binding.
 *
 * Conventions: plain pointers and sizes only.  Functions returning `int` status return
 * GSM_OK (0) or a negative GSM_E* code, with the text available from gsm_last_error()
 * (thread-local).  The reference's convention — print "<stmt> returned <msg>(<code>) at
 * file:line" and exit (/root/reference/src/cuda_tools.hpp:20-28) — is re-created by the C++
 * shim, not here.  Input arrays stay owned by the caller and may be freed after the call
 * returns; result pointers are library-owned pinned host memory, valid until the next
 * gsm_run*() on the same stream slot or gsm_destroy().
 *
 * Threading contract (same as the reference): all calls that take a
 * stream id `sid` come from one thread at a time per sid; results of slot `sid` may be
 * read by another thread once gsm_query(ctx, sid) returned 1.
 *
 * There is no CPU fallback: every compute entry point needs a CUDA device of compute
 * capability 10.0 (sm_100a) and fails with GSM_ENODEVICE otherwise.
 */
#ifndef GSMFRET_C_H
#define GSMFRET_C_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSM_ABI_VERSION 5

#define GSM_OK 0
#define GSM_EINVAL (-1)    /* bad argument */
#define GSM_ECUDA (-2)     /* a CUDA runtime call failed */
#define GSM_ENODEVICE (-3) /* no usable sm_100 device */
#define GSM_ESTATE (-4)    /* call out of order (e.g. run before init_data) */
#define GSM_ENCCL (-5)     /* an NCCL call failed */
#define GSM_ERANGE (-6)    /* input outside the draw contract (e.g. background lambda >= 1000) */

/* gsm_create flags */
#define GSM_KEEP_MCE 1u     /* also return the per-sample efficiencies (mc::hmcE) */
#define GSM_KEEP_COUNTS 2u  /* also return per-sample integer acceptor counts and dwell steps */
#define GSM_FORCE_RAW64 4u  /* do not pack photons: read burst_ad/burst_dd as uploaded */
#define GSM_NO_INDEX 8u     /* packed photons without the window index: photon-by-photon scan */
#define GSM_NO_TIMING 16u   /* do not bracket every run with CUDA events (gsm_last_elapsed_ms returns -1):
                               two driver calls less per evaluation, for nodes that evaluate small bursts sets */
#define GSM_PEER_SHARED_DEVICE 32u /* the caller vouches that ranks of one peer exchange that share a GPU run
                               concurrently there (see gsm_peer_connect); without it such a layout is refused */

typedef struct gsm_ctx gsm_ctx;

int gsm_abi_version(void);
const char* gsm_last_error(void);
/* number of CUDA devices of compute capability 10.x visible to the process */
int gsm_device_count(void);

/* mc::mc(devid, streamNum, debug, hdf5size, profiler)   mc.cu:212-324
 * stream_num == 0 selects 2 x asyncEngineCount as the reference does (mc.cu:61-66,219-225). */
int gsm_create(int devid, int stream_num, unsigned flags, gsm_ctx** out);
/* mc::~mc()   mc.cu:618-660  (does NOT call cudaDeviceReset: the process may own other contexts) */
void gsm_destroy(gsm_ctx* ctx);
/* the profiler switch of mc::mc / mc::~mc (cudaProfilerStart / cudaProfilerStop when the node runs with
 * -P/--profiler, mc.cu:207-210,655-658): on != 0 starts, 0 stops the collection of an attached profiler */
int gsm_profiler(int on);
/* mc::set_gpuid()   mc.cu:199-211 : bind the calling thread to the context's device */
int gsm_set_device(gsm_ctx* ctx);
/* mc::getStreams(&n)   mc.cu:661-664 */
int gsm_stream_count(const gsm_ctx* ctx);
void* gsm_stream(const gsm_ctx* ctx, int sid); /* cudaStream_t */
int gsm_device_id(const gsm_ctx* ctx);
/* mc::gpuuuid : 32 hex digits + NUL */
int gsm_device_uuid(const gsm_ctx* ctx, char out[33]);

/* mc::set_reSampleTimes(n)   mc.cu:316-318 ; any R >= 1 (the reference needs a power of two) */
int gsm_set_resample_times(gsm_ctx* ctx, int R);
int gsm_resample_times(const gsm_ctx* ctx);
/* key of the counter-based generator; the reference seeds from cuRAND tables (mc.cu:366-455) */
int gsm_set_seed(gsm_ctx* ctx, uint32_t seed);
/* Arithmetic mode of the evaluation (no counterpart in the reference, whose only mode is
 * libdevice logf/powf + FP64 rk_binomial on cuRAND Sobol draws, mc.cu:83-197):
 *   GSM_MODE_REPLAY  deterministic replay: explicit-rounding arithmetic, FP64 binomial on the
 *                    xor128 stream; integer counts, bins and chi-square bits equal the oracle's
 *   GSM_MODE_NORMAL  FP32 arithmetic on the hardware lg2/ex2/rcp units, FP32 binomial inversion;
 *                    same distributions (KS p > 0.01, chi-square within 1 %), about 1.3x faster
 * Results of either mode are a pure function of (data, parameters, seed, epoch). */
#define GSM_MODE_REPLAY 0
#define GSM_MODE_NORMAL 1
/*   GSM_MODE_REFERENCE  replay arithmetic on the REFERENCE'S OWN generators: one cuRAND scrambled
 *                    Sobol64 sequence per (burst, repeat) thread (dimension tidx % 20000, offset 1234,
 *                    mc.cu:366-455, gen_rand.cuh:7-25) drawn in the reference's order, and its xor128
 *                    stream; both persist across evaluations of a stream slot and restart when N
 *                    changes, as gpuWorker drives them (gpuWorker.cpp:74-77).  Per-sample results,
 *                    histogram and chi-square then equal the reference's class mc bit for bit.
 *                    One vector per launch; needs libcurand (tables only) at run time. */
#define GSM_MODE_REFERENCE 2
int gsm_set_mode(gsm_ctx* ctx, int mode);
int gsm_mode(const gsm_ctx* ctx);
/* fretHistNum / leftPad of streamWorker (streamWorker.cpp:41,62-70; main.cpp:155); defaults 200, 2 */
int gsm_set_hist(gsm_ctx* ctx, int n_bins, int left_pad);

/* mc::init_data_gpu(...)   mc.cu:456-521 : one-time upload of the burst arrays.
 * g_offset is the global index of burst 0 when a dataset is split across GPUs (0 otherwise).
 * Also builds the observed histogram of SgDivSr (streamWorker.cpp:62-70). */
int gsm_init_data(gsm_ctx* ctx, int64_t n_bursts, const int64_t* istart, const int64_t* start,
                  const int64_t* stop, const int32_t* ph_count, int64_t n_photons,
                  const int64_t* burst_ad, const int64_t* burst_dd, const float* T,
                  const float* SgDivSr, float clk_p, float bg_ad_rate, float bg_dd_rate,
                  int64_t g_offset);
/* mc::sz_burst */
int64_t gsm_n_bursts(const gsm_ctx* ctx);
/* 1 if the photons were packed to 4 bytes each, 0 if the raw int64 layout is in use */
int gsm_is_packed(const gsm_ctx* ctx);
/* photon layout in HBM: 0 raw int64 ticks (16 B/photon), 1 packed words (4 B/photon),
 * 2 packed words + running-maximum + next-photon words (12 B/photon, O(log n) windows) */
int gsm_layout(const gsm_ctx* ctx);
/* observed histogram, n_bins floats (streamWorker's ohist) */
int gsm_observed_hist(gsm_ctx* ctx, float* out, int n_bins);
/* replace the observed histogram (burst-split runs hand every rank the global one) */
int gsm_set_observed_hist(gsm_ctx* ctx, const float* ohist, int n_bins);
/* The observed histogram without any allocation.  Has no counterpart in the reference (which allocates
 * wherever it likes, mc.cu:456-521); it exists for hosts that cannot let this library do that: a process
 * whose device memory is already committed to its own pool or arena, so that a cudaMalloc from a plug-in
 * fails or fragments it; a generation loop or acquisition callback that cannot afford the device-wide
 * synchronisation cudaMalloc / cudaFree bring; a worker that is handed buffers and a stream and nothing else.
 *   d_SgDivSr  n_bursts floats   DEVICE memory of the current device (input)
 *   d_out      n_bins floats     DEVICE memory of the current device (output: the same values
 *                                gsm_observed_hist returns after gsm_init_data on the same SgDivSr)
 *   stream     a cudaStream_t of the current device (NULL = the default stream); the work is enqueued
 *              there and the call returns without waiting for it
 * No context is needed.  The 64-bit bin counters the kernel accumulates into live in an array built into
 * the library image, (GSM_STATIC_MAX_BINS + 2) x 8 bytes = 32 784 bytes on every GPU that loads
 * libgsmfret.so, held until the process exits.  n_bins beyond GSM_STATIC_MAX_BINS is refused with
 * GSM_ERANGE (use the allocating path for such an axis).  The call makes no cudaMalloc / cudaFree and uses
 * no allocating library routine.
 * First call on a device: it looks the array's address up once (cudaGetSymbolAddress, with that device
 * current, under a lock) and keeps it; make this first call outside any stream capture.
 * The array is ONE per GPU and keeps what the last call left: every call clears it first, on `stream`.
 * Two calls whose work overlaps on the same GPU (different streams, different threads or different
 * contexts of this process) add into the same counters and corrupt each other: order them on one stream
 * or with events. */
#define GSM_STATIC_MAX_BINS 4096
int gsm_observed_hist_static(const float* d_SgDivSr, int64_t n_bursts, int n_bins, float* d_out, void* stream);

/* mc::get_max_conversion_capacity()   mc.cu:326-365 : dead buffer grab in the reference; no-op, returns 1 */
int gsm_get_max_conversion_capacity(gsm_ctx* ctx, int max_state_num);
/* mc::set_nstates(n, sid) -> previous n   mc.cu:723-782 */
int gsm_set_nstates(gsm_ctx* ctx, int n, int sid);
/* mc::set_params(n, sid, args) -> 1, or 0 when n_args != n*n+n (eigenhelper.cpp:5-8)   mc.cu:790-829
 * args = [E_0..E_{n-1} | k (n*n-n, row-major without the diagonal) | v_0..v_{n-1}] */
int gsm_set_params(gsm_ctx* ctx, int n, int sid, const float* args, int n_args);
/* mc::setBurstBd(cstart, cstop, sid) -> N (-1 if cstart == -1)   mc.cu:523-566
 * As in the reference the evaluated bursts are [0, N): the range start is not applied (quirk Q7). */
int gsm_set_burst_bd(gsm_ctx* ctx, int cstart, int cstop, int sid);
/* mc::init_randstate(N, sid)   mc.cu:366-455 : nothing to allocate or seed with a
 * counter-based generator (the draw epoch advances with every run); no-op */
int gsm_init_randstate(gsm_ctx* ctx, int N, int sid);
/* mc::set_params_buff(oldS_n, N_sid, sid)   mc.cu:783-789 : no scratch needed; no-op */
int gsm_set_params_buff(gsm_ctx* ctx, int old_n, int N_sid, int sid);
/* explicit draw epoch of the next run on the slot (replay); otherwise it advances by one per run */
int gsm_set_epoch(gsm_ctx* ctx, int sid, uint32_t epoch);

/* mc::run_kernel(N, sid)   mc.cu:567-617 : asynchronous; enqueues the evaluation of bursts
 * [0, N) x R repeats with the slot's parameters and the copy of the results to pinned host
 * memory on the slot's stream, and returns. */
int gsm_run(gsm_ctx* ctx, int N, int sid);
/* batch entry: P parameter vectors of n states in one launch (args = P rows of n*n+n floats);
 * vector p draws with epoch + p.  Results via gsm_result_chi2v / gsm_result_histv. */
int gsm_run_batch(gsm_ctx* ctx, int N, int sid, int n, int P, const float* args);
/* The single vectors of several stream slots in ONE launch on a carrier slot (not in the reference:
 * its gpuWorker issues one launch per stream, gpuWorker.cpp:38-87, which bounds a node at the
 * reference's own job size by the driver calls of the issuing thread).  args = P rows of n*n+n
 * floats, member_sids[p] = the stream vector p belongs to.  Vector p is drawn with member p's own
 * epoch — the results are the bits the member's own gsm_run would have produced — and that epoch
 * advances by one.  Until a member runs by itself again, gsm_query / gsm_sync / gsm_result_chi2 /
 * gsm_result_hist(…, 0) / gsm_result_flags on the MEMBER read vector p of the carrier.  The carrier
 * must not be a member and must not be launched again before the members' results have been read.
 * Not with GSM_KEEP_MCE / GSM_KEEP_COUNTS (per-sample outputs are kept per stream). */
int gsm_run_coalesced(gsm_ctx* ctx, int N, int carrier_sid, int n, int P, const float* args, const int* member_sids);
/* same, the vectors of the previous gsm_run_batch / gsm_set_params stay resident: only launches */
int gsm_run_resident(gsm_ctx* ctx, int N, int sid);
/* mc::streamQuery(sid)   mc.cu:665-672 : 1 finished, 0 still running, <0 error */
int gsm_query(gsm_ctx* ctx, int sid);
int gsm_sync(gsm_ctx* ctx, int sid);

/* results of the last run on the slot (valid after gsm_query == 1 / gsm_sync) */
float gsm_result_chi2(const gsm_ctx* ctx, int sid);
const float* gsm_result_chi2v(const gsm_ctx* ctx, int sid, int* P);
/* histogram of vector p: n_bins inner counts, then [n_bins] overflow (E >= 1, NaN), [n_bins+1] underflow */
const uint64_t* gsm_result_hist(const gsm_ctx* ctx, int sid, int p);
/* mc::hmcE[sid] : N*R floats per vector (needs GSM_KEEP_MCE) */
const float* gsm_result_mcE(const gsm_ctx* ctx, int sid);
/* per-sample integer acceptor counts / dwell steps (need GSM_KEEP_COUNTS) */
const int64_t* gsm_result_counts(const gsm_ctx* ctx, int sid);
const int32_t* gsm_result_steps(const gsm_ctx* ctx, int sid);
/* bit 0: a sample hit the dwell-step cap (2^20; rates <= 0 hang the reference)
 * bit 1: the peer exchange timed out waiting for another rank (the histogram is partial) */
uint64_t gsm_result_flags(const gsm_ctx* ctx, int sid);
/* %globaltimer stamps (ns, this GPU's clock) the last block of the last run left behind:
 * [0] it started finalising, [1] its partial histogram and arrival flags were posted to every peer,
 * [2] every peer's flag had arrived, [3] results written.  [1] = [2] = [0] without a peer exchange.
 * [2] - [1] on the rank that arrived last is the cost of the exchange itself; on the other ranks it
 * also holds the time they were ahead (skew).  Returns GSM_ESTATE before the first run.  The stamps are
 * those of the launch whose results the slot reads (its carrier's after gsm_run_coalesced); read them
 * only after gsm_query returned 1 / gsm_sync — earlier they still belong to the previous run. */
int gsm_result_timing(const gsm_ctx* ctx, int sid, uint64_t out_ns[4]);
/* device time of the last run on the slot between its first and last enqueued operation, ms */
float gsm_last_elapsed_ms(gsm_ctx* ctx, int sid);
/* number of kernels this library launched on behalf of the context so far */
uint64_t gsm_launch_count(const gsm_ctx* ctx);

/* burst-split evaluation across GPUs (new capability): every rank holds a
 * contiguous range of the bursts; each run all-reduces the [P x (n_bins+2)] histograms with
 * NCCL over NVLink before chi-square is taken, so every rank reports the global result. */
int gsm_comm_unique_id(void* out128);               /* ncclGetUniqueId, 128 bytes */
int gsm_comm_init(gsm_ctx* ctx, const void* id128, int rank, int world);
int gsm_comm_destroy(gsm_ctx* ctx);

/* The same reduction fused into the evaluation kernel (no library call, no second launch): the
 * last block of every rank's launch stores its partial histogram into an inbox on EVERY rank —
 * peer memory over NVLink / NVSwitch, mapped with CUDA IPC between processes or used directly
 * between contexts of one process — raises an arrival flag there, waits on its own flags and
 * sums its own inboxes, then takes chi-square.  Integer bins: every rank ends with identical bits.
 *   1. every rank: gsm_peer_export(ctx, world, max_P, handle)  (allocates the symmetric block;
 *      max_P = most parameter vectors per launch; call after gsm_set_hist)
 *   2. the caller gathers the world handles in rank order (any transport)
 *   3. every rank: gsm_peer_connect(ctx, handles, rank, world)
 * Every rank must then issue the same sequence of gsm_run* calls per stream slot.  A rank whose
 * peer does not arrive within the timeout (default 20 s) reports flag bit 1 in gsm_result_flags.
 * Exclusive with gsm_comm_init.
 * Forward progress: the last block of a launch WAITS inside the kernel for the other ranks'
 * launches, so all ranks of an exchange must be able to execute at the same time.  One rank per
 * GPU (the intended layout) guarantees that.  Ranks that share a GPU — several contexts of one
 * process on one device, or several processes on one device — depend on the GPU running their
 * kernels concurrently, which CUDA does not promise: gsm_peer_connect refuses such a layout
 * (GSM_ESTATE) unless the context was created with GSM_PEER_SHARED_DEVICE, and even then only when
 * co-resident ranks x stream slots fit the device's hardware queues (CUDA_DEVICE_MAX_CONNECTIONS,
 * default 8) and CUDA_LAUNCH_BLOCKING is not set; ranks of different processes on one GPU then
 * advance by the driver's time-slicing (or MPS) — live, but slow.  Under a kernel-serialising
 * profiler (ncu) use one rank per GPU or the NCCL path (gsm_comm_init). */
#define GSM_PEER_HANDLE_BYTES 128
#define GSM_MAX_PEERS 8
int gsm_peer_export(gsm_ctx* ctx, int world, int max_P, void* handle_out128);
int gsm_peer_connect(gsm_ctx* ctx, const void* handles /* world x 128 bytes */, int rank, int world);
int gsm_peer_disconnect(gsm_ctx* ctx);
int gsm_peer_set_timeout(gsm_ctx* ctx, uint32_t milliseconds);

/* primitive probes: one element per input through the device code, for the unit tests
 * (role of the reference's test kernels, /root/reference/tests/GenRandTest.cu:107-135,272-287).
 * what: 0 philox, 1 logf, 2 log, 3 exp, 4 normal icdf, 5 sixth root, 6 u01, 7 binomial,
 * 8 poisson, 9 drawE, 10 drawTau, 11 multinomial, 12 categorical, 13 histogram index,
 * 14 the hand-scheduled division of the binomial inversion (d0[2i] / d0[2i+1] -> od[i]). */
int gsm_probe(gsm_ctx* ctx, int what, int64_t n, int ni, int nj, const uint32_t* w0,
              const uint32_t* w1, const float* f0, const float* f1, const double* d0,
              const int32_t* i0, uint32_t* ou, float* of, double* od, int64_t* ol, int32_t* oi);

/* Measured instruction-issue peak of this GPU at its current clocks: thread-level 32-bit
 * multiply-adds (kind 0: integer IMAD, kind 1: FP32 FFMA) per second over a grid that fills every
 * SM, timed with CUDA events.  The denominator of the issue roofline bench.py reports
 * ("INT32-only peak is to be measured by an IMAD micro-benchmark"). */
int gsm_issue_peak(gsm_ctx* ctx, int kind, double* thread_ops_per_s);

/* Host only (no device needed): how a launch of N bursts x R repeats x P vectors on a persistent grid of
 * `resident_blocks` blocks is cut into tickets (role of the <<<grid, block>>> arithmetic of
 * mc::set_params_buff / run_kernel, /root/reference/src/mc.cu:543,571).  guided: the tail of small chunks (what the dense
 * build's workloads get).  out6 = chunk (samples per ticket chunk), span (chunks per ticket), log2 of the tail
 * chunk, big chunks per vector, chunks per vector, blocks launched.  Sample t of a vector belongs to chunk
 * c < big ? [c * chunk, (c + 1) * chunk) : [big * chunk + (c - big) << tail_log2, ...), clipped to N * R. */
int gsm_launch_shape(int64_t N, int R, int P, int resident_blocks, int guided, int64_t* out6);

#ifdef __cplusplus
}
#endif
#endif
