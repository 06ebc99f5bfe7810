// pda_kernels.h — launch interface between the C-ABI runtime (gsmfret_c.cu) and the
// sm_100a kernels (pda_kernels.cu).  Internal; the public boundary is include/gsmfret_c.h.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace pda {

constexpr int kMaxStates = 8;
constexpr int kMaxPeers = 8;      // GPUs of one NVSwitch domain that can split one dataset
constexpr int kPeerFlagWords = 2 * kMaxPeers;  // arrival flags [parity][sender] in front of the inboxes

// One burst, 32 bytes, read with two 16-byte loads (role of the reference's five parallel
// per-burst arrays start/stop/istart/phCount/T, /root/reference/src/mc.cu:456-521).
struct __align__(16) BurstRec {
    int64_t start;
    int64_t stop;
    float T;
    int32_t ph;
    union {
        int64_t off;          // raw layout: first photon in burst_ad / burst_dd
        struct {
            uint32_t off4;    // packed layouts: first photon word / 4 (bursts are 16-byte aligned)
            uint32_t scale;   // cell index: cell(word) = min(Q-1, umulhi(max(word,0), scale))
        };
    };
};
static_assert(sizeof(BurstRec) == 32, "BurstRec must be 32 bytes");

// One parameter vector as mc::set_params leaves it on the device
// (/root/reference/src/mc.cu:790-829): E, v, K, P_i2j, p.
struct ParamRaw {
    int32_t n;
    float E[8];
    float v[8];
    float K[64];      // column-major n x n, K(i,j) = rate j -> i
    float P_i2j[64];  // column-major n x n
    float p[8];       // unnormalised inflow rates
    uint32_t epoch;   // draw epoch of this vector when EvalArgs::epoch_per_vector (coalesced launches)
};

// Cross-GPU histogram exchange fused into the evaluation kernel (burst-split sharding,
// SURVEY.md §8e).  Every rank owns one symmetric block per stream slot,
//   u64 flag[2][kMaxPeers] | u64 inbox[2][world][stride]
// mapped into every peer's address space (CUDA IPC over NVLink, or the same pointer inside one
// process).  The last block of a rank's launch stores its partial histogram into inbox[parity]
// [rank] of EVERY rank, releases flag[parity][rank] = seq there, then waits on its own flags and
// sums its own inboxes: a one-shot all-reduce without a second launch or a library call.
struct PeerArgs {
    unsigned long long* block[kMaxPeers];  // rank r's block for this slot, in this process' mapping
    unsigned long long stride;             // u64 words per sender section (capacity P x (nb+2))
    uint32_t seq;                          // evaluation number on the slot (>= 1); parity selects the buffer
    int32_t world;
    int32_t rank;
    uint32_t timeout_ms;                   // give up waiting for a peer after this long (flag bit 1)
};

struct EvalArgs {
    const BurstRec* bursts;
    const int32_t* pk;         // packed photons, (rel_tick << 2) | acceptor << 1 | donor ; or null
    const int32_t* pm;         // running maximum of pk & ~3 inside each burst ; or null
    const uint32_t* nx;        // distance to the next acceptor (low 16) / donor (high 16) photon ; or null
    const uint32_t* cw;        // cell index: photons in cells below c (low 16) / below c+1 (high 16) ; or null
    const int64_t* ad;         // raw burst_ad / burst_dd (fallback layout); or null
    const int64_t* dd;
    const ParamRaw* params;    // [P]
    const float* ohist;        // [nb] observed counts as float
    unsigned long long* whist; // [P][nb+2] working histogram, zero on entry, zero on exit if finalize
    unsigned long long* hist_out;  // [P][nb+2]
    float* chi2_out;           // [P]
    unsigned long long* flags_out; // [1] error flags of the run
    unsigned long long* counters;  // [0] tile ticket, [1] finished blocks, [2] error flags
    float* mcE;                // optional [P][N*R]
    long long* sum_ai;         // optional [P][N*R]
    int32_t* n_steps;          // optional [P][N*R]
    int64_t b0;                // first burst evaluated
    int64_t N;                 // bursts evaluated
    int64_t g_offset;          // global burst id of local burst 0 (burst-split sharding)
    int32_t R;
    int32_t P;
    int32_t nb;
    int32_t left_pad;
    int32_t chunk;             // consecutive samples per ticket (32 << k; set by the launcher)
    int32_t chunk_log2;
    int32_t span;              // consecutive chunks a warp draws per ticket (set by the launcher)
    int32_t tail_log2;         // guided schedule: after big_chunks chunks of `chunk` samples a vector goes out in
    int64_t big_chunks;        //   chunks of 1 << tail_log2 samples; chunks_per_p = both kinds together
    int64_t chunks_per_p;
    // launch-shape overrides for tests and measurements (GSM_CHUNK / GSM_SPAN / GSM_TAIL at gsm_create;
    // 0 / 0 / -1: the launcher decides).  Results do not depend on any of them.
    int32_t force_chunk_log2, force_span, force_tail;
    int32_t finalize;          // 1: the last block reduces chi2 and resets the working set
                               // 2: the same after the peer exchange (PeerArgs); 0: leave whist for a library all-reduce
    uint32_t seed;
    uint32_t epoch0;           // parameter vector p draws with epoch0 + p ...
    int32_t epoch_per_vector;  // ... or, when set, with its own ParamRaw::epoch
    float clk_p, bg_ad_rate, bg_dd_rate;
    PeerArgs peer;
    // reference-draws mode (GSM_MODE_REFERENCE): the reference's per-thread generators
    const unsigned long long* sob_dv;   // [20000][64] direction vectors (CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6)
    ulonglong2* sob_state;              // [N*R] (index i, word x) of every thread's sequence
    uint4* xs_state;                    // [N*R] xor128 state of rk_state
    // single-vector launches carry the vector in the launch parameters: no upload on the stream
    int32_t use_inl;
    ParamRaw inl;
};
static_assert(sizeof(EvalArgs) <= 4096, "EvalArgs travels as a kernel parameter");

struct LaunchCfg {
    int grid;        // SMs x resident blocks of the 64-register build
    int block;
    int grid_dense;  // the same for the dense build (more resident warps, fewer registers)
};

// how many blocks to launch (persistent grid: SMs x resident blocks)
// layout: 0 raw int64 ticks, 1 packed words scanned linearly, 2 packed words + index
// mode: 0 replay (bit-exact draw contract), 1 normal (FP32 fast arithmetic)
LaunchCfg eval_launch_cfg(int device, int layout, bool with_bg, int nb, int mode);
// dense: the build with more resident warps per SM (pda_eval.cuh, PDA_MIN_BLOCKS_DENSE)
cudaError_t launch_eval(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, int mode, bool dense);
// reference-draws instantiation (pda_kernels_ref.cu) and the role of setup_kernel (gen_rand.cuh:7-25):
// thread t starts dimension t % 20000 at offset 1234 and seeds its xor128 stream from the first word
LaunchCfg eval_launch_cfg_ref(int device, int layout, bool with_bg, int nb);
cudaError_t launch_eval_ref(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, bool dense);
cudaError_t launch_setup_ref(const unsigned long long* dv, const unsigned long long* scramble, long long NN,
                             ulonglong2* sob_state, uint4* xs_state, cudaStream_t s);
// the normal-mode instantiation lives in its own translation unit (pda_kernels_fast.cu, fmad on)
LaunchCfg eval_launch_cfg_fast(int device, int layout, bool with_bg, int nb);
cudaError_t launch_eval_fast(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, bool dense);
// the launcher's ticket arithmetic on its own (gsm_launch_shape): fills a.chunk .. a.chunks_per_p, returns blocks
int eval_launch_shape(EvalArgs& a, int resident_blocks, bool guided);
// chi2 + copy-out + reset of the working histogram (used after a cross-GPU reduction)
cudaError_t launch_finalize(const EvalArgs& a, cudaStream_t s);

// one-time data preparation
struct PackArgs {
    const int64_t* start; const int64_t* stop; const int64_t* istart; const int32_t* ph; const float* T;
    const int64_t* ad; const int64_t* dd;
    const int64_t* poff;     // [B] packed offset of each burst (multiple of 4)
    BurstRec* bursts;        // out
    int32_t* pk;             // out, may be null (raw layout: bursts only)
    int32_t* pm;             // out, may be null (no index)
    uint32_t* nx;            // out, may be null
    uint32_t* cw;            // out, may be null
    unsigned int* bad;       // out: number of photons that do not fit the packed layout
    int64_t B;
};
cudaError_t launch_pack(const PackArgs& a, cudaStream_t s);
cudaError_t launch_hist_f32(const float* x, int64_t n, int nb, unsigned long long* hist, cudaStream_t s);
cudaError_t launch_u64_to_f32(const unsigned long long* in, float* out, int n, cudaStream_t s);

// primitive probes for the test-suite (one draw per element, through the device code)
struct ProbeArgs {
    int what; int64_t n; int ni; int nj;
    const uint32_t* w0; const uint32_t* w1; const float* f0; const float* f1; const double* d0;
    const int32_t* i0;
    uint32_t* ou; float* of; double* od; long long* ol; int32_t* oi;
};
cudaError_t launch_probe(const ProbeArgs& a, cudaStream_t s);
// 16 independent multiply-adds per trip per thread, 256 threads per block (issue-rate roofline)
cudaError_t launch_issue_probe(unsigned int* out, int grid, int iters, bool integer, cudaStream_t s);

}  // namespace pda
