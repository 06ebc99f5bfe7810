// pda_eval.cuh — the persistent evaluation kernel template, shared by the two translation
// units that instantiate it:
//   kMode: kModeReplay (0), kModeNormal (1, "kFast"), kModeReference (2, "kSobol": replay arithmetic
//   on the reference's own generator — per-thread cuRAND scrambled Sobol64 sequences and a
//   persistent xor128 stream, pda_kernels_ref.cu)
//   pda_kernels.cu       kFast = false  replay mode: explicit-rounding arithmetic, FP64 binomial on
//                                       the xor128 stream, compiled -fmad=false (bit-exact contract)
//   pda_kernels_fast.cu  kFast = true   normal mode: FP32 arithmetic with hardware lg2/ex2/rcp and
//                                       fused multiply-adds, binomial by FP32 inversion on a Philox
//                                       word; statistically equivalent, not bit-equal
// Role of the reference's mc_kernel (/root/reference/src/mc.cu:83-197) fused with the host
// histogram and chi-square of streamWorker.cpp:187-206.
#pragma once
#include "pda_kernels.h"
#include "pda_device.cuh"

namespace pda {

constexpr int kModeReplay = 0, kModeNormal = 1, kModeReference = 2;
#ifndef PDA_BLOCK
#define PDA_BLOCK 256
#endif
#ifndef PDA_CHUNK_MAX
#define PDA_CHUNK_MAX 256
#endif
constexpr int kBlock = PDA_BLOCK;
constexpr int kChunkMax = PDA_CHUNK_MAX;  // most consecutive samples a warp draws per ticket (EvalArgs::chunk)
constexpr int kMaxBins = 1024;          // nb + 2 <= kMaxBins
constexpr uint32_t kStepCap = 1u << 20; // guards the GPU against rates <= 0 (a hang in the reference)

// parameter vector staged in shared memory, with the per-state terms hoisted
struct SmemParams {
    int n;
    float E[kMaxStates];
    float v[kMaxStates];
    float rootE[kMaxStates];            // r0 * (1/E - 1)^(1/6)
    float K[kMaxStates * kMaxStates];
    float negInvK[kMaxStates * kMaxStates];  // -1/K (normal mode: tau = lg(u) * negInvK)
    float cumJ[kMaxStates][kMaxStates]; // running sums of column i of P_i2j without its diagonal
    float cumP[kMaxStates];             // running sums of p
};

// executed by one warp (tid = lane)
__device__ __forceinline__ void stage_params(SmemParams& sp, const ParamRaw* __restrict__ pr, int tid)
{
    const int n = pr->n;
    if (tid == 0) sp.n = n;
    if (tid < n) {
        float e = pr->E[tid];
        sp.E[tid] = e;
        sp.v[tid] = pr->v[tid];
        sp.rootE[tid] = kR0 * det_root6f(1.0f / e - 1.0f);
        // column tid of P_i2j without the diagonal, summed in drawDisIdx's order
        float a = 0.0f;
        for (int ii = 0; ii < n - 1; ++ii) {
            a += (ii >= tid) ? pr->P_i2j[tid * n + ii + 1] : pr->P_i2j[tid * n + ii];
            sp.cumJ[tid][ii] = a;
        }
    }
    for (int i = tid; i < n * n; i += 32) {
        const float k = pr->K[i];
        sp.K[i] = k;
        sp.negInvK[i] = -1.0f / k;
    }
    if (tid == 31) {
        float a = 0.0f;
        for (int i = 0; i < n; ++i) {
            a += pr->p[i];
            sp.cumP[i] = a;
        }
    }
}

// first index whose running sum exceeds pv, else m-1 (role of drawDisIdx, gen_rand.cuh:35-50)
__device__ __forceinline__ int pick(const float* cum, int m, float pv)
{
    // m is warp-uniform (one parameter vector per warp): the small cases skip the loop
    if (m <= 1) return 0;
    if (m == 2) return cum[0] > pv ? 0 : 1;
    if (m == 3) return cum[0] > pv ? 0 : (cum[1] > pv ? 1 : 2);
    // at most 7 entries (kMaxStates = 8): a rolled loop — unrolled 16-fold it was 600 instructions
    // of the hot path's code at each of its two call sites
    int r = m - 1;
#pragma unroll 1
    for (int i = m - 2; i >= 0; --i)
        if (cum[i] > pv) r = i;
    return r;
}

// ------------------------------------------------------------------------------------
// One lane carries one burst-sample at a time; every trip of the warp's main loop advances
// each lane by ONE dwell window, and a lane whose sample is finished takes the next sample of
// the warp's chunk.  Lanes therefore stay converged on the expensive part (Philox, the
// draws, the FP64 binomial) whatever the number of windows of their samples.
// ------------------------------------------------------------------------------------
struct Lane {
    // sample identity: local burst index and repeat (sample index in the vector = bl * R + rep,
    // global burst id of the draw contract = b0 + bl + g_offset)
    uint32_t bl, rep;
    // burst record (stop and, in the raw layout, the photon offset are re-read when needed)
    int64_t start;
    uint32_t off4;      // first photon of the burst in the packed stream, in units of 4 words
    uint32_t scale;     // cell index scale of the burst (indexed layout)
    float T;
    int ph;
    // walk state
    Xor128 xs;          // replay mode only
    float F, acc, spent;
    // left edge of the current window as the float the reference converts it to every step
    // ((float)bin0, mc.cu:132).  After the first window bin0 is itself the truncation of a float
    // sum, so its float image is known without a 64-bit integer conversion.
    float fb;
    int first, si;
    uint32_t step;
    // reference-draws mode: position in the thread's Sobol64 sequence
    unsigned long long sob_i, sob_x;
    const unsigned long long* sob_dv;
};

// next 64-bit word of the thread's scrambled Sobol64 sequence (cuRAND's update: flip the
// direction vector of the lowest zero bit of the index)
__device__ __forceinline__ unsigned long long sobol_next(Lane& L)
{
    const unsigned long long r = L.sob_x;
    L.sob_x ^= __ldg(L.sob_dv + (__ffsll((long long)~L.sob_i) - 1));
    L.sob_i++;
    return r;
}

__device__ __forceinline__ uint32_t lane_g(const EvalArgs& a, const Lane& L)
{
    // only the low 32 bits enter the draw counter: one three-input add on constant-bank operands
    return L.bl + (uint32_t)a.b0 + (uint32_t)a.g_offset;
}

template <bool kPacked>
__device__ __forceinline__ int64_t lane_off(const EvalArgs& a, const Lane& L)
{
    if (kPacked) return (int64_t)L.off4 << 2;
    return __ldg(&a.bursts[a.b0 + L.bl].off);
}

// warp-uniform bookkeeping of the persistent loop (one per warp, in shared memory)
struct WarpCtl {
    unsigned long long pend_item;   // a fetched chunk of another vector waits for the lanes to drain
    uint32_t c_bl, c_rep;           // next unassigned sample of the warp's chunk = (burst, repeat)
    int c_cnt;                      // samples left in the chunk
    int cur_p;                      // vector whose parameters / histogram this warp holds
    uint32_t epoch;
    int have_pending, exhausted;
    unsigned long long span_next;   // next chunk of the span of consecutive chunks this warp drew ...
    int span_left;                  // ... and how many of them are left (EvalArgs::span per ticket)
};

struct Window {
    int end;        // index of the closing photon; >= ph: the window saw no photon at all
    int f_ia, f_id; // counted acceptor / donor photons
};

// The reference's photon scan (/root/reference/src/mc.cu:140-181), photon by photon.
template <bool kPacked>
__device__ __forceinline__ Window scan_linear(const EvalArgs& a, const Lane& L, int64_t bin1, int thr)
{
    Window w{L.ph, 0, 0};
    int flags = 0;
    const int64_t off = lane_off<kPacked>(a, L);
    for (int i = L.first; i < L.ph; ++i) {
        bool close;
        int fl;
        if (kPacked) {
            int pw = __ldg(a.pk + off + i);
            close = pw >= thr;
            fl = pw & 3;
        } else {
            int64_t ddx = __ldg(a.dd + off + i), adx = __ldg(a.ad + off + i);
            close = ddx >= bin1 || adx >= bin1;
            fl = (ddx > 0 ? 1 : 0) | (adx > 0 ? 2 : 0);
        }
        if (close || i == L.ph - 1) {
            w.end = i;
            break;
        }
        flags |= fl;
        if (flags == 3) continue;
        w.f_ia += flags >> 1;
        w.f_id += flags & 1;
    }
    return w;
}

// Same result in O(1) expected: the cell index brackets the closing photon — photons in cells
// below cell(thr) are before the edge, photons in cells above it are past it — a binary search
// over the running maximum inside the bracket (0 or 1 photons on average) finishes it, and the
// counts follow from the distance to the next acceptor / donor photon.  Exact whenever no photon
// before `first` already reaches the new edge; otherwise (a window edge that moved backwards,
// unsorted ticks) the photon-by-photon scan decides.
// The two loads a window search starts with, issued by the caller BEFORE the arithmetic that does not
// depend on them (the efficiency draw of the window): the search then finds them arrived.
struct ScanLoads { uint32_t nx, cw; };
__device__ __forceinline__ ScanLoads scan_indexed_issue(const EvalArgs& a, const Lane& L, int thr)
{
    ScanLoads s{0u, 0u};
    if (L.first < L.ph) {
        const int64_t off = lane_off<true>(a, L);
        s.nx = __ldg(a.nx + off + L.first);   // independent of the edge
        const int Q = ((L.ph + 3) & ~3) - 1;
        s.cw = __ldg(a.cw + off + cell_of(thr, L.scale, Q));
    }
    return s;
}

__device__ __forceinline__ Window scan_indexed(const EvalArgs& a, const Lane& L, int64_t bin1, int thr, const ScanLoads ld)
{
    if (L.first >= L.ph) return Window{L.ph, 0, 0};
    const int64_t off = lane_off<true>(a, L);
    const uint32_t nx = ld.nx, cw = ld.cw;
    int lo = (int)(cw & 0xffffu), hi = (int)(cw >> 16);
    const int32_t* __restrict__ pm = a.pm + off;
    // (loading the last four candidates of the bracket together instead of two or three dependent
    // loads was tried: no gain, profiles/r03b_* — the other warps hide this latency)
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(pm + mid) >= thr) hi = mid;
        else lo = mid + 1;
    }
    if (lo < L.first) return scan_linear<true>(a, L, bin1, thr);
    lo = lo < L.ph - 1 ? lo : L.ph - 1;                // the last photon closes any window that reaches it
    Window w{lo, 0, 0};
    const int ia = L.first + (int)(nx & 0xffffu);  // next acceptor photon at or after `first`
    const int id = L.first + (int)(nx >> 16);      // next donor photon
    if (ia < id) {
        int stop_at = id < lo ? id : lo;
        w.f_ia = stop_at > ia ? stop_at - ia : 0;
    } else if (id < ia) {
        int stop_at = ia < lo ? ia : lo;
        w.f_id = stop_at > id ? stop_at - id : 0;
    }
    return w;
}

// start a sample: burst record, binomial stream (role of setup_kernel, gen_rand.cuh:7-25)
template <int kMode>
__device__ __forceinline__ void lane_begin(Lane& L, const SmemParams& sp, const EvalArgs& a, uint32_t bl,
                                           uint32_t rep, uint32_t epoch)
{
    const int64_t b = a.b0 + (int64_t)bl;
    const int4 r0 = __ldg(reinterpret_cast<const int4*>(a.bursts + b));
    const int4 r1 = __ldg(reinterpret_cast<const int4*>(a.bursts + b) + 1);
    L.bl = bl;
    L.rep = rep;
    L.start = ((int64_t)(uint32_t)r0.y << 32) | (uint32_t)r0.x;
    L.T = __int_as_float(r1.x);
    L.ph = r1.y;
    L.off4 = (uint32_t)r1.z;      // packed layouts; the raw layout re-reads its 64-bit offset
    L.scale = (uint32_t)r1.w;
    L.si = 0;
    if (kMode == kModeReference) {
        // the reference's thread tidx = burst * R + repeat resumes its Sobol sequence and its
        // xor128 stream where the previous evaluation left them (mc.cu:366-455, quirk Q8)
        const size_t t = (size_t)bl * (size_t)a.R + rep;
        const ulonglong2 st = a.sob_state[t];
        L.sob_i = st.x; L.sob_x = st.y;
        L.sob_dv = a.sob_dv + 64u * (unsigned)(t % 20000u);
        const uint4 xs = a.xs_state[t];
        L.xs.s0 = xs.x; L.xs.s1 = xs.y; L.xs.s2 = xs.z; L.xs.s3 = xs.w;
    }
    // replay mode: the xor128 stream of the sample is seeded from its step-0 block (lane_step)
    L.F = 0.0f; L.acc = 0.0f; L.spent = 0.0f;
    L.fb = (float)L.start;
    L.first = 0;
    L.step = 0;
}

// reference-draws mode: leave the generators where the next evaluation resumes them
__device__ __forceinline__ void lane_save_reference(const Lane& L, const EvalArgs& a)
{
    const size_t t = (size_t)L.bl * (size_t)a.R + L.rep;
    a.sob_state[t] = make_ulonglong2(L.sob_i, L.sob_x);
    a.xs_state[t] = make_uint4(L.xs.s0, L.xs.s1, L.xs.s2, L.xs.s3);
}

// drawE (/root/reference/src/gen_rand.cuh:63-68) with the per-state root hoisted: E of the window from one normal word
__device__ __forceinline__ float draw_efficiency(uint32_t word, const SmemParams& sp, int si)
{
    const float rd = det_norm_icdf(word) * sp.v[si] + sp.rootE[si];
    return 1.0f / (1.0f + det_pow6f(div_by_r0(rd)));
}

__device__ __forceinline__ float draw_efficiency_fast(uint32_t word, const SmemParams& sp, int si)
{
    const float rd = fmaf(fast_norm_icdf(word), sp.v[si], sp.rootE[si]);
    const float x = rd * (1.0f / kR0);
    const float x2 = x * x;
    return fast_rcp(fmaf(x2 * x2, x2, 1.0f));
}

// one dwell window (one trip of the reference's while loop, mc.cu:117-191)
template <int kLayout, bool kBg, int kMode>
__device__ __forceinline__ void lane_step(Lane& L, const SmemParams& sp, const EvalArgs& a, uint32_t epoch,
                                          const LogTabShared ltab, uint32_t rcp_sa)
{
    constexpr bool kFast = kMode == kModeNormal;
    constexpr bool kSobol = kMode == kModeReference;
    const int n = sp.n;
    const uint32_t g = lane_g(a, L);
    u32x4 w;
    if (kSobol) {
        // the reference's draw order: initial state (first step only), next state, dwell;
        // the E draw and the background draws are taken further down, only when they happen
        w.w = L.step == 0 ? (uint32_t)(sobol_next(L) >> 32) : 0u;
        w.x = (uint32_t)(sobol_next(L) >> 32);
        w.y = (uint32_t)(sobol_next(L) >> 32);
        w.z = 0u;
    } else {
        w = philox4x32_10(g, L.rep, L.step, kSlotStep, a.seed, epoch);
    }
    if (L.step == 0) {
        // initial state (mc.cu:110).  Counter modes: the low bytes of the words — the uniforms of
        // this block are built from the top 24 bits of each word, so those bits are an independent
        // word that costs no Philox call.  Replay: word 3 seeds the sample's xor128 stream (the
        // role of setup_kernel's rk_seed(curand()), gen_rand.cuh:19-21); normal mode spends it on
        // the binomial and takes a fourth low byte from it.
        uint32_t iw = w.w;
        if (kFast) iw = (w.x & 0xffu) | ((w.y & 0xffu) << 8) | ((w.z & 0xffu) << 16) | (w.w << 24);
        else if (!kSobol) {
            iw = ((w.x & 0xffu) << 24) | ((w.y & 0xffu) << 16) | ((w.z & 0xffu) << 8) | 0x80u;
            L.xs.seed((uint64_t)w.w);
        }
        L.si = pick(sp.cumP, n, u01(iw));
    }
    const int si = L.si;
    int sj = pick(sp.cumJ[si], n - 1, u01(w.x));
    if (sj >= si) ++sj;
    float tau;
    if (kFast) tau = __logf(u01(w.y)) * sp.negInvK[sj + si * n];
    else tau = det_logf(u01(w.y)) / (-sp.K[sj + si * n]);
    L.spent = L.spent + tau;
    int64_t bin1;
    // bg: the left edge as an integer again (exact: fb holds (float)bin0, and past the first
    // window bin0 is the truncation of that very float)
    const int64_t bin0 = kBg ? (L.step == 0 ? L.start : (int64_t)L.fb) : 0;
    if (L.spent >= L.T) {
        bin1 = __ldg(&a.bursts[a.b0 + L.bl].stop);   // last window of the sample: fb is not read again
    } else {
        const float f = L.fb + (kFast ? __fdividef(L.spent, a.clk_p) : L.spent / a.clk_p);
        bin1 = (int64_t)f;
        // (float)bin1: truncation toward zero is exact in float while the conversion does not saturate
        L.fb = fabsf(f) < 9.0e18f ? truncf(f) : (float)bin1;
    }

    constexpr bool kHoistE = !kSobol && kLayout == 2;
    float de_early = 0.0f;
    Window win;
    if (kLayout == 0) {
        win = scan_linear<false>(a, L, bin1, 0);
    } else {
        // photon word = rel*4 + flags compares against rel1*4; a window edge at or below
        // tick 0 closes on any photon (0 >= bin1 holds for the empty channels too)
        int64_t rel = bin1 - L.start;
        rel = rel < -(1ll << 29) + 1 ? -(1ll << 29) + 1 : rel;
        rel = rel > (1ll << 29) - 1 ? (1ll << 29) - 1 : rel;
        const int thr = bin1 <= 0 ? INT32_MIN : (int)rel * 4;
        if (kLayout == 2) {
            const ScanLoads ld = scan_indexed_issue(a, L, thr);
            // the efficiency of this window (if it closes on a photon: nearly always) does not depend on
            // the search: drawn here, it runs while the index loads are in flight.  Reference-draws mode
            // takes its word from the Sobol stream only when the window closes, so it cannot.
            if (kHoistE) de_early = kFast ? draw_efficiency_fast(w.z, sp, si) : draw_efficiency(w.z, sp, si);
            // (the Philox block of the background draws could go here too: measured, no gain —
            // profiles/r03d_*)
            win = scan_indexed(a, L, bin1, thr, ld);
        } else {
            win = scan_linear<true>(a, L, bin1, thr);
        }
    }

    int ai = 0;
    if (win.end < L.ph) {
        L.first = win.end;
        const float f_if = (kGamma - kDch2Ach) * (float)win.f_id + (1.0f - kDexDirAem) * (float)win.f_ia;
        float f_i;
        L.F += f_if;
        if (!kBg) {
            f_i = floorf(f_if + 0.5f);
        } else if (kFast) {
            const float t_diff = (float)(bin1 - bin0) * a.clk_p;
            const float rf_ia = (1.0f - kDexDirAem) * (float)win.f_ia;
            const float rf_id = (kGamma - kDch2Ach) * (float)win.f_id;
            // one Philox block serves both channels: x/y acceptor (Poisson word, fall-back
            // uniform), z/w donor
            const u32x4 bw = philox4x32_10(g, L.rep, L.step, kSlotBg, a.seed, epoch);
            float sv = (float)poisson_fast(bw.x, t_diff * a.bg_ad_rate);
            const float bg_a = (sv <= rf_ia) ? sv : floorf(rf_ia * u01(bw.y));
            sv = (float)poisson_fast(bw.z, t_diff * a.bg_dd_rate);
            const float bg_d = (sv <= rf_id) ? sv : floorf(rf_id * u01(bw.w));
            f_i = floorf(f_if - bg_d - bg_a + 0.5f);
        } else if (kSobol) {
            // draw_P_B_Tr (gen_rand.cuh:92-107): one 64-bit word per Poisson draw, a further
            // word only when the draw exceeds the photons there are
            const float t_diff = (float)(bin1 - bin0) * a.clk_p;
            const float rf_ia = (1.0f - kDexDirAem) * (float)win.f_ia;
            unsigned long long x = sobol_next(L);
            uint32_t sv = poisson_itr((uint32_t)(x >> 32), (uint32_t)x, (double)(t_diff * a.bg_ad_rate));
            const float bg_a = ((float)(int32_t)sv <= rf_ia) ? (float)(int32_t)sv
                                                             : floorf(rf_ia * u01((uint32_t)(sobol_next(L) >> 32)));
            const float rf_id = (kGamma - kDch2Ach) * (float)win.f_id;
            x = sobol_next(L);
            sv = poisson_itr((uint32_t)(x >> 32), (uint32_t)x, (double)(t_diff * a.bg_dd_rate));
            const float bg_d = ((float)(int32_t)sv <= rf_id) ? (float)(int32_t)sv
                                                             : floorf(rf_id * u01((uint32_t)(sobol_next(L) >> 32)));
            f_i = floorf(f_if - bg_d - bg_a + 0.5f);
        } else {
            const float t_diff = (float)(bin1 - bin0) * a.clk_p;
            const float rf_ia = (1.0f - kDexDirAem) * (float)win.f_ia;
            // one block for both channels: x / z the high word of each Poisson draw (low word 0),
            // y / w the fall-back uniforms
            const u32x4 bw = philox4x32_10(g, L.rep, L.step, kSlotBg, a.seed, epoch);
            uint32_t sv = poisson_itr_smem(bw.x, 0u, (double)(t_diff * a.bg_ad_rate), rcp_sa);
            const float bg_a = ((float)(int32_t)sv <= rf_ia) ? (float)(int32_t)sv : floorf(rf_ia * u01(bw.y));
            const float rf_id = (kGamma - kDch2Ach) * (float)win.f_id;
            sv = poisson_itr_smem(bw.z, 0u, (double)(t_diff * a.bg_dd_rate), rcp_sa);
            const float bg_d = ((float)(int32_t)sv <= rf_id) ? (float)(int32_t)sv : floorf(rf_id * u01(bw.w));
            f_i = floorf(f_if - bg_d - bg_a + 0.5f);
        }
        if (kFast) {
            const float de = kHoistE ? de_early : draw_efficiency_fast(w.z, sp, si);
            ai = binomial_fast((int)f_i, de, w.w, g, L.rep, L.step, a.seed, epoch);
        } else {
            if (kSobol) w.z = (uint32_t)(sobol_next(L) >> 32);
            const float de = kHoistE ? de_early : draw_efficiency(w.z, sp, si);
            ai = (int)binomial(L.xs, (int)f_i, (double)de, ltab);
        }
    }
    L.acc += (float)ai;
    L.si = sj;
    L.step++;
}

// reduced chi-square for parameter vector p, one thread, bins in the host loop's order
__device__ __forceinline__ float chi2_of(const unsigned long long* __restrict__ mh,
                                         const float* __restrict__ oh, int nb, int R,
                                         int left_pad, int n_states)
{
    int effN = nb - left_pad;
    float chisqr = 0.0f;
    const float fR = (float)R;
    for (int i = left_pad; i < nb; ++i) {
        float M = __ull2float_rn(__ldcg(mh + i));
        if (M > 0.0f) {
            float mr = M / fR;
            float d = oh[i] - mr;
            double num = (double)d * (double)d;
            chisqr = (float)((double)chisqr + num / (double)mr);
        } else {
            effN--;
        }
    }
    return chisqr / (float)(effN - n_states * (n_states + 1));
}

// The same number by the whole block, for launches of a few vectors: every thread prepares the
// term of one bin (the L2 round trips and the FP64 division overlap instead of queueing behind
// each other: the one-thread loop costs ~50 us for 200 bins, most of a small evaluation), then
// one thread folds the terms in the host loop's order.  An empty bin contributes +0.0, which
// leaves the float-rounded running sum unchanged, so the bits equal chi2_of's.
__device__ __forceinline__ float chi2_block(const unsigned long long* __restrict__ mh,
                                            const float* __restrict__ oh, int nb, int R,
                                            int left_pad, int n_states, double* s_term, int* s_empty)
{
    const float fR = (float)R;
    if (threadIdx.x == 0) *s_empty = 0;
    __syncthreads();
    int empty = 0;
    for (int i = left_pad + (int)threadIdx.x; i < nb; i += blockDim.x) {
        const float M = __ull2float_rn(__ldcg(mh + i));
        double term = 0.0;
        if (M > 0.0f) {
            const float mr = M / fR;
            const float d = oh[i] - mr;
            const double num = (double)d * (double)d;
            term = num / (double)mr;
        } else {
            empty++;
        }
        s_term[i] = term;
    }
    if (empty) atomicAdd(s_empty, empty);
    __syncthreads();
    float chisqr = 0.0f;
    if (threadIdx.x == 0) {
#pragma unroll 8
        for (int i = left_pad; i < nb; ++i) chisqr = (float)((double)chisqr + s_term[i]);
        const int effN = nb - left_pad - *s_empty;
        chisqr = chisqr / (float)(effN - n_states * (n_states + 1));
    }
    __syncthreads();
    return chisqr;   // valid in thread 0
}

constexpr int kChi2BlockMaxP = 16;   // beyond this one thread per vector is faster than one block per vector

// chi2 for every p, copy the histogram out, re-zero the working set (whole block).
// s_term: nb doubles of shared memory.
__device__ __forceinline__ void finalize_block(const EvalArgs& a, double* s_term)
{
    __shared__ int s_empty;
    const int nbp = a.nb + 2;
    if (a.P <= kChi2BlockMaxP) {
        for (int p = 0; p < a.P; ++p) {
            const float c = chi2_block(a.whist + (size_t)p * nbp, a.ohist, a.nb, a.R, a.left_pad,
                                       a.use_inl ? a.inl.n : a.params[p].n, s_term, &s_empty);
            if (threadIdx.x == 0) a.chi2_out[p] = c;
        }
    } else {
        for (int p = threadIdx.x; p < a.P; p += blockDim.x)
            a.chi2_out[p] = chi2_of(a.whist + (size_t)p * nbp, a.ohist, a.nb, a.R, a.left_pad,
                                    a.params[p].n);
    }
    __syncthreads();
    const size_t tot = (size_t)a.P * nbp;
    for (size_t i = threadIdx.x; i < tot; i += blockDim.x) {
        a.hist_out[i] = __ldcg(a.whist + i);
        a.whist[i] = 0ull;
    }
    if (threadIdx.x == 0) {
        a.flags_out[0] = __ldcg(a.counters + 2);
        a.counters[0] = 0ull;
        a.counters[1] = 0ull;
        a.counters[2] = 0ull;
    }
}

// ---- fused cross-GPU reduction (burst-split sharding) ---------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// One-shot all-reduce of the working histogram by the last block of every rank: push the
// partial bins into every rank's inbox over NVLink, release an arrival flag, wait for the
// flags of all senders in the LOCAL block (no polling across the link), sum in rank order.
// Integer bins: the sum does not depend on arrival order, so every rank holds identical bits.
// stamp[1] / stamp[2]: %globaltimer when this rank's flags were posted / when every flag had arrived
__device__ __forceinline__ void peer_exchange_block(const EvalArgs& a, unsigned long long* stamp)
{
    const PeerArgs& pe = a.peer;
    const int nbp = a.nb + 2;
    const size_t tot = (size_t)a.P * nbp;
    const unsigned par = pe.seq & 1u;
    const size_t my_section = kPeerFlagWords + ((size_t)par * pe.world + pe.rank) * pe.stride;
    for (int r = 0; r < pe.world; ++r) {
        unsigned long long* dst = pe.block[r] + my_section;
        for (size_t i = threadIdx.x; i < tot; i += blockDim.x) dst[i] = __ldcg(a.whist + i);
    }
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < pe.world)
        st_release_sys(pe.block[threadIdx.x] + par * kMaxPeers + pe.rank, (unsigned long long)pe.seq);
    if (threadIdx.x == 0) stamp[1] = global_ns();
    if ((int)threadIdx.x < pe.world) {
        const unsigned long long* f = pe.block[pe.rank] + par * kMaxPeers + threadIdx.x;
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(f) != (unsigned long long)pe.seq) {
            __nanosleep(200);
            if (global_ns() - t0 > (unsigned long long)pe.timeout_ms * 1000000ull) {
                atomicOr(a.counters + 2, 2ull);   // a peer never arrived: results are partial
                break;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) stamp[2] = global_ns();
    const unsigned long long* in = pe.block[pe.rank] + kPeerFlagWords + (size_t)par * pe.world * pe.stride;
    for (size_t i = threadIdx.x; i < tot; i += blockDim.x) {
        unsigned long long s = 0ull;
        for (int r = 0; r < pe.world; ++r) s += ld_relaxed_sys(in + (size_t)r * pe.stride + i);
        a.whist[i] = s;
    }
    __syncthreads();
}

// Persistent, warp-autonomous evaluation kernel.
//   kLayout 0: raw int64 burst_ad/burst_dd (fallback), 1: packed words scanned photon by
//   photon, 2: packed words + running-max / next-photon index (O(log ph) windows).
// Every warp owns a copy of the parameter vector and a private histogram in shared memory,
// draws chunks of a.chunk consecutive samples from one global ticket, and only meets the other
// warps of its block once, at the very end (last-block finalisation).
// Two register budgets per instantiation: 4 resident blocks per SM (64 registers: the two-state
// walk without background is arithmetic-bound and pays for every spill) or 6 (40 registers, 48 warps
// per SM: the background path and walks over three or more states spend their time in divergent FP64
// loops and dependent loads, which the extra warps hide — config 5 +9 %, config 3 +4 %, config 2 -1 %,
// profiles/r02r_variants_*.txt).  The launcher picks (launch_eval_t, `dense`).
#ifndef PDA_MIN_BLOCKS
#define PDA_MIN_BLOCKS 4
#endif
#ifndef PDA_MIN_BLOCKS_DENSE
#define PDA_MIN_BLOCKS_DENSE 6
#endif
template <int kLayout, bool kBg, int kMode, bool kDense>
__global__ void __launch_bounds__(kBlock, kDense ? PDA_MIN_BLOCKS_DENSE : PDA_MIN_BLOCKS)
pda_eval_kernel(const __grid_constant__ EvalArgs a)
{
    constexpr int kWarps = kBlock / 32;
    // per warp: the staged parameter vector and the loop's bookkeeping, one block of shared memory
    // reached through ONE pinned 32-bit shared address (see `pin`: with 40 registers the compiler
    // otherwise rebuilds every such pointer from %tid and the shared window base, 5-7 instructions a
    // time, two dozen times over the loop)
    struct WarpArea { SmemParams sp; WarpCtl ctl; };
    __shared__ WarpArea s_warp[kWarps];
    extern __shared__ __align__(16) unsigned int s_hist_all[];  // [kWarps][nbp]; the last block reuses it for chi2 terms
    __shared__ int s_last;
    // the {c_j, L_j} table of det_log (FP64 binomial): divergent lookups are bank conflicts here
    // instead of L1 sectors
    __shared__ double2 s_logtab[kMode == kModeNormal ? 1 : 128];
    // RN(1/k), k <= 64, of the Poisson inversion (background path of the replay arithmetic)
    __shared__ double s_rcp[(kBg && kMode == kModeReplay) ? 65 : 1];
    if (kMode != kModeNormal) {
        for (int i = threadIdx.x; i < 128; i += kBlock) s_logtab[i] = kLogTab[i];
        if (kBg && kMode == kModeReplay)
            for (int i = threadIdx.x; i < 65; i += kBlock) s_rcp[i] = kRcpSmallInt[i];
        __syncthreads();
    }

    // shared address of the reciprocal table, taken once (see lds_f64)
    const uint32_t rcp_sa = pin((uint32_t)__cvta_generic_to_shared(s_rcp));
    const uint32_t ltab_sa = pin((uint32_t)__cvta_generic_to_shared(s_logtab));
    const int lane = threadIdx.x & 31;
    const int wid = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int nbp = a.nb + 2;
    const uint32_t warp_sa = pin((uint32_t)__cvta_generic_to_shared(&s_warp[wid]));
    WarpArea& wa = *reinterpret_cast<WarpArea*>(__cvta_shared_to_generic(warp_sa));
    SmemParams& sp = wa.sp;
    unsigned int* hist = s_hist_all + wid * nbp;
    const uint32_t hist_sa = pin((uint32_t)__cvta_generic_to_shared(hist));   // the hot add goes by shared address

    const int64_t per_p = a.N * (int64_t)a.R;  // samples per parameter vector
    // per vector: a.big_chunks chunks of a.chunk samples, then chunks of 1 << a.tail_log2 (shape_launch)
    const unsigned long long chunks_per_p = (unsigned long long)a.chunks_per_p;
    const unsigned long long items = chunks_per_p * (unsigned long long)a.P;
    const bool small = per_p < (1ll << 31) && items < (1ull << 32);

    for (int i = lane; i < nbp; i += 32) hist[i] = 0u;
    __syncwarp();

    // floor(x / R) for the cursor arithmetic (x < R + 32): one multiply-high for R < 2^15 (exact while
    // x * R < 2^32); for the larger R the quotient is 0 or 1
    const bool smallR = a.R < (1 << 15);
    const uint32_t magicR = smallR ? (uint32_t)((0x100000000ull + (uint32_t)a.R - 1u) / (uint32_t)a.R) : 0u;

    // warp-uniform bookkeeping lives in shared memory (volatile: every access is a broadcast
    // LDS / a lane-0 STS), which keeps it out of the per-thread register budget
    volatile WarpCtl& wc = wa.ctl;
    if (lane == 0) {
        wc.cur_p = -1; wc.epoch = 0u; wc.c_bl = 0u; wc.c_rep = 0u; wc.c_cnt = 0;
        wc.have_pending = 0; wc.exhausted = 0; wc.pend_item = 0ull;
        wc.span_next = 0ull; wc.span_left = 0;
    }
    __syncwarp();

    Lane L;
    bool active = false;

    for (;;) {
        // ---- hand samples to idle lanes -------------------------------------------------
        for (;;) {
            const unsigned need = __ballot_sync(0xffffffffu, !active);
            if (need == 0u) break;
            int c_cnt = wc.c_cnt;
            if (c_cnt <= 0) {
                unsigned long long item;
                if (wc.have_pending) {
                    if (need != 0xffffffffu) break;  // lanes of the previous vector still walking
                    item = wc.pend_item;
                } else {
                    if (wc.exhausted) break;
                    // a ticket is a span of consecutive chunks (one chunk for a single vector; dozens
                    // for a population, so that a warp stays on one vector instead of switching —
                    // drain, histogram flush, parameter staging — after every chunk or two)
                    const int span_left = wc.span_left;
                    if (span_left > 0) {
                        item = wc.span_next;
                        __syncwarp();
                        if (lane == 0) { wc.span_next = item + 1ull; wc.span_left = span_left - 1; }
                        __syncwarp();
                    } else {
                        item = 0;
                        if (lane == 0) item = atomicAdd(a.counters + 0, (unsigned long long)a.span);
                        item = __shfl_sync(0xffffffffu, item, 0);
                        if (item >= items) {
                            __syncwarp();
                            if (lane == 0) wc.exhausted = 1;
                            __syncwarp();
                            break;
                        }
                        const unsigned long long rest = items - item;
                        __syncwarp();
                        if (lane == 0) {
                            wc.span_next = item + 1ull;
                            wc.span_left = (int)(rest < (unsigned long long)a.span ? rest : (unsigned long long)a.span) - 1;
                        }
                        __syncwarp();
                    }
                }
                const int p = small ? (int)((uint32_t)item / (uint32_t)chunks_per_p) : (int)(item / chunks_per_p);
                const int cur_p = wc.cur_p;
                if (p != cur_p) {
                    if (need != 0xffffffffu) {   // park the chunk until every lane has drained
                        __syncwarp();
                        if (lane == 0) { wc.have_pending = 1; wc.pend_item = item; }
                        __syncwarp();
                        break;
                    }
                    // all lanes idle: switch vector
                    if (cur_p >= 0) {
                        unsigned long long* gh = a.whist + (size_t)cur_p * nbp;
                        for (int i = lane; i < nbp; i += 32) {
                            unsigned int c = hist[i];
                            if (c) atomicAdd(gh + i, (unsigned long long)c);
                            hist[i] = 0u;
                        }
                    }
                    __syncwarp();
                    stage_params(sp, a.use_inl ? &a.inl : a.params + p, lane);
                    if (lane == 0) {
                        wc.cur_p = p;
                        wc.epoch = a.epoch_per_vector ? (a.use_inl ? a.inl.epoch : a.params[p].epoch) : a.epoch0 + (uint32_t)p;
                    }
                    __syncwarp();
                }
                const unsigned long long c = item - (unsigned long long)p * chunks_per_p;
                // guided schedule: the last stretch of a vector goes out in small chunks, so that the
                // warps run dry within one SMALL chunk of each other (the tail of the persistent grid)
                const bool tail = c >= (unsigned long long)a.big_chunks;
                const int64_t t0 = tail ? (a.big_chunks << a.chunk_log2) + (int64_t)((c - (unsigned long long)a.big_chunks) << a.tail_log2)
                                        : (int64_t)(c << a.chunk_log2);
                const int len = tail ? (1 << a.tail_log2) : a.chunk;
                const int64_t left = per_p - t0;
                const int64_t bl0 = small ? (int64_t)((uint32_t)t0 / (uint32_t)a.R) : t0 / a.R;
                __syncwarp();
                if (lane == 0) {
                    wc.have_pending = 0;
                    wc.c_cnt = left < len ? (int)left : len;
                    wc.c_bl = (uint32_t)bl0;
                    wc.c_rep = (uint32_t)(t0 - bl0 * a.R);
                }
                __syncwarp();
                continue;
            }
            const uint32_t c_bl = wc.c_bl, c_rep = wc.c_rep;
            const int rank = __popc(need & lt_mask);
            const bool take = !active && rank < c_cnt;
            const int want = __popc(need);
            if (take) {
                const uint32_t x = c_rep + (uint32_t)rank;
                const uint32_t q = smallR ? __umulhi(x, magicR) : (x >= (uint32_t)a.R ? 1u : 0u);   // x < R + 32
                lane_begin<kMode>(L, sp, a, c_bl + q, x - q * (uint32_t)a.R, wc.epoch);
                active = true;
            }
            __syncwarp();
            if (lane == 0) {
                const int adv = want < c_cnt ? want : c_cnt;
                const uint32_t x = c_rep + (uint32_t)adv;
                const uint32_t q = smallR ? __umulhi(x, magicR) : (x >= (uint32_t)a.R ? 1u : 0u);
                wc.c_cnt = c_cnt - adv;
                wc.c_bl = c_bl + q;
                wc.c_rep = x - q * (uint32_t)a.R;
            }
            __syncwarp();
            // a burst of zero duration never enters the reference's while loop: E = 0
            const bool dead = take && !(L.T > 0.0f);
            if (dead) {
                if (kMode == kModeReference) {   // the initial-state draw precedes the loop (mc.cu:110)
                    (void)sobol_next(L);
                    lane_save_reference(L, a);
                }
                const size_t oi = (size_t)wc.cur_p * (size_t)per_p + (size_t)L.bl * (size_t)a.R + L.rep;
                if (a.mcE) a.mcE[oi] = 0.0f;
                if (a.sum_ai) a.sum_ai[oi] = 0;
                if (a.n_steps) a.n_steps[oi] = 0;
                atomicAdd(&hist[hist_index(0.0f, a.nb)], 1u);
                active = false;
            }
        }
        if (__ballot_sync(0xffffffffu, active) == 0u) {
            if (wc.exhausted && !wc.have_pending) break;
            continue;
        }

        // ---- one dwell window for every active lane -------------------------------------
        if (active) {
            lane_step<kLayout, kBg, kMode>(L, sp, a, wc.epoch, LogTabShared{ltab_sa}, rcp_sa);
            bool fin = !(L.T > L.spent);
            if (!fin && L.step >= kStepCap) {
                atomicOr(a.counters + 2, 1ull);
                fin = true;
            }
            if (fin) {
                if (kMode == kModeReference) lane_save_reference(L, a);
                // IEEE division in both modes: E takes few distinct values (small integer
                // counts), so a 1-ulp difference would move whole atoms of the distribution.
                // 0 / F = +0 is spelled out: a zero numerator (no acceptor photon drawn — a
                // large share of the samples) sends the division through its slow path.
                const float e = (L.F > 0.0f && L.acc != 0.0f) ? __fdiv_rn(L.acc, L.F) : 0.0f;
                if (a.mcE || a.sum_ai || a.n_steps) {
                    const size_t oi = (size_t)wc.cur_p * (size_t)per_p + (size_t)L.bl * (size_t)a.R + L.rep;
                    if (a.mcE) a.mcE[oi] = e;
                    if (a.sum_ai) a.sum_ai[oi] = (long long)L.acc;   // integer-valued, exact below 2^24
                    if (a.n_steps) a.n_steps[oi] = (int32_t)L.step;
                }
                // warp-aggregated add: one shared-memory atomic per distinct bin among the
                // lanes that finish in this trip
                const int bin = hist_index(e, a.nb);
                const unsigned fm = __activemask();
                const unsigned peers = __match_any_sync(fm, bin);
                if (lane == __ffs(peers) - 1)
                    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(hist_sa + (uint32_t)bin * 4u), "r"((unsigned)__popc(peers)) : "memory");
                active = false;
            }
        }
        __syncwarp();
    }
    const int cur_p = wc.cur_p;

    // queue drained: flush this warp's bins, then the last block to arrive finalises
    if (cur_p >= 0) {
        unsigned long long* gh = a.whist + (size_t)cur_p * nbp;
        for (int i = lane; i < nbp; i += 32) {
            unsigned int c = hist[i];
            if (c) atomicAdd(gh + i, (unsigned long long)c);
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long done = atomicAdd(a.counters + 1, 1ull);
        s_last = (done == (unsigned long long)gridDim.x - 1ull);
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        __shared__ unsigned long long s_stamp[4];
        if (threadIdx.x == 0) s_stamp[0] = s_stamp[1] = s_stamp[2] = global_ns();
        if (a.finalize) {
            if (a.finalize == 2) peer_exchange_block(a, s_stamp);
            finalize_block(a, reinterpret_cast<double*>(s_hist_all));   // every warp has flushed its bins
            if (threadIdx.x == 0) {
                a.flags_out[1] = s_stamp[0];
                a.flags_out[2] = s_stamp[1];
                a.flags_out[3] = s_stamp[2];
                a.flags_out[4] = global_ns();
            }
        } else if (threadIdx.x == 0) {
            a.counters[0] = 0ull;
            a.counters[1] = 0ull;
        }
    }
}

// ------------------------------------------------------------------------------------
// launch helpers, instantiated once per translation unit (kFast fixed by the unit)
// ------------------------------------------------------------------------------------
// the reference-draws instantiation keeps one register budget (it is a parity tool, not a fast path)
template <int kMode> struct HasDense { static constexpr bool value = kMode != kModeReference; };

template <int kLayout, bool kBg, int kMode, bool kDense>
static int resident_blocks(size_t dyn)
{
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, pda_eval_kernel<kLayout, kBg, kMode, kDense && HasDense<kMode>::value>,
                                                  kBlock, dyn);
    return nb < 1 ? 1 : nb;
}

static inline size_t eval_dyn_smem(int nb) { return (size_t)(kBlock / 32) * (size_t)(nb + 2) * sizeof(unsigned int); }

template <int kMode>
static LaunchCfg eval_launch_cfg_t(int device, int layout, bool with_bg, int nb)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t dyn = eval_dyn_smem(nb);
    int per, perd;
#define PDA_RES(LAY, BG) (per = resident_blocks<LAY, BG, kMode, false>(dyn), perd = resident_blocks<LAY, BG, kMode, true>(dyn))
    if (layout == 0) { if (with_bg) PDA_RES(0, true); else PDA_RES(0, false); }
    else if (layout == 1) { if (with_bg) PDA_RES(1, true); else PDA_RES(1, false); }
    else { if (with_bg) PDA_RES(2, true); else PDA_RES(2, false); }
#undef PDA_RES
    return LaunchCfg{sms * per, kBlock, sms * perd};
}

// Chunk and grid for one launch.  Large jobs: 256-sample chunks (lanes of a warp stay on the same
// few bursts) on the full persistent grid.  Small jobs (the reference's own scale: a few thousand
// bursts x 4 repeats) are latency-bound: hand out chunks as small as one sample per lane so that
// every warp of the machine gets work, and do not launch blocks that would find the queue empty.
static inline void shape_launch(EvalArgs& a, LaunchCfg& cfg, bool guided)
{
    const long long per_p = (long long)a.N * a.R;
    const long long total = per_p * a.P;
    const long long warps = (long long)cfg.grid * (kBlock / 32);
    long long chunk = total / (warps * 4);
    int lg = 5;                                   // 32, 64, 128 or 256 samples
    while ((2ll << lg) <= chunk && (2 << lg) <= kChunkMax) ++lg;
    if (a.force_chunk_log2 >= 5 && (1 << a.force_chunk_log2) <= kChunkMax) lg = a.force_chunk_log2;
    chunk = 1ll << lg;
    a.chunk = (int)chunk;
    a.chunk_log2 = lg;
    // Guided schedule for one big vector: the last stretch (half of what the grid holds at any moment)
    // goes out in 64-sample chunks, so that the warps run dry within ~one small chunk of each other
    // instead of one big one: config 3 +1.5 %, config 5 +1.9 %, but config 2 -0.8 % and its normal mode
    // -2.2 % (profiles/r02w_*, r02x_*) — so only where samples are long (`guided` = the dense build's
    // rule).  Small chunks all the way cost up to 9 % (ticket round trips, profiles/r02t_*).
    constexpr int kTailLog2 = 6;   // 64-sample chunks ...
    constexpr int kTailDiv = 2;    // ... over the last half of what the grid holds
    long long big = (per_p + chunk - 1) / chunk, small_n = 0;
    a.tail_log2 = lg;
    if (a.force_tail >= 0) guided = a.force_tail != 0;
    if (guided && a.P == 1 && lg > kTailLog2) {
        const long long grid_holds = warps * chunk / kTailDiv;
        const long long tail_samples = grid_holds < per_p / 8 ? grid_holds : per_p / 8;
        const long long tc = 1ll << kTailLog2;
        big = (per_p - tail_samples) / chunk;
        a.tail_log2 = kTailLog2;
        small_n = (per_p - big * chunk + tc - 1) / tc;
    }
    a.big_chunks = big;
    a.chunks_per_p = big + small_n;
    const long long items = (big + small_n) * a.P;
    // chunks per ticket: at least 64 tickets per warp keep the tail short
    long long span = items / (warps * 64);
    a.span = (int)(span < 1 ? 1 : (span > 32 ? 32 : span));
    if (a.force_span >= 1) a.span = a.force_span;
    long long blocks = (items + kBlock / 32 - 1) / (kBlock / 32);
    if (blocks < 1) blocks = 1;
    if (blocks < cfg.grid) cfg.grid = (int)blocks;
}

template <int kMode>
static cudaError_t launch_eval_t(const EvalArgs& a_in, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, bool dense)
{
    if (a_in.nb + 2 > kMaxBins) return cudaErrorInvalidValue;
    EvalArgs a = a_in;
    constexpr bool kHasDense = HasDense<kMode>::value;
    if (dense) cfg.grid = cfg.grid_dense;
    // the guided tail pays where samples are long (the workloads that also take the dense build)
    shape_launch(a, cfg, dense);
    const size_t dyn = eval_dyn_smem(a.nb);
#define PDA_LAUNCH(LAY, BG)                                                                   \
    do {                                                                                      \
        if (dense) pda_eval_kernel<LAY, BG, kMode, kHasDense><<<cfg.grid, cfg.block, dyn, s>>>(a); \
        else pda_eval_kernel<LAY, BG, kMode, false><<<cfg.grid, cfg.block, dyn, s>>>(a);      \
    } while (0)
    if (layout == 0) { if (with_bg) PDA_LAUNCH(0, true); else PDA_LAUNCH(0, false); }
    else if (layout == 1) { if (with_bg) PDA_LAUNCH(1, true); else PDA_LAUNCH(1, false); }
    else { if (with_bg) PDA_LAUNCH(2, true); else PDA_LAUNCH(2, false); }
#undef PDA_LAUNCH
    return cudaGetLastError();
}

}  // namespace pda
