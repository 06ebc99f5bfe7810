// pda_kernels.cu — the Monte-Carlo PDA evaluation for sm_100a (compile with -fmad=false).
//
// One thread walks one burst-sample (burst b, repeat r, parameter vector p): the role of
// the reference's mc_kernel (/root/reference/src/mc.cu:83-197), fused with the E histogram
// (streamWorker.cpp:187-193) and the reduced chi-square (streamWorker.cpp:197-206) that
// the reference computes on the host after copying N*R floats back.
//
// Design:
//  * persistent grid of autonomous warps; one ticket counter hands out chunks of consecutive
//    samples, a lane that finishes its sample takes the next one of the warp's chunk, so the
//    warp stays converged on the expensive part whatever the window count per sample, and
//    lanes of a warp are repeats of the same / neighbouring bursts (shared photon words);
//  * photons are one packed 32-bit word each, (tick - burst start) << 2 | channel flags,
//    instead of the reference's two int64, plus a running-maximum word and a next-photon
//    word per photon that turn the per-window photon scan into a binary search;
//  * RNG is Philox4x32-10 keyed by (global burst, repeat, step, slot): no state in memory;
//  * per-block shared-memory histogram with warp-aggregated atomics, one 64-bit global
//    atomic per non-empty bin per block, and the last block to finish reduces chi-square,
//    copies the histogram out and re-zeroes the working set: one launch per evaluation.
#include "pda_kernels.h"
#include "pda_device.cuh"

namespace pda {

constexpr int kBlock = 256;
constexpr int kChunk = 256;             // consecutive samples a warp draws per ticket
constexpr int kMaxBins = 1024;          // nb + 2 <= kMaxBins
constexpr uint32_t kStepCap = 1u << 20; // guards the GPU against rates <= 0 (a hang in the reference)

// parameter vector staged in shared memory, with the per-state terms hoisted
struct SmemParams {
    int n;
    float E[kMaxStates];
    float v[kMaxStates];
    float rootE[kMaxStates];            // r0 * (1/E - 1)^(1/6)
    float K[kMaxStates * kMaxStates];
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
    for (int i = tid; i < n * n; i += 32) sp.K[i] = pr->K[i];
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
    int r = m - 1;
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
    // sample identity
    int64_t t;          // sample index inside the parameter vector: burst_local * R + repeat
    uint32_t g, rep;
    // burst record
    int64_t start, stop, off;
    float T;
    int ph;
    // walk state
    Xor128 xs;
    float F, acc, spent;
    long long sum_ai;
    int64_t bin0;
    int first, si;
    uint32_t step;
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
    for (int i = L.first; i < L.ph; ++i) {
        bool close;
        int fl;
        if (kPacked) {
            int pw = __ldg(a.pk + L.off + i);
            close = pw >= thr;
            fl = pw & 3;
        } else {
            int64_t ddx = __ldg(a.dd + L.off + i), adx = __ldg(a.ad + L.off + i);
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

// Same result in O(log ph): the closing photon by binary search over the running maximum of
// the photon words, the counts from the distance to the next acceptor / donor photon.
// Exact whenever no photon before `first` already reaches the new edge; otherwise (a window
// edge that moved backwards, unsorted ticks) the photon-by-photon scan decides.
__device__ __forceinline__ Window scan_indexed(const EvalArgs& a, const Lane& L, int64_t bin1, int thr)
{
    if (L.first >= L.ph) return Window{L.ph, 0, 0};
    const int32_t* __restrict__ pm = a.pm + L.off;
    if (L.first > 0 && __ldg(pm + L.first - 1) >= thr) return scan_linear<true>(a, L, bin1, thr);
    int lo = L.first, hi = L.ph - 1;  // the last photon closes any window that reaches it
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (__ldg(pm + mid) >= thr) hi = mid;
        else lo = mid + 1;
    }
    Window w{lo, 0, 0};
    const uint32_t nx = __ldg(a.nx + L.off + L.first);
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
__device__ __forceinline__ void lane_begin(Lane& L, const EvalArgs& a, int64_t t, int64_t bl,
                                           uint32_t rep, uint32_t epoch)
{
    const int64_t b = a.b0 + bl;
    const int4 r0 = __ldg(reinterpret_cast<const int4*>(a.bursts + b));
    const int4 r1 = __ldg(reinterpret_cast<const int4*>(a.bursts + b) + 1);
    L.t = t;
    L.g = (uint32_t)(b + a.g_offset);
    L.rep = rep;
    L.start = ((int64_t)(uint32_t)r0.y << 32) | (uint32_t)r0.x;
    L.stop = ((int64_t)(uint32_t)r0.w << 32) | (uint32_t)r0.z;
    L.T = __int_as_float(r1.x);
    L.ph = r1.y;
    L.off = ((int64_t)(uint32_t)r1.w << 32) | (uint32_t)r1.z;
    u32x4 sd = philox4x32_10(L.g, rep, 0u, kSlotSeed, a.seed, epoch);
    L.xs.seed((uint64_t)sd.x | ((uint64_t)sd.y << 32));
    L.F = 0.0f; L.acc = 0.0f; L.spent = 0.0f;
    L.sum_ai = 0;
    L.bin0 = L.start;
    L.first = 0;
    L.si = 0;
    L.step = 0;
}

// one dwell window (one trip of the reference's while loop, mc.cu:117-191)
template <int kLayout, bool kBg>
__device__ __forceinline__ void lane_step(Lane& L, const SmemParams& sp, const EvalArgs& a, uint32_t epoch)
{
    const int n = sp.n;
    const u32x4 w = philox4x32_10(L.g, L.rep, L.step, kSlotStep, a.seed, epoch);
    if (L.step == 0) L.si = pick(sp.cumP, n, u01(w.w));
    const int si = L.si;
    int sj = pick(sp.cumJ[si], n - 1, u01(w.x));
    if (sj >= si) ++sj;
    const float tau = det_logf(u01(w.y)) / (-sp.K[sj + si * n]);
    L.spent = L.spent + tau;
    int64_t bin1;
    if (L.spent >= L.T)
        bin1 = L.stop;
    else
        bin1 = (int64_t)((float)L.bin0 + L.spent / a.clk_p);

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
        win = (kLayout == 2) ? scan_indexed(a, L, bin1, thr) : scan_linear<true>(a, L, bin1, thr);
    }

    long long ai = 0;
    if (win.end < L.ph) {
        L.first = win.end;
        const float f_if = (kGamma - kDch2Ach) * (float)win.f_id + (1.0f - kDexDirAem) * (float)win.f_ia;
        float f_i;
        L.F += f_if;
        if (!kBg) {
            f_i = floorf(f_if + 0.5f);
        } else {
            const float t_diff = (float)(bin1 - L.bin0) * a.clk_p;
            const float rf_ia = (1.0f - kDexDirAem) * (float)win.f_ia;
            const u32x4 ba = philox4x32_10(L.g, L.rep, L.step, kSlotBgA, a.seed, epoch);
            uint32_t sv = poisson_itr(ba.x, ba.y, (double)(t_diff * a.bg_ad_rate));
            const float bg_a = ((float)(int32_t)sv <= rf_ia) ? (float)(int32_t)sv : floorf(rf_ia * u01(ba.z));
            const float rf_id = (kGamma - kDch2Ach) * (float)win.f_id;
            const u32x4 bq = philox4x32_10(L.g, L.rep, L.step, kSlotBgD, a.seed, epoch);
            sv = poisson_itr(bq.x, bq.y, (double)(t_diff * a.bg_dd_rate));
            const float bg_d = ((float)(int32_t)sv <= rf_id) ? (float)(int32_t)sv : floorf(rf_id * u01(bq.z));
            f_i = floorf(f_if - bg_d - bg_a + 0.5f);
        }
        const float rd = det_norm_icdf(w.z) * sp.v[si] + sp.rootE[si];
        const float de = 1.0f / (1.0f + det_pow6f(rd / kR0));
        ai = binomial(L.xs, (int)f_i, (double)de);
    }
    L.acc += (float)ai;
    L.sum_ai += ai;
    L.bin0 = bin1;
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

// chi2 for every p, copy the histogram out, re-zero the working set (whole block)
__device__ __forceinline__ void finalize_block(const EvalArgs& a)
{
    const int nbp = a.nb + 2;
    for (int p = threadIdx.x; p < a.P; p += blockDim.x)
        a.chi2_out[p] = chi2_of(a.whist + (size_t)p * nbp, a.ohist, a.nb, a.R, a.left_pad,
                                a.params[p].n);
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

// Persistent, warp-autonomous evaluation kernel.
//   kLayout 0: raw int64 burst_ad/burst_dd (fallback), 1: packed words scanned photon by
//   photon, 2: packed words + running-max / next-photon index (O(log ph) windows).
// Every warp owns a copy of the parameter vector and a private histogram in shared memory,
// draws chunks of kChunk consecutive samples from one global ticket, and only meets the other
// warps of its block once, at the very end (last-block finalisation).
#ifndef PDA_MIN_BLOCKS
#define PDA_MIN_BLOCKS 4
#endif
template <int kLayout, bool kBg>
__global__ void __launch_bounds__(kBlock, PDA_MIN_BLOCKS) pda_eval_kernel(const EvalArgs a)
{
    constexpr int kWarps = kBlock / 32;
    __shared__ SmemParams s_params[kWarps];
    extern __shared__ unsigned int s_hist_all[];  // [kWarps][nbp]
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int wid = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int nbp = a.nb + 2;
    SmemParams& sp = s_params[wid];
    unsigned int* hist = s_hist_all + wid * nbp;

    const int64_t per_p = a.N * (int64_t)a.R;  // samples per parameter vector
    const unsigned long long chunks_per_p = (unsigned long long)((per_p + kChunk - 1) / kChunk);
    const unsigned long long items = chunks_per_p * (unsigned long long)a.P;
    const bool small = per_p < (1ll << 31) && items < (1ull << 32);

    for (int i = lane; i < nbp; i += 32) hist[i] = 0u;
    __syncwarp();

    int cur_p = -1;           // vector whose parameters / histogram this warp holds
    uint32_t epoch = 0;
    int64_t c_next = 0;       // next unassigned sample of the current chunk (index inside the vector)
    int64_t c_end = 0;
    bool have_pending = false;  // a fetched chunk of another vector waits for the lanes to drain
    int pend_p = 0;
    int64_t pend_next = 0, pend_end = 0;
    bool exhausted = false;

    Lane L;
    bool active = false;

    for (;;) {
        // ---- hand samples to idle lanes -------------------------------------------------
        for (;;) {
            const unsigned need = __ballot_sync(0xffffffffu, !active);
            if (need == 0u) break;
            if (c_next >= c_end) {
                if (have_pending) {
                    if (need != 0xffffffffu) break;  // lanes of the previous vector still walking
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
                    stage_params(sp, a.params + pend_p, lane);
                    __syncwarp();
                    cur_p = pend_p;
                    epoch = a.epoch0 + (uint32_t)cur_p;
                    c_next = pend_next;
                    c_end = pend_end;
                    have_pending = false;
                    continue;
                }
                if (exhausted) break;
                unsigned long long item = 0;
                if (lane == 0) item = atomicAdd(a.counters + 0, 1ull);
                item = __shfl_sync(0xffffffffu, item, 0);
                if (item >= items) {
                    exhausted = true;
                    break;
                }
                const int p = small ? (int)((uint32_t)item / (uint32_t)chunks_per_p) : (int)(item / chunks_per_p);
                const unsigned long long c = item - (unsigned long long)p * chunks_per_p;
                const int64_t nx0 = (int64_t)c * kChunk;
                const int64_t nx1 = nx0 + kChunk < per_p ? nx0 + kChunk : per_p;
                if (p != cur_p) {
                    have_pending = true;
                    pend_p = p;
                    pend_next = nx0;
                    pend_end = nx1;
                } else {
                    c_next = nx0;
                    c_end = nx1;
                }
                continue;
            }
            const int avail = (int)(c_end - c_next);
            const int rank = __popc(need & lt_mask);
            const bool take = !active && rank < avail;
            const int want = __popc(need);
            if (take) {
                const int64_t t = c_next + rank;
                const int64_t bl = small ? (int64_t)((uint32_t)t / (uint32_t)a.R) : t / a.R;
                const uint32_t rep = (uint32_t)(t - bl * a.R);
                lane_begin(L, a, t, bl, rep, epoch);
                active = true;
            }
            c_next += want < avail ? want : avail;
            // a burst of zero duration never enters the reference's while loop: E = 0
            const bool dead = take && !(L.T > 0.0f);
            if (dead) {
                const size_t oi = (size_t)cur_p * (size_t)per_p + (size_t)L.t;
                if (a.mcE) a.mcE[oi] = 0.0f;
                if (a.sum_ai) a.sum_ai[oi] = 0;
                if (a.n_steps) a.n_steps[oi] = 0;
                atomicAdd(&hist[hist_index(0.0f, a.nb)], 1u);
                active = false;
            }
        }
        if (__ballot_sync(0xffffffffu, active) == 0u) {
            if (exhausted && !have_pending) break;
            continue;
        }

        // ---- one dwell window for every active lane -------------------------------------
        if (active) {
            lane_step<kLayout, kBg>(L, sp, a, epoch);
            bool fin = !(L.T > L.spent);
            if (!fin && L.step >= kStepCap) {
                atomicOr(a.counters + 2, 1ull);
                fin = true;
            }
            if (fin) {
                const float e = (L.F > 0.0f) ? L.acc / L.F : 0.0f;
                const size_t oi = (size_t)cur_p * (size_t)per_p + (size_t)L.t;
                if (a.mcE) a.mcE[oi] = e;
                if (a.sum_ai) a.sum_ai[oi] = L.sum_ai;
                if (a.n_steps) a.n_steps[oi] = (int32_t)L.step;
                // warp-aggregated add: one shared-memory atomic per distinct bin among the
                // lanes that finish in this trip
                const int bin = hist_index(e, a.nb);
                const unsigned fm = __activemask();
                const unsigned peers = __match_any_sync(fm, bin);
                if (lane == __ffs(peers) - 1) atomicAdd(&hist[bin], (unsigned)__popc(peers));
                active = false;
            }
        }
        __syncwarp();
    }

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
        if (a.finalize) {
            finalize_block(a);
        } else if (threadIdx.x == 0) {
            a.counters[0] = 0ull;
            a.counters[1] = 0ull;
        }
    }
}

__global__ void __launch_bounds__(kBlock) pda_finalize_kernel(const EvalArgs a) { finalize_block(a); }

// ------------------------------------------------------------------------------------
// one-time data preparation: burst records + packed photon words (role of the uploads in
// mc::init_data_gpu, /root/reference/src/mc.cu:456-521)
// ------------------------------------------------------------------------------------
__global__ void pda_pack_kernel(const PackArgs a)
{
    // one warp per burst, lanes stride over its photons
    const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (b >= a.B) return;
    const int64_t start = a.start[b];
    const int ph = a.ph[b];
    const int64_t is = a.istart[b];
    if (lane == 0) {
        BurstRec r;
        r.start = start;
        r.stop = a.stop[b];
        r.T = a.T[b];
        r.ph = ph;
        r.off = a.pk ? a.poff[b] : is;
        a.bursts[b] = r;
    }
    if (!a.pk) return;
    const int64_t po = a.poff[b];
    const int padded = (ph + 3) & ~3;
    unsigned int bad = 0;
    for (int i = lane; i < padded; i += 32) {
        int word = INT32_MAX;
        if (i < ph) {
            const int64_t adx = a.ad[is + i], ddx = a.dd[is + i];
            const int64_t m = adx > ddx ? adx : ddx;
            const int fl = (ddx > 0 ? 1 : 0) | (adx > 0 ? 2 : 0);
            int64_t rel = m - start;
            if (rel < -(1ll << 29) + 1 || rel > (1ll << 29) - 2) {
                if (m == 0) rel = -(1ll << 29);  // empty channels of a burst far from tick 0
                else { rel = 0; bad++; }
            }
            word = (int)(rel * 4) | fl;
        }
        a.pk[po + i] = word;
    }
    if (bad) atomicAdd(a.bad, bad);
}

// per-burst index over the packed words: running maximum (binary-searchable window edge) and
// distance to the next acceptor / donor photon (window counts in O(1)); one thread per burst
__global__ void pda_index_kernel(const PackArgs a)
{
    const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= a.B) return;
    const int ph = a.ph[b];
    const int64_t po = a.poff[b];
    const int padded = (ph + 3) & ~3;
    int m = INT32_MIN;
    for (int i = 0; i < ph; ++i) {
        const int key = a.pk[po + i] & ~3;
        m = key > m ? key : m;
        a.pm[po + i] = m;
    }
    for (int i = ph; i < padded; ++i) { a.pm[po + i] = INT32_MAX; a.nx[po + i] = 0u; }
    int na = ph, nd = ph;
    for (int i = ph - 1; i >= 0; --i) {
        const int f = a.pk[po + i] & 3;
        if (f & 2) na = i;
        if (f & 1) nd = i;
        a.nx[po + i] = (uint32_t)(na - i) | ((uint32_t)(nd - i) << 16);
    }
}

// boost regular-axis histogram of a float array (observed SgDivSr histogram,
// role of streamWorker.cpp:62-70)
__global__ void pda_hist_f32_kernel(const float* __restrict__ x, int64_t n, int nb,
                                    unsigned long long* __restrict__ hist)
{
    extern __shared__ unsigned int sh[];
    const int nbp = nb + 2;
    for (int i = threadIdx.x; i < nbp; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&sh[hist_index(x[i], nb)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < nbp; i += blockDim.x)
        if (sh[i]) atomicAdd(hist + i, (unsigned long long)sh[i]);
}

__global__ void pda_u64_to_f32_kernel(const unsigned long long* in, float* out, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __ull2float_rn(in[i]);
}

// ------------------------------------------------------------------------------------
// primitive probes (tests only call these through the C ABI)
// ------------------------------------------------------------------------------------
__global__ void pda_probe_kernel(const ProbeArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    switch (a.what) {
    case 0: {  // philox: ctr = (w0[4i..4i+3]), key = (w1[2i], w1[2i+1]) -> ou[4i..]
        u32x4 r = philox4x32_10(a.w0[4 * i], a.w0[4 * i + 1], a.w0[4 * i + 2], a.w0[4 * i + 3],
                                a.w1[2 * i], a.w1[2 * i + 1]);
        a.ou[4 * i] = r.x; a.ou[4 * i + 1] = r.y; a.ou[4 * i + 2] = r.z; a.ou[4 * i + 3] = r.w;
        break;
    }
    case 1: a.of[i] = det_logf(a.f0[i]); break;
    case 2: a.od[i] = det_log(a.d0[i]); break;
    case 3: a.od[i] = det_exp(a.d0[i]); break;
    case 4: a.of[i] = det_norm_icdf(a.w0[i]); break;
    case 5: a.of[i] = det_root6f(a.f0[i]); break;
    case 6: a.of[i] = u01(a.w0[i]); break;
    case 7: {  // binomial: seed words (w0[2i], w0[2i+1]) n = i0[i], p = d0[i]; ni draws chained
        Xor128 g;
        g.seed((uint64_t)a.w0[2 * i] | ((uint64_t)a.w0[2 * i + 1] << 32));
        long long s = 0;
        for (int k = 0; k < a.ni; ++k) s = s * 31 + binomial(g, a.i0[i], a.d0[i]);
        a.ol[i] = s;
        break;
    }
    case 8: a.ou[i] = poisson_itr(a.w0[2 * i], a.w0[2 * i + 1], a.d0[i]); break;
    case 9: {  // drawE(e = f0[i], v = f1[i], word w0[i])
        float rd = det_norm_icdf(a.w0[i]) * a.f1[i] + kR0 * det_root6f(1.0f / a.f0[i] - 1.0f);
        a.of[i] = 1.0f / (1.0f + det_pow6f(rd / kR0));
        break;
    }
    case 10: a.of[i] = det_logf(u01(a.w0[i])) / (-a.f0[i]); break;  // drawTau(k = f0[i])
    case 11: {  // multinomial: K = ni classes p = f0[0..ni), N = nj trials -> oi[i*ni..]
        Xor128 g;
        g.seed((uint64_t)a.w0[2 * i] | ((uint64_t)a.w0[2 * i + 1] << 32));
        int out[16];
        int_multinomial(g, a.ni, a.nj, a.f0, out);
        for (int k = 0; k < a.ni; ++k) a.oi[i * a.ni + k] = out[k];
        break;
    }
    case 12: {  // categorical over f0[0..ni) with word w0[i] (drawDisIdx)
        float cum[16];
        float s = 0.0f;
        for (int k = 0; k < a.ni; ++k) { s += a.f0[k]; cum[k] = s; }
        a.oi[i] = pick(cum, a.ni, u01(a.w0[i]));
        break;
    }
    case 13: a.oi[i] = hist_index(a.f0[i], a.ni); break;
    default: break;
    }
}

// ------------------------------------------------------------------------------------
// host-side launchers
// ------------------------------------------------------------------------------------
template <int kLayout, bool kBg>
static int resident_blocks(size_t dyn)
{
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, pda_eval_kernel<kLayout, kBg>, kBlock, dyn);
    return nb < 1 ? 1 : nb;
}

static size_t eval_dyn_smem(int nb) { return (size_t)(kBlock / 32) * (size_t)(nb + 2) * sizeof(unsigned int); }

LaunchCfg eval_launch_cfg(int device, int layout, bool with_bg, int nb)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t dyn = eval_dyn_smem(nb);
    int per;
    if (layout == 0) per = with_bg ? resident_blocks<0, true>(dyn) : resident_blocks<0, false>(dyn);
    else if (layout == 1) per = with_bg ? resident_blocks<1, true>(dyn) : resident_blocks<1, false>(dyn);
    else per = with_bg ? resident_blocks<2, true>(dyn) : resident_blocks<2, false>(dyn);
    return LaunchCfg{sms * per, kBlock};
}

cudaError_t launch_eval(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s)
{
    if (a.nb + 2 > kMaxBins) return cudaErrorInvalidValue;
    const size_t dyn = eval_dyn_smem(a.nb);
#define PDA_LAUNCH(LAY, BG) pda_eval_kernel<LAY, BG><<<cfg.grid, cfg.block, dyn, s>>>(a)
    if (layout == 0) { if (with_bg) PDA_LAUNCH(0, true); else PDA_LAUNCH(0, false); }
    else if (layout == 1) { if (with_bg) PDA_LAUNCH(1, true); else PDA_LAUNCH(1, false); }
    else { if (with_bg) PDA_LAUNCH(2, true); else PDA_LAUNCH(2, false); }
#undef PDA_LAUNCH
    return cudaGetLastError();
}

cudaError_t launch_finalize(const EvalArgs& a, cudaStream_t s)
{
    pda_finalize_kernel<<<1, kBlock, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_pack(const PackArgs& a, cudaStream_t s)
{
    if (a.B == 0) return cudaSuccess;
    const int64_t threads = a.B * 32;
    const int64_t blocks = (threads + 255) / 256;
    pda_pack_kernel<<<(unsigned)blocks, 256, 0, s>>>(a);
    if (a.pk && a.pm && a.nx) pda_index_kernel<<<(unsigned)((a.B + 127) / 128), 128, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_hist_f32(const float* x, int64_t n, int nb, unsigned long long* hist, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pda_hist_f32_kernel<<<(unsigned)blocks, 256, (nb + 2) * sizeof(unsigned int), s>>>(x, n, nb, hist);
    return cudaGetLastError();
}

cudaError_t launch_u64_to_f32(const unsigned long long* in, float* out, int n, cudaStream_t s)
{
    pda_u64_to_f32_kernel<<<(n + 255) / 256, 256, 0, s>>>(in, out, n);
    return cudaGetLastError();
}

cudaError_t launch_probe(const ProbeArgs& a, cudaStream_t s)
{
    if (a.n == 0) return cudaSuccess;
    pda_probe_kernel<<<(unsigned)((a.n + 127) / 128), 128, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace pda
