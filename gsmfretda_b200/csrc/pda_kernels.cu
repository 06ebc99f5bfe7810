// pda_kernels.cu — the Monte-Carlo PDA evaluation for sm_100a (compile with -fmad=false).
//
// One thread walks one burst-sample (burst b, repeat r, parameter vector p): the role of
// the reference's mc_kernel (/root/reference/src/mc.cu:83-197), fused with the E histogram
// (streamWorker.cpp:187-193) and the reduced chi-square (streamWorker.cpp:197-206) that
// the reference computes on the host after copying N*R floats back.
//
// Design:
//  * persistent grid, one ticket counter hands out tiles of blockDim consecutive samples,
//    so that lanes of a warp are repeats of the same / neighbouring bursts and read the
//    same photon words (one L1 transaction per warp-wide load);
//  * photons are one packed 32-bit word each, (tick - burst start) << 2 | channel flags,
//    instead of the reference's two int64 (16 B -> 4 B per photon);
//  * RNG is Philox4x32-10 keyed by (global burst, repeat, step, slot): no state in memory;
//  * per-block shared-memory histogram with warp-aggregated atomics, one 64-bit global
//    atomic per non-empty bin per block, and the last block to finish reduces chi-square,
//    copies the histogram out and re-zeroes the working set: one launch per evaluation.
#include "pda_kernels.h"
#include "pda_device.cuh"

namespace pda {

constexpr int kBlock = 256;
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

__device__ __forceinline__ void stage_params(SmemParams& sp, const ParamRaw* __restrict__ pr)
{
    const int tid = threadIdx.x;
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
    if (tid < n * n) sp.K[tid] = pr->K[tid];
    if (tid == 32) {
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

struct SampleOut {
    float mcE;
    long long sum_ai;
    uint32_t steps;
};

template <bool kPacked, bool kBg>
__device__ __forceinline__ SampleOut walk_sample(const SmemParams& sp, const EvalArgs& a,
                                                 int64_t b, uint32_t g, uint32_t rep,
                                                 uint32_t epoch)
{
    const int n = sp.n;
    const int4 r0 = __ldg(reinterpret_cast<const int4*>(a.bursts + b));
    const int4 r1 = __ldg(reinterpret_cast<const int4*>(a.bursts + b) + 1);
    const int64_t start = ((int64_t)(uint32_t)r0.y << 32) | (uint32_t)r0.x;
    const int64_t stop = ((int64_t)(uint32_t)r0.w << 32) | (uint32_t)r0.z;
    const float T = __int_as_float(r1.x);
    const int ph = r1.y;
    const int64_t off = ((int64_t)(uint32_t)r1.w << 32) | (uint32_t)r1.z;
    const int32_t* __restrict__ pk = kPacked ? a.pk + off : nullptr;
    const int64_t* __restrict__ ad = kPacked ? nullptr : a.ad + off;
    const int64_t* __restrict__ dd = kPacked ? nullptr : a.dd + off;
    const float clk_p = a.clk_p;

    Xor128 xs;
    {
        u32x4 sd = philox4x32_10(g, rep, 0u, kSlotSeed, a.seed, epoch);
        xs.seed((uint64_t)sd.x | ((uint64_t)sd.y << 32));
    }
    u32x4 w = philox4x32_10(g, rep, 0u, kSlotStep, a.seed, epoch);

    float F = 0.0f, acc = 0.0f, spent = 0.0f;
    long long sum_ai = 0;
    int si = pick(sp.cumP, n, u01(w.w));
    int64_t bin0 = start, bin1 = start;
    int first = 0;
    uint32_t step = 0;
    while (T > spent) {
        if (step) w = philox4x32_10(g, rep, step, kSlotStep, a.seed, epoch);
        int sj = pick(sp.cumJ[si], n - 1, u01(w.x));
        if (sj >= si) ++sj;
        float tau = det_logf(u01(w.y)) / (-sp.K[sj + si * n]);
        spent = spent + tau;
        if (spent >= T)
            bin1 = stop;
        else
            bin1 = (int64_t)((float)bin0 + spent / clk_p);

        int flags = 0, f_id = 0, f_ia = 0;
        long long ai = 0;
        int thr = 0;
        if (kPacked) {
            // photon word w = rel*4 + flags compares against rel1*4; a window edge at or
            // below tick 0 closes on any photon (0 >= bin1 holds for the empty channels too)
            int64_t rel = bin1 - start;
            rel = rel < -(1ll << 29) + 1 ? -(1ll << 29) + 1 : rel;
            rel = rel > (1ll << 29) - 1 ? (1ll << 29) - 1 : rel;
            thr = bin1 <= 0 ? INT32_MIN : (int)rel * 4;
        }
        for (int i = first; i < ph; ++i) {
            bool close;
            int fl;
            if (kPacked) {
                int pw = __ldg(pk + i);
                close = pw >= thr;
                fl = pw & 3;
            } else {
                int64_t ddx = __ldg(dd + i), adx = __ldg(ad + i);
                close = ddx >= bin1 || adx >= bin1;
                fl = (ddx > 0 ? 1 : 0) | (adx > 0 ? 2 : 0);
            }
            if (close || i == ph - 1) {
                first = i;
                float f_if = (kGamma - kDch2Ach) * (float)f_id + (1.0f - kDexDirAem) * (float)f_ia;
                float f_i;
                F += f_if;
                if (!kBg) {
                    f_i = floorf(f_if + 0.5f);
                } else {
                    float t_diff = (float)(bin1 - bin0) * clk_p;
                    float rf_ia = (1.0f - kDexDirAem) * (float)f_ia;
                    u32x4 ba = philox4x32_10(g, rep, step, kSlotBgA, a.seed, epoch);
                    uint32_t sv = poisson_itr(ba.x, ba.y, (double)(t_diff * a.bg_ad_rate));
                    float bg_a = ((float)(int32_t)sv <= rf_ia) ? (float)(int32_t)sv
                                                               : floorf(rf_ia * u01(ba.z));
                    float rf_id = (kGamma - kDch2Ach) * (float)f_id;
                    u32x4 bq = philox4x32_10(g, rep, step, kSlotBgD, a.seed, epoch);
                    sv = poisson_itr(bq.x, bq.y, (double)(t_diff * a.bg_dd_rate));
                    float bg_d = ((float)(int32_t)sv <= rf_id) ? (float)(int32_t)sv
                                                               : floorf(rf_id * u01(bq.z));
                    f_i = floorf(f_if - bg_d - bg_a + 0.5f);
                }
                float rd = det_norm_icdf(w.z) * sp.v[si] + sp.rootE[si];
                float de = 1.0f / (1.0f + det_pow6f(rd / kR0));
                ai = binomial(xs, (int)f_i, (double)de);
                break;
            }
            flags |= fl;
            if (flags == 3) continue;
            f_ia += flags >> 1;
            f_id += flags & 1;
        }
        acc += (float)ai;
        sum_ai += ai;
        bin0 = bin1;
        si = sj;
        if (++step >= kStepCap) {
            atomicOr(a.counters + 2, 1ull);
            break;
        }
    }
    SampleOut o;
    o.mcE = (F > 0.0f) ? acc / F : 0.0f;
    o.sum_ai = sum_ai;
    o.steps = step;
    return o;
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

template <bool kPacked, bool kBg>
__global__ void __launch_bounds__(kBlock) pda_eval_kernel(const EvalArgs a)
{
    __shared__ SmemParams sp;
    __shared__ unsigned int shist[kMaxBins];
    __shared__ unsigned long long s_item;
    __shared__ int s_last;

    const int tid = threadIdx.x;
    const int nbp = a.nb + 2;
    const int64_t per_p = a.N * (int64_t)a.R;                      // samples per parameter vector
    const unsigned long long tiles_per_p = (unsigned long long)((per_p + kBlock - 1) / kBlock);
    const unsigned long long items = tiles_per_p * (unsigned long long)a.P;
    const bool small = per_p < (1ll << 31) && items < (1ull << 32);

    for (int i = tid; i < nbp; i += kBlock) shist[i] = 0u;
    int cur_p = -1;

    for (;;) {
        __syncthreads();  // s_item consumed by everyone, previous tile's histogram adds done
        if (tid == 0) s_item = atomicAdd(a.counters + 0, 1ull);
        __syncthreads();
        const unsigned long long item = s_item;
        if (item >= items) break;
        const int p = small ? (int)((uint32_t)item / (uint32_t)tiles_per_p) : (int)(item / tiles_per_p);
        const unsigned long long tile = item - (unsigned long long)p * tiles_per_p;
        if (p != cur_p) {
            if (cur_p >= 0) {  // flush the finished vector's bins
                unsigned long long* gh = a.whist + (size_t)cur_p * nbp;
                for (int i = tid; i < nbp; i += kBlock) {
                    unsigned int c = shist[i];
                    if (c) atomicAdd(gh + i, (unsigned long long)c);
                    shist[i] = 0u;
                }
            }
            stage_params(sp, a.params + p);
            cur_p = p;
            __syncthreads();
        }
        const int64_t t = (int64_t)tile * kBlock + tid;
        int bin = -1;
        if (t < per_p) {
            const int64_t bl = small ? (int64_t)((uint32_t)t / (uint32_t)a.R) : t / a.R;
            const uint32_t rep = (uint32_t)(t - bl * a.R);
            const int64_t b = a.b0 + bl;
            SampleOut o = walk_sample<kPacked, kBg>(sp, a, b, (uint32_t)(b + a.g_offset), rep,
                                                    a.epoch0 + (uint32_t)p);
            const size_t oi = (size_t)p * (size_t)per_p + (size_t)t;
            if (a.mcE) a.mcE[oi] = o.mcE;
            if (a.sum_ai) a.sum_ai[oi] = o.sum_ai;
            if (a.n_steps) a.n_steps[oi] = (int32_t)o.steps;
            bin = hist_index(o.mcE, a.nb);
        }
        __syncwarp();
        const unsigned peers = __match_any_sync(0xffffffffu, bin);
        if (bin >= 0 && (tid & 31) == __ffs(peers) - 1) atomicAdd(&shist[bin], (unsigned)__popc(peers));
    }

    // queue drained: flush, then the last block to arrive finalises
    if (cur_p >= 0) {
        unsigned long long* gh = a.whist + (size_t)cur_p * nbp;
        for (int i = tid; i < nbp; i += kBlock) {
            unsigned int c = shist[i];
            if (c) atomicAdd(gh + i, (unsigned long long)c);
        }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        unsigned long long done = atomicAdd(a.counters + 1, 1ull);
        s_last = (done == (unsigned long long)gridDim.x - 1ull);
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        if (a.finalize) {
            finalize_block(a);
        } else if (tid == 0) {
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
template <bool kPacked, bool kBg>
static int resident_blocks()
{
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, pda_eval_kernel<kPacked, kBg>, kBlock, 0);
    return nb < 1 ? 1 : nb;
}

LaunchCfg eval_launch_cfg(int device, bool packed, bool with_bg)
{
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    int per = packed ? (with_bg ? resident_blocks<true, true>() : resident_blocks<true, false>())
                     : (with_bg ? resident_blocks<false, true>() : resident_blocks<false, false>());
    return LaunchCfg{sms * per, kBlock};
}

cudaError_t launch_eval(const EvalArgs& a, bool packed, bool with_bg, LaunchCfg cfg, cudaStream_t s)
{
    if (a.nb + 2 > kMaxBins) return cudaErrorInvalidValue;
    if (packed) {
        if (with_bg) pda_eval_kernel<true, true><<<cfg.grid, cfg.block, 0, s>>>(a);
        else pda_eval_kernel<true, false><<<cfg.grid, cfg.block, 0, s>>>(a);
    } else {
        if (with_bg) pda_eval_kernel<false, true><<<cfg.grid, cfg.block, 0, s>>>(a);
        else pda_eval_kernel<false, false><<<cfg.grid, cfg.block, 0, s>>>(a);
    }
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
