// mc.hpp — the reference's `class mc` (/root/reference/src/mc.hpp:78-179) re-created on top of
// the C ABI of include/gsmfret_c.h.  Same public members, argument meaning and error
// behaviour, so that code written against the reference — gpuWorker::run
// (/root/reference/src/gpuWorker.cpp:38-87, compiled UNCHANGED against this header by
// `make -C host dropin`), the result-reading part of streamWorker::run
// (/root/reference/src/streamWorker.cpp:186-206), main (/root/reference/src/main.cpp:129-160)
// — compiles against it.  Header-only; link with libgsmfret.so.
//
// Differences, all additive:
//   * hHist(sid, p) / hChi2(sid, p): the histogram and reduced chi-square the fused kernel already
//     produced, so a caller no longer needs hmcE and the host histogram loop;
//   * mc::Options chooses whether hmcE is kept (default: yes, for drop-in compatibility);
//   * set_reSampleTimes keeps the reference's rounding down to a power of two; use
//     set_reSampleTimes_exact for any R (the new kernel has no such restriction);
//   * set_params_batch: P vectors for one launch (gsm_run_batch);
//   * ONE `mc` MAY SPAN SEVERAL GPUS (constructor taking a list of device ids): init_data_gpu cuts
//     the bursts into one contiguous range per GPU, balanced by photon count; every call fans out
//     to one library context per GPU; the histograms meet inside the evaluation kernels (fused
//     peer exchange, gsm_peer_*), so hHist / hChi2 are the results over the WHOLE dataset and
//     gpuWorker / streamWorker drive it exactly like a one-GPU `mc`.  This replaces the
//     reference's one-process-per-GPU launcher (/root/reference/src/call_main.cpp:195-227) for the
//     case of one huge dataset (SURVEY.md §8e "burst split", §8 f3).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#include "../../include/gsmfret_c.h"

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

typedef float retype;  // /root/reference/src/loadHdf5.hpp:8

// The reference's convention (/root/reference/src/cuda_tools.hpp:20-28): print
// "<stmt> returned <msg> at file:line" and leave the process.
#define GSM_CHECK_RETURN(stmt)                                                              \
    do {                                                                                    \
        int gsm_rc_ = (stmt);                                                               \
        if (gsm_rc_ < 0) {                                                                  \
            std::fprintf(stderr, "%s returned %s(%d) at %s:%d\n", #stmt, gsm_last_error(), \
                         gsm_rc_, __FILE__, __LINE__);                                      \
            std::exit(EXIT_FAILURE);                                                        \
        }                                                                                   \
    } while (0)

class mc {
  public:
    struct Options {
        bool keep_mce = true;      // fill hmcE[sid] like the reference (costs an N*R-float copy per run)
        bool force_raw64 = false;  // keep the reference's 16 B/photon layout
        bool timing = false;       // bracket every run with CUDA events (gsm_last_elapsed_ms); the reference has none
        bool peer_shared_device = false;  // several parts of one dataset on one GPU (tests on a one-GPU box)
        uint32_t seed = 0x9E3779B9u;
        int fret_hist_num = 200;   // cmdline.ggo:14
        int left_pad = 2;          // main.cpp:155
        int max_batch = 1;         // most vectors per launch when the `mc` spans GPUs (sizes the exchange)
        int carriers = 0;          // extra stream slots that carry coalesced launches (run_coalesced); needs !keep_mce
    };

    int devid;
    char gpuuuid[33];
    std::atomic_int workerNum;
    retype** hmcE;
    int* s_n;
    int sz_burst;
    int reSampleTimes;

    mc(int id, int _streamNum = 16, unsigned char debug = 0, std::uintmax_t hdf5size = 0,
       bool profiler = false)
        : mc(std::vector<int>(1, id), _streamNum, debug, hdf5size, profiler, Options()) {}

    mc(int id, int _streamNum, unsigned char debug, std::uintmax_t hdf5size, bool profiler, const Options& opt)
        : mc(std::vector<int>(1, id), _streamNum, debug, hdf5size, profiler, opt) {}

    // one context per entry of `ids` (a GPU may be listed more than once only with
    // Options::peer_shared_device); devid / gpuuuid describe the first one
    mc(const std::vector<int>& ids, int _streamNum, unsigned char /*debug*/, std::uintmax_t /*hdf5size*/,
       bool profiler, const Options& opt)
        : devid(ids.empty() ? 0 : ids[0]), hmcE(nullptr), s_n(nullptr), sz_burst(0), reSampleTimes(4), ctx(nullptr),
          streamNum(0), streams(nullptr), opt_(opt), profiler_(profiler)
    {
        gpuuuid[0] = 0;
        const unsigned flags = (opt.keep_mce ? GSM_KEEP_MCE : 0u) | (opt.force_raw64 ? GSM_FORCE_RAW64 : 0u) |
                               (opt.timing ? 0u : GSM_NO_TIMING) | (opt.peer_shared_device ? GSM_PEER_SHARED_DEVICE : 0u);
        for (size_t r = 0; r < ids.size(); ++r) {
            gsm_ctx* c = nullptr;
            int rc = gsm_create(ids[r], _streamNum + (opt.keep_mce ? 0 : opt.carriers), flags, &c);
            if (rc == GSM_EINVAL && r == 0) {  // "gpu id set error!" (mc.cu:228-231): the reference returns quietly
                std::printf("gpu id set error!\n");
                workerNum.store(_streamNum);
                return;
            }
            GSM_CHECK_RETURN(rc);
            part_.push_back(c);
            GSM_CHECK_RETURN(gsm_set_seed(c, opt.seed));
            GSM_CHECK_RETURN(gsm_set_hist(c, opt.fret_hist_num, opt.left_pad));
            GSM_CHECK_RETURN(gsm_set_resample_times(c, reSampleTimes));
        }
        ctx = part_[0];
        if (profiler_) GSM_CHECK_RETURN(gsm_profiler(1));   // cudaProfilerStart, as mc::mc does (mc.cu:207-210)
        carriers_ = opt.keep_mce ? 0 : opt.carriers;
        streamNum = gsm_stream_count(ctx) - carriers_;   // the carrier slots are not streams of the node
        workerNum.store(streamNum);
        // the reference keeps the 16 raw bytes of the device uuid in gpuuuid (mc.cu:43-48); the C
        // ABI hands out 32 hex digits
        char hex[33] = {0};
        gsm_device_uuid(ctx, hex);
        for (int i = 0; i < 33; ++i) gpuuuid[i] = 0;
        for (int i = 0; i < 16 && hex[2 * i] && hex[2 * i + 1]; ++i) {
            auto nib = [](char c) { return c <= '9' ? c - '0' : (c | 0x20) - 'a' + 10; };
            gpuuuid[i] = (char)((nib(hex[2 * i]) << 4) | nib(hex[2 * i + 1]));
        }
        streams = new cudaStream_t[streamNum];
        bad_.assign((size_t)streamNum, 0);
        batch_.assign((size_t)streamNum, 0);
        batch_args_.resize((size_t)streamNum);
        hmcE = new retype*[streamNum]();
        s_n = new int[streamNum]();
        for (int i = 0; i < streamNum; ++i) streams[i] = (cudaStream_t)gsm_stream(ctx, i);
    }

    ~mc()
    {
        // peers have each other's exchange blocks mapped: release all mappings before any block goes
        if (part_.size() > 1)
            for (gsm_ctx* c : part_) gsm_peer_disconnect(c);
        if (profiler_ && !part_.empty()) gsm_profiler(0);   // cudaProfilerStop (mc.cu:655-658)
        for (gsm_ctx* c : part_) gsm_destroy(c);
        delete[] streams;
        delete[] hmcE;
        delete[] s_n;
    }
    mc(const mc&) = delete;
    mc& operator=(const mc&) = delete;

    void set_gpuid() { GSM_CHECK_RETURN(gsm_set_device(ctx)); }

    cudaStream_t* getStreams(int* numStream = nullptr)
    {
        if (numStream) *numStream = streamNum;
        return streams;
    }

    bool streamQuery(int sid)
    {
        if (sid < 0 || sid >= streamNum) return false;
        for (gsm_ctx* c : part_) {
            int r = gsm_query(c, sid);
            GSM_CHECK_RETURN(r);
            if (r != 1) return false;
        }
        return true;
    }

    int set_nstates(int n, int sid)
    {
        int prev = 0;
        for (gsm_ctx* c : part_) {
            prev = gsm_set_nstates(c, n, sid);
            GSM_CHECK_RETURN(prev);
        }
        s_n[sid] = n;
        return prev;  // the previous n (== n when unchanged), as mc::set_nstates
    }

    // false when args has the wrong count for n states (eigenhelper.cpp:5-8); the slot is then
    // marked failed: run_kernel launches nothing and failed(sid) tells the reader of the results
    bool set_params(int n, int sid, std::vector<float>& args)
    {
        bool ok = true;
        for (gsm_ctx* c : part_) {
            int r = gsm_set_params(c, n, sid, args.data(), (int)args.size());
            GSM_CHECK_RETURN(r);
            ok = ok && r == 1;
        }
        bad_[sid] = !ok;
        batch_[sid] = 0;
        return ok;
    }

    // N of the whole dataset; every part then evaluates what falls into its own range (the
    // reference starts at burst 0 whatever cstart says, quirk Q7 — so does every part's share)
    int setBurstBd(int cstart, int cstop, int sid)
    {
        if (part_.size() == 1) {
            int r = gsm_set_burst_bd(ctx, cstart, cstop, sid);
            if (r < -1) GSM_CHECK_RETURN(r);
            return r;
        }
        int64_t rstart = cstart, rstop = cstop;
        if (cstop >= sz_burst || cstop < 0) rstop = sz_burst;
        if (cstart >= rstop) rstart = 0;
        return cstart == -1 ? -1 : (int)(rstop - rstart);
    }

    // reSampleTimes = 2^floor(log2 t), /root/reference/src/mc.cu:316
    void set_reSampleTimes(int t)
    {
        int r = 1;
        while (r * 2 <= t) r *= 2;
        set_reSampleTimes_exact(r);
    }
    void set_reSampleTimes_exact(int t)
    {
        for (gsm_ctx* c : part_) GSM_CHECK_RETURN(gsm_set_resample_times(c, t));
        reSampleTimes = t;
    }

    void init_randstate(int N, int sid)
    {
        for (gsm_ctx* c : part_) GSM_CHECK_RETURN(gsm_init_randstate(c, N, sid));
    }
    void set_params_buff(int oldS_n, int N_sid, int sid)
    {
        for (gsm_ctx* c : part_) GSM_CHECK_RETURN(gsm_set_params_buff(c, oldS_n, N_sid, sid));
    }

    void run_kernel(int N, int sid)
    {
        if (bad_[sid] || N < 0) return;   // refused vector / nothing asked for: the stream stays idle
        for (size_t r = 0; r < part_.size(); ++r) {
            int Nr = N;
            if (part_.size() > 1) {       // the first N bursts of the dataset, as seen from part r
                int64_t in_part = (int64_t)N - cuts_[r];
                const int64_t size = cuts_[r + 1] - cuts_[r];
                Nr = (int)(in_part < 0 ? 0 : (in_part > size ? size : in_part));
            }
            if (batch_[sid] > 0)
                GSM_CHECK_RETURN(gsm_run_batch(part_[r], Nr, sid, s_n[sid], batch_[sid], batch_args_[sid].data()));
            else
                GSM_CHECK_RETURN(gsm_run(part_[r], Nr, sid));
        }
        hmcE[sid] = const_cast<retype*>(gsm_result_mcE(ctx, sid));  // pinned; valid once streamQuery(sid)
    }

    void get_res(int /*sid*/, int /*N*/) {}  // debug print in the reference (mc.cu:602-617)
    void free_data_gpu() {}                  // all device memory is released by ~mc

    void init_data_gpu(std::vector<int64_t>& istart, std::vector<int64_t>& start, std::vector<int64_t>& stop,
                       std::vector<int>& phCount, long _sz_tag, int64_t* burst_ad, int64_t* burst_dd,
                       std::vector<float>& T_burst_duration, std::vector<float>& SgDivSr, float& iclk_p,
                       float& ibg_ad_rate, float& ibg_dd_rate)
    {
        const int64_t B = (int64_t)start.size();
        sz_burst = (int)B;
        const int G = (int)part_.size();
        if (G == 1) {
            GSM_CHECK_RETURN(gsm_init_data(ctx, B, istart.data(), start.data(), stop.data(), phCount.data(),
                                           (int64_t)_sz_tag, burst_ad, burst_dd, T_burst_duration.data(),
                                           SgDivSr.data(), iclk_p, ibg_ad_rate, ibg_dd_rate, 0));
            return;
        }
        cuts_ = plan_split(phCount, G);
        const int nb = opt_.fret_hist_num;
        std::vector<float> observed((size_t)nb, 0.0f), mine((size_t)nb);
        for (int r = 0; r < G; ++r) {
            const int64_t b0 = cuts_[(size_t)r], b1 = cuts_[(size_t)r + 1];
            const int64_t p0 = b0 < B ? istart[(size_t)b0] : (int64_t)_sz_tag, p1 = b1 < B ? istart[(size_t)b1] : (int64_t)_sz_tag;
            std::vector<int64_t> rebased(istart.begin() + b0, istart.begin() + b1);
            for (int64_t& v : rebased) v -= p0;
            GSM_CHECK_RETURN(gsm_init_data(part_[(size_t)r], b1 - b0, rebased.data(), start.data() + b0, stop.data() + b0,
                                           phCount.data() + b0, p1 - p0, burst_ad + p0, burst_dd + p0,
                                           T_burst_duration.data() + b0, SgDivSr.data() + b0, iclk_p, ibg_ad_rate,
                                           ibg_dd_rate, b0));
            GSM_CHECK_RETURN(gsm_observed_hist(part_[(size_t)r], mine.data(), nb));
            for (int i = 0; i < nb; ++i) observed[(size_t)i] += mine[(size_t)i];   // counts: exact in float
        }
        std::vector<unsigned char> handles((size_t)G * GSM_PEER_HANDLE_BYTES);
        for (int r = 0; r < G; ++r)
            GSM_CHECK_RETURN(gsm_peer_export(part_[(size_t)r], G, opt_.max_batch, handles.data() + (size_t)r * GSM_PEER_HANDLE_BYTES));
        for (int r = 0; r < G; ++r) {
            GSM_CHECK_RETURN(gsm_set_observed_hist(part_[(size_t)r], observed.data(), nb));   // chi-square needs the global one
            GSM_CHECK_RETURN(gsm_peer_connect(part_[(size_t)r], handles.data(), r, G));
        }
    }

    // Contiguous burst ranges, one per part, balanced by sum(phCount + 4) — photons plus a floor
    // per burst record; the rule of gsmfretda_b200/sharding.py::split_bursts (the two are held
    // equal by tests/test_host_cpp.py).  -> parts + 1 cut points.
    static std::vector<int64_t> plan_split(const std::vector<int>& ph, int parts)
    {
        const int64_t B = (int64_t)ph.size();
        std::vector<int64_t> cost((size_t)B);
        int64_t acc = 0;
        for (int64_t i = 0; i < B; ++i) { acc += (int64_t)ph[(size_t)i] + 4; cost[(size_t)i] = acc; }
        std::vector<int64_t> cuts((size_t)parts + 1, 0);
        for (int r = 1; r < parts; ++r) {
            const double target = (double)acc * r / parts;
            int64_t lo = 0, hi = B;                       // first i with cost[i] >= target
            while (lo < hi) {
                const int64_t mid = (lo + hi) / 2;
                if ((double)cost[(size_t)mid] >= target) hi = mid; else lo = mid + 1;
            }
            cuts[(size_t)r] = B ? lo : 0;
        }
        cuts[(size_t)parts] = B;
        for (int r = 1; r <= parts; ++r) if (cuts[(size_t)r] < cuts[(size_t)r - 1]) cuts[(size_t)r] = cuts[(size_t)r - 1];
        return cuts;
    }

    bool get_max_conversion_capacity(int max_stateNum = 8) { return gsm_get_max_conversion_capacity(ctx, max_stateNum) == 1; }

    // ---- additions -------------------------------------------------------------------
    const uint64_t* hHist(int sid, int p = 0) const { return gsm_result_hist(ctx, sid, p); }
    float hChi2(int sid, int p = 0) const
    {
        int P = 0;
        const float* v = gsm_result_chi2v(ctx, sid, &P);
        return (v && p >= 0 && p < P) ? v[p] : NAN;
    }
    bool failed(int sid) const { return bad_[sid] != 0; }
    // P vectors (each n*n + n values, back to back) for ONE launch: the next run_kernel(N, sid)
    // evaluates them all (gsm_run_batch); results per vector through hChi2(sid, p) / hHist(sid, p),
    // hmcE[sid] holds P blocks of N * reSampleTimes values
    bool set_params_batch(int n, int sid, int P, const std::vector<float>& args)
    {
        bad_[sid] = n < 2 || P < 1 || (long long)args.size() != (long long)P * (n * n + n) ||
                    (part_.size() > 1 && P > opt_.max_batch);
        batch_[sid] = bad_[sid] ? 0 : P;
        if (!bad_[sid]) batch_args_[sid] = args;
        return !bad_[sid];
    }
    // The single vectors of the streams `sids` (args: their n*n + n values back to back, in the same
    // order) in ONE launch on carrier slot `carrier` (0 .. carriers()-1).  Every vector is drawn with
    // its own stream's epoch, so the results equal the per-stream launches bit for bit; afterwards
    // streamQuery(sid) / hChi2(sid) / hHist(sid) of a member read its share of the carrier, until the
    // stream runs by itself again.  A carrier is free again once carrierQuery says so AND its members
    // have read their results.  Lifts the per-launch driver cost off a node that evaluates many small
    // jobs (the reference's gpuWorker: one launch per stream, gpuWorker.cpp:38-87).
    // Reference-draws mode evaluates one vector per launch (gsm_run_coalesced answers GSM_ESTATE there): it has no carriers.
    int carriers() const { return mode_ == GSM_MODE_REFERENCE ? 0 : carriers_; }
    bool run_coalesced(int carrier, int n, int N, const std::vector<int>& sids, const std::vector<float>& args)
    {
        const int P = (int)sids.size();
        if (carrier < 0 || carrier >= carriers() || P < 1 || n < 2 || (long long)args.size() != (long long)P * (n * n + n) ||
            (part_.size() > 1 && P > opt_.max_batch))
            return false;
        for (size_t r = 0; r < part_.size(); ++r) {
            int Nr = N;
            if (part_.size() > 1) {
                int64_t in_part = (int64_t)N - cuts_[r];
                const int64_t size = cuts_[r + 1] - cuts_[r];
                Nr = (int)(in_part < 0 ? 0 : (in_part > size ? size : in_part));
            }
            const int rc = gsm_run_coalesced(part_[r], Nr, streamNum + carrier, n, P, args.data(), sids.data());
            // A vector the library refuses (GSM_EINVAL) or a context that does not coalesce (GSM_ESTATE: reference
            // draws, per-sample outputs): nothing was launched, the caller falls back to one launch per stream.
            // Every part holds the same flags, mode and vectors, so a refusal can only come from part 0.
            if ((rc == GSM_EINVAL || rc == GSM_ESTATE) && r == 0) return false;
            GSM_CHECK_RETURN(rc);
        }
        for (int sid : sids) { bad_[(size_t)sid] = 0; batch_[(size_t)sid] = 0; s_n[sid] = n; }
        return true;
    }
    bool carrierQuery(int carrier)
    {
        bool done = true;
        for (gsm_ctx* c : part_) done = (gsm_query(c, streamNum + carrier) == 1) && done;
        return done;
    }
    // draws of the next evaluation on the slot are keyed by `epoch` (reproducible runs)
    void set_epoch(int sid, uint32_t epoch)
    {
        for (gsm_ctx* c : part_) GSM_CHECK_RETURN(gsm_set_epoch(c, sid, epoch));
    }
    void set_mode(int mode)
    {
        for (gsm_ctx* c : part_) GSM_CHECK_RETURN(gsm_set_mode(c, mode));
        mode_ = mode;
    }
    gsm_ctx* handle() { return ctx; }
    int gpus() const { return (int)part_.size(); }
    gsm_ctx* part(int r) { return part_[(size_t)r]; }
    // the bursts [first, last) part r holds
    void part_range(int r, int64_t* first, int64_t* last) const
    {
        *first = part_.size() > 1 ? cuts_[(size_t)r] : 0;
        *last = part_.size() > 1 ? cuts_[(size_t)r + 1] : sz_burst;
    }
    // 0 when every part finished cleanly (bit 0: step cap, bit 1: a peer never arrived)
    uint64_t hFlags(int sid) const
    {
        uint64_t f = 0;
        for (gsm_ctx* c : part_) f |= gsm_result_flags(c, sid);
        return f;
    }
    // burst-split across processes (one `mc` per process): every rank peer_export, the handles are
    // gathered in rank order over any transport, peer_connect.  An `mc` that spans GPUs itself does
    // this in init_data_gpu.
    void peer_export(int world, int max_P, void* handle128) { GSM_CHECK_RETURN(gsm_peer_export(ctx, world, max_P, handle128)); }
    void peer_connect(const void* handles, int rank, int world) { GSM_CHECK_RETURN(gsm_peer_connect(ctx, handles, rank, world)); }
    void peer_disconnect() { GSM_CHECK_RETURN(gsm_peer_disconnect(ctx)); }

  protected:
    gsm_ctx* ctx;
    int streamNum;
    cudaStream_t* streams;
    Options opt_;
    std::vector<gsm_ctx*> part_;
    std::vector<int64_t> cuts_;
    int carriers_ = 0;
    int mode_ = GSM_MODE_REPLAY;
    bool profiler_ = false;
    std::vector<char> bad_;
    std::vector<int> batch_;
    std::vector<std::vector<float>> batch_args_;
};
