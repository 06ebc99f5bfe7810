// mc.hpp — the reference's `class mc` (/root/reference/src/mc.hpp:78-179) re-created on top of
// the C ABI of include/gsmfret_c.h.  Same public members, argument meaning and error
// behaviour, so that code written against the reference — gpuWorker::run
// (/root/reference/src/gpuWorker.cpp:38-87), the result-reading part of streamWorker::run
// (/root/reference/src/streamWorker.cpp:186-206), main (/root/reference/src/main.cpp:129-160)
// — compiles against it unchanged.  Header-only; link with libgsmfret.so.
//
// Differences, all additive:
//   * hHist(sid) / hChi2(sid): the histogram and reduced chi-square the fused kernel already
//     produced, so a caller no longer needs hmcE and the host histogram loop;
//   * mc::Options chooses whether hmcE is kept (default: yes, for drop-in compatibility);
//   * set_reSampleTimes keeps the reference's rounding down to a power of two; use
//     set_reSampleTimes_exact for any R (the new kernel has no such restriction).
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
        uint32_t seed = 0x9E3779B9u;
        int fret_hist_num = 200;   // cmdline.ggo:14
        int left_pad = 2;          // main.cpp:155
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
        : mc(id, _streamNum, debug, hdf5size, profiler, Options()) {}

    mc(int id, int _streamNum, unsigned char /*debug*/, std::uintmax_t /*hdf5size*/, bool /*profiler*/,
       const Options& opt)
        : devid(id), hmcE(nullptr), s_n(nullptr), sz_burst(0), reSampleTimes(4), ctx(nullptr),
          streamNum(0), streams(nullptr)
    {
        gpuuuid[0] = 0;
        unsigned flags = (opt.keep_mce ? GSM_KEEP_MCE : 0u) | (opt.force_raw64 ? GSM_FORCE_RAW64 : 0u) |
                         (opt.timing ? 0u : GSM_NO_TIMING);
        int rc = gsm_create(id, _streamNum, flags, &ctx);
        if (rc == GSM_EINVAL) {  // "gpu id set error!" (mc.cu:228-231): the reference returns quietly
            std::printf("gpu id set error!\n");
            workerNum.store(_streamNum);
            return;
        }
        GSM_CHECK_RETURN(rc);
        streamNum = gsm_stream_count(ctx);
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
        GSM_CHECK_RETURN(gsm_set_seed(ctx, opt.seed));
        GSM_CHECK_RETURN(gsm_set_hist(ctx, opt.fret_hist_num, opt.left_pad));
        GSM_CHECK_RETURN(gsm_set_resample_times(ctx, reSampleTimes));
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
        gsm_destroy(ctx);
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
        int r = gsm_query(ctx, sid);
        GSM_CHECK_RETURN(r);
        return r == 1;
    }

    int set_nstates(int n, int sid)
    {
        int r = gsm_set_nstates(ctx, n, sid);
        GSM_CHECK_RETURN(r);
        s_n[sid] = n;
        return r;  // the previous n (== n when unchanged), as mc::set_nstates
    }

    // false when args has the wrong count for n states (eigenhelper.cpp:5-8); the slot is then
    // marked failed: run_kernel launches nothing and failed(sid) tells the reader of the results
    bool set_params(int n, int sid, std::vector<float>& args)
    {
        int r = gsm_set_params(ctx, n, sid, args.data(), (int)args.size());
        GSM_CHECK_RETURN(r);
        bad_[sid] = r != 1;
        batch_[sid] = 0;
        return r == 1;
    }

    int setBurstBd(int cstart, int cstop, int sid)
    {
        int r = gsm_set_burst_bd(ctx, cstart, cstop, sid);
        if (r < -1) GSM_CHECK_RETURN(r);
        return r;
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
        GSM_CHECK_RETURN(gsm_set_resample_times(ctx, t));
        reSampleTimes = t;
    }

    void init_randstate(int N, int sid) { GSM_CHECK_RETURN(gsm_init_randstate(ctx, N, sid)); }
    void set_params_buff(int oldS_n, int N_sid, int sid) { GSM_CHECK_RETURN(gsm_set_params_buff(ctx, oldS_n, N_sid, sid)); }

    void run_kernel(int N, int sid)
    {
        if (bad_[sid]) return;   // the vector was refused: nothing to evaluate, the stream stays idle
        if (batch_[sid] > 0) {
            GSM_CHECK_RETURN(gsm_run_batch(ctx, N, sid, s_n[sid], batch_[sid], batch_args_[sid].data()));
        } else {
            GSM_CHECK_RETURN(gsm_run(ctx, N, sid));
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
        GSM_CHECK_RETURN(gsm_init_data(ctx, (int64_t)start.size(), istart.data(), start.data(), stop.data(),
                                       phCount.data(), (int64_t)_sz_tag, burst_ad, burst_dd,
                                       T_burst_duration.data(), SgDivSr.data(), iclk_p, ibg_ad_rate,
                                       ibg_dd_rate, 0));
        sz_burst = (int)start.size();
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
    bool failed(int sid) const { return bad_[sid]; }
    // P vectors (each n*n + n values, back to back) for ONE launch: the next run_kernel(N, sid)
    // evaluates them all (gsm_run_batch); results per vector through hChi2(sid, p) / hHist(sid, p),
    // hmcE[sid] holds P blocks of N * reSampleTimes values
    bool set_params_batch(int n, int sid, int P, const std::vector<float>& args)
    {
        bad_[sid] = n < 2 || P < 1 || (long long)args.size() != (long long)P * (n * n + n);
        batch_[sid] = bad_[sid] ? 0 : P;
        if (!bad_[sid]) batch_args_[sid] = args;
        return !bad_[sid];
    }
    gsm_ctx* handle() { return ctx; }
    // burst-split across GPUs, the all-reduce of the histogram fused into the evaluation kernel
    // (include/gsmfret_c.h: gsm_peer_*).  Every rank: peer_export, gather the handles in rank
    // order over any transport, peer_connect; init_data_split gives the rank its burst range.
    void peer_export(int world, int max_P, void* handle128) { GSM_CHECK_RETURN(gsm_peer_export(ctx, world, max_P, handle128)); }
    void peer_connect(const void* handles, int rank, int world) { GSM_CHECK_RETURN(gsm_peer_connect(ctx, handles, rank, world)); }
    void peer_disconnect() { GSM_CHECK_RETURN(gsm_peer_disconnect(ctx)); }
    uint64_t hFlags(int sid) const { return gsm_result_flags(ctx, sid); }

  protected:
    gsm_ctx* ctx;
    int streamNum;
    cudaStream_t* streams;
    std::vector<char> bad_;
    std::vector<int> batch_;
    std::vector<std::vector<float>> batch_args_;
};
