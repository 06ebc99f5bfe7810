// oracle/ref_gpu_shim.cpp — TEST INFRASTRUCTURE, never the product.
//
// extern "C" handle around the REFERENCE'S OWN `class mc`, compiled UNMODIFIED for sm_100a
// from where it lies (/root/reference/src/{mc.cu,eigenhelper.cpp,mrImp.cpp,tools.cpp},
// recipe: oracle/Makefile target `refgpu` -> oracle/_ref/libref_gpu.so).  It gives the GPU
// box the reference's real kernel (cuRAND scrambled Sobol64 + libdevice logf/powf/erfcinvf +
// FP64 rk_binomial), used for
//   * the statistical parity tests (KS on E, chi-square within 1 %) of tests/test_gpu_vs_refgpu.py,
//   * the GPU-side baseline SURVEY.md §8(d) asks for (`bench.py` key `ref_gpu`).
// The call sequence of one evaluation is gpuWorker::run's (/root/reference/src/gpuWorker.cpp:71-80).
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "mc.hpp"   // the reference's header (-I/root/reference/src)

namespace {
struct RefGpu {
    mc* m = nullptr;
    int N[64];
    std::vector<int64_t> istart, start, stop;
    std::vector<int> ph;
    std::vector<float> T, sg;
    float clk = 0, bga = 0, bgd = 0;
};
}  // namespace

extern "C" {

void* refgpu_create(int devid, int streams)
{
    RefGpu* r = new RefGpu;
    r->m = new mc(devid, streams, 0, 0, false);   // mc.cu:218
    for (int& n : r->N) n = -12345;               // main.cpp:36 leaves N[] uninitialised (Q9)
    return r;
}

// the reference's destructor ends with cudaDeviceReset() (mc.cu:660): the caller must not hold
// any other CUDA context state in this process afterwards
void refgpu_destroy(void* h)
{
    RefGpu* r = (RefGpu*)h;
    delete r->m;
    delete r;
}

void refgpu_set_resample(void* h, int R) { ((RefGpu*)h)->m->set_reSampleTimes(R); }   // mc.cu:316
int refgpu_resample(void* h) { return ((RefGpu*)h)->m->reSampleTimes; }

void refgpu_init_data(void* h, int64_t B, const int64_t* istart, const int64_t* start, const int64_t* stop,
                      const int* ph, int64_t n_ph, int64_t* burst_ad, int64_t* burst_dd, const float* T,
                      const float* sg, float clk_p, float bg_ad, float bg_dd)
{
    RefGpu* r = (RefGpu*)h;
    r->istart.assign(istart, istart + B);
    r->start.assign(start, start + B);
    r->stop.assign(stop, stop + B);
    r->ph.assign(ph, ph + B);
    r->T.assign(T, T + B);
    r->sg.assign(sg, sg + B);
    r->clk = clk_p; r->bga = bg_ad; r->bgd = bg_dd;
    r->m->set_gpuid();
    r->m->init_data_gpu(r->istart, r->start, r->stop, r->ph, (long)n_ph, burst_ad, burst_dd, r->T, r->sg,
                        r->clk, r->bga, r->bgd);   // mc.cu:456
}

// one evaluation exactly as gpuWorker::run drives it; returns N (bursts), results in hmcE[sid]
int refgpu_launch(void* h, int n, const float* args, int sid)
{
    RefGpu* r = (RefGpu*)h;
    mc* m = r->m;
    std::vector<float> a(args, args + n * n + n);
    int old_n = m->set_nstates(n, sid);
    m->set_params(n, sid, a);
    int Ns = m->setBurstBd(0, -1, sid);
    if (Ns != r->N[sid]) {
        m->init_randstate(Ns, sid);
        r->N[sid] = Ns;
        // gpuWorker.cpp:78 tests N_sid != N[sid] AFTER the assignment above, so the buffer is
        // only (re)allocated when n changes; do it here too whenever N changed, otherwise the
        // first evaluation of a stream whose n is already set would write through a stale pointer
        m->set_params_buff(old_n, Ns, sid);
    } else if (old_n != n) {
        m->set_params_buff(old_n, Ns, sid);
    }
    m->run_kernel(r->N[sid], sid);
    return Ns;
}

int refgpu_query(void* h, int sid) { return ((RefGpu*)h)->m->streamQuery(sid) ? 1 : 0; }

const float* refgpu_hmcE(void* h, int sid) { return ((RefGpu*)h)->m->hmcE[sid]; }

// K evaluations back to back on one stream, host clock around launch + wait (the reference has
// no device timers); returns seconds
double refgpu_time(void* h, int n, const float* args, int sid, int K)
{
    auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < K; ++k) {
        refgpu_launch(h, n, args, sid);
        while (!refgpu_query(h, sid)) {}
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// the HOST half of one reference evaluation lives in ref_host_half.cpp (plain C++: nvcc's front end
// does not take boost::histogram); this hands it the reference's own buffers
double ref_host_half(const float* mcE, size_t count, const float* sg, size_t n_sg, int R, int fret_hist_num, int left_pad,
                     int n_states, int K, float* chi2_out);
double refgpu_host_half(void* h, int sid, int N, int fret_hist_num, int left_pad, int n_states, int K, float* chi2_out)
{
    RefGpu* r = (RefGpu*)h;
    return ref_host_half(r->m->hmcE[sid], (size_t)N * r->m->reSampleTimes, r->sg.data(), r->sg.size(), r->m->reSampleTimes,
                         fret_hist_num, left_pad, n_states, K, chi2_out);
}

}  // extern "C"
