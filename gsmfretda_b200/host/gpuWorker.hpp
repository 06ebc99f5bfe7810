// gpuWorker.hpp — the single GPU-side scheduling thread, same constructor and state machine
// as the reference's gpuWorker (/root/reference/src/gpuWorker.hpp:9-27, gpuWorker.cpp:38-87):
//   dataready[sid]: 0 need-capacity, 1 need-params, 2 params-ready, 3 kernel-in-flight,
//   4 results-ready.  The net/param side moves 0->1->2 and 4->0; this thread moves 2->3->4.
#pragma once
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "mc.hpp"

class gpuWorker {
  public:
    gpuWorker(mc* _pdamc, int _streamNum, std::vector<float>* _d, int _fretHistNum, std::mutex* m,
              std::condition_variable* cv, int* _dataready, int* _sn, std::vector<float>* _params,
              int* _ga_start, int* _ga_stop, int* _N, unsigned char debugbool = 0)
        : pdamc(_pdamc), streamNum(_streamNum), SgDivSr(_d), fretHistNum(_fretHistNum), _m(m), _cv(cv),
          dataready(_dataready), s_n(_sn), params(_params), N(_N), ga_start(_ga_start),
          ga_stop(_ga_stop), debug(debugbool) {}

    void run(int /*sz_burst*/)
    {
        pdamc->set_gpuid();
        do {
            for (int sid = 0; sid < streamNum; ++sid) {
                std::unique_lock<std::mutex> lck(_m[sid], std::defer_lock);
                if (!lck.try_lock()) continue;
                const int st = dataready[sid];
                if (st == 3) {
                    const bool done = pdamc->streamQuery(sid);
                    if (done) dataready[sid] = 4;
                    lck.unlock();
                    if (done) _cv[sid].notify_one();
                } else if (st == 2) {
                    launch(sid);
                    dataready[sid] = 3;
                    lck.unlock();
                } else {
                    lck.unlock();
                }
            }
        } while (pdamc->workerNum.load() > 0);
    }

  private:
    // the six calls of gpuWorker.cpp:71-80, in the reference's order.  params[sid] holding a
    // whole number P > 1 of vectors is a batch (streamWorker::fetch): staged with
    // set_params_batch and evaluated by the same run_kernel in one launch.  A vector the library
    // refuses (wrong count, n < 2) launches nothing; streamWorker reports it with chi2 = +inf.
    void launch(int sid)
    {
        const int n = s_n[sid];
        const int oldS_n = pdamc->set_nstates(n, sid);
        const size_t per_vector = (size_t)n * (size_t)(n + 1);
        const size_t have = params[sid].size();
        if (per_vector > 0 && have > per_vector && have % per_vector == 0)
            pdamc->set_params_batch(n, sid, (int)(have / per_vector), params[sid]);
        else
            pdamc->set_params(n, sid, params[sid]);
        const int N_sid = pdamc->setBurstBd(ga_start[sid], ga_stop[sid], sid);
        const bool n_changed = N_sid != N[sid];
        if (n_changed) {
            pdamc->init_randstate(N_sid, sid);
            N[sid] = N_sid;
        }
        if (oldS_n != n || n_changed) pdamc->set_params_buff(oldS_n, N_sid, sid);
        pdamc->run_kernel(N[sid], sid);
    }

    mc* pdamc;
    int streamNum;
    std::vector<float>* SgDivSr;
    int fretHistNum;
    std::mutex* _m;
    std::condition_variable* _cv;
    int* dataready;
    int* s_n;
    std::vector<float>* params;
    int* N;
    int* ga_start;
    int* ga_stop;
    unsigned char debug;
};
