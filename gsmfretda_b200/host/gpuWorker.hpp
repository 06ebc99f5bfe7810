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

    // Coalescing (not in the reference): when at least `min_group` streams hold a single vector of
    // the same model over the same burst range at the same moment, they are evaluated by ONE launch
    // on a carrier slot of the `mc` (mc::run_coalesced; every vector keeps its own stream's draw
    // epoch, so the results are the bits of the per-stream launches).  0 switches it off.
    // While other evaluations are still in flight, vectors that are ready wait up to `linger_us` for
    // company (until `target` of them are ready): the launches a node saves this way are what bounds it
    // at the reference's own job size (~9 us of driver calls per launch on the one gpuWorker thread).
    void set_coalesce(int min_group, int target = 16, int linger_us = 0)
    {
        coalesce_min = (pdamc->carriers() > 0) ? min_group : 0;
        coalesce_target = target;
        coalesce_linger_us = linger_us;
    }
    // what the coalescing loop did: passes over the streams, launches that carried a group, vectors in
    // those groups, launches of a single stream
    struct Stats { unsigned long long passes = 0, group_launches = 0, grouped_vectors = 0, single_launches = 0; };
    const Stats& stats() const { return stats_; }

    void run(int sz_burst)
    {
        if (coalesce_min > 0) { run_coalescing(sz_burst); return; }
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
    // One pass over the streams keeps the locks of those with a vector ready, groups them by
    // (states, bursts), launches each group of min_group or more on a free carrier and the rest one
    // by one, and only then lets the streams go.  A carrier is free again when its launch has
    // finished and every member has been seen past "results ready" (its results are read).
    void run_coalescing(int /*sz_burst*/)
    {
        pdamc->set_gpuid();
        const int C = pdamc->carriers();
        std::vector<std::vector<int>> riding((size_t)C);
        std::vector<int> seen((size_t)streamNum, -1), ready, group, on_carrier((size_t)streamNum, -1);
        std::vector<char> landed((size_t)C, 0);
        bool lingering = false;
        std::chrono::steady_clock::time_point linger_since{};
        std::vector<std::unique_lock<std::mutex>> held;
        std::vector<float> args;
        do {
            ++stats_.passes;
            ready.clear();
            held.clear();
            // one query per carrier and pass answers for all its members
            for (int c = 0; c < C; ++c) landed[(size_t)c] = !riding[(size_t)c].empty() && pdamc->carrierQuery(c);
            for (int sid = 0; sid < streamNum; ++sid) {
                std::unique_lock<std::mutex> lck(_m[sid], std::defer_lock);
                seen[(size_t)sid] = -1;
                if (!lck.try_lock()) continue;
                const int st = dataready[sid];
                seen[(size_t)sid] = st;
                if (st == 3) {
                    const int c = on_carrier[(size_t)sid];
                    const bool done = c >= 0 ? landed[(size_t)c] != 0 : pdamc->streamQuery(sid);
                    if (done) dataready[sid] = 4;
                    lck.unlock();
                    if (done) _cv[sid].notify_one();
                } else if (st == 2) {
                    ready.push_back(sid);
                    held.push_back(std::move(lck));
                }
            }
            // company is worth waiting for only while the GPU has other work
            if (!ready.empty()) {
                int inflight = 0;
                for (int sid = 0; sid < streamNum; ++sid) inflight += seen[(size_t)sid] == 3;
                const auto now = std::chrono::steady_clock::now();
                if (!lingering) { lingering = true; linger_since = now; }
                const bool wait_more = inflight > 0 && (int)ready.size() < coalesce_target &&
                                       std::chrono::duration_cast<std::chrono::microseconds>(now - linger_since).count() < coalesce_linger_us;
                if (wait_more) {
                    held.clear();   // the vectors stay ready (state 2); look again
                    std::this_thread::yield();   // the parameter threads need these mutexes to join the group
                    continue;
                }
                lingering = false;
            }
            for (int c = 0; c < C; ++c) {
                if (riding[(size_t)c].empty() || !landed[(size_t)c]) continue;
                bool read = true;
                for (int sid : riding[(size_t)c]) read = read && seen[(size_t)sid] >= 0 && seen[(size_t)sid] != 3 && seen[(size_t)sid] != 4;
                if (read) riding[(size_t)c].clear();
            }
            std::vector<char> taken(ready.size(), 0);
            for (size_t i = 0; i < ready.size(); ++i) {
                if (taken[i]) continue;
                const int sid = ready[i], n = s_n[sid];
                const size_t per_vector = (size_t)n * (size_t)(n + 1);
                int carrier = -1;
                for (int c = 0; c < C && carrier < 0; ++c)
                    if (riding[(size_t)c].empty()) carrier = c;
                group.clear();
                if (carrier >= 0 && n >= 2 && params[sid].size() == per_vector) {
                    for (size_t k = i; k < ready.size(); ++k) {
                        const int o = ready[k];
                        if (!taken[k] && s_n[o] == n && params[o].size() == per_vector && ga_start[o] == ga_start[sid] &&
                            ga_stop[o] == ga_stop[sid])
                            group.push_back(o);
                    }
                }
                if ((int)group.size() >= coalesce_min) {
                    args.clear();
                    for (int o : group) args.insert(args.end(), params[o].begin(), params[o].end());
                    const int Ng = pdamc->setBurstBd(ga_start[sid], ga_stop[sid], sid);
                    if (pdamc->run_coalesced(carrier, n, Ng, group, args)) {
                        for (int o : group) {
                            N[o] = Ng;
                            on_carrier[(size_t)o] = carrier;
                            dataready[o] = 3;
                            for (size_t k = i; k < ready.size(); ++k)
                                if (ready[k] == o) taken[k] = 1;
                        }
                        riding[(size_t)carrier] = group;
                        ++stats_.group_launches;
                        stats_.grouped_vectors += group.size();
                        continue;
                    }
                }
                launch(sid);
                ++stats_.single_launches;
                on_carrier[(size_t)sid] = -1;
                dataready[sid] = 3;
                taken[i] = 1;
            }
            held.clear();   // unlocks
        } while (pdamc->workerNum.load() > 0);
    }

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
    int coalesce_min = 0, coalesce_target = 16, coalesce_linger_us = 0;
    Stats stats_;
};
