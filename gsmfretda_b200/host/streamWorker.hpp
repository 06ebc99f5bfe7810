// streamWorker.hpp — the network side of one stream slot: the REQ/REP conversation with
// pySMFRETda's parameter server.
//
// Kept from the reference because other code binds to it: the class name, the constructor
// (/root/reference/src/streamWorker.hpp:10-14), `run(sid, sz_burst, leftPad)`, the meaning of the
// shared per-stream arrays and of the `dataready` values (0 need-capacity, 1 need-parameters,
// 2 parameters-ready, 3 in flight, 4 results-ready; the gpuWorker thread moves 2 -> 3 -> 4), and
// every byte on the wire (opcode + args.proto message, /root/reference/src/streamWorker.cpp:83-238):
//
//   'c' + p_cap{cap}                       -> p_n{s_n}
//   'p' + p_str{str, histNum, ohist[]}     -> p_ga{idx, start, stop, bestfv, params[]}
//                                             idx -1: nothing to do now, ask again; idx -2: leave
//   'k' + p_sid{sid = current idx}         -> anything            (keep-alive while the GPU works)
//   'r' + res{ridx, e, hist, idx, shist[]} -> p_sid{sid}; sid -1: leave
//
// The body is this node's own: the conversation is a table of four handlers, one per state this
// thread owns, over a small typed request/reply helper (`Link`).  Two things go beyond the
// reference.  (1) The model histogram and the reduced chi-square come from the fused evaluation
// kernel (mc::hHist / mc::hChi2); `host_hist = true` recomputes them on the host from hmcE with
// the reference's arithmetic (tests/test_gpu_node.py holds both to the same wire bytes).
// (2) Batches: a p_ga whose params[] holds P > 1 vectors (P (n*n + n) values, indices idx .. idx +
// P - 1) is evaluated by ONE launch (mc::run_batch) and answered by one `res` that adds es[] and
// best (args_pb.hpp); a single vector produces the reference's bytes exactly.
#pragma once
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <limits>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "args_pb.hpp"
#include "mc.hpp"
#include "node_log.hpp"
#include "sp_req.hpp"

namespace node {

// One REQ socket and typed round trips over it.
class Link {
  public:
    explicit Link(const std::string& url)
    {
        fd_ = nn_socket(AF_SP, NN_REQ);
        if (fd_ < 0) return;
        int ms = 200;   // the reference's reconnect interval (streamWorker.cpp:48-49)
        nn_setsockopt(fd_, NN_SOL_SOCKET, NN_RECONNECT_IVL, &ms, sizeof ms);
        if (nn_connect(fd_, url.c_str()) < 0) {
            nn_close(fd_);
            fd_ = -1;
        }
    }
    ~Link() { if (fd_ >= 0) nn_close(fd_); }
    Link(const Link&) = delete;
    Link& operator=(const Link&) = delete;
    bool ok() const { return fd_ >= 0; }

    // opcode byte + serialized request out, one message in; false = try the exchange again
    template <class Request, class Reply>
    bool ask(char opcode, const Request& rq, Reply* rp)
    {
        frame_.assign(1, opcode);
        frame_ += rq.SerializeAsString();
        if (nn_send(fd_, frame_.data(), frame_.size(), 0) < 0) return false;
        char* in = nullptr;
        const int got = nn_recv(fd_, &in, NN_MSG, 0);
        if (got < 0) return false;
        if (rp) rp->ParseFromArray(in, got);
        nn_freemsg(in);
        return true;
    }

  private:
    int fd_ = -1;
    std::string frame_;
};

// regular axis [lo, hi) with `bins` inner bins, underflow/overflow dropped: what the reference
// builds with boost::histogram + indexed(coverage::inner) (streamWorker.cpp:33-38,62-67; bin rule
// boost/histogram/axis/regular.hpp:292-308: z = (x - lo) / (hi - lo) in double, bin (int)(z bins))
inline std::vector<float> inner_histogram(const float* x, size_t count, int bins, float lo, float hi)
{
    std::vector<float> h((size_t)bins, 0.0f);
    const double span = (double)hi - (double)lo;
    for (const float* it = x; it != x + count; ++it) {
        const double z = ((double)*it - (double)lo) / span;
        if (z >= 0.0 && z < 1.0) h[(size_t)(int)(z * bins)] += 1.0f;
    }
    return h;
}

// reduced chi-square of a model histogram against the observed one, the reference's host loop
// (streamWorker.cpp:197-206), float accumulation included
inline float reduced_chi2(const std::vector<float>& observed, const std::vector<float>& model, int R, int left_pad,
                          int n_states)
{
    const int bins = (int)model.size();
    int dof = bins - left_pad;
    float sum = 0;
    for (int b = left_pad; b < bins; ++b) {
        if (model[(size_t)b] > 0)
            sum += std::pow((float(observed[(size_t)b] - model[(size_t)b] / R)), 2) / (float(model[(size_t)b]) / R);
        else
            --dof;
    }
    return sum / (dof - n_states * (n_states + 1));
}

}  // namespace node

class streamWorker {
  public:
    streamWorker(mc* _pdamc, std::string* _url, std::vector<float>* _d, int _fretHistNum, std::mutex* m,
                 std::condition_variable* cv, int* _dataready, int* _sn, std::vector<float>* _params,
                 int* _ga_start, int* _ga_stop, int* _N, unsigned char debugl = 0)
        : engine_(_pdamc), url_(_url), observed_values_(_d), bins_(_fretHistNum), slot_mutex_(m), slot_cv_(cv),
          state_(_dataready), n_states_(_sn), vectors_(_params), range_lo_(_ga_start), range_hi_(_ga_stop),
          bursts_(_N), log_mask_(debugl) {}

    bool host_hist = false;       // recompute histogram + chi-square on the host from hmcE (the reference's path)
    int keepalive_every = 60;     // 500 ms waits between two 'k' messages (streamWorker.cpp:165)
    bool epoch_from_idx = false;  // key the draws of an evaluation by the server's vector index (reproducible runs)

    // the reference's name for the observed-histogram helper, kept for callers and tests
    static std::vector<float> mkhist(const float* x, size_t n, int binnum, float lv, float uv)
    {
        return node::inner_histogram(x, n, binnum, lv, uv);
    }

    void run(int sid, int sz_burst, int leftPad)
    {
        Session s(*url_);
        s.sid = sid;
        s.capacity = sz_burst;
        s.left_pad = leftPad;
        s.name = node::stream_name(engine_->devid, sid, engine_->gpuuuid);
        s.observed = node::inner_histogram(observed_values_->data(), observed_values_->size(), bins_, 0.0f, 1.0f);
        if (s.link.ok()) {
            // one handler per value of dataready[sid]; 2 and 3 belong to the GPU thread, this thread waits
            static constexpr Handler table[5] = {&streamWorker::announce, &streamWorker::fetch, &streamWorker::wait_gpu,
                                                 &streamWorker::wait_gpu, &streamWorker::report};
            while (!s.leaving) {
                std::unique_lock<std::mutex> hold(slot_mutex_[sid], std::try_to_lock);
                if (!hold.owns_lock()) continue;   // the GPU thread is looking at the slot
                const int st = state_[sid];
                if (st < 0 || st > 4) break;
                (this->*table[st])(s, hold);
            }
        }
        engine_->workerNum--;   // the GPU thread leaves its loop when every stream has left
    }

  private:
    struct Session {
        explicit Session(const std::string& url) : link(url) {}
        node::Link link;
        int sid = 0, capacity = 0, left_pad = 0;
        std::string name;
        std::vector<float> observed;
        bool observed_sent = false;
        bool leaving = false;
        int32_t job = 0;           // index of the (first) vector in flight
        int batch = 1;             // vectors in flight
        float best_so_far = std::numeric_limits<float>::max();
        unsigned long idle_waits = 0;
    };
    using Handler = void (streamWorker::*)(Session&, std::unique_lock<std::mutex>&);

    // state 0: tell the server how many bursts this node holds, learn the number of states
    void announce(Session& s, std::unique_lock<std::mutex>&)
    {
        gSMFRETda::pb::p_cap mine;
        mine.set_cap(s.capacity);
        gSMFRETda::pb::p_n model;
        if (!s.link.ask('c', mine, &model)) return;
        n_states_[s.sid] = model.s_n();
        state_[s.sid] = 1;
    }

    // state 1: ask for work; the observed histogram rides on the first request only
    void fetch(Session& s, std::unique_lock<std::mutex>& hold)
    {
        gSMFRETda::pb::p_str hello;
        hello.set_str(s.name);
        hello.set_histnum(s.observed_sent ? 0 : bins_);
        if (!s.observed_sent)
            for (float v : s.observed) hello.add_ohist(v);
        gSMFRETda::pb::p_ga work;
        const bool answered = s.link.ask('p', hello, &work);
        s.observed_sent = true;       // written once, whether or not the reply arrived (as the reference)
        if (!answered) return;
        node::say(log_mask_, node::kCpu, "asked for work as %s\n", s.name.c_str());
        s.best_so_far = work.bestfv();
        if (work.idx() == -1) return;                     // nothing for this stream right now
        if (work.idx() == -2) { s.leaving = true; return; }   // the optimisation is over
        const int n = n_states_[s.sid];
        const int per_vector = n * (n + 1);
        // a batch is a whole number (> 1) of vectors; anything else is one vector, zero-padded or
        // cut to size (the reference reads params(i) for i < n(n+1) whatever arrived)
        int count = 1;
        if (per_vector > 0 && work.params_size() > per_vector && work.params_size() % per_vector == 0)
            count = work.params_size() / per_vector;
        std::vector<float>& v = vectors_[s.sid];
        v.assign((size_t)per_vector * (size_t)count, 0.0f);
        for (int i = 0; i < (int)v.size() && i < work.params_size(); ++i) v[(size_t)i] = work.params(i);
        s.job = work.idx();
        s.batch = count;
        range_lo_[s.sid] = work.start();
        range_hi_[s.sid] = work.stop();
        node::say(log_mask_, node::kNet, "vector %d (+%d) received\n", (int)s.job, count - 1);
        // a host-side field of the slot: this thread holds the slot and the GPU thread only touches
        // it in states 2 and 3
        if (epoch_from_idx) engine_->set_epoch(s.sid, (uint32_t)s.job + 1u);
        state_[s.sid] = 2;
        hold.unlock();
        slot_cv_[s.sid].notify_one();
    }

    // states 2 and 3: the GPU thread owns the slot; sleep until it posts the results, and let the
    // server know now and then that this stream is alive
    void wait_gpu(Session& s, std::unique_lock<std::mutex>& hold)
    {
        const int sid = s.sid;
        if (slot_cv_[sid].wait_for(hold, std::chrono::milliseconds(500), [this, sid] { return state_[sid] == 4; })) return;
        node::say(log_mask_, node::kCpu, "stream %d still waiting (state %d)\n", sid, state_[sid]);
        if (++s.idle_waits % (unsigned long)keepalive_every != 0) return;
        gSMFRETda::pb::p_sid alive;
        alive.set_sid(s.job);
        s.link.ask('k', alive, (gSMFRETda::pb::p_sid*)nullptr);
    }

    // state 4: send chi-square (and the model histogram when it beats the server's best)
    void report(Session& s, std::unique_lock<std::mutex>&)
    {
        const int sid = s.sid;
        const int R = engine_->reSampleTimes;
        const int n = n_states_[sid];
        const bool failed = engine_->failed(sid);   // the vector was refused (wrong size / n < 2): nothing ran
        std::vector<float> chi(s.batch, std::numeric_limits<float>::infinity());
        int best = 0;
        if (!failed) {
            for (int p = 0; p < s.batch; ++p) {
                if (host_hist) {
                    const size_t per = (size_t)bursts_[sid] * (size_t)R;
                    chi[p] = node::reduced_chi2(s.observed, node::inner_histogram(engine_->hmcE[sid] + p * per, per, bins_, 0.0f, 1.0f),
                                                R, s.left_pad, n);
                } else {
                    chi[p] = engine_->hChi2(sid, p);
                }
                if (chi[p] < chi[best]) best = p;
            }
        }
        gSMFRETda::pb::res out;
        out.set_idx(s.name);
        out.set_ridx(s.job);
        out.set_e(chi[best]);
        const bool attach = !failed && chi[best] < s.best_so_far;
        if (attach) {
            if (host_hist) {
                const size_t per = (size_t)bursts_[sid] * (size_t)R;
                for (float c : node::inner_histogram(engine_->hmcE[sid] + best * per, per, bins_, 0.0f, 1.0f)) out.add_shist(c / R);
            } else {
                const uint64_t* counts = engine_->hHist(sid, best);
                for (int b = 0; b < bins_; ++b) out.add_shist((float)counts[b] / R);
            }
        }
        out.set_hist(attach);
        if (s.batch > 1) {
            for (float c : chi) out.add_es(c);
            out.set_best(best);
        }
        gSMFRETda::pb::p_sid verdict;
        if (!s.link.ask('r', out, &verdict)) return;     // same state: the result is sent again
        node::say(log_mask_, node::kNet, "vector %d: chi2 %g sent, server says %d\n", (int)s.job, (double)chi[best],
                  (int)verdict.sid());
        if (verdict.sid() == -1) s.leaving = true;
        state_[sid] = 0;
    }

    mc* engine_;
    std::string* url_;
    std::vector<float>* observed_values_;
    int bins_;
    std::mutex* slot_mutex_;
    std::condition_variable* slot_cv_;
    int* state_;
    int* n_states_;
    std::vector<float>* vectors_;
    int* range_lo_;
    int* range_hi_;
    int* bursts_;
    unsigned char log_mask_;
};
