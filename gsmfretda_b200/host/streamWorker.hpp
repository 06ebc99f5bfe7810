// streamWorker.hpp — the network side of one stream of the node: the REQ/REP conversation with
// pySMFRETda's parameter server, same constructor, same `run(sid, sz_burst, leftPad)` and same
// wire bytes as the reference's streamWorker (/root/reference/src/streamWorker.hpp:8-31,
// streamWorker.cpp:39-245).  Every request is one opcode byte followed by a protobuf message
// (args.proto):
//
//   dataready 0   'c' + p_cap{cap = sz_burst}            ->  p_n{s_n}                 -> 1
//   dataready 1   'p' + p_str{str = node id, histNum,    ->  p_ga{idx, start, stop,   -> 2
//                  ohist[] on the first request only}         bestfv, params[s_n(s_n+1)]}
//                                                             idx == -1: nothing to do, ask again
//                                                             idx == -2: leave
//   dataready 2/3  (the gpuWorker thread launches and polls the evaluation; every 60 waits of
//                  500 ms: 'k' + p_sid{sid = params_idx} -> any reply, keep-alive)
//   dataready 4   'r' + res{ridx = params_idx, e = chi2, ->  p_sid{sid}; sid == -1: leave  -> 0
//                  hist, idx = node id, shist[] = model
//                  histogram / reSampleTimes when chi2 < bestfv}
//
// What differs from the reference is where the numbers come from, not what is sent: the model
// histogram and the reduced chi-square are produced by the evaluation kernel itself
// (mc::hHist / mc::hChi2, fused), so the N*R-float copy and the host histogram loop of
// streamWorker.cpp:187-206 are gone.  `streamWorker::host_hist = true` brings that loop back —
// hmcE -> regular-axis histogram -> chi-square on the host, the reference's exact arithmetic —
// and tests/test_gpu_node.py checks both produce the same bytes on the wire.
#pragma once
#include <cassert>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "args_pb.hpp"
#include "mc.hpp"
#include "sp_req.hpp"
#include "tools.hpp"

class streamWorker {
  public:
    streamWorker(mc* _pdamc, std::string* _url, std::vector<float>* _d, int _fretHistNum, std::mutex* m,
                 std::condition_variable* cv, int* _dataready, int* _sn, std::vector<float>* _params,
                 int* _ga_start, int* _ga_stop, int* _N, unsigned char debugl = 0)
        : pdamc(_pdamc), SgDivSr(_d), fretHistNum(_fretHistNum), url(_url), _m(m), _cv(cv),
          dataready(_dataready), s_n(_sn), params(_params), debug(debugl), N(_N), ga_start(_ga_start),
          ga_stop(_ga_stop) {}

    bool host_hist = false;     // true: the reference's host path (hmcE -> histogram -> chi-square)
    int keepalive_every = 60;   // waits of 500 ms between two 'k' messages (streamWorker.cpp:165)
    bool epoch_from_idx = false;  // draws of an evaluation keyed by the server's params_idx (reproducible)

    // boost::histogram regular axis [lv, uv) with binnum bins, inner bins only (role of mkhist +
    // indexed(coverage::inner), streamWorker.cpp:33-38,62-67; bin rule of
    // boost/histogram/axis/regular.hpp:292-308: z = (x - lv) / (uv - lv) in double,
    // 0 <= z < 1 -> (int)(z * binnum); z >= 1 or NaN -> overflow, z < 0 -> underflow, both dropped)
    static std::vector<float> mkhist(const float* x, size_t n, int binnum, float lv, float uv)
    {
        std::vector<float> h((size_t)binnum, 0.0f);
        const double delta = (double)uv - (double)lv;
        for (size_t i = 0; i < n; ++i) {
            const double z = ((double)x[i] - (double)lv) / delta;
            if (z < 1.0) {
                if (z >= 0.0) h[(size_t)(int)(z * binnum)] += 1.0f;
            }
        }
        return h;
    }

    void run(int sid, int sz_burst, int leftPad)
    {
        bool send_sHist = false;
        bool sent_oHist = false;
        float msg_bestcs = 3.1E32f;
        unsigned long keepalivecount = 0;
        int sock = nn_socket(AF_SP, NN_REQ);
        assert(sock >= 0);
        int reconnect_ivl = 200;
        nn_setsockopt(sock, NN_SOL_SOCKET, NN_RECONNECT_IVL, &reconnect_ivl, sizeof(reconnect_ivl));
        int nneo = nn_connect(sock, url->c_str());
        assert(nneo >= 0);
        (void)nneo;
        std::string gpuNodeId;
        genuid(&gpuNodeId, pdamc->devid, sid, pdamc->gpuuuid);
        // observed histogram of SgDivSr, raw counts (streamWorker.cpp:61-67)
        const std::vector<float> vOEHist = mkhist(SgDivSr->data(), SgDivSr->size(), fretHistNum, 0.0f, 1.0f);
        bool ending = false;
        int32_t params_idx = 0;
        do {
            std::unique_lock<std::mutex> lck(_m[sid], std::defer_lock);
            if (!lck.try_lock()) continue;
            char* rbuf = nullptr;
            int bytes;
            if (dataready[sid] == 0) {
                gSMFRETda::pb::p_cap cap;
                cap.set_cap(sz_burst);
                std::string scap;
                cap.SerializeToString(&scap);
                scap = "c" + scap;
                bytes = nn_send(sock, scap.c_str(), scap.length(), 0);
                if (bytes == -1) continue;
                bytes = nn_recv(sock, &rbuf, NN_MSG, 0);
                if (bytes < 0) continue;
                gSMFRETda::pb::p_n sn;
                sn.ParseFromArray(rbuf, bytes);
                nn_freemsg(rbuf);
                s_n[sid] = sn.s_n();
                dataready[sid] = 1;
                continue;
            }
            if (dataready[sid] == 1) {
                gSMFRETda::pb::p_str gpuidStr;
                gpuidStr.set_str(gpuNodeId);
                if (!sent_oHist) {
                    gpuidStr.set_histnum(fretHistNum);
                    sent_oHist = true;
                    for (float o : vOEHist) gpuidStr.add_ohist(o);
                } else {
                    gpuidStr.set_histnum(0);
                }
                std::string sgpuid;
                gpuidStr.SerializeToString(&sgpuid);
                sgpuid = "p" + sgpuid;
                bytes = nn_send(sock, sgpuid.c_str(), sgpuid.length(), 0);
                if (bytes == -1) continue;
                AtomicWriter(debug, debugLevel::cpu) << "p" + gpuNodeId << '\n';
                gSMFRETda::pb::p_ga ga;
                bytes = nn_recv(sock, &rbuf, NN_MSG, 0);
                if (bytes < 0) continue;
                ga.ParseFromArray(rbuf, bytes);
                nn_freemsg(rbuf);
                const int pidx = ga.idx();
                msg_bestcs = ga.bestfv();
                if (pidx == -1) continue;   // the server has nothing for this stream right now
                if (pidx == -2) break;      // the server is done
                params_idx = pidx;
                const int ps_n = s_n[sid] * (s_n[sid] + 1);
                params[sid].resize((size_t)ps_n);
                AtomicWriter(debug, debugLevel::net) << "params_idx recv: " << params_idx << '\n';
                // a short params[] reads as zeros, like protobuf's repeated-field access would not:
                // the reference indexes past the end; here the evaluation is skipped by set_params
                for (int pi = 0; pi < ps_n; pi++) params[sid][(size_t)pi] = pi < ga.params_size() ? ga.params(pi) : 0.0f;
                ga_start[sid] = ga.start();
                ga_stop[sid] = ga.stop();
                // (a host-side field of the stream slot; this thread holds _m[sid] and the GPU thread
                // only touches the slot in states 2 and 3)
                if (epoch_from_idx) gsm_set_epoch(pdamc->handle(), sid, (uint32_t)params_idx + 1u);
                dataready[sid] = 2;
                lck.unlock();
                _cv[sid].notify_one();
                continue;
            }
            using namespace std::chrono_literals;
            if (!_cv[sid].wait_for(lck, 500ms, [this, sid] { return dataready[sid] == 4; })) {
                AtomicWriter(debug, debugLevel::cpu) << dataready[sid] << " dataready !=4\n";
                if (++keepalivecount % (unsigned long)keepalive_every == 0) {
                    gSMFRETda::pb::p_sid pidx;
                    pidx.set_sid(params_idx);
                    std::string spidx;
                    pidx.SerializeToString(&spidx);
                    spidx = "k" + spidx;
                    bytes = nn_send(sock, spidx.c_str(), spidx.length(), 0);
                    if (bytes == -1) continue;
                    if (nn_recv(sock, &rbuf, NN_MSG, 0) >= 0) nn_freemsg(rbuf);
                }
                continue;
            }
            AtomicWriter(debug, debugLevel::cpu) << dataready[sid] << " dataready ==4\n";
            const int R = pdamc->reSampleTimes;
            std::vector<float> vMcHist;
            float chisqr;
            if (host_hist) {
                // the reference's host path, operation for operation (streamWorker.cpp:186-206)
                vMcHist = mkhist(pdamc->hmcE[sid], (size_t)N[sid] * (size_t)R, fretHistNum, 0.0f, 1.0f);
                int effN = fretHistNum - leftPad;
                chisqr = 0;
                for (int ihist = leftPad; ihist < fretHistNum; ihist++) {
                    if (vMcHist[(size_t)ihist] > 0)
                        chisqr += std::pow((float(vOEHist[(size_t)ihist] - vMcHist[(size_t)ihist] / R)), 2) /
                                  (float(vMcHist[(size_t)ihist]) / R);
                    else
                        effN--;
                }
                chisqr = chisqr / (effN - s_n[sid] * (s_n[sid] + 1));
            } else {
                // fused: the kernel's histogram (u64 counts, inner bins first) and chi-square
                const uint64_t* h = pdamc->hHist(sid);
                vMcHist.resize((size_t)fretHistNum);
                for (int i = 0; i < fretHistNum; ++i) vMcHist[(size_t)i] = (float)h[i];
                chisqr = pdamc->hChi2(sid);
            }
            gSMFRETda::pb::res chi2res;
            chi2res.set_idx(gpuNodeId);
            chi2res.set_ridx(params_idx);
            chi2res.set_e(chisqr);
            if (chisqr < msg_bestcs) {
                send_sHist = true;
                for (float s : vMcHist) chi2res.add_shist(s / R);
            } else {
                send_sHist = false;
            }
            chi2res.set_hist(send_sHist);
            std::string sres;
            chi2res.SerializeToString(&sres);
            bytes = nn_send(sock, ("r" + sres).c_str(), sres.length() + 1, 0);
            if (bytes == -1) continue;
            gSMFRETda::pb::p_sid gas;
            AtomicWriter(debug, debugLevel::net) << chisqr << "\t params_idx send: " << params_idx << '\n';
            bytes = nn_recv(sock, &rbuf, NN_MSG, 0);
            if (bytes < 0) continue;
            gas.ParseFromArray(rbuf, bytes);
            AtomicWriter(debug, debugLevel::net) << params_idx << " p_sid recv: " << gas.sid() << '\n';
            if (gas.sid() == -1) ending = true;
            nn_freemsg(rbuf);
            dataready[sid] = 0;
        } while (!ending);
        pdamc->workerNum--;   // tells the gpuWorker loop this stream has left (streamWorker.cpp:240)
        nn_close(sock);
    }

  private:
    mc* pdamc;
    std::vector<float>* SgDivSr;
    int fretHistNum;
    std::string* url;
    std::mutex* _m;
    std::condition_variable* _cv;
    int* dataready;
    int* s_n;
    std::vector<float>* params;
    unsigned char debug;
    int* N;
    int* ga_start;
    int* ga_stop;
};
