// local_node.cpp — a worker node without the network: the reference's thread structure
// (main.cpp:129-160: one parameter thread per stream + one gpuWorker thread around one `mc`)
// with the parameter server replaced by a list of vectors read from a file.  It exercises the
// reference-shaped host API end to end — share_var_init's arrays, the dataready state machine,
// the six mc calls per evaluation — and reports chi-square per vector from the fused kernel.
//
//   local_node <bursts.bin> <params.txt> <out.txt> [streams=4] [reSampleTimes=4] [seed] [coalesce=0] [spin=0] [target=16] [linger_us=0] [batch=1]
//
// batch = P > 1: every parameter thread asks for P vectors at a time and they are evaluated by ONE launch
// (mc::set_params_batch, the batched p_ga of streamWorker.hpp); the output is the same, bit for bit.
//
// coalesce = K > 0: the gpuWorker thread evaluates K or more vectors that are ready at the same moment
// with ONE launch (gpuWorker::set_coalesce); the output is the same, bit for bit.
//
// bursts.bin (little endian), written by gsmfretda_b200.synth.save_bin:
//   "GSMB" u32 version=1, i64 B, i64 Nph, f32 clk_p, f32 bg_ad, f32 bg_dd, f32 pad,
//   i64 start[B], i64 stop[B], i64 istart[B], i32 phCount[B], f32 T[B], f32 SgDivSr[B],
//   i64 burst_ad[Nph], i64 burst_dd[Nph]
// params.txt: one vector per line: n then n*n+n floats.
// out.txt: per vector "idx sid epoch chi2 hist[0..nb+1]".
#include <algorithm>
#include <chrono>
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

// GSM_GPUWORKER_HEADER: the drop-in proof (host/Makefile `dropin`, tests/test_host_cpp.py) points
// this at a scratch copy of the REFERENCE's gpuWorker.hpp, whose gpuWorker.cpp is then compiled
// unchanged against host/mc.hpp and linked in the place of this node's own gpuWorker
#ifndef GSM_GPUWORKER_HEADER
#define GSM_GPUWORKER_HEADER "gpuWorker.hpp"
#define GSM_NODE_OWN_GPUWORKER 1
#endif
#include GSM_GPUWORKER_HEADER

struct Job { int idx; int n; std::vector<float> args; };

template <class T>
static bool rd(FILE* f, std::vector<T>& v, size_t n) { v.resize(n); return n == 0 || fread(v.data(), sizeof(T), n, f) == n; }

int main(int argc, char** argv)
{
    if (argc < 4) { std::fprintf(stderr, "usage: %s bursts.bin params.txt out.txt [streams] [R] [seed] [coalesce] [spin] [target] [linger_us] [batch]\n", argv[0]); return 2; }
    const int nstreams = argc > 4 ? std::atoi(argv[4]) : 4;
    const int R = argc > 5 ? std::atoi(argv[5]) : 4;
    const uint32_t seed = argc > 6 ? (uint32_t)std::strtoul(argv[6], nullptr, 10) : 1u;
    const int coalesce = argc > 7 ? std::atoi(argv[7]) : 0;
    const int spin_looks = argc > 8 ? std::atoi(argv[8]) : 0;   // looks at the result flag before sleeping on the condition variable
    const int batch = std::max(1, argc > 11 ? std::atoi(argv[11]) : 1);   // vectors per request of a parameter thread

    FILE* f = std::fopen(argv[1], "rb");
    if (!f) { std::perror(argv[1]); return 1; }
    char magic[4]; uint32_t ver; int64_t B, Nph; float hdr[4];
    if (fread(magic, 1, 4, f) != 4 || std::memcmp(magic, "GSMB", 4) || fread(&ver, 4, 1, f) != 1 || ver != 1 ||
        fread(&B, 8, 1, f) != 1 || fread(&Nph, 8, 1, f) != 1 || fread(hdr, 4, 4, f) != 4) { std::fprintf(stderr, "bad header\n"); return 1; }
    std::vector<int64_t> start, stop, istart, ad, dd; std::vector<int> ph; std::vector<float> T, sg;
    if (!rd(f, start, B) || !rd(f, stop, B) || !rd(f, istart, B) || !rd(f, ph, B) || !rd(f, T, B) || !rd(f, sg, B) ||
        !rd(f, ad, Nph) || !rd(f, dd, Nph)) { std::fprintf(stderr, "short file\n"); return 1; }
    std::fclose(f);
    float clk_p = hdr[0], bg_ad = hdr[1], bg_dd = hdr[2];

    std::vector<Job> jobs;
    { std::ifstream in(argv[2]); std::string line;
      while (std::getline(in, line)) { std::istringstream ss(line); Job j; if (!(ss >> j.n)) continue; float x; while (ss >> x) j.args.push_back(x); j.idx = (int)jobs.size(); jobs.push_back(j); } }

    mc::Options opt; opt.keep_mce = false; opt.seed = seed;
    opt.carriers = coalesce > 0 ? 4 : 0;
    mc pdamc(0, nstreams, 0, 0, false, opt);
    int streamNum = 0; pdamc.getStreams(&streamNum);
    pdamc.set_reSampleTimes_exact(R);
    pdamc.init_data_gpu(istart, start, stop, ph, (long)Nph, ad.data(), dd.data(), T, sg, clk_p, bg_ad, bg_dd);
    pdamc.get_max_conversion_capacity();

    // share_var_init (main.cpp:25-40)
    std::mutex* _m = new std::mutex[streamNum];
    std::condition_variable* _cv = new std::condition_variable[streamNum];
    int* s_n = new int[streamNum](); int* ga_start = new int[streamNum](); int* ga_stop = new int[streamNum]();
    int* N = new int[streamNum]; int* dataready = new int[streamNum]();
    for (int i = 0; i < streamNum; ++i) N[i] = -2;
    std::vector<float>* params = new std::vector<float>[streamNum];
    const int nbp = opt.fret_hist_num + 2;

    std::atomic<int> next{0};
    std::vector<std::string> lines(jobs.size());
    auto param_thread = [&](int sid) {      // the role of streamWorker::run without the socket
        for (;;) {
            // `batch` consecutive vectors per request (the batched p_ga of streamWorker.hpp); runs of one
            // model inside the claim go out as one launch each, vector p with epoch 1000 + its index
            const int j0 = next.fetch_add(batch);
            if (j0 >= (int)jobs.size()) break;
            const int j1 = std::min((int)jobs.size(), j0 + batch);
            for (int ja = j0; ja < j1;) {
                int jb = ja + 1;
                while (jb < j1 && jobs[jb].n == jobs[ja].n && jobs[jb].args.size() == jobs[ja].args.size()) ++jb;
                { std::unique_lock<std::mutex> lck(_m[sid]);
                  s_n[sid] = jobs[ja].n; ga_start[sid] = 0; ga_stop[sid] = -1;
                  params[sid].clear();
                  for (int j = ja; j < jb; ++j) params[sid].insert(params[sid].end(), jobs[j].args.begin(), jobs[j].args.end());
                  pdamc.set_epoch(sid, (uint32_t)(1000 + ja));
                  dataready[sid] = 2; }
                _cv[sid].notify_one();
                // a short look before going to sleep: an evaluation at this scale takes tens of microseconds,
                // a futex sleep + wake costs the gpuWorker thread more than that
                // (dataready is a plain int the gpuWorker thread writes under the mutex, as in the reference: the look
                // takes the mutex too, when it is free, instead of reading the int behind its back)
                for (int spin = 0; spin < spin_looks; ++spin) {
                    bool done = false;
                    if (_m[sid].try_lock()) { done = dataready[sid] == 4; _m[sid].unlock(); }
                    if (done) break;
                    std::this_thread::yield();
                }
                { std::unique_lock<std::mutex> lck(_m[sid]);
                  _cv[sid].wait(lck, [&] { return dataready[sid] == 4; });
                  for (int j = ja; j < jb; ++j) {
                      std::ostringstream os; os.precision(9);
                      os << j << ' ' << sid << ' ' << 1000 + j << ' ' << pdamc.hChi2(sid, j - ja);
                      const uint64_t* h = pdamc.hHist(sid, j - ja);
                      for (int i = 0; i < nbp; ++i) os << ' ' << h[i];
                      lines[j] = os.str();
                  }
                  dataready[sid] = 0; }
                ja = jb;
            }
        }
        pdamc.workerNum--;                    // streamWorker.cpp:239-245
    };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int sid = 0; sid < streamNum; ++sid) th.emplace_back(param_thread, sid);
    gpuWorker gw(&pdamc, streamNum, &sg, opt.fret_hist_num, _m, _cv, dataready, s_n, params, ga_start, ga_stop, N, 0);
#ifdef GSM_NODE_OWN_GPUWORKER
    gw.set_coalesce(coalesce, argc > 9 ? std::atoi(argv[9]) : 16, argc > 10 ? std::atoi(argv[10]) : 0);
#endif
    std::thread tg(&gpuWorker::run, &gw, (int)B);
    for (auto& t : th) t.join();
    tg.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::ofstream out(argv[3]);
    for (auto& l : lines) out << l << '\n';
    std::printf("%zu evaluations of %" PRId64 " bursts x %d repeats on %d streams in %.3f s: %.3g burst-samples/s, %.1f param-evals/s\n",
                jobs.size(), B, R, streamNum, sec, (double)jobs.size() * B * R / sec, jobs.size() / sec);
#ifdef GSM_NODE_OWN_GPUWORKER
    if (coalesce > 0) {
        const gpuWorker::Stats& st = gw.stats();
        std::printf("coalescing: %llu passes, %llu group launches carrying %llu vectors (%.1f each), %llu single launches\n", st.passes,
                    st.group_launches, st.grouped_vectors, st.group_launches ? (double)st.grouped_vectors / st.group_launches : 0.0,
                    st.single_launches);
    }
#endif
    delete[] _m; delete[] _cv; delete[] s_n; delete[] ga_start; delete[] ga_stop; delete[] N; delete[] dataready; delete[] params;
    return 0;
}
