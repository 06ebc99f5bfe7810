// split_node.cpp — ONE dataset evaluated on several GPUs from ONE process, without a server:
// the multi-GPU `mc` of host/mc.hpp (one library context per GPU, contiguous burst ranges
// balanced by photon count, histograms meeting inside the evaluation kernels) driven by the six
// calls of gpuWorker (gpuWorker.cpp:71-80) for a list of parameter vectors read from a file.
// The worker node does the same behind `gSMFRETda -g 0,1,...` (main.cpp); this program is the
// server-less harness of tests/test_host_cpp.py.
//
//   split_node <bursts.bin> <params.txt> <out.txt> [parts=#GPUs] [reSampleTimes=4] [seed=1] [--oversubscribe]
//   split_node --plan <bursts.bin> <parts>      print the burst ranges only (no GPU needed)
//
// `parts` beyond the number of GPUs would put several parts on one GPU, whose kernels would then
// wait for each other inside the device (include/gsmfret_c.h, gsm_peer_connect): the parts that
// share a GPU are merged instead (parts = #GPUs) — the result does not depend on the number of
// parts.  --oversubscribe keeps them apart (the test-suite's way to exercise the exchange on a
// one-GPU box): the contexts are created with GSM_PEER_SHARED_DEVICE and the process asks for one
// hardware queue per part before CUDA starts.
// File formats: host/local_node.cpp.  out.txt: per vector "idx 0 epoch chi2 hist[0..nb+1]".
#include <chrono>
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "mc.hpp"

struct Job { int n; std::vector<float> args; };

template <class T>
static bool rd(FILE* f, std::vector<T>& v, size_t n) { v.resize(n); return n == 0 || fread(v.data(), sizeof(T), n, f) == n; }

static int plan_only(const char* path, int parts)
{
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::perror(path); return 1; }
    char magic[4]; uint32_t ver; int64_t B, Nph; float hdr[4];
    if (fread(magic, 1, 4, f) != 4 || std::memcmp(magic, "GSMB", 4) || fread(&ver, 4, 1, f) != 1 || ver != 1 ||
        fread(&B, 8, 1, f) != 1 || fread(&Nph, 8, 1, f) != 1 || fread(hdr, 4, 4, f) != 4) { std::fprintf(stderr, "bad header\n"); return 1; }
    std::vector<int64_t> skip; std::vector<int> ph;
    if (!rd(f, skip, B) || !rd(f, skip, B) || !rd(f, skip, B) || !rd(f, ph, B)) { std::fprintf(stderr, "short file\n"); return 1; }
    std::fclose(f);
    const std::vector<int64_t> cuts = mc::plan_split(ph, parts);
    for (int r = 0; r < parts; ++r) std::printf("%" PRId64 " %" PRId64 "\n", cuts[r], cuts[r + 1]);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 4 && std::strcmp(argv[1], "--plan") == 0) return plan_only(argv[2], std::atoi(argv[3]));
    bool oversubscribe = false;
    if (argc > 1 && std::strcmp(argv[argc - 1], "--oversubscribe") == 0) { oversubscribe = true; --argc; }
    if (argc < 4) { std::fprintf(stderr, "usage: %s bursts.bin params.txt out.txt [parts] [R] [seed] [--oversubscribe]\n", argv[0]); return 2; }
    if (oversubscribe) setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);   // before the first CUDA call
    const int ngpu = gsm_device_count();
    if (ngpu < 1) { std::fprintf(stderr, "no sm_100 device: this node has no CPU path\n"); return 1; }
    int parts = argc > 4 ? std::atoi(argv[4]) : ngpu;
    const int R = argc > 5 ? std::atoi(argv[5]) : 4;
    const uint32_t seed = argc > 6 ? (uint32_t)std::strtoul(argv[6], nullptr, 10) : 1u;
    if (parts < 1 || parts > GSM_MAX_PEERS) { std::fprintf(stderr, "parts must be 1..%d\n", GSM_MAX_PEERS); return 2; }
    const int asked = parts;
    if (parts > ngpu && !oversubscribe) parts = ngpu;   // parts that would share a GPU are one part

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
      while (std::getline(in, line)) { std::istringstream ss(line); Job j; if (!(ss >> j.n)) continue; float x; while (ss >> x) j.args.push_back(x); jobs.push_back(j); } }

    mc::Options opt; opt.keep_mce = false; opt.seed = seed; opt.peer_shared_device = oversubscribe;
    const int nbp = opt.fret_hist_num + 2;
    std::vector<int> ids;
    for (int r = 0; r < parts; ++r) ids.push_back(r % ngpu);
    mc engine(ids, 1, 0, 0, false, opt);
    engine.set_reSampleTimes_exact(R);
    engine.init_data_gpu(istart, start, stop, ph, (long)Nph, ad.data(), dd.data(), T, sg, clk_p, bg_ad, bg_dd);

    std::ofstream out(argv[3]);
    out.precision(9);
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t ji = 0; ji < jobs.size(); ++ji) {
        engine.set_gpuid();
        engine.set_nstates(jobs[ji].n, 0);
        if (!engine.set_params(jobs[ji].n, 0, jobs[ji].args)) { std::fprintf(stderr, "vector %zu: wrong argument count\n", ji); return 1; }
        const int N = engine.setBurstBd(0, -1, 0);
        engine.set_epoch(0, (uint32_t)(1000 + ji));
        engine.run_kernel(N, 0);
        while (!engine.streamQuery(0)) {}
        if (engine.hFlags(0)) { std::fprintf(stderr, "vector %zu: result flags %llu\n", ji, (unsigned long long)engine.hFlags(0)); return 1; }
        for (int r = 1; r < engine.gpus(); ++r)          // every part holds the same bits
            if (std::memcmp(gsm_result_hist(engine.part(r), 0, 0), engine.hHist(0), nbp * sizeof(uint64_t)) != 0 ||
                gsm_result_chi2(engine.part(r), 0) != engine.hChi2(0)) {
                std::fprintf(stderr, "vector %zu: part %d disagrees with part 0\n", ji, r); return 1; }
        out << ji << " 0 " << 1000 + ji << ' ' << engine.hChi2(0);
        const uint64_t* h = engine.hHist(0);
        for (int i = 0; i < nbp; ++i) out << ' ' << h[i];
        out << '\n';
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%zu evaluations of %" PRId64 " bursts x %d repeats split into %d part(s) (%d asked) on %d GPU(s) in %.3f s: %.3g burst-samples/s, %.1f param-evals/s\n",
                jobs.size(), B, R, parts, asked, ngpu < parts ? ngpu : parts, sec, (double)jobs.size() * B * R / sec, jobs.size() / sec);
    return 0;
}
