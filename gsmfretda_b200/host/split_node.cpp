// split_node.cpp — ONE dataset evaluated by all the GPUs of a box from one process (SURVEY.md
// §8e "burst split"; the reference can only give different vectors to different GPUs,
// /root/reference/src/call_main.cpp:195-227).
//
//   split_node <bursts.bin> <params.txt> <out.txt> [parts=#GPUs] [reSampleTimes=4] [seed=1]
//   split_node --plan <bursts.bin> <parts>      print the burst ranges only (no GPU needed)
//
// The bursts are cut into `parts` contiguous ranges balanced by photon count (the rule of
// gsmfretda_b200/sharding.py); part r lives in a context of its own on GPU r % #GPUs with the
// global burst id of its first burst (draws are keyed by global burst id), the contexts are
// connected with gsm_peer_export / gsm_peer_connect, and every parameter vector is launched once
// on every part: the last block of each launch stores its partial histogram into every part's
// inbox over NVLink (or through the device's own memory when parts share a GPU) and finishes
// chi-square itself.  Every part ends up with the same global histogram; part 0's is written.
// Input/output formats are local_node's.
#include <chrono>
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "mc.hpp"

struct Job { int n; std::vector<float> args; };

template <class T>
static bool rd(FILE* f, std::vector<T>& v, size_t n) { v.resize(n); return n == 0 || fread(v.data(), sizeof(T), n, f) == n; }

// contiguous ranges balanced by sum(phCount + 4): sharding.split_bursts
static std::vector<int64_t> split_bursts(const std::vector<int>& ph, int parts)
{
    const int64_t B = (int64_t)ph.size();
    std::vector<int64_t> cost(B);
    int64_t acc = 0;
    for (int64_t i = 0; i < B; ++i) { acc += (int64_t)ph[i] + 4; cost[i] = acc; }
    std::vector<int64_t> cuts(parts + 1, 0);
    for (int r = 1; r < parts; ++r) {
        const double target = (double)acc * r / parts;
        int64_t lo = 0, hi = B;                       // first i with cost[i] >= target
        while (lo < hi) { const int64_t mid = (lo + hi) / 2; if ((double)cost[mid] >= target) hi = mid; else lo = mid + 1; }
        cuts[r] = B ? lo : 0;
    }
    cuts[parts] = B;
    for (int r = 1; r <= parts; ++r) if (cuts[r] < cuts[r - 1]) cuts[r] = cuts[r - 1];
    return cuts;
}

static int plan_only(const char* path, int parts)
{
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::perror(path); return 1; }
    char magic[4]; uint32_t ver; int64_t B, Nph; float hdr[4];
    if (fread(magic, 1, 4, f) != 4 || std::memcmp(magic, "GSMB", 4) || fread(&ver, 4, 1, f) != 1 || ver != 1 ||
        fread(&B, 8, 1, f) != 1 || fread(&Nph, 8, 1, f) != 1 || fread(hdr, 4, 4, f) != 4) { std::fprintf(stderr, "bad header\n"); return 1; }
    std::vector<int> ph;
    if (std::fseek(f, (long)(3 * B * 8), SEEK_CUR) != 0 || !rd(f, ph, B)) { std::fprintf(stderr, "short file\n"); return 1; }
    std::fclose(f);
    const std::vector<int64_t> cuts = split_bursts(ph, parts);
    for (int r = 0; r < parts; ++r) std::printf("%" PRId64 " %" PRId64 "\n", cuts[r], cuts[r + 1]);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 4 && std::strcmp(argv[1], "--plan") == 0) return plan_only(argv[2], std::atoi(argv[3]));
    if (argc < 4) { std::fprintf(stderr, "usage: %s bursts.bin params.txt out.txt [parts] [R] [seed]\n", argv[0]); return 2; }
    const int ngpu = gsm_device_count();
    if (ngpu < 1) { std::fprintf(stderr, "no sm_100 device: this node has no CPU path\n"); return 1; }
    const int parts = argc > 4 ? std::atoi(argv[4]) : ngpu;
    const int R = argc > 5 ? std::atoi(argv[5]) : 4;
    const uint32_t seed = argc > 6 ? (uint32_t)std::strtoul(argv[6], nullptr, 10) : 1u;
    if (parts < 1 || parts > GSM_MAX_PEERS) { std::fprintf(stderr, "parts must be 1..%d\n", GSM_MAX_PEERS); return 2; }

    FILE* f = std::fopen(argv[1], "rb");
    if (!f) { std::perror(argv[1]); return 1; }
    char magic[4]; uint32_t ver; int64_t B, Nph; float hdr[4];
    if (fread(magic, 1, 4, f) != 4 || std::memcmp(magic, "GSMB", 4) || fread(&ver, 4, 1, f) != 1 || ver != 1 ||
        fread(&B, 8, 1, f) != 1 || fread(&Nph, 8, 1, f) != 1 || fread(hdr, 4, 4, f) != 4) { std::fprintf(stderr, "bad header\n"); return 1; }
    std::vector<int64_t> start, stop, istart, ad, dd; std::vector<int> ph; std::vector<float> T, sg;
    if (!rd(f, start, B) || !rd(f, stop, B) || !rd(f, istart, B) || !rd(f, ph, B) || !rd(f, T, B) || !rd(f, sg, B) ||
        !rd(f, ad, Nph) || !rd(f, dd, Nph)) { std::fprintf(stderr, "short file\n"); return 1; }
    std::fclose(f);
    const float clk_p = hdr[0], bg_ad = hdr[1], bg_dd = hdr[2];

    std::vector<Job> jobs;
    { std::ifstream in(argv[2]); std::string line;
      while (std::getline(in, line)) { std::istringstream ss(line); Job j; if (!(ss >> j.n)) continue; float x; while (ss >> x) j.args.push_back(x); jobs.push_back(j); } }

    // one context per part; gsm_init_data takes the part's arrays with istart rebased and the
    // global id of its first burst
    mc::Options opt; opt.keep_mce = false; opt.seed = seed;
    const int nb = opt.fret_hist_num, nbp = nb + 2;
    const std::vector<int64_t> cuts = split_bursts(ph, parts);
    std::vector<std::unique_ptr<mc>> ctx;
    std::vector<float> ohist(nb, 0.0f), part_hist(nb);
    for (int r = 0; r < parts; ++r) {
        const int64_t b0 = cuts[r], b1 = cuts[r + 1];
        const int64_t p0 = b0 < B ? istart[b0] : Nph, p1 = b1 < B ? istart[b1] : Nph;
        std::vector<int64_t> is(istart.begin() + b0, istart.begin() + b1);
        for (auto& v : is) v -= p0;
        ctx.emplace_back(new mc(r % ngpu, 1, 0, 0, false, opt));
        mc& m = *ctx.back();
        m.set_reSampleTimes_exact(R);
        GSM_CHECK_RETURN(gsm_init_data(m.handle(), b1 - b0, is.data(), start.data() + b0, stop.data() + b0, ph.data() + b0,
                                       p1 - p0, ad.data() + p0, dd.data() + p0, T.data() + b0, sg.data() + b0, clk_p,
                                       bg_ad, bg_dd, b0));
        m.sz_burst = (int)(b1 - b0);
        GSM_CHECK_RETURN(gsm_observed_hist(m.handle(), part_hist.data(), nb));
        for (int i = 0; i < nb; ++i) ohist[i] += part_hist[i];    // counts: exact in float
    }
    std::vector<unsigned char> handles((size_t)parts * GSM_PEER_HANDLE_BYTES);
    for (int r = 0; r < parts; ++r) ctx[r]->peer_export(parts, 1, handles.data() + (size_t)r * GSM_PEER_HANDLE_BYTES);
    for (int r = 0; r < parts; ++r) {
        GSM_CHECK_RETURN(gsm_set_observed_hist(ctx[r]->handle(), ohist.data(), nb));   // chi-square needs the global one
        ctx[r]->peer_connect(handles.data(), r, parts);
    }

    std::ofstream out(argv[3]);
    out.precision(9);
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t ji = 0; ji < jobs.size(); ++ji) {
        for (int r = 0; r < parts; ++r) {              // the six calls of gpuWorker.cpp:71-80 on every part
            mc& m = *ctx[r];
            m.set_gpuid();
            m.set_nstates(jobs[ji].n, 0);
            if (!m.set_params(jobs[ji].n, 0, jobs[ji].args)) { std::fprintf(stderr, "vector %zu: wrong argument count\n", ji); return 1; }
            const int N = m.setBurstBd(0, -1, 0);
            GSM_CHECK_RETURN(gsm_set_epoch(m.handle(), 0, (uint32_t)(1000 + ji)));
            m.run_kernel(N, 0);
        }
        for (int r = 0; r < parts; ++r) while (!ctx[r]->streamQuery(0)) {}
        for (int r = 0; r < parts; ++r)
            if (ctx[r]->hFlags(0)) { std::fprintf(stderr, "vector %zu part %d: result flags %llu\n", ji, r, (unsigned long long)ctx[r]->hFlags(0)); return 1; }
        for (int r = 1; r < parts; ++r)                  // every part holds the same bits
            if (std::memcmp(ctx[r]->hHist(0), ctx[0]->hHist(0), nbp * sizeof(uint64_t)) != 0 || ctx[r]->hChi2(0) != ctx[0]->hChi2(0)) {
                std::fprintf(stderr, "vector %zu: part %d disagrees with part 0\n", ji, r); return 1; }
        out << ji << " 0 " << 1000 + ji << ' ' << ctx[0]->hChi2(0);
        const uint64_t* h = ctx[0]->hHist(0);
        for (int i = 0; i < nbp; ++i) out << ' ' << h[i];
        out << '\n';
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%zu evaluations of %" PRId64 " bursts x %d repeats split into %d parts on %d GPU(s) in %.3f s: %.3g burst-samples/s, %.1f param-evals/s\n",
                jobs.size(), B, R, parts, ngpu < parts ? ngpu : parts, sec, (double)jobs.size() * B * R / sec, jobs.size() / sec);
    for (auto& m : ctx) m->peer_disconnect();
    return 0;
}
