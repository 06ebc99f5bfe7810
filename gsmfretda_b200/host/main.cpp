// main.cpp — the worker node `gSMFRETda`: same command line, same start-up sequence and same
// thread structure as the reference's main (/root/reference/src/main.cpp:66-180; options of
// /root/reference/3rdparty/gengetopt/cmdline.ggo:8-21, parsed with getopt_long since gengetopt is
// not part of this build): load the burst file, build `mc`, upload, then one streamWorker
// thread per stream talking to the parameter server and one gpuWorker thread driving the GPU.
//
// Additions (long options only, none changes the defaults):
//   --resample-times R   Monte-Carlo repeats per burst (reference: fixed 4, mc.hpp:88)
//   --mode replay|normal|reference evaluation arithmetic / draw source; default replay
//          (bit-exact against the checker; normal = FP32 fast arithmetic, statistically equivalent, ~1.3x faster);
//          reference = the reference's own Sobol64 + xor128 generators: results equal the reference's bit for bit
//   --seed S             seed of the counter-based generator
//   --epoch-from-idx     draw stream of an evaluation = f(seed, params_idx) instead of the order
//                        in which streams happened to evaluate: results reproducible whatever
//                        stream / GPU / node gets which vector
//   --host-hist          histogram + chi-square on the host from hmcE, the reference's path
//   -p E..,k..,v..       (reference: parsed, unused) evaluate this one vector locally, print
//                        chi-square, exit — no server needed
#include <cassert>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <getopt.h>
#include <iostream>
#include <sstream>
#include <thread>
#include <vector>

#include "gpuWorker.hpp"
#include "loadHdf5.hpp"
#include "streamWorker.hpp"
#include "node_log.hpp"

static void usage()
{
    std::cout <<
        "Usage: gSMFRETda [OPTIONS]... [FILES]...\nCompute smFRET PDA by GPU\n\n"
        "  -h, --help                Print help and exit\n"
        "  -V, --version             Print version and exit\n"
        "  -u, --url=STRING          params server url tcp://ip:port  (default=`tcp://127.0.0.1:7777')\n"
        "  -i, --input=STRING        Input HDF5 filename\n"
        "  -g, --gpuid=INT[,INT...]  The index of the GPU will be used  (default=`0'); a list = one node\n"
        "                            over several GPUs, the bursts split between them\n"
        "  -f, --fret_hist_num=INT   fret hist Number  (default=`200')\n"
        "  -D, --debug               Turn on all debug info  (default=off)\n"
        "  -C, --debugcpu            CPU debug info  (default=off)\n"
        "  -G, --debuggpu            GPU debug info  (default=off)\n"
        "  -N, --debugnet            Network debug info  (default=off)\n"
        "  -K, --debugkernel         GPU Kernel debug info  (default=off)\n"
        "  -q, --gpuinfo             Show GPUs' info  (default=off)\n"
        "  -P, --profiler            Enable CUDA profiler  (default=off)\n"
        "  -s, --snum=INT            Stream Number, when it's 0 automatically determine  (default=`0')\n"
        "  -p, --parameters=FLOAT    Calculate with the parameters. Give them like -p 0.34,0.56,455,673,2,4\n"
        "  -I, --pid=STRING          Appoint a pid file\n"
        "      --resample-times=INT  Monte-Carlo repeats per burst  (default=`4')\n"
        "      --mode=STRING         replay|normal|reference  (default=`replay')\n"
        "      --seed=INT            generator seed\n"
        "      --epoch-from-idx      reproducible draws per parameter index  (default=off)\n"
        "      --host-hist           host histogram and chi-square, as the reference  (default=off)\n"
        "      --max-batch=INT       most vectors per batched p_ga when the node spans GPUs  (default=`64')\n"
        "      --coalesce=INT        evaluate INT or more vectors that are ready together in one launch, 0 = off  (default=`0')\n"
        "      --oversubscribe       allow a GPU to appear twice in -g (test rigs)  (default=off)\n";
}

static void writepid(const std::string& fn)   // main.cpp:53-64
{
    if (FILE* f = std::fopen(fn.c_str(), "a")) {
        std::fputs((std::to_string(getpid()) + '\n').c_str(), f);
        std::fclose(f);
    }
}

static int show_gpus(int only)   // role of showGPUsInfo (mc.cu:26-59) through the C ABI
{
    const int n = gsm_device_count();
    std::printf("%d CUDA device(s)\n", n < 0 ? 0 : n);
    for (int i = 0; i < n; ++i) {
        if (only >= 0 && i != only) continue;
        gsm_ctx* c = nullptr;
        if (gsm_create(i, 1, 0, &c) < 0) {
            std::printf("GPU %d: %s\n", i, gsm_last_error());
            continue;
        }
        char uuid[33] = {0};
        gsm_device_uuid(c, uuid);
        std::printf("GPU %d: uuid %s, %d stream(s) by default\n", i, uuid, gsm_stream_count(c));
        gsm_destroy(c);
    }
    return n;
}

int main(int argc, char* argv[])
{
    std::string url = "tcp://127.0.0.1:7777", input, pidfile, mode = "replay";
    int gpuid = 0, fretHistNum = 200, streamNum = 0, R = 4, max_batch = 64, coalesce = 0;
    std::vector<int> gpu_list;   // -g 0,1,2: ONE node over several GPUs (the dataset is split)
    bool gpuid_given = false, gpuinfo = false, profiler = false, epoch_from_idx = false, host_hist = false, oversubscribe = false;
    unsigned char debuglevel = 0;
    uint32_t seed = 0x9E3779B9u;
    std::vector<float> one_vector;
    static const struct option lo[] = {
        {"help", no_argument, 0, 'h'}, {"version", no_argument, 0, 'V'}, {"url", required_argument, 0, 'u'},
        {"input", required_argument, 0, 'i'}, {"gpuid", required_argument, 0, 'g'},
        {"fret_hist_num", required_argument, 0, 'f'}, {"debug", no_argument, 0, 'D'},
        {"debugcpu", no_argument, 0, 'C'}, {"debuggpu", no_argument, 0, 'G'}, {"debugnet", no_argument, 0, 'N'},
        {"debugkernel", no_argument, 0, 'K'}, {"gpuinfo", no_argument, 0, 'q'}, {"profiler", no_argument, 0, 'P'},
        {"snum", required_argument, 0, 's'}, {"parameters", required_argument, 0, 'p'},
        {"pid", required_argument, 0, 'I'}, {"resample-times", required_argument, 0, 1000},
        {"mode", required_argument, 0, 1001}, {"seed", required_argument, 0, 1002},
        {"epoch-from-idx", no_argument, 0, 1003}, {"host-hist", no_argument, 0, 1004},
        {"oversubscribe", no_argument, 0, 1005}, {"max-batch", required_argument, 0, 1006},
        {"coalesce", required_argument, 0, 1007}, {0, 0, 0, 0}};
    int c;
    while ((c = getopt_long(argc, argv, "hVu:i:g:f:DCGNKqPs:p:I:", lo, nullptr)) != -1) {
        switch (c) {
            case 'h': usage(); return EXIT_SUCCESS;
            case 'V': std::cout << "gSMFRETda 1.0.0\n"; return EXIT_SUCCESS;
            case 'u': url = optarg; break;
            case 'i': input = optarg; break;
            case 'g': {   // one id as in the reference, or a comma-separated list: one node spanning those GPUs
                gpu_list.clear();
                std::stringstream ss(optarg);
                std::string t;
                while (std::getline(ss, t, ',')) if (!t.empty()) gpu_list.push_back(std::atoi(t.c_str()));
                gpuid = gpu_list.empty() ? 0 : gpu_list[0];
                gpuid_given = true;
                break;
            }
            case 'f': fretHistNum = std::atoi(optarg); break;
            case 'D': debuglevel = (unsigned char)~0; break;
            case 'C': debuglevel |= node::kCpu; break;
            case 'G': debuglevel |= node::kGpu; break;
            case 'N': debuglevel |= node::kNet; break;
            case 'K': debuglevel |= node::kKernel; break;
            case 'q': gpuinfo = true; break;
            case 'P': profiler = true; break;
            case 's': streamNum = std::atoi(optarg); break;
            case 'p': {   // comma separated, option may repeat (gengetopt "multiple")
                std::stringstream ss(optarg);
                std::string t;
                while (std::getline(ss, t, ',')) if (!t.empty()) one_vector.push_back(std::strtof(t.c_str(), nullptr));
                break;
            }
            case 'I': pidfile = optarg; break;
            case 1000: R = std::atoi(optarg); break;
            case 1001: mode = optarg; break;
            case 1002: seed = (uint32_t)std::strtoul(optarg, nullptr, 0); break;
            case 1003: epoch_from_idx = true; break;
            case 1004: host_hist = true; break;
            case 1005:   // a GPU listed twice in -g: its parts wait for each other inside the device (gsm_peer_connect)
                oversubscribe = true;
                setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);   // before the first CUDA call
                break;
            case 1006: max_batch = std::atoi(optarg); break;
            case 1007: coalesce = std::atoi(optarg); break;
            default: return EXIT_FAILURE;
        }
    }
    if (!pidfile.empty()) writepid(pidfile);
    if (gpuinfo) {
        show_gpus(gpuid_given ? gpuid : -1);
        return EXIT_SUCCESS;
    }
    if (input.empty() && optind >= argc) {
        std::cout << "You need either appoint -i or add hdf5 filename in the end of cmdline!" << std::endl;
        return EXIT_FAILURE;
    }
    const std::string H5FILE_NAME = input.empty() ? argv[optind] : input;
    if (mode != "replay" && mode != "normal" && mode != "reference") {
        std::cout << "--mode is replay, normal or reference" << std::endl;
        return EXIT_FAILURE;
    }
    std::cout << "Using " << H5FILE_NAME << std::endl;

    std::vector<int64_t> istart, istop, stop, start, times_ms;
    std::vector<unsigned char> mask_ad, mask_dd;
    std::vector<float> T_burst_duration, SgDivSr;
    float clk_p, bg_ad_rate, bg_dd_rate;
    std::uintmax_t hdf5size = loadhdf5(H5FILE_NAME, start, stop, istart, istop, times_ms, mask_ad, mask_dd,
                                       T_burst_duration, SgDivSr, clk_p, bg_ad_rate, bg_dd_rate);
    std::cout << "HD5 size: " << hdf5size << std::endl;
    std::cout << H5FILE_NAME << " loaded." << std::endl;
    if (mask_ad.size() != times_ms.size() || mask_dd.size() != times_ms.size() ||
        T_burst_duration.size() != SgDivSr.size() || start.size() != istart.size() || stop.size() != start.size() ||
        istop.size() != istart.size() || T_burst_duration.size() != start.size()) {
        std::cerr << "the datasets of " << H5FILE_NAME << " do not have matching lengths" << std::endl;
        return EXIT_FAILURE;   // the reference asserts (main.cpp:117-119)
    }
    std::cout << T_burst_duration.size() << std::endl;
    std::vector<int> phCount(start.size());
    int64_t *burst_ad, *burst_dd;
    burst_data(istart, istop, times_ms, mask_ad, mask_dd, phCount, &burst_ad, &burst_dd);

    mc::Options opt;
    opt.keep_mce = host_hist;          // the N*R-float copy per evaluation only when someone reads it
    opt.seed = seed;
    opt.fret_hist_num = fretHistNum;
    opt.left_pad = 2;                  // main.cpp:155
    opt.peer_shared_device = oversubscribe;
    opt.max_batch = max_batch;         // most vectors one batched p_ga may carry when the node spans GPUs
    opt.carriers = coalesce > 0 ? 4 : 0;   // slots for launches that evaluate several streams' vectors at once
    if (gpu_list.empty()) gpu_list.push_back(gpuid);
    if (gpu_list.size() > 1 && host_hist) {
        std::cout << "--host-hist reads per-sample efficiencies of one GPU: not available with a GPU list" << std::endl;
        return EXIT_FAILURE;
    }
    mc pdamc(gpu_list, streamNum, debuglevel, hdf5size, profiler, opt);
    pdamc.getStreams(&streamNum);      // streamNum = 0 asked for the automatic choice
    std::cout << "streamNum: " << streamNum << std::endl;
    if (streamNum <= 0) return EXIT_FAILURE;
    pdamc.set_reSampleTimes_exact(R);
    pdamc.set_mode(mode == "replay" ? GSM_MODE_REPLAY : (mode == "reference" ? GSM_MODE_REFERENCE : GSM_MODE_NORMAL));
    std::cout << "mode: " << mode << (mode == "replay" ? " (explicit-rounding arithmetic on counter-based draws: results are a pure function of data, parameters, seed and draw epoch, bit-equal to the checker)"
                                      : mode == "normal" ? " (FP32 fast arithmetic on the same draws: statistically equivalent to replay, not bit-equal)"
                                                         : " (the reference's own Sobol64/xor128 generators: bit-equal to the reference's GPU run)")
              << "; GPUs:";
    for (int g : gpu_list) std::cout << ' ' << g;
    std::cout << std::endl;
    pdamc.init_data_gpu(istart, start, stop, phCount, (long)times_ms.size(), burst_ad, burst_dd, T_burst_duration,
                        SgDivSr, clk_p, bg_ad_rate, bg_dd_rate);

    if (!one_vector.empty()) {
        // n*n + n numbers: E (n), k (n*n - n), v (n)
        int n = 1;
        while (n * n + n < (int)one_vector.size()) ++n;
        pdamc.set_gpuid();
        pdamc.set_nstates(n, 0);
        if (n * n + n != (int)one_vector.size() || !pdamc.set_params(n, 0, one_vector)) {
            std::cout << "-p needs n*n+n numbers (E, k row-major without the diagonal, v)" << std::endl;
            return EXIT_FAILURE;
        }
        const int N = pdamc.setBurstBd(0, -1, 0);
        pdamc.run_kernel(N, 0);
        while (!pdamc.streamQuery(0)) std::this_thread::yield();
        std::cout.precision(9);
        std::cout << "chi2: " << pdamc.hChi2(0) << std::endl;
        delete[] burst_ad;
        delete[] burst_dd;
        return EXIT_SUCCESS;
    }

    // share_var_init (main.cpp:25-40)
    std::mutex* _m = new std::mutex[streamNum]();
    std::condition_variable* _cv = new std::condition_variable[streamNum]();
    int* s_n = new int[streamNum]();
    std::vector<float>* params = new std::vector<float>[streamNum]();
    int* ga_start = new int[streamNum]();
    int* ga_stop = new int[streamNum]();
    int* N = new int[streamNum]();     // the reference leaves it uninitialised
    int* dataready = new int[streamNum]();
    for (int i = 0; i < streamNum; ++i) N[i] = -2;

    streamWorker worker(&pdamc, &url, &SgDivSr, fretHistNum, _m, _cv, dataready, s_n, params, ga_start, ga_stop, N,
                        debuglevel);
    worker.host_hist = host_hist;
    worker.epoch_from_idx = epoch_from_idx;
    gpuWorker gpuworker(&pdamc, streamNum, &SgDivSr, fretHistNum, _m, _cv, dataready, s_n, params, ga_start, ga_stop,
                        N, debuglevel);
    if (coalesce > 0 && pdamc.carriers() == 0)
        std::cout << "--coalesce is off: " << (mode == "reference" ? "reference mode evaluates one vector per launch"
                                                                    : "per-sample outputs (--host-hist) are kept per stream") << std::endl;
    gpuworker.set_coalesce(coalesce);
    pdamc.get_max_conversion_capacity();
    std::cout << "workers created\n";
    std::vector<std::thread> threads;
    for (int i = 0; i < streamNum; i++) threads.push_back(std::thread(&streamWorker::run, &worker, i, pdamc.sz_burst, 2));
    std::cout << "net threads looped\n";
    std::thread thGpu(&gpuWorker::run, &gpuworker, pdamc.sz_burst);
    std::cout << "gpu thread looped\n";
    for (auto& th : threads) th.join();
    std::cout << "net threads joined\n";
    thGpu.join();
    std::cout << "gpu thread joined\n";
    delete[] burst_ad;
    delete[] burst_dd;
    delete[] _m; delete[] _cv; delete[] s_n; delete[] params; delete[] ga_start; delete[] ga_stop; delete[] N;
    delete[] dataready;
    return EXIT_SUCCESS;
}
