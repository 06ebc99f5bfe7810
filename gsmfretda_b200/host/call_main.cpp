// call_main.cpp — `call_gSMFRETda`: one worker-node process per GPU of this machine, same
// command line as the reference's launcher (/root/reference/src/call_main.cpp:160-260, options
// of /root/reference/3rdparty/gengetopt/call_gSMFRETda.ggo:8-13): starts `gSMFRETda` (found
// next to this executable) once per selected GPU with -g <id>, forwards -u/-i/-f/-s and the pid
// file, waits for all of them, kills them on SIGINT, and returns the first non-zero exit code.
// The population of a GA server is thereby sharded across the GPUs with no communication
// between them (SURVEY.md §8e "population" mode).  The node's own long options
// (--mode, --resample-times, --seed, --epoch-from-idx, --host-hist) are passed through.
#include <algorithm>
#include <chrono>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <getopt.h>
#include <iostream>
#include <string>
#include <sys/types.h>
#include <sys/wait.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include "../../include/gsmfret_c.h"

namespace {
volatile std::sig_atomic_t gSignalStatus = 0;
void signal_handler(int s) { gSignalStatus = s; }

std::string self_dir()
{
    char buf[4096];
    ssize_t n = ::readlink("/proc/self/exe", buf, sizeof buf - 1);
    if (n <= 0) return ".";
    buf[n] = 0;
    std::string p(buf);
    const size_t k = p.rfind('/');
    return k == std::string::npos ? "." : p.substr(0, k);
}
}  // namespace

int main(int argc, char** argv)
{
    std::signal(SIGINT, signal_handler);
    std::string url, input, fret, snum, pidfile = "daGPUs.pid";
    std::vector<int> gpuids;
    std::vector<std::string> passthrough;
    static const struct option lo[] = {
        {"help", no_argument, 0, 'h'}, {"version", no_argument, 0, 'V'}, {"url", required_argument, 0, 'u'},
        {"input", required_argument, 0, 'i'}, {"fret_hist_num", required_argument, 0, 'f'},
        {"snum", required_argument, 0, 's'}, {"pid", required_argument, 0, 'I'}, {"gpuids", required_argument, 0, 'g'},
        {"resample-times", required_argument, 0, 1000}, {"mode", required_argument, 0, 1001},
        {"seed", required_argument, 0, 1002}, {"epoch-from-idx", no_argument, 0, 1003},
        {"host-hist", no_argument, 0, 1004}, {"node", required_argument, 0, 1005}, {0, 0, 0, 0}};
    std::string node = self_dir() + "/gSMFRETda";
    int c;
    while ((c = getopt_long(argc, argv, "hVu:i:f:s:I:g:", lo, nullptr)) != -1) {
        switch (c) {
            case 'h':
                std::cout << "Usage: call_gSMFRETda [OPTIONS]... [FILES]...\nCompute wapper smFRET PDA by GPU\n\n"
                             "  -u, --url=STRING         params server url tcp://ip:port\n"
                             "  -i, --input=STRING       Input HDF5 filename\n"
                             "  -f, --fret_hist_num=INT  fret hist Number\n"
                             "  -s, --snum=INT           Stream Number, when it's 0 automatically determine\n"
                             "  -I, --pid=STRING         Appoint a pid file  (default=`daGPUs.pid')\n"
                             "  -g, --gpuids=INT         The indices of the GPUs will be used. Give -1 to use All GPUs on\n"
                             "                           the node. Or like '0,2,3' to use GPU 0 2 3.\n"
                             "      --node=PATH          the node executable (default: gSMFRETda next to this file)\n";
                return EXIT_SUCCESS;
            case 'V': std::cout << "call_gSMFRETda 1.0.0\n"; return EXIT_SUCCESS;
            case 'u': url = optarg; break;
            case 'i': input = optarg; break;
            case 'f': fret = optarg; break;
            case 's': snum = optarg; break;
            case 'I': pidfile = optarg; break;
            case 'g': {
                char* s = optarg;
                while (*s) {
                    gpuids.push_back((int)std::strtol(s, &s, 10));
                    while (*s == ',' || *s == ' ') ++s;
                }
                break;
            }
            case 1000: passthrough.push_back(std::string("--resample-times=") + optarg); break;
            case 1001: passthrough.push_back(std::string("--mode=") + optarg); break;
            case 1002: passthrough.push_back(std::string("--seed=") + optarg); break;
            case 1003: passthrough.push_back("--epoch-from-idx"); break;
            case 1004: passthrough.push_back("--host-hist"); break;
            case 1005: node = optarg; break;
            default: return EXIT_FAILURE;
        }
    }
    if (input.empty() && optind >= argc) {   // call_main.cpp:46-49
        std::cout << "You need either appoint -i or add hdf5 filename in the end of cmdline!" << std::endl;
        return EXIT_FAILURE;
    }
    if (input.empty()) input = argv[optind];
    std::vector<std::string> cmd{node};
    if (!url.empty()) { cmd.push_back("-u"); cmd.push_back(url); }
    cmd.push_back("-i"); cmd.push_back(input);
    if (!fret.empty()) { cmd.push_back("-f"); cmd.push_back(fret); }
    if (!snum.empty()) { cmd.push_back("-s"); cmd.push_back(snum); }
    if (std::remove(pidfile.c_str()) != 0) std::perror("Error deleting file");   // as the reference: not fatal
    cmd.push_back("-I"); cmd.push_back(pidfile);
    for (auto& p : passthrough) cmd.push_back(p);

    const bool useAll = gpuids.empty() || *std::min_element(gpuids.begin(), gpuids.end()) == -1;
    const int nDevices = gsm_device_count();
    if (nDevices <= 0) {
        std::fprintf(stderr, "no CUDA device: %s\n", gsm_last_error());
        return EXIT_FAILURE;
    }
    if (useAll) {
        gpuids.clear();
        for (int i = 0; i < nDevices; ++i) gpuids.push_back(i);
    } else {
        std::sort(gpuids.begin(), gpuids.end());
        gpuids.erase(std::unique(gpuids.begin(), gpuids.end()), gpuids.end());
    }
    std::vector<pid_t> pids;
    for (int g : gpuids) {
        if (g < 0 || g >= nDevices) continue;
        pid_t pid = ::fork();
        if (pid == 0) {
            std::vector<std::string> cv = cmd;
            cv.push_back("-g");
            cv.push_back(std::to_string(g));
            std::vector<char*> av;
            for (auto& s : cv) av.push_back(const_cast<char*>(s.c_str()));
            av.push_back(nullptr);
            ::execv(av[0], av.data());
            std::fprintf(stderr, "an error occurred in execv %s: %s\n", av[0], std::strerror(errno));
            std::_Exit(EXIT_FAILURE);
        }
        if (pid > 0) {
            std::cout << "pid: " << pid << " gpu " << g << std::endl;
            pids.push_back(pid);
        }
    }
    int retcode = EXIT_SUCCESS;
    size_t running = pids.size();
    std::vector<bool> done(pids.size(), false);
    while (running > 0) {
        if (gSignalStatus == SIGINT) {
            std::printf("SIGINT got\n");
            for (size_t i = 0; i < pids.size(); ++i) if (!done[i]) ::kill(pids[i], SIGKILL);
        }
        for (size_t i = 0; i < pids.size(); ++i) {
            if (done[i]) continue;
            int st = 0;
            pid_t w = ::waitpid(pids[i], &st, WNOHANG);
            if (w > 0) {
                done[i] = true;
                --running;
                const int code = WIFEXITED(st) ? WEXITSTATUS(st) : EXIT_FAILURE;
                std::printf("pid %d exit code:%d\n", (int)w, code);
                if (code != EXIT_SUCCESS) retcode = code;
            }
        }
        if (running > 0) std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    return retcode;
}
