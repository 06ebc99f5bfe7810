// wire_probe.cpp — command-line access to the node's wire layer for the test-suite
// (tests/test_wire_protocol.py): the args.proto codec of args_pb.hpp and the SP REQ client of
// sp_req.hpp, exactly the code streamWorker.hpp uses.
//
//   wire_probe encode <message> [field=value ...]     -> hex of SerializeToString
//        repeated fields: params=0.25,0.75   strings: str=abc   bool: hist=1
//   wire_probe decode <message> <hex>                 -> one "field=value" per line (floats as
//                                                        IEEE-754 hex words, exact)
//   wire_probe req <url> <hex> [<hex> ...]            -> sends every payload as one request,
//                                                        prints the reply payloads as hex
//        optional env: WIRE_RCVTIMEO_MS, WIRE_RESEND_MS, WIRE_RECONNECT_MS
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "args_pb.hpp"
#include "sp_req.hpp"

using namespace gSMFRETda::pb;

static std::string to_hex(const std::string& s)
{
    static const char* d = "0123456789abcdef";
    std::string o;
    for (unsigned char c : s) {
        o.push_back(d[c >> 4]);
        o.push_back(d[c & 15]);
    }
    return o;
}
static std::string from_hex(const char* h)
{
    std::string o;
    size_t n = std::strlen(h);
    for (size_t i = 0; i + 1 < n; i += 2) o.push_back((char)std::strtoul(std::string(h + i, 2).c_str(), nullptr, 16));
    return o;
}
static uint32_t fbits(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}
// floats are given either as decimal text or as 0x<8 hex digits> (exact bit pattern)
static float parse_float(const std::string& s)
{
    if (s.size() > 2 && s[0] == '0' && (s[1] == 'x' || s[1] == 'X')) {
        uint32_t u = (uint32_t)std::strtoul(s.c_str(), nullptr, 16);
        float f;
        std::memcpy(&f, &u, 4);
        return f;
    }
    return std::strtof(s.c_str(), nullptr);
}
static std::vector<std::string> split(const std::string& s, char c)
{
    std::vector<std::string> v;
    std::stringstream ss(s);
    std::string t;
    while (std::getline(ss, t, c)) v.push_back(t);
    return v;
}

template <class M, class F>
static int enc(int argc, char** argv, F set)
{
    M m;
    for (int i = 3; i < argc; ++i) {
        const char* eq = std::strchr(argv[i], '=');
        if (!eq) return 2;
        set(m, std::string(argv[i], eq - argv[i]), std::string(eq + 1));
    }
    std::string s;
    m.SerializeToString(&s);
    std::printf("%s\n", to_hex(s).c_str());
    return 0;
}

static void print_floats(const char* name, const std::vector<float>& v)
{
    std::printf("%s=", name);
    for (size_t i = 0; i < v.size(); ++i) std::printf("%s0x%08x", i ? "," : "", fbits(v[i]));
    std::printf("\n");
}

int main(int argc, char** argv)
{
    if (argc < 3) {
        std::fprintf(stderr, "usage: wire_probe encode|decode|req ...\n");
        return 2;
    }
    const std::string cmd = argv[1], what = argv[2];
    if (cmd == "encode") {
        if (what == "p_ga")
            return enc<p_ga>(argc, argv, [](p_ga& m, const std::string& k, const std::string& v) {
                if (k == "idx") m.set_idx(std::atoi(v.c_str()));
                else if (k == "start") m.set_start(std::atoi(v.c_str()));
                else if (k == "stop") m.set_stop(std::atoi(v.c_str()));
                else if (k == "bestfv") m.set_bestfv(parse_float(v));
                else if (k == "params") for (auto& t : split(v, ',')) m.add_params(parse_float(t));
            });
        if (what == "p_n") return enc<p_n>(argc, argv, [](p_n& m, const std::string&, const std::string& v) { m.set_s_n(std::atoi(v.c_str())); });
        if (what == "p_sid") return enc<p_sid>(argc, argv, [](p_sid& m, const std::string&, const std::string& v) { m.set_sid(std::atoi(v.c_str())); });
        if (what == "p_cap") return enc<p_cap>(argc, argv, [](p_cap& m, const std::string&, const std::string& v) { m.set_cap(std::atoi(v.c_str())); });
        if (what == "p_str")
            return enc<p_str>(argc, argv, [](p_str& m, const std::string& k, const std::string& v) {
                if (k == "histNum") m.set_histnum(std::atoi(v.c_str()));
                else if (k == "str") m.set_str(v);
                else if (k == "ohist") for (auto& t : split(v, ',')) m.add_ohist(parse_float(t));
            });
        if (what == "res")
            return enc<res>(argc, argv, [](res& m, const std::string& k, const std::string& v) {
                if (k == "ridx") m.set_ridx(std::atoi(v.c_str()));
                else if (k == "e") m.set_e(parse_float(v));
                else if (k == "hist") m.set_hist(std::atoi(v.c_str()) != 0);
                else if (k == "idx") m.set_idx(v);
                else if (k == "shist") for (auto& t : split(v, ',')) m.add_shist(parse_float(t));
                else if (k == "es") for (auto& t : split(v, ',')) m.add_es(parse_float(t));
                else if (k == "best") m.set_best(std::atoi(v.c_str()));
            });
        return 2;
    }
    if (cmd == "decode") {
        if (argc < 4) return 2;
        const std::string b = from_hex(argv[3]);
        bool ok = false;
        if (what == "p_ga") {
            p_ga m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("idx=%d\nstart=%d\nstop=%d\nbestfv=0x%08x\n", m.idx(), m.start(), m.stop(), fbits(m.bestfv()));
            print_floats("params", m.params());
        } else if (what == "p_n") {
            p_n m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("s_n=%d\n", m.s_n());
        } else if (what == "p_sid") {
            p_sid m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("sid=%d\n", m.sid());
        } else if (what == "p_cap") {
            p_cap m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("cap=%d\n", m.cap());
        } else if (what == "p_str") {
            p_str m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("histNum=%d\nstr=%s\n", m.histnum(), m.str().c_str());
            print_floats("ohist", m.ohist());
        } else if (what == "res") {
            res m;
            ok = m.ParseFromArray(b.data(), (int)b.size());
            std::printf("ridx=%d\ne=0x%08x\nhist=%d\nidx=%s\n", m.ridx(), fbits(m.e()), m.hist() ? 1 : 0, m.idx().c_str());
            print_floats("shist", m.shist());
            if (m.es_size()) {
                std::vector<float> es;
                for (int i = 0; i < m.es_size(); ++i) es.push_back(m.es(i));
                print_floats("es", es);
                std::printf("best=%d\n", m.best());
            }
        } else {
            return 2;
        }
        std::printf("ok=%d\n", ok ? 1 : 0);
        return 0;
    }
    if (cmd == "req") {
        int sock = nn_socket(AF_SP, NN_REQ);
        if (sock < 0) return 1;
        auto envopt = [&](const char* name, int level, int opt) {
            if (const char* e = std::getenv(name)) {
                int v = std::atoi(e);
                nn_setsockopt(sock, level, opt, &v, sizeof v);
            }
        };
        int ivl = 200;
        nn_setsockopt(sock, NN_SOL_SOCKET, NN_RECONNECT_IVL, &ivl, sizeof ivl);
        envopt("WIRE_RCVTIMEO_MS", NN_SOL_SOCKET, NN_RCVTIMEO);
        envopt("WIRE_RECONNECT_MS", NN_SOL_SOCKET, NN_RECONNECT_IVL);
        envopt("WIRE_RESEND_MS", NN_REQ, NN_REQ_RESEND_IVL);
        if (nn_connect(sock, what.c_str()) < 0) {
            std::fprintf(stderr, "nn_connect: %s\n", nn_strerror(errno));
            return 1;
        }
        for (int i = 3; i < argc; ++i) {
            const std::string b = from_hex(argv[i]);
            if (nn_send(sock, b.data(), b.size(), 0) < 0) {
                std::printf("send-error %s\n", nn_strerror(errno));
                return 1;
            }
            char* r = nullptr;
            int n = nn_recv(sock, &r, NN_MSG, 0);
            if (n < 0) {
                std::printf("recv-error %s\n", nn_strerror(errno));
                return 1;
            }
            std::printf("%s\n", to_hex(std::string(r, (size_t)n)).c_str());
            std::fflush(stdout);
            nn_freemsg(r);
        }
        nn_close(sock);
        return 0;
    }
    return 2;
}
