// sp_req.hpp — the REQ side of nanomsg's scalability protocols over tcp:// and ipc://, with the
// slice of the nanomsg C API the node uses (nn_socket, nn_setsockopt, nn_connect, nn_send,
// nn_recv with NN_MSG, nn_freemsg, nn_close, nn_strerror).
//
// The reference links nng 1.5.2 through its nanomsg compatibility layer
// (/root/reference/src/streamWorker.cpp:3-4,51-59,87-95,241); nng is not part of this build, so
// the two wire documents it implements are written out here:
//   * sp-tcp-mapping-01: after connect() both ends send 8 bytes, 0x00 'S' 'P' 0x00, the 16-bit
//     big-endian protocol number of the sender (REQ = 0x30, REP = 0x31), 16 reserved zero bits;
//     then every message is a 64-bit big-endian length followed by the payload.  The ipc://
//     mapping is the same with one extra byte 0x01 in front of each length;
//   * sp-request-reply-01: a request starts with a 32-bit big-endian request id whose top bit is
//     set (the bottom of the backtrace stack); the reply echoes it; a REQ socket has one
//     outstanding request, drops replies whose id does not match, and sends the request again
//     when the connection is lost or when no reply came within the resend interval (60 s).
// Blocking sockets, one connection per socket, reconnect every NN_RECONNECT_IVL ms — what
// streamWorker::run needs, nothing more (no REP/PUB/..., no device, no poll).
#pragma once
#include <arpa/inet.h>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <random>
#include <string>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/un.h>
#include <thread>
#include <unistd.h>
#include <vector>

#define AF_SP 1
#define NN_PROTO_REQREP 3
#define NN_REQ (NN_PROTO_REQREP * 16 + 0)
#define NN_REP (NN_PROTO_REQREP * 16 + 1)
#define NN_SOL_SOCKET 0
#define NN_SNDTIMEO 4
#define NN_RCVTIMEO 5
#define NN_RECONNECT_IVL 6
#define NN_REQ_RESEND_IVL 1
#define NN_MSG ((size_t)-1)
#ifndef EFSM
#define EFSM (156384712 + 54)   // nanomsg's NN_HAUSNUMERO + 54
#endif

namespace sp {

struct ReqSocket {
    std::string url;
    int fd = -1;
    bool ipc = false;
    int reconnect_ivl = 100;      // ms (nanomsg default); the node sets 200
    int resend_ivl = 60000;       // ms
    int rcvtimeo = -1, sndtimeo = -1;
    uint32_t next_id = 0;
    uint32_t cur_id = 0;
    std::string pending;          // the outstanding request (body), kept for resends
    bool outstanding = false;
    bool closed = false;
};

inline std::vector<ReqSocket*>& table()
{
    static std::vector<ReqSocket*> t;
    return t;
}
inline std::mutex& table_mutex()
{
    static std::mutex m;
    return m;
}
inline ReqSocket* lookup(int s)
{
    std::lock_guard<std::mutex> g(table_mutex());
    if (s < 0 || s >= (int)table().size()) return nullptr;
    return table()[(size_t)s];
}

inline bool write_all(int fd, const void* buf, size_t n)
{
    const char* p = (const char*)buf;
    while (n > 0) {
        ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
        if (w < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        p += w;
        n -= (size_t)w;
    }
    return true;
}

// 1 = read, 0 = timed out (nothing consumed), -1 = connection lost
inline int read_all(int fd, void* buf, size_t n, int timeout_ms, bool first_byte_only_times_out)
{
    char* p = (char*)buf;
    size_t got = 0;
    while (got < n) {
        struct pollfd pf = {fd, POLLIN, 0};
        const int tmo = (got == 0 || !first_byte_only_times_out) ? timeout_ms : 30000;
        int pr = ::poll(&pf, 1, tmo);
        if (pr < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        if (pr == 0) return got == 0 ? 0 : -1;
        ssize_t r = ::recv(fd, p + got, n - got, 0);
        if (r < 0) {
            if (errno == EINTR || errno == EAGAIN) continue;
            return -1;
        }
        if (r == 0) return -1;
        got += (size_t)r;
    }
    return 1;
}

inline void drop(ReqSocket& s)
{
    if (s.fd >= 0) ::close(s.fd);
    s.fd = -1;
}

// one connection attempt + SP handshake
inline bool try_connect(ReqSocket& s)
{
    int fd = -1;
    if (s.url.compare(0, 6, "tcp://") == 0) {
        std::string hp = s.url.substr(6);
        // nanomsg allows "tcp://iface;host:port": the local interface part is ignored here
        const size_t semi = hp.find(';');
        if (semi != std::string::npos) hp = hp.substr(semi + 1);
        const size_t colon = hp.rfind(':');
        if (colon == std::string::npos) return false;
        std::string host = hp.substr(0, colon), port = hp.substr(colon + 1);
        if (!host.empty() && host.front() == '[' && host.back() == ']') host = host.substr(1, host.size() - 2);
        struct addrinfo hints;
        std::memset(&hints, 0, sizeof hints);
        hints.ai_family = AF_UNSPEC;
        hints.ai_socktype = SOCK_STREAM;
        struct addrinfo* res = nullptr;
        if (::getaddrinfo(host.c_str(), port.c_str(), &hints, &res) != 0) return false;
        for (struct addrinfo* ai = res; ai; ai = ai->ai_next) {
            fd = ::socket(ai->ai_family, ai->ai_socktype | SOCK_CLOEXEC, ai->ai_protocol);
            if (fd < 0) continue;
            if (::connect(fd, ai->ai_addr, ai->ai_addrlen) == 0) break;
            ::close(fd);
            fd = -1;
        }
        ::freeaddrinfo(res);
        if (fd < 0) return false;
        int one = 1;
        ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        s.ipc = false;
    } else if (s.url.compare(0, 6, "ipc://") == 0) {
        struct sockaddr_un sa;
        std::memset(&sa, 0, sizeof sa);
        sa.sun_family = AF_UNIX;
        const std::string path = s.url.substr(6);
        if (path.size() >= sizeof sa.sun_path) return false;
        std::memcpy(sa.sun_path, path.c_str(), path.size());
        fd = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (fd < 0) return false;
        if (::connect(fd, (struct sockaddr*)&sa, sizeof sa) != 0) {
            ::close(fd);
            return false;
        }
        s.ipc = true;
    } else {
        return false;
    }
    const unsigned char hello[8] = {0x00, 'S', 'P', 0x00, 0x00, NN_REQ, 0x00, 0x00};
    unsigned char peer[8];
    if (!write_all(fd, hello, 8) || read_all(fd, peer, 8, 5000, false) != 1 || peer[0] != 0x00 || peer[1] != 'S' ||
        peer[2] != 'P' || peer[3] != 0x00 || ((peer[4] << 8) | peer[5]) != NN_REP) {
        ::close(fd);
        return false;
    }
    s.fd = fd;
    return true;
}

inline bool send_frame(ReqSocket& s)
{
    std::string f;
    const uint64_t len = 4u + s.pending.size();
    if (s.ipc) f.push_back((char)0x01);
    for (int i = 7; i >= 0; --i) f.push_back((char)(uint8_t)(len >> (8 * i)));
    for (int i = 3; i >= 0; --i) f.push_back((char)(uint8_t)(s.cur_id >> (8 * i)));
    f.append(s.pending);
    return write_all(s.fd, f.data(), f.size());
}

// (re)connect until the outstanding request is on the wire; false on deadline
inline bool deliver(ReqSocket& s, std::chrono::steady_clock::time_point deadline, bool has_deadline)
{
    for (;;) {
        if (s.closed) return false;
        if (s.fd < 0 && !try_connect(s)) {
            if (has_deadline && std::chrono::steady_clock::now() >= deadline) return false;
            std::this_thread::sleep_for(std::chrono::milliseconds(s.reconnect_ivl > 0 ? s.reconnect_ivl : 100));
            continue;
        }
        if (send_frame(s)) return true;
        drop(s);
    }
}

}  // namespace sp

inline int nn_socket(int domain, int protocol)
{
    if (domain != AF_SP || protocol != NN_REQ) {
        errno = EINVAL;
        return -1;
    }
    sp::ReqSocket* s = new sp::ReqSocket();
    std::random_device rd;
    s->next_id = (uint32_t)rd();
    std::lock_guard<std::mutex> g(sp::table_mutex());
    sp::table().push_back(s);
    return (int)sp::table().size() - 1;
}

inline int nn_setsockopt(int sock, int level, int option, const void* val, size_t len)
{
    sp::ReqSocket* s = sp::lookup(sock);
    if (!s || len != sizeof(int)) {
        errno = EINVAL;
        return -1;
    }
    int v;
    std::memcpy(&v, val, sizeof v);
    if (level == NN_SOL_SOCKET && option == NN_RECONNECT_IVL) s->reconnect_ivl = v;
    else if (level == NN_SOL_SOCKET && option == NN_RCVTIMEO) s->rcvtimeo = v;
    else if (level == NN_SOL_SOCKET && option == NN_SNDTIMEO) s->sndtimeo = v;
    else if (level == NN_REQ && option == NN_REQ_RESEND_IVL) s->resend_ivl = v;
    else {
        errno = ENOPROTOOPT;
        return -1;
    }
    return 0;
}

// like nanomsg: the endpoint is recorded, the connection is made (and re-made) in the background
// of the calls that need it
inline int nn_connect(int sock, const char* url)
{
    sp::ReqSocket* s = sp::lookup(sock);
    if (!s || !url || (std::strncmp(url, "tcp://", 6) != 0 && std::strncmp(url, "ipc://", 6) != 0)) {
        errno = EINVAL;
        return -1;
    }
    s->url = url;
    return 1;   // endpoint id
}

// a new request cancels the outstanding one (REQ state machine)
inline int nn_send(int sock, const void* buf, size_t len, int /*flags*/)
{
    sp::ReqSocket* s = sp::lookup(sock);
    if (!s || s->url.empty() || len == NN_MSG) {
        errno = EINVAL;
        return -1;
    }
    s->pending.assign((const char*)buf, len);
    s->cur_id = 0x80000000u | (s->next_id++);
    s->outstanding = true;
    const bool has_deadline = s->sndtimeo >= 0;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(has_deadline ? s->sndtimeo : 0);
    if (!sp::deliver(*s, deadline, has_deadline)) {
        s->outstanding = false;
        errno = ETIMEDOUT;
        return -1;
    }
    return (int)len;
}

// NN_MSG only: *(void**)buf receives a malloc'ed message, to be released with nn_freemsg
inline int nn_recv(int sock, void* buf, size_t len, int /*flags*/)
{
    sp::ReqSocket* s = sp::lookup(sock);
    if (!s || len != NN_MSG || !buf) {
        errno = EINVAL;
        return -1;
    }
    if (!s->outstanding) {
        errno = EFSM;   // no request to answer (nanomsg: "operation cannot be performed in this state")
        return -1;
    }
    using clock = std::chrono::steady_clock;
    const bool has_deadline = s->rcvtimeo >= 0;
    const auto deadline = clock::now() + std::chrono::milliseconds(has_deadline ? s->rcvtimeo : 0);
    auto resend_at = clock::now() + std::chrono::milliseconds(s->resend_ivl);
    for (;;) {
        if (s->fd < 0) {
            if (!sp::deliver(*s, deadline, has_deadline)) {
                errno = ETIMEDOUT;
                return -1;
            }
            resend_at = clock::now() + std::chrono::milliseconds(s->resend_ivl);
        }
        auto now = clock::now();
        if (has_deadline && now >= deadline) {
            errno = ETIMEDOUT;
            return -1;
        }
        if (now >= resend_at) {
            if (!sp::send_frame(*s)) {
                sp::drop(*s);
                continue;
            }
            resend_at = now + std::chrono::milliseconds(s->resend_ivl);
        }
        auto until = resend_at;
        if (has_deadline && deadline < until) until = deadline;
        int wait_ms = (int)std::chrono::duration_cast<std::chrono::milliseconds>(until - now).count() + 1;
        unsigned char hdr[9];
        const size_t hl = s->ipc ? 9 : 8;
        int r = sp::read_all(s->fd, hdr, hl, wait_ms, true);
        if (r == 0) continue;
        if (r < 0 || (s->ipc && hdr[0] != 0x01)) {
            sp::drop(*s);
            continue;   // deliver() sends the request again on the new connection
        }
        uint64_t n = 0;
        for (size_t i = hl - 8; i < hl; ++i) n = (n << 8) | hdr[i];
        if (n > (1ull << 31)) {
            sp::drop(*s);
            continue;
        }
        std::string body((size_t)n, '\0');
        if (n > 0 && sp::read_all(s->fd, &body[0], (size_t)n, 30000, false) != 1) {
            sp::drop(*s);
            continue;
        }
        if (n < 4) continue;
        const uint32_t id = ((uint32_t)(uint8_t)body[0] << 24) | ((uint32_t)(uint8_t)body[1] << 16) |
                            ((uint32_t)(uint8_t)body[2] << 8) | (uint32_t)(uint8_t)body[3];
        if (id != s->cur_id) continue;   // a reply to a cancelled request
        const size_t m = (size_t)n - 4;
        char* out = (char*)std::malloc(m ? m : 1);
        if (!out) {
            errno = ENOMEM;
            return -1;
        }
        std::memcpy(out, body.data() + 4, m);
        *(void**)buf = out;
        s->outstanding = false;
        return (int)m;
    }
}

inline int nn_freemsg(void* msg)
{
    std::free(msg);
    return 0;
}

inline int nn_close(int sock)
{
    sp::ReqSocket* s = sp::lookup(sock);
    if (!s) {
        errno = EBADF;
        return -1;
    }
    s->closed = true;
    sp::drop(*s);
    {
        std::lock_guard<std::mutex> g(sp::table_mutex());
        sp::table()[(size_t)sock] = nullptr;
    }
    delete s;
    return 0;
}

inline const char* nn_strerror(int errnum) { return std::strerror(errnum); }
