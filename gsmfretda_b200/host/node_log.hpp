// node_log.hpp — levelled diagnostics and stream naming of the worker node.
//
// The reference switches its debug output per subsystem with the -C/-G/-N/-K/-D flags
// (/root/reference/3rdparty/gengetopt/cmdline.ggo:15-19); the bit assigned to each flag is part of
// that command line and is kept (cpu 1, gpu 2, net 4, info 8, parameter 16, kernel 32).  Everything
// else here is this node's own: one printf-style call per message, formatted into a stack buffer
// and written with a single fwrite so that lines of different threads do not interleave.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <functional>
#include <string>
#include <thread>
#include <unistd.h>

namespace node {

enum Channel : unsigned char { kCpu = 0x01, kGpu = 0x02, kNet = 0x04, kInfo = 0x08, kParameter = 0x10, kKernel = 0x20 };

#if defined(__GNUC__)
__attribute__((format(printf, 3, 4)))
#endif
inline void say(unsigned char enabled, Channel ch, const char* fmt, ...)
{
    if (!(enabled & ch)) return;
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    int n = std::vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (n < 0) return;
    if (n >= (int)sizeof buf) n = (int)sizeof buf - 1;
    std::fwrite(buf, 1, (size_t)n, stdout);
}

// The name a stream of this node carries at the parameter server:
// "<32 hex digits of the GPU uuid>-<pid>-<thread>-<gpu>-<stream>" — the fields the reference
// puts into it (/root/reference/src/tools.cpp:150-167); the server treats it as an opaque key.
inline std::string stream_name(int gpu, int stream, const char uuid16[16])
{
    char hex[33];
    for (int i = 0; i < 16; ++i) std::snprintf(hex + 2 * i, 3, "%02x", (unsigned)(unsigned char)uuid16[i]);
    const unsigned long long tid = (unsigned long long)std::hash<std::thread::id>()(std::this_thread::get_id());
    char out[128];
    std::snprintf(out, sizeof out, "%s-%ld-%llu-%d-%d", hex, (long)getpid(), tid, gpu, stream);
    return out;
}

}  // namespace node
