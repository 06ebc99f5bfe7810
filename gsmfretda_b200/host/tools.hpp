// tools.hpp — debug levels, the atomic debug writer and the node id of the reference
// (/root/reference/src/tools.hpp:8-41, tools.cpp:150-167), header-only.
#pragma once
#include <cstdio>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <thread>
#include <unistd.h>

enum debugLevel : unsigned char {
    cpu = 1,
    gpu = 1 << 1,
    net = 1 << 2,
    info = 1 << 3,
    parameter = 1 << 4,
    kernel = 1 << 5
};

// collects a message and writes it to stdout in one piece when it goes out of scope, if the
// wanted level is switched on (tools.hpp:17-41)
class AtomicWriter {
    std::ostringstream st;
    unsigned char debug = 0;

  public:
    AtomicWriter(unsigned char used, debugLevel wanted) { debug = wanted & used; }
    template <typename T>
    AtomicWriter& operator<<(T const& t)
    {
        if (debug) st << t;
        return *this;
    }
    ~AtomicWriter()
    {
        if (debug) std::cout << st.str();
    }
};

// "<32 hex digits of the GPU uuid>-<pid>-<thread id>-<gpu>-<stream>" (tools.cpp:150-167): the
// name a stream of this node carries at the parameter server
inline void genuid(std::string* id, int gid, int sid, const char* gpuuuid)
{
    std::stringstream ss, sss;
    for (int i = 0; i < 16; i++) {
        unsigned char b;
        std::memcpy(&b, gpuuuid + i, 1);
        sss << std::setfill('0') << std::setw(2) << std::hex << int(b);
    }
    ss << sss.str() << '-' << std::to_string(getpid()) << "-" << std::this_thread::get_id() << "-"
       << std::to_string(gid) << "-" << std::to_string(sid);
    *id = ss.str();
}
