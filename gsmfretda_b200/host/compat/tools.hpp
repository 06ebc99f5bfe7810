// compat/tools.hpp — lets code written against the REFERENCE's tools.hpp
// (/root/reference/src/tools.hpp: `AtomicWriter(debug, debugLevel::gpu) << ...`, `genuid`) compile
// against this node.  Not used by the node itself (it logs through node_log.hpp); it exists for
// the drop-in proof of tests/test_host_cpp.py, which compiles the reference's unmodified
// gpuWorker.cpp against host/mc.hpp, and for a maintainer who keeps reference-side callers.
#pragma once
#include <sstream>
#include <string>

#include "../node_log.hpp"

// the reference's names for the debug channels (bit values = the -C/-G/-N/-K command-line flags)
enum debugLevel : unsigned char {
    cpu = node::kCpu, gpu = node::kGpu, net = node::kNet, info = node::kInfo,
    parameter = node::kParameter, kernel = node::kKernel
};

// stream-style front end of node::say: the pieces are joined and handed over as one message
struct AtomicWriter {
    AtomicWriter(unsigned char enabled, debugLevel channel) : on_((enabled & channel) != 0), mask_(enabled), channel_((node::Channel)channel) {}
    ~AtomicWriter() { if (on_) node::say(mask_, channel_, "%s", text_.str().c_str()); }
    template <class Piece>
    AtomicWriter& operator<<(const Piece& piece) { return on_ ? (text_ << piece, *this) : *this; }

  private:
    bool on_;
    unsigned char mask_;
    node::Channel channel_;
    std::ostringstream text_;
};

inline void genuid(std::string* id, int gid, int sid, char* gpuuuid) { *id = node::stream_name(gid, sid, gpuuuid); }
