// args_pb.hpp — the six messages of the parameter-server protocol
// (/root/reference/3rdparty/protobuf/args.proto:1-46), encoded and decoded by hand.
//
// The reference generates args.pb.h with protoc and links libprotobuf 3.13; neither is part of
// this build.  The wire format of these messages needs four things only — base-128 varints,
// little-endian fixed32, length-delimited fields and proto3's "default values are not written"
// rule — so the codec is written out here, with the accessor names protoc would have generated
// (set_cap, s_n, add_ohist, params(i), SerializeToString, ParseFromArray ...), so that code
// written against args.pb.h (streamWorker.cpp:83-238) reads the same.
//
// Byte compatibility with libprotobuf is pinned by tests/test_wire_protocol.py against the
// Python protobuf runtime: same bytes out for the same fields (fields in tag order, packed
// repeated floats, int32 < 0 as a 10-byte varint), and everything that runtime writes parses
// (packed and unpacked repeated fields, unknown fields skipped, last value of a scalar wins).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace gSMFRETda {
namespace pb {

namespace wire {

inline void put_varint(std::string& o, uint64_t v)
{
    while (v >= 0x80u) {
        o.push_back((char)(uint8_t)(v | 0x80u));
        v >>= 7;
    }
    o.push_back((char)(uint8_t)v);
}
inline void put_tag(std::string& o, int field, int wt) { put_varint(o, ((uint64_t)field << 3) | (uint64_t)wt); }
// proto3 scalars: the default value is not written
inline void put_int32(std::string& o, int field, int32_t v)
{
    if (v == 0) return;
    put_tag(o, field, 0);
    put_varint(o, (uint64_t)(int64_t)v);   // negative values are sign-extended to 64 bits: 10 bytes
}
inline void put_bool(std::string& o, int field, bool v)
{
    if (!v) return;
    put_tag(o, field, 0);
    o.push_back((char)1);
}
inline void put_fixed32(std::string& o, uint32_t u)
{
    for (int i = 0; i < 4; ++i) o.push_back((char)(uint8_t)(u >> (8 * i)));
}
inline void put_float(std::string& o, int field, float v)
{
    if (v <= 0.0f && v >= 0.0f) return;    // libprotobuf 3.13's test: +0.0 and -0.0 are "default", NaN is not
    uint32_t u;
    std::memcpy(&u, &v, 4);
    put_tag(o, field, 5);
    put_fixed32(o, u);
}
inline void put_string(std::string& o, int field, const std::string& v)
{
    if (v.empty()) return;
    put_tag(o, field, 2);
    put_varint(o, v.size());
    o.append(v);
}
inline void put_packed_floats(std::string& o, int field, const std::vector<float>& v)
{
    if (v.empty()) return;
    put_tag(o, field, 2);
    put_varint(o, v.size() * 4u);
    for (float f : v) {
        uint32_t u;
        std::memcpy(&u, &f, 4);
        put_fixed32(o, u);
    }
}

struct Reader {
    const uint8_t* p;
    const uint8_t* end;
    bool ok = true;
    Reader(const void* d, size_t n) : p((const uint8_t*)d), end((const uint8_t*)d + n) {}
    bool done() const { return p >= end; }
    uint64_t varint()
    {
        uint64_t v = 0;
        for (int shift = 0; shift < 70; shift += 7) {
            if (p >= end) { ok = false; return 0; }
            const uint8_t b = *p++;
            if (shift < 64) v |= (uint64_t)(b & 0x7fu) << shift;
            if (!(b & 0x80u)) return v;
        }
        ok = false;
        return 0;
    }
    uint32_t fixed32()
    {
        if (end - p < 4) { ok = false; p = end; return 0; }
        uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
        p += 4;
        return u;
    }
    float f32()
    {
        uint32_t u = fixed32();
        float f;
        std::memcpy(&f, &u, 4);
        return f;
    }
    // length-delimited payload
    Reader sub()
    {
        uint64_t n = varint();
        if (!ok || n > (uint64_t)(end - p)) { ok = false; p = end; return Reader(end, 0); }
        Reader r(p, (size_t)n);
        p += n;
        return r;
    }
    void skip(int wt)
    {
        switch (wt) {
            case 0: varint(); break;
            case 1: if (end - p < 8) { ok = false; p = end; } else p += 8; break;
            case 2: sub(); break;
            case 5: fixed32(); break;
            default: ok = false; p = end;   // groups (3, 4) do not occur in proto3 output
        }
    }
    // repeated float: packed (wire type 2) or one element (wire type 5)
    void floats(int wt, std::vector<float>& out)
    {
        if (wt == 2) {
            Reader r = sub();
            while (ok && !r.done()) out.push_back(r.f32());
            if (!r.ok) ok = false;
        } else if (wt == 5) {
            out.push_back(f32());
        } else {
            skip(wt);
        }
    }
    void str(int wt, std::string& out)
    {
        if (wt != 2) { skip(wt); return; }
        Reader r = sub();
        out.assign((const char*)r.p, (size_t)(r.end - r.p));
    }
};

}  // namespace wire

// Common surface of protoc's generated classes that the node uses
#define GSM_PB_COMMON(NAME)                                                              \
    bool SerializeToString(std::string* out) const { out->clear(); append_to(*out); return true; } \
    std::string SerializeAsString() const { std::string s; append_to(s); return s; }              \
    bool ParseFromArray(const void* data, int size)                                               \
    {                                                                                             \
        Clear();                                                                                  \
        if (size < 0) return false;                                                               \
        wire::Reader r(data, (size_t)size);                                                       \
        while (r.ok && !r.done()) {                                                               \
            const uint64_t tag = r.varint();                                                      \
            if (!r.ok) break;                                                                     \
            if ((tag >> 3) == 0) { r.ok = false; break; }                                         \
            field((int)(tag >> 3), (int)(tag & 7u), r);                                           \
        }                                                                                         \
        return r.ok;                                                                              \
    }                                                                                             \
    bool ParseFromString(const std::string& s) { return ParseFromArray(s.data(), (int)s.size()); }

// message p_ga { int32 idx = 1; int32 start = 2; int32 stop = 3; float bestfv = 4;
//                repeated float params = 5 [packed=true]; }            (args.proto:4-13)
class p_ga {
  public:
    int32_t idx() const { return idx_; }
    int32_t start() const { return start_; }
    int32_t stop() const { return stop_; }
    float bestfv() const { return bestfv_; }
    int params_size() const { return (int)params_.size(); }
    float params(int i) const { return params_[(size_t)i]; }
    const std::vector<float>& params() const { return params_; }
    void set_idx(int32_t v) { idx_ = v; }
    void set_start(int32_t v) { start_ = v; }
    void set_stop(int32_t v) { stop_ = v; }
    void set_bestfv(float v) { bestfv_ = v; }
    void add_params(float v) { params_.push_back(v); }
    void Clear() { idx_ = start_ = stop_ = 0; bestfv_ = 0.0f; params_.clear(); }
    GSM_PB_COMMON(p_ga)
  private:
    void append_to(std::string& o) const
    {
        wire::put_int32(o, 1, idx_);
        wire::put_int32(o, 2, start_);
        wire::put_int32(o, 3, stop_);
        wire::put_float(o, 4, bestfv_);
        wire::put_packed_floats(o, 5, params_);
    }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) idx_ = (int32_t)r.varint();
        else if (f == 2 && wt == 0) start_ = (int32_t)r.varint();
        else if (f == 3 && wt == 0) stop_ = (int32_t)r.varint();
        else if (f == 4 && wt == 5) bestfv_ = r.f32();
        else if (f == 5) r.floats(wt, params_);
        else r.skip(wt);
    }
    int32_t idx_ = 0, start_ = 0, stop_ = 0;
    float bestfv_ = 0.0f;
    std::vector<float> params_;
};

// message p_n { int32 s_n = 1; }                                         (args.proto:15-17)
class p_n {
  public:
    int32_t s_n() const { return s_n_; }
    void set_s_n(int32_t v) { s_n_ = v; }
    void Clear() { s_n_ = 0; }
    GSM_PB_COMMON(p_n)
  private:
    void append_to(std::string& o) const { wire::put_int32(o, 1, s_n_); }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) s_n_ = (int32_t)r.varint();
        else r.skip(wt);
    }
    int32_t s_n_ = 0;
};

// message p_sid { int32 sid = 1; }   sid == -1 stops the node            (args.proto:22-27)
class p_sid {
  public:
    int32_t sid() const { return sid_; }
    void set_sid(int32_t v) { sid_ = v; }
    void Clear() { sid_ = 0; }
    GSM_PB_COMMON(p_sid)
  private:
    void append_to(std::string& o) const { wire::put_int32(o, 1, sid_); }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) sid_ = (int32_t)r.varint();
        else r.skip(wt);
    }
    int32_t sid_ = 0;
};

// message p_cap { int32 cap = 1; }                                       (args.proto:29-31)
class p_cap {
  public:
    int32_t cap() const { return cap_; }
    void set_cap(int32_t v) { cap_ = v; }
    void Clear() { cap_ = 0; }
    GSM_PB_COMMON(p_cap)
  private:
    void append_to(std::string& o) const { wire::put_int32(o, 1, cap_); }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) cap_ = (int32_t)r.varint();
        else r.skip(wt);
    }
    int32_t cap_ = 0;
};

// message p_str { int32 histNum = 1; string str = 2; repeated float ohist = 3 [packed=true]; }
//                                                                        (args.proto:33-37)
class p_str {
  public:
    int32_t histnum() const { return histnum_; }
    const std::string& str() const { return str_; }
    int ohist_size() const { return (int)ohist_.size(); }
    float ohist(int i) const { return ohist_[(size_t)i]; }
    const std::vector<float>& ohist() const { return ohist_; }
    void set_histnum(int32_t v) { histnum_ = v; }
    void set_str(const std::string& v) { str_ = v; }
    void add_ohist(float v) { ohist_.push_back(v); }
    void Clear() { histnum_ = 0; str_.clear(); ohist_.clear(); }
    GSM_PB_COMMON(p_str)
  private:
    void append_to(std::string& o) const
    {
        wire::put_int32(o, 1, histnum_);
        wire::put_string(o, 2, str_);
        wire::put_packed_floats(o, 3, ohist_);
    }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) histnum_ = (int32_t)r.varint();
        else if (f == 2) r.str(wt, str_);
        else if (f == 3) r.floats(wt, ohist_);
        else r.skip(wt);
    }
    int32_t histnum_ = 0;
    std::string str_;
    std::vector<float> ohist_;
};

// message res { int32 ridx = 1; float e = 2; bool hist = 3; string idx = 4;
//               repeated float shist = 5 [packed=true]; }               (args.proto:39-45)
// Batch extension: when a p_ga reply carried P > 1 vectors
// (params_size = P (n*n + n), indices idx .. idx + P - 1), the result message adds
//               repeated float es = 6 [packed=true];   chi-square of every vector of the batch
//               int32 best = 7;                         offset of the smallest one; e, hist, shist describe it
// Both are absent for a single vector (proto3 writes no defaults), so the unbatched bytes are the
// reference's; a server that does not know them skips them as unknown fields.
class res {
  public:
    int32_t ridx() const { return ridx_; }
    float e() const { return e_; }
    bool hist() const { return hist_; }
    const std::string& idx() const { return idx_; }
    int shist_size() const { return (int)shist_.size(); }
    float shist(int i) const { return shist_[(size_t)i]; }
    const std::vector<float>& shist() const { return shist_; }
    void set_ridx(int32_t v) { ridx_ = v; }
    void set_e(float v) { e_ = v; }
    void set_hist(bool v) { hist_ = v; }
    void set_idx(const std::string& v) { idx_ = v; }
    void add_shist(float v) { shist_.push_back(v); }
    int es_size() const { return (int)es_.size(); }
    float es(int i) const { return es_[(size_t)i]; }
    void add_es(float v) { es_.push_back(v); }
    int32_t best() const { return best_; }
    void set_best(int32_t v) { best_ = v; }
    void Clear() { ridx_ = 0; e_ = 0.0f; hist_ = false; idx_.clear(); shist_.clear(); es_.clear(); best_ = 0; }
    GSM_PB_COMMON(res)
  private:
    void append_to(std::string& o) const
    {
        wire::put_int32(o, 1, ridx_);
        wire::put_float(o, 2, e_);
        wire::put_bool(o, 3, hist_);
        wire::put_string(o, 4, idx_);
        wire::put_packed_floats(o, 5, shist_);
        wire::put_packed_floats(o, 6, es_);
        wire::put_int32(o, 7, best_);
    }
    void field(int f, int wt, wire::Reader& r)
    {
        if (f == 1 && wt == 0) ridx_ = (int32_t)r.varint();
        else if (f == 2 && wt == 5) e_ = r.f32();
        else if (f == 3 && wt == 0) hist_ = r.varint() != 0;
        else if (f == 4) r.str(wt, idx_);
        else if (f == 5) r.floats(wt, shist_);
        else if (f == 6) r.floats(wt, es_);
        else if (f == 7 && wt == 0) best_ = (int32_t)r.varint();
        else r.skip(wt);
    }
    int32_t ridx_ = 0;
    float e_ = 0.0f;
    bool hist_ = false;
    std::string idx_;
    std::vector<float> shist_;
    std::vector<float> es_;
    int32_t best_ = 0;
};

#undef GSM_PB_COMMON

}  // namespace pb
}  // namespace gSMFRETda
