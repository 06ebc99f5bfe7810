// h5_probe.cpp — runs loadhdf5 + burst_data (loadHdf5.hpp) on a file and prints what they
// produced, for tests/test_hdf5_loader.py.  No GPU library behind it.
//   h5_probe load <file>                 -> one line per array: name n v0 v1 ... (floats as hex words)
//   h5_probe read <file> <dataset> i64|f32|u8   -> one dataset through h5lite alone
//   h5_probe info <file> <dataset>       -> dims, type, layout of a dataset
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include "loadHdf5.hpp"

template <class T> static void dump_i(const char* name, const std::vector<T>& v)
{
    std::printf("%s %zu", name, v.size());
    for (auto x : v) std::printf(" %" PRId64, (int64_t)x);
    std::printf("\n");
}
static void dump_f(const char* name, const std::vector<float>& v)
{
    std::printf("%s %zu", name, v.size());
    for (float x : v) { uint32_t u; std::memcpy(&u, &x, 4); std::printf(" %08x", u); }
    std::printf("\n");
}

int main(int argc, char** argv)
{
    if (argc < 3) { std::fprintf(stderr, "usage: h5_probe load|read|info <file> ...\n"); return 2; }
    const std::string cmd = argv[1];
    try {
        if (cmd == "load") {
            std::vector<int64_t> start, stop, istart, istop, times;
            std::vector<unsigned char> mad, mdd;
            std::vector<float> T, S;
            float clk, bga, bgd;
            std::uintmax_t sz = loadhdf5(argv[2], start, stop, istart, istop, times, mad, mdd, T, S, clk, bga, bgd);
            std::vector<int> ph;
            int64_t *ad = nullptr, *dd = nullptr;
            burst_data(istart, istop, times, mad, mdd, ph, &ad, &dd);
            std::printf("fsize 1 %ju\n", sz);
            dump_i("start", start); dump_i("stop", stop); dump_i("istart", istart); dump_i("istop", istop);
            dump_i("times", times); dump_i("mask_ad", mad); dump_i("mask_dd", mdd);
            dump_f("T_burst_duration", T); dump_f("SgDivSr", S);
            dump_f("scalars", std::vector<float>{clk, bga, bgd});
            dump_i("phCount", ph);
            dump_i("burst_ad", std::vector<int64_t>(ad, ad + times.size()));
            dump_i("burst_dd", std::vector<int64_t>(dd, dd + times.size()));
            delete[] ad; delete[] dd;
            return 0;
        }
        if (argc < 4) return 2;
        h5lite::File f(argv[2]);
        if (cmd == "read") {
            const std::string t = argc > 4 ? argv[4] : "i64";
            if (t == "i64") { std::vector<int64_t> v; f.read(argv[3], v); dump_i(argv[3], v); }
            else if (t == "f32") { std::vector<float> v; f.read(argv[3], v); dump_f(argv[3], v); }
            else if (t == "u8") { std::vector<unsigned char> v; f.read(argv[3], v); dump_i(argv[3], v); }
            else if (t == "i32") { std::vector<int> v; f.read(argv[3], v); dump_i(argv[3], v); }
            else return 2;
            return 0;
        }
        if (cmd == "info") {
            h5lite::Dataset d = f.dataset(argv[3]);
            std::printf("superblock %d rank %zu count %" PRIu64 " class %d size %d be %d signed %d layout %d layout_version %d index %d filters %zu\n",
                        f.superblock_version(), d.dims.size(), d.count(), d.type.cls, d.type.size, (int)d.type.big_endian,
                        (int)d.type.is_signed, d.layout.cls, d.layout.version, d.layout.index_type, d.filters.size());
            return 0;
        }
    } catch (const std::exception& e) {
        std::printf("error %s\n", e.what());
        return 1;
    }
    return 2;
}
