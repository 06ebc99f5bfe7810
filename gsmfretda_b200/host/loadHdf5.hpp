// loadHdf5.hpp — input side of the node: the burst file -> the host arrays mc::init_data_gpu takes.
// Same functions and argument lists as the reference (/root/reference/src/loadHdf5.hpp:9-19,
// loadHdf5.cpp:17-128):
//   loadhdf5   reads /sub_bursts_l/{start,stop,istart,istop} (int64), times (int64),
//              mask_ad, mask_dd (uint8), T_burst_duration, SgDivSr (float32) and the scalars
//              clk_p, bg_ad_rate, bg_dd_rate (float32); returns the file size, exits on error;
//   burst_data phCount[b] = istop[b] - istart[b] + 1 and the two per-photon arrays
//              burst_ad = mask_ad * times, burst_dd = mask_dd * times (0 where the photon is
//              not of that channel), filled burst by burst at istart[b].
// The reference reads through HighFive + libhdf5; here the file is parsed by h5lite.hpp, a
// reader for the subset of the HDF5 file format these files use.  The same entry point also
// accepts the flat "GSMB" file of gsmfretda_b200.synth.save_bin (first four bytes decide).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <sys/stat.h>
#include <vector>

#include "h5lite.hpp"

typedef float retype;   // loadHdf5.hpp:8

namespace gsmio {

template <class T>
inline bool rd(FILE* f, std::vector<T>& v, size_t n)
{
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}

// "GSMB" u32 version=1, i64 B, i64 Nph, f32 clk_p, f32 bg_ad, f32 bg_dd, f32 pad, i64 start[B],
// i64 stop[B], i64 istart[B], i32 phCount[B], f32 T[B], f32 SgDivSr[B], i64 burst_ad[Nph],
// i64 burst_dd[Nph] — rebuilt into the HDF5 file's arrays (times, masks, istop)
inline bool load_bin(const std::string& fn, std::vector<int64_t>& start, std::vector<int64_t>& stop,
                     std::vector<int64_t>& istart, std::vector<int64_t>& istop, std::vector<int64_t>& times,
                     std::vector<unsigned char>& mask_ad, std::vector<unsigned char>& mask_dd,
                     std::vector<float>& T, std::vector<float>& SgDivSr, float& clk_p, float& bg_ad, float& bg_dd)
{
    FILE* f = std::fopen(fn.c_str(), "rb");
    if (!f) return false;
    char magic[4];
    uint32_t ver;
    int64_t B, Nph;
    float hdr[4];
    std::vector<int> ph;
    std::vector<int64_t> ad, dd;
    bool ok = fread(magic, 1, 4, f) == 4 && !std::memcmp(magic, "GSMB", 4) && fread(&ver, 4, 1, f) == 1 && ver == 1 &&
              fread(&B, 8, 1, f) == 1 && fread(&Nph, 8, 1, f) == 1 && fread(hdr, 4, 4, f) == 4 && B >= 0 && Nph >= 0;
    if (ok) {
        // the header may promise no more than the file holds: 36 bytes per burst, 16 per photon
        struct stat st;
        ok = ::fstat(fileno(f), &st) == 0 && B <= (int64_t)st.st_size / 36 && Nph <= (int64_t)st.st_size / 16 &&
             40 + 36 * B + 16 * Nph <= (int64_t)st.st_size;
    }
    ok = ok && rd(f, start, (size_t)B) && rd(f, stop, (size_t)B) && rd(f, istart, (size_t)B) && rd(f, ph, (size_t)B) &&
              rd(f, T, (size_t)B) && rd(f, SgDivSr, (size_t)B) && rd(f, ad, (size_t)Nph) && rd(f, dd, (size_t)Nph);
    std::fclose(f);
    if (!ok) return false;
    clk_p = hdr[0];
    bg_ad = hdr[1];
    bg_dd = hdr[2];
    istop.resize((size_t)B);
    for (int64_t b = 0; b < B; ++b) {
        if (ph[(size_t)b] < 0 || istart[(size_t)b] < 0 || istart[(size_t)b] > Nph - ph[(size_t)b]) return false;
        istop[(size_t)b] = istart[(size_t)b] + ph[(size_t)b] - 1;
    }
    times.resize((size_t)Nph);
    mask_ad.resize((size_t)Nph);
    mask_dd.resize((size_t)Nph);
    for (int64_t i = 0; i < Nph; ++i) {
        mask_ad[(size_t)i] = ad[(size_t)i] != 0;
        mask_dd[(size_t)i] = dd[(size_t)i] != 0;
        times[(size_t)i] = ad[(size_t)i] != 0 ? ad[(size_t)i] : dd[(size_t)i];
    }
    return true;
}

}  // namespace gsmio

inline std::uintmax_t loadhdf5(std::string H5FILE_NAME, std::vector<int64_t>& start, std::vector<int64_t>& stop,
                               std::vector<int64_t>& istart, std::vector<int64_t>& istop,
                               std::vector<int64_t>& times_ms, std::vector<unsigned char>& mask_ad,
                               std::vector<unsigned char>& mask_dd, std::vector<float>& T_burst_duration,
                               std::vector<float>& SgDivSr, float& clk_p, float& bg_ad_rate, float& bg_dd_rate)
{
    struct stat st;
    if (::stat(H5FILE_NAME.c_str(), &st) != 0) {   // loadHdf5.cpp:26-33
        std::cout << "file_size: " << H5FILE_NAME << ": " << std::strerror(errno) << '\n';
        std::cerr << EXIT_FAILURE << std::endl;
        std::exit(EXIT_FAILURE);
    }
    const std::uintmax_t fsize = (std::uintmax_t)st.st_size;
    char magic[8] = {0};
    if (FILE* f = std::fopen(H5FILE_NAME.c_str(), "rb")) {
        if (fread(magic, 1, 8, f) != 8) magic[0] = 0;
        std::fclose(f);
    }
    if (!std::memcmp(magic, "GSMB", 4)) {
        if (!gsmio::load_bin(H5FILE_NAME, start, stop, istart, istop, times_ms, mask_ad, mask_dd, T_burst_duration,
                             SgDivSr, clk_p, bg_ad_rate, bg_dd_rate)) {
            std::cerr << H5FILE_NAME << ": malformed GSMB file" << std::endl;
            std::exit(EXIT_FAILURE);
        }
        return fsize;
    }
    try {
        h5lite::File file(H5FILE_NAME);
        file.read("/sub_bursts_l/start", start);
        file.read("/sub_bursts_l/stop", stop);
        file.read("/sub_bursts_l/istart", istart);
        file.read("/sub_bursts_l/istop", istop);
        file.read("times", times_ms);
        file.read("mask_ad", mask_ad);
        file.read("mask_dd", mask_dd);
        file.read("T_burst_duration", T_burst_duration);
        file.read("SgDivSr", SgDivSr);
        file.read_scalar("clk_p", clk_p);
        file.read_scalar("bg_ad_rate", bg_ad_rate);
        file.read_scalar("bg_dd_rate", bg_dd_rate);
    } catch (const std::exception& err) {   // loadHdf5.cpp:73-77
        std::cerr << err.what() << std::endl;
        std::exit(EXIT_FAILURE);
    }
    return fsize;
}

inline void burst_data(std::vector<int64_t>& istart, std::vector<int64_t>& istop, std::vector<int64_t>& times_ms,
                       std::vector<unsigned char>& mask_ad, std::vector<unsigned char>& mask_dd,
                       std::vector<int>& phCount, int64_t** burst_ad, int64_t** burst_dd)
{
    const size_t sz_burst = istart.size();
    if (phCount.size() != sz_burst) phCount.resize(sz_burst);
    // istart / istop come from the file: every burst must lie inside the photon arrays (all three
    // of them) and its count must fit the int the reference keeps it in, before anything is indexed
    const int64_t n_ph = (int64_t)times_ms.size();
    if (istop.size() != sz_burst || (int64_t)mask_ad.size() != n_ph || (int64_t)mask_dd.size() != n_ph) {
        std::cerr << "burst_data: istart/istop or times/mask_ad/mask_dd differ in length" << std::endl;
        std::exit(EXIT_FAILURE);
    }
    for (size_t idx = 0; idx < sz_burst; idx++) {
        const int64_t a = istart[idx], z = istop[idx];
        if (a < 0 || z < a - 1 || z >= n_ph || z - a + 1 > 0x7fffffff) {
            std::cerr << "burst_data: burst " << idx << " claims photons [" << a << ", " << z << "] of " << n_ph << std::endl;
            std::exit(EXIT_FAILURE);
        }
        phCount[idx] = (int)(z - a + 1);
    }
    // value-initialised: photons between bursts (not covered by any istart..istop) read as 0,
    // where the reference leaves them uninitialised (loadHdf5.cpp:112-113)
    *burst_ad = new int64_t[times_ms.size()]();
    *burst_dd = new int64_t[times_ms.size()]();
    for (size_t idx = 0; idx < sz_burst; idx++) {
        const int64_t o = istart[idx];
        for (int i = 0; i < phCount[idx]; ++i) {
            (*burst_ad)[o + i] = (int64_t)mask_ad[(size_t)(o + i)] * times_ms[(size_t)(o + i)];
            (*burst_dd)[o + i] = (int64_t)mask_dd[(size_t)(o + i)] * times_ms[(size_t)(o + i)];
        }
    }
}
