// oracle/ref_host_half.cpp — TEST INFRASTRUCTURE, never the product.
//
// The HOST half of one reference evaluation (/root/reference/src/streamWorker.cpp:33-38,62-67,
// 187-206): copy hmcE into a vector, fill a boost::histogram regular axis [0,1) with fretHistNum
// bins (the reference's vendored boost), read the inner bins back, reduced chi-square against the
// observed histogram — K times, seconds returned.  This is what the reference pays per evaluation
// on top of kernel + D2H; bench.py adds it to the reference-on-the-GPU baseline.
#include <boost/histogram.hpp>
#include <chrono>
#include <cmath>
#include <cstddef>
#include <vector>

namespace bh = boost::histogram;

static std::vector<float> inner_bins(const std::vector<float>& v, int bins)
{
    auto hist = bh::make_histogram(bh::axis::regular<>(bins, 0, 1));
    for (float x : v) hist(x);
    std::vector<float> out((size_t)bins);
    int i = 0;
    for (auto&& x : bh::indexed(hist, bh::coverage::inner)) out[(size_t)i++] = (float)*x;
    return out;
}

extern "C" double ref_host_half(const float* mcE, size_t count, const float* sg, size_t n_sg, int R, int fret_hist_num,
                                int left_pad, int n_states, int K, float* chi2_out)
{
    const std::vector<float> observed = inner_bins(std::vector<float>(sg, sg + n_sg), fret_hist_num);
    float chisqr = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < K; ++k) {
        std::vector<float> sample(mcE, mcE + count);
        const std::vector<float> model = inner_bins(sample, fret_hist_num);
        int effN = fret_hist_num - left_pad;
        chisqr = 0;
        for (int b = left_pad; b < fret_hist_num; ++b) {
            if (model[(size_t)b] > 0)
                chisqr += std::pow((float(observed[(size_t)b] - model[(size_t)b] / R)), 2) / (float(model[(size_t)b]) / R);
            else
                effN--;
        }
        chisqr = chisqr / (effN - n_states * (n_states + 1));
    }
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (chi2_out) *chi2_out = chisqr;
    return s;
}
