// blocking.cpp -- LD blocking of one chromosome from the row sums of its banded |correlation| matrix
// (`mps block`).  Host only: O(m * window) multiply-adds per window tried, a handful of windows.
//
// Behaviour follows the reference's src/blocking.cpp:
//   hann / hanning_smoothing  :8-36   Hann window (cosine in single precision), valid part of the convolution,
//                                     zeros in the margins
//   local_minima              :38-55  a minimum needs a larger value since the previous minimum on its left
//   blocks_from_minima        :57-70  [start, minimum] ... last block ends at the last marker
//   block_chr                 :97-136 bisection on the odd window size in [3, m] until the largest block is at
//                                     most max_block_size and within 100 of it
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "host.hpp"

namespace cuskhost {

namespace {

const int BLOCK_SIZE_SLACK = 100;  // MAX_BLOCK_SIZE_TOL, src/blocking.cpp:6

inline int odd_at_most(int v) { return (v & 1) ? v : v - 1; }

}  // namespace

std::vector<double> hanning_smoothing(const float *v, size_t nv, int window)
{
    std::vector<double> out(nv, 0.0);
    if (window <= 0) return out;
    std::vector<double> w((size_t)window);
    const double denom = (double)window - 1.0;  // window 1: 0/0, NaN weights, exactly as the reference
    for (int i = 0; i < window; i++) w[(size_t)i] = 0.5 - 0.5 * cosf(2.0 * M_PI * (double)i / denom);
    const long long margin = window / 2, stop = (long long)nv - margin;
#pragma omp parallel for schedule(static) if (nv * (size_t)window > (1u << 20))
    for (long long c = margin; c < stop; c++) {
        const float *src = v + (c - margin);
        double acc = 0.0;
        for (int i = 0; i < window; i++) acc += w[(size_t)i] * (double)src[i];  // same order as the reference
        out[(size_t)c] = acc;
    }
    return out;
}

BlockPartition blocks_for_window(const float *row_sums, size_t m, int window)
{
    BlockPartition bp;
    bp.window = window;
    const std::vector<double> s = hanning_smoothing(row_sums, m, window);
    size_t start = 0;
    double high = 0.0;  // largest smoothed value seen since the last cut
    for (size_t i = 1; i + 1 < m; i++) {
        const double x = s[i];
        if (high > x && x < s[i + 1]) {
            bp.first.push_back(start);
            bp.last.push_back(i);
            start = i + 1;
            high = 0.0;
        } else if (x > high) {
            high = x;
        }
    }
    bp.first.push_back(start);
    bp.last.push_back(m - 1);
    for (size_t b = 0; b < bp.first.size(); b++) bp.largest = std::max(bp.largest, bp.last[b] - bp.first[b] + 1);
    return bp;
}

BlockPartition block_chr(const float *row_sums, size_t m, int max_block_size)
{
    int hi = (int)m, lo = 3;  // a window of hi gives blocks that are too large, one of lo too small
    int window = odd_at_most((hi + lo) / 2);
    BlockPartition bp = blocks_for_window(row_sums, m, window);
    for (;;) {
        const long long largest = (long long)bp.largest;
        const bool over = largest > max_block_size;
        if (!over && std::llabs(largest - (long long)max_block_size) <= BLOCK_SIZE_SLACK) break;
        if (over) hi = std::min(hi, window);
        else lo = std::max(lo, window);
        const int next = odd_at_most((hi + lo) / 2);
        if (next == window) break;
        window = next;
        bp = blocks_for_window(row_sums, m, window);
    }
    return bp;
}

}  // namespace cuskhost
