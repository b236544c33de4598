// TEST INFRASTRUCTURE, never part of the product: compiles the bit-plane routines of the banded correlation kernel
// (ci-gwas_b200/csrc/banded_core.h) for the HOST, so that tests/test_block_cpu.py can check them against the oracle's
// sample-by-sample contingency tables in a container without a GPU.  The loop below does per marker pair what
// banded_npn_kernel (corr_banded.cu) does per thread: AND + popcount over the plane words, nine joint counts when a
// marker has a missing call, otherwise four joint counts completed from the markers' own genotype counts.
#include <vector>

#include "../ci-gwas_b200/csrc/banded_core.h"

extern "C" void bandc_host_pair_table(const unsigned char *bed, size_t n, size_t a, size_t b, int force_general,
                                      long long *s9, int *used_general)
{
    const size_t col_bytes = (n + 3) / 4, kw = (n + 31) / 32 + 3;  // a few padding words: they must come out empty
    std::vector<uint32_t> pa(3 * kw), pb(3 * kw);
    int ca[4] = {0, 0, 0, 0}, cb[4] = {0, 0, 0, 0};
    for (size_t w = 0; w < kw; w++) {
        uint32_t x[3], y[3];
        bandc_planes(bed + a * col_bytes, col_bytes, n, w, x);
        bandc_planes(bed + b * col_bytes, col_bytes, n, w, y);
        for (int g = 0; g < 3; g++) {
            pa[g * kw + w] = x[g];
            pb[g * kw + w] = y[g];
            ca[g] += __builtin_popcount(x[g]);
            cb[g] += __builtin_popcount(y[g]);
        }
    }
    ca[3] = (int)n - ca[0] - ca[1] - ca[2];
    cb[3] = (int)n - cb[0] - cb[1] - cb[2];
    const bool general = force_general || ca[3] != 0 || cb[3] != 0;
    *used_general = general;
    if (general) {
        for (int ga = 0; ga < 3; ga++)
            for (int gb = 0; gb < 3; gb++) {
                long long c = 0;
                for (size_t w = 0; w < kw; w++) c += __builtin_popcount(pa[ga * kw + w] & pb[gb * kw + w]);
                s9[3 * ga + gb] = c;
            }
        return;
    }
    long long j[2][2] = {{0, 0}, {0, 0}};
    for (int ga = 1; ga < 3; ga++)
        for (int gb = 1; gb < 3; gb++)
            for (size_t w = 0; w < kw; w++) j[ga - 1][gb - 1] += __builtin_popcount(pa[ga * kw + w] & pb[gb * kw + w]);
    double s[9];
    bandc_table_from_joint(j[0][0], j[0][1], j[1][0], j[1][1], ca, cb, (long long)n, s);
    for (int i = 0; i < 9; i++) s9[i] = (long long)s[i];
}

// tile list of the tensor-core route: number of tiles, (tm, tn) pairs into out
extern "C" size_t bandc_host_gram_tiles(size_t Mc, size_t R, size_t W, size_t MI, int *out, size_t cap)
{
    return bandc_gram_tiles(Mc, R, W, MI, out, cap);
}
