// banded_core.h -- arithmetic core of the banded marker x marker correlations (`mps block`),
// shared by the CUDA kernels (corr_banded.cu) and by a host compile in the tests (tests/banded_host_core.cpp),
// so that the bit tricks are checked on CPU against the oracle's sample-by-sample contingency tables.
//
// A BED column (include/mps/bed_lut_gpu.h:3,1030: code 00 -> 2, 01 -> missing, 10 -> 1, 11 -> 0; sample j in bits
// 2(j&3).. of byte j>>2) is re-expressed as three INDICATOR BIT PLANES, one bit per sample: plane g has a 1 where the
// genotype is g.  Samples are permuted inside every 32-sample word (the same permutation for every marker, so
// joint counts are unaffected); samples >= n and missing calls are 0 in all three planes.  The 3x3 table of a
// marker pair (src/corr_kernels.cu:920-948, valid-valid samples only) is then s[3a+b] = popcount(plane_a & plane_b).
#pragma once
#include <cstddef>
#include <cstdint>

#if defined(__CUDACC__)
#define BANDC_HD __host__ __device__
#else
#define BANDC_HD
#endif

// Plane words of 32-sample word `w` of one column: out[g], g = 0, 1, 2.
BANDC_HD inline void bandc_planes(const unsigned char *col, size_t col_bytes, size_t n, size_t w, uint32_t out[3])
{
    unsigned long long x = 0;
    for (int b = 0; b < 8; b++) {
        const size_t at = 8 * w + (size_t)b;
        if (at < col_bytes) x |= (unsigned long long)col[at] << (8 * b);
    }
    const unsigned long long M5 = 0x5555555555555555ull;
    const size_t s0 = 32 * w;
    unsigned long long live = 0;  // even bit 2j set for samples s0 + j < n
    if (s0 < n) {
        const size_t cnt = n - s0;
        live = cnt >= 32 ? M5 : (((1ull << (2 * cnt)) - 1ull) & M5);
    }
    const unsigned long long lo = x & M5, hi = (x >> 1) & M5;
    const unsigned long long i2 = ~lo & ~hi & live;  // 00
    const unsigned long long i1 = ~lo & hi & live;   // 10
    const unsigned long long i0 = lo & hi & live;    // 11
    // the indicators sit on even bits: fold the upper 16 samples onto the odd bits of the lower half
    out[0] = (uint32_t)i0 | ((uint32_t)(i0 >> 32) << 1);
    out[1] = (uint32_t)i1 | ((uint32_t)(i1 >> 32) << 1);
    out[2] = (uint32_t)i2 | ((uint32_t)(i2 >> 32) << 1);
}

// Full table from the four joint counts of genotypes {1,2} x {1,2} when NEITHER marker has a missing call:
// the remaining five cells follow from the markers' own genotype counts ca/cb = (n0, n1, n2, nmissing = 0).
BANDC_HD inline void bandc_table_from_joint(long long s11, long long s12, long long s21, long long s22, const int *ca,
                                            const int *cb, long long n, double s[9])
{
    const long long s10 = ca[1] - s11 - s12, s20 = ca[2] - s21 - s22;
    const long long s01 = cb[1] - s11 - s21, s02 = cb[2] - s12 - s22;
    const long long s00 = n - (s01 + s02 + s10 + s11 + s12 + s20 + s21 + s22);
    s[0] = (double)s00;
    s[1] = (double)s01;
    s[2] = (double)s02;
    s[3] = (double)s10;
    s[4] = (double)s11;
    s[5] = (double)s12;
    s[6] = (double)s20;
    s[7] = (double)s21;
    s[8] = (double)s22;
}

// Tensor-core route (CUSK_BLOCK_TC=1): tiles of the indicator Gram kernel (corr_gram.cu) that hold at least one pair
// of the band.  Markers sit MI to a 128-row group; tile (tm, tn) pairs the markers of group tm (A side) with those of
// groups 2 tn, 2 tn + 1 (B side).  A pair (i, j), i < j, lies in exactly one tile (tm = i / MI, tn = j / (2 MI)), so
// every band pair 0 < j - i <= W with i < R, j < Mc is computed once when the tiles below are run.
// Returns the number of tiles; writes (tm, tn) pairs to out when out != nullptr (capacity cap, in tiles).
inline size_t bandc_gram_tiles(size_t Mc, size_t R, size_t W, size_t MI, int *out, size_t cap)
{
    const size_t groups = (Mc + MI - 1) / MI, n_tn = (groups + 1) / 2;
    size_t cnt = 0;
    for (size_t tm = 0; tm < groups; tm++) {
        const size_t a_min = tm * MI;
        if (a_min >= R) break;
        size_t a_max = a_min + MI - 1;
        if (a_max > R - 1) a_max = R - 1;
        // B blocks holding some j in (a_min, a_max + W]
        const size_t tn_lo = (a_min + 1) / (2 * MI);
        size_t j_hi = a_max + W;
        if (j_hi > Mc - 1) j_hi = Mc - 1;
        const size_t tn_hi = j_hi / (2 * MI);
        for (size_t tn = tn_lo; tn <= tn_hi && tn < n_tn; tn++) {
            if (out && cnt < cap) {
                out[2 * cnt] = (int)tm;
                out[2 * cnt + 1] = (int)tn;
            }
            cnt++;
        }
    }
    return cnt;
}
