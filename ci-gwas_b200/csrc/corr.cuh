// corr.cuh -- shared definitions of the correlation front end (BED decode, marker statistics,
// indicator Gram on tcgen05, marker x trait and trait x trait correlations).
#pragma once
#include "common.cuh"

// PLINK BED 2-bit codes (include/mps/bed_lut_gpu.h:3,1030): 00 -> genotype 2, 01 -> missing,
// 10 -> genotype 1, 11 -> genotype 0; sample j of a marker sits in bits 2(j&3).. of byte j>>2.

#define CUSK_E_UNSUPPORTED_PAIR (-1000)  // internal: launch_gram_pair cannot place the grid; the caller takes the one-CTA form

struct GramParams {
    int parts;        // indicator rows per marker: 2 (no missing call in the block) or 3
    int MI;           // markers per 128-row group: 64 (parts 2) or 42 (parts 3)
    int m;            // markers
    int n_samples;
    int num_k;        // K stages of 128 bytes
    int ntiles;
    const int2 *tiles;   // (tm, tn): A rows [128 tm, +128), B rows [256 tn, +256)
    const int *counts;   // [m][4]: n0, n1, n2, nmissing
    float *C;            // dense output, row-major, leading dimension ldc; both triangles written
    int ldc;
    int mode;            // 0: sin(pi/2 tau_b) (cusk), 1: Pearson with the given means / stds
    const float *mean, *stdv;
    long long *tables;   // optional [m*(m-1)/2][9] exact contingency tables (tests), or nullptr
    // marker x trait on the same tiles: B block `trait_tn` of the operand matrix holds, per trait t, the
    // rows 8t..8t+3 = byte limbs of the trait values in offset fixed point (u = round(y 2^S) + 2^31, 0 where
    // y is NaN) and row 8t+4 = [y not NaN]; -1: no trait block
    int trait_tn;
    int p;                      // traits (<= 32)
    const double *trait_scale;  // [p] 2^-S
    const double *Ty;           // [p] sum of valid y
    const int *cnty;            // [p] number of valid y
    float *mxp;                 // [m*p] out
    // banded output (`mps block`, corr_banded.cu): when band != nullptr the pairs 0 < j - i <= band_w with
    // i < band_rows go to band[i * band_w + (j - i - 1)] and nothing is written to C
    float *band;
    int band_w, band_rows;
    // level 0 of the individual-level skeleton fused into the epilogue (north_star item 2; cal_Indepl0,
    // src/cuPC-S.cu:458-484): when l0_bits != nullptr every marker pair whose |atanh r| is not below l0_th sets its
    // two bits of the adjacency bitmap (l0_rw words per row, zeroed by the caller), so that the skeleton engine
    // starts at level 1 without reading the matrix back
    uint32_t *l0_bits;
    int l0_rw;
    double l0_th;
    // K-panel synchronisation of a wave (the `grid` consecutive tiles that run together): every ksync k-steps the TMA
    // producers of the wave's CTAs meet at a counter, so that the CTAs sharing A / B panels read them within the same
    // L2 window.  wave_tiles[w] = real tiles of wave w (entries with tile.x < 0 are padding), sync[w * ksync_blocks +
    // kb / ksync] the counters (zeroed by the caller); ksync == 0: off
    int ksync, ksync_blocks;
    const int *wave_tiles;
    int *sync;
};

struct Level0Fuse {  // corr_block_device: the level-0 adjacency comes out of the correlation pass
    uint32_t *bits;  // [nvar][rw], written whole
    int rw;
    double th;       // Th[0] of threshold_array
};

// the level-0 statistic exactly as level0_kernel evaluates it (skeleton_kernels.cu)
__host__ __device__ inline double corr_fisher_z(double rho) { return fabs(0.5 * log(fabs((1.0 + rho) / (1.0 - rho)))); }

#define CUSK_TC_MXP_MAX_TRAITS 32

// row of (marker i, part a) in the padded indicator matrix
__host__ __device__ inline size_t ind_row(int i, int a, int parts, int MI)
{
    return (size_t)(i / MI) * 128 + (size_t)(i % MI) * parts + a;
}

// src/corr_kernels.cu:548-563: Kendall tau_b from the 3x3 table s[3a+b], then sin(pi/2 tau).
__host__ __device__ inline double npn_from_table(const double *s)
{
    const double p = s[0] * (s[4] + s[5] + s[7] + s[8]) + s[1] * (s[5] + s[8]) + s[3] * (s[7] + s[8]) + s[4] * s[8];
    const double q = s[1] * (s[3] + s[6]) + s[2] * (s[3] + s[4] + s[6] + s[7]) + s[4] * s[6] + s[5] * (s[6] + s[7]);
    const double t = s[0] * (s[1] + s[2]) + s[1] * s[2] + s[3] * (s[4] + s[5]) + s[4] * s[5] + s[6] * (s[7] + s[8]) +
                     s[7] * s[8];
    const double u = s[0] * (s[3] + s[6]) + s[1] * (s[4] + s[7]) + s[2] * (s[5] + s[8]) + s[3] * s[6] + s[4] * s[7] +
                     s[5] * s[8];
    const double kendall = (p - q) / sqrt((p + q + t) * (p + q + u));
    return sin(1.57079632679489661923 * kendall);
}

// src/corr_kernels.cu:346-403 (bed_marker_corr_pearson): (sum v a b / sum v - mu_a mu_b) / (sd_a sd_b)
__host__ __device__ inline double pearson_from_table(const double *s, double mu_a, double mu_b, double sd_a,
                                                     double sd_b)
{
    const double sab = s[4] + 2.0 * (s[5] + s[7]) + 4.0 * s[8];
    double nv = 0;
    for (int i = 0; i < 9; i++) nv += s[i];
    return (sab / nv - mu_a * mu_b) / (sd_a * sd_b);
}
