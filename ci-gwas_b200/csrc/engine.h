// engine.h -- internal C++ entry points of the skeleton engines (device-resident inputs).
#pragma once
#include <vector>

#include "common.cuh"

// Individual-level engine.  seps != nullptr: all separating-set records, sorted by (x, y).
// level0_done: dbits already holds the level-0 adjacency (the correlation pass wrote it, corr.cuh Level0Fuse); the
// engine then starts at level 1 (not with d_pMax: the level-0 statistics of the removed pairs are not kept)
int skeleton_engine(cusk_ctx *ctx, const float *dC, int n, uint32_t *dbits, const float *Th, int maxlevel,
                    int *l_out, float *d_pMax, std::vector<cusk_sepset> *seps, cusk_stats *stats,
                    bool level0_done = false);

// Summary-statistic engine.  Candidate edges: dG_dense (int n*n, ==1 is a candidate), or the
// bitmap itself (bits_are_input), or every pair when both are absent.  dN == nullptr: uniform
// effective sample size n_uniform.
int hetcor_engine(cusk_ctx *ctx, const float *dC, int n, uint32_t *dbits, const int *dG_dense, bool bits_are_input,
                  const float *dN, float n_uniform, float th, int maxlevel, const int *d_time, int *l_out,
                  cusk_stats *stats);

void launch_expand_bits(cudaStream_t st, const uint32_t *bits, int n, int rw, int *G);

// correlation front end on device pointers (corr.cu); scratch from ctx->arena (caller resets it)
int corr_block_device(cusk_ctx *ctx, const unsigned char *d_bed, const float *d_phen, size_t m, size_t n, size_t p,
                      const float *d_mean, const float *d_std, int mode, float *d_Csq, float *d_mxp_out,
                      float *d_pxp_out, long long *d_tables, bool want_mxm, cusk_corr_stats *st, bool mxp_ready,
                      const struct Level0Fuse *l0 = nullptr);
void launch_bed_counts(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int *counts);
void launch_bed_stats(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, const int *counts,
                      float *means, float *stds, int *modes);
