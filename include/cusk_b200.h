/* cusk_b200.h -- C ABI of libcusk_b200.so, the B200 (sm_100a) implementation of the CI-GWAS
 * skeleton hot path.  Plain pointers and sizes only; no C++ or torch types.
 *
 * Every entry point names the reference interface it replaces (paths relative to
 * /root/reference/cusk/).  Two families:
 *
 *   1. the reference's own host-function signatures (drop-in: same names, same argument
 *      meaning, host buffers in / host buffers out, device 0, blocking), and
 *   2. an explicit-context API (device id, reusable workspace, status codes, statistics,
 *      sparse separating-set output, device-resident inputs) which the drop-in wrappers and
 *      the `mps`-compatible host binary are built on.
 *
 * Error convention: family 2 returns 0 on success and a negative CUSK_E_* code otherwise;
 * cusk_last_error() returns a thread-local message.  Family 1 mirrors the reference
 * (include/mps/gpuerrors.h:6-35): print the message to stderr and exit(EXIT_FAILURE).
 * There is no CPU fallback anywhere: without a CUDA device every compute call fails.
 */
#ifndef CUSK_B200_H_
#define CUSK_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUSK_ML 14 /* include/mps/cuPC-S.h:49  (#define ML 14) */

enum {
    CUSK_OK = 0,
    CUSK_E_CUDA = -1,       /* a CUDA runtime call failed */
    CUSK_E_ARG = -2,        /* invalid argument */
    CUSK_E_INFEASIBLE = -3, /* a level needs >= 2^62 conditioning sets for one row */
    CUSK_E_NODEVICE = -4    /* no CUDA device / wrong architecture */
};

typedef struct cusk_ctx cusk_ctx; /* opaque: one device, one stream, reusable workspace */

/* Per-call statistics (all levels 0..14).  tests_canonical[l] = sum_X C(d_X,l)(d_X-l), the
 * schedule independent full PC-stable enumeration; tests_executed[l] = CI tests
 * actually evaluated by this implementation (early termination skips the rest). */
typedef struct cusk_stats {
    int64_t tests_canonical[CUSK_ML + 1];
    int64_t tests_executed[CUSK_ML + 1];
    float level_ms[CUSK_ML + 1]; /* device time of each level (CUDA events) */
    float ci_kernel_ms[CUSK_ML + 1]; /* device time of the level's CI-test kernel alone */
    int32_t max_degree[CUSK_ML + 1]; /* max degree at the START of level l (l >= 1) */
    int64_t edges_after[CUSK_ML + 1]; /* undirected edges alive after level l */
    int64_t singular_pivots; /* conditioning sets with a (near-)singular leading block */
    int32_t last_level;      /* == *l of the reference API */
    int32_t levels_run;      /* number of levels that executed CI tests (incl. level 0) */
    float total_ms;          /* device time, level 0 .. last level, excluding copies */
    float h2d_ms, d2h_ms;    /* copy times of the host-buffer entry points */
    int64_t h2d_bytes, d2h_bytes; /* bytes the host-buffer entry points actually moved over PCIe (an all-ones
                                     candidate matrix and a constant N are recognised on the host and not uploaded;
                                     the adjacency comes back as the bitmap and is expanded on the host) */
    int64_t kernel_launches;      /* kernels the engine launched for this call (counted at the launch sites) */
} cusk_stats;

/* One separating set of the individual-level engine: slot [x,y] of the reference's dense
 * SepSet array (src/cuPC-S.cu:568-577), in sparse form. */
typedef struct cusk_sepset {
    int32_t x, y;
    int32_t level;        /* |S| */
    float pmax;           /* Z statistic of the recorded test (pMax[x*n+y]) */
    int32_t s[CUSK_ML];   /* variable indices, ascending, -1 padded */
} cusk_sepset;

/* ---- context ---------------------------------------------------------------------------- */
int cusk_ctx_create(int device_id, cusk_ctx **out);
void cusk_ctx_destroy(cusk_ctx *ctx);
const char *cusk_last_error(void);
const char *cusk_version(void);
int cusk_device_count(void);
/* the CUDA stream the context launches on (a cudaStream_t), for callers that time with events */
void *cusk_ctx_stream(cusk_ctx *ctx);
int cusk_ctx_sync(cusk_ctx *ctx);
/* Per-level instrumentation (cusk_stats.level_ms / ci_kernel_ms: three event records per level between the launches).
 * OFF by default -- the records cost ~4.5 us each, 0.2 ms of the 1.1 ms C2 step -- (CUSK_LEVEL_TIMING=1 turns it on for
 * new contexts); total_ms and the counters are always filled.
 * The reference's counterpart are its per-level "spent seconds" host timers (src/cuPC-S.cu:139-159). */
void cusk_ctx_set_level_timing(cusk_ctx *ctx, int on);

/* ---- thresholds  (replaces src/cuPC_call_prep.cpp:13-28) --------------------------------- */
/* Th[l] = |qnorm(alpha/2)| / sqrt(n-l-3), l = 0..14, same float/double type flow. */
void cusk_threshold_array(int n, float alpha, float *out15);
float cusk_hetcor_threshold(float alpha);

/* ---- skeleton, individual-level engine  (replaces Skeleton(), src/cuPC-S.cu:61-450) ------ */
/* Host buffers.  C[n*n] in; G[n*n] out (0/1, symmetric, input ignored); Th[15];
 * pMax[n*n] out or NULL; SepSet[n*n*14] out (-1 padded) or NULL; *l_out = last level.
 * sep/sep_cap/sep_n: optional sparse separating-set list (NULL/0/NULL to skip). */
int cusk_skeleton(cusk_ctx *ctx, const float *C, int n, int *G, const float *Th, int maxlevel,
                  int *l_out, float *pMax, int *SepSet, cusk_sepset *sep, int64_t sep_cap,
                  int64_t *sep_n, cusk_stats *stats);

/* ---- skeleton, summary-statistic engine  (replaces hetcor_skeleton(),
 *      src/hetcor-cuPC-S.cu:75-341) -------------------------------------------------------- */
/* Host buffers.  G[n*n] in/out (1 = candidate edge, only cleared), or G_in_all_ones != 0 to
 * ignore the input values; N[n*n] effective sample sizes (may hold NaN); th = |qnorm(a/2)|;
 * time_index[n] or NULL (all equal). */
int cusk_hetcor_skeleton(cusk_ctx *ctx, const float *C, int n, int *G, int G_in_all_ones,
                         const float *N, float th, int maxlevel, const int *time_index,
                         int *l_out, cusk_stats *stats);

/* Device-resident variants: dC, dN are device pointers (float, n*n); dG_bits is the bit-packed
 * adjacency in/out on the device: n rows of cusk_row_words(n) uint32 words, bit (y & 31) of word
 * y >> 5 in row x is edge x-y.  Nothing is copied to or from the host except the statistics.
 * These are what bench.py's device-resident `value` times, and what corr -> skeleton uses. */
size_t cusk_row_words(int n);
int cusk_hetcor_skeleton_dev(cusk_ctx *ctx, const float *dC, int n, uint32_t *dG_bits,
                             int G_in_all_ones, const float *dN, float n_uniform, float th,
                             int maxlevel, const int *d_time_index, int *l_out,
                             cusk_stats *stats);
int cusk_skeleton_dev(cusk_ctx *ctx, const float *dC, int n, uint32_t *dG_bits, const float *Th,
                      int maxlevel, int *l_out, float *d_pMax, cusk_sepset *sep, int64_t sep_cap,
                      int64_t *sep_n, cusk_stats *stats);

/* ---- correlation front end ------------------------------------------------------------------ */
/* bed: packed PLINK SNP-major genotype columns WITHOUT the 3 magic bytes, m * ceil(n/4) bytes
 * (what read_block_from_bed returns, src/io.cpp:238); phen: float [p][n] column-major per trait,
 * NaN = missing (src/phen.cpp:27-74); mean/std: per-marker values as read from .means/.stds. */
typedef struct cusk_corr_stats {
    float h2d_ms, counts_ms, decode_ms, gram_ms, mxp_ms, pxp_ms, d2h_ms, total_ms;
    int32_t parts;        /* indicator rows per marker in the tensor-core Gram: 2 or 3 */
    int32_t gram_tiles;   /* 128 x 256 indicator-row tiles computed (upper triangle) */
    int64_t gram_int8_ops;/* 2 * 128 * 256 * K_padded per tile, summed */
    int64_t bed_bytes;    /* packed genotype bytes of the block */
} cusk_corr_stats;

/* prep-bed statistics (replaces compute_bed_col_stats_no_impute, src/prep.cpp:15-77): means, stds
 * (float, population, missing excluded, the reference's sequential float accumulation order),
 * modes, and exact genotype counts counts4[m][4] = n0, n1, n2, nmissing (may be NULL). */
int cusk_bed_marker_stats(cusk_ctx *ctx, const unsigned char *bed, size_t m, size_t n, float *means,
                          float *stds, int *modes, int64_t *counts4);

/* Allocation-free variant of cusk_bed_marker_stats.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * d_bed [m][ceil(n/4)], d_means[m], d_stds[m], d_modes[m] and the optional d_counts4[m][4] (int32: n0, n1, n2, nmissing;
 * may be NULL) are DEVICE buffers of the caller on the current device; `stream` is a cudaStream_t (NULL = the default
 * stream).  The work is queued on `stream` and the call returns without synchronising.  The only scratch, the counts
 * table between the two kernels, is a fixed __device__ array inside the library (1 MiB per GPU, held until exit), so
 * m <= CUSK_NOALLOC_MAX_MARKERS (CUSK_E_ARG beyond it) and no cudaMalloc / cudaFree / allocating library call is made.
 * All calls share that array: calls that overlap on the GPU (two streams, concurrent graph replays) race -- serialise
 * them.  The first call on a device looks the array's address up (cudaGetSymbolAddress); make it outside a capture. */
#define CUSK_NOALLOC_MAX_MARKERS 65536
int cusk_bed_marker_stats_noalloc(const unsigned char *d_bed, size_t m, size_t n, float *d_means, float *d_stds,
                                  int *d_modes, int *d_counts4, void *stream);

/* marker x trait Pearson (replaces cu_marker_phen_corr_pearson, src/corr_host.cu:1023). mxp[m*p] */
int cusk_marker_phen_corr(cusk_ctx *ctx, const unsigned char *bed, const float *phen, size_t m, size_t n,
                          size_t p, const float *mean, const float *std, float *mxp,
                          cusk_corr_stats *stats);

/* all three correlation sets of a block (replaces cu_corr_pearson_npn, src/corr_host.cu:1094):
 * mxm[m(m-1)/2] upper triangle row-major, sin(pi/2 tau_b) from the 3x3 genotype table (mode 0) or
 * Pearson with the given means/stds (mode 1; bed_marker_corr_pearson, src/corr_kernels.cu:346);
 * mxp[m*p]; pxp[p(p-1)/2].  Any output may be NULL.  tables (may be NULL): the exact contingency
 * tables [m(m-1)/2][9] the correlations were computed from (integer dot products; for tests). */
int cusk_corr_block(cusk_ctx *ctx, const unsigned char *bed, const float *phen, size_t m, size_t n,
                    size_t p, const float *mean, const float *std, int mode, float *mxm, float *mxp,
                    float *pxp, int64_t *tables, cusk_corr_stats *stats);

/* Device-resident variant: d_bed, d_phen, d_mean, d_std are device pointers; writes the dense
 * symmetric (m+p) x (m+p) correlation matrix d_Csq (unit diagonal, markers first then traits --
 * the layout cli.cpp:597-649 assembles on the host) that cusk_skeleton_dev consumes directly. */
int cusk_corr_block_dev(cusk_ctx *ctx, const unsigned char *d_bed, const float *d_phen, size_t m,
                        size_t n, size_t p, const float *d_mean, const float *d_std, int mode,
                        float *d_Csq, cusk_corr_stats *stats);

/* ---- LD blocking (`mps block`) ----------------------------------------------------------
 * cusk_mcorrk_banded replaces cal_mcorrk_banded (src/corr_host.cu:64-110; launch loop
 * cu_marker_corr_pearson_npn_batched_sparse, :1199-1319) and marker_corr_banded_mat_row_abs_sums (:112-128):
 * corrs[row * corr_width + k] = sin(pi/2 tau_b) of markers (row, row + 1 + k), 0 where row + 1 + k >= m;
 * row_abs_sums[row] = sum_k |corrs[row][k]| accumulated in float, left to right.  Either output may be NULL.
 * bed: the chromosome's packed columns (read_chr_from_bed, src/io.cpp:251).  device_mem_gb > 0 bounds the device
 * memory used (the reference's <device-mem-gb>; 0 = what is free): the chromosome is then processed in chunks of
 * rows.  Needs 0 < corr_width < m (the reference exits or reads out of bounds otherwise). */
typedef struct cusk_band_stats {
    float h2d_ms, kernel_ms, d2h_ms, total_ms; /* summed over chunks (CUDA events) */
    int32_t chunks;
    int64_t pairs;     /* marker pairs inside the band */
    int64_t bed_bytes; /* packed genotype bytes copied to the device (chunk overlaps included) */
} cusk_band_stats;
int cusk_mcorrk_banded(cusk_ctx *ctx, const unsigned char *bed, size_t m, size_t n, size_t corr_width,
                       float device_mem_gb, float *corrs, float *row_abs_sums, cusk_band_stats *stats);
/* host only.  hanning_smoothing (src/blocking.cpp:13-36): out[nv] doubles.  block_chr (src/blocking.cpp:97-136):
 * blocks [first[i], last[i]] (0-based, inclusive) of one chromosome from its row sums; first/last need room for m
 * entries; returns the number of blocks (< 0 on error), *window (may be NULL) the Hann window finally used. */
int cusk_hanning_smoothing(const float *v, size_t nv, int window_size, double *out);
int64_t cusk_block_chr(const float *row_sums, size_t m, int max_block_size, int64_t *first, int64_t *last,
                       int *window);
/* make_blocks (src/cli.cpp:362-411) = `mps block <bfiles> <max-block-size> <device-mem-gb> <corr-width>`:
 * appends every chromosome's blocks to <bfiles>_m<max-block-size>.blocks ("chr\tfirst\tlast" lines). */
int cusk_make_blocks(cusk_ctx *ctx, const char *bfiles, int max_block_size, float device_mem_gb, size_t corr_width);

/* ---- multi-GPU: one skeleton sharded over several GPUs (cuskss --marker-indices) ------
 * The level-l work items are dealt round-robin to the ranks; after each level the ranks exchange what
 * they removed (NCCL all-gather of one removed-flag byte per CSR slot for the summary-statistic engine,
 * all-reduce MIN of the separating-set ranks for the individual-level engine), so every rank returns
 * the same graph as a single-GPU call.  The reference has no multi-GPU path (nothing to cite). */
typedef struct cusk_comm cusk_comm;
int cusk_comm_unique_id(char *out128);                       /* ncclGetUniqueId; share it with all ranks */
int cusk_comm_create(cusk_ctx *ctx, int rank, int world, const char *id128, cusk_comm **out); /* one process per GPU */
int cusk_comm_create_all(cusk_ctx **ctxs, int world, cusk_comm **out);   /* one process, NCCL, distinct devices */
int cusk_comm_create_local(cusk_ctx **ctxs, int world, cusk_comm **out); /* one process, ranks on ONE device (testing); CUSK_E_ARG for distinct devices */
void cusk_comm_destroy(cusk_comm *c);
int cusk_ctx_set_comm(cusk_ctx *ctx, cusk_comm *c);          /* attach (NULL detaches): later skeleton calls are sharded */
int cusk_comm_stats(const cusk_comm *c, double *exchange_ms, int64_t *exchange_bytes); /* of the last call */
int cusk_nccl_version(void);                                 /* 0 when libnccl cannot be loaded */

/* ---- host layer: the `mps` process contract (apps/mps.cpp:17-121, src/cli.cpp:194,432,696) --------
 * argv exactly as the reference's binary takes it (`prep`, `cusk`, `cuskss`; plus `cusk-all`,
 * `cuskss-all`); returns the process exit status.  prep_bed_no_impute (src/prep.cpp:157) is
 * `mps prep <bfiles>`. */
int cusk_mps_main(int argc, char **argv);
int cusk_prep_bed(cusk_ctx *ctx, const char *bfiles);
/* host-only pieces (no device needed): the cuskss input assembly make_square_cuskss_inputs
 * (src/cli.cpp:89-173) over the loaders of src/{marker,marker_trait,trait}_summary_stats.cpp
 * ("NULL" = absent path; *N_out = NULL when the sample size is uniform; release with cusk_free);
 * subset_variables (src/parent_set.cpp:8-53; out_sel[i] = 1 keeps variable i); load_phen
 * (src/phen.cpp:27-74; column-major). */
int cusk_cuskss_square_inputs(const char *mxm, const char *mxp, const char *mxp_se, const char *pxp,
                              const char *pxp_se, const char *blockfile, int block_index, const char *marker_ixs,
                              float num_samples, int *num_var, int *num_phen, float **C_out, float **N_out);
int cusk_subset_variables(const int *G, int num_var, int num_markers, int max_depth, int *out_sel);
int cusk_load_phen(const char *path, int64_t *num_samples, int64_t *num_phen, float **data_out);
void cusk_free(void *p);
/* post-processing of the block outputs, host only.
 * cusk_merge_block_outputs = merge_block_outputs(blockfile, outdir).write_mm(out_base)
 * (cusk_postprocessing/merge_blocks.py:361-403, :300-325; `ci-gwas.py merge-block-outputs`, ci-gwas.py:459-464):
 * writes <out_base>_sam.mtx, <out_base>_scm.mtx, <out_base>.mdim, <out_base>.ixs, byte-identical to the reference.
 * cusk_reformat_cuskss_merged = reformat_cuskss_merged_output(cusk_dir).write_mm(cusk_dir/cuskss_merged)
 * (merge_blocks.py:406-425; the step after `mps cuskss` with marker indices, ci-gwas.py:452-456). */
int cusk_merge_block_outputs(const char *blockfile, const char *outdir, const char *out_base);
int cusk_reformat_cuskss_merged(const char *cusk_dir);

/* ---- separating-set selection on a merged skeleton -----------------------------------
 * cusk_sepselect = orient_v_structures_merged(stem, alpha, num_samples, prior).to_file(out_stem)
 * (cusk_postprocessing/sepselect.py:576-584, :545-573; `ci-gwas.py sepselect` / `orient-v-structs`,
 * ci-gwas.py:467-476): reads <stem>.mdim, <stem>_sam.mtx, <stem>_scm.mtx (what merge-block-outputs writes) and an
 * optional orientation prior (int32 [P x P], NULL = none); writes <out_stem>.mdim, _sam.mtx, _scm.mtx, _spm.mtx,
 * .atr, .ut, .ssm.  The greedy search of find_maximal_and_min_pcorr_sepsets_incr (sepselect.py:262-329) runs on
 * the device, one thread per outer pair; fewer than 32 traits.
 * cusk_sepselect_pairs is that search alone on host buffers: Ct = rows 0..P of the symmetric correlation matrix
 * ([(P+1) x ld] doubles), pairs = [npairs][2] variable indices (i, j) with at least one end <= P, tmask[v] = bit t
 * set when trait t < P is a neighbour of variable v (never v itself, and never j for a pair (v, j): CUSK_E_ARG),
 * thr[l] = qnorm(1 - alpha/2) / sqrt(num_samples - l - 3) for l = 0..P.  cusk_qnorm: the normal quantile in double precision (scipy.stats.norm.ppf, sepselect.py:26-27). */
#define CUSK_SS_MAXP 32
typedef struct cusk_sepselect_result {
    int32_t len;                /* size of the maximal separating set */
    int32_t found_minimum;      /* the pair also has a "minimal partial correlation" set (sepselect.py:287-289) */
    int32_t set[CUSK_SS_MAXP];  /* traits in the order they were added, -1 padded */
} cusk_sepselect_result;
int cusk_sepselect(cusk_ctx *ctx, const char *stem, double alpha, int64_t num_samples, const char *prior_path,
                   const char *out_stem);
int cusk_sepselect_pairs(cusk_ctx *ctx, const double *Ct, int64_t ld, int P, int n_var, const int *pairs,
                         int64_t npairs, const uint32_t *tmask, const double *thr, cusk_sepselect_result *out);
double cusk_qnorm(double p);

/* ---- drop-in signatures of the reference -------------------------------------------------- */
/* include/mps/corr_host.h: cu_marker_phen_corr_pearson / cu_corr_pearson_npn (C++ linkage in the
 * reference; exported here with C linkage and the same argument order) */
void cu_marker_phen_corr_pearson(const unsigned char *marker_vals, const float *phen_vals,
                                 const size_t num_markers, const size_t num_individuals,
                                 const size_t num_phen, const float *marker_mean,
                                 const float *marker_std, float *marker_phen_corrs);
void cu_corr_pearson_npn(const unsigned char *marker_vals, const float *phen_vals,
                         const size_t num_markers, const size_t num_individuals, const size_t num_phen,
                         const float *marker_mean, const float *marker_std, float *marker_corrs,
                         float *marker_phen_corrs, float *phen_corrs);
/* include/mps/corr_host.h: cu_marker_corr_pearson_npn_batched_sparse (src/corr_host.cu:1199); batch_size is ignored */
void cu_marker_corr_pearson_npn_batched_sparse(const unsigned char *marker_vals, const size_t num_markers,
                                               const size_t num_individuals, const size_t corr_width,
                                               const size_t batch_size, float *corrs);
/* include/mps/cuPC-S.h:192-194 */
void Skeleton(float *C, int *P, int *G, float *Th, int *l, const int *maxlevel, float *pMax,
              int *SepSet);
/* include/mps/hetcor-cuPC-S.h:46 */
void hetcor_skeleton(float *C, int *P, int *G, float *N, float *Th, int *l, const int *maxlevel,
                     const int *time_index);

#ifdef __cplusplus
}
#endif
#endif /* CUSK_B200_H_ */
