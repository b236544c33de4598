// host.hpp -- thin C++ host layer above the CUDA engines: the file formats and the orchestration of
// `mps prep | cusk | cuskss` (cusk/src/cli.cpp, io.cpp, phen.cpp, bim.cpp, *_summary_stats.cpp,
// parent_set.cpp in the reference), re-written around device-resident data and sparse outputs.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/cusk_b200.h"

namespace cuskhost {

// An input problem the caller can act on (bad path, inconsistent dimensions).  The mps front end
// prints the message and exits with status 1 like the reference's check_path()/exit(1) sites.
struct HostError {
    std::string msg;
    int code;  // process exit code
};
[[noreturn]] void fail(const std::string &msg, int code = 1);

// ---- text / binary files -------------------------------------------------------------------
bool path_exists(const std::string &p);
void check_path(const std::string &p);                                   // io.cpp:399-406
std::string make_path(const std::string &dir, const std::string &stem, const std::string &suffix);  // io.cpp:50-72
std::string slurp(const std::string &path);                              // whole file
size_t count_lines(const std::string &path);                             // io.cpp:38-48
std::vector<std::string> split_ws(const char *b, const char *e);         // io.cpp:14-24
std::vector<float> read_floats_from_line_range(const std::string &path, size_t first, size_t last);  // io.cpp:137
std::vector<int> read_ints_from_lines(const std::string &path);          // io.cpp:160
std::vector<int> read_ints_from_binary(const std::string &path);         // io.cpp:216
std::vector<float> read_floats_from_binary(const std::string &path);     // io.cpp:227
void write_ints_to_binary(const int *d, size_t n, const std::string &path);      // io.cpp:279
void write_floats_to_binary(const float *d, size_t n, const std::string &path);  // io.cpp:301
void write_column(const std::vector<float> &v, const std::string &path);         // io.cpp:350-362
void write_column(const std::vector<int> &v, const std::string &path);

// ---- genotype file set -----------------------------------------------------------------------
struct MarkerBlock {  // include/mps/marker_block.h
    std::string chr;
    size_t first = 0, last = 0, chr_global_offset = 0;
    size_t size() const { return last - first + 1; }
    size_t first_global() const { return first + chr_global_offset; }
    size_t last_global() const { return last + chr_global_offset; }
    std::string stem() const { return chr + "_" + std::to_string(first) + "_" + std::to_string(last); }
};
std::vector<MarkerBlock> read_blocks(const std::string &path);  // io.cpp:74-101

struct BimInfo {  // include/mps/bim.h, src/bim.cpp:20-47
    size_t num_lines = 0;
    std::vector<std::string> chr_ids;
    std::vector<size_t> num_on_chr, chr_start;
    std::unordered_map<std::string, size_t> ix;
    explicit BimInfo(const std::string &path);
    size_t chr_index(const std::string &chr) const;  // exits 1 when absent (bim.cpp:12-18)
};

struct BedDims {
    size_t num_samples = 0, num_markers = 0;
    size_t col_bytes() const { return (num_samples + 3) / 4; }
};
BedDims read_dims(const std::string &path);  // io.h:30-39
bool has_valid_bed_prefix(const std::string &bed_path);  // bfiles_base.h:39-52

struct Phen {  // src/phen.cpp:27-74: column-major [num_phen][num_samples], NA -> NaN
    size_t num_samples = 0, num_phen = 0;
    std::vector<float> data;
};
Phen load_phen(const std::string &path);
void read_file_range(const std::string &path, size_t offset, size_t bytes, unsigned char *dst);  // several threads; fails on a short read

// ---- summary statistics (cuskss inputs) -----------------------------------------------------
struct SquareInputs {
    size_t num_markers = 0, num_phen = 0;
    std::vector<float> C;  // [nv*nv]
    std::vector<float> N;  // [nv*nv]; empty when uniform
    float n_uniform = 0.f;
    bool uniform = true;
};
// marker_summary_stats.cpp:8-24 (lower triangle incl. diagonal, row-major, NaN -> 0)
std::vector<float> load_mxm_square(const std::string &path, size_t *num_markers);
struct MxpRows {
    size_t num_markers = 0, num_phen = 0;
    std::vector<float> corr, ess;  // [m*p]; ess empty without standard errors
};
// marker_trait_summary_stats.cpp:40-299: rows [first,last] (block) or the sorted marker indices
MxpRows load_mxp(const std::string &corr_path, const std::string &se_path /* "" = none */, bool by_block,
                 size_t first_global, size_t last_global, const std::vector<int> &marker_ixs);
struct Pxp {
    size_t num_phen = 0;
    std::vector<float> corr, ess;  // [p*p]
};
Pxp load_pxp(const std::string &corr_path, const std::string &se_path /* "" = none */, float sample_size);  // trait_summary_stats.cpp
// cli.cpp:89-173
SquareInputs make_square_cuskss_inputs(const std::vector<float> &mxm_sq, size_t m, const MxpRows &mxp, const Pxp &pxp,
                                       float sample_size, bool het);

// ---- variable selection and reduction (parent_set.cpp) -----------------------------------------
// adjacency as a bitmap: nv rows of rw = ceil(nv/32) words
std::vector<int> subset_variables(const uint32_t *bits, size_t nv, size_t rw, size_t num_markers, int max_depth);

struct Reduced {  // ReducedGC / ReducedGCS
    size_t num_var = 0, num_phen = 0;
    int max_level = CUSK_ML;
    std::vector<int> new_to_old;
    std::vector<int> G;    // [nv*nv]
    std::vector<float> C;  // [nv*nv]
    std::vector<float> N;  // ReducedGC only: [nv*nv], or empty + n_uniform
    float n_uniform = 0.f;
    bool uniform = true;
    std::vector<int> S;    // ReducedGCS only: [nv*nv*max_level]
    bool has_sep = false;
    size_t num_markers() const { return num_var - num_phen; }
    void to_file(const std::string &base) const;  // parent_set.h:42-52 / :99-108
};

// ---- orchestration ------------------------------------------------------------------------------
struct CuskssArgs {  // include/mps/cli.h:10-31
    bool merged = false, hetcor = false, trait_only = false, two_stage = false, time_indexed = false;
    float alpha = 1e-4f, pearson_sample_size = 0.f;
    int max_level_one = 14, max_level_two = 0, depth = 1, block_ix = 0;
    std::string block_path, marker_ixs_path, mxm_path, mxp_path, mxp_se_path, pxp_path, pxp_se_path, time_index_path,
        outdir;
    bool write_output = true;  // edge-sharded runs: only rank 0 writes
};
struct CuskArgs {
    std::string phen_path, bfiles, block_path, outdir;
    float alpha = 1e-4f;
    int max_level = 14, max_level_two = 0, depth = 1, block_index = 0;
};

struct RunReport {  // what a run did, for the machine-readable log line
    bool skipped = false;
    size_t num_var_in = 0, num_var_out = 0, retained_markers = 0;
    cusk_stats stage[2];
    int stages = 0;
    cusk_corr_stats corr;
    double load_s = 0, device_s = 0, write_s = 0;
};

// Pre-loaded inputs shared by all blocks of a multi-block run (the scheduler loads them once).
struct CuskShared {
    Phen phen;
    BedDims dims;
    std::vector<float> means, stds;  // whole genome, global marker order
    std::vector<MarkerBlock> blocks;
    std::vector<size_t> block_chr_start;  // global start of each block's chromosome (from the .bim)
    std::vector<float> Th;  // 15 thresholds
};
void load_cusk_shared(const CuskArgs &a, CuskShared *s, bool verbose);
int run_cusk_block(cusk_ctx *ctx, const CuskArgs &a, const CuskShared &s, size_t block_index, bool verbose,
                   RunReport *rep);
// Inputs shared by all blocks of a multi-block cuskss run: pxp, the whole mxp table, block list.
struct CuskssShared {
    Pxp pxp;
    MxpRows mxp_all;
    std::vector<MarkerBlock> blocks;
    std::vector<int> time_index_traits;
};
void load_cuskss_shared(const CuskssArgs &a, CuskssShared *s);
// shared == nullptr: everything is read from the files named in the arguments (the mps contract)
int run_cuskss(cusk_ctx *ctx, const CuskssArgs &a, bool verbose, RunReport *rep, const CuskssShared *shared = nullptr);
int run_prep(cusk_ctx *ctx, const std::string &bfiles, bool verbose);

// multi-GPU block scheduler (sched.cpp): all blocks of a .blocks file over the given devices
int run_cusk_all(const CuskArgs &a, const std::vector<int> &devices, bool resume);
int run_cuskss_all(const CuskssArgs &a, const std::string &mxm_pattern, const std::vector<int> &devices, bool resume);
std::vector<int> devices_from_env();

// post-processing of the block outputs (merge.cpp; cusk_postprocessing/merge_blocks.py:361-425): host only
int merge_block_outputs(const std::string &blockfile, const std::string &outdir, const std::string &out_base);
int reformat_cuskss_merged(const std::string &cusk_dir);
// separating-set selection + v-structure orientation on a merged skeleton (sepselect.cpp;
// cusk_postprocessing/sepselect.py:429-584); the search runs on ctx's device
int run_sepselect(cusk_ctx *ctx, const std::string &stem, double alpha, long num_samples, const std::string &prior_path,
                  const std::string &out_stem);

// LD blocking (blocking.cpp; src/blocking.cpp:8-136): host only
struct BlockPartition {
    std::vector<size_t> first, last;  // inclusive marker ranges, in order
    size_t largest = 0;
    int window = 0;  // Hann window that produced them
};
std::vector<double> hanning_smoothing(const float *v, size_t nv, int window);
BlockPartition blocks_for_window(const float *row_sums, size_t m, int window);
BlockPartition block_chr(const float *row_sums, size_t m, int max_block_size);
// `mps block` (block.cu; src/cli.cpp:362-411): banded correlations on ctx's device, blocking on the host
int run_block(cusk_ctx *ctx, const std::string &bfiles, int max_block_size, float device_mem_gb, size_t corr_width,
              bool verbose);
// the chromosomes of the file set spread over several devices (largest first, no exchange); same output file
int run_block_all(const std::string &bfiles, int max_block_size, float device_mem_gb, size_t corr_width,
                  const std::vector<int> &devices);

int mps_main(int argc, char **argv);

}  // namespace cuskhost
