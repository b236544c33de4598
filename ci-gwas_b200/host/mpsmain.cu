// mpsmain.cu -- the `mps` command line (cusk/apps/mps.cpp:17-121, cusk/src/cli.cpp:413-454,681-708):
// same sub-commands, same positional arguments, "NULL" for absent optional paths, exit status 0 on
// success and on "block skipped", 1 on input problems, EXIT_FAILURE on CUDA errors.
// Additions: `cusk-all`, `cuskss-all` (all blocks over all visible GPUs) and, for cuskss with
// marker indices, CUSK_SHARD_GPUS=N (level-l tests of the one skeleton sharded over N GPUs).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <thread>

#include "../csrc/common.cuh"
#include "host.hpp"

extern "C" int cusk_comm_create_all(cusk_ctx **ctxs, int world, struct cusk_comm **out);
extern "C" int cusk_comm_create_local(cusk_ctx **ctxs, int world, struct cusk_comm **out);
extern "C" void cusk_comm_destroy(struct cusk_comm *c);
extern "C" int cusk_ctx_set_comm(cusk_ctx *ctx, struct cusk_comm *c);

namespace cuskhost {

namespace {

const char *MPS_USAGE = R"(
usage: mps <command> [<args>]

commands:
    prep                    Prepare input (PLINK) .bed file for mps
    block                   Group markers into approximately unlinked blocks (banded correlations + LD blocking)
    cusk                    Run cuda-skeleton on a single block of block diagonal genomic covariance matrix
    cuskss                  Run cuda-skeleton on a block of markers and traits with pre-computed correlations.
    cusk-all                cusk for every block of a .blocks file, scheduled over all visible GPUs
    cuskss-all              cuskss for every block of a .blocks file, scheduled over all visible GPUs
    merge                   Merge the outputs of all blocks (ci-gwas.py merge-block-outputs): <blockfile> <outdir>
    reformat-merged         Rewrite <outdir>/cuskss_merged.* in the merged-blocks format: <outdir>
    sepselect               Maximal separating sets and v-structures of a merged skeleton (ci-gwas.py sepselect,
                            orient-v-structs): <merged-stem> <alpha> <num-samples> [<orientation-prior>|NULL [<out-stem>]]
)";

const char *CUSK_USAGE = R"(
usage: mps cusk <.phen> <bfiles> <.blocks> <alpha> <max-level> <max-level-two> <depth> <outdir> <block-index>
)";

cusk_ctx *make_ctx(int dev)
{
    cusk_ctx *ctx = nullptr;
    if (cusk_ctx_create(dev, &ctx) != CUSK_OK) {
        fprintf(stderr, "cusk_b200: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);  // include/mps/gpuerrors.h:6-35
    }
    return ctx;
}

int finish(int rc)
{
    if (rc != CUSK_OK) {
        fprintf(stderr, "cusk_b200: %s\n", cusk_last_error());
        return EXIT_FAILURE;
    }
    return EXIT_SUCCESS;
}

CuskssArgs parse_cuskss(char **v)  // v[0] = mxm ... v[14] = outdir  (mps.cpp:33-53)
{
    CuskssArgs a;
    a.mxm_path = v[0];
    a.mxp_path = v[1];
    a.mxp_se_path = v[2];
    a.pxp_path = v[3];
    a.pxp_se_path = v[4];
    a.time_index_path = v[5];
    a.block_ix = std::stoi(v[6]);
    a.block_path = v[7];
    a.marker_ixs_path = v[8];
    a.alpha = std::stof(v[9]);
    a.max_level_one = std::stoi(v[10]);
    a.max_level_two = std::stoi(v[11]);
    a.depth = std::stoi(v[12]);
    a.pearson_sample_size = (float)std::stoi(v[13]);
    a.outdir = v[14];
    a.merged = a.marker_ixs_path != "NULL";
    a.hetcor = a.mxp_se_path != "NULL";
    a.trait_only = a.mxm_path == "NULL";
    a.two_stage = a.max_level_two > 0;
    a.time_indexed = a.time_index_path != "NULL";
    return a;
}

void check_cuskss_paths(const CuskssArgs &a, bool need_mxm)  // mps.cpp:55-78
{
    check_path(a.pxp_path);
    check_path(a.outdir);
    if (a.merged)
        check_path(a.marker_ixs_path);
    else
        check_path(a.block_path);
    if (a.hetcor || a.merged || !a.trait_only) {
        if (need_mxm) check_path(a.mxm_path);
        check_path(a.mxp_path);
    }
    if (a.hetcor) {
        check_path(a.mxp_se_path);
        check_path(a.pxp_se_path);
    }
    if (a.time_indexed) check_path(a.time_index_path);
}

// one skeleton over `world` GPUs of this process: one host thread per GPU, NCCL between levels
int cuskss_sharded(const CuskssArgs &a, const std::vector<int> &devs)
{
    const int world = (int)devs.size();
    std::vector<cusk_ctx *> ctxs(world);
    for (int i = 0; i < world; i++) ctxs[i] = make_ctx(devs[i]);
    std::vector<cusk_comm *> comms(world, nullptr);
    bool distinct = true;
    for (int i = 0; i < world; i++)
        for (int j = 0; j < i; j++) distinct &= devs[i] != devs[j];
    int rc = (distinct && !getenv("CUSK_LOCAL_EXCHANGE")) ? cusk_comm_create_all(ctxs.data(), world, comms.data()) : 1;
    if (rc != CUSK_OK) rc = cusk_comm_create_local(ctxs.data(), world, comms.data());
    if (rc != CUSK_OK) return finish(rc);
    std::vector<int> rcs(world, 0);
    std::vector<std::string> errs(world);
    std::vector<std::thread> th;
    for (int i = 0; i < world; i++) {
        th.emplace_back([&, i]() {
            cusk_ctx_set_comm(ctxs[i], comms[i]);
            CuskssArgs b = a;
            b.write_output = (i == 0);
            try {
                rcs[i] = run_cuskss(ctxs[i], b, i == 0, nullptr);
                if (rcs[i]) errs[i] = cusk_last_error();
            } catch (const HostError &e) {
                rcs[i] = 1000 + e.code;
                errs[i] = e.msg;
            }
        });
    }
    for (auto &t : th) t.join();
    int out = EXIT_SUCCESS;
    for (int i = 0; i < world; i++)
        if (rcs[i]) {
            std::cout << errs[i] << std::endl;
            out = rcs[i] >= 1000 ? rcs[i] - 1000 : EXIT_FAILURE;
            break;
        }
    for (int i = 0; i < world; i++) {
        cusk_ctx_set_comm(ctxs[i], nullptr);
        cusk_comm_destroy(comms[i]);
        cusk_ctx_destroy(ctxs[i]);
    }
    return out;
}

int dispatch(int argc, char **argv)
{
    if (argc == 1) {
        std::cout << MPS_USAGE << std::endl;
        return EXIT_SUCCESS;
    }
    const std::string cmd = argv[1];
    if (cmd == "--help" || cmd == "-h") {
        std::cout << MPS_USAGE << std::endl;
        return EXIT_SUCCESS;
    }
    // host-only post-processing (cusk_postprocessing/merge_blocks.py; no device is touched)
    if (cmd == "merge") {
        if (argc != 4) {
            std::cout << "\nusage: mps merge <.blocks> <cusk-output-dir>\n" << std::endl;
            return 1;
        }
        std::string dir = argv[3];
        return merge_block_outputs(argv[2], dir, dir + "/merged_blocks");  // ci-gwas.py:459-464
    }
    if (cmd == "reformat-merged") {
        if (argc != 3) {
            std::cout << "\nusage: mps reformat-merged <cusk-output-dir>\n" << std::endl;
            return 1;
        }
        return reformat_cuskss_merged(argv[2]);  // ci-gwas.py:452-456
    }
    const std::vector<int> devs = devices_from_env();
    const int dev0 = devs.empty() ? 0 : devs[0];
    if (cmd == "cuskss") {
        if (argc < 17) {
            std::cout << MPS_USAGE << std::endl;
            return 1;
        }
        const CuskssArgs a = parse_cuskss(argv + 2);
        check_cuskss_paths(a, true);
        const char *sg = getenv("CUSK_SHARD_GPUS");
        const int shard = sg ? atoi(sg) : 1;
        if (a.merged && shard > 1) {
            std::vector<int> use;
            for (int i = 0; i < shard; i++) use.push_back(devs.empty() ? 0 : devs[(size_t)i % devs.size()]);
            return cuskss_sharded(a, use);
        }
        cusk_ctx *ctx = make_ctx(dev0);
        const int rc = run_cuskss(ctx, a, true, nullptr);
        cusk_ctx_destroy(ctx);
        return finish(rc);
    }
    if (cmd == "sepselect") {
        if (argc < 5 || argc > 7) {
            std::cout << "\nusage: mps sepselect <merged-stem> <alpha> <num-samples> [<orientation-prior>|NULL [<out-stem>]]\n"
                      << std::endl;
            return 1;
        }
        const std::string stem = argv[2];
        const std::string prior = (argc > 5 && strcmp(argv[5], "NULL")) ? argv[5] : "";
        // ci-gwas.py:469,475: <dirname of the stem>/max_sep_min_pc
        const size_t sl = stem.rfind('/');
        const std::string out = argc > 6 ? argv[6] : (sl == std::string::npos ? std::string(".") : stem.substr(0, sl)) + "/max_sep_min_pc";
        cusk_ctx *ctx = make_ctx(dev0);
        const int rc = run_sepselect(ctx, stem, std::stod(argv[3]), std::stol(argv[4]), prior, out);
        cusk_ctx_destroy(ctx);
        if (rc == CUSK_OK) std::cout << "Sepselect / v-structs done." << std::endl;  // ci-gwas.py:476
        return finish(rc);
    }
    if (cmd == "prep") {
        if (argc != 3 || !strcmp(argv[2], "--help") || !strcmp(argv[2], "-h")) {
            std::cout << "\nusage: mps prep <.bfiles>\n" << std::endl;
            return 1;
        }
        for (const char *sfx : {".bed", ".bim", ".fam"}) check_path(std::string(argv[2]) + sfx);  // io.cpp:417-424
        cusk_ctx *ctx = make_ctx(dev0);
        const int rc = run_prep(ctx, argv[2], true);
        cusk_ctx_destroy(ctx);
        return finish(rc);
    }
    if (cmd == "cusk" || cmd == "cusk-all") {
        const bool all = cmd == "cusk-all";
        if (argc < (all ? 10 : 11)) {
            std::cout << CUSK_USAGE << std::endl;
            return 1;
        }
        CuskArgs a;
        a.phen_path = argv[2];
        a.bfiles = argv[3];
        a.block_path = argv[4];
        a.alpha = std::stof(argv[5]);
        a.max_level = std::stoi(argv[6]);
        a.max_level_two = std::stoi(argv[7]);
        a.depth = std::stoi(argv[8]);
        a.outdir = argv[9];
        if (all) return run_cusk_all(a, devs, getenv("CUSK_NO_RESUME") == nullptr);
        a.block_index = std::stoi(argv[10]);
        std::cout << "Got args: " << std::endl
                  << ".phen: " << a.phen_path << std::endl
                  << "bfiles: " << a.bfiles << std::endl
                  << ".blocks: " << a.block_path << std::endl
                  << "alpha: " << a.alpha << std::endl
                  << "max_level: " << a.max_level << std::endl
                  << "max_level: " << a.max_level << std::endl
                  << "depth: " << a.depth << std::endl
                  << "outdir: " << a.outdir << std::endl
                  << "block-index: " << a.block_index << std::endl
                  << "Checking paths" << std::endl;
        CuskShared shared;
        load_cusk_shared(a, &shared, true);
        std::cout << "Number of levels: " << a.max_level << std::endl << "Setting level thr for cuPC: " << std::endl;
        for (int i = 0; i <= a.max_level && i <= CUSK_ML; ++i)
            std::cout << "\t Level: " << i << " thr: " << shared.Th[(size_t)i] << std::endl;
        cusk_ctx *ctx = make_ctx(dev0);
        const int rc = run_cusk_block(ctx, a, shared, (size_t)a.block_index, true, nullptr);
        cusk_ctx_destroy(ctx);
        return finish(rc);
    }
    if (cmd == "cuskss-all") {
        // mps cuskss-all <mxm-pattern> <mxp> <mxp_se> <pxp> <pxp_se> <time_index> <blockfile> <alpha> <max_level_one>
        //                <max_level_two> <max_depth> <num_samples> <outdir>     (mxm-pattern: "{}" = block stem)
        if (argc < 15) {
            std::cout << MPS_USAGE << std::endl;
            return 1;
        }
        char *v[15] = {argv[2], argv[3], argv[4], argv[5], argv[6], argv[7], (char *)"0", argv[8], (char *)"NULL",
                       argv[9], argv[10], argv[11], argv[12], argv[13], argv[14]};
        const CuskssArgs a = parse_cuskss(v);
        check_cuskss_paths(a, false);
        return run_cuskss_all(a, argv[2], devs, getenv("CUSK_NO_RESUME") == nullptr);
    }
    if (cmd == "block") {  // make_blocks, src/cli.cpp:362-411 (6 argv)
        if (argc != 6 || !strcmp(argv[2], "--help") || !strcmp(argv[2], "-h")) {
            std::cout << "\nGroup markers into approximately unlinked blocks\n\n"
                         "usage: mps block <bfiles> <max-block-size> <device-mem-gb> <corr-width>\n\n"
                         "arguments:\n"
                         "    bfiles          stem of .bed, .bim, .fam files\n"
                         "    max-block-size  maximum number of markers per block\n"
                         "    device-mem-gb   maximum memory available on gpu in GB (0: all that is free)\n"
                         "    corr-width      max distance at which to compute correlations\n"
                      << std::endl;
            return 1;
        }
        int max_block_size = 0;
        float device_mem_gb = 0.f;
        long corr_width = 0;
        try {
            max_block_size = std::stoi(argv[3]);
            device_mem_gb = std::stof(argv[4]);
            corr_width = std::stol(argv[5]);
        } catch (const std::exception &) {
            fail("mps block: max-block-size, device-mem-gb and corr-width must be numbers");
        }
        if (max_block_size < 1 || corr_width < 1) fail("mps block: max-block-size and corr-width must be positive");
        std::cout << "Checking paths" << std::endl;
        for (const char *sfx : {".bed", ".bim", ".fam"}) check_path(std::string(argv[2]) + sfx);
        if (devs.size() > 1)  // chromosomes are independent: one worker per device (CUSK_DEVICES selects them)
            return finish(run_block_all(argv[2], max_block_size, device_mem_gb, (size_t)corr_width, devs));
        cusk_ctx *ctx = make_ctx(dev0);
        const int rc = run_block(ctx, argv[2], max_block_size, device_mem_gb, (size_t)corr_width, true);
        cusk_ctx_destroy(ctx);
        return finish(rc);
    }
    std::cout << MPS_USAGE << std::endl;
    return EXIT_SUCCESS;
}

}  // namespace

int mps_main(int argc, char **argv)
{
    // ranks as threads of this process (CUSK_SHARD_GPUS): load every kernel up front, so that no module load
    // (a context-wide operation) can fall between two ranks' collectives
    if (getenv("CUSK_SHARD_GPUS")) setenv("CUDA_MODULE_LOADING", "EAGER", 0);
    try {
        return dispatch(argc, argv);
    } catch (const HostError &e) {
        std::cout << e.msg << std::endl;
        return e.code;
    } catch (const std::exception &e) {
        std::cout << "mps: " << e.what() << std::endl;
        return 1;
    }
}

}  // namespace cuskhost

extern "C" int cusk_mps_main(int argc, char **argv) { return cuskhost::mps_main(argc, argv); }

// prep_bed_no_impute(BfilesBase) (src/prep.cpp:157-201) with a status return
extern "C" int cusk_prep_bed(cusk_ctx *ctx, const char *bfiles)
{
    if (!ctx || !bfiles) {
        cusk_set_error("cusk_prep_bed: NULL argument");
        return CUSK_E_ARG;
    }
    try {
        return cuskhost::run_prep(ctx, bfiles, false);
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

// make_blocks (src/cli.cpp:362-411) with a status return, and its host-only pieces (src/blocking.cpp)
extern "C" int cusk_make_blocks(cusk_ctx *ctx, const char *bfiles, int max_block_size, float device_mem_gb,
                                size_t corr_width)
{
    if (!ctx || !bfiles || max_block_size < 1 || corr_width < 1) {
        cusk_set_error("cusk_make_blocks: invalid argument");
        return CUSK_E_ARG;
    }
    try {
        return cuskhost::run_block(ctx, bfiles, max_block_size, device_mem_gb, corr_width, false);
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

extern "C" int cusk_hanning_smoothing(const float *v, size_t nv, int window_size, double *out)
{
    if (!v || !out || window_size < 1) {
        cusk_set_error("cusk_hanning_smoothing: invalid argument");
        return CUSK_E_ARG;
    }
    const std::vector<double> s = cuskhost::hanning_smoothing(v, nv, window_size);
    std::copy(s.begin(), s.end(), out);
    return CUSK_OK;
}

extern "C" int64_t cusk_block_chr(const float *row_sums, size_t m, int max_block_size, int64_t *first, int64_t *last,
                                  int *window)
{
    if (!row_sums || !first || !last || m == 0 || max_block_size < 1) {
        cusk_set_error("cusk_block_chr: invalid argument");
        return CUSK_E_ARG;
    }
    const cuskhost::BlockPartition bp = cuskhost::block_chr(row_sums, m, max_block_size);
    for (size_t i = 0; i < bp.first.size(); i++) {
        first[i] = (int64_t)bp.first[i];
        last[i] = (int64_t)bp.last[i];
    }
    if (window) *window = bp.window;
    return (int64_t)bp.first.size();
}

// merge_block_outputs(blockfile, outdir).write_mm(out_base) and reformat_cuskss_merged_output(dir).write_mm(...)
// (cusk_postprocessing/merge_blocks.py:361-425, :300-325); host only
extern "C" int cusk_merge_block_outputs(const char *blockfile, const char *outdir, const char *out_base)
{
    if (!blockfile || !outdir || !out_base) {
        cusk_set_error("cusk_merge_block_outputs: NULL argument");
        return CUSK_E_ARG;
    }
    try {
        return cuskhost::merge_block_outputs(blockfile, outdir, out_base);
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

extern "C" int cusk_sepselect(cusk_ctx *ctx, const char *stem, double alpha, int64_t num_samples, const char *prior_path,
                              const char *out_stem)
{
    if (!stem || !out_stem) {
        cusk_set_error("cusk_sepselect: NULL argument");
        return CUSK_E_ARG;
    }
    if (!ctx) {
        cusk_set_error("cusk_sepselect: no context (there is no CPU path)");
        return CUSK_E_NODEVICE;
    }
    try {
        return cuskhost::run_sepselect(ctx, stem, alpha, (long)num_samples, prior_path ? prior_path : "", out_stem);
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

extern "C" int cusk_reformat_cuskss_merged(const char *cusk_dir)
{
    if (!cusk_dir) {
        cusk_set_error("cusk_reformat_cuskss_merged: NULL argument");
        return CUSK_E_ARG;
    }
    try {
        return cuskhost::reformat_cuskss_merged(cusk_dir);
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

// ---- host-side pieces exposed for callers and tests that have no device ------------------------------
// make_square_cuskss_inputs (src/cli.cpp:89-173) over the reference's loaders: builds the dense
// correlation matrix C and, when standard errors are given, the effective-sample-size matrix N.
// "NULL" or NULL = absent path.  *C_out / *N_out are malloc'ed (release with cusk_free); *N_out is
// NULL when the sample size is uniform (= num_samples).
extern "C" int cusk_cuskss_square_inputs(const char *mxm, const char *mxp, const char *mxp_se, const char *pxp,
                                         const char *pxp_se, const char *blockfile, int block_index,
                                         const char *marker_ixs, float num_samples, int *num_var, int *num_phen,
                                         float **C_out, float **N_out)
{
    using namespace cuskhost;
    auto absent = [](const char *p) { return !p || !strcmp(p, "NULL"); };
    try {
        const bool het = !absent(mxp_se) && !absent(pxp_se), merged = !absent(marker_ixs);
        const Pxp px = load_pxp(pxp, het ? pxp_se : "", num_samples);
        size_t m = 0;
        const std::vector<float> mm = load_mxm_square(mxm, &m);
        std::vector<int> ixs;
        MarkerBlock block;
        if (merged) {
            ixs = read_ints_from_binary(marker_ixs);
        } else {
            const std::vector<MarkerBlock> blocks = read_blocks(blockfile);
            if (block_index < 0 || (size_t)block_index >= blocks.size()) fail("block index out of range");
            block = blocks[(size_t)block_index];
        }
        const MxpRows mp = load_mxp(mxp, het ? mxp_se : "", !merged, merged ? 0 : block.first_global(),
                                    merged ? 0 : block.last_global(), ixs);
        if (px.num_phen != mp.num_phen) fail("Numbers of traits seem to differ between pxp and mxp");
        if (m != mp.num_markers) fail("Numbers of markers seem to differ between mxm and mxp");
        const SquareInputs si = make_square_cuskss_inputs(mm, m, mp, px, num_samples, het);
        const size_t nv = m + px.num_phen;
        *num_var = (int)nv;
        *num_phen = (int)px.num_phen;
        *C_out = (float *)malloc(nv * nv * sizeof(float));
        memcpy(*C_out, si.C.data(), nv * nv * sizeof(float));
        *N_out = nullptr;
        if (!si.uniform) {
            *N_out = (float *)malloc(nv * nv * sizeof(float));
            memcpy(*N_out, si.N.data(), nv * nv * sizeof(float));
        }
        return CUSK_OK;
    } catch (const HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}

extern "C" void cusk_free(void *p) { free(p); }

// subset_variables (src/parent_set.cpp:8-53) on a dense 0/1 adjacency: out_sel[i] = 1 when variable i is kept
extern "C" int cusk_subset_variables(const int *G, int num_var, int num_markers, int max_depth, int *out_sel)
{
    if (!G || !out_sel || num_var < 0 || num_markers > num_var) return CUSK_E_ARG;
    const size_t rw = cusk_row_words(num_var);
    std::vector<uint32_t> bits((size_t)num_var * rw, 0u);
    for (int i = 0; i < num_var; i++)
        for (int j = 0; j < num_var; j++)
            if (G[(size_t)i * num_var + j] == 1) bits[(size_t)i * rw + (j >> 5)] |= 1u << (j & 31);
    const std::vector<int> sel = cuskhost::subset_variables(bits.data(), (size_t)num_var, rw, (size_t)num_markers, max_depth);
    for (int i = 0; i < num_var; i++) out_sel[i] = 0;
    for (int v : sel) out_sel[v] = 1;
    return CUSK_OK;
}

// load_phen (src/phen.cpp:27-74): column-major values, malloc'ed
extern "C" int cusk_load_phen(const char *path, int64_t *num_samples, int64_t *num_phen, float **data_out)
{
    try {
        const cuskhost::Phen ph = cuskhost::load_phen(path);
        *num_samples = (int64_t)ph.num_samples;
        *num_phen = (int64_t)ph.num_phen;
        *data_out = (float *)malloc(std::max<size_t>(ph.data.size(), 1) * sizeof(float));
        memcpy(*data_out, ph.data.data(), ph.data.size() * sizeof(float));
        return CUSK_OK;
    } catch (const cuskhost::HostError &e) {
        cusk_set_error("%s", e.msg.c_str());
        return CUSK_E_ARG;
    }
}
