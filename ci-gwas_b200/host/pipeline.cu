// pipeline.cu -- orchestration of one `mps cusk` / `mps cuskss` / `mps prep` run on one context
// (replaces cusk(), cuskss(), run_cusk(), reduced_gcs_cusk(), prep_bed_no_impute(): cusk/src/cli.cpp:29-87,
// 194-346, 432-678; cusk/src/prep.cpp:157-201).  Differences from the reference's flow, none of which
// change a written byte: correlations stay on the device and feed the skeleton directly; the
// adjacency travels as a bitmap; separating sets come back sparse; the marker x trait screen is
// computed once instead of twice.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <iostream>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "../csrc/corr.cuh"
#include "../csrc/engine.h"
#include "host.hpp"

namespace cuskhost {

namespace {

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

#define HOST_CUDA(call)                                                                          \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            cusk_set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return CUSK_E_CUDA;                                                                  \
        }                                                                                        \
    } while (0)

// out[i][j] = in[sel[i]][sel[j]]
__global__ void __launch_bounds__(256) gather_square_kernel(const float *__restrict__ in, int n,
                                                            const int *__restrict__ sel, int ns,
                                                            float *__restrict__ out)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= ns) return;
    out[(size_t)i * ns + j] = in[(size_t)sel[i] * n + sel[j]];
}

// Square-matrix assembly on the device (make_square_cuskss_inputs, cli.cpp:89-173, for a block): C = [mxm, mxp; mxp',
// pxp] from the block's lower-triangular mxm file (NaN -> 0, diagonal as stored), its rows of the marker x trait table
// and the trait x trait table; N likewise when standard errors were given (marker x marker pairs get the sample size).
__global__ void __launch_bounds__(256) assemble_square_kernel(const float *__restrict__ tril, int m,
                                                              const float *__restrict__ mxp, const float *__restrict__ pxp,
                                                              int p, float *__restrict__ C,
                                                              const float *__restrict__ mxp_ess,
                                                              const float *__restrict__ pxp_ess, float sample_size,
                                                              float *__restrict__ N)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    const int nv = m + p;
    if (j >= nv) return;
    float c, nn = sample_size;
    if (i < m && j < m) {
        const size_t hi = (size_t)max(i, j), lo = (size_t)min(i, j);
        const float v = tril[hi * (hi + 1) / 2 + lo];
        c = isnan(v) ? 0.0f : v;
    } else if (i < m) {
        c = mxp[(size_t)i * p + (j - m)];
        if (mxp_ess) nn = mxp_ess[(size_t)i * p + (j - m)];
    } else if (j < m) {
        c = mxp[(size_t)j * p + (i - m)];
        if (mxp_ess) nn = mxp_ess[(size_t)j * p + (i - m)];
    } else {
        c = pxp[(size_t)(i - m) * p + (j - m)];
        if (pxp_ess) nn = pxp_ess[(size_t)(i - m) * p + (j - m)];
    }
    C[(size_t)i * nv + j] = c;
    if (N) N[(size_t)i * nv + j] = nn;
}

std::vector<int> expand_bits(const std::vector<uint32_t> &bits, size_t nv, size_t rw, const std::vector<int> &sel)
{
    const size_t ns = sel.size();
    std::vector<int> G(ns * ns);
    for (size_t i = 0; i < ns; i++) {
        const uint32_t *row = bits.data() + (size_t)sel[i] * rw;
        for (size_t j = 0; j < ns; j++) G[i * ns + j] = (row[sel[j] >> 5] >> (sel[j] & 31)) & 1u;
    }
    (void)nv;
    return G;
}

struct Pinned {  // page-locked staging buffer, falls back to pageable memory when pinning fails
    void *p = nullptr;
    bool pinned = false;
    size_t bytes = 0;
    ~Pinned() { release(); }
    void release()
    {
        if (!p) return;
        if (pinned)
            cudaFreeHost(p);
        else
            free(p);
        p = nullptr;
    }
    void *get(size_t n)
    {
        if (n <= bytes && p) return p;
        release();
        if (cudaMallocHost(&p, n) == cudaSuccess) {
            pinned = true;
        } else {
            cudaGetLastError();
            p = malloc(n);
            pinned = false;
        }
        bytes = n;
        return p;
    }
};

}  // namespace

// parent_set.cpp:8-53.  The reference runs one BFS per trait over marker nodes; the union over
// traits of "within max_depth hops" equals one multi-source BFS from all traits.
std::vector<int> subset_variables(const uint32_t *bits, size_t nv, size_t rw, size_t num_markers, int max_depth)
{
    std::vector<char> in(nv, 0);
    std::vector<int> frontier, next;
    for (size_t t = num_markers; t < nv; t++) {
        in[t] = 1;
        frontier.push_back((int)t);
    }
    for (int depth = 0; depth < max_depth && !frontier.empty(); depth++) {
        next.clear();
        for (int node : frontier) {
            const uint32_t *row = bits + (size_t)node * rw;
            for (size_t w = 0; w < rw; w++) {
                uint32_t word = row[w];
                while (word) {
                    const int b = __builtin_ctz(word);
                    word &= word - 1;
                    const size_t c = w * 32 + (size_t)b;
                    if (c < num_markers && !in[c]) {
                        in[c] = 1;
                        next.push_back((int)c);
                    }
                }
            }
        }
        frontier.swap(next);
    }
    std::vector<int> sel;
    for (size_t i = 0; i < nv; i++)
        if (in[i]) sel.push_back((int)i);
    return sel;
}

// ------------------------------------------------------------------------------------------------
// cuskss: one stage of run_cusk() (cli.cpp:29-60) on a ReducedGC held on the host
// ------------------------------------------------------------------------------------------------
static int cuskss_stage(cusk_ctx *ctx, Reduced &gc, bool first_stage, float th, int depth, int max_level,
                        const std::vector<int> &time_index_traits, cusk_stats *stats)
{
    const size_t nv = gc.num_var, m = gc.num_markers();
    const size_t rw = cusk_row_words((int)nv);
    cudaStream_t st = (cudaStream_t)cusk_ctx_stream(ctx);
    HOST_CUDA(cudaSetDevice(ctx->device));
    arena_reset(ctx->io);
    float *dC = (float *)arena_take(ctx->io, nv * nv * 4);
    float *dN = gc.uniform ? nullptr : (float *)arena_take(ctx->io, nv * nv * 4);
    uint32_t *dbits = (uint32_t *)arena_take(ctx->io, nv * rw * 4);
    int *dG = first_stage ? nullptr : (int *)arena_take(ctx->io, nv * nv * 4);
    int *dT = (int *)arena_take(ctx->io, nv * 4);
    if (!dC || (!gc.uniform && !dN) || !dbits || (!first_stage && !dG) || !dT) return CUSK_E_CUDA;
    // markers are time 0, traits carry their time index (cli.cpp:37-42)
    std::vector<int> time_index(nv, 0);
    for (size_t i = m, r = 0; i < nv; i++, r++) time_index[i] = time_index_traits[r];
    bool use_time = false;
    for (size_t i = 1; i < nv; i++) use_time |= time_index[i] != time_index[0];
    HOST_CUDA(cudaMemcpyAsync(dC, gc.C.data(), nv * nv * 4, cudaMemcpyHostToDevice, st));
    if (dN) HOST_CUDA(cudaMemcpyAsync(dN, gc.N.data(), nv * nv * 4, cudaMemcpyHostToDevice, st));
    if (dG) HOST_CUDA(cudaMemcpyAsync(dG, gc.G.data(), nv * nv * 4, cudaMemcpyHostToDevice, st));
    if (use_time) HOST_CUDA(cudaMemcpyAsync(dT, time_index.data(), nv * 4, cudaMemcpyHostToDevice, st));
    int last = 0;
    int rc = hetcor_engine(ctx, dC, (int)nv, dbits, dG, false, dN, gc.n_uniform, th, max_level, use_time ? dT : nullptr,
                           &last, stats);
    if (rc) return rc;
    std::vector<uint32_t> bits(nv * rw);
    HOST_CUDA(cudaMemcpyAsync(bits.data(), dbits, nv * rw * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    // subset_variables + reduce_gc(G, C, S, P, num_var, num_phen, ML, new_to_old)  (cli.cpp:56-59)
    const std::vector<int> sel = subset_variables(bits.data(), nv, rw, m, depth);
    Reduced out;
    const size_t ns = sel.size();
    out.num_var = ns;
    out.num_phen = gc.num_phen;
    out.max_level = CUSK_ML;
    out.uniform = gc.uniform;
    out.n_uniform = gc.n_uniform;
    out.new_to_old.resize(ns);
    for (size_t i = 0; i < ns; i++) out.new_to_old[i] = gc.new_to_old[sel[i]];
    out.G = expand_bits(bits, nv, rw, sel);
    out.C.resize(ns * ns);
    if (!gc.uniform) out.N.resize(ns * ns);
    for (size_t i = 0; i < ns; i++)
        for (size_t j = 0; j < ns; j++) {
            out.C[i * ns + j] = gc.C[(size_t)sel[i] * nv + sel[j]];
            if (!gc.uniform) out.N[i * ns + j] = gc.N[(size_t)sel[i] * nv + sel[j]];
        }
    gc = std::move(out);
    return CUSK_OK;
}

// First stage of a BLOCK run with the square matrix built on the device: the host reads the block's mxm file into
// page-locked memory and uploads the triangle (half the bytes of the square), its mxp rows and the trait table; the
// reduced matrices come back gathered (a few hundred variables), the full n_var^2 matrix never exists on the host.
// Same bytes in the output files as the host-assembled path (cuskss_stage on make_square_cuskss_inputs).
static int cuskss_block_first_stage(cusk_ctx *ctx, const CuskssArgs &a, const float *mxp_corr, const float *mxp_ess,
                                    size_t mxp_rows, const Pxp &pxp, const std::vector<int> &time_index_traits, float th,
                                    Reduced *out, cusk_stats *stats, double *load_s, double *device_s)
{
    const double t0 = now_s();
    static thread_local Pinned tril_stage;
    FILE *f = fopen(a.mxm_path.c_str(), "rb");
    if (!f) fail("file or directory not found: " + a.mxm_path);
    fseeko(f, 0, SEEK_END);
    const size_t bytes = (size_t)ftello(f);
    fseeko(f, 0, SEEK_SET);
    const size_t cnt = bytes / sizeof(float);
    float *h_tril = (float *)tril_stage.get(std::max<size_t>(bytes, 4));
    const size_t got = fread(h_tril, 1, cnt * sizeof(float), f);
    fclose(f);
    if (got != cnt * sizeof(float)) fail("short read of " + a.mxm_path);
    const size_t m = (size_t)((std::sqrt((double)(4 * 2 * cnt + 1)) - 1) / 2);  // marker_summary_stats.cpp:8-24
    const size_t p = pxp.num_phen, nv = m + p;
    if (m != mxp_rows) {
        fail("Numbers of markers seem to differ between mxm and mxp\nmxp: " + std::to_string(mxp_rows) + " x " +
             std::to_string(p) + "\nmxm: " + std::to_string(m) + " x " + std::to_string(m));
    }
    *load_s = now_s() - t0;
    const double t1 = now_s();
    const bool het = a.hetcor;
    const size_t rw = cusk_row_words((int)nv);
    cudaStream_t st = (cudaStream_t)cusk_ctx_stream(ctx);
    HOST_CUDA(cudaSetDevice(ctx->device));
    arena_reset(ctx->io);
    float *d_tril = (float *)arena_take(ctx->io, std::max<size_t>(m * (m + 1) / 2, 1) * 4);
    float *d_mxp = (float *)arena_take(ctx->io, std::max<size_t>(m * p, 1) * 4);
    float *d_pxp = (float *)arena_take(ctx->io, std::max<size_t>(p * p, 1) * 4);
    float *d_mess = het ? (float *)arena_take(ctx->io, std::max<size_t>(m * p, 1) * 4) : nullptr;
    float *d_pess = het ? (float *)arena_take(ctx->io, std::max<size_t>(p * p, 1) * 4) : nullptr;
    float *dC = (float *)arena_take(ctx->io, nv * nv * 4);
    float *dN = het ? (float *)arena_take(ctx->io, nv * nv * 4) : nullptr;
    uint32_t *dbits = (uint32_t *)arena_take(ctx->io, nv * rw * 4);
    int *dT = (int *)arena_take(ctx->io, nv * 4);
    if (!d_tril || !d_mxp || !d_pxp || !dC || !dbits || !dT || (het && (!d_mess || !d_pess || !dN))) return CUSK_E_CUDA;
    std::vector<int> time_index(nv, 0);  // markers are time 0, traits carry their time index (cli.cpp:37-42)
    for (size_t i = m, r = 0; i < nv; i++, r++) time_index[i] = time_index_traits[r];
    bool use_time = false;
    for (size_t i = 1; i < nv; i++) use_time |= time_index[i] != time_index[0];
    HOST_CUDA(cudaMemcpyAsync(d_tril, h_tril, m * (m + 1) / 2 * 4, cudaMemcpyHostToDevice, st));
    HOST_CUDA(cudaMemcpyAsync(d_mxp, mxp_corr, m * p * 4, cudaMemcpyHostToDevice, st));
    HOST_CUDA(cudaMemcpyAsync(d_pxp, pxp.corr.data(), p * p * 4, cudaMemcpyHostToDevice, st));
    if (het) {
        HOST_CUDA(cudaMemcpyAsync(d_mess, mxp_ess, m * p * 4, cudaMemcpyHostToDevice, st));
        HOST_CUDA(cudaMemcpyAsync(d_pess, pxp.ess.data(), p * p * 4, cudaMemcpyHostToDevice, st));
    }
    if (use_time) HOST_CUDA(cudaMemcpyAsync(dT, time_index.data(), nv * 4, cudaMemcpyHostToDevice, st));
    assemble_square_kernel<<<dim3((unsigned)((nv + 255) / 256), (unsigned)nv), 256, 0, st>>>(
        d_tril, (int)m, d_mxp, d_pxp, (int)p, dC, d_mess, d_pess, a.pearson_sample_size, dN);
    HOST_CUDA(cudaGetLastError());
    int last = 0;
    int rc = hetcor_engine(ctx, dC, (int)nv, dbits, nullptr, false, dN, a.pearson_sample_size, th, a.max_level_one,
                           use_time ? dT : nullptr, &last, stats);
    if (rc) return rc;
    std::vector<uint32_t> bits(nv * rw);
    HOST_CUDA(cudaMemcpyAsync(bits.data(), dbits, nv * rw * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    // subset_variables + reduce_gc(G, C, S, P, num_var, num_phen, ML, new_to_old)  (cli.cpp:56-59)
    const std::vector<int> sel = subset_variables(bits.data(), nv, rw, m, a.depth);
    const size_t ns = sel.size();
    out->num_var = ns;
    out->num_phen = p;
    out->max_level = CUSK_ML;
    out->uniform = !het;
    out->n_uniform = a.pearson_sample_size;
    out->new_to_old = sel;
    out->G = expand_bits(bits, nv, rw, sel);
    out->C.resize(ns * ns);
    if (het) out->N.resize(ns * ns);
    int *d_sel = (int *)arena_take(ctx->io, std::max<size_t>(ns, 1) * 4);
    float *d_Cs = (float *)arena_take(ctx->io, std::max<size_t>(ns * ns, 1) * 4);
    float *d_Ns = het ? (float *)arena_take(ctx->io, std::max<size_t>(ns * ns, 1) * 4) : nullptr;
    if (!d_sel || !d_Cs || (het && !d_Ns)) return CUSK_E_CUDA;
    HOST_CUDA(cudaMemcpyAsync(d_sel, sel.data(), ns * 4, cudaMemcpyHostToDevice, st));
    const dim3 gg((unsigned)((ns + 255) / 256), (unsigned)ns);
    gather_square_kernel<<<gg, 256, 0, st>>>(dC, (int)nv, d_sel, (int)ns, d_Cs);
    if (het) gather_square_kernel<<<gg, 256, 0, st>>>(dN, (int)nv, d_sel, (int)ns, d_Ns);
    HOST_CUDA(cudaGetLastError());
    HOST_CUDA(cudaMemcpyAsync(out->C.data(), d_Cs, ns * ns * 4, cudaMemcpyDeviceToHost, st));
    if (het) HOST_CUDA(cudaMemcpyAsync(out->N.data(), d_Ns, ns * ns * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    *device_s = now_s() - t1;
    return CUSK_OK;
}

void load_cuskss_shared(const CuskssArgs &a, CuskssShared *s)
{
    s->blocks = read_blocks(a.block_path);
    s->pxp = load_pxp(a.pxp_path, a.hetcor ? a.pxp_se_path : std::string(), a.pearson_sample_size);
    s->time_index_traits.assign(s->pxp.num_phen, 1);
    if (a.time_indexed) {
        s->time_index_traits = read_ints_from_lines(a.time_index_path);
        if (s->time_index_traits.size() < s->pxp.num_phen) fail("time index file has fewer entries than traits");
    }
    s->mxp_all = load_mxp(a.mxp_path, a.hetcor ? a.mxp_se_path : std::string(), true, 0, (size_t)-1, {});
}

int run_cuskss(cusk_ctx *ctx, const CuskssArgs &a, bool verbose, RunReport *rep, const CuskssShared *shared)
{
    RunReport local;
    if (!rep) rep = &local;
    *rep = RunReport();
    const double t0 = now_s();
    auto say = [&](const char *s) {
        if (verbose) std::cout << s << std::endl;
    };
    say("Loading input files");
    MarkerBlock block;
    std::vector<int> marker_ixs;
    if (a.merged) {
        say("Loading marker indices");
        marker_ixs = read_ints_from_binary(a.marker_ixs_path);
    } else {
        say("Loading block file");
        const std::vector<MarkerBlock> blocks_local = shared ? std::vector<MarkerBlock>() : read_blocks(a.block_path);
        const std::vector<MarkerBlock> &blocks = shared ? shared->blocks : blocks_local;
        if (a.block_ix < 0 || (size_t)a.block_ix >= blocks.size()) fail("block index out of range");
        block = blocks[(size_t)a.block_ix];
    }
    say("Loading pxp");
    const Pxp pxp_local = shared ? Pxp() : load_pxp(a.pxp_path, a.hetcor ? a.pxp_se_path : std::string(), a.pearson_sample_size);
    const Pxp &pxp = shared ? shared->pxp : pxp_local;
    const size_t num_phen = pxp.num_phen;
    std::vector<int> time_index_traits(num_phen, 1);
    if (shared) {
        time_index_traits = shared->time_index_traits;
    } else if (a.time_indexed) {
        say("Loading time_indices");
        time_index_traits = read_ints_from_lines(a.time_index_path);
        if (time_index_traits.size() < num_phen) fail("time index file has fewer entries than traits");
    }
    const float th = cusk_hetcor_threshold(a.alpha);
    Reduced gc;
    std::string stem;
    bool stage_one_done = false;
    if (a.trait_only) {
        gc.num_var = num_phen;
        gc.num_phen = num_phen;
        gc.max_level = a.max_level_one;
        gc.C = pxp.corr;
        gc.uniform = !a.hetcor;
        gc.n_uniform = a.pearson_sample_size;
        if (a.hetcor) gc.N = pxp.ess;
        stem = "trait_only";
    } else {
        say("Loading mxp summary stats");
        MxpRows mxp;
        const float *mxp_corr = nullptr, *mxp_ess = nullptr;
        size_t mxp_rows = 0;
        if (shared && !a.merged) {
            // rows [first_global, last_global] of the preloaded table
            const MxpRows &all = shared->mxp_all;
            const size_t f0 = std::min(block.first_global(), all.num_markers);
            const size_t f1 = std::min(block.last_global() + 1, all.num_markers);
            if (pxp.num_phen != all.num_phen) fail("Numbers of traits seem to differ between pxp and mxp");
            mxp_rows = f1 - f0;
            mxp_corr = all.corr.data() + f0 * all.num_phen;
            mxp_ess = all.ess.empty() ? nullptr : all.ess.data() + f0 * all.num_phen;
        } else {
            mxp = load_mxp(a.mxp_path, a.hetcor ? a.mxp_se_path : std::string(), !a.merged,
                           a.merged ? 0 : block.first_global(), a.merged ? 0 : block.last_global(), marker_ixs);
            if (pxp.num_phen != mxp.num_phen) fail("Numbers of traits seem to differ between pxp and mxp");
            mxp_rows = mxp.num_markers;
            mxp_corr = mxp.corr.data();
            mxp_ess = mxp.ess.empty() ? nullptr : mxp.ess.data();
        }
        static const bool host_assembly = getenv("CUSK_HOST_ASSEMBLY") != nullptr;
        if (!a.merged && !(ctx->shard && ctx->shard->world > 1) && !host_assembly) {
            // a block: square matrix built on the device from the uploaded triangle (cuskss_block_first_stage)
            say("Loading mxm");
            say("Merging correlations into single matrix");
            say("Starting first cusk stage");
            double load_mxm_s = 0, dev_s = 0;
            int rc1 = cuskss_block_first_stage(ctx, a, mxp_corr, mxp_ess, mxp_rows, pxp, time_index_traits, th, &gc,
                                               &rep->stage[0], &load_mxm_s, &dev_s);
            if (rc1) return rc1;
            rep->stages = 1;
            rep->num_var_in = mxp_rows + num_phen;
            rep->load_s = (now_s() - t0) - dev_s;
            rep->device_s = dev_s;
            stage_one_done = true;
            stem = block.stem();
        } else {
            say("Loading mxm");
            size_t m = 0;
            const std::vector<float> mxm = load_mxm_square(a.mxm_path, &m);
            if (m != mxp_rows) {
                fail("Numbers of markers seem to differ between mxm and mxp\nmxp: " + std::to_string(mxp_rows) + " x " +
                     std::to_string(pxp.num_phen) + "\nmxm: " + std::to_string(m) + " x " + std::to_string(m));
            }
            say("Merging correlations into single matrix");
            MxpRows view;
            view.num_markers = mxp_rows;
            view.num_phen = pxp.num_phen;
            view.corr.assign(mxp_corr, mxp_corr + mxp_rows * pxp.num_phen);
            if (mxp_ess) view.ess.assign(mxp_ess, mxp_ess + mxp_rows * pxp.num_phen);
            SquareInputs si = make_square_cuskss_inputs(mxm, m, view, pxp, a.pearson_sample_size, a.hetcor);
            gc.num_var = m + num_phen;
            gc.num_phen = num_phen;
            gc.max_level = a.max_level_one;
            gc.C = std::move(si.C);
            gc.N = std::move(si.N);
            gc.uniform = si.uniform;
            gc.n_uniform = si.n_uniform;
            stem = a.merged ? std::string("cuskss_merged") : block.stem();
        }
    }
    if (!stage_one_done) {
        gc.new_to_old.resize(gc.num_var);
        for (size_t i = 0; i < gc.num_var; i++) gc.new_to_old[i] = (int)i;
        rep->num_var_in = gc.num_var;
        rep->load_s = now_s() - t0;
    }

    const double t1 = now_s();
    int rc = CUSK_OK;
    if (!stage_one_done) {
        say("Starting first cusk stage");
        rc = cuskss_stage(ctx, gc, true, th, a.depth, a.max_level_one, time_index_traits, &rep->stage[0]);
        if (rc) return rc;
        rep->stages = 1;
    }
    if (a.two_stage && !a.trait_only) {
        say("Starting second cusk stage");
        rc = cuskss_stage(ctx, gc, false, th, a.depth, a.max_level_two, time_index_traits, &rep->stage[1]);
        if (rc) return rc;
        rep->stages = 2;
    }
    rep->device_s += now_s() - t1;
    if (verbose) std::cout << "Retained " << gc.num_markers() << " markers" << std::endl;
    const double t2 = now_s();
    if (a.write_output) gc.to_file(make_path(a.outdir, stem, ""));
    rep->write_s = now_s() - t2;
    rep->num_var_out = gc.num_var;
    rep->retained_markers = gc.num_markers();
    return CUSK_OK;
}

// ------------------------------------------------------------------------------------------------
// cusk: individual-level data (cli.cpp:432-678)
// ------------------------------------------------------------------------------------------------
void load_cusk_shared(const CuskArgs &a, CuskShared *s, bool verbose)
{
    for (const char *sfx : {".bed", ".dim", ".means", ".stds", ".bim", ".fam"}) check_path(a.bfiles + sfx);  // io.cpp:408-415
    check_path(a.phen_path);
    check_path(a.block_path);
    check_path(a.outdir);
    s->phen = load_phen(a.phen_path);
    s->dims = read_dims(a.bfiles + ".dim");
    if (s->phen.num_samples != s->dims.num_samples) {
        std::cerr << "different num samples in phen and dims" << std::endl;
        fail("different num samples in phen and dims");
    }
    BimInfo bim(a.bfiles + ".bim");
    if (verbose) std::cout << "Found " << s->phen.num_phen << " phenotypes" << std::endl;
    if (verbose) std::cout << "Loading blocks" << std::endl;
    s->blocks = read_blocks(a.block_path);
    if (verbose) std::cout << "Found " << s->blocks.size() << " blocks" << std::endl;
    s->block_chr_start.resize(s->blocks.size());
    for (size_t i = 0; i < s->blocks.size(); i++) {
        const MarkerBlock &b = s->blocks[i];
        const size_t ci = bim.chr_index(b.chr);
        if (b.first >= bim.num_on_chr[ci] || b.last >= bim.num_on_chr[ci]) {
            std::cerr << "block out of bounds with first_ix: " << b.first << " last_ix: " << b.last << std::endl;
            fail("block out of bounds");
        }
        s->block_chr_start[i] = bim.chr_start[ci];
    }
    s->Th.resize(15);
    cusk_threshold_array((int)s->dims.num_samples, a.alpha, s->Th.data());
    s->means = read_floats_from_line_range(a.bfiles + ".means", 0, (size_t)-1);
    s->stds = read_floats_from_line_range(a.bfiles + ".stds", 0, (size_t)-1);
}

int run_cusk_block(cusk_ctx *ctx, const CuskArgs &a, const CuskShared &s, size_t bid, bool verbose, RunReport *rep)
{
    RunReport local;
    if (!rep) rep = &local;
    *rep = RunReport();
    if (bid >= s.blocks.size()) fail("block index out of range");
    const MarkerBlock &block = s.blocks[bid];
    const size_t m = block.size(), n = s.dims.num_samples, p = s.phen.num_phen, nv = m + p;
    const size_t col_bytes = s.dims.col_bytes();
    const size_t g0 = s.block_chr_start[bid] + block.first;
    if (verbose) {
        std::cout << std::endl << "Processing block " << bid + 1 << " / " << s.blocks.size() << std::endl;
        std::cout << "Block size: " << m << std::endl << "Loading bed data" << std::endl;
    }
    if (g0 + m > s.means.size() || g0 + m > s.stds.size()) fail("block size and number of means or stds differ");
    const double t0 = now_s();
    cudaStream_t st = (cudaStream_t)cusk_ctx_stream(ctx);
    HOST_CUDA(cudaSetDevice(ctx->device));
    // ---- load the block (io.cpp:238-250) straight into page-locked memory -----------------------
    static thread_local Pinned bed_stage;
    unsigned char *h_bed = (unsigned char *)bed_stage.get(std::max<size_t>(m * col_bytes, 1));
    // a short read must fail: code 00 is a valid genotype, so zero fill would give plausible wrong correlations
    read_file_range(a.bfiles + ".bed", 3 + g0 * col_bytes, m * col_bytes, h_bed);
    rep->load_s = now_s() - t0;
    const double t1 = now_s();
    arena_reset(ctx->io);
    unsigned char *d_bed = (unsigned char *)arena_take(ctx->io, m * col_bytes);
    float *d_phen = (float *)arena_take(ctx->io, std::max<size_t>(p * n, 1) * 4);
    float *d_mean = (float *)arena_take(ctx->io, m * 4);
    float *d_std = (float *)arena_take(ctx->io, m * 4);
    float *d_mxp = (float *)arena_take(ctx->io, std::max<size_t>(m * p, 1) * 4);
    if (!d_bed || !d_phen || !d_mean || !d_std || !d_mxp) return CUSK_E_CUDA;
    HOST_CUDA(cudaMemcpyAsync(d_bed, h_bed, m * col_bytes, cudaMemcpyHostToDevice, st));
    HOST_CUDA(cudaMemcpyAsync(d_phen, s.phen.data.data(), p * n * 4, cudaMemcpyHostToDevice, st));
    HOST_CUDA(cudaMemcpyAsync(d_mean, s.means.data() + g0, m * 4, cudaMemcpyHostToDevice, st));
    HOST_CUDA(cudaMemcpyAsync(d_std, s.stds.data() + g0, m * 4, cudaMemcpyHostToDevice, st));
    // ---- marker x trait screen (cli.cpp:548-576) ---------------------------------------------------
    if (verbose) std::cout << "Checking for significant marker - phen correlations" << std::endl;
    arena_reset(ctx);
    int rc = corr_block_device(ctx, d_bed, d_phen, m, n, p, d_mean, d_std, 0, nullptr, d_mxp, nullptr, nullptr, false,
                               nullptr, false);
    if (rc) return rc;
    std::vector<float> mxp(m * p);
    HOST_CUDA(cudaMemcpyAsync(mxp.data(), d_mxp, m * p * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    int num_sig = 0;
    for (float c : mxp) num_sig += (fabs(0.5 * (log(fabs((1 + c))) - log(fabs(1 - c)))) >= s.Th[0]);
    if (num_sig > 0) {
        if (verbose) std::cout << "Found " << num_sig << " marker - phen correlations. Proceeding." << std::endl;
    } else {
        if (verbose) std::cout << "No significant correlations found. Skipping block." << std::endl;
        rep->skipped = true;
        rep->device_s = now_s() - t1;
        return CUSK_OK;
    }
    // ---- all correlations, dense on the device (cli.cpp:581-649) -------------------------------------
    if (verbose) std::cout << "Computing all correlations" << std::endl;
    float *d_C = (float *)arena_take(ctx->io, nv * nv * 4);
    const size_t rw = cusk_row_words((int)nv);
    uint32_t *d_bits = (uint32_t *)arena_take(ctx->io, nv * rw * 4);
    if (!d_C || !d_bits) return CUSK_E_CUDA;
    arena_reset(ctx);
    // level 0 comes out of the correlation pass (the Gram epilogue holds r in a register; the trait strip is a
    // few thousand pairs): the engine below starts at level 1.  CUSK_FUSE_LEVEL0=0 runs level 0 as its own pass.
    const char *fuse_env = getenv("CUSK_FUSE_LEVEL0");
    const bool fuse0 = !(fuse_env && atoi(fuse_env) == 0);
    Level0Fuse l0 = {d_bits, (int)rw, (double)s.Th[0]};
    rc = corr_block_device(ctx, d_bed, d_phen, m, n, p, d_mean, d_std, 0, d_C, d_mxp, nullptr, nullptr, true, &rep->corr,
                           true, fuse0 ? &l0 : nullptr);
    if (rc) return rc;
    // ---- first skeleton stage (cli.cpp:662-673) -------------------------------------------------------
    if (verbose) std::cout << "Running cuPC" << std::endl;
    int last = 0;
    rc = skeleton_engine(ctx, d_C, (int)nv, d_bits, s.Th.data(), a.max_level, &last, nullptr, nullptr, &rep->stage[0],
                         fuse0);
    if (rc) return rc;
    std::vector<uint32_t> bits(nv * rw);
    HOST_CUDA(cudaMemcpyAsync(bits.data(), d_bits, nv * rw * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    const std::vector<int> sel1 = subset_variables(bits.data(), nv, rw, m, a.depth);
    const size_t n1 = sel1.size(), m1 = n1 - p;
    // reduce_gcs #1: only C (gathered on the device) and the index map survive into stage two
    int *d_sel = (int *)arena_take(ctx->io, n1 * 4);
    float *d_C1 = (float *)arena_take(ctx->io, n1 * n1 * 4);
    const size_t rw1 = cusk_row_words((int)n1);
    uint32_t *d_bits1 = (uint32_t *)arena_take(ctx->io, n1 * rw1 * 4);
    if (!d_sel || !d_C1 || !d_bits1) return CUSK_E_CUDA;
    HOST_CUDA(cudaMemcpyAsync(d_sel, sel1.data(), n1 * 4, cudaMemcpyHostToDevice, st));
    gather_square_kernel<<<dim3((unsigned)((n1 + 255) / 256), (unsigned)n1), 256, 0, st>>>(d_C, (int)nv, d_sel, (int)n1, d_C1);
    HOST_CUDA(cudaGetLastError());
    // ---- second stage (reduced_gcs_cusk, cli.cpp:62-87): level 0 re-initialises G -----------------------
    if (verbose) std::cout << "Starting second cusk stage" << std::endl;
    std::vector<cusk_sepset> seps;
    rc = skeleton_engine(ctx, d_C1, (int)n1, d_bits1, s.Th.data(), a.max_level_two, &last, nullptr, &seps, &rep->stage[1]);
    if (rc) return rc;
    rep->stages = 2;
    std::vector<uint32_t> bits1(n1 * rw1);
    std::vector<float> C1(n1 * n1);
    HOST_CUDA(cudaMemcpyAsync(bits1.data(), d_bits1, n1 * rw1 * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaMemcpyAsync(C1.data(), d_C1, n1 * n1 * 4, cudaMemcpyDeviceToHost, st));
    HOST_CUDA(cudaStreamSynchronize(st));
    rep->device_s = now_s() - t1;
    const double t2 = now_s();
    const std::vector<int> sel2 = subset_variables(bits1.data(), n1, rw1, m1, a.depth);
    const size_t n2 = sel2.size();
    // reduce_gcs #2 with index_map = stage one's new_to_old (parent_set.cpp:130-175)
    Reduced out;
    out.num_var = n2;
    out.num_phen = p;
    out.max_level = CUSK_ML;
    out.has_sep = true;
    out.new_to_old.resize(n2);
    for (size_t i = 0; i < n2; i++) out.new_to_old[i] = sel1[(size_t)sel2[i]];
    out.G = expand_bits(bits1, n1, rw1, sel2);
    out.C.resize(n2 * n2);
    for (size_t i = 0; i < n2; i++)
        for (size_t j = 0; j < n2; j++) out.C[i * n2 + j] = C1[(size_t)sel2[i] * n1 + sel2[j]];
    out.S.assign(n2 * n2 * CUSK_ML, -1);
    // The reference translates separating-set members through VariableSubsetIndices(P, index_map)
    // (parent_set.h:156-166), whose old_to_new map is keyed by the ORIGINAL index
    // index_map[sorted P[i]], while it is queried with the stage-two index of the member; a key that
    // is absent yields 0 (std::unordered_map::operator[]).  Reproduced as is: .sep is a file contract.
    std::unordered_map<int, int> old_to_new;
    for (size_t i = 0; i < n2; i++) old_to_new[out.new_to_old[i]] = (int)i;
    std::vector<int> pos_in_sel2(n1, -1);
    for (size_t i = 0; i < n2; i++) pos_in_sel2[(size_t)sel2[i]] = (int)i;
    for (const cusk_sepset &r : seps) {
        const int i = pos_in_sel2[(size_t)r.x], j = pos_in_sel2[(size_t)r.y];
        if (i < 0 || j < 0) continue;
        int *slot = out.S.data() + ((size_t)i * n2 + (size_t)j) * CUSK_ML;
        int k = 0;
        for (int q = 0; q < r.level; q++) {
            const int e = r.s[q];
            if (e != -1 && pos_in_sel2[(size_t)e] >= 0) {
                auto it = old_to_new.find(e);
                slot[k++] = (it == old_to_new.end()) ? 0 : it->second;
            }
        }
    }
    if (verbose) std::cout << "Retained " << out.num_markers() << " markers" << std::endl;
    out.to_file(make_path(a.outdir, block.stem(), ""));
    rep->write_s = now_s() - t2;
    rep->num_var_in = nv;
    rep->num_var_out = n2;
    rep->retained_markers = out.num_markers();
    return CUSK_OK;
}

// ------------------------------------------------------------------------------------------------
// prep-bed (prep.cpp:157-201): .dim, .means, .stds, .modes
// ------------------------------------------------------------------------------------------------
int run_prep(cusk_ctx *ctx, const std::string &bfiles, bool verbose)
{
    for (const char *sfx : {".bed", ".bim", ".fam"}) check_path(bfiles + sfx);
    if (!has_valid_bed_prefix(bfiles + ".bed")) fail("Invalid prefix bytes in bed");
    const size_t n = count_lines(bfiles + ".fam");
    BimInfo bim(bfiles + ".bim");
    const size_t m_total = bim.num_lines;
    const size_t col_bytes = (n + 3) / 4;
    if (verbose) std::cout << "Writing dim file." << std::endl;
    {
        FILE *f = fopen((bfiles + ".dim").c_str(), "w");
        if (!f) fail("cannot write " + bfiles + ".dim");
        fprintf(f, "%zu\t%zu\n", n, m_total);
        fclose(f);
    }
    if (verbose) std::cout << "Computing means, stds, modes." << std::endl;
    std::vector<float> means, stds;
    std::vector<int> modes;
    cudaStream_t st = (cudaStream_t)cusk_ctx_stream(ctx);
    HOST_CUDA(cudaSetDevice(ctx->device));
    FILE *f = fopen((bfiles + ".bed").c_str(), "rb");
    if (!f) fail("file or directory not found: " + bfiles + ".bed");
    fseek(f, 3, SEEK_SET);
    // stream the file in chunks of markers; only whole columns count (the reference's read loop
    // stops at the first short read, prep.cpp:185)
    const size_t chunk_markers = std::max<size_t>(1, std::min<size_t>(m_total ? m_total : 1, ((size_t)256 << 20) / std::max<size_t>(col_bytes, 1)));
    Pinned stage;
    unsigned char *h = (unsigned char *)stage.get(chunk_markers * col_bytes);
    for (;;) {
        const size_t got = fread(h, 1, chunk_markers * col_bytes, f);
        const size_t mk = got / col_bytes;
        if (mk == 0) break;
        arena_reset(ctx->io);
        unsigned char *d_bed = (unsigned char *)arena_take(ctx->io, mk * col_bytes);
        int *d_counts = (int *)arena_take(ctx->io, mk * 16);
        float *d_means = (float *)arena_take(ctx->io, mk * 4);
        float *d_stds = (float *)arena_take(ctx->io, mk * 4);
        int *d_modes = (int *)arena_take(ctx->io, mk * 4);
        if (!d_bed || !d_counts || !d_means || !d_stds || !d_modes) {
            fclose(f);
            return CUSK_E_CUDA;
        }
        const size_t base = means.size();
        means.resize(base + mk);
        stds.resize(base + mk);
        modes.resize(base + mk);
        cudaMemcpyAsync(d_bed, h, mk * col_bytes, cudaMemcpyHostToDevice, st);
        launch_bed_counts(st, d_bed, (int)mk, (int)n, col_bytes, d_counts);
        launch_bed_stats(st, d_bed, (int)mk, (int)n, col_bytes, d_counts, d_means, d_stds, d_modes);
        cudaMemcpyAsync(means.data() + base, d_means, mk * 4, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(stds.data() + base, d_stds, mk * 4, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(modes.data() + base, d_modes, mk * 4, cudaMemcpyDeviceToHost, st);
        const cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) {
            fclose(f);
            cusk_set_error("prep: %s", cudaGetErrorString(e));
            return CUSK_E_CUDA;
        }
        if (got < chunk_markers * col_bytes) break;
    }
    fclose(f);
    if (verbose) std::cout << "Writing stats to files." << std::endl;
    write_column(means, bfiles + ".means");
    write_column(stds, bfiles + ".stds");
    write_column(modes, bfiles + ".modes");
    if (verbose) std::cout << "Done." << std::endl;
    return CUSK_OK;
}

// ---------------------------------------------------------------------------------------------------------
// `mps block <bfiles> <max-block-size> <device-mem-gb> <corr-width>` (make_blocks, src/cli.cpp:362-411): per
// chromosome of the .bim, banded correlations + row |r| sums on the device (cusk_mcorrk_banded; only the m row
// sums come back), Hann smoothing / minima / window bisection on the host (blocking.cpp), blocks APPENDED to
// <bfiles>_m<max-block-size>.blocks (write_marker_blocks_to_file, src/io.cpp:266-277).
// ---------------------------------------------------------------------------------------------------------
// One chromosome: its columns from the open .bed, banded correlations + row sums on ctx's device, blocks on the host.
int block_chromosome(cusk_ctx *ctx, FILE *bed, const std::string &bfiles, const BimInfo &bim, size_t c, size_t n,
                     int max_block_size, float device_mem_gb, size_t corr_width, bool verbose, BlockPartition *out,
                     cusk_band_stats *bs)
{
    const std::string &cid = bim.chr_ids[c];
    const size_t m = bim.num_on_chr[c], col_bytes = (n + 3) / 4;
    if (verbose) std::cout << "[Chr " << cid << "]: Loading bed data for " << m << " markers." << std::endl;
    Pinned stage;
    unsigned char *h = (unsigned char *)stage.get(m * col_bytes);
    if (!h) fail("out of host memory for chromosome " + cid);
    if (fseeko(bed, (off_t)(3 + bim.chr_start[c] * col_bytes), SEEK_SET) != 0 ||
        fread(h, 1, m * col_bytes, bed) != m * col_bytes)
        fail("short read of " + bfiles + ".bed (chromosome " + cid + ")");
    if (verbose) std::cout << "[Chr " << cid << "]: Computing correlations and row sums." << std::endl;
    std::vector<float> sums(m, 0.f);
    const int rc = cusk_mcorrk_banded(ctx, h, m, n, corr_width, device_mem_gb, nullptr, sums.data(), bs);
    if (rc == CUSK_E_ARG) fail(cusk_last_error());  // exit(1) like the reference's batch-size check
    if (rc != CUSK_OK) return rc;
    if (verbose) std::cout << "[Chr " << cid << "]: Making blocks." << std::endl;
    *out = block_chr(sums.data(), m, max_block_size);
    return CUSK_OK;
}

namespace {

void append_blocks(const std::string &out_path, const std::string &cid, const BlockPartition &bp)
{
    FILE *out = fopen(out_path.c_str(), "a");
    if (!out) fail("cannot write " + out_path);
    for (size_t b = 0; b < bp.first.size(); b++) fprintf(out, "%s\t%zu\t%zu\n", cid.c_str(), bp.first[b], bp.last[b]);
    fclose(out);
}

void log_chromosome(const std::string &cid, size_t m, int device, const cusk_band_stats &bs, const BlockPartition &bp)
{
    printf("{\"event\": \"block_chr\", \"chr\": \"%s\", \"device\": %d, \"markers\": %zu, \"pairs\": %lld, "
           "\"chunks\": %d, \"h2d_ms\": %.3f, \"kernel_ms\": %.3f, \"window\": %d, \"blocks\": %zu, \"largest\": %zu}\n",
           cid.c_str(), device, m, (long long)bs.pairs, bs.chunks, bs.h2d_ms, bs.kernel_ms, bp.window, bp.first.size(),
           bp.largest);
    fflush(stdout);
}

struct FileCloser {
    FILE *f;
    ~FileCloser()
    {
        if (f) fclose(f);
    }
};

}  // namespace

int run_block(cusk_ctx *ctx, const std::string &bfiles, int max_block_size, float device_mem_gb, size_t corr_width,
              bool verbose)
{
    for (const char *sfx : {".bed", ".bim", ".fam"}) check_path(bfiles + sfx);
    if (!has_valid_bed_prefix(bfiles + ".bed")) fail("Invalid prefix bytes in bed");
    const size_t n = count_lines(bfiles + ".fam");
    BimInfo bim(bfiles + ".bim");
    const std::string out_path = bfiles + "_m" + std::to_string(max_block_size) + ".blocks";
    FileCloser bed{fopen((bfiles + ".bed").c_str(), "rb")};
    if (!bed.f) fail("file or directory not found: " + bfiles + ".bed");
    for (size_t c = 0; c < bim.chr_ids.size(); c++) {
        const std::string &cid = bim.chr_ids[c];
        BlockPartition bp;
        cusk_band_stats bs;
        const int rc = block_chromosome(ctx, bed.f, bfiles, bim, c, n, max_block_size, device_mem_gb, corr_width, verbose,
                                        &bp, &bs);
        if (rc != CUSK_OK) return rc;
        if (verbose) {
            std::cout << "[Chr " << cid << "]: Partitioned into " << bp.first.size() << " blocks." << std::endl;
            std::cout << "[Chr " << cid << "]: Writing blocks to output file." << std::endl;
            log_chromosome(cid, bim.num_on_chr[c], ctx->device, bs, bp);
        }
        append_blocks(out_path, cid, bp);
    }
    if (verbose) std::cout << "Done." << std::endl;
    return CUSK_OK;
}

// Chromosomes are independent units (no collective): one worker thread per device takes them largest
// first; the blocks are appended in .bim order once all are done, so the file equals the one-device run's.
int run_block_all(const std::string &bfiles, int max_block_size, float device_mem_gb, size_t corr_width,
                  const std::vector<int> &devices)
{
    for (const char *sfx : {".bed", ".bim", ".fam"}) check_path(bfiles + sfx);
    if (!has_valid_bed_prefix(bfiles + ".bed")) fail("Invalid prefix bytes in bed");
    if (devices.empty()) fail("no CUDA device available");
    const size_t n = count_lines(bfiles + ".fam");
    BimInfo bim(bfiles + ".bim");
    const size_t nchr = bim.chr_ids.size();
    std::vector<size_t> order(nchr);
    for (size_t i = 0; i < nchr; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return bim.num_on_chr[a] > bim.num_on_chr[b]; });
    std::vector<BlockPartition> parts(nchr);
    std::atomic<size_t> next{0};
    std::mutex mu;
    int first_rc = CUSK_OK, exit_code = 0;
    std::string first_error;
    std::vector<std::thread> workers;
    for (int dev : devices) {
        workers.emplace_back([&, dev]() {
            cusk_ctx *ctx = nullptr;
            FileCloser bed{fopen((bfiles + ".bed").c_str(), "rb")};
            int rc = bed.f ? cusk_ctx_create(dev, &ctx) : CUSK_E_ARG;
            std::string err = rc ? (bed.f ? cusk_last_error() : "file or directory not found: " + bfiles + ".bed") : "";
            int code = 0;
            while (rc == CUSK_OK) {
                const size_t k = next.fetch_add(1);
                if (k >= nchr) break;
                const size_t c = order[k];
                cusk_band_stats bs;
                try {
                    rc = block_chromosome(ctx, bed.f, bfiles, bim, c, n, max_block_size, device_mem_gb, corr_width, false,
                                          &parts[c], &bs);
                    if (rc) err = cusk_last_error();
                } catch (const HostError &e) {
                    rc = CUSK_E_ARG;
                    err = e.msg;
                    code = e.code;
                }
                if (rc == CUSK_OK) {
                    std::lock_guard<std::mutex> lk(mu);
                    log_chromosome(bim.chr_ids[c], bim.num_on_chr[c], dev, bs, parts[c]);
                }
            }
            if (ctx) cusk_ctx_destroy(ctx);
            if (rc != CUSK_OK) {
                next.store(nchr);  // let the other workers finish what they hold and stop
                std::lock_guard<std::mutex> lk(mu);
                if (first_rc == CUSK_OK) {
                    first_rc = rc;
                    first_error = err;
                    exit_code = code;
                }
            }
        });
    }
    for (auto &w : workers) w.join();
    if (first_rc != CUSK_OK) {
        if (exit_code) fail(first_error, exit_code);
        cusk_set_error("%s", first_error.c_str());
        return first_rc;
    }
    const std::string out_path = bfiles + "_m" + std::to_string(max_block_size) + ".blocks";
    for (size_t c = 0; c < nchr; c++) append_blocks(out_path, bim.chr_ids[c], parts[c]);
    std::cout << "Done." << std::endl;
    return CUSK_OK;
}

}  // namespace cuskhost
