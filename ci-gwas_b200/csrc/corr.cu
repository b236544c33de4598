// corr.cu -- host drivers of the correlation front end (replaces cu_marker_phen_corr_pearson and
// cu_corr_pearson_npn, src/corr_host.cu:1023-1197, and the statistics pass of prep-bed).
#include <cuda.h>

#include <algorithm>
#include <vector>

#include "corr.cuh"

// corr_kernels.cu
void launch_bed_counts(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int *counts);
void launch_bed_stats(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, const int *counts,
                      float *means, float *stds, int *modes);
void launch_decode(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int parts, int MI,
                   size_t Kp, unsigned char *ind);
cudaError_t clear_indicator_padding(cudaStream_t st, unsigned char *ind, size_t m, int parts, int MI, size_t Kp,
                                    size_t rows_total);
int launch_mxp(cudaStream_t st, const unsigned char *bed, const float *phen, int m, int n, int p, size_t col_bytes,
               const float *mean, const float *stdv, float *out, bool any_missing, double *d_Ty, int *d_cnty);
void launch_trait_rows(cudaStream_t st, const float *phen, int n, int p, size_t Kp, unsigned char *rows, double *d_Ty,
                       int *d_cnty, double *d_scale_inv, int *d_sexp);
void launch_pxp(cudaStream_t st, const float *phen, int n, int p, float *out_upper, float *Csq, int ldc, int m_off);
void launch_pack_upper(cudaStream_t st, const float *Csq, int m, int ldc, float *upper);
void launch_set_diag(cudaStream_t st, float *Csq, int nvar, int ldc);
// corr_gram.cu
int launch_gram(cudaStream_t st, const CUtensorMap &map_a, const CUtensorMap &map_b, const GramParams &p, int grid);
int launch_gram_pair(cudaStream_t st, const CUtensorMap &map_a, const CUtensorMap &map_b, const GramParams &p, int grid);
int gram_make_maps(const unsigned char *ind, size_t rows, size_t Kp, CUtensorMap *map_a, CUtensorMap *map_b);
cusk_ctx *cusk_default_ctx();

namespace {

__global__ void scatter_mxp_kernel(const float *__restrict__ mxp, int m, int p, float *__restrict__ Csq, int ldc)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= m * p) return;
    const int i = idx / p, t = idx - i * p;
    const float r = mxp[idx];
    Csq[(size_t)i * ldc + (m + t)] = r;
    Csq[(size_t)(m + t) * ldc + i] = r;
}

// level 0 of the pairs the Gram epilogue does not see: marker x trait and trait x trait (the strip of C right of /
// below the markers); same expression and the same upper-triangle entry as level0_kernel
__global__ void __launch_bounds__(256) level0_strip_kernel(const float *__restrict__ Csq, int nvar, int m, int p, double th,
                                                           uint32_t *bits, int rw)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int r, c;
    if (idx < m * p) {
        r = idx / p;
        c = m + (idx - r * p);
    } else {
        int k = idx - m * p;  // trait pairs a < b, row-major
        if (k >= p * (p - 1) / 2) return;
        int a = 0;
        while (k >= p - 1 - a) {
            k -= p - 1 - a;
            a++;
        }
        r = m + a;
        c = m + a + 1 + k;
    }
    if (!(corr_fisher_z((double)Csq[(size_t)r * nvar + c]) < th)) {
        atomicOr(&bits[(size_t)r * rw + (c >> 5)], 1u << (c & 31));
        atomicOr(&bits[(size_t)c * rw + (r >> 5)], 1u << (r & 31));
    }
}

struct EvTimer {
    cudaEvent_t a, b;
    cudaStream_t st;
    bool used = false;
    EvTimer(cusk_ctx *ctx, int slot) : a(ctx->ev[2 * slot]), b(ctx->ev[2 * slot + 1]), st(ctx->stream) {}
    void start()
    {
        cudaEventRecord(a, st);
        used = true;
    }
    void stop() { cudaEventRecord(b, st); }
    float ms()
    {
        if (!used) return 0.f;
        float t = 0.f;
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&t, a, b);
        return t;
    }
};

}  // namespace

// Device-resident block: everything below works on device pointers; scratch from ctx->arena.
// d_Csq may be nullptr when only mxp is wanted (the level-0 screen of cusk, cli.cpp:548-576).
// mxp_ready: d_mxp_out already holds the block's marker x trait correlations (the screen computed
// them; the reference runs that kernel twice, corr_host.cu:1070,1164).
int corr_block_device(cusk_ctx *ctx, const unsigned char *d_bed, const float *d_phen, size_t m, size_t n, size_t p,
                      const float *d_mean, const float *d_std, int mode, float *d_Csq, float *d_mxp_out,
                      float *d_pxp_out, long long *d_tables, bool want_mxm, cusk_corr_stats *st, bool mxp_ready,
                      const Level0Fuse *l0)
{
    cudaStream_t s = ctx->stream;
    if (l0 && (!d_Csq || !want_mxm || mode != 0)) {
        cusk_set_error("corr: the fused level 0 needs the dense matrix of the npn correlations");
        return CUSK_E_ARG;
    }
    if (l0) CUSK_CUDA(cudaMemsetAsync(l0->bits, 0, sizeof(uint32_t) * (size_t)(m + p) * (size_t)l0->rw, s));
    const size_t col_bytes = (n + 3) / 4;
    const int nvar = (int)(m + p);
    EvTimer t_counts(ctx, 0), t_decode(ctx, 1), t_gram(ctx, 2), t_mxp(ctx, 3), t_pxp(ctx, 4);
    if (m > 0x3fffffff || n > 0x7fffffff || p > 65535) {
        cusk_set_error("corr: dimensions out of range");
        return CUSK_E_ARG;
    }
    float *d_mxp = d_mxp_out;
    if (p > 0 && !d_mxp) {
        d_mxp = (float *)arena_take(ctx, sizeof(float) * m * p);
        if (!d_mxp) return CUSK_E_CUDA;
    }
    // ---- exact genotype counts (also tell whether the block has any missing call) ------------------
    bool any_missing = false;
    int *d_counts = nullptr;
    if (m > 0) {
        d_counts = (int *)arena_take(ctx, sizeof(int) * 4 * m);
        if (!d_counts) return CUSK_E_CUDA;
        t_counts.start();
        launch_bed_counts(s, d_bed, (int)m, (int)n, col_bytes, d_counts);
        t_counts.stop();
        std::vector<int> h_counts(4 * m);
        CUSK_CUDA(cudaMemcpyAsync(h_counts.data(), d_counts, sizeof(int) * 4 * m, cudaMemcpyDeviceToHost, s));
        CUSK_CUDA(cudaStreamSynchronize(s));
        for (size_t i = 0; i < m; i++) any_missing |= h_counts[4 * i + 3] != 0;
    }
    // ---- marker x trait -------------------------------------------------------------------------------
    // Default: on the tensor cores, as extra tiles of the Gram kernel below (trait values as uint8 limb
    // rows).  CUDA-core kernel (corr_mxp.cu) for more than 32 traits or with CUSK_MXP_CUDA_CORES=1.
    static const bool force_cuda_cores = getenv("CUSK_MXP_CUDA_CORES") != nullptr;
    const bool tc_mxp = p > 0 && m > 0 && p <= CUSK_TC_MXP_MAX_TRAITS && !force_cuda_cores;
    const bool tc_traits = tc_mxp && !mxp_ready;  // trait tiles to compute in this call
    double *d_Ty = nullptr, *d_scale = nullptr;
    int *d_cnty = nullptr, *d_sexp = nullptr;
    if (p > 0 && m > 0) {
        d_Ty = (double *)arena_take(ctx, sizeof(double) * p);
        d_cnty = (int *)arena_take(ctx, sizeof(int) * p);
        d_scale = (double *)arena_take(ctx, sizeof(double) * p);
        d_sexp = (int *)arena_take(ctx, sizeof(int) * p);
        if (!d_Ty || !d_cnty || !d_scale || !d_sexp) return CUSK_E_CUDA;
        if (!tc_mxp) {
            t_mxp.start();
            if (!mxp_ready) {
                int rc = launch_mxp(s, d_bed, d_phen, (int)m, (int)n, (int)p, col_bytes, d_mean, d_std, d_mxp,
                                    any_missing, d_Ty, d_cnty);
                if (rc) return rc;
            }
            if (d_Csq)
                scatter_mxp_kernel<<<(unsigned)((m * p + 255) / 256), 256, 0, s>>>(d_mxp, (int)m, (int)p, d_Csq, nvar);
            t_mxp.stop();
        }
    }
    if (p > 1 && (d_pxp_out || d_Csq)) {
        t_pxp.start();
        launch_pxp(s, d_phen, (int)n, (int)p, d_pxp_out, d_Csq, nvar, (int)m);
        t_pxp.stop();
    }
    CUSK_CUDA(cudaGetLastError());
    if (st) st->bed_bytes = (int64_t)(m * col_bytes);
    // ---- marker x marker: indicator Gram on the tensor cores ----------------------------------------
    const bool mxm_tiles = want_mxm && m > 1;
    if (mxm_tiles || tc_traits) {
        const int parts = any_missing ? 3 : 2;
        const int MI = 128 / parts;
        const size_t groups = (m + MI - 1) / MI;
        const size_t rows = ((groups * 128 + 255) / 256) * 256;
        const size_t Kp = ((n + 127) / 128) * 128;
        // operand matrix: the genotype indicator rows, then (tensor-core mxp) one 256-row block of trait rows
        const size_t rows_total = rows + (tc_traits ? 256 : 0);
        const int trait_tn = tc_traits ? (int)(rows / 256) : -1;
        unsigned char *d_ind = (unsigned char *)arena_take(ctx, rows_total * Kp);
        if (!d_ind) return CUSK_E_CUDA;
        t_decode.start();
        // only the rows no marker owns are cleared (clearing all 9.4 GB of the C3 operand matrix first cost 1.3 ms of the
        // 3.3 ms stage); the trait block is among them: launch_trait_rows fills five of a trait's eight rows
        CUSK_CUDA(clear_indicator_padding(s, d_ind, m, parts, MI, Kp, rows_total));
        launch_decode(s, d_bed, (int)m, (int)n, col_bytes, parts, MI, Kp, d_ind);
        t_decode.stop();
        if (tc_traits) {
            t_mxp.start();
            launch_trait_rows(s, d_phen, (int)n, (int)p, Kp, d_ind + rows * Kp, d_Ty, d_cnty, d_scale, d_sexp);
            t_mxp.stop();
        }
        // upper-triangle tile list: A block tm = group tm; B block tn = groups 2tn, 2tn+1
        std::vector<int2> tiles;
        const int n_tm = (int)groups, n_tn = (int)((groups + 1) / 2);
        if (tc_traits)
            for (int tm = 0; tm < n_tm; tm++)
                if ((size_t)tm * MI < m) tiles.push_back({tm, trait_tn});
        // Tile order = the order CTAs take them (tile t runs on CTA t % grid, so 148 consecutive tiles
        // are co-resident).  The triangle is walked in super-blocks of SB_M x SB_N tiles.  Default
        // SB_M = all, SB_N = 1: a wave shares ONE B block through L2 and streams distinct A blocks from
        // HBM.  Squarer super-blocks (16 x 9) should cut HBM traffic 4x on paper but measured worse
        // (ncu: 518 GB read vs 374 GB, L2 hit 39 % vs 50 %, profiles/r1_gram_tile_order.txt): the CTAs
        // that share a block drift apart by more k-steps than the L2 window holds.
        static const int SB_M_ENV = [] { const char *e = getenv("CUSK_GRAM_SB_M"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 0; }();
        static const int SB_N_ENV = [] { const char *e = getenv("CUSK_GRAM_SB_N"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 0; }();
        // K-panel synchronisation (GramParams.ksync): with it a wave can be a SQUARE super-block of tiles whose A and B
        // panels are both shared through L2: (16 x 128 + 9 x 256) K bytes from HBM per 144 tiles instead of 128 K per
        // tile.  CUSK_GRAM_KSYNC = k-steps between meetings (default 128 = 16 KB of K per operand row; 0 = off)
        static const int KSYNC = [] { const char *e = getenv("CUSK_GRAM_KSYNC"); return e ? atoi(e) : 128; }();
        const int grid_full = ctx->sm_count;
        const bool ksync_on = KSYNC > 0 && mxm_tiles && (size_t)n_tm * (size_t)n_tn > (size_t)2 * grid_full;
        const int SB_M = SB_M_ENV ? SB_M_ENV : (ksync_on ? 16 : 1 << 20);
        const int SB_N = SB_N_ENV ? SB_N_ENV : (ksync_on ? 9 : 1);
        // CTA pairs (cta_group::2, CUSK_GRAM_PAIR=0 for one CTA per tile): a pair of CTAs takes the two A groups 2 tm2,
        // 2 tm2 + 1 against one B block -- a 256 x 256 tile whose B rows each CTA loads only half of -- so a CTA fills
        // 32 KB of shared memory per k-step instead of 48 KB for the same 128 x 256 outputs.  The list carries the
        // pair's two entries next to each other; the grid is even, so entries 2c, 2c + 1 go to CTAs 2c, 2c + 1 = one
        // cluster.  Used with the K-panel meeting (large blocks); small blocks keep the one-CTA form.
        const char *pair_e = getenv("CUSK_GRAM_PAIR");  // read on every call: the tests compare the two forms
        const bool pair_on = !(pair_e && atoi(pair_e) == 0) && ksync_on && (grid_full % 2 == 0);
        std::vector<int> wave_tiles;
        if (pair_on) {
            tiles.clear();
            const int n_tm2 = (n_tm + 1) / 2;
            if (tc_traits)
                for (int tm2 = 0; tm2 < n_tm2; tm2++) {
                    tiles.push_back({2 * tm2, trait_tn});
                    tiles.push_back({2 * tm2 + 1, trait_tn});
                }
            const int SB_M2 = std::max(1, SB_M / 2);
            for (int sn = ((n_tn - 1) / SB_N) * SB_N; sn >= 0; sn -= SB_N)
                for (int sm = 0; sm < n_tm2; sm += SB_M2)
                    for (int tn = sn; tn < std::min(sn + SB_N, n_tn); tn++)
                        for (int tm2 = sm; tm2 < std::min(sm + SB_M2, n_tm2); tm2++)
                            if ((size_t)(2 * tn + 2) * MI > (size_t)(2 * tm2) * MI + 1 && (size_t)(2 * tm2) * MI < m) {
                                tiles.push_back({2 * tm2, tn});
                                tiles.push_back({2 * tm2 + 1, tn});
                            }
            for (size_t w = 0; w * (size_t)grid_full < tiles.size(); w++)
                wave_tiles.push_back((int)std::min<size_t>((size_t)grid_full, tiles.size() - w * (size_t)grid_full));
        } else if (!ksync_on) {
            for (int sn = ((n_tn - 1) / SB_N) * SB_N; mxm_tiles && sn >= 0; sn -= SB_N)  // long rows of the triangle first
                for (int sm = 0; sm < n_tm; sm += SB_M)
                    for (int tn = sn; tn < std::min(sn + SB_N, n_tn); tn++)
                        for (int tm = sm; tm < std::min(sm + SB_M, n_tm); tm++)
                            if ((size_t)(2 * tn + 2) * MI > (size_t)tm * MI + 1 && (size_t)tm * MI < m)
                                tiles.push_back({tm, tn});
        } else {
            // tiles in super-block order; a wave is simply `grid_full` consecutive tiles (the tail of one super-block
            // and the head of the next may share a wave: two super-blocks' panels still fit the L2 window, and no
            // slot of a wave idles -- padding the waves to whole super-blocks left 14 % of the slots empty)
            for (int sn = ((n_tn - 1) / SB_N) * SB_N; sn >= 0; sn -= SB_N)
                for (int sm = 0; sm < n_tm; sm += SB_M)
                    for (int tn = sn; tn < std::min(sn + SB_N, n_tn); tn++)
                        for (int tm = sm; tm < std::min(sm + SB_M, n_tm); tm++)
                            if ((size_t)(2 * tn + 2) * MI > (size_t)tm * MI + 1 && (size_t)tm * MI < m)
                                tiles.push_back({tm, tn});
            for (size_t w = 0; w * (size_t)grid_full < tiles.size(); w++)
                wave_tiles.push_back((int)std::min<size_t>((size_t)grid_full, tiles.size() - w * (size_t)grid_full));
        }
        int2 *d_tiles = (int2 *)arena_take(ctx, sizeof(int2) * std::max<size_t>(tiles.size(), 1));
        if (!d_tiles) return CUSK_E_CUDA;
        CUSK_CUDA(cudaMemcpyAsync(d_tiles, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, s));
        float *d_out = d_Csq;
        int ldc = nvar;
        CUtensorMap map_a, map_b;
        int rc = gram_make_maps(d_ind, rows_total, Kp, &map_a, &map_b);
        if (rc) return rc;
        GramParams gp = {};
        gp.parts = parts;
        gp.MI = MI;
        gp.m = (int)m;
        gp.n_samples = (int)n;
        gp.num_k = (int)(Kp / 128);
        gp.ntiles = (int)tiles.size();
        gp.tiles = d_tiles;
        gp.counts = d_counts;
        gp.C = d_out;
        gp.ldc = ldc;
        gp.mode = mode;
        gp.mean = d_mean;
        gp.stdv = d_std;
        gp.tables = d_tables;
        gp.trait_tn = trait_tn;
        gp.p = (int)p;
        gp.trait_scale = d_scale;
        gp.Ty = d_Ty;
        gp.cnty = d_cnty;
        gp.mxp = d_mxp;
        const int launch_grid = ksync_on ? grid_full : std::min<int>(ctx->sm_count, (int)tiles.size());
        if (ksync_on) {
            gp.ksync = KSYNC;
            gp.ksync_blocks = (gp.num_k + KSYNC - 1) / KSYNC;
            int *d_wave = (int *)arena_take(ctx, sizeof(int) * std::max<size_t>(wave_tiles.size(), 1));
            int *d_sync = (int *)arena_take(ctx, sizeof(int) * std::max<size_t>(wave_tiles.size() * (size_t)gp.ksync_blocks, 1));
            if (!d_wave || !d_sync) return CUSK_E_CUDA;
            CUSK_CUDA(cudaMemcpyAsync(d_wave, wave_tiles.data(), sizeof(int) * wave_tiles.size(), cudaMemcpyHostToDevice, s));
            CUSK_CUDA(cudaMemsetAsync(d_sync, 0, sizeof(int) * wave_tiles.size() * (size_t)gp.ksync_blocks, s));
            gp.wave_tiles = d_wave;
            gp.sync = d_sync;
        }
        if (l0) {
            gp.l0_bits = l0->bits;
            gp.l0_rw = l0->rw;
            gp.l0_th = l0->th;
        }
        t_gram.start();
        rc = pair_on ? launch_gram_pair(s, map_a, map_b, gp, launch_grid) : launch_gram(s, map_a, map_b, gp, launch_grid);
        // a device that cannot hold the grid's clusters at once: the pair list is also a valid list of 128 x 256 tiles
        if (rc == CUSK_E_UNSUPPORTED_PAIR) rc = launch_gram(s, map_a, map_b, gp, launch_grid);
        if (rc) return rc;
        t_gram.stop();
        CUSK_CUDA(cudaGetLastError());
        CUSK_CUDA(cudaStreamSynchronize(s));  // the host tile list must outlive the copy
        if (st) {
            st->parts = parts;
            size_t real_tiles = 0;
            for (const int2 &t : tiles) real_tiles += t.x >= 0;
            st->gram_tiles = (int)real_tiles;
            st->gram_int8_ops = (int64_t)real_tiles * 2 * 128 * 256 * (int64_t)Kp;
        }
    }
    // tensor-core mxp route: d_mxp is valid here whether this call computed it (trait tiles above) or the caller's
    // screen did (mxp_ready) -- also for a one-marker block, which has no Gram tile at all
    if (tc_mxp && d_Csq)
        scatter_mxp_kernel<<<(unsigned)((m * p + 255) / 256), 256, 0, s>>>(d_mxp, (int)m, (int)p, d_Csq, nvar);
    if (d_Csq) launch_set_diag(s, d_Csq, nvar, nvar);
    if (l0 && p > 0) {
        const size_t pairs = m * p + p * (p - 1) / 2;
        level0_strip_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, s>>>(d_Csq, nvar, (int)m, (int)p, l0->th, l0->bits,
                                                                           l0->rw);
    }
    CUSK_CUDA(cudaGetLastError());
    if (st) {
        CUSK_CUDA(cudaStreamSynchronize(s));
        st->counts_ms = t_counts.ms();
        st->decode_ms = t_decode.ms();
        st->gram_ms = t_gram.ms();
        st->mxp_ms = t_mxp.ms();
        st->pxp_ms = t_pxp.ms();
    }
    return CUSK_OK;
}

extern "C" int cusk_corr_block_dev(cusk_ctx *ctx, const unsigned char *d_bed, const float *d_phen, size_t m, size_t n,
                                   size_t p, const float *d_mean, const float *d_std, int mode, float *d_Csq,
                                   cusk_corr_stats *stats)
{
    if (!ctx || !d_bed || !d_mean || !d_std || !d_Csq || (p > 0 && !d_phen)) {
        cusk_set_error("cusk_corr_block_dev: NULL argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    arena_reset(ctx);
    if (stats) memset(stats, 0, sizeof(*stats));
    cudaEvent_t e0 = ctx->ev[20], e1 = ctx->ev[21];
    cudaEventRecord(e0, ctx->stream);
    int rc = corr_block_device(ctx, d_bed, d_phen, m, n, p, d_mean, d_std, mode, d_Csq, nullptr, nullptr, nullptr, true,
                               stats, false, nullptr);
    if (rc) return rc;
    cudaEventRecord(e1, ctx->stream);
    CUSK_CUDA(cudaStreamSynchronize(ctx->stream));
    if (stats) cudaEventElapsedTime(&stats->total_ms, e0, e1);
    return CUSK_OK;
}

extern "C" int cusk_corr_block(cusk_ctx *ctx, const unsigned char *bed, const float *phen, size_t m, size_t n, size_t p,
                               const float *mean, const float *stdv, int mode, float *mxm, float *mxp, float *pxp,
                               int64_t *tables, cusk_corr_stats *stats)
{
    if (!ctx || !bed || !mean || !stdv || (p > 0 && !phen)) {
        cusk_set_error("cusk_corr_block: NULL argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    arena_reset(ctx);
    arena_reset(ctx->io);
    cusk_corr_stats local;
    cusk_corr_stats *st = stats ? stats : &local;
    memset(st, 0, sizeof(*st));
    const size_t col_bytes = (n + 3) / 4;
    const size_t npairs = m * (m - (m > 0)) / 2, ppairs = p * (p - (p > 0)) / 2;
    const int nvar = (int)(m + p);
    const bool want_mxm = mxm != nullptr || tables != nullptr;
    unsigned char *d_bed = (unsigned char *)arena_take(ctx->io, m * col_bytes);
    float *d_phen = p ? (float *)arena_take(ctx->io, sizeof(float) * p * n) : nullptr;
    float *d_mean = (float *)arena_take(ctx->io, sizeof(float) * m);
    float *d_std = (float *)arena_take(ctx->io, sizeof(float) * m);
    float *d_Csq = want_mxm ? (float *)arena_take(ctx->io, sizeof(float) * (size_t)nvar * nvar) : nullptr;
    float *d_mxp = (p && mxp) ? (float *)arena_take(ctx->io, sizeof(float) * m * p) : nullptr;
    float *d_pxp = (pxp && ppairs) ? (float *)arena_take(ctx->io, sizeof(float) * ppairs) : nullptr;
    float *d_upper = (mxm && npairs) ? (float *)arena_take(ctx->io, sizeof(float) * npairs) : nullptr;
    long long *d_tables = (tables && npairs) ? (long long *)arena_take(ctx->io, sizeof(long long) * 9 * npairs) : nullptr;
    if (!d_bed || !d_mean || !d_std || (p && !d_phen) || (want_mxm && !d_Csq)) return CUSK_E_CUDA;
    cudaEvent_t e0 = ctx->ev[20], e1 = ctx->ev[21], e2 = ctx->ev[22], e3 = ctx->ev[23];
    cudaEventRecord(e0, s);
    CUSK_CUDA(cudaMemcpyAsync(d_bed, bed, m * col_bytes, cudaMemcpyHostToDevice, s));
    if (p) CUSK_CUDA(cudaMemcpyAsync(d_phen, phen, sizeof(float) * p * n, cudaMemcpyHostToDevice, s));
    CUSK_CUDA(cudaMemcpyAsync(d_mean, mean, sizeof(float) * m, cudaMemcpyHostToDevice, s));
    CUSK_CUDA(cudaMemcpyAsync(d_std, stdv, sizeof(float) * m, cudaMemcpyHostToDevice, s));
    cudaEventRecord(e1, s);
    int rc = corr_block_device(ctx, d_bed, d_phen, m, n, p, d_mean, d_std, mode, d_Csq, d_mxp, d_pxp, d_tables, want_mxm,
                               st, false, nullptr);
    if (rc) return rc;
    cudaEventRecord(e2, s);
    if (d_upper) {
        launch_pack_upper(s, d_Csq, (int)m, nvar, d_upper);
        CUSK_CUDA(cudaMemcpyAsync(mxm, d_upper, sizeof(float) * npairs, cudaMemcpyDeviceToHost, s));
    }
    if (d_mxp) CUSK_CUDA(cudaMemcpyAsync(mxp, d_mxp, sizeof(float) * m * p, cudaMemcpyDeviceToHost, s));
    if (d_pxp) CUSK_CUDA(cudaMemcpyAsync(pxp, d_pxp, sizeof(float) * ppairs, cudaMemcpyDeviceToHost, s));
    if (d_tables)
        CUSK_CUDA(cudaMemcpyAsync(tables, d_tables, sizeof(long long) * 9 * npairs, cudaMemcpyDeviceToHost, s));
    cudaEventRecord(e3, s);
    CUSK_CUDA(cudaStreamSynchronize(s));
    cudaEventElapsedTime(&st->h2d_ms, e0, e1);
    cudaEventElapsedTime(&st->d2h_ms, e2, e3);
    cudaEventElapsedTime(&st->total_ms, e0, e3);
    return CUSK_OK;
}

extern "C" int cusk_marker_phen_corr(cusk_ctx *ctx, const unsigned char *bed, const float *phen, size_t m, size_t n,
                                     size_t p, const float *mean, const float *stdv, float *mxp,
                                     cusk_corr_stats *stats)
{
    return cusk_corr_block(ctx, bed, phen, m, n, p, mean, stdv, 0, nullptr, mxp, nullptr, nullptr, stats);
}

extern "C" int cusk_bed_marker_stats(cusk_ctx *ctx, const unsigned char *bed, size_t m, size_t n, float *means,
                                     float *stds, int *modes, int64_t *counts4)
{
    if (!ctx || !bed || !means || !stds) {
        cusk_set_error("cusk_bed_marker_stats: NULL argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    arena_reset(ctx->io);
    const size_t col_bytes = (n + 3) / 4;
    unsigned char *d_bed = (unsigned char *)arena_take(ctx->io, m * col_bytes);
    int *d_counts = (int *)arena_take(ctx->io, sizeof(int) * 4 * m);
    float *d_means = (float *)arena_take(ctx->io, sizeof(float) * m);
    float *d_stds = (float *)arena_take(ctx->io, sizeof(float) * m);
    int *d_modes = (int *)arena_take(ctx->io, sizeof(int) * m);
    if (!d_bed || !d_counts || !d_means || !d_stds || !d_modes) return CUSK_E_CUDA;
    CUSK_CUDA(cudaMemcpyAsync(d_bed, bed, m * col_bytes, cudaMemcpyHostToDevice, s));
    launch_bed_counts(s, d_bed, (int)m, (int)n, col_bytes, d_counts);
    launch_bed_stats(s, d_bed, (int)m, (int)n, col_bytes, d_counts, d_means, d_stds, d_modes);
    CUSK_CUDA(cudaGetLastError());
    std::vector<int> hc(4 * m), hm(m);
    CUSK_CUDA(cudaMemcpyAsync(means, d_means, sizeof(float) * m, cudaMemcpyDeviceToHost, s));
    CUSK_CUDA(cudaMemcpyAsync(stds, d_stds, sizeof(float) * m, cudaMemcpyDeviceToHost, s));
    CUSK_CUDA(cudaMemcpyAsync(hm.data(), d_modes, sizeof(int) * m, cudaMemcpyDeviceToHost, s));
    CUSK_CUDA(cudaMemcpyAsync(hc.data(), d_counts, sizeof(int) * 4 * m, cudaMemcpyDeviceToHost, s));
    CUSK_CUDA(cudaStreamSynchronize(s));
    if (modes) std::copy(hm.begin(), hm.end(), modes);
    if (counts4)
        for (size_t i = 0; i < 4 * m; i++) counts4[i] = hc[i];
    return CUSK_OK;
}

// ---- the reference's own names (src/corr_host.cu:1023, :1094) ---------------------------------------
extern "C" void cu_marker_phen_corr_pearson(const unsigned char *marker_vals, const float *phen_vals,
                                            const size_t num_markers, const size_t num_individuals,
                                            const size_t num_phen, const float *marker_mean, const float *marker_std,
                                            float *marker_phen_corrs)
{
    if (cusk_marker_phen_corr(cusk_default_ctx(), marker_vals, phen_vals, num_markers, num_individuals, num_phen,
                              marker_mean, marker_std, marker_phen_corrs, nullptr) != CUSK_OK) {
        fprintf(stderr, "cusk_b200 cu_marker_phen_corr_pearson: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);
    }
}

extern "C" void cu_corr_pearson_npn(const unsigned char *marker_vals, const float *phen_vals, const size_t num_markers,
                                    const size_t num_individuals, const size_t num_phen, const float *marker_mean,
                                    const float *marker_std, float *marker_corrs, float *marker_phen_corrs,
                                    float *phen_corrs)
{
    if (cusk_corr_block(cusk_default_ctx(), marker_vals, phen_vals, num_markers, num_individuals, num_phen, marker_mean,
                        marker_std, 0, marker_corrs, marker_phen_corrs, phen_corrs, nullptr, nullptr) != CUSK_OK) {
        fprintf(stderr, "cusk_b200 cu_corr_pearson_npn: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);
    }
}
