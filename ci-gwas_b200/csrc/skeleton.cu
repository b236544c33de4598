// skeleton.cu -- host drivers of the two skeleton engines: the level loop of Skeleton()
// (src/cuPC-S.cu:61-450) and hetcor_skeleton() (src/hetcor-cuPC-S.cu:75-341), re-designed so that
// the whole call is one stream of asynchronous launches with a single mid-call synchronisation
// (the directed edge count after level 0, which sizes the compacted adjacency).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <vector>

#include "skeleton.cuh"

// launch wrappers (skeleton_kernels.cu, ci_level.cu)
void launch_level0(cudaStream_t st, bool het, const float *C, const float *N, float n_uniform, const int *G_in,
                   int n, int rw, double th, uint32_t *bits, float *pMax);
cudaError_t launch_degree_plan(cudaStream_t st, const uint32_t *bits, int n, int rw, int *deg, int l, int *rowptr,
                               int *itemstart, u64 *chunk, LevelCtl *ctl, const LevelCtl *prev, double th_level,
                               double th_het, float n_uniform, int *work_counter, int *work_counter_small,
                               int *work_counter_mid, int world, int sm_count, u64 shard_min_tests, bool fused = false,
                               int *adj = nullptr, u64 *minkey = nullptr, CiItem *items = nullptr,
                               CiItem *items_small = nullptr, CiItem *items_mid = nullptr, int item_cap = 0);
bool plan_compact_fusable(int n, int sm_count);
unsigned ci_classes_possible(int l, int maxdeg);
void launch_compact(cudaStream_t st, const uint32_t *bits, int n, int rw, const int *rowptr, int *adj,
                    u64 *minkey, const LevelCtl *ctl, const int *deg, const int *itemstart, const u64 *chunk, int l,
                    CiItem *items, CiItem *items_small, CiItem *items_mid, int item_cap);
void launch_resolve(cudaStream_t st, const float *C, int n, int rw, int l, uint32_t *bits, const int *deg,
                    const int *rowptr, const int *adj, const u64 *minkey, float *pMax, cusk_sepset *seps,
                    unsigned long long *sep_count, u64 sep_cap, const LevelCtl *ctl);
void launch_pmax_final(cudaStream_t st, const uint32_t *bits, int n, int rw, float *pMax);
void launch_expand_bits(cudaStream_t st, const uint32_t *bits, int n, int rw, int *G);
void launch_pack_bits(cudaStream_t st, const int *G, int n, int rw, uint32_t *bits);
void launch_ess_minmax(cudaStream_t st, const float *N, int n, int *mn, int *mx, int sm_count);
int launch_ci_level(cudaStream_t st, bool het, const CiParams &p, int grid);
int launch_ci_small_multi(cudaStream_t st, const CiParams &p, int l_first, int l_last, int maxl, bool coop, int *launches);
int launch_ci_small(cudaStream_t st, bool het, const CiParams &p);
int launch_ci_mid(cudaStream_t st, bool het, const CiParams &p);
int launch_ci_dfs(cudaStream_t st, bool het, const CiParams &p, bool wide);
void launch_and_reduce(cudaStream_t st, const uint32_t *gathered, int world, size_t words, uint32_t *bits);
void launch_apply_removed(cudaStream_t st, const unsigned char *flags, int world, size_t stride, int n, int rw,
                          const int *rowptr, const int *deg, const int *adj, uint32_t *bits);

extern "C" size_t cusk_row_words(int n) { return (size_t)(n + 31) / 32; }

namespace {

struct EngineArgs {
    bool het;
    const float *dC;
    int n;
    uint32_t *dbits;       // in/out
    const int *dG_in;      // het: dense candidate matrix on the device, or nullptr
    bool bits_are_input;   // het: dbits already holds the candidate edges
    bool all_ones;         // het: every pair is a candidate
    const float *dN;       // het: nullptr -> uniform n_uniform
    float n_uniform;
    const float *Th;       // non-het: 15 thresholds (host)
    float th_het;
    int maxlevel;
    const int *d_time;     // het
    float *d_pMax;         // non-het, may be nullptr
    bool want_seps;
    bool level0_done;      // non-het: dbits already holds the level-0 adjacency (fused into the correlation epilogue)
};

struct EngineOut {
    int last_level = 0;
    cusk_sepset *d_seps = nullptr;
    u64 sep_count = 0;
};

// Runs levels 0..maxlevel on device-resident inputs.  Scratch comes from ctx's arena (taken after
// whatever the caller already took).
int run_engine(cusk_ctx *ctx, const EngineArgs &a, EngineOut *out, cusk_stats *stats)
{
    const int n = a.n;
    const int rw = (int)cusk_row_words(n);
    cudaStream_t st = ctx->stream;
    int64_t nl = 0;  // kernels launched by this call (counted where they are launched, reported in the statistics)
    const int maxl = std::min(a.maxlevel, CUSK_ML);
    if (n < 1 || a.maxlevel < 0) {
        cusk_set_error("skeleton: invalid n=%d maxlevel=%d", n, a.maxlevel);
        return CUSK_E_ARG;
    }

    int *d_deg = (int *)arena_take(ctx, sizeof(int) * (size_t)n);
    int *d_rowptr = (int *)arena_take(ctx, sizeof(int) * ((size_t)n + 1));
    int *d_itemstart = (int *)arena_take(ctx, sizeof(int) * ((size_t)n + 1));
    u64 *d_chunk = (u64 *)arena_take(ctx, sizeof(u64) * (size_t)n);
    LevelCtl *d_ctl = (LevelCtl *)arena_take(ctx, sizeof(LevelCtl) * (CUSK_ML + 2));
    int *d_work = (int *)arena_take(ctx, sizeof(int) * 4 * (CUSK_ML + 2));  // queues of the CI kernels, one per level
    int *d_work_small = d_work ? d_work + (CUSK_ML + 2) : nullptr;
    int *d_work_mid = d_work ? d_work + 2 * (CUSK_ML + 2) : nullptr;
    int *d_work_dfs2 = d_work ? d_work + 3 * (CUSK_ML + 2) : nullptr;
    unsigned long long *d_sepcount = (unsigned long long *)arena_take(ctx, sizeof(unsigned long long));
    int *d_minmax = (int *)arena_take(ctx, 2 * sizeof(int));
    if (!d_deg || !d_rowptr || !d_itemstart || !d_chunk || !d_ctl || !d_work || !d_sepcount || !d_minmax)
        return CUSK_E_CUDA;
    CUSK_CUDA(cudaMemsetAsync(d_ctl, 0, sizeof(LevelCtl) * (CUSK_ML + 2), st));
    CUSK_CUDA(cudaMemsetAsync(d_work, 0, sizeof(int) * 4 * (CUSK_ML + 2), st));
    CUSK_CUDA(cudaMemsetAsync(d_sepcount, 0, sizeof(unsigned long long), st));

    // het engine: uniform effective sample sizes take the constant-threshold fast path
    const float *dN = a.dN;
    float n_uniform = a.n_uniform;
    if (a.het && dN) {
        int init[2] = {2147483647, -2147483647 - 1};
        CUSK_CUDA(cudaMemcpyAsync(d_minmax, init, sizeof(init), cudaMemcpyHostToDevice, st));
        nl++, launch_ess_minmax(st, dN, n, d_minmax, d_minmax + 1, ctx->sm_count);
        int mm[2];
        CUSK_CUDA(cudaMemcpyAsync(mm, d_minmax, sizeof(mm), cudaMemcpyDeviceToHost, st));
        CUSK_CUDA(cudaStreamSynchronize(st));
        if (mm[0] == mm[1] && n > 1) {
            memcpy(&n_uniform, &mm[0], sizeof(float));
            dN = nullptr;
        }
    }

    // edge-sharded mode: every rank holds the whole problem and takes every world-th work item
    ShardLink *sh = (ctx->shard && ctx->shard->world > 1) ? ctx->shard : nullptr;
    // het engine: what a rank removed in a level travels as one byte per slot of the level's CSR
    // (nnz_l bytes per rank instead of the n * rw word bitmap), all-gathered and applied by every rank
    unsigned char *d_removed = nullptr, *d_gather = nullptr;
    size_t flag_stride = 0, nnz1 = 0;
    int exchanged_levels = 0;
    static const u64 shard_min_tests = [] {
        const char *e = getenv("CUSK_SHARD_MIN_TESTS");
        const long long v = e ? atoll(e) : -1;
        return (u64)(v >= 0 ? v : 4000000);
    }();
    if (sh) {
        sh->exchange_ms = 0.0;
        sh->exchange_bytes = 0;
    }
    // plan_kernel aims at up to sm_count * CI_SMALL_IPSM_MAX * world items per level, plus at most one short item per row
    const int item_cap = n + ctx->sm_count * CI_SMALL_IPSM_MAX * ((sh ? sh->world : 1) + 1);
    CiItem *d_items = (CiItem *)arena_take(ctx, sizeof(CiItem) * (size_t)item_cap);
    CiItem *d_items_small = (CiItem *)arena_take(ctx, sizeof(CiItem) * (size_t)item_cap);
    CiItem *d_items_mid = (CiItem *)arena_take(ctx, sizeof(CiItem) * (size_t)item_cap);
    if (!d_items || !d_items_small || !d_items_mid) return CUSK_E_CUDA;

    // self-planned deep levels (ci_small.cu, multi-level kernel): het engine on one GPU, skeletons of at most CI_SELF_NMAX
    // variables whose rows are all short by level CI_DFS_LMIN
    const char *side_e = getenv("CUSK_SIDE_STREAM");  // level 2: ci_small and ci_mid on two streams (0: one after the other)
    const bool side_env = !(side_e && atoi(side_e) == 0);
    const char *coop_e = getenv("CUSK_SELF_COOP");  // 0: the multi-level launches as plain (programmatic dependent) launches
    const bool self_coop = !(coop_e && atoi(coop_e) == 0);
    const char *self_e = getenv("CUSK_SELF_PLAN");  // read on every call (tests compare the two routes)
    const bool self_env = !(self_e && atoi(self_e) == 0);
    const bool self_ok = self_env && a.het && !sh && n <= CI_SELF_NMAX && maxl > CI_DFS_LMIN;
    // items of a self-planned level: at most one partial item per row plus (sets / chunk) <= lanes / 8 full ones
    const int self_cap = n + ctx->sm_count * 64 + 64;
    int *d_self_adj = nullptr;
    CiItem *d_self_items = nullptr;
    if (self_ok) {
        d_self_adj = (int *)arena_take(ctx, sizeof(int) * (size_t)n * CI_SELF_STRIDE);
        d_self_items = (CiItem *)arena_take(ctx, sizeof(CiItem) * (size_t)self_cap);
        if (!d_self_adj || !d_self_items) return CUSK_E_CUDA;
    }
    int self_first = 0, self_last = -1;  // levels that ran inside the multi-level launches

    cudaEvent_t *ev = ctx->ev;
    // per-level and per-kernel-group events are instrumentation (cusk_stats.level_ms / ci_kernel_ms): three records per
    // level sit between back-to-back launches, so a caller that only wants the result can switch them off
    const bool timing = ctx->level_timing;
    CUSK_CUDA(cudaEventRecord(ev[0], st));
    // ---- level 0 -----------------------------------------------------------------------------
    if (a.het && a.bits_are_input) {
        // candidates given as a bitmap: expand to the level-0 kernel's input via a dense pass
        // (only used by the device-resident API; the bitmap is both input and output)
        int *dG_tmp = (int *)arena_take(ctx, sizeof(int) * (size_t)n * n);
        if (!dG_tmp) return CUSK_E_CUDA;
        nl++, launch_expand_bits(st, a.dbits, n, rw, dG_tmp);
        nl++, launch_level0(st, true, a.dC, dN, n_uniform, dG_tmp, n, rw, (double)a.th_het, a.dbits, nullptr);
    } else if (a.het) {
        nl++, launch_level0(st, true, a.dC, dN, n_uniform, a.all_ones ? nullptr : a.dG_in, n, rw, (double)a.th_het,
                      a.dbits, nullptr);
    } else if (!a.level0_done) {
        nl++, launch_level0(st, false, a.dC, nullptr, 0.f, nullptr, n, rw, (double)a.Th[0], a.dbits, a.d_pMax);
    }
    CUSK_CUDA(cudaGetLastError());

    int *d_adj = nullptr;
    u64 *d_minkey = nullptr;
    cusk_sepset *d_seps = nullptr;
    u64 sep_cap = 0;
    const int grid = ctx->sm_count;
    const int smem_bytes = ctx->max_smem_optin - 1024;  // leave room for static shared memory

    // Upper bound of every row's degree from the level it was read at (degrees only fall): decides which CI kernels a
    // level can need at all, and ends the loop at the first level no row can reach (src/cuPC-S.cu:154-159).
    int maxdeg_bound = 2147483647;
    int l_stop = maxl + 1;  // first level whose kernels were not launched
    const bool fusable = plan_compact_fusable(n, ctx->sm_count);
    for (int l = 1; l <= maxl; l++) {
        if (maxdeg_bound - 1 < l) {
            l_stop = l;
            break;
        }
        if (timing) CUSK_CUDA(cudaEventRecord(ev[l], st));
        LevelCtl *ctl = d_ctl + l;
        const LevelCtl *prev = (l > 1) ? d_ctl + (l - 1) : nullptr;
        // from level 2 on (the CSR arrays exist) the plan and the compaction are one launch when the grid fits the device
        // deep levels of a light block: no separate plan / compaction / launch per level, the thread-per-set kernel plans
        // its levels itself (degrees only fall, so the bound read at level 3 holds for the rest of the call)
        if (self_ok && l > CI_DFS_LMIN && maxdeg_bound <= CI_SELF_DMAX) {
            const int l_last = std::min(maxl, maxdeg_bound - 1);
            CiParams p = {};
            p.C = a.dC;
            p.N = dN;
            p.time_index = a.d_time;
            p.n = n;
            p.l = l;
            p.rw = rw;
            p.bits = a.dbits;
            p.ctl = d_ctl;                       // bases of the per-level arrays
            p.work_counter_small = d_work_small;
            p.th_het = (double)a.th_het;
            p.smem_bytes = smem_bytes;
            p.shard_world = 1;
            p.sm_count = ctx->sm_count;
            p.self_adj = d_self_adj;
            p.self_items = d_self_items;
            p.item_cap = self_cap;
            p.deg_out = d_deg;
            p.n_uniform = n_uniform;
            {
                const char *e = getenv("CUSK_WARP_SETS");
                p.warp_sets = !(e && atoi(e) == 0);
            }
            int launched = 0;
            const int rc = launch_ci_small_multi(st, p, l, l_last, maxl, self_coop, &launched);
            nl += launched;
            if (rc) return rc;
            CUSK_CUDA(cudaGetLastError());
            self_first = l;
            self_last = l_last;
            l_stop = l_last + 1;
            break;
        }
        const bool fused = l >= 2 && fusable;
        nl++;
        CUSK_CUDA(launch_degree_plan(st, a.dbits, n, rw, d_deg, l, d_rowptr, d_itemstart, d_chunk, ctl, prev,
                                     a.het ? -1.0 : (double)a.Th[l], (double)a.th_het, n_uniform, d_work + l,
                                     d_work_small + l, d_work_mid + l, sh ? sh->world : 1, ctx->sm_count, shard_min_tests,
                                     fused, d_adj, d_minkey, d_items, d_items_small, d_items_mid, item_cap));
        if (l == 1) {
            // the only mid-call synchronisation: directed edge count after level 0
            LevelCtl h;
            CUSK_CUDA(cudaMemcpyAsync(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost, st));
            CUSK_CUDA(cudaStreamSynchronize(st));
            const size_t nnz = (size_t)h.nnz;
            nnz1 = nnz;
            maxdeg_bound = h.maxdeg;

            d_adj = (int *)arena_take(ctx, sizeof(int) * std::max<size_t>(nnz, 1));
            if (!d_adj) return CUSK_E_CUDA;
            if (!a.het) {
                d_minkey = (u64 *)arena_take(ctx, sizeof(u64) * std::max<size_t>(nnz, 1));
                if (!d_minkey) return CUSK_E_CUDA;
                if (a.want_seps) {
                    sep_cap = nnz;
                    d_seps = (cusk_sepset *)arena_take(ctx, sizeof(cusk_sepset) * std::max<size_t>(nnz, 1));
                    if (!d_seps) return CUSK_E_CUDA;
                }
            }
            if (sh && a.het) {
                flag_stride = (nnz + 15) & ~(size_t)15;
                d_removed = (unsigned char *)arena_take(ctx, std::max<size_t>(flag_stride, 16));
                d_gather = (unsigned char *)arena_take(ctx, std::max<size_t>(flag_stride, 16) * (size_t)sh->world);
                if (!d_removed || !d_gather) return CUSK_E_CUDA;
            }
        }
        if (sh && l == 1 && sh->host_barrier) sh->host_barrier(sh->user);  // every rank has done its allocations
        if (l == CI_DFS_LMIN && maxdeg_bound > CI_DFS_LMIN) {
            // one more look at the plan: after level 2 most rows are short, and the bound read here lets the remaining
            // levels launch only the kernels some row can need (and stop at the first level none can reach).  Every
            // rank of an edge-sharded run reads the same numbers, so the ranks keep launching the same sequence.
            LevelCtl h;
            CUSK_CUDA(cudaMemcpyAsync(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost, st));
            run_idle_job(ctx);  // levels 1-2 are in flight: the caller's host work hides behind them
            CUSK_CUDA(cudaStreamSynchronize(st));
            if (h.done) {
                l_stop = l;
                break;
            }
            maxdeg_bound = h.maxdeg;
        }
        // edge-sharded mode: whether a level is dealt out is decided on the device (LevelCtl.shard_on, the same on every
        // rank); the flags of the level start at zero and are exchanged after its kernels whatever the decision was, so
        // that the host never has to read the plan back: one fixed-size stream-ordered collective per level.
        if (sh && d_removed) CUSK_CUDA(cudaMemsetAsync(d_removed, 0, std::max<size_t>(flag_stride, 16), st));
        const unsigned classes = ci_classes_possible(l, maxdeg_bound);
        if (!fused)
            nl++, launch_compact(st, a.dbits, n, rw, d_rowptr, d_adj, d_minkey, ctl, d_deg, d_itemstart, d_chunk, l, d_items,
                                 d_items_small, d_items_mid, item_cap);
        CiParams p = {};
        p.C = a.dC;
        p.N = a.het ? dN : nullptr;
        p.time_index = a.het ? a.d_time : nullptr;
        p.n = n;
        p.l = l;
        p.rw = rw;
        p.bits = a.dbits;
        p.deg = d_deg;
        p.rowptr = d_rowptr;
        p.adj = d_adj;
        p.ctl = ctl;
        p.work_counter = d_work + l;
        p.work_counter_small = d_work_small + l;
        p.work_counter_mid = d_work_mid + l;
        p.work_counter_dfs2 = d_work_dfs2 + l;
        p.sm_count = ctx->sm_count;
        p.minkey = d_minkey;
        p.th_het = (double)a.th_het;
        p.smem_bytes = smem_bytes;
        p.items = d_items;
        p.items_small = d_items_small;
        p.items_mid = d_items_mid;
        p.item_cap = item_cap;
        p.removed = sh ? d_removed : nullptr;
        p.shard_rank = sh ? sh->rank : 0;  // used by the kernels only when the level's plan set shard_on
        p.shard_world = sh ? sh->world : 1;
        if (timing) CUSK_CUDA(cudaEventRecord(ev[(CUSK_ML + 2) + 2 * l], st));
        int rc = CUSK_OK;
        // level 2: the short rows' kernel and the medium rows' kernel side by side (each alone keeps a quarter to a third
        // of the warp slots busy, profiles/r2_ci_kernels_full_summary.txt); the medium rows go first, on the side stream
        const bool side = side_env && !sh && l == CI_MID_LMAX && (classes & 2u) && (classes & 4u) && ctx->stream2;
        if (side) {
            CUSK_CUDA(cudaEventRecord(ctx->ev_fork, st));
            CUSK_CUDA(cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0));
            rc = (nl++, launch_ci_mid(ctx->stream2, a.het, p));
            if (rc) return rc;
            CUSK_CUDA(cudaEventRecord(ctx->ev_join, ctx->stream2));
        }
        if (classes & 1u) rc = (nl++, launch_ci_level(st, a.het, p, grid));  // hub rows
        if (rc) return rc;
        if (classes & 2u) rc = (nl++, launch_ci_small(st, a.het, p));  // the level's short rows (thread per conditioning set)
        if (rc) return rc;
        if ((classes & 4u) && l <= CI_MID_LMAX && !side) rc = (nl++, launch_ci_mid(st, a.het, p));  // levels 1-2: the medium rows
        if (rc) return rc;
        if (side) CUSK_CUDA(cudaStreamWaitEvent(st, ctx->ev_join, 0));
        if ((classes & 4u) && l > CI_MID_LMAX) {
            // deeper levels: rows of up to 63 neighbours whose sets serve many tests (lanes over the neighbours);
            // narrow (<= 31) and wide rows share the item table, each kernel skips the other's
            rc = (nl++, launch_ci_dfs(st, a.het, p, false));
            if (!rc && (classes & 256u)) rc = (nl++, launch_ci_dfs(st, a.het, p, true));
            if (rc) return rc;
        }
        if (timing) CUSK_CUDA(cudaEventRecord(ev[(CUSK_ML + 2) + 2 * l + 1], st));
        if (sh) {
            // exchange step of the level: removed edges (het: one flag byte per slot of the level-1 CSR, all-gathered;
            // individual-level: minimum of the separating-set ranks), after which every rank holds the same graph
            // again.  Stream-ordered behind the level's kernels, no host synchronisation; timed with events.
            cudaEvent_t *xe = ev + 3 * (CUSK_ML + 2) + 4 + 2 * l;
            CUSK_CUDA(cudaEventRecord(xe[0], st));
            int xrc;
            if (a.het) {
                const size_t cnt = std::max<size_t>(flag_stride, 16);
                xrc = sh->allgather(sh->user, d_removed, d_gather, cnt, st);
                if (!xrc)
                    nl++, launch_apply_removed(st, d_gather, sh->world, cnt, n, rw, d_rowptr, d_deg, d_adj, a.dbits);
                sh->exchange_bytes += cnt * (size_t)sh->world;
            } else {
                xrc = sh->allreduce_min_u64(sh->user, d_minkey, std::max<size_t>(nnz1, 1), st);
                sh->exchange_bytes += nnz1 * 8;
            }
            if (xrc) {
                cusk_set_error("skeleton: exchange step of level %d failed (rc %d)", l, xrc);
                return CUSK_E_CUDA;
            }
            CUSK_CUDA(cudaEventRecord(xe[1], st));
            exchanged_levels = l;
        }
        if (!a.het)
            nl++, launch_resolve(st, a.dC, n, rw, l, a.dbits, d_deg, d_rowptr, d_adj, d_minkey, a.d_pMax, d_seps,
                           d_sepcount, sep_cap, ctl);
        CUSK_CUDA(cudaGetLastError());
    }
    // final degrees (edge count) and pMax post-processing
    nl++, launch_degree_plan(st, a.dbits, n, rw, d_deg, CUSK_ML + 1, d_rowptr, d_itemstart, d_chunk, d_ctl + CUSK_ML + 1,
                       nullptr, 1.0, 1.0, n_uniform, d_work + CUSK_ML + 1, d_work_small + CUSK_ML + 1,
                       d_work_mid + CUSK_ML + 1, 1, ctx->sm_count, shard_min_tests);
    if (!a.het && a.d_pMax) launch_pmax_final(st, a.dbits, n, rw, a.d_pMax);
    const int l_end = std::min(l_stop, maxl + 1);  // event slot that closes the last launched level
    CUSK_CUDA(cudaEventRecord(ev[l_end], st));
    CUSK_CUDA(cudaGetLastError());

    // ---- read back the control blocks --------------------------------------------------------
    LevelCtl *h_ctl = (LevelCtl *)ctx->h_pinned;
    unsigned long long *h_sepcount = (unsigned long long *)(h_ctl + CUSK_ML + 2);
    CUSK_CUDA(cudaMemcpyAsync(h_ctl, d_ctl, sizeof(LevelCtl) * (CUSK_ML + 2), cudaMemcpyDeviceToHost, st));
    CUSK_CUDA(cudaMemcpyAsync(h_sepcount, d_sepcount, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    run_idle_job(ctx);
    CUSK_CUDA(cudaStreamSynchronize(st));

    if (sh) {
        for (int l = 1; l <= exchanged_levels; l++) {
            float ms = 0.f;
            cudaEvent_t *xe = ev + 3 * (CUSK_ML + 2) + 4 + 2 * l;
            if (cudaEventElapsedTime(&ms, xe[0], xe[1]) == cudaSuccess) sh->exchange_ms += ms;
        }
    }
    if (self_last >= self_first && getenv("CUSK_SELF_TRACE")) {
        // phases of the levels inside the multi-level launches, microseconds (CTA 0's view): row phase + barrier, plan,
        // CTA 0's tests, wait for the other CTAs; and the gap to the next level's start (launch boundaries show up here)
        for (int l = self_first; l <= self_last && !h_ctl[l].done; l++) {
            const LevelCtl &c = h_ctl[l];
            const double gap = (l < self_last && !h_ctl[l + 1].done && h_ctl[l + 1].t_begin > c.t_end)
                                   ? (double)(h_ctl[l + 1].t_begin - c.t_end) * 1e-3 : 0.0;
            fprintf(stderr, "cusk self-trace: level %2d rows %6.2f plan %6.2f tests(cta0) %6.2f tail %6.2f | gap %6.2f | items %d\n", l,
                    (double)(c.t_rows - c.t_begin) * 1e-3, (double)(c.t_planned - c.t_rows) * 1e-3,
                    (double)(c.t_body - c.t_planned) * 1e-3, (double)(c.t_end - c.t_body) * 1e-3, gap, c.nitems_small);
        }
    }
    // last level as the reference reports it (src/cuPC-S.cu:99,154-159): the first level that
    // finds max_degree - 1 < l stops the loop with l - 1; otherwise the loop counter runs out.
    int last = maxl + 1;
    int levels_run = 1;
    int infeasible = 0;
    for (int l = 1; l <= maxl; l++) {
        if (l < l_stop && h_ctl[l].infeasible) infeasible = std::max(infeasible, h_ctl[l].infeasible);
        if (l >= l_stop || h_ctl[l].done) {
            last = l - 1;
            break;
        }
        levels_run++;
    }
    out->last_level = last;
    out->d_seps = d_seps;
    out->sep_count = std::min<u64>(*h_sepcount, sep_cap);
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->tests_canonical[0] = stats->tests_executed[0] = (int64_t)n * (n - 1) / 2;
        stats->edges_after[0] = (int64_t)(h_ctl[1].nnz / 2);
        for (int l = 1; l <= maxl; l++) {
            if (l >= l_stop || h_ctl[l].done) break;
            stats->tests_canonical[l] = (int64_t)h_ctl[l].canon;
            stats->tests_executed[l] = (int64_t)h_ctl[l].executed;
            stats->max_degree[l] = h_ctl[l].maxdeg;
            stats->singular_pivots += (int64_t)h_ctl[l].singular;
            // plan(l+1) ran (level l was not "done") and recorded nnz before deciding about l+1
            stats->edges_after[l] = (int64_t)(((l + 1 < l_stop && l < maxl) ? h_ctl[l + 1].nnz : h_ctl[CUSK_ML + 1].nnz) / 2);
        }
        if (maxl == 0) stats->edges_after[0] = (int64_t)(h_ctl[CUSK_ML + 1].nnz / 2);
        for (int l = 0; timing && l < l_end; l++) {
            float ms = 0.f;
            if (l >= self_first && l <= self_last) {
                // levels inside a multi-level launch: %globaltimer stamps of the kernel instead of events
                if (!h_ctl[l].done && h_ctl[l].t_end > h_ctl[l].t_begin) {
                    stats->level_ms[l] = (float)((double)(h_ctl[l].t_end - h_ctl[l].t_begin) * 1e-6);
                    stats->ci_kernel_ms[l] = (float)((double)(h_ctl[l].t_end - h_ctl[l].t_planned) * 1e-6);
                }
                continue;
            }
            if (cudaEventElapsedTime(&ms, ev[l], ev[l + 1]) == cudaSuccess) stats->level_ms[l] = ms;
            if (l >= 1 && !h_ctl[l].done &&
                cudaEventElapsedTime(&ms, ev[(CUSK_ML + 2) + 2 * l], ev[(CUSK_ML + 2) + 2 * l + 1]) == cudaSuccess)
                stats->ci_kernel_ms[l] = ms;
        }
        float tot = 0.f;
        if (cudaEventElapsedTime(&tot, ev[0], ev[l_end]) == cudaSuccess) stats->total_ms = tot;
        stats->last_level = last;
        stats->levels_run = levels_run;
        stats->kernel_launches = nl;
    }
    if (infeasible == 1) {
        cusk_set_error("skeleton: a row needs >= 2^48 conditioning sets at some level (degree too large)");
        return CUSK_E_INFEASIBLE;
    }
    if (infeasible > 2) {
        cusk_set_error("skeleton: internal bound violated at some level (code %d)", infeasible);
        return CUSK_E_INFEASIBLE;
    }
    if (infeasible == 2) {
        cusk_set_error("skeleton: a row's per-warp state does not fit in shared memory at some level");
        return CUSK_E_INFEASIBLE;
    }
    return CUSK_OK;
}

}  // namespace

// ---- engine entry points used by api.cu and the correlation -> skeleton pipeline ----------------
int skeleton_engine(cusk_ctx *ctx, const float *dC, int n, uint32_t *dbits, const float *Th, int maxlevel,
                    int *l_out, float *d_pMax, std::vector<cusk_sepset> *seps, cusk_stats *stats, bool level0_done)
{
    if (level0_done && d_pMax) {
        cusk_set_error("skeleton: a fused level 0 does not produce pMax");
        return CUSK_E_ARG;
    }
    arena_reset(ctx);
    EngineArgs a = {};
    a.het = false;
    a.dC = dC;
    a.n = n;
    a.dbits = dbits;
    a.Th = Th;
    a.maxlevel = maxlevel;
    a.d_pMax = d_pMax;
    a.want_seps = seps != nullptr;
    a.level0_done = level0_done;
    EngineOut eo;
    int rc = run_engine(ctx, a, &eo, stats);
    if (rc) return rc;
    if (l_out) *l_out = eo.last_level;
    if (seps) {
        seps->resize((size_t)eo.sep_count);
        if (eo.sep_count) {
            CUSK_CUDA(cudaMemcpyAsync(seps->data(), eo.d_seps, (size_t)eo.sep_count * sizeof(cusk_sepset),
                                      cudaMemcpyDeviceToHost, ctx->stream));
            CUSK_CUDA(cudaStreamSynchronize(ctx->stream));
            // the device appends in scheduling order; the contents are deterministic, the order is made so
            std::sort(seps->begin(), seps->end(), [](const cusk_sepset &p, const cusk_sepset &q) {
                return p.x != q.x ? p.x < q.x : p.y < q.y;
            });
        }
    }
    return CUSK_OK;
}

int hetcor_engine(cusk_ctx *ctx, const float *dC, int n, uint32_t *dbits, const int *dG_dense, bool bits_are_input,
                  const float *dN, float n_uniform, float th, int maxlevel, const int *d_time, int *l_out,
                  cusk_stats *stats)
{
    arena_reset(ctx);
    EngineArgs a = {};
    a.het = true;
    a.dC = dC;
    a.n = n;
    a.dbits = dbits;
    a.dG_in = dG_dense;
    a.bits_are_input = bits_are_input;
    a.all_ones = !bits_are_input && dG_dense == nullptr;
    a.dN = dN;
    a.n_uniform = n_uniform;
    a.th_het = th;
    a.maxlevel = maxlevel;
    a.d_time = d_time;
    EngineOut eo;
    int rc = run_engine(ctx, a, &eo, stats);
    if (rc) return rc;
    if (l_out) *l_out = eo.last_level;
    return CUSK_OK;
}

extern "C" int cusk_skeleton_dev(cusk_ctx *ctx, const float *dC, int n, uint32_t *dG_bits, const float *Th,
                                 int maxlevel, int *l_out, float *d_pMax, cusk_sepset *sep, int64_t sep_cap,
                                 int64_t *sep_n, cusk_stats *stats)
{
    if (!ctx || !dC || !dG_bits || !Th) {
        cusk_set_error("cusk_skeleton_dev: NULL argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    std::vector<cusk_sepset> seps;
    const bool want = sep != nullptr || sep_n != nullptr;
    int rc = skeleton_engine(ctx, dC, n, dG_bits, Th, maxlevel, l_out, d_pMax, want ? &seps : nullptr, stats, false);
    if (rc) return rc;
    if (sep_n) *sep_n = (int64_t)seps.size();
    if (sep && sep_cap > 0)
        memcpy(sep, seps.data(), std::min<size_t>(seps.size(), (size_t)sep_cap) * sizeof(cusk_sepset));
    return CUSK_OK;
}

extern "C" int cusk_hetcor_skeleton_dev(cusk_ctx *ctx, const float *dC, int n, uint32_t *dG_bits,
                                        int G_in_all_ones, const float *dN, float n_uniform, float th,
                                        int maxlevel, const int *d_time_index, int *l_out, cusk_stats *stats)
{
    if (!ctx || !dC || !dG_bits) {
        cusk_set_error("cusk_hetcor_skeleton_dev: NULL argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    return hetcor_engine(ctx, dC, n, dG_bits, nullptr, !G_in_all_ones, dN, n_uniform, th, maxlevel, d_time_index,
                         l_out, stats);
}
