// skeleton_kernels.cu -- level-0 tests, adjacency compaction, level planning, separating-set
// resolution and output expansion kernels of the two skeleton engines (sm_100a).
//
// Device data layout:
//   C      float [n][n]     correlations, symmetric, row-major (as the reference passes them)
//   bits   u32   [n][rw]    adjacency bitmap, rw = ceil(n/32); bit (y&31) of word y>>5 of row x
//   deg    int   [n]        row degrees at level start
//   rowptr int   [n+1]      CSR offsets of the compacted adjacency (the reference's GPrime)
//   adj    int   [nnz]      sorted neighbour ids per row
//   minkey u64   [nnz]      individual-level engine: lowest separating-set rank found per slot
#include <algorithm>

#include "skeleton.cuh"

__device__ __forceinline__ double fisher_z(double rho)
{
    return fabs(0.5 * log(fabs((1.0 + rho) / (1.0 - rho))));
}

// ---------------------------------------------------------------------------------------------
// Level 0 (src/cuPC-S.cu:458-484, src/hetcor-cuPC-S.cu:343-377).  One warp produces one 32-bit
// adjacency word with a ballot: lane c of the warp tests pair (row, 32*word + c).
// Algorithmic bytes: 4 n^2 (C) [+ 4 n^2 N, + 4 n^2 G_in for the het engine] read, n^2/8 written.
// ---------------------------------------------------------------------------------------------
template <bool HET>
__global__ void __launch_bounds__(256) level0_kernel(const float *__restrict__ C,
                                                     const float *__restrict__ N, float n_uniform,
                                                     const int *__restrict__ G_in, int n, int rw,
                                                     double th, uint32_t *__restrict__ bits,
                                                     float *__restrict__ pMax)
{
    const int lane = threadIdx.x & 31;
    const int word = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int r = blockIdx.y;
    if (word >= rw) return;
    const int c = word * 32 + lane;
    bool keep = false;
    if (c < n && c != r) {
        // the reference evaluates the pair once, from the upper-triangle entry
        const size_t ix = (r < c) ? (size_t)r * n + c : (size_t)c * n + r;
        const double z = fisher_z((double)C[ix]);
        if (HET) {
            const bool cand = G_in ? (G_in[(size_t)r * n + c] == 1) : true;
            const double nn = N ? (double)N[ix] : (double)n_uniform;
            const double loc_th = th / sqrt(nn - 3.0);
            keep = cand && !(z < loc_th);
        } else {
            keep = !(z < th);
            if (pMax) pMax[(size_t)r * n + c] = keep ? 0.f : (float)z;
        }
    } else if (!HET && pMax && c == r && c < n) {
        pMax[(size_t)r * n + c] = 0.f;
    }
    const uint32_t w = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) bits[(size_t)r * rw + word] = w;
}

// ---------------------------------------------------------------------------------------------
// Level 0, tiled over the upper triangle (no candidate matrix).  One warp owns a 32 x 32 tile
// (rows R0.., columns C0.., R0 <= C0): lane c reads C[R0+i][C0+c] for i = 0..31 (coalesced 128 B
// rows), the ballot of row i is the adjacency word (R0+i, C0/32) and lane c collects its own
// decisions into the mirrored word (C0+c, R0/32), so every pair is read and tested once and the
// lower triangle is never touched: 2 n^2 bytes of C [+ 2 n^2 of N] read, n^2/8 written.
// Het engine: the decision is taken in FP32 when z and the threshold differ by more than 1e-3
// relative (FP32 evaluation error is < 1e-6), otherwise - and for any non-finite operand - by the
// FP64 expression of level0_kernel, so the result is that of the FP64 expression everywhere.
// ---------------------------------------------------------------------------------------------
template <bool HET>
__global__ void __launch_bounds__(256) level0_tile_kernel(const float *__restrict__ C,
                                                          const float *__restrict__ N, float n_uniform, int n,
                                                          int rw, double th, uint32_t *__restrict__ bits,
                                                          float *__restrict__ pMax)
{
    __shared__ float s_z[HET ? 1 : 8][HET ? 1 : 32][HET ? 1 : 33];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int tr = blockIdx.y;
    const int tc = blockIdx.x * 8 + wid;
    if (tc < tr || tc >= rw) return;
    const int R0 = tr * 32, C0 = tc * 32;
    const int c = C0 + lane;
    const bool diag = (tr == tc);
    const float thf = (float)th;
    const float t_uni = HET ? thf * rsqrtf(n_uniform - 3.0f) : 0.f;
    const double th_uni = HET ? th / sqrt((double)n_uniform - 3.0) : th;
    uint32_t my_row_word = 0;  // word (R0 + lane, tc)
    uint32_t my_col_word = 0;  // word (C0 + lane, tr)
#pragma unroll 8
    for (int i = 0; i < 32; i++) {
        const int r = R0 + i;
        bool keep = false;
        float zf_out = 0.f;
        if (r < n && c < n && r < c) {
            const size_t ix = (size_t)r * n + c;
            const float rho = __ldg(&C[ix]);
            if (HET) {
                const float nn = N ? __ldg(&N[ix]) : n_uniform;
                const float zf = fabsf(0.5f * logf(fabsf((1.0f + rho) / (1.0f - rho))));
                const float tf = N ? thf * rsqrtf(nn - 3.0f) : t_uni;
                if (fabsf(zf - tf) > 1e-3f * tf + 1e-6f) {
                    keep = !(zf < tf);
                } else {
                    const double z = fisher_z((double)rho);
                    const double loc_th = N ? th / sqrt((double)nn - 3.0) : th_uni;
                    keep = !(z < loc_th);
                }
            } else {
                const double z = fisher_z((double)rho);
                keep = !(z < th);
                zf_out = keep ? 0.f : (float)z;
            }
        }
        if (!HET && pMax) {
            if (r < n && c < n && (r < c || (diag && r == c))) pMax[(size_t)r * n + c] = zf_out;
            s_z[wid][i][lane] = zf_out;
        }
        const uint32_t w = __ballot_sync(0xffffffffu, keep);
        if (lane == i) my_row_word = w;
        my_col_word |= (keep ? 1u : 0u) << i;
    }
    if (diag) {
        // both triangles of the diagonal tile come from its upper entries
        if (R0 + lane < n) bits[(size_t)(R0 + lane) * rw + tc] = my_row_word | my_col_word;
    } else {
        if (R0 + lane < n) bits[(size_t)(R0 + lane) * rw + tc] = my_row_word;
        if (c < n) bits[(size_t)c * rw + tr] = my_col_word;
    }
    if (!HET && pMax) {
        __syncwarp();
        // mirrored pMax entries, written row-wise: row C0 + j, columns R0 + lane
        for (int j = 0; j < 32; j++) {
            const int rr = C0 + j, cc = R0 + lane;
            if (rr < n && cc < n && cc < rr) pMax[(size_t)rr * n + cc] = s_z[wid][lane][j];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Level plan: a single CTA scans the degrees into CSR offsets, decides whether the level runs
// (src/cuPC-S.cu:154-159: a level l runs iff max_degree - 1 >= l), sizes the work items of the
// CI kernel and computes the level's decision constant.
//   th_level >= 0 : fixed threshold of the level (individual-level engine)
//   th_level <  0 : het engine; uniform N -> threshold th_het / sqrt(trunc(N) - l - 3)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 block_scan_excl_u64(u64 v, u64 *s_warp, u64 *total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    u64 inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        const u64 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        u64 w = (lane < (int)(blockDim.x >> 5)) ? s_warp[lane] : 0;
        u64 winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            const u64 t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        s_warp[lane] = winc - w;
        if (lane == 31) s_warp[32] = winc;
    }
    __syncthreads();
    *total = s_warp[32];
    return s_warp[wid] + inc - v;
}

// two exclusive block scans behind one set of barriers (s_warp: 2 x 33 entries)
__device__ __forceinline__ void block_scan_excl_2u64(u64 v0, u64 v1, u64 *s_warp, u64 *ex0, u64 *ex1, u64 *tot0,
                                                     u64 *tot1)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    u64 i0 = v0, i1 = v1;
    for (int o = 1; o < 32; o <<= 1) {
        const u64 t0 = __shfl_up_sync(0xffffffffu, i0, o);
        const u64 t1 = __shfl_up_sync(0xffffffffu, i1, o);
        if (lane >= o) {
            i0 += t0;
            i1 += t1;
        }
    }
    __syncthreads();
    if (lane == 31) {
        s_warp[wid] = i0;
        s_warp[33 + wid] = i1;
    }
    __syncthreads();
    if (wid < 2) {
        u64 *sw = s_warp + 33 * wid;
        const u64 w = (lane < (int)(blockDim.x >> 5)) ? sw[lane] : 0;
        u64 winc = w;
        for (int o = 1; o < 32; o <<= 1) {
            const u64 t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        sw[lane] = winc - w;
        if (lane == 31) sw[32] = winc;
    }
    __syncthreads();
    *tot0 = s_warp[32];
    *tot1 = s_warp[33 + 32];
    *ex0 = s_warp[wid] + i0 - v0;
    *ex1 = s_warp[33 + wid] + i1 - v1;
}

// ceil(a / b) with the 32-bit divider when both operands fit (they nearly always do; a 64-bit division is
// ~4x the instructions, and the plan phase below is one CTA on the critical path of every level)
// C(n, k) for n < 64, k < 16 from a table (Pascal's rule at compile time): binom64's chain of 64-bit divisions on one lane
// per row was a visible part of the per-level planning time; larger arguments fall back to it
struct BinomTab {
    u64 v[64 * 16];
};
constexpr BinomTab make_binom_tab()
{
    BinomTab t{};
    for (int n = 0; n < 64; n++)
        for (int k = 0; k < 16; k++)
            t.v[n * 16 + k] = (k > n) ? 0 : ((k == 0 || k == n) ? 1 : t.v[(n - 1) * 16 + k - 1] + t.v[(n - 1) * 16 + k]);
    return t;
}
__device__ const BinomTab g_binom_tab = make_binom_tab();
__device__ __forceinline__ u64 binom_dev(int n, int k)
{
    if ((unsigned)n < 64u && (unsigned)k < 16u) return g_binom_tab.v[n * 16 + k];
    return binom64(n, k);
}

__device__ __forceinline__ u64 ceil_div_u64(u64 a, u64 b)
{
    if (a == 0) return 0;
    if (((a | b) >> 32) == 0) return (u64)(((unsigned)a - 1u) / (unsigned)b + 1u);
    return (a - 1) / b + 1;
}

// One row of the compaction and its work items (a warp): shared by compact_kernel and the fused tail of
// degree_plan_kernel.  The plan's arrays are read through L2 (another CTA wrote them in the fused kernel).
__device__ __forceinline__ void compact_row(int r, int lane, const uint32_t *__restrict__ bits, int rw,
                                            const int *rowptr, int *__restrict__ adj, u64 *__restrict__ minkey,
                                            const int *deg, const int *itemstart, const u64 *chunk, int l,
                                            CiItem *__restrict__ items, CiItem *__restrict__ items_small,
                                            CiItem *__restrict__ items_mid, int item_cap)
{
    const int row0 = __ldcg(&rowptr[r]);
    // the row's work items (row, degree, CSR offset, rank range) into the table of its CI kernel
    {
        const u64 chx = __ldcg(&chunk[r]);
        if (chx != 0) {
            const int dd = __ldcg(&deg[r]);
            const u64 nc = binom_dev(dd, l);
            const u64 cnt = ceil_div_u64(nc, chx);
            const int is = __ldcg(&itemstart[r]);
            const int cls = is >> 24;
            CiItem *table = (cls == 1) ? items_small : ((cls == 2) ? items_mid : items);
            const u64 off = (u64)(is & 0xffffff);
            for (u64 k = lane; k < cnt; k += 32) {
                if (off + k >= (u64)item_cap) break;
                const u64 r0 = k * chx;
                CiItem it;
                it.X = r;
                it.d = dd;
                it.row0 = row0;
                it.pad = 0;
                it.R0 = r0;
                it.R1 = (r0 + chx < nc) ? r0 + chx : nc;
                table[off + k] = it;
            }
        }
    }
    int out = row0;
    for (int w0 = 0; w0 < rw; w0 += 32) {
        const int w = w0 + lane;
        uint32_t word = (w < rw) ? bits[(size_t)r * rw + w] : 0u;
        const int cnt = __popc(word);
        int inc = cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        int pos = out + inc - cnt;
        while (word) {
            const int b = __ffs(word) - 1;
            word &= word - 1;
            adj[pos] = w * 32 + b;
            if (minkey) minkey[pos] = CUSK_KEY_NONE;
            pos++;
        }
        out += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// Degrees and level plan in one launch.  Row phase (every CTA, a warp per row): popcount of the bitmap
// row, C(d, l) and the row's share of the level totals.  The CTA that finishes last then plans the level:
// CSR offsets, whether the level runs (src/cuPC-S.cu:154-159: iff max_degree - 1 >= l), the work items of
// the two CI kernels and the level's decision constant.
__global__ void __launch_bounds__(1024) degree_plan_kernel(const uint32_t *__restrict__ bits, int n, int rw,
                                                           int *__restrict__ deg, int l, int *__restrict__ rowptr,
                                                           int *__restrict__ itemstart, u64 *__restrict__ chunk,
                                                           LevelCtl *ctl, const LevelCtl *prev, double th_level,
                                                           double th_het, float n_uniform, int *work_counter,
                                                           int *work_counter_small, int *work_counter_mid, int world,
                                                           int ipsm_shallow, int ipsm_deep, int ipsm_small, int ipsm_mid,
                                                           int dfs_span, int dfs_dmin, int dfs_lnarrow, int ipsm_dfs, int sm_count,
                                                           u64 shard_min_tests, int fuse_compact, int *__restrict__ adj,
                                                           u64 *__restrict__ minkey, CiItem *__restrict__ items,
                                                           CiItem *__restrict__ items_small, CiItem *__restrict__ items_mid,
                                                           int item_cap)
{
    __shared__ u64 s_warp2[66];
    __shared__ u64 s_acc[4];
    __shared__ int s_maxdeg, s_infeasible, s_last;
    __shared__ u64 s_target, s_target_small, s_target_mid;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    pdl_enter();
    if (prev && prev->done) {
        if (blockIdx.x == 0 && tid == 0) {
            ctl->done = 1;
            ctl->nitems = 0;
            ctl->nitems_small = 0;
            ctl->nitems_mid = 0;
        }
        return;
    }
    if (tid == 0) {
        s_acc[0] = s_acc[1] = s_acc[2] = s_acc[3] = 0;
        s_maxdeg = 0;
        s_infeasible = 0;
    }
    __syncthreads();
    // rows of at most this degree go to the thread-per-set kernel (ci_small.cu), the others to ci_level.cu
    const int dsmall = (ipsm_small > 0 && l <= CI_SMALL_LMAX) ? CI_SMALL_DMAX : 0;
    // ... and at levels 1-2 the rows up to this degree to ci_mid.cu
    const int dmid = (ipsm_mid > 0 && l <= CI_MID_LMAX) ? (l == 1 ? CI_MID_DMAX1 : CI_MID_DMAX) : 0;
    // ---- row phase -----------------------------------------------------------------------------------
    {
        const int r = blockIdx.x * 32 + wid;
        if (r < n) {
            int c = 0;
            for (int w = lane; w < rw; w += 32) c += __popc(bits[(size_t)r * rw + w]);
            for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
            if (lane == 0) {
                const int d = c;
                deg[r] = d;
                u64 cb = 0;
                if (d > l) {
                    cb = binom_dev(d, l);
                    const u64 canon = (cb > ((u64)1 << 40) ? ((u64)1 << 40) : cb) * (u64)(d - l);
                    atomicAdd((unsigned long long *)&s_acc[2], (unsigned long long)canon);
                    if (cb >= CI_MAX_COMB)
                        atomicExch(&s_infeasible, 1);
                    else {
                        const int rc = ci_row_class(d, l, dsmall, dmid, dfs_span, dfs_dmin, dfs_lnarrow);
                        atomicAdd((unsigned long long *)&s_acc[rc == 1 ? 1 : (rc == 2 ? 3 : 0)],
                                  (unsigned long long)(cb * (u64)(d - l)));
                    }
                }
                chunk[r] = cb;  // C(d, l) for the plan phase, which replaces it by the row's chunk size
                atomicMax(&s_maxdeg, d);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (s_acc[0]) atomicAdd((unsigned long long *)&ctl->acc_tests, (unsigned long long)s_acc[0]);
        if (s_acc[1]) atomicAdd((unsigned long long *)&ctl->acc_small, (unsigned long long)s_acc[1]);
        if (s_acc[2]) atomicAdd((unsigned long long *)&ctl->acc_canon, (unsigned long long)s_acc[2]);
        if (s_acc[3]) atomicAdd((unsigned long long *)&ctl->acc_mid, (unsigned long long)s_acc[3]);
        atomicMax(&ctl->acc_maxdeg, s_maxdeg);
        if (s_infeasible) atomicExch(&ctl->acc_infeasible, 1);
        __threadfence();
        s_last = (atomicAdd(&ctl->ticket, 1u) == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) {
        if (!fuse_compact) return;
        // Fused compaction (the host only asks for it when the whole grid is resident at once): wait for the
        // planning CTA, then compact this CTA's 32 rows -- one launch per level instead of two.
        if (tid == 0) {
            while (atomicAdd(&ctl->plan_ready, 0) == 0) __nanosleep(200);
            __threadfence();
        }
        __syncthreads();
        if (*(volatile int *)&ctl->done) return;
        const int r = blockIdx.x * 32 + wid;
        if (r < n)
            compact_row(r, lane, bits, rw, rowptr, adj, minkey, deg, itemstart, chunk, l, items, items_small, items_mid,
                        item_cap);
        return;
    }
    __threadfence();
    // ---- plan phase (one CTA) ----------------------------------------------------------------------------
    const int maxdeg = *(volatile int *)&ctl->acc_maxdeg;
    const int infeasible = *(volatile int *)&ctl->acc_infeasible;
    const u64 tot_tests = *(volatile u64 *)&ctl->acc_tests;
    const u64 tot_small = *(volatile u64 *)&ctl->acc_small;
    const u64 tot_canon = *(volatile u64 *)&ctl->acc_canon;
    const u64 tot_mid = *(volatile u64 *)&ctl->acc_mid;
    const bool runs = (maxdeg - 1 >= l) && !infeasible;
    if (tid == 0) {
        // Items per SM of ci_level_kernel: the first two levels are throughput-bound with very uneven rows
        // and want many items for balance; from level 3 on an item's cost is dominated by the per-prefix
        // factorisation, which consecutive conditioning sets share, so few large items win (sweep on the
        // C2 block: profiles/r1_item_floor_sweep.txt).
        const u64 per_sm = (u64)(l <= 2 ? ipsm_shallow : ipsm_deep);
        u64 target = tot_tests / ((u64)sm_count * per_sm * (u64)world);
        if (target < 32) target = 32;
        s_target = target;
        // ci_small_kernel: a warp per item, a lane per rank, so many items of >= 32 ranks
        u64 ts = tot_small / ((u64)sm_count * (u64)(ipsm_small > 0 ? ipsm_small : 1) * (u64)world);
        if (ts < 32) ts = 32;
        s_target_small = ts;
        // ci_mid_kernel (levels 1-2) / ci_dfs_kernel (levels >= 3): a CTA per item
        const int ipsm2 = (l <= CI_MID_LMAX) ? ipsm_mid : ipsm_dfs;
        u64 tm = tot_mid / ((u64)sm_count * (u64)(ipsm2 > 0 ? ipsm2 : 1) * (u64)world);
        if (tm < 32) tm = 32;
        s_target_mid = tm;
    }
    __syncthreads();
    // One pass over the rows: CSR offsets (scan of the degrees) and work items.  A row's C(d,l) conditioning
    // sets are cut into items of `chunk` ranks.  The three kernels have separate item tables; itemstart[r] =
    // class * 2^24 + offset in the class's table (class 0: ci_level_kernel, 1: ci_small_kernel, 2:
    // ci_mid_kernel).  The next block of rows is loaded before the current one is scanned.
    const u64 target = s_target, target_small = s_target_small, target_mid = s_target_mid;
    u64 carry = 0;        // directed edges so far
    u64 carry_items = 0;  // three 21-bit fields: items of ci_level_kernel, ci_small_kernel, ci_mid_kernel
    int d_next = (tid < n) ? __ldcg(&deg[tid]) : 0;
    u64 c_next = (tid < n) ? __ldcg(&chunk[tid]) : 0;
    for (int base = 0; base < n; base += blockDim.x) {
        const int r = base + tid;
        const int d = d_next;
        const u64 c = c_next;
        {
            const int rn = r + (int)blockDim.x;
            d_next = (rn < n) ? __ldcg(&deg[rn]) : 0;
            c_next = (rn < n) ? __ldcg(&chunk[rn]) : 0;
        }
        u64 items = 0;
        const int cls = ci_row_class(d, l, dsmall, dmid, dfs_span, dfs_dmin, dfs_lnarrow);
        if (runs && r < n && d > l) {
            const u64 tests = c * (u64)(d - l);
            u64 tgt = (cls == 1) ? target_small : ((cls == 2) ? target_mid : target);
            // an item of ci_level_kernel re-stages the row's d x d submatrix: keep at least ~4 tests per
            // staged entry (rows too large to stage gather from L2 per test and have no such cost; rows
            // with d <= 64 stage in a couple of round trips whatever the item size)
            const u64 stage_cost = (cls == 0 && d > 64 && d <= 200) ? 4 * (u64)d * (u64)d : 0;
            if (tgt < stage_cost) tgt = stage_cost;
            items = ceil_div_u64(tests, tgt);
            if (items > c) items = c;
            if (items < 1) items = 1;
            u64 ch = ceil_div_u64(c, items);
            if (cls == 1 && ch < 32) ch = (c < 32) ? c : 32;  // a lane per rank
            if (cls == 2 && l <= CI_MID_LMAX) {
                // a thread per (rank, 32-neighbour chunk): at least a CTA's worth of them per item
                const u64 minr = (u64)((256 + ((d + 31) >> 5) - 1) / ((d + 31) >> 5));
                if (ch < minr) ch = (c < minr) ? c : minr;
            }
            if (cls == 2 && l > CI_MID_LMAX) {
                // ci_dfs stages (d + 1)^2 correlations per item and gives each of its 4 warps a rank sub-range:
                // at least ~4 tests per staged entry
                const u64 minr = ceil_div_u64(4 * (u64)(d + 1) * (u64)(d + 1), (u64)(d - l));
                if (ch < minr) ch = (c < minr) ? c : minr;
            }
            items = ceil_div_u64(c, ch);
            chunk[r] = ch;
        } else if (runs && r < n) {
            chunk[r] = 0;
        }
        u64 ex_d, ex_i, tot_d, tot_i;
        block_scan_excl_2u64((u64)d, items << (21 * cls), s_warp2, &ex_d, &ex_i, &tot_d, &tot_i);
        if (r < n) {
            rowptr[r] = (int)(carry + ex_d);
            const u64 at = carry_items + ex_i;
            itemstart[r] = (cls << 24) | (int)((at >> (21 * cls)) & 0x1fffffu);
        }
        carry += tot_d;
        carry_items += tot_i;
    }
    if (tid == 0) {
        rowptr[n] = (int)carry;
        ctl->maxdeg = maxdeg;
        ctl->nnz = carry;
        ctl->canon = (maxdeg - 1 >= l) ? tot_canon : 0;
        ctl->executed = 0;
        ctl->singular = 0;
        ctl->infeasible = infeasible;
        double th = th_level;
        if (th_level < 0) {
            const double m = (double)__float2int_rz(n_uniform);
            th = th_het / sqrt(m - (double)l - 3.0);
        }
        const double t = tanh(th);
        ctl->T2 = t * t;
        *work_counter = 0;
        *work_counter_small = 0;
        *work_counter_mid = 0;
        ctl->nitems = runs ? (int)(carry_items & 0x1fffffu) : 0;
        ctl->nitems_small = runs ? (int)((carry_items >> 21) & 0x1fffffu) : 0;
        ctl->nitems_mid = runs ? (int)((carry_items >> 42) & 0x1fffffu) : 0;
        // a level with few tests is cheaper run whole by every rank (same result everywhere) than dealt out
        ctl->shard_on = (runs && world > 1 && tot_tests + tot_small + tot_mid >= shard_min_tests) ? 1 : 0;
        ctl->done = runs ? 0 : 1;
    }
    if (!fuse_compact) return;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        atomicExch(&ctl->plan_ready, 1);
    }
    if (!runs) return;
    {
        const int r = blockIdx.x * 32 + wid;
        if (r < n)
            compact_row(r, lane, bits, rw, rowptr, adj, minkey, deg, itemstart, chunk, l, items, items_small, items_mid,
                        item_cap);
    }
}

// ---------------------------------------------------------------------------------------------
// Compaction (replaces scan_compact, src/cuPC-S.cu:6355-6432) and work-item decoding: one warp per row; each lane
// popcounts one bitmap word, a warp prefix sum gives its output offset, set bits are expanded
// to ascending neighbour ids.  Bytes: n^2/8 read + 4 nnz written (+ 8 nnz for minkey).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) compact_kernel(const uint32_t *__restrict__ bits, int n, int rw,
                                                      const int *__restrict__ rowptr, int *__restrict__ adj,
                                                      u64 *__restrict__ minkey, const LevelCtl *ctl,
                                                      const int *__restrict__ deg, const int *__restrict__ itemstart,
                                                      const u64 *__restrict__ chunk, int l, CiItem *__restrict__ items,
                                                      CiItem *__restrict__ items_small, CiItem *__restrict__ items_mid,
                                                      int item_cap)
{
    pdl_enter();
    if (ctl->done) return;
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= n) return;
    compact_row(r, lane, bits, rw, rowptr, adj, minkey, deg, itemstart, chunk, l, items, items_small, items_mid, item_cap);
}

// ---------------------------------------------------------------------------------------------
// Individual-level engine, end of a level: every slot with a recorded rank becomes a separating
// set (lowest lexicographic rank of row X; the deterministic reading of src/cuPC-S.cu:568-577),
// its Z statistic is recomputed directly and both adjacency bits are cleared.
// ---------------------------------------------------------------------------------------------
__device__ double partial_corr_z_direct(const float *__restrict__ C, int n, int X, int Y, const int *S, int l)
{
    // incremental LDL^T over S, tracking only the later pivots and the two targets x, y
    double u[CUSK_ML][CUSK_ML + 2];  // u[j][t]: t < l -> S[t], t = l -> X, t = l+1 -> Y
    double dinv[CUSK_ML];
    for (int j = 0; j < l; j++) {
        double D = 1.0;
        for (int i = 0; i < j; i++) D -= u[i][j] * u[i][j] * dinv[i];
        const bool drop = fabs(D) < CI_PIVOT_TOL;
        for (int t = 0; t < l + 2; t++) {
            const int v = (t < l) ? S[t] : (t == l ? X : Y);
            double acc = (t == j) ? 1.0 : (double)C[(size_t)S[j] * n + v];
            for (int i = 0; i < j; i++) acc -= u[i][j] * u[i][t] * dinv[i];
            u[j][t] = drop ? 0.0 : acc;
        }
        dinv[j] = drop ? 0.0 : 1.0 / D;
    }
    double h00 = 1.0, h11 = 1.0, h01 = (double)C[(size_t)X * n + Y];
    for (int j = 0; j < l; j++) {
        h00 -= u[j][l] * u[j][l] * dinv[j];
        h11 -= u[j][l + 1] * u[j][l + 1] * dinv[j];
        h01 -= u[j][l] * u[j][l + 1] * dinv[j];
    }
    const double rho = h01 / sqrt(fabs(h00 * h11));
    return fisher_z(rho);
}

__global__ void __launch_bounds__(128) resolve_kernel(const float *__restrict__ C, int n, int rw, int l,
                                                      uint32_t *bits, const int *__restrict__ deg,
                                                      const int *__restrict__ rowptr,
                                                      const int *__restrict__ adj,
                                                      const u64 *__restrict__ minkey, float *pMax,
                                                      cusk_sepset *seps, unsigned long long *sep_count,
                                                      u64 sep_cap, const LevelCtl *ctl)
{
    pdl_enter();
    if (ctl->done) return;
    const int lane = threadIdx.x & 31;
    const int X = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (X >= n) return;
    const int d = deg[X];
    if (d <= l) return;
    const int *a = adj + rowptr[X];
    for (int yp = lane; yp < d; yp += 32) {
        const u64 key = minkey[rowptr[X] + yp];
        if (key == CUSK_KEY_NONE) continue;
        int pos[CUSK_ML], S[CUSK_ML];
        unrank_comb_fast(key, d, l, pos);
        for (int i = 0; i < l; i++) S[i] = a[pos[i]];
        const int Y = a[yp];
        const double z = partial_corr_z_direct(C, n, X, Y, S, l);
        if (pMax) pMax[(size_t)X * n + Y] = (float)z;
        atomicAnd(&bits[(size_t)X * rw + (Y >> 5)], ~(1u << (Y & 31)));
        atomicAnd(&bits[(size_t)Y * rw + (X >> 5)], ~(1u << (X & 31)));
        if (seps) {
            const u64 slot = atomicAdd(sep_count, 1ULL);
            if (slot < sep_cap) {
                cusk_sepset rec;
                rec.x = X;
                rec.y = Y;
                rec.level = l;
                rec.pmax = (float)z;
                for (int i = 0; i < CUSK_ML; i++) rec.s[i] = (i < l) ? S[i] : -1;
                seps[slot] = rec;
            }
        }
    }
}

// pMax post-processing of src/cuPC-S.cu:424-442: diagonal 1, kept edges -100000, removed edges
// the larger of the two directed entries.
__global__ void __launch_bounds__(256) pmax_final_kernel(const uint32_t *__restrict__ bits, int n, int rw,
                                                         float *pMax)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= n || j < i) return;
    if (i == j) {
        pMax[(size_t)i * n + i] = 1.f;
        return;
    }
    const bool alive = (bits[(size_t)i * rw + (j >> 5)] >> (j & 31)) & 1u;
    float v;
    if (alive)
        v = -100000.f;
    else
        v = fmaxf(pMax[(size_t)j * n + i], pMax[(size_t)i * n + j]);
    pMax[(size_t)j * n + i] = v;
    pMax[(size_t)i * n + j] = v;
}

// bitmap -> dense int adjacency (the reference's G[n*n] output)
__global__ void __launch_bounds__(256) expand_bits_kernel(const uint32_t *__restrict__ bits, int n, int rw,
                                                          int *__restrict__ G)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    if (c >= n) return;
    G[(size_t)r * n + c] = (bits[(size_t)r * rw + (c >> 5)] >> (c & 31)) & 1u;
}

// dense int adjacency -> bitmap (G == 1 is an edge, src/cuPC-S.cu:6379)
__global__ void __launch_bounds__(256) pack_bits_kernel(const int *__restrict__ G, int n, int rw,
                                                        uint32_t *__restrict__ bits)
{
    const int lane = threadIdx.x & 31;
    const int word = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int r = blockIdx.y;
    if (word >= rw) return;
    const int c = word * 32 + lane;
    const bool e = (c < n) && (G[(size_t)r * n + c] == 1);
    const uint32_t w = __ballot_sync(0xffffffffu, e);
    if (lane == 0) bits[(size_t)r * rw + word] = w;
}

// reduce trunc(N[i,j]) over i != j to (min, max): uniform sample sizes take the fast path
__global__ void __launch_bounds__(256) ess_minmax_kernel(const float *__restrict__ N, int n, int *mn, int *mx)
{
    // min / max of the float bit patterns (ordered like the values for f >= 0); a NaN or a
    // negative entry poisons the result so that the general path is taken.  Uniform <=> mn == mx.
    int lo = 2147483647, hi = -2147483647 - 1;
    for (int r = blockIdx.x; r < n; r += gridDim.x) {
        const float *row = N + (size_t)r * n;
        for (int c = threadIdx.x; c < n; c += blockDim.x) {
            if (c == r) continue;
            const float f = __ldg(row + c);
            if (!(f >= 0.f)) {
                lo = -2147483647 - 1;
                hi = 2147483647;
            } else {
                const int v = __float_as_int(f);
                lo = min(lo, v);
                hi = max(hi, v);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(mn, lo);
        atomicMax(mx, hi);
    }
}

// ---- host-callable launch wrappers (defined here so that skeleton.cu stays host-only logic) ---
void launch_level0(cudaStream_t st, bool het, const float *C, const float *N, float n_uniform, const int *G_in,
                   int n, int rw, double th, uint32_t *bits, float *pMax)
{
    if (!G_in) {
        dim3 tg((rw + 7) / 8, rw);
        if (het)
            level0_tile_kernel<true><<<tg, 256, 0, st>>>(C, N, n_uniform, n, rw, th, bits, pMax);
        else
            level0_tile_kernel<false><<<tg, 256, 0, st>>>(C, N, n_uniform, n, rw, th, bits, pMax);
        return;
    }
    dim3 grid((rw + 7) / 8, n);
    if (het)
        level0_kernel<true><<<grid, 256, 0, st>>>(C, N, n_uniform, G_in, n, rw, th, bits, pMax);
    else
        level0_kernel<false><<<grid, 256, 0, st>>>(C, N, n_uniform, G_in, n, rw, th, bits, pMax);
}

// ci_dfs.cu routing knobs, read on every call (tests switch them between calls of one process):
//   CUSK_DFS=0 switches the kernel off; for rows of at most CI_SMALL_DMAX neighbours: CUSK_DFS_NARROW_LEVEL, the first
//   level at which they may leave ci_small (default 8), CUSK_DFS_MIN_SPAN: least d - l (default 8), CUSK_DFS_MIN_D:
//   least degree (default 20); CUSK_ITEMS_PER_SM_DFS: work items per SM the plan aims at (default 48)
DfsTuning dfs_tuning()
{
    DfsTuning t;
    const char *off = getenv("CUSK_DFS");
    const char *sp = getenv("CUSK_DFS_MIN_SPAN");
    const char *dm = getenv("CUSK_DFS_MIN_D");
    const char *ip = getenv("CUSK_ITEMS_PER_SM_DFS");
    const char *ln = getenv("CUSK_DFS_NARROW_LEVEL");
    t.span = (off && atoi(off) == 0) ? 0 : (sp && atoi(sp) > 0 ? atoi(sp) : 8);
    t.dmin = (dm && atoi(dm) > 0) ? atoi(dm) : 20;
    t.lnarrow = (ln && atoi(ln) > 0) ? atoi(ln) : 8;
    const int v = ip ? atoi(ip) : 0;
    t.ipsm = v > 0 ? (v > CI_SMALL_IPSM_MAX ? CI_SMALL_IPSM_MAX : v) : 48;
    return t;
}

namespace {
struct PlanKnobs {
    int ipsm_shallow, ipsm_deep, ipsm_small, ipsm_mid;
};
int env_pos(const char *name, int dflt)
{
    const char *e = getenv(name);
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : dflt;
}
// tuning knobs, read once: work items per SM of ci_level_kernel at levels 1-2 (SHALLOW) and from level 3 on (DEEP),
// of the thread-per-set kernel (ci_small.cu; CUSK_SMALL_KERNEL=0 switches it off) and of the level-1/2 kernel for
// medium rows (ci_mid.cu; CUSK_MID_KERNEL=0 switches it off)
const PlanKnobs &plan_knobs()
{
    static const PlanKnobs k = [] {
        PlanKnobs q;
        q.ipsm_shallow = env_pos("CUSK_ITEMS_PER_SM_SHALLOW", 8);
        q.ipsm_deep = env_pos("CUSK_ITEMS_PER_SM_DEEP", 2);
        const char *off = getenv("CUSK_SMALL_KERNEL");
        q.ipsm_small = (off && atoi(off) == 0) ? 0 : std::min(env_pos("CUSK_ITEMS_PER_SM_SMALL", 32), CI_SMALL_IPSM_MAX);
        off = getenv("CUSK_MID_KERNEL");
        q.ipsm_mid = (off && atoi(off) == 0) ? 0 : std::min(env_pos("CUSK_ITEMS_PER_SM_MID", 4), CI_SMALL_IPSM_MAX);
        return q;
    }();
    return k;
}
}  // namespace

// Host-side mirror of the plan's routing: which CI kernels can get a row at level l when no row has more than
// maxdeg neighbours (bit c = class c of ci_row_class; bit 8 = a class-2 row with 32 or more neighbours, i.e.
// ci_dfs_kernel<NS = 2>).  Degrees only fall from level to level, so the maximum read at one level bounds the later ones.
unsigned ci_classes_possible(int l, int maxdeg)
{
    const PlanKnobs &k = plan_knobs();
    const DfsTuning dt = dfs_tuning();
    const int dsmall = (k.ipsm_small > 0 && l <= CI_SMALL_LMAX) ? CI_SMALL_DMAX : 0;
    const int dmid = (k.ipsm_mid > 0 && l <= CI_MID_LMAX) ? (l == 1 ? CI_MID_DMAX1 : CI_MID_DMAX) : 0;
    unsigned mask = 0;
    const int top = std::max(CI_DFS_DMAX, std::max(dsmall, dmid)) + 1;  // every larger degree is class 0
    if (maxdeg > top) mask |= 1u;
    for (int d = l + 1; d <= std::min(maxdeg, top); d++) {
        const int c = ci_row_class(d, l, dsmall, dmid, dt.span, dt.dmin, dt.lnarrow);
        mask |= 1u << c;
        if (c == 2 && l > CI_MID_LMAX && d >= 32) mask |= 1u << 8;
    }
    return mask;
}

// The level's degrees + plan, and (fused = true) its compaction too, as one launch.  The fused form makes CTAs wait
// for the planning CTA, so it is a cooperative launch (the driver only starts it when the whole grid is resident: two
// host slots driving the same GPU cannot wedge each other) and only offered when the grid fits the device at once.
bool pdl_enabled()
{
    static const bool on = [] {
        const char *e = getenv("CUSK_PDL");
        return !(e && atoi(e) == 0);  // on unless CUSK_PDL=0
    }();
    return on;
}

bool plan_compact_fusable(int n, int sm_count)
{
    static int occ = -1;
    static int coop = -1;
    if (occ < 0) {
        const char *e = getenv("CUSK_FUSE_PLAN");
        int dev = 0, o = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        // a cooperative launch cannot be a programmatic dependent: with PDL on (the default) two plain launches are
        // faster than the fused one (C2 step 0.84 against 0.89 ms), so the fused form is the default only without PDL
        if (e ? atoi(e) == 0 : pdl_enabled()) coop = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, degree_plan_kernel, 1024, 0);
        occ = o;
    }
    return coop > 0 && occ > 0 && (n + 31) / 32 <= occ * (sm_count > 0 ? sm_count : 148);
}

cudaError_t launch_degree_plan(cudaStream_t st, const uint32_t *bits, int n, int rw, int *deg, int l, int *rowptr,
                               int *itemstart, u64 *chunk, LevelCtl *ctl, const LevelCtl *prev, double th_level,
                               double th_het, float n_uniform, int *work_counter, int *work_counter_small,
                               int *work_counter_mid, int world, int sm_count, u64 shard_min_tests, bool fused, int *adj,
                               u64 *minkey, CiItem *items, CiItem *items_small, CiItem *items_mid, int item_cap)
{
    const PlanKnobs &k = plan_knobs();
    const DfsTuning dt = dfs_tuning();
    int ipsm_shallow = k.ipsm_shallow, ipsm_deep = k.ipsm_deep, ipsm_small = k.ipsm_small, ipsm_mid = k.ipsm_mid;
    int span = dt.span, dmin = dt.dmin, lnarrow = dt.lnarrow, ipsm_dfs = dt.ipsm;
    int smc = sm_count > 0 ? sm_count : 148, fuse = fused ? 1 : 0;
    if (!fused) {
        return launch_k(degree_plan_kernel, dim3((n + 31) / 32), dim3(1024), 0, st, bits, n, rw, deg, l, rowptr, itemstart, chunk, ctl,
                        prev, th_level, th_het, n_uniform, work_counter, work_counter_small, work_counter_mid, world,
                        ipsm_shallow, ipsm_deep, ipsm_small, ipsm_mid, span, dmin, lnarrow, ipsm_dfs, smc, shard_min_tests, 0,
                        (int *)nullptr, (u64 *)nullptr, (CiItem *)nullptr, (CiItem *)nullptr, (CiItem *)nullptr, 0);
    }
    void *args[] = {&bits, &n, &rw, &deg, &l, &rowptr, &itemstart, &chunk, &ctl, &prev, &th_level, &th_het, &n_uniform,
                    &work_counter, &work_counter_small, &work_counter_mid, &world, &ipsm_shallow, &ipsm_deep, &ipsm_small,
                    &ipsm_mid, &span, &dmin, &lnarrow, &ipsm_dfs, &smc, &shard_min_tests, &fuse, &adj, &minkey, &items,
                    &items_small, &items_mid, &item_cap};
    return cudaLaunchCooperativeKernel((const void *)degree_plan_kernel, dim3((n + 31) / 32), dim3(1024), args, 0, st);
}

void launch_compact(cudaStream_t st, const uint32_t *bits, int n, int rw, const int *rowptr, int *adj,
                    u64 *minkey, const LevelCtl *ctl, const int *deg, const int *itemstart, const u64 *chunk, int l,
                    CiItem *items, CiItem *items_small, CiItem *items_mid, int item_cap)
{
    launch_k(compact_kernel, dim3((n + 7) / 8), dim3(256), 0, st, bits, n, rw, rowptr, adj, minkey, ctl, deg, itemstart, chunk, l,
             items, items_small, items_mid, item_cap);
}

void launch_resolve(cudaStream_t st, const float *C, int n, int rw, int l, uint32_t *bits, const int *deg,
                    const int *rowptr, const int *adj, const u64 *minkey, float *pMax, cusk_sepset *seps,
                    unsigned long long *sep_count, u64 sep_cap, const LevelCtl *ctl)
{
    launch_k(resolve_kernel, dim3((n + 3) / 4), dim3(128), 0, st, C, n, rw, l, bits, deg, rowptr, adj, minkey, pMax, seps, sep_count,
             sep_cap, ctl);
}

void launch_pmax_final(cudaStream_t st, const uint32_t *bits, int n, int rw, float *pMax)
{
    dim3 grid((n + 255) / 256, n);
    pmax_final_kernel<<<grid, 256, 0, st>>>(bits, n, rw, pMax);
}

void launch_expand_bits(cudaStream_t st, const uint32_t *bits, int n, int rw, int *G)
{
    dim3 grid((n + 255) / 256, n);
    expand_bits_kernel<<<grid, 256, 0, st>>>(bits, n, rw, G);
}

void launch_pack_bits(cudaStream_t st, const int *G, int n, int rw, uint32_t *bits)
{
    dim3 grid((rw + 7) / 8, n);
    pack_bits_kernel<<<grid, 256, 0, st>>>(G, n, rw, bits);
}

void launch_ess_minmax(cudaStream_t st, const float *N, int n, int *mn, int *mx, int sm_count)
{
    ess_minmax_kernel<<<std::min(n, sm_count * 8), 256, 0, st>>>(N, n, mn, mx);
}

// edge-sharded mode: bits = AND over the ranks' bitmaps (each rank only ever clears bits)
__global__ void __launch_bounds__(256) and_reduce_kernel(const uint32_t *__restrict__ gathered, int world,
                                                         size_t words, uint32_t *__restrict__ bits)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t v = 0xffffffffu;
        for (int r = 0; r < world; r++) v &= gathered[(size_t)r * words + i];
        bits[i] = v;
    }
}
// edge-sharded het engine: flags[r][slot] = 1 where rank r removed the edge of CSR slot `slot` of this level;
// one warp per row clears both adjacency bits of every flagged slot (idempotent on the removing rank).
__global__ void __launch_bounds__(256) apply_removed_kernel(const unsigned char *__restrict__ flags, int world,
                                                            size_t stride, int n, int rw,
                                                            const int *__restrict__ rowptr, const int *__restrict__ deg,
                                                            const int *__restrict__ adj, uint32_t *__restrict__ bits)
{
    const int lane = threadIdx.x & 31;
    const int X = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (X >= n) return;
    const int r0 = rowptr[X], d = deg[X];
    for (int a = lane; a < d; a += 32) {
        unsigned f = 0;
        for (int r = 0; r < world; r++) f |= flags[(size_t)r * stride + r0 + a];
        if (f) {
            const int Y = adj[r0 + a];
            atomicAnd(&bits[(size_t)X * rw + (Y >> 5)], ~(1u << (Y & 31)));
            atomicAnd(&bits[(size_t)Y * rw + (X >> 5)], ~(1u << (X & 31)));
        }
    }
}
void launch_apply_removed(cudaStream_t st, const unsigned char *flags, int world, size_t stride, int n, int rw,
                          const int *rowptr, const int *deg, const int *adj, uint32_t *bits)
{
    apply_removed_kernel<<<(n + 7) / 8, 256, 0, st>>>(flags, world, stride, n, rw, rowptr, deg, adj, bits);
}

void launch_and_reduce(cudaStream_t st, const uint32_t *gathered, int world, size_t words, uint32_t *bits)
{
    const int grid = (int)std::min<size_t>((words + 255) / 256, 148 * 8);
    and_reduce_kernel<<<grid, 256, 0, st>>>(gathered, world, words, bits);
}
