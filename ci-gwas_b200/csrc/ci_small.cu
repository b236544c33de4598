// ci_small.cu -- level-l CI tests of the short rows (degree <= CS_DMAX) at levels l <= CI_SMALL_LMAX: one
// THREAD per conditioning set (replaces cal_Indepl1..14, src/cuPC-S.cu:486-3020, and the _ess
// variants, src/hetcor-cuPC-S.cu:379-3052, which are also thread-per-set; sm_100a).
//
// ci_level.cu gives a whole CTA to one row, shares one incremental factorisation between the sets of
// a warp and spreads the neighbours over the lanes: right for hub rows with hundreds of neighbours,
// but with d ~ 16-30 most lanes idle and every step waits on the previous one.  plan_kernel therefore
// routes every row by its degree; for the short ones the parallelism is ACROSS sets instead:
//   * a warp pulls a work item (row X, rank range) from the level's queue, stages the row's
//     neighbourhood submatrix C[adj(X), adj(X)] (<= 32 x 32 floats) in its own shared-memory slice,
//   * every lane takes a contiguous sub-range of the ranks, unranks its first set with a Pascal
//     table in shared memory and steps to the successor set afterwards,
//   * per set S the lane runs an LDL^T (unit lower L, signed pivots D) of C[S,S] in registers - the
//     kernel is compiled once per level, every loop unrolled (l > 10: L in shared memory) - carrying X
//     along (tx = L^-1 C[S,X]); per neighbour Y it forward-substitutes ty = L^-1 C[S,Y] and gets
//       h01 = Cxy - sum_j tx_j ty_j / D_j, h00 = 1 - sum_j tx_j^2 / D_j, h11 = 1 - sum_j ty_j^2 / D_j,
//     the same quantities, pivot rule (|D| < CI_PIVOT_TOL drops the direction) and decision
//     rho^2 < tanh(th)^2 [or > coth(th)^2] as ci_level.cu, in FP64.
// PC-stable semantics are those of ci_level.cu: sets come from the adjacency compacted at level
// start, removals clear bits (het engine) or lower the slot's minimum rank (individual-level engine).
//
// Algorithmic bytes per test: 4 (l + 2).
#include <atomic>

#include "skeleton.cuh"

#define CS_DMAX CI_SMALL_DMAX  // rows up to this degree (plan_kernel picks the kernel per level)
#define CS_LREG 10      // up to this level the factor L lives in registers; above, in shared memory
#define CS_NWARPS(LT) ((LT) <= CS_LREG ? 8 : 4)
#define CS_NTHREADS(LT) (CS_NWARPS(LT) * 32)
#define CS_STRIDE (CS_DMAX + 1)
#define FULLMASK 0xffffffffu
#define CS_WARP_SETS 1  // multi-level kernel: a level with at most this many sets per warp of the launch runs a warp per set (measured: 21 against 37 us at one set per warp, a tie at two)
static_assert(CI_SELF_DMAX < 32, "the warp-per-set form keeps X in lane d");

struct CsWarp {
    float csub[CS_DMAX * CS_STRIDE];
    float cx[CS_DMAX];
    int adj[CS_DMAX];
    int time[CS_DMAX];
    unsigned long long minkey[CS_DMAX];
    unsigned skipmask;  // het engine: bit y set = edge (X, adj[y]) already removed
    unsigned pad[3];
};

// factor storage: registers (static indices after unrolling) or shared memory, element e of thread tid at
// s[e * NT + tid] (conflict-free: the lanes of a warp run in lockstep over e)
template <bool REG, int NT>
struct CsL {
    __device__ __forceinline__ static double get(const double *r, const double *s, int e, int tid)
    {
        if constexpr (REG)
            return r[e];
        else
            return s[e * NT + tid];
    }
    __device__ __forceinline__ static void set(double *r, double *s, int e, int tid, double v)
    {
        if constexpr (REG)
            r[e] = v;
        else
            s[e * NT + tid] = v;
    }
};

template <int LT>
struct CsIdx {
    static constexpr int LM = LT;
    // W[j][a], j < a < LM, packed row by row
    __device__ __forceinline__ static constexpr int w(int j, int a) { return j * LM - (j * (j + 1)) / 2 + (a - j - 1); }
};

// ---- the level's tests: shared by the one-level kernel and by the self-planned multi-level kernel ---------------------------
// Everything another CTA may have written during the same launch (the multi-level kernel rebuilds adjacency and items per
// level) is read with ld.cg: L1 is not coherent across SMs.
//   STATIC1 (multi-level kernel): warp w of the launch takes item w first and only then pulls from the queue -- a light level
//   has about one item per warp, and a few thousand fetches of one counter at the same instant cost microseconds
template <bool HET, int LT, bool STATIC1 = false>
__device__ __forceinline__ void cs_body(const CiParams &p, LevelCtl *ctl, int *work_counter, const CiItem *items,
                                        const int *adj, int item_cap, int nitems, double T2u, int srank, int sworld,
                                        CsWarp *s_w, const unsigned *s_binom, double *s_L)
{
    constexpr int LM = CsIdx<LT>::LM;
    constexpr int NT = CS_NTHREADS(LT);
    constexpr bool LREG = (LT <= CS_LREG);
    using LS = CsL<LREG, NT>;
    constexpr int l = LT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = p.n;
    const bool het_general = HET && p.N != nullptr;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    CsWarp &S = s_w[warp];
    u64 my_tests = 0;
    unsigned my_sing = 0;
    __shared__ unsigned long long s_tot[2];  // tests and dropped pivots of the CTA (one global atomic each at the end)
    if (tid == 0) s_tot[0] = s_tot[1] = 0ull;
    __syncthreads();
    const int warps_total = (int)gridDim.x * (NT / 32);
    bool first = STATIC1;

    for (;;) {
        int item = 0;
        if (STATIC1 && first) {
            first = false;
            item = (int)blockIdx.x * (NT / 32) + warp;
        } else {
            if (STATIC1 && nitems <= warps_total) break;
            if (lane == 0) item = (STATIC1 ? warps_total : 0) + srank + sworld * atomicAdd(work_counter, 1);
            item = __shfl_sync(FULLMASK, item, 0);
        }
        if (item >= nitems) break;
        if (item >= item_cap) {  // cannot happen: plan_kernel bounds the item count of a small level
            if (lane == 0) ctl->infeasible = 3;
            break;
        }
        const int4 a4 = __ldcg((const int4 *)&items[item]);
        const ulonglong2 r2 = __ldcg((const ulonglong2 *)&items[item].R0);
        const int X = a4.x, d = a4.y, row0 = a4.z;
        const unsigned R0 = (unsigned)r2.x, R1 = (unsigned)r2.y;
        const int tX = use_time ? p.time_index[X] : 0;

        // ---- stage the row (this warp's slice) ------------------------------------------------
        __syncwarp();
        if (lane < d) {
            const int v = __ldcg(&adj[row0 + lane]);
            S.adj[lane] = v;
            S.cx[lane] = p.C[(size_t)X * n + v];
            if (use_time) S.time[lane] = p.time_index[v];
            if (!HET) S.minkey[lane] = __ldcg(&p.minkey[row0 + lane]);
        }
        if (HET) {
            bool gone = false;
            if (lane < d) {
                const int v = __ldcg(&adj[row0 + lane]);
                const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (v >> 5)]);
                gone = !((wbits >> (v & 31)) & 1u);
            }
            const unsigned m = __ballot_sync(FULLMASK, gone);
            if (lane == 0) S.skipmask = m;
        }
        __syncwarp();
        {
            const int myv = (lane < d) ? S.adj[lane] : 0;
            for (int i0 = 0; i0 < d; i0 += 16) {  // sixteen independent gathers in flight per lane: two round trips per row
                float v[16];
#pragma unroll
                for (int u = 0; u < 16; u++) {
                    const int i = i0 + u;
                    v[u] = 1.0f;
                    if (i < d && lane < d && i != lane) v[u] = __ldg(&p.C[(size_t)S.adj[i] * n + myv]);
                }
#pragma unroll
                for (int u = 0; u < 16; u++)
                    if (i0 + u < d && lane < d) S.csub[(i0 + u) * CS_STRIDE + lane] = v[u];
            }
        }
        __syncwarp();

        // ---- this lane's ranks: [ra, rb) ---------------------------------------------------------
        const unsigned cnt = R1 - R0;
        const unsigned per = (cnt + 31u) / 32u;
        const unsigned ra = R0 + (unsigned)lane * per;
        const unsigned rb = (ra + per < R1) ? ra + per : R1;
        if (ra < rb) {
            int pos[LM];
            {
                unsigned rank = ra;
                int x = 0;
#pragma unroll
                for (int i = 0; i < LM; i++) {
                    for (;; x++) {
                        const unsigned c = s_binom[(d - 1 - x) * (CUSK_ML + 1) + (l - 1 - i)];
                        if (rank < c) break;
                        rank -= c;
                    }
                    pos[i] = x;
                    x++;
                }
            }
            for (unsigned R = ra; R < rb; R++) {
                // ---- LDL^T of C[S,S] (unit lower L, signed pivots D), X carried along ---------------
                //   G_ji = A_ji - sum_{k<i} G_jk L_ik,  L_ji = G_ji / D_i,  D_j = 1 - sum_{i<j} G_ji L_ji
                //   t = L^-1 c  =>  c' A^-1 c'' = sum_j t'_j t''_j / D_j
                double Lm[LREG ? LM * (LM - 1) / 2 + 1 : 1], invD[LM], tx[LM], txd[LM];
                int tS = -2147483647 - 1;
                unsigned inS = 0;
                bool any_drop = false;
                double h00 = 1.0;
#pragma unroll
                for (int j = 0; j < LM; j++) {
                    double G[LM];
                    double D = 1.0;
                    double t = (double)S.cx[pos[j]];
#pragma unroll
                    for (int i = 0; i < j; i++) {
                        double g = (double)S.csub[pos[j] * CS_STRIDE + pos[i]];
#pragma unroll
                        for (int k = 0; k < i; k++) g -= G[k] * LS::get(Lm, s_L, CsIdx<LT>::w(k, i), tid);
                        G[i] = g;
                        const double lji = g * invD[i];
                        LS::set(Lm, s_L, CsIdx<LT>::w(i, j), tid, lji);
                        D -= g * lji;
                        t -= lji * tx[i];
                    }
                    const bool drop = fabs(D) < CI_PIVOT_TOL;
                    invD[j] = drop ? 0.0 : 1.0 / D;
                    if (drop && j < LM - 1) any_drop = true;
                    tx[j] = t;
                    txd[j] = t * invD[j];
                    h00 -= t * txd[j];
                    inS |= 1u << pos[j];
                    if (use_time) tS = max(tS, S.time[pos[j]]);
                }
                if (any_drop) my_sing++;
                const double c2 = fabs(h00);
                double SP = 0.0;
                if (het_general) {
                    // pairs inside S and S x X (src/hetcor-cuPC-S.cu:3068-3089: N[later * n + earlier],
                    // tuple ordered X, Y, S_1..S_l)
#pragma unroll
                    for (int j = 0; j < LM; j++) {
                        const size_t rowj = (size_t)S.adj[pos[j]] * n;
                        SP += (double)__float2int_rz(p.N[rowj + X]);
#pragma unroll
                        for (int i = 0; i < j; i++) SP += (double)__float2int_rz(p.N[rowj + S.adj[pos[i]]]);
                    }
                }
                // ---- this set's neighbours Y: every member of adj(X) outside S ----------------------
                unsigned todo = ~inS & ((d >= 32) ? 0xffffffffu : ((1u << d) - 1u));
                if (HET) todo &= ~*(volatile unsigned *)&S.skipmask;
                while (todo) {
                    const int y = __ffs(todo) - 1;
                    todo &= todo - 1;
                    if (HET) {
                        if (use_time && tS > max(tX, S.time[y])) continue;
                    } else {
                        if (S.minkey[y] < (unsigned long long)R) continue;
                    }
                    my_tests++;
                    double ty[LM];
                    double h01 = (double)S.cx[y], h11 = 1.0;
#pragma unroll
                    for (int j = 0; j < LM; j++) {
                        double u = (double)S.csub[y * CS_STRIDE + pos[j]];
#pragma unroll
                        for (int i = 0; i < j; i++) u -= LS::get(Lm, s_L, CsIdx<LT>::w(i, j), tid) * ty[i];
                        ty[j] = u;
                        h01 -= u * txd[j];
                        h11 -= u * u * invD[j];
                    }
                    const double lhs = h01 * h01;
                    const double den = c2 * fabs(h11);
                    double T2 = T2u;
                    if (het_general) {
                        const int Y = S.adj[y];
                        double sum = SP + (double)__float2int_rz(p.N[(size_t)Y * n + X]);
#pragma unroll
                        for (int j = 0; j < LM; j++) sum += (double)__float2int_rz(p.N[(size_t)S.adj[pos[j]] * n + Y]);
                        const double m = sum / npairs;
                        const double th = p.th_het / sqrt(m - (double)l - 3.0);
                        const double t = tanh(th);
                        T2 = t * t;
                    }
                    // |rho| < tanh(th)  or, on the |rho| > 1 branch, |rho| > coth(th)
                    const bool rem = (lhs < T2 * den) || (den > 0.0 && lhs * T2 > den);
                    if (rem) {
                        const int Y = S.adj[y];
                        if (HET) {
                            atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                            atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                            if (p.removed) p.removed[row0 + y] = 1;
                            atomicOr(&S.skipmask, 1u << y);
                        } else {
                            atomicMin(&p.minkey[row0 + y], (u64)R);
                            atomicMin(&S.minkey[y], (unsigned long long)R);
                        }
                    }
                }
                // ---- successor set -----------------------------------------------------------------
                int ii = -1;
#pragma unroll
                for (int j = 0; j < LM; j++)
                    if (pos[j] < d - l + j) ii = j;
#pragma unroll
                for (int j = 0; j < LM; j++) {
                    if (j == ii)
                        pos[j]++;
                    else if (j > ii && ii >= 0 && j > 0)
                        pos[j] = pos[j - 1] + 1;
                }
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        my_tests += __shfl_xor_sync(FULLMASK, my_tests, o);
        my_sing += __shfl_xor_sync(FULLMASK, my_sing, o);
    }
    if (lane == 0) {
        if (my_tests) atomicAdd(&s_tot[0], (unsigned long long)my_tests);
        if (my_sing) atomicAdd(&s_tot[1], (unsigned long long)my_sing);
    }
    __syncthreads();
    if (tid == 0) {
        if (s_tot[0]) atomicAdd((unsigned long long *)&ctl->executed, s_tot[0]);
        if (s_tot[1]) atomicAdd((unsigned long long *)&ctl->singular, s_tot[1]);
    }
}

// Pascal's triangle in shared memory, one row per step: lane r holds C(m, r) (first warp of the CTA; barrier by the caller)
__device__ __forceinline__ void cs_fill_binom(unsigned *s_binom)
{
    const int lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {
        unsigned prev = (lane == 0) ? 1u : 0u;  // row m = 0
        for (int m = 0; m <= CS_DMAX; m++) {
            if (lane <= CUSK_ML) s_binom[m * (CUSK_ML + 1) + lane] = prev;
            const unsigned left = __shfl_up_sync(FULLMASK, prev, 1);
            prev = (lane == 0) ? 1u : prev + left;  // C(m+1, r) = C(m, r) + C(m, r-1); < 2^32 for m <= 33, r <= 14
        }
    }
}

template <bool HET, int LT>
__global__ void __launch_bounds__(CS_NTHREADS(LT), (LT <= 6 || LT > CS_LREG) ? 2 : 1) ci_small_kernel(const CiParams p)
{
    extern __shared__ double s_L[];  // LT > CS_LREG: LT (LT - 1) / 2 doubles per thread
    __shared__ CsWarp s_w[CS_NWARPS(LT)];
    __shared__ unsigned s_binom[(CS_DMAX + 1) * (CUSK_ML + 1)];  // C(m, r), m <= 32, r <= 14
    LevelCtl *ctl = p.ctl;
    pdl_enter();
    if (ctl->done) return;
    // edge-sharded mode: the level plan decides (identically on every rank) whether this level's items are dealt out
    const int srank = ctl->shard_on ? p.shard_rank : 0, sworld = ctl->shard_on ? p.shard_world : 1;
    cs_fill_binom(s_binom);
    __syncthreads();
    cs_body<HET, LT>(p, ctl, p.work_counter_small, p.items_small, p.adj, p.item_cap, ctl->nitems_small, ctl->T2, srank,
                     sworld, s_w, s_binom, s_L);
}

// ---- self-planned deep levels (multi-level kernel) -------------------------------------------------------------------------
// Deep levels of a light block run a few hundred thousand tests each, and the degree / plan / compaction launches in front of
// them plus the launches themselves cost more than the tests (profiles/r2_launches_bench_steps1_call_p.csv: ~17 us of plan and
// ~20 us of CI kernel per level for ~15 us of work).  When every row is known to be short (host: max degree <= CI_SELF_DMAX at
// level 3, het engine, one GPU) consecutive levels run inside ONE cooperative launch, each level as
//   A. a warp per row (grid-stride): popcount and expansion of the bitmap row into self_adj[row][0..d) -- the adjacency as of
//      the level's start, fixed stride, so no CSR offsets and no prefix sums -- and the level totals (max degree, edges,
//      canonical tests, conditioning sets) by atomics;                                                       [grid barrier]
//   B. the rows' conditioning sets are cut into work items of `chunk` sets, chunk chosen from the level's set total so that
//      the items cover the launch's warps (one atomicAdd per row reserves its slots: item order is irrelevant);  [grid barrier]
//   C. the tests, exactly as the one-level kernel runs them (cs_body);                                       [grid barrier]
// and the launch ends at the first level no row can reach (src/cuPC-S.cu:154-159) or at its last level.  Three launches
// cover levels 4..14 (4-6, 7-10, 11-14: the groups share a block size and a register budget with the one-level kernels).
__device__ __forceinline__ void cs_grid_barrier(unsigned *bar, unsigned target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        unsigned seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
        } while (seen < target);
    }
    __syncthreads();
}

__device__ __forceinline__ unsigned long long cs_globaltimer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// conditioning sets per work item of a level with `sets` sets on a launch of `lanes` threads: one item per warp of the
// launch while the level is light (fewest row stagings), four per warp when it is not
__device__ __forceinline__ u64 cs_self_chunk(u64 sets, u64 lanes)
{
    u64 per = (sets + lanes - 1) / lanes;
    if (sets > 8 * lanes) per = (sets + 4 * lanes - 1) / (4 * lanes);
    return 32 * (per ? per : 1);
}

// the work items of row r (d neighbours, nsets conditioning sets at this level), `chunk` sets each
__device__ __forceinline__ void cs_self_emit(const CiParams &p, LevelCtl *ctl, int r, int d, u64 nsets, u64 chunk, int lane)
{
    const int cnt = (int)((nsets + chunk - 1) / chunk);
    int base = 0;
    if (lane == 0) base = atomicAdd(&ctl->nitems_small, cnt);
    base = __shfl_sync(FULLMASK, base, 0);
    for (int k = lane; k < cnt; k += 32) {
        if (base + k >= p.item_cap) break;  // cannot happen: at most n + total sets / chunk items (see the host)
        const u64 R0 = (u64)k * chunk;
        const u64 R1 = (R0 + chunk < nsets) ? R0 + chunk : nsets;
        int4 *dst = (int4 *)&p.self_items[base + k];
        __stcg(dst, make_int4(r, d, r * CI_SELF_STRIDE, 0));
        __stcg((ulonglong2 *)(dst + 1), make_ulonglong2(R0, R1));
    }
}

// phase A of a level: degrees, fixed-stride adjacency, level totals; the row's bitmap words are fetched in one batch (the
// next row's while this one is expanded), so a warp pays one L2 round trip per row instead of one per 32 words.
//   EMIT: the items are cut here, `chunk` sets each (chunk from the bound the previous level left in its acc_mid)
template <bool EMIT>
__device__ __forceinline__ void cs_self_rows(const CiParams &p, int l, LevelCtl *ctl, const unsigned *s_binom, u64 chunk,
                                             unsigned long long *s_acc)
{
    constexpr int KW = (CI_SELF_NMAX / 32 + 31) / 32;  // bitmap words of a row per lane
    const int tid = threadIdx.x, lane = tid & 31;
    const int nw = blockDim.x >> 5;
    const int n = p.n, rw = p.rw;
    const int stride = gridDim.x * nw;
    u64 canon = 0, nnz = 0, sets = 0, next = 0;
    int maxd = 0, bad = 0;
    uint32_t cur[KW], nxt[KW];
    int r = blockIdx.x * nw + (tid >> 5);
#pragma unroll
    for (int k = 0; k < KW; k++) {
        const int w = k * 32 + lane;
        nxt[k] = (r < n && w < rw) ? __ldcg(&p.bits[(size_t)r * rw + w]) : 0u;
    }
    for (; r < n; r += stride) {
#pragma unroll
        for (int k = 0; k < KW; k++) {
            cur[k] = nxt[k];
            const int w = k * 32 + lane;
            nxt[k] = (r + stride < n && w < rw) ? __ldcg(&p.bits[(size_t)(r + stride) * rw + w]) : 0u;
        }
        int out = 0;
        int *arow = p.self_adj + (size_t)r * CI_SELF_STRIDE;
#pragma unroll
        for (int k = 0; k < KW; k++) {
            if (k * 32 >= rw) break;
            uint32_t word = cur[k];
            const int cnt = __popc(word);
            int inc = cnt;
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(FULLMASK, inc, o);
                if (lane >= o) inc += t;
            }
            int pos = out + inc - cnt;
            while (word) {
                const int b = __ffs(word) - 1;
                word &= word - 1;
                if (pos < CI_SELF_STRIDE) __stcg(&arow[pos], (k * 32 + lane) * 32 + b);
                pos++;
            }
            out += __shfl_sync(FULLMASK, inc, 31);
        }
        const int d = out;
        if (lane == 0) __stcg(&p.deg_out[r], d);
        maxd = max(maxd, d);
        nnz += (u64)d;
        if (d <= l) continue;
        if (d > CI_SELF_STRIDE) {  // cannot happen: degrees only fall and the host checked the bound
            bad = 1;
            continue;
        }
        const u64 nsets = (u64)s_binom[d * (CUSK_ML + 1) + l];
        sets += nsets;
        canon += nsets * (u64)(d - l);
        if (l < CUSK_ML && d > l + 1) next += (u64)s_binom[d * (CUSK_ML + 1) + l + 1];
        if (EMIT) cs_self_emit(p, ctl, r, d, nsets, chunk, lane);
    }
    // level totals: per CTA in shared memory, then one global atomic per CTA and counter (thousands of warps adding to the
    // same five words would queue up in L2 for microseconds); s_acc is zero at entry (the kernel clears it per level)
    if (lane == 0) {
        if (maxd) atomicMax((unsigned long long *)&s_acc[0], (unsigned long long)maxd);
        if (nnz) atomicAdd(&s_acc[1], (unsigned long long)nnz);
        if (canon) atomicAdd(&s_acc[2], (unsigned long long)canon);
        if (sets) atomicAdd(&s_acc[3], (unsigned long long)sets);
        if (next) atomicAdd(&s_acc[4], (unsigned long long)next);
        if (bad) atomicExch(&ctl->infeasible, 4);
    }
    __syncthreads();
    if (tid == 0) {
        if (s_acc[0]) atomicMax(&ctl->maxdeg, (int)s_acc[0]);
        if (s_acc[1]) atomicAdd((unsigned long long *)&ctl->nnz, s_acc[1]);
        if (s_acc[2]) atomicAdd((unsigned long long *)&ctl->canon, s_acc[2]);
        if (s_acc[3]) atomicAdd((unsigned long long *)&ctl->acc_small, s_acc[3]);
        if (s_acc[4]) atomicAdd((unsigned long long *)&ctl->acc_mid, s_acc[4]);  // bound of level l + 1's sets
        s_acc[0] = s_acc[1] = s_acc[2] = s_acc[3] = s_acc[4] = 0ull;  // for the next level (a grid barrier follows)
    }
}

// phase B (first self-planned level of a call only: no bound from a previous level): work items of `chunk` sets
__device__ __forceinline__ void cs_self_items(const CiParams &p, int l, LevelCtl *ctl, const unsigned *s_binom, u64 chunk)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int nw = blockDim.x >> 5;
    for (int r = blockIdx.x * nw + (tid >> 5); r < p.n; r += gridDim.x * nw) {
        const int d = __ldcg(&p.deg_out[r]);
        if (d <= l || d > CI_SELF_STRIDE) continue;
        cs_self_emit(p, ctl, r, d, (u64)s_binom[d * (CUSK_ML + 1) + l], chunk, lane);
    }
}

// ---- a WARP per conditioning set (multi-level kernel, levels that hold fewer sets than a few per warp) ---------------------
// With one set per lane the level waits for a single thread's chain of l^3/6 + (d - l) l^2/2 dependent FP64 operations
// (~30 us at l = 14).  Here the lanes are the COLUMNS of a right-looking elimination on M = C[S, {adj(X), X}]: lane c < d
// holds neighbour c, lane d holds X; after the pivots s_0..s_{j-1}
//     w_j[c] = M[s_j][c] - sum_{i<j} (w_i[s_j] / D_i) w_i[c],   D_j = w_j[s_j]
// (w_i[s_j] comes from lane s_j by a shuffle), so w_j[c] = (L^-1 C[S, c])_j -- the forward-substituted column of cs_body,
// accumulated in the same order: the two forms take the same decisions bit for bit -- and the chain is ~l^2/2 steps.
// A set costs ~320 warp instructions here against ~19 per set there (32 sets per instruction), so the kernel takes this
// form only when the level is short of sets (uniform N only: the per-pair form gathers l^2/2 sample sizes per set).
template <int LT>
__device__ __forceinline__ void cs_warp_body(const CiParams &p, LevelCtl *ctl, int *work_counter, const CiItem *items,
                                             const int *adj, int item_cap, int nitems, double T2u, CsWarp *s_w,
                                             const unsigned *s_binom, double *s_L)
{
    static_assert(LT > CS_LREG, "the warp-per-set form keeps its columns in the dynamic shared memory of the deep levels");
    constexpr int LM = LT;
    constexpr int NT = CS_NTHREADS(LT);
    constexpr int l = LT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = p.n;
    const bool use_time = p.time_index != nullptr;
    CsWarp &S = s_w[warp];
    // per warp: W[LM][32] doubles, Dv[LM] doubles, P[LM] ints -- inside the launch's LT (LT - 1) / 2 doubles per thread
    constexpr int WSTRIDE = 32 * CUSK_ML + CUSK_ML + (CUSK_ML + 1) / 2;
    static_assert(WSTRIDE <= 32 * ((CS_LREG + 1) * CS_LREG / 2), "per-warp scratch must fit the shallowest deep level's share");
    double *W = s_L + warp * WSTRIDE;
    double *Dv = W + 32 * CUSK_ML;
    int *P = (int *)(Dv + CUSK_ML);
    unsigned my_tests = 0, my_sing = 0;  // per warp (lane 0's copy is the warp's)
    __shared__ unsigned long long s_tot[2];
    if (tid == 0) s_tot[0] = s_tot[1] = 0ull;
    __syncthreads();
    const int warps_total = (int)gridDim.x * (NT / 32);
    bool first = true;

    for (;;) {
        int item = 0;
        if (first) {
            first = false;
            item = (int)blockIdx.x * (NT / 32) + warp;
        } else {
            if (nitems <= warps_total) break;
            if (lane == 0) item = warps_total + atomicAdd(work_counter, 1);
            item = __shfl_sync(FULLMASK, item, 0);
        }
        if (item >= nitems) break;
        if (item >= item_cap) {
            if (lane == 0) ctl->infeasible = 3;
            break;
        }
        const int4 a4 = __ldcg((const int4 *)&items[item]);
        const ulonglong2 r2 = __ldcg((const ulonglong2 *)&items[item].R0);
        const int X = a4.x, d = a4.y, row0 = a4.z;
        const unsigned R0 = (unsigned)r2.x, R1 = (unsigned)r2.y;
        const int tX = use_time ? p.time_index[X] : 0;

        // ---- stage the row (as cs_body does) ------------------------------------------------------
        __syncwarp();
        {
            bool gone = false;
            if (lane < d) {
                const int v = __ldcg(&adj[row0 + lane]);
                S.adj[lane] = v;
                S.cx[lane] = p.C[(size_t)X * n + v];
                if (use_time) S.time[lane] = p.time_index[v];
                const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (v >> 5)]);
                gone = !((wbits >> (v & 31)) & 1u);
            }
            const unsigned m = __ballot_sync(FULLMASK, gone);
            if (lane == 0) S.skipmask = m;
        }
        __syncwarp();
        {
            const int myv = (lane < d) ? S.adj[lane] : 0;
            for (int i0 = 0; i0 < d; i0 += 16) {
                float v[16];
#pragma unroll
                for (int u = 0; u < 16; u++) {
                    const int i = i0 + u;
                    v[u] = 1.0f;
                    if (i < d && lane < d && i != lane) v[u] = __ldg(&p.C[(size_t)S.adj[i] * n + myv]);
                }
#pragma unroll
                for (int u = 0; u < 16; u++)
                    if (i0 + u < d && lane < d) S.csub[(i0 + u) * CS_STRIDE + lane] = v[u];
            }
        }
        __syncwarp();

        // ---- the item's sets, one after the other; every lane carries the same positions ------------
        int pos[LM];
        {
            unsigned rank = R0;
            int x = 0;
#pragma unroll
            for (int i = 0; i < LM; i++) {
                for (;; x++) {
                    const unsigned c = s_binom[(d - 1 - x) * (CUSK_ML + 1) + (l - 1 - i)];
                    if (rank < c) break;
                    rank -= c;
                }
                pos[i] = x;
                x++;
            }
        }
        const bool is_nb = lane < d, is_x = lane == d;
        const double cxy = is_nb ? (double)S.cx[lane] : 0.0;
        for (unsigned R = R0; R < R1; R++) {
            // The elimination as ROLLED loops over shared memory (W[j][c] = w_j[c], Dv[j] = 1 / D_j, P[j] = s_j): a set is
            // run once per warp, so straight-line code unrolled over l (l + 1) / 2 steps would mostly wait for its own
            // instruction fetches.
            int tS = -2147483647 - 1;
            unsigned inS = 0;
#pragma unroll
            for (int j = 0; j < LM; j++) {
                if (lane == j) P[j] = pos[j];
                inS |= 1u << pos[j];
                if (use_time) tS = max(tS, S.time[pos[j]]);
            }
            __syncwarp();
            bool any_drop = false;
#pragma unroll 1
            for (int j = 0; j < LM; j++) {
                const int sj = P[j];
                double acc = is_nb ? (double)S.csub[sj * CS_STRIDE + lane] : (is_x ? (double)S.cx[sj] : 0.0);
#pragma unroll 2
                for (int i = 0; i < j; i++) {
                    const double g = W[i * 32 + sj];  // w_i[s_j]
                    const double lji = g * Dv[i];
                    acc -= lji * W[i * 32 + lane];
                }
                W[j * 32 + lane] = acc;
                __syncwarp();
                const double D = W[j * 32 + sj];
                const bool drop = fabs(D) < CI_PIVOT_TOL;
                if (lane == 0) Dv[j] = drop ? 0.0 : 1.0 / D;
                if (drop && j < LM - 1) any_drop = true;
                __syncwarp();
            }
            if (any_drop) my_sing++;
            // X's column (lane d) gives t_x; every lane then has its own column's h01 / h11 against it
            double h00 = 1.0, h01 = cxy, h11 = 1.0;
#pragma unroll 1
            for (int j = 0; j < LM; j++) {
                const double t = W[j * 32 + d], inv = Dv[j], wj = W[j * 32 + lane];
                const double txd = t * inv;
                h00 -= t * txd;
                h01 -= wj * txd;
                h11 -= wj * wj * inv;
            }
            const double c2 = fabs(h00);
            bool todo = is_nb && !((inS >> lane) & 1u) && !((*(volatile unsigned *)&S.skipmask >> lane) & 1u);
            if (todo && use_time && tS > max(tX, S.time[lane])) todo = false;
            my_tests += __popc(__ballot_sync(FULLMASK, todo));
            if (todo) {
                const double lhs = h01 * h01;
                const double den = c2 * fabs(h11);
                const bool rem = (lhs < T2u * den) || (den > 0.0 && lhs * T2u > den);
                if (rem) {
                    const int Y = S.adj[lane];
                    atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                    atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                    atomicOr(&S.skipmask, 1u << lane);
                }
            }
            __syncwarp();
            // ---- successor set ---------------------------------------------------------------------
            int ii = -1;
#pragma unroll
            for (int j = 0; j < LM; j++)
                if (pos[j] < d - l + j) ii = j;
#pragma unroll
            for (int j = 0; j < LM; j++) {
                if (j == ii)
                    pos[j]++;
                else if (j > ii && ii >= 0 && j > 0)
                    pos[j] = pos[j - 1] + 1;
            }
        }
    }
    if (lane == 0) {
        if (my_tests) atomicAdd(&s_tot[0], (unsigned long long)my_tests);
        if (my_sing) atomicAdd(&s_tot[1], (unsigned long long)my_sing);
    }
    __syncthreads();
    if (tid == 0) {
        if (s_tot[0]) atomicAdd((unsigned long long *)&ctl->executed, s_tot[0]);
        if (s_tot[1]) atomicAdd((unsigned long long *)&ctl->singular, s_tot[1]);
    }
}

template <int L0, int L1>
__device__ __forceinline__ void cs_multi_dispatch(int l, bool warp_mode, const CiParams &p, LevelCtl *ctl, int nitems, double T2u,
                                                  CsWarp *s_w, const unsigned *s_binom, double *s_L)
{
    if constexpr (L0 <= L1) {
        if (l == L0) {
            bool done_here = false;
            if constexpr (L0 > CS_LREG) {
                if (warp_mode) {
                    cs_warp_body<L0>(p, ctl, p.work_counter_small + L0, p.self_items, p.self_adj, p.item_cap, nitems, T2u, s_w,
                                     s_binom, s_L);
                    done_here = true;
                }
            }
            if (!done_here)
                cs_body<true, L0, true>(p, ctl, p.work_counter_small + L0, p.self_items, p.self_adj, p.item_cap, nitems, T2u, 0,
                                        1, s_w, s_binom, s_L);
        } else {
            cs_multi_dispatch<L0 + 1, L1>(l, warp_mode, p, ctl, nitems, T2u, s_w, s_binom, s_L);
        }
    }
}

// p.ctl and p.work_counter_small are the BASES of the per-level arrays here; levels l_first..l_last (within L0..L1) run.
template <int L0, int L1>
__global__ void __launch_bounds__(CS_NTHREADS(L0), (L1 <= 6 || L0 > CS_LREG) ? 2 : 1)
    ci_small_multi_kernel(const CiParams p, int l_first, int l_last, int maxl)
{
    static_assert(CS_NTHREADS(L0) == CS_NTHREADS(L1), "one block size per launch");
    extern __shared__ double s_L[];  // levels above CS_LREG: l (l - 1) / 2 doubles per thread
    __shared__ CsWarp s_w[CS_NWARPS(L0)];
    __shared__ unsigned s_binom[(CS_DMAX + 1) * (CUSK_ML + 1)];
    pdl_enter();
    LevelCtl *base = p.ctl;
    const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
    if (base[l_first - 1].done) {  // an earlier launch ended the run (complete before this one started: the same for every CTA)
        if (lead)
            for (int k = l_first; k <= maxl; k++) base[k].done = 1;
        return;
    }
    __shared__ unsigned long long s_acc[5];
    if (threadIdx.x < 5) s_acc[threadIdx.x] = 0ull;
    cs_fill_binom(s_binom);
    __syncthreads();
    unsigned *bar = &base[l_first].ticket;  // zero at the start of a call; counts arrivals of all barriers of this launch
    unsigned epoch = 0;
    const u64 lanes = (u64)gridDim.x * blockDim.x, warps = lanes / 32;
    // uniform N only, levels above CS_LREG (their launch has the dynamic shared memory); CUSK_WARP_SETS=0: a thread per set
    const bool warp_ok = p.N == nullptr && p.warp_sets && L0 > CS_LREG;
    for (int l = l_first; l <= l_last; l++) {
        LevelCtl *ctl = base + l;
        if (lead) ctl->t_begin = cs_globaltimer();
        // every level but the call's first self-planned one (p.l) knows an upper bound of its set total: the level before
        // it summed C(d, l) over its own rows (degrees only fall), so the items are cut in the row phase itself
        const bool emit = l > p.l;
        // a level short of sets (at most CS_WARP_SETS per warp of the launch, by the same estimate that cuts the items) takes
        // the warp-per-set form, items of est / warps sets; otherwise a thread per set
        u64 est = emit ? *(volatile u64 *)&base[l - 1].acc_mid : 0;
        bool warp_mode = emit && warp_ok && est <= (u64)CS_WARP_SETS * warps;
        if (emit)
            cs_self_rows<true>(p, l, ctl, s_binom, warp_mode ? (est + warps - 1) / warps + (est == 0) : cs_self_chunk(est, lanes),
                               s_acc);
        else
            cs_self_rows<false>(p, l, ctl, s_binom, 0, s_acc);
        cs_grid_barrier(bar, ++epoch * gridDim.x);
        if (lead) ctl->t_rows = cs_globaltimer();
        const u64 sets = *(volatile u64 *)&ctl->acc_small;
        const bool stop = sets == 0 || *(volatile int *)&ctl->infeasible != 0;
        double T2u;
        {
            const double m = (double)__float2int_rz(p.n_uniform);
            const double t = tanh(p.th_het / sqrt(m - (double)l - 3.0));
            T2u = t * t;
        }
        if (stop) {  // no row has more than l neighbours: this level and all later ones do not run
            if (lead) {
                ctl->T2 = T2u;
                for (int k = l; k <= maxl; k++) base[k].done = 1;
            }
            return;
        }
        if (!emit) {
            est = sets;
            warp_mode = warp_ok && est <= (u64)CS_WARP_SETS * warps;
            cs_self_items(p, l, ctl, s_binom, warp_mode ? (est + warps - 1) / warps : cs_self_chunk(est, lanes));
            cs_grid_barrier(bar, ++epoch * gridDim.x);
        }
        int nitems = *(volatile int *)&ctl->nitems_small;
        if (nitems > p.item_cap) nitems = p.item_cap;
        if (lead) {
            ctl->T2 = T2u;
            ctl->done = 0;
            ctl->t_planned = cs_globaltimer();
        }
        cs_multi_dispatch<L0, L1>(l, warp_mode, p, ctl, nitems, T2u, s_w, s_binom, s_L);
        if (lead) ctl->t_body = cs_globaltimer();
        cs_grid_barrier(bar, ++epoch * gridDim.x);  // the level's removals are in memory
        if (lead) ctl->t_end = cs_globaltimer();
    }
}

template <int L0, int L1>
static int cs_launch_multi(cudaStream_t st, const CiParams &p, int l_first, int l_last, int maxl, bool coop)
{
    constexpr int NT = CS_NTHREADS(L0);
    constexpr int dyn = (L1 <= CS_LREG) ? 0 : L1 * (L1 - 1) / 2 * 8 * NT;
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        if (dyn > 0 && cudaFuncSetAttribute(ci_small_multi_kernel<L0, L1>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn) !=
                           cudaSuccess) {
            cusk_set_error("cudaFuncSetAttribute(ci_small_multi_kernel<%d, %d>, %d B): %s", L0, L1, dyn,
                           cudaGetErrorString(cudaGetLastError()));
            return CUSK_E_CUDA;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_small_multi_kernel<L0, L1>, NT, dyn) != cudaSuccess || nb < 1)
            nb = 1;
        ctas_per_sm = nb > 2 ? 2 : nb;
        ctas_of_device[dev].store(ctas_per_sm, std::memory_order_release);
    }
    const dim3 grid((p.sm_count > 0 ? p.sm_count : 148) * ctas_per_sm);
    if (!coop) {
        // plain (programmatic dependent) launch: the grid fits the device, so its CTAs are co-resident unless ANOTHER
        // barrier-synchronised grid runs beside it -- the caller serialises those (skeleton.cu)
        const cudaError_t e = launch_k(ci_small_multi_kernel<L0, L1>, grid, dim3(NT), dyn, st, p, l_first, l_last, maxl);
        if (e != cudaSuccess) {
            cusk_set_error("launch of ci_small_multi_kernel<%d, %d>: %s", L0, L1, cudaGetErrorString(e));
            return CUSK_E_CUDA;
        }
        return CUSK_OK;
    }
    CiParams q = p;
    void *args[] = {&q, &l_first, &l_last, &maxl};
    const cudaError_t e = cudaLaunchCooperativeKernel((const void *)ci_small_multi_kernel<L0, L1>, grid, dim3(NT), args, dyn, st);
    if (e != cudaSuccess) {
        cusk_set_error("cooperative launch of ci_small_multi_kernel<%d, %d>: %s", L0, L1, cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    return CUSK_OK;
}

// Levels l_first..l_last (4 <= l_first) of the het engine as up to three cooperative launches; returns the launch count in
// *launches.  p.ctl / p.work_counter_small: bases of the per-level arrays.
int launch_ci_small_multi(cudaStream_t st, const CiParams &p, int l_first, int l_last, int maxl, bool coop, int *launches)
{
    int rc = CUSK_OK, nl = 0;
    auto span = [&](int lo, int hi, auto fn) {
        const int a = l_first > lo ? l_first : lo, b = l_last < hi ? l_last : hi;
        if (a <= b && rc == CUSK_OK) rc = (nl++, fn(a, b));
    };
    span(CI_DFS_LMIN + 1, 6, [&](int a, int b) { return cs_launch_multi<CI_DFS_LMIN + 1, 6>(st, p, a, b, maxl, coop); });
    span(7, CS_LREG, [&](int a, int b) { return cs_launch_multi<7, CS_LREG>(st, p, a, b, maxl, coop); });
    span(CS_LREG + 1, CUSK_ML, [&](int a, int b) { return cs_launch_multi<CS_LREG + 1, CUSK_ML>(st, p, a, b, maxl, coop); });
    if (launches) *launches = nl;
    return rc;
}

template <bool HET, int LT>
static int cs_launch(cudaStream_t st, const CiParams &p)
{
    constexpr int NT = CS_NTHREADS(LT);
    constexpr int dyn = (LT <= CS_LREG) ? 0 : LT * (LT - 1) / 2 * 8 * NT;
    // per instantiation AND per device: function attributes belong to the device's context, and the block
    // scheduler / the threaded sharded mode drive several devices from one process
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        if (dyn > 0 && cudaFuncSetAttribute(ci_small_kernel<HET, LT>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn) !=
                           cudaSuccess) {
            cusk_set_error("cudaFuncSetAttribute(ci_small_kernel<%d>, %d B): %s", LT, dyn,
                           cudaGetErrorString(cudaGetLastError()));
            return CUSK_E_CUDA;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_small_kernel<HET, LT>, NT, dyn) != cudaSuccess || nb < 1)
            nb = 1;
        ctas_per_sm = nb;
        ctas_of_device[dev].store(nb, std::memory_order_release);
    }
    launch_k(ci_small_kernel<HET, LT>, dim3((p.sm_count > 0 ? p.sm_count : 148) * ctas_per_sm), dim3(NT), dyn, st, p);
    return CUSK_OK;
}

template <bool HET>
static int cs_dispatch(cudaStream_t st, const CiParams &p)
{
    switch (p.l) {
        case 1: return cs_launch<HET, 1>(st, p);
        case 2: return cs_launch<HET, 2>(st, p);
        case 3: return cs_launch<HET, 3>(st, p);
        case 4: return cs_launch<HET, 4>(st, p);
        case 5: return cs_launch<HET, 5>(st, p);
        case 6: return cs_launch<HET, 6>(st, p);
        case 7: return cs_launch<HET, 7>(st, p);
        case 8: return cs_launch<HET, 8>(st, p);
        case 9: return cs_launch<HET, 9>(st, p);
        case 10: return cs_launch<HET, 10>(st, p);
        case 11: return cs_launch<HET, 11>(st, p);
        case 12: return cs_launch<HET, 12>(st, p);
        case 13: return cs_launch<HET, 13>(st, p);
        case 14: return cs_launch<HET, 14>(st, p);
        default: return CUSK_OK;
    }
}

int launch_ci_small(cudaStream_t st, bool het, const CiParams &p)
{
    return het ? cs_dispatch<true>(st, p) : cs_dispatch<false>(st, p);
}
