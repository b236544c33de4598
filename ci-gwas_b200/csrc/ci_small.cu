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

// ---- self-planned level (SELF) -------------------------------------------------------------------------------------------
// Deep levels of a light block run a few hundred thousand tests in ~15 us, and the separate degree / plan / compaction
// launch in front of them costs as much again.  When every row is known to be short (host: max degree <= CI_SELF_DMAX, het
// engine, one GPU) the CI kernel does that work itself in a prologue, without prefix sums:
//   * a warp per row (grid-stride): popcount and expansion of the bitmap row into self_adj[row][0..d) -- the adjacency as
//     of the level's start, fixed stride, so no CSR offsets are needed;
//   * the row's conditioning sets are cut into at most CI_SELF_MAXI items of at least CI_SELF_ITEM_TESTS tests, which the
//     warp appends to the level's item table (one atomicAdd reserves the slots: the order of items is irrelevant);
//   * level totals (max degree, edges, canonical test count) by atomics; then ONE grid barrier (cooperative launch) and the
//     kernel goes on as usual, pulling items from the table.  The level runs iff some row has more than l neighbours, i.e.
//     iff the table is not empty (src/cuPC-S.cu:154-159).
// Returns the number of items; *T2 = tanh(threshold)^2 of the level.
__device__ __forceinline__ int cs_self_plan(const CiParams &p, int l, LevelCtl *ctl, const unsigned *s_binom)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int nw = blockDim.x >> 5;
    const int n = p.n, rw = p.rw;
    u64 canon = 0, nnz = 0;
    int maxd = 0, bad = 0;
    for (int r = blockIdx.x * nw + (tid >> 5); r < n; r += gridDim.x * nw) {
        int out = 0;
        int *arow = p.self_adj + (size_t)r * CI_SELF_STRIDE;
        for (int w0 = 0; w0 < rw; w0 += 32) {
            const int w = w0 + lane;
            uint32_t word = (w < rw) ? p.bits[(size_t)r * rw + w] : 0u;
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
                if (pos < CI_SELF_STRIDE) arow[pos] = w * 32 + b;
                pos++;
            }
            out += __shfl_sync(FULLMASK, inc, 31);
        }
        const int d = out;
        if (lane == 0) p.deg_out[r] = d;
        maxd = max(maxd, d);
        nnz += (u64)d;
        if (d <= l) continue;
        if (d > CI_SELF_STRIDE) {  // cannot happen: degrees only fall and the host checked the bound
            bad = 1;
            continue;
        }
        const u64 nsets = (u64)s_binom[d * (CUSK_ML + 1) + l];
        const u64 tests = nsets * (u64)(d - l);
        canon += tests;
        u64 cnt = (tests + CI_SELF_ITEM_TESTS - 1) / CI_SELF_ITEM_TESTS;
        if (cnt > CI_SELF_MAXI) cnt = CI_SELF_MAXI;
        if (cnt > nsets) cnt = nsets;
        const u64 chunk = (nsets + cnt - 1) / cnt;
        cnt = (nsets + chunk - 1) / chunk;
        int base = 0;
        if (lane == 0) base = atomicAdd(&ctl->nitems_small, (int)cnt);
        base = __shfl_sync(FULLMASK, base, 0);
        for (int k = lane; k < (int)cnt; k += 32) {
            if (base + k >= p.item_cap) break;  // cannot happen: cap = n * CI_SELF_MAXI
            CiItem it;
            it.X = r;
            it.d = d;
            it.row0 = r * CI_SELF_STRIDE;
            it.pad = 0;
            it.R0 = (u64)k * chunk;
            it.R1 = (it.R0 + chunk < nsets) ? it.R0 + chunk : nsets;
            p.self_items[base + k] = it;
        }
    }
    if (lane == 0) {
        if (maxd) atomicMax(&ctl->maxdeg, maxd);
        if (nnz) atomicAdd((unsigned long long *)&ctl->nnz, (unsigned long long)nnz);
        if (canon) atomicAdd((unsigned long long *)&ctl->canon, (unsigned long long)canon);
        if (bad) atomicExch(&ctl->infeasible, 4);
    }
    // grid barrier: every CTA is resident (cooperative launch)
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        atomicAdd(&ctl->ticket, 1u);
        while (*(volatile unsigned *)&ctl->ticket < gridDim.x) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
    int nitems = *(volatile int *)&ctl->nitems_small;
    if (*(volatile int *)&ctl->infeasible) nitems = 0;
    if (nitems > p.item_cap) nitems = p.item_cap;
    return nitems;
}

template <bool HET, int LT, bool SELF = false>
__global__ void __launch_bounds__(CS_NTHREADS(LT), (LT <= 6 || LT > CS_LREG) ? 2 : 1) ci_small_kernel(const CiParams p)
{
    constexpr int LM = CsIdx<LT>::LM;
    constexpr int NT = CS_NTHREADS(LT);
    constexpr bool LREG = (LT <= CS_LREG);
    using LS = CsL<LREG, NT>;
    extern __shared__ double s_L[];  // LT > CS_LREG: LT (LT - 1) / 2 doubles per thread
    __shared__ CsWarp s_w[CS_NWARPS(LT)];
    __shared__ unsigned s_binom[(CS_DMAX + 1) * (CUSK_ML + 1)];  // C(m, r), m <= 32, r <= 14
    LevelCtl *ctl = p.ctl;
    pdl_enter();
    if constexpr (SELF) {
        if (p.prev && p.prev->done) {  // an earlier level ended the run (the same for every CTA: that kernel is complete)
            if (blockIdx.x == 0 && threadIdx.x == 0) ctl->done = 1;
            return;
        }
    } else {
        if (ctl->done) return;
    }
    // edge-sharded mode: the level plan decides (identically on every rank) whether this level's items are dealt out
    const int srank = ctl->shard_on ? p.shard_rank : 0, sworld = ctl->shard_on ? p.shard_world : 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int l = LT;
    const int n = p.n;
    int nitems = SELF ? 0 : ctl->nitems_small;
    double T2u = SELF ? 0.0 : ctl->T2;
    const bool het_general = HET && p.N != nullptr;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    if (tid < 32) {
        // Pascal's triangle, one row per step: lane r holds C(m, r)
        unsigned prev = (lane == 0) ? 1u : 0u;  // row m = 0
        for (int m = 0; m <= CS_DMAX; m++) {
            if (lane <= CUSK_ML) s_binom[m * (CUSK_ML + 1) + lane] = prev;
            const unsigned left = __shfl_up_sync(FULLMASK, prev, 1);
            prev = (lane == 0) ? 1u : prev + left;  // C(m+1, r) = C(m, r) + C(m, r-1); < 2^32 for m <= 33, r <= 14
        }
    }
    __syncthreads();
    if constexpr (SELF) {
        nitems = cs_self_plan(p, l, ctl, s_binom);
        {
            const double m = (double)__float2int_rz(p.n_uniform);
            const double t = tanh(p.th_het / sqrt(m - (double)l - 3.0));
            T2u = t * t;
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            ctl->T2 = T2u;
            ctl->done = (nitems == 0) ? 1 : 0;
        }
        if (nitems == 0) return;
    }
    CsWarp &S = s_w[warp];
    u64 my_tests = 0;
    unsigned my_sing = 0;

    for (;;) {
        int item = 0;
        if (lane == 0) item = srank + sworld * atomicAdd(p.work_counter_small, 1);
        item = __shfl_sync(FULLMASK, item, 0);
        if (item >= nitems) break;
        if (item >= p.item_cap) {  // cannot happen: plan_kernel bounds the item count of a small level
            if (lane == 0) ctl->infeasible = 3;
            break;
        }
        const int4 a4 = *(const int4 *)&p.items_small[item];
        const ulonglong2 r2 = *(const ulonglong2 *)&p.items_small[item].R0;
        const int X = a4.x, d = a4.y, row0 = a4.z;
        const unsigned R0 = (unsigned)r2.x, R1 = (unsigned)r2.y;
        const int tX = use_time ? p.time_index[X] : 0;

        // ---- stage the row (this warp's slice) ------------------------------------------------
        __syncwarp();
        if (lane < d) {
            const int v = p.adj[row0 + lane];
            S.adj[lane] = v;
            S.cx[lane] = p.C[(size_t)X * n + v];
            if (use_time) S.time[lane] = p.time_index[v];
            if (!HET) S.minkey[lane] = __ldcg(&p.minkey[row0 + lane]);
        }
        if (HET) {
            bool gone = false;
            if (lane < d) {
                const int v = p.adj[row0 + lane];
                const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (v >> 5)]);
                gone = !((wbits >> (v & 31)) & 1u);
            }
            const unsigned m = __ballot_sync(FULLMASK, gone);
            if (lane == 0) S.skipmask = m;
        }
        __syncwarp();
        {
            const int myv = (lane < d) ? S.adj[lane] : 0;
            for (int i0 = 0; i0 < d; i0 += 8) {  // eight independent gathers in flight per lane
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int i = i0 + u;
                    v[u] = 1.0f;
                    if (i < d && lane < d && i != lane) v[u] = __ldg(&p.C[(size_t)S.adj[i] * n + myv]);
                }
#pragma unroll
                for (int u = 0; u < 8; u++)
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
        if (my_tests) atomicAdd((unsigned long long *)&ctl->executed, (unsigned long long)my_tests);
        if (my_sing) atomicAdd((unsigned long long *)&ctl->singular, (unsigned long long)my_sing);
    }
}

// self-planned form: cooperative launch (the prologue ends in a grid barrier)
template <int LT>
static int cs_launch_self(cudaStream_t st, const CiParams &p)
{
    constexpr int NT = CS_NTHREADS(LT);
    constexpr int dyn = (LT <= CS_LREG) ? 0 : LT * (LT - 1) / 2 * 8 * NT;
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        if (dyn > 0 && cudaFuncSetAttribute(ci_small_kernel<true, LT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            dyn) != cudaSuccess) {
            cusk_set_error("cudaFuncSetAttribute(ci_small_kernel<%d, self>, %d B): %s", LT, dyn,
                           cudaGetErrorString(cudaGetLastError()));
            return CUSK_E_CUDA;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_small_kernel<true, LT, true>, NT, dyn) != cudaSuccess ||
            nb < 1)
            nb = 1;
        ctas_per_sm = nb;
        ctas_of_device[dev].store(nb, std::memory_order_release);
    }
    CiParams q = p;
    void *args[] = {&q};
    const cudaError_t e = cudaLaunchCooperativeKernel((const void *)ci_small_kernel<true, LT, true>,
                                                      dim3((p.sm_count > 0 ? p.sm_count : 148) * ctas_per_sm), dim3(NT), args,
                                                      dyn, st);
    if (e != cudaSuccess) {
        cusk_set_error("cooperative launch of ci_small_kernel<%d, self>: %s", LT, cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    return CUSK_OK;
}

int launch_ci_small_self(cudaStream_t st, const CiParams &p)
{
    switch (p.l) {
        case 4: return cs_launch_self<4>(st, p);
        case 5: return cs_launch_self<5>(st, p);
        case 6: return cs_launch_self<6>(st, p);
        case 7: return cs_launch_self<7>(st, p);
        case 8: return cs_launch_self<8>(st, p);
        case 9: return cs_launch_self<9>(st, p);
        case 10: return cs_launch_self<10>(st, p);
        case 11: return cs_launch_self<11>(st, p);
        case 12: return cs_launch_self<12>(st, p);
        case 13: return cs_launch_self<13>(st, p);
        case 14: return cs_launch_self<14>(st, p);
        default: cusk_set_error("launch_ci_small_self: level %d", p.l); return CUSK_E_ARG;
    }
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
