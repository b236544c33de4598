// ci_dfs.cu -- level-l (l >= 3) CI tests of the rows of medium degree (up to CI_DFS_DMAX = 63 neighbours) whose
// conditioning sets each serve many tests (d - l large): one WARP walks the conditioning sets of its rank range in
// lexicographic order as a depth-first search, with the LANES spread over the row's neighbours (replaces
// cal_Indepl3..14, src/cuPC-S.cu:716-3020, and cal_Indepl3_ess..14_ess, src/hetcor-cuPC-S.cu:632-3052, for those
// rows; sm_100a).
//
// Why a third mapping.  ci_small.cu (a thread per set) spends l^2/2 FMAs per test on the forward substitution
// t_y = L^-1 C[S,y] and holds the whole factor per thread (172-255 registers for l >= 7), so its rate falls like
// 1/l^2 (measured on trait rows with ~29 neighbours: 141 G tests/s at level 6, 16 G at level 14); ci_level.cu (a CTA
// per row, 16 warps per SM) is built for hub rows with hundreds of neighbours.  Trait rows with a few dozen associated
// markers carry almost all tests of a dense block and never lose an edge, so every conditioning set costs d - l tests.
//
// Here the elimination is right-looking on the (d+1) x (d+1) matrix M = C[{adj(X), X}, {adj(X), X}]: after the pivots
// s_0..s_{j-1} the row  w_j[a] = M[s_j][a] - sum_{i<j} (w_i[s_j] / D_i) w_i[a]  (D_i = w_i[s_i]) lives in ONE REGISTER
// PER DEPTH AND COLUMN: lane (a mod 32) holds column a (two columns per lane when d + 1 > 32).  The scaled rows
// T[a][i] = w_i[a] / D_i are kept in shared memory, column-major per warp, so the coefficients of any pivot column are
// one broadcast read away (LDS.128: two depths per instruction).  A new row costs j FMAs per column; consecutive sets
// share their leading pivots, so only the rows from the first changed depth on are rebuilt.  With the prefix P = S
// minus its last element z fixed,
//     A[y] = M[x][y | P],  B[y] = M[y][y | P]  (per prefix),      u[y] = M[z][y | P]  (l - 1 FMAs per leaf z)
// give the test (X, y | P + z) by the first-order formula
//     h01 = A[y] - (A[z] / B[z]) u[y],   h11 = B[y] - u[y]^2 / B[z],   h00 = A[x] - A[z]^2 / B[z]
// (the per-column factors 1 / B[z], A[z] / B[z] and |h00| are tabulated once per prefix), i.e. l + 10 FP64 operations
// per test instead of l^2/2 + 3l, no division, barrier or shuffle per leaf, and 2 (l - 1) + 12 FP64 registers per lane.
// Same quantities, pivot rule (|D| < CI_PIVOT_TOL drops the direction) and decision rho^2 < tanh(th)^2 [or >
// coth(th)^2] as ci_small.cu / ci_level.cu; PC-stable semantics as there (sets from the adjacency compacted at level
// start; removals clear bits or lower the slot's minimum rank).
//
// A work item (row, rank range; class 2 of the level plan at l >= 3) goes to a CTA of 4 warps, which stages M once
// (<= 64 x 64 doubles) and splits the ranks evenly.  Rows with up to 31 neighbours run on the NS = 1 instantiation (one
// column per lane), rows with 32..63 on NS = 2; both walk the level's one item table with a queue each and skip the
// other's rows.  Algorithmic bytes per test: 4 (l + 2).
#include <atomic>
#include <type_traits>

#include "skeleton.cuh"

#define CD_NWARPS 4
#define CD_NTHREADS (CD_NWARPS * 32)
#define FULLMASK 0xffffffffu

// 1 / x to about an ulp (MUFU.RCP64H seed, two Newton steps) for |x| >= CI_PIVOT_TOL: the pivots' reciprocals
__device__ __forceinline__ double cd_rcp(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    r = fma(fma(-x, r, 1.0), r, r);
    r = fma(fma(-x, r, 1.0), r, r);
    return r;
}

template <int LT, int NS, bool GEN>
struct CdLayout {
    static constexpr int NC = 32 * NS;       // columns: neighbours 0..d-1, X at column d, zero padding above
    static constexpr int K = LT - 1;         // prefix length
    static constexpr int KP = (K + 1) & ~1;  // doubles per column of T (16-byte aligned columns)
    static constexpr size_t m_bytes = sizeof(double) * NC * NC;
    static constexpr size_t t_bytes = sizeof(double) * CD_NWARPS * NC * KP;
    static constexpr size_t p_bytes = sizeof(double) * CD_NWARPS * NC * 4;
    static constexpr size_t n_bytes = GEN ? sizeof(int) * NC * NC : 0;
    static constexpr size_t bytes = m_bytes + t_bytes + p_bytes + n_bytes + sizeof(int) * 2 * NC;
};

// MODE 0: individual-level engine (minimum separating-set rank per slot); 1: summary-statistic engine with one
// sample size for every pair; 2: the same with per-pair effective sample sizes N (own instantiation: the staged
// integer matrix and the per-test threshold cost registers and shared memory the other two should not carry)
template <int MODE, int LT, int NS>
__global__ void __launch_bounds__(CD_NTHREADS) ci_dfs_kernel(const CiParams p)
{
    constexpr bool HET = MODE != 0;
    constexpr bool GEN = MODE == 2;
    using LY = CdLayout<LT, NS, GEN>;
    constexpr int NC = LY::NC, K = LY::K, KP = LY::KP;
    static_assert(LT >= 3 && LT <= CUSK_ML, "levels 3..14");
    extern __shared__ __align__(16) unsigned char cd_smem[];
    double *s_M = (double *)cd_smem;
    double *s_T = (double *)(cd_smem + LY::m_bytes);
    double *s_P = (double *)(cd_smem + LY::m_bytes + LY::t_bytes);
    int *s_N = (int *)(cd_smem + LY::m_bytes + LY::t_bytes + LY::p_bytes);
    int *s_adj = (int *)(cd_smem + LY::m_bytes + LY::t_bytes + LY::p_bytes + LY::n_bytes);
    int *s_time = s_adj + NC;
    __shared__ int s_item;
    LevelCtl *ctl = p.ctl;
    pdl_enter();
    if (ctl->done) return;
    // edge-sharded mode: the level plan decides (identically on every rank) whether this level's items are dealt out
    const int srank = ctl->shard_on ? p.shard_rank : 0, sworld = ctl->shard_on ? p.shard_world : 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int l = LT;
    const int n = p.n;
    const int nitems = ctl->nitems_mid;
    const double T2u = ctl->T2;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    double *Tw = s_T + (size_t)warp * NC * KP;  // T[a * KP + i] = w_i[a] / D_i
    double *Pw = s_P + (size_t)warp * NC * 4;   // per column z: 1 / B[z], A[z] / B[z], |h00(z)|
    u64 my_tests = 0;
    unsigned my_sing = 0;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_item = srank + sworld * atomicAdd(NS == 1 ? p.work_counter_mid : p.work_counter_dfs2, 1);
        __syncthreads();
        const int item = s_item;
        if (item >= nitems) break;
        if (item >= p.item_cap) {  // cannot happen: the plan bounds the item counts
            if (tid == 0) ctl->infeasible = 3;
            break;
        }
        const int4 a4 = *(const int4 *)&p.items_mid[item];
        const ulonglong2 r2 = *(const ulonglong2 *)&p.items_mid[item].R0;
        const int X = a4.x, d = a4.y, row0 = a4.z;
        const u64 R0 = r2.x, R1 = r2.y;
        if ((d >= 32) != (NS == 2)) continue;  // the other instantiation's row (both walk the one item table)
        const int tX = use_time ? p.time_index[X] : 0;

        // ---- stage the row: columns, then M (and trunc(N)) ---------------------------------------------
        for (int i = tid; i < NC; i += CD_NTHREADS) {
            const int v = (i < d) ? p.adj[row0 + i] : (i == d ? X : -1);
            s_adj[i] = v;
            s_time[i] = (use_time && v >= 0) ? p.time_index[v] : 0;
        }
        __syncthreads();
        {
            const int total = (d + 1) * NC;
            for (int base = tid; base < total; base += CD_NTHREADS * 8) {
                float v[8];
                float w[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CD_NTHREADS;
                    v[u] = 0.0f;
                    w[u] = 0.0f;
                    if (idx < total) {
                        const int r = idx / NC, a = idx - r * NC;
                        if (a <= d) {
                            v[u] = (r == a) ? 1.0f : __ldg(&p.C[(size_t)s_adj[r] * n + s_adj[a]]);
                            // sample sizes are read as N[later][earlier] of the tuple X, Y, S_1..S_l
                            // (src/hetcor-cuPC-S.cu:3068-3089): rows 0..d-1 as they lie, row d (X) transposed,
                            // s_N[d][a] = N[adj a][X], so that every later read is along a row
                            if (GEN && r != a)
                                w[u] = (r == d) ? __ldg(&p.N[(size_t)s_adj[a] * n + s_adj[r]])
                                                : __ldg(&p.N[(size_t)s_adj[r] * n + s_adj[a]]);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CD_NTHREADS;
                    if (idx < total) {
                        s_M[idx] = (double)v[u];
                        if (GEN) s_N[idx] = __float2int_rz(w[u]);
                    }
                }
            }
        }
        __syncthreads();

        // ---- this warp's ranks ---------------------------------------------------------------------------
        const u64 cnt = R1 - R0;
        const int nw = (cnt < (u64)CD_NWARPS) ? (int)cnt : CD_NWARPS;
        if (warp >= nw) continue;
        const u64 per = (cnt + (u64)nw - 1) / (u64)nw;
        const u64 Rw0 = R0 + (u64)warp * per;
        const u64 Rw1 = (Rw0 + per < R1) ? Rw0 + per : R1;
        if (Rw0 >= Rw1) continue;

        int pos[LT];
        {
            int tmp[CUSK_ML];
            unrank_comb_fast(Rw0, d, l, tmp);
#pragma unroll
            for (int i = 0; i < LT; i++) pos[i] = tmp[i];
        }
        int zstart = pos[K];
        int zfirst = pos[K - 1] + 1;
        u64 Rbase = Rw0 - (u64)(zstart - zfirst);
        int changed = 0;
        unsigned nprefix = 0;
        double W[K][NS];        // raw elimination rows w_i[a] of this lane's columns
        double Ac[NS], Bc[NS];  // A, B after the first K - 1 pivots (stay valid while only the last prefix element moves)
        // per-column state of this lane
        bool alive[NS];
        u64 mk[NS];
        int tmaxcol[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) {
            alive[s] = false;
            mk[s] = CUSK_KEY_NONE;
            tmaxcol[s] = use_time ? max(tX, s_time[lane + 32 * s]) : 2147483647;
            Ac[s] = 0.0;
            Bc[s] = 0.0;
        }
        const int xl = d & 31;     // lane of column X
        const bool xhi = d >= 32;  // ... in the second slot

        // row j of the elimination (pivot pos[j]) for every column of this lane; j is a compile-time constant so that
        // W[j][.] are registers
        auto rebuild = [&](auto Jc) {
            constexpr int j = decltype(Jc)::value;
            if constexpr (j < K) {
                const int pj = pos[j];
                const double *Tp = Tw + pj * KP;  // coefficients of the pivot column: w_i[pj] / D_i, i < j
                double u[NS];
#pragma unroll
                for (int s = 0; s < NS; s++) u[s] = s_M[pj * NC + lane + 32 * s];
#pragma unroll
                for (int i = 0; i + 1 < j; i += 2) {
                    const double2 c = *(const double2 *)(Tp + i);
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = fma(-c.y, W[i + 1][s], fma(-c.x, W[i][s], u[s]));
                }
                if (j & 1) {
                    const double c = Tp[j - 1];
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = fma(-c, W[j - 1][s], u[s]);
                }
                double D;
                if (NS > 1 && pj >= 32)
                    D = __shfl_sync(FULLMASK, u[NS - 1], pj & 31);
                else
                    D = __shfl_sync(FULLMASK, u[0], pj & 31);
                const bool drop = fabs(D) < CI_PIVOT_TOL;
                const double invD = drop ? 0.0 : cd_rcp(D);
                if (drop && lane == 0) my_sing++;
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    W[j][s] = u[s];
                    Tw[(lane + 32 * s) * KP + j] = u[s] * invD;
                }
                __syncwarp();
                if (j == K - 2) {
                    // checkpoint: A, B over the first K - 1 pivots
                    const double *Tx = Tw + d * KP;
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        Ac[s] = s_M[d * NC + lane + 32 * s];
                        Bc[s] = (lane + 32 * s <= d) ? 1.0 : 0.0;
                    }
#pragma unroll
                    for (int i = 0; i < K - 1; i++) {
                        const double cx = Tx[i];
#pragma unroll
                        for (int s = 0; s < NS; s++) {
                            const double t = Tw[(lane + 32 * s) * KP + i];
                            Ac[s] = fma(-cx, W[i][s], Ac[s]);
                            Bc[s] = fma(-t, W[i][s], Bc[s]);
                        }
                    }
                }
            }
        };

        for (;;) {
            // ---- what other warps / rows removed meanwhile (refreshed every 8th prefix) -------------------
            if ((nprefix & 7u) == 0) {
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    if (a < d) {
                        if (HET) {
                            const int Y = s_adj[a];
                            const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (Y >> 5)]);
                            alive[s] = (wbits >> (Y & 31)) & 1u;
                        } else {
                            alive[s] = true;
                            const u64 g = __ldcg(&p.minkey[row0 + a]);
                            mk[s] = g < mk[s] ? g : mk[s];
                        }
                    }
                }
            }
            nprefix++;
            // ---- (re)build the elimination rows changed..K-1 (one jump into the unrolled sequence) ----------
            switch (changed) {
                case 0: rebuild(std::integral_constant<int, 0>{}); [[fallthrough]];
                case 1: rebuild(std::integral_constant<int, 1>{}); [[fallthrough]];
                case 2: rebuild(std::integral_constant<int, 2>{}); [[fallthrough]];
                case 3: rebuild(std::integral_constant<int, 3>{}); [[fallthrough]];
                case 4: rebuild(std::integral_constant<int, 4>{}); [[fallthrough]];
                case 5: rebuild(std::integral_constant<int, 5>{}); [[fallthrough]];
                case 6: rebuild(std::integral_constant<int, 6>{}); [[fallthrough]];
                case 7: rebuild(std::integral_constant<int, 7>{}); [[fallthrough]];
                case 8: rebuild(std::integral_constant<int, 8>{}); [[fallthrough]];
                case 9: rebuild(std::integral_constant<int, 9>{}); [[fallthrough]];
                case 10: rebuild(std::integral_constant<int, 10>{}); [[fallthrough]];
                case 11: rebuild(std::integral_constant<int, 11>{}); [[fallthrough]];
                default: rebuild(std::integral_constant<int, 12>{});
            }
            // ---- prefix quantities: A[y] = M[x][y | P], B[y] = M[y][y | P], and the per-column leaf factors ------
            // The leaf formulas are used multiplied through by B[z] (rho^2 is a ratio, B[z]^2 cancels), so that no
            // division is needed:  h01' = A[y] B[z] - A[z] u[y],  h11' = B[y] B[z] - u[y]^2,  h00' = A[x] B[z] - A[z]^2;
            // a dropped pivot (|B[z]| < CI_PIVOT_TOL) becomes B[z] = 1, A[z] = 0, u = 0.
            double A[NS], B[NS];
            bool ok[NS];
            {
                const double cx = Tw[d * KP + (K - 1)];
                unsigned long long inPmask = 0;
#pragma unroll
                for (int i = 0; i < K; i++) inPmask |= 1ull << pos[i];
                int tP = -2147483647 - 1;
                if (use_time) {
#pragma unroll
                    for (int i = 0; i < K; i++) tP = max(tP, s_time[pos[i]]);
                }
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    const double t = Tw[a * KP + (K - 1)];
                    A[s] = fma(-cx, W[K - 1][s], Ac[s]);
                    B[s] = fma(-t, W[K - 1][s], Bc[s]);
                    // the time rule (src/hetcor-cuPC-S.cu:3055-3065): max over S <= max(time X, time Y); the prefix part here
                    ok[s] = alive[s] && !((inPmask >> a) & 1ull) && (a < d) && (tP <= tmaxcol[s]);
                }
                double rxx;
                if (NS > 1 && xhi)
                    rxx = __shfl_sync(FULLMASK, A[NS - 1], xl);
                else
                    rxx = __shfl_sync(FULLMASK, A[0], xl);
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    const bool dropz = fabs(B[s]) < CI_PIVOT_TOL;
                    const double Bz = dropz ? 1.0 : B[s], Az = dropz ? 0.0 : A[s];
                    *(double2 *)(Pw + a * 4) = make_double2(Bz, Az);
                    *(double2 *)(Pw + a * 4 + 2) = make_double2(fabs(fma(rxx, Bz, -(Az * Az))), dropz ? 0.0 : 1.0);
                }
            }
            long long SP = 0;  // GEN: pairs inside the prefix and prefix x X
            long long NY[NS];  // GEN: N[y][x] + sum_j N[s_j][y]
#pragma unroll
            for (int s = 0; s < NS; s++) NY[s] = 0;
            if (GEN) {
                // the mean over all pairs of {X, Y} + S of trunc(N) (src/hetcor-cuPC-S.cu:3068-3089); every addend is
                // an integer below 2^31 and there are at most 120 of them: exact in 64-bit integers, any order
#pragma unroll
                for (int j = 0; j < K; j++) {
                    const int *rowj = s_N + pos[j] * NC;
                    SP += rowj[d];
#pragma unroll
                    for (int i = 0; i < j; i++) SP += rowj[pos[i]];
#pragma unroll
                    for (int s = 0; s < NS; s++) NY[s] += rowj[lane + 32 * s];
                }
#pragma unroll
                for (int s = 0; s < NS; s++) NY[s] += s_N[d * NC + lane + 32 * s];
            }
            __syncwarp();  // P table complete

            // ---- leaves: the last element z = zstart .. zend-1 -----------------------------------------------
            const u64 remain = Rw1 - Rbase;
            const int zend = (remain >= (u64)(d - zfirst)) ? d : zfirst + (int)remain;
            unsigned ntests = 0;
            const double *Mz = s_M + zstart * NC + lane;
            const double *Tz = Tw + zstart * KP;
            const double *Pz = Pw + zstart * 4;
            for (int z = zstart; z < zend; z++, Mz += NC, Tz += KP, Pz += 4) {
                double u[NS];
#pragma unroll
                for (int s = 0; s < NS; s++) u[s] = Mz[32 * s];
#pragma unroll
                for (int i = 0; i + 1 < K; i += 2) {
                    const double2 c = *(const double2 *)(Tz + i);
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = fma(-c.y, W[i + 1][s], fma(-c.x, W[i][s], u[s]));
                }
                if (K & 1) {
                    const double c = Tz[K - 1];
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = fma(-c, W[K - 1][s], u[s]);
                }
                const double2 pz = *(const double2 *)Pz;        // B[z], A[z]
                const double2 qz = *(const double2 *)(Pz + 2);  // |h00'|, 0 when the pivot z is dropped
                if (qz.y == 0.0) {
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = 0.0;
                }
                const int tz = use_time ? s_time[z] : -2147483647 - 1;
                long long SPz = 0;
                if (GEN) {
                    const int *rowz = s_N + z * NC;
                    SPz = SP + rowz[d];
#pragma unroll
                    for (int j = 0; j < K; j++) SPz += rowz[pos[j]];
                }
                bool valid[NS], rem[NS];
                bool hit = false;
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    valid[s] = ok[s] && (a != z) && (tz <= tmaxcol[s]);
                    if (!HET) valid[s] = valid[s] && !(mk[s] < Rbase + (u64)(z - zfirst));
                    const double syz = u[s];
                    const double h01 = fma(A[s], pz.x, -(pz.y * syz));
                    const double h11 = fma(B[s], pz.x, -(syz * syz));
                    const double lhs = h01 * h01;
                    const double den = qz.x * fabs(h11);
                    if (GEN) {
                        rem[s] = false;
                        if (valid[s]) {
                            const long long sum = SPz + NY[s] + (long long)s_N[z * NC + a];
                            const double m = (double)sum / npairs;
                            const double th = p.th_het / sqrt(m - (double)l - 3.0);
                            const double t = tanh(th);
                            const double T2 = t * t;
                            rem[s] = (lhs < T2 * den) || (den > 0.0 && lhs * T2 > den);
                        }
                    } else {
                        // |rho| < tanh(th)  or, on the |rho| > 1 branch, |rho| > coth(th)
                        rem[s] = (lhs < T2u * den) || (den > 0.0 && lhs * T2u > den);
                    }
                    ntests += valid[s] ? 1u : 0u;
                    hit |= valid[s] && rem[s];
                }
                if (__any_sync(FULLMASK, hit)) {
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        if (valid[s] && rem[s]) {
                            const int a = lane + 32 * s;
                            const int Y = s_adj[a];
                            if (HET) {
                                atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                                atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                                if (p.removed) p.removed[row0 + a] = 1;
                                alive[s] = false;
                                ok[s] = false;
                            } else {
                                const u64 R = Rbase + (u64)(z - zfirst);
                                atomicMin(&p.minkey[row0 + a], R);
                                mk[s] = R < mk[s] ? R : mk[s];
                            }
                        }
                    }
                }
            }
            my_tests += ntests;

            // ---- next prefix ---------------------------------------------------------------------------------
            Rbase += (u64)(d - zfirst);
            if (Rbase >= Rw1) break;
            int ii = -1;
#pragma unroll
            for (int j = 0; j < K; j++)
                if (pos[j] < d - 1 - K + j) ii = j;
            if (ii < 0) break;
#pragma unroll
            for (int j = 0; j < K; j++) {
                if (j == ii)
                    pos[j]++;
                else if (j > ii && j > 0)
                    pos[j] = pos[j - 1] + 1;
            }
            changed = ii;
            zfirst = pos[K - 1] + 1;
            zstart = zfirst;
            __syncwarp();  // the leaves' reads of T and P are done before the next prefix rewrites them
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

template <int MODE, int LT, int NS>
static int cd_launch(cudaStream_t st, const CiParams &p)
{
    constexpr int dyn = (int)CdLayout<LT, NS, MODE == 2>::bytes;
    // per instantiation AND per device (see ci_small.cu)
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        if (cudaFuncSetAttribute(ci_dfs_kernel<MODE, LT, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn) !=
            cudaSuccess) {
            cusk_set_error("cudaFuncSetAttribute(ci_dfs_kernel<%d,%d,%d>, %d B): %s", MODE, LT, NS, dyn,
                           cudaGetErrorString(cudaGetLastError()));
            return CUSK_E_CUDA;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_dfs_kernel<MODE, LT, NS>, CD_NTHREADS, dyn) !=
                cudaSuccess ||
            nb < 1)
            nb = 1;
        ctas_per_sm = nb;
        ctas_of_device[dev].store(nb, std::memory_order_release);
    }
    launch_k(ci_dfs_kernel<MODE, LT, NS>, dim3((p.sm_count > 0 ? p.sm_count : 148) * ctas_per_sm), dim3(CD_NTHREADS), dyn, st, p);
    return CUSK_OK;
}

template <int MODE, int NS>
static int cd_dispatch(cudaStream_t st, const CiParams &p)
{
    switch (p.l) {
        case 3: return cd_launch<MODE, 3, NS>(st, p);
        case 4: return cd_launch<MODE, 4, NS>(st, p);
        case 5: return cd_launch<MODE, 5, NS>(st, p);
        case 6: return cd_launch<MODE, 6, NS>(st, p);
        case 7: return cd_launch<MODE, 7, NS>(st, p);
        case 8: return cd_launch<MODE, 8, NS>(st, p);
        case 9: return cd_launch<MODE, 9, NS>(st, p);
        case 10: return cd_launch<MODE, 10, NS>(st, p);
        case 11: return cd_launch<MODE, 11, NS>(st, p);
        case 12: return cd_launch<MODE, 12, NS>(st, p);
        case 13: return cd_launch<MODE, 13, NS>(st, p);
        case 14: return cd_launch<MODE, 14, NS>(st, p);
        default: return CUSK_OK;
    }
}

// wide = the level may hold rows with more than 31 neighbours in this kernel's class (two columns per lane)
int launch_ci_dfs(cudaStream_t st, bool het, const CiParams &p, bool wide)
{
    const int mode = !het ? 0 : (p.N ? 2 : 1);
    if (wide) return mode == 0 ? cd_dispatch<0, 2>(st, p) : (mode == 1 ? cd_dispatch<1, 2>(st, p) : cd_dispatch<2, 2>(st, p));
    return mode == 0 ? cd_dispatch<0, 1>(st, p) : (mode == 1 ? cd_dispatch<1, 1>(st, p) : cd_dispatch<2, 1>(st, p));
}
