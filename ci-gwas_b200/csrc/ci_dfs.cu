// ci_dfs.cu -- level-l (l >= 2) CI tests of the rows of medium degree (up to CI_DFS_DMAX = 63 neighbours) whose
// conditioning sets each serve many tests (d - l large): one WARP walks the conditioning sets of its rank range in
// lexicographic order as a depth-first search, with the LANES spread over the row's neighbours (replaces
// cal_Indepl3..14, src/cuPC-S.cu:716-3020, and cal_Indepl3_ess..14_ess, src/hetcor-cuPC-S.cu:632-3052, for those
// rows; sm_100a).
//
// Why a third mapping.  ci_small.cu (a thread per set) spends l^2/2 FMAs per test on the forward substitution
// t_y = L^-1 C[S,y] and needs the whole factor per thread (172-255 registers for l >= 7: 9-12 % of the warp slots
// busy); ci_level.cu (a CTA per row, incremental factorisation in shared memory, 16 warps per SM) is built for hub
// rows with hundreds of neighbours.  Rows with 33..63 neighbours at l >= 3 -- traits with a few dozen associated
// markers, the rows that carry almost all tests of a dense block -- fell to ci_level.cu and ran at 5 G tests/s.
//
// Here the elimination is right-looking on the (d+1) x (d+1) matrix M = C[{adj(X), X}, {adj(X), X}]: after the
// pivots s_0..s_{j-1} the row  w_j[a] = M[s_j][a] - sum_{i<j} w_i[s_j] w_i[a] / D_i  (D_i = w_i[s_i]) is kept
// normalised, W_j[a] = w_j[a] / sqrt|D_j| with the sign of D_j in a bit mask, ONE REGISTER PER DEPTH AND COLUMN: lane
// (a mod 32) holds column a (two columns per lane when d + 1 > 32).  A new row costs j FMAs per column; the
// coefficients W_i[s_j] (held by the lane of column s_j) reach the other lanes through a 128-byte shared-memory
// buffer (one predicated store, broadcast loads).  Consecutive sets share their leading pivots, so only the rows from
// the first changed depth on are rebuilt.  With the prefix P = S minus its last element z fixed,
//     A[y] = M[x][y | P],  B[y] = M[y][y | P]   (once per prefix),   u[y] = M[z][y | P]   (l - 1 FMAs per leaf z)
// give the test (X, y | P + z) by the first-order formula
//     h01 = A[y] - A[z] u[y] / B[z],  h11 = B[y] - u[y]^2 / B[z],  h00 = A[x] - A[z]^2 / B[z],
// i.e. ~l + 10 FP64 operations per test instead of ~l^2/2 + 3l, and 2 (l - 1) + 8 FP64 registers per lane.
// Same quantities, pivot rule (|D| < CI_PIVOT_TOL drops the direction) and decision rho^2 < tanh(th)^2 [or >
// coth(th)^2] as ci_small.cu / ci_level.cu; PC-stable semantics as there (sets from the adjacency compacted at level
// start; removals clear bits or lower the slot's minimum rank).
//
// A work item (row, rank range; class 2 of the level plan at l >= 3) goes to a CTA of 4 warps, which stages M once
// (<= 64 x 64 floats) and splits the ranks evenly.  Rows with up to 31 neighbours run on the NS = 1 instantiation (one
// column per lane), rows with 32..63 on NS = 2; both walk the level's one item table with a queue each and skip the
// other's rows.  Algorithmic bytes per test: 4 (l + 2).
#include <atomic>

#include "skeleton.cuh"

#define CD_NWARPS 4
#define CD_NTHREADS (CD_NWARPS * 32)
#define FULLMASK 0xffffffffu

// MODE 0: individual-level engine (minimum separating-set rank per slot); 1: summary-statistic engine with one
// sample size for every pair; 2: the same with per-pair effective sample sizes N (own instantiation: its gathers
// cost ~80 registers that the other two should not carry)
template <int MODE, int LT, int NS>
__global__ void __launch_bounds__(CD_NTHREADS) ci_dfs_kernel(const CiParams p)
{
    constexpr bool HET = MODE != 0;
    constexpr int NC = 32 * NS;  // columns: neighbours 0..d-1, X at column d, zero padding above
    constexpr int K = LT - 1;    // prefix length
    static_assert(LT >= 2 && LT <= CUSK_ML, "levels 2..14");
    __shared__ float s_M[NC * NC];
    __shared__ int s_adj[NC];
    __shared__ int s_time[NC];
    __shared__ __align__(16) double s_cb[CD_NWARPS][16];
    __shared__ int s_item;
    LevelCtl *ctl = p.ctl;
    if (ctl->done) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int l = LT;
    const int n = p.n;
    const int nitems = ctl->nitems_mid;
    const double T2u = ctl->T2;
    constexpr bool het_general = MODE == 2;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    double *cb = s_cb[warp];
    u64 my_tests = 0;
    unsigned my_sing = 0;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_item = p.shard_rank + p.shard_world * atomicAdd(NS == 1 ? p.work_counter_mid : p.work_counter_dfs2, 1);
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

        // ---- stage the row: columns, then M ------------------------------------------------------------
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
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CD_NTHREADS;
                    v[u] = 0.0f;
                    if (idx < total) {
                        const int r = idx / NC, a = idx - r * NC;
                        if (a <= d) v[u] = (r == a) ? 1.0f : __ldg(&p.C[(size_t)s_adj[r] * n + s_adj[a]]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CD_NTHREADS;
                    if (idx < total) s_M[idx] = v[u];
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
        int zfirst = (K > 0) ? pos[K - 1] + 1 : 0;
        u64 Rbase = Rw0 - (u64)(zstart - zfirst);
        int changed = 0;
        unsigned negmask = 0;
        unsigned nprefix = 0;
        double W[K > 0 ? K : 1][NS];
        // per-column state of this lane
        bool alive[NS];
        u64 mk[NS];
        int tcol[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) {
            alive[s] = false;
            mk[s] = CUSK_KEY_NONE;
            tcol[s] = use_time ? s_time[lane + 32 * s] : 0;
        }
        const int xl = d & 31, xs = d >> 5;  // lane and slot of column X

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
            // ---- (re)build the elimination rows changed..K-1 ----------------------------------------------
#pragma unroll
            for (int j = 0; j < K; j++) {
                if (j >= changed) {
                    const int pj = pos[j];
                    if (lane == (pj & 31)) {
#pragma unroll
                        for (int i = 0; i < j; i++) {
                            double v = (NS > 1 && (pj >> 5)) ? W[i][NS - 1] : W[i][0];
                            if ((negmask >> i) & 1u) v = -v;
                            cb[i] = v;
                        }
                    }
                    __syncwarp();
                    double u[NS];
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = (double)s_M[pj * NC + lane + 32 * s];
#pragma unroll
                    for (int i = 0; i < j; i++) {
                        const double c = cb[i];
#pragma unroll
                        for (int s = 0; s < NS; s++) u[s] = fma(-c, W[i][s], u[s]);
                    }
                    const double dsel = (NS > 1 && (pj >> 5)) ? u[NS - 1] : u[0];
                    const double D = __shfl_sync(FULLMASK, dsel, pj & 31);
                    const bool drop = fabs(D) < CI_PIVOT_TOL;
                    const double sc = drop ? 0.0 : rsqrt(fabs(D));
                    if (D < 0 && !drop)
                        negmask |= (1u << j);
                    else
                        negmask &= ~(1u << j);
                    if (drop && lane == 0) my_sing++;
#pragma unroll
                    for (int s = 0; s < NS; s++) W[j][s] = u[s] * sc;
                    __syncwarp();
                }
            }
            // ---- prefix quantities: A[y] = M[x][y | P], B[y] = M[y][y | P] -----------------------------------
            if (lane == xl) {
#pragma unroll
                for (int i = 0; i < K; i++) {
                    double v = (NS > 1 && xs) ? W[i][NS - 1] : W[i][0];
                    if ((negmask >> i) & 1u) v = -v;
                    cb[i] = v;
                }
            }
            __syncwarp();
            double A[NS], B[NS];
            bool inP[NS];
#pragma unroll
            for (int s = 0; s < NS; s++) {
                A[s] = (double)s_M[d * NC + lane + 32 * s];
                B[s] = 1.0;
                inP[s] = false;
            }
            int tP = -2147483647 - 1;
#pragma unroll
            for (int i = 0; i < K; i++) {
                const double c = cb[i];
                const bool ng = (negmask >> i) & 1u;
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const double t = W[i][s];
                    A[s] = fma(-c, t, A[s]);
                    B[s] = fma(ng ? t : -t, t, B[s]);
                    inP[s] |= (pos[i] == lane + 32 * s);
                }
                if (use_time) tP = max(tP, s_time[pos[i]]);
            }
            const double rxx = __shfl_sync(FULLMASK, (NS > 1 && xs) ? A[NS - 1] : A[0], xl);
            double SP = 0.0;
            double NY[NS];
#pragma unroll
            for (int s = 0; s < NS; s++) NY[s] = 0.0;
            if (het_general) {
                // pairs inside the prefix and prefix x X (src/hetcor-cuPC-S.cu:3068-3089: N[later * n + earlier], the
                // tuple ordered X, Y, S_1..S_l); every addend is an integer below 2^31, so any order gives the same sum
#pragma unroll
                for (int j = 0; j < K; j++) {
                    const size_t rowj = (size_t)s_adj[pos[j]] * n;
                    SP += (double)__float2int_rz(__ldg(&p.N[rowj + X]));
#pragma unroll
                    for (int i = 0; i < j; i++) SP += (double)__float2int_rz(__ldg(&p.N[rowj + s_adj[pos[i]]]));
                }
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    if (a < d) {
                        const int Y = s_adj[a];
                        double sy = (double)__float2int_rz(__ldg(&p.N[(size_t)Y * n + X]));
#pragma unroll
                        for (int j = 0; j < K; j++)
                            sy += (double)__float2int_rz(__ldg(&p.N[(size_t)s_adj[pos[j]] * n + Y]));
                        NY[s] = sy;
                    }
                }
            }
            __syncwarp();

            // ---- leaves: the last element z = zstart .. zend-1 -----------------------------------------------
            const u64 remain = Rw1 - Rbase;
            const int zend = (remain >= (u64)(d - zfirst)) ? d : zfirst + (int)remain;
            for (int z = zstart; z < zend; z++) {
                const int zl = z & 31, zs = z >> 5;
                if (lane == zl) {
#pragma unroll
                    for (int i = 0; i < K; i++) {
                        double v = (NS > 1 && zs) ? W[i][NS - 1] : W[i][0];
                        if ((negmask >> i) & 1u) v = -v;
                        cb[i] = v;
                    }
                    cb[K] = (NS > 1 && zs) ? A[NS - 1] : A[0];
                    cb[K + 1] = (NS > 1 && zs) ? B[NS - 1] : B[0];
                }
                __syncwarp();
                double u[NS];
#pragma unroll
                for (int s = 0; s < NS; s++) u[s] = (double)s_M[z * NC + lane + 32 * s];
#pragma unroll
                for (int i = 0; i < K; i++) {
                    const double c = cb[i];
#pragma unroll
                    for (int s = 0; s < NS; s++) u[s] = fma(-c, W[i][s], u[s]);
                }
                const double xz = cb[K], Dz = cb[K + 1];
                __syncwarp();
                const double iz = (fabs(Dz) < CI_PIVOT_TOL) ? 0.0 : 1.0 / Dz;
                const double g = xz * iz;
                const double c2 = fabs(rxx - xz * g);
                const int tS = use_time ? max(tP, s_time[z]) : 0;
                const u64 R = Rbase + (u64)(z - zfirst);
                double SPz = 0.0;
                if (het_general) {
                    const size_t rowz = (size_t)s_adj[z] * n;
                    SPz = SP + (double)__float2int_rz(__ldg(&p.N[rowz + X]));
#pragma unroll
                    for (int j = 0; j < K; j++) SPz += (double)__float2int_rz(__ldg(&p.N[rowz + s_adj[pos[j]]]));
                }
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    const int a = lane + 32 * s;
                    bool valid = alive[s] && !inP[s] && (a != z) && (a < d);
                    if (HET) {
                        if (use_time) valid = valid && (tS <= max(tX, tcol[s]));
                    } else {
                        valid = valid && !(mk[s] < R);
                    }
                    if (!valid) continue;
                    my_tests++;
                    const double syz = u[s];
                    const double h01 = A[s] - g * syz;
                    const double h11 = B[s] - syz * syz * iz;
                    const double lhs = h01 * h01;
                    const double den = c2 * fabs(h11);
                    double T2 = T2u;
                    if (het_general) {
                        const double sum =
                            SPz + NY[s] + (double)__float2int_rz(__ldg(&p.N[(size_t)s_adj[z] * n + s_adj[a]]));
                        const double m = sum / npairs;
                        const double th = p.th_het / sqrt(m - (double)l - 3.0);
                        const double t = tanh(th);
                        T2 = t * t;
                    }
                    // |rho| < tanh(th)  or, on the |rho| > 1 branch, |rho| > coth(th)
                    const bool rem = (lhs < T2 * den) || (den > 0.0 && lhs * T2 > den);
                    if (rem) {
                        const int Y = s_adj[a];
                        if (HET) {
                            atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                            atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                            if (p.removed) p.removed[row0 + a] = 1;
                            alive[s] = false;
                        } else {
                            atomicMin(&p.minkey[row0 + a], R);
                            mk[s] = R < mk[s] ? R : mk[s];
                        }
                    }
                }
            }

            // ---- next prefix ---------------------------------------------------------------------------------
            Rbase += (u64)(d - zfirst);
            if (K == 0 || Rbase >= Rw1) break;
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
    // per instantiation AND per device (see ci_small.cu)
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_dfs_kernel<MODE, LT, NS>, CD_NTHREADS, 0) != cudaSuccess ||
            nb < 1)
            nb = 1;
        ctas_per_sm = nb;
        ctas_of_device[dev].store(nb, std::memory_order_release);
    }
    ci_dfs_kernel<MODE, LT, NS><<<p.sm_count * ctas_per_sm, CD_NTHREADS, 0, st>>>(p);
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
