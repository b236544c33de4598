// ci_mid.cu -- levels 1 and 2 for the rows of medium degree (CI_SMALL_DMAX < d <= 1024 / 160): one CTA
// per work item, one thread per (conditioning set, 32-neighbour chunk) (replaces cal_Indepl1/2,
// src/cuPC-S.cu:486-760, and cal_Indepl1_ess/2_ess, src/hetcor-cuPC-S.cu:379-640; sm_100a).
//
// At levels 1-2 the graph is still dense (an LD block's markers are nearly all marginally dependent:
// d ~ 40-150) and almost all of a skeleton's tests are here.  A test needs l + 2 correlations and ~10
// FP64 operations, so what matters is that every lane carries a test and that nothing waits on a
// factorisation: with l <= 2 the LDL^T of C[S,S] is a couple of operations (D_0 = 1, L_10 = C[s1,s0],
// D_1 = 1 - L_10^2) and is simply recomputed by each thread.
//   * level 1: every entry C[y, z] of the row's neighbourhood submatrix is used by exactly one test, so
//     nothing is staged: adjacent lanes hold adjacent z and gather C[adj(y), adj(z)] straight from L2
//     (sorted neighbours -> a few sectors per warp load), four loads in flight per thread.
//   * level 2: the submatrix is staged in shared memory once per item (each entry serves ~d tests).
// Quantities, pivot rule and decision are those of ci_small.cu / ci_level.cu.
#include <atomic>

#include "skeleton.cuh"

#define CM_NTHREADS 256
#define FULLMASK 0xffffffffu

// Level 1 of the het engine with one sample size for all pairs (the cuskss case): the decision
//   (r_xy - r_yz r_xz)^2 < T2 (1 - r_xz^2)(1 - r_yz^2)
// is taken in FP32 when it is not close -- every FP32 quantity here is one or two correctly rounded operations away from
// the exact value (the FMAs round once), so a relative margin of 2e-6 on both sides leaves the FP64 expression of the
// general loop only the tests within that band (same decisions everywhere, ~40 % fewer instructions per test).  One
// 16-byte shared-memory entry per neighbour y: {adj(y) n, latest time a conditioning variable may have (INT_MIN: edge
// gone), r_xy, adj(y)}.
__device__ __forceinline__ bool level1_exact(float cxy, float cxz, float cyz, double T2u)
{
    const double tx0 = (double)cxz, ty0 = (double)cyz;
    const double h01 = (double)cxy - ty0 * tx0;
    const double h11 = 1.0 - ty0 * ty0;
    const double c2 = fabs(1.0 - tx0 * tx0);
    const double lhs = h01 * h01;
    const double den = c2 * fabs(h11);
    return (lhs < T2u * den) || (den > 0.0 && lhs * T2u > den);
}

__device__ __forceinline__ u64 level1_fast(const CiParams &p, int X, int d, int row0, unsigned R0, unsigned R1, int tX,
                                           bool use_time, double T2u, unsigned char *smem)
{
    constexpr int INT_LOW = -2147483647 - 1;
    const int tid = threadIdx.x;
    const int n = p.n;
    int4 *s_t = (int4 *)smem;                     // CI_MID_DMAX1 entries
    int *s_tz = (int *)(s_t + CI_MID_DMAX1);      // time of each neighbour (as a conditioning variable)
    for (int i = tid; i < d; i += CM_NTHREADS) {
        const int v = p.adj[row0 + i];
        const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (v >> 5)]);
        const bool gone = !((wbits >> (v & 31)) & 1u);
        const int tv = use_time ? p.time_index[v] : 0;
        int4 t;
        t.x = (int)((unsigned)v * (unsigned)n);
        t.y = gone ? INT_LOW : (use_time ? max(tX, tv) : 2147483647);
        t.z = __float_as_int(p.C[(size_t)X * n + v]);
        t.w = v;
        s_t[i] = t;
        s_tz[i] = tv;
    }
    __syncthreads();
    const float T2f = (float)T2u;
    unsigned tests = 0;
    const unsigned nranks = R1 - R0;
    const unsigned nych = (unsigned)(d + 31) >> 5;
    const unsigned units = nranks * nych;
    const float *__restrict__ C = p.C;
    for (unsigned u = tid; u < units; u += CM_NTHREADS) {
        const unsigned ych = u / nranks;
        const int z = (int)(R0 + (u - ych * nranks));
        const int y0 = (int)ych * 32, y1 = min(d, y0 + 32);
        const int4 tz = s_t[z];
        const unsigned adjz = (unsigned)tz.w;
        const float cxz = __int_as_float(tz.z);
        const int tS = s_tz[z];
        const float c2 = fabsf(fmaf(-cxz, cxz, 1.0f));
        const float k = T2f * c2;
        const float k_lo = k * (1.0f - 2e-6f), k_hi = k * (1.0f + 2e-6f);
        const float k2 = (c2 / T2f) * (1.0f - 2e-6f);  // lhs T2 > den cannot hold below this multiple of |h11|
        for (int yb = y0; yb < y1; yb += 4) {
            int4 ty[4];
            float cyz[4];
            bool live[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int y = min(yb + q, y1 - 1);
                ty[q] = s_t[y];
                live[q] = (yb + q < y1) && (y != z) && (tS <= ty[q].y);
                cyz[q] = 0.f;
                if (live[q]) cyz[q] = __ldg(C + ((unsigned)ty[q].x + adjz));
            }
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (!live[q]) continue;
                tests++;
                const float cxy = __int_as_float(ty[q].z);
                const float a = fmaf(-cyz[q], cxz, cxy);
                const float ah = fabsf(fmaf(-cyz[q], cyz[q], 1.0f));
                const float lhs = a * a;
                bool rem;
                if (lhs < k_lo * ah)
                    rem = true;
                else if (lhs > k_hi * ah && lhs < k2 * ah)
                    rem = false;
                else
                    rem = level1_exact(cxy, cxz, cyz[q], T2u);
                if (rem) {
                    const int Y = ty[q].w;
                    atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                    atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                    if (p.removed) p.removed[row0 + yb + q] = 1;
                    s_t[yb + q].y = INT_LOW;
                }
            }
        }
    }
    return (u64)tests;
}

template <bool HET, int LT>
__global__ void __launch_bounds__(CM_NTHREADS) ci_mid_kernel(const CiParams p)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_item;
    LevelCtl *ctl = p.ctl;
    pdl_enter();
    if (ctl->done) return;
    // edge-sharded mode: the level plan decides (identically on every rank) whether this level's items are dealt out
    const int srank = ctl->shard_on ? p.shard_rank : 0, sworld = ctl->shard_on ? p.shard_world : 1;
    const int tid = threadIdx.x, lane = tid & 31;
    constexpr int l = LT;
    const int n = p.n;
    const int nitems = ctl->nitems_mid;
    const double T2u = ctl->T2;
    const bool het_general = HET && p.N != nullptr;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    // shared memory: adj, cx, time, skip (int each, DM), minkey (u64), then csub (level 2)
    constexpr int DM = (LT == 1) ? CI_MID_DMAX1 : CI_MID_DMAX;
    int *s_adj = (int *)smem;
    float *s_cx = (float *)(s_adj + DM);
    int *s_time = (int *)(s_cx + DM);
    int *s_skip = s_time + DM;
    unsigned long long *s_minkey = (unsigned long long *)(s_skip + DM);
    float *csub = (float *)(s_minkey + DM);
    u64 my_tests = 0;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_item = srank + sworld * atomicAdd(p.work_counter_mid, 1);
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
        const unsigned R0 = (unsigned)r2.x, R1 = (unsigned)r2.y;
        const int tX = use_time ? p.time_index[X] : 0;
        const int stride = d | 1;
        if constexpr (HET && LT == 1) {
            if (!het_general && (unsigned)n < 65536u) {
                my_tests += level1_fast(p, X, d, row0, R0, R1, tX, use_time, T2u, smem);
                continue;
            }
        }
        for (int i = tid; i < d; i += CM_NTHREADS) {
            const int v = p.adj[row0 + i];
            s_adj[i] = v;
            s_cx[i] = p.C[(size_t)X * n + v];
            if (use_time) s_time[i] = p.time_index[v];
            if (HET) {
                const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (v >> 5)]);
                s_skip[i] = !((wbits >> (v & 31)) & 1u);
            } else {
                s_minkey[i] = __ldcg(&p.minkey[row0 + i]);
            }
        }
        __syncthreads();
        if (LT == 2) {
            const int total = d * d;
            for (int base = tid; base < total; base += CM_NTHREADS * 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CM_NTHREADS;
                    v[u] = 1.0f;
                    if (idx < total) {
                        const int i = idx / d, j = idx - i * d;
                        if (i != j) v[u] = __ldg(&p.C[(size_t)s_adj[i] * n + s_adj[j]]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CM_NTHREADS;
                    if (idx < total) {
                        const int i = idx / d, j = idx - i * d;
                        csub[i * stride + j] = v[u];
                    }
                }
            }
            __syncthreads();
        }
        const unsigned nranks = R1 - R0;
        const unsigned nych = (unsigned)(d + 31) >> 5;
        const unsigned units = nranks * nych;
        for (unsigned u = tid; u < units; u += CM_NTHREADS) {
            const unsigned ych = u / nranks;
            const unsigned R = R0 + (u - ych * nranks);
            const int y0 = (int)ych * 32, y1 = min(d, y0 + 32);
            // ---- the set and its factor -----------------------------------------------------------
            int p0, p1 = -1;
            if (LT == 1) {
                p0 = (int)R;
            } else {
                // lexicographic rank of {p0 < p1}: d - 1 - p0 sets start with p0
                unsigned rank = R;
                int x = 0;
                for (;; x++) {
                    const unsigned c = (unsigned)(d - 1 - x);
                    if (rank < c) break;
                    rank -= c;
                }
                p0 = x;
                p1 = x + 1 + (int)rank;
            }
            const double tx0 = (double)s_cx[p0];
            double h00 = 1.0 - tx0 * tx0;
            double l10 = 0.0, invD1 = 0.0, txd1 = 0.0;
            int tS = use_time ? s_time[p0] : 0;
            if (LT == 2) {
                l10 = (double)csub[p1 * stride + p0];
                const double D1 = 1.0 - l10 * l10;
                invD1 = (fabs(D1) < CI_PIVOT_TOL) ? 0.0 : 1.0 / D1;
                const double tx1 = (double)s_cx[p1] - l10 * tx0;
                txd1 = tx1 * invD1;
                h00 -= tx1 * txd1;
                if (use_time) tS = max(tS, s_time[p1]);
            }
            const double c2 = fabs(h00);
            double SP = 0.0;
            if (het_general) {
                const size_t row0n = (size_t)s_adj[p0] * n;
                SP = (double)__float2int_rz(p.N[row0n + X]);
                if (LT == 2) {
                    const size_t row1n = (size_t)s_adj[p1] * n;
                    SP += (double)__float2int_rz(p.N[row1n + X]) + (double)__float2int_rz(p.N[row1n + s_adj[p0]]);
                }
            }
            const float *ccol = p.C + s_adj[p0];  // level 1: C[adj(y), adj(z)] read from L2
            for (int yb = y0; yb < y1; yb += 4) {
                float cy0[4];
                bool live[4];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int y = yb + q;
                    live[q] = (y < y1) && (y != p0) && (y != p1);
                    if (live[q]) {
                        if (HET) {
                            if (s_skip[y]) live[q] = false;
                            if (use_time && tS > max(tX, s_time[y])) live[q] = false;
                        } else {
                            if (s_minkey[y] < (unsigned long long)R) live[q] = false;
                        }
                    }
                    cy0[q] = 0.f;
                    if (live[q]) cy0[q] = (LT == 1) ? __ldg(ccol + (size_t)s_adj[y] * n) : csub[y * stride + p0];
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    if (!live[q]) continue;
                    const int y = yb + q;
                    my_tests++;
                    const double ty0 = (double)cy0[q];
                    double h01 = (double)s_cx[y] - ty0 * tx0;
                    double h11 = 1.0 - ty0 * ty0;
                    if (LT == 2) {
                        const double ty1 = (double)csub[y * stride + p1] - l10 * ty0;
                        h01 -= ty1 * txd1;
                        h11 -= ty1 * ty1 * invD1;
                    }
                    const double lhs = h01 * h01;
                    const double den = c2 * fabs(h11);
                    double T2 = T2u;
                    if (het_general) {
                        const int Y = s_adj[y];
                        double sum = SP + (double)__float2int_rz(p.N[(size_t)Y * n + X]) +
                                     (double)__float2int_rz(p.N[(size_t)s_adj[p0] * n + Y]);
                        if (LT == 2) sum += (double)__float2int_rz(p.N[(size_t)s_adj[p1] * n + Y]);
                        const double m = sum / npairs;
                        const double th = p.th_het / sqrt(m - (double)l - 3.0);
                        const double t = tanh(th);
                        T2 = t * t;
                    }
                    // |rho| < tanh(th)  or, on the |rho| > 1 branch, |rho| > coth(th)
                    const bool rem = (lhs < T2 * den) || (den > 0.0 && lhs * T2 > den);
                    if (rem) {
                        const int Y = s_adj[y];
                        if (HET) {
                            atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                            atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                            if (p.removed) p.removed[row0 + y] = 1;
                            s_skip[y] = 1;
                        } else {
                            atomicMin(&p.minkey[row0 + y], (u64)R);
                            atomicMin(&s_minkey[y], (unsigned long long)R);
                        }
                    }
                }
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) my_tests += __shfl_xor_sync(FULLMASK, my_tests, o);
    if (lane == 0 && my_tests) atomicAdd((unsigned long long *)&ctl->executed, (unsigned long long)my_tests);
}

template <bool HET, int LT>
static int cm_launch(cudaStream_t st, const CiParams &p)
{
    constexpr int base = ((LT == 1) ? CI_MID_DMAX1 : CI_MID_DMAX) * (4 + 4 + 4 + 4 + 8);
    constexpr int dyn = base + (LT == 2 ? CI_MID_DMAX * (CI_MID_DMAX | 1) * 4 : 0);
    // per instantiation AND per device (see ci_small.cu)
    static std::atomic<int> ctas_of_device[CUSK_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = (dev >= 0 && dev < CUSK_MAX_DEVICES) ? dev : 0;
    int ctas_per_sm = ctas_of_device[dev].load(std::memory_order_acquire);
    if (!ctas_per_sm) {
        if (cudaFuncSetAttribute(ci_mid_kernel<HET, LT>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn) != cudaSuccess) {
            cusk_set_error("cudaFuncSetAttribute(ci_mid_kernel<%d>, %d B): %s", LT, dyn,
                           cudaGetErrorString(cudaGetLastError()));
            return CUSK_E_CUDA;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ci_mid_kernel<HET, LT>, CM_NTHREADS, dyn) != cudaSuccess ||
            nb < 1)
            nb = 1;
        ctas_per_sm = nb;
        ctas_of_device[dev].store(nb, std::memory_order_release);
    }
    launch_k(ci_mid_kernel<HET, LT>, dim3((p.sm_count > 0 ? p.sm_count : 148) * ctas_per_sm), dim3(CM_NTHREADS), dyn, st, p);
    return CUSK_OK;
}

int launch_ci_mid(cudaStream_t st, bool het, const CiParams &p)
{
    if (p.l == 1) return het ? cm_launch<true, 1>(st, p) : cm_launch<false, 1>(st, p);
    if (p.l == 2) return het ? cm_launch<true, 2>(st, p) : cm_launch<false, 2>(st, p);
    return CUSK_OK;
}
