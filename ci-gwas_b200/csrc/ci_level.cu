// ci_level.cu -- the level-l (l = 1..14) conditional-independence test kernel of both skeleton
// engines (replaces cal_Indepl1..14, src/cuPC-S.cu:486-3020, and cal_Indepl1_ess..14_ess,
// src/hetcor-cuPC-S.cu:379-3052; sm_100a).
//
// Design (B200-first, not the reference's thread-per-S scheme):
//   * persistent CTAs (one per SM, 16 warps, ~227 KB shared memory) pull work items from an
//     atomic queue; an item is a contiguous range of conditioning-set ranks of ONE row X.
//   * the CTA stages X's neighbourhood correlation submatrix C[adj(X), adj(X)] in shared memory
//     once (float, d x d), so that the inner loops never touch global memory.
//   * every warp walks its rank range in lexicographic order as (prefix P of l-1 positions,
//     last element z).  For the prefix it maintains an incremental LDL^T factorisation over ALL
//     neighbours: rows W[j][a] = j-th elimination vector, so that the conditional covariance
//     Sigma[a,b | P] = C[a,b] - sum_j sg_j W[j][a] W[j][b] costs one length-(l-1) dot product.
//     Consecutive prefixes share their leading rows; only the changed depths are rebuilt.
//   * the test (X,Y | P + z) is then the first-order formula on Sigma[. | P]:
//        h01 = Sxy - Sxz Syz / Szz, h00 = Sxx - Sxz^2 / Szz, h11 = Syy - Syz^2 / Szz
//     lanes run over Y, z is register-blocked by 4, all arithmetic is FP64.
//   * Z = |atanh(rho)| < th is decided as rho^2 < tanh(th)^2 (and rho^2 > coth(th)^2 for the
//     |rho| > 1 branch of the reference's |.| formula): no log, sqrt or division per test.
//   * PC-stable: conditioning sets come from the adjacency compacted at level start; removals
//     only clear bits (het engine) or lower a per-slot minimum rank (individual-level engine),
//     so the surviving graph and the recorded separating sets do not depend on scheduling.
//
// Algorithmic bytes per test: 4 (l + 2); with the staged submatrix the actual
// global traffic per item is 4 d^2 + 12 d.
#include "skeleton.cuh"

#define FULLMASK 0xffffffffu

template <bool HET>
__global__ void __launch_bounds__(CI_NTHREADS, 1) ci_level_kernel(const CiParams p)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ struct {
        int item, X, d, row0;
        u64 R0, R1;
    } s_it;
    LevelCtl *ctl = p.ctl;
    pdl_enter();
    if (ctl->done) return;  // (short rows are not in this kernel's item table: ci_small.cu)
    // edge-sharded mode: the level plan decides (identically on every rank) whether this level's items are dealt out
    const int srank = ctl->shard_on ? p.shard_rank : 0, sworld = ctl->shard_on ? p.shard_world : 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int l = p.l, k = l - 1, n = p.n;
    const int nitems = ctl->nitems;
    const double T2u = ctl->T2;
    const bool het_general = HET && p.N != nullptr;
    const bool use_time = HET && p.time_index != nullptr;
    const double npairs = 0.5 * (double)((l + 2) * (l + 1));
    u64 my_tests = 0;
    unsigned my_sing = 0;

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            // one thread decodes the item (row lookup, C(d,l), rank range) for the whole CTA
            const int item = srank + sworld * atomicAdd(p.work_counter, 1);
            s_it.item = item;
            if (item < nitems && item < p.item_cap) {
                const int4 a4 = *(const int4 *)&p.items[item];
                const ulonglong2 r2 = *(const ulonglong2 *)&p.items[item].R0;
                s_it.X = a4.x;
                s_it.d = a4.y;
                s_it.row0 = a4.z;
                s_it.R0 = r2.x;
                s_it.R1 = r2.y;
            } else if (item < nitems) {  // cannot happen: plan_kernel bounds the item count
                ctl->infeasible = 3;
                s_it.item = nitems;
            }
        }
        __syncthreads();
        if (s_it.item >= nitems) break;
        const int X = s_it.X;
        const int d = s_it.d;
        const int row0 = s_it.row0;
        const int *a_g = p.adj + row0;
        const u64 R0 = s_it.R0;
        const u64 R1 = s_it.R1;
        const int tX = use_time ? p.time_index[X] : 0;

        // ---- shared memory carve-up -------------------------------------------------------
        const int dpad = (d + 3) & ~3;
        size_t off = 0;
        double *s_cx = (double *)(smem + off);
        off += (size_t)dpad * 8;
        int *s_adj = (int *)(smem + off);
        off += (size_t)dpad * 4;
        int *s_time = (int *)(smem + off);
        if (use_time) off += (size_t)dpad * 4;
        off = (off + 15) & ~(size_t)15;
        const int nrowsW = k + 3 + (het_general ? 1 : 0);
        size_t perwarp = ((size_t)nrowsW * dpad + 16) * 8 + 64 + dpad;
        perwarp = (perwarp + 15) & ~(size_t)15;
        const size_t avail = (size_t)p.smem_bytes - off;
        const size_t csub_bytes = (size_t)d * dpad * 4;
        int nw;
        bool staged;
        // staging costs d^2 gathers spread over the CTA's 512 threads: for d <= 64 that is a couple of L2
        // round trips; for larger rows it pays when the item runs a few times as many tests
        const bool worth_staging = d <= 64 || (R1 - R0) * (u64)(d - l) >= 3 * (u64)d * (u64)d;
        if (worth_staging && csub_bytes + 8 * perwarp <= avail) {
            staged = true;
            const size_t q = (avail - csub_bytes) / perwarp;
            nw = q < CI_NWARPS ? (int)q : CI_NWARPS;
        } else {
            staged = false;
            const size_t q = avail / perwarp;
            nw = q < CI_NWARPS ? (int)q : CI_NWARPS;
        }
        if (nw == 0) {
            if (tid == 0) ctl->infeasible = 2;
            continue;
        }
        if ((u64)nw > R1 - R0) nw = (int)(R1 - R0);
        float *Csub = (float *)(smem + off);
        unsigned char *warp_area = smem + off + (staged ? csub_bytes : 0);

        // ---- stage the row ------------------------------------------------------------------
        for (int i = tid; i < dpad; i += CI_NTHREADS) {
            if (i < d) {
                const int v = a_g[i];
                s_adj[i] = v;
                s_cx[i] = (double)p.C[(size_t)X * n + v];
                if (use_time) s_time[i] = p.time_index[v];
            } else {
                s_adj[i] = 0;
                s_cx[i] = 0.0;
                if (use_time) s_time[i] = 0;
            }
        }
        __syncthreads();
        if (staged) {
            // d x dpad gathers from L2: eight independent loads in flight per thread (the row is idle until
            // they land, so the memory-level parallelism here is what a short item costs)
            const int total = d * dpad;
            for (int base = tid; base < total; base += CI_NTHREADS * 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CI_NTHREADS;
                    v[u] = 0.0f;
                    if (idx < total) {
                        const int i = idx / dpad, j = idx - i * dpad;
                        if (j < d) v[u] = (i == j) ? 1.0f : __ldg(&p.C[(size_t)s_adj[i] * n + s_adj[j]]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int idx = base + u * CI_NTHREADS;
                    if (idx < total) Csub[idx] = v[u];
                }
            }
            __syncthreads();
        }
        if (warp >= nw) continue;
        const u64 per = (R1 - R0 + nw - 1) / nw;
        const u64 Rw0 = R0 + (u64)warp * per;
        const u64 Rw1 = (Rw0 + per < R1) ? Rw0 + per : R1;
        if (Rw0 >= Rw1) continue;

        // ---- per-warp state ----------------------------------------------------------------
        double *W = (double *)(warp_area + (size_t)warp * perwarp);
        double *Rxa = W + (size_t)k * dpad;
        double *Raa = Rxa + dpad;
        double *IRaa = Raa + dpad;
        double *NY = IRaa + dpad;  // only when het_general
        double *wx = W + (size_t)nrowsW * dpad;
        int *pos = (int *)(wx + 16);
        unsigned char *skip = (unsigned char *)(pos + 16);

        if (lane == 0) {
            int tmp[CUSK_ML];
            unrank_comb_fast(Rw0, d, l, tmp);
            for (int i = 0; i < l; i++) pos[i] = tmp[i];
        }
        __syncwarp();
        int zstart = pos[k];
        int zfirst = (k > 0) ? pos[k - 1] + 1 : 0;
        u64 Rbase = Rw0 - (u64)(zstart - zfirst);
        int changed = 0;
        uint32_t negmask = 0;

        for (;;) {
            // ---- (re)build elimination rows changed..k-1 ---------------------------------
            for (int j = changed; j < k; j++) {
                const int pj = pos[j];
                double D = 1.0, xacc = s_cx[pj];
                for (int i = 0; i < j; i++) {
                    const double w = W[(size_t)i * dpad + pj];
                    const double sw = ((negmask >> i) & 1u) ? -w : w;
                    D -= sw * w;
                    xacc -= sw * wx[i];
                }
                const bool drop = fabs(D) < CI_PIVOT_TOL;
                const double sc = drop ? 0.0 : rsqrt(fabs(D));
                if (D < 0 && !drop)
                    negmask |= (1u << j);
                else
                    negmask &= ~(1u << j);
                if (drop && lane == 0) my_sing++;
                const float *crow_g = p.C + (size_t)s_adj[pj] * n;
                for (int a = lane; a < dpad; a += 32) {
                    double u = 0.0;
                    if (a < d) {
                        u = (a == pj) ? 1.0
                                      : (staged ? (double)Csub[(size_t)pj * dpad + a] : (double)crow_g[s_adj[a]]);
                        for (int i = 0; i < j; i++) {
                            const double w = W[(size_t)i * dpad + pj];
                            const double sw = ((negmask >> i) & 1u) ? -w : w;
                            u -= sw * W[(size_t)i * dpad + a];
                        }
                    }
                    W[(size_t)j * dpad + a] = u * sc;
                }
                __syncwarp();
                if (lane == 0) wx[j] = xacc * sc;
                __syncwarp();
            }
            // ---- prefix residuals --------------------------------------------------------
            double rxx = 1.0;
            int tP = -2147483647 - 1;
            for (int j = 0; j < k; j++) {
                const double w = wx[j];
                rxx -= (((negmask >> j) & 1u) ? -w : w) * w;
                if (use_time) tP = max(tP, s_time[pos[j]]);
            }
            double SP = 0.0;
            if (het_general) {
                // pairs inside the prefix and prefix x X (src/hetcor-cuPC-S.cu:3068-3089 order:
                // N[later * n + earlier] with the tuple ordered X, Y, S_1..S_l)
                double part = 0.0;
                if (lane < k) {
                    const size_t rowj = (size_t)s_adj[pos[lane]] * n;
                    part = (double)__float2int_rz(p.N[rowj + X]);
                    for (int i = 0; i < lane; i++) part += (double)__float2int_rz(p.N[rowj + s_adj[pos[i]]]);
                }
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULLMASK, part, o);
                SP = part;
            }
            for (int a = lane; a < dpad; a += 32) {
                double rx = s_cx[a], ra = 1.0;
                bool sk = (a >= d);
                for (int j = 0; j < k; j++) {
                    const double w = W[(size_t)j * dpad + a];
                    const double sw = ((negmask >> j) & 1u) ? -w : w;
                    rx -= sw * wx[j];
                    ra -= sw * w;
                    sk |= (pos[j] == a);
                }
                Rxa[a] = rx;
                Raa[a] = ra;
                IRaa[a] = (fabs(ra) < CI_PIVOT_TOL) ? 0.0 : 1.0 / ra;
                if (a < d) {
                    const int Y = s_adj[a];
                    if (HET) {
                        const uint32_t wbits = __ldcg(&p.bits[(size_t)X * p.rw + (Y >> 5)]);
                        sk |= !((wbits >> (Y & 31)) & 1u);
                        if (het_general) {
                            double s = (double)__float2int_rz(p.N[(size_t)Y * n + X]);
                            for (int j = 0; j < k; j++)
                                s += (double)__float2int_rz(p.N[(size_t)s_adj[pos[j]] * n + Y]);
                            NY[a] = s;
                        }
                    } else {
                        sk |= (__ldcg(&p.minkey[row0 + a]) < Rbase);
                    }
                }
                skip[a] = sk ? 1 : 0;
            }
            __syncwarp();

            // ---- leaves: z = zstart .. zend-1 ----------------------------------------------
            const u64 remain = Rw1 - Rbase;
            const int zend = (remain >= (u64)(d - zfirst)) ? d : zfirst + (int)remain;
            for (int z4 = zstart & ~3; z4 < zend; z4 += CI_ZB) {
                double g[CI_ZB], iz[CI_ZB], c1[CI_ZB], c2[CI_ZB], SPz[CI_ZB];
                int tS[CI_ZB];
                bool act[CI_ZB];
#pragma unroll
                for (int q = 0; q < CI_ZB; q++) {
                    const int zq = z4 + q;
                    act[q] = (zq >= zstart) && (zq < zend);
                    const double xz = Rxa[zq];
                    iz[q] = IRaa[zq];
                    g[q] = xz * iz[q];
                    c2[q] = fabs(rxx - xz * g[q]);
                    c1[q] = T2u * c2[q];
                    tS[q] = 0;
                    SPz[q] = 0.0;
                    if (use_time) tS[q] = max(tP, s_time[zq]);
                    if (het_general && act[q]) {
                        const size_t rowz = (size_t)s_adj[zq] * n;
                        double part = (lane == 0) ? (double)__float2int_rz(p.N[rowz + X]) : 0.0;
                        if (lane < k) part += (double)__float2int_rz(p.N[rowz + s_adj[pos[lane]]]);
                        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULLMASK, part, o);
                        SPz[q] = SP + part;
                    }
                }
                for (int y0 = 0; y0 < d; y0 += 32) {
                    const int y = y0 + lane;
                    const bool inb = y < d;
                    const int yc = inb ? y : d - 1;
                    const bool sk = inb ? (skip[yc] != 0) : true;
                    if (__all_sync(FULLMASK, sk)) continue;
                    double acc[CI_ZB];
                    if (staged) {
#pragma unroll
                        for (int q = 0; q < CI_ZB; q++) {
                            const int zr = min(z4 + q, d - 1);
                            acc[q] = (double)Csub[(size_t)zr * dpad + yc];
                        }
                    } else {
                        const int ya = s_adj[yc];
#pragma unroll
                        for (int q = 0; q < CI_ZB; q++) {
                            const int zr = min(z4 + q, d - 1);
                            acc[q] = (double)__ldg(&p.C[(size_t)s_adj[zr] * n + ya]);
                        }
                    }
                    for (int j = 0; j < k; j++) {
                        const double *wrow = W + (size_t)j * dpad;
                        double wy = wrow[yc];
                        if ((negmask >> j) & 1u) wy = -wy;
                        const double2 s01 = *(const double2 *)(wrow + z4);
                        const double2 s23 = *(const double2 *)(wrow + z4 + 2);
                        acc[0] = fma(-s01.x, wy, acc[0]);
                        acc[1] = fma(-s01.y, wy, acc[1]);
                        acc[2] = fma(-s23.x, wy, acc[2]);
                        acc[3] = fma(-s23.y, wy, acc[3]);
                    }
                    const double rxy = Rxa[yc], ryy = Raa[yc];
                    const int ty = use_time ? max(tX, s_time[yc]) : 0;
                    bool removed_any = false;
#pragma unroll
                    for (int q = 0; q < CI_ZB; q++) {
                        bool valid = act[q] && !sk && (y != z4 + q);
                        if (use_time) valid = valid && (tS[q] <= ty);
                        if (!valid) continue;
                        my_tests++;
                        const double syz = acc[q];
                        const double h01 = rxy - g[q] * syz;
                        const double h11 = ryy - syz * syz * iz[q];
                        const double lhs = h01 * h01;
                        const double ah11 = fabs(h11);
                        double T2 = T2u, thr = c1[q] * ah11;
                        if (het_general) {
                            const double sum =
                                SPz[q] + NY[yc] +
                                (double)__float2int_rz(p.N[(size_t)s_adj[z4 + q] * n + s_adj[yc]]);
                            const double m = sum / npairs;
                            const double th = p.th_het / sqrt(m - (double)l - 3.0);
                            const double t = tanh(th);
                            T2 = t * t;
                            thr = T2 * c2[q] * ah11;
                        }
                        const double den = c2[q] * ah11;
                        // |rho| < tanh(th)  or, on the |rho| > 1 branch, |rho| > coth(th)
                        const bool rem = (lhs < thr) || (den > 0.0 && lhs * T2 > den);
                        if (rem) {
                            removed_any = true;
                            const int Y = s_adj[yc];
                            if (HET) {
                                atomicAnd(&p.bits[(size_t)X * p.rw + (Y >> 5)], ~(1u << (Y & 31)));
                                atomicAnd(&p.bits[(size_t)Y * p.rw + (X >> 5)], ~(1u << (X & 31)));
                                if (p.removed) p.removed[row0 + yc] = 1;
                            } else {
                                const u64 R = Rbase + (u64)(z4 + q - zfirst);
                                atomicMin(&p.minkey[row0 + yc], R);
                            }
                        }
                    }
                    if (removed_any) skip[yc] = 1;
                }
            }

            // ---- next prefix -------------------------------------------------------------
            Rbase += (u64)(d - zfirst);
            if (k == 0 || Rbase >= Rw1) break;
            int chg = -1;
            if (lane == 0) {
                int i = k - 1;
                while (i >= 0 && pos[i] == d - 1 - k + i) i--;
                if (i >= 0) {
                    pos[i]++;
                    for (int jj = i + 1; jj < k; jj++) pos[jj] = pos[jj - 1] + 1;
                    chg = i;
                }
            }
            chg = __shfl_sync(FULLMASK, chg, 0);
            __syncwarp();
            if (chg < 0) break;
            changed = chg;
            zfirst = pos[k - 1] + 1;
            zstart = zfirst;
        }
    }
    // ---- statistics ----------------------------------------------------------------------------
    for (int o = 16; o > 0; o >>= 1) {
        my_tests += __shfl_xor_sync(FULLMASK, my_tests, o);
        my_sing += __shfl_xor_sync(FULLMASK, my_sing, o);
    }
    if (lane == 0) {
        if (my_tests) atomicAdd((unsigned long long *)&ctl->executed, (unsigned long long)my_tests);
        if (my_sing) atomicAdd((unsigned long long *)&ctl->singular, (unsigned long long)my_sing);
    }
}

int launch_ci_level(cudaStream_t st, bool het, const CiParams &p, int grid)
{
    cudaError_t e;
    if (het)
        e = cudaFuncSetAttribute(ci_level_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 p.smem_bytes);
    else
        e = cudaFuncSetAttribute(ci_level_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 p.smem_bytes);
    if (e != cudaSuccess) {
        cusk_set_error("cudaFuncSetAttribute(ci_level_kernel, %d B): %s", p.smem_bytes, cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    if (het)
        launch_k(ci_level_kernel<true>, dim3(grid), dim3(CI_NTHREADS), p.smem_bytes, st, p);
    else
        launch_k(ci_level_kernel<false>, dim3(grid), dim3(CI_NTHREADS), p.smem_bytes, st, p);
    return CUSK_OK;
}
