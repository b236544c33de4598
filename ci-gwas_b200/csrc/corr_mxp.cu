// corr_mxp.cu -- marker x trait Pearson correlations, tiled (replaces
// bed_marker_phen_corr_pearson_scan, src/corr_kernels.cu:157-238, which re-reads every trait column
// once per marker):
//     r = (sum v g y - mean_g sum v y) / (sum v * std_g),   v = genotype valid and y not NaN.
// Rewritten over genotype classes: sum v g y = S1 + 2 S2 with S_a = sum_{g = a, y valid} y, and
// sum v y = T_y - S_miss, sum v = n_y - c_miss (T_y, n_y: per-trait totals over valid y; S_miss, c_miss:
// the same sums over samples whose genotype is missing -- identically zero for a block without
// missing calls, which then needs S1 and S2 only).
//
// A CTA owns 8 * MW markers and walks the samples in chunks of 1024; the chunk's trait values are
// staged once in shared memory (NaN -> 0) and reused by all its markers, so trait traffic drops from
// m * p * n * 4 bytes to (m / (8 MW)) * p * n * 4.  Class sums are accumulated in float over at most
// 128 samples per lane, warp-reduced and added to FP64 accumulators (relative error of the sums
// ~1e-7, far inside the 1e-6 bar; the old all-FP64 kernel is kept in corr_kernels.cu for A/B tests).
#include "corr.cuh"

namespace {

constexpr int CH = 1024;        // samples per chunk (256 packed bytes per marker)
constexpr int PT = 17;          // traits per pass
constexpr int FLUSH_EVERY = 4;  // chunks between float -> double flushes

__global__ void __launch_bounds__(256) trait_totals_kernel(const float *__restrict__ phen, int n, int p,
                                                           double *__restrict__ Ty, int *__restrict__ cnty)
{
    const int t = blockIdx.x;
    double s = 0.0;
    int c = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float y = phen[(size_t)t * n + i];
        if (y == y) {
            s += (double)y;
            c++;
        }
    }
    __shared__ double ss[8];
    __shared__ int sc[8];
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    if ((threadIdx.x & 31) == 0) {
        ss[threadIdx.x >> 5] = s;
        sc[threadIdx.x >> 5] = c;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double S = 0;
        int Cn = 0;
        for (int w = 0; w < 8; w++) {
            S += ss[w];
            Cn += sc[w];
        }
        Ty[t] = S;
        cnty[t] = Cn;
    }
}

// MISSING = false: MW = 2 markers per warp, classes {g=1, g=2}
// MISSING = true : MW = 1 marker per warp,  classes {g=1, g=2, missing-sum, missing-count}
template <bool MISSING>
__global__ void __launch_bounds__(256) mxp_tiled_kernel(const unsigned char *__restrict__ bed,
                                                        const float *__restrict__ phen, int m, int n, int p, int t0,
                                                        size_t col_bytes, const float *__restrict__ mean,
                                                        const float *__restrict__ stdv, const double *__restrict__ Ty,
                                                        const int *__restrict__ cnty, float *__restrict__ out)
{
    constexpr int MW = MISSING ? 1 : 2;
    constexpr int NA = MISSING ? 4 : 2;
    constexpr int NCOMBO = MW * PT * NA;
    constexpr int NSLOT = (NCOMBO + 31) / 32;
    extern __shared__ __align__(16) float sm[];
    float *ys = sm;                               // [PT][CH]
    float *yv = sm + (MISSING ? PT * CH : 0);     // [PT][CH], MISSING only
    __shared__ double s_fin[8][NCOMBO];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nt = min(PT, p - t0);
    const int mk0 = (blockIdx.x * 8 + warp) * MW;
    const unsigned char *col[MW];
#pragma unroll
    for (int w = 0; w < MW; w++) col[w] = bed + (size_t)min(mk0 + w, m - 1) * col_bytes;

    float acc[MW][PT][NA];
#pragma unroll
    for (int w = 0; w < MW; w++)
#pragma unroll
        for (int t = 0; t < PT; t++)
#pragma unroll
            for (int a = 0; a < NA; a++) acc[w][t][a] = 0.f;
    double dacc[NSLOT];
#pragma unroll
    for (int i = 0; i < NSLOT; i++) dacc[i] = 0.0;

    const int nchunks = (n + CH - 1) / CH;
    for (int c = 0; c < nchunks; c++) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < nt * CH; idx += 256) {
            const int t = idx / CH, s = idx - t * CH;
            const size_t gs = (size_t)c * CH + s;
            float y = (gs < (size_t)n) ? phen[(size_t)(t0 + t) * n + gs] : __int_as_float(0x7fc00000);
            const bool ok = (y == y);
            ys[t * CH + s] = ok ? y : 0.f;
            if (MISSING) yv[t * CH + s] = ok ? 1.f : 0.f;
        }
        __syncthreads();
        // this lane's 8 packed bytes (32 samples) of each marker: byte index lane + 32 r
        uint32_t vb[MW][2];
#pragma unroll
        for (int w = 0; w < MW; w++) {
            vb[w][0] = vb[w][1] = 0xffffffffu;  // code 11 (genotype 0) contributes nothing
#pragma unroll
            for (int r = 0; r < 8; r++) {
                const size_t bi = (size_t)c * (CH / 4) + lane + 32 * r;
                const uint32_t b = (bi < col_bytes) ? (uint32_t)col[w][bi] : 0xffu;
                vb[w][r >> 2] = (vb[w][r >> 2] & ~(0xffu << (8 * (r & 3)))) | (b << (8 * (r & 3)));
            }
        }
#pragma unroll
        for (int t = 0; t < PT; t++) {
            if (t < nt) {
                const float4 *y4 = (const float4 *)(ys + t * CH);
                const float4 *v4 = (const float4 *)(yv + t * CH);
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const float4 f = y4[lane + 32 * r];
                    const float yy[4] = {f.x, f.y, f.z, f.w};
                    float vv[4] = {0.f, 0.f, 0.f, 0.f};
                    if (MISSING) {
                        const float4 g = v4[lane + 32 * r];
                        vv[0] = g.x;
                        vv[1] = g.y;
                        vv[2] = g.z;
                        vv[3] = g.w;
                    }
#pragma unroll
                    for (int w = 0; w < MW; w++) {
                        const uint32_t byte = (vb[w][r >> 2] >> (8 * (r & 3))) & 0xffu;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint32_t code = (byte >> (2 * j)) & 3u;
                            acc[w][t][0] += (code == 2u) ? yy[j] : 0.f;  // genotype 1
                            acc[w][t][1] += (code == 0u) ? yy[j] : 0.f;  // genotype 2
                            if (MISSING) {
                                acc[w][t][2] += (code == 1u) ? yy[j] : 0.f;
                                acc[w][t][3] += (code == 1u) ? vv[j] : 0.f;
                            }
                        }
                    }
                }
            }
        }
        if ((c % FLUSH_EVERY) == FLUSH_EVERY - 1 || c == nchunks - 1) {
#pragma unroll
            for (int w = 0; w < MW; w++)
#pragma unroll
                for (int t = 0; t < PT; t++)
#pragma unroll
                    for (int a = 0; a < NA; a++) {
                        float v = acc[w][t][a];
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                        constexpr int dummy = 0;
                        (void)dummy;
                        const int cidx = (w * PT + t) * NA + a;
                        if (lane == (cidx & 31)) dacc[cidx >> 5] += (double)v;
                        acc[w][t][a] = 0.f;
                    }
        }
    }
    // publish the FP64 sums and finish
#pragma unroll
    for (int i = 0; i < NSLOT; i++) {
        const int cidx = i * 32 + lane;
        if (cidx < NCOMBO) s_fin[warp][cidx] = dacc[i];
    }
    __syncwarp();
    for (int q = lane; q < MW * nt; q += 32) {
        const int w = q / nt, t = q - w * nt;
        const int mk = mk0 + w;
        if (mk >= m) continue;
        const double *f = &s_fin[warp][(w * PT + t) * NA];
        const double s1 = f[0], s2 = f[1];
        double sy = Ty[t0 + t], sv = (double)cnty[t0 + t];
        if (MISSING) {
            sy -= f[2];
            sv -= f[3];
        }
        // genotype-0 samples contribute y to sum v y but nothing to sum v g y
        const double r = (s1 + 2.0 * s2 - (double)mean[mk] * sy) / (sv * (double)stdv[mk]);
        out[(size_t)mk * p + (t0 + t)] = (float)r;
    }
}

}  // namespace

// out[m*p] row-major (marker major), the reference's layout (src/corr_host.cu:1070)
int launch_mxp(cudaStream_t st, const unsigned char *bed, const float *phen, int m, int n, int p, size_t col_bytes,
               const float *mean, const float *stdv, float *out, bool any_missing, double *d_Ty, int *d_cnty)
{
    trait_totals_kernel<<<p, 256, 0, st>>>(phen, n, p, d_Ty, d_cnty);
    for (int t0 = 0; t0 < p; t0 += PT) {
        if (any_missing) {
            const int smem = 2 * PT * CH * 4;
            cudaError_t e = cudaFuncSetAttribute(mxp_tiled_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) {
                cusk_set_error("cudaFuncSetAttribute(mxp_tiled_kernel): %s", cudaGetErrorString(e));
                return CUSK_E_CUDA;
            }
            mxp_tiled_kernel<true><<<(m + 7) / 8, 256, smem, st>>>(bed, phen, m, n, p, t0, col_bytes, mean, stdv, d_Ty,
                                                                  d_cnty, out);
        } else {
            const int smem = PT * CH * 4;
            cudaError_t e = cudaFuncSetAttribute(mxp_tiled_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) {
                cusk_set_error("cudaFuncSetAttribute(mxp_tiled_kernel): %s", cudaGetErrorString(e));
                return CUSK_E_CUDA;
            }
            mxp_tiled_kernel<false><<<(m + 15) / 16, 256, smem, st>>>(bed, phen, m, n, p, t0, col_bytes, mean, stdv, d_Ty,
                                                                    d_cnty, out);
        }
    }
    return CUSK_OK;
}

// ---------------------------------------------------------------------------------------------
// Tensor-core route (north_star item 1: marker x trait as a dense contraction on tcgen05): the trait
// values become uint8 operand rows next to the genotype indicator rows, so that
//     sum_{k in class} y_k  =  2^-S ( sum_j 2^(8j) <I_class, limb_j>  -  2^31 <I_class, valid> )
// with exact int32 dot products.  Fixed point: q = round(y 2^S), |q| < 2^31, S = 29 - floor(log2 max|y|)
// per trait: the rounding error of a class sum is below 2^-S sqrt(n) ~ 1e-6 absolute on sums of order
// n E|y| ~ 1e5, i.e. ~1e-11 relative -- five orders inside the 1e-6 bar, and tighter than the float
// partial sums of mxp_tiled_kernel above.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) trait_scale_kernel(const float *__restrict__ phen, int n, int p,
                                                          double *__restrict__ Ty, int *__restrict__ cnty,
                                                          double *__restrict__ scale_inv, int *__restrict__ sexp)
{
    const int t = blockIdx.x;
    double s = 0.0;
    int c = 0;
    float mx = 0.f;
    // 1024 threads, eight samples in flight per thread (17 CTAs of 256 threads reading one value at a time took 0.84 ms)
    const int B = blockDim.x;
    const float *col = phen + (size_t)t * n;
    int i = threadIdx.x;
    for (; i + 7 * B < n; i += 8 * B) {
        float y[8];
#pragma unroll
        for (int u = 0; u < 8; u++) y[u] = __ldg(col + i + u * B);
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (y[u] == y[u]) {
                s += (double)y[u];
                c++;
                mx = fmaxf(mx, fabsf(y[u]));
            }
    }
    for (; i < n; i += B) {
        const float y = col[i];
        if (y == y) {
            s += (double)y;
            c++;
            mx = fmaxf(mx, fabsf(y));
        }
    }
    __shared__ double ss[32];
    __shared__ int sc[32];
    __shared__ float sm_[32];
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        ss[threadIdx.x >> 5] = s;
        sc[threadIdx.x >> 5] = c;
        sm_[threadIdx.x >> 5] = mx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double S = 0;
        int Cn = 0;
        float M = 0.f;
        for (int w = 0; w < (B >> 5); w++) {
            S += ss[w];
            Cn += sc[w];
            M = fmaxf(M, sm_[w]);
        }
        Ty[t] = S;
        cnty[t] = Cn;
        const int e = (M > 0.f && M < 3.0e38f) ? ilogbf(M) : 0;
        const int S_ = 29 - e;
        sexp[t] = S_;
        scale_inv[t] = ldexp(1.0, -S_);
    }
}

// thread = 16 samples of one trait -> one 16-byte store into each of the trait's 5 operand rows
__global__ void __launch_bounds__(256) trait_limbs_kernel(const float *__restrict__ phen, int n, size_t Kp,
                                                          const int *__restrict__ sexp,
                                                          unsigned char *__restrict__ rows)
{
    const int t = blockIdx.y;
    const size_t chunk = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (chunk * 16 >= Kp) return;
    const int S_ = sexp[t];
    uint32_t o[5][4];
#pragma unroll
    for (int j = 0; j < 5; j++)
#pragma unroll
        for (int w = 0; w < 4; w++) o[j][w] = 0u;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const size_t s = chunk * 16 + i;
        if (s >= (size_t)n) continue;
        const float y = phen[(size_t)t * n + s];
        if (!(y == y)) continue;
        const long long q = llrint(ldexp((double)y, S_));
        const uint32_t u = (uint32_t)(q + 2147483648LL);
        const int w = i >> 2, sh = 8 * (i & 3);
        o[0][w] |= (u & 0xffu) << sh;
        o[1][w] |= ((u >> 8) & 0xffu) << sh;
        o[2][w] |= ((u >> 16) & 0xffu) << sh;
        o[3][w] |= ((u >> 24) & 0xffu) << sh;
        o[4][w] |= 1u << sh;
    }
#pragma unroll
    for (int j = 0; j < 5; j++)
        *(uint4 *)(rows + ((size_t)t * 8 + j) * Kp + chunk * 16) = make_uint4(o[j][0], o[j][1], o[j][2], o[j][3]);
}

// rows: first row of the trait block (256 zero-initialised rows of Kp bytes)
void launch_trait_rows(cudaStream_t st, const float *phen, int n, int p, size_t Kp, unsigned char *rows, double *d_Ty,
                       int *d_cnty, double *d_scale_inv, int *d_sexp)
{
    trait_scale_kernel<<<p, 1024, 0, st>>>(phen, n, p, d_Ty, d_cnty, d_scale_inv, d_sexp);
    dim3 grid((unsigned)((Kp / 16 + 255) / 256), p);
    trait_limbs_kernel<<<grid, 256, 0, st>>>(phen, n, Kp, d_sexp, rows);
}
