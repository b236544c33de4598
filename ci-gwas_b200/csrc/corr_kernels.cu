// corr_kernels.cu -- CUDA-core kernels of the correlation front end: BED marker statistics
// (prep-bed), 2-bit -> indicator decode, marker x trait and trait x trait correlations.
#include "corr.cuh"

// ---------------------------------------------------------------------------------------------
// Genotype counts per marker (exact integers).  One warp per marker; 16 samples per 32-bit word.
// Replaces the counting pass of compute_bed_col_stats_no_impute (src/prep.cpp:28-50).
// Bytes: m * ceil(n/4) read once (HBM bound).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bed_counts_kernel(const unsigned char *__restrict__ bed, int m, int n,
                                                         size_t col_bytes, int *__restrict__ counts)
{
    const int lane = threadIdx.x & 31;
    const int mk = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (mk >= m) return;
    const unsigned char *col = bed + (size_t)mk * col_bytes;
    int c0 = 0, c1 = 0, c2 = 0, cm = 0;
    for (size_t b = lane; b < col_bytes; b += 32) {
        const unsigned v = col[b];
        const int ns = (int)min((size_t)4, (size_t)n - b * 4);  // samples in this byte
        const unsigned valid = (1u << (2 * ns)) - 1u;
        const unsigned lo = v & 0x55u, hi = (v >> 1) & 0x55u, smask = valid & 0x55u;
        c2 += __popc(~lo & ~hi & smask);  // 00
        cm += __popc(lo & ~hi & smask);   // 01
        c1 += __popc(~lo & hi & smask);   // 10
        c0 += __popc(lo & hi & smask);    // 11
    }
    for (int o = 16; o > 0; o >>= 1) {
        c0 += __shfl_xor_sync(0xffffffffu, c0, o);
        c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        c2 += __shfl_xor_sync(0xffffffffu, c2, o);
        cm += __shfl_xor_sync(0xffffffffu, cm, o);
    }
    if (lane == 0) {
        counts[4 * mk + 0] = c0;
        counts[4 * mk + 1] = c1;
        counts[4 * mk + 2] = c2;
        counts[4 * mk + 3] = cm;
    }
}

// ---------------------------------------------------------------------------------------------
// Means / standard deviations / modes exactly as prep-bed computes them (src/prep.cpp:52-76):
// mean = sum / (float)(n - nan); std = sqrt(sum_squares / (float)(n - nan)) where sum_squares is a
// float accumulated SEQUENTIALLY in sample order.  The sequential float rounding is part of the
// on-disk contract (.stds is what cusk re-reads), so one thread per marker replays it: the only
// per-sample work is one float add of one of three precomputed squares.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) bed_stats_kernel(const unsigned char *__restrict__ bed, int m, int n,
                                                        size_t col_bytes, const int *__restrict__ counts,
                                                        float *__restrict__ means, float *__restrict__ stds,
                                                        int *__restrict__ modes)
{
    const int mk = blockIdx.x * blockDim.x + threadIdx.x;
    if (mk >= m) return;
    const int *c = counts + 4 * mk;
    const size_t gt_sum = (size_t)c[1] + 2 * (size_t)c[2];
    const float denom = (float)((size_t)n - (size_t)c[3]);
    const float mean = (float)gt_sum / denom;  // size_t / float: the sum converts to float first
    const float sq_g0 = (0.0f - mean) * (0.0f - mean);  // code 11 -> 0
    const float sq_g1 = (1.0f - mean) * (1.0f - mean);  // code 10 -> 1
    const float sq_g2 = (2.0f - mean) * (2.0f - mean);  // code 00 -> 2 (code 01 = missing: skipped)
    const unsigned char *col = bed + (size_t)mk * col_bytes;
    float sum_squares = 0.0f;
    size_t s = 0;
    for (size_t b = 0; b < col_bytes; b++) {
        const unsigned v = col[b];
#pragma unroll
        for (int j = 0; j < 4; j++, s++) {
            const unsigned code = (v >> (2 * j)) & 3u;
            const float add = (code == 0u) ? sq_g2 : (code == 2u ? sq_g1 : sq_g0);
            if (s < (size_t)n && code != 1u) sum_squares = __fadd_rn(sum_squares, add);
        }
    }
    means[mk] = mean;
    stds[mk] = sqrtf(sum_squares / denom);
    int mode = 0;
    if (c[1] > c[mode]) mode = 1;
    if (c[2] > c[mode]) mode = 2;
    modes[mk] = mode;
}

// ---------------------------------------------------------------------------------------------
// 2-bit BED -> uint8 indicator rows for the tensor-core Gram.  Thread = 4 packed bytes (16 samples)
// of one marker -> one 16-byte store per part.  Bytes: m ceil(n/4) read, parts * m * Kp written.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) decode_indicators_kernel(const unsigned char *__restrict__ bed, int m, int n,
                                                                size_t col_bytes, int parts, int MI, size_t Kp,
                                                                unsigned char *__restrict__ ind)
{
    const int mk = blockIdx.y;
    const size_t chunk = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // 16 samples each
    if (chunk * 16 >= Kp) return;
    const unsigned char *col = bed + (size_t)mk * col_bytes;
    unsigned v = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const size_t bi = chunk * 4 + b;
        if (bi < col_bytes) v |= (unsigned)col[bi] << (8 * b);
    }
    uint32_t o1[4] = {0, 0, 0, 0}, o2[4] = {0, 0, 0, 0}, ov[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const size_t s = chunk * 16 + j;
        const unsigned code = (v >> (2 * j)) & 3u;
        if (s < (size_t)n) {
            o1[j >> 2] |= (uint32_t)(code == 2u) << (8 * (j & 3));
            o2[j >> 2] |= (uint32_t)(code == 0u) << (8 * (j & 3));
            ov[j >> 2] |= (uint32_t)(code != 1u) << (8 * (j & 3));
        }
    }
    const size_t off = chunk * 16;
    if (parts == 2) {
        *(uint4 *)(ind + ind_row(mk, 0, 2, MI) * Kp + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
        *(uint4 *)(ind + ind_row(mk, 1, 2, MI) * Kp + off) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
    } else {
        *(uint4 *)(ind + ind_row(mk, 0, 3, MI) * Kp + off) = make_uint4(ov[0], ov[1], ov[2], ov[3]);
        *(uint4 *)(ind + ind_row(mk, 1, 3, MI) * Kp + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
        *(uint4 *)(ind + ind_row(mk, 2, 3, MI) * Kp + off) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// Marker x trait Pearson correlation (replaces bed_marker_phen_corr_pearson_scan,
// src/corr_kernels.cu:157-238): r = (sum v g y - mean_g sum v y) / (sum v * std_g), samples with NaN
// y skipped, FP64 accumulation.  One warp per marker, lanes stride over the packed bytes; the 8
// warps of a CTA walk the same sample range so the trait values are shared through L1.
// PMAX traits per pass keep the accumulators in registers.
// ---------------------------------------------------------------------------------------------
template <int PT>
__global__ void __launch_bounds__(256) mxp_kernel(const unsigned char *__restrict__ bed, const float *__restrict__ phen,
                                                  int m, int n, int p, int t0, size_t col_bytes,
                                                  const float *__restrict__ mean, const float *__restrict__ stdv,
                                                  float *__restrict__ out, int ldo, size_t out_stride_marker)
{
    const int lane = threadIdx.x & 31;
    const int mk = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (mk >= m) return;
    const unsigned char *col = bed + (size_t)mk * col_bytes;
    double sgy[PT], sy[PT];
    int sv[PT];
#pragma unroll
    for (int t = 0; t < PT; t++) {
        sgy[t] = 0.0;
        sy[t] = 0.0;
        sv[t] = 0;
    }
    const int nt = min(PT, p - t0);
    for (size_t b = lane; b < col_bytes; b += 32) {
        const unsigned v = col[b];
        const size_t s0 = b * 4;
#pragma unroll
        for (int t = 0; t < PT; t++) {
            if (t < nt) {
                const float *y = phen + (size_t)(t0 + t) * n + s0;
                float yy[4];
                if (s0 + 3 < (size_t)n && (((size_t)(t0 + t) * n + s0) & 3) == 0) {
                    const float4 f = *(const float4 *)y;
                    yy[0] = f.x;
                    yy[1] = f.y;
                    yy[2] = f.z;
                    yy[3] = f.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 4; j++) yy[j] = (s0 + j < (size_t)n) ? y[j] : __int_as_float(0x7fc00000);
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const unsigned code = (v >> (2 * j)) & 3u;
                    const float yv = yy[j];
                    if (yv == yv && code != 1u) {  // y not NaN, genotype valid
                        const double g = (code == 0u) ? 2.0 : (code == 2u ? 1.0 : 0.0);
                        sgy[t] += g * (double)yv;
                        sy[t] += (double)yv;
                        sv[t] += 1;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < PT; t++) {
        for (int o = 16; o > 0; o >>= 1) {
            sgy[t] += __shfl_xor_sync(0xffffffffu, sgy[t], o);
            sy[t] += __shfl_xor_sync(0xffffffffu, sy[t], o);
            sv[t] += __shfl_xor_sync(0xffffffffu, sv[t], o);
        }
    }
    if (lane == 0) {
        for (int t = 0; t < nt; t++) {
            const double r = (sgy[t] - (double)mean[mk] * sy[t]) / ((double)sv[t] * (double)stdv[mk]);
            out[(size_t)mk * out_stride_marker + (size_t)(t0 + t) * ldo] = (float)r;
        }
    }
}

// Trait x trait pairwise-complete correlation of pre-standardised columns (replaces
// phen_corr_pearson_scan, src/corr_kernels.cu:285-343): sum y_a y_b / #{both non-NaN}.
__global__ void __launch_bounds__(256) pxp_kernel(const float *__restrict__ phen, int n, int p,
                                                  float *__restrict__ out_upper, float *__restrict__ Csq, int ldc,
                                                  int m_off)
{
    // one CTA per unordered pair (a < b), linear index = blockIdx.x in upper-triangle row-major order
    int a = 0, rem = blockIdx.x;
    while (rem >= p - 1 - a) {
        rem -= p - 1 - a;
        a++;
    }
    const int b = a + 1 + rem;
    const float *ya = phen + (size_t)a * n, *yb = phen + (size_t)b * n;
    double s = 0.0;
    int cnt = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float va = ya[i], vb = yb[i];
        if (va == va && vb == vb) {
            s += (double)va * (double)vb;
            cnt++;
        }
    }
    __shared__ double ss[8];
    __shared__ int sc[8];
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if ((threadIdx.x & 31) == 0) {
        ss[threadIdx.x >> 5] = s;
        sc[threadIdx.x >> 5] = cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double S = 0;
        int Cn = 0;
        for (int w = 0; w < 8; w++) {
            S += ss[w];
            Cn += sc[w];
        }
        const float r = (float)(S / (double)Cn);
        if (out_upper) out_upper[blockIdx.x] = r;
        if (Csq) {
            Csq[(size_t)(m_off + a) * ldc + (m_off + b)] = r;
            Csq[(size_t)(m_off + b) * ldc + (m_off + a)] = r;
        }
    }
}

// dense square -> upper triangle, row-major, diagonal excluded (the reference's mxm layout,
// src/corr_host.cu:1132-1148); also sets the diagonal of the square matrix to 1.
__global__ void __launch_bounds__(256) pack_upper_kernel(const float *__restrict__ Csq, int m, int ldc,
                                                         float *__restrict__ upper)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= m || j <= i) return;
    const size_t lin = (size_t)i * (2 * (size_t)m - i - 1) / 2 + (size_t)(j - i - 1);
    upper[lin] = Csq[(size_t)i * ldc + j];
}

__global__ void __launch_bounds__(256) set_diag_kernel(float *Csq, int nvar, int ldc)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nvar) Csq[(size_t)i * ldc + i] = 1.0f;
}

// ---- launch wrappers --------------------------------------------------------------------------------
void launch_bed_counts(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int *counts)
{
    bed_counts_kernel<<<(m + 7) / 8, 256, 0, st>>>(bed, m, n, col_bytes, counts);
}
void launch_bed_stats(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, const int *counts,
                      float *means, float *stds, int *modes)
{
    bed_stats_kernel<<<(m + 127) / 128, 128, 0, st>>>(bed, m, n, col_bytes, counts, means, stds, modes);
}
void launch_decode(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int parts, int MI,
                   size_t Kp, unsigned char *ind)
{
    dim3 grid((unsigned)((Kp / 16 + 255) / 256), m);
    decode_indicators_kernel<<<grid, 256, 0, st>>>(bed, m, n, col_bytes, parts, MI, Kp, ind);
}
// first (simple) version, kept for A/B checks; the product path uses corr_mxp.cu
void launch_mxp_simple(cudaStream_t st, const unsigned char *bed, const float *phen, int m, int n, int p,
                       size_t col_bytes, const float *mean, const float *stdv, float *out, int ldo,
                       size_t out_stride_marker)
{
    constexpr int PT = 6;
    for (int t0 = 0; t0 < p; t0 += PT)
        mxp_kernel<PT><<<(m + 7) / 8, 256, 0, st>>>(bed, phen, m, n, p, t0, col_bytes, mean, stdv, out, ldo,
                                                   out_stride_marker);
}
void launch_pxp(cudaStream_t st, const float *phen, int n, int p, float *out_upper, float *Csq, int ldc, int m_off)
{
    if (p < 2) return;
    pxp_kernel<<<p * (p - 1) / 2, 256, 0, st>>>(phen, n, p, out_upper, Csq, ldc, m_off);
}
void launch_pack_upper(cudaStream_t st, const float *Csq, int m, int ldc, float *upper)
{
    dim3 grid((m + 255) / 256, m);
    pack_upper_kernel<<<grid, 256, 0, st>>>(Csq, m, ldc, upper);
}
void launch_set_diag(cudaStream_t st, float *Csq, int nvar, int ldc)
{
    set_diag_kernel<<<(nvar + 255) / 256, 256, 0, st>>>(Csq, nvar, ldc);
}
