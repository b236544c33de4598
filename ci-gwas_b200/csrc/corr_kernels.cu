// corr_kernels.cu -- CUDA-core kernels of the correlation front end: BED marker statistics
// (prep-bed), 2-bit -> indicator decode, marker x trait and trait x trait correlations.
#include "corr.cuh"

// ---------------------------------------------------------------------------------------------
// Genotype counts per marker (exact integers).  One warp per marker.  Replaces the counting pass of
// compute_bed_col_stats_no_impute (src/prep.cpp:28-50).  Bytes: m * ceil(n/4) read once (HBM bound).
// Round 2: 16-byte loads (two in flight per lane) after a byte-wise head up to the column's first 16-byte boundary --
// columns start at mk * ceil(n/4), any alignment -- and three popcounts per 32-bit word instead of four per byte:
// with lo / hi the low / high bits of the 2-bit codes, #11 = popc(lo & hi), #01 = popc(lo) - #11, #10 = popc(hi) - #11,
// #00 = samples - popc(lo | hi).  (The byte-per-lane first version ran at 0.8 TB/s, profiles/r2_corr_helpers_full_summary.txt.)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void bed_acc_word(uint32_t w, int &plo, int &phi, int &pand)
{
    const uint32_t lo = w & 0x55555555u, hi = (w >> 1) & 0x55555555u;
    plo += __popc(lo);
    phi += __popc(hi);
    pand += __popc(lo & hi);
}

__global__ void __launch_bounds__(256) bed_counts_kernel(const unsigned char *__restrict__ bed, int m, int n,
                                                         size_t col_bytes, int *__restrict__ counts)
{
    const int lane = threadIdx.x & 31;
    const int mk = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (mk >= m) return;
    const unsigned char *col = bed + (size_t)mk * col_bytes;
    const size_t full = (size_t)n / 4;  // bytes whose four samples all exist
    int plo = 0, phi = 0, pand = 0;
    size_t head = (size_t)((16u - (unsigned)((uintptr_t)col & 15u)) & 15u);
    if (head > full) head = full;
    if ((size_t)lane < head) bed_acc_word(col[lane], plo, phi, pand);
    const uint4 *v = (const uint4 *)(col + head);
    const size_t nvec = (full - head) / 16;
    size_t i = lane;
    for (; i + 32 < nvec; i += 64) {
        const uint4 a = __ldg(v + i), b = __ldg(v + i + 32);
        bed_acc_word(a.x, plo, phi, pand);
        bed_acc_word(a.y, plo, phi, pand);
        bed_acc_word(a.z, plo, phi, pand);
        bed_acc_word(a.w, plo, phi, pand);
        bed_acc_word(b.x, plo, phi, pand);
        bed_acc_word(b.y, plo, phi, pand);
        bed_acc_word(b.z, plo, phi, pand);
        bed_acc_word(b.w, plo, phi, pand);
    }
    if (i < nvec) {
        const uint4 a = __ldg(v + i);
        bed_acc_word(a.x, plo, phi, pand);
        bed_acc_word(a.y, plo, phi, pand);
        bed_acc_word(a.z, plo, phi, pand);
        bed_acc_word(a.w, plo, phi, pand);
    }
    const size_t done = head + nvec * 16;
    if (done + lane < full) bed_acc_word(col[done + lane], plo, phi, pand);  // fewer than 16 bytes left
    for (int o = 16; o > 0; o >>= 1) {
        plo += __shfl_xor_sync(0xffffffffu, plo, o);
        phi += __shfl_xor_sync(0xffffffffu, phi, o);
        pand += __shfl_xor_sync(0xffffffffu, pand, o);
    }
    if (lane == 0) {
        int c0 = pand, cm = plo - pand, c1 = phi - pand, c2 = (int)(4 * full) - (plo + phi - pand);
        const int ns = n - (int)(4 * full);  // samples in the last, partial byte
        if (ns > 0) {
            const unsigned v8 = col[full];
            const unsigned valid = (1u << (2 * ns)) - 1u;
            const unsigned lo = v8 & 0x55u, hi = (v8 >> 1) & 0x55u, smask = valid & 0x55u;
            c2 += __popc(~lo & ~hi & smask);  // 00
            cm += __popc(lo & ~hi & smask);   // 01
            c1 += __popc(~lo & hi & smask);   // 10
            c0 += __popc(lo & hi & smask);    // 11
        }
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

// The same statistics with the columns staged through shared memory (round 2).  In the kernel above every thread walks
// its own column byte by byte: 32 sectors per warp load, 40 GB/s at the C3 shape (28.9 ms for 1.15 GB,
// profiles/r2_corr_helpers_full_summary.txt).  Here a warp owns 32 markers: it fetches 128 bytes of each column with one
// coalesced load per column (32 independent loads in flight per lane, the next tile's while this one is summed), parks
// them in a padded shared tile and every lane then replays its own column's samples IN ORDER from there -- the float
// sum is the same sequence of additions (a missing call adds +0.0f to a sum that is never -0, which leaves it as it is),
// so the result is bit-identical; what remains is the chain of dependent float additions itself (~5 cycles per sample).
// Needs 4-byte aligned columns (ceil(n/4) % 4 == 0); other shapes take the kernel above.
__global__ void __launch_bounds__(32) bed_stats_tiled_kernel(const unsigned char *__restrict__ bed, int m, int n,
                                                            size_t col_bytes, const int *__restrict__ counts,
                                                            float *__restrict__ means, float *__restrict__ stds,
                                                            int *__restrict__ modes)
{
    __shared__ uint32_t s_t[32 * 33];
    const int lane = threadIdx.x;
    const int mk0 = blockIdx.x * 32, mk = mk0 + lane;
    const int rows = min(32, m - mk0);
    const size_t words = col_bytes / 4;
    float mean = 0.f, denom = 1.f;
    int c0 = 0, c1 = 0, c2 = 0;
    if (mk < m) {
        const int *c = counts + 4 * mk;
        c0 = c[0], c1 = c[1], c2 = c[2];
        const size_t gt_sum = (size_t)c1 + 2 * (size_t)c2;
        denom = (float)((size_t)n - (size_t)c[3]);
        mean = (float)gt_sum / denom;  // size_t / float: the sum converts to float first
    }
    const float sq_g0 = (0.0f - mean) * (0.0f - mean);  // code 11 -> 0
    const float sq_g1 = (1.0f - mean) * (1.0f - mean);  // code 10 -> 1
    const float sq_g2 = (2.0f - mean) * (2.0f - mean);  // code 00 -> 2 (code 01 = missing: adds +0)
    const uint32_t *base = (const uint32_t *)(bed + (size_t)mk0 * col_bytes);
    uint32_t nxt[32];
#pragma unroll
    for (int r = 0; r < 32; r++) nxt[r] = (r < rows && (size_t)lane < words) ? __ldg(base + (size_t)r * words + lane) : 0x55555555u;
    float sum_squares = 0.0f;
    for (size_t t0 = 0; t0 < words; t0 += 32) {
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 32; r++) s_t[r * 33 + lane] = nxt[r];
        __syncwarp();
        const size_t t1 = t0 + 32 + lane;
#pragma unroll
        for (int r = 0; r < 32; r++) nxt[r] = (r < rows && t1 < words) ? __ldg(base + (size_t)r * words + t1) : 0x55555555u;
        for (int j = 0; j < 32; j++) {
            uint32_t w = s_t[lane * 33 + j];
            const long long left = (long long)n - (long long)(t0 + j) * 16;  // samples of this word that exist
            if (left < 16) {
                const uint32_t keep = left <= 0 ? 0u : ((1u << (2 * (int)left)) - 1u);
                w = (w & keep) | (0x55555555u & ~keep);  // what does not exist counts as missing
            }
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const bool b0 = (w >> (2 * k)) & 1u, b1 = (w >> (2 * k + 1)) & 1u;
                const float add = b1 ? (b0 ? sq_g0 : sq_g1) : (b0 ? 0.0f : sq_g2);
                sum_squares = __fadd_rn(sum_squares, add);
            }
        }
    }
    if (mk < m) {
        means[mk] = mean;
        stds[mk] = sqrtf(sum_squares / denom);
        int mode = 0, best = c0;
        if (c1 > best) mode = 1, best = c1;
        if (c2 > best) mode = 2;
        modes[mk] = mode;
    }
}

// ---------------------------------------------------------------------------------------------
// 2-bit BED -> uint8 indicator rows for the tensor-core Gram.  Thread = 4 packed bytes (16 samples)
// of one marker -> one 16-byte store per part.  Bytes: m ceil(n/4) read, parts * m * Kp written.
// Round 2: one 32-bit load per thread where the columns are 4-byte aligned, and the indicators of four samples at a time
// by byte-parallel logic (the 2-bit codes of a packed byte spread to one byte each, then lo / hi bit planes): the
// bit-at-a-time first version was issue-bound at 2.5 TB/s of its 10.3 GB (profiles/r2_corr_helpers_full_summary.txt).
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
    if ((((uintptr_t)col) & 3u) == 0 && chunk * 4 + 4 <= col_bytes) {
        v = __ldg((const uint32_t *)(col + chunk * 4));
    } else {
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const size_t bi = chunk * 4 + b;
            if (bi < col_bytes) v |= (unsigned)col[bi] << (8 * b);
        }
    }
    uint32_t o1[4], o2[4], ov[4];
    const long long left = (long long)n - (long long)chunk * 16;  // samples of this chunk that exist
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const unsigned byte = (v >> (8 * k)) & 0xffu;
        const unsigned x = (byte | (byte << 6) | (byte << 12) | (byte << 18)) & 0x03030303u;  // one code per byte
        const unsigned lo = x & 0x01010101u, hi = (x >> 1) & 0x01010101u;
        uint32_t is2 = hi & (lo ^ 0x01010101u);          // code 10
        uint32_t is0 = (hi | lo) ^ 0x01010101u;          // code 00
        uint32_t ok = (lo & (hi ^ 0x01010101u)) ^ 0x01010101u;  // code != 01
        if (left < 16) {
            const long long e = left - 4 * k;  // samples of this byte that exist
            const uint32_t keep = e >= 4 ? 0xffffffffu : (e <= 0 ? 0u : ((1u << (8 * (int)e)) - 1u));
            is2 &= keep, is0 &= keep, ok &= keep;
        }
        o1[k] = is2, o2[k] = is0, ov[k] = ok;
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
__global__ void __launch_bounds__(1024) pxp_kernel(const float *__restrict__ phen, int n, int p,
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
    // 1024 threads, four samples of both columns in flight per thread (the 256-thread, one-at-a-time first version was
    // latency-bound: 0.82 ms for 136 pairs x 458 000 samples at 5 % issue utilisation, profiles/r2_corr_helpers_full_summary.txt)
    const int B = blockDim.x;
    int i = threadIdx.x;
    for (; i + 3 * B < n; i += 4 * B) {
        float va[4], vb[4];
#pragma unroll
        for (int u = 0; u < 4; u++) va[u] = __ldg(ya + i + u * B), vb[u] = __ldg(yb + i + u * B);
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (va[u] == va[u] && vb[u] == vb[u]) {
                s += (double)va[u] * (double)vb[u];
                cnt++;
            }
    }
    for (; i < n; i += B) {
        const float va = ya[i], vb = yb[i];
        if (va == va && vb == vb) {
            s += (double)va * (double)vb;
            cnt++;
        }
    }
    __shared__ double ss[32];
    __shared__ int sc[32];
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
        for (int w = 0; w < (B >> 5); w++) {
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
    if (col_bytes % 4 == 0 && ((uintptr_t)bed & 3u) == 0)
        bed_stats_tiled_kernel<<<(m + 31) / 32, 32, 0, st>>>(bed, m, n, col_bytes, counts, means, stds, modes);
    else
        bed_stats_kernel<<<(m + 127) / 128, 128, 0, st>>>(bed, m, n, col_bytes, counts, means, stds, modes);
}
// The rows of the padded indicator matrix that no marker owns must be zero: the padding rows of every 128-row group
// (parts = 3 uses 126), the unused part of the last group and every row from groups * 128 on (up to the multiple of 256,
// plus whatever block the caller appended).  All other rows -- the tail of K included -- are written by
// decode_indicators_kernel itself.
cudaError_t clear_indicator_padding(cudaStream_t st, unsigned char *ind, size_t m, int parts, int MI, size_t Kp,
                                    size_t rows_total)
{
    const size_t groups = (m + MI - 1) / MI, used = (size_t)MI * parts;
    cudaError_t e = cudaSuccess;
    if (groups && used < 128) e = cudaMemset2DAsync(ind + used * Kp, 128 * Kp, 0, (128 - used) * Kp, groups, st);
    const size_t last_markers = groups ? m - (groups - 1) * (size_t)MI : (size_t)MI;
    if (e == cudaSuccess && last_markers < (size_t)MI)
        e = cudaMemsetAsync(ind + ((groups - 1) * 128 + last_markers * parts) * Kp, 0, (used - last_markers * parts) * Kp, st);
    if (e == cudaSuccess && rows_total > groups * 128)
        e = cudaMemsetAsync(ind + groups * 128 * Kp, 0, (rows_total - groups * 128) * Kp, st);
    return e;
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
    pxp_kernel<<<p * (p - 1) / 2, 1024, 0, st>>>(phen, n, p, out_upper, Csq, ldc, m_off);
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
