// corr_banded.cu -- banded marker x marker correlations of one chromosome and their row |r| sums: the device part
// of `mps block`.  Replaces cal_mcorrk_banded (src/corr_host.cu:64-110), its per-row launch loop
// cu_marker_corr_pearson_npn_batched_sparse (:1199-1319, ONE launch of bed_marker_corr_pearson_npn_sparse_scan per
// marker, src/corr_kernels.cu:920-1010) and marker_corr_banded_mat_row_abs_sums (:112-128).
//
// Layout in HBM, per chunk of markers held on the device:
//   bed     [Mc][col_bytes]      packed columns as in the file
//   planes  [Mc][3][Kw] uint32   indicator bit planes (banded_core.h), Kw = plane words padded to a multiple of 32
//   counts  [Mc][4] int32        n0, n1, n2, nmissing per marker (bed_counts_kernel)
//   out     [R][W] float         out[row][k] = correlation of (row, row + 1 + k); 0 beyond the last marker
//   sums    [R] float            sum_k |out[row][k]|, accumulated left to right in float like the reference
// Work: integer AND + POPC on CUDA cores, 4 (no missing call in the tile) or 9 joint counts per 32 samples of a pair,
// operands staged through shared memory as 32 x 32 marker tiles so that every plane word fetched from L2 serves 32
// pairs.  Algorithmic bytes: the chunk's planes once per tile row/column it takes part in (L2 resident), 4 W bytes
// written per row.  The tensor-core route (band-restricted tile list on the tcgen05 indicator Gram of
// corr_gram.cu, banded_chunk_device_tc below) gives the same floats 4.7x faster (8 192 markers x 65 536 samples,
// corr_width 512: 1.90 ms against 8.88 ms, profiles/r2_bench_block_tc.json) and is the default from
// CUSK_BLOCK_TC_MIN_MARKERS (512) markers on; CUSK_BLOCK_TC=0 / 1 force one route.
#include "banded_core.h"
#include "corr.cuh"

#include <cuda.h>

#include <vector>

void launch_bed_counts(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int *counts);
// tensor-core route (corr_kernels.cu, corr_gram.cu)
void launch_decode(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int parts, int MI,
                   size_t Kp, unsigned char *ind);
cudaError_t clear_indicator_padding(cudaStream_t st, unsigned char *ind, size_t m, int parts, int MI, size_t Kp,
                                    size_t rows_total);
int launch_gram(cudaStream_t st, const CUtensorMap &map_a, const CUtensorMap &map_b, const GramParams &p, int grid);
int gram_make_maps(const unsigned char *ind, size_t rows, size_t Kp, CUtensorMap *map_a, CUtensorMap *map_b);
cusk_ctx *cusk_default_ctx();

#define BT 32        // markers per tile edge
#define BKC 32       // plane words per shared-memory stage (1024 samples)
#define BPSTR 36     // words between the planes of one marker in shared memory (16-byte aligned rows)
#define BMSTR (3 * BPSTR)  // words between markers: 108 = 12 mod 32 -> 8 lanes x 16 B hit 8 distinct bank groups

__global__ void __launch_bounds__(256) bed_planes_kernel(const unsigned char *__restrict__ bed, size_t col_bytes,
                                                         size_t n, int Mc, size_t Kw, uint32_t *__restrict__ planes)
{
    const size_t total = (size_t)Mc * Kw;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t mk = i / Kw, w = i % Kw;
        uint32_t pl[3];
        bandc_planes(bed + mk * col_bytes, col_bytes, n, w, pl);
        uint32_t *dst = planes + mk * 3 * Kw + w;
        dst[0] = pl[0];
        dst[Kw] = pl[1];
        dst[2 * Kw] = pl[2];
    }
}

__device__ __forceinline__ int popc4(const uint4 &a, const uint4 &b)
{
    return __popc(a.x & b.x) + __popc(a.y & b.y) + __popc(a.z & b.z) + __popc(a.w & b.w);
}

// One 32 x 32 tile: rows a0.., columns b0.. (marker indices local to the chunk).  Thread (tx, ty) of the 16 x 16
// layout owns rows {ty, ty + 16} x columns {tx, tx + 16}.  MISS: some marker of the tile has a missing call -> all
// nine joint counts; otherwise the four counts of genotypes {1,2} x {1,2} and the markers' own genotype counts.
template <bool MISS>
__device__ __forceinline__ void band_tile(const uint32_t *__restrict__ planes, size_t Kw, const int *__restrict__ counts,
                                          int Mc, int R, int W, int n_samples, int a0, int b0, uint32_t *sA, uint32_t *sB,
                                          float *__restrict__ out)
{
    constexpr int NP = MISS ? 3 : 2;   // planes staged: (0,1,2) or (1,2)
    constexpr int NC = NP * NP;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    int cnt[2][2][NC];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int c = 0; c < NC; c++) cnt[i][j][c] = 0;

    for (size_t k0 = 0; k0 < Kw; k0 += BKC) {
        __syncthreads();  // the previous stage has been consumed
        for (int it = threadIdx.x; it < 2 * BT * NP * (BKC / 4); it += 256) {
            const int q = it & 7;
            const int pl = (it >> 3) % NP;
            const int mkl = it / (8 * NP);  // 0..63: rows then columns
            const int r = mkl & (BT - 1);
            const bool colside = mkl >= BT;
            const int mk = (colside ? b0 : a0) + r;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (mk < Mc) {
                const int plane = MISS ? pl : pl + 1;
                v = __ldg(reinterpret_cast<const uint4 *>(planes + ((size_t)mk * 3 + plane) * Kw + k0) + q);
            }
            *reinterpret_cast<uint4 *>((colside ? sB : sA) + r * BMSTR + pl * BPSTR + 4 * q) = v;
        }
        __syncthreads();
#pragma unroll 2
        for (int q = 0; q < BKC / 4; q++) {
            uint4 a[2][NP], b[2][NP];
#pragma unroll
            for (int i = 0; i < 2; i++)
#pragma unroll
                for (int p = 0; p < NP; p++) {
                    a[i][p] = *reinterpret_cast<const uint4 *>(sA + (ty + 16 * i) * BMSTR + p * BPSTR + 4 * q);
                    b[i][p] = *reinterpret_cast<const uint4 *>(sB + (tx + 16 * i) * BMSTR + p * BPSTR + 4 * q);
                }
#pragma unroll
            for (int i = 0; i < 2; i++)
#pragma unroll
                for (int j = 0; j < 2; j++)
#pragma unroll
                    for (int pa = 0; pa < NP; pa++)
#pragma unroll
                        for (int pb = 0; pb < NP; pb++) cnt[i][j][pa * NP + pb] += popc4(a[i][pa], b[j][pb]);
        }
    }

#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int row = a0 + ty + 16 * i, col = b0 + tx + 16 * j;
            if (row >= R || col >= Mc || col <= row || col - row > W) continue;
            double s[9];
            if (MISS) {
#pragma unroll
                for (int c = 0; c < 9; c++) s[c] = (double)cnt[i][j][c];  // plane index = genotype: s[3a+b]
            } else {
                bandc_table_from_joint(cnt[i][j][0], cnt[i][j][1], cnt[i][j][2], cnt[i][j][3], counts + 4 * row,
                                       counts + 4 * col, n_samples, s);
            }
            out[(size_t)row * W + (col - row - 1)] = (float)npn_from_table(s);
        }
}

// grid (ndj, row tiles): CTA (dj, ti) computes the tile of rows [32 ti, +32) x columns [32 (ti + dj), +32)
__global__ void __launch_bounds__(256, 2) banded_npn_kernel(const uint32_t *__restrict__ planes, size_t Kw,
                                                         const int *__restrict__ counts, int Mc, int R, int W,
                                                         int n_samples, float *__restrict__ out)
{
    __shared__ __align__(16) uint32_t sA[BT * BMSTR];
    __shared__ __align__(16) uint32_t sB[BT * BMSTR];
    __shared__ int s_miss;
    const int a0 = blockIdx.y * BT, b0 = (blockIdx.y + blockIdx.x) * BT;
    if (b0 >= Mc) return;
    if (threadIdx.x == 0) s_miss = 0;
    __syncthreads();
    if (threadIdx.x < 2 * BT) {
        const int mk = threadIdx.x < BT ? a0 + threadIdx.x : b0 + threadIdx.x - BT;
        if (mk < Mc && counts[4 * mk + 3] != 0) s_miss = 1;  // every writer stores the same value
    }
    __syncthreads();
    if (s_miss)
        band_tile<true>(planes, Kw, counts, Mc, R, W, n_samples, a0, b0, sA, sB, out);
    else
        band_tile<false>(planes, Kw, counts, Mc, R, W, n_samples, a0, b0, sA, sB, out);
}

// float accumulation in column order, exactly marker_corr_banded_mat_row_abs_sums (src/corr_host.cu:112-128)
__global__ void __launch_bounds__(128) band_row_abs_sums_kernel(const float *__restrict__ out, int R, int W,
                                                                float *__restrict__ sums)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= R) return;
    const float *p = out + (size_t)row * W;
    float s = 0.0f;
    for (int k = 0; k < W; k++) s = __fadd_rn(s, fabsf(p[k]));
    sums[row] = s;
}

namespace {

struct DevBuf {  // chromosome-sized scratch: allocated per call, not from the context's grow-only arenas
    void *p = nullptr;
    ~DevBuf()
    {
        if (p) cudaFree(p);
    }
};

struct Events4 {
    cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr};
    ~Events4()
    {
        for (auto &x : e)
            if (x) cudaEventDestroy(x);
    }
};

}  // namespace

// Markers [0, Mc) of the chunk are on the device in d_bed; computes rows [0, R).
static int banded_chunk_device(cusk_ctx *ctx, const unsigned char *d_bed, int Mc, int R, size_t n, int W, size_t Kw,
                               uint32_t *d_planes, int *d_counts, float *d_out, float *d_sums)
{
    cudaStream_t s = ctx->stream;
    const size_t col_bytes = (n + 3) / 4;
    launch_bed_counts(s, d_bed, Mc, (int)n, col_bytes, d_counts);
    const size_t words = (size_t)Mc * Kw;
    const int pgrid = (int)std::min<size_t>((words + 255) / 256, (size_t)ctx->sm_count * 32);
    bed_planes_kernel<<<pgrid, 256, 0, s>>>(d_bed, col_bytes, n, Mc, Kw, d_planes);
    CUSK_CUDA(cudaMemsetAsync(d_out, 0, sizeof(float) * (size_t)R * W, s));
    const dim3 grid((unsigned)((W + BT - 1) / BT + 1), (unsigned)((R + BT - 1) / BT));
    banded_npn_kernel<<<grid, 256, 0, s>>>(d_planes, Kw, d_counts, Mc, R, W, (int)n, d_out);
    if (d_sums) band_row_abs_sums_kernel<<<(R + 127) / 128, 128, 0, s>>>(d_out, R, W, d_sums);
    CUSK_CUDA(cudaGetLastError());
    return CUSK_OK;
}

// Tensor-core route (default for 512 markers and more, see the file header): the chunk's columns become uint8 indicator rows (decode_indicators_kernel), the tcgen05 int8 Gram kernel
// of corr_gram.cu runs over the tiles that intersect the band (bandc_gram_tiles) and its epilogue writes band
// coordinates.  Same exact integer tables and the same FP64 npn_from_table as banded_npn_kernel -> same floats.
// Operand rows per marker: 2 when the chunk has no missing call, else 3 (a host look at the counts decides).
static size_t tc_ind_rows(size_t markers, size_t MI) { return ((((markers + MI - 1) / MI) * 128 + 255) / 256) * 256; }

static int banded_chunk_device_tc(cusk_ctx *ctx, const unsigned char *d_bed, int Mc, int R, size_t n, int W,
                                  unsigned char *d_ind, int *d_counts, int2 *d_tiles, size_t tiles_cap,
                                  std::vector<int> &h_scratch, float *d_out, float *d_sums)
{
    cudaStream_t s = ctx->stream;
    const size_t col_bytes = (n + 3) / 4, Kp = align_up(n, 128);
    launch_bed_counts(s, d_bed, Mc, (int)n, col_bytes, d_counts);
    h_scratch.resize(std::max<size_t>(4 * (size_t)Mc, 2 * tiles_cap));
    CUSK_CUDA(cudaMemcpyAsync(h_scratch.data(), d_counts, sizeof(int) * 4 * (size_t)Mc, cudaMemcpyDeviceToHost, s));
    CUSK_CUDA(cudaStreamSynchronize(s));
    bool any_missing = false;
    for (int i = 0; i < Mc; i++) any_missing |= h_scratch[4 * (size_t)i + 3] != 0;
    const int parts = any_missing ? 3 : 2, MI = 128 / parts;
    const size_t rows = tc_ind_rows((size_t)Mc, (size_t)MI);
    CUSK_CUDA(clear_indicator_padding(s, d_ind, (size_t)Mc, parts, MI, Kp, rows));  // only the rows no marker owns
    launch_decode(s, d_bed, Mc, (int)n, col_bytes, parts, MI, Kp, d_ind);
    const size_t ntiles = bandc_gram_tiles((size_t)Mc, (size_t)R, (size_t)W, (size_t)MI, h_scratch.data(), tiles_cap);
    if (ntiles > tiles_cap) {
        cusk_set_error("cusk_mcorrk_banded: tile list overflow (%zu > %zu)", ntiles, tiles_cap);
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaMemcpyAsync(d_tiles, h_scratch.data(), sizeof(int2) * ntiles, cudaMemcpyHostToDevice, s));
    CUSK_CUDA(cudaMemsetAsync(d_out, 0, sizeof(float) * (size_t)R * W, s));
    CUtensorMap map_a, map_b;
    int rc = gram_make_maps(d_ind, rows, Kp, &map_a, &map_b);
    if (rc) return rc;
    GramParams gp = {};
    gp.parts = parts;
    gp.MI = MI;
    gp.m = Mc;
    gp.n_samples = (int)n;
    gp.num_k = (int)(Kp / 128);
    gp.ntiles = (int)ntiles;
    gp.tiles = d_tiles;
    gp.counts = d_counts;
    gp.mode = 0;
    gp.trait_tn = -1;
    gp.band = d_out;
    gp.band_w = W;
    gp.band_rows = R;
    rc = launch_gram(s, map_a, map_b, gp, std::min<int>(ctx->sm_count, (int)ntiles));
    if (rc) return rc;
    if (d_sums) band_row_abs_sums_kernel<<<(R + 127) / 128, 128, 0, s>>>(d_out, R, W, d_sums);
    CUSK_CUDA(cudaGetLastError());
    CUSK_CUDA(cudaStreamSynchronize(s));  // h_scratch (the host tile list) is reused by the next chunk
    return CUSK_OK;
}

extern "C" int cusk_mcorrk_banded(cusk_ctx *ctx, const unsigned char *bed, size_t m, size_t n, size_t corr_width,
                                  float device_mem_gb, float *corrs, float *row_abs_sums, cusk_band_stats *stats)
{
    if (!ctx || !bed || (!corrs && !row_abs_sums)) {
        cusk_set_error("cusk_mcorrk_banded: NULL argument");
        return CUSK_E_ARG;
    }
    if (m == 0 || n == 0 || corr_width == 0 || corr_width >= m || m > (size_t)1 << 30 || n > (size_t)1 << 30) {
        // the reference exits when the batch (at most m markers) is smaller than corr_width (src/corr_host.cu:82-91) and
        // reads past its buffers when corr_width == m (:1224, size_t underflow of num_markers - corr_width - 1)
        cusk_set_error("cusk_mcorrk_banded: need 0 < corr_width < num_markers (got corr_width %zu, %zu markers, %zu samples)",
                       corr_width, m, n);
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    cusk_band_stats local;
    cusk_band_stats *st = stats ? stats : &local;
    memset(st, 0, sizeof(*st));
    const size_t col_bytes = (n + 3) / 4;
    const size_t Kw = align_up((n + 31) / 32, BKC);
    const size_t W = corr_width;
    // device bytes per marker held (bed + planes + counts) and per row computed (band + sum)
    const char *tc_env = getenv("CUSK_BLOCK_TC");
    const char *tc_min = getenv("CUSK_BLOCK_TC_MIN_MARKERS");
    const size_t tc_min_markers = (tc_min && atoll(tc_min) > 0) ? (size_t)atoll(tc_min) : 512;
    const bool tc = (tc_env && *tc_env) ? atoi(tc_env) != 0 : m >= tc_min_markers;
    const size_t Kp = align_up(n, 128);
    // tensor-core route: uint8 indicator rows, at most 3 per marker with 42 markers to a 128-row group, plus one
    // 256-row block of padding; 8 bytes of tile list per marker are more than the band ever needs per group
    const size_t per_marker = tc ? col_bytes + (128 * Kp + 41) / 42 + 16 + 8 * (W / 84 + 4) : col_bytes + 12 * Kw + 16;
    const size_t per_row = 4 * W + 4;
    size_t free_b = 0, total_b = 0;
    CUSK_CUDA(cudaMemGetInfo(&free_b, &total_b));
    double budget = 0.9 * (double)free_b;
    if (device_mem_gb > 0.f) budget = std::min(budget, (double)device_mem_gb * 1e9);
    if (tc) budget -= 384.0 * (double)Kp;
    size_t batch = budget > 0 ? (size_t)(budget / (double)(per_marker + per_row)) : 0;
    if (batch > m) batch = m;
    if (tc && batch > 65535) batch = 65535;  // decode_indicators_kernel takes the marker from blockIdx.y
    if (batch < m && batch <= W) {  // a chunk holds its rows plus the W markers to their right
        cusk_set_error("Maximal batch size (%zu) <= corr width (%zu). Decrease distance threshold or increase device memory.",
                       batch, W);
        return CUSK_E_ARG;
    }
    const size_t rows_per_chunk = batch == m ? m : batch - W;
    DevBuf b_bed, b_planes, b_counts, b_out, b_sums;
    CUSK_CUDA(cudaMalloc(&b_bed.p, batch * col_bytes));
    const size_t tiles_cap = tc ? batch * (W / 84 + 4) : 0;  // per 42-marker group at most W / 84 + 3 B blocks
    DevBuf b_tiles;
    std::vector<int> h_scratch;
    if (tc) {
        CUSK_CUDA(cudaMalloc(&b_planes.p, tc_ind_rows(batch, 42) * Kp));
        CUSK_CUDA(cudaMalloc(&b_tiles.p, std::max<size_t>(tiles_cap, 1) * sizeof(int2)));
    } else
        CUSK_CUDA(cudaMalloc(&b_planes.p, batch * 12 * Kw));
    CUSK_CUDA(cudaMalloc(&b_counts.p, batch * 16));
    CUSK_CUDA(cudaMalloc(&b_out.p, rows_per_chunk * W * sizeof(float)));
    CUSK_CUDA(cudaMalloc(&b_sums.p, rows_per_chunk * sizeof(float)));
    Events4 evs;
    cudaEvent_t *ev = evs.e;
    for (auto &e : evs.e) CUSK_CUDA(cudaEventCreate(&e));
    int rc = CUSK_OK;
    for (size_t r0 = 0; r0 < m && rc == CUSK_OK; r0 += rows_per_chunk) {
        const size_t R = std::min(rows_per_chunk, m - r0);
        const size_t Mc = std::min(m - r0, R + W);
        cudaEventRecord(ev[0], s);
        rc = cudaMemcpyAsync(b_bed.p, bed + r0 * col_bytes, Mc * col_bytes, cudaMemcpyHostToDevice, s) == cudaSuccess
                 ? CUSK_OK
                 : CUSK_E_CUDA;
        if (rc) {
            cusk_set_error("cusk_mcorrk_banded: host-to-device copy failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        cudaEventRecord(ev[1], s);
        if (tc)
            rc = banded_chunk_device_tc(ctx, (const unsigned char *)b_bed.p, (int)Mc, (int)R, n, (int)W,
                                        (unsigned char *)b_planes.p, (int *)b_counts.p, (int2 *)b_tiles.p, tiles_cap,
                                        h_scratch, (float *)b_out.p, row_abs_sums ? (float *)b_sums.p : nullptr);
        else
            rc = banded_chunk_device(ctx, (const unsigned char *)b_bed.p, (int)Mc, (int)R, n, (int)W, Kw,
                                     (uint32_t *)b_planes.p, (int *)b_counts.p, (float *)b_out.p,
                                     row_abs_sums ? (float *)b_sums.p : nullptr);
        if (rc) break;
        cudaEventRecord(ev[2], s);
        if (corrs) cudaMemcpyAsync(corrs + r0 * W, b_out.p, sizeof(float) * R * W, cudaMemcpyDeviceToHost, s);
        if (row_abs_sums) cudaMemcpyAsync(row_abs_sums + r0, b_sums.p, sizeof(float) * R, cudaMemcpyDeviceToHost, s);
        cudaEventRecord(ev[3], s);
        if (cudaStreamSynchronize(s) != cudaSuccess) {
            cusk_set_error("cusk_mcorrk_banded: %s", cudaGetErrorString(cudaGetLastError()));
            rc = CUSK_E_CUDA;
            break;
        }
        float t = 0.f;
        cudaEventElapsedTime(&t, ev[0], ev[1]);
        st->h2d_ms += t;
        cudaEventElapsedTime(&t, ev[1], ev[2]);
        st->kernel_ms += t;
        cudaEventElapsedTime(&t, ev[2], ev[3]);
        st->d2h_ms += t;
        cudaEventElapsedTime(&t, ev[0], ev[3]);
        st->total_ms += t;
        st->chunks += 1;
        st->bed_bytes += (int64_t)(Mc * col_bytes);
    }
    // pairs inside the band: row r has min(W, m - 1 - r) of them
    st->pairs = (int64_t)((m - W) * W + W * (W - 1) / 2);
    return rc;
}

// src/corr_host.cu:1199 (C++ linkage in the reference): batch_size only bounded the reference's device buffers
extern "C" void cu_marker_corr_pearson_npn_batched_sparse(const unsigned char *marker_vals, const size_t num_markers,
                                                          const size_t num_individuals, const size_t corr_width,
                                                          const size_t batch_size, float *corrs)
{
    (void)batch_size;
    if (cusk_mcorrk_banded(cusk_default_ctx(), marker_vals, num_markers, num_individuals, corr_width, 0.f, corrs, nullptr,
                           nullptr) != CUSK_OK) {
        fprintf(stderr, "cu_marker_corr_pearson_npn_batched_sparse: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);  // the reference's HANDLE_ERROR / CudaCheckError exit the process (include/mps/gpuerrors.h:12,34)
    }
}
