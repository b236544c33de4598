// corr_gram.cu -- genotype indicator Gram on the 5th-generation tensor cores (sm_100a):
//   D[r][c] = sum_k Ind[r][k] * Ind[c][k]      (uint8 0/1 operands, exact int32 accumulation)
// where Ind holds, per marker, `parts` indicator rows over the samples (parts = 2: [g==1],[g==2]
// when the block has no missing call; parts = 3: [valid],[g==1],[g==2] otherwise).  The 3x3
// contingency table of every marker pair -- what bed_marker_corr_pearson_npn_scan accumulates with
// LUT decodes and float adds (src/corr_kernels.cu:478-565) -- follows from 4 (or 9) entries of D
// and the marker marginals, and the epilogue turns it into Kendall tau_b and sin(pi/2 tau) in FP64.
//
// Kernel anatomy (one CTA per SM, persistent over upper-triangle tiles of 128 x 256 indicator rows):
//   warp 0   TMA producer: cp.async.bulk.tensor 2D boxes (128 B of K x 128 / 256 rows, SWIZZLE_128B)
//            into a 4-stage shared-memory ring, completion on mbarriers
//   warp 1   MMA issuer: one elected thread issues tcgen05.mma.cta_group::1.kind::i8 (M=128, N=256,
//            K=32 per instruction), accumulators in TMEM (2 x 256 columns, double buffered),
//            tcgen05.commit releases ring slots / publishes the accumulator
//   warp 2   TMEM allocation
//   (CTA-pair form, gram_i8_kernel<true>: two CTAs of a cluster on a 256 x 256 tile, cta_group::2 -- see the kernel)
//   warps 4-7 epilogue: tcgen05.ld the accumulator, stage 16 markers' columns through shared
//            memory, contingency table -> tau_b -> sin, write both triangles of the dense matrix
#include <cuda.h>

#include "corr.cuh"

namespace {

constexpr int BK = 128;                 // bytes of K per stage (= one 128 B swizzle span)
constexpr int TILE_M = 128, TILE_N = 256;
constexpr int A_BYTES = TILE_M * BK;
constexpr int EPI_MARKERS = 16;         // markers (column side) staged per epilogue chunk
constexpr int EPI_STRIDE = 16 * 3 + 1;  // ints per staged row (worst case parts = 3), padded
constexpr int EPI_BYTES = TILE_M * EPI_STRIDE * 4;
constexpr int GRAM_THREADS = 256;
// One CTA per tile (PAIR = false): a stage holds the tile's 128 A rows and 256 B rows, 4 stages.  CTA pair
// (PAIR = true, cta_group::2): a stage holds this CTA's 128 A rows and ITS HALF of the 256 B rows, 6 stages.
template <bool PAIR>
struct GramGeom {
    static constexpr int STAGES = PAIR ? 6 : 4;
    static constexpr int B_BYTES = (PAIR ? TILE_N / 2 : TILE_N) * BK;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // shared::cluster address of the same location in the pair's even CTA

// ---- PTX helpers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// CTA-pair forms: the load lands in THIS CTA's shared memory and completes on the even CTA's barrier
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar) & PEER_MASK), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_even_cta(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & PEER_MASK) : "memory");
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// the same arrival on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void tc_commit_pair(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
}
// D[tmem of both CTAs] (+)= A[128 rows in each CTA] * B[128 rows in each CTA]^T: M = 256, N = 256 over the pair
__device__ __forceinline__ void mma_i8_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, uint8 x uint8 -> int32
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                       uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile, 128 B rows, SWIZZLE_128B: 8-row groups are 1024 B apart.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);      // start address, bits [0,14)
    d |= (uint64_t)0 << 16;                       // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                       // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int *v)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace

// ---------------------------------------------------------------------------------------------
// The kernel
// ---------------------------------------------------------------------------------------------
template <bool PAIR>
__global__ void __launch_bounds__(GRAM_THREADS, 1)
gram_i8_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const GramParams p)
{
    constexpr int STAGES = GramGeom<PAIR>::STAGES, STAGE_BYTES = GramGeom<PAIR>::STAGE_BYTES;
    // CTA pair: the even CTA of the cluster issues the MMAs for both and owns the barriers the producers / the
    // accumulator hand-back complete on; tile entries 2c, 2c + 1 of the list are the pair's two A groups of one B block
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    extern __shared__ unsigned char smem_raw[];
    // SWIZZLE_128B operand tiles need 1024 B alignment
    unsigned char *smem = (unsigned char *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *stage_base = smem;
    int *epi = (int *)(smem + STAGES * STAGE_BYTES);
    uint64_t *bars = (uint64_t *)(smem + STAGES * STAGE_BYTES + EPI_BYTES);
    uint64_t *full = bars;                 // [STAGES]
    uint64_t *empty = bars + STAGES;       // [STAGES]
    uint64_t *tfull = bars + 2 * STAGES;   // [2]
    uint64_t *tempty = tfull + 2;          // [2]
    uint32_t *tmem_slot = (uint32_t *)(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; i++) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], PAIR ? 8 : 4);  // one arrival per epilogue warp (of both CTAs)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        if constexpr (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "n"(512)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "n"(512)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // both CTAs' barriers exist before a load or an arrival crosses over
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // instruction descriptor: D = S32, A/B = unsigned 8 bit, both K-major, N = 256, M = 128 (256 over a CTA pair)
    const uint32_t idesc = (2u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(TILE_N >> 3) << 17) |
                           ((uint32_t)((PAIR ? 2 * TILE_M : TILE_M) >> 4) << 24);

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
                const int2 tile = p.tiles[t];
                if (tile.x < 0) continue;  // padding entry of a partial wave
                const int wave = t / (int)gridDim.x;
                for (int kb = 0; kb < p.num_k; kb++) {
                    if (p.ksync && (kb % p.ksync) == 0) {
                        // every producer of the wave is about to start this K panel: keeps the wave's shared operand
                        // panels inside the L2 window (without it the CTAs drift apart by thousands of k-steps)
                        int *c = p.sync + (size_t)wave * p.ksync_blocks + kb / p.ksync;
                        const int need = p.wave_tiles[wave];
                        atomicAdd(c, 1);
                        while (*(volatile int *)c < need) __nanosleep(64);
                    }
                    mbar_wait(&empty[stage], phase ^ 1);
                    unsigned char *sa = stage_base + stage * STAGE_BYTES;
                    if constexpr (PAIR) {
                        // both CTAs load (their A group, their half of the B block: the 128-row box of map_a serves
                        // both); the bytes of all four loads complete on the even CTA's barrier
                        if (leader) mbar_arrive_expect_tx(&full[stage], 2 * STAGE_BYTES);
                        tma_load_2d_pair(sa, &map_a, &full[stage], kb * BK, tile.x * TILE_M);
                        tma_load_2d_pair(sa + A_BYTES, &map_a, &full[stage], kb * BK,
                                         tile.y * TILE_N + (int)rank * (TILE_N / 2));
                    } else {
                        mbar_arrive_expect_tx(&full[stage], STAGE_BYTES);
                        tma_load_2d(sa, &map_a, &full[stage], kb * BK, tile.x * TILE_M);
                        tma_load_2d(sa + A_BYTES, &map_b, &full[stage], kb * BK, tile.y * TILE_N);
                    }
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0 && leader) {
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t aphase = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
                if (p.tiles[t].x < 0) continue;
                mbar_wait(&tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(as * TILE_N);
                for (int kb = 0; kb < p.num_k; kb++) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(stage_base + stage * STAGE_BYTES);
                    const uint64_t da = make_desc(sa), db = make_desc(sa + A_BYTES);
#pragma unroll
                    for (int k4 = 0; k4 < BK / 32; k4++) {
                        // advance 32 bytes of K inside the 128 B swizzle span: +2 in 16 B units
                        if constexpr (PAIR)
                            mma_i8_pair(tmem_d, da + (uint64_t)(k4 * 2), db + (uint64_t)(k4 * 2), idesc,
                                        (uint32_t)((kb | k4) != 0));
                        else
                            mma_i8(tmem_d, da + (uint64_t)(k4 * 2), db + (uint64_t)(k4 * 2), idesc,
                                   (uint32_t)((kb | k4) != 0));
                    }
                    // frees the slot (in both CTAs of a pair) when these MMAs have read it
                    if constexpr (PAIR) tc_commit_pair(&empty[stage]); else tc_commit(&empty[stage]);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if constexpr (PAIR) tc_commit_pair(&tfull[as]); else tc_commit(&tfull[as]);  // accumulator complete
                if (++as == 2) {
                    as = 0;
                    aphase ^= 1;
                }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: 4 warps, warp q owns TMEM lanes [32q, 32q+32) =====
        const int q = warp - 4;
        const int et = threadIdx.x - 128;  // 0..127
        const int row = q * 32 + lane;     // accumulator row handled by this thread in tcgen05.ld
        const int parts = p.parts, MI = p.MI;
        int as = 0;
        uint32_t aphase = 0;
        for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
            const int2 tile = p.tiles[t];
            if (tile.x < 0) continue;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * TILE_N) + ((uint32_t)(q * 32) << 16);
            if (tile.y == p.trait_tn) {
                // ---- marker x trait tile: columns 8t..8t+4 = <row, limb_0..3>, <row, valid> of trait t ----
                // (replaces bed_marker_phen_corr_pearson_scan, src/corr_kernels.cu:157-238:
                //  r = (sum v g y - mean_g sum v y) / (sum v * std_g))
                double *sS = (double *)epi;           // [128 rows][2 traits] class sums of y
                int *sC = (int *)(sS + 2 * TILE_M);   // [128 rows][2 traits] class counts
                for (int c = 0; 2 * c < p.p; c++) {
                    int v[16];
                    tmem_ld16(tmem_d + (uint32_t)(c * 16), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int tt = 0; tt < 2; tt++) {
                        const int t = 2 * c + tt;
                        const long long sq = (long long)v[8 * tt] + ((long long)v[8 * tt + 1] << 8) +
                                             ((long long)v[8 * tt + 2] << 16) + ((long long)v[8 * tt + 3] << 24) -
                                             ((long long)v[8 * tt + 4] << 31);
                        sS[row * 2 + tt] = (t < p.p) ? (double)sq * p.trait_scale[t] : 0.0;
                        sC[row * 2 + tt] = v[8 * tt + 4];
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    {
                        const int il = et >> 1, tt = et & 1;
                        const int t = 2 * c + tt, i = tile.x * MI + il;
                        if (il < MI && i < p.m && t < p.p) {
                            double s1, s2, sy, sv;
                            if (parts == 2) {  // no missing call in the block: rows [g==1], [g==2]
                                s1 = sS[(2 * il) * 2 + tt];
                                s2 = sS[(2 * il + 1) * 2 + tt];
                                sy = p.Ty[t];
                                sv = (double)p.cnty[t];
                            } else {           // rows [valid], [g==1], [g==2]
                                sy = sS[(3 * il) * 2 + tt];
                                sv = (double)sC[(3 * il) * 2 + tt];
                                s1 = sS[(3 * il + 1) * 2 + tt];
                                s2 = sS[(3 * il + 2) * 2 + tt];
                            }
                            const double r = (s1 + 2.0 * s2 - (double)p.mean[i] * sy) / (sv * (double)p.stdv[i]);
                            p.mxp[(size_t)i * p.p + t] = (float)r;
                        }
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
            } else
            for (int g = 0; g < 2; g++) {                       // the two 128-row groups of the B block
                const int jbase_group = (2 * tile.y + g) * MI;  // first marker of the group
                if (jbase_group >= p.m) continue;
                for (int mc = 0; mc * EPI_MARKERS < MI; mc++) {  // 16 markers at a time
                    const int col0 = mc * EPI_MARKERS * parts;   // group-relative first column
                    // stage this thread's row: up to `parts` pieces of 16 columns
                    for (int pc = 0; pc < parts; pc++) {
                        const int c = col0 + pc * 16;
                        int v[16];
                        if (c < 128) {
                            tmem_ld16(tmem_d + (uint32_t)(g * 128 + c), v);
                            tmem_ld_wait();
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; i++) v[i] = 0;
                        }
#pragma unroll
                        for (int i = 0; i < 16; i++) epi[row * EPI_STRIDE + pc * 16 + i] = v[i];
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    // pairs (i in the A group, j in this chunk), j > i only
                    const int i0 = tile.x * MI;
                    for (int pq = et; pq < MI * EPI_MARKERS; pq += 128) {
                        const int il = pq / EPI_MARKERS, jl = pq % EPI_MARKERS;
                        const int i = i0 + il, j = jbase_group + mc * EPI_MARKERS + jl;
                        if (mc * EPI_MARKERS + jl >= MI || i >= p.m || j >= p.m || j <= i) continue;
                        if (p.band && (j - i > p.band_w || i >= p.band_rows)) continue;  // outside the band
                        const int *e = epi + (il * parts) * EPI_STRIDE + jl * parts;
                        double s[9];
                        if (parts == 2) {
                            const double s11 = e[0], s12 = e[1], s21 = e[EPI_STRIDE], s22 = e[EPI_STRIDE + 1];
                            const int *ci = p.counts + 4 * i, *cj = p.counts + 4 * j;
                            s[4] = s11;
                            s[5] = s12;
                            s[7] = s21;
                            s[8] = s22;
                            s[3] = (double)ci[1] - s11 - s12;
                            s[6] = (double)ci[2] - s21 - s22;
                            s[1] = (double)cj[1] - s11 - s21;
                            s[2] = (double)cj[2] - s12 - s22;
                            s[0] = (double)p.n_samples - (s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7] + s[8]);
                        } else {
                            const double vv = e[0], v1 = e[1], v2 = e[2];
                            const double a1v = e[EPI_STRIDE], s11 = e[EPI_STRIDE + 1], s12 = e[EPI_STRIDE + 2];
                            const double a2v = e[2 * EPI_STRIDE], s21 = e[2 * EPI_STRIDE + 1],
                                         s22 = e[2 * EPI_STRIDE + 2];
                            s[4] = s11;
                            s[5] = s12;
                            s[7] = s21;
                            s[8] = s22;
                            s[3] = a1v - s11 - s12;
                            s[6] = a2v - s21 - s22;
                            s[1] = v1 - s11 - s21;
                            s[2] = v2 - s12 - s22;
                            s[0] = vv - (s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7] + s[8]);
                        }
                        double r;
                        if (p.mode == 0)
                            r = npn_from_table(s);
                        else
                            r = pearson_from_table(s, (double)p.mean[i], (double)p.mean[j], (double)p.stdv[i],
                                                   (double)p.stdv[j]);
                        const float rf = (float)r;
                        if (p.band) {  // `mps block`: out[row][k] = correlation of (row, row + 1 + k)
                            p.band[(size_t)i * p.band_w + (j - i - 1)] = rf;
                            continue;
                        }
                        p.C[(size_t)i * p.ldc + j] = rf;
                        p.C[(size_t)j * p.ldc + i] = rf;
                        if (p.l0_bits && !(corr_fisher_z((double)rf) < p.l0_th)) {  // level 0: the edge stays
                            atomicOr(&p.l0_bits[(size_t)i * p.l0_rw + (j >> 5)], 1u << (j & 31));
                            atomicOr(&p.l0_bits[(size_t)j * p.l0_rw + (i >> 5)], 1u << (i & 31));
                        }
                        if (p.tables) {
                            const size_t lin = (size_t)i * (2 * (size_t)p.m - i - 1) / 2 + (size_t)(j - i - 1);
                            for (int k = 0; k < 9; k++) p.tables[lin * 9 + k] = (long long)s[k];
                        }
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (PAIR) mbar_arrive_even_cta(&tempty[as]); else mbar_arrive(&tempty[as]);
            }
            if (++as == 2) {
                as = 0;
                aphase ^= 1;
            }
        }
    }
    // ---- teardown ------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // the peer may still read its accumulator / receive an arrival
    if (warp == 2) {
        if constexpr (PAIR)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ---- host side ----------------------------------------------------------------------------------
// cuTensorMapEncodeTiled is a driver entry point; resolve it through the runtime so that the
// library does not link against libcuda (the build container has no driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !ptr) {
        cudaGetLastError();
        return nullptr;
    }
    fn = (EncodeTiledFn)ptr;
    return fn;
}

// Indicator matrix: `rows` x Kp bytes, row-major.  Boxes: 128 B of K x 128 rows (A) / 256 rows (B).
int gram_make_maps(const unsigned char *ind, size_t rows, size_t Kp, CUtensorMap *map_a, CUtensorMap *map_b)
{
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        cusk_set_error("cuTensorMapEncodeTiled is not available from this driver");
        return CUSK_E_CUDA;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)Kp, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)Kp};
    const cuuint32_t estr[2] = {1, 1};
    const cuuint32_t box_a[2] = {BK, TILE_M}, box_b[2] = {BK, TILE_N};
    CUresult r1 = enc(map_a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)ind, dims, strides, box_a, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = enc(map_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)ind, dims, strides, box_b, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS) {
        cusk_set_error("cuTensorMapEncodeTiled failed (%d, %d): rows=%zu Kp=%zu", (int)r1, (int)r2, rows, Kp);
        return CUSK_E_CUDA;
    }
    return CUSK_OK;
}

int launch_gram(cudaStream_t st, const CUtensorMap &map_a, const CUtensorMap &map_b, const GramParams &p, int grid)
{
    if (grid < 1) return CUSK_OK;
    constexpr int SMEM = GramGeom<false>::SMEM_BYTES;
    cudaError_t e = cudaFuncSetAttribute(gram_i8_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
        cusk_set_error("cudaFuncSetAttribute(gram_i8_kernel, %d B): %s", SMEM, cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    gram_i8_kernel<false><<<grid, GRAM_THREADS, SMEM, st>>>(map_a, map_b, p);
    return CUSK_OK;
}

// CTA-pair form: clusters of two CTAs (one TPC), `grid` even; tile entries 2c, 2c + 1 are the A groups 2 tm2, 2 tm2 + 1 of
// one B block (see the kernel).  map_b is unused: both operands come in 128-row boxes.
int launch_gram_pair(cudaStream_t st, const CUtensorMap &map_a, const CUtensorMap &map_b, const GramParams &p, int grid)
{
    if (grid < 2) return CUSK_OK;
    if (grid & 1) {
        cusk_set_error("launch_gram_pair: odd grid %d", grid);
        return CUSK_E_ARG;
    }
    constexpr int SMEM = GramGeom<true>::SMEM_BYTES;
    cudaError_t e = cudaFuncSetAttribute(gram_i8_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
        cusk_set_error("cudaFuncSetAttribute(gram_i8_kernel<pair>, %d B): %s", SMEM, cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GRAM_THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    // the wave's producers meet at counters (K-panel meeting): every cluster of the grid must be resident at once
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, gram_i8_kernel<true>, &cfg) != cudaSuccess || clusters * 2 < grid) {
        cudaGetLastError();
        return CUSK_E_UNSUPPORTED_PAIR;
    }
    e = cudaLaunchKernelEx(&cfg, gram_i8_kernel<true>, map_a, map_b, p);
    if (e != cudaSuccess) {
        cusk_set_error("launch of gram_i8_kernel<pair>: %s", cudaGetErrorString(e));
        return CUSK_E_CUDA;
    }
    return CUSK_OK;
}
