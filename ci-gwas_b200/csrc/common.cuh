// common.cuh -- shared host/device helpers of libcusk_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "../../include/cusk_b200.h"

typedef unsigned long long u64;

#define CUSK_SAT ((u64)1 << 62)  // saturation value of 64-bit binomials
#define CUSK_KEY_NONE (~(u64)0)

// ---- error plumbing ------------------------------------------------------------------------
void cusk_set_error(const char *fmt, ...);

#define CUSK_CUDA(call)                                                                         \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            cusk_set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return CUSK_E_CUDA;                                                                 \
        }                                                                                       \
    } while (0)

// ---- context -------------------------------------------------------------------------------
// A grow-only device arena: the skeleton engines and the correlation front end carve their
// buffers from it so that repeated calls (one per LD block) never cudaMalloc in steady state.
struct ArenaBlock {
    char *base = nullptr;
    size_t cap = 0;
    size_t off = 0;
};
struct Arena {
    std::vector<ArenaBlock> blocks;
};

// Edge-sharded execution of ONE skeleton over several GPUs (cuskss-merged): set on the
// context for the duration of a call.  The level-l work items are dealt round-robin to the ranks;
// after the CI kernel of a level the ranks exchange what they removed.
struct ShardLink {
    int rank = 0, world = 1;
    // gather `bytes` from every rank's d_send into d_recv[world][bytes] (ncclAllGather semantics)
    int (*allgather)(void *user, const void *d_send, void *d_recv, size_t bytes, cudaStream_t st) = nullptr;
    // element-wise minimum over ranks, in place (ncclAllReduce, ncclUint64, ncclMin)
    int (*allreduce_min_u64)(void *user, void *d_buf, size_t count, cudaStream_t st) = nullptr;
    // ranks that are threads of ONE process: host barrier taken once per call, after the call's last device
    // allocation and before its first exchange (cudaMalloc synchronises peer devices, so it must not run
    // while another rank's collective kernel is already waiting for this rank); nullptr for one process per GPU
    void (*host_barrier)(void *user) = nullptr;
    void *user = nullptr;
    double exchange_ms = 0.0;   // device time between the end of a level's CI kernels and the end of its exchange,
                                // summed over the levels of the last call (includes waiting for the slowest rank)
    size_t exchange_bytes = 0;  // bytes received by this rank in the exchanges of the last call
};

struct cusk_ctx {
    int device = 0;
    ShardLink *shard = nullptr;
    int sm_count = 148;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    Arena arena;  // engine scratch, rewound by every engine call
    Arena io;     // device copies of host-buffer arguments, rewound by every host-buffer entry point
    // pinned staging for small control blocks / statistics
    void *h_pinned = nullptr;
    size_t h_pinned_bytes = 0;
    // pinned, grow-only: the adjacency bitmap on its way back to a host-buffer caller
    void *h_stage = nullptr;
    size_t h_stage_bytes = 0;
    // [0, ML+2): level boundaries; [ML+2, 3(ML+2)): CI kernel start/stop per level; next 4: copies;
    // [3(ML+2)+4, 5(ML+2)+4): exchange start/stop per level (edge-sharded mode)
    cudaEvent_t ev[5 * (CUSK_ML + 2) + 4];
    cudaEvent_t ev_inputs = nullptr;  // the host-buffer arguments of the current call have reached the device
    // side stream of the engine: at level 2 the thread-per-set kernel (short rows) and ci_mid_kernel (medium rows) are
    // independent and each fills about a third of the warp slots, so they run side by side (fork / join by events)
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool ev_ready = false;
    // host work of the calling entry point that does not depend on the device's result (zero-filling the caller's output
    // matrix): the engine runs it once, just before one of its stream synchronisations, while the device is busy
    std::function<void()> idle_job;
    bool level_timing = false;  // record the per-level / per-kernel-group events behind cusk_stats.level_ms and ci_kernel_ms
    // helper thread of the host-buffer entry points (api.cu): host scans and fills that run BESIDE the engine; created on
    // first use, joined when the context goes
    std::shared_ptr<void> host_helper;
};

inline void run_idle_job(cusk_ctx *ctx)
{
    if (ctx->idle_job) {
        std::function<void()> f = std::move(ctx->idle_job);
        ctx->idle_job = nullptr;
        f();
    }
}

void arena_reset(Arena &a);                // rewind: all earlier takes become invalid
void *arena_take(Arena &a, size_t bytes);  // 256-byte aligned; grows by whole blocks; nullptr + error on failure
void arena_free(Arena &a);
inline void arena_reset(cusk_ctx *ctx) { arena_reset(ctx->arena); }
inline void *arena_take(cusk_ctx *ctx, size_t bytes) { return arena_take(ctx->arena, bytes); }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
void *host_stage(cusk_ctx *ctx, size_t bytes);  // ctx->h_stage grown to at least `bytes`; nullptr + error on failure

// ---- combinatorics (64-bit, saturating; the reference's BINOM / IthCombination are int32,
//      src/cuPC-S.cu:6453-6506) --------------------------------------------------------------
__host__ __device__ inline u64 binom64(int n, int k)
{
    if (k < 0 || k > n) return 0;
    if (k > n - k) k = n - k;
    u64 c = 1;
    for (int i = 1; i <= k; i++) {
        // c * (n - k + i) / i is always integral
        if (c < ((u64)1 << 32) && n < (1 << 30)) {  // product fits in 64 bits: plain division
            c = c * (u64)(n - k + i) / (u64)i;
            continue;
        }
        // widen the product
        const unsigned __int128 t = (unsigned __int128)c * (unsigned)(n - k + i);
        const unsigned __int128 q = t / (unsigned)i;
        if (q >= (unsigned __int128)CUSK_SAT) return CUSK_SAT;
        c = (u64)q;
    }
    return c;
}

// Same mapping as unrank_comb below, for C(N,k) < 2^48 and N < 2^15 (what the CI kernel
// guarantees): the binomial is updated incrementally with exact 64-bit mul/div instead of being
// recomputed with 128-bit arithmetic for every candidate position.
__host__ __device__ inline void unrank_comb_fast(u64 rank, int N, int k, int *pos)
{
    if (k <= 0) return;
    int m = N - 1;      // current "top" of C(m, r): sets whose slot i holds x use C(N-1-x, k-1-i)
    int r = k - 1;
    u64 c = binom64(m, r);
    int x = 0;
    for (int i = 0; i < k; i++) {
        if (r == 0) {  // last slot: C(m, 0) = 1 per candidate, so the remaining rank is the offset
            pos[i] = x + (int)rank;
            break;
        }
        while (rank >= c) {  // skip x: C(m-1, r) = C(m, r) (m - r) / m
            rank -= c;
            c = (m > 0) ? c * (u64)(m - r) / (u64)m : 0;
            m--;
            x++;
        }
        pos[i] = x;
        // take x: next slot counts C(m-1, r-1) = C(m, r) r / m
        c = (m > 0) ? c * (u64)r / (u64)m : 0;
        m--;
        r--;
        x++;
        if (r < 0) break;
    }
}

// 0-based lexicographic rank -> strictly increasing positions pos[0..k) out of {0..N-1}.
__host__ __device__ inline void unrank_comb(u64 rank, int N, int k, int *pos)
{
    int x = 0;
    for (int i = 0; i < k; i++) {
        for (;; x++) {
            const u64 c = binom64(N - 1 - x, k - 1 - i);  // sets that start with x at slot i
            if (rank < c) break;
            rank -= c;
        }
        pos[i] = x;
        x++;
    }
}
