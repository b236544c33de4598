// skeleton.cuh -- device-side control block and kernel parameter structs of the skeleton engines.
#pragma once
#include "common.cuh"

#define CUSK_MAX_DEVICES 64     // per-device launch state of the CI kernels
#define CI_NWARPS 16
#define CI_NTHREADS (CI_NWARPS * 32)
#define CI_ZB 4                 // leaf (z) register blocking
#define CI_PIVOT_TOL 1e-10      // |pivot| below this: the direction is dropped (pseudo-inverse)
#define CI_SMALL_DMAX 32        // ci_small.cu: thread-per-set kernel for the rows whose degree is at most this
#define CI_SMALL_LMAX 14        // ... at levels up to this one
#define CI_MID_DMAX 160         // ci_mid.cu: level 2, rows with CI_SMALL_DMAX < degree <= this (CTA per item)
#define CI_MID_DMAX1 1024       // ... level 1 (nothing staged, so only the row's index arrays bound it)
#define CI_MID_LMAX 2
#define CI_DFS_DMAX 63          // ci_dfs.cu: levels above CI_MID_LMAX, rows up to this degree whose sets serve enough tests
#define CI_DFS_LMIN (CI_MID_LMAX + 1)
#define CI_SMALL_IPSM_MAX 64    // upper bound of its items per SM (sizes the decoded item table)
#define CI_MAX_COMB ((u64)1 << 48)  // per-row conditioning-set count above which a level is infeasible

// One per level, written by plan_kernel, read by every other kernel of the level and copied
// back to the host at the end of the call.
struct LevelCtl {
    int done;        // 1: this level (and all later ones) does not run
    int maxdeg;      // max degree at level start
    int nitems;      // work items of the CI kernel
    int infeasible;  // 1: some row has >= CI_MAX_COMB sets; 2: a row does not fit in shared memory
    u64 canon;       // canonical test count (per-row C(d,l) capped at 2^40 as the oracle does)
    u64 executed;    // tests evaluated
    u64 nnz;         // directed edges at level start
    u64 singular;    // dropped pivots
    double T2;       // tanh(threshold)^2 of the level (non-het and uniform-N het)
    int nitems_small;  // work items of the thread-per-set kernel (rows with degree <= CI_SMALL_DMAX)
    int nitems_mid;    // work items of ci_mid_kernel (levels 1-2, medium rows) / ci_dfs_kernel (levels >= 3)
    // accumulators of degree_plan_kernel's row phase (zero at the start of a call)
    u64 acc_tests, acc_small, acc_canon;
    int acc_maxdeg, acc_infeasible;
    unsigned ticket;
    int shard_on;  // edge-sharded mode: 1 = this level's work items are dealt over the ranks (enough tests to pay for the exchange)
    u64 acc_mid;
    int plan_ready;  // fused plan + compaction: set by the planning CTA once the level's plan is in memory
    int pad3;
    u64 t_begin, t_planned, t_end;  // levels inside a multi-level launch (ci_small.cu): %globaltimer at start / after the plan / at the end
    u64 t_rows, t_body;             // ... after the row phase's barrier / when CTA 0 has run out of items (CUSK_SELF_TRACE=1 prints them)
};

// One work item of the CI kernel, decoded once per level by items_kernel (the CI kernel's CTAs then
// fetch an item with one atomic and one 32-byte load instead of a binary search over the rows).
struct __align__(16) CiItem {
    int X, d, row0, pad;
    u64 R0, R1;
};

struct CiParams {
    const float *C;
    const float *N;          // het, general N; nullptr when uniform
    const int *time_index;   // het; nullptr when all equal
    int n, l, rw;
    uint32_t *bits;
    const int *deg;
    const int *rowptr;
    const int *adj;
    LevelCtl *ctl;
    int *work_counter;
    int *work_counter_small;  // queue of ci_small_kernel
    int *work_counter_mid;    // queue of ci_mid_kernel (levels 1-2) / of ci_dfs_kernel<NS = 1> (levels >= 3)
    int *work_counter_dfs2;   // queue of ci_dfs_kernel<NS = 2>: the same item table, each kernel skips the other's rows
    u64 *minkey;             // non-het
    double th_het;
    int smem_bytes;
    const CiItem *items;        // decoded items of ci_level_kernel [0, ctl->nitems)
    const CiItem *items_small;  // decoded items of ci_small_kernel [0, ctl->nitems_small)
    const CiItem *items_mid;    // decoded items of ci_mid_kernel [0, ctl->nitems_mid)
    int item_cap;               // capacity of either table (plan_kernel keeps the counts below it)
    unsigned char *removed;       // edge-sharded het engine: removed[slot of (X,Y) in the level's CSR] = 1
    int shard_rank, shard_world;  // this launch takes items rank, rank + world, ... (1 GPU: 0, 1)
    int sm_count;                 // persistent grids: CTAs per SM (occupancy) x this
    // self-planned levels (ci_small.cu, multi-level kernel): the level's degrees, fixed-stride adjacency and work items are
    // produced inside the launch; ctl and work_counter_small are the BASES of the per-level arrays there
    int *self_adj;         // [n][CI_SELF_STRIDE] neighbours at level start, rows of at most CI_SELF_STRIDE neighbours
    CiItem *self_items;    // [item_cap]
    int *deg_out;          // [n]
    float n_uniform;       // het engine, uniform N (threshold of the level)
    int warp_sets;         // 1: levels short of conditioning sets run a warp per set (ci_small.cu, cs_warp_body)
};

#define CI_SELF_STRIDE 32    // = CI_SMALL_DMAX: every row of a self-planned level fits the thread-per-set kernel
#define CI_SELF_DMAX 22      // the host takes the self-planned form when no row has more neighbours than this ...
#define CI_SELF_NMAX 8192    // ... and the skeleton has at most this many variables

struct DfsTuning {
    int span, dmin, ipsm, lnarrow;
};
DfsTuning dfs_tuning();  // skeleton_kernels.cu

// Which CI kernel takes a row of degree d at level l (degree_plan_kernel and its host-side mirror in skeleton.cu):
// 0 ci_level_kernel, 1 ci_small_kernel, 2 ci_mid_kernel (l <= 2) / ci_dfs_kernel (l >= 3).
//   dsmall: CI_SMALL_DMAX or 0 (kernel off); dmid: level-1/2 bound of ci_mid or 0; dfs_span > 0: ci_dfs is on.
//   At l >= CI_DFS_LMIN it takes every row with more than CI_SMALL_DMAX neighbours (up to CI_DFS_DMAX: 4.5x the rate
//   of ci_level there) and, from level dfs_lnarrow on, the shorter rows with d >= dfs_dmin and d - l >= dfs_span
//   (every set then serves at least dfs_span tests; measured on trait rows with ~29 neighbours the two kernels tie up
//   to level 12 and ci_dfs is 1.5-2.2x faster at 13-14, profiles/r2_heavy_*.json; below level 8 the thread-per-set
//   kernel wins on the small items of a light block)
__host__ __device__ inline int ci_row_class(int d, int l, int dsmall, int dmid, int dfs_span, int dfs_dmin, int dfs_lnarrow)
{
    if (l >= CI_DFS_LMIN && dfs_span > 0 && d <= CI_DFS_DMAX &&
        (d > dsmall || (l >= dfs_lnarrow && d >= dfs_dmin && d - l >= dfs_span)))
        return 2;
    if (d <= dsmall) return 1;
    if (l <= CI_MID_LMAX && d <= dmid) return 2;
    return 0;
}

// ---- programmatic dependent launch (sm_90+) ---------------------------------------------------------------------------
// The level loop is a chain of short dependent kernels.  Unless CUSK_PDL=0 they are launched with the programmatic
// stream-serialisation attribute: a kernel lets its successor's CTAs become resident at once (launch_dependents) and
// itself waits for its predecessor's completion and memory flush (wait) before it touches anything, so the successor's
// launch latency hides behind the predecessor instead of following it.  Without the attribute both are no-ops.
bool pdl_enabled();  // skeleton_kernels.cu
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_enter()
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename K, typename... A>
inline cudaError_t launch_k(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, args...);
}
#endif
