// noalloc.cu -- an allocation-free variant of the prep-bed statistics (cusk_bed_marker_stats, corr.cu) for callers that
// cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  No counterpart in
// the reference.
//
// The one piece of scratch the two kernels need -- the exact genotype counts int[m][4] that bed_counts_kernel writes
// and bed_stats_kernel reads -- lives in a file-scope __device__ array of fixed size instead of the context's arena:
// CUSK_NOALLOC_MAX_MARKERS markers = 1 MiB, held by every GPU that loads the library until the process exits.  Inputs and
// outputs are the caller's DEVICE buffers; the entry point calls no cudaMalloc / cudaFree / cudaMallocHost and no library
// routine that allocates.
//
// In host code the array's name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current, and that pointer is what the kernels get.
//
// All calls share the array and it keeps the last call's contents.  Nothing in it has to start at zero (the first kernel
// of a call writes every slot the second one reads); the used part is cleared anyway so that a short call after a long one
// never sees stale counts.  Two calls that overlap on the GPU (different streams, or graphs replayed concurrently) race on
// it: serialise them.
#include <mutex>

#include "common.cuh"

// launch wrappers of the two kernels (corr_kernels.cu), the same ones cusk_bed_marker_stats uses
void launch_bed_counts(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, int *counts);
void launch_bed_stats(cudaStream_t st, const unsigned char *bed, int m, int n, size_t col_bytes, const int *counts,
                      float *means, float *stds, int *modes);

#define CUSK_NOALLOC_MAX_MARKERS_IMPL 65536
static_assert(CUSK_NOALLOC_MAX_MARKERS_IMPL == CUSK_NOALLOC_MAX_MARKERS, "include/cusk_b200.h states the limit");

static __device__ __align__(256) int g_noalloc_counts[4 * CUSK_NOALLOC_MAX_MARKERS_IMPL];

namespace
{
std::mutex g_mu;
int *g_counts_of_device[64] = {};  // device address of g_noalloc_counts on each device, 0 = not asked yet

// device address of the scratch on the CURRENT device (first call per device: one cudaGetSymbolAddress, no allocation)
int scratch_address(int **out)
{
    int dev = -1;
    CUSK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        cusk_set_error("cusk_bed_marker_stats_noalloc: device %d out of range", dev);
        return CUSK_E_ARG;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_counts_of_device[dev]) {
        int *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_counts);
        if (e != cudaSuccess || !ptr) {
            cusk_set_error("cusk_bed_marker_stats_noalloc: cudaGetSymbolAddress on device %d: %s", dev,
                           cudaGetErrorString(e));
            cudaGetLastError();
            return CUSK_E_CUDA;
        }
        g_counts_of_device[dev] = ptr;
    }
    *out = g_counts_of_device[dev];
    return CUSK_OK;
}
}  // namespace

extern "C" int cusk_bed_marker_stats_noalloc(const unsigned char *d_bed, size_t m, size_t n, float *d_means, float *d_stds,
                                             int *d_modes, int *d_counts4, void *stream)
{
    if (!d_bed || !d_means || !d_stds || !d_modes) {
        cusk_set_error("cusk_bed_marker_stats_noalloc: NULL argument");
        return CUSK_E_ARG;
    }
    if (m == 0 || n == 0) {
        cusk_set_error("cusk_bed_marker_stats_noalloc: empty input (m = %zu, n = %zu)", m, n);
        return CUSK_E_ARG;
    }
    if (m > (size_t)CUSK_NOALLOC_MAX_MARKERS) {
        cusk_set_error("cusk_bed_marker_stats_noalloc: %zu markers exceed the fixed scratch of %d markers "
                       "(CUSK_NOALLOC_MAX_MARKERS); split the call or use cusk_bed_marker_stats",
                       m, (int)CUSK_NOALLOC_MAX_MARKERS);
        return CUSK_E_ARG;
    }
    if (n > (size_t)2147483647) {
        cusk_set_error("cusk_bed_marker_stats_noalloc: n = %zu does not fit the kernels' int sample index", n);
        return CUSK_E_ARG;
    }
    int *counts = nullptr;
    const int rc = scratch_address(&counts);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t col_bytes = (n + 3) / 4;
    CUSK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * 4 * m, s));
    launch_bed_counts(s, d_bed, (int)m, (int)n, col_bytes, counts);
    launch_bed_stats(s, d_bed, (int)m, (int)n, col_bytes, counts, d_means, d_stds, d_modes);
    CUSK_CUDA(cudaGetLastError());
    if (d_counts4)
        CUSK_CUDA(cudaMemcpyAsync(d_counts4, counts, sizeof(int) * 4 * m, cudaMemcpyDeviceToDevice, s));
    return CUSK_OK;  // asynchronous on `stream`, like a kernel launch: the caller synchronises
}
