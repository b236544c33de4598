// sepselect.cu -- batched separating-set selection: one thread per outer pair runs the greedy
// search of cusk_postprocessing/sepselect.py:262-304 (sepselect_core.h) on the device.
//
// Work: an outer pair costs at most P rounds x P candidates x one forward substitution of length <= P, all on
// (P+1) x n_var doubles of correlations that every pair shares (P traits <= 32; 17 in the reference's data), so
// the block is L2-resident and the kernel is FP64-latency bound; a merged genome has 1e4..1e6 outer pairs, far
// more than the 148 x 2048 resident threads.  The reference spends its time in one numpy.linalg.inv call per
// candidate set.
#include "common.cuh"
#include "sepselect_core.h"

namespace {

__global__ void __launch_bounds__(128) sepselect_pairs_kernel(const double *__restrict__ Ct, int64_t ld, int P,
                                                              const int2 *__restrict__ pairs, int64_t npairs,
                                                              const uint32_t *__restrict__ tmask,
                                                              const double *__restrict__ thr,
                                                              cusk_sepselect_result *__restrict__ out)
{
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npairs; p += (int64_t)gridDim.x * blockDim.x) {
        const int2 ij = pairs[p];
        cusk_sepselect_result r;
        cusk_sepselect_pair(Ct, ld, P, ij.x, ij.y, tmask[ij.x], thr, &r);
        out[p] = r;
    }
}

}  // namespace

// Host buffers in, host buffers out (include/cusk_b200.h).  Ct: [(P+1) x ld] doubles, rows 0..P of the correlation
// matrix; pairs: [npairs][2] variable indices (i, j); tmask[n_var]: bit t set when trait t is a neighbour of the
// variable; thr[P+1]; out[npairs].
extern "C" int cusk_sepselect_pairs(cusk_ctx *ctx, const double *Ct, int64_t ld, int P, int n_var, const int *pairs,
                                    int64_t npairs, const uint32_t *tmask, const double *thr,
                                    cusk_sepselect_result *out)
{
    if (!ctx) {
        cusk_set_error("cusk_sepselect_pairs: no context (there is no CPU path)");
        return CUSK_E_NODEVICE;
    }
    if (!Ct || !pairs || !tmask || !thr || !out || P < 0 || P >= CUSK_SS_MAXP || n_var <= P || ld < n_var || npairs < 0) {
        cusk_set_error("cusk_sepselect_pairs: invalid argument (P = %d must be below %d, n_var = %d above P)", P,
                       CUSK_SS_MAXP, n_var);
        return CUSK_E_ARG;
    }
    for (int64_t p = 0; p < npairs; p++) {
        const int i = pairs[2 * p], j = pairs[2 * p + 1];
        if (i < 0 || j < 0 || i >= n_var || j >= n_var || (i > P && j > P)) {
            cusk_set_error("cusk_sepselect_pairs: pair %lld = (%d, %d) has no end among variables 0..%d", (long long)p, i, j, P);
            return CUSK_E_ARG;
        }
        // an outer pair of an unshielded triple: i is not its own neighbour and j is not adjacent to i.  With an end of
        // the pair among the candidates the reference inverts a matrix with a repeated index (LinAlgError, or a
        // result decided by rounding noise)
        const uint32_t mi = tmask[i];
        if ((i < P && (mi >> i & 1u)) || (j < P && (mi >> j & 1u))) {
            cusk_set_error("cusk_sepselect_pairs: pair %lld = (%d, %d): an end of the pair is among the trait neighbours of %d",
                           (long long)p, i, j, i);
            return CUSK_E_ARG;
        }
    }
    if (npairs == 0) return CUSK_OK;
    CUSK_CUDA(cudaSetDevice(ctx->device));
    arena_reset(ctx->io);
    const size_t ct_bytes = (size_t)(P + 1) * (size_t)ld * sizeof(double);
    double *d_ct = (double *)arena_take(ctx->io, ct_bytes);
    int2 *d_pairs = (int2 *)arena_take(ctx->io, (size_t)npairs * sizeof(int2));
    uint32_t *d_mask = (uint32_t *)arena_take(ctx->io, (size_t)n_var * sizeof(uint32_t));
    double *d_thr = (double *)arena_take(ctx->io, (size_t)(P + 1) * sizeof(double));
    cusk_sepselect_result *d_out = (cusk_sepselect_result *)arena_take(ctx->io, (size_t)npairs * sizeof(cusk_sepselect_result));
    if (!d_ct || !d_pairs || !d_mask || !d_thr || !d_out) return CUSK_E_CUDA;
    cudaStream_t st = ctx->stream;
    CUSK_CUDA(cudaMemcpyAsync(d_ct, Ct, ct_bytes, cudaMemcpyHostToDevice, st));
    CUSK_CUDA(cudaMemcpyAsync(d_pairs, pairs, (size_t)npairs * sizeof(int2), cudaMemcpyHostToDevice, st));
    CUSK_CUDA(cudaMemcpyAsync(d_mask, tmask, (size_t)n_var * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CUSK_CUDA(cudaMemcpyAsync(d_thr, thr, (size_t)(P + 1) * sizeof(double), cudaMemcpyHostToDevice, st));
    const int threads = 128;
    const int64_t want = (npairs + threads - 1) / threads;
    const int64_t cap = (int64_t)ctx->sm_count * 16;  // a few waves of 128-thread CTAs, grid-stride beyond that
    const int blocks = (int)(want < cap ? want : cap);
    sepselect_pairs_kernel<<<blocks, threads, 0, st>>>(d_ct, ld, P, d_pairs, npairs, d_mask, d_thr, d_out);
    CUSK_CUDA(cudaGetLastError());
    CUSK_CUDA(cudaMemcpyAsync(out, d_out, (size_t)npairs * sizeof(cusk_sepselect_result), cudaMemcpyDeviceToHost, st));
    CUSK_CUDA(cudaStreamSynchronize(st));
    return CUSK_OK;
}
