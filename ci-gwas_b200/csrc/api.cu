// api.cu -- host-buffer entry points and the reference's own signatures (drop-in boundary).
#include <algorithm>
#include <mutex>

#include "engine.h"

namespace {
struct Timer {
    cudaEvent_t a, b;
    cudaStream_t st;
    Timer(cusk_ctx *ctx, int slot)
        : a(ctx->ev[3 * (CUSK_ML + 2) + slot * 2]), b(ctx->ev[3 * (CUSK_ML + 2) + slot * 2 + 1]), st(ctx->stream)
    {
    }
    void start() { cudaEventRecord(a, st); }
    void stop() { cudaEventRecord(b, st); }
    float ms()
    {
        float t = 0.f;
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&t, a, b);
        return t;
    }
};
}  // namespace

extern "C" int cusk_skeleton(cusk_ctx *ctx, const float *C, int n, int *G, const float *Th, int maxlevel,
                             int *l_out, float *pMax, int *SepSet, cusk_sepset *sep, int64_t sep_cap,
                             int64_t *sep_n, cusk_stats *stats)
{
    if (!ctx || !C || !G || !Th || n < 1) {
        cusk_set_error("cusk_skeleton: invalid argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nn = (size_t)n * n;
    const int rw = (int)cusk_row_words(n);
    arena_reset(ctx->io);
    float *dC = (float *)arena_take(ctx->io, nn * 4);
    uint32_t *dbits = (uint32_t *)arena_take(ctx->io, (size_t)n * rw * 4);
    int *dG = (int *)arena_take(ctx->io, nn * 4);
    float *dP = pMax ? (float *)arena_take(ctx->io, nn * 4) : nullptr;
    if (!dC || !dbits || !dG || (pMax && !dP)) return CUSK_E_CUDA;
    Timer th2d(ctx, 0), td2h(ctx, 1);
    th2d.start();
    CUSK_CUDA(cudaMemcpyAsync(dC, C, nn * 4, cudaMemcpyHostToDevice, st));
    th2d.stop();
    cusk_stats local;
    cusk_stats *sp = stats ? stats : &local;
    std::vector<cusk_sepset> seps;
    const bool want_seps = SepSet || sep || sep_n;
    int rc = skeleton_engine(ctx, dC, n, dbits, Th, maxlevel, l_out, dP, want_seps ? &seps : nullptr, sp);
    if (rc) return rc;
    td2h.start();
    launch_expand_bits(st, dbits, n, rw, dG);
    CUSK_CUDA(cudaMemcpyAsync(G, dG, nn * 4, cudaMemcpyDeviceToHost, st));
    if (pMax) CUSK_CUDA(cudaMemcpyAsync(pMax, dP, nn * 4, cudaMemcpyDeviceToHost, st));
    td2h.stop();
    CUSK_CUDA(cudaStreamSynchronize(st));
    if (sep_n) *sep_n = (int64_t)seps.size();
    if (sep && sep_cap > 0)
        memcpy(sep, seps.data(), std::min<size_t>(seps.size(), (size_t)sep_cap) * sizeof(cusk_sepset));
    if (SepSet) {
        // dense form of the reference: SepSet[(x*n + y)*14 + i], -1 padded (src/cuPC-S.cu:452-456)
        std::fill(SepSet, SepSet + nn * CUSK_ML, -1);
        for (const cusk_sepset &r : seps) {
            int *slot = SepSet + ((size_t)r.x * n + r.y) * CUSK_ML;
            for (int j = 0; j < r.level; j++) slot[j] = r.s[j];
        }
    }
    sp->h2d_ms = th2d.ms();
    sp->d2h_ms = td2h.ms();
    return CUSK_OK;
}

extern "C" int cusk_hetcor_skeleton(cusk_ctx *ctx, const float *C, int n, int *G, int G_in_all_ones,
                                    const float *N, float th, int maxlevel, const int *time_index, int *l_out,
                                    cusk_stats *stats)
{
    if (!ctx || !C || !G || !N || n < 1) {
        cusk_set_error("cusk_hetcor_skeleton: invalid argument");
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nn = (size_t)n * n;
    const int rw = (int)cusk_row_words(n);
    // a time index only matters when it holds more than one distinct value
    bool use_time = false;
    if (time_index)
        for (int i = 1; i < n; i++)
            if (time_index[i] != time_index[0]) use_time = true;
    arena_reset(ctx->io);
    float *dC = (float *)arena_take(ctx->io, nn * 4);
    float *dN = (float *)arena_take(ctx->io, nn * 4);
    uint32_t *dbits = (uint32_t *)arena_take(ctx->io, (size_t)n * rw * 4);
    int *dG = (int *)arena_take(ctx->io, nn * 4);
    int *dT = use_time ? (int *)arena_take(ctx->io, (size_t)n * 4) : nullptr;
    if (!dC || !dN || !dbits || !dG || (use_time && !dT)) return CUSK_E_CUDA;
    Timer th2d(ctx, 0), td2h(ctx, 1);
    th2d.start();
    CUSK_CUDA(cudaMemcpyAsync(dC, C, nn * 4, cudaMemcpyHostToDevice, st));
    CUSK_CUDA(cudaMemcpyAsync(dN, N, nn * 4, cudaMemcpyHostToDevice, st));
    if (!G_in_all_ones) CUSK_CUDA(cudaMemcpyAsync(dG, G, nn * 4, cudaMemcpyHostToDevice, st));
    if (use_time) CUSK_CUDA(cudaMemcpyAsync(dT, time_index, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    th2d.stop();
    cusk_stats local;
    cusk_stats *sp = stats ? stats : &local;
    int rc = hetcor_engine(ctx, dC, n, dbits, G_in_all_ones ? nullptr : dG, false, dN, 0.f, th, maxlevel, dT, l_out,
                           sp);
    if (rc) return rc;
    td2h.start();
    launch_expand_bits(st, dbits, n, rw, dG);
    CUSK_CUDA(cudaMemcpyAsync(G, dG, nn * 4, cudaMemcpyDeviceToHost, st));
    td2h.stop();
    CUSK_CUDA(cudaStreamSynchronize(st));
    sp->h2d_ms = th2d.ms();
    sp->d2h_ms = td2h.ms();
    return CUSK_OK;
}

// ---- the reference's own signatures ------------------------------------------------------------
// One lazily created context on device 0, as the reference uses device 0's default stream.
static cusk_ctx *g_default_ctx = nullptr;
static std::mutex g_default_mu;

cusk_ctx *cusk_default_ctx()
{
    std::lock_guard<std::mutex> lk(g_default_mu);
    if (!g_default_ctx) {
        if (cusk_ctx_create(0, &g_default_ctx) != CUSK_OK) {
            fprintf(stderr, "cusk_b200: %s\n", cusk_last_error());
            exit(EXIT_FAILURE);  // include/mps/gpuerrors.h:6-35: errors end the process
        }
    }
    return g_default_ctx;
}

// include/mps/cuPC-S.h:192-194, src/cuPC-S.cu:61
extern "C" void Skeleton(float *C, int *P, int *G, float *Th, int *l, const int *maxlevel, float *pMax,
                         int *SepSet)
{
    cusk_ctx *ctx = cusk_default_ctx();
    int last = 0;
    const int rc = cusk_skeleton(ctx, C, *P, G, Th, *maxlevel, &last, pMax, SepSet, nullptr, 0, nullptr, nullptr);
    if (rc != CUSK_OK) {
        fprintf(stderr, "cusk_b200 Skeleton: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);
    }
    *l = last;
}

// include/mps/hetcor-cuPC-S.h:46, src/hetcor-cuPC-S.cu:75
extern "C" void hetcor_skeleton(float *C, int *P, int *G, float *N, float *Th, int *l, const int *maxlevel,
                                const int *time_index)
{
    cusk_ctx *ctx = cusk_default_ctx();
    int last = 0;
    const int rc = cusk_hetcor_skeleton(ctx, C, *P, G, 0, N, *Th, *maxlevel, time_index, &last, nullptr);
    if (rc != CUSK_OK) {
        fprintf(stderr, "cusk_b200 hetcor_skeleton: %s\n", cusk_last_error());
        exit(EXIT_FAILURE);
    }
    *l = last;
}
