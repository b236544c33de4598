// ctx.cu -- context, error state, thresholds.
#include <algorithm>
#include <cmath>

#include "common.cuh"

static thread_local char g_err[1024] = "";

void cusk_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *cusk_last_error(void) { return g_err; }
extern "C" const char *cusk_version(void) { return "cusk-b200 0.1 (sm_100a)"; }

extern "C" int cusk_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" void cusk_ctx_set_level_timing(cusk_ctx *ctx, int on)
{
    if (ctx) ctx->level_timing = on != 0;
}

extern "C" int cusk_ctx_create(int device_id, cusk_ctx **out)
{
    if (!out) {
        cusk_set_error("cusk_ctx_create: out is NULL");
        return CUSK_E_ARG;
    }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cusk_set_error("no CUDA device available (%s); libcusk_b200 has no CPU path",
                       e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        cudaGetLastError();
        return CUSK_E_NODEVICE;
    }
    if (device_id < 0 || device_id >= ndev) {
        cusk_set_error("device %d out of range (0..%d)", device_id, ndev - 1);
        return CUSK_E_ARG;
    }
    CUSK_CUDA(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    CUSK_CUDA(cudaGetDeviceProperties(&prop, device_id));
    if (prop.major != 10) {
        cusk_set_error("device %d is sm_%d%d; libcusk_b200 is built for sm_100a only", device_id,
                       prop.major, prop.minor);
        return CUSK_E_NODEVICE;
    }
    cusk_ctx *ctx = new cusk_ctx();
    ctx->device = device_id;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    CUSK_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    ctx->h_pinned_bytes = 1 << 16;
    CUSK_CUDA(cudaMallocHost(&ctx->h_pinned, ctx->h_pinned_bytes));
    for (auto &ev : ctx->ev) CUSK_CUDA(cudaEventCreate(&ev));
    CUSK_CUDA(cudaEventCreateWithFlags(&ctx->ev_inputs, cudaEventDisableTiming));
    CUSK_CUDA(cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
    CUSK_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    CUSK_CUDA(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    ctx->ev_ready = true;
    {
        const char *e = getenv("CUSK_LEVEL_TIMING");
        ctx->level_timing = e && atoi(e) != 0;  // instrumentation: off unless asked for
    }
    *out = ctx;
    return CUSK_OK;
}

extern "C" void cusk_ctx_destroy(cusk_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    arena_free(ctx->arena);
    arena_free(ctx->io);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->ev_ready)
    {
        for (auto &ev : ctx->ev) cudaEventDestroy(ev);
        if (ctx->ev_inputs) cudaEventDestroy(ctx->ev_inputs);
        if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
        if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    }
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" void *cusk_ctx_stream(cusk_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

extern "C" int cusk_ctx_sync(cusk_ctx *ctx)
{
    if (!ctx) return CUSK_E_ARG;
    CUSK_CUDA(cudaSetDevice(ctx->device));
    CUSK_CUDA(cudaStreamSynchronize(ctx->stream));
    return CUSK_OK;
}

void *host_stage(cusk_ctx *ctx, size_t bytes)
{
    if (bytes <= ctx->h_stage_bytes && ctx->h_stage) return ctx->h_stage;
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    ctx->h_stage = nullptr;
    ctx->h_stage_bytes = 0;
    const size_t cap = align_up(bytes, (size_t)1 << 20);
    const cudaError_t e = cudaMallocHost(&ctx->h_stage, cap);
    if (e != cudaSuccess) {
        cusk_set_error("cudaMallocHost(%zu bytes): %s", cap, cudaGetErrorString(e));
        ctx->h_stage = nullptr;
        return nullptr;
    }
    ctx->h_stage_bytes = cap;
    return ctx->h_stage;
}

void arena_reset(Arena &a)
{
    for (auto &b : a.blocks) b.off = 0;
}

void arena_free(Arena &a)
{
    for (auto &b : a.blocks) cudaFree(b.base);
    a.blocks.clear();
}

void *arena_take(Arena &a, size_t bytes)
{
    if (bytes == 0) bytes = 1;
    for (auto &b : a.blocks) {
        const size_t off = align_up(b.off, 256);
        if (off + bytes <= b.cap) {
            b.off = off + bytes;
            return b.base + off;
        }
    }
    ArenaBlock nb;
    nb.cap = align_up(std::max<size_t>(bytes, (size_t)32 << 20), (size_t)2 << 20);
    cudaError_t e = cudaMalloc((void **)&nb.base, nb.cap);
    if (e != cudaSuccess) {
        cusk_set_error("cudaMalloc(%zu bytes): %s", nb.cap, cudaGetErrorString(e));
        return nullptr;
    }
    nb.off = bytes;
    a.blocks.push_back(nb);
    return nb.base;
}

// ---- thresholds ------------------------------------------------------------------------------
// Standard normal quantile, Wichura AS 241 (PPND16).  Replaces the Boost call of
// src/cuPC_call_prep.cpp:7-11.
static double qnorm(double p)
{
    if (!(p > 0.0 && p < 1.0)) {
        if (p == 0.0) return -INFINITY;
        if (p == 1.0) return INFINITY;
        return NAN;
    }
    static const double a[8] = {3.3871328727963666080e0,  1.3314166789178437745e2,
                                1.9715909503065514427e3,  1.3731693765509461125e4,
                                4.5921953931549871457e4,  6.7265770927008700853e4,
                                3.3430575583588128105e4,  2.5090809287301226727e3};
    static const double b[8] = {1.0,
                                4.2313330701600911252e1,
                                6.8718700749205790830e2,
                                5.3941960214247511077e3,
                                2.1213794301586595867e4,
                                3.9307895800092710610e4,
                                2.8729085735721942674e4,
                                5.2264952788528545610e3};
    static const double c[8] = {1.42343711074968357734e0, 4.63033784615654529590e0,
                                5.76949722146069140550e0, 3.64784832476320460504e0,
                                1.27045825245236838258e0, 2.41780725177450611770e-1,
                                2.27238449892691845833e-2, 7.74545014278341407640e-4};
    static const double d[8] = {1.0,
                                2.05319162663775882187e0,
                                1.67638483018380384940e0,
                                6.89767334985100004550e-1,
                                1.48103976427480074590e-1,
                                1.51986665636164571966e-2,
                                5.47593808499534494600e-4,
                                1.05075007164441684324e-9};
    static const double e[8] = {6.65790464350110377720e0, 5.46378491116411436990e0,
                                1.78482653991729133580e0, 2.96560571828504891230e-1,
                                2.65321895265761230930e-2, 1.24266094738807843860e-3,
                                2.71155556874348757815e-5, 2.01033439929228813265e-7};
    static const double f[8] = {1.0,
                                5.99832206555887937690e-1,
                                1.36929880922735805310e-1,
                                1.48753612908506148525e-2,
                                7.86869131145613259100e-4,
                                1.84631831751005468180e-5,
                                1.42151175831644588870e-7,
                                2.04426310338993978564e-15};
    auto horner = [](const double *k, double r) {
        double v = k[7];
        for (int i = 6; i >= 0; i--) v = v * r + k[i];
        return v;
    };
    const double q = p - 0.5;
    if (fabs(q) <= 0.425) {
        const double r = 0.180625 - q * q;
        return q * horner(a, r) / horner(b, r);
    }
    double r = (q < 0) ? p : 1.0 - p;
    r = sqrt(-log(r));
    double val;
    if (r <= 5.0) {
        r -= 1.6;
        val = horner(c, r) / horner(d, r);
    } else {
        r -= 5.0;
        val = horner(e, r) / horner(f, r);
    }
    return q < 0 ? -val : val;
}

// src/cuPC_call_prep.cpp:13-23
// scipy.stats.norm.ppf in double precision (cusk_postprocessing/sepselect.py:26-27)
extern "C" double cusk_qnorm(double p) { return qnorm(p); }

extern "C" void cusk_threshold_array(int n, float alpha, float *out15)
{
    const float p = 0.5f * alpha;
    const float q = (float)qnorm((double)p);
    for (int i = 0; i <= CUSK_ML; i++) {
        const size_t dd = (size_t)n - (size_t)i - (size_t)3;  // size_t arithmetic as the reference
        out15[i] = (float)((double)fabsf(q) / sqrt((double)dd));
    }
}

// src/cuPC_call_prep.cpp:25-28
extern "C" float cusk_hetcor_threshold(float alpha)
{
    const float p = 0.5f * alpha;
    return fabsf((float)qnorm((double)p));
}
