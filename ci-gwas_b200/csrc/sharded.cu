// sharded.cu -- communicator of the edge-sharded mode (one large skeleton over several GPUs,
// SURVEY 8e): NCCL all-gather / all-reduce over NVLink between the levels of the skeleton.
// libnccl is resolved at run time (dlopen) so that libcusk_b200.so itself has no NCCL link
// dependency; inside a torch process this binds to the NCCL that torch already loaded.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace {

struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

NcclApi g_nccl;
std::mutex g_nccl_mu;

bool load_nccl()
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.h) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *nm : names) {
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        cusk_set_error("libnccl.so.2 not found: %s", dlerror());
        return false;
    }
#define CUSK_SYM(field, name)                                     \
    *(void **)(&g_nccl.field) = dlsym(h, name);                   \
    if (!g_nccl.field) {                                          \
        cusk_set_error("symbol %s missing from libnccl", name);   \
        return false;                                             \
    }
    CUSK_SYM(GetUniqueId, "ncclGetUniqueId")
    CUSK_SYM(CommInitRank, "ncclCommInitRank")
    CUSK_SYM(CommInitAll, "ncclCommInitAll")
    CUSK_SYM(CommDestroy, "ncclCommDestroy")
    CUSK_SYM(AllGather, "ncclAllGather")
    CUSK_SYM(AllReduce, "ncclAllReduce")
    CUSK_SYM(GetErrorString, "ncclGetErrorString")
    CUSK_SYM(GetVersion, "ncclGetVersion")
#undef CUSK_SYM
    g_nccl.h = h;
    return true;
}

}  // namespace

// In-process exchange for ranks that live in one process (one host thread per rank): peers publish
// their send pointers, meet at a barrier and pull each other's buffers with cudaMemcpyPeerAsync
// (NVLink P2P between devices; a plain device copy when ranks share a device, which is how the
// sharded mode is tested on a single GPU).
struct LocalGroup {
    int world = 0;
    std::mutex mu;
    std::condition_variable cv;
    int arrived = 0;
    unsigned long long generation = 0;
    std::vector<const void *> send;
    std::vector<int> device;
    int refs = 0;
    void barrier()
    {
        std::unique_lock<std::mutex> lk(mu);
        const unsigned long long gen = generation;
        if (++arrived == world) {
            arrived = 0;
            generation++;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return generation != gen; });
        }
    }
};

struct cusk_comm {
    ncclComm_t comm = nullptr;
    LocalGroup *group = nullptr;
    int device = 0;
    ShardLink link;
    u64 *scratch = nullptr;  // in-process all-reduce: the peers' buffers, grown on demand and kept
    size_t scratch_count = 0;
};

__global__ void __launch_bounds__(256) min_u64_kernel(u64 *__restrict__ dst, const u64 *__restrict__ src, size_t count)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const u64 a = dst[i], b = src[i];
        dst[i] = a < b ? a : b;
    }
}

static int local_allgather(void *user, const void *d_send, void *d_recv, size_t bytes, cudaStream_t st)
{
    cusk_comm *c = (cusk_comm *)user;
    LocalGroup *g = c->group;
    const int me = c->link.rank;
    g->send[me] = d_send;
    // this exchange is host-synchronous (unlike NCCL's, which is ordered on the stream): the level's kernels of this
    // rank must have finished before a peer copies the buffer
    cudaError_t e0 = cudaStreamSynchronize(st);
    g->barrier();  // every rank's send buffer is complete
    cudaError_t e = e0;
    for (int r = 0; r < g->world && e == cudaSuccess; r++)
        e = cudaMemcpyPeerAsync((char *)d_recv + (size_t)r * bytes, c->device, g->send[r], g->device[r], bytes, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    g->barrier();  // nobody may overwrite its send buffer before all peers have read it
    if (e != cudaSuccess) {
        cusk_set_error("local all-gather: %s", cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

static int local_allreduce_min_u64(void *user, void *d_buf, size_t count, cudaStream_t st)
{
    cusk_comm *c = (cusk_comm *)user;
    LocalGroup *g = c->group;
    const int me = c->link.rank;
    // pull every peer's buffer into scratch and fold; two barriers keep reads and the write-back apart.  A rank
    // whose scratch allocation fails still takes part in both barriers (its peers would wait for ever otherwise).
    cudaError_t e = cudaSuccess;
    if (c->scratch_count < (size_t)g->world * count) {
        if (c->scratch) cudaFree(c->scratch);
        c->scratch = nullptr;
        c->scratch_count = 0;
        e = cudaMalloc((void **)&c->scratch, (size_t)g->world * count * sizeof(u64));
        if (e == cudaSuccess) c->scratch_count = (size_t)g->world * count;
    }
    u64 *scratch = c->scratch;
    g->send[me] = d_buf;
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);  // host-synchronous exchange: this rank's level is complete
    g->barrier();
    for (int r = 0; r < g->world && e == cudaSuccess; r++)
        e = cudaMemcpyPeerAsync(scratch + (size_t)r * count, c->device, g->send[r], g->device[r], count * sizeof(u64), st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    g->barrier();
    if (e == cudaSuccess) {
        const int grid = (int)std::min<size_t>((count + 255) / 256, 148 * 8);
        for (int r = 0; r < g->world; r++)
            if (r != me && count) min_u64_kernel<<<grid, 256, 0, st>>>((u64 *)d_buf, scratch + (size_t)r * count, count);
        e = cudaStreamSynchronize(st);
    }
    if (e != cudaSuccess) {
        cusk_set_error("local all-reduce: %s", cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

static int nccl_allgather(void *user, const void *d_send, void *d_recv, size_t bytes, cudaStream_t st)
{
    cusk_comm *c = (cusk_comm *)user;
    const ncclResult_t r = g_nccl.AllGather(d_send, d_recv, bytes, ncclUint8, c->comm, st);
    if (r != ncclSuccess) {
        cusk_set_error("ncclAllGather: %s", g_nccl.GetErrorString(r));
        return (int)r;
    }
    return 0;
}

static int nccl_allreduce_min_u64(void *user, void *d_buf, size_t count, cudaStream_t st)
{
    cusk_comm *c = (cusk_comm *)user;
    const ncclResult_t r = g_nccl.AllReduce(d_buf, d_buf, count, ncclUint64, ncclMin, c->comm, st);
    if (r != ncclSuccess) {
        cusk_set_error("ncclAllReduce: %s", g_nccl.GetErrorString(r));
        return (int)r;
    }
    return 0;
}

static void group_barrier(void *user)
{
    cusk_comm *c = (cusk_comm *)user;
    if (c->group) c->group->barrier();
}

static void fill_link(cusk_comm *c, int rank, int world)
{
    c->link.rank = rank;
    c->link.world = world;
    c->link.allgather = nccl_allgather;
    c->link.allreduce_min_u64 = nccl_allreduce_min_u64;
    c->link.user = c;
}

extern "C" int cusk_comm_unique_id(char *out128)
{
    if (!out128) return CUSK_E_ARG;
    if (!load_nccl()) return CUSK_E_CUDA;
    ncclUniqueId id;
    const ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != ncclSuccess) {
        cusk_set_error("ncclGetUniqueId: %s", g_nccl.GetErrorString(r));
        return CUSK_E_CUDA;
    }
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
    return CUSK_OK;
}

extern "C" int cusk_comm_create(cusk_ctx *ctx, int rank, int world, const char *id128, cusk_comm **out)
{
    if (!ctx || !id128 || !out || world < 1 || rank < 0 || rank >= world) {
        cusk_set_error("cusk_comm_create: invalid argument");
        return CUSK_E_ARG;
    }
    *out = nullptr;
    if (!load_nccl()) return CUSK_E_CUDA;
    CUSK_CUDA(cudaSetDevice(ctx->device));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    cusk_comm *c = new cusk_comm();
    const ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        cusk_set_error("ncclCommInitRank: %s", g_nccl.GetErrorString(r));
        delete c;
        return CUSK_E_CUDA;
    }
    fill_link(c, rank, world);
    *out = c;
    return CUSK_OK;
}

// one process driving several GPUs (the mps binary): out[i] belongs to ctxs[i]
extern "C" int cusk_comm_create_all(cusk_ctx **ctxs, int world, cusk_comm **out)
{
    if (!ctxs || !out || world < 1 || world > 64) {
        cusk_set_error("cusk_comm_create_all: invalid argument");
        return CUSK_E_ARG;
    }
    if (!load_nccl()) return CUSK_E_CUDA;
    int devs[64];
    ncclComm_t comms[64];
    for (int i = 0; i < world; i++) devs[i] = ctxs[i]->device;
    const ncclResult_t r = g_nccl.CommInitAll(comms, world, devs);
    if (r != ncclSuccess) {
        cusk_set_error("ncclCommInitAll: %s", g_nccl.GetErrorString(r));
        return CUSK_E_CUDA;
    }
    LocalGroup *g = new LocalGroup();  // only its host barrier is used (the exchanges go through NCCL)
    g->world = world;
    g->refs = world;
    for (int i = 0; i < world; i++) {
        out[i] = new cusk_comm();
        out[i]->comm = comms[i];
        out[i]->group = g;
        out[i]->device = devs[i];
        fill_link(out[i], i, world);
        out[i]->link.host_barrier = group_barrier;
    }
    return CUSK_OK;
}

// ranks of one process without NCCL, all on ONE device (device copies between the ranks' buffers): how the sharded
// mode is tested on a single GPU.  Each rank must run on its own host thread.  Contexts on distinct devices are
// refused: between GPUs use cusk_comm_create_all / cusk_comm_create (NCCL); the host-synchronous copy exchange
// of this group hung on the 8-GPU NVSwitch box in round 1.
extern "C" int cusk_comm_create_local(cusk_ctx **ctxs, int world, cusk_comm **out)
{
    if (!ctxs || !out || world < 1) {
        cusk_set_error("cusk_comm_create_local: invalid argument");
        return CUSK_E_ARG;
    }
    for (int i = 1; i < world; i++)
        if (!ctxs[i] || !ctxs[0] || ctxs[i]->device != ctxs[0]->device) {
            cusk_set_error("cusk_comm_create_local: every context must be on the same device (got devices %d and %d); "
                           "use cusk_comm_create_all (NCCL) for ranks on distinct GPUs",
                           ctxs[0] ? ctxs[0]->device : -1, ctxs[i] ? ctxs[i]->device : -1);
            return CUSK_E_ARG;
        }
    LocalGroup *g = new LocalGroup();
    g->world = world;
    g->send.assign(world, nullptr);
    g->device.resize(world);
    g->refs = world;
    for (int i = 0; i < world; i++) g->device[i] = ctxs[i]->device;
    for (int i = 0; i < world; i++) {
        out[i] = new cusk_comm();
        out[i]->group = g;
        out[i]->device = g->device[i];
        out[i]->link.rank = i;
        out[i]->link.world = world;
        out[i]->link.host_barrier = group_barrier;
        out[i]->link.allgather = local_allgather;
        out[i]->link.allreduce_min_u64 = local_allreduce_min_u64;
        out[i]->link.user = out[i];
    }
    return CUSK_OK;
}

extern "C" void cusk_comm_destroy(cusk_comm *c)
{
    if (!c) return;
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->scratch) cudaFree(c->scratch);
    if (c->group) {
        bool last;
        {
            std::lock_guard<std::mutex> lk(c->group->mu);
            last = --c->group->refs == 0;
        }
        if (last) delete c->group;
    }
    delete c;
}

// attach (or detach with NULL): every later skeleton call on ctx runs edge-sharded over the comm
extern "C" int cusk_ctx_set_comm(cusk_ctx *ctx, cusk_comm *c)
{
    if (!ctx) return CUSK_E_ARG;
    ctx->shard = c ? &c->link : nullptr;
    return CUSK_OK;
}

extern "C" int cusk_comm_stats(const cusk_comm *c, double *exchange_ms, int64_t *exchange_bytes)
{
    if (!c) return CUSK_E_ARG;
    if (exchange_ms) *exchange_ms = c->link.exchange_ms;
    if (exchange_bytes) *exchange_bytes = (int64_t)c->link.exchange_bytes;
    return CUSK_OK;
}

extern "C" int cusk_nccl_version(void)
{
    if (!load_nccl()) return 0;
    int v = 0;
    g_nccl.GetVersion(&v);
    return v;
}
