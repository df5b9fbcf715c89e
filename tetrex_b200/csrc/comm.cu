// comm.cu — the three collectives of the column-sharded query path over NCCL / NVLink, issued on the
// context's own stream so that they queue behind the probe kernel that produced their input and in front of the copy
// that reads their output: no host round trip between "probe my 64-bin columns" and "does the path live anywhere?".
//
//   any_alive   in-place all-reduce(MAX) of one byte per probed item — `path_.none()` (include/otf_collector.h:383) is
//               a statement about ALL bins, a shard only knows its own columns
//   allgather   the final candidate words (bin_words / S per rank), hit counts, hit tuples
//   bcast       a k-mer batch from the rank that enumerated it
//
// NCCL is bound at run time (dlopen of libnccl.so.2): inside a torchrun worker that resolves to the NCCL torch already
// loaded (one NCCL per process), in the stand-alone `tetrex --gpus N` binary to the system library.  The library has no
// link-time dependency on NCCL, single-GPU use never touches it.
#include "common.cuh"

#include <dlfcn.h>
#include <strings.h>

using namespace trx;

namespace
{

// the handful of NCCL declarations used here (nccl.h 2.x; enum values are ABI-stable across 2.x)
typedef struct ncclComm * ncclComm_t;
struct nccl_unique_id
{
    char internal[128];
};
enum : int { kNcclSuccess = 0 };
enum : int { kNcclUint8 = 1, kNcclUint64 = 5 }; // ncclDataType_t: ncclInt8 0, ncclUint8 1, …, ncclInt64 4, ncclUint64 5
enum : int { kNcclMax = 2 };                     // ncclRedOp_t: sum 0, prod 1, max 2, min 3

struct nccl_api
{
    void * lib{nullptr};
    int (*GetUniqueId)(nccl_unique_id *){};
    int (*CommInitRank)(ncclComm_t *, int, nccl_unique_id, int){};
    int (*CommInitAll)(ncclComm_t *, int, int const *){};
    int (*CommDestroy)(ncclComm_t){};
    int (*AllReduce)(void const *, void *, size_t, int, int, ncclComm_t, cudaStream_t){};
    int (*AllGather)(void const *, void *, size_t, int, ncclComm_t, cudaStream_t){};
    int (*Broadcast)(void const *, void *, size_t, int, int, ncclComm_t, cudaStream_t){};
    char const * (*GetErrorString)(int){};
    int (*GetVersion)(int *){};
    std::string error;
};

nccl_api & api()
{
    static nccl_api a = [] {
        nccl_api x;
        // NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO) to stdout, where `tetrex query` prints its hits: unless
        // the caller chose a file, the banner goes to stderr
        setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);
        // (NCCL honours NCCL_DEBUG_FILE only above the VERSION level; WARN prints the same banner and nothing else)
        if (char const * lvl = getenv("NCCL_DEBUG"); lvl && strcasecmp(lvl, "VERSION") == 0)
            setenv("NCCL_DEBUG", "WARN", 1);
        for (char const * name : {"libnccl.so.2", "libnccl.so"})
            if ((x.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL)))
                break;
        if (!x.lib)
        {
            x.error = std::string("cannot load NCCL: ") + dlerror();
            return x;
        }
        auto sym = [&](char const * n) -> void * {
            void * p = dlsym(x.lib, n);
            if (!p && x.error.empty())
                x.error = std::string("NCCL symbol missing: ") + n;
            return p;
        };
        x.GetUniqueId = reinterpret_cast<decltype(x.GetUniqueId)>(sym("ncclGetUniqueId"));
        x.CommInitRank = reinterpret_cast<decltype(x.CommInitRank)>(sym("ncclCommInitRank"));
        x.CommInitAll = reinterpret_cast<decltype(x.CommInitAll)>(sym("ncclCommInitAll"));
        x.CommDestroy = reinterpret_cast<decltype(x.CommDestroy)>(sym("ncclCommDestroy"));
        x.AllReduce = reinterpret_cast<decltype(x.AllReduce)>(sym("ncclAllReduce"));
        x.AllGather = reinterpret_cast<decltype(x.AllGather)>(sym("ncclAllGather"));
        x.Broadcast = reinterpret_cast<decltype(x.Broadcast)>(sym("ncclBroadcast"));
        x.GetErrorString = reinterpret_cast<decltype(x.GetErrorString)>(sym("ncclGetErrorString"));
        x.GetVersion = reinterpret_cast<decltype(x.GetVersion)>(sym("ncclGetVersion"));
        return x;
    }();
    return a;
}

int need_api()
{
    if (!api().error.empty())
        return set_error(TRX_ENODEV, "%s", api().error.c_str());
    return TRX_OK;
}

#define TRX_NCCL(expr)                                                                                                 \
    do                                                                                                                 \
    {                                                                                                                  \
        int r__ = (expr);                                                                                              \
        if (r__ != kNcclSuccess)                                                                                       \
            return set_error(TRX_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr, api().GetErrorString(r__));         \
    } while (0)

} // namespace

struct trx_comm
{
    trx_ctx * ctx{nullptr};
    ncclComm_t comm{nullptr};
    int rank{0}, world{1};
    uint64_t collectives{0};
};

extern "C"
{

int trx_comm_unique_id(uint8_t id[TRX_COMM_ID_BYTES])
{
    if (!id)
        return set_error(TRX_EINVAL, "trx_comm_unique_id: NULL");
    TRX_TRY(need_api());
    nccl_unique_id u{};
    TRX_NCCL(api().GetUniqueId(&u));
    static_assert(sizeof u == TRX_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    memcpy(id, &u, sizeof u);
    return TRX_OK;
}

int trx_comm_create(trx_ctx_t * ctx, int rank, int world, const uint8_t id[TRX_COMM_ID_BYTES], trx_comm_t ** out)
{
    if (!ctx || !id || !out || world < 1 || rank < 0 || rank >= world)
        return set_error(TRX_EINVAL, "trx_comm_create: bad argument");
    *out = nullptr;
    TRX_TRY(need_api());
    TRX_TRY(ctx->use());
    nccl_unique_id u{};
    memcpy(&u, id, sizeof u);
    auto * c = new (std::nothrow) trx_comm;
    if (!c)
        return set_error(TRX_ENOMEM, "out of host memory");
    c->ctx = ctx;
    c->rank = rank;
    c->world = world;
    int const rc = api().CommInitRank(&c->comm, world, u, rank);
    if (rc != kNcclSuccess)
    {
        delete c;
        return set_error(TRX_ECUDA, "ncclCommInitRank(rank %d of %d): %s", rank, world, api().GetErrorString(rc));
    }
    *out = c;
    return TRX_OK;
}

int trx_comm_create_all(trx_ctx_t * const * ctxs, int n, trx_comm_t ** out)
{
    if (!ctxs || !out || n < 1)
        return set_error(TRX_EINVAL, "trx_comm_create_all: bad argument");
    TRX_TRY(need_api());
    std::vector<int> devs(n);
    for (int i = 0; i < n; ++i)
    {
        if (!ctxs[i])
            return set_error(TRX_EINVAL, "trx_comm_create_all: NULL context");
        devs[i] = ctxs[i]->device;
        out[i] = nullptr;
    }
    std::vector<ncclComm_t> comms(n);
    TRX_NCCL(api().CommInitAll(comms.data(), n, devs.data()));
    for (int i = 0; i < n; ++i)
    {
        auto * c = new trx_comm;
        c->ctx = ctxs[i];
        c->comm = comms[i];
        c->rank = i;
        c->world = n;
        out[i] = c;
    }
    return TRX_OK;
}

void trx_comm_destroy(trx_comm_t * c)
{
    if (!c)
        return;
    cudaSetDevice(c->ctx->device);
    cudaStreamSynchronize(c->ctx->stream);
    if (c->comm)
        api().CommDestroy(c->comm);
    delete c;
}

int trx_comm_rank(const trx_comm_t * c)
{
    return c ? c->rank : -1;
}

int trx_comm_world(const trx_comm_t * c)
{
    return c ? c->world : 0;
}

uint64_t trx_comm_collectives(const trx_comm_t * c)
{
    return c ? c->collectives : 0;
}

int trx_comm_version(void)
{
    if (need_api() != TRX_OK)
        return -1;
    int v = 0;
    return api().GetVersion(&v) == kNcclSuccess ? v : -1;
}

int trx_comm_any_alive_dev(trx_comm_t * c, uint8_t * d_alive, uint64_t n)
{
    if (!c || (n && !d_alive))
        return set_error(TRX_EINVAL, "trx_comm_any_alive_dev: NULL argument");
    if (n == 0 || c->world == 1)
        return TRX_OK;
    TRX_TRY(c->ctx->use());
    TRX_NCCL(api().AllReduce(d_alive, d_alive, n, kNcclUint8, kNcclMax, c->comm, c->ctx->stream));
    ++c->collectives;
    return TRX_OK;
}

int trx_comm_allgather_dev(trx_comm_t * c, const void * d_send, void * d_recv, uint64_t bytes_per_rank)
{
    if (!c || (bytes_per_rank && (!d_send || !d_recv)))
        return set_error(TRX_EINVAL, "trx_comm_allgather_dev: NULL argument");
    if (bytes_per_rank == 0)
        return TRX_OK;
    TRX_TRY(c->ctx->use());
    TRX_NCCL(api().AllGather(d_send, d_recv, bytes_per_rank, kNcclUint8, c->comm, c->ctx->stream));
    ++c->collectives;
    return TRX_OK;
}

int trx_comm_bcast_dev(trx_comm_t * c, void * d_buf, uint64_t bytes, int root)
{
    if (!c || (bytes && !d_buf) || root < 0 || root >= c->world)
        return set_error(TRX_EINVAL, "trx_comm_bcast_dev: bad argument");
    if (bytes == 0 || c->world == 1)
        return TRX_OK;
    TRX_TRY(c->ctx->use());
    TRX_NCCL(api().Broadcast(d_buf, d_buf, bytes, kNcclUint8, root, c->comm, c->ctx->stream));
    ++c->collectives;
    return TRX_OK;
}

// host-buffer convenience for small control messages (final candidate words): stage, all-gather, read back, sync
int trx_comm_allgather(trx_comm_t * c, const void * send, void * recv, uint64_t bytes_per_rank)
{
    if (!c || (bytes_per_rank && (!send || !recv)))
        return set_error(TRX_EINVAL, "trx_comm_allgather: NULL argument");
    if (bytes_per_rank == 0)
        return TRX_OK;
    trx_ctx * x = c->ctx;
    TRX_TRY(x->use());
    uint64_t const total = bytes_per_rank * static_cast<uint64_t>(c->world);
    TRX_TRY(x->d_in[3].reserve(bytes_per_rank));
    TRX_TRY(x->d_out[1].reserve(total));
    TRX_CUDA(cudaMemcpyAsync(x->d_in[3].ptr, send, bytes_per_rank, cudaMemcpyHostToDevice, x->stream));
    TRX_TRY(trx_comm_allgather_dev(c, x->d_in[3].ptr, x->d_out[1].ptr, bytes_per_rank));
    TRX_CUDA(cudaMemcpyAsync(recv, x->d_out[1].ptr, total, cudaMemcpyDeviceToHost, x->stream));
    TRX_CUDA(cudaStreamSynchronize(x->stream));
    return TRX_OK;
}

} // extern "C"
