// ctx.cu — context, error text, device-memory helpers and the host-side sizing helpers of the C ABI.
#include "common.cuh"

#include <cmath>

namespace trx
{

static thread_local std::string g_error;

int set_error(int code, char const * fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

int scratch::reserve(size_t want)
{
    if (want <= bytes)
        return TRX_OK;
    release();
    size_t grow = want + want / 4 + 256;
    if (pinned_host)
        TRX_CUDA(cudaMallocHost(&ptr, grow));
    else
        TRX_CUDA(cudaMalloc(&ptr, grow));
    bytes = grow;
    return TRX_OK;
}

void scratch::release()
{
    if (ptr)
    {
        if (pinned_host)
            cudaFreeHost(ptr);
        else
            cudaFree(ptr);
    }
    ptr = nullptr;
    bytes = 0;
}

} // namespace trx

int trx_ctx::use() const
{
    TRX_CUDA(cudaSetDevice(device));
    return TRX_OK;
}

extern "C"
{

const char * trx_last_error(void)
{
    return trx::g_error.c_str();
}

int trx_abi_version(void)
{
    return TRX_ABI_VERSION;
}

int trx_ctx_create(int device, void * stream, trx_ctx_t ** out)
{
    if (!out)
        return trx::set_error(TRX_EINVAL, "trx_ctx_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t err = cudaGetDeviceCount(&n);
    if (err != cudaSuccess || n == 0)
        return trx::set_error(TRX_ENODEV, "no CUDA device (%s); tetrex_b200 has no CPU fallback",
                              err == cudaSuccess ? "device count 0" : cudaGetErrorString(err));
    if (device < 0 || device >= n)
        return trx::set_error(TRX_EINVAL, "device %d out of range [0,%d)", device, n);
    cudaDeviceProp prop{};
    TRX_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return trx::set_error(TRX_ENODEV, "device %d is sm_%d%d; this library carries sm_100a code only", device,
                              prop.major, prop.minor);
    TRX_CUDA(cudaSetDevice(device));
    auto * ctx = new (std::nothrow) trx_ctx;
    if (!ctx)
        return trx::set_error(TRX_ENOMEM, "out of host memory");
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    if (stream)
    {
        ctx->stream = static_cast<cudaStream_t>(stream);
        ctx->owns_stream = false;
    }
    else
    {
        cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess)
        {
            delete ctx;
            return trx::set_error(TRX_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
        }
        ctx->owns_stream = true;
    }
    cudaEventCreate(&ctx->ev0);
    cudaEventCreate(&ctx->ev1);
    for (auto & s : ctx->h_pin)
        s.pinned_host = true;
    *out = ctx;
    return TRX_OK;
}

void trx_ctx_destroy(trx_ctx_t * ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto & s : ctx->d_in)
        s.release();
    for (auto & s : ctx->d_out)
        s.release();
    for (auto & s : ctx->h_pin)
        s.release();
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    for (cudaEvent_t e : ctx->copy_events)
        cudaEventDestroy(e);
    if (ctx->copy_stream)
    {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaStreamDestroy(ctx->copy_stream);
    }
    for (cudaEvent_t e : ctx->in_events)
        cudaEventDestroy(e);
    if (ctx->in_stream)
    {
        cudaStreamSynchronize(ctx->in_stream);
        cudaStreamDestroy(ctx->in_stream);
    }
    if (ctx->owns_stream)
        cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int trx_ctx_sync(trx_ctx_t * ctx)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaStreamSynchronize(ctx->stream));
    return TRX_OK;
}

void * trx_ctx_stream(trx_ctx_t * ctx)
{
    return ctx ? ctx->stream : nullptr;
}

uint64_t trx_ctx_launches(trx_ctx_t * ctx)
{
    return ctx ? ctx->launches : 0;
}

int trx_ctx_device(const trx_ctx_t * ctx)
{
    return ctx ? ctx->device : -1;
}

int trx_dev_alloc(trx_ctx_t * ctx, uint64_t bytes, void ** out)
{
    if (!ctx || !out)
        return trx::set_error(TRX_EINVAL, "trx_dev_alloc: NULL argument");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaMalloc(out, bytes ? bytes : 1));
    return TRX_OK;
}

int trx_dev_free(trx_ctx_t * ctx, void * ptr)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaStreamSynchronize(ctx->stream));
    TRX_CUDA(cudaFree(ptr));
    return TRX_OK;
}

int trx_dev_memset(trx_ctx_t * ctx, void * dst, int value, uint64_t bytes)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaMemsetAsync(dst, value, bytes, ctx->stream));
    return TRX_OK;
}

int trx_copy_h2d(trx_ctx_t * ctx, void * dst_dev, const void * src_host, uint64_t bytes)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    TRX_CUDA(cudaStreamSynchronize(ctx->stream));
    return TRX_OK;
}

int trx_copy_d2h(trx_ctx_t * ctx, void * dst_host, const void * src_dev, uint64_t bytes)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    TRX_CUDA(cudaStreamSynchronize(ctx->stream));
    return TRX_OK;
}

int trx_timer_start(trx_ctx_t * ctx)
{
    if (!ctx)
        return trx::set_error(TRX_EINVAL, "ctx is NULL");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
    return TRX_OK;
}

int trx_timer_stop_ms(trx_ctx_t * ctx, float * ms)
{
    if (!ctx || !ms)
        return trx::set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(ctx->use());
    TRX_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
    TRX_CUDA(cudaEventSynchronize(ctx->ev1));
    TRX_CUDA(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return TRX_OK;
}

// ---------------------------------------------------------------------------- host-side sizing helpers

uint64_t trx_emit_count(int molecule, uint32_t k, uint64_t record_len)
{
    if (k == 0 || record_len < k)
        return 0; // include/index_ibf.h:114-118: records shorter than k are skipped
    return molecule == TRX_MOL_NA ? record_len + 1 : record_len - k + 1;
}

void trx_peptide_maps(int reduction, uint8_t aamap[256], uint8_t redmap[256])
{
    // Written as "letters of one class" strings; class index = code (peptide_decomposer.h:59-151).
    static char const * const base[] = {"A", "C", "DB", "EZ", "F", "G", "H", "I", "K", "LJ", "M",
                                        "N", "P", "Q", "R", "S", "T", "V", "W", "Y", "XOU"};
    static char const * const murphy[] = {"A", "BDENQZ", "CU", "FWY", "G", "H", "IJLMV", "KOR", "P", "STX"};
    static char const * const li[] = {"ASTX", "BDEQZ", "CU", "FWY", "G", "HN", "IV", "JLM", "KOR", "P"};
    memset(aamap, 0, 256);
    memset(redmap, 0, 256);
    char const * const * classes = reduction == TRX_RED_MURPHY ? murphy : reduction == TRX_RED_LI ? li : base;
    int n = reduction == TRX_RED_BASE ? 21 : 10;
    for (int c = 0; c < n; ++c)
        for (char const * p = classes[c]; *p; ++p)
            aamap[static_cast<uint8_t>(*p)] = static_cast<uint8_t>(c);
    // redmap_: residue → representative letter of its reduced class (peptide_decomposer.h:153-213).
    // For Base the reference fills the Li table (its else-branch) and never consults it.
    static char const * const murphy_rep = "ABCFGHIKPS";
    static char const * const li_rep = "ABCFGHIJKP";
    bool use_murphy = reduction == TRX_RED_MURPHY;
    char const * const * rc = use_murphy ? murphy : li;
    char const * rep = use_murphy ? murphy_rep : li_rep;
    for (int c = 0; c < 10; ++c)
        for (char const * p = rc[c]; *p; ++p)
            redmap[static_cast<uint8_t>(*p)] = static_cast<uint8_t>(rep[c]);
}

uint64_t trx_compute_bitcount(uint64_t n, float fpr)
{
    // include/index_ibf.h:133-139: std::log(float) is the single-precision overload.
    double numerator = -static_cast<double>(n) * static_cast<double>(std::log(fpr));
    double denominator = std::pow(std::log(2.0), 2);
    return static_cast<uint64_t>(std::ceil(numerator / denominator));
}

} // extern "C"
