// common.cuh — shared by every translation unit of libtetrex_cuda.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/tetrex_b200.h"

namespace trx
{

int set_error(int code, char const * fmt, ...);

#define TRX_CUDA(expr)                                                                                                 \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t err__ = (expr);                                                                                    \
        if (err__ != cudaSuccess)                                                                                      \
            return ::trx::set_error(TRX_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(err__));  \
    } while (0)

#define TRX_TRY(expr)                                                                                                  \
    do                                                                                                                 \
    {                                                                                                                  \
        int rc__ = (expr);                                                                                             \
        if (rc__ != TRX_OK)                                                                                            \
            return rc__;                                                                                               \
    } while (0)

constexpr int kSMs = 148; // B200

// Grow-only scratch buffer (device or pinned host) owned by a context.
struct scratch
{
    void * ptr{nullptr};
    size_t bytes{0};
    bool pinned_host{false};
    int reserve(size_t want);
    void release();
};

} // namespace trx

struct trx_ctx
{
    int device{0};
    cudaStream_t stream{nullptr};
    bool owns_stream{false};
    uint64_t launches{0};
    int sm_count{trx::kSMs};
    cudaEvent_t ev0{nullptr}, ev1{nullptr};
    // second stream + events for the host-buffer entry points that overlap result copies with compute
    cudaStream_t copy_stream{nullptr};
    std::vector<cudaEvent_t> copy_events;
    // third stream: the k-mers of piece n + 1 arrive while piece n is computed and piece n - 1 leaves (PCIe is full duplex)
    cudaStream_t in_stream{nullptr};
    std::vector<cudaEvent_t> in_events;
    // staging for the host-buffer entry points
    trx::scratch d_in[4];
    trx::scratch d_out[2];
    trx::scratch h_pin[2];
    // h_pin[1] doubles as an arena for small index lists that kernels read straight from pinned host memory without a
    // sync of their own (trx_paths_or); the offset goes back to 0 at every stream synchronisation of the path algebra
    size_t pin_arena_off{0};
    int use() const; // cudaSetDevice
};

struct trx_ibf
{
    trx_ctx * ctx{nullptr};
    uint64_t bins{}, technical_bins{}, bin_size{}, hash_shift{}, bin_words{}, hash_funs{}, n_words{};
    uint64_t shard_word_begin{}, shard_words{}, shard_n_words{};
    uint64_t * d_words{nullptr}; // shard_n_words words, row r at r*shard_words
    bool owns_words{true};
    trx::scratch slab;           // K1: dense copy of a few word columns while a slab of bins is being emplaced
};

struct trx_paths
{
    trx_ctx * ctx{nullptr};
    uint64_t words_per_slot{}, capacity{};
    uint64_t * d_words{nullptr};
};

namespace trx
{

// seqan::hibf::interleaved_bloom_filter::hash_seeds (lib/hibf/include/hibf/interleaved_bloom_filter.hpp:164-168)
__host__ __device__ __forceinline__ constexpr uint64_t hash_seed(int i)
{
    return i == 0   ? 13572355802537770549ULL
           : i == 1 ? 13043817825332782213ULL
           : i == 2 ? 10650232656628343401ULL
           : i == 3 ? 16499269484942379435ULL
                    : 4893150838803335377ULL;
}

// hash_and_fit without the final "* technical_bins" (…hpp:177-187): the row index r in [0, bin_size).
__device__ __forceinline__ uint64_t hash_row(uint64_t v, int i, uint32_t hash_shift, uint64_t bin_size)
{
    uint64_t h = v * hash_seed(i);
    h ^= h >> hash_shift;
    h *= 11400714819323198485ULL;
    return __umul64hi(h, bin_size);
}

inline unsigned grid_for(uint64_t work_items, unsigned per_block, int sm_count, unsigned blocks_per_sm)
{
    uint64_t need = (work_items + per_block - 1) / per_block;
    uint64_t cap = static_cast<uint64_t>(sm_count) * blocks_per_sm;
    if (need < 1)
        need = 1;
    return static_cast<unsigned>(need < cap ? need : cap);
}

} // namespace trx
