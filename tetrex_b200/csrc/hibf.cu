// hibf.cu — K3: HIBF membership (threshold 1, one value per query) as a level-synchronous frontier.
//
// Reference: membership_agent_type::membership_for_impl, lib/hibf/include/hibf/
// hierarchical_interleaved_bloom_filter.hpp:321-355, called by HIBFIndex::query
// (include/index_hibf.h:132-147) with {kmer} and threshold 1.  With one value and threshold 1 the
// running `sum` over a split bin degenerates to OR: a user bin is reported iff any of its technical
// bins hits in an IBF that is reached through hitting merged bins.  The recursion becomes a frontier
// of (k-mer index, IBF id) items; one kernel launch per tree level, no host round trip in between
// (the next level's item count stays in device memory).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace trx
{

struct hibf_ibf_desc
{
    uint64_t word_off;  // into the word arena
    uint64_t bin_size;
    uint32_t bin_words;
    uint32_t hash_shift;
    uint32_t hash_funs;
    uint32_t table_off; // into next_ibf / user_bin tables (padded to bin_words*64 per IBF); /64 → mask index
};

constexpr uint32_t kMerged = 0xFFFFFFFFu;
constexpr uint32_t kDeleted = 0xFFFFFFFEu;

struct hibf_dev
{
    uint64_t const * words;
    hibf_ibf_desc const * ibfs;
    uint32_t const * next_ibf;
    uint32_t const * user_bin;
    uint64_t const * merged_mask; // per (ibf, word): technical bins that are merged bins
    uint64_t const * user_mask;   // per (ibf, word): technical bins that carry a user bin
    uint32_t out_words;           // ceil(n_user_bins/64)
};

// counters[0] = overflow flag, counters[1 + L] = number of items of level L
template <bool ROOT>
__global__ void __launch_bounds__(256)
    hibf_level_kernel(hibf_dev h, uint64_t const * __restrict__ kmers, uint64_t n_root,
                      uint64_t const * __restrict__ frontier_in, uint64_t * __restrict__ frontier_out,
                      unsigned long long * __restrict__ counters, int level, uint64_t capacity,
                      uint64_t * __restrict__ out)
{
    uint64_t const n_in = ROOT ? n_root : min(static_cast<uint64_t>(counters[1 + level]), capacity);
    unsigned const lane = threadIdx.x & 31;
    uint64_t const tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    uint64_t const stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    // whole warps iterate together so that the warp-aggregated append below sees all 32 lanes
    uint64_t const n_rounded = (n_in + 31) & ~uint64_t{31};
    for (uint64_t i = tid; i < n_rounded; i += stride)
    {
        bool const active = i < n_in;
        uint32_t kidx = 0, ibf_id = 0;
        if (active)
        {
            if constexpr (ROOT)
            {
                kidx = static_cast<uint32_t>(i);
            }
            else
            {
                uint64_t item = frontier_in[i];
                kidx = static_cast<uint32_t>(item);
                ibf_id = static_cast<uint32_t>(item >> 32);
            }
        }
        hibf_ibf_desc const d = h.ibfs[ibf_id];
        uint64_t const v = active ? __ldg(kmers + kidx) : 0;
        uint64_t rowbase[5];
#pragma unroll
        for (int f = 0; f < 5; ++f)
            rowbase[f] = (static_cast<uint32_t>(f) < d.hash_funs)
                             ? d.word_off + hash_row(v, f, d.hash_shift, d.bin_size) * d.bin_words
                             : 0;
        // the trip count must be warp-uniform (shuffles below): iterate to the widest IBF in the warp
        uint32_t wmax = active ? d.bin_words : 0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
            wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, off));
        for (uint32_t w = 0; w < wmax; ++w) // 1 for the usual 64-technical-bin IBFs
        {
            bool const live = active && w < d.bin_words;
            uint64_t hit = 0;
            if (live)
            {
                hit = ~0ULL;
#pragma unroll
                for (int f = 0; f < 5; ++f)
                    if (static_cast<uint32_t>(f) < d.hash_funs)
                        hit &= __ldg(h.words + rowbase[f] + w);
            }
            uint32_t const mask_idx = d.table_off / 64 + w;
            uint64_t children = live ? (hit & __ldg(h.merged_mask + mask_idx)) : 0;
            uint64_t users = live ? (hit & __ldg(h.user_mask + mask_idx)) : 0;
            while (users)
            {
                int b = __ffsll(static_cast<long long>(users)) - 1;
                users &= users - 1;
                uint32_t ub = __ldg(h.user_bin + d.table_off + w * 64 + b);
                atomicOr(reinterpret_cast<unsigned long long *>(out + static_cast<uint64_t>(kidx) * h.out_words + (ub >> 6)),
                         1ULL << (ub & 63));
            }
            // warp-aggregated append of the children: one atomicAdd per warp
            unsigned const cnt = __popcll(children);
            unsigned incl = cnt;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1)
            {
                unsigned t = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= static_cast<unsigned>(off))
                    incl += t;
            }
            unsigned const total = __shfl_sync(0xffffffffu, incl, 31);
            if (total == 0)
                continue;
            unsigned long long base = 0;
            if (lane == 31)
                base = atomicAdd(counters + 2 + level, static_cast<unsigned long long>(total));
            base = __shfl_sync(0xffffffffu, base, 31);
            uint64_t pos = base + (incl - cnt);
            while (children)
            {
                int b = __ffsll(static_cast<long long>(children)) - 1;
                children &= children - 1;
                if (pos < capacity)
                    frontier_out[pos] = (static_cast<uint64_t>(__ldg(h.next_ibf + d.table_off + w * 64 + b)) << 32) | kidx;
                else
                    counters[0] = 1; // overflow: the host re-runs this chunk in smaller pieces
                ++pos;
            }
        }
    }
}

} // namespace trx

struct trx_hibf
{
    trx_ctx * ctx{nullptr};
    uint64_t n_ibfs{}, n_user_bins{}, out_words{};
    int depth{1};
    uint64_t * d_words{nullptr};
    trx::hibf_ibf_desc * d_ibfs{nullptr};
    uint32_t * d_next{nullptr};
    uint32_t * d_user{nullptr};
    uint64_t * d_merged_mask{nullptr};
    uint64_t * d_user_mask{nullptr};
    // frontier scratch, grown on demand
    trx::scratch frontier[2];
    trx::scratch counters;
    trx::scratch d_kmers, d_out, sort_tmp;
    int ibf_bits{1}; // bits needed for an IBF id (radix-sort range)
    bool sort_frontier{true};
    trx::hibf_dev dev() const
    {
        return trx::hibf_dev{d_words, d_ibfs, d_next, d_user, d_merged_mask, d_user_mask, static_cast<uint32_t>(out_words)};
    }
};

using namespace trx;

static constexpr uint64_t kChunk = 1ull << 20;   // k-mers per frontier pass
static constexpr uint64_t kFanoutBudget = 32;    // frontier capacity = kFanoutBudget × chunk items
static constexpr int kMaxLevels = 30;

extern "C"
{

int trx_hibf_upload(trx_ctx_t * ctx, const trx_hibf_desc * desc, trx_hibf_t ** out)
{
    if (!ctx || !desc || !out)
        return set_error(TRX_EINVAL, "NULL argument");
    *out = nullptr;
    uint64_t const n = desc->n_ibfs;
    if (n == 0 || desc->n_user_bins == 0)
        return set_error(TRX_EINVAL, "empty HIBF");
    std::vector<hibf_ibf_desc> ibfs(n);
    uint64_t word_total = 0, table_total = 0;
    for (uint64_t i = 0; i < n; ++i)
    {
        uint64_t bins = desc->bins[i], bs = desc->bin_size[i], hf = desc->hash_funs[i];
        if (bins == 0 || bs == 0 || hf == 0 || hf > 5)
            return set_error(TRX_EINVAL, "IBF %llu: invalid geometry", (unsigned long long)i);
        uint64_t bw = (bins + 63) / 64;
        if (desc->table_len[i] > bw * 64)
            return set_error(TRX_EINVAL, "IBF %llu: table longer than its technical bins", (unsigned long long)i);
        ibfs[i] = hibf_ibf_desc{word_total, bs, static_cast<uint32_t>(bw), static_cast<uint32_t>(__builtin_clzll(bs)),
                                static_cast<uint32_t>(hf), static_cast<uint32_t>(table_total)};
        word_total += bw * bs;
        table_total += bw * 64;
    }
    if (table_total >= (1ull << 32))
        return set_error(TRX_EINVAL, "HIBF too large for 32-bit table offsets");
    std::vector<uint32_t> next(table_total, 0), user(table_total, kDeleted);
    std::vector<uint64_t> mmask(table_total / 64, 0), umask(table_total / 64, 0);
    std::vector<int> level(n, -1);
    level[0] = 1;
    int depth = 1;
    for (uint64_t i = 0; i < n; ++i)
    {
        for (uint64_t t = 0; t < desc->table_len[i]; ++t)
        {
            uint64_t ub = desc->user_bin[i][t];
            uint64_t idx = ibfs[i].table_off + t;
            if (ub == UINT64_MAX)
            {
                uint64_t nx = desc->next_ibf[i][t];
                if (nx >= n)
                    return set_error(TRX_EINVAL, "IBF %llu bin %llu: next_ibf_id out of range", (unsigned long long)i,
                                     (unsigned long long)t);
                // an unassigned technical bin keeps {merged, self} (hierarchical_interleaved_bloom_filter.cpp:50-56);
                // it never has bits, and following it would loop: treat as deleted.
                if (nx == i)
                    continue;
                user[idx] = kMerged;
                next[idx] = static_cast<uint32_t>(nx);
                mmask[idx / 64] |= 1ULL << (idx % 64);
            }
            else if (ub == UINT64_MAX - 1)
            {
                continue;
            }
            else
            {
                if (ub >= desc->n_user_bins)
                    return set_error(TRX_EINVAL, "IBF %llu bin %llu: user bin id out of range", (unsigned long long)i,
                                     (unsigned long long)t);
                user[idx] = static_cast<uint32_t>(ub);
                umask[idx / 64] |= 1ULL << (idx % 64);
            }
        }
    }
    // depth of the tree = number of level launches
    {
        std::vector<uint64_t> cur{0}, nxt;
        std::vector<char> seen(n, 0);
        seen[0] = 1;
        while (!cur.empty() && depth < kMaxLevels)
        {
            nxt.clear();
            for (uint64_t i : cur)
                for (uint64_t t = 0; t < static_cast<uint64_t>(ibfs[i].bin_words) * 64; ++t)
                    if (user[ibfs[i].table_off + t] == kMerged)
                    {
                        uint32_t c = next[ibfs[i].table_off + t];
                        if (!seen[c])
                        {
                            seen[c] = 1;
                            nxt.push_back(c);
                        }
                    }
            if (nxt.empty())
                break;
            ++depth;
            cur.swap(nxt);
        }
    }
    TRX_TRY(ctx->use());
    auto * h = new (std::nothrow) trx_hibf;
    if (!h)
        return set_error(TRX_ENOMEM, "out of host memory");
    h->ctx = ctx;
    h->n_ibfs = n;
    h->n_user_bins = desc->n_user_bins;
    h->out_words = (desc->n_user_bins + 63) / 64;
    h->depth = depth;
    while ((1ull << h->ibf_bits) < n)
        ++h->ibf_bits;
    {
        char const * e = getenv("TRX_HIBF_SORT"); // TRX_HIBF_SORT=off: A/B evidence for the frontier sort
        h->sort_frontier = !(e && strcmp(e, "off") == 0);
    }
    auto fail = [&](cudaError_t e)
    {
        trx_hibf_destroy(h);
        return set_error(TRX_ECUDA, "trx_hibf_upload: %s", cudaGetErrorString(e));
    };
    cudaError_t e;
    if ((e = cudaMalloc(&h->d_words, word_total * 8)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&h->d_ibfs, n * sizeof(hibf_ibf_desc))) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&h->d_next, table_total * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&h->d_user, table_total * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&h->d_merged_mask, table_total / 8)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&h->d_user_mask, table_total / 8)) != cudaSuccess) return fail(e);
    for (uint64_t i = 0; i < n; ++i)
    {
        uint64_t nw = static_cast<uint64_t>(ibfs[i].bin_words) * ibfs[i].bin_size;
        if ((e = cudaMemcpyAsync(h->d_words + ibfs[i].word_off, desc->words[i], nw * 8, cudaMemcpyHostToDevice,
                                 ctx->stream)) != cudaSuccess)
            return fail(e);
    }
    if ((e = cudaMemcpyAsync(h->d_ibfs, ibfs.data(), n * sizeof(hibf_ibf_desc), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(h->d_next, next.data(), table_total * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(h->d_user, user.data(), table_total * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(h->d_merged_mask, mmask.data(), table_total / 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(h->d_user_mask, umask.data(), table_total / 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e);
    *out = h;
    return TRX_OK;
}

void trx_hibf_destroy(trx_hibf_t * h)
{
    if (!h)
        return;
    cudaSetDevice(h->ctx->device);
    cudaStreamSynchronize(h->ctx->stream);
    cudaFree(h->d_words);
    cudaFree(h->d_ibfs);
    cudaFree(h->d_next);
    cudaFree(h->d_user);
    cudaFree(h->d_merged_mask);
    cudaFree(h->d_user_mask);
    h->frontier[0].release();
    h->frontier[1].release();
    h->counters.release();
    h->d_kmers.release();
    h->d_out.release();
    h->sort_tmp.release();
    delete h;
}

uint64_t trx_hibf_user_bin_words(const trx_hibf_t * h)
{
    return h ? h->out_words : 0;
}

// one chunk of k-mers through all levels; returns TRX_ECAPACITY on frontier overflow
static int run_chunk(trx_hibf * h, uint64_t const * d_kmers, uint64_t n, uint64_t * d_out, uint64_t * visited)
{
    trx_ctx * c = h->ctx;
    uint64_t const capacity = kFanoutBudget * (n < 4096 ? 4096 : n);
    TRX_TRY(h->frontier[0].reserve(capacity * 8));
    TRX_TRY(h->frontier[1].reserve(capacity * 8));
    TRX_TRY(h->counters.reserve((kMaxLevels + 3) * 8));
    auto * ctr = static_cast<unsigned long long *>(h->counters.ptr);
    TRX_CUDA(cudaMemsetAsync(ctr, 0, (kMaxLevels + 3) * 8, c->stream));
    TRX_CUDA(cudaMemsetAsync(d_out, 0, n * h->out_words * 8, c->stream));
    unsigned const grid = grid_for(n, 256, c->sm_count, 8);
    hibf_level_kernel<true><<<grid, 256, 0, c->stream>>>(h->dev(), d_kmers, n, nullptr,
                                                         static_cast<uint64_t *>(h->frontier[0].ptr), ctr, 0, capacity, d_out);
    ++c->launches;
    unsigned const full_grid = static_cast<unsigned>(c->sm_count) * 8;
    // ping-pong: `cur` holds the items of the level about to run, `other` is free
    int cur = 0;
    for (int level = 1; level < h->depth; ++level)
    {
        if (h->sort_frontier)
        {
            // Group the frontier by IBF id (high word of an item) so that the probes of one lower-level IBF
            // (a few MB) run together and hit L2 instead of spreading 3 random 64-byte DRAM fetches per item
            // over the whole HIBF.  The item count comes back with one 8-byte read.
            unsigned long long n_items = 0;
            TRX_CUDA(cudaMemcpyAsync(&n_items, ctr + 1 + level, 8, cudaMemcpyDeviceToHost, c->stream));
            TRX_CUDA(cudaStreamSynchronize(c->stream));
            if (n_items == 0)
                break;
            if (n_items > capacity)
                return TRX_ECAPACITY;
            auto * in = static_cast<uint64_t *>(h->frontier[cur].ptr);
            auto * out_buf = static_cast<uint64_t *>(h->frontier[cur ^ 1].ptr);
            size_t tmp = 0;
            TRX_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp, in, out_buf, static_cast<int64_t>(n_items), 32,
                                                    32 + h->ibf_bits, c->stream));
            TRX_TRY(h->sort_tmp.reserve(tmp));
            TRX_CUDA(cub::DeviceRadixSort::SortKeys(h->sort_tmp.ptr, tmp, in, out_buf, static_cast<int64_t>(n_items), 32,
                                                    32 + h->ibf_bits, c->stream));
            c->launches += 2;
            cur ^= 1;
        }
        hibf_level_kernel<false><<<full_grid, 256, 0, c->stream>>>(
            h->dev(), d_kmers, 0, static_cast<uint64_t *>(h->frontier[cur].ptr),
            static_cast<uint64_t *>(h->frontier[cur ^ 1].ptr), ctr, level, capacity, d_out);
        ++c->launches;
        cur ^= 1;
    }
    TRX_CUDA(cudaGetLastError());
    unsigned long long host_ctr[kMaxLevels + 3];
    TRX_CUDA(cudaMemcpyAsync(host_ctr, ctr, sizeof host_ctr, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    if (host_ctr[0])
        return TRX_ECAPACITY;
    uint64_t v = n;
    for (int level = 1; level < h->depth; ++level)
        v += host_ctr[1 + level];
    *visited += v;
    return TRX_OK;
}

int trx_hibf_membership_dev(trx_hibf_t * h, const uint64_t * d_kmers, uint64_t n, uint64_t * d_out_words,
                            uint64_t * stats)
{
    if (!h || (n && (!d_kmers || !d_out_words)))
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(h->ctx->use());
    uint64_t visited = 0;
    uint64_t done = 0;
    uint64_t chunk = kChunk;
    while (done < n)
    {
        uint64_t m = n - done < chunk ? n - done : chunk;
        int rc = run_chunk(h, d_kmers + done, m, d_out_words + done * h->out_words, &visited);
        if (rc == TRX_ECAPACITY)
        {
            if (chunk <= 64)
                return set_error(TRX_ECAPACITY, "HIBF frontier overflow even at 64 k-mers per pass");
            chunk /= 4; // fan-out larger than budgeted: shrink the pass and redo this chunk
            continue;
        }
        TRX_TRY(rc);
        done += m;
    }
    if (stats)
    {
        stats[0] = visited;
        stats[1] = static_cast<uint64_t>(h->depth);
    }
    return TRX_OK;
}

int trx_hibf_membership(trx_hibf_t * h, const uint64_t * kmers, uint64_t n, uint64_t * out_words, uint64_t * stats)
{
    if (!h || (n && (!kmers || !out_words)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
    {
        if (stats)
            stats[0] = 0, stats[1] = static_cast<uint64_t>(h->depth);
        return TRX_OK;
    }
    trx_ctx * c = h->ctx;
    TRX_TRY(c->use());
    TRX_TRY(h->d_kmers.reserve(n * 8));
    TRX_TRY(h->d_out.reserve(n * h->out_words * 8));
    TRX_CUDA(cudaMemcpyAsync(h->d_kmers.ptr, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_TRY(trx_hibf_membership_dev(h, static_cast<uint64_t *>(h->d_kmers.ptr), n, static_cast<uint64_t *>(h->d_out.ptr), stats));
    TRX_CUDA(cudaMemcpyAsync(out_words, h->d_out.ptr, n * h->out_words * 8, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

} // extern "C"
