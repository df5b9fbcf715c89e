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

#include <algorithm>

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

// ---------------------------------------------------------------------------------------------------------------
// Root + second level without a frontier: "windowed" schedule.
//
// The HIBF (0.37 GB at C2) does not fit the 126 MB L2, but a WINDOW of a few neighbouring second-level IBFs does.
// hibf_root_kernel probes the root for every k-mer of the batch and keeps the merged-bin hits (one 64-bit word per
// root word) in a scratch array.  Then one hibf_window_kernel launch per window walks ALL k-mers of the batch
// against the children of that window only: the window's matrices are read from DRAM once per batch and stay in L2
// (loads carry an evict_last policy) while millions of probes hit them.  Everything that merely streams through —
// k-mers, root words, results — is read and written with evict_first so that it does not push the window out: a
// scattered 8-byte atomic into the 512 MB result array allocates a whole 128-byte L2 line, and ~10 M of those per
// pass were measured to flush L2 ten times over.  Results therefore never go to the result array directly: a
// window's children report into a narrow run of user-bin words [word_lo, word_lo+span); the warp accumulates those
// words in shared memory and stores them, coalesced, into a per-window strip; hibf_merge_kernel finally ORs the
// strips into the caller's rows.  The frontier kernels above get their locality from a radix sort and pay 16 B of
// frontier traffic per (k-mer, child) pair plus the sort; here a pair costs nothing but its probe.
// Inside the window kernel a warp owns 32 k-mers (lane = k-mer); a k-mer reaches only about a third of the
// children, so the warp gathers its (k-mer, child) pairs through ballots into a shared-memory queue and probes 32
// pairs at a time with all lanes busy.  Merged bins hit on the second level (HIBFs with three or more levels) are
// appended to the frontier and handled by hibf_level_kernel<false> after the merge.
constexpr int kWinThreads = 256;
constexpr uint32_t kMaxSpan = 24;    // user-bin words one window (or the root) may report into
constexpr uint32_t kMaxWindowBins = 16; // root technical bins per window (shared-memory tables and pair queue)

struct hibf_window
{
    uint32_t bin_begin, bin_end; // root technical bins of this window
    uint32_t word_lo, span;      // user-bin words [word_lo, word_lo + span) its children report into
    uint64_t strip_off;          // offset of the window's strip, in units of n words (Σ span of earlier windows)
};

__device__ __forceinline__ uint64_t policy_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t ld_hint(uint64_t const * p, uint64_t policy)
{
    uint64_t v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(policy));
    return v;
}
__device__ __forceinline__ void st_hint(uint64_t * p, uint64_t v, uint64_t policy)
{
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.u64 [%0], %1, %2;" ::"l"(p), "l"(v), "l"(policy) : "memory");
}

// FAST: every IBF has one word per row (≤ 64 technical bins) and fewer than 2^32 rows — the shape seqan's layout
// produces for tmax = 64; the fastrange product then needs two 32x32 multiplies instead of a 64x64 high multiply
// calls f(bit index) for every set bit, low half first: 32-bit find-first-set is one FLO, the 64-bit one a dozen instructions
template <class F>
__device__ __forceinline__ void for_each_bit(uint64_t x, F f)
{
    uint32_t lo = static_cast<uint32_t>(x), hi = static_cast<uint32_t>(x >> 32);
    while (lo)
    {
        int b = __ffs(static_cast<int>(lo)) - 1;
        lo &= lo - 1;
        f(b);
    }
    while (hi)
    {
        int b = __ffs(static_cast<int>(hi)) - 1;
        hi &= hi - 1;
        f(b + 32);
    }
}

template <int H, bool FAST>
__device__ __forceinline__ uint64_t probe_word(hibf_dev const & h, hibf_ibf_desc const & d, uint64_t kmer, uint32_t w,
                                               uint64_t policy)
{
    uint64_t hit = ~0ULL;
#pragma unroll
    for (int f = 0; f < 5; ++f)
    {
        if (H ? f < H : static_cast<uint32_t>(f) < d.hash_funs)
        {
            if constexpr (FAST)
            {
                uint64_t x = kmer * hash_seed(f);
                x ^= x >> d.hash_shift;
                x *= 11400714819323198485ULL;
                uint64_t const bs = static_cast<uint32_t>(d.bin_size);
                uint64_t const row = ((x >> 32) * bs + (((x & 0xffffffffULL) * bs) >> 32)) >> 32;
                hit &= ld_hint(h.words + d.word_off + row, policy);
            }
            else
                hit &= ld_hint(h.words + d.word_off + hash_row(kmer, f, d.hash_shift, d.bin_size) * d.bin_words + w, policy);
        }
    }
    return hit;
}

// root_words[i * bin_words(root) + w] = merged-bin hits of k-mer i in root word w.  root_strip (n × root.span words,
// zeroed) receives the user bins that sit directly on the root.
template <int H, bool FAST>
__global__ void __launch_bounds__(256)
    hibf_root_kernel(hibf_dev h, uint64_t const * __restrict__ kmers, uint64_t n, uint64_t * __restrict__ root_words,
                     hibf_window root, uint64_t * __restrict__ root_strip, unsigned long long * __restrict__ counters)
{
    hibf_ibf_desc const d0 = h.ibfs[0];
    uint64_t const keep = policy_evict_last(), stream = policy_evict_first();
    unsigned long long reached = 0;
    uint64_t const stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        uint64_t const v = ld_hint(kmers + i, stream);
        for (uint32_t w = 0; w < d0.bin_words; ++w)
        {
            uint64_t const hit = probe_word<H, FAST>(h, d0, v, w, keep);
            uint32_t const mask_idx = d0.table_off / 64 + w;
            uint64_t users = hit & __ldg(h.user_mask + mask_idx);
            uint64_t const children = hit & __ldg(h.merged_mask + mask_idx);
            st_hint(root_words + i * d0.bin_words + w, children, stream);
            reached += __popcll(children);
            while (users)
            {
                int b = __ffsll(static_cast<long long>(users)) - 1;
                users &= users - 1;
                uint32_t ub = __ldg(h.user_bin + d0.table_off + w * 64 + b);
                atomicOr(reinterpret_cast<unsigned long long *>(root_strip + i * root.span + ((ub >> 6) - root.word_lo)),
                         1ULL << (ub & 63));
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
        reached += __shfl_xor_sync(0xffffffffu, reached, off);
    if ((threadIdx.x & 31) == 0 && reached)
        atomicAdd(counters + 2, reached); // = number of level-1 items, what the frontier path counts there
}

// Everything a pair needs about its child, for the (at most kMaxWindowBins) root bins of the window, in shared
// memory: with the tables in global memory every pair made three more L1 requests (masks, user-bin ids) that missed
// L1 most of the time because the random row sectors stream through it.
struct window_child
{
    uint64_t word_off;
    uint64_t bin_size;
    uint32_t bin_words, hash_shift, hash_funs, table_off;
    uint64_t user_mask, merged_mask; // of word 0 (FAST shape: the only word)
};

template <int H, bool FAST>
__global__ void __launch_bounds__(kWinThreads)
    hibf_window_kernel(hibf_dev h, uint64_t const * __restrict__ kmers, uint64_t n, uint64_t const * __restrict__ root_words,
                       hibf_window win, uint64_t * __restrict__ strip, uint64_t * __restrict__ frontier_out,
                       unsigned long long * __restrict__ counters, uint64_t capacity)
{
    extern __shared__ uint64_t win_acc[]; // kWinThreads rows of (span + 1) words
    __shared__ window_child s_child[kMaxWindowBins];
    __shared__ uint16_t s_ub[FAST ? kMaxWindowBins : 1][64]; // user bin of (window bin, technical bin); FAST: ids < 65536
    __shared__ uint16_t s_queue[kWinThreads / 32][32 * kMaxWindowBins];
    hibf_ibf_desc const d0 = h.ibfs[0];
    uint32_t const win_bins = min(win.bin_end - win.bin_begin, kMaxWindowBins);
    if (threadIdx.x < win_bins)
    {
        uint32_t const bin = win.bin_begin + threadIdx.x;
        uint32_t const idx = d0.table_off + bin;
        window_child c{};
        if (h.merged_mask[idx / 64] >> (idx % 64) & 1)
        {
            hibf_ibf_desc const d = h.ibfs[h.next_ibf[idx]];
            c = window_child{d.word_off, d.bin_size, d.bin_words, d.hash_shift, d.hash_funs, d.table_off,
                             h.user_mask[d.table_off / 64], h.merged_mask[d.table_off / 64]};
        }
        s_child[threadIdx.x] = c;
    }
    __syncthreads();
    if constexpr (FAST)
    {
        for (uint32_t t = threadIdx.x; t < win_bins * 64; t += kWinThreads)
            s_ub[t >> 6][t & 63] = static_cast<uint16_t>(h.user_bin[s_child[t >> 6].table_off + (t & 63)]);
        __syncthreads();
    }
    uint64_t const keep = policy_evict_last(), stream = policy_evict_first();
    unsigned const lane = threadIdx.x & 31;
    unsigned const warp = threadIdx.x >> 5;
    uint32_t const win_mask = win_bins >= 32 ? ~0u : ((1u << win_bins) - 1);
    uint32_t const lt_mask = (1u << lane) - 1;
    uint32_t const span = win.span, pitch = span + 1;
    uint32_t const inv_span = span ? (65536 + span - 1) / span : 0;
    uint16_t * const queue = s_queue[warp];
    uint64_t * const warp_acc = win_acc + static_cast<size_t>(warp * 32) * pitch;
    uint64_t const warps = static_cast<uint64_t>(gridDim.x) * (kWinThreads / 32);
    uint64_t const n_groups = (n + 31) / 32;
    for (uint64_t g = static_cast<uint64_t>(blockIdx.x) * (kWinThreads / 32) + warp; g < n_groups; g += warps)
    {
        uint64_t const base = g * 32;
        uint64_t const i = base + lane;
        bool const active = i < n;
        uint64_t const v = active ? ld_hint(kmers + i, stream) : 0;
        for (uint32_t w = 0; w < span; ++w)
            warp_acc[lane * pitch + w] = 0;
        __syncwarp();
        // second level for one queued pair per lane; every lane takes part in the shuffle
        auto process = [&](bool valid, uint32_t entry) {
            unsigned const src = entry & 31;
            uint32_t const wb = entry >> 5; // bin inside the window
            uint64_t const kmer = __shfl_sync(0xffffffffu, v, src);
            if (!valid)
                return;
            window_child const & c = s_child[wb];
            hibf_ibf_desc const dc{c.word_off, c.bin_size, c.bin_words, c.hash_shift, c.hash_funs, c.table_off};
            unsigned * const row32 = reinterpret_cast<unsigned *>(warp_acc + src * pitch) - 2 * win.word_lo;
            for (uint32_t w2 = 0; w2 < (FAST ? 1u : dc.bin_words); ++w2)
            {
                uint64_t const hit2 = probe_word<H, FAST>(h, dc, kmer, w2, keep);
                if (hit2 == 0)
                    continue;
                uint64_t const users2 = hit2 & (FAST ? c.user_mask : __ldg(h.user_mask + dc.table_off / 64 + w2));
                uint64_t const deeper = hit2 & (FAST ? c.merged_mask : __ldg(h.merged_mask + dc.table_off / 64 + w2));
                for_each_bit(users2, [&](int b2) {
                    uint32_t ub;
                    if constexpr (FAST)
                        ub = s_ub[wb][b2];
                    else
                        ub = __ldg(h.user_bin + dc.table_off + w2 * 64 + b2);
                    // two lanes may hold pairs of the same k-mer: shared-memory atomic, on the 32-bit half (64-bit
                    // shared atomics other than CAS compile to a CAS loop)
                    atomicOr(row32 + (ub >> 5), 1u << (ub & 31));
                });
                for_each_bit(deeper, [&](int b2) { // third level and below: hand over to the frontier kernels
                    unsigned long long pos = atomicAdd(counters + 3, 1ULL);
                    if (pos < capacity)
                        frontier_out[pos] = (static_cast<uint64_t>(__ldg(h.next_ibf + dc.table_off + w2 * 64 + b2)) << 32) |
                                            static_cast<uint32_t>(base + src);
                    else
                        counters[0] = 1;
                });
            }
        };
        // gather the warp's (k-mer, root bin) pairs, child-major: all lanes that reach bin b, then bin b + 1, …  A
        // window never crosses a root word and has at most kMaxWindowBins bins.
        uint64_t const root_word = active ? ld_hint(root_words + i * d0.bin_words + (win.bin_begin >> 6), stream) : 0;
        uint32_t const children = static_cast<uint32_t>(root_word >> (win.bin_begin & 63)) & win_mask;
        uint32_t total = 0;
        for (uint32_t b = 0; b < win_bins; ++b)
        {
            bool const want = (children >> b) & 1;
            unsigned const m = __ballot_sync(0xffffffffu, want);
            if (want)
                queue[total + __popc(m & lt_mask)] = static_cast<uint16_t>((b << 5) | lane);
            total += __popc(m);
        }
        __syncwarp();
        for (uint32_t r = 0; r < total; r += 32)
        {
            bool const valid = r + lane < total;
            process(valid, valid ? queue[r + lane] : lane);
        }
        __syncwarp();
        // the warp's 32 strip rows are contiguous: consecutive lanes store consecutive words
        uint64_t const rows = base + 32 <= n ? 32 : n - base;
        uint64_t * const dst = strip + base * span;
        for (uint32_t idx = lane; idx < rows * span; idx += 32)
        {
            uint32_t const r = (idx * inv_span) >> 16; // idx / span for idx < 32 * kMaxSpan
            st_hint(dst + idx, warp_acc[r * pitch + (idx - r * span)], stream);
        }
        __syncwarp();
    }
}

// out[i][w] = OR of every strip that covers user-bin word w (windows[0] is the root's own strip, span 0 = none).
// A thread keeps one word column w for all its rows, so the covering strips are looked up once.
__global__ void __launch_bounds__(256)
    hibf_merge_kernel(hibf_window const * __restrict__ windows, uint32_t n_windows, uint64_t const * __restrict__ strips,
                      uint64_t n, uint32_t out_words, uint64_t * __restrict__ out)
{
    uint64_t const stream = policy_evict_first();
    uint32_t const cols = out_words < 256 ? out_words : 256;
    uint32_t const rows_per_block = 256 / cols;
    uint32_t const r_in_block = threadIdx.x / cols;
    if (r_in_block >= rows_per_block)
        return;
    for (uint32_t w = threadIdx.x % cols; w < out_words; w += cols)
    {
        constexpr int kMaxCover = 4;
        uint64_t const * src[kMaxCover];
        uint32_t span[kMaxCover];
        int n_cover = 0;
        bool overflow = false;
        for (uint32_t g = 0; g < n_windows; ++g)
        {
            hibf_window const win = windows[g];
            if (w - win.word_lo < win.span) // unsigned: also false for w < word_lo
            {
                if (n_cover < kMaxCover)
                {
                    src[n_cover] = strips + win.strip_off * n + (w - win.word_lo);
                    span[n_cover] = win.span;
                    ++n_cover;
                }
                else
                    overflow = true;
            }
        }
        for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * rows_per_block + r_in_block; i < n;
             i += static_cast<uint64_t>(gridDim.x) * rows_per_block)
        {
            uint64_t acc = 0;
#pragma unroll
            for (int c = 0; c < kMaxCover; ++c)
                if (c < n_cover)
                    acc |= ld_hint(src[c] + i * span[c], stream);
            if (overflow) // more than kMaxCover strips cover this word: the general loop
            {
                acc = 0;
                for (uint32_t g = 0; g < n_windows; ++g)
                {
                    hibf_window const win = windows[g];
                    if (w - win.word_lo < win.span)
                        acc |= ld_hint(strips + win.strip_off * n + i * win.span + (w - win.word_lo), stream);
                }
            }
            st_hint(out + i * out_words + w, acc, stream);
        }
    }
}

// Same result through a shared-memory tile: a warp takes 32 rows, pulls every strip's 32 x span block in with
// fully coalesced loads, ORs it into the tile and writes the 32 finished rows (contiguous in `out`) with
// consecutive lanes on consecutive words.  Used when a 32-row tile fits (out_words <= kMergeTileWords).
constexpr uint32_t kMergeWarps = 4;
constexpr uint32_t kMergeTileWords = 64;

__global__ void __launch_bounds__(kMergeWarps * 32)
    hibf_merge_tile_kernel(hibf_window const * __restrict__ windows, uint32_t n_windows, uint64_t const * __restrict__ strips,
                           uint64_t n, uint32_t out_words, uint64_t * __restrict__ out, uint32_t const * __restrict__ inv_perm)
{
    extern __shared__ uint64_t merge_tile[];
    uint64_t const stream = policy_evict_first();
    unsigned const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t const pitch = out_words + 1;
    uint64_t * const tile = merge_tile + static_cast<size_t>(warp) * 32 * pitch;
    // idx / out_words for idx < 32 * kMergeTileWords = 2048: a 20-bit reciprocal is exact there (error < 2048 / 2^20 <
    // 1 / 64); the 16-bit one the window kernel uses for its narrow spans is not (63 words: 2015 * 1041 >> 16 = 32)
    uint32_t const inv_w = ((1u << 20) + out_words - 1) / out_words;
    uint64_t const n_groups = (n + 31) / 32;
    for (uint64_t g = static_cast<uint64_t>(blockIdx.x) * kMergeWarps + warp; g < n_groups; g += static_cast<uint64_t>(gridDim.x) * kMergeWarps)
    {
        uint64_t const base = g * 32;
        uint32_t const rows = base + 32 <= n ? 32 : static_cast<uint32_t>(n - base);
        for (uint32_t idx = lane; idx < 32 * pitch; idx += 32)
            tile[idx] = 0;
        __syncwarp();
        for (uint32_t k = 0; k < n_windows; ++k)
        {
            hibf_window const win = windows[k];
            if (win.span == 0)
                continue;
            uint32_t const inv = (65536 + win.span - 1) / win.span;
            uint64_t const * const src = strips + win.strip_off * n + base * win.span;
            if (!inv_perm)
            {
                for (uint32_t idx = lane; idx < rows * win.span; idx += 32)
                {
                    uint32_t const r = (idx * inv) >> 16;
                    tile[r * pitch + win.word_lo + (idx - r * win.span)] |= ld_hint(src + idx, stream);
                }
            }
            else
            {
                // strips live in the PERMUTED user-bin order (layout order: the bins of one child are neighbours, so a
                // window's strip is a few words wide whatever ids the layout gave them); hits are sparse — a few dozen
                // of the row's bits — so every set bit is moved to its real id on the way into the tile
                for (uint32_t idx = lane; idx < rows * win.span; idx += 32)
                {
                    uint64_t const v = ld_hint(src + idx, stream);
                    if (v == 0)
                        continue;
                    uint32_t const r = (idx * inv) >> 16;
                    uint32_t const pos0 = (win.word_lo + (idx - r * win.span)) * 64;
                    unsigned * const row32 = reinterpret_cast<unsigned *>(tile + r * pitch);
                    for_each_bit(v, [&](int b) {
                        uint32_t const ub = __ldg(inv_perm + pos0 + b);
                        atomicOr(row32 + (ub >> 5), 1u << (ub & 31)); // two lanes may hold words of the same row
                    });
                }
            }
            __syncwarp(); // strips may overlap in user-bin words: one at a time
        }
        uint64_t * const dst = out + base * out_words;
        for (uint32_t idx = lane; idx < rows * out_words; idx += 32)
        {
            uint32_t const r = (idx * inv_w) >> 20;
            st_hint(dst + idx, tile[r * pitch + (idx - r * out_words)], stream);
        }
        __syncwarp();
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
    trx::scratch d_id_counts, d_id_blocks, d_ids, h_id_total; // sparse result (trx_hibf_membership_ids)
    int ibf_bits{1}; // bits needed for an IBF id (radix-sort range)
    bool sort_frontier{true};
    bool fused{true};        // root + second level in one kernel (TRX_HIBF_FUSED=off → frontier kernels only)
    uint64_t small_batch{2048}; // up to this many k-mers a pass takes the unsorted frontier route: 3 launches instead of the
                                // windowed schedule's ~12 (collector batches are a handful of k-mers); 0 when TRX_HIBF_FUSED is set
    int uniform_hash_funs{0}; // > 0 if every IBF uses the same number of hash functions
    uint64_t root_bin_words{1};
    trx::scratch root_words; // merged-bin hits of the root per k-mer
    std::vector<trx::hibf_window> windows; // [0] = the root's own user bins, then root technical-bin ranges whose children share L2
    trx::hibf_window * d_windows{nullptr};
    // user-bin ids in layout (DFS) order for the root / window kernels' strips, and the way back for the merge;
    // nullptr = the ids already are in that order (or the permuted route does not apply): strips hold real ids
    uint32_t * d_user_perm{nullptr};
    uint32_t * d_inv_perm{nullptr};
    uint64_t strip_words{0}; // Σ span
    bool window_attr_set{false}, merge_attr_set{false};
    unsigned window_ctas{8}; // resident CTAs per SM of the window kernel's persistent grid
    bool fast_shape{false}; // every IBF: one word per row and fewer than 2^32 rows
    trx::scratch strips;
    trx::hibf_dev dev() const
    {
        return trx::hibf_dev{d_words, d_ibfs, d_next, d_user, d_merged_mask, d_user_mask, static_cast<uint32_t>(out_words)};
    }
    trx::hibf_dev dev_strips() const // what the kernels that write strips see
    {
        trx::hibf_dev d = dev();
        if (d_user_perm)
            d.user_bin = d_user_perm;
        return d;
    }
};

using namespace trx;

static constexpr uint64_t kChunk = 1ull << 20;   // k-mers per frontier pass
static constexpr uint64_t kFanoutBudget = 32;    // frontier capacity = kFanoutBudget × chunk items
static constexpr int kMaxLevels = 30;

// ---- sparse result: result rows → (count per k-mer, ascending user-bin ids) ---------------------------------------------
constexpr int kIdsBlock = 256;

// counts[i] = popcount of row i; block_tot[b] = ids of block b's kIdsBlock k-mers.  A group of 16 lanes takes one row at a
// time, one word per lane and round: the rows are read in whole 128-byte lines (a thread per row read them 128 bytes apart).
constexpr int kIdsGroup = 16;
__global__ void __launch_bounds__(kIdsBlock) ids_count_kernel(uint64_t const * __restrict__ rows, uint64_t n, uint32_t W,
                                                               uint32_t * __restrict__ counts, unsigned long long * __restrict__ block_tot)
{
    __shared__ unsigned int s_total;
    if (threadIdx.x == 0)
        s_total = 0;
    __syncthreads();
    unsigned const lane = threadIdx.x % kIdsGroup, group = threadIdx.x / kIdsGroup;
    uint64_t const first = static_cast<uint64_t>(blockIdx.x) * kIdsBlock;
    unsigned int mine = 0;
    for (unsigned j = group; j < kIdsBlock; j += kIdsBlock / kIdsGroup)
    {
        uint64_t const i = first + j;
        unsigned int cnt = 0;
        if (i < n)
            for (uint32_t w = lane; w < W; w += kIdsGroup)
                cnt += __popcll(rows[i * W + w]);
        for (int o = kIdsGroup / 2; o > 0; o >>= 1)
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0 && i < n)
        {
            counts[i] = cnt;
            mine += cnt;
        }
    }
    if (mine)
        atomicAdd(&s_total, mine);
    __syncthreads();
    if (threadIdx.x == 0)
        block_tot[blockIdx.x] = s_total;
}

// in place: block_tot[b] → ids before block b; block_tot[blocks] = all ids of the piece.  One CTA (≤ 4096 blocks per piece).
__global__ void __launch_bounds__(1024) ids_scan_blocks_kernel(unsigned long long * __restrict__ block_tot, unsigned blocks)
{
    __shared__ unsigned long long s_part[1024];
    unsigned const per = (blocks + 1023) / 1024, lo = threadIdx.x * per, hi = min(lo + per, blocks);
    unsigned long long sum = 0;
    for (unsigned b = lo; b < hi; ++b)
        sum += block_tot[b];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        unsigned long long run = 0;
        for (int t = 0; t < 1024; ++t)
        {
            unsigned long long const v = s_part[t];
            s_part[t] = run;
            run += v;
        }
        block_tot[blocks] = run;
    }
    __syncthreads();
    unsigned long long run = s_part[threadIdx.x];
    for (unsigned b = lo; b < hi; ++b)
    {
        unsigned long long const v = block_tot[b];
        block_tot[b] = run;
        run += v;
    }
}

// ids of k-mer i go to out[block base + ids of the block's earlier k-mers ...): a CTA-wide exclusive scan of the counts
// (already there from ids_count_kernel), then a group of 16 lanes per row — a word per lane, the lanes' ids placed by a
// prefix sum inside the group, so that a row's ids and the rows of a block leave as one contiguous run.
template <class id_t>
__global__ void __launch_bounds__(kIdsBlock) ids_write_kernel(uint64_t const * __restrict__ rows, uint64_t n, uint32_t W,
                                                               uint32_t const * __restrict__ counts,
                                                               unsigned long long const * __restrict__ block_base, id_t * __restrict__ out)
{
    __shared__ unsigned int s_off[kIdsBlock];
    __shared__ unsigned int s_warp[kIdsBlock / 32];
    uint64_t const first = static_cast<uint64_t>(blockIdx.x) * kIdsBlock;
    {
        uint64_t const i = first + threadIdx.x;
        unsigned int const cnt = i < n ? counts[i] : 0;
        unsigned int incl = cnt;
        unsigned const wl = threadIdx.x & 31;
        for (int o = 1; o < 32; o <<= 1)
        {
            unsigned int const v = __shfl_up_sync(0xffffffffu, incl, o);
            if (wl >= static_cast<unsigned>(o))
                incl += v;
        }
        if (wl == 31)
            s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned int warp_base = 0;
        for (unsigned w = 0; w < (threadIdx.x >> 5); ++w)
            warp_base += s_warp[w];
        s_off[threadIdx.x] = warp_base + incl - cnt;
    }
    __syncthreads();
    unsigned const lane = threadIdx.x % kIdsGroup, group = threadIdx.x / kIdsGroup;
    unsigned const group_mask = 0xFFFFu << (threadIdx.x & 16); // the 16 lanes of this group inside the warp
    id_t * const base = out + block_base[blockIdx.x];
    for (unsigned j = group; j < kIdsBlock; j += kIdsBlock / kIdsGroup)
    {
        uint64_t const i = first + j;
        if (i >= n)
            break; // the whole group leaves together
        unsigned int run = s_off[j];
        for (uint32_t w0 = 0; w0 < W; w0 += kIdsGroup)
        {
            uint32_t const w = w0 + lane;
            uint64_t bits = w < W ? rows[i * W + w] : 0;
            unsigned int const cnt = __popcll(bits);
            unsigned int incl = cnt;
            for (int o = 1; o < kIdsGroup; o <<= 1)
            {
                unsigned int const v = __shfl_up_sync(group_mask, incl, o, kIdsGroup);
                if (lane >= static_cast<unsigned>(o))
                    incl += v;
            }
            id_t * dst = base + run + (incl - cnt);
            while (bits)
            {
                *dst++ = static_cast<id_t>(w * 64 + (__ffsll(static_cast<long long>(bits)) - 1));
                bits &= bits - 1;
            }
            run += __shfl_sync(group_mask, incl, kIdsGroup - 1, kIdsGroup);
        }
    }
}

template <int H, bool FAST>
static int launch_fused(trx_hibf * h, uint64_t const * d_kmers, uint64_t n, uint64_t * d_out, uint64_t * frontier,
                        unsigned long long * ctr, uint64_t capacity)
{
    trx_ctx * c = h->ctx;
    uint64_t const bw0 = h->root_bin_words;
    TRX_TRY(h->root_words.reserve(n * bw0 * 8));
    TRX_TRY(h->strips.reserve(n * h->strip_words * 8 + 8));
    auto * rw = static_cast<uint64_t *>(h->root_words.ptr);
    auto * strips = static_cast<uint64_t *>(h->strips.ptr);
    hibf_window const & root = h->windows[0];
    if (root.span)
        TRX_CUDA(cudaMemsetAsync(strips, 0, n * root.span * 8, c->stream));
    hibf_root_kernel<H, FAST><<<grid_for(n, 256, c->sm_count, 8), 256, 0, c->stream>>>(h->dev_strips(), d_kmers, n, rw, root, strips, ctr);
    ++c->launches;
    if (!h->window_attr_set)
    {
        TRX_CUDA(cudaFuncSetAttribute(hibf_window_kernel<H, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(kWinThreads * (kMaxSpan + 1) * 8)));
        h->window_attr_set = true;
    }
    unsigned const grid = grid_for((n + 31) / 32, kWinThreads / 32, c->sm_count, h->window_ctas);
    for (size_t g = 1; g < h->windows.size(); ++g)
    {
        hibf_window const & win = h->windows[g];
        size_t const smem = static_cast<size_t>(kWinThreads) * (win.span + 1) * 8;
        hibf_window_kernel<H, FAST><<<grid, kWinThreads, smem, c->stream>>>(h->dev_strips(), d_kmers, n, rw, win, strips + win.strip_off * n,
                                                                     frontier, ctr, capacity);
        ++c->launches;
    }
    if (h->out_words <= kMergeTileWords)
    {
        size_t const tile = static_cast<size_t>(kMergeWarps) * 32 * (h->out_words + 1) * 8;
        if (!h->merge_attr_set)
        {
            TRX_CUDA(cudaFuncSetAttribute(hibf_merge_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(kMergeWarps * 32 * (kMergeTileWords + 1) * 8)));
            h->merge_attr_set = true;
        }
        hibf_merge_tile_kernel<<<grid_for((n + 31) / 32, kMergeWarps, c->sm_count, 12), kMergeWarps * 32, tile, c->stream>>>(
            h->d_windows, static_cast<uint32_t>(h->windows.size()), strips, n, static_cast<uint32_t>(h->out_words), d_out,
            h->d_inv_perm);
    }
    else
    {
        uint64_t const rows_per_block = 256 / (h->out_words < 256 ? h->out_words : 256);
        hibf_merge_kernel<<<grid_for(n, static_cast<unsigned>(rows_per_block), c->sm_count, 8), 256, 0, c->stream>>>(
            h->d_windows, static_cast<uint32_t>(h->windows.size()), strips, n, static_cast<uint32_t>(h->out_words), d_out);
    }
    ++c->launches;
    return TRX_OK;
}

namespace
{
// k-mers of a host batch reach the device piece by piece on the context's inbound stream; events[k] fires when piece k
// is there.  Everything is enqueued at once: the copies run back to back at PCIe speed, compute of piece k waits for
// events[k] only.
int stage_kmers_in_pieces(trx_ctx * c, uint64_t * d_kmers, const uint64_t * kmers, uint64_t n, uint64_t piece, bool & staged)
{
    staged = n > piece;
    if (!staged) // one piece (a query's handful of k-mers): in order on the compute stream, no second stream to wait for
    {
        TRX_CUDA(cudaMemcpyAsync(d_kmers, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
        return TRX_OK;
    }
    TRX_CUDA(cudaStreamSynchronize(c->stream)); // d_kmers may still be read by work enqueued before this call
    if (!c->in_stream)
        TRX_CUDA(cudaStreamCreateWithFlags(&c->in_stream, cudaStreamNonBlocking));
    size_t k = 0;
    for (uint64_t done = 0; done < n; done += piece, ++k)
    {
        uint64_t const m = n - done < piece ? n - done : piece;
        TRX_CUDA(cudaMemcpyAsync(d_kmers + done, kmers + done, m * 8, cudaMemcpyHostToDevice, c->in_stream));
        if (k >= c->in_events.size())
        {
            cudaEvent_t e;
            TRX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            c->in_events.push_back(e);
        }
        TRX_CUDA(cudaEventRecord(c->in_events[k], c->in_stream));
    }
    return TRX_OK;
}
} // namespace

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
    std::vector<uint32_t> user_strip, inv_perm; // strip-order ids per technical bin / the way back (see below)
    bool use_perm = false;
    {
        char const * e = getenv("TRX_HIBF_SORT"); // TRX_HIBF_SORT=off: A/B evidence for the frontier sort
        h->sort_frontier = !(e && strcmp(e, "off") == 0);
        e = getenv("TRX_HIBF_FUSED"); // TRX_HIBF_FUSED=off: frontier kernels for every level (A/B evidence)
        h->fused = !(e && strcmp(e, "off") == 0) && depth >= 2;
        if (e)
            h->small_batch = 0; // an explicit route request is honoured for every batch size
        if ((e = getenv("TRX_HIBF_CTAS")) && atoi(e) > 0)
            h->window_ctas = static_cast<unsigned>(atoi(e));
        h->fast_shape = true;
        for (uint64_t i = 0; i < n; ++i)
            if (ibfs[i].bin_words != 1 || ibfs[i].bin_size >= (1ull << 32))
                h->fast_shape = false;
        if (desc->n_user_bins >= 65536)
            h->fast_shape = false; // 16-bit user-bin ids in the window kernel's shared-memory table
        h->uniform_hash_funs = static_cast<int>(desc->hash_funs[0]);
        for (uint64_t i = 1; i < n; ++i)
            if (desc->hash_funs[i] != desc->hash_funs[0])
                h->uniform_hash_funs = 0;
        // Strip order of the user bins = the order in which a walk of the layout meets them (an IBF's own user bins, then
        // its children's, recursively).  seqan's layout deals user bins to merged bins by size, so the ids inside one
        // child are scattered over the whole range: in id order every window would report into ALL result words and the
        // strips would be as large as the result (measured on the seqan-built C2 index: 16 words per k-mer and window,
        // 10 GB of strip traffic per 4 Mi k-mers, 1.17 instead of 2.4 G probes/s); in walk order a child's bins are
        // neighbours and a window reports into 2-3 words.  The merge kernel moves the (sparse) hit bits back to real ids.
        std::vector<uint32_t> perm(desc->n_user_bins, UINT32_MAX);
        inv_perm.assign(h->out_words * 64, 0);
        {
            uint32_t pos = 0;
            std::vector<char> walked(n, 0);
            std::vector<uint64_t> stack{0};
            walked[0] = 1;
            while (!stack.empty())
            {
                uint64_t const i = stack.back();
                stack.pop_back();
                uint64_t const tbs = static_cast<uint64_t>(ibfs[i].bin_words) * 64;
                for (uint64_t t = 0; t < tbs; ++t)
                {
                    uint32_t const ub = user[ibfs[i].table_off + t];
                    if (ub < kDeleted && ub != kMerged && perm[ub] == UINT32_MAX)
                        perm[ub] = pos++;
                }
                for (uint64_t t = tbs; t-- > 0;) // pushed in reverse: children are walked in technical-bin order
                    if (user[ibfs[i].table_off + t] == kMerged && !walked[next[ibfs[i].table_off + t]])
                    {
                        walked[next[ibfs[i].table_off + t]] = 1;
                        stack.push_back(next[ibfs[i].table_off + t]);
                    }
            }
            for (uint64_t ub = 0; ub < desc->n_user_bins; ++ub) // bins the layout never placed
                if (perm[ub] == UINT32_MAX)
                    perm[ub] = pos++;
        }
        bool identity = true;
        for (uint64_t ub = 0; ub < desc->n_user_bins; ++ub)
        {
            identity = identity && perm[ub] == ub;
            inv_perm[perm[ub]] = static_cast<uint32_t>(ub);
        }
        char const * pe = getenv("TRX_HIBF_PERMUTE"); // TRX_HIBF_PERMUTE=off: strips in id order (A/B evidence)
        use_perm = !identity && h->fused && h->out_words <= kMergeTileWords && !(pe && strcmp(pe, "off") == 0);
        user_strip = user;
        if (use_perm)
            for (auto & ub : user_strip)
                if (ub < kDeleted && ub != kMerged)
                    ub = perm[ub];
        // windows: consecutive root technical bins whose children together stay below the L2 budget; for each the run of
        // user-bin words its children report into (a window that would span more than kMaxSpan words is cut earlier)
        h->root_bin_words = ibfs[0].bin_words;
        uint64_t budget = 44ull << 20;
        if ((e = getenv("TRX_HIBF_WINDOW_MB")) && atoll(e) > 0)
            budget = static_cast<uint64_t>(atoll(e)) << 20;
        auto word_range = [&](uint64_t ibf, uint32_t & lo, uint32_t & hi) {
            for (uint64_t t = 0; t < static_cast<uint64_t>(ibfs[ibf].bin_words) * 64; ++t)
            {
                uint64_t idx = ibfs[ibf].table_off + t;
                if (umask[idx / 64] >> (idx % 64) & 1)
                {
                    lo = std::min(lo, user_strip[idx] >> 6);
                    hi = std::max(hi, user_strip[idx] >> 6);
                }
            }
        };
        {
            uint32_t lo = UINT32_MAX, hi = 0;
            word_range(0, lo, hi);
            hibf_window root{0, 0, lo == UINT32_MAX ? 0 : lo, lo == UINT32_MAX ? 0 : hi - lo + 1, 0};
            h->windows.push_back(root);
            if (root.span > kMaxSpan)
                h->fused = false;
        }
        uint32_t const root_bins = ibfs[0].bin_words * 64;
        uint32_t begin = 0, lo = UINT32_MAX, hi = 0;
        uint64_t bytes = 0;
        auto close = [&](uint32_t end) {
            if (lo != UINT32_MAX || bytes > 0)
                h->windows.push_back(hibf_window{begin, end, lo == UINT32_MAX ? 0 : lo, lo == UINT32_MAX ? 0 : hi - lo + 1, 0});
            begin = end;
            lo = UINT32_MAX;
            hi = 0;
            bytes = 0;
        };
        for (uint32_t b = 0; b < root_bins && h->fused; ++b)
        {
            uint64_t const idx = ibfs[0].table_off + b;
            if (!(mmask[idx / 64] >> (idx % 64) & 1))
                continue;
            uint64_t const child = next[idx];
            uint64_t const child_bytes = ibfs[child].bin_size * ibfs[child].bin_words * 8;
            uint32_t clo = UINT32_MAX, chi = 0;
            word_range(child, clo, chi);
            if (clo != UINT32_MAX && chi - clo + 1 > kMaxSpan)
            {
                h->fused = false; // one child alone reports into too wide a run: frontier kernels
                break;
            }
            uint32_t const nlo = std::min(lo, clo), nhi = std::max(hi, chi);
            bool const too_wide = nlo != UINT32_MAX && nhi - nlo + 1 > kMaxSpan;
            if (b > begin && (bytes + child_bytes > budget || too_wide || b - begin >= kMaxWindowBins || (b >> 6) != (begin >> 6)))
                close(b);
            lo = std::min(lo, clo);
            hi = std::max(hi, chi);
            bytes += child_bytes;
        }
        close(root_bins);
        h->strip_words = 0;
        for (auto & win : h->windows)
        {
            win.strip_off = h->strip_words;
            h->strip_words += win.span;
        }
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
    if (use_perm && h->fused)
    {
        if ((e = cudaMalloc(&h->d_user_perm, table_total * 4)) != cudaSuccess) return fail(e);
        if ((e = cudaMalloc(&h->d_inv_perm, inv_perm.size() * 4)) != cudaSuccess) return fail(e);
        if ((e = cudaMemcpyAsync(h->d_user_perm, user_strip.data(), table_total * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
        if ((e = cudaMemcpyAsync(h->d_inv_perm, inv_perm.data(), inv_perm.size() * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    }
    if ((e = cudaMalloc(&h->d_windows, h->windows.size() * sizeof(hibf_window))) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(h->d_windows, h->windows.data(), h->windows.size() * sizeof(hibf_window), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
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
    h->root_words.release();
    h->strips.release();
    cudaFree(h->d_windows);
    cudaFree(h->d_user_perm);
    cudaFree(h->d_inv_perm);
    delete h;
}

uint64_t trx_hibf_user_bin_words(const trx_hibf_t * h)
{
    return h ? h->out_words : 0;
}

// one chunk of k-mers through all levels; returns TRX_ECAPACITY on frontier overflow
// deferred_ctr (pinned host memory, kMaxLevels + 3 counters): for a small batch the level counters are copied there and
// NOT waited for — the caller checks them after its own synchronisation ([0] != 0: frontier overflow)
static int run_chunk(trx_hibf * h, uint64_t const * d_kmers, uint64_t n, uint64_t * d_out, uint64_t * visited,
                     unsigned long long * deferred_ctr = nullptr)
{
    trx_ctx * c = h->ctx;
    bool const small = n <= h->small_batch;
    bool const fused = h->fused && !small, sort_frontier = h->sort_frontier && !small;
    bool const need_frontier = !fused || h->depth > 2;
    uint64_t const capacity = need_frontier ? kFanoutBudget * (n < 4096 ? 4096 : n) : 0;
    if (need_frontier)
    {
        TRX_TRY(h->frontier[0].reserve(capacity * 8));
        TRX_TRY(h->frontier[1].reserve(capacity * 8));
    }
    TRX_TRY(h->counters.reserve((kMaxLevels + 3) * 8));
    auto * ctr = static_cast<unsigned long long *>(h->counters.ptr);
    TRX_CUDA(cudaMemsetAsync(ctr, 0, (kMaxLevels + 3) * 8, c->stream));
    int first_level = 1; // first level that still has to run through the frontier kernels
    if (fused)
    {
        auto * fr = static_cast<uint64_t *>(h->frontier[0].ptr);
        int const key = h->uniform_hash_funs * 2 + (h->fast_shape ? 1 : 0);
        switch (key)
        {
        case 4: TRX_TRY((launch_fused<2, false>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        case 5: TRX_TRY((launch_fused<2, true>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        case 6: TRX_TRY((launch_fused<3, false>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        case 7: TRX_TRY((launch_fused<3, true>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        case 8: TRX_TRY((launch_fused<4, false>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        case 9: TRX_TRY((launch_fused<4, true>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        default: TRX_TRY((launch_fused<0, false>(h, d_kmers, n, d_out, fr, ctr, capacity))); break;
        }
        first_level = 2;
    }
    else
    {
        TRX_CUDA(cudaMemsetAsync(d_out, 0, n * h->out_words * 8, c->stream));
        unsigned const grid = grid_for(n, 256, c->sm_count, 8);
        hibf_level_kernel<true><<<grid, 256, 0, c->stream>>>(h->dev(), d_kmers, n, nullptr,
                                                             static_cast<uint64_t *>(h->frontier[0].ptr), ctr, 0, capacity, d_out);
        ++c->launches;
    }
    unsigned const full_grid = static_cast<unsigned>(c->sm_count) * 8;
    // ping-pong: `cur` holds the items of the level about to run, `other` is free
    int cur = 0;
    for (int level = first_level; level < h->depth; ++level)
    {
        if (sort_frontier)
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
    if (deferred_ctr && small && !sort_frontier)
    {
        TRX_CUDA(cudaMemcpyAsync(deferred_ctr, ctr, (kMaxLevels + 3) * 8, cudaMemcpyDeviceToHost, c->stream));
        return TRX_OK;
    }
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
    // two-level HIBFs on the fused kernel need no frontier: the whole batch is one persistent launch
    uint64_t chunk = (h->fused && h->depth <= 2) ? (n > kChunk ? n : kChunk) : kChunk;
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

// seqan's own result form (membership_for returns user-bin ids): counts[i] ids per k-mer, ids concatenated in k-mer order,
// ascending inside a k-mer.  The rows are computed like trx_hibf_membership's and compacted on the device before they
// cross PCIe: at C2 a k-mer hits ~17 of 1024 user bins, so 2-byte ids are 38 B per k-mer instead of a 128-byte row.
int trx_hibf_membership_ids(trx_hibf_t * h, const uint64_t * kmers, uint64_t n, uint32_t * counts, void * ids, uint32_t id_bytes,
                            uint64_t cap, uint64_t * n_ids, uint64_t * stats)
{
    if (!h || !n_ids || (n && (!kmers || !counts)) || (cap && !ids))
        return set_error(TRX_EINVAL, "NULL argument");
    if (id_bytes != 2 && id_bytes != 4)
        return set_error(TRX_EINVAL, "id_bytes must be 2 or 4");
    if (id_bytes == 2 && h->n_user_bins > 65536)
        return set_error(TRX_ERANGE, "%llu user bins do not fit 2-byte ids", (unsigned long long)h->n_user_bins);
    *n_ids = 0;
    if (stats)
        stats[0] = 0, stats[1] = static_cast<uint64_t>(h->depth);
    if (n == 0)
        return TRX_OK;
    trx_ctx * c = h->ctx;
    TRX_TRY(c->use());
    static uint64_t const piece = [] {
        char const * e = getenv("TRX_IDS_PIECE"); // k-mers per piece of the pipeline (A/B)
        uint64_t const v = e ? strtoull(e, nullptr, 10) : 0;
        return v >= 1024 ? v : (1ull << 20);
    }();
    uint64_t const W = h->out_words;
    TRX_TRY(h->d_kmers.reserve(n * 8));
    TRX_TRY(h->d_out.reserve(piece * W * 8 * 2)); // two pieces of rows: one being compacted while the next is computed
    TRX_TRY(h->d_id_counts.reserve(n * 4));
    TRX_TRY(h->d_id_blocks.reserve(2 * ((piece + kIdsBlock - 1) / kIdsBlock + 2) * 8));
    TRX_TRY(h->d_ids.reserve((cap ? cap : 1) * id_bytes));
    h->h_id_total.pinned_host = true;
    TRX_TRY(h->h_id_total.reserve(((n + piece - 1) / piece + 1) * 8));
    auto * d_kmers = static_cast<uint64_t *>(h->d_kmers.ptr);
    auto * d_rows = static_cast<uint64_t *>(h->d_out.ptr);
    auto * d_counts = static_cast<uint32_t *>(h->d_id_counts.ptr);
    auto * d_ids = static_cast<uint8_t *>(h->d_ids.ptr);
    auto * h_totals = static_cast<unsigned long long *>(h->h_id_total.ptr);
    bool staged = false;
    TRX_TRY(stage_kmers_in_pieces(c, d_kmers, kmers, n, piece, staged));
    if (!c->copy_stream)
        TRX_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    uint64_t visited = 0, written = 0;
    size_t k = 0;
    for (uint64_t done = 0; done < n; done += piece, ++k)
    {
        uint64_t const m = n - done < piece ? n - done : piece;
        uint64_t * const rows = d_rows + (k & 1) * piece * W;
        uint64_t st[2] = {0, 0};
        if (staged)
            TRX_CUDA(cudaStreamWaitEvent(c->stream, c->in_events[k], 0));
        TRX_TRY(trx_hibf_membership_dev(h, d_kmers + done, m, rows, st));
        visited += st[0];
        unsigned const blocks = static_cast<unsigned>((m + kIdsBlock - 1) / kIdsBlock);
        auto * d_block_tot = static_cast<unsigned long long *>(h->d_id_blocks.ptr) + (k & 1) * (blocks + 2);
        ids_count_kernel<<<blocks, kIdsBlock, 0, c->stream>>>(rows, m, static_cast<uint32_t>(W), d_counts + done, d_block_tot);
        ids_scan_blocks_kernel<<<1, 1024, 0, c->stream>>>(d_block_tot, blocks);
        c->launches += 2;
        TRX_CUDA(cudaMemcpyAsync(h_totals + k, d_block_tot + blocks, 8, cudaMemcpyDeviceToHost, c->stream));
        TRX_CUDA(cudaStreamSynchronize(c->stream)); // how many ids this piece has (and whether they still fit)
        uint64_t const piece_ids = h_totals[k];
        if (written + piece_ids > cap)
        {
            *n_ids = written + piece_ids;
            cudaStreamSynchronize(c->copy_stream);
            return set_error(TRX_ECAPACITY, "trx_hibf_membership_ids: at least %llu ids, room for %llu",
                             (unsigned long long)(written + piece_ids), (unsigned long long)cap);
        }
        if (piece_ids)
        {
            if (id_bytes == 2)
                ids_write_kernel<uint16_t><<<blocks, kIdsBlock, 0, c->stream>>>(rows, m, static_cast<uint32_t>(W), d_counts + done, d_block_tot,
                                                                               reinterpret_cast<uint16_t *>(d_ids) + written);
            else
                ids_write_kernel<uint32_t><<<blocks, kIdsBlock, 0, c->stream>>>(rows, m, static_cast<uint32_t>(W), d_counts + done, d_block_tot,
                                                                               reinterpret_cast<uint32_t *>(d_ids) + written);
            ++c->launches;
        }
        TRX_CUDA(cudaGetLastError());
        if (k >= c->copy_events.size())
        {
            cudaEvent_t e;
            TRX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            c->copy_events.push_back(e);
        }
        TRX_CUDA(cudaEventRecord(c->copy_events[k], c->stream));
        TRX_CUDA(cudaStreamWaitEvent(c->copy_stream, c->copy_events[k], 0));
        TRX_CUDA(cudaMemcpyAsync(counts + done, d_counts + done, m * 4, cudaMemcpyDeviceToHost, c->copy_stream));
        if (piece_ids)
            TRX_CUDA(cudaMemcpyAsync(static_cast<uint8_t *>(ids) + written * id_bytes, d_ids + written * id_bytes, piece_ids * id_bytes,
                                     cudaMemcpyDeviceToHost, c->copy_stream));
        written += piece_ids;
        // the rows buffer (k & 1) is reused two pieces later: its compaction must have finished by then — it has, the
        // stream was synchronised above before the next piece's kernels were enqueued
    }
    TRX_CUDA(cudaStreamSynchronize(c->copy_stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    *n_ids = written;
    if (stats)
        stats[0] = visited;
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
    auto * d_kmers = static_cast<uint64_t *>(h->d_kmers.ptr);
    auto * d_out = static_cast<uint64_t *>(h->d_out.ptr);
    if (n <= 64 && n <= h->small_batch && n * h->out_words * 8 <= (64u << 10))
    {
        // a query's handful of k-mers (the collector's one-batch walk): everything through ONE pinned block and one
        // synchronisation — k-mers in, rows and level counters out — instead of three staged copies and two waits
        size_t const row_bytes = n * h->out_words * 8;
        size_t const off_rows = (n * 8 + 255) & ~size_t{255}, off_ctr = (off_rows + row_bytes + 255) & ~size_t{255};
        c->h_pin[0].pinned_host = true;
        TRX_TRY(c->h_pin[0].reserve(off_ctr + (kMaxLevels + 3) * 8));
        auto * pin = static_cast<uint8_t *>(c->h_pin[0].ptr);
        auto * pin_ctr = reinterpret_cast<unsigned long long *>(pin + off_ctr);
        memcpy(pin, kmers, n * 8);
        TRX_CUDA(cudaMemcpyAsync(d_kmers, pin, n * 8, cudaMemcpyHostToDevice, c->stream));
        uint64_t visited_small = 0;
        TRX_TRY(run_chunk(h, d_kmers, n, d_out, &visited_small, pin_ctr));
        TRX_CUDA(cudaMemcpyAsync(pin + off_rows, d_out, row_bytes, cudaMemcpyDeviceToHost, c->stream));
        TRX_CUDA(cudaStreamSynchronize(c->stream));
        if (pin_ctr[0] == 0) // no frontier overflow (else: the general path below shrinks its passes)
        {
            memcpy(out_words, pin + off_rows, row_bytes);
            if (stats)
            {
                uint64_t v = n;
                for (int level = 1; level < h->depth; ++level)
                    v += pin_ctr[1 + level];
                stats[0] = v;
                stats[1] = static_cast<uint64_t>(h->depth);
            }
            return TRX_OK;
        }
    }
    // The result (out_words x 8 B per k-mer, 128 B at C2) dwarfs the input: it leaves in pieces on a second stream
    // while the next piece is being computed, so the call costs about max(compute, PCIe) instead of their sum; the k-mers
    // arrive piece by piece on a third.
    uint64_t const piece = 1ull << 20;
    bool staged = false;
    TRX_TRY(stage_kmers_in_pieces(c, d_kmers, kmers, n, piece, staged));
    if (!c->copy_stream)
        TRX_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    uint64_t visited = 0;
    size_t k = 0;
    for (uint64_t done = 0; done < n; done += piece, ++k)
    {
        uint64_t const m = n - done < piece ? n - done : piece;
        uint64_t st[2] = {0, 0};
        if (staged)
            TRX_CUDA(cudaStreamWaitEvent(c->stream, c->in_events[k], 0));
        TRX_TRY(trx_hibf_membership_dev(h, d_kmers + done, m, d_out + done * h->out_words, st));
        visited += st[0];
        if (k >= c->copy_events.size())
        {
            cudaEvent_t e;
            TRX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            c->copy_events.push_back(e);
        }
        TRX_CUDA(cudaEventRecord(c->copy_events[k], c->stream));
        TRX_CUDA(cudaStreamWaitEvent(c->copy_stream, c->copy_events[k], 0));
        TRX_CUDA(cudaMemcpyAsync(out_words + done * h->out_words, d_out + done * h->out_words, m * h->out_words * 8,
                                 cudaMemcpyDeviceToHost, c->copy_stream));
    }
    TRX_CUDA(cudaStreamSynchronize(c->copy_stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    if (stats)
    {
        stats[0] = visited;
        stats[1] = static_cast<uint64_t>(h->depth);
    }
    return TRX_OK;
}

} // extern "C"
