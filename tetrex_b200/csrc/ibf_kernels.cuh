// ibf_kernels.cuh — K1 (hash + emplace) and K2 (bulk_contains / path AND) device code.
//
// Reference arithmetic: lib/hibf/include/hibf/interleaved_bloom_filter.hpp:164-187 (hash_and_fit),
// lib/hibf/src/interleaved_bloom_filter.cpp:119-140 (emplace), :203-290 (bulk_contains),
// include/nucleotide_decomposer.h:81-121, include/peptide_decomposer.h:261-299 (k-mer packing).
#pragma once

#include "common.cuh"

namespace trx
{

struct ibf_geom
{
    uint64_t * words;       // this shard's matrix, row r at r*shard_words
    uint64_t bin_size;      // rows
    uint64_t shard_words;   // words per row held here
    uint64_t word_begin;    // first global word column held here
    uint32_t hash_shift;
    uint32_t hash_funs;
    uint64_t row_lo, row_hi; // K1 only: rows this pass may touch (see trx_emplace_records_dev); [0, bin_size) otherwise
    // K1 only, optional: a dense stand-in for the word columns [slab_begin, slab_begin + slab_words) of every row
    // (row r at r*slab_words).  Bits of bins in those columns are set there instead of in `words`; the caller ORs the
    // slab into the matrix afterwards (slab_merge_kernel).  nullptr = everything goes to `words`.
    uint64_t * slab;
    uint64_t slab_words, slab_begin;
    // N3, optional: HyperLogLog sketching instead of emplacing (trx_hll_sketch_records): one sketch of 2^hll_bits 32-bit
    // registers per bin; every emitted k-mer updates a register of its bin's sketch.  nullptr = emplace.
    unsigned int * hll;
    uint32_t hll_bits;
};

// ------------------------------------------------------------------------------------------- K2
// One group of G lanes per k-mer; a lane owns V consecutive words (V=2 → 16-byte LDG.128) of every
// row, striding by G*V.  All H row loads of a lane are issued before the first use so that every
// thread has H independent random sector requests in flight; with 2048 resident threads per SM that
// is the memory-level parallelism the random-probe roofline needs.  Nothing is staged in shared
// memory: a row word is used exactly once, so registers are the right home for it.
//
// MODE 0: out[i*shard_words + w] = AND_h row_h[w]                       (bulk_contains)
// MODE 1: slot[slots[i]][w] &= AND_h row_h[w]; alive[i] = any(slot)     (update_path + none())
// MODE 2: slot[slots[i]][w] &= hits[hit_index[i]][w]; alive[i] = any    (HIBF route)
template <int H, int V, int G, int MODE>
__global__ void __launch_bounds__(256)
    probe_kernel(ibf_geom g, uint64_t const * __restrict__ kmers, uint64_t n, uint64_t * __restrict__ out,
                 uint32_t const * __restrict__ slots, uint8_t * __restrict__ alive,
                 uint64_t const * __restrict__ hits, uint32_t const * __restrict__ hit_index)
{
    using vec_t = typename std::conditional<V == 2, ulonglong2, uint64_t>::type;
    constexpr int groups_per_block = 256 / G;
    unsigned const lane_in_group = threadIdx.x % G;
    uint64_t const group0 = static_cast<uint64_t>(blockIdx.x) * groups_per_block + threadIdx.x / G;
    uint64_t const stride = static_cast<uint64_t>(gridDim.x) * groups_per_block;
    uint64_t const W = g.shard_words;

    for (uint64_t i = group0; i < n; i += stride)
    {
        uint64_t rowbase[H > 0 ? H : 1];
        if constexpr (MODE != 2)
        {
            uint64_t const v = __ldg(kmers + i);
#pragma unroll
            for (int h = 0; h < H; ++h)
                rowbase[h] = hash_row(v, h, g.hash_shift, g.bin_size) * W;
        }
        else
        {
            rowbase[0] = static_cast<uint64_t>(__ldg(hit_index + i)) * W;
        }
        uint64_t * dst;
        if constexpr (MODE == 0)
            dst = out + i * W;
        else
            dst = out + static_cast<uint64_t>(__ldg(slots + i)) * W;

        uint64_t any = 0;
        for (uint64_t w = static_cast<uint64_t>(lane_in_group) * V; w < W; w += static_cast<uint64_t>(G) * V)
        {
            vec_t r[H > 0 ? H : 1];
            if constexpr (MODE != 2)
            {
#pragma unroll
                for (int h = 0; h < H; ++h)
                    r[h] = *reinterpret_cast<vec_t const *>(g.words + rowbase[h] + w);
            }
            else
            {
                r[0] = *reinterpret_cast<vec_t const *>(hits + rowbase[0] + w);
            }
            vec_t acc = r[0];
            if constexpr (MODE != 2)
            {
#pragma unroll
                for (int h = 1; h < H; ++h)
                {
                    if constexpr (V == 2)
                    {
                        acc.x &= r[h].x;
                        acc.y &= r[h].y;
                    }
                    else
                        acc &= r[h];
                }
            }
            if constexpr (MODE != 0)
            {
                vec_t cur = *reinterpret_cast<vec_t *>(dst + w);
                if constexpr (V == 2)
                {
                    acc.x &= cur.x;
                    acc.y &= cur.y;
                    any |= acc.x | acc.y;
                }
                else
                {
                    acc &= cur;
                    any |= acc;
                }
            }
            *reinterpret_cast<vec_t *>(dst + w) = acc;
        }
        if constexpr (MODE != 0)
        {
            // group-wide none(): OR-reduce over the G lanes of this k-mer with shuffles
            unsigned flag = any != 0;
#pragma unroll
            for (int off = G / 2; off > 0; off >>= 1)
                flag |= __shfl_xor_sync(0xffffffffu, flag, off, G);
            if (lane_in_group == 0)
                alive[i] = static_cast<uint8_t>(flag);
        }
    }
}

// slot[dst[i]] |= slot[src[i]] — push/absorb/split and `path_matrix |= path` of the collector.
template <int G>
__global__ void __launch_bounds__(256)
    paths_or_kernel(uint64_t * __restrict__ words, uint64_t W, uint32_t const * __restrict__ dst,
                    uint32_t const * __restrict__ src, uint64_t n)
{
    constexpr int groups_per_block = 256 / G;
    unsigned const lane_in_group = threadIdx.x % G;
    uint64_t const group0 = static_cast<uint64_t>(blockIdx.x) * groups_per_block + threadIdx.x / G;
    uint64_t const stride = static_cast<uint64_t>(gridDim.x) * groups_per_block;
    for (uint64_t i = group0; i < n; i += stride)
    {
        uint64_t const * s = words + static_cast<uint64_t>(src[i]) * W;
        unsigned long long * d = reinterpret_cast<unsigned long long *>(words + static_cast<uint64_t>(dst[i]) * W);
        for (uint64_t w = lane_in_group; w < W; w += G)
        {
            uint64_t v = s[w];
            if (v)
                atomicOr(d + w, static_cast<unsigned long long>(v));
        }
    }
}

__global__ void __launch_bounds__(256)
    paths_fill_kernel(uint64_t * __restrict__ words, uint64_t W, uint64_t first, uint64_t n, uint64_t bins)
{
    uint64_t const total = n * W;
    for (uint64_t t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t const w = t % W;
        uint64_t v;
        if (bins >= (w + 1) * 64)
            v = ~0ULL;
        else if (bins <= w * 64)
            v = 0;
        else
            v = (1ULL << (bins - w * 64)) - 1;
        words[first * W + t] = v;
    }
}

// ------------------------------------------------------------------------------------------- K1a
// emplace for explicit (value, bin) pairs: H hashes, H 64-bit atomic ORs (RED.OR, no return value).
__global__ void __launch_bounds__(256)
    emplace_values_kernel(ibf_geom g, uint64_t const * __restrict__ values, uint32_t const * __restrict__ bins,
                          uint64_t n)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t const v = __ldg(values + i);
        uint32_t const b = __ldg(bins + i);
        uint64_t const col = (b >> 6);
        if (col < g.word_begin || col >= g.word_begin + g.shard_words)
            continue; // another shard owns this 64-bin block
        unsigned long long const bit = 1ULL << (b & 63);
        for (uint32_t h = 0; h < g.hash_funs; ++h)
        {
            uint64_t const row = hash_row(v, static_cast<int>(h), g.hash_shift, g.bin_size);
            atomicOr(reinterpret_cast<unsigned long long *>(g.words + row * g.shard_words + (col - g.word_begin)), bit);
        }
    }
}

// ------------------------------------------------------------------------------------------- K1b
__device__ __forceinline__ uint64_t dna_code(uint8_t c)
{
    return (c >> 1) & 3u; // nucleotide_decomposer.h:83-89
}

// NucleotideDecomposer::revComplement (nucleotide_decomposer.h:44-79): complement = code^2, reverse the
// 2-bit groups.  Quirk kept for bit-exactness: the reference shifts a SIGNED 64-bit vector lane right by
// 4 (`kmer2 >>= 4`), so when bit 63 is set (k = 32, first base T/G) the top byte's shuffle index has bit 7
// set, pshufb yields 0 and the two leading bases drop out: the lowest nibble of the result is 0.
__device__ __forceinline__ uint64_t dna_revcomp(uint64_t fwd, uint32_t k)
{
    uint64_t x = fwd ^ 0xAAAAAAAAAAAAAAAAULL;
    x = __brevll(x);                                                            // reverses single bits …
    x = ((x & 0x5555555555555555ULL) << 1) | ((x >> 1) & 0x5555555555555555ULL); // … so swap each pair back
    x >>= (64 - 2 * k);
    if (fwd >> 63)
        x &= ~0xFULL;
    return x;
}

// rollover_nuc_hash (nucleotide_decomposer.h:92-101): `(fb^0b10) << left_shift_` is evaluated on an int.
// For k >= 16 (left_shift_ >= 30) that overflows; as compiled for x86 the count is taken mod 32 and the
// int is sign-extended when ORed into the 64-bit reverse store.  The reference's INDEX side therefore
// does not hold true reverse complements for k >= 16; the bits it sets are what we must reproduce.
__device__ __forceinline__ uint64_t dna_comp_bits(uint64_t f, uint32_t left_shift)
{
    return static_cast<uint64_t>(static_cast<int64_t>(static_cast<int32_t>(static_cast<uint32_t>(f ^ 2) << (left_shift & 31))));
}

struct dna_state
{
    uint64_t fwd, rev;
    __device__ __forceinline__ void roll(uint8_t c, uint64_t mask, uint32_t left_shift)
    {
        uint64_t const f = dna_code(c);
        fwd = ((fwd << 2) & mask) | f;
        rev = ((rev >> 2) & mask) | dna_comp_bits(f, left_shift);
    }
    __device__ __forceinline__ uint64_t canonical() const
    {
        return fwd <= rev ? fwd : rev;
    }
};

// how many rolled symbols determine the (fwd, rev) stores completely: every contribution to `rev` enters at
// bit <= 63 and moves down 2 bits per step, so 33 steps flush anything older; for k <= 15 the stores are the
// exact k-mer and its reverse complement after k steps.
__device__ __forceinline__ uint32_t dna_warmup(uint32_t k)
{
    return k >= 16 ? 33u : k;
}

// State of decompose_record (nucleotide_decomposer.h:103-111) BEFORE rolling rec[t] in, i.e. after the
// initial k-mer and the symbols rec[0..t): replayed from at most dna_warmup(k) symbols back.
__device__ __forceinline__ dna_state dna_state_before(uint8_t const * __restrict__ rec, uint64_t t, uint32_t k,
                                                       uint64_t mask, uint32_t left_shift)
{
    dna_state s{0, 0};
    uint32_t const warm = dna_warmup(k);
    uint64_t t0 = t > warm ? t - warm : 0;
    if (t0 == 0)
    {
        for (uint32_t i = 0; i < k; ++i)
            s.fwd = (s.fwd << 2) | dna_code(__ldg(rec + i));
        s.rev = dna_revcomp(s.fwd, k);
    }
    for (uint64_t j = t0; j < t; ++j)
        s.roll(__ldg(rec + j), mask, left_shift);
    return s;
}

struct emplace_sink
{
    ibf_geom g;
    uint64_t col;
    unsigned long long bit;
    bool mine;
    uint64_t * base;
    uint64_t stride;
    unsigned int * sketch; // HLL mode: this bin's registers
    __device__ __forceinline__ void set_bin(uint32_t b)
    {
        if (g.hll)
        {
            sketch = g.hll + (static_cast<uint64_t>(b) << g.hll_bits);
            mine = true;
            return;
        }
        uint64_t c = b >> 6;
        mine = c >= g.word_begin && c < g.word_begin + g.shard_words;
        col = c - g.word_begin;
        base = g.words;
        stride = g.shard_words;
        if (g.slab && c - g.slab_begin < g.slab_words) // unsigned: also false below slab_begin
        {
            base = g.slab;
            stride = g.slab_words;
            col = c - g.slab_begin;
        }
        bit = 1ULL << (b & 63);
    }
    // What bounds K1 (measured on the C4 slab route, profiles/README.md): a pass costs 1.2 ms of text + hashing per 234 M
    // k-mers plus 10.7 ps per atomic that fires — about 94 G red.or per second at L2, the same for 64- and 32-bit operands
    // and whether or not the target is L2 resident; a target far larger than L2 with one bit per sector (the interleaved
    // matrix itself) pays 21 ps, the DRAM read-modify-write.  The hash loop is unrolled so that the seeds are immediates,
    // and for filters with fewer than 2^32 rows (every configuration of BASELINE.json) the 64x64 high multiply of the
    // fastrange step is two 32x32 wide multiplies: mulhi64(x, bs) = (hi(x)*bs + (lo(x)*bs >> 32)) >> 32 exactly.
    __device__ __forceinline__ void operator()(uint64_t v) const
    {
        if (!mine)
            return;
        if (g.hll)
        {
            // seqan::hibf::sketch::hyperloglog::add (lib/hibf/src/sketch/hyperloglog.cpp:54-70): 64x64→128 multiply folded
            // to 64 bits, top bits = register, rank = leading zeros of the rest + 1
            uint64_t const lo = v * 0x9E3779B97F4A7C15ULL, hi = __umul64hi(v, 0x9E3779B97F4A7C15ULL);
            uint64_t const hash = lo ^ hi;
            uint64_t const rest = (hash << g.hll_bits) | ((1ULL << g.hll_bits) - 1);
            atomicMax(sketch + (hash >> (64 - g.hll_bits)), static_cast<unsigned int>(__clzll(static_cast<long long>(rest)) + 1));
            return;
        }
        if (g.bin_size >> 32 == 0)
        {
            uint64_t const bs = static_cast<uint32_t>(g.bin_size);
            uint32_t const lo = static_cast<uint32_t>(g.row_lo), n_rows = static_cast<uint32_t>(g.row_hi - g.row_lo);
#pragma unroll
            for (int h = 0; h < 5; ++h)
            {
                if (static_cast<uint32_t>(h) < g.hash_funs)
                {
                    uint64_t x = v * hash_seed(h);
                    x ^= x >> g.hash_shift;
                    x *= 11400714819323198485ULL;
                    uint32_t const row = static_cast<uint32_t>(((x >> 32) * bs + (((x & 0xffffffffULL) * bs) >> 32)) >> 32);
                    if (row - lo < n_rows) // this pass's slice of the matrix (the L2-resident one)
                        atomicOr(reinterpret_cast<unsigned long long *>(base + static_cast<uint64_t>(row) * stride + col), bit);
                }
            }
            return;
        }
#pragma unroll
        for (int h = 0; h < 5; ++h)
        {
            if (static_cast<uint32_t>(h) < g.hash_funs)
            {
                uint64_t const row = hash_row(v, h, g.hash_shift, g.bin_size);
                if (row - g.row_lo < g.row_hi - g.row_lo)
                    atomicOr(reinterpret_cast<unsigned long long *>(base + row * stride + col), bit);
            }
        }
    }
};

// matrix[r][col0 + j] |= slab[r][j] for the rows of the slab: plain read-modify-write (stream order keeps emplaces apart),
// untouched words are skipped.  One 32-byte sector is read and written back per touched word: what is left of the
// 64 B per atomic that K1 pays when it sets bits in a matrix far larger than L2 one at a time.
__global__ void __launch_bounds__(256)
    slab_merge_kernel(uint64_t * __restrict__ matrix, uint64_t shard_words, uint64_t col0, uint64_t const * __restrict__ slab,
                      uint64_t slab_words, uint64_t rows)
{
    uint64_t const total = rows * slab_words;
    uint64_t const stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride)
    {
        uint64_t const v = __ldcs(slab + i);
        if (v == 0)
            continue;
        uint64_t const r = i / slab_words, j = i - r * slab_words;
        uint64_t * const dst = matrix + r * shard_words + col0 + j;
        *dst |= v;
    }
}

// 32-bit sketch registers → the bytes seqan's hyperloglog keeps
__global__ void __launch_bounds__(256) hll_pack_kernel(unsigned int const * __restrict__ regs, uint8_t * __restrict__ out, uint64_t n)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
        out[i] = static_cast<uint8_t>(regs[i]);
}

constexpr int kRun = 16;          // residues per thread = one LDG.128
constexpr int kWindowBytes = 48;  // run + halo (k-1 <= 31) rounded up to 3 × 16 B

// Residue-parallel decomposition + emplace.  Thread t owns the windows that START in
// [16t, 16t+16); it reads those 16 bytes plus k-1 halo bytes (three 16-byte loads, the second and
// third hit L1 because the neighbouring lanes load them as their own run), rolls the packed k-mer
// across its run and fires H atomic ORs per window.  Record boundaries come from rec_offsets; the
// record of the first byte is found by binary search, later boundaries by walking forward.
template <int MOL>
__global__ void __launch_bounds__(256)
    emplace_records_kernel(ibf_geom g, uint8_t const * __restrict__ text, uint64_t n_residues,
                           uint64_t const * __restrict__ rec_offsets, uint32_t const * __restrict__ rec_bin,
                           uint64_t n_records, uint32_t k, uint8_t const * __restrict__ aamap)
{
    __shared__ uint8_t s_map[256];
    if constexpr (MOL == TRX_MOL_AA)
    {
        s_map[threadIdx.x] = aamap[threadIdx.x];
        __syncthreads();
    }
    constexpr uint32_t bits = MOL == TRX_MOL_NA ? 2 : 5;
    uint64_t const mask = (bits * k >= 64) ? ~0ULL : ((1ULL << (bits * k)) - 1);
    uint64_t const n_runs = (n_residues + kRun - 1) / kRun;

    for (uint64_t run = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; run < n_runs;
         run += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t const p0 = run * kRun;
        uint64_t const run_end = min(p0 + kRun, n_residues);
        uint64_t const last = min(run_end + k - 1, n_residues); // exclusive: bytes this thread may consume

        // 48 bytes into registers
        uint8_t buf[kWindowBytes];
        if (p0 + kWindowBytes <= n_residues && (reinterpret_cast<uintptr_t>(text) & 15) == 0)
        {
            uint4 const * src = reinterpret_cast<uint4 const *>(text + p0);
            uint4 * dstv = reinterpret_cast<uint4 *>(buf);
            dstv[0] = __ldg(src);
            dstv[1] = __ldg(src + 1);
            dstv[2] = __ldg(src + 2);
        }
        else
        {
#pragma unroll
            for (int j = 0; j < kWindowBytes; ++j)
                buf[j] = (p0 + j < n_residues) ? text[p0 + j] : 0;
        }

        // record containing p0: largest r with rec_offsets[r] <= p0
        uint64_t lo = 0, hi = n_records;
        while (hi - lo > 1)
        {
            uint64_t mid = (lo + hi) >> 1;
            if (__ldg(rec_offsets + mid) <= p0)
                lo = mid;
            else
                hi = mid;
        }
        uint64_t r = lo;
        uint64_t rec_begin = __ldg(rec_offsets + r);
        uint64_t rec_end = __ldg(rec_offsets + r + 1);
        while (rec_end <= p0 && r + 1 < n_records) // empty records sharing the offset
        {
            ++r;
            rec_begin = rec_end;
            rec_end = __ldg(rec_offsets + r + 1);
        }
        emplace_sink sink{g, 0, 0, false};
        sink.set_bin(__ldg(rec_bin + r));

        uint64_t fwd = 0;
        uint32_t cnt = 0;
#pragma unroll 1
        for (uint64_t p = p0; p < last; ++p)
        {
            if (p >= rec_end)
            {
                if (p >= run_end)
                    break; // windows starting in the next record belong to another thread
                do
                {
                    ++r;
                    rec_begin = rec_end;
                    rec_end = __ldg(rec_offsets + r + 1);
                } while (rec_end <= p);
                sink.set_bin(__ldg(rec_bin + r));
                cnt = 0;
                fwd = 0;
            }
            fwd = ((fwd << 5) & mask) | s_map[buf[p - p0]]; // peptide_decomposer.h:261-291
            if (cnt < k)
                ++cnt;
            if (cnt == k)
                sink(fwd);
        }
    }
}

// DNA flavour (run_len steps per thread: the warm-up below costs up to 33 rolls per thread whatever the run length, so DNA
// threads take longer runs than the peptide kernel's 16).  One emit per rolled symbol (plus one per record for the initial k-mer), so a thread owns
// the STEPS whose byte lies in its 16-byte run: it rebuilds the rolling stores from the preceding
// dna_warmup(k) symbols (or from the record start) and then rolls through its run.  That reproduces the
// reference's emits exactly, including the wrap-around k-mers of nucleotide_decomposer.h:110 (the whole
// record is rolled in after the first k-mer) and the k >= 16 reverse-store quirk.
__global__ void __launch_bounds__(256)
    emplace_dna_kernel(ibf_geom g, uint8_t const * __restrict__ text, uint64_t n_residues,
                       uint64_t const * __restrict__ rec_offsets, uint32_t const * __restrict__ rec_bin,
                       uint64_t n_records, uint32_t k, uint32_t run_len)
{
    uint64_t const mask = (2 * k >= 64) ? ~0ULL : ((1ULL << (2 * k)) - 1);
    uint32_t const left_shift = 2 * k - 2;
    uint64_t const n_runs = (n_residues + run_len - 1) / run_len;
    for (uint64_t run = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; run < n_runs;
         run += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t const p0 = run * run_len;
        uint64_t const run_end = min(p0 + run_len, n_residues);
        uint64_t lo = 0, hi = n_records;
        while (hi - lo > 1)
        {
            uint64_t mid = (lo + hi) >> 1;
            if (__ldg(rec_offsets + mid) <= p0)
                lo = mid;
            else
                hi = mid;
        }
        uint64_t r = lo;
        uint64_t rec_begin = __ldg(rec_offsets + r);
        uint64_t rec_end = __ldg(rec_offsets + r + 1);
        emplace_sink sink{g, 0, 0, false};
        sink.set_bin(__ldg(rec_bin + r));
        bool usable = rec_end - rec_begin >= k; // shorter records are skipped (include/index_ibf.h:114-118)
        dna_state s{0, 0};
        if (usable)
        {
            s = dna_state_before(text + rec_begin, p0 - rec_begin, k, mask, left_shift);
            if (p0 == rec_begin)
                sink(s.canonical()); // the initial k-mer
        }
#pragma unroll 1
        for (uint64_t p = p0; p < run_end; ++p)
        {
            if (p >= rec_end)
            {
                do
                {
                    ++r;
                    rec_begin = rec_end;
                    rec_end = __ldg(rec_offsets + r + 1);
                } while (rec_end <= p);
                sink.set_bin(__ldg(rec_bin + r));
                usable = rec_end - rec_begin >= k;
                if (usable)
                {
                    s = dna_state_before(text + rec_begin, 0, k, mask, left_shift);
                    sink(s.canonical());
                }
            }
            if (!usable)
                continue;
            s.roll(__ldg(text + p), mask, left_shift);
            sink(s.canonical());
        }
    }
}

// ---------------------------------------------------------------------------------------- DGram (gapped 3+3-mers)
// DGramIndex::process_sequence (include/dGramIndex.h:194-243) for every record at once.  The reference walks a record
// with i = 2 .. while i + min_gap + 3 < L: the three residues ending at i, a gap of min_gap..max_gap residues, the three
// residues starting at j = i + gap + 1; all six must be letters of "ACDEFGHIKLMNPQRSTVWY" (upper case only: alpha_map_
// holds nothing else, :119-126), the inner loop stops at the first gap whose right side would leave the record
// (`if (j + 2 >= L) break`), and records shorter than min_gap + 7 give nothing.  code = gap·20^6 + a1·20^5 + a2·20^4 +
// a3·20^3 + b1·400 + b2·20 + b3.  One thread per position i; the block's residues plus a halo of max_gap + 3 are turned
// into 0..19 / 0xFF codes in shared memory once, so the 3 + 3·gaps reads per position never leave the SM.
//   COUNT = true : per_bin[bin] += number of codes (duplicates included — dgram_buffer_ keeps them, and the largest
//                  buffer sizes the filter, :128-148)
//   COUNT = false: every code goes through emplace (init_ibf, :149-153) — H atomic ORs each.
constexpr int kDgramTile = 256;

__device__ __forceinline__ uint8_t dgram_letter(uint8_t c)
{
    // "ACDEFGHIKLMNPQRSTVWY" → 0..19 (the alphabet is in alphabetical order, so the code of a letter is the number of
    // alphabet letters below it); B J O U X Z and everything that is not an upper-case letter → 0xFF
    constexpr uint32_t letters = 0x16fbdfdu;
    uint32_t const d = static_cast<uint32_t>(c) - 'A';
    if (d >= 26 || !(letters >> d & 1))
        return 0xFF;
    return static_cast<uint8_t>(__popc(letters & ((1u << d) - 1)));
}

// residue range of every 64-bin slab (records are grouped by ascending bin): range[s] = offset of the first record whose
// bin is >= 64 s, range[n_slabs] = n_residues.  One thread per slab, binary search over rec_bin.
__global__ void dgram_slab_ranges_kernel(uint64_t const * __restrict__ rec_offsets, uint32_t const * __restrict__ rec_bin,
                                         uint64_t n_records, uint64_t n_slabs, uint64_t * __restrict__ range)
{
    uint64_t const s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s > n_slabs)
        return;
    uint64_t lo = 0, hi = n_records; // first record with rec_bin >= 64 s
    while (lo < hi)
    {
        uint64_t const mid = (lo + hi) >> 1;
        if (static_cast<uint64_t>(rec_bin[mid]) < 64 * s)
            lo = mid + 1;
        else
            hi = mid;
    }
    range[s] = rec_offsets[s == n_slabs ? n_records : lo];
}

// `range` (optional, device): only positions in [range[0], range[1]) are worked on — the residues of one slab of bins
template <bool COUNT>
__global__ void __launch_bounds__(kDgramTile)
    dgram_kernel(ibf_geom g, uint8_t const * __restrict__ text, uint64_t n_residues,
                 uint64_t const * __restrict__ rec_offsets, uint32_t const * __restrict__ rec_bin, uint64_t n_records,
                 uint32_t min_gap, uint32_t max_gap, unsigned long long * __restrict__ per_bin, uint64_t const * __restrict__ range)
{
    extern __shared__ uint8_t s_code[]; // codes of positions [t0 - 2, t0 + kDgramTile + max_gap + 3)
    uint32_t const span = kDgramTile + 2 + max_gap + 3;
    uint64_t const p_lo = range ? range[0] : 0, p_hi = range ? range[1] : n_residues;
    uint64_t const n_tiles = (p_hi + kDgramTile - 1) / kDgramTile;
    for (uint64_t tile = p_lo / kDgramTile + blockIdx.x; tile < n_tiles; tile += gridDim.x)
    {
        uint64_t const t0 = tile * kDgramTile;
        for (uint32_t j = threadIdx.x; j < span; j += kDgramTile)
        {
            uint64_t const p = t0 + j; // position + 2
            s_code[j] = (p >= 2 && p - 2 < n_residues) ? dgram_letter(__ldg(text + p - 2)) : 0xFF;
        }
        __syncthreads();
        uint64_t const p = t0 + threadIdx.x;
        unsigned long long cnt = 0;
        uint32_t bin = 0;
        if (p >= p_lo && p < p_hi)
        {
            // record containing p: largest r with rec_offsets[r] <= p, stepping over empty records that share the offset
            uint64_t lo = 0, hi = n_records;
            while (hi - lo > 1)
            {
                uint64_t mid = (lo + hi) >> 1;
                if (__ldg(rec_offsets + mid) <= p)
                    lo = mid;
                else
                    hi = mid;
            }
            uint64_t const rec_begin = __ldg(rec_offsets + lo), rec_end = __ldg(rec_offsets + lo + 1);
            uint64_t const L = rec_end - rec_begin, i = p - rec_begin;
            uint8_t const * const w = s_code + threadIdx.x; // w[0..2] = residues i-2, i-1, i
            bin = __ldg(rec_bin + lo);
            if (L >= static_cast<uint64_t>(min_gap) + 7 && i >= 2 && i + min_gap + 3 < L && w[0] != 0xFF && w[1] != 0xFF &&
                w[2] != 0xFF)
            {
                emplace_sink sink{g, 0, 0, false};
                if constexpr (!COUNT)
                    sink.set_bin(bin);
                uint64_t const left = w[0] * 3200000ULL + w[1] * 160000ULL + w[2] * 8000ULL;
                for (uint32_t gap = min_gap; gap <= max_gap; ++gap)
                {
                    if (i + gap + 3 >= L) // j + 2 >= L with j = i + gap + 1
                        break;
                    uint8_t const b1 = w[gap + 3], b2 = w[gap + 4], b3 = w[gap + 5];
                    if (b1 == 0xFF || b2 == 0xFF || b3 == 0xFF)
                        continue;
                    if constexpr (COUNT)
                        ++cnt;
                    else
                        sink(gap * 64000000ULL + left + b1 * 400ULL + b2 * 20ULL + b3);
                }
            }
        }
        if constexpr (COUNT)
        {
            // lanes of a warp nearly always sit in the same record: one atomic per warp then
            unsigned const active = __activemask();
            unsigned const same = __match_any_sync(active, bin);
            if (same == active)
            {
                unsigned long long total = cnt;
                for (int o = 16; o > 0; o >>= 1)
                    total += __shfl_xor_sync(active, total, o);
                if ((threadIdx.x & 31) == 0 && total)
                    atomicAdd(per_bin + bin, total);
            }
            else if (cnt)
                atomicAdd(per_bin + bin, cnt);
        }
        __syncthreads();
    }
}

// Ordered emission (decompose_record's exact call sequence) for hand-off to HIBF construction and for
// parity tests of rows a1/a2.  One thread per emitted value; packs its k symbols directly.
template <int MOL>
__global__ void __launch_bounds__(256)
    decompose_kernel(uint8_t const * __restrict__ text, uint64_t const * __restrict__ rec_offsets,
                     uint64_t const * __restrict__ out_offsets, uint64_t n_records, uint32_t k,
                     uint8_t const * __restrict__ aamap, uint64_t * __restrict__ out, uint64_t n_out)
{
    for (uint64_t e = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n_out;
         e += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t lo = 0, hi = n_records;
        while (hi - lo > 1)
        {
            uint64_t mid = (lo + hi) >> 1;
            if (out_offsets[mid] <= e)
                lo = mid;
            else
                hi = mid;
        }
        // skip records with zero emits that share the offset: pick the one that actually contains e
        while (out_offsets[lo + 1] <= e)
            ++lo;
        uint8_t const * rec = text + rec_offsets[lo];
        uint64_t const j = e - out_offsets[lo];
        uint64_t v = 0;
        if constexpr (MOL == TRX_MOL_NA)
        {
            // emit 0 = initial k-mer; emit j >= 1 = state after rolling rec[j-1]
            uint64_t const mask = (2 * k >= 64) ? ~0ULL : ((1ULL << (2 * k)) - 1);
            dna_state st = dna_state_before(rec, j == 0 ? 0 : j - 1, k, mask, 2 * k - 2);
            if (j > 0)
                st.roll(rec[j - 1], mask, 2 * k - 2);
            out[e] = st.canonical();
            (void)v;
        }
        else
        {
            for (uint32_t s = 0; s < k; ++s)
                v = (v << 5) | aamap[rec[j + s]];
            out[e] = v;
        }
    }
}

} // namespace trx
