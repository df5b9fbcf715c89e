// regex_kernels.cuh — K4 device code: streaming Glushkov scan (phase 1) and exact match resolution
// (phase 2).  Reference semantics: src/query.cpp:167-315 (FindAndConsume loops), include/query.h:97-188.
//
// Phase 1 streams every residue of the candidate bins once.  The text is cut into 128-byte chunks on
// a fixed global grid; a CTA of 256 threads takes a 32 KB tile (256 chunks) that one elected thread
// fetches with a single 2-D TMA tensor copy (SWIZZLE_128B, so that thread t reading 16-byte piece i of
// row t touches bank group i^(t&7): conflict-free LDS.128) into a 2-stage shared-memory ring guarded
// by mbarriers.  Thread t runs the position automaton over chunk t, injecting the start state only
// inside its own chunk and running on past the chunk end (no injection) until the state dies, so every
// match is seen by exactly the thread whose chunk holds its START.  Output: one flag bit per chunk and
// one byte per record ("a match starts here").
//
// Phase 2 runs only on flagged records (rare) and reproduces the reference's non-overlapping
// leftmost-longest (POSIX) or leftmost-first (Perl) match sequence exactly.
#pragma once

#include <cuda.h>

#include "common.cuh"
#include "regex.cuh"

namespace trx
{

constexpr int kChunkBytes = 128;
constexpr int kScanThreads = 256;
constexpr int kTileBytes = kChunkBytes * kScanThreads; // 32 KB
constexpr int kStages = 2;              // exact automaton kernel (phase 1)
constexpr int kFilterSplitDefault = 1;  // prefilter kernels: threads per text row; see filter_split()
constexpr unsigned long long kSmallHits = 1024;

// device copy of the automaton; B already has the per-scan byte map (redmap / comp_tab) folded in
struct regex_dev
{
    uint64_t first[kRegexWords];
    uint64_t last[kRegexWords];
    uint64_t B[256][kRegexWords];
    uint64_t follow[64 * kRegexWords][kRegexWords];
    // ordered lists for leftmost-first
    uint16_t first_len;
    uint16_t follow_off[64 * kRegexWords + 1];
    uint8_t first_list[64 * kRegexWords];
    uint8_t follow_list[4096]; // flattened; compile rejects patterns whose lists do not fit
    int32_t positions;
    int32_t posix_longest;
    int32_t anchor_begin;
    int32_t anchor_end;
};

// necessary-factor prefilter: up to 8 literal bytes inside an 8-byte window that EVERY match carries at a
// fixed offset from its first byte (derived on the host from the automaton's level sets).  Literals are
// ordered by address; the first n_a lie in bytes 0..3 of the window, the following n_b in bytes 4..7.
struct filter_spec
{
    int32_t n_a, n_b;         // n_a + n_b = 0: no usable factor, run the exact automaton kernel instead
    int32_t start_delta;      // address of a match's first byte (reading direction) minus address of window byte 0
    uint32_t delta[8];        // byte offset of literal j inside its 4-byte half (0..3)
    uint32_t splat[8];        // literal byte replicated into all four lanes
};

// gapped necessary-factor prefilter ("atoms"): up to 4 distinct literal bytes and up to 8 atoms.  An atom says "one of
// the atom's bytes occurs at address anchor + d for some d in [lo, hi]" (0 <= lo <= hi <= 31); it is stored as one TERM
// per byte of its class (1-3 bytes), the terms of an atom are consecutive and the last one carries end_group.  One atom
// is the anchor itself (lo = hi = 0).  Every match contains such a constellation, and its first byte (reading
// direction) lies at an address in [anchor + start_lo, anchor + start_hi].  Derived on the host from the automaton's
// dominator chain.
constexpr int kMaxAtomTerms = 16;
constexpr int kMaxConfirm = 6;
constexpr int kWindowLen = 6;     // positions of a class window (regex_window_multi)
constexpr int kMultiLetters = 24; // alphabet of a multi-pattern batch (regex_filter_multi); also the room in atom_spec::splat
// Second stage of the gapped prefilter, evaluated only at the anchors the SIMD stage lets through: "the byte at address
// anchor + d belongs to `mask` for some d in [lo, hi]" (offsets may be negative; classes of any size).  These are the
// necessary positions that do not fit the SIMD stage's budget of 4 distinct bytes.
struct confirm_check
{
    int8_t lo, hi;
    uint8_t pad[2];
    uint32_t mask[8]; // bit c set: byte c is in the class
};
struct atom_spec
{
    int32_t n_lit;            // 0: unused
    int32_t n_atoms;
    int32_t n_terms;
    int32_t start_lo, start_hi;
    uint32_t splat[kMultiLetters];   // the single-pattern kernel uses the first n_lit <= 4
    uint8_t term_lit[kMaxAtomTerms]; // index into splat
    uint8_t term_lo[kMaxAtomTerms], term_hi[kMaxAtomTerms];
    uint8_t term_end[kMaxAtomTerms]; // 1 = last term of its atom
    int32_t n_confirm;
    confirm_check confirm[kMaxConfirm]; // most selective first
};

struct scan_params
{
    uint8_t const * text;         // padded to a tile multiple, zero filled
    uint64_t n_text;              // real bytes
    uint64_t n_text_padded;
    uint64_t const * rec_offsets; // n_records + 1
    uint32_t const * rec_bin;     // n_records
    uint64_t n_records;
    uint64_t const * cand_words;  // candidate bin bitmap (device)
    uint32_t const * tiles;       // tile ids to scan
    uint32_t n_tiles;
    uint32_t * chunk_flags;       // bitmap over global chunk ids: "a match may start in this chunk"
    uint32_t * rec_flags;         // bitmap over records already listed
    uint32_t * rec_list;          // records that may hold a match (each once), for phase 2
    unsigned int * rec_count;
    unsigned int * tile_counter;  // next entry of `tiles` to hand out (prefilter kernels schedule tiles dynamically)
    regex_dev const * re;
    filter_spec filter;
    atom_spec atoms;
    // reverse searcher (regex_scan_starts_dfa): exact match starts, one bit per text byte
    uint32_t * start_bits;         // n_text_padded / 32 words, zeroed per scan
    uint16_t const * r_trans;      // states pre-multiplied by r_classes, like the matcher's table
    uint8_t const * r_cls;         // 256 bytes
    uint32_t r_classes, r_states, r_first_accept, r_max_len;
};

// first flagging of a record appends it to the phase-2 work list
__device__ __forceinline__ void list_record(scan_params const & p, uint64_t r)
{
    uint32_t const bit = 1u << (r & 31);
    uint32_t const old = atomicOr(p.rec_flags + (r >> 5), bit);
    if (!(old & bit))
        p.rec_list[atomicAdd(p.rec_count, 1u)] = static_cast<uint32_t>(r);
}

template <int W>
struct state
{
    uint64_t w[W];
    __device__ __forceinline__ bool any() const
    {
        uint64_t x = w[0];
#pragma unroll
        for (int k = 1; k < W; ++k)
            x |= w[k];
        return x != 0;
    }
};

// shared-memory image of the tables phase 1 needs
template <int W>
struct scan_tables
{
    uint64_t Bf[256][W]; // B & first: the whole step when the state is empty
    uint64_t B[256][W];
    uint64_t follow[64 * W][W];
    uint64_t first[W];
    uint64_t last[W];
};

template <int W>
__device__ __forceinline__ void step(state<W> & S, scan_tables<W> const & T, uint8_t c, bool inject)
{
    if (!S.any())
    {
        if (inject)
        {
#pragma unroll
            for (int k = 0; k < W; ++k)
                S.w[k] = T.Bf[c][k];
        }
        return;
    }
    uint64_t F[W];
#pragma unroll
    for (int k = 0; k < W; ++k)
        F[k] = inject ? T.first[k] : 0;
#pragma unroll
    for (int k = 0; k < W; ++k)
    {
        // 32-bit halves: find-first-set on 32 bits is one FLO, on 64 bits a dozen instructions
#pragma unroll
        for (int half = 0; half < 2; ++half)
        {
            uint32_t s = static_cast<uint32_t>(S.w[k] >> (32 * half));
            while (s)
            {
                int i = __ffs(static_cast<int>(s)) - 1 + 32 * half;
                s &= s - 1;
#pragma unroll
                for (int q = 0; q < W; ++q)
                    F[q] |= T.follow[k * 64 + i][q];
            }
        }
    }
#pragma unroll
    for (int k = 0; k < W; ++k)
        S.w[k] = F[k] & T.B[c][k];
}

template <int W>
__device__ __forceinline__ bool accepts(state<W> const & S, scan_tables<W> const & T)
{
    uint64_t x = S.w[0] & T.last[0];
#pragma unroll
    for (int k = 1; k < W; ++k)
        x |= S.w[k] & T.last[k];
    return x != 0;
}

__device__ __forceinline__ uint32_t smem_u32(void const * p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity)
{
    uint32_t ok;
    do
    {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}

// one elected thread: arm the barrier with the byte count, then launch the 2-D tensor copy of one tile
__device__ __forceinline__ void tma_load_tile(CUtensorMap const * tmap, void * dst, uint64_t * bar, int32_t row)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(kTileBytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :
                 : "r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(0), "r"(row), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ bool is_candidate(uint64_t const * cand, uint32_t bin)
{
    return (cand[bin >> 6] >> (bin & 63)) & 1;
}

// largest r in [0, n) with off[r] <= g, skipping forward over empty records so that off[r+1] > g
__device__ __forceinline__ uint64_t record_of(uint64_t const * __restrict__ off, uint64_t n, uint64_t g)
{
    uint64_t lo = 0, hi = n;
    while (hi - lo > 1)
    {
        uint64_t mid = (lo + hi) >> 1;
        if (__ldg(off + mid) <= g)
            lo = mid;
        else
            hi = mid;
    }
    return lo;
}

// DIR = +1: forward strand.  DIR = -1: the record is read back to front (reverse strand; the
// complement is folded into B).  "Position" below always means position in reading direction.
template <int W, int DIR, bool USE_TMA>
__global__ void __launch_bounds__(kScanThreads)
    regex_scan_phase1(scan_params p, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // layout: [tile stage 0][tile stage 1][tables][barriers]   (tiles first: SWIZZLE_128B wants 1024-B alignment)
    uint8_t * tiles = smem_raw;
    auto & T = *reinterpret_cast<scan_tables<W> *>(smem_raw + (USE_TMA ? kStages * kTileBytes : 0));
    uint64_t * bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(&T) + sizeof(scan_tables<W>));

    int const tid = threadIdx.x;
    regex_dev const & re = *p.re;
    for (int c = tid; c < 256; c += kScanThreads)
#pragma unroll
        for (int k = 0; k < W; ++k)
        {
            T.B[c][k] = re.B[c][k];
            T.Bf[c][k] = re.B[c][k] & re.first[k];
        }
    for (int i = tid; i < 64 * W; i += kScanThreads)
#pragma unroll
        for (int k = 0; k < W; ++k)
            T.follow[i][k] = re.follow[i][k];
    if (tid < W)
    {
        T.first[tid] = re.first[tid];
        T.last[tid] = re.last[tid];
    }
    bool const anchor_begin = re.anchor_begin != 0;
    bool const anchor_end = re.anchor_end != 0;
    if constexpr (USE_TMA)
    {
        if (tid == 0)
        {
            mbar_init(&bars[0], 1);
            mbar_init(&bars[1], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();

    uint32_t const first_tile = blockIdx.x;
    if constexpr (USE_TMA)
    {
        if (tid == 0 && first_tile < p.n_tiles)
            tma_load_tile(&tmap, tiles, &bars[0], static_cast<int32_t>(p.tiles[first_tile]) * kScanThreads);
    }

    uint32_t it = 0;
    for (uint32_t t = first_tile; t < p.n_tiles; t += gridDim.x, ++it)
    {
        uint32_t const stage = it & 1;
        if constexpr (USE_TMA)
        {
            // prefetch the next tile into the other stage; every thread left it at the previous __syncthreads
            uint32_t const tn = t + gridDim.x;
            if (tid == 0 && tn < p.n_tiles)
            {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                tma_load_tile(&tmap, tiles + (stage ^ 1) * kTileBytes, &bars[stage ^ 1],
                              static_cast<int32_t>(p.tiles[tn]) * kScanThreads);
            }
            mbar_wait(&bars[stage], (it >> 1) & 1);
        }
        uint64_t const chunk_id = static_cast<uint64_t>(p.tiles[t]) * kScanThreads + tid;
        uint64_t const g_lo = chunk_id * kChunkBytes;
        if (g_lo < p.n_text)
        {
            uint64_t const g_hi = min(g_lo + kChunkBytes, p.n_text); // exclusive
            // Positions inside the chunk are counted as STEPS in reading direction — step j reads address g_lo + j
            // forwards, g_lo + 127 - j backwards — so the per-byte bookkeeping is 32-bit: `boundary` = first step
            // that no longer belongs to the current record, `first_step` = step of the record's first byte (negative
            // if the record began before this chunk).  Only the last chunk of the text is short: forwards it ends at
            // step n_here, backwards its first 128 - n_here steps are padding.
            int32_t const n_here = static_cast<int32_t>(g_hi - g_lo);
            int32_t const j_lo = DIR > 0 ? 0 : kChunkBytes - n_here;
            int32_t const j_hi = DIR > 0 ? n_here : kChunkBytes;
            uint64_t const g_first = DIR > 0 ? g_lo : g_hi - 1;
            uint64_t r = record_of(p.rec_offsets, p.n_records, g_first);
            uint64_t rec_begin = __ldg(p.rec_offsets + r), rec_end = __ldg(p.rec_offsets + r + 1);
            while (rec_end <= g_first) // empty records sharing this offset
            {
                ++r;
                rec_begin = rec_end;
                rec_end = __ldg(p.rec_offsets + r + 1);
            }
            bool cand = is_candidate(p.cand_words, __ldg(p.rec_bin + r));
            state<W> S;
#pragma unroll
            for (int k = 0; k < W; ++k)
                S.w[k] = 0;
            bool chunk_hit = false, rec_hit = false;
            int32_t boundary = 0, first_step = 0;
            auto set_bounds = [&](int32_t j, uint64_t gb)
            {
                uint64_t const left = DIR > 0 ? rec_end - gb : gb - rec_begin + 1; // steps until the record is left
                uint64_t const done = DIR > 0 ? gb - rec_begin : rec_end - 1 - gb; // steps since its first byte
                boundary = j + static_cast<int32_t>(left < (1u << 20) ? left : (1u << 20));
                first_step = done < (1u << 20) ? j - static_cast<int32_t>(done) : -(1 << 20);
            };
            set_bounds(j_lo, g_first);
            bool pending_start = j_lo > 0; // backwards, short last chunk: nothing to read before step j_lo
            bool const cand_at_start = cand;
            if (pending_start)
            {
                cand = false;
                boundary = j_lo;
            }
            auto flush_record = [&]()
            {
                if (rec_hit)
                    list_record(p, r);
                rec_hit = false;
            };
            // step j == boundary: the next record in reading direction begins, or the text ends
            auto on_boundary = [&](int32_t j)
            {
                if (pending_start)
                {
                    pending_start = false;
                    cand = cand_at_start;
                    set_bounds(j_lo, g_first);
                    return;
                }
                flush_record();
#pragma unroll
                for (int k = 0; k < W; ++k)
                    S.w[k] = 0;
                if (j >= j_hi)
                {
                    cand = false; // padding behind the last residue
                    boundary = 1 << 30;
                    return;
                }
                uint64_t const gb = DIR > 0 ? g_lo + j : g_lo + (kChunkBytes - 1 - j);
                if constexpr (DIR > 0)
                {
                    do
                    {
                        ++r;
                        rec_begin = rec_end;
                        rec_end = __ldg(p.rec_offsets + r + 1);
                    } while (rec_end <= gb);
                }
                else
                {
                    do
                    {
                        --r;
                        rec_end = rec_begin;
                        rec_begin = __ldg(p.rec_offsets + r);
                    } while (rec_begin > gb);
                }
                cand = is_candidate(p.cand_words, __ldg(p.rec_bin + r));
                set_bounds(j, gb);
            };
            auto feed = [&](uint8_t c, bool inject, bool at_first, bool at_last)
            {
                step<W>(S, T, c, inject && (!anchor_begin || at_first));
                if (S.any() && accepts<W>(S, T) && (!anchor_end || at_last))
                {
                    chunk_hit = true;
                    rec_hit = true;
                }
            };

            uint8_t const * row = tiles + stage * kTileBytes + tid * kChunkBytes;
#pragma unroll 1
            for (int piece = 0; piece < kChunkBytes / 16; ++piece)
            {
                int32_t const j0 = piece * 16;
                if (j0 >= j_hi)
                    break; // beyond the end of the text (forwards, last chunk)
                if (j0 + 16 <= j_lo)
                    continue; // padding before the last residue (backwards, last chunk)
                int const pi = DIR > 0 ? piece : kChunkBytes / 16 - 1 - piece;
                uint4 v;
                if constexpr (USE_TMA)
                    v = *reinterpret_cast<uint4 const *>(row + ((pi ^ (tid & 7)) << 4));
                else
                    v = __ldg(reinterpret_cast<uint4 const *>(p.text + g_lo + static_cast<uint64_t>(pi) * 16));
                uint32_t const words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int b = 0; b < 16; ++b)
                {
                    int32_t const j = j0 + b;
                    if (j == boundary)
                        on_boundary(j);
                    if (!cand)
                        continue;
                    int const bi = DIR > 0 ? b : 15 - b;
                    uint8_t const c = static_cast<uint8_t>(words[bi >> 2] >> ((bi & 3) * 8));
                    feed(c, true, j == first_step, j + 1 == boundary);
                }
            }
            // run on past the chunk while a match that started inside it is still alive
            if (cand && S.any() && (DIR < 0 || n_here == kChunkBytes))
            {
                if constexpr (DIR > 0)
                {
                    for (uint64_t gb = g_hi; gb < rec_end && S.any(); ++gb)
                        feed(__ldg(p.text + gb), false, false, gb + 1 == rec_end);
                }
                else
                {
                    for (int64_t gb = static_cast<int64_t>(g_lo) - 1; gb >= static_cast<int64_t>(rec_begin) && S.any(); --gb)
                        feed(__ldg(p.text + gb), false, false, static_cast<uint64_t>(gb) == rec_begin);
                }
            }
            flush_record();
            if (chunk_hit)
                atomicOr(p.chunk_flags + (chunk_id >> 5), 1u << (chunk_id & 31));
        }
        __syncthreads(); // everyone is done with this stage before it is refilled
    }
}

// Phase 1 for patterns without a selective necessary factor: the reverse searcher DFA (regex.cuh) reads every chunk
// BACKWARDS, one table lookup per byte, after warming up on the max_len - 1 bytes that follow the chunk in its record, and
// writes the exact set of match starts as one bit per text byte.  Same tile ring as the exact automaton kernel (thread per
// 128-byte row of a TMA tile).  The position-automaton kernel above costs 43 thread-instructions per residue; this one a
// class lookup and a table lookup.
__global__ void __launch_bounds__(kScanThreads)
    regex_scan_starts_dfa(scan_params p, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // layout: [tile stage 0][tile stage 1][256 class bytes][barriers][table]
    uint8_t * tiles = smem_raw;
    uint8_t * s_cls = smem_raw + kStages * kTileBytes;
    uint64_t * bars = reinterpret_cast<uint64_t *>(s_cls + 256);
    uint16_t * s_trans = reinterpret_cast<uint16_t *>(s_cls + 256 + 16);
    int const tid = threadIdx.x;
    for (int c = tid; c < 256; c += kScanThreads)
        s_cls[c] = p.r_cls[c];
    for (unsigned i = tid; i < p.r_states * p.r_classes; i += kScanThreads)
        s_trans[i] = p.r_trans[i];
    uint32_t const accept_from = p.r_first_accept * p.r_classes;
    if (tid == 0)
    {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t const first_tile = blockIdx.x;
    if (tid == 0 && first_tile < p.n_tiles)
        tma_load_tile(&tmap, tiles, &bars[0], static_cast<int32_t>(p.tiles[first_tile]) * kScanThreads);
    uint32_t it = 0;
    for (uint32_t t = first_tile; t < p.n_tiles; t += gridDim.x, ++it)
    {
        uint32_t const stage = it & 1;
        uint32_t const tn = t + gridDim.x;
        if (tid == 0 && tn < p.n_tiles)
        {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tma_load_tile(&tmap, tiles + (stage ^ 1) * kTileBytes, &bars[stage ^ 1], static_cast<int32_t>(p.tiles[tn]) * kScanThreads);
        }
        mbar_wait(&bars[stage], (it >> 1) & 1);
        uint64_t const chunk_id = static_cast<uint64_t>(p.tiles[t]) * kScanThreads + tid;
        uint64_t const g_lo = chunk_id * kChunkBytes;
        if (g_lo < p.n_text)
        {
            uint64_t const g_hi = min(g_lo + kChunkBytes, p.n_text); // exclusive
            uint64_t const r_last = [&] {
                uint64_t r = record_of(p.rec_offsets, p.n_records, g_hi - 1);
                while (__ldg(p.rec_offsets + r + 1) <= g_hi - 1) // empty records sharing this offset
                    ++r;
                return r;
            }();
            // Everything that depends on record boundaries is decided BEFORE the byte loop, as two 128-bit masks over the
            // chunk, so that the loop itself is the same instruction stream for every lane (with the bookkeeping inside
            // the loop the warps ran at 13 of 32 threads): `ends` = last byte of a record (the state is cleared before
            // such a byte is consumed), `valid` = byte of a record in a candidate bin (only those may report a start).
            uint32_t ends[4] = {0, 0, 0, 0}, valid[4] = {0, 0, 0, 0};
            auto or_range = [](uint32_t (&m)[4], uint32_t lo, uint32_t hi) { // bits [lo, hi) of a 128-bit mask
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    uint32_t const a = lo > 32u * k ? lo - 32u * k : 0u, b = hi > 32u * k ? hi - 32u * k : 0u;
                    if (a < 32 && b > a)
                        m[k] |= (b >= 32 ? ~0u : ((1u << b) - 1)) & (~0u << a);
                }
            };
            {
                uint64_t r = r_last, rb = __ldg(p.rec_offsets + r), re = __ldg(p.rec_offsets + r + 1);
                while (true)
                {
                    uint32_t const lo = rb > g_lo ? static_cast<uint32_t>(rb - g_lo) : 0u;
                    uint32_t const hi = static_cast<uint32_t>(min(re, g_hi) - g_lo);
                    if (hi > lo)
                    {
                        if (is_candidate(p.cand_words, __ldg(p.rec_bin + r)))
                            or_range(valid, lo, hi);
                        if (re <= g_hi)
                            ends[(hi - 1) >> 5] |= 1u << ((hi - 1) & 31);
                    }
                    if (rb <= g_lo || r == 0)
                        break;
                    --r;
                    re = rb;
                    rb = __ldg(p.rec_offsets + r);
                }
            }
            // warm-up on the bytes that follow the chunk in its last record: the same number of steps for every lane
            uint32_t st = 0;
            {
                uint64_t const re = __ldg(p.rec_offsets + r_last + 1);
                uint64_t const w_end = min(g_hi + p.r_max_len - 1, re);
                for (uint32_t k = p.r_max_len - 1; k-- > 0;)
                {
                    uint64_t const g = g_hi + k;
                    if (g < w_end)
                        st = s_trans[st + s_cls[__ldg(p.text + g)]];
                }
            }
            uint32_t mask[4] = {0, 0, 0, 0};
            uint8_t const * row = tiles + stage * kTileBytes + tid * kChunkBytes;
#pragma unroll
            for (int piece = kChunkBytes / 16 - 1; piece >= 0; --piece)
            {
                uint4 const v = *reinterpret_cast<uint4 const *>(row + ((piece ^ (tid & 7)) << 4));
                uint32_t const words[4] = {v.x, v.y, v.z, v.w};
                uint32_t const end16 = (ends[piece >> 1] >> ((piece & 1) * 16)) & 0xFFFFu;
                uint32_t piece_mask = 0;
#pragma unroll
                for (int b = 15; b >= 0; --b)
                {
                    uint32_t const cl = s_cls[(words[b >> 2] >> ((b & 3) * 8)) & 0xFF]; // independent of the state
                    st = (end16 >> b) & 1 ? 0u : st;
                    st = s_trans[st + cl];
                    piece_mask |= st >= accept_from ? 1u << b : 0u;
                }
                mask[piece >> 1] |= piece_mask << ((piece & 1) * 16);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                mask[k] &= valid[k];
            if (mask[0] | mask[1] | mask[2] | mask[3])
            {
                *reinterpret_cast<uint4 *>(p.start_bits + chunk_id * 4) = make_uint4(mask[0], mask[1], mask[2], mask[3]);
                atomicOr(p.chunk_flags + (chunk_id >> 5), 1u << (chunk_id & 31));
                // list the records that own a start
                uint64_t r = r_last, rb = __ldg(p.rec_offsets + r), re = __ldg(p.rec_offsets + r + 1);
                while (true)
                {
                    uint32_t const lo = rb > g_lo ? static_cast<uint32_t>(rb - g_lo) : 0u;
                    uint32_t const hi = static_cast<uint32_t>(min(re, g_hi) - g_lo);
                    if (hi > lo)
                    {
                        uint32_t range[4] = {0, 0, 0, 0};
                        or_range(range, lo, hi);
                        if ((range[0] & mask[0]) | (range[1] & mask[1]) | (range[2] & mask[2]) | (range[3] & mask[3]))
                            list_record(p, r);
                    }
                    if (rb <= g_lo || r == 0)
                        break;
                    --r;
                    re = rb;
                    rb = __ldg(p.rec_offsets + r);
                }
            }
        }
        __syncthreads(); // everyone is done with this stage before it is refilled
    }
}

} // namespace trx

namespace trx
{

struct resolve_params
{
    uint8_t const * text;
    uint64_t const * rec_offsets;
    uint32_t const * rec_list;         // records flagged by phase 1
    unsigned int const * rec_count;
    uint32_t const * chunk_flags;
    regex_dev const * re;
    unsigned long long * hit_count;
    uint64_t * hit_keys;               // record << 32 | start
    uint32_t * hit_ends;
    uint64_t * small_keys;             // the first kSmallHits hits go here: they sit next to the counters, so that
    uint32_t * small_ends;             // one device-to-host copy brings the count and (almost always) every hit
    uint64_t cap;
    // multi-pattern scans (regex_scan_phase2_multi): hits of all patterns share one list, tagged with the pattern;
    // nullptr = single-pattern scan.  rec_cap bounds the listed records (0 = no bound); chunk_flags may be nullptr
    // (no chunk is skipped).
    uint32_t * hit_pats;
    uint32_t pattern_id;
    uint32_t rec_cap;
    // Determinised anchored matcher (POSIX leftmost-longest patterns whose subset automaton stays small): state 0 = dead,
    // state 1 = before the first byte, states >= dfa_first_accept hold an accepting position.  The device copy of the table
    // holds states PRE-MULTIPLIED by dfa_classes: next = trans[state * dfa_classes + cls[byte]] is again such a product.
    // nullptr = simulate the position automaton (match_longest / match_first).
    uint16_t const * dfa_trans;
    uint8_t const * dfa_cls; // 256 bytes
    uint32_t dfa_classes;
    uint32_t dfa_states;
    uint32_t dfa_first_accept;
    uint32_t anchor_begin, anchor_end; // copies of the automaton's flags: the DFA route never stages regex_dev
    uint32_t const * start_bits;       // exact match starts from regex_scan_starts_dfa (nullptr: every position is tried)
};

// Anchored leftmost-longest match starting at position s of a record read through `at`: the automaton
// runs without re-injection; the last position at which an accepting state was live wins.
template <int W, class At>
__device__ int64_t match_longest(regex_dev const & re, At at, int64_t s, int64_t L)
{
    uint64_t S[W];
    uint8_t c = at(s);
#pragma unroll
    for (int k = 0; k < W; ++k)
        S[k] = re.first[k] & re.B[c][k];
    int64_t best = -1;
    int64_t pos = s + 1; // exclusive end of what has been consumed
    while (true)
    {
        uint64_t any = 0, acc = 0;
#pragma unroll
        for (int k = 0; k < W; ++k)
        {
            any |= S[k];
            acc |= S[k] & re.last[k];
        }
        if (!any)
            break;
        if (acc && (!re.anchor_end || pos == L))
            best = pos;
        if (pos >= L)
            break;
        c = at(pos);
        uint64_t F[W];
#pragma unroll
        for (int k = 0; k < W; ++k)
            F[k] = 0;
#pragma unroll
        for (int k = 0; k < W; ++k)
        {
            uint64_t sbits = S[k];
            while (sbits)
            {
                int i = __ffsll(static_cast<long long>(sbits)) - 1;
                sbits &= sbits - 1;
#pragma unroll
                for (int q = 0; q < W; ++q)
                    F[q] |= re.follow[k * 64 + i][q];
            }
        }
#pragma unroll
        for (int k = 0; k < W; ++k)
            S[k] = F[k] & re.B[c][k];
        ++pos;
    }
    return best;
}

// Anchored leftmost-first match (Perl / default RE2): Pike VM over the priority-ordered position lists.
// A thread = a position waiting to consume the next byte; list order = priority.  Reaching ACCEPT
// records the end and cuts every lower-priority thread; higher-priority threads run on and may
// still override it.
template <class At>
__device__ int64_t match_first(regex_dev const & re, At at, int64_t s, int64_t L)
{
    uint8_t cur[64 * kRegexWords], nxt[64 * kRegexWords];
    int ncur = re.first_len;
    for (int i = 0; i < ncur; ++i)
        cur[i] = re.first_list[i];
    int64_t best = -1;
    for (int64_t pos = s; pos < L && ncur > 0; ++pos)
    {
        uint8_t const c = at(pos);
        uint64_t in_next[kRegexWords] = {};
        int nn = 0;
        bool cut = false;
        for (int t = 0; t < ncur && !cut; ++t)
        {
            int const q = cur[t];
            if (!((re.B[c][q >> 6] >> (q & 63)) & 1))
                continue;
            for (int j = re.follow_off[q]; j < re.follow_off[q + 1]; ++j)
            {
                uint8_t const item = re.follow_list[j];
                if (item == kAccept)
                {
                    if (!re.anchor_end || pos + 1 == L)
                    {
                        best = pos + 1;
                        cut = true;
                        break;
                    }
                    continue;
                }
                if (!((in_next[item >> 6] >> (item & 63)) & 1))
                {
                    in_next[item >> 6] |= 1ULL << (item & 63);
                    nxt[nn++] = item;
                }
            }
        }
        for (int i = 0; i < nn; ++i)
            cur[i] = nxt[i];
        ncur = nn;
    }
    return best;
}

// One WARP per record that phase 1 listed (the list length is read from device memory, so no host round trip
// separates the two phases).  The 32 lanes try 32 consecutive start positions at once — almost all die on the
// first byte — and a ballot picks the leftmost start that really matches; the warp then jumps to the end of
// that match (FindAndConsume) and goes on.  Chunks phase 1 did not flag are skipped whole.
template <int W, int DIR>
__device__ __forceinline__ void phase2_body(resolve_params const & p, regex_dev & s_re)
{
    // usually a handful of records are listed: CTAs without work leave before staging anything
    if (static_cast<uint64_t>(blockIdx.x) * (blockDim.x >> 5) >= *p.rec_count)
        return;
    {
        uint32_t const * src = reinterpret_cast<uint32_t const *>(p.re);
        uint32_t * dst = reinterpret_cast<uint32_t *>(&s_re);
        for (unsigned i = threadIdx.x; i < sizeof(regex_dev) / 4; i += blockDim.x)
            dst[i] = src[i];
    }
    __syncthreads();
    regex_dev const & re = s_re;
    bool const anchor_begin = p.anchor_begin != 0;
    unsigned const lane = threadIdx.x & 31;
    uint64_t n_listed = *p.rec_count;
    if (p.rec_cap && n_listed > p.rec_cap)
        n_listed = 0; // the list overflowed: the host repeats this pattern on the single-pattern path, nothing is reported here
    uint64_t const warp0 = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    uint64_t const n_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
    for (uint64_t i = warp0; i < n_listed; i += n_warps)
    {
        uint64_t const r = p.rec_list[i];
        uint64_t const rb = p.rec_offsets[r], re_off = p.rec_offsets[r + 1];
        int64_t const L = static_cast<int64_t>(re_off - rb);
        auto at = [&](int64_t pos) -> uint8_t
        { return DIR > 0 ? p.text[rb + pos] : p.text[re_off - 1 - pos]; };
        auto addr_of = [&](int64_t pos) -> uint64_t
        { return DIR > 0 ? rb + pos : re_off - 1 - pos; };

        int64_t s = 0; // warp-uniform
        while (s < L)
        {
            uint64_t const ch = addr_of(s) / kChunkBytes;
            // first position (reading direction) that lies in the next chunk
            int64_t const next_chunk = DIR > 0 ? static_cast<int64_t>((ch + 1) * kChunkBytes - rb)
                                               : static_cast<int64_t>(re_off) - static_cast<int64_t>(ch * kChunkBytes);
            if (!anchor_begin && p.chunk_flags && !((p.chunk_flags[ch >> 5] >> (ch & 31)) & 1))
            {
                s = next_chunk; // no match starts anywhere in this chunk
                continue;
            }
            int64_t const limit = next_chunk < L ? next_chunk : L;
            int64_t const my_s = s + lane;
            int64_t e = -1;
            if (my_s < limit && (!anchor_begin || lane == 0))
            {
                uint8_t const c = at(my_s);
                uint64_t quick = 0;
#pragma unroll
                for (int k = 0; k < W; ++k)
                    quick |= re.first[k] & re.B[c][k];
                if (quick)
                    e = re.posix_longest ? match_longest<W>(re, at, my_s, L) : match_first(re, at, my_s, L);
            }
            unsigned const found = __ballot_sync(0xffffffffu, e > my_s);
            if (found)
            {
                int const leader = __ffs(found) - 1;
                int64_t const e_hit = __shfl_sync(0xffffffffu, e, leader);
                if (lane == 0)
                {
                    unsigned long long idx = atomicAdd(p.hit_count, 1ULL);
                    if (idx < p.cap)
                    {
                        uint64_t const key = (r << 32) | static_cast<uint64_t>(s + leader);
                        if (p.hit_pats)
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = static_cast<uint32_t>(e_hit);
                            p.hit_pats[idx] = p.pattern_id;
                        }
                        else if (idx < kSmallHits)
                        {
                            p.small_keys[idx] = key;
                            p.small_ends[idx] = static_cast<uint32_t>(e_hit);
                        }
                        else
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = static_cast<uint32_t>(e_hit);
                        }
                    }
                }
                s = e_hit; // FindAndConsume: the next search starts where this match ended
            }
            else
            {
                if (anchor_begin)
                    break; // '^' can only match at the resume point
                s = (s + 32 < limit) ? s + 32 : limit;
            }
        }
    }
}

// The DFA route of phase 2, written for instruction count (it is issue-bound: ~1 G table steps per 204 M start positions):
// the lanes of a warp try 32 consecutive starts, so the warp first turns the 48 bytes those starts can see in their first
// 16 steps into byte classes in a warp-private shared-memory window (every text byte is loaded and classified once, not
// once per lane and step), each lane pulls its 16 classes into registers with aligned word loads + funnel shifts, and a
// step is then: add the class to the (pre-multiplied) state, one shared-memory load, two compares.  The table holds
// next_state * classes, so there is no multiply on the chain; positions are 32-bit offsets into the record; the warp
// leaves the unrolled walk as soon as no lane is alive.  Walks that outlive 16 steps continue byte by byte.
template <int DIR, bool ANCHOR_END>
__device__ __forceinline__ void phase2_dfa_body(resolve_params const & p, uint16_t const * __restrict__ trans,
                                                uint8_t const * __restrict__ s_cls, uint32_t * __restrict__ win)
{
    unsigned const lane = threadIdx.x & 31;
    bool const anchor_begin = p.anchor_begin != 0;
    uint32_t const ncls = p.dfa_classes, accept_from = p.dfa_first_accept * ncls; // states are stored pre-multiplied
    uint64_t n_listed = *p.rec_count;
    if (p.rec_cap && n_listed > p.rec_cap)
        n_listed = 0; // the list overflowed: the host repeats this pattern on the single-pattern path, nothing is reported here
    uint64_t const warp0 = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    uint64_t const n_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
    uint8_t * const win8 = reinterpret_cast<uint8_t *>(win);
    unsigned const w0 = lane >> 2, sh = (lane & 3) * 8;
    for (uint64_t i = warp0; i < n_listed; i += n_warps)
    {
        uint64_t const r = p.rec_list[i];
        uint64_t const rb = p.rec_offsets[r], re_off = p.rec_offsets[r + 1];
        uint32_t const L = static_cast<uint32_t>(re_off - rb);
        uint8_t const * const base = DIR > 0 ? p.text + rb : p.text + re_off - 1;
        auto at = [&](uint32_t pos) -> uint8_t { return DIR > 0 ? __ldg(base + pos) : __ldg(base - pos); };
        uint32_t s = 0; // warp-uniform
        while (s < L)
        {
            uint32_t limit = L;
            if (p.chunk_flags)
            {
                uint64_t const ch = (DIR > 0 ? rb + s : re_off - 1 - s) / kChunkBytes;
                uint32_t const next_chunk = DIR > 0 ? static_cast<uint32_t>((ch + 1) * kChunkBytes - rb)
                                                    : static_cast<uint32_t>(re_off - ch * kChunkBytes);
                if (!anchor_begin && !((p.chunk_flags[ch >> 5] >> (ch & 31)) & 1))
                {
                    s = next_chunk; // no match starts anywhere in this chunk
                    continue;
                }
                limit = next_chunk < L ? next_chunk : L;
            }
            if (p.start_bits) // exact starts are known (regex_scan_starts_dfa): most windows of 32 have none
            {
                uint64_t const g = (DIR > 0 ? rb + s : re_off - 1 - s) + (DIR > 0 ? lane : 0) - (DIR > 0 ? 0 : lane);
                bool const bit = s + lane < limit && ((__ldg(p.start_bits + (g >> 5)) >> (g & 31)) & 1);
                unsigned const any = __ballot_sync(0xffffffffu, bit);
                if (!any)
                {
                    s = (s + 32 < limit) ? s + 32 : limit;
                    continue;
                }
                s += __ffs(any) - 1; // the leftmost start of this window becomes lane 0: it is the next match
            }
            // classes of positions [s, s + 48); beyond the record: class 0, which every state leaves for the dead state
            uint32_t const my_cl = s + lane < L ? s_cls[at(s + lane)] : 0u;
            // most windows of 32 hold no byte a match can start with: one lookup and one ballot settle those
            if (!__ballot_sync(0xffffffffu, s + lane < limit && trans[ncls + my_cl] != 0))
            {
                if (anchor_begin)
                    break;
                s = (s + 32 < limit) ? s + 32 : limit;
                continue;
            }
            __syncwarp();
            win8[lane] = static_cast<uint8_t>(my_cl);
            if (lane < 16)
                win8[32 + lane] = s + 32 + lane < L ? s_cls[at(s + 32 + lane)] : 0;
            __syncwarp();
            uint32_t const my_s = s + lane;
            uint32_t e = 0; // 0 = no match (a match is never empty)
            // with exact starts lane 0 sits on the leftmost one: its longest match is the next hit, nobody else needs to walk
            bool const active = my_s < limit && ((!anchor_begin && !p.start_bits) || lane == 0);
            uint32_t st = active ? ncls : 0; // state 1, pre-multiplied
            {
                uint32_t const a = win[w0], b = win[w0 + 1], c2 = win[w0 + 2], d = win[w0 + 3], f = win[w0 + 4];
                uint32_t const reg[4] = {__funnelshift_r(a, b, sh), __funnelshift_r(b, c2, sh), __funnelshift_r(c2, d, sh),
                                         __funnelshift_r(d, f, sh)};
#pragma unroll
                for (int j = 0; j < 16; ++j)
                {
                    if ((j & 3) == 0 && j > 0 && !__any_sync(0xffffffffu, st != 0))
                        break;
                    if (st != 0)
                    {
                        st = trans[st + ((reg[j >> 2] >> ((j & 3) * 8)) & 0xFF)];
                        if (st >= accept_from && (!ANCHOR_END || my_s + j + 1 == L))
                            e = my_s + j + 1;
                    }
                }
                uint32_t pos = my_s + 16;
                while (st != 0 && pos < L) // long walks: byte by byte
                {
                    st = trans[st + s_cls[at(pos)]];
                    ++pos;
                    if (st >= accept_from && (!ANCHOR_END || pos == L))
                        e = pos;
                }
            }
            unsigned const found = __ballot_sync(0xffffffffu, e != 0);
            if (found)
            {
                int const leader = __ffs(found) - 1;
                uint32_t const e_hit = __shfl_sync(0xffffffffu, e, leader);
                if (lane == 0)
                {
                    unsigned long long idx = atomicAdd(p.hit_count, 1ULL);
                    if (idx < p.cap)
                    {
                        uint64_t const key = (r << 32) | static_cast<uint64_t>(s + leader);
                        if (p.hit_pats)
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = e_hit;
                            p.hit_pats[idx] = p.pattern_id;
                        }
                        else if (idx < kSmallHits)
                        {
                            p.small_keys[idx] = key;
                            p.small_ends[idx] = e_hit;
                        }
                        else
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = e_hit;
                        }
                    }
                }
                s = e_hit; // FindAndConsume: the next search starts where this match ended
            }
            else
            {
                if (anchor_begin)
                    break; // '^' can only match at the resume point
                s = (s + 32 < limit) ? s + 32 : limit;
            }
        }
    }
}

// Phase 2 when phase 1 delivered the EXACT starts (regex_scan_starts_dfa; forward strand, no anchors): a warp looks at
// 1024 start bits of its record per step (one word per lane, one ballot), jumps to the first start at or behind the
// resume point, lane 0 walks the matcher from there — that start IS the next match — and the search resumes at its end.
__device__ __forceinline__ void phase2_starts_body(resolve_params const & p, uint16_t const * __restrict__ trans,
                                                   uint8_t const * __restrict__ s_cls)
{
    unsigned const lane = threadIdx.x & 31;
    uint32_t const ncls = p.dfa_classes, accept_from = p.dfa_first_accept * ncls;
    uint64_t n_listed = *p.rec_count;
    if (p.rec_cap && n_listed > p.rec_cap)
        n_listed = 0; // the list overflowed: the host repeats this pattern on the single-pattern path, nothing is reported here
    uint64_t const warp0 = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    uint64_t const n_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
    for (uint64_t i = warp0; i < n_listed; i += n_warps)
    {
        uint64_t const r = p.rec_list[i];
        uint64_t const rb = p.rec_offsets[r], re_off = p.rec_offsets[r + 1];
        uint64_t g = rb; // resume point (global byte index), warp-uniform
        while (g < re_off)
        {
            uint64_t const w = (g >> 5) + lane;
            uint32_t bits = w * 32 < re_off ? __ldg(p.start_bits + w) : 0;
            if (lane == 0)
                bits &= ~0u << (g & 31); // starts before the resume point lie inside the previous match
            if (w * 32 + 32 > re_off && w * 32 < re_off)
                bits &= (re_off & 31) ? ~(~0u << (re_off & 31)) : ~0u; // the next record's starts share this word
            unsigned const any = __ballot_sync(0xffffffffu, bits != 0);
            if (!any)
            {
                g = ((g >> 5) + 32) << 5;
                continue;
            }
            int const src = __ffs(any) - 1;
            uint32_t const word = __shfl_sync(0xffffffffu, bits, src);
            uint64_t const q = (((g >> 5) + src) << 5) + (__ffs(word) - 1); // the next match starts here
            // the walk: every lane fetches and classifies ONE byte of the 32 behind q (one memory latency for the whole
            // match instead of one per byte), then all lanes run the same chain on classes handed round by shuffles
            uint32_t e = 0;
            {
                uint32_t const my_cl = q + lane < re_off ? s_cls[__ldg(p.text + q + lane)] : 0u; // class 0: every state dies
                uint32_t st = ncls;
#pragma unroll 4
                for (int j = 0; j < 32; ++j)
                {
                    st = trans[st + __shfl_sync(0xffffffffu, my_cl, j)];
                    if (st == 0)
                        break; // warp-uniform
                    if (st >= accept_from)
                        e = static_cast<uint32_t>(q - rb) + j + 1;
                }
                if (st != 0) // a match longer than 32 bytes: byte by byte
                {
                    uint64_t pos = q + 32;
                    while (pos < re_off)
                    {
                        st = trans[st + s_cls[__ldg(p.text + pos)]];
                        if (st == 0)
                            break;
                        ++pos;
                        if (st >= accept_from)
                            e = static_cast<uint32_t>(pos - rb);
                    }
                }
                if (lane == 0 && e != 0)
                {
                    unsigned long long idx = atomicAdd(p.hit_count, 1ULL);
                    if (idx < p.cap)
                    {
                        uint64_t const key = (r << 32) | (q - rb);
                        if (p.hit_pats)
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = e;
                            p.hit_pats[idx] = p.pattern_id;
                        }
                        else if (idx < kSmallHits)
                        {
                            p.small_keys[idx] = key;
                            p.small_ends[idx] = e;
                        }
                        else
                        {
                            p.hit_keys[idx] = key;
                            p.hit_ends[idx] = e;
                        }
                    }
                }
            }
            g = e != 0 ? rb + e : q + 1; // (e == 0 cannot happen for an exact start; stay safe and move on)
        }
    }
}

// dynamic shared memory: [256 class bytes | 16 window words per warp (4 warps) | transition table if <= kDfaSmemBytes]
constexpr unsigned kDfaSmemBytes = 32 * 1024;
constexpr unsigned kDfaSmemFixed = 256 + 4 * 64;

__host__ __device__ inline unsigned dfa_smem_bytes(unsigned states, unsigned classes)
{
    unsigned const table = states * classes * 2;
    return kDfaSmemFixed + (table <= kDfaSmemBytes ? ((table + 15) & ~15u) : 0);
}

template <int DIR>
__device__ __forceinline__ void phase2_dfa_block(resolve_params const & p, uint8_t * smem)
{
    if (static_cast<uint64_t>(blockIdx.x) * (blockDim.x >> 5) >= *p.rec_count)
        return;
    unsigned const entries = p.dfa_states * p.dfa_classes;
    bool const staged = entries * 2 <= kDfaSmemBytes;
    uint8_t * const s_cls = smem;
    uint32_t * const s_win_all = reinterpret_cast<uint32_t *>(smem + 256);
    uint16_t * const s_trans = reinterpret_cast<uint16_t *>(smem + kDfaSmemFixed);
    for (unsigned i = threadIdx.x; i < 256; i += blockDim.x)
        s_cls[i] = p.dfa_cls[i];
    if (staged)
        for (unsigned i = threadIdx.x; i < entries; i += blockDim.x)
            s_trans[i] = p.dfa_trans[i];
    if (threadIdx.x < 64)
        s_win_all[threadIdx.x] = 0;
    __syncthreads();
    if (p.start_bits && DIR > 0)
        phase2_starts_body(p, staged ? s_trans : p.dfa_trans, s_cls);
    else if (p.anchor_end)
        phase2_dfa_body<DIR, true>(p, staged ? s_trans : p.dfa_trans, s_cls, s_win_all + (threadIdx.x >> 5) * 16);
    else
        phase2_dfa_body<DIR, false>(p, staged ? s_trans : p.dfa_trans, s_cls, s_win_all + (threadIdx.x >> 5) * 16);
}

template <int DIR>
__global__ void __launch_bounds__(128) regex_scan_phase2_dfa(resolve_params p)
{
    extern __shared__ __align__(16) uint8_t dfa_smem[];
    phase2_dfa_block<DIR>(p, dfa_smem);
}

template <int W, int DIR>
__global__ void __launch_bounds__(128) regex_scan_phase2(resolve_params p)
{
    __shared__ regex_dev s_re; // the automaton, staged once per CTA: every attempt starts with a B[] lookup
    phase2_body<W, DIR>(p, s_re);
}

// Patterns without any selective necessary factor and with a determinised matcher skip phase 1: every record of a
// candidate bin is resolved directly (one table lookup per byte and start position is cheaper than streaming the text
// through the position automaton first).
__global__ void __launch_bounds__(256)
    list_candidate_records_kernel(uint32_t const * __restrict__ rec_bin, uint64_t const * __restrict__ cand_words, uint64_t n_records,
                                  uint32_t * __restrict__ rec_list, unsigned int * __restrict__ rec_count)
{
    for (uint64_t r = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r < n_records;
         r += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        bool const take = is_candidate(cand_words, __ldg(rec_bin + r));
        unsigned const m = __ballot_sync(__activemask(), take);
        if (!m)
            continue;
        unsigned const lane = threadIdx.x & 31;
        unsigned const leader = __ffs(m) - 1;
        unsigned base = 0;
        if (lane == leader)
            base = atomicAdd(rec_count, __popc(m));
        base = __shfl_sync(__activemask(), base, leader);
        if (take)
            rec_list[base + __popc(m & ((1u << lane) - 1))] = static_cast<uint32_t>(r);
    }
}

// sorted (record << 32 | start, end) pairs → the trx_hit records the caller receives: one copy instead of a host loop
__global__ void __launch_bounds__(256)
    format_hits_kernel(uint64_t const * __restrict__ keys, uint32_t const * __restrict__ ends, uint32_t const * __restrict__ rec_bin,
                       uint64_t n, uint4 * __restrict__ out)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint64_t const k = keys[i];
        uint32_t const rec = static_cast<uint32_t>(k >> 32);
        out[i] = make_uint4(__ldg(rec_bin + rec), rec, static_cast<uint32_t>(k), ends[i]); // {bin, record, start, end}
    }
}

// trx_regex_scan_multi's closing sort: (end, pattern) travel as one 64-bit value through the sort by key; the sort by
// pattern then moves 32-bit indices only, and the formatting kernel gathers through them
__global__ void __launch_bounds__(256)
    multi_pack_kernel(uint32_t const * __restrict__ ends, uint32_t const * __restrict__ pats, uint64_t n, uint64_t * __restrict__ vals)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
        vals[i] = static_cast<uint64_t>(ends[i]) | (static_cast<uint64_t>(pats[i]) << 32);
}

__global__ void __launch_bounds__(256)
    multi_unpack_kernel(uint64_t const * __restrict__ vals, uint64_t n, uint32_t * __restrict__ pat, uint32_t * __restrict__ idx)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        pat[i] = static_cast<uint32_t>(vals[i] >> 32);
        idx[i] = static_cast<uint32_t>(i);
    }
}

__global__ void __launch_bounds__(256)
    multi_format_kernel(uint64_t const * __restrict__ keys, uint64_t const * __restrict__ vals, uint32_t const * __restrict__ idx,
                        uint32_t const * __restrict__ rec_bin, uint64_t n, uint4 * __restrict__ out)
{
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint32_t const j = idx[i];
        uint64_t const k = keys[j];
        uint32_t const rec = static_cast<uint32_t>(k >> 32);
        out[i] = make_uint4(__ldg(rec_bin + rec), rec, static_cast<uint32_t>(k), static_cast<uint32_t>(vals[j])); // {bin, record, start, end}
    }
}

// one launch for a batch of patterns: blockIdx.y = pattern; its CTAs take that pattern's automaton and work list
template <int W, int DIR>
__global__ void __launch_bounds__(128) regex_scan_phase2_multi(resolve_params const * __restrict__ params)
{
    __shared__ regex_dev s_re;
    __shared__ resolve_params s_p;
    extern __shared__ __align__(16) uint8_t dfa_smem[]; // sized by the host for the largest DFA of the batch
    if (threadIdx.x == 0)
        s_p = params[blockIdx.y];
    __syncthreads();
    if (s_p.dfa_trans)
        phase2_dfa_block<DIR>(s_p, dfa_smem);
    else
        phase2_body<W, DIR>(s_p, s_re);
}

} // namespace trx

namespace trx
{

// ------------------------------------------------------------------------------------------------------
// Phase 1, fast variant: SIMD-within-register necessary-factor filter.
//
// A byte-at-a-time automaton costs >= 9 thread instructions per residue, which caps a B200 at ~40 % of its
// HBM bandwidth.  Phase 1 only has to produce a SUPERSET of the places where a match can start (phase 2 is
// exact), so when every match carries NL literal bytes at fixed distances (filter_spec), the filter tests
// four text positions per 32-bit operation:
//      u  = (w ^ X0) | (shift(w, d1) ^ X1) | …        byte lane = 0  ⇔  all NL literals present there
//      z  = (u - 0x01010101) & ~u & 0x80808080        ≠ 0  ⇔  some lane of u is 0
// ≈ 2 instructions per residue for 3 literals, ≈ 5 for 8.  Record boundaries, candidate bins and reading direction do not
// matter here: a factor that straddles two records, or lies in a non-candidate bin that shares a tile, only
// creates a false candidate that phase 2 never visits or rejects.
// Shared skeleton of the two prefilter kernels: a warp-specialised ring of 32 KB tiles (256 rows x 128 B, one 2-D TMA
// copy each).  The LAST warp of the CTA is the producer: one lane takes the next tile from a device-wide counter
// (dynamic scheduling — with static striding the SMs were idle for 29 % of the kernel: ramp, 14-vs-15-tile imbalance
// and drain), waits until every consumer warp has released the stage (`empty` mbarrier, one arrival per warp) and
// issues the TMA copy, which completes on the stage's `full` mbarrier.  Consumer warps never meet at a CTA-wide
// barrier: ncu showed 3.1 warps per issue slot parked at the closing __syncthreads of the previous design.
// SPLIT threads share a 128-byte row, so a CTA has 256 * SPLIT consumer threads + 32.  Ring depth 2: three and four
// stages with fewer CTAs per SM were measured 20-30 % slower.
constexpr uint32_t kNoTile = 0xFFFFFFFFu;

__device__ __forceinline__ void mbar_arrive(uint64_t * bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int STAGES, int CONSUMERS>
struct tile_ring
{
    static constexpr int kRingBytes = STAGES * kTileBytes + STAGES * (2 * sizeof(uint64_t) + sizeof(uint32_t));
    uint8_t * tiles;
    uint64_t * full;
    uint64_t * empty;
    uint32_t * slot; // tile id held by each stage (kNoTile = no more work)
    uint32_t n{0};

    __device__ __forceinline__ void init(uint8_t * smem)
    {
        tiles = smem;
        full = reinterpret_cast<uint64_t *>(smem + STAGES * kTileBytes);
        empty = full + STAGES;
        slot = reinterpret_cast<uint32_t *>(empty + STAGES);
        if (threadIdx.x == 0)
        {
#pragma unroll
            for (int st = 0; st < STAGES; ++st)
            {
                mbar_init(&full[st], 1);
                mbar_init(&empty[st], CONSUMERS / 32);
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
    __device__ __forceinline__ bool is_producer() const { return threadIdx.x >= CONSUMERS; }

    // producer warp, one lane: feeds the ring until the device-wide tile counter runs out
    __device__ __forceinline__ void produce(scan_params const & p, CUtensorMap const * tmap)
    {
        if (threadIdx.x != CONSUMERS)
            return;
        for (uint32_t i = 0;; ++i)
        {
            uint32_t const stage = i % STAGES;
            if (i >= STAGES)
                mbar_wait(&empty[stage], ((i / STAGES) - 1) & 1);
            uint32_t const idx = atomicAdd(p.tile_counter, 1u);
            if (idx >= p.n_tiles)
            {
                slot[stage] = kNoTile;
                mbar_arrive(&full[stage]); // release: the slot write is visible to whoever sees this phase complete
                return;
            }
            uint32_t const tile = __ldg(p.tiles + idx);
            slot[stage] = tile;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tma_load_tile(tmap, tiles + stage * kTileBytes, &full[stage], static_cast<int32_t>(tile) * kScanThreads);
        }
    }
    // consumers: next tile (false = done).  release() hands the stage back, one arrival per warp.
    __device__ __forceinline__ bool acquire(uint8_t const *& tile_ptr, uint32_t & tile_id)
    {
        uint32_t const stage = n % STAGES;
        mbar_wait(&full[stage], (n / STAGES) & 1);
        tile_id = slot[stage];
        tile_ptr = tiles + stage * kTileBytes;
        return tile_id != kNoTile;
    }
    __device__ __forceinline__ void release()
    {
        __syncwarp();
        if ((threadIdx.x & 31) == 0)
            mbar_arrive(&empty[n % STAGES]);
        ++n;
    }
};

template <int NA, int NB, int STAGES, int SPLIT>
__global__ void __launch_bounds__(kScanThreads * SPLIT + 32)
    regex_filter_simd(scan_params p, const __grid_constant__ CUtensorMap tmap)
{
    constexpr int NL = NA + NB;
    constexpr int BYTES = kChunkBytes / SPLIT; // text bytes per thread
    constexpr int WORDS = BYTES / 4, PIECES = BYTES / 16;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    tile_ring<STAGES, kScanThreads * SPLIT> ring;
    ring.init(smem_raw);
    if (ring.is_producer())
    {
        ring.produce(p, &tmap);
        return;
    }
    int const tid = threadIdx.x;
    int const row_i = tid / SPLIT, sub = tid % SPLIT;

    uint32_t X[NL > 0 ? NL : 1], D[NL > 0 ? NL : 1];
#pragma unroll
    for (int j = 0; j < NL; ++j)
    {
        X[j] = p.filter.splat[j];
        D[j] = p.filter.delta[j] * 8;
    }

    uint8_t const * tile;
    uint32_t tile_id;
    while (ring.acquire(tile, tile_id))
    {
        uint64_t const g_lo = (static_cast<uint64_t>(tile_id) * kScanThreads + row_i) * kChunkBytes + sub * BYTES;
        uint8_t const * row = tile + row_i * kChunkBytes;
        uint32_t w[WORDS + 2];
#pragma unroll
        for (int piece = 0; piece < PIECES; ++piece)
        {
            uint4 v = *reinterpret_cast<uint4 const *>(row + (((sub * PIECES + piece) ^ (row_i & 7)) << 4));
            w[4 * piece + 0] = v.x;
            w[4 * piece + 1] = v.y;
            w[4 * piece + 2] = v.z;
            w[4 * piece + 3] = v.w;
        }
        // look-ahead = the 8 bytes that follow: same row, the next row of the tile, or global memory
        if (sub + 1 < SPLIT)
        {
            uint2 la = *reinterpret_cast<uint2 const *>(row + ((((sub + 1) * PIECES) ^ (row_i & 7)) << 4));
            w[WORDS] = la.x;
            w[WORDS + 1] = la.y;
        }
        else if (row_i + 1 < kScanThreads)
        {
            uint2 la = *reinterpret_cast<uint2 const *>(row + kChunkBytes + ((0 ^ ((row_i + 1) & 7)) << 4));
            w[WORDS] = la.x;
            w[WORDS + 1] = la.y;
        }
        else if (g_lo + BYTES + 8 <= p.n_text_padded)
        {
            uint2 la = __ldg(reinterpret_cast<uint2 const *>(p.text + g_lo + BYTES));
            w[WORDS] = la.x;
            w[WORDS + 1] = la.y;
        }
        else
        {
            w[WORDS] = w[WORDS + 1] = 0;
        }
        uint32_t piece_hits = 0; // bit i = piece i holds a candidate
#pragma unroll
        for (int piece = 0; piece < PIECES; ++piece)
        {
            uint32_t acc = 0;
#pragma unroll
            for (int q = 4 * piece; q < 4 * piece + 4; ++q)
            {
                uint32_t u = 0;
#pragma unroll
                for (int j = 0; j < NL; ++j)
                {
                    int const base = q + (j >= NA ? 1 : 0); // second half of the window = next word
                    u |= __funnelshift_r(w[base], w[base + 1], D[j]) ^ X[j];
                }
                acc |= (u - 0x01010101u) & ~u & 0x80808080u;
            }
            if (acc)
                piece_hits |= 1u << piece;
        }
        if (piece_hits && g_lo < p.n_text)
        {
            // rare path: translate candidate pieces into chunk flags and listed records
            while (piece_hits)
            {
                int const piece = __ffs(piece_hits) - 1;
                piece_hits &= piece_hits - 1;
                int64_t lo = static_cast<int64_t>(g_lo) + piece * 16 + p.filter.start_delta;
                int64_t hi = lo + 15;
                if (hi < 0 || lo >= static_cast<int64_t>(p.n_text))
                    continue;
                if (lo < 0)
                    lo = 0;
                if (hi >= static_cast<int64_t>(p.n_text))
                    hi = static_cast<int64_t>(p.n_text) - 1;
                uint64_t const c0 = static_cast<uint64_t>(lo) / kChunkBytes, c1 = static_cast<uint64_t>(hi) / kChunkBytes;
                atomicOr(p.chunk_flags + (c0 >> 5), 1u << (c0 & 31));
                if (c1 != c0)
                    atomicOr(p.chunk_flags + (c1 >> 5), 1u << (c1 & 31));
                uint64_t r = record_of(p.rec_offsets, p.n_records, static_cast<uint64_t>(lo));
                while (r < p.n_records && __ldg(p.rec_offsets + r) <= static_cast<uint64_t>(hi))
                {
                    if (__ldg(p.rec_offsets + r + 1) > static_cast<uint64_t>(lo) &&
                        is_candidate(p.cand_words, __ldg(p.rec_bin + r)))
                        list_record(p, r);
                    ++r;
                }
            }
        }
        ring.release();
    }
}

// ------------------------------------------------------------------------------------------------------
// Phase 1, gapped variant: literal bit masks + shifted mask algebra.
//
// PROSITE-style patterns (C-x(2,4)-C-x(3)-[LIVMFYWC]-x(8)-H-x(3,5)-H) have no window of fixed-offset literals,
// but they do have literals that EVERY match passes through, at distances that vary inside small ranges.  The
// thread turns its 128-byte chunk into one 128-bit occurrence mask per literal — per 32-bit text word: XOR with
// the splat, zero-byte test, and one multiply that gathers the four lane flags into a nibble — fetches the mask
// of the following 32 bytes from its neighbour, and evaluates
//      cand = AND over atoms ( OR over d in [lo, hi] ( mask[lit] >> d ) )
// with funnel shifts on 32-bit pieces.  A surviving bit is the anchor of a possible match; the chunks that can
// hold the match's first byte are flagged and the records they belong to are listed for phase 2, exactly like
// the fixed-window filter does.  Cost ≈ 1.25 ALU instructions per residue and literal, independent of the
// number of atoms and of the pattern's length.
__device__ __forceinline__ uint32_t literal_nibble(uint32_t word, uint32_t splat)
{
    uint32_t const t = word ^ splat;
    uint32_t z; // (t - 0x01010101) & ~t & 0x80808080: bit 7 of a lane set if the lane (or a lower one) is 0.  One LOP3
                // with the mask as immediate — left to itself ptxas folds the XOR in again and spends a fourth LOP3.
    asm("lop3.b32 %0, %1, %2, 0x80808080, 0x20;" : "=r"(z) : "r"(t - 0x01010101u), "r"(t));
    return (z * 0x00204081u) >> 28; // lane flags 7,15,23,31 → bits 0..3
}

template <int NLIT, int STAGES, int SPLIT>
__global__ void __launch_bounds__(kScanThreads * SPLIT + 32)
    regex_filter_atoms(scan_params p, const __grid_constant__ CUtensorMap tmap)
{
    static_assert(SPLIT == 1 || SPLIT == 2, "a thread owns 128 or 64 bytes (256 * SPLIT + 32 threads per CTA)");
    constexpr int THREADS = kScanThreads * SPLIT; // consumer threads
    constexpr int BYTES = kChunkBytes / SPLIT;
    constexpr int MW = BYTES / 32; // mask words per literal (one per 32 text bytes)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using ring_t = tile_ring<STAGES, THREADS>;
    ring_t ring;
    ring.init(smem_raw);
    if (ring.is_producer())
    {
        ring.produce(p, &tmap);
        return;
    }
    // [2][NLIT][THREADS]: mask of each thread's first 32 bytes, double-buffered by tile parity (no closing barrier)
    uint32_t * halo_base = reinterpret_cast<uint32_t *>(smem_raw + ((ring_t::kRingBytes + 15) & ~15));
    int const tid = threadIdx.x;
    int const row_i = tid / SPLIT, sub = tid % SPLIT;

    uint32_t X[NLIT];
#pragma unroll
    for (int l = 0; l < NLIT; ++l)
        X[l] = p.atoms.splat[l];

    uint8_t const * tile;
    uint32_t tile_id;
    while (ring.acquire(tile, tile_id))
    {
        uint32_t * halo = halo_base + (ring.n & 1) * (NLIT * THREADS);
        uint64_t const g_lo = (static_cast<uint64_t>(tile_id) * kScanThreads + row_i) * kChunkBytes + sub * BYTES;
        uint8_t const * row = tile + row_i * kChunkBytes;
        uint32_t m[NLIT][MW + 1];
#pragma unroll
        for (int q = 0; q < MW; ++q) // 32 bytes of text → one mask word per literal
        {
            int const piece = sub * (BYTES / 16) + 2 * q;
            uint4 const v0 = *reinterpret_cast<uint4 const *>(row + ((piece ^ (row_i & 7)) << 4));
            uint4 const v1 = *reinterpret_cast<uint4 const *>(row + (((piece + 1) ^ (row_i & 7)) << 4));
            uint32_t const w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
            for (int l = 0; l < NLIT; ++l)
            {
                uint32_t mm = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mm |= literal_nibble(w[j], X[l]) << (4 * j);
                m[l][q] = mm;
            }
        }
#pragma unroll
        for (int l = 0; l < NLIT; ++l)
            halo[l * THREADS + tid] = m[l][0];
        asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory"); // consumers only: the producer warp is elsewhere
        if (tid + 1 < THREADS)
        {
#pragma unroll
            for (int l = 0; l < NLIT; ++l)
                m[l][MW] = halo[l * THREADS + tid + 1];
        }
        else if (g_lo + BYTES + 32 <= p.n_text_padded)
        {
            // last thread of the tile: the following 32 bytes belong to another tile — read them from global memory
            uint4 const v0 = __ldg(reinterpret_cast<uint4 const *>(p.text + g_lo + BYTES));
            uint4 const v1 = __ldg(reinterpret_cast<uint4 const *>(p.text + g_lo + BYTES + 16));
            uint32_t const w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
            for (int l = 0; l < NLIT; ++l)
            {
                uint32_t mm = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mm |= literal_nibble(w[j], X[l]) << (4 * j);
                m[l][MW] = mm;
            }
        }
        else
        {
#pragma unroll
            for (int l = 0; l < NLIT; ++l)
                m[l][MW] = 0;
        }

        // cand = AND over atoms ( OR over the atom's bytes and offsets ( mask[byte] >> d ) ).  The term loop is uniform
        // across the CTA; the literal index picks a register array, hence the if-chain instead of an indexed access.
        uint32_t acc[MW], r[MW];
#pragma unroll
        for (int q = 0; q < MW; ++q)
        {
            acc[q] = ~0u;
            r[q] = 0;
        }
        for (int i = 0; i < p.atoms.n_terms; ++i)
        {
            uint32_t const lit = p.atoms.term_lit[i], d_lo = p.atoms.term_lo[i], d_hi = p.atoms.term_hi[i];
#pragma unroll
            for (int l = 0; l < NLIT; ++l)
            {
                if (lit == static_cast<uint32_t>(l))
                {
                    for (uint32_t d = d_lo; d <= d_hi; ++d)
                    {
#pragma unroll
                        for (int q = 0; q < MW; ++q)
                            r[q] |= __funnelshift_r(m[l][q], m[l][q + 1], d);
                    }
                }
            }
            if (p.atoms.term_end[i])
            {
#pragma unroll
                for (int q = 0; q < MW; ++q)
                {
                    acc[q] &= r[q];
                    r[q] = 0;
                }
            }
        }
        uint32_t any = 0;
#pragma unroll
        for (int q = 0; q < MW; ++q)
            any |= acc[q];
        if (any && g_lo < p.n_text)
        {
            // rare path: every surviving bit is an anchor; flag the chunks its match can start in, list their records
#pragma unroll
            for (int q = 0; q < MW; ++q)
            {
                uint32_t bits = acc[q];
                while (bits)
                {
                    int const b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    int64_t const anchor = static_cast<int64_t>(g_lo) + 32 * q + b;
                    // second stage: the remaining necessary positions, straight from the text (L2: the tile just went by)
                    bool confirmed = true;
                    for (int c = 0; c < p.atoms.n_confirm && confirmed; ++c)
                    {
                        bool found = false;
                        for (int d = p.atoms.confirm[c].lo; d <= p.atoms.confirm[c].hi && !found; ++d)
                        {
                            int64_t const a = anchor + d;
                            if (a >= 0 && a < static_cast<int64_t>(p.n_text))
                            {
                                uint32_t const byte = __ldg(p.text + a);
                                found = (p.atoms.confirm[c].mask[byte >> 5] >> (byte & 31)) & 1;
                            }
                        }
                        confirmed = found;
                    }
                    if (!confirmed)
                        continue;
                    int64_t lo = anchor + p.atoms.start_lo, hi = anchor + p.atoms.start_hi;
                    if (hi < 0 || lo >= static_cast<int64_t>(p.n_text))
                        continue;
                    if (lo < 0)
                        lo = 0;
                    if (hi >= static_cast<int64_t>(p.n_text))
                        hi = static_cast<int64_t>(p.n_text) - 1;
                    for (uint64_t c = static_cast<uint64_t>(lo) / kChunkBytes; c <= static_cast<uint64_t>(hi) / kChunkBytes; ++c)
                        atomicOr(p.chunk_flags + (c >> 5), 1u << (c & 31));
                    uint64_t r = record_of(p.rec_offsets, p.n_records, static_cast<uint64_t>(lo));
                    while (r < p.n_records && __ldg(p.rec_offsets + r) <= static_cast<uint64_t>(hi))
                    {
                        if (__ldg(p.rec_offsets + r + 1) > static_cast<uint64_t>(lo) &&
                            is_candidate(p.cand_words, __ldg(p.rec_bin + r)))
                            list_record(p, r);
                        ++r;
                    }
                }
            }
        }
        ring.release();
    }
}

// byte histogram of the residue store (once per upload): the host uses it to pick the rarest literals
__global__ void __launch_bounds__(256) byte_histogram_kernel(uint8_t const * __restrict__ text, uint64_t n,
                                                             unsigned long long * __restrict__ hist)
{
    __shared__ unsigned int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t const n16 = n / 16;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n16;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        uint4 v = __ldg(reinterpret_cast<uint4 const *>(text) + i);
        uint32_t const ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int b = 0; b < 4; ++b)
                atomicAdd(&h[(ws[k] >> (8 * b)) & 0xFF], 1u);
    }
    if (blockIdx.x == 0)
        for (uint64_t i = n16 * 16 + threadIdx.x; i < n; i += blockDim.x)
            atomicAdd(&h[text[i]], 1u);
    __syncthreads();
    if (h[threadIdx.x])
        atomicAdd(hist + threadIdx.x, static_cast<unsigned long long>(h[threadIdx.x]));
}

// ------------------------------------------------------------------------------------------------------
// Phase 1 for a BATCH of patterns: one pass over the text for all of them (run_multiple_queries,
// include/query.h:330-340, loops over a file of motifs and re-reads every candidate bin per motif; the pattern-at-a-time
// kernels above do the same with HBM).
//
// What is shared between patterns is everything that depends on the text only: the TMA tile, and the OCCURRENCE MASKS —
// for every byte of the batch's alphabet (the union of the literals its prefilters test, <= kMultiLetters) one bit per
// text position.  A thread owns a 128-byte row, builds its 4 mask words per letter once per tile and parks them in
// shared memory, plane-wise (conflict-free).  After that a pattern's necessary condition — the same window / gapped-atom
// plans the single-pattern kernels use, as one list of terms "letter l at anchor + d, d in [lo, hi]", atoms ANDed, terms
// of an atom ORed — costs 5 LDS + a few funnel shifts per term and 128 anchors: ~0.5 instructions per residue and
// pattern instead of 2.75-3.75, and no HBM traffic at all after the first pattern.  Survivors are confirmed (classes,
// signed offsets: the plan's second stage) and listed per pattern for regex_scan_phase2_multi.
constexpr int kMultiPatterns = 128; // per launch
constexpr uint32_t kMultiListCap = 32768; // listed records per pattern; more = a pattern for the single-pattern path

struct multi_terms // the part every thread walks for every tile: kept in shared memory
{
    int32_t n_terms;
    uint8_t letter[kMaxAtomTerms], lo[kMaxAtomTerms], hi[kMaxAtomTerms], end[kMaxAtomTerms];
};

struct multi_program // per pattern, in global memory; the rest is only touched where an anchor survives
{
    multi_terms terms;
    int32_t start_lo, start_hi; // a match's first byte lies in [anchor + start_lo, anchor + start_hi]
    int32_t n_confirm;
    confirm_check confirm[kMaxConfirm];
    uint64_t const * cand_words; // candidate bins of THIS pattern
    uint32_t * chunk_flags;      // bitmap over 128-byte chunks: a match of this pattern may start here (phase 2 skips the rest)
    uint32_t * rec_flags;        // bitmap over records already listed for this pattern
    uint32_t * rec_list;         // kMultiListCap entries
    unsigned int * rec_count;
};

struct multi_params
{
    multi_program const * programs;
    uint32_t n_patterns;
    uint32_t n_letters;
    uint32_t splat[kMultiLetters]; // letter byte replicated into the four lanes
    uint32_t const * terms;        // the batch's terms, packed (multi_pack_term), pattern after pattern
    uint32_t n_terms_total;
};

// rare path of regex_filter_multi: every surviving bit is an anchor of the pattern; confirm, flag its chunks, list its records
__device__ __noinline__ void multi_survivors(scan_params const & p, multi_program const & P, uint64_t g_lo, uint32_t a0, uint32_t a1,
                                             uint32_t a2, uint32_t a3)
{
    uint32_t const acc[4] = {a0, a1, a2, a3};
    for (int q = 0; q < 4; ++q)
    {
        uint32_t bits = acc[q];
        while (bits)
        {
            int const b = __ffs(bits) - 1;
            bits &= bits - 1;
            int64_t const anchor = static_cast<int64_t>(g_lo) + 32 * q + b;
            bool confirmed = true;
            for (int c = 0; c < P.n_confirm && confirmed; ++c)
            {
                bool found = false;
                for (int d = P.confirm[c].lo; d <= P.confirm[c].hi && !found; ++d)
                {
                    int64_t const a = anchor + d;
                    if (a >= 0 && a < static_cast<int64_t>(p.n_text))
                    {
                        uint32_t const byte = __ldg(p.text + a);
                        found = (P.confirm[c].mask[byte >> 5] >> (byte & 31)) & 1;
                    }
                }
                confirmed = found;
            }
            if (!confirmed)
                continue;
            int64_t lo = anchor + P.start_lo, hi = anchor + P.start_hi;
            if (hi < 0 || lo >= static_cast<int64_t>(p.n_text))
                continue;
            lo = lo < 0 ? 0 : lo;
            hi = hi >= static_cast<int64_t>(p.n_text) ? static_cast<int64_t>(p.n_text) - 1 : hi;
            for (uint64_t c = static_cast<uint64_t>(lo) / kChunkBytes; c <= static_cast<uint64_t>(hi) / kChunkBytes; ++c)
                atomicOr(P.chunk_flags + (c >> 5), 1u << (c & 31));
            uint64_t r = record_of(p.rec_offsets, p.n_records, static_cast<uint64_t>(lo));
            while (r < p.n_records && __ldg(p.rec_offsets + r) <= static_cast<uint64_t>(hi))
            {
                if (__ldg(p.rec_offsets + r + 1) > static_cast<uint64_t>(lo) && is_candidate(P.cand_words, __ldg(p.rec_bin + r)))
                {
                    uint32_t const bit = 1u << (r & 31);
                    uint32_t const old = atomicOr(P.rec_flags + (r >> 5), bit);
                    if (!(old & bit))
                    {
                        unsigned const at = atomicAdd(P.rec_count, 1u);
                        if (at < kMultiListCap)
                            P.rec_list[at] = static_cast<uint32_t>(r);
                    }
                }
                ++r;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Class-window first stage for batches of patterns: BIT-PARALLEL OVER THE PATTERNS.  Every pattern of the batch brings a
// run of up to kWindowLen consecutive necessary positions with their byte classes (class_window, derived on the host);
// tables[k][byte] is the 128-bit set of patterns whose k-th window position accepts the byte (patterns with a shorter
// window accept everything there), kept as four 32-bit rows.  The patterns whose window can begin at text position j are
//      tables[0][t[j]] & tables[1][t[j+1]] & ... & tables[5][t[j+5]]
// — twenty-four conflict-free shared-memory loads and twenty ANDs per position for up to 128 patterns, against one term list per pattern
// in regex_filter_multi (1.5 thread-instructions per residue and pattern).  Survivors (about one in 10^6 per pattern) flag
// the chunks the match may start in and list the record, like regex_filter_multi's do.
struct window_params
{
    multi_program const * programs; // start_lo / start_hi and the confirm checks relative to the window's first byte
    uint32_t n_patterns;
    uint32_t const * tables;        // [kWindowLen][4][256]: window position, group of 32 patterns, byte
};

__device__ __noinline__ void window_survivors(scan_params const & p, window_params const & wp, uint64_t position, uint4 cand)
{
    uint32_t const words[4] = {cand.x, cand.y, cand.z, cand.w};
    for (int q = 0; q < 4; ++q)
    {
        uint32_t bits = words[q];
        while (bits)
        {
            int const b = __ffs(bits) - 1;
            bits &= bits - 1;
            multi_program const & P = wp.programs[32 * q + b];
            bool confirmed = true;
            for (int c = 0; c < P.n_confirm && confirmed; ++c)
            {
                bool found = false;
                for (int d = P.confirm[c].lo; d <= P.confirm[c].hi && !found; ++d)
                {
                    int64_t const a = static_cast<int64_t>(position) + d;
                    if (a >= 0 && a < static_cast<int64_t>(p.n_text))
                    {
                        uint32_t const byte = __ldg(p.text + a);
                        found = (P.confirm[c].mask[byte >> 5] >> (byte & 31)) & 1;
                    }
                }
                confirmed = found;
            }
            if (!confirmed)
                continue;
            int64_t lo = static_cast<int64_t>(position) + P.start_lo, hi = static_cast<int64_t>(position) + P.start_hi;
            if (hi < 0 || lo >= static_cast<int64_t>(p.n_text))
                continue;
            lo = lo < 0 ? 0 : lo;
            hi = hi >= static_cast<int64_t>(p.n_text) ? static_cast<int64_t>(p.n_text) - 1 : hi;
            for (uint64_t c = static_cast<uint64_t>(lo) / kChunkBytes; c <= static_cast<uint64_t>(hi) / kChunkBytes; ++c)
                atomicOr(P.chunk_flags + (c >> 5), 1u << (c & 31));
            uint64_t r = record_of(p.rec_offsets, p.n_records, static_cast<uint64_t>(lo));
            while (r < p.n_records && __ldg(p.rec_offsets + r) <= static_cast<uint64_t>(hi))
            {
                if (__ldg(p.rec_offsets + r + 1) > static_cast<uint64_t>(lo) && is_candidate(P.cand_words, __ldg(p.rec_bin + r)))
                {
                    uint32_t const bit = 1u << (r & 31);
                    uint32_t const old = atomicOr(P.rec_flags + (r >> 5), bit);
                    if (!(old & bit))
                    {
                        unsigned const at = atomicAdd(P.rec_count, 1u);
                        if (at < kMultiListCap)
                            P.rec_list[at] = static_cast<uint32_t>(r);
                    }
                }
                ++r;
            }
        }
    }
}

template <int STAGES>
__global__ void __launch_bounds__(kScanThreads + 32)
    regex_window_multi(scan_params p, window_params wp, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using ring_t = tile_ring<STAGES, kScanThreads>;
    ring_t ring;
    ring.init(smem_raw);
    // 32-bit rows: the 20-odd letters of a sequence alphabet are consecutive byte values, so the rows a warp asks for lie
    // in different banks (or are the same row: a broadcast).  With 16-byte rows more than half of the shared-memory
    // wavefronts were bank conflicts (ncu: 172 M of 329 M).
    uint32_t * const tables = reinterpret_cast<uint32_t *>(smem_raw + ((ring_t::kRingBytes + 15) & ~15));
    for (uint32_t i = threadIdx.x; i < kWindowLen * 4 * 256; i += blockDim.x)
        tables[i] = wp.tables[i];
    __syncthreads();
    if (ring.is_producer())
    {
        ring.produce(p, &tmap);
        return;
    }
    int const tid = threadIdx.x;
    uint8_t const * tile;
    uint32_t tile_id;
    while (ring.acquire(tile, tile_id))
    {
        uint64_t const g_lo = (static_cast<uint64_t>(tile_id) * kScanThreads + tid) * kChunkBytes;
        uint32_t const * const row = reinterpret_cast<uint32_t const *>(tile + tid * kChunkBytes);
        auto word = [&](int w) { return row[(((w >> 2) ^ (tid & 7)) << 2) + (w & 3)]; }; // SWIZZLE_128B: 16-byte pieces
        // the 8 bytes behind the row: the next row of the tile, or (last row) the text behind the tile
        uint32_t h0, h1;
        if (tid + 1 < kScanThreads)
        {
            uint32_t const * const nxt = reinterpret_cast<uint32_t const *>(tile + (tid + 1) * kChunkBytes);
            h0 = nxt[(((tid + 1) & 7) << 2) + 0];
            h1 = nxt[(((tid + 1) & 7) << 2) + 1];
        }
        else
        {
            h0 = h1 = 0;
            if (g_lo + kChunkBytes + 8 <= p.n_text_padded)
            {
                uint2 const v = __ldg(reinterpret_cast<uint2 const *>(p.text + g_lo + kChunkBytes));
                h0 = v.x;
                h1 = v.y;
            }
        }
        if (g_lo < p.n_text)
        {
            uint32_t a = word(0), b = word(1);
#pragma unroll 1
            for (int q = 0; q < 32; ++q)
            {
                uint32_t const c = q + 2 < 32 ? word(q + 2) : (q + 2 == 32 ? h0 : h1);
                // bytes 4q .. 4q+8 of the row
                uint32_t const by[9] = {a & 0xFF, (a >> 8) & 0xFF, (a >> 16) & 0xFF, a >> 24, b & 0xFF, (b >> 8) & 0xFF, (b >> 16) & 0xFF,
                                        b >> 24, c & 0xFF};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    uint32_t const * row = tables + by[j];
                    uint4 m = make_uint4(row[0], row[256], row[512], row[768]);
#pragma unroll
                    for (int k = 1; k < kWindowLen; ++k)
                    {
                        row = tables + k * 1024 + by[j + k];
                        m.x &= row[0];
                        m.y &= row[256];
                        m.z &= row[512];
                        m.w &= row[768];
                    }
                    if (m.x | m.y | m.z | m.w)
                        window_survivors(p, wp, g_lo + 4 * q + j, m);
                }
                a = b;
                b = c;
            }
        }
        ring.release();
    }
}

// packed term: letter | offset << 8 | rounds left in this pattern after this one << 16 | end-of-atom << 24 |
// end-of-pattern << 25.  Offset ranges are expanded on the host into one term per offset (terms of an atom are ORed
// anyway), so the kernel's term is "r |= mask[letter] >> offset"; every pattern's list is padded to whole rounds of four
// terms (with copies of its first term: ORing the same mask twice changes nothing), most selective atoms first, so that a
// warp can drop a pattern at a round boundary once none of its 4096 anchors survives the atoms merged so far.
__host__ __device__ __forceinline__ uint32_t multi_pack_term(uint32_t letter, uint32_t offset, uint32_t rounds_left, bool end_atom,
                                                             bool end_pattern)
{
    return letter | (offset << 8) | (rounds_left << 16) | (end_atom ? 1u << 24 : 0u) | (end_pattern ? 1u << 25 : 0u);
}
constexpr int kMultiMaxTermsPerPattern = 40; // after range expansion and padding; patterns that need more take the single-pattern path

template <int STAGES>
__global__ void __launch_bounds__(kScanThreads + 32)
    regex_filter_multi(scan_params p, multi_params mp, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using ring_t = tile_ring<STAGES, kScanThreads>;
    ring_t ring;
    ring.init(smem_raw);
    // masks4[letter][row] = the row's four 32-position mask words; next0[letter][row] = word 0 of the FOLLOWING row (the
    // 32 positions a shifted window runs into; row 255's comes from global memory); terms[] = the batch's packed terms
    uint4 * const masks4 = reinterpret_cast<uint4 *>(smem_raw + ((ring_t::kRingBytes + 15) & ~15));
    uint32_t * const next0 = reinterpret_cast<uint32_t *>(masks4 + kMultiLetters * kScanThreads);
    uint32_t * const terms = next0 + kMultiLetters * kScanThreads;
    uint32_t const n_terms4 = (mp.n_terms_total + 3) & ~3u;
    for (uint32_t i = threadIdx.x; i < n_terms4 + 4; i += blockDim.x)
        terms[i] = i < mp.n_terms_total ? mp.terms[i] : 0u; // padding after the last pattern: harmless ORs into r
    __syncthreads();
    if (ring.is_producer())
    {
        ring.produce(p, &tmap);
        return;
    }
    int const tid = threadIdx.x;
    uint4 const * const my_masks = masks4 + tid;
    uint32_t const * const my_next = next0 + tid;
    uint8_t const * tile;
    uint32_t tile_id;
    while (ring.acquire(tile, tile_id))
    {
        uint64_t const g_lo = (static_cast<uint64_t>(tile_id) * kScanThreads + tid) * kChunkBytes;
        uint32_t w[32];
        {
            uint8_t const * row = tile + tid * kChunkBytes;
#pragma unroll
            for (int piece = 0; piece < 8; ++piece)
            {
                uint4 const v = *reinterpret_cast<uint4 const *>(row + ((piece ^ (tid & 7)) << 4));
                w[4 * piece + 0] = v.x;
                w[4 * piece + 1] = v.y;
                w[4 * piece + 2] = v.z;
                w[4 * piece + 3] = v.w;
            }
        }
        ring.release(); // the row is in registers: the producer may refill the stage while the batch is evaluated
        for (uint32_t l = 0; l < mp.n_letters; ++l)
        {
            uint32_t const x = mp.splat[l];
            uint32_t mm[4];
#pragma unroll
            for (int q = 0; q < 4; ++q)
            {
                mm[q] = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mm[q] |= literal_nibble(w[8 * q + j], x) << (4 * j);
            }
            masks4[l * kScanThreads + tid] = make_uint4(mm[0], mm[1], mm[2], mm[3]);
            if (tid > 0)
                next0[l * kScanThreads + tid - 1] = mm[0];
        }
        if (tid == kScanThreads - 1)
        {
            // the 32 bytes after the tile belong to another tile: straight from global memory
            uint32_t la[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (g_lo + kChunkBytes + 32 <= p.n_text_padded)
            {
                uint4 const v0 = __ldg(reinterpret_cast<uint4 const *>(p.text + g_lo + kChunkBytes));
                uint4 const v1 = __ldg(reinterpret_cast<uint4 const *>(p.text + g_lo + kChunkBytes + 16));
                la[0] = v0.x, la[1] = v0.y, la[2] = v0.z, la[3] = v0.w, la[4] = v1.x, la[5] = v1.y, la[6] = v1.z, la[7] = v1.w;
            }
            for (uint32_t l = 0; l < mp.n_letters; ++l)
            {
                uint32_t mm = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mm |= literal_nibble(la[j], mp.splat[l]) << (4 * j);
                next0[l * kScanThreads + tid] = mm; // zero padding never equals a letter (no pattern literal is byte 0)
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kScanThreads) : "memory"); // consumers only; once per tile and BATCH
        bool const real = g_lo < p.n_text;
        uint32_t acc0 = ~0u, acc1 = ~0u, acc2 = ~0u, acc3 = ~0u, r0 = 0, r1 = 0, r2 = 0, r3 = 0;
        uint32_t pat = 0;
        // four terms per round, all eight operand loads issued before the first use: there are only eight consumer warps
        // per SM (the masks take the shared memory), so the latency has to be hidden inside the thread
        for (uint32_t k = 0; k < n_terms4; k += 4)
        {
            uint4 const tw4 = *reinterpret_cast<uint4 const *>(terms + k);
            uint32_t const tw[4] = {tw4.x, tw4.y, tw4.z, tw4.w};
            uint32_t const rounds_left = (tw4.x >> 16) & 0xFF;
            uint4 m[4];
            uint32_t nx[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                uint32_t const l = tw[u] & 0xFF;
                m[u] = my_masks[l * kScanThreads];
                nx[u] = my_next[l * kScanThreads];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                uint32_t const d = (tw[u] >> 8) & 0xFF;
                r0 |= __funnelshift_r(m[u].x, m[u].y, d);
                r1 |= __funnelshift_r(m[u].y, m[u].z, d);
                r2 |= __funnelshift_r(m[u].z, m[u].w, d);
                r3 |= __funnelshift_r(m[u].w, nx[u], d);
                if (tw[u] & (3u << 24))
                {
                    if (tw[u] & (1u << 24))
                    {
                        acc0 &= r0, acc1 &= r1, acc2 &= r2, acc3 &= r3;
                        r0 = r1 = r2 = r3 = 0;
                    }
                    if (tw[u] & (1u << 25))
                    {
                        if ((acc0 | acc1 | acc2 | acc3) && real)
                            multi_survivors(p, mp.programs[pat], g_lo, acc0, acc1, acc2, acc3);
                        acc0 = acc1 = acc2 = acc3 = ~0u;
                        ++pat;
                    }
                }
            }
            // nothing left of this pattern in the whole warp: its remaining rounds cannot bring anything back
            if (rounds_left && !__any_sync(0xffffffffu, (acc0 | acc1 | acc2 | acc3) != 0))
            {
                k += 4 * rounds_left;
                acc0 = acc1 = acc2 = acc3 = ~0u;
                r0 = r1 = r2 = r3 = 0;
                ++pat;
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kScanThreads) : "memory"); // nobody still reads the masks of this tile
    }
}

} // namespace trx
