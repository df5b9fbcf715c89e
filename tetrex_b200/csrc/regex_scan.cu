// regex_scan.cu — C ABI for the HBM-resident residue store and the K4 verify scan.
// Reference: iter_disk_search + verify_* (include/query.h:97-188, src/query.cpp:167-315).
#include <algorithm>
#include <chrono>
#include <atomic>
#include <thread>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <memory>
#include <unordered_map>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "regex_kernels.cuh"

struct trx_text
{
    trx_ctx * ctx{nullptr};
    uint64_t n_text{}, n_text_padded{}, n_records{}, n_bins{};
    uint8_t * d_text{nullptr};
    uint64_t * d_rec_offsets{nullptr};
    uint32_t * d_rec_bin{nullptr};
    uint64_t * d_bin_rec_begin{nullptr};
    // Per-scan state lives in SLOTS so that several scans can be in flight on the context's stream (trx_regex_scan_begin /
    // _end): the kernels of consecutive scans then run back to back, without the host round trip (plan, enqueue, sync,
    // read-back — ~50 us) between them.  One allocation per slot, so that a scan needs ONE memset and ONE read-back:
    //   [chunk flags | record flags | counters (16 B) | first kSmallHits keys | first kSmallHits ends]
    //    <------------- cleared per scan ------------->
    //                                <------------------ copied to the host --------------------------->
    struct scan_slot
    {
        uint8_t * d_block{nullptr};
        uint32_t * d_chunk_flags{nullptr};
        uint32_t * d_rec_flags{nullptr};   // bitmap
        uint8_t * d_counters{nullptr};     // hit count (u64), listed-record count (u32), next tile (u32)
        uint32_t * d_rec_list{nullptr};
        trx::scratch d_stage, h_stage, h_small, d_keys, d_ends, d_tiles, d_start_bits;
        uint64_t tiles_serial{0};          // plan_serial of the tile list in d_tiles
        cudaStream_t stream{nullptr};      // a stream per slot: the copies and the clear of one scan overlap the kernels of another
        cudaEvent_t done{nullptr};
        bool busy{false};
        uint64_t cap{}, residues{}, filter_units{};
    };
    static constexpr int kScanSlots = 4;
    scan_slot slots[kScanSlots];
    uint64_t clear_bytes{};            // chunk flags + record flags + counters
    uint64_t n_chunk_flag_words{};
    uint64_t n_rec_flag_words{};
    double byte_freq[256]{};           // relative byte frequencies of the residue store
    std::vector<uint64_t> h_rec_offsets;   // host copies: tile planning and hit formatting
    std::vector<uint32_t> h_rec_bin;
    std::vector<uint64_t> h_bin_rec_begin; // n_bins + 1
    CUtensorMap tmap{};
    bool has_tmap{false};
    // the last tile plan (candidate bins → 32 KB tiles) stays on the device: queries that verify the same candidate set
    // (bench loops, strand pairs, -c conjunctions) skip the planning loop and the upload of the tile list
    std::vector<uint64_t> plan_cand;
    std::vector<uint32_t> plan_tiles;
    uint64_t plan_residues{};
    bool plan_valid{false};
    uint64_t plan_serial{0};           // bumped whenever plan_tiles changes
    // phase-1 choice per (regex, strand, byte map, TRX_REGEX_FILTER): derived from the automaton once, not per scan
    struct cached_choice
    {
        uint64_t serial, map_hash;
        int reverse, mode, kind;
        trx::filter_spec window;
        trx::atom_spec atoms;
        // the plan for multi-pattern passes (whole alphabet, larger classes, as selective as the pattern allows)
        int multi_kind{-1}; // -1 = not derived yet
        trx::atom_spec multi_atoms;
        // class window for regex_window_multi (derive_class_window): -1 = not derived yet, 0 = none, 1 = usable
        int win_state{-1};
        int win_len{0}, win_off_lo{0}, win_off_hi{0};
        uint32_t win_cls[trx::kWindowLen][8];
        int win_n_confirm{0};
        trx::confirm_check win_confirm[trx::kMaxConfirm];
    };
    std::vector<cached_choice> choices;
    std::unordered_map<uint64_t, uint32_t> plain_choice; // serial -> index in choices (forward, unmapped, default mode); entries are verified on use
    // per-scan scratch
    trx::scratch d_keys2, d_ends2, d_sort_tmp, d_hits_fmt;
    // multi-pattern scans (trx_regex_scan_multi): one staging block up (programs, automata, candidate words, phase-2
    // parameters), one block of flags / lists / counters, one hit list
    // two sets of per-pass blocks: the host stages pass n + 1 while the device works on pass n
    struct multi_set
    {
        trx::scratch h_stage, d_stage, d_flags, d_lists, d_counters, h_back;
        cudaEvent_t done{nullptr};
        bool in_flight{false};
        std::vector<uint32_t> ids; // the patterns of the pass in flight
    };
    multi_set m_sets[2];
    // the hits of a whole call stay on the device (one list for all patterns, one counter) until they are sorted there
    trx::scratch m_d_hit_keys, m_d_hit_ends, m_d_hit_pats, m_d_total, m_h_total, m_d_sort, m_h_out;
    bool multi_attr_set{false};
};

using namespace trx;

namespace
{

using encode_fn_t = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, cuuint64_t const *,
                                 cuuint64_t const *, cuuint32_t const *, cuuint32_t const *, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// libcuda is reached through the runtime (no link-time dependency on the driver library)
int make_tensor_map(trx_text * t)
{
    void * fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    TRX_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess)
        return set_error(TRX_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    auto encode = reinterpret_cast<encode_fn_t>(fn);
    // the text as a 2-D byte tensor: [rows][128]; one TMA box = 256 rows × 128 B = one 32 KB tile
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(kChunkBytes), static_cast<cuuint64_t>(t->n_text_padded / kChunkBytes)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(kChunkBytes)};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kChunkBytes), static_cast<cuuint32_t>(kScanThreads)};
    cuuint32_t elem_strides[2] = {1, 1};
    CUresult r = encode(&t->tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, t->d_text, dims, strides, box, elem_strides,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return set_error(TRX_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    t->has_tmap = true;
    return TRX_OK;
}

} // namespace

extern "C"
{

int trx_text_upload(trx_ctx_t * ctx, const uint8_t * residues, const uint64_t * rec_offsets, const uint32_t * rec_bin,
                    uint64_t n_records, uint64_t n_bins, trx_text_t ** out)
{
    if (!ctx || !out || !rec_offsets || !rec_bin || n_records == 0 || n_bins == 0)
        return set_error(TRX_EINVAL, "trx_text_upload: bad argument");
    *out = nullptr;
    if (rec_offsets[0] != 0)
        return set_error(TRX_EINVAL, "rec_offsets[0] must be 0");
    if (n_records >= (1ull << 32))
        return set_error(TRX_EINVAL, "more than 2^32 records");
    for (uint64_t i = 0; i < n_records; ++i)
    {
        if (rec_offsets[i + 1] < rec_offsets[i])
            return set_error(TRX_EINVAL, "rec_offsets must be non-decreasing");
        if (rec_offsets[i + 1] - rec_offsets[i] >= (1ull << 32))
            return set_error(TRX_EINVAL, "record %llu longer than 2^32 bytes", (unsigned long long)i);
        if (rec_bin[i] >= n_bins)
            return set_error(TRX_ERANGE, "record %llu: bin %u >= %llu", (unsigned long long)i, rec_bin[i],
                             (unsigned long long)n_bins);
        if (i && rec_bin[i] < rec_bin[i - 1])
            return set_error(TRX_EINVAL, "records must be grouped by ascending bin");
    }
    uint64_t const n_text = rec_offsets[n_records];
    if (n_text && !residues)
        return set_error(TRX_EINVAL, "residues is NULL");
    if (n_text / kTileBytes >= (1ull << 31) / kScanThreads)
        return set_error(TRX_EINVAL, "text too large for 32-bit TMA row coordinates");
    TRX_TRY(ctx->use());
    auto * t = new (std::nothrow) trx_text;
    if (!t)
        return set_error(TRX_ENOMEM, "out of host memory");
    t->ctx = ctx;
    t->n_text = n_text;
    t->n_text_padded = (n_text / kTileBytes + 1) * kTileBytes;
    t->n_records = n_records;
    t->n_bins = n_bins;
    t->h_rec_offsets.assign(rec_offsets, rec_offsets + n_records + 1);
    t->h_rec_bin.assign(rec_bin, rec_bin + n_records);
    t->h_bin_rec_begin.assign(n_bins + 1, 0);
    {
        uint64_t r = 0;
        for (uint64_t b = 0; b <= n_bins; ++b)
        {
            while (r < n_records && rec_bin[r] < b)
                ++r;
            t->h_bin_rec_begin[b] = r;
        }
    }
    t->n_chunk_flag_words = (t->n_text_padded / kChunkBytes + 31) / 32;
    auto fail = [&](cudaError_t e)
    {
        trx_text_destroy(t);
        return set_error(e == cudaErrorMemoryAllocation ? TRX_ENOMEM : TRX_ECUDA, "trx_text_upload: %s",
                         cudaGetErrorString(e));
    };
    cudaError_t e;
    if ((e = cudaMalloc(&t->d_text, t->n_text_padded)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&t->d_rec_offsets, (n_records + 1) * 8)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&t->d_rec_bin, n_records * 4)) != cudaSuccess) return fail(e);
    if ((e = cudaMalloc(&t->d_bin_rec_begin, (n_bins + 1) * 8)) != cudaSuccess) return fail(e);
    t->n_rec_flag_words = (n_records + 31) / 32;
    {
        uint64_t const flag_bytes = (t->n_chunk_flag_words * 4 + t->n_rec_flag_words * 4 + 15) & ~uint64_t{15};
        t->clear_bytes = flag_bytes + 16;
        for (trx_text::scan_slot & sl : t->slots)
        {
            if ((e = cudaMalloc(&sl.d_block, t->clear_bytes + kSmallHits * 12)) != cudaSuccess) return fail(e);
            sl.d_chunk_flags = reinterpret_cast<uint32_t *>(sl.d_block);
            sl.d_rec_flags = sl.d_chunk_flags + t->n_chunk_flag_words;
            sl.d_counters = sl.d_block + flag_bytes;
            if ((e = cudaMalloc(&sl.d_rec_list, n_records * 4)) != cudaSuccess) return fail(e);
            if ((e = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming)) != cudaSuccess) return fail(e);
            if ((e = cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e);
        }
    }
    if ((e = cudaMemsetAsync(t->d_text + (t->n_text_padded - kTileBytes), 0, kTileBytes, ctx->stream)) != cudaSuccess) return fail(e);
    if (n_text && (e = cudaMemcpyAsync(t->d_text, residues, n_text, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(t->d_rec_offsets, rec_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(t->d_rec_bin, rec_bin, n_records * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(t->d_bin_rec_begin, t->h_bin_rec_begin.data(), (n_bins + 1) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e);
    {
        // byte histogram (one pass over the text on the device): lets the scan pick the rarest literals
        unsigned long long * d_hist = nullptr;
        unsigned long long h_hist[256];
        if ((e = cudaMalloc(&d_hist, 256 * 8)) != cudaSuccess) return fail(e);
        cudaMemsetAsync(d_hist, 0, 256 * 8, ctx->stream);
        byte_histogram_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(t->d_text, n_text, d_hist);
        ++ctx->launches;
        cudaMemcpyAsync(h_hist, d_hist, sizeof h_hist, cudaMemcpyDeviceToHost, ctx->stream);
        e = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_hist);
        if (e != cudaSuccess) return fail(e);
        for (int i = 0; i < 256; ++i)
            t->byte_freq[i] = n_text ? static_cast<double>(h_hist[i]) / static_cast<double>(n_text) : 0.0;
    }
    int rc = make_tensor_map(t);
    if (rc != TRX_OK)
    {
        trx_text_destroy(t);
        return rc;
    }
    *out = t;
    return TRX_OK;
}

void trx_text_destroy(trx_text_t * t)
{
    if (!t)
        return;
    cudaSetDevice(t->ctx->device);
    cudaStreamSynchronize(t->ctx->stream);
    cudaFree(t->d_text);
    cudaFree(t->d_rec_offsets);
    cudaFree(t->d_rec_bin);
    cudaFree(t->d_bin_rec_begin);
    for (trx_text::scan_slot & sl : t->slots)
    {
        cudaFree(sl.d_block);
        cudaFree(sl.d_rec_list);
        if (sl.stream)
        {
            cudaStreamSynchronize(sl.stream);
            cudaStreamDestroy(sl.stream);
        }
        if (sl.done)
            cudaEventDestroy(sl.done);
        for (trx::scratch * s : {&sl.d_stage, &sl.h_stage, &sl.h_small, &sl.d_keys, &sl.d_ends, &sl.d_tiles, &sl.d_start_bits})
            s->release();
    }
    for (trx::scratch * s : {&t->d_keys2, &t->d_ends2, &t->d_sort_tmp, &t->d_hits_fmt, &t->m_d_hit_keys, &t->m_d_hit_ends, &t->m_d_hit_pats, &t->m_d_total,
                             &t->m_h_total, &t->m_d_sort, &t->m_h_out})
        s->release();
    for (trx_text::multi_set & ms : t->m_sets)
    {
        if (ms.done)
            cudaEventDestroy(ms.done);
        for (trx::scratch * s : {&ms.h_stage, &ms.d_stage, &ms.d_flags, &ms.d_lists, &ms.d_counters, &ms.h_back})
            s->release();
    }
    delete t;
}

} // extern "C"

namespace
{

template <int W, int DIR, bool USE_TMA>
int launch_phase1(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    size_t smem = (USE_TMA ? kStages * kTileBytes : 0) + sizeof(scan_tables<W>) + 2 * sizeof(uint64_t);
    auto kern = regex_scan_phase1<W, DIR, USE_TMA>;
    // per template instantiation AND per device (function attributes belong to a device; `tetrex query --gpus N` drives
    // several from one process): attribute + occupancy are queried once
    static int per_sm_of[64] = {};
    int & per_sm = per_sm_of[c->device & 63];
    if (per_sm == 0)
    {
        TRX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int q = 0;
        TRX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, kern, kScanThreads, smem));
        per_sm = q < 1 ? 1 : q;
    }
    unsigned grid = static_cast<unsigned>(std::min<uint64_t>(sp.n_tiles, static_cast<uint64_t>(c->sm_count) * per_sm));
    kern<<<grid, kScanThreads, smem, st>>>(sp, tmap);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

template <int W>
int launch_phase1_w(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap, bool reverse, bool use_tma)
{
    if (reverse)
        return use_tma ? launch_phase1<W, -1, true>(c, st, sp, tmap) : launch_phase1<W, -1, false>(c, st, sp, tmap);
    return use_tma ? launch_phase1<W, 1, true>(c, st, sp, tmap) : launch_phase1<W, 1, false>(c, st, sp, tmap);
}

// threads per 128-byte text row in the prefilter kernels (TRX_FILTER_SPLIT = 1 | 2; 1 was measured 2-8 % faster).
// The ring depth is fixed at 2: 3 and 4 stages were measured 20-30 % slower (fewer resident CTAs per SM).
int filter_split()
{
    static int split = []
    {
        char const * e = getenv("TRX_FILTER_SPLIT");
        int v = e ? atoi(e) : kFilterSplitDefault;
        return v <= 1 ? 1 : 2;
    }();
    return split;
}

template <int NA, int NB, int STAGES, int SPLIT>
int launch_filter_s(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    constexpr int kThreads = kScanThreads * SPLIT + 32; // + the producer warp
    size_t smem = tile_ring<STAGES, kScanThreads * SPLIT>::kRingBytes;
    auto kern = regex_filter_simd<NA, NB, STAGES, SPLIT>;
    static int per_sm_of[64] = {};
    int & per_sm = per_sm_of[c->device & 63];
    if (per_sm == 0)
    {
        TRX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int q = 0;
        TRX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, kern, kThreads, smem));
        per_sm = q < 1 ? 1 : q;
    }
    unsigned grid = static_cast<unsigned>(std::min<uint64_t>(sp.n_tiles, static_cast<uint64_t>(c->sm_count) * per_sm));
    kern<<<grid, kThreads, smem, st>>>(sp, tmap);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

template <int NA, int NB>
int launch_filter(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    return filter_split() == 1 ? launch_filter_s<NA, NB, 2, 1>(c, st, sp, tmap) : launch_filter_s<NA, NB, 2, 2>(c, st, sp, tmap);
}

template <int NA>
int launch_filter_a(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    switch (sp.filter.n_b)
    {
    case 0: return launch_filter<NA, 0>(c, st, sp, tmap);
    case 1: return launch_filter<NA, 1>(c, st, sp, tmap);
    case 2: return launch_filter<NA, 2>(c, st, sp, tmap);
    case 3: return launch_filter<NA, 3>(c, st, sp, tmap);
    default: return launch_filter<NA, 4>(c, st, sp, tmap);
    }
}

int launch_filter_any(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    switch (sp.filter.n_a)
    {
    case 0: return launch_filter_a<0>(c, st, sp, tmap);
    case 1: return launch_filter_a<1>(c, st, sp, tmap);
    case 2: return launch_filter_a<2>(c, st, sp, tmap);
    case 3: return launch_filter_a<3>(c, st, sp, tmap);
    default: return launch_filter_a<4>(c, st, sp, tmap);
    }
}

template <int NLIT, int STAGES, int SPLIT>
int launch_atoms_s(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    constexpr int kThreads = kScanThreads * SPLIT + 32; // + the producer warp
    size_t smem = ((tile_ring<STAGES, kScanThreads * SPLIT>::kRingBytes + 15) & ~15) +
                  2 * NLIT * kScanThreads * SPLIT * sizeof(uint32_t);
    auto kern = regex_filter_atoms<NLIT, STAGES, SPLIT>;
    static int per_sm_of[64] = {};
    int & per_sm = per_sm_of[c->device & 63];
    if (per_sm == 0)
    {
        TRX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int q = 0;
        TRX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, kern, kThreads, smem));
        per_sm = q < 1 ? 1 : q;
    }
    unsigned grid = static_cast<unsigned>(std::min<uint64_t>(sp.n_tiles, static_cast<uint64_t>(c->sm_count) * per_sm));
    kern<<<grid, kThreads, smem, st>>>(sp, tmap);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

template <int NLIT>
int launch_atoms(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    return filter_split() == 1 ? launch_atoms_s<NLIT, 2, 1>(c, st, sp, tmap) : launch_atoms_s<NLIT, 2, 2>(c, st, sp, tmap);
}

int launch_atoms_any(trx_ctx * c, cudaStream_t st, scan_params const & sp, CUtensorMap const & tmap)
{
    switch (sp.atoms.n_lit)
    {
    case 1: return launch_atoms<1>(c, st, sp, tmap);
    case 2: return launch_atoms<2>(c, st, sp, tmap);
    case 3: return launch_atoms<3>(c, st, sp, tmap);
    default: return launch_atoms<4>(c, st, sp, tmap);
    }
}

template <int W>
int launch_phase2_w(trx_ctx * c, cudaStream_t st, resolve_params const & rp, bool reverse)
{
    unsigned grid = static_cast<unsigned>(c->sm_count) * 4;
    if (reverse)
        regex_scan_phase2<W, -1><<<grid, 128, 0, st>>>(rp);
    else
        regex_scan_phase2<W, 1><<<grid, 128, 0, st>>>(rp);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

// comp_tab of src/query.cpp:7-16 as a byte map: IUPAC complement, everything else maps to itself.
void complement_table(uint8_t tab[256])
{
    for (int i = 0; i < 256; ++i)
        tab[i] = static_cast<uint8_t>(i);
    char const * from = "ACGTUMRWSYKVHDBNacgtumrwsykvhdbn";
    char const * to = "TGCAAKYWSRMBDHVNtgcaakywsrmbdhvn";
    for (int i = 0; from[i]; ++i)
        tab[static_cast<uint8_t>(from[i])] = static_cast<uint8_t>(to[i]);
}

constexpr double kDefaultMaxPass = 1e-3;
constexpr double kMaxSimdPass = 2e-2; // admission of the gapped prefilter's SIMD stage when confirm checks follow

struct bits128 // a set of positions (the name is from the days of two words)
{
    uint64_t w[kRegexWords]{};
    bool any() const
    {
        uint64_t x = 0;
        for (int k = 0; k < kRegexWords; ++k)
            x |= w[k];
        return x != 0;
    }
    bool intersects(uint64_t const * o) const
    {
        uint64_t x = 0;
        for (int k = 0; k < kRegexWords; ++k)
            x |= w[k] & o[k];
        return x != 0;
    }
};

// Derives the necessary factor of filter_spec from the automaton's level sets: L_0 = First,
// L_{j+1} = Follow(L_j); every match has a byte of class(L_j) at offset j as long as no shorter match exists.
// Offsets whose class is a single raw byte are literals; the best window of 4 consecutive offsets (by the
// text's measured byte frequencies) becomes the filter.  n_a = n_b = 0 if nothing selective enough exists.
// A prefilter whose expected pass rate (fraction of text positions it lets through) exceeds this is not used: phase 2
// would then visit most chunks anyway and the exact automaton kernel is the better phase 1.  TRX_FILTER_MAX_PASS
// overrides the measured default.
double max_pass_rate()
{
    static double const v = []
    {
        char const * e = getenv("TRX_FILTER_MAX_PASS");
        double x = e ? atof(e) : kDefaultMaxPass;
        return x > 0 ? x : kDefaultMaxPass;
    }();
    return v;
}

filter_spec derive_filter(regex_dev const & img, double const * freq, bool reverse, double * pass_out)
{
    filter_spec fs{};
    *pass_out = 1.0;
    constexpr int kMaxDepth = 48;
    int literal[kMaxDepth];
    int depths = 0;
    bits128 L;
    for (int k = 0; k < kRegexWords; ++k)
        L.w[k] = img.first[k];
    for (; depths < kMaxDepth && L.any(); ++depths)
    {
        int count = 0, byte = -1;
        for (int c = 0; c < 256 && count < 2; ++c)
            if (L.intersects(img.B[c]))
            {
                ++count;
                byte = c;
            }
        literal[depths] = count == 1 ? byte : -1;
        if (L.intersects(img.last))
        {
            ++depths; // this offset is still necessary, deeper ones are not
            break;
        }
        bits128 next;
        for (int k = 0; k < kRegexWords; ++k)
            for (uint64_t s = L.w[k]; s; s &= s - 1)
            {
                int pos = k * 64 + __builtin_ctzll(s);
                for (int j = 0; j < kRegexWords; ++j)
                    next.w[j] |= img.follow[pos][j];
            }
        L = next;
    }
    // best window of 8 consecutive offsets: take literals rarest-first until the expected candidate rate is low
    constexpr double kGoodEnough = 2e-5;
    double const kTooWeak = max_pass_rate();
    double best = 1.0;
    int best_a = -1;
    uint32_t best_mask = 0;
    for (int a = 0; a < depths; ++a)
    {
        int order[8], n = 0;
        for (int j = a; j < a + 8 && j < depths; ++j)
            if (literal[j] >= 0)
                order[n++] = j;
        std::sort(order, order + n, [&](int x, int y) { return freq[literal[x]] < freq[literal[y]]; });
        double prod = 1.0;
        uint32_t mask = 0;
        for (int i = 0; i < n && prod > kGoodEnough; ++i)
        {
            prod *= std::max(freq[literal[order[i]]], 1e-4);
            mask |= 1u << (order[i] - a);
        }
        // prefer the cheaper filter among equally good ones
        if (mask && (prod < best && (best > kGoodEnough || __builtin_popcount(mask) < __builtin_popcount(best_mask))))
        {
            best = prod;
            best_a = a;
            best_mask = mask;
        }
    }
    if (best_a < 0 || best > kTooWeak) // weaker than ~1 candidate per 8 chunks: the exact automaton is the better filter
        return fs;
    // window byte 0 = lowest address: offset best_a when reading forward, best_a + 7 when reading backwards
    int n = 0;
    for (int off = 0; off < 8; ++off)
    {
        int const d = reverse ? best_a + 7 - off : best_a + off;
        if (d < best_a || d >= best_a + 8 || !((best_mask >> (d - best_a)) & 1))
            continue;
        fs.delta[n] = static_cast<uint32_t>(off & 3);
        fs.splat[n] = 0x01010101u * static_cast<uint32_t>(literal[d]);
        if (off < 4)
            ++fs.n_a;
        else
            ++fs.n_b;
        ++n;
    }
    fs.start_delta = reverse ? best_a + 7 : -best_a;
    *pass_out = best;
    return fs;
}

// ---- gapped factor ("atoms") derivation -----------------------------------------------------------------
// Works on the position automaton as the scan sees it (byte maps already folded into B).  Nodes: positions
// 0..P-1, S = P (before the first symbol), F = P + 1 (accept).  A position is *necessary* when F cannot be
// reached from S without it; necessary positions form a chain every match walks through in the same order.
// For neighbours on that chain the shortest and the longest walk between them bound their distance in the
// text (the {m,n} expansions of the Glushkov construction are loop-free, so the bounds are finite there; a
// '*' or '+' in between makes the distance unbounded and ends the usable part of the chain).
struct atom_plan
{
    atom_spec spec{};
    double pass{1.0};      // expected fraction of text positions that survive the SIMD stage
    double confirmed{1.0}; // … and the confirm checks: what phase 2 gets to see
};

struct automaton_graph
{
    int P{0};
    std::vector<std::vector<int>> succ; // P + 2 nodes
    int S() const { return P; }
    int F() const { return P + 1; }
};

automaton_graph build_graph(regex_dev const & img)
{
    automaton_graph g;
    g.P = img.positions;
    g.succ.assign(g.P + 2, {});
    auto has = [](uint64_t const * w, int i) { return (w[i >> 6] >> (i & 63)) & 1; };
    for (int q = 0; q < g.P; ++q)
        if (has(img.first, q))
            g.succ[g.S()].push_back(q);
    for (int pos = 0; pos < g.P; ++pos)
    {
        for (int q = 0; q < g.P; ++q)
            if (has(img.follow[pos], q))
                g.succ[pos].push_back(q);
        if (has(img.last, pos))
            g.succ[pos].push_back(g.F());
    }
    return g;
}

// is `to` reachable from `from` when node `without` may not be entered?
bool reachable_without(automaton_graph const & g, int from, int to, int without)
{
    std::vector<char> seen(g.succ.size(), 0);
    std::vector<int> stack{from};
    seen[from] = 1;
    while (!stack.empty())
    {
        int u = stack.back();
        stack.pop_back();
        if (u == to)
            return true;
        for (int v : g.succ[u])
            if (v != without && !seen[v])
            {
                seen[v] = 1;
                stack.push_back(v);
            }
    }
    return false;
}

constexpr int kUnbounded = 1 << 20;

// Number of symbols strictly between leaving `src` and the first arrival at `dst`, minimum and maximum over all
// walks (dst is not left again).  hi = kUnbounded if a walk can loop before reaching dst.
bool walk_range(automaton_graph const & g, int src, int dst, int & lo, int & hi)
{
    size_t const n = g.succ.size();
    // nodes that can reach dst without passing through dst
    std::vector<char> co(n, 0);
    co[dst] = 1;
    for (bool grew = true; grew;)
    {
        grew = false;
        for (size_t u = 0; u < n; ++u)
            if (!co[u] && static_cast<int>(u) != dst)
                for (int v : g.succ[u])
                    if (co[v])
                    {
                        co[u] = 1;
                        grew = true;
                        break;
                    }
    }
    if (!co[src])
        return false;
    // shortest: BFS over co-reachable nodes (edges = one symbol each; the edge into dst counts as its own symbol)
    std::vector<int> dist(n, -1);
    std::vector<int> queue;
    for (int v : g.succ[src])
        if (co[v] && dist[v] < 0)
        {
            dist[v] = 0;
            queue.push_back(v);
        }
    for (size_t h = 0; h < queue.size(); ++h)
    {
        int u = queue[h];
        if (u == dst)
            continue;
        for (int v : g.succ[u])
            if (co[v] && dist[v] < 0)
            {
                dist[v] = dist[u] + 1;
                queue.push_back(v);
            }
    }
    if (dist[dst] < 0)
        return false;
    lo = dist[dst];
    // longest: memoised DFS, a back edge among co-reachable nodes = unbounded
    std::vector<int> best(n, -2); // -2 unvisited, -3 on stack
    bool cyclic = false;
    std::function<int(int)> longest = [&](int u) -> int
    {
        if (u == dst)
            return 0;
        if (best[u] == -3)
        {
            cyclic = true;
            return 0;
        }
        if (best[u] >= 0)
            return best[u];
        best[u] = -3;
        int b = 0;
        for (int v : g.succ[u])
            if (co[v])
                b = std::max(b, 1 + longest(v));
        best[u] = b;
        return b;
    };
    int b = 0;
    for (int v : g.succ[src])
        if (co[v])
            b = std::max(b, longest(v));
    hi = cyclic ? kUnbounded : b;
    return true;
}

// what the SIMD stage of a gapped plan may spend.  The single-pattern kernel builds its occurrence masks itself (4 distinct
// bytes, classes of <= 3); the multi-pattern kernel finds the masks of the whole batch alphabet ready in shared memory, so a
// term costs it three memory instructions whatever the byte: it takes every necessary position it can get.
struct atom_budget
{
    int max_lits{4};
    int max_class{3};
    int max_atoms{8};
    double good_enough{2e-5};
};
constexpr int kMaxClassBytes = 6;

atom_plan derive_atoms(regex_dev const & img, double const * freq, bool reverse, atom_budget const & budget = atom_budget{})
{
    atom_plan none;
    automaton_graph const g = build_graph(img);
    int const P = g.P;
    if (P <= 0)
        return none;
    // class of a position = the raw bytes it accepts.  At most 3 of them: usable as a SIMD atom; any other class that
    // is not nearly everything: usable as a confirm check at the anchors the SIMD stage lets through.
    int const kMaxClass = budget.max_class;
    struct small_class
    {
        int n{0};          // 1..max_class: SIMD-capable; 0: not
        int byte[kMaxClassBytes]{};
        double freq{0.0};  // expected fraction of text bytes in the class
        uint32_t mask[8]{};
        bool usable{false}; // as a confirm check
    };
    std::vector<small_class> cls(P);
    {
        std::vector<int> count(P, 0);
        for (int c = 0; c < 256; ++c)
            for (int pos = 0; pos < P; ++pos)
                if ((img.B[c][pos >> 6] >> (pos & 63)) & 1)
                {
                    if (count[pos] < kMaxClassBytes)
                        cls[pos].byte[count[pos]] = c;
                    ++count[pos];
                    cls[pos].mask[c >> 5] |= 1u << (c & 31);
                    cls[pos].freq += c == 0 ? 1.0 : std::max(freq[c], 1e-4); // the text is zero padded: NUL is "everywhere"
                }
        for (int pos = 0; pos < P; ++pos)
        {
            cls[pos].n = (count[pos] <= kMaxClass && !(cls[pos].mask[0] & 1u)) ? count[pos] : 0;
            cls[pos].usable = count[pos] > 0 && cls[pos].freq <= 0.6;
        }
    }
    // chain of necessary positions with a usable class, ordered by their shortest distance from S
    struct link
    {
        int pos;
        int off_lo, off_hi; // offset of the symbol inside the match (first symbol = 0)
    };
    std::vector<link> chain;
    {
        std::vector<std::pair<int, int>> order; // (min offset, position)
        for (int pos = 0; pos < P; ++pos)
        {
            if (!cls[pos].usable || reachable_without(g, g.S(), g.F(), pos))
                continue;
            int lo = 0, hi = 0;
            if (!walk_range(g, g.S(), pos, lo, hi))
                continue;
            order.emplace_back(lo, pos);
        }
        std::sort(order.begin(), order.end());
        int prev = g.S(), off_lo = 0, off_hi = 0;
        for (auto const & [min_off, pos] : order)
        {
            (void)min_off;
            int lo = 0, hi = 0;
            if (!walk_range(g, prev, pos, lo, hi) || hi >= kUnbounded)
                break; // distance unbounded from here on: later atoms cannot be tied to the match start
            if (prev != g.S())
            {
                ++lo; // symbols strictly between the two + the step onto `pos`
                ++hi;
            }
            off_lo += lo;
            off_hi += hi;
            chain.push_back(link{pos, off_lo, off_hi});
            prev = pos;
        }
    }
    if (chain.empty())
        return none;

    double const kGoodEnough = budget.good_enough;
    constexpr int kMaxStartSpread = 1024;
    int const kMaxAtoms = budget.max_atoms;
    atom_plan best;
    int best_lits = 99;
    int const n = static_cast<int>(chain.size());
    struct cand
    {
        int pos, lo, hi;
        double factor;
    };
    for (int anchor = 0; anchor < n; ++anchor)
    {
        if (chain[anchor].off_hi - chain[anchor].off_lo > kMaxStartSpread || cls[chain[anchor].pos].n == 0)
            continue;
        // address offsets of the other atoms relative to the anchor: forward scans look at later symbols (higher
        // addresses), reverse scans read downwards, so EARLIER symbols lie at higher addresses
        std::vector<cand> cands;
        for (int i = 0; i < n; ++i)
        {
            if (i == anchor || (reverse ? i > anchor : i < anchor) || cls[chain[i].pos].n == 0)
                continue;
            int lo, hi;
            if (!reverse)
            {
                lo = chain[i].off_lo - chain[anchor].off_lo; // sum of the gaps in between, gap-wise minimum / maximum
                hi = chain[i].off_hi - chain[anchor].off_hi;
            }
            else
            {
                lo = chain[anchor].off_lo - chain[i].off_lo;
                hi = chain[anchor].off_hi - chain[i].off_hi;
            }
            if (lo < 1 || hi < lo || hi > 31)
                continue;
            double const f = 1.0 - std::pow(1.0 - std::min(cls[chain[i].pos].freq, 1.0), hi - lo + 1); // any of the offsets
            if (f > 0.7)
                continue;
            cands.push_back(cand{chain[i].pos, lo, hi, f});
        }
        std::sort(cands.begin(), cands.end(), [](cand const & a, cand const & b) { return a.factor < b.factor; });
        int lits[kMultiLetters], n_lit = 0, n_terms = 0;
        auto lits_after = [&](small_class const & c)
        {
            int extra = 0;
            for (int j = 0; j < c.n; ++j)
            {
                bool known = false;
                for (int l = 0; l < n_lit; ++l)
                    known |= lits[l] == c.byte[j];
                extra += known ? 0 : 1;
            }
            return n_lit + extra;
        };
        auto take = [&](small_class const & c)
        {
            for (int j = 0; j < c.n; ++j)
            {
                bool known = false;
                for (int l = 0; l < n_lit; ++l)
                    known |= lits[l] == c.byte[j];
                if (!known)
                    lits[n_lit++] = c.byte[j];
            }
            n_terms += c.n;
        };
        small_class const & ac = cls[chain[anchor].pos];
        take(ac);
        std::vector<cand> chosen{cand{chain[anchor].pos, 0, 0, ac.freq}};
        double pass = ac.freq;
        for (cand const & cd : cands)
        {
            if (pass <= kGoodEnough || static_cast<int>(chosen.size()) >= kMaxAtoms)
                break;
            small_class const & c = cls[cd.pos];
            if (lits_after(c) > budget.max_lits || n_terms + c.n > kMaxAtomTerms)
                continue;
            take(c);
            chosen.push_back(cd);
            pass *= cd.factor;
        }
        // second stage: every other necessary position within reach of the anchor, most selective first
        struct check
        {
            int pos, lo, hi;
            double factor;
        };
        std::vector<check> checks;
        if (pass > kGoodEnough)
            for (int i = 0; i < n; ++i)
            {
                bool used = false;
                for (cand const & cd : chosen)
                    used |= cd.pos == chain[i].pos;
                if (used)
                    continue;
                // signed offset (in reading direction) of symbol i relative to the anchor's symbol
                int lo = chain[i].off_lo - chain[anchor].off_lo, hi = chain[i].off_hi - chain[anchor].off_hi;
                if (i < anchor)
                    std::swap(lo, hi); // before the anchor the gap-wise bounds come out reversed
                if (lo > hi)
                    continue;
                if (reverse) // reading direction = descending addresses
                {
                    int const t_lo = -hi, t_hi = -lo;
                    lo = t_lo;
                    hi = t_hi;
                }
                if (lo < -120 || hi > 120 || hi - lo > 7)
                    continue;
                double const f = 1.0 - std::pow(1.0 - std::min(cls[chain[i].pos].freq, 1.0), hi - lo + 1);
                if (f > 0.7)
                    continue;
                checks.push_back(check{chain[i].pos, lo, hi, f});
            }
        std::sort(checks.begin(), checks.end(), [](check const & a, check const & b) { return a.factor < b.factor; });
        double confirmed = pass;
        size_t n_checks = 0;
        while (n_checks < checks.size() && n_checks < static_cast<size_t>(kMaxConfirm) && confirmed > kGoodEnough)
            confirmed *= checks[n_checks++].factor;
        // The cost of a plan is its SIMD stage plus a few dependent L2 reads per surviving anchor: a usable plan first of
        // all, then the more selective SIMD stage, then — good enough is good enough — the cheaper one.
        auto usable = [](double simd, double conf) { return simd <= kMaxSimdPass && conf <= max_pass_rate(); };
        bool better;
        if (usable(pass, confirmed) != usable(best.pass, best.confirmed))
            better = usable(pass, confirmed);
        else if (!usable(pass, confirmed))
            better = confirmed < best.confirmed;
        else if (pass > kGoodEnough || best.pass > kGoodEnough)
            better = pass < best.pass;
        else
            better = n_lit < best_lits || (n_lit == best_lits && n_terms < best.spec.n_terms);
        if (!better)
            continue;
        atom_plan plan;
        plan.pass = pass;
        plan.confirmed = confirmed;
        plan.spec.n_confirm = static_cast<int32_t>(n_checks);
        for (size_t ci = 0; ci < n_checks; ++ci)
        {
            plan.spec.confirm[ci].lo = static_cast<int8_t>(checks[ci].lo);
            plan.spec.confirm[ci].hi = static_cast<int8_t>(checks[ci].hi);
            memcpy(plan.spec.confirm[ci].mask, cls[checks[ci].pos].mask, sizeof plan.spec.confirm[ci].mask);
        }
        plan.spec.n_lit = n_lit;
        plan.spec.n_atoms = static_cast<int32_t>(chosen.size());
        for (int l = 0; l < n_lit; ++l)
            plan.spec.splat[l] = 0x01010101u * static_cast<uint32_t>(lits[l]);
        int a = 0;
        for (cand const & cd : chosen)
        {
            small_class const & c = cls[cd.pos];
            for (int j = 0; j < c.n; ++j, ++a)
            {
                int l = 0;
                while (lits[l] != c.byte[j])
                    ++l;
                plan.spec.term_lit[a] = static_cast<uint8_t>(l);
                plan.spec.term_lo[a] = static_cast<uint8_t>(cd.lo);
                plan.spec.term_hi[a] = static_cast<uint8_t>(cd.hi);
                plan.spec.term_end[a] = j + 1 == c.n ? 1 : 0;
            }
        }
        plan.spec.n_terms = a;
        if (!reverse)
        {
            plan.spec.start_lo = -chain[anchor].off_hi;
            plan.spec.start_hi = -chain[anchor].off_lo;
        }
        else
        {
            plan.spec.start_lo = chain[anchor].off_lo;
            plan.spec.start_hi = chain[anchor].off_hi;
        }
        best = plan;
        best_lits = n_lit;
    }
    return best;
}

// the automaton as one scan sees it: comp_tab (reverse strand) and the reduced-alphabet map folded into B
void fill_image(trx_regex const & rx, bool reverse_strand, uint8_t const * map256, regex_dev * img)
{
    trx_regex const * re = &rx;
    memset(img, 0, sizeof(regex_dev));
    uint8_t comp[256];
    if (reverse_strand)
        complement_table(comp);
    for (int k = 0; k < kRegexWords; ++k)
    {
        img->first[k] = re->masks.first[k];
        img->last[k] = re->masks.last[k];
    }
    for (int ch = 0; ch < 256; ++ch)
    {
        uint8_t seen = static_cast<uint8_t>(ch);
        if (reverse_strand)
            seen = comp[seen];
        if (map256)
            seen = map256[seen];
        for (int k = 0; k < kRegexWords; ++k)
            img->B[ch][k] = re->masks.B[seen][k];
    }
    memcpy(img->follow, re->masks.follow, sizeof(img->follow));
    img->first_len = static_cast<uint16_t>(re->first_list.size());
    memcpy(img->first_list, re->first_list.data(), re->first_list.size());
    memcpy(img->follow_off, re->follow_off.data(), re->follow_off.size() * sizeof(uint16_t));
    memcpy(img->follow_list, re->follow_list.data(), re->follow_list.size());
    img->positions = re->positions;
    img->posix_longest = re->posix_longest;
    img->anchor_begin = re->anchor_begin;
    img->anchor_end = re->anchor_end;
}

// Subset construction over the position automaton of `img` WITHOUT re-injection: the matcher phase 2 runs from one start
// position.  Bytes with the same B column share a class; class 0 = bytes no position accepts.  Capped: a pattern whose
// subset automaton outgrows max_states keeps the position-automaton matcher.
constexpr uint32_t kDfaMaxStates = 2048;
constexpr size_t kDfaMaxBytes = 96 * 1024; // entries = bytes / 2 < 65536: pre-multiplied states fit 16 bits

constexpr uint32_t kReverseDfaMaxEntries = 16384; // 32 KB of shared memory

// The reverse searcher of dfa_table (see regex.cuh): subsets of positions reachable by reading backwards,
// X' = (pred(X) | last) & B[c], accepting when X meets `first`.
void build_reverse_dfa(regex_dev const & img, std::vector<std::pair<uint64_t, uint64_t>> const & cols, dfa_table & out)
{
    out.r_usable = false;
    out.r_trans.clear();
    if (img.anchor_begin || img.anchor_end || img.positions > 128) // subsets are kept as two words here
        return;
    int const n = img.positions;
    // longest match = longest path first → last in the follow graph; a cycle means unbounded
    std::vector<int> longest(n, 0), mark(n, 0); // mark: 0 new, 1 on stack, 2 done
    bool cyclic = false;
    std::function<int(int)> dfs = [&](int i) -> int {
        if (mark[i] == 2)
            return longest[i];
        if (mark[i] == 1)
        {
            cyclic = true;
            return 0;
        }
        mark[i] = 1;
        int best = 1;
        for (int j = 0; j < n && !cyclic; ++j)
            if ((img.follow[i][j >> 6] >> (j & 63)) & 1)
                best = std::max(best, 1 + dfs(j));
        mark[i] = 2;
        return longest[i] = best;
    };
    int max_len = 0;
    for (int i = 0; i < n && !cyclic; ++i)
        if ((img.first[i >> 6] >> (i & 63)) & 1)
            max_len = std::max(max_len, dfs(i));
    if (cyclic || max_len == 0 || max_len > 1024)
        return;
    uint64_t pred[128][2] = {};
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)
            if ((img.follow[i][j >> 6] >> (j & 63)) & 1)
                pred[j][i >> 6] |= 1ull << (i & 63);
    uint32_t const nc = static_cast<uint32_t>(cols.size());
    struct key
    {
        uint64_t a, b;
        bool operator==(key const & o) const { return a == o.a && b == o.b; }
    };
    struct key_hash
    {
        size_t operator()(key const & k) const { return static_cast<size_t>((k.a * 0x9E3779B97F4A7C15ull) ^ (k.b + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full); }
    };
    std::unordered_map<key, uint32_t, key_hash> id_of;
    std::vector<key> sets{key{0, 0}};
    id_of.emplace(key{0, 0}, 0);
    std::vector<uint32_t> trans(nc, 0);
    for (uint32_t s = 0; s < sets.size(); ++s)
    {
        if (sets.size() * nc > kReverseDfaMaxEntries)
            return;
        key const cur = sets[s];
        uint64_t p0 = img.last[0], p1 = img.last[1];
        for (int w = 0; w < 2; ++w)
        {
            uint64_t bits = w == 0 ? cur.a : cur.b;
            while (bits)
            {
                int const j = __builtin_ctzll(bits) + 64 * w;
                bits &= bits - 1;
                p0 |= pred[j][0];
                p1 |= pred[j][1];
            }
        }
        for (uint32_t c = 0; c < nc; ++c)
        {
            key const to{p0 & cols[c].first, p1 & cols[c].second};
            auto it = id_of.find(to);
            uint32_t id;
            if (it != id_of.end())
                id = it->second;
            else
            {
                id = static_cast<uint32_t>(sets.size());
                id_of.emplace(to, id);
                sets.push_back(to);
                trans.resize(trans.size() + nc, 0);
            }
            trans[static_cast<size_t>(s) * nc + c] = id;
        }
    }
    if (sets.size() * nc > kReverseDfaMaxEntries)
        return;
    uint32_t const ns = static_cast<uint32_t>(sets.size());
    std::vector<uint32_t> new_id(ns, 0);
    uint32_t next = 1; // state 0 (nothing live) keeps its number
    auto accepting = [&](uint32_t s) { return ((sets[s].a & img.first[0]) | (sets[s].b & img.first[1])) != 0; };
    for (uint32_t s = 1; s < ns; ++s)
        if (!accepting(s))
            new_id[s] = next++;
    out.r_first_accept = next;
    for (uint32_t s = 1; s < ns; ++s)
        if (accepting(s))
            new_id[s] = next++;
    out.r_states = ns;
    out.max_len = static_cast<uint32_t>(max_len);
    out.r_trans.assign(static_cast<size_t>(ns) * nc, 0);
    for (uint32_t s = 0; s < ns; ++s)
        for (uint32_t c = 0; c < nc; ++c)
            out.r_trans[static_cast<size_t>(new_id[s]) * nc + c] = static_cast<uint16_t>(new_id[trans[static_cast<size_t>(s) * nc + c]]);
    out.r_usable = true;
}

void build_dfa(regex_dev const & img, dfa_table & out)
{
    out.usable = false;
    out.trans.clear();
    if (!img.posix_longest || img.positions > 128) // subsets are kept as two words here; wider patterns run the position automaton
        return;
    // byte classes
    std::vector<std::pair<uint64_t, uint64_t>> cols; // class → column
    cols.push_back({0, 0});
    for (int ch = 0; ch < 256; ++ch)
    {
        std::pair<uint64_t, uint64_t> const col{img.B[ch][0], img.B[ch][1]};
        size_t k = 0;
        while (k < cols.size() && cols[k] != col)
            ++k;
        if (k == cols.size())
            cols.push_back(col);
        out.cls[ch] = static_cast<uint8_t>(k);
    }
    uint32_t const nc = static_cast<uint32_t>(cols.size());
    if (nc > 64)
        return;
    struct key
    {
        uint64_t a, b;
        bool operator==(key const & o) const { return a == o.a && b == o.b; }
    };
    struct key_hash
    {
        size_t operator()(key const & k) const { return static_cast<size_t>((k.a * 0x9E3779B97F4A7C15ull) ^ (k.b + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full); }
    };
    std::unordered_map<key, uint32_t, key_hash> id_of;
    std::vector<key> sets;      // provisional id → subset (ids 0 = dead, 1 = start are not subsets)
    sets.push_back({0, 0});
    sets.push_back({0, 0});
    std::vector<uint32_t> trans; // provisional ids
    trans.assign(2 * nc, 0);
    auto id_for = [&](key const & k) -> uint32_t {
        if ((k.a | k.b) == 0)
            return 0;
        auto it = id_of.find(k);
        if (it != id_of.end())
            return it->second;
        uint32_t const id = static_cast<uint32_t>(sets.size());
        id_of.emplace(k, id);
        sets.push_back(k);
        trans.resize(trans.size() + nc, 0);
        return id;
    };
    // start: S = first & B[c]
    for (uint32_t c = 1; c < nc; ++c)
    {
        uint32_t const to = id_for(key{img.first[0] & cols[c].first, img.first[1] & cols[c].second});
        trans[1 * nc + c] = to;
    }
    for (uint32_t s = 2; s < sets.size(); ++s)
    {
        if (sets.size() > kDfaMaxStates || sets.size() * nc * 2 > kDfaMaxBytes)
            return;
        key const cur = sets[s];
        uint64_t f0 = 0, f1 = 0;
        for (int w = 0; w < 2; ++w)
        {
            uint64_t bits = w == 0 ? cur.a : cur.b;
            while (bits)
            {
                int const i = __builtin_ctzll(bits);
                bits &= bits - 1;
                f0 |= img.follow[w * 64 + i][0];
                f1 |= img.follow[w * 64 + i][1];
            }
        }
        for (uint32_t c = 1; c < nc; ++c)
        {
            uint32_t const to = id_for(key{f0 & cols[c].first, f1 & cols[c].second}); // may grow sets / trans
            trans[static_cast<size_t>(s) * nc + c] = to;
        }
    }
    if (sets.size() > kDfaMaxStates || sets.size() * nc * 2 > kDfaMaxBytes)
        return;
    // renumber: accepting subsets last
    uint32_t const n = static_cast<uint32_t>(sets.size());
    std::vector<uint32_t> new_id(n, 0);
    uint32_t next = 2;
    new_id[0] = 0;
    new_id[1] = 1;
    auto accepting = [&](uint32_t s) { return ((sets[s].a & img.last[0]) | (sets[s].b & img.last[1])) != 0; };
    for (uint32_t s = 2; s < n; ++s)
        if (!accepting(s))
            new_id[s] = next++;
    out.first_accept = next;
    for (uint32_t s = 2; s < n; ++s)
        if (accepting(s))
            new_id[s] = next++;
    out.n_states = n;
    out.n_classes = nc;
    out.trans.assign(static_cast<size_t>(n) * nc, 0);
    for (uint32_t s = 0; s < n; ++s)
        for (uint32_t c = 0; c < nc; ++c)
            out.trans[static_cast<size_t>(new_id[s]) * nc + c] = static_cast<uint16_t>(new_id[trans[static_cast<size_t>(s) * nc + c]]);
    out.start_class.assign(nc, 0);
    for (uint32_t c = 0; c < nc; ++c)
        out.start_class[c] = out.trans[static_cast<size_t>(1) * nc + c] != 0;
    out.usable = true;
    build_reverse_dfa(img, cols, out);
}

// device form of the table: every state pre-multiplied by the class count (n_states * n_classes < 65536 by construction)
void stage_dfa_table(dfa_table const & d, uint16_t * dst)
{
    for (size_t i = 0; i < d.trans.size(); ++i)
        dst[i] = static_cast<uint16_t>(d.trans[i] * d.n_classes);
}

uint64_t hash_map256(uint8_t const * map256)
{
    if (!map256)
        return 0;
    uint64_t h = 1469598103934665603ull; // FNV-1a
    for (int i = 0; i < 256; ++i)
        h = (h ^ map256[i]) * 1099511628211ull;
    return h;
}

// the regex's matcher for this (strand, byte map), built on first use.  TRX_REGEX_DFA=off: never (A/B evidence, tests)
dfa_table const * dfa_for(trx_regex & re, regex_dev const & img, int reverse, uint64_t map_hash)
{
    char const * const env = getenv("TRX_REGEX_DFA");
    if ((env && strcmp(env, "off") == 0) || !re.posix_longest)
        return nullptr;
    for (dfa_table const & d : re.dfas)
        if (d.reverse == reverse && d.map_hash == map_hash)
            return d.usable ? &d : nullptr;
    if (re.dfas.size() >= 8)
        re.dfas.clear();
    re.dfas.emplace_back();
    dfa_table & d = re.dfas.back();
    d.reverse = reverse;
    d.map_hash = map_hash;
    build_dfa(img, d);
    return d.usable ? &d : nullptr;
}

enum filter_kind : int { kFilterNone = 0, kFilterWindow = 1, kFilterAtoms = 2 };

// Chooses phase 1 for one scan.  TRX_REGEX_FILTER = off | window | atoms restricts the choice (A/B evidence, tests).
// The fixed window costs ~0.5 ALU instructions per residue and literal, the gapped atoms ~1.5: the window is kept
// when it is selective enough on its own.
filter_kind choose_filter(regex_dev const & img, double const * freq, bool reverse, filter_spec & window, atom_spec & atoms)
{
    char const * filt = getenv("TRX_REGEX_FILTER");
    bool const filter_off = filt && strcmp(filt, "off") == 0;
    bool const only_window = filt && strcmp(filt, "window") == 0;
    bool const only_atoms = filt && strcmp(filt, "atoms") == 0;
    window = filter_spec{};
    atoms = atom_spec{};
    if (filter_off)
        return kFilterNone;
    double window_pass = 1.0;
    if (!only_atoms)
        window = derive_filter(img, freq, reverse, &window_pass);
    bool const have_window = window.n_a + window.n_b >= 1;
    if (!only_window && !(have_window && window_pass <= 1e-4))
    {
        atom_plan plan = derive_atoms(img, freq, reverse);
        // the SIMD stage alone may be weak (every surviving anchor costs a few dependent L2 reads in the confirm stage),
        // what phase 2 gets to see must not be
        if (plan.spec.n_lit >= 1 &&
            (only_atoms || (plan.pass <= kMaxSimdPass && plan.confirmed <= max_pass_rate() &&
                            (!have_window || plan.confirmed < 0.25 * window_pass))))
        {
            atoms = plan.spec;
            return kFilterAtoms;
        }
    }
    return have_window ? kFilterWindow : kFilterNone;
}

size_t align16(size_t x)
{
    return (x + 15) & ~size_t{15};
}

} // namespace

extern "C"
{

// trx_regex_scan_multi's pattern-at-a-time scans report into the call's shared hit list instead of the slot's own
struct hit_sink
{
    unsigned long long * count;
    uint64_t * keys;
    uint32_t * ends;
    uint32_t * pats;
    uint32_t pattern_id;
    uint64_t cap;
};

static int scan_begin_impl(trx_text_t * t, trx_regex_t * re, const uint64_t * candidate_bin_words, int reverse_strand,
                           const uint8_t * map256, uint64_t cap, int * ticket, hit_sink const * sink)
{
    if (!t || !re || !candidate_bin_words || !ticket)
        return set_error(TRX_EINVAL, "trx_regex_scan_begin: NULL argument");
    *ticket = -1;
    trx_ctx * c = t->ctx;
    TRX_TRY(c->use());
    int slot_index = -1;
    for (int i = 0; i < trx_text::kScanSlots && slot_index < 0; ++i)
        if (!t->slots[i].busy)
            slot_index = i;
    if (slot_index < 0)
        return set_error(TRX_ECAPACITY, "trx_regex_scan_begin: %d scans already in flight on this text (call trx_regex_scan_end)",
                         trx_text::kScanSlots);
    trx_text::scan_slot & sl = t->slots[slot_index];
    cudaStream_t const st = sl.stream;
    sl.cap = cap;
    sl.residues = 0;
    sl.filter_units = 0;
    // TRX_SCAN_TRACE=1: host-clock breakdown of this call on stderr (diagnostic)
    static bool const trace = getenv("TRX_SCAN_TRACE") != nullptr;
    auto const now = [] { return std::chrono::steady_clock::now(); };
    auto const t_begin = now();
    auto t_plan = t_begin, t_image = t_begin, t_choice = t_begin;
    if (re->follow_list.size() > sizeof(regex_dev::follow_list))
        return set_error(TRX_EREGEX, "pattern '%s': ordered follow lists need %zu bytes (max %zu)", re->pattern.c_str(),
                         re->follow_list.size(), sizeof(regex_dev::follow_list));

    // the clear of flags and counters does not depend on anything below: let it run while the host plans
    TRX_CUDA(cudaMemsetAsync(sl.d_block, 0, t->clear_bytes, st));

    // ---- plan: candidate bins → tiles of the text
    uint64_t const cand_words = (t->n_bins + 63) / 64;
    bool const same_plan = t->plan_valid && memcmp(t->plan_cand.data(), candidate_bin_words, cand_words * 8) == 0;
    if (!same_plan)
    {
        t->plan_valid = false;
        t->plan_cand.assign(candidate_bin_words, candidate_bin_words + cand_words);
        std::vector<uint32_t> & tl = t->plan_tiles;
        tl.clear();
        uint64_t residues = 0;
        for (uint64_t b = 0; b < t->n_bins; ++b)
        {
            if (!((candidate_bin_words[b >> 6] >> (b & 63)) & 1))
                continue;
            uint64_t r0 = t->h_bin_rec_begin[b], r1 = t->h_bin_rec_begin[b + 1];
            if (r0 == r1)
                continue;
            uint64_t g0 = t->h_rec_offsets[r0], g1 = t->h_rec_offsets[r1];
            if (g1 == g0)
                continue;
            residues += g1 - g0;
            for (uint64_t tile = g0 / kTileBytes; tile <= (g1 - 1) / kTileBytes; ++tile)
                if (tl.empty() || tl.back() != tile)
                    tl.push_back(static_cast<uint32_t>(tile));
        }
        t->plan_residues = residues;
        ++t->plan_serial;
        t->plan_valid = true;
    }
    if (sl.tiles_serial != t->plan_serial && !t->plan_tiles.empty())
    {
        // every slot keeps its own device copy of the tile list (a scan in flight on another slot may still read the old one);
        // pageable source: the copy is staged by the driver before the call returns
        TRX_TRY(sl.d_tiles.reserve(t->plan_tiles.size() * 4));
        TRX_CUDA(cudaMemcpyAsync(sl.d_tiles.ptr, t->plan_tiles.data(), t->plan_tiles.size() * 4, cudaMemcpyHostToDevice, st));
    }
    sl.tiles_serial = t->plan_serial;
    std::vector<uint32_t> const & tiles = t->plan_tiles;
    sl.residues = t->plan_residues;
    // even an empty plan goes through the whole protocol (the memset above left the slot's counters at zero)

    t_plan = now();
    // ---- one pinned staging block: [candidate words | automaton image], one H2D copy
    static_assert(sizeof(regex_dev) < 24576, "regex_dev grew unexpectedly");
    size_t const off_cand = 0;
    size_t const off_regex = align16(off_cand + cand_words * 8);
    size_t const off_dfa = align16(off_regex + sizeof(regex_dev)); // [256 class bytes | transitions] when there is a DFA
    sl.h_stage.pinned_host = true;
    TRX_TRY(sl.h_stage.reserve(off_dfa + 256 + kDfaMaxBytes + 16 + kReverseDfaMaxEntries * 2));
    TRX_TRY(sl.d_stage.reserve(off_dfa + 256 + kDfaMaxBytes + 16 + kReverseDfaMaxEntries * 2));
    auto * hs = static_cast<uint8_t *>(sl.h_stage.ptr);
    memcpy(hs + off_cand, candidate_bin_words, cand_words * 8);
    auto * img = reinterpret_cast<regex_dev *>(hs + off_regex);
    fill_image(*re, reverse_strand != 0, map256, img);
    uint64_t const map_hash = hash_map256(map256);
    dfa_table const * const dfa = dfa_for(*re, *img, reverse_strand, map_hash);
    size_t stage_bytes = off_dfa;
    size_t off_rdfa = 0;
    if (dfa)
    {
        memcpy(hs + off_dfa, dfa->cls, 256);
        stage_dfa_table(*dfa, reinterpret_cast<uint16_t *>(hs + off_dfa + 256));
        stage_bytes = off_dfa + 256 + dfa->trans.size() * 2;
        if (dfa->r_usable) // the reverse searcher: a few KB, staged whether or not this scan ends up using it
        {
            off_rdfa = align16(stage_bytes);
            uint16_t * dst = reinterpret_cast<uint16_t *>(hs + off_rdfa);
            for (size_t i = 0; i < dfa->r_trans.size(); ++i)
                dst[i] = static_cast<uint16_t>(dfa->r_trans[i] * dfa->n_classes);
            stage_bytes = off_rdfa + dfa->r_trans.size() * 2;
        }
    }
    t_image = now();

    if (cap > (1ull << 31))
        cap = 1ull << 31;
    uint64_t const dev_cap = cap ? cap : 1;
    TRX_TRY(sl.d_keys.reserve(dev_cap * 8));
    TRX_TRY(sl.d_ends.reserve(dev_cap * 4));
    auto * ds = static_cast<uint8_t *>(sl.d_stage.ptr);
    TRX_CUDA(cudaMemcpyAsync(ds, hs, stage_bytes, cudaMemcpyHostToDevice, st));
    scan_params sp{};
    sp.text = t->d_text;
    sp.n_text = t->n_text;
    sp.n_text_padded = t->n_text_padded;
    sp.rec_offsets = t->d_rec_offsets;
    sp.rec_bin = t->d_rec_bin;
    sp.n_records = t->n_records;
    sp.cand_words = reinterpret_cast<uint64_t *>(ds + off_cand);
    sp.tiles = static_cast<uint32_t *>(sl.d_tiles.ptr);
    sp.n_tiles = static_cast<uint32_t>(tiles.size());
    sp.chunk_flags = sl.d_chunk_flags;
    sp.rec_flags = sl.d_rec_flags;
    sp.rec_list = sl.d_rec_list;
    sp.rec_count = reinterpret_cast<unsigned int *>(sl.d_counters + 8);
    sp.tile_counter = reinterpret_cast<unsigned int *>(sl.d_counters + 12);
    sp.re = reinterpret_cast<regex_dev *>(ds + off_regex);

    // TRX_REGEX_STAGING=ldg  → exact automaton kernel with direct loads (A/B evidence for the TMA staging)
    // TRX_REGEX_FILTER=off   → never use a SIMD prefilter (A/B evidence for the filters)
    char const * staging = getenv("TRX_REGEX_STAGING");
    bool const use_tma = !(staging && strcmp(staging, "ldg") == 0);
    bool const reverse = reverse_strand != 0;
    filter_kind kind = kFilterNone;
    if (use_tma)
    {
        char const * filt = getenv("TRX_REGEX_FILTER");
        int const mode = !filt ? 0 : (filt[0] == 'o' ? 1 : (filt[0] == 'w' ? 2 : (filt[0] == 'a' ? 3 : 4)));
        trx_text::cached_choice const * found = nullptr;
        for (auto const & cc : t->choices)
            if (cc.serial == re->serial && cc.reverse == reverse_strand && cc.mode == mode && cc.map_hash == map_hash)
                found = &cc;
        if (!found)
        {
            trx_text::cached_choice cc{};
            cc.serial = re->serial;
            cc.map_hash = map_hash;
            cc.reverse = reverse_strand;
            cc.mode = mode;
            cc.kind = choose_filter(*img, t->byte_freq, reverse, cc.window, cc.atoms);
            if (t->choices.size() >= 4096)
                t->choices.clear();
            t->choices.push_back(cc);
            found = &t->choices.back();
        }
        kind = static_cast<filter_kind>(found->kind);
        sp.filter = found->window;
        sp.atoms = found->atoms;
    }
    bool const use_atoms = kind == kFilterAtoms, use_window = kind == kFilterWindow;
    // no selective necessary factor (or TRX_REGEX_FILTER=off), but determinised automata: phase 1 need not be the position
    // automaton.  TRX_REGEX_DFA=off keeps the position-automaton kernels reachable for A/B runs and tests.
    char const * const filt_env = getenv("TRX_REGEX_FILTER");
    bool const skip_phase1 = kind == kFilterNone && dfa != nullptr && use_tma && (!filt_env || strcmp(filt_env, "off") == 0);
    bool use_start_bits = false;
    // Which no-prefilter route.  Measured on the C2 text (profiles/time_nofilter.py): the reverse searcher costs 0.32 ms per
    // 204 MB whatever the pattern, as long as its table is small (several CTAs per SM) and the warm-up behind each chunk
    // short; the position automaton 0.75-1.0 ms; trying every start in phase 2 1.0-3.4 ms (one dependent global load per
    // window of 32 starts).  So: reverse searcher when it qualifies, otherwise the position automaton in phase 1 with the
    // DFA in phase 2; "every start" stays reachable for tests (TRX_REGEX_STARTS=off).
    bool const starts_off = getenv("TRX_REGEX_STARTS") && strcmp(getenv("TRX_REGEX_STARTS"), "off") == 0;
    bool const starts_route = skip_phase1 && !starts_off && dfa->r_usable && dfa->r_trans.size() * 2 <= 8192 && dfa->max_len <= 64;
    bool const every_start_route = skip_phase1 && starts_off;
    t_choice = now();
    sl.filter_units = use_atoms ? static_cast<uint64_t>(sp.atoms.n_atoms)
                                : (use_window ? static_cast<uint64_t>(sp.filter.n_a + sp.filter.n_b) : 0);
    if (tiles.empty())
        ; // nothing to scan: no kernel, the read-back below returns the zeroed counters
    else if (starts_route)
    {
        // exact match starts from the reverse searcher, one bit per text byte
        uint64_t const bits_bytes = t->n_text_padded / 8;
        TRX_TRY(sl.d_start_bits.reserve(bits_bytes));
        TRX_CUDA(cudaMemsetAsync(sl.d_start_bits.ptr, 0, bits_bytes, st));
        sp.start_bits = static_cast<uint32_t *>(sl.d_start_bits.ptr);
        sp.r_trans = reinterpret_cast<uint16_t const *>(ds + off_rdfa);
        sp.r_cls = ds + off_dfa;
        sp.r_classes = dfa->n_classes;
        sp.r_states = dfa->r_states;
        sp.r_first_accept = dfa->r_first_accept;
        sp.r_max_len = dfa->max_len;
        use_start_bits = true;
        size_t const smem = kStages * kTileBytes + 256 + 16 + ((dfa->r_trans.size() * 2 + 15) & ~size_t{15});
        static size_t attr_smem_of[64] = {};
        size_t & attr_smem = attr_smem_of[c->device & 63];
        if (smem > attr_smem)
        {
            TRX_CUDA(cudaFuncSetAttribute(regex_scan_starts_dfa, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
            attr_smem = smem;
        }
        unsigned const grid = static_cast<unsigned>(std::min<uint64_t>(sp.n_tiles, static_cast<uint64_t>(c->sm_count) * 2));
        regex_scan_starts_dfa<<<grid, kScanThreads, smem, st>>>(sp, t->tmap);
        ++c->launches;
        TRX_CUDA(cudaGetLastError());
    }
    else if (every_start_route)
    {
        list_candidate_records_kernel<<<grid_for(t->n_records, 256, c->sm_count, 8), 256, 0, st>>>(
            t->d_rec_bin, sp.cand_words, t->n_records, sl.d_rec_list, sp.rec_count);
        ++c->launches;
        TRX_CUDA(cudaGetLastError());
    }
    else if (use_atoms)
        TRX_TRY(launch_atoms_any(c, st, sp, t->tmap));
    else if (use_window)
        TRX_TRY(launch_filter_any(c, st, sp, t->tmap));
    else if (re->words > 2)
        TRX_TRY(launch_phase1_w<4>(c, st, sp, t->tmap, reverse, use_tma));
    else if (re->words == 1)
        TRX_TRY(launch_phase1_w<1>(c, st, sp, t->tmap, reverse, use_tma));
    else
        TRX_TRY(launch_phase1_w<2>(c, st, sp, t->tmap, reverse, use_tma));

    resolve_params rp{};
    rp.text = t->d_text;
    rp.rec_offsets = t->d_rec_offsets;
    rp.rec_list = sl.d_rec_list;
    rp.rec_count = sp.rec_count;
    rp.chunk_flags = every_start_route ? nullptr : sl.d_chunk_flags;
    rp.start_bits = use_start_bits ? static_cast<uint32_t const *>(sl.d_start_bits.ptr) : nullptr;
    rp.re = sp.re;
    rp.anchor_begin = re->anchor_begin;
    rp.anchor_end = re->anchor_end;
    if (dfa)
    {
        rp.dfa_cls = ds + off_dfa;
        rp.dfa_trans = reinterpret_cast<uint16_t const *>(ds + off_dfa + 256);
        rp.dfa_classes = dfa->n_classes;
        rp.dfa_states = dfa->n_states;
        rp.dfa_first_accept = dfa->first_accept;
    }
    rp.hit_count = reinterpret_cast<unsigned long long *>(sl.d_counters);
    rp.hit_keys = static_cast<uint64_t *>(sl.d_keys.ptr);
    rp.hit_ends = static_cast<uint32_t *>(sl.d_ends.ptr);
    rp.small_keys = reinterpret_cast<uint64_t *>(sl.d_counters + 16);
    rp.small_ends = reinterpret_cast<uint32_t *>(sl.d_counters + 16 + kSmallHits * 8);
    rp.cap = cap;
    if (sink)
    {
        rp.hit_count = sink->count;
        rp.hit_keys = sink->keys;
        rp.hit_ends = sink->ends;
        rp.hit_pats = sink->pats;
        rp.pattern_id = sink->pattern_id;
        rp.cap = sink->cap;
    }
    if (tiles.empty())
        ;
    else if (dfa)
    {
        unsigned const smem = dfa_smem_bytes(dfa->n_states, dfa->n_classes);
        unsigned const grid = static_cast<unsigned>(c->sm_count) * 16;
        if (reverse)
            regex_scan_phase2_dfa<-1><<<grid, 128, smem, st>>>(rp);
        else
            regex_scan_phase2_dfa<1><<<grid, 128, smem, st>>>(rp);
        ++c->launches;
        TRX_CUDA(cudaGetLastError());
    }
    else if (re->words > 2)
        TRX_TRY(launch_phase2_w<4>(c, st, rp, reverse));
    else if (re->words == 1)
        TRX_TRY(launch_phase2_w<1>(c, st, rp, reverse));
    else
        TRX_TRY(launch_phase2_w<2>(c, st, rp, reverse));

    // ---- read back: the counters and the first kSmallHits hits in ONE copy
    constexpr uint64_t kSmall = kSmallHits;
    sl.h_small.pinned_host = true;
    TRX_TRY(sl.h_small.reserve(16 + kSmall * 12));
    auto * h_total = static_cast<unsigned long long *>(sl.h_small.ptr);
    TRX_CUDA(cudaMemcpyAsync(h_total, sl.d_counters, 16 + kSmall * 12, cudaMemcpyDeviceToHost, st));
    TRX_CUDA(cudaEventRecord(sl.done, st));
    sl.busy = true;
    *ticket = slot_index;
    if (trace)
    {
        auto us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
        fprintf(stderr, "trx_regex_scan_begin: plan %.1f  image %.1f  filter choice %.1f  enqueue %.1f us (%zu tiles, slot %d)\n",
                us(t_begin, t_plan), us(t_plan, t_image), us(t_image, t_choice), us(t_choice, now()), tiles.size(), slot_index);
    }
    return TRX_OK;
}

int trx_regex_scan_begin(trx_text_t * t, trx_regex_t * re, const uint64_t * candidate_bin_words, int reverse_strand,
                         const uint8_t * map256, uint64_t cap, int * ticket)
{
    return scan_begin_impl(t, re, candidate_bin_words, reverse_strand, map256, cap, ticket, nullptr);
}

int trx_regex_scan_end(trx_text_t * t, int ticket, trx_hit * hits, uint64_t * n_hits, uint64_t * stats)
{
    if (!t || ticket < 0 || ticket >= trx_text::kScanSlots || !t->slots[ticket].busy)
        return set_error(TRX_EINVAL, "trx_regex_scan_end: bad ticket");
    trx_text::scan_slot & sl = t->slots[ticket];
    if (!hits && !n_hits)
    {
        // abandon: the caller gave up on this scan (error path); wait it out and free the slot
        cudaEventSynchronize(sl.done);
        sl.busy = false;
        return TRX_OK;
    }
    if (!n_hits)
        return set_error(TRX_EINVAL, "trx_regex_scan_end: n_hits is NULL");
    uint64_t const cap = sl.cap;
    if (cap && !hits)
        return set_error(TRX_EINVAL, "trx_regex_scan_end: hits is NULL");
    trx_ctx * c = t->ctx;
    TRX_TRY(c->use());
    cudaStream_t const st = sl.stream;
    *n_hits = 0;
    if (stats)
    {
        stats[0] = sl.residues;
        stats[1] = sl.filter_units;
    }
    cudaError_t const waited = cudaEventSynchronize(sl.done);
    sl.busy = false; // whatever happens below, the slot is free again
    TRX_CUDA(waited);
    constexpr uint64_t kSmall = kSmallHits;
    uint64_t const spec = cap < kSmall ? cap : kSmall;
    auto * h_total = static_cast<unsigned long long *>(sl.h_small.ptr);
    auto * h_keys = reinterpret_cast<uint64_t *>(static_cast<uint8_t *>(sl.h_small.ptr) + 16);
    auto * h_ends = reinterpret_cast<uint32_t *>(static_cast<uint8_t *>(sl.h_small.ptr) + 16 + kSmall * 8);
    uint64_t * const d_hit_keys = static_cast<uint64_t *>(sl.d_keys.ptr);
    uint32_t * const d_hit_ends = static_cast<uint32_t *>(sl.d_ends.ptr);
    uint64_t * const d_small_keys = reinterpret_cast<uint64_t *>(sl.d_counters + 16);
    uint32_t * const d_small_ends = reinterpret_cast<uint32_t *>(sl.d_counters + 16 + kSmallHits * 8);
    unsigned long long const total = *h_total;
    *n_hits = total;
    if (total > cap)
        return set_error(TRX_ECAPACITY, "trx_regex_scan: %llu hits, room for %llu", total, (unsigned long long)cap);
    if (total == 0)
        return TRX_OK;
    if (total <= spec)
    {
        // a handful of hits: put them into print order right here instead of launching a device sort
        std::vector<std::pair<uint64_t, uint32_t>> v(total);
        for (uint64_t i = 0; i < total; ++i)
            v[i] = {h_keys[i], h_ends[i]};
        std::sort(v.begin(), v.end());
        for (uint64_t i = 0; i < total; ++i)
        {
            uint32_t rec = static_cast<uint32_t>(v[i].first >> 32);
            hits[i] = trx_hit{t->h_rec_bin[rec], rec, static_cast<uint32_t>(v[i].first), v[i].second};
        }
        return TRX_OK;
    }

    // ---- order = (bin, record, start): records are grouped by ascending bin, so (record, start) suffices
    uint64_t const n = total;
    TRX_CUDA(cudaMemcpyAsync(d_hit_keys, d_small_keys, kSmall * 8, cudaMemcpyDeviceToDevice, st));
    TRX_CUDA(cudaMemcpyAsync(d_hit_ends, d_small_ends, kSmall * 4, cudaMemcpyDeviceToDevice, st));
    TRX_TRY(t->d_keys2.reserve(n * 8));
    TRX_TRY(t->d_ends2.reserve(n * 4));
    size_t tmp_bytes = 0;
    TRX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_hit_keys, static_cast<uint64_t *>(t->d_keys2.ptr),
                                             d_hit_ends, static_cast<uint32_t *>(t->d_ends2.ptr), static_cast<int>(n), 0, 64,
                                             st));
    TRX_TRY(t->d_sort_tmp.reserve(tmp_bytes));
    TRX_CUDA(cub::DeviceRadixSort::SortPairs(t->d_sort_tmp.ptr, tmp_bytes, d_hit_keys,
                                             static_cast<uint64_t *>(t->d_keys2.ptr), d_hit_ends,
                                             static_cast<uint32_t *>(t->d_ends2.ptr), static_cast<int>(n), 0, 64, st));
    c->launches += 3; // cub's histogram + onesweep passes
    // the sorted pairs become trx_hit records on the device (the unsorted key buffer is free again: n x 8 B of the n x 16 B
    // needed, the rest comes from the ends buffer's neighbour) and reach the caller's array in ONE copy
    static_assert(sizeof(trx_hit) == 16, "format_hits_kernel writes uint4");
    TRX_TRY(t->d_hits_fmt.reserve(n * sizeof(trx_hit)));
    format_hits_kernel<<<grid_for(n, 256, c->sm_count, 8), 256, 0, st>>>(static_cast<uint64_t const *>(t->d_keys2.ptr),
                                                                           static_cast<uint32_t const *>(t->d_ends2.ptr), t->d_rec_bin,
                                                                           n, static_cast<uint4 *>(t->d_hits_fmt.ptr));
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    TRX_CUDA(cudaMemcpyAsync(hits, t->d_hits_fmt.ptr, n * sizeof(trx_hit), cudaMemcpyDeviceToHost, st));
    TRX_CUDA(cudaStreamSynchronize(st));
    return TRX_OK;
}

int trx_regex_scan(trx_text_t * t, trx_regex_t * re, const uint64_t * candidate_bin_words, int reverse_strand,
                   const uint8_t * map256, trx_hit * hits, uint64_t cap, uint64_t * n_hits, uint64_t * stats)
{
    if (!t || !re || !candidate_bin_words || !n_hits || (cap && !hits))
        return set_error(TRX_EINVAL, "trx_regex_scan: NULL argument");
    *n_hits = 0;
    int ticket = -1;
    TRX_TRY(trx_regex_scan_begin(t, re, candidate_bin_words, reverse_strand, map256, cap, &ticket));
    return trx_regex_scan_end(t, ticket, hits, n_hits, stats);
}

// host-only (tests): the reverse searcher run over text[0, len) as ONE record: starts[p] = 1 iff some match starts at p.
// Returns the longest possible match length, 0 if the pattern has no reverse searcher.
uint32_t trx_regex_dfa_starts(trx_regex_t * re, const uint8_t * text, uint64_t len, uint8_t * starts)
{
    if (!re || (len && (!text || !starts)))
        return 0;
    auto img = std::make_unique<regex_dev>();
    fill_image(*re, false, nullptr, img.get());
    dfa_table const * d = dfa_for(*re, *img, 0, 0);
    if (!d || !d->r_usable)
        return 0;
    uint32_t st = 0;
    for (uint64_t p = len; p-- > 0;)
    {
        st = d->r_trans[static_cast<size_t>(st) * d->n_classes + d->cls[text[p]]];
        starts[p] = st >= d->r_first_accept;
    }
    return d->max_len;
}

int trx_regex_dfa_info(trx_regex_t * re, uint32_t info[3])
{
    if (!re || !info)
        return 0;
    auto img = std::make_unique<regex_dev>();
    fill_image(*re, false, nullptr, img.get());
    dfa_table const * d = dfa_for(*re, *img, 0, 0);
    info[0] = d ? d->n_states : 0;
    info[1] = d ? d->n_classes : 0;
    info[2] = d ? d->first_accept : 0;
    return d ? 1 : 0;
}

// the device's match_longest_dfa (regex_kernels.cuh) on the host, for CPU-side parity tests of the subset construction
int64_t trx_regex_dfa_match_end(trx_regex_t * re, const uint8_t * text, uint64_t len, uint64_t start)
{
    if (!re || (len && !text))
        return -2;
    auto img = std::make_unique<regex_dev>();
    fill_image(*re, false, nullptr, img.get());
    dfa_table const * d = dfa_for(*re, *img, 0, 0);
    if (!d)
        return -2;
    uint32_t st = 1;
    int64_t best = -1;
    for (uint64_t pos = start; pos < len;)
    {
        st = d->trans[static_cast<size_t>(st) * d->n_classes + d->cls[text[pos]]];
        if (st == 0)
            break;
        ++pos;
        if (st >= d->first_accept && (!re->anchor_end || pos == len))
            best = static_cast<int64_t>(pos);
    }
    return best;
}

int trx_regex_filter_plan(const trx_regex_t * re, const double * byte_freq, int reverse_strand, const uint8_t * map256,
                          int32_t out[136])
{
    if (!re || !byte_freq || !out)
        return set_error(TRX_EINVAL, "trx_regex_filter_plan: NULL argument");
    auto img = std::make_unique<regex_dev>();
    fill_image(*re, reverse_strand != 0, map256, img.get());
    filter_spec window;
    atom_spec atoms;
    filter_kind const kind = choose_filter(*img, byte_freq, reverse_strand != 0, window, atoms);
    memset(out, 0, 136 * sizeof(int32_t));
    out[0] = kind;
    if (kind == kFilterWindow)
    {
        out[1] = window.n_a;
        out[2] = window.n_b;
        out[3] = window.start_delta;
        for (int j = 0; j < window.n_a + window.n_b; ++j)
        {
            out[8 + j] = static_cast<int32_t>(window.delta[j]) + (j >= window.n_a ? 4 : 0); // byte offset inside the window
            out[16 + j] = static_cast<int32_t>(window.splat[j] & 0xFF);
        }
    }
    else if (kind == kFilterAtoms)
    {
        out[1] = atoms.n_lit;
        out[2] = atoms.n_terms;
        out[3] = atoms.start_lo;
        out[4] = atoms.start_hi;
        out[5] = atoms.n_atoms;
        int group = 0;
        for (int a = 0; a < atoms.n_terms; ++a)
        {
            out[8 + a] = static_cast<int32_t>(atoms.splat[atoms.term_lit[a]] & 0xFF);
            out[24 + a] = atoms.term_lo[a];
            out[40 + a] = atoms.term_hi[a];
            out[56 + a] = group;
            group += atoms.term_end[a];
        }
        out[6] = atoms.n_confirm;
        for (int c = 0; c < atoms.n_confirm; ++c)
        {
            out[72 + 10 * c] = atoms.confirm[c].lo;
            out[73 + 10 * c] = atoms.confirm[c].hi;
            for (int w = 0; w < 8; ++w)
                out[74 + 10 * c + w] = static_cast<int32_t>(atoms.confirm[c].mask[w]);
        }
    }
    return TRX_OK;
}

} // extern "C"


// ------------------------------------------------------------------------------------------------ multi-pattern scan
namespace
{

// the cached phase-1 choice of a pattern scanned forward, unmapped, in the default mode — without a walk over the cache
trx_text::cached_choice * plain_choice_of(trx_text * t, uint64_t serial)
{
    auto const is_it = [&](trx_text::cached_choice const & cc) {
        return cc.serial == serial && cc.reverse == 0 && cc.mode == 0 && cc.map_hash == 0;
    };
    auto const it = t->plain_choice.find(serial);
    if (it != t->plain_choice.end() && it->second < t->choices.size() && is_it(t->choices[it->second]))
        return &t->choices[it->second];
    for (size_t i = t->choices.size(); i-- > 0;) // made by trx_regex_scan_begin, or the cache was cleared since
        if (is_it(t->choices[i]))
        {
            t->plain_choice[serial] = static_cast<uint32_t>(i);
            return &t->choices[i];
        }
    return nullptr;
}

// Class window of a pattern for regex_window_multi: up to kWindowLen CONSECUTIVE positions of the automaton that every
// match passes through (dominators, each with its successor as the only way on), with the byte class of each.  "The text
// carries class_0 .. class_{len-1} at p .. p+len-1" is then necessary for a match whose first byte lies in
// [p - off_hi, p - off_lo].  The most selective such run (byte frequencies of the text) is taken; ok = selective enough to
// hand its survivors to phase 2 directly.
struct class_window
{
    bool ok{false};
    int len{0};
    int off_lo{0}, off_hi{0};
    uint32_t cls[kWindowLen][8]; // 256-bit byte sets
    double pass{1.0};            // expected fraction of text positions the window lets through …
    double confirmed{1.0};       // … and of those the confirm checks keep
    int n_confirm{0};
    confirm_check confirm[kMaxConfirm]; // other necessary positions, offsets relative to the window's first byte
};

constexpr size_t kFewLeftovers = 24;         // patterns without a class window, when most of the batch has one
constexpr double kMaxWindowPass = 2e-3;      // survivors per text position and pattern the window stage may hand to its confirm checks
constexpr double kMaxWindowConfirmed = 1.2e-4; // … and to phase 2 (its record lists hold kMultiListCap entries; a longer list sends the pattern to the single-pattern path)

class_window derive_class_window(regex_dev const & img, double const * freq)
{
    class_window none;
    automaton_graph const g = build_graph(img);
    int const P = g.P;
    if (P <= 0 || img.anchor_begin || img.anchor_end)
        return none;
    std::vector<double> p_of(P, 0.0);
    for (int q = 0; q < P; ++q)
        for (int ch = 0; ch < 256; ++ch)
            if ((img.B[ch][q >> 6] >> (q & 63)) & 1)
                p_of[q] += freq[ch];
    std::vector<char> dom(P, 0);
    for (int q = 0; q < P; ++q)
        dom[q] = !reachable_without(g, g.S(), g.F(), q);
    struct cand
    {
        double pass;
        int first, len;
    };
    std::vector<cand> cands;
    for (int q = 0; q < P; ++q)
    {
        if (!dom[q])
            continue;
        double pass = 1.0;
        int cur = q;
        for (int len = 1; len <= kWindowLen; ++len)
        {
            pass *= p_of[cur];
            cands.push_back({pass, q, len});
            if (g.succ[cur].size() != 1 || g.succ[cur][0] == g.F() || g.succ[cur][0] == cur || !dom[g.succ[cur][0]])
                break;
            cur = g.succ[cur][0];
        }
    }
    std::sort(cands.begin(), cands.end(), [](cand const & a, cand const & b) { return a.pass != b.pass ? a.pass < b.pass : a.len < b.len; });
    // Confirm checks for a window whose first position is f (entered for the first time at text position p): a dominator
    // d that every walk reaches only through f has its first visit after that, 1 + k bytes behind p with k in
    // walk_range(f, d); a dominator d that every walk to f passes has its last visit before p at p - 1 - k, k in
    // walk_range(d, f).  Both are necessary for a match — evaluated only where the window holds.
    struct check
    {
        int pos, lo, hi;
        double factor;
    };
    auto checks_for = [&](int first, int len) {
        std::vector<char> in_window(P, 0);
        int cur = first;
        for (int i = 0; i < len; ++i)
        {
            in_window[cur] = 1;
            if (i + 1 < len)
                cur = g.succ[cur][0];
        }
        std::vector<check> out;
        for (int d = 0; d < P; ++d)
        {
            if (!dom[d] || in_window[d] || p_of[d] > 0.6)
                continue;
            int lo = 0, hi = 0;
            check c{d, 0, 0, 1.0};
            if (!reachable_without(g, g.S(), d, first))
            {
                if (!walk_range(g, first, d, lo, hi) || hi >= kUnbounded)
                    continue;
                c.lo = 1 + lo;
                c.hi = 1 + hi;
            }
            else if (!reachable_without(g, g.S(), first, d))
            {
                if (!walk_range(g, d, first, lo, hi) || hi >= kUnbounded)
                    continue;
                c.lo = -1 - hi;
                c.hi = -1 - lo;
            }
            else
                continue;
            if (c.hi - c.lo > 23 || c.lo < -120 || c.hi > 120)
                continue;
            c.factor = std::min(1.0, (c.hi - c.lo + 1) * p_of[d]);
            if (c.factor < 0.7)
                out.push_back(c);
        }
        std::sort(out.begin(), out.end(), [](check const & a, check const & b) { return a.factor < b.factor; });
        return out;
    };
    for (cand const & c : cands)
    {
        if (c.pass > kMaxWindowPass)
            break;
        int lo = 0, hi = 0;
        if (!walk_range(g, g.S(), c.first, lo, hi) || hi >= kUnbounded || hi - lo > 31 || hi > 200)
            continue;
        class_window w;
        w.confirmed = c.pass;
        if (c.pass > kMaxWindowConfirmed)
        {
            std::vector<check> const checks = checks_for(c.first, c.len);
            for (check const & k : checks)
            {
                if (w.n_confirm == kMaxConfirm || w.confirmed <= 1e-6)
                    break;
                confirm_check & out = w.confirm[w.n_confirm++];
                memset(&out, 0, sizeof out);
                out.lo = static_cast<int8_t>(k.lo);
                out.hi = static_cast<int8_t>(k.hi);
                for (int ch = 0; ch < 256; ++ch)
                    if ((img.B[ch][k.pos >> 6] >> (k.pos & 63)) & 1)
                        out.mask[ch >> 5] |= 1u << (ch & 31);
                w.confirmed *= k.factor;
            }
            if (w.confirmed > kMaxWindowConfirmed)
                continue;
        }
        w.ok = true;
        w.len = c.len;
        w.off_lo = lo;
        w.off_hi = hi;
        w.pass = c.pass;
        memset(w.cls, 0, sizeof w.cls);
        int cur = c.first;
        for (int i = 0; i < c.len; ++i)
        {
            for (int ch = 0; ch < 256; ++ch)
                if ((img.B[ch][cur >> 6] >> (cur & 63)) & 1)
                    w.cls[i][ch >> 5] |= 1u << (ch & 31);
            if (i + 1 < c.len)
                cur = g.succ[cur][0];
        }
        return w;
    }
    return none;
}

struct multi_plan
{
    multi_program prog{};           // terms / start range / confirm checks filled in, pointers not yet
    bool usable{false};
};

// One pattern's phase-1 plan (the cached window / atoms choice of the single-pattern path) as a term list over the
// batch's alphabet.  `letter_of[byte]` = index into the alphabet, growing as new bytes appear; false = the alphabet is
// full, or the pattern has no prefilter at all (exact automaton only): such patterns take the single-pattern path.
bool plan_for_multi(filter_kind kind, filter_spec const & window, atom_spec const & atoms, int letter_of[256],
                    uint32_t splat[kMultiLetters], uint32_t & n_letters, multi_plan & out)
{
    int local_of[256];
    memcpy(local_of, letter_of, sizeof local_of);
    uint32_t local_n = n_letters;
    uint32_t local_splat[kMultiLetters];
    memcpy(local_splat, splat, sizeof local_splat);
    auto letter = [&](uint32_t byte) -> int {
        if (local_of[byte] < 0)
        {
            if (local_n >= static_cast<uint32_t>(kMultiLetters) || byte == 0)
                return -1;
            local_of[byte] = static_cast<int>(local_n);
            local_splat[local_n++] = byte * 0x01010101u;
        }
        return local_of[byte];
    };
    multi_program & P = out.prog;
    P = multi_program{};
    if (kind == kFilterWindow)
    {
        int const n = window.n_a + window.n_b;
        for (int j = 0; j < n; ++j)
        {
            int const l = letter(window.splat[j] & 0xFF);
            if (l < 0)
                return false;
            uint32_t const off = window.delta[j] + (j >= window.n_a ? 4 : 0); // byte offset inside the 8-byte window
            P.terms.letter[j] = static_cast<uint8_t>(l);
            P.terms.lo[j] = P.terms.hi[j] = static_cast<uint8_t>(off);
            P.terms.end[j] = 1;
        }
        P.terms.n_terms = n;
        P.start_lo = P.start_hi = window.start_delta;
        P.n_confirm = 0;
    }
    else if (kind == kFilterAtoms)
    {
        for (int i = 0; i < atoms.n_terms; ++i)
        {
            int const l = letter(atoms.splat[atoms.term_lit[i]] & 0xFF);
            if (l < 0)
                return false;
            P.terms.letter[i] = static_cast<uint8_t>(l);
            P.terms.lo[i] = atoms.term_lo[i];
            P.terms.hi[i] = atoms.term_hi[i];
            P.terms.end[i] = atoms.term_end[i];
        }
        P.terms.n_terms = atoms.n_terms;
        P.start_lo = atoms.start_lo;
        P.start_hi = atoms.start_hi;
        P.n_confirm = atoms.n_confirm;
        memcpy(P.confirm, atoms.confirm, sizeof P.confirm);
    }
    else
        return false;
    if (P.terms.n_terms <= 0)
        return false;
    int expanded = 0;
    for (int i = 0; i < P.terms.n_terms; ++i)
        expanded += P.terms.hi[i] >= P.terms.lo[i] ? P.terms.hi[i] - P.terms.lo[i] + 1 : 0;
    if (expanded + 3 > kMultiMaxTermsPerPattern)
        return false;
    memcpy(letter_of, local_of, sizeof local_of);
    memcpy(splat, local_splat, sizeof local_splat);
    n_letters = local_n;
    out.usable = true;
    return true;
}

} // namespace

extern "C"
{

int trx_regex_window_plan(const trx_regex_t * re, const double * byte_freq, int32_t out[5 + 10 * 6], uint32_t classes[48], double pass[2])
{
    if (!re || !byte_freq || !out)
        return set_error(TRX_EINVAL, "trx_regex_window_plan: NULL argument");
    auto img = std::make_unique<regex_dev>();
    fill_image(*re, false, nullptr, img.get());
    class_window const w = derive_class_window(*img, byte_freq);
    out[0] = w.ok ? 1 : 0;
    out[1] = w.len;
    out[2] = w.off_lo;
    out[3] = w.off_hi;
    out[4] = w.ok ? w.n_confirm : 0;
    for (int c = 0; c < out[4]; ++c)
    {
        out[5 + 10 * c] = w.confirm[c].lo;
        out[6 + 10 * c] = w.confirm[c].hi;
        for (int k = 0; k < 8; ++k)
            out[7 + 10 * c + k] = static_cast<int32_t>(w.confirm[c].mask[k]);
    }
    if (classes)
    {
        memset(classes, 0, 48 * sizeof(uint32_t));
        if (w.ok)
            memcpy(classes, w.cls, sizeof w.cls);
    }
    if (pass)
    {
        pass[0] = w.pass;
        pass[1] = w.confirmed;
    }
    return TRX_OK;
}


int trx_regex_scan_multi(trx_text_t * t, trx_regex_t * const * regexes, uint32_t n_patterns,
                         const uint64_t * candidate_bin_words, uint64_t cand_stride_words, trx_hit * hits,
                         uint32_t * hit_pattern, uint64_t cap, uint64_t * n_hits, uint64_t * per_pattern_hits,
                         uint64_t * stats)
{
    if (!t || !regexes || !candidate_bin_words || !n_hits || (cap && (!hits || !hit_pattern)))
        return set_error(TRX_EINVAL, "trx_regex_scan_multi: NULL argument");
    *n_hits = 0;
    for (uint32_t i = 0; i < n_patterns; ++i)
        if (!regexes[i])
            return set_error(TRX_EINVAL, "trx_regex_scan_multi: pattern %u is NULL", i);
    uint64_t const cand_words = (t->n_bins + 63) / 64;
    if (cand_stride_words != 0 && cand_stride_words < cand_words)
        return set_error(TRX_EINVAL, "trx_regex_scan_multi: candidate stride %llu < %llu words per pattern",
                         (unsigned long long)cand_stride_words, (unsigned long long)cand_words);
    if (per_pattern_hits)
        memset(per_pattern_hits, 0, n_patterns * sizeof(uint64_t));
    if (stats)
        stats[0] = stats[1] = stats[2] = stats[3] = 0;
    if (n_patterns == 0)
        return TRX_OK;
    trx_ctx * c = t->ctx;
    TRX_TRY(c->use());
    cudaStream_t const st = c->stream;
    auto cand_of = [&](uint32_t i) { return candidate_bin_words + static_cast<uint64_t>(i) * cand_stride_words; };

    // Hits of all patterns land in ONE device list (key = record << 32 | start, end, pattern) behind one counter, whichever
    // path found them; nothing comes back before the list is sorted into print order on the device at the end.  The shared
    // passes are enqueued back to back (two sets of staging blocks); the host only looks at a pass's record-list counters
    // when it needs the set again.
    std::vector<uint32_t> single; // patterns for the pattern-at-a-time path
    uint64_t const dev_cap = cap ? cap : 1;
    TRX_TRY(t->m_d_hit_keys.reserve(dev_cap * 8));
    TRX_TRY(t->m_d_hit_ends.reserve(dev_cap * 4));
    TRX_TRY(t->m_d_hit_pats.reserve(dev_cap * 4));
    TRX_TRY(t->m_d_total.reserve(16));
    t->m_h_total.pinned_host = true;
    TRX_TRY(t->m_h_total.reserve(16));
    TRX_CUDA(cudaMemsetAsync(t->m_d_total.ptr, 0, 16, st));
    for (trx_text::multi_set & ms : t->m_sets)
    {
        if (!ms.done)
            TRX_CUDA(cudaEventCreateWithFlags(&ms.done, cudaEventDisableTiming));
        if (ms.in_flight) // a call that ended in an error may have left a pass behind
            cudaEventSynchronize(ms.done);
        ms.in_flight = false;
    }
    // a pass is over: patterns whose record list overflowed (they reported nothing) go to the pattern-at-a-time path
    auto retire = [&](trx_text::multi_set & ms) -> int {
        if (!ms.in_flight)
            return TRX_OK;
        ms.in_flight = false;
        TRX_CUDA(cudaEventSynchronize(ms.done));
        auto const * listed = reinterpret_cast<uint32_t const *>(static_cast<uint8_t const *>(ms.h_back.ptr) + 16);
        for (size_t k = 0; k < ms.ids.size(); ++k)
            if (listed[k] > kMultiListCap)
            {
                single.push_back(ms.ids[k]);
                if (stats)
                    --stats[1];
            }
        return TRX_OK;
    };
    uint32_t n_passes = 0;
    // pattern-at-a-time scans run on the scan slots' streams and report into the shared list: the counter must be clear
    TRX_CUDA(cudaStreamSynchronize(st));
    std::vector<int> pending; // tickets of scans in flight
    struct abandon_pending
    {
        trx_text_t * t;
        std::vector<int> & tickets;
        ~abandon_pending()
        {
            for (int ticket : tickets) // only an error leaves any
                trx_regex_scan_end(t, ticket, nullptr, nullptr, nullptr);
        }
    } const on_exit{t, pending};
    size_t single_next = 0;
    auto finish_oldest = [&]() -> int {
        int const ticket = pending.front();
        pending.erase(pending.begin());
        uint64_t n = 0;
        return trx_regex_scan_end(t, ticket, nullptr, &n, nullptr); // waits for the scan; its hits are in the shared list
    };
    // begin up to `limit` of the patterns queued for the single path, at most three in flight
    auto pump_singles = [&](size_t limit) -> int {
        for (; limit && single_next < single.size(); --limit)
        {
            if (pending.size() == 3)
                TRX_TRY(finish_oldest());
            uint32_t const i = single[single_next];
            hit_sink sink{static_cast<unsigned long long *>(t->m_d_total.ptr), static_cast<uint64_t *>(t->m_d_hit_keys.ptr),
                          static_cast<uint32_t *>(t->m_d_hit_ends.ptr), static_cast<uint32_t *>(t->m_d_hit_pats.ptr), i, cap};
            int ticket = -1;
            TRX_TRY(scan_begin_impl(t, regexes[i], cand_of(i), 0, nullptr, 0, &ticket, &sink));
            pending.push_back(ticket);
            ++single_next;
        }
        return TRX_OK;
    };

    // ---- every pattern's phase-1 plan (cached per text like the single-pattern path does)
    char const * filt = getenv("TRX_REGEX_FILTER");
    char const * multi_off = getenv("TRX_REGEX_MULTI"); // TRX_REGEX_MULTI=off: every pattern on the single path (A/B, tests)
    bool const shared_pass = t->has_tmap && !filt && !(multi_off && strcmp(multi_off, "off") == 0);
    auto img = std::make_unique<regex_dev>();
    static bool const trace = getenv("TRX_SCAN_TRACE") != nullptr;
    auto const now = [] { return std::chrono::steady_clock::now(); };
    auto const us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
    double us_plan = 0, us_stage = 0, us_device = 0, us_single = 0, us_wait = 0;
    uint32_t next = 0;
    // ---- patterns with a class window (derive_class_window) go first, 128 at a time through regex_window_multi: one pass
    // over the tiles whose cost does not grow with the number of patterns.  TRX_MULTI_WINDOW=off: term lists for all (A/B)
    struct staged_window
    {
        int len, off_lo, off_hi;
        uint32_t cls[kWindowLen][8];
        int n_confirm;
        confirm_check confirm[kMaxConfirm];
    };
    std::vector<uint32_t> eligible;
    std::vector<staged_window> wins;
    std::vector<char> in_window(n_patterns, 0);
    {
        char const * e = getenv("TRX_MULTI_WINDOW");
        bool const use_window = shared_pass && !(e && strcmp(e, "off") == 0);
        auto const t_win = now();
        for (uint32_t i = 0; use_window && i < n_patterns; ++i)
        {
            trx_regex_t * re = regexes[i];
            if (re->follow_list.size() > sizeof(regex_dev::follow_list))
                continue;
            trx_text::cached_choice * found = plain_choice_of(t, re->serial);
            if (!found)
            {
                fill_image(*re, false, nullptr, img.get());
                trx_text::cached_choice cc{};
                cc.serial = re->serial;
                cc.kind = choose_filter(*img, t->byte_freq, false, cc.window, cc.atoms);
                if (t->choices.size() >= 4096)
                    t->choices.clear();
                t->choices.push_back(cc);
                t->plain_choice[re->serial] = static_cast<uint32_t>(t->choices.size() - 1);
                found = &t->choices.back();
            }
            if (found->win_state < 0)
            {
                fill_image(*re, false, nullptr, img.get());
                class_window const w = derive_class_window(*img, t->byte_freq);
                found->win_state = w.ok ? 1 : 0;
                found->win_len = w.len;
                found->win_off_lo = w.off_lo;
                found->win_off_hi = w.off_hi;
                found->win_n_confirm = w.n_confirm;
                if (w.ok)
                {
                    memcpy(found->win_cls, w.cls, sizeof w.cls);
                    memcpy(found->win_confirm, w.confirm, sizeof w.confirm);
                }
            }
            if (found->win_state != 1)
                continue;
            staged_window sw;
            sw.len = found->win_len;
            sw.off_lo = found->win_off_lo;
            sw.off_hi = found->win_off_hi;
            memcpy(sw.cls, found->win_cls, sizeof sw.cls);
            sw.n_confirm = found->win_n_confirm;
            memcpy(sw.confirm, found->win_confirm, sizeof sw.confirm);
            wins.push_back(sw);
            eligible.push_back(i);
            in_window[i] = 1;
        }
        // a term-list pass for a handful of leftovers costs more than the patterns do one at a time (its fixed cost is
        // one pass over the text with every leftover's weakest atoms): they join the single-pattern path
        if (use_window && !eligible.empty() && n_patterns - eligible.size() <= kFewLeftovers)
            for (uint32_t i = 0; i < n_patterns; ++i)
                if (!in_window[i])
                {
                    in_window[i] = 1;
                    single.push_back(i);
                }
        us_plan += us(t_win, now());
    }
    size_t wcur = 0;
    while (shared_pass && (wcur < eligible.size() || next < n_patterns))
    {
        auto const t_batch = now();
        bool const window_mode = wcur < eligible.size();
        std::vector<uint32_t> wtab; // [kWindowLen][4][256]
        // ---- a batch: up to kMultiPatterns patterns whose prefilter literals fit one alphabet of kMultiLetters bytes
        int letter_of[256];
        for (int & x : letter_of)
            x = -1;
        multi_params mp{};
        std::vector<multi_plan> plans;
        std::vector<uint32_t> ids;
        if (window_mode)
        {
            wtab.assign(static_cast<size_t>(kWindowLen) * 4 * 256, 0u);
            uint32_t beyond[kWindowLen][4] = {};
            while (wcur < eligible.size() && ids.size() < static_cast<size_t>(kMultiPatterns))
            {
                staged_window const & w = wins[wcur];
                uint32_t const k = static_cast<uint32_t>(ids.size());
                multi_plan plan;
                plan.prog = multi_program{};
                plan.prog.start_lo = -w.off_hi;
                plan.prog.start_hi = -w.off_lo;
                plan.prog.n_confirm = w.n_confirm;
                memcpy(plan.prog.confirm, w.confirm, sizeof w.confirm);
                plan.usable = true;
                for (int pos = 0; pos < w.len; ++pos)
                    for (int word = 0; word < 8; ++word)
                        for (uint32_t bits = w.cls[pos][word]; bits; bits &= bits - 1)
                            wtab[(static_cast<size_t>(pos) * 4 + (k >> 5)) * 256 + 32 * word + __builtin_ctz(bits)] |= 1u << (k & 31);
                for (int pos = w.len; pos < kWindowLen; ++pos) // a shorter window accepts every byte behind its end
                    beyond[pos][k >> 5] |= 1u << (k & 31);
                plans.push_back(plan);
                ids.push_back(eligible[wcur++]);
            }
            for (int pos = 0; pos < kWindowLen; ++pos)
                for (int g = 0; g < 4; ++g)
                    if (beyond[pos][g])
                        for (int ch = 0; ch < 256; ++ch)
                            wtab[(static_cast<size_t>(pos) * 4 + g) * 256 + ch] |= beyond[pos][g];
        }
        while (!window_mode && next < n_patterns && ids.size() < static_cast<size_t>(kMultiPatterns))
        {
            if (in_window[next])
            {
                ++next;
                continue;
            }
            trx_regex_t * re = regexes[next];
            trx_text::cached_choice const * found = plain_choice_of(t, re->serial);
            if (!found)
            {
                fill_image(*re, false, nullptr, img.get());
                trx_text::cached_choice cc{};
                cc.serial = re->serial;
                cc.kind = choose_filter(*img, t->byte_freq, false, cc.window, cc.atoms);
                if (t->choices.size() >= 4096)
                    t->choices.clear();
                t->choices.push_back(cc);
                t->plain_choice[re->serial] = static_cast<uint32_t>(t->choices.size() - 1);
                found = &t->choices.back();
            }
            if (found->multi_kind < 0)
            {
                fill_image(*re, false, nullptr, img.get());
                atom_budget budget;
                budget.max_lits = kMultiLetters - 2;
                budget.max_class = kMaxClassBytes;
                budget.max_atoms = 12;
                budget.good_enough = 1e-7;
                if (char const * e = getenv("TRX_MULTI_GOOD_ENOUGH")) // A/B: SIMD-stage terms against survivors
                    budget.good_enough = atof(e);
                if (char const * e = getenv("TRX_MULTI_MAX_ATOMS"))
                    budget.max_atoms = atoi(e);
                atom_plan const strong = derive_atoms(*img, t->byte_freq, false, budget);
                auto & mut = const_cast<trx_text::cached_choice &>(*found);
                if (strong.spec.n_lit >= 1 && (found->kind != kFilterWindow || strong.confirmed < 1e-6))
                {
                    mut.multi_kind = kFilterAtoms;
                    mut.multi_atoms = strong.spec;
                }
                else
                {
                    mut.multi_kind = found->kind; // the fixed window, the single-pattern atoms, or nothing
                    mut.multi_atoms = found->atoms;
                }
            }
            trx::atom_spec const & multi_atoms = found->multi_atoms;
            multi_plan plan;
            uint32_t const letters_before = mp.n_letters;
            bool const ok = re->follow_list.size() <= sizeof(regex_dev::follow_list) &&
                            plan_for_multi(static_cast<filter_kind>(found->multi_kind), found->window, multi_atoms, letter_of, mp.splat,
                                           mp.n_letters, plan);
            if (!ok)
            {
                bool const alphabet_full = found->multi_kind != kFilterNone && letters_before > 0 && !ids.empty() &&
                                           re->follow_list.size() <= sizeof(regex_dev::follow_list);
                if (alphabet_full)
                {
                    // does it fit an empty alphabet?  then it opens the next batch; otherwise it is a single-path pattern
                    int fresh_of[256];
                    for (int & x : fresh_of)
                        x = -1;
                    uint32_t fresh_splat[kMultiLetters], fresh_n = 0;
                    multi_plan probe;
                    if (plan_for_multi(static_cast<filter_kind>(found->multi_kind), found->window, multi_atoms, fresh_of, fresh_splat,
                                       fresh_n, probe))
                        break; // close this batch, the pattern is retried as the first of the next one
                }
                single.push_back(next++);
                continue;
            }
            plans.push_back(plan);
            ids.push_back(next++);
        }
        auto const t_planned = now();
        us_plan += us(t_batch, t_planned);
        if (ids.empty())
            continue;
        uint32_t const nb = static_cast<uint32_t>(ids.size());
        mp.n_patterns = nb;

        // ---- tiles: the union of the batch's candidate bins
        std::vector<uint64_t> uni(cand_words, 0);
        for (uint32_t k = 0; k < nb; ++k)
            for (uint64_t w = 0; w < cand_words; ++w)
                uni[w] |= cand_of(ids[k])[w];
        std::vector<uint32_t> tiles;
        uint64_t residues = 0;
        for (uint64_t b = 0; b < t->n_bins; ++b)
        {
            if (!((uni[b >> 6] >> (b & 63)) & 1))
                continue;
            uint64_t const r0 = t->h_bin_rec_begin[b], r1 = t->h_bin_rec_begin[b + 1];
            if (r0 == r1)
                continue;
            uint64_t const g0 = t->h_rec_offsets[r0], g1 = t->h_rec_offsets[r1];
            if (g1 == g0)
                continue;
            residues += g1 - g0;
            for (uint64_t tile = g0 / kTileBytes; tile <= (g1 - 1) / kTileBytes; ++tile)
                if (tiles.empty() || tiles.back() != tile)
                    tiles.push_back(static_cast<uint32_t>(tile));
        }
        if (stats)
        {
            stats[0] += residues;
            stats[1] += nb;
            stats[3] += 1;
        }
        if (tiles.empty())
            continue; // no candidate bin holds a residue: no hits for this batch

        // ---- device blocks.  staging (one H2D): [programs | automata | candidate words | phase-2 parameters | terms | DFAs]
        uint64_t const rfw = t->n_rec_flag_words;
        size_t const off_prog = 0;
        size_t const off_img = align16(off_prog + nb * sizeof(multi_program));
        size_t const off_cand = align16(off_img + nb * sizeof(regex_dev));
        size_t const off_rp = align16(off_cand + static_cast<size_t>(nb) * cand_words * 8);
        size_t const off_terms = align16(off_rp + nb * sizeof(resolve_params));
        trx_text::multi_set & ms = t->m_sets[n_passes++ & 1];
        {
            auto const t_retire = now();
            TRX_TRY(retire(ms));
            us_wait += us(t_retire, now());
        }
        size_t const off_dfa = align16(off_terms + std::max((static_cast<size_t>(nb) * kMultiMaxTermsPerPattern + 8) * 4,
                                                            window_mode ? sizeof(uint4) * kWindowLen * 256 : size_t{0}));
        std::vector<dfa_table const *> dfas(nb, nullptr);
        std::vector<size_t> dfa_at(nb, 0);
        size_t const off_tiles = off_dfa;
        size_t stage_bytes = align16(off_tiles + tiles.size() * 4);
        for (uint32_t k = 0; k < nb; ++k)
        {
            trx_regex & rx = *regexes[ids[k]];
            bool cached = false;
            for (dfa_table const & d : rx.dfas)
                cached = cached || (d.reverse == 0 && d.map_hash == 0);
            if (!cached)
                fill_image(rx, false, nullptr, img.get());
            dfas[k] = dfa_for(rx, *img, 0, 0);
            if (dfas[k])
            {
                dfa_at[k] = stage_bytes;
                stage_bytes = align16(stage_bytes + 256 + dfas[k]->trans.size() * 2);
            }
        }
        uint64_t const cfw = t->n_chunk_flag_words;
        ms.h_stage.pinned_host = true;
        TRX_TRY(ms.h_stage.reserve(stage_bytes));
        TRX_TRY(ms.d_stage.reserve(stage_bytes));
        TRX_TRY(ms.d_flags.reserve(static_cast<size_t>(nb) * (rfw + cfw) * 4));
        TRX_TRY(ms.d_lists.reserve(static_cast<size_t>(nb) * kMultiListCap * 4));
        size_t const counter_bytes = 16 + static_cast<size_t>(nb) * 4; // unused (u64), tile counter (u32), pad, listed records per pattern
        TRX_TRY(ms.d_counters.reserve(counter_bytes));
        ms.h_back.pinned_host = true;
        TRX_TRY(ms.h_back.reserve(counter_bytes));
        auto * hs = static_cast<uint8_t *>(ms.h_stage.ptr);
        auto * ds = static_cast<uint8_t *>(ms.d_stage.ptr);
        auto * d_counters = static_cast<uint8_t *>(ms.d_counters.ptr);
        memcpy(hs + off_tiles, tiles.data(), tiles.size() * 4);
        auto * h_prog = reinterpret_cast<multi_program *>(hs + off_prog);
        auto * h_rp = reinterpret_cast<resolve_params *>(hs + off_rp);
        for (uint32_t k = 0; k < nb; ++k)
        {
            fill_image(*regexes[ids[k]], false, nullptr, reinterpret_cast<regex_dev *>(hs + off_img) + k);
            memcpy(hs + off_cand + static_cast<size_t>(k) * cand_words * 8, cand_of(ids[k]), cand_words * 8);
            multi_program & P = h_prog[k];
            P = plans[k].prog;
            P.cand_words = reinterpret_cast<uint64_t const *>(ds + off_cand) + static_cast<size_t>(k) * cand_words;
            P.rec_flags = static_cast<uint32_t *>(ms.d_flags.ptr) + static_cast<size_t>(k) * rfw;
            P.chunk_flags = static_cast<uint32_t *>(ms.d_flags.ptr) + static_cast<size_t>(nb) * rfw + static_cast<size_t>(k) * cfw;
            {
                // a range = one term per offset (ORed like the atom's other terms); the list is padded in front to whole
                // rounds of four with copies of its first term, and every round's first word says how many rounds follow
                struct flat
                {
                    uint32_t letter, d;
                    bool end_atom, end_pattern;
                };
                std::vector<flat> fl;
                for (int i = 0; i < P.terms.n_terms; ++i)
                    for (uint32_t d = P.terms.lo[i]; d <= P.terms.hi[i]; ++d)
                        fl.push_back(flat{P.terms.letter[i], d, P.terms.end[i] != 0 && d == P.terms.hi[i],
                                          i + 1 == P.terms.n_terms && d == P.terms.hi[i]});
                while (fl.size() % 4)
                    fl.insert(fl.begin(), flat{fl[0].letter, fl[0].d, false, false});
                uint32_t const rounds = static_cast<uint32_t>(fl.size() / 4);
                auto * out = reinterpret_cast<uint32_t *>(hs + off_terms);
                for (size_t i = 0; i < fl.size(); ++i)
                    out[mp.n_terms_total++] = multi_pack_term(fl[i].letter, fl[i].d, rounds - 1 - static_cast<uint32_t>(i / 4),
                                                              fl[i].end_atom, fl[i].end_pattern);
            }
            P.rec_list = static_cast<uint32_t *>(ms.d_lists.ptr) + static_cast<size_t>(k) * kMultiListCap;
            P.rec_count = reinterpret_cast<unsigned int *>(d_counters + 16) + k;
            resolve_params & rp = h_rp[k];
            rp = resolve_params{};
            rp.text = t->d_text;
            rp.rec_offsets = t->d_rec_offsets;
            rp.rec_list = P.rec_list;
            rp.rec_count = P.rec_count;
            rp.chunk_flags = P.chunk_flags;
            rp.re = reinterpret_cast<regex_dev const *>(ds + off_img) + k;
            rp.hit_count = static_cast<unsigned long long *>(t->m_d_total.ptr);
            rp.hit_keys = static_cast<uint64_t *>(t->m_d_hit_keys.ptr);
            rp.hit_ends = static_cast<uint32_t *>(t->m_d_hit_ends.ptr);
            rp.hit_pats = static_cast<uint32_t *>(t->m_d_hit_pats.ptr);
            rp.pattern_id = ids[k];
            rp.rec_cap = kMultiListCap;
            rp.cap = cap;
            rp.anchor_begin = regexes[ids[k]]->anchor_begin;
            rp.anchor_end = regexes[ids[k]]->anchor_end;
            if (dfas[k])
            {
                memcpy(hs + dfa_at[k], dfas[k]->cls, 256);
                stage_dfa_table(*dfas[k], reinterpret_cast<uint16_t *>(hs + dfa_at[k] + 256));
                rp.dfa_cls = ds + dfa_at[k];
                rp.dfa_trans = reinterpret_cast<uint16_t const *>(ds + dfa_at[k] + 256);
                rp.dfa_classes = dfas[k]->n_classes;
                rp.dfa_states = dfas[k]->n_states;
                rp.dfa_first_accept = dfas[k]->first_accept;
            }
        }
        if (window_mode) // the term region carries the class tables instead
            memcpy(hs + off_terms, wtab.data(), wtab.size() * sizeof(uint32_t));
        auto const t_staged = now();
        us_stage += us(t_planned, t_staged);
        TRX_CUDA(cudaMemsetAsync(ms.d_flags.ptr, 0, static_cast<size_t>(nb) * (rfw + cfw) * 4, st));
        TRX_CUDA(cudaMemsetAsync(d_counters, 0, counter_bytes, st));
        TRX_CUDA(cudaMemcpyAsync(ds, hs, stage_bytes, cudaMemcpyHostToDevice, st));
        mp.programs = reinterpret_cast<multi_program const *>(ds + off_prog);
        mp.terms = reinterpret_cast<uint32_t const *>(ds + off_terms);
        scan_params sp{};
        sp.text = t->d_text;
        sp.n_text = t->n_text;
        sp.n_text_padded = t->n_text_padded;
        sp.rec_offsets = t->d_rec_offsets;
        sp.rec_bin = t->d_rec_bin;
        sp.n_records = t->n_records;
        sp.tiles = reinterpret_cast<uint32_t *>(ds + off_tiles);
        sp.n_tiles = static_cast<uint32_t>(tiles.size());
        sp.tile_counter = reinterpret_cast<unsigned int *>(d_counters + 8);
        if (window_mode)
        {
            constexpr int STAGES = 2;
            size_t const smem = ((tile_ring<STAGES, kScanThreads>::kRingBytes + 15) & ~size_t{15}) + sizeof(uint4) * kWindowLen * 256 + 1024;
            static bool attr_set[64] = {};
            int const dev = c->device >= 0 && c->device < 64 ? c->device : 0;
            if (!attr_set[dev])
            {
                TRX_CUDA(cudaFuncSetAttribute(regex_window_multi<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              static_cast<int>(smem)));
                attr_set[dev] = true;
            }
            window_params wp{};
            wp.programs = mp.programs;
            wp.n_patterns = nb;
            wp.tables = reinterpret_cast<uint32_t const *>(ds + off_terms);
            unsigned const grid = std::min<unsigned>(2u * static_cast<unsigned>(c->sm_count), sp.n_tiles); // two CTAs fit an SM
            regex_window_multi<STAGES><<<grid, kScanThreads + 32, smem, st>>>(sp, wp, t->tmap);
            ++c->launches;
            TRX_CUDA(cudaGetLastError());
        }
        else
        {
            constexpr int STAGES = 2;
            size_t const smem = ((tile_ring<STAGES, kScanThreads>::kRingBytes + 15) & ~size_t{15}) +
                                static_cast<size_t>(kMultiLetters) * 5 * kScanThreads * 4 +
                                (static_cast<size_t>(kMultiPatterns) * kMultiMaxTermsPerPattern + 8) * 4 + 1024;
            if (!t->multi_attr_set)
            {
                TRX_CUDA(cudaFuncSetAttribute(regex_filter_multi<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              static_cast<int>(smem)));
                t->multi_attr_set = true;
            }
            unsigned const grid = std::min<unsigned>(static_cast<unsigned>(c->sm_count), sp.n_tiles);
            regex_filter_multi<STAGES><<<grid, kScanThreads + 32, smem, st>>>(sp, mp, t->tmap);
            ++c->launches;
            TRX_CUDA(cudaGetLastError());
        }
        unsigned dyn_smem = 0;
        for (uint32_t k = 0; k < nb; ++k)
            if (dfas[k])
                dyn_smem = std::max(dyn_smem, dfa_smem_bytes(dfas[k]->n_states, dfas[k]->n_classes));
        // CTAs beyond a pattern's record list leave at once: the grid is sized for the longest lists (tens of thousands of
        // records), which otherwise sit on 128 warps long after the other patterns are done
        bool wide = false; // a pattern of more than 128 positions in the batch: the four-word resolver for all of it
        for (uint32_t k = 0; k < nb; ++k)
            wide = wide || regexes[ids[k]]->words > 2;
        if (wide)
            regex_scan_phase2_multi<4, 1><<<dim3(window_mode ? 128 : 32, nb), 128, dyn_smem, st>>>(reinterpret_cast<resolve_params const *>(ds + off_rp));
        else
            regex_scan_phase2_multi<2, 1><<<dim3(window_mode ? 128 : 32, nb), 128, dyn_smem, st>>>(reinterpret_cast<resolve_params const *>(ds + off_rp));
        ++c->launches;
        TRX_CUDA(cudaGetLastError());

        // ---- the pass's record-list counters follow it to the host; nobody waits for them here
        TRX_CUDA(cudaMemcpyAsync(ms.h_back.ptr, d_counters, counter_bytes, cudaMemcpyDeviceToHost, st));
        TRX_CUDA(cudaEventRecord(ms.done, st));
        ms.ids = ids;
        ms.in_flight = true;
        TRX_TRY(pump_singles(1)); // one of the waiting single-path patterns per pass, on the slots' streams beside it
        us_device += us(t_staged, now());
    }
    {
        auto const t_retire = now();
        for (trx_text::multi_set & ms : t->m_sets)
            TRX_TRY(retire(ms));
        us_wait += us(t_retire, now());
    }
    auto const t_single = now();
    if (!shared_pass)
        for (uint32_t i = 0; i < n_patterns; ++i)
            single.push_back(i);

    // ---- patterns without a usable prefilter, with an alphabet of their own, or with too many candidate records:
    // pattern at a time, pipelined over the scan slots like a caller of trx_regex_scan_begin/_end would
    if (stats)
        stats[2] = single.size();
    TRX_TRY(pump_singles(single.size()));
    while (!pending.empty())
        TRX_TRY(finish_oldest());

    us_single = us(t_single, now());
    auto const t_assemble = now();
    struct trace_line
    {
        bool on;
        std::function<void()> print;
        ~trace_line()
        {
            if (on)
                print();
        }
    } const at_return{trace, [&] {
        fprintf(stderr, "trx_regex_scan_multi: %u patterns: plans %.0f us, staging %.0f us, shared passes (enqueue .. hits back) %.0f us "
                        "(of it %.0f waiting for the device), %zu single-path patterns %.0f us, %zu class-window patterns, output %.0f us\n",
                n_patterns, us_plan, us_stage, us_device, us_wait, single.size(), us_single, eligible.size(), us(t_assemble, now()));
    }};
    // ---- print order: (pattern, bin, record, start); records are grouped by ascending bin, so (pattern, record, start).
    // Two stable radix sorts on the device — by key, then by pattern — one kernel that writes trx_hit records, and the
    // caller's arrays are filled by two copies.
    auto * h_total = static_cast<unsigned long long *>(t->m_h_total.ptr);
    TRX_CUDA(cudaMemcpyAsync(h_total, t->m_d_total.ptr, 8, cudaMemcpyDeviceToHost, st));
    TRX_CUDA(cudaStreamSynchronize(st));
    uint64_t const n = *h_total;
    *n_hits = n;
    if (n > cap)
        return set_error(TRX_ECAPACITY, "trx_regex_scan_multi: %llu hits, room for %llu", (unsigned long long)n, (unsigned long long)cap);
    if (n == 0)
        return TRX_OK;
    {
        int pat_bits = 1;
        while ((1ull << pat_bits) < n_patterns)
            ++pat_bits;
        int rec_bits = 1;
        while ((1ull << rec_bits) < t->n_records)
            ++rec_bits;
        size_t tmp_a = 0, tmp_b = 0;
        TRX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_a, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr),
                                                 static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<int>(n),
                                                 0, 32 + rec_bits, st));
        TRX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_b, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                                 static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<int>(n),
                                                 0, pat_bits, st));
        size_t const tmp_bytes = align16(std::max(tmp_a, tmp_b));
        // [keys2 8n | vals 8n | vals2 8n | pat 4n | pat2 4n | idx 4n | idx2 4n | hits 16n | temp]
        size_t const n16 = align16(n * 4) / 4; // 16-byte aligned pieces
        TRX_TRY(t->m_d_sort.reserve(n16 * 56 + tmp_bytes));
        auto * base = static_cast<uint8_t *>(t->m_d_sort.ptr);
        auto * keys2 = reinterpret_cast<uint64_t *>(base);
        auto * vals = reinterpret_cast<uint64_t *>(base + n16 * 8);
        auto * vals2 = reinterpret_cast<uint64_t *>(base + n16 * 16);
        auto * pat = reinterpret_cast<uint32_t *>(base + n16 * 24);
        auto * pat2 = reinterpret_cast<uint32_t *>(base + n16 * 28);
        auto * idx = reinterpret_cast<uint32_t *>(base + n16 * 32);
        auto * idx2 = reinterpret_cast<uint32_t *>(base + n16 * 36);
        auto * fmt = reinterpret_cast<uint4 *>(base + n16 * 40);
        void * tmp = base + n16 * 56;
        unsigned const grid = grid_for(n, 256, c->sm_count, 8);
        multi_pack_kernel<<<grid, 256, 0, st>>>(static_cast<uint32_t const *>(t->m_d_hit_ends.ptr),
                                                static_cast<uint32_t const *>(t->m_d_hit_pats.ptr), n, vals);
        size_t tb = tmp_bytes;
        TRX_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, static_cast<uint64_t const *>(t->m_d_hit_keys.ptr), keys2,
                                                 static_cast<uint64_t const *>(vals), vals2, static_cast<int>(n), 0, 32 + rec_bits, st));
        multi_unpack_kernel<<<grid, 256, 0, st>>>(vals2, n, pat, idx);
        tb = tmp_bytes;
        TRX_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, static_cast<uint32_t const *>(pat), pat2, static_cast<uint32_t const *>(idx), idx2,
                                                 static_cast<int>(n), 0, pat_bits, st));
        multi_format_kernel<<<grid, 256, 0, st>>>(keys2, vals2, idx2, t->d_rec_bin, n, fmt);
        c->launches += 3 + 6; // three kernels of ours + cub's histogram / onesweep passes
        TRX_CUDA(cudaGetLastError());
        if (trace)
        {
            cudaStreamSynchronize(st);
            fprintf(stderr, "trx_regex_scan_multi: %llu hits sorted and formatted %.0f us after the last scan\n", (unsigned long long)n,
                    us(t_assemble, now()));
        }
        if (n < (1u << 16))
        {
            TRX_CUDA(cudaMemcpyAsync(hits, fmt, n * sizeof(trx_hit), cudaMemcpyDeviceToHost, st));
            TRX_CUDA(cudaMemcpyAsync(hit_pattern, pat2, n * 4, cudaMemcpyDeviceToHost, st));
            TRX_CUDA(cudaStreamSynchronize(st));
        }
        else
        {
            // long lists: a copy into the caller's pageable arrays runs at a few GB/s.  Pinned staging at PCIe speed, in four
            // pieces, each moved on by a host thread of its own as soon as it has landed.
            constexpr int kPieces = 4;
            t->m_h_out.pinned_host = true;
            TRX_TRY(t->m_h_out.reserve(n16 * 20));
            auto * h_fmt = static_cast<uint8_t *>(t->m_h_out.ptr);
            auto * h_pat = h_fmt + n16 * 16;
            uint64_t const step = ((n + kPieces - 1) / kPieces + 3) & ~uint64_t{3};
            cudaEvent_t landed[kPieces] = {};
            int n_events = 0;
            cudaError_t err = cudaSuccess;
            for (int i = 0; i < kPieces && err == cudaSuccess; ++i)
            {
                uint64_t const lo = std::min<uint64_t>(n, i * step), hi = std::min<uint64_t>(n, lo + step);
                if (hi > lo)
                {
                    err = cudaMemcpyAsync(h_fmt + lo * 16, reinterpret_cast<uint8_t const *>(fmt) + lo * 16, (hi - lo) * 16,
                                          cudaMemcpyDeviceToHost, st);
                    if (err == cudaSuccess)
                        err = cudaMemcpyAsync(h_pat + lo * 4, reinterpret_cast<uint8_t const *>(pat2) + lo * 4, (hi - lo) * 4,
                                              cudaMemcpyDeviceToHost, st);
                }
                if (err == cudaSuccess)
                    err = cudaEventCreateWithFlags(&landed[i], cudaEventDisableTiming);
                if (err == cudaSuccess)
                {
                    ++n_events;
                    err = cudaEventRecord(landed[i], st);
                }
            }
            std::vector<std::thread> movers;
            std::atomic<int> failed{0};
            for (int i = 0; i < n_events; ++i)
                movers.emplace_back([&, i] {
                    cudaSetDevice(c->device);
                    uint64_t const lo = std::min<uint64_t>(n, i * step), hi = std::min<uint64_t>(n, lo + step);
                    if (cudaEventSynchronize(landed[i]) != cudaSuccess)
                    {
                        failed = 1;
                        return;
                    }
                    memcpy(reinterpret_cast<uint8_t *>(hits) + lo * 16, h_fmt + lo * 16, (hi - lo) * 16);
                    memcpy(reinterpret_cast<uint8_t *>(hit_pattern) + lo * 4, h_pat + lo * 4, (hi - lo) * 4);
                });
            for (std::thread & th : movers)
                th.join();
            for (int i = 0; i < n_events; ++i)
                cudaEventDestroy(landed[i]);
            TRX_CUDA(err);
            if (failed)
                return set_error(TRX_ECUDA, "trx_regex_scan_multi: the copy of the hit list failed");
        }
    }
    if (per_pattern_hits) // the list is sorted by pattern: boundaries by bisection, not a pass over a million entries
    {
        uint32_t const * at = hit_pattern;
        for (uint32_t i = 0; i < n_patterns && at < hit_pattern + n; ++i)
        {
            uint32_t const * const next_at = std::upper_bound(at, static_cast<uint32_t const *>(hit_pattern + n), i);
            per_pattern_hits[i] = static_cast<uint64_t>(next_at - at);
            at = next_at;
        }
    }
    return TRX_OK;
}

} // extern "C"
