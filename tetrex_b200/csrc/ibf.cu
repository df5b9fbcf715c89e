// ibf.cu — C ABI for the flat IBF (rows a1–a7 of SURVEY.md §8): create/upload/download, K1 emplace,
// K2 bulk_contains and the collector's path algebra.
#include <mutex>
#include <type_traits>

#include "ibf_kernels.cuh"

namespace trx
{

static ibf_geom geom_of(trx_ibf const * f)
{
    return ibf_geom{f->d_words, f->bin_size, f->shard_words, f->shard_word_begin,
                    static_cast<uint32_t>(f->hash_shift), static_cast<uint32_t>(f->hash_funs), 0, f->bin_size, nullptr, 0, 0,
                    nullptr, 0};
}

// host-buffer entry points of the path algebra go zero-copy (pinned host memory read by the kernel) up to this many items
constexpr uint64_t kZeroCopyItems = 4096;

static int pow2_ceil(uint64_t x)
{
    int g = 1;
    while (static_cast<uint64_t>(g) < x && g < 32)
        g <<= 1;
    return g;
}

template <int H, int V, int MODE>
static int launch_probe_g(trx_ctx * ctx, ibf_geom g, int G, uint64_t const * kmers, uint64_t n, uint64_t * out,
                          uint32_t const * slots, uint8_t * alive, uint64_t const * hits, uint32_t const * hit_index)
{
    auto go = [&](auto gtag) -> int
    {
        constexpr int GG = decltype(gtag)::value;
        unsigned grid = grid_for(n, 256 / GG, ctx->sm_count, 8);
        probe_kernel<H, V, GG, MODE><<<grid, 256, 0, ctx->stream>>>(g, kmers, n, out, slots, alive, hits, hit_index);
        ++ctx->launches;
        TRX_CUDA(cudaGetLastError());
        return TRX_OK;
    };
    switch (G)
    {
    case 1: return go(std::integral_constant<int, 1>{});
    case 2: return go(std::integral_constant<int, 2>{});
    case 4: return go(std::integral_constant<int, 4>{});
    case 8: return go(std::integral_constant<int, 8>{});
    case 16: return go(std::integral_constant<int, 16>{});
    default: return go(std::integral_constant<int, 32>{});
    }
}

template <int MODE>
static int launch_probe(trx_ctx * ctx, ibf_geom g, uint64_t const * kmers, uint64_t n, uint64_t * out,
                        uint32_t const * slots, uint8_t * alive, uint64_t const * hits = nullptr,
                        uint32_t const * hit_index = nullptr)
{
    if (n == 0)
        return TRX_OK;
    bool const vec2 = (g.shard_words % 2 == 0);
    int const G = pow2_ceil(vec2 ? g.shard_words / 2 : g.shard_words);
#define TRX_PROBE_H(HH)                                                                                                \
    return vec2 ? launch_probe_g<HH, 2, MODE>(ctx, g, G, kmers, n, out, slots, alive, hits, hit_index)                 \
                : launch_probe_g<HH, 1, MODE>(ctx, g, G, kmers, n, out, slots, alive, hits, hit_index)
    if constexpr (MODE == 2)
    {
        TRX_PROBE_H(1);
    }
    else
    {
        switch (g.hash_funs)
        {
        case 1: TRX_PROBE_H(1);
        case 2: TRX_PROBE_H(2);
        case 3: TRX_PROBE_H(3);
        case 4: TRX_PROBE_H(4);
        default: TRX_PROBE_H(5);
        }
    }
#undef TRX_PROBE_H
}

// exposed to hibf.cu
int paths_and_hits_launch(trx_ctx * ctx, uint64_t * slot_words, uint64_t W, uint64_t const * d_hits,
                          uint32_t const * d_hit_index, uint32_t const * d_slots, uint64_t n, uint8_t * d_alive)
{
    ibf_geom g{nullptr, 0, W, 0, 0, 1};
    return launch_probe<2>(ctx, g, nullptr, n, slot_words, d_slots, d_alive, d_hits, d_hit_index);
}

} // namespace trx

using namespace trx;

extern "C"
{

int trx_ibf_create(trx_ctx_t * ctx, uint64_t bins, uint64_t bin_size, uint32_t hash_funs, uint32_t shard_index,
                   uint32_t shard_count, trx_ibf_t ** out)
{
    if (!ctx || !out)
        return set_error(TRX_EINVAL, "trx_ibf_create: NULL argument");
    *out = nullptr;
    // messages follow lib/hibf/src/interleaved_bloom_filter.cpp:41-46
    if (bins == 0)
        return set_error(TRX_EINVAL, "The number of bins must be > 0.");
    if (hash_funs == 0 || hash_funs > 5)
        return set_error(TRX_EINVAL, "The number of hash functions must be > 0 and <= 5.");
    if (bin_size == 0)
        return set_error(TRX_EINVAL, "The size of a bin must be > 0.");
    if (shard_count == 0)
        shard_count = 1;
    auto * f = new (std::nothrow) trx_ibf;
    if (!f)
        return set_error(TRX_ENOMEM, "out of host memory");
    f->ctx = ctx;
    f->bins = bins;
    f->bin_size = bin_size;
    f->hash_funs = hash_funs;
    f->hash_shift = static_cast<uint64_t>(__builtin_clzll(bin_size));
    f->bin_words = (bins + 63) / 64;
    f->technical_bins = f->bin_words * 64;
    f->n_words = f->bin_words * bin_size; // technical_bins*bin_size bits / 64
    if (shard_index >= shard_count || f->bin_words % shard_count != 0)
    {
        delete f;
        return set_error(TRX_EINVAL, "shard %u/%u: shard_count must divide bin_words (%llu)", shard_index, shard_count,
                         static_cast<unsigned long long>((bins + 63) / 64));
    }
    f->shard_words = f->bin_words / shard_count;
    f->shard_word_begin = f->shard_words * shard_index;
    f->shard_n_words = f->shard_words * bin_size;
    if (ctx->use() != TRX_OK)
    {
        delete f;
        return TRX_ECUDA;
    }
    cudaError_t e = cudaMalloc(&f->d_words, f->shard_n_words * sizeof(uint64_t));
    if (e != cudaSuccess)
    {
        delete f;
        return set_error(TRX_ENOMEM, "cudaMalloc of %llu matrix bytes: %s",
                         static_cast<unsigned long long>(f->shard_n_words * 8), cudaGetErrorString(e));
    }
    e = cudaMemsetAsync(f->d_words, 0, f->shard_n_words * sizeof(uint64_t), ctx->stream);
    if (e != cudaSuccess)
    {
        cudaFree(f->d_words);
        delete f;
        return set_error(TRX_ECUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
    }
    *out = f;
    return TRX_OK;
}

void trx_ibf_destroy(trx_ibf_t * f)
{
    if (!f)
        return;
    cudaSetDevice(f->ctx->device);
    cudaStreamSynchronize(f->ctx->stream);
    if (f->owns_words)
        cudaFree(f->d_words);
    f->slab.release();
    delete f;
}

int trx_ibf_geometry(const trx_ibf_t * f, uint64_t out[10])
{
    if (!f || !out)
        return set_error(TRX_EINVAL, "NULL argument");
    out[0] = f->bins;
    out[1] = f->technical_bins;
    out[2] = f->bin_size;
    out[3] = f->hash_shift;
    out[4] = f->bin_words;
    out[5] = f->hash_funs;
    out[6] = f->n_words;
    out[7] = f->shard_word_begin;
    out[8] = f->shard_words;
    out[9] = f->shard_n_words;
    return TRX_OK;
}

void * trx_ibf_device_words(trx_ibf_t * f)
{
    return f ? f->d_words : nullptr;
}

int trx_ibf_clear(trx_ibf_t * f)
{
    if (!f)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(f->ctx->use());
    TRX_CUDA(cudaMemsetAsync(f->d_words, 0, f->shard_n_words * sizeof(uint64_t), f->ctx->stream));
    return TRX_OK;
}

int trx_ibf_upload(trx_ibf_t * f, const uint64_t * words, uint64_t n_words)
{
    if (!f || !words)
        return set_error(TRX_EINVAL, "NULL argument");
    if (n_words != f->n_words)
        return set_error(TRX_EINVAL, "trx_ibf_upload: expected %llu words, got %llu",
                         static_cast<unsigned long long>(f->n_words), static_cast<unsigned long long>(n_words));
    TRX_TRY(f->ctx->use());
    // strided 2-D copy picks this shard's columns out of every row
    TRX_CUDA(cudaMemcpy2DAsync(f->d_words, f->shard_words * 8, words + f->shard_word_begin, f->bin_words * 8,
                               f->shard_words * 8, f->bin_size, cudaMemcpyHostToDevice, f->ctx->stream));
    TRX_CUDA(cudaStreamSynchronize(f->ctx->stream));
    return TRX_OK;
}

int trx_ibf_download(trx_ibf_t * f, uint64_t * words, uint64_t n_words)
{
    if (!f || !words)
        return set_error(TRX_EINVAL, "NULL argument");
    if (n_words != f->n_words)
        return set_error(TRX_EINVAL, "trx_ibf_download: expected %llu words, got %llu",
                         static_cast<unsigned long long>(f->n_words), static_cast<unsigned long long>(n_words));
    TRX_TRY(f->ctx->use());
    TRX_CUDA(cudaMemcpy2DAsync(words + f->shard_word_begin, f->bin_words * 8, f->d_words, f->shard_words * 8,
                               f->shard_words * 8, f->bin_size, cudaMemcpyDeviceToHost, f->ctx->stream));
    TRX_CUDA(cudaStreamSynchronize(f->ctx->stream));
    return TRX_OK;
}

// ------------------------------------------------------------------------------------------- K1
int trx_emplace_values_dev(trx_ibf_t * f, const uint64_t * d_values, const uint32_t * d_bins, uint64_t n)
{
    if (!f)
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    TRX_TRY(f->ctx->use());
    unsigned grid = grid_for(n, 256, f->ctx->sm_count, 8);
    emplace_values_kernel<<<grid, 256, 0, f->ctx->stream>>>(geom_of(f), d_values, d_bins, n);
    ++f->ctx->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

int trx_emplace_values(trx_ibf_t * f, const uint64_t * values, const uint32_t * bins, uint64_t n)
{
    if (!f || (n && (!values || !bins)))
        return set_error(TRX_EINVAL, "NULL argument");
    for (uint64_t i = 0; i < n; ++i)
        if (bins[i] >= f->bins)
            return set_error(TRX_ERANGE, "bin %u >= bin count %llu", bins[i], static_cast<unsigned long long>(f->bins));
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    TRX_TRY(c->d_in[0].reserve(n * 8));
    TRX_TRY(c->d_in[1].reserve(n * 4));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, values, n * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[1].ptr, bins, n * 4, cudaMemcpyHostToDevice, c->stream));
    TRX_TRY(trx_emplace_values_dev(f, static_cast<uint64_t *>(c->d_in[0].ptr), static_cast<uint32_t *>(c->d_in[1].ptr), n));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

static int check_kmer_args(int molecule, uint32_t k)
{
    if (molecule != TRX_MOL_NA && molecule != TRX_MOL_AA)
        return set_error(TRX_EINVAL, "molecule must be TRX_MOL_NA or TRX_MOL_AA");
    if (k == 0 || (molecule == TRX_MOL_NA && k > 32) || (molecule == TRX_MOL_AA && k > 12))
        return set_error(TRX_EINVAL, "k=%u out of range (DNA 1..32, peptide 1..12, src/main.cpp:28-32)", k);
    return TRX_OK;
}

// aamap lives in the context's 4th input scratch (256 bytes)
static int upload_aamap(trx_ctx * c, int reduction, uint8_t ** d_map)
{
    uint8_t aamap[256], redmap[256];
    trx_peptide_maps(reduction, aamap, redmap);
    TRX_TRY(c->d_in[3].reserve(256));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[3].ptr, aamap, 256, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream)); // aamap is a stack buffer
    *d_map = static_cast<uint8_t *>(c->d_in[3].ptr);
    return TRX_OK;
}

static int emplace_records_impl(trx_ibf_t * f, int molecule, uint32_t k, int reduction, const uint8_t * d_residues,
                                uint64_t n_residues, const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin,
                                uint64_t n_records, uint64_t hint_first_bin, uint64_t hint_n_bins);

int trx_emplace_records_dev(trx_ibf_t * f, int molecule, uint32_t k, int reduction, const uint8_t * d_residues,
                            uint64_t n_residues, const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin,
                            uint64_t n_records)
{
    return emplace_records_impl(f, molecule, k, reduction, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, 0, 0);
}

int trx_emplace_records_slab_dev(trx_ibf_t * f, int molecule, uint32_t k, int reduction, const uint8_t * d_residues,
                                 uint64_t n_residues, const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin,
                                 uint64_t n_records, uint64_t first_bin, uint64_t n_bins)
{
    if (f && (n_bins == 0 || first_bin + n_bins > f->bins))
        return set_error(TRX_ERANGE, "bins [%llu, %llu) are not bins of this filter (%llu)", (unsigned long long)first_bin,
                         (unsigned long long)(first_bin + n_bins), (unsigned long long)f->bins);
    return emplace_records_impl(f, molecule, k, reduction, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records,
                                first_bin, n_bins);
}

static int emplace_records_impl(trx_ibf_t * f, int molecule, uint32_t k, int reduction, const uint8_t * d_residues,
                                uint64_t n_residues, const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin,
                                uint64_t n_records, uint64_t hint_first_bin, uint64_t hint_n_bins)
{
    if (!f)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_kmer_args(molecule, k));
    if (n_records == 0 || n_residues == 0)
        return TRX_OK;
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    uint32_t dna_run = 64; // steps per DNA thread (TRX_DNA_RUN: A/B evidence)
    if (char const * e = getenv("TRX_DNA_RUN"))
        dna_run = atoi(e) > 0 ? static_cast<uint32_t>(atoi(e)) : dna_run;
    uint64_t const per_thread = molecule == TRX_MOL_NA ? dna_run : static_cast<uint32_t>(kRun);
    uint64_t n_runs = (n_residues + per_thread - 1) / per_thread;
    unsigned grid = grid_for(n_runs, 256, c->sm_count, 8);
    // A matrix that does not fit L2 is filled in row slices that do: every pass recomputes all hashes (cheap: the
    // kernel is bound by the atomics, not the arithmetic) but only fires the atomics that fall into its slice, so the
    // read-modify-writes stay in L2 instead of costing a 32-byte DRAM read and write-back each.  TRX_EMPLACE_WINDOW_MB
    // sets the slice size (0 = one pass, the measured baseline).
    uint64_t window = 64ull << 20;
    if (char const * e = getenv("TRX_EMPLACE_WINDOW_MB"))
        window = static_cast<uint64_t>(atoll(e)) << 20;
    uint64_t const row_bytes = f->shard_words * 8;
    uint64_t passes = window ? (f->bin_size * row_bytes + window - 1) / window : 1;
    if (passes < 1)
        passes = 1;
    // Far beyond L2 (C4: 23 GB, 2.9 GB per GPU at 8 shards) slicing the whole matrix would mean re-hashing the text dozens
    // of times.  But the records of one call usually belong to a few neighbouring bins (one FASTA file = one bin, a slab of
    // files = one 64-bin word column): when the caller names that bin range, the bits go into a DENSE copy of just those
    // word columns (C4, 64 bins: 22.8 M rows x 8 B = 183 MB instead of one 8-byte word every 1 KiB), which IS sliceable
    // into L2-sized passes, and one strided OR pass folds it into the interleaved matrix.  Bins outside the named range
    // still go straight to the matrix, so the hint is about speed only.
    uint64_t slab_words = 0, slab_begin = 0;
    if (passes > 10)
    {
        passes = 1;
        if (hint_n_bins)
        {
            uint64_t w0 = hint_first_bin / 64, w1 = (hint_first_bin + hint_n_bins + 63) / 64;
            w0 = w0 < f->shard_word_begin ? f->shard_word_begin : w0;
            w1 = w1 > f->shard_word_begin + f->shard_words ? f->shard_word_begin + f->shard_words : w1;
            uint64_t const slab_bytes = w1 > w0 ? (w1 - w0) * 8 * f->bin_size : 0;
            char const * off = getenv("TRX_EMPLACE_SLAB"); // TRX_EMPLACE_SLAB=off: A/B evidence
            if (slab_bytes && slab_bytes <= (1ull << 30) && (w1 - w0) * 4 <= f->shard_words && window &&
                !(off && strcmp(off, "off") == 0))
            {
                slab_words = w1 - w0;
                slab_begin = w0;
                // a dense slab is hit by ~100 atomics per 32-byte sector, so even at 1.5x the L2 size nearly all of them
                // find their sector in L2: one pass (8.7 ms per 64-bin C4 slab) beats three L2-sized slices (11.3 ms),
                // whose extra text + hashing passes cost more than the few misses; slice only slabs far beyond L2
                passes = slab_bytes <= (256ull << 20) ? 1 : (slab_bytes + window - 1) / window;
            }
        }
    }
    uint64_t * d_slab = nullptr;
    if (slab_words)
    {
        TRX_TRY(f->slab.reserve(slab_words * 8 * f->bin_size));
        d_slab = static_cast<uint64_t *>(f->slab.ptr);
        TRX_CUDA(cudaMemsetAsync(d_slab, 0, slab_words * 8 * f->bin_size, c->stream));
    }
    uint64_t const rows_per_pass = (f->bin_size + passes - 1) / passes;
    uint8_t * d_map = nullptr;
    if (molecule != TRX_MOL_NA)
        TRX_TRY(upload_aamap(c, reduction, &d_map));
    for (uint64_t p = 0; p < passes; ++p)
    {
        ibf_geom g = geom_of(f);
        g.slab = d_slab;
        g.slab_words = slab_words;
        g.slab_begin = slab_begin;
        g.row_lo = p * rows_per_pass;
        g.row_hi = g.row_lo + rows_per_pass < f->bin_size ? g.row_lo + rows_per_pass : f->bin_size;
        if (g.row_lo >= g.row_hi)
            break;
        if (molecule == TRX_MOL_NA)
            emplace_dna_kernel<<<grid, 256, 0, c->stream>>>(g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, k,
                                                            dna_run);
        else
            emplace_records_kernel<TRX_MOL_AA><<<grid, 256, 0, c->stream>>>(g, d_residues, n_residues, d_rec_offsets, d_rec_bin,
                                                                            n_records, k, d_map);
        ++c->launches;
    }
    if (slab_words)
    {
        uint64_t const total = slab_words * f->bin_size;
        slab_merge_kernel<<<grid_for(total, 256, c->sm_count, 16), 256, 0, c->stream>>>(
            f->d_words, f->shard_words, slab_begin - f->shard_word_begin, d_slab, slab_words, f->bin_size);
        ++c->launches;
    }
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

int trx_emplace_records(trx_ibf_t * f, int molecule, uint32_t k, int reduction, const uint8_t * residues,
                        const uint64_t * rec_offsets, const uint32_t * rec_bin, uint64_t n_records)
{
    if (!f || (n_records && (!residues || !rec_offsets || !rec_bin)))
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_kmer_args(molecule, k));
    if (n_records == 0)
        return TRX_OK;
    if (rec_offsets[0] != 0)
        return set_error(TRX_EINVAL, "rec_offsets[0] must be 0");
    for (uint64_t i = 0; i < n_records; ++i)
    {
        if (rec_offsets[i + 1] < rec_offsets[i])
            return set_error(TRX_EINVAL, "rec_offsets must be non-decreasing");
        if (rec_bin[i] >= f->bins)
            return set_error(TRX_ERANGE, "record %llu: bin %u >= bin count %llu", static_cast<unsigned long long>(i),
                             rec_bin[i], static_cast<unsigned long long>(f->bins));
    }
    uint32_t bin_lo = rec_bin[0], bin_hi = rec_bin[0];
    for (uint64_t i = 1; i < n_records; ++i)
    {
        bin_lo = rec_bin[i] < bin_lo ? rec_bin[i] : bin_lo;
        bin_hi = rec_bin[i] > bin_hi ? rec_bin[i] : bin_hi;
    }
    uint64_t const n_res = rec_offsets[n_records];
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    TRX_TRY(c->d_in[0].reserve(n_res + 64));
    TRX_TRY(c->d_in[1].reserve((n_records + 1) * 8));
    TRX_TRY(c->d_in[2].reserve(n_records * 4));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, residues, n_res, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[1].ptr, rec_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[2].ptr, rec_bin, n_records * 4, cudaMemcpyHostToDevice, c->stream));
    TRX_TRY(emplace_records_impl(f, molecule, k, reduction, static_cast<uint8_t *>(c->d_in[0].ptr), n_res,
                                 static_cast<uint64_t *>(c->d_in[1].ptr), static_cast<uint32_t *>(c->d_in[2].ptr), n_records,
                                 bin_lo, static_cast<uint64_t>(bin_hi) - bin_lo + 1));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

static int check_record_table(const uint64_t * rec_offsets, const uint32_t * rec_bin, uint64_t n_records, uint64_t n_bins);
static int stage_records(trx_ctx * c, const uint8_t * residues, const uint64_t * rec_offsets, const uint32_t * rec_bin,
                         uint64_t n_records);

// N3 — compute_sketches (lib/hibf/src/sketch/compute_sketches.cpp:27-43): the HyperLogLog sketch of every user bin,
// straight from the residues with K1's decomposition (so a sketch sees exactly the k-mers HIBFIndex::emplace would buffer,
// duplicates and wrap-around k-mers included).  d_registers: n_bins << sketch_bits bytes, seqan's register layout.
int trx_hll_sketch_records_dev(trx_ctx_t * c, int molecule, uint32_t k, int reduction, const uint8_t * d_residues,
                               uint64_t n_residues, const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin, uint64_t n_records,
                               uint64_t n_bins, uint32_t sketch_bits, uint8_t * d_registers)
{
    if (!c || !d_registers || n_bins == 0)
        return set_error(TRX_EINVAL, "NULL argument");
    if (sketch_bits < 5 || sketch_bits > 20)
        return set_error(TRX_ERANGE, "sketch_bits %u outside [5, 20]", sketch_bits);
    TRX_TRY(check_kmer_args(molecule, k));
    TRX_TRY(c->use());
    uint64_t const n_regs = n_bins << sketch_bits;
    TRX_TRY(c->d_out[1].reserve(n_regs * 4));
    auto * d_regs = static_cast<unsigned int *>(c->d_out[1].ptr);
    TRX_CUDA(cudaMemsetAsync(d_regs, 0, n_regs * 4, c->stream));
    if (n_records && n_residues)
    {
        if (!d_residues || !d_rec_offsets || !d_rec_bin)
            return set_error(TRX_EINVAL, "NULL argument");
        ibf_geom g{};
        g.bin_size = 1;
        g.shard_words = (n_bins + 63) / 64;
        g.hash_funs = 1;
        g.row_hi = 1;
        g.hll = d_regs;
        g.hll_bits = sketch_bits;
        uint32_t const dna_run = 64;
        uint64_t const per_thread = molecule == TRX_MOL_NA ? dna_run : static_cast<uint32_t>(kRun);
        unsigned const grid = grid_for((n_residues + per_thread - 1) / per_thread, 256, c->sm_count, 8);
        if (molecule == TRX_MOL_NA)
            emplace_dna_kernel<<<grid, 256, 0, c->stream>>>(g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, k, dna_run);
        else
        {
            uint8_t * d_map = nullptr;
            TRX_TRY(upload_aamap(c, reduction, &d_map));
            emplace_records_kernel<TRX_MOL_AA><<<grid, 256, 0, c->stream>>>(g, d_residues, n_residues, d_rec_offsets, d_rec_bin,
                                                                            n_records, k, d_map);
        }
        ++c->launches;
        TRX_CUDA(cudaGetLastError());
    }
    hll_pack_kernel<<<grid_for(n_regs, 256, c->sm_count, 8), 256, 0, c->stream>>>(d_regs, d_registers, n_regs);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

int trx_hll_sketch_records(trx_ctx_t * c, int molecule, uint32_t k, int reduction, const uint8_t * residues,
                           const uint64_t * rec_offsets, const uint32_t * rec_bin, uint64_t n_records, uint64_t n_bins,
                           uint32_t sketch_bits, uint8_t * registers)
{
    if (!c || !registers || n_bins == 0 || (n_records && (!residues || !rec_offsets || !rec_bin)))
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(c->use());
    uint64_t const n_regs = n_bins << sketch_bits;
    TRX_TRY(c->d_out[0].reserve(n_regs));
    if (n_records)
    {
        TRX_TRY(check_record_table(rec_offsets, rec_bin, n_records, n_bins));
        TRX_TRY(stage_records(c, residues, rec_offsets, rec_bin, n_records));
    }
    TRX_TRY(trx_hll_sketch_records_dev(c, molecule, k, reduction, static_cast<uint8_t *>(c->d_in[0].ptr),
                                       n_records ? rec_offsets[n_records] : 0, static_cast<uint64_t *>(c->d_in[1].ptr),
                                       static_cast<uint32_t *>(c->d_in[2].ptr), n_records, n_bins, sketch_bits,
                                       static_cast<uint8_t *>(c->d_out[0].ptr)));
    TRX_CUDA(cudaMemcpyAsync(registers, c->d_out[0].ptr, n_regs, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

int trx_decompose_records(trx_ctx_t * c, int molecule, uint32_t k, int reduction, const uint8_t * residues,
                          const uint64_t * rec_offsets, uint64_t n_records, uint64_t * out_values,
                          uint64_t out_capacity, uint64_t * out_offsets)
{
    if (!c || !rec_offsets || !out_offsets)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_kmer_args(molecule, k));
    out_offsets[0] = 0;
    for (uint64_t i = 0; i < n_records; ++i)
        out_offsets[i + 1] = out_offsets[i] + trx_emit_count(molecule, k, rec_offsets[i + 1] - rec_offsets[i]);
    uint64_t const n_out = out_offsets[n_records];
    if (n_out > out_capacity)
        return set_error(TRX_ECAPACITY, "trx_decompose_records: need room for %llu values, got %llu",
                         static_cast<unsigned long long>(n_out), static_cast<unsigned long long>(out_capacity));
    if (n_out == 0)
        return TRX_OK;
    if (!residues || !out_values)
        return set_error(TRX_EINVAL, "NULL argument");
    uint64_t const n_res = rec_offsets[n_records];
    TRX_TRY(c->use());
    TRX_TRY(c->d_in[0].reserve(n_res + 64));
    TRX_TRY(c->d_in[1].reserve((n_records + 1) * 8));
    TRX_TRY(c->d_in[2].reserve((n_records + 1) * 8));
    TRX_TRY(c->d_out[0].reserve(n_out * 8));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, residues, n_res, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[1].ptr, rec_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[2].ptr, out_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    unsigned grid = grid_for(n_out, 256, c->sm_count, 8);
    auto * d_text = static_cast<uint8_t *>(c->d_in[0].ptr);
    auto * d_ro = static_cast<uint64_t *>(c->d_in[1].ptr);
    auto * d_oo = static_cast<uint64_t *>(c->d_in[2].ptr);
    auto * d_out = static_cast<uint64_t *>(c->d_out[0].ptr);
    if (molecule == TRX_MOL_NA)
    {
        decompose_kernel<TRX_MOL_NA><<<grid, 256, 0, c->stream>>>(d_text, d_ro, d_oo, n_records, k, nullptr, d_out, n_out);
    }
    else
    {
        uint8_t * d_map = nullptr;
        TRX_TRY(upload_aamap(c, reduction, &d_map));
        decompose_kernel<TRX_MOL_AA><<<grid, 256, 0, c->stream>>>(d_text, d_ro, d_oo, n_records, k, d_map, d_out, n_out);
    }
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    TRX_CUDA(cudaMemcpyAsync(out_values, d_out, n_out * 8, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

// ------------------------------------------------------------------------------ DGram (gapped 3+3-mers)
static int check_dgram_args(uint32_t min_gap, uint32_t max_gap)
{
    // the halo of one tile is max_gap + 5 bytes of shared memory next to the tile itself
    if (max_gap > TRX_DGRAM_MAX_GAP)
        return set_error(TRX_ERANGE, "DGram upper gap bound %u exceeds %u", max_gap, (unsigned)TRX_DGRAM_MAX_GAP);
    (void)min_gap; // min_gap > max_gap is legal in the reference: the gap loop never runs, no code is produced
    return TRX_OK;
}

extern "C++" {
template <bool COUNT>
static int launch_dgram(trx_ctx * c, ibf_geom g, const uint8_t * d_residues, uint64_t n_residues,
                        const uint64_t * d_rec_offsets, const uint32_t * d_rec_bin, uint64_t n_records, uint32_t min_gap,
                        uint32_t max_gap, unsigned long long * d_per_bin, const uint64_t * d_range = nullptr, uint64_t range_tiles = 0)
{
    uint64_t const n_tiles = d_range ? range_tiles : (n_residues + kDgramTile - 1) / kDgramTile;
    unsigned const grid = grid_for(n_tiles, 1, c->sm_count, 8);
    size_t const smem = kDgramTile + 2 + max_gap + 3;
    dgram_kernel<COUNT><<<grid, kDgramTile, smem, c->stream>>>(g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records,
                                                               min_gap, max_gap, d_per_bin, d_range);
    ++c->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}
} // extern "C++"

int trx_dgram_count_dev(trx_ctx_t * c, const uint8_t * d_residues, uint64_t n_residues, const uint64_t * d_rec_offsets,
                        const uint32_t * d_rec_bin, uint64_t n_records, uint32_t min_gap, uint32_t max_gap, uint64_t n_bins,
                        uint64_t * d_per_bin)
{
    if (!c || !d_per_bin || n_bins == 0)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_dgram_args(min_gap, max_gap));
    TRX_TRY(c->use());
    TRX_CUDA(cudaMemsetAsync(d_per_bin, 0, n_bins * 8, c->stream));
    if (n_records == 0 || n_residues == 0)
        return TRX_OK;
    if (!d_residues || !d_rec_offsets || !d_rec_bin)
        return set_error(TRX_EINVAL, "NULL argument");
    ibf_geom g{};
    return launch_dgram<true>(c, g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, min_gap, max_gap,
                              reinterpret_cast<unsigned long long *>(d_per_bin));
}

int trx_dgram_emplace_dev(trx_ibf_t * f, const uint8_t * d_residues, uint64_t n_residues, const uint64_t * d_rec_offsets,
                          const uint32_t * d_rec_bin, uint64_t n_records, uint32_t min_gap, uint32_t max_gap)
{
    if (!f)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_dgram_args(min_gap, max_gap));
    if (n_records == 0 || n_residues == 0)
        return TRX_OK;
    if (!d_residues || !d_rec_offsets || !d_rec_bin)
        return set_error(TRX_EINVAL, "NULL argument");
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    uint64_t const matrix_bytes = f->bin_size * f->shard_words * 8;
    uint64_t const n_slabs = f->shard_words; // one 64-bin word column each
    char const * const slab_env = getenv("TRX_EMPLACE_SLAB");
    uint64_t slab_min = 512ull << 20; // TRX_DGRAM_SLAB_MIN_MB: tests take the slab route on small matrices
    if (char const * e = getenv("TRX_DGRAM_SLAB_MIN_MB"))
        slab_min = static_cast<uint64_t>(atoll(e)) << 20;
    if (matrix_bytes > slab_min && n_slabs > 1 && f->bin_size * 8 <= (1ull << 30) && f->shard_words == f->bin_words &&
        !(slab_env && strcmp(slab_env, "off") == 0))
    {
        // Far beyond L2 every atomic is a DRAM read-modify-write (46 ps each measured here).  Records are grouped by bin,
        // so the codes of one 64-bin slab come from one contiguous stretch of the text: its bits are collected in a DENSE
        // copy of the slab's word column (rows x 8 B — C3 with gaps 3..21: 189 MB, hit by ~10 atomics per sector, served
        // from L2) and folded into the interleaved matrix by one strided pass — K1's slab route.  The stretches are found
        // on the device (one binary search per slab), so nothing here waits for the host.
        TRX_TRY(f->slab.reserve(f->bin_size * 8));
        auto * d_slab = static_cast<uint64_t *>(f->slab.ptr);
        TRX_TRY(c->d_out[2].reserve((n_slabs + 1) * 8));
        auto * d_ranges = static_cast<uint64_t *>(c->d_out[2].ptr);
        dgram_slab_ranges_kernel<<<static_cast<unsigned>((n_slabs + 1 + 127) / 128), 128, 0, c->stream>>>(d_rec_offsets, d_rec_bin,
                                                                                                         n_records, n_slabs, d_ranges);
        ++c->launches;
        uint64_t const tiles_per_slab = (n_residues / n_slabs) / kDgramTile + 2; // grid sizing only: the kernel strides
        for (uint64_t s = 0; s < n_slabs; ++s)
        {
            TRX_CUDA(cudaMemsetAsync(d_slab, 0, f->bin_size * 8, c->stream));
            ibf_geom g = geom_of(f);
            g.slab = d_slab;
            g.slab_words = 1;
            g.slab_begin = f->shard_word_begin + s;
            TRX_TRY(launch_dgram<false>(c, g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, min_gap, max_gap, nullptr,
                                        d_ranges + s, tiles_per_slab));
            slab_merge_kernel<<<grid_for(f->bin_size, 256, c->sm_count, 16), 256, 0, c->stream>>>(f->d_words, f->shard_words, s, d_slab,
                                                                                                 1, f->bin_size);
            ++c->launches;
        }
        TRX_CUDA(cudaGetLastError());
        return TRX_OK;
    }
    // like K1: a matrix beyond L2 (but not far beyond) is filled in row slices that fit: every pass re-derives the codes —
    // reading shared memory — and fires only the atomics of its slice
    uint64_t window = 64ull << 20;
    if (char const * e = getenv("TRX_EMPLACE_WINDOW_MB"))
        window = static_cast<uint64_t>(atoll(e)) << 20;
    uint64_t passes = window ? (matrix_bytes + window - 1) / window : 1;
    if (passes < 1 || passes > 10)
        passes = 1;
    uint64_t const rows_per_pass = (f->bin_size + passes - 1) / passes;
    for (uint64_t p = 0; p < passes; ++p)
    {
        ibf_geom g = geom_of(f);
        g.row_lo = p * rows_per_pass;
        g.row_hi = g.row_lo + rows_per_pass < f->bin_size ? g.row_lo + rows_per_pass : f->bin_size;
        if (g.row_lo >= g.row_hi)
            break;
        TRX_TRY(launch_dgram<false>(c, g, d_residues, n_residues, d_rec_offsets, d_rec_bin, n_records, min_gap, max_gap, nullptr));
    }
    return TRX_OK;
}

static int check_record_table(const uint64_t * rec_offsets, const uint32_t * rec_bin, uint64_t n_records, uint64_t n_bins)
{
    if (rec_offsets[0] != 0)
        return set_error(TRX_EINVAL, "rec_offsets[0] must be 0");
    for (uint64_t i = 0; i < n_records; ++i)
    {
        if (rec_offsets[i + 1] < rec_offsets[i])
            return set_error(TRX_EINVAL, "rec_offsets must be non-decreasing");
        if (rec_bin[i] >= n_bins)
            return set_error(TRX_ERANGE, "record %llu: bin %u >= bin count %llu", static_cast<unsigned long long>(i),
                             rec_bin[i], static_cast<unsigned long long>(n_bins));
    }
    return TRX_OK;
}

static int stage_records(trx_ctx * c, const uint8_t * residues, const uint64_t * rec_offsets, const uint32_t * rec_bin,
                         uint64_t n_records)
{
    uint64_t const n_res = rec_offsets[n_records];
    TRX_TRY(c->d_in[0].reserve(n_res + 64));
    TRX_TRY(c->d_in[1].reserve((n_records + 1) * 8));
    TRX_TRY(c->d_in[2].reserve(n_records * 4));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, residues, n_res, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[1].ptr, rec_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[2].ptr, rec_bin, n_records * 4, cudaMemcpyHostToDevice, c->stream));
    return TRX_OK;
}

int trx_dgram_count(trx_ctx_t * c, const uint8_t * residues, const uint64_t * rec_offsets, const uint32_t * rec_bin,
                    uint64_t n_records, uint32_t min_gap, uint32_t max_gap, uint64_t n_bins, uint64_t * per_bin)
{
    if (!c || !per_bin || n_bins == 0 || (n_records && (!residues || !rec_offsets || !rec_bin)))
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_dgram_args(min_gap, max_gap));
    for (uint64_t b = 0; b < n_bins; ++b)
        per_bin[b] = 0;
    if (n_records == 0)
        return TRX_OK;
    TRX_TRY(check_record_table(rec_offsets, rec_bin, n_records, n_bins));
    TRX_TRY(c->use());
    TRX_TRY(stage_records(c, residues, rec_offsets, rec_bin, n_records));
    TRX_TRY(c->d_out[0].reserve(n_bins * 8));
    TRX_TRY(trx_dgram_count_dev(c, static_cast<uint8_t *>(c->d_in[0].ptr), rec_offsets[n_records],
                                static_cast<uint64_t *>(c->d_in[1].ptr), static_cast<uint32_t *>(c->d_in[2].ptr), n_records,
                                min_gap, max_gap, n_bins, static_cast<uint64_t *>(c->d_out[0].ptr)));
    TRX_CUDA(cudaMemcpyAsync(per_bin, c->d_out[0].ptr, n_bins * 8, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

int trx_dgram_emplace(trx_ibf_t * f, const uint8_t * residues, const uint64_t * rec_offsets, const uint32_t * rec_bin,
                      uint64_t n_records, uint32_t min_gap, uint32_t max_gap)
{
    if (!f || (n_records && (!residues || !rec_offsets || !rec_bin)))
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(check_dgram_args(min_gap, max_gap));
    if (n_records == 0)
        return TRX_OK;
    TRX_TRY(check_record_table(rec_offsets, rec_bin, n_records, f->bins));
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    TRX_TRY(stage_records(c, residues, rec_offsets, rec_bin, n_records));
    TRX_TRY(trx_dgram_emplace_dev(f, static_cast<uint8_t *>(c->d_in[0].ptr), rec_offsets[n_records],
                                  static_cast<uint64_t *>(c->d_in[1].ptr), static_cast<uint32_t *>(c->d_in[2].ptr), n_records,
                                  min_gap, max_gap));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

// ------------------------------------------------------------------------------------------- K2
int trx_bulk_contains_dev(trx_ibf_t * f, const uint64_t * d_kmers, uint64_t n, uint64_t * d_out_words)
{
    if (!f)
        return set_error(TRX_EINVAL, "NULL argument");
    TRX_TRY(f->ctx->use());
    return launch_probe<0>(f->ctx, geom_of(f), d_kmers, n, d_out_words, nullptr, nullptr);
}

int trx_bulk_contains(trx_ibf_t * f, const uint64_t * kmers, uint64_t n, uint64_t * out_words)
{
    if (!f || (n && (!kmers || !out_words)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    uint64_t const out_bytes = n * f->shard_words * 8;
    if (n <= 64 && out_bytes <= (64u << 10))
    {
        // a handful of k-mers (one query's k-graph): the kernel reads them from pinned host memory and writes the hit
        // vectors back into it — one launch and one synchronisation instead of two staged copies around them
        size_t const off_out = (n * 8 + 255) & ~size_t{255};
        c->h_pin[0].pinned_host = true;
        TRX_TRY(c->h_pin[0].reserve(off_out + out_bytes));
        auto * pin = static_cast<uint8_t *>(c->h_pin[0].ptr);
        memcpy(pin, kmers, n * 8);
        TRX_TRY(trx_bulk_contains_dev(f, reinterpret_cast<uint64_t *>(pin), n, reinterpret_cast<uint64_t *>(pin + off_out)));
        TRX_CUDA(cudaStreamSynchronize(c->stream));
        memcpy(out_words, pin + off_out, out_bytes);
        return TRX_OK;
    }
    TRX_TRY(c->d_in[0].reserve(n * 8));
    TRX_TRY(c->d_out[0].reserve(out_bytes));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_TRY(trx_bulk_contains_dev(f, static_cast<uint64_t *>(c->d_in[0].ptr), n, static_cast<uint64_t *>(c->d_out[0].ptr)));
    TRX_CUDA(cudaMemcpyAsync(out_words, c->d_out[0].ptr, out_bytes, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

// --------------------------------------------------------------------------------- path algebra
int trx_paths_create(trx_ctx_t * ctx, uint64_t words_per_slot, uint64_t capacity, trx_paths_t ** out)
{
    if (!ctx || !out || words_per_slot == 0 || capacity == 0)
        return set_error(TRX_EINVAL, "trx_paths_create: bad argument");
    TRX_TRY(ctx->use());
    auto * p = new (std::nothrow) trx_paths;
    if (!p)
        return set_error(TRX_ENOMEM, "out of host memory");
    p->ctx = ctx;
    p->words_per_slot = words_per_slot;
    p->capacity = capacity;
    cudaError_t e = cudaMalloc(&p->d_words, words_per_slot * capacity * 8);
    if (e != cudaSuccess)
    {
        delete p;
        return set_error(TRX_ENOMEM, "cudaMalloc paths: %s", cudaGetErrorString(e));
    }
    cudaMemsetAsync(p->d_words, 0, words_per_slot * capacity * 8, ctx->stream);
    *out = p;
    return TRX_OK;
}

void trx_paths_destroy(trx_paths_t * p)
{
    if (!p)
        return;
    cudaSetDevice(p->ctx->device);
    cudaStreamSynchronize(p->ctx->stream);
    cudaFree(p->d_words);
    delete p;
}

void * trx_paths_device_words(trx_paths_t * p)
{
    return p ? p->d_words : nullptr;
}

int trx_paths_fill(trx_paths_t * p, uint64_t first, uint64_t n, uint64_t bins)
{
    if (!p)
        return set_error(TRX_EINVAL, "NULL argument");
    if (first + n > p->capacity)
        return set_error(TRX_ERANGE, "slots [%llu,%llu) exceed capacity %llu", (unsigned long long)first,
                         (unsigned long long)(first + n), (unsigned long long)p->capacity);
    if (bins > p->words_per_slot * 64)
        return set_error(TRX_EINVAL, "bins exceed slot width");
    if (n == 0)
        return TRX_OK;
    TRX_TRY(p->ctx->use());
    unsigned grid = grid_for(n * p->words_per_slot, 256, p->ctx->sm_count, 8);
    paths_fill_kernel<<<grid, 256, 0, p->ctx->stream>>>(p->d_words, p->words_per_slot, first, n, bins);
    ++p->ctx->launches;
    TRX_CUDA(cudaGetLastError());
    return TRX_OK;
}

static int stage_u32(trx_ctx * c, int which, uint32_t const * host, uint64_t n, uint32_t ** dev)
{
    TRX_TRY(c->d_in[which].reserve(n * 4));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[which].ptr, host, n * 4, cudaMemcpyHostToDevice, c->stream));
    *dev = static_cast<uint32_t *>(c->d_in[which].ptr);
    return TRX_OK;
}

static int check_slots(trx_paths const * p, uint32_t const * slots, uint64_t n)
{
    uint32_t worst = 0; // branch-free maximum first (vectorises: 4 Mi slots in well under a millisecond), report second
    for (uint64_t i = 0; i < n; ++i)
        worst = slots[i] > worst ? slots[i] : worst;
    if (n && worst >= p->capacity)
        return set_error(TRX_ERANGE, "slot %u >= capacity %llu", worst, (unsigned long long)p->capacity);
    return TRX_OK;
}

int trx_paths_step(trx_ibf_t * f, trx_paths_t * p, const uint64_t * kmers, const uint32_t * slots, uint64_t n,
                   uint8_t * alive)
{
    if (!f || !p || (n && (!kmers || !slots || !alive)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (p->words_per_slot != f->shard_words || p->ctx != f->ctx)
        return set_error(TRX_EINVAL, "path slots (%llu words) do not match the IBF shard (%llu words)",
                         (unsigned long long)p->words_per_slot, (unsigned long long)f->shard_words);
    if (n == 0)
        return TRX_OK;
    TRX_TRY(check_slots(p, slots, n));
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    if (n <= kZeroCopyItems)
    {
        // the collector's usual case — a handful of items per rank of the k-graph: the kernel reads k-mers and slots
        // straight from pinned host memory and writes the alive bytes there; one launch + one sync instead of two
        // staged uploads, a launch, a staged download and a sync (~35 us -> ~10 us per step)
        c->h_pin[0].pinned_host = true;
        TRX_TRY(c->h_pin[0].reserve(kZeroCopyItems * 13 + 64));
        auto * base = static_cast<uint8_t *>(c->h_pin[0].ptr);
        auto * h_kmers = reinterpret_cast<uint64_t *>(base);
        auto * h_slots = reinterpret_cast<uint32_t *>(base + kZeroCopyItems * 8);
        uint8_t * h_alive = base + kZeroCopyItems * 12;
        memcpy(h_kmers, kmers, n * 8);
        memcpy(h_slots, slots, n * 4);
        TRX_TRY(launch_probe<1>(c, geom_of(f), h_kmers, n, p->d_words, h_slots, h_alive));
        TRX_CUDA(cudaStreamSynchronize(c->stream));
        c->pin_arena_off = 0;
        memcpy(alive, h_alive, n);
        return TRX_OK;
    }
    TRX_TRY(c->d_in[0].reserve(n * 8));
    TRX_TRY(c->d_out[1].reserve(n));
    TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
    uint32_t * d_slots = nullptr;
    TRX_TRY(stage_u32(c, 1, slots, n, &d_slots));
    TRX_TRY(launch_probe<1>(c, geom_of(f), static_cast<uint64_t *>(c->d_in[0].ptr), n, p->d_words, d_slots,
                            static_cast<uint8_t *>(c->d_out[1].ptr)));
    TRX_CUDA(cudaMemcpyAsync(alive, c->d_out[1].ptr, n, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

int trx_paths_step_dev(trx_ibf_t * f, trx_paths_t * p, const uint64_t * d_kmers, const uint32_t * d_slots, uint64_t n,
                       uint8_t * d_alive)
{
    if (!f || !p || (n && (!d_kmers || !d_slots || !d_alive)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (p->words_per_slot != f->shard_words || p->ctx != f->ctx)
        return set_error(TRX_EINVAL, "path slots (%llu words) do not match the IBF shard (%llu words)",
                         (unsigned long long)p->words_per_slot, (unsigned long long)f->shard_words);
    if (n == 0)
        return TRX_OK;
    TRX_TRY(f->ctx->use());
    return launch_probe<1>(f->ctx, geom_of(f), d_kmers, n, p->d_words, d_slots, d_alive);
}

int trx_paths_step_sharded(trx_ibf_t * f, trx_paths_t * p, trx_comm_t * comm, const uint64_t * kmers,
                           const uint32_t * slots, uint64_t n, uint8_t * alive)
{
    if (!f || !p || !comm || (n && (!kmers || !slots || !alive)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    TRX_TRY(check_slots(p, slots, n));
    trx_ctx * c = f->ctx;
    TRX_TRY(c->use());
    // k-mers and slots are read from pinned host memory (small steps) or staged; the alive bytes are written to
    // DEVICE memory, OR-ed over the shards there (NCCL, same stream) and only then copied out: one sync per step
    uint64_t const * src_kmers = nullptr;
    uint32_t const * src_slots = nullptr;
    if (n <= kZeroCopyItems)
    {
        c->h_pin[0].pinned_host = true;
        TRX_TRY(c->h_pin[0].reserve(kZeroCopyItems * 13 + 64));
        auto * base = static_cast<uint8_t *>(c->h_pin[0].ptr);
        memcpy(base, kmers, n * 8);
        memcpy(base + kZeroCopyItems * 8, slots, n * 4);
        src_kmers = reinterpret_cast<uint64_t *>(base);
        src_slots = reinterpret_cast<uint32_t *>(base + kZeroCopyItems * 8);
    }
    else
    {
        TRX_TRY(c->d_in[0].reserve(n * 8));
        TRX_CUDA(cudaMemcpyAsync(c->d_in[0].ptr, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
        uint32_t * d_slots = nullptr;
        TRX_TRY(stage_u32(c, 1, slots, n, &d_slots));
        src_kmers = static_cast<uint64_t *>(c->d_in[0].ptr);
        src_slots = d_slots;
    }
    TRX_TRY(c->d_out[1].reserve(n));
    auto * d_alive = static_cast<uint8_t *>(c->d_out[1].ptr);
    TRX_TRY(trx_paths_step_dev(f, p, src_kmers, src_slots, n, d_alive));
    TRX_TRY(trx_comm_any_alive_dev(comm, d_alive, n));
    TRX_CUDA(cudaMemcpyAsync(alive, d_alive, n, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    c->pin_arena_off = 0;
    return TRX_OK;
}

int trx_paths_and_hits(trx_paths_t * p, const uint64_t * d_hits, const uint32_t * hit_index, const uint32_t * slots,
                       uint64_t n, uint8_t * alive)
{
    if (!p || (n && (!d_hits || !hit_index || !slots || !alive)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    TRX_TRY(check_slots(p, slots, n));
    trx_ctx * c = p->ctx;
    TRX_TRY(c->use());
    if (n <= kZeroCopyItems)
    {
        c->h_pin[0].pinned_host = true;
        TRX_TRY(c->h_pin[0].reserve(kZeroCopyItems * 13 + 64));
        auto * base = static_cast<uint8_t *>(c->h_pin[0].ptr);
        auto * h_idx = reinterpret_cast<uint32_t *>(base);
        auto * h_slots = reinterpret_cast<uint32_t *>(base + kZeroCopyItems * 8);
        uint8_t * h_alive = base + kZeroCopyItems * 12;
        memcpy(h_idx, hit_index, n * 4);
        memcpy(h_slots, slots, n * 4);
        TRX_TRY(paths_and_hits_launch(c, p->d_words, p->words_per_slot, d_hits, h_idx, h_slots, n, h_alive));
        TRX_CUDA(cudaStreamSynchronize(c->stream));
        c->pin_arena_off = 0;
        memcpy(alive, h_alive, n);
        return TRX_OK;
    }
    uint32_t *d_idx = nullptr, *d_slots = nullptr;
    TRX_TRY(stage_u32(c, 0, hit_index, n, &d_idx));
    TRX_TRY(stage_u32(c, 1, slots, n, &d_slots));
    TRX_TRY(c->d_out[1].reserve(n));
    TRX_TRY(paths_and_hits_launch(c, p->d_words, p->words_per_slot, d_hits, d_idx, d_slots, n,
                                  static_cast<uint8_t *>(c->d_out[1].ptr)));
    TRX_CUDA(cudaMemcpyAsync(alive, c->d_out[1].ptr, n, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

int trx_paths_or(trx_paths_t * p, const uint32_t * dst, const uint32_t * src, uint64_t n)
{
    if (!p || (n && (!dst || !src)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    TRX_TRY(check_slots(p, dst, n));
    TRX_TRY(check_slots(p, src, n));
    trx_ctx * c = p->ctx;
    TRX_TRY(c->use());
    uint32_t *d_dst = nullptr, *d_src = nullptr;
    if (n <= kZeroCopyItems)
    {
        // fire and forget: the lists go into the pinned arena and the kernel reads them from there whenever it runs
        constexpr size_t kArenaBytes = 64 * kZeroCopyItems * 8;
        c->h_pin[1].pinned_host = true;
        TRX_TRY(c->h_pin[1].reserve(kArenaBytes));
        if (c->pin_arena_off + n * 8 > kArenaBytes)
        {
            TRX_CUDA(cudaStreamSynchronize(c->stream)); // every earlier reader is done
            c->pin_arena_off = 0;
        }
        d_dst = reinterpret_cast<uint32_t *>(static_cast<uint8_t *>(c->h_pin[1].ptr) + c->pin_arena_off);
        d_src = d_dst + n;
        memcpy(d_dst, dst, n * 4);
        memcpy(d_src, src, n * 4);
        c->pin_arena_off += (n * 8 + 15) & ~size_t{15};
    }
    else
    {
        TRX_TRY(stage_u32(c, 0, dst, n, &d_dst));
        TRX_TRY(stage_u32(c, 1, src, n, &d_src));
    }
    uint64_t const W = p->words_per_slot;
    int const G = pow2_ceil(W);
    auto go = [&](auto gtag) -> int
    {
        constexpr int GG = decltype(gtag)::value;
        unsigned grid = grid_for(n, 256 / GG, c->sm_count, 8);
        paths_or_kernel<GG><<<grid, 256, 0, c->stream>>>(p->d_words, W, d_dst, d_src, n);
        ++c->launches;
        TRX_CUDA(cudaGetLastError());
        return TRX_OK;
    };
    int rc;
    switch (G)
    {
    case 1: rc = go(std::integral_constant<int, 1>{}); break;
    case 2: rc = go(std::integral_constant<int, 2>{}); break;
    case 4: rc = go(std::integral_constant<int, 4>{}); break;
    case 8: rc = go(std::integral_constant<int, 8>{}); break;
    case 16: rc = go(std::integral_constant<int, 16>{}); break;
    default: rc = go(std::integral_constant<int, 32>{}); break;
    }
    TRX_TRY(rc);
    // arena path: nothing to wait for — the lists stay where they are until the arena wraps (which synchronises) or a
    // later call that reads results back has synchronised and reset it.  Staged path: the caller's arrays may have been
    // pinned (then the copies above are truly asynchronous) and may be rewritten as soon as this returns.
    if (n > kZeroCopyItems)
        TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

int trx_paths_read(trx_paths_t * p, uint64_t first, uint64_t n, uint64_t * out_words)
{
    if (!p || !out_words)
        return set_error(TRX_EINVAL, "NULL argument");
    if (first + n > p->capacity)
        return set_error(TRX_ERANGE, "slots out of range");
    TRX_TRY(p->ctx->use());
    TRX_CUDA(cudaMemcpyAsync(out_words, p->d_words + first * p->words_per_slot, n * p->words_per_slot * 8,
                             cudaMemcpyDeviceToHost, p->ctx->stream));
    TRX_CUDA(cudaStreamSynchronize(p->ctx->stream));
    p->ctx->pin_arena_off = 0;
    return TRX_OK;
}

// ------------------------------------------------------------------------------- no-allocation probe
// trx_bulk_contains for callers that must not allocate (CUDA graph capture around them, tight serving loops, harnesses
// that forbid cudaMalloc): the staging that trx_bulk_contains keeps in grow-on-demand cudaMalloc'ed buffers (k-mers in,
// hit vectors out) lives in ONE file-scope device array of fixed size instead.  Every GPU that loads this library holds
// the array (TRX_NOALLOC_SCRATCH_BYTES) until the process exits; the
// reference has nothing of the kind.
}  // extern "C" (the array and its address cache are file-local)

static __device__ __align__(256) unsigned char g_noalloc_scratch[TRX_NOALLOC_SCRATCH_BYTES];

namespace
{
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mutex;
void * g_noalloc_ptr[kNoallocMaxDevices] = {}; // per device: the DEVICE address of g_noalloc_scratch on that device

// In host code `g_noalloc_scratch` is only a stand-in symbol, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, asked once per device with that device current.
int noalloc_scratch_for(int device, unsigned char ** out)
{
    if (device < 0 || device >= kNoallocMaxDevices)
        return set_error(TRX_EINVAL, "device %d: the no-allocation scratch is kept for devices 0..%d", device,
                         kNoallocMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device])
    {
        TRX_CUDA(cudaSetDevice(device));
        void * ptr = nullptr;
        TRX_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (!ptr)
            return set_error(TRX_ECUDA, "cudaGetSymbolAddress returned a null address for the no-allocation scratch");
        g_noalloc_ptr[device] = ptr;
    }
    *out = static_cast<unsigned char *>(g_noalloc_ptr[device]);
    return TRX_OK;
}
} // namespace

extern "C"
{

uint64_t trx_bulk_contains_noalloc_max_kmers(const trx_ibf_t * f)
{
    if (!f)
        return 0;
    return TRX_NOALLOC_SCRATCH_BYTES / (8 + f->shard_words * 8);
}

int trx_bulk_contains_noalloc(trx_ibf_t * f, const uint64_t * kmers, uint64_t n, uint64_t * out_words)
{
    if (!f || (n && (!kmers || !out_words)))
        return set_error(TRX_EINVAL, "NULL argument");
    if (n == 0)
        return TRX_OK;
    uint64_t const out_bytes = n * f->shard_words * 8;
    uint64_t const in_bytes = (n * 8 + 255) & ~uint64_t{255}; // the hit vectors start 256-byte aligned
    if (n > trx_bulk_contains_noalloc_max_kmers(f) || in_bytes + out_bytes > TRX_NOALLOC_SCRATCH_BYTES)
        return set_error(TRX_ECAPACITY,
                         "trx_bulk_contains_noalloc: %llu k-mers x %llu words need %llu bytes of staging, the fixed scratch "
                         "holds %llu (at most %llu k-mers per call for this filter); split the batch or use trx_bulk_contains",
                         (unsigned long long)n, (unsigned long long)f->shard_words,
                         (unsigned long long)(in_bytes + out_bytes), (unsigned long long)TRX_NOALLOC_SCRATCH_BYTES,
                         (unsigned long long)trx_bulk_contains_noalloc_max_kmers(f));
    trx_ctx * c = f->ctx;
    unsigned char * scratch = nullptr;
    TRX_TRY(noalloc_scratch_for(c->device, &scratch));
    TRX_TRY(c->use());
    auto * d_kmers = reinterpret_cast<uint64_t *>(scratch);
    auto * d_out = reinterpret_cast<uint64_t *>(scratch + in_bytes);
    // nothing here has to start at zero: the k-mers are copied in, and the probe STORES every word of every hit vector
    // (MODE 0 writes the AND, it never accumulates), so whatever the previous call left in the array is overwritten
    TRX_CUDA(cudaMemcpyAsync(d_kmers, kmers, n * 8, cudaMemcpyHostToDevice, c->stream));
    TRX_TRY(launch_probe<0>(c, geom_of(f), d_kmers, n, d_out, nullptr, nullptr));
    TRX_CUDA(cudaMemcpyAsync(out_words, d_out, out_bytes, cudaMemcpyDeviceToHost, c->stream));
    TRX_CUDA(cudaStreamSynchronize(c->stream));
    return TRX_OK;
}

} // extern "C"
