// capi.cpp — C facade of the host layer (libtrxhost.so) for drivers that are not C++: one process per GPU under
// torchrun opens its shard of an index file, runs process_query with the global pruning decision delegated to a
// callback (torch.distributed all-reduce in tetrex_b200/sharded.py) and verifies its own bins.  Plain pointers and
// sizes only; every call returns 0 or -1 with the message in trxh_last_error().
#include <csignal>
#include <cstring>
#include <sstream>

#include <execinfo.h>
#include <unistd.h>

#include "../../include/tetrex_host.h"
#include "rh_order.hpp"
#include "hibf_layout.hpp"
#include "trx_host.hpp"

namespace
{
thread_local std::string g_error;

struct handle
{
    trx_ctx_t * ctx;
    trxh::DeviceIndex index;
    trx_comm_t * comm{nullptr}; // not owned
    bool augment{false};        // `-a`
    std::unique_ptr<trxh::DGramIndex> dibf; // `-g`
    explicit handle(trx_ctx_t * c) : ctx(c), index(c) {}
};

template <class F>
int guarded(F && f)
{
    try
    {
        f();
        return 0;
    }
    catch (std::exception const & e)
    {
        g_error = e.what();
        return -1;
    }
}
} // namespace

namespace
{
void segv_backtrace(int)
{
    void * frames[64];
    int const n = backtrace(frames, 64);
    backtrace_symbols_fd(frames, n, 2);
    _exit(139);
}
} // namespace

extern "C"
{

// diagnostic: native backtrace (module + offset, for addr2line) on SIGSEGV instead of a silent death of the driver
void trxh_debug_backtrace_on_segv(void)
{
    signal(SIGSEGV, segv_backtrace);
}

const char * trxh_last_error(void)
{
    return g_error.c_str();
}

// load_ibf (include/index_base.h:189-195) of shard `shard_index` of `shard_count` onto the device of `ctx`
int trxh_open(trx_ctx_t * ctx, const char * path, uint32_t shard_index, uint32_t shard_count, void ** out)
{
    return guarded([&] {
        auto h = std::make_unique<handle>(ctx);
        h->index.set_shard(shard_index, shard_count);
        h->index.load(path);
        *out = h.release();
    });
}

void trxh_close(void * h)
{
    delete static_cast<handle *>(h);
}

// From now on a column-sharded trxh_process_query reduces its pruning flags and gathers its candidate words through this
// communicator (NCCL, on the device); pass NULL to detach.  The communicator stays the caller's.
int trxh_attach_comm(void * hv, trx_comm_t * comm)
{
    return guarded([&] { static_cast<handle *>(hv)->comm = comm; });
}

// resident processes: read and upload every bin's text on the first verify instead of growing query by query
int trxh_preload_text(void * hv, int on)
{
    return guarded([&] { static_cast<handle *>(hv)->index.preload_all_text(on != 0); });
}

// `tetrex query -a`: trxh_process_query augments the k-graph (OTFCollector::augment) before the walk
int trxh_set_augment(void * hv, int on)
{
    return guarded([&] { static_cast<handle *>(hv)->augment = on != 0; });
}

// `tetrex query -g FILE`: load_dindex (include/dGramIndex.h:320-326) onto the handle's device; the collector's gapped items
// probe it from now on.  NULL or "" detaches.
int trxh_attach_dgram(void * hv, const char * path)
{
    return guarded([&] {
        auto * h = static_cast<handle *>(hv);
        h->index.attach_dgram(nullptr);
        h->dibf.reset();
        if (!path || !*path)
            return;
        auto d = std::make_unique<trxh::DGramIndex>(h->ctx);
        d->load(path);
        h->index.attach_dgram(d.get());
        h->dibf = std::move(d);
    });
}

// drive_dindex's work (src/dGramIndex.cpp:35-37) on ctx's device: DGramIndex(min, max, 1, hash_count, fpr, files),
// populate(), save_dindex(out_path).  The file names are stored as given.  per_bin (nullable, n_files entries) receives
// dgram_buffer_[b].size().
int trxh_dindex_build(trx_ctx_t * ctx, const char * const * files, uint64_t n_files, uint64_t min_gap, uint64_t max_gap,
                      uint32_t hash_count, float fpr, const char * out_path, uint64_t * per_bin)
{
    return guarded([&] {
        trxh::dindex_arguments args;
        args.min = min_gap;
        args.max = max_gap;
        args.hash_count = static_cast<uint8_t>(hash_count);
        args.fpr = fpr;
        args.acid_libs.assign(files, files + n_files);
        trxh::DGramIndex d(ctx);
        d.build(args);
        if (per_bin)
            for (uint64_t b = 0; b < n_files; ++b)
                per_bin[b] = d.codes_per_bin()[b];
        if (out_path && *out_path)
            d.store(out_path);
    });
}

// Host only: seqan::hibf's layout for user bins given as k-mer lists (values[bin_offsets[b] .. bin_offsets[b+1]) in the
// order HIBFIndex::emplace collected them): HyperLogLog sketches, k-mer count estimates, compute_layout
// (lib/hibf/src/hierarchical_interleaved_bloom_filter.cpp:212-236).  Dump (uint64 words): tmax, top_level_max_bin_id,
// #max_bins, per max bin {len, indices..., id}, #user_bins, per user bin {len, prev..., storage_TB_id,
// number_of_technical_bins, idx}, then the n_bins estimates.  Returns the number of words (may exceed cap), 0 on error.
uint64_t trxh_hibf_layout(const uint64_t * values, const uint64_t * bin_offsets, uint64_t n_bins, uint32_t hash_funs, float fpr,
                          uint64_t * out, uint64_t cap)
{
    uint64_t n = 0;
    int const rc = guarded([&] {
        namespace h = trxh::hibf;
        h::config cfg;
        cfg.number_of_user_bins = n_bins;
        cfg.number_of_hash_functions = hash_funs;
        cfg.maximum_fpr = fpr; // float → double, like `.maximum_fpr = fpr_` (include/index_hibf.h:127)
        cfg.validate_and_set_defaults();
        std::vector<h::hyperloglog> sketches(n_bins, h::hyperloglog{cfg.sketch_bits});
        std::vector<size_t> counts(n_bins);
        for (uint64_t b = 0; b < n_bins; ++b)
        {
            for (uint64_t i = bin_offsets[b]; i < bin_offsets[b + 1]; ++i)
                sketches[b].add(values[i]);
            counts[b] = sketches[b].estimate();
        }
        h::layout const l = h::compute_layout(cfg, counts, sketches);
        std::vector<uint64_t> w{cfg.tmax, l.top_level_max_bin_id, l.max_bins.size()};
        for (auto const & mb : l.max_bins)
        {
            w.push_back(mb.previous_TB_indices.size());
            w.insert(w.end(), mb.previous_TB_indices.begin(), mb.previous_TB_indices.end());
            w.push_back(mb.id);
        }
        w.push_back(l.user_bins.size());
        for (auto const & ub : l.user_bins)
        {
            w.push_back(ub.previous_TB_indices.size());
            w.insert(w.end(), ub.previous_TB_indices.begin(), ub.previous_TB_indices.end());
            w.push_back(ub.storage_TB_id);
            w.push_back(ub.number_of_technical_bins);
            w.push_back(ub.idx);
        }
        w.insert(w.end(), counts.begin(), counts.end());
        for (size_t i = 0; i < w.size() && i < cap; ++i)
            out[i] = w[i];
        n = w.size();
    });
    return rc == 0 ? n : 0;
}

// out[0..8) = bin_count, bin_words, is_hibf, is_dna, k, shard_word_begin, shard_words, reduction
int trxh_info(void * hv, uint64_t out[8])
{
    return guarded([&] {
        auto & ix = static_cast<handle *>(hv)->index;
        out[0] = ix.bin_count();
        out[1] = ix.bin_words();
        out[2] = ix.is_hibf();
        out[3] = ix.is_dna();
        out[4] = ix.k();
        out[5] = ix.shard_word_begin();
        out[6] = ix.shard_words();
        out[7] = ix.reduction();
    });
}

// process_query (include/query.h:226-247).  out_words: bin_words entries; a column shard fills only its own words.
// alive (nullable when unsharded): ORs n bytes over all shards in place.  stats (nullable): probes, ranks, items.
int trxh_process_query(void * hv, const char * regex, trxh_alive_fn alive, void * user, uint64_t * out_words, uint64_t * stats)
{
    return guarded([&] {
        auto & ix = static_cast<handle *>(hv)->index;
        trxh::shard_hooks hooks;
        if (alive)
            hooks.any_alive = [alive, user](uint8_t * a, size_t n) { alive(a, n, user); };
        hooks.comm = static_cast<handle *>(hv)->comm;
        trxh::collect_stats st;
        trxh::bitvector got = trxh::process_query(regex, ix, &st, (alive || hooks.comm) ? &hooks : nullptr,
                                                   static_cast<handle *>(hv)->augment);
        // run_collection (include/query.h:256-262): the answer is ANDed into a vector of the real bins
        for (uint64_t w = 0; w < got.size(); ++w)
        {
            uint64_t const first = w * 64, bins = ix.bin_count();
            uint64_t const mask = bins >= first + 64 ? ~0ull : (bins > first ? (1ull << (bins - first)) - 1 : 0);
            out_words[w] = got[w] & mask;
        }
        if (stats)
        {
            stats[0] = st.probes;
            stats[1] = st.ranks;
            stats[2] = st.items;
        }
    });
}

// process_query for n patterns with ONE membership batch for all their k-mers (what `tetrex query -f` does for its motif
// file).  out_words: n x bin_words; status[i] = 1 answered, 0 = process_query would have thrown for that pattern (its words
// are zero); stats (nullable): n x 3 like trxh_process_query.
int trxh_process_query_batch(void * hv, const char * const * regexes, uint64_t n, uint64_t * out_words, uint8_t * status, uint64_t * stats)
{
    return guarded([&] {
        auto * h = static_cast<handle *>(hv);
        auto & ix = h->index;
        std::vector<std::string> patterns(regexes, regexes + n);
        std::vector<trxh::batch_answer> got = trxh::process_query_batch(patterns, ix, h->augment);
        uint64_t const words = ix.bin_words(), bins = ix.bin_count();
        for (uint64_t i = 0; i < n; ++i)
        {
            status[i] = got[i].ok ? 1 : 0;
            for (uint64_t w = 0; w < words; ++w)
            {
                uint64_t const first = w * 64;
                uint64_t const mask = bins >= first + 64 ? ~0ull : (bins > first ? (1ull << (bins - first)) - 1 : 0);
                out_words[i * words + w] = got[i].ok && w < got[i].bins.size() ? got[i].bins[w] & mask : 0;
            }
            if (stats)
            {
                stats[3 * i + 0] = got[i].stats.probes;
                stats[3 * i + 1] = got[i].stats.ranks;
                stats[3 * i + 2] = got[i].stats.items;
            }
        }
    });
}

// process_query over membership rows the caller supplies (trxh::process_query_with_rows): no index file, no device.
// molecule 0 = DNA, 1 = peptide; rows(kmers, n, out_words, user) fills n x ceil(bins / 64) words.
int trxh_process_query_with_rows(int molecule, uint32_t k, int reduction, uint64_t bins, const char * regex, int augment,
                                 trxh_rows_fn rows, void * user, uint64_t * out_words, uint64_t * stats)
{
    return guarded([&] {
        if (!regex || !rows || !out_words)
            throw trxh::error("trxh_process_query_with_rows: NULL argument");
        trxh::DeviceIndex ix(nullptr);
        ix.describe_only(molecule == 0 ? "na" : "aa", static_cast<uint8_t>(k), static_cast<uint8_t>(reduction), bins);
        trxh::collect_stats st;
        trxh::bitvector got = trxh::process_query_with_rows(
            regex, ix, [rows, user](uint64_t const * kmers, size_t n, uint64_t * out) { rows(kmers, n, out, user); }, &st, augment != 0);
        for (uint64_t w = 0; w < got.size(); ++w)
        {
            uint64_t const first = w * 64;
            uint64_t const mask = bins >= first + 64 ? ~0ull : (bins > first ? (1ull << (bins - first)) - 1 : 0);
            out_words[w] = got[w] & mask;
        }
        if (stats)
        {
            stats[0] = st.probes;
            stats[1] = st.ranks;
            stats[2] = st.items;
        }
    });
}

// iter_disk_search (include/query.h:97-188) over this shard's bins; the lines come back as one malloc'ed buffer
// (free with trxh_free).
int trxh_verify(void * hv, const uint64_t * cand_words, const char * regex, char ** out_text, uint64_t * out_len, uint64_t * n_lines)
{
    return guarded([&] {
        auto & ix = static_cast<handle *>(hv)->index;
        trxh::bitvector cand(cand_words, cand_words + ix.bin_words());
        std::ostringstream os;
        *n_lines = trxh::iter_disk_search(cand, regex, ix, os);
        std::string const text = os.str();
        *out_text = static_cast<char *>(malloc(text.size() + 1));
        memcpy(*out_text, text.c_str(), text.size() + 1);
        *out_len = text.size();
    });
}

void trxh_free(void * p)
{
    free(p);
}

// host only: the order in which the collector drains a rank whose items arrived with these keys (duplicates are
// absorbed).  out[j] = index of the j-th item popped; returns how many.
uint64_t trxh_rank_pop_order(const uint64_t * keys, uint64_t n, uint32_t * out)
{
    trxh::rh_slots t;
    for (uint64_t i = 0; i < n; ++i)
        if (t.find(keys[i]) == trxh::rh_slots::npos)
            t.insert_new(keys[i], static_cast<uint32_t>(i));
    uint64_t j = 0;
    t.for_each([&](uint32_t v) { out[j++] = v; });
    return j;
}

// Host only: hibf's k-mer set (robin_hood::unordered_flat_set<uint64_t>) under a script — op[i] = 0: insert(keys[i]),
// 1: clear() — restated by rh_order.hpp; writes the final iteration order, returns the number of elements.
uint64_t trxh_flat_set_script(const uint64_t * keys, const uint8_t * op, uint64_t n, uint64_t * out)
{
    trxh::rh_slots t;
    for (uint64_t i = 0; i < n; ++i)
    {
        if (op[i] == 1)
            t.clear_keep_capacity();
        else
            t.insert_key(keys[i]);
    }
    uint64_t j = 0;
    t.for_each_key([&](uint64_t k) { out[j++] = k; });
    return j;
}

} // extern "C"
