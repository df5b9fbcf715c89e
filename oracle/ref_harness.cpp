// oracle/_ref harness — TEST INFRASTRUCTURE ONLY.
//
// This translation unit is compiled TOGETHER WITH THE UNMODIFIED REFERENCE SOURCES
// where they lie under /root/reference (recipe: oracle/Makefile, target `ref`) into
// oracle/_ref/libtetrex_ref.so.  It contains no algorithm of its own: every entry
// point forwards to the reference's classes (TetRex include/*.h, lib/hibf, lib/re2)
// and flattens their results into plain C arrays so that tests/ can compare our
// CUDA path and our C restatement (oracle/tetrex_oracle.c) with the real thing.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load the resulting library.  The product never links it.
//
// Known reference wrinkles worked around here:
//   * IBFIndex's 4-arg ctor reads an uninitialised member and throws; we default-
//     construct the index and fill the same members by hand (private → public for
//     this TU only), then run the reference's own populate_index().
//   * index_arguments::molecule/ibf are frozen before parsing; we call the
//     create_* paths through the C++ API instead of the CLI.

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <functional>
#include <iostream>
#include <list>
#include <memory>
#include <optional>
#include <set>
#include <sstream>
#include <stack>
#include <string>
#include <string_view>
#include <syncstream>
#include <unordered_map>
#include <vector>

#include <fcntl.h>
#include <omp.h>
#include <unistd.h>
#include <zlib.h>

#include <any>
#include <atomic>
#include <barrier>
#include <charconv>
#include <chrono>
#include <compare>
#include <condition_variable>
#include <deque>
#include <exception>
#include <forward_list>
#include <future>
#include <iomanip>
#include <ios>
#include <limits>
#include <locale>
#include <new>
#include <queue>
#include <regex>
#include <shared_mutex>
#include <stdexcept>
#include <thread>
#include <typeindex>
#include <typeinfo>
#include <unordered_set>
#include <valarray>
#include <bit>
#include <cassert>
#include <cinttypes>
#include <climits>
#include <concepts>
#include <iterator>
#include <map>
#include <mutex>
#include <numeric>
#include <random>
#include <ranges>
#include <span>
#include <tuple>
#include <type_traits>
#include <utility>
#include <variant>

#include <re2/re2.h>
#include <re2/set.h>

#include <hibf/config.hpp>
#include <hibf/hierarchical_interleaved_bloom_filter.hpp>
#include <hibf/interleaved_bloom_filter.hpp>
#include <hibf/misc/bit_vector.hpp>
#include <hibf/misc/divide_and_ceil.hpp>
#include <hibf/layout/compute_layout.hpp>
#include <hibf/layout/layout.hpp>
#include <hibf/sketch/compute_sketches.hpp>
#include <hibf/sketch/estimate_kmer_counts.hpp>
#include <hibf/sketch/hyperloglog.hpp>
#include <sharg/all.hpp>
#include <seqan3/core/debug_stream.hpp>
#include <lemon/smart_graph.h>
#include <lemon/connectivity.h>
#include <lemon/dfs.h>
#include <lemon/adaptors.h>

// TetRex's own headers only: IBFIndex / HIBFIndex / OTFCollector keep state private and the
// harness must read and seed it.  hibf is included above WITHOUT this define (the IBF
// inherits privately from bit_vector and cereal's dispatch depends on that).
#define private public
#include "utils.h"
#include "arg_parse.h"
#include "index_base.h"
#include "query.h"
#undef private

namespace
{

// Collects what decompose_record() hands to base_ref.emplace(); mirrors the three
// members of TetrexIndex the decomposers touch (include/index_base.h:32-33,99-102,136-140).
struct emit_sink
{
    uint64_t forward_store_{};
    uint64_t reverse_store_{};
    std::vector<uint64_t> values;
    void set_stores(uint64_t f, uint64_t r)
    {
        forward_store_ = f;
        reverse_store_ = r;
    }
    void emplace(uint64_t v, size_t)
    {
        values.push_back(v);
    }
};

struct ref_ibf
{
    seqan::hibf::interleaved_bloom_filter ibf;
};

struct ref_hibf
{
    seqan::hibf::hierarchical_interleaved_bloom_filter hibf;
    size_t user_bins{};
};

thread_local std::string g_err;

// Geometry as the IBF ctor derives it (lib/hibf/src/interleaved_bloom_filter.cpp:48-52).
void fill_params(seqan::hibf::interleaved_bloom_filter const & ibf, uint64_t * out)
{
    uint64_t const bins = ibf.bin_count();
    uint64_t const bin_words = (bins + 63) / 64;
    out[0] = bins;
    out[1] = bin_words * 64;
    out[2] = ibf.bin_size();
    out[3] = static_cast<uint64_t>(std::countl_zero(static_cast<uint64_t>(ibf.bin_size())));
    out[4] = bin_words;
    out[5] = ibf.hash_function_count();
    out[6] = (ibf.bit_size() + 63) / 64;
}

} // namespace

extern "C"
{

char const * ref_last_error()
{
    return g_err.c_str();
}

// ---------------------------------------------------------------- a1 / a2
// include/nucleotide_decomposer.h:81-111, include/peptide_decomposer.h:261-291
uint64_t ref_decompose(int molecule, int k, int reduction, char const * seq, uint64_t len, uint64_t * out, uint64_t cap)
{
    emit_sink sink;
    std::string_view rec(seq, len);
    if (molecule == 0)
    {
        molecules::NucleotideDecomposer d(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
        d.decompose_record(rec, size_t{0}, sink);
    }
    else
    {
        molecules::PeptideDecomposer d(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
        d.decompose_record(rec, size_t{0}, sink);
    }
    uint64_t n = sink.values.size();
    for (uint64_t i = 0; i < n && i < cap; ++i)
        out[i] = sink.values[i];
    return n;
}

// update_kmer: returns the canonical value, updates *kmer to the forward value
// (include/nucleotide_decomposer.h:113-121, include/peptide_decomposer.h:293-299)
uint64_t ref_update_kmer(int molecule, int k, int reduction, int symbol, uint64_t * kmer)
{
    if (molecule == 0)
    {
        molecules::NucleotideDecomposer d(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
        return d.update_kmer(symbol, *kmer);
    }
    molecules::PeptideDecomposer d(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
    return d.update_kmer(symbol, *kmer);
}

void ref_peptide_maps(int k, int reduction, uint8_t * aamap256, char * redmap256)
{
    molecules::PeptideDecomposer d(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
    std::memcpy(aamap256, d.aamap_.data(), 256);
    std::memcpy(redmap256, d.redmap_.data(), 256);
}

// IBFIndex::compute_bitcount (include/index_ibf.h:133-139)
uint64_t ref_compute_bitcount(uint64_t n, float fpr)
{
    IBFIndex idx;
    idx.fpr_ = fpr;
    return idx.compute_bitcount(n);
}

// ---------------------------------------------------------------- a4 / a5
void * ref_ibf_new(uint64_t bins, uint64_t bin_size, uint32_t hash_funs)
{
    try
    {
        auto * p = new ref_ibf{seqan::hibf::interleaved_bloom_filter{seqan::hibf::bin_count{bins},
                                                                      seqan::hibf::bin_size{bin_size},
                                                                      seqan::hibf::hash_function_count{hash_funs}}};
        return p;
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return nullptr;
    }
}

void ref_ibf_free(void * h)
{
    delete static_cast<ref_ibf *>(h);
}

// out[0..7) = bins, technical_bins, bin_size, hash_shift, bin_words, hash_funs, n_words
void ref_ibf_params(void * h, uint64_t * out)
{
    auto & ibf = static_cast<ref_ibf *>(h)->ibf;
    fill_params(ibf, out);
}

uint64_t const * ref_ibf_words(void * h)
{
    return static_cast<ref_ibf *>(h)->ibf.data();
}

void ref_ibf_emplace(void * h, uint64_t const * values, uint32_t const * bin_ids, uint64_t n)
{
    auto & ibf = static_cast<ref_ibf *>(h)->ibf;
    for (uint64_t i = 0; i < n; ++i)
        ibf.emplace(values[i], seqan::hibf::bin_index{bin_ids[i]});
}

// One containment agent per thread (hibf's thread-safety rule); out = n × bin_words words.
void ref_ibf_bulk_contains(void * h, uint64_t const * values, uint64_t n, uint64_t * out, int threads)
{
    auto & ibf = static_cast<ref_ibf *>(h)->ibf;
    size_t const bw = (ibf.bin_count() + 63) / 64;
    if (threads < 1)
        threads = 1;
#pragma omp parallel num_threads(threads)
    {
        auto agent = ibf.containment_agent();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < static_cast<int64_t>(n); ++i)
        {
            auto const & res = agent.bulk_contains(values[i]);
            std::memcpy(out + static_cast<size_t>(i) * bw, res.data(), bw * sizeof(uint64_t));
        }
    }
}

// Same loop without storing results (CPU-baseline timing without an n × bin_words buffer);
// returns a checksum so the work cannot be optimised away.
uint64_t ref_ibf_bulk_contains_checksum(void * h, uint64_t const * values, uint64_t n, int threads)
{
    auto & ibf = static_cast<ref_ibf *>(h)->ibf;
    size_t const bw = (ibf.bin_count() + 63) / 64;
    if (threads < 1)
        threads = 1;
    uint64_t total = 0;
#pragma omp parallel num_threads(threads) reduction(+ : total)
    {
        auto agent = ibf.containment_agent();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < static_cast<int64_t>(n); ++i)
        {
            auto const & res = agent.bulk_contains(values[i]);
            uint64_t const * w = res.data();
            for (size_t j = 0; j < bw; ++j)
                total += w[j] * (j + 1);
        }
    }
    return total;
}


// ---------------------------------------------------------------- a8 (HIBF)
// Build through the reference's own call (include/index_hibf.h:110-129): the user-bin
// k-mer buffers go to seqan::hibf::config::input_fn, everything else is hibf defaults.
void * ref_hibf_build(uint64_t const * values, uint64_t const * bin_offsets, uint64_t n_bins, uint32_t hash_funs,
                      float fpr, int threads)
{
    try
    {
        auto get_user_bin_data = [&](size_t const user_bin_id, seqan::hibf::insert_iterator it)
        {
            for (uint64_t i = bin_offsets[user_bin_id]; i < bin_offsets[user_bin_id + 1]; ++i)
                it = values[i];
        };
        seqan::hibf::config config{.input_fn = get_user_bin_data,
                                   .number_of_user_bins = n_bins,
                                   .number_of_hash_functions = hash_funs,
                                   .maximum_fpr = fpr,
                                   .threads = static_cast<size_t>(threads < 1 ? 1 : threads)};
        auto * p = new ref_hibf{seqan::hibf::hierarchical_interleaved_bloom_filter{config}, n_bins};
        return p;
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return nullptr;
    }
}

// The layout alone: compute_sketches → estimate_kmer_counts → compute_layout exactly as the HIBF constructor runs them
// (lib/hibf/src/hierarchical_interleaved_bloom_filter.cpp:212-245).  Dump (uint64 words): tmax, top_level_max_bin_id,
// #max_bins, then per max bin {len, indices..., id}, #user_bins, then per user bin {len, prev..., storage_TB_id,
// number_of_technical_bins, idx}, then n_bins k-mer count estimates.  Returns the number of words (may exceed cap), 0 on error.
uint64_t ref_hibf_layout(uint64_t const * values, uint64_t const * bin_offsets, uint64_t n_bins, uint32_t hash_funs, float fpr,
                         int threads, uint64_t * out, uint64_t cap)
{
    try
    {
        auto get_user_bin_data = [&](size_t const user_bin_id, seqan::hibf::insert_iterator it)
        {
            for (uint64_t i = bin_offsets[user_bin_id]; i < bin_offsets[user_bin_id + 1]; ++i)
                it = values[i];
        };
        seqan::hibf::config config{.input_fn = get_user_bin_data,
                                   .number_of_user_bins = n_bins,
                                   .number_of_hash_functions = hash_funs,
                                   .maximum_fpr = fpr,
                                   .threads = static_cast<size_t>(threads < 1 ? 1 : threads)};
        config.validate_and_set_defaults();
        std::vector<seqan::hibf::sketch::hyperloglog> sketches{};
        std::vector<size_t> kmer_counts{};
        seqan::hibf::sketch::compute_sketches(config, sketches);
        seqan::hibf::sketch::estimate_kmer_counts(sketches, kmer_counts);
        auto layout = seqan::hibf::layout::compute_layout(config, kmer_counts, sketches);
        std::vector<uint64_t> w;
        w.push_back(config.tmax);
        w.push_back(layout.top_level_max_bin_id);
        w.push_back(layout.max_bins.size());
        for (auto const & mb : layout.max_bins)
        {
            w.push_back(mb.previous_TB_indices.size());
            for (size_t x : mb.previous_TB_indices)
                w.push_back(x);
            w.push_back(mb.id);
        }
        w.push_back(layout.user_bins.size());
        for (auto const & ub : layout.user_bins)
        {
            w.push_back(ub.previous_TB_indices.size());
            for (size_t x : ub.previous_TB_indices)
                w.push_back(x);
            w.push_back(ub.storage_TB_id);
            w.push_back(ub.number_of_technical_bins);
            w.push_back(ub.idx);
        }
        for (size_t c : kmer_counts)
            w.push_back(c);
        for (size_t i = 0; i < w.size() && i < cap; ++i)
            out[i] = w[i];
        return w.size();
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return 0;
    }
}

void ref_hibf_free(void * h)
{
    delete static_cast<ref_hibf *>(h);
}

uint64_t ref_hibf_ibf_count(void * h)
{
    return static_cast<ref_hibf *>(h)->hibf.ibf_vector.size();
}

// Geometry of IBF i, same 7 slots as ref_ibf_params.
void ref_hibf_ibf_params(void * h, uint64_t i, uint64_t * out)
{
    auto & ibf = static_cast<ref_hibf *>(h)->hibf.ibf_vector[i];
    fill_params(ibf, out);
}

uint64_t const * ref_hibf_ibf_words(void * h, uint64_t i)
{
    return static_cast<ref_hibf *>(h)->hibf.ibf_vector[i].data();
}

// Copies next_ibf_id[i] and ibf_bin_to_user_bin_id[i]; returns their length (must be equal).
uint64_t ref_hibf_tables(void * h, uint64_t i, uint64_t * next_ibf, uint64_t * user_bin, uint64_t cap)
{
    auto & hibf = static_cast<ref_hibf *>(h)->hibf;
    auto const & nx = hibf.next_ibf_id[i];
    auto const & ub = hibf.ibf_bin_to_user_bin_id[i];
    if (nx.size() != ub.size())
        return UINT64_MAX;
    for (uint64_t j = 0; j < nx.size() && j < cap; ++j)
    {
        next_ibf[j] = nx[j];
        user_bin[j] = ub[j];
    }
    return nx.size();
}

// HIBFIndex::query (include/index_hibf.h:132-147): membership_for({kmer}, 1) → bitvector.
// out = n × ceil(user_bins/64) words.
void ref_hibf_membership(void * h, uint64_t const * values, uint64_t n, uint64_t * out, int threads)
{
    auto * self = static_cast<ref_hibf *>(h);
    size_t const words = (self->user_bins + 63) / 64;
    if (threads < 1)
        threads = 1;
#pragma omp parallel num_threads(threads)
    {
        auto agent = self->hibf.membership_agent();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < static_cast<int64_t>(n); ++i)
        {
            std::vector<uint64_t> one{values[i]};
            auto & res = agent.membership_for(one, 1u);
            uint64_t * dst = out + static_cast<size_t>(i) * words;
            std::memset(dst, 0, words * sizeof(uint64_t));
            for (uint64_t id : res)
                dst[id >> 6] |= uint64_t{1} << (id & 63);
        }
    }
}


} // extern "C" (templates below)

namespace
{

template <class flavor_t, class mol_t>
void harness_build_hibf_index(std::vector<std::string> const & files, uint8_t k, float fpr, uint8_t hc,
                              std::string const & molecule, uint8_t reduction, uint8_t threads,
                              std::string const & out_path)
{
    // src/index_base.cpp:38-70 (create_hibf_*_index) without the debug prints
    TetrexIndex<flavor_t, mol_t> ibf(k, fpr, hc, molecule, files, reduction);
    ibf.populate_index(threads);
    store_ibf(ibf, std::filesystem::path{out_path});
}

template <class mol_t>
void harness_build_flat_index(std::vector<std::string> const & files, uint8_t k, float fpr, uint8_t hc,
                              std::string const & molecule, uint8_t reduction, uint8_t threads,
                              std::string const & out_path)
{
    // src/index_base.cpp:25-35,51-58 (create_ibf_*_index).  The TetrexIndex/IBFIndex ctors are
    // bypassed because IBFIndex's reads bin_size_ before it is initialised;
    // the members are set to exactly what those ctors would have set.
    TetrexIndex<index_structure::IBF, mol_t> ibf;
    ibf.k_ = k;
    ibf.molecule_ = molecule;
    ibf.acid_libs_ = files;
    ibf.reduction_ = reduction;
    ibf.decomposer_ = MoleculeDecomposer<mol_t>(k, reduction);
    ibf.permanent_hibf_status_ = false;
    ibf.ibf_.bin_count_ = files.size();
    ibf.ibf_.fpr_ = fpr;
    ibf.ibf_.hash_count_ = hc;
    ibf.ibf_.tech_bins_ = files;
    ibf.ibf_.user_bin_data_.resize(files.size());
    ibf.ibf_.hits_ = bitvector(files.size(), true);
    ibf.populate_index(threads);
    store_ibf(ibf, std::filesystem::path{out_path});
}

struct fd_capture
{
    int saved{-1};
    explicit fd_capture(char const * path)
    {
        std::cout.flush();
        fflush(stdout);
        saved = dup(1);
        int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
        dup2(fd, 1);
        close(fd);
    }
    ~fd_capture()
    {
        std::cout.flush();
        fflush(stdout);
        dup2(saved, 1);
        close(saved);
    }
};

template <class flavor_t, class mol_t>
int harness_query(std::string const & index_path, std::string regex, bool augment, bool verify, int threads,
                  uint64_t * hit_words, uint64_t cap_words, char const * stdout_path, char const * dibf_path = nullptr)
{
    // src/query.cpp:375-475 (query_*), include/query.h:250-290 (run_collection)
    omp_set_num_threads(threads < 1 ? 1 : threads);
    TetrexIndex<flavor_t, mol_t> ibf;
    load_ibf(ibf, std::filesystem::path{index_path});
    ibf.spawn_agent();
    bitvector hit_vector(ibf.getBinCount(), true);
    DGramIndex dibf{};
    bool has_dibf = false;
    if (dibf_path && *dibf_path) // run_collection, include/query.h:258-264
    {
        load_dindex(dibf, std::filesystem::path{dibf_path});
        has_dibf = true;
    }
    if (ibf.getBinCount() > 1u)
        hit_vector &= process_query(regex, ibf, false, augment, false, has_dibf, dibf);
    size_t const words = (ibf.getBinCount() + 63) / 64;
    for (size_t i = 0; i < words && i < cap_words; ++i)
        hit_words[i] = hit_vector.data()[i];
    if (verify && !hit_vector.none())
    {
        fd_capture cap(stdout_path);
        iter_disk_search(hit_vector, regex, ibf, std::string{"-"});
    }
    return static_cast<int>(words);
}

} // namespace

extern "C"
{

// flavor: 0 = flat IBF, 1 = HIBF; molecule: 0 = "na", 1 = "aa".  Writes <out_path> (full name).
int ref_index_build(char const * const * files, uint64_t n_files, int k, float fpr, int hash_count, int molecule,
                    int flavor, int reduction, int threads, char const * out_path)
{
    try
    {
        std::vector<std::string> f(files, files + n_files);
        std::string mol = molecule == 0 ? "na" : "aa";
        auto K = static_cast<uint8_t>(k);
        auto H = static_cast<uint8_t>(hash_count);
        auto R = static_cast<uint8_t>(reduction);
        auto T = static_cast<uint8_t>(threads);
        if (flavor == 0 && molecule == 0)
            harness_build_flat_index<molecules::nucleotide>(f, K, fpr, H, mol, R, T, out_path);
        else if (flavor == 0)
            harness_build_flat_index<molecules::peptide>(f, K, fpr, H, mol, R, T, out_path);
        else if (molecule == 0)
            harness_build_hibf_index<index_structure::HIBF, molecules::nucleotide>(f, K, fpr, H, mol, R, T, out_path);
        else
            harness_build_hibf_index<index_structure::HIBF, molecules::peptide>(f, K, fpr, H, mol, R, T, out_path);
        return 0;
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// Runs the reference's filter (process_query) and, if `verify`, iter_disk_search with
// everything it prints to stdout captured into stdout_path.  Returns #words or <0.
int ref_index_query(char const * index_path, char const * regex, int augment, int verify, int threads,
                    uint64_t * hit_words, uint64_t cap_words, char const * stdout_path)
{
    try
    {
        index_params params;
        load_params(params, std::filesystem::path{index_path});
        bool const na = params.molecule_ == "na";
        if (!params.is_hibf_ && na)
            return harness_query<index_structure::IBF, molecules::nucleotide>(index_path, regex, augment, verify,
                                                                              threads, hit_words, cap_words, stdout_path);
        if (params.is_hibf_ && na)
            return harness_query<index_structure::HIBF, molecules::nucleotide>(index_path, regex, augment, verify,
                                                                               threads, hit_words, cap_words, stdout_path);
        if (!params.is_hibf_)
            return harness_query<index_structure::IBF, molecules::peptide>(index_path, regex, augment, verify,
                                                                           threads, hit_words, cap_words, stdout_path);
        return harness_query<index_structure::HIBF, molecules::peptide>(index_path, regex, augment, verify, threads,
                                                                        hit_words, cap_words, stdout_path);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// The same with a DGram index (`-g`) next to the index: run_collection's has_dibf path.
int ref_index_query_dibf(char const * index_path, char const * regex, int augment, char const * dibf_path, int verify,
                         int threads, uint64_t * hit_words, uint64_t cap_words, char const * stdout_path)
{
    try
    {
        index_params params;
        load_params(params, std::filesystem::path{index_path});
        bool const na = params.molecule_ == "na";
        if (!params.is_hibf_ && na)
            return harness_query<index_structure::IBF, molecules::nucleotide>(index_path, regex, augment, verify, threads,
                                                                              hit_words, cap_words, stdout_path, dibf_path);
        if (params.is_hibf_ && na)
            return harness_query<index_structure::HIBF, molecules::nucleotide>(index_path, regex, augment, verify, threads,
                                                                               hit_words, cap_words, stdout_path, dibf_path);
        if (!params.is_hibf_)
            return harness_query<index_structure::IBF, molecules::peptide>(index_path, regex, augment, verify, threads,
                                                                           hit_words, cap_words, stdout_path, dibf_path);
        return harness_query<index_structure::HIBF, molecules::peptide>(index_path, regex, augment, verify, threads,
                                                                        hit_words, cap_words, stdout_path, dibf_path);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// drive_dindex's work (src/dGramIndex.cpp:35-37): DGramIndex(min, max, pad, hc, fpr, files).populate() + save_dindex.
// per_bin[b] = dgram_buffer_[b].size().
int ref_dindex_build(char const * const * files, uint64_t n_files, uint64_t min_gap, uint64_t max_gap, int hash_count, float fpr,
                     char const * out_path, uint64_t * per_bin)
{
    try
    {
        std::vector<std::string> f(files, files + n_files);
        DGramIndex dindex(min_gap, max_gap, 1, static_cast<size_t>(hash_count), fpr, f);
        dindex.populate();
        for (uint64_t b = 0; b < n_files && per_bin; ++b)
            per_bin[b] = dindex.dgrams()[b].size();
        save_dindex(dindex, std::filesystem::path{out_path});
        return 0;
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

} // extern "C"

namespace
{

template <class mol_t>
TetrexIndex<index_structure::IBF, mol_t> harness_dummy_flat(int k, int reduction)
{
    TetrexIndex<index_structure::IBF, mol_t> ibf;
    ibf.k_ = static_cast<uint8_t>(k);
    ibf.molecule_ = molecules::is_dna<mol_t> ? "na" : "aa";
    ibf.reduction_ = static_cast<uint8_t>(reduction);
    ibf.decomposer_ = MoleculeDecomposer<mol_t>(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
    ibf.permanent_hibf_status_ = false;
    return ibf;
}

// include/query.h:226-247 up to (not including) collect(): build the k-graph and its ranks.
template <class mol_t>
int harness_kgraph(std::string regex, int k, int reduction, int augment, int32_t * symbol, int32_t * first,
                   int32_t * second, int32_t * rank, uint64_t cap, char * postfix_out, uint64_t postfix_cap)
{
    auto ibf = harness_dummy_flat<mol_t>(k, reduction);
    std::string query;
    preprocess_query(regex, query, ibf);
    if (postfix_out && postfix_cap)
    {
        std::snprintf(postfix_out, postfix_cap, "%s", query.c_str());
    }
    std::unique_ptr<nfa_t> NFA = std::make_unique<nfa_t>();
    std::unique_ptr<lmap_t> nfa_map = std::make_unique<lmap_t>(*NFA);
    amap_t arc_map;
    catsites_t catsites;
    if (ibf.reduction_ == Base)
        catsites = construct_kgraph(query, *NFA, *nfa_map, arc_map, ibf.k_, false);
    else
        catsites = construct_reduced_kgraph(query, *NFA, *nfa_map, arc_map, ibf.k_);
    std::unique_ptr<gmap_t> gap_map = std::make_unique<gmap_t>(*NFA);
    DGramIndex dibf{};
    OTFCollector<index_structure::IBF, mol_t> collector(std::move(NFA), std::move(nfa_map), ibf, std::move(arc_map),
                                                        std::move(gap_map), false, dibf);
    collector.analyze_complexity();
    if (augment && catsites.size() > 0)
        collector.augment(catsites);
    size_t const n = collector.node_count_;
    for (size_t i = 0; i < n && i < cap; ++i)
    {
        node_t node = collector.NFA_->nodeFromId(static_cast<int>(i));
        symbol[i] = (*collector.nfa_map_)[node];
        auto it = collector.arc_map_.find(static_cast<int>(i));
        if (it == collector.arc_map_.end())
        {
            first[i] = -1;
            second[i] = -1;
        }
        else
        {
            first[i] = collector.NFA_->id(it->second.first);
            second[i] = collector.NFA_->id(it->second.second);
        }
        rank[i] = collector.rank_map_[i];
        if (symbol[i] == Gap)
            second[i] = static_cast<int32_t>((*collector.gap_map_)[node]); // gap length rides in slot 2
    }
    return static_cast<int>(n);
}

template <class mol_t>
int harness_collect_flat(ref_ibf * h, int k, int reduction, std::string regex, int augment, uint64_t * out)
{
    auto ibf = harness_dummy_flat<mol_t>(k, reduction);
    ibf.ibf_.bin_count_ = h->ibf.bin_count();
    ibf.ibf_.hash_count_ = static_cast<uint8_t>(h->ibf.hash_function_count());
    ibf.ibf_.ibf_ = h->ibf;
    DGramIndex dibf{};
    bitvector res = process_query(regex, ibf, false, augment, false, false, dibf);
    size_t const words = (ibf.getBinCount() + 63) / 64;
    for (size_t i = 0; i < words; ++i)
        out[i] = res.data()[i];
    return static_cast<int>(words);
}

template <class mol_t>
int harness_collect_hibf(ref_hibf * h, int k, int reduction, std::string regex, int augment, uint64_t * out)
{
    TetrexIndex<index_structure::HIBF, mol_t> ibf;
    ibf.k_ = static_cast<uint8_t>(k);
    ibf.molecule_ = molecules::is_dna<mol_t> ? "na" : "aa";
    ibf.reduction_ = static_cast<uint8_t>(reduction);
    ibf.decomposer_ = MoleculeDecomposer<mol_t>(static_cast<uint8_t>(k), static_cast<uint8_t>(reduction));
    ibf.permanent_hibf_status_ = true;
    ibf.ibf_.bin_count_ = h->user_bins;
    ibf.ibf_.hibf_ = h->hibf;
    DGramIndex dibf{};
    bitvector res = process_query(regex, ibf, false, augment, false, false, dibf);
    size_t const words = (ibf.getBinCount() + 63) / 64;
    for (size_t i = 0; i < words; ++i)
        out[i] = res.data()[i];
    return static_cast<int>(words);
}

} // namespace

extern "C"
{

// Dumps the k-graph the reference builds for `regex` (after its own pre-processing):
// per node id: symbol (byte, or 256 Match / 257 Ghost / 258 Split / 259 Gap), the two
// arc_map successors (-1 = none), and the collector's rank.  Returns node count or <0.
int ref_kgraph(char const * regex, int molecule, int k, int reduction, int augment, int32_t * symbol, int32_t * first,
               int32_t * second, int32_t * rank, uint64_t cap, char * postfix_out, uint64_t postfix_cap)
{
    try
    {
        if (molecule == 0)
            return harness_kgraph<molecules::nucleotide>(regex, k, reduction, augment, symbol, first, second, rank, cap,
                                                         postfix_out, postfix_cap);
        return harness_kgraph<molecules::peptide>(regex, k, reduction, augment, symbol, first, second, rank, cap,
                                                  postfix_out, postfix_cap);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// process_query() against an in-memory flat IBF (ref_ibf_new handle). out: ceil(bins/64) words.
int ref_collect_flat(void * ibf_handle, int molecule, int k, int reduction, char const * regex, int augment,
                     uint64_t * out)
{
    try
    {
        auto * h = static_cast<ref_ibf *>(ibf_handle);
        if (molecule == 0)
            return harness_collect_flat<molecules::nucleotide>(h, k, reduction, regex, augment, out);
        return harness_collect_flat<molecules::peptide>(h, k, reduction, regex, augment, out);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// process_query() against an in-memory HIBF (ref_hibf_build handle).
int ref_collect_hibf(void * hibf_handle, int molecule, int k, int reduction, char const * regex, int augment,
                     uint64_t * out)
{
    try
    {
        auto * h = static_cast<ref_hibf *>(hibf_handle);
        if (molecule == 0)
            return harness_collect_hibf<molecules::nucleotide>(h, k, reduction, regex, augment, out);
        return harness_collect_hibf<molecules::peptide>(h, k, reduction, regex, augment, out);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

// RE2 exactly as iter_disk_search uses it (include/query.h:101-104,146-149; src/query.cpp:194-237):
// "(" + regex + ")", POSIX (longest) for peptides / default for DNA, FindAndConsume loop over one
// record.  Writes (start,end) pairs; returns the number of matches (may exceed cap).
int64_t ref_re2_find_all(char const * regex, int posix_longest, char const * text, uint64_t len, uint64_t * starts,
                         uint64_t * ends, uint64_t cap)
{
    std::string wrapped = std::string("(") + regex + ")";
    std::unique_ptr<re2::RE2> re;
    if (posix_longest)
        re = std::make_unique<re2::RE2>(wrapped, re2::RE2::POSIX);
    else
        re = std::make_unique<re2::RE2>(wrapped);
    if (!re->ok())
    {
        g_err = re->error();
        return -1;
    }
    re2::StringPiece content(text, len);
    re2::StringPiece match;
    char const * base = content.data();
    int64_t n = 0;
    while (RE2::FindAndConsume(&content, *re, &match))
    {
        if (match.size() == 0)
        {
            g_err = "pattern matches the empty string (reference would loop forever)";
            return -2;
        }
        if (static_cast<uint64_t>(n) < cap)
        {
            starts[n] = static_cast<uint64_t>(match.data() - base);
            ends[n] = starts[n] + match.size();
        }
        ++n;
    }
    return n;
}

} // extern "C"

extern "C"
{

// Wraps caller-provided HIBF tables (e.g. a synthetic HIBF built on the GPU) in the reference's own
// seqan::hibf::hierarchical_interleaved_bloom_filter so that its membership_agent can be timed /
// compared on exactly the same structure.  All members touched are public in hibf
// (hierarchical_interleaved_bloom_filter.hpp:209-238).
void * ref_hibf_from_tables(uint64_t n_ibfs, uint64_t n_user_bins, uint64_t const * bins, uint64_t const * bin_size,
                            uint64_t const * hash_funs, uint64_t const * const * words,
                            uint64_t const * const * next_ibf, uint64_t const * const * user_bin,
                            uint64_t const * table_len)
{
    try
    {
        auto * p = new ref_hibf{};
        p->user_bins = n_user_bins;
        auto & h = p->hibf;
        for (uint64_t i = 0; i < n_ibfs; ++i)
        {
            h.ibf_vector.emplace_back(seqan::hibf::bin_count{bins[i]}, seqan::hibf::bin_size{bin_size[i]},
                                      seqan::hibf::hash_function_count{hash_funs[i]});
            auto & ibf = h.ibf_vector.back();
            std::memcpy(const_cast<uint64_t *>(ibf.data()), words[i], ((ibf.bit_size() + 63) / 64) * sizeof(uint64_t));
            h.next_ibf_id.emplace_back(next_ibf[i], next_ibf[i] + table_len[i]);
            h.ibf_bin_to_user_bin_id.emplace_back(user_bin[i], user_bin[i] + table_len[i]);
            h.prev_ibf_id.push_back({0, 0});
        }
        return p;
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return nullptr;
    }
}

// The verify inner loop of src/query.cpp:194-237 over records held in memory (no gz / kseq), OpenMP over
// records like the reference's `omp parallel for` over bins.  Returns the total number of matches.
int64_t ref_re2_scan_records(char const * regex, int posix_longest, char const * text, uint64_t const * rec_offsets,
                             uint64_t n_records, int threads)
{
    std::string wrapped = std::string("(") + regex + ")";
    std::unique_ptr<re2::RE2> re;
    if (posix_longest)
        re = std::make_unique<re2::RE2>(wrapped, re2::RE2::POSIX);
    else
        re = std::make_unique<re2::RE2>(wrapped);
    if (!re->ok())
    {
        g_err = re->error();
        return -1;
    }
    int64_t total = 0;
    if (threads < 1)
        threads = 1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads) reduction(+ : total)
    for (int64_t r = 0; r < static_cast<int64_t>(n_records); ++r)
    {
        re2::StringPiece content(text + rec_offsets[r], rec_offsets[r + 1] - rec_offsets[r]);
        re2::StringPiece match;
        while (RE2::FindAndConsume(&content, *re, &match))
        {
            ++total;
            if (match.size() == 0)
                break;
        }
    }
    return total;
}

} // extern "C"

extern "C"
{

// The order in which OTFCollector::collect drains one rank (include/otf_collector.h:355-357: `comp_table_[i].begin()`
// + scrub): keys are inserted like push() does (find, then operator[]), then begin() is erased until the map is empty.
// out[j] = index (into keys) of the j-th popped entry; duplicates of a key are absorbed (not inserted).  Returns the
// number of entries popped.  Checks the host layer's restatement of the container's slot placement (rh_order.hpp).
uint64_t ref_rh_pop_order(uint64_t const * keys, uint64_t n, uint32_t * out)
{
    robin_hood::unordered_map<uint64_t, uint32_t> m;
    for (uint64_t i = 0; i < n; ++i)
        if (m.find(keys[i]) == m.end())
            m[keys[i]] = static_cast<uint32_t>(i);
    uint64_t j = 0;
    while (!m.empty())
    {
        out[j++] = m.begin()->second;
        m.erase(m.begin());
    }
    return j;
}

// The k-mer sets of seqan::hibf's build (robin_hood::unordered_flat_set<uint64_t>, lib/hibf/include/hibf/contrib/
// robin_hood.hpp) under a script of operations: op[i] = 0 → insert(keys[i]), 1 → clear().  Writes the final iteration
// order (which decides how insert_into_ibf chunks a split user bin) and returns the number of elements.
uint64_t ref_flat_set_script(uint64_t const * keys, uint8_t const * op, uint64_t n, uint64_t * out)
{
    robin_hood::unordered_flat_set<uint64_t> set;
    for (uint64_t i = 0; i < n; ++i)
    {
        if (op[i] == 1)
            set.clear();
        else
            set.insert(keys[i]);
    }
    uint64_t j = 0;
    for (uint64_t v : set)
        out[j++] = v;
    return j;
}

} // extern "C"

extern "C"
{

// ref_re2_scan_records with the matches returned: (record, start, end) triples in (record, start) order — the order
// iter_disk_search prints them with -t 1.  Returns the total number of matches (may exceed cap; then only the first
// cap triples are written).
int64_t ref_re2_scan_records_hits(char const * regex, int posix_longest, char const * text, uint64_t const * rec_offsets,
                                  uint64_t n_records, int threads, uint32_t * out_triples, uint64_t cap)
{
    std::string wrapped = std::string("(") + regex + ")";
    std::unique_ptr<re2::RE2> re;
    if (posix_longest)
        re = std::make_unique<re2::RE2>(wrapped, re2::RE2::POSIX);
    else
        re = std::make_unique<re2::RE2>(wrapped);
    if (!re->ok())
    {
        g_err = re->error();
        return -1;
    }
    if (threads < 1)
        threads = 1;
    std::vector<std::vector<uint32_t>> per(n_records);
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
    for (int64_t r = 0; r < static_cast<int64_t>(n_records); ++r)
    {
        char const * base = text + rec_offsets[r];
        re2::StringPiece content(base, rec_offsets[r + 1] - rec_offsets[r]);
        re2::StringPiece match;
        while (RE2::FindAndConsume(&content, *re, &match))
        {
            if (match.size() == 0)
                break;
            per[r].push_back(static_cast<uint32_t>(match.data() - base));
            per[r].push_back(static_cast<uint32_t>(match.data() - base + match.size()));
        }
    }
    int64_t total = 0;
    for (uint64_t r = 0; r < n_records; ++r)
        for (size_t i = 0; i + 1 < per[r].size(); i += 2)
        {
            if (static_cast<uint64_t>(total) < cap)
            {
                out_triples[3 * total] = static_cast<uint32_t>(r);
                out_triples[3 * total + 1] = per[r][i];
                out_triples[3 * total + 2] = per[r][i + 1];
            }
            ++total;
        }
    return total;
}

} // extern "C"

namespace
{
template <class flavor_t, class mol_t>
int harness_conjunction(std::string const & index_path, std::string const & regex_list, int threads, uint64_t * hit_words,
                        uint64_t cap_words, char const * stdout_path)
{
    // run_conjunction (include/query.h:293-326) without the timing / verbose prints: AND of process_query over the ':'-
    // separated queries (src/query.cpp:463-466), then iter_disk_search_set (include/query.h:191-224)
    omp_set_num_threads(threads < 1 ? 1 : threads);
    TetrexIndex<flavor_t, mol_t> ibf;
    load_ibf(ibf, std::filesystem::path{index_path});
    ibf.spawn_agent();
    std::vector<std::string> queries = split_string(regex_list, ':');
    bitvector hit_vector(ibf.getBinCount(), true);
    DGramIndex dibf{};
    if (ibf.getBinCount() > 1u)
        for (auto rx : queries)
            hit_vector &= process_query(rx, ibf, false, false, false, false, dibf);
    size_t const words = (ibf.getBinCount() + 63) / 64;
    for (size_t i = 0; i < words && i < cap_words; ++i)
        hit_words[i] = hit_vector.data()[i];
    if (!hit_vector.none())
    {
        fd_capture cap(stdout_path);
        iter_disk_search_set(hit_vector, queries, ibf);
    }
    return static_cast<int>(words);
}
} // namespace

extern "C"
{

// `tetrex query -c idx "q1:q2:..."` as the reference runs it (peptide HIBF is the only flavour its CLI wires up; the
// templates work for all four).  Returns #words or <0; stdout of the verify step goes to stdout_path.
int ref_index_conjunction(char const * index_path, char const * regex_list, int threads, uint64_t * hit_words,
                          uint64_t cap_words, char const * stdout_path)
{
    try
    {
        index_params params;
        load_params(params, std::filesystem::path{index_path});
        bool const na = params.molecule_ == "na";
        if (!params.is_hibf_ && na)
            return harness_conjunction<index_structure::IBF, molecules::nucleotide>(index_path, regex_list, threads, hit_words,
                                                                                    cap_words, stdout_path);
        if (params.is_hibf_ && na)
            return harness_conjunction<index_structure::HIBF, molecules::nucleotide>(index_path, regex_list, threads, hit_words,
                                                                                     cap_words, stdout_path);
        if (!params.is_hibf_)
            return harness_conjunction<index_structure::IBF, molecules::peptide>(index_path, regex_list, threads, hit_words,
                                                                                 cap_words, stdout_path);
        return harness_conjunction<index_structure::HIBF, molecules::peptide>(index_path, regex_list, threads, hit_words,
                                                                              cap_words, stdout_path);
    }
    catch (std::exception const & e)
    {
        g_err = e.what();
        return -1;
    }
}

} // extern "C"
