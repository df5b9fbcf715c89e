// hibf_build.cpp — seqan::hibf's hierarchical build (lib/hibf/src/hierarchical_interleaved_bloom_filter.cpp:41-210,
// lib/hibf/src/build/construct_ibf.cpp, insert_into_ibf.cpp, compute_kmers.cpp, update_parent_kmers.hpp,
// update_user_bins.hpp) on top of the layout of hibf_layout.cpp, with K1 (trx_emplace_values) as the emplace step.
//
// What has to be reproduced for the index to answer like the reference's — false positives included — is WHICH k-mers go
// into WHICH technical bin of WHICH size:
//  * sizes come from the number of DISTINCT k-mers of each IBF's fullest bin (a user bin, or the union below a merged
//    bin) through bin_size_in_bits and the split-bin correction;
//  * a user bin that is split over several technical bins is cut into equal chunks along the ITERATION ORDER of the
//    robin_hood::unordered_flat_set the reference collects it in (insert_into_ibf.cpp:28-45).  That order depends on the
//    table's whole history — the same `kmers` object is cleared and reused for every record of an IBF and is also what a
//    favourite child merges its k-mers into — so the build below runs the reference's sequence of set operations on a
//    restatement of the container (rh_order.hpp, checked against the real one by tests/test_hibf_layout_cpu.py).
// What is NOT reproduced is the numbering of the IBFs: the reference hands the root's children to OpenMP threads in an
// order shuffled with std::random_device (…cpp:118-121), so its own numbering differs from run to run; membership
// answers do not depend on it.  Here the children are built in layout order.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "hibf_layout.hpp"
#include "rh_order.hpp"
#include "trx_host.hpp"

namespace trxh
{

namespace
{

using hibf::graph_node;
using hibf::layout;

// robin_hood::unordered_flat_set<uint64_t> as far as the build uses it
class kmer_set
{
public:
    void insert(uint64_t v) { slots_.insert_key(v); }
    void insert_all(kmer_set const & other) // insert(other.begin(), other.end()): in other's iteration order
    {
        other.slots_.for_each_key([&](uint64_t v) { slots_.insert_key(v); });
    }
    void clear() { slots_.clear_keep_capacity(); }
    size_t size() const { return slots_.size(); }
    template <class F>
    void for_each(F && f) const
    {
        slots_.for_each_key(f);
    }

private:
    rh_slots slots_;
};

struct builder
{
    trx_ctx_t * ctx;
    hibf::config const & cfg;
    std::vector<uint64_t> const & values;  // k-mers of all user bins, in HIBFIndex::emplace order
    std::vector<uint64_t> const & offsets; // user bin b = values[offsets[b], offsets[b + 1])
    std::vector<double> const & fpr_correction;
    hibf_image & out;
    // position of every node's IBF, fixed before anything is built: the order in which ONE thread would have reached the
    // nodes (a node, its favourite child's subtree, then its other children in layout order)
    std::unordered_map<graph_node const *, size_t> const & ibf_pos_of;
    std::vector<uint64_t> stage_values;
    std::vector<uint32_t> stage_bins;

    // compute_kmers (build/compute_kmers.cpp:21-31): input_fn inserts every value of the user bin
    void compute_kmers(kmer_set & kmers, layout::user_bin const & record) const
    {
        for (uint64_t i = offsets[record.idx]; i < offsets[record.idx + 1]; ++i)
            kmers.insert(values[i]);
    }

    // insert_into_ibf (build/insert_into_ibf.cpp:26-45): chunks of ceil(size / number_of_bins) along the iteration order
    void insert_into_ibf(kmer_set const & kmers, size_t number_of_bins, size_t bin_index, trx_ibf_t * ibf)
    {
        if (kmers.size() == 0)
            return;
        size_t const chunk_size = (kmers.size() + number_of_bins - 1) / number_of_bins;
        stage_values.clear();
        stage_bins.clear();
        stage_values.reserve(kmers.size());
        stage_bins.reserve(kmers.size());
        size_t i = 0;
        kmers.for_each([&](uint64_t v) {
            stage_values.push_back(v);
            stage_bins.push_back(static_cast<uint32_t>(bin_index + i / chunk_size));
            ++i;
        });
        check(trx_emplace_values(ibf, stage_values.data(), stage_bins.data(), stage_values.size()));
    }

    // insert_into_ibf(data, record, ibf) (:69-83): a root-level single-bin record goes in unfiltered
    void insert_record_direct(layout::user_bin const & record, trx_ibf_t * ibf)
    {
        uint64_t const n = offsets[record.idx + 1] - offsets[record.idx];
        if (n == 0)
            return;
        stage_bins.assign(n, static_cast<uint32_t>(record.storage_TB_id));
        check(trx_emplace_values(ibf, values.data() + offsets[record.idx], stage_bins.data(), n));
    }

    static void update_user_bins(std::vector<uint64_t> & tb_to_ub, layout::user_bin const & record)
    {
        std::fill_n(tb_to_ub.begin() + record.storage_TB_id, record.number_of_technical_bins, record.idx);
    }

    // hierarchical_build (…cpp:41-172).  Returns the position of the node's IBF; the node's k-mers are merged into
    // parent_kmers unless it is the root.
    size_t build_node(kmer_set & parent_kmers, graph_node const & node, bool is_root, size_t parent_ibf_idx)
    {
        size_t const ibf_pos = ibf_pos_of.at(&node);
        out.next_ibf_id[ibf_pos].assign(node.number_of_technical_bins, ibf_pos);
        out.ibf_bin_to_user_bin_id[ibf_pos].assign(node.number_of_technical_bins, UINT64_MAX); // bin_kind::merged
        kmer_set kmers;

        size_t max_bin_tbs;
        if (node.max_bin_is_merged())
        {
            // the favourite child first: its k-mers arrive in `kmers`
            size_t const pos = build_node(kmers, node.children[*node.favourite_child_idx], false, ibf_pos);
            out.next_ibf_id[ibf_pos][node.max_bin_index] = pos;
            max_bin_tbs = 1;
        }
        else
        {
            layout::user_bin const & record = node.remaining_records[0]; // the fullest bin's record leads the list
            compute_kmers(kmers, record);
            update_user_bins(out.ibf_bin_to_user_bin_id[ibf_pos], record);
            max_bin_tbs = record.number_of_technical_bins;
        }

        // construct_ibf (build/construct_ibf.cpp:27-64)
        size_t const kmers_per_bin = (kmers.size() + max_bin_tbs - 1) / max_bin_tbs;
        double const fpr = node.max_bin_is_merged() ? cfg.relaxed_fpr : cfg.maximum_fpr;
        size_t const bin_bits = hibf::bin_size_in_bits(fpr, cfg.number_of_hash_functions, kmers_per_bin);
        size_t const bin_size = max_bin_tbs == 1u ? bin_bits : static_cast<size_t>(std::ceil(bin_bits * fpr_correction[max_bin_tbs]));
        if (bin_size == 0)
            throw error("The size of a bin must be > 0."); // seqan::hibf::interleaved_bloom_filter's constructor
        trx_ibf_t * ibf = nullptr;
        check(trx_ibf_create(ctx, node.number_of_technical_bins, bin_size, static_cast<uint32_t>(cfg.number_of_hash_functions), 0, 1,
                             &ibf));
        try
        {
            insert_into_ibf(kmers, max_bin_tbs, node.max_bin_index, ibf);
            if (!is_root)
                parent_kmers.insert_all(kmers);
            kmers.clear();

            // the other merged bins: one lower-level IBF each
            std::vector<size_t> todo;
            for (size_t c = 0; c < node.children.size(); ++c)
                if (!(node.max_bin_is_merged() && c == *node.favourite_child_idx))
                    todo.push_back(c);
            unsigned const workers = is_root ? static_cast<unsigned>(std::min<size_t>({cfg.threads, todo.size(), size_t{32}})) : 1u;
            if (workers <= 1)
            {
                for (size_t c : todo)
                {
                    graph_node const & child = node.children[c];
                    kmer_set local_kmers;
                    size_t const pos = build_node(local_kmers, child, false, ibf_pos);
                    out.prev_ibf_id[pos] = {parent_ibf_idx, child.parent_bin_index}; // as written: the caller's parent (…cpp:135)
                    out.next_ibf_id[ibf_pos][child.parent_bin_index] = pos;
                    insert_into_ibf(local_kmers, 1, child.parent_bin_index, ibf);
                    if (!is_root)
                        parent_kmers.insert_all(local_kmers);
                }
            }
            else
            {
                // The root's children do not share anything but the root's matrix (…cpp:109-142 hands them to OpenMP threads
                // for the same reason): every worker builds whole subtrees with a context of its own on the same device and
                // takes a lock only to emplace a finished child's k-mers into the root's bin.  Nothing a worker computes
                // depends on the others, so the index is the same as with one thread.
                std::atomic<size_t> next{0};
                std::mutex root_mutex;
                std::vector<std::string> errors(workers);
                std::vector<std::thread> pool;
                int const device = trx_ctx_device(ctx);
                for (unsigned w = 0; w < workers; ++w)
                    pool.emplace_back([&, w] {
                        trx_ctx_t * own = nullptr;
                        try
                        {
                            check(trx_ctx_create(device, nullptr, &own));
                            builder local{own, cfg, values, offsets, fpr_correction, out, ibf_pos_of};
                            for (size_t i = next.fetch_add(1); i < todo.size(); i = next.fetch_add(1))
                            {
                                graph_node const & child = node.children[todo[i]];
                                kmer_set local_kmers;
                                size_t const pos = local.build_node(local_kmers, child, false, ibf_pos);
                                out.prev_ibf_id[pos] = {parent_ibf_idx, child.parent_bin_index};
                                std::lock_guard<std::mutex> guard(root_mutex);
                                out.next_ibf_id[ibf_pos][child.parent_bin_index] = pos;
                                local.insert_into_ibf(local_kmers, 1, child.parent_bin_index, ibf);
                            }
                        }
                        catch (std::exception const & e)
                        {
                            errors[w] = e.what();
                            next.store(todo.size());
                        }
                        if (own)
                            trx_ctx_destroy(own);
                    });
                for (std::thread & t : pool)
                    t.join();
                for (std::string const & e : errors)
                    if (!e.empty())
                        throw error(e);
            }

            // the user bins that live in this IBF
            for (size_t i = node.max_bin_is_merged() ? 0u : 1u; i < node.remaining_records.size(); ++i)
            {
                layout::user_bin const & record = node.remaining_records[i];
                if (is_root && record.number_of_technical_bins == 1)
                    insert_record_direct(record, ibf);
                else
                {
                    compute_kmers(kmers, record);
                    insert_into_ibf(kmers, record.number_of_technical_bins, record.storage_TB_id, ibf);
                    if (!is_root)
                        parent_kmers.insert_all(kmers);
                }
                update_user_bins(out.ibf_bin_to_user_bin_id[ibf_pos], record);
                kmers.clear();
            }

            // the finished matrix comes back for the index file
            uint64_t geo[10];
            check(trx_ibf_geometry(ibf, geo));
            ibf_image & f = out.ibf_vector[ibf_pos];
            f.bins = geo[0];
            f.technical_bins = geo[1];
            f.bin_size = geo[2];
            f.hash_shift = geo[3];
            f.bin_words = geo[4];
            f.hash_funs = geo[5];
            f.words.resize(f.bin_words * f.bin_size);
            check(trx_ibf_download(ibf, f.words.data(), f.words.size()));
            f.occupancy.assign(f.technical_bins, 0);
            f.track_occupancy = false;
        }
        catch (...)
        {
            trx_ibf_destroy(ibf);
            throw;
        }
        trx_ibf_destroy(ibf);
        return ibf_pos;
    }
};

} // namespace

// hierarchical_interleaved_bloom_filter{config} (…cpp:212-245) for user bins given as k-mer lists
hibf_image build_hibf_like_seqan(trx_ctx_t * ctx, std::vector<uint64_t> const & values, std::vector<uint64_t> const & offsets,
                                 size_t hash_count, float fpr, size_t threads, std::vector<uint8_t> const * sketch_registers)
{
    hibf::config cfg;
    cfg.number_of_user_bins = offsets.size() - 1;
    cfg.number_of_hash_functions = hash_count;
    cfg.maximum_fpr = fpr; // float → double like `.maximum_fpr = fpr_` (include/index_hibf.h:127)
    // -t decides how many host threads build the root's children side by side; the index does not depend on it (unlike the
    // reference's, whose clustering breaks ties by thread id), so the default is every core
    cfg.threads = threads > 1 ? threads : std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    cfg.validate_and_set_defaults();

    std::vector<hibf::hyperloglog> sketches(cfg.number_of_user_bins, hibf::hyperloglog{cfg.sketch_bits});
    std::vector<size_t> kmer_counts(cfg.number_of_user_bins);
    size_t const sketch_size = size_t{1} << cfg.sketch_bits;
    if (sketch_registers && sketch_registers->size() != cfg.number_of_user_bins * sketch_size)
        throw error("sketch registers: expected " + std::to_string(cfg.number_of_user_bins * sketch_size) + " bytes");
    for (size_t b = 0; b < cfg.number_of_user_bins; ++b)
    {
        if (sketch_registers) // sketched on the device (trx_hll_sketch_records_dev)
            std::copy(sketch_registers->begin() + b * sketch_size, sketch_registers->begin() + (b + 1) * sketch_size,
                      sketches[b].registers().begin());
        else
            for (uint64_t i = offsets[b]; i < offsets[b + 1]; ++i)
                sketches[b].add(values[i]);
        kmer_counts[b] = sketches[b].estimate();
        if (kmer_counts[b] == 0)
            throw error("user bin " + std::to_string(b) + " holds no k-mer: the HIBF layout needs at least one per bin");
    }
    bool const trace = getenv("TRX_BUILD_TRACE") != nullptr; // phase times on stderr
    auto const t0 = std::chrono::steady_clock::now();
    auto const since = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    hibf::layout const l = hibf::compute_layout(cfg, kmer_counts, sketches);
    hibf::graph_node const root = hibf::make_graph(l);
    if (trace)
        fprintf(stderr, "hibf build: layout of %zu user bins (%zu lower-level IBFs) %.2f s\n", cfg.number_of_user_bins, l.max_bins.size(), since());

    hibf_image img;
    img.number_of_user_bins = cfg.number_of_user_bins;
    size_t const number_of_ibfs = l.max_bins.size() + 1;
    img.ibf_vector.resize(number_of_ibfs);
    img.ibf_bin_to_user_bin_id.resize(number_of_ibfs);
    img.prev_ibf_id.resize(number_of_ibfs, {0, 0});
    img.next_ibf_id.resize(number_of_ibfs);
    std::vector<double> const fpr_correction =
        hibf::compute_fpr_correction(cfg.maximum_fpr, cfg.number_of_hash_functions, root.number_of_technical_bins);
    std::unordered_map<graph_node const *, size_t> ibf_pos_of;
    {
        size_t next = 0;
        std::function<void(graph_node const &)> number = [&](graph_node const & n) {
            ibf_pos_of[&n] = next++;
            if (n.max_bin_is_merged())
                number(n.children[*n.favourite_child_idx]);
            for (size_t c = 0; c < n.children.size(); ++c)
                if (!(n.max_bin_is_merged() && c == *n.favourite_child_idx))
                    number(n.children[c]);
        };
        number(root);
    }
    builder b{ctx, cfg, values, offsets, fpr_correction, img, ibf_pos_of};
    kmer_set root_kmers;
    b.build_node(root_kmers, root, true, 0);
    if (trace)
        fprintf(stderr, "hibf build: hierarchical build (set replay + emplace) done at %.2f s\n", since());
    return img;
}

} // namespace trxh
