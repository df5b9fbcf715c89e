// collector.cpp — OTFCollector::collect (include/otf_collector.h:341-393) as a level-synchronous batch producer.
//
// The reference walks the k-graph in topological rank order; all items of one rank sit on the same node, are
// independent of each other and only write to higher ranks.  That is the batch boundary used here: per rank the
// host decides WHAT happens (roll the k-mer, probe, push, absorb, split, match — identical rules: items are keyed
// by (rank, kmer & submask), the first arrival's kmer/shift_count win, later arrivals only OR their path in) and
// the device does it: one trx_paths_step launch probes every k-mer of the rank, ANDs the hit vectors into the
// path slots and reports which paths died; one trx_paths_or launch performs every push-copy, absorb and match OR.
// Path bitvectors never leave HBM; only the 1-byte alive flags travel.
//
// Order inside a rank: the reference drains a rank through `comp_table_[i].begin()` of a robin_hood map, i.e. in the
// SLOT order of that table, and the order is observable — when two items of one rank collide on a later key
// (e.g. `(XAC|C)DEF`, k = 4: "XACD" and "CD" both key as "ACD" because A packs to 0) the first one pushed keeps its
// kmer_/shift_count_ (absorb, otf_collector.h:204-208) and decides whether the next symbol is probed.  rh_order.hpp
// restates the container's slot placement, every rank is processed in exactly that order, and
// tests/test_collector_vs_ref.py fuzzes the resulting candidate sets against the compiled reference's process_query.
#include <algorithm>
#include <memory>
#include <cstdio>
#include <chrono>
#include <thread>
#include <atomic>
#include <cstdlib>
#include <string>
#include <unordered_map>

#include "rh_order.hpp"
#include "trx_host.hpp"

namespace trxh
{

namespace
{

struct item
{
    uint8_t shift_count;
    uint64_t kmer;
    uint32_t slot;
    // `-a` (CollectorsItem::gapped_ / res_cache_1 / res_cache_2, otf_collector.h:26-28): the item has just jumped a Gap
    // node; kmer holds the gapped code under construction and the next three residues complete it
    bool gapped{false};
    int res1{0}, res2{0};
};

// DGramTools::aa_to_num (include/dGramIndex.h:22-59): note X = 20 and the B/J/O/U/Z aliases — not the index side's
// 20-letter map; anything that is not an upper-case letter is 0
uint64_t dgram_aa_to_num(int symbol)
{
    char const aa = static_cast<char>(symbol);
    if (aa < 'A' || aa > 'Z')
        return 0;
    switch (aa)
    {
    case 'A': return 0;
    case 'C': return 1;
    case 'D': case 'B': return 2;
    case 'E': case 'Z': return 3;
    case 'F': return 4;
    case 'G': return 5;
    case 'H': return 6;
    case 'I': return 7;
    case 'K': return 8;
    case 'L': case 'J': return 9;
    case 'M': return 10;
    case 'N': return 11;
    case 'P': return 12;
    case 'Q': return 13;
    case 'R': return 14;
    case 'S': return 15;
    case 'T': return 16;
    case 'V': return 17;
    case 'W': return 18;
    case 'Y': return 19;
    case 'X': case 'O': case 'U': return 20;
    default: return 0;
    }
}

struct bucket
{
    std::vector<item> items; // in arrival order; drained in the slot order of by_key
    rh_slots by_key;         // kmer & submask → index into items, placed like comp_table_[rank] places it
    // the items in the order OTFCollector::collect pops them (include/otf_collector.h:355-357)
    void sort_into_pop_order()
    {
        std::vector<item> ordered;
        ordered.reserve(items.size());
        by_key.for_each([&](uint32_t i) { ordered.push_back(items[i]); });
        items.swap(ordered);
    }
};

// view of the index's persistent slot pool for one query
struct slot_pool
{
    DeviceIndex::collector_scratch & sc;
    explicit slot_pool(DeviceIndex::collector_scratch & s) : sc(s) {}
    trx_paths_t * paths() const { return sc.paths; }
    uint32_t fresh()
    {
        if (sc.used >= sc.capacity)
            throw error("regular expression too complex: more than " + std::to_string(sc.capacity) +
                        " live paths (try a more specific query)");
        return static_cast<uint32_t>(sc.used++);
    }
};

// (re)creates the pool for slots of W words and hands it out all-zero
void prepare_pool(DeviceIndex & index, uint64_t W)
{
    DeviceIndex::collector_scratch & sc = index.scratch();
    uint64_t const capacity = std::max<uint64_t>(1ull << 12, std::min<uint64_t>(1ull << 20, (1ull << 28) / (W * 8)));
    if (!sc.paths || sc.words != W || sc.capacity != capacity)
    {
        if (sc.paths)
            trx_paths_destroy(sc.paths);
        sc.paths = nullptr;
        check(trx_paths_create(index.ctx(), W, capacity, &sc.paths)); // zeroed
        sc.words = W;
        sc.capacity = capacity;
    }
    else if (sc.used)
        check(trx_dev_memset(index.ctx(), trx_paths_device_words(sc.paths), 0, sc.used * W * 8));
    sc.used = 0;
}

void reserve_dev(trx_ctx_t * ctx, void *& ptr, uint64_t & have, uint64_t want)
{
    if (want <= have)
        return;
    if (ptr)
        check(trx_dev_free(ctx, ptr));
    ptr = nullptr;
    have = 0;
    check(trx_dev_alloc(ctx, want + want / 2, &ptr));
    have = want + want / 2;
}

} // namespace

// HBM-resident path slots (a6): one launch per rank probes and ANDs, one performs the ORs; only alive bytes travel
struct device_paths
{
    DeviceIndex & index;
    shard_hooks const * hooks;
    trx_comm_t * comm;
    bool sharded;
    uint64_t W;
    slot_pool pool;
    device_paths(DeviceIndex & ix, shard_hooks const * h, trx_comm_t * c, bool sh, uint64_t words)
        : index(ix), hooks(h), comm(c), sharded(sh), W(words), pool(ix.scratch())
    {}
    uint32_t fresh() { return pool.fresh(); }
    void fill(uint32_t slot, uint64_t bins) { check(trx_paths_fill(pool.paths(), slot, 1, bins)); }
    void step(std::vector<uint64_t> const & kmers, std::vector<uint32_t> const & slots, std::vector<uint8_t> & res)
    {
        DeviceIndex::collector_scratch & sc = index.scratch();
        if (comm) // probe + all-reduce(MAX) of the alive bytes on the device + one read-back
            check(trx_paths_step_sharded(index.flat(), pool.paths(), comm, kmers.data(), slots.data(), kmers.size(), res.data()));
        else if (!index.is_hibf())
            check(trx_paths_step(index.flat(), pool.paths(), kmers.data(), slots.data(), kmers.size(), res.data()));
        else
        {
            uint64_t const n = kmers.size();
            reserve_dev(index.ctx(), sc.d_kmers, sc.kmers_bytes, n * 8);
            reserve_dev(index.ctx(), sc.d_hits, sc.hits_bytes, n * W * 8);
            check(trx_copy_h2d(index.ctx(), sc.d_kmers, kmers.data(), n * 8));
            check(trx_hibf_membership_dev(index.hibf(), static_cast<uint64_t *>(sc.d_kmers), n,
                                          static_cast<uint64_t *>(sc.d_hits), nullptr));
            std::vector<uint32_t> idx(n);
            for (uint64_t i = 0; i < n; ++i)
                idx[i] = static_cast<uint32_t>(i);
            check(trx_paths_and_hits(pool.paths(), static_cast<uint64_t *>(sc.d_hits), idx.data(), slots.data(), n, res.data()));
        }
        if (sharded && !comm)
            hooks->any_alive(res.data(), res.size());
    }
    void step_dgram(std::vector<uint64_t> const & codes, std::vector<uint32_t> const & slots, std::vector<uint8_t> & res)
    {
        check(trx_paths_step(index.dgram()->flat(), pool.paths(), codes.data(), slots.data(), codes.size(), res.data()));
    }
    void or_pairs(std::vector<uint32_t> const & dst, std::vector<uint32_t> const & src)
    {
        check(trx_paths_or(pool.paths(), dst.data(), src.data(), dst.size()));
    }
};

// Dry run: nothing dies, so the walk asks for every k-mer ANY pruning of it could ask for (an item's k-mer does not depend
// on which paths are alive).  Over budget = the graph is too wide to probe speculatively.
struct budget_exceeded
{};
struct recording_paths
{
    size_t max_probes{1024}, max_slots{1u << 15}; // one query: beyond this the rank-by-rank device walk is no slower
    recording_paths() = default;
    recording_paths(size_t probes, size_t slots) : max_probes(probes), max_slots(slots) {}
    std::vector<uint64_t> kmers, codes; // distinct, in first-use order
    std::unordered_map<uint64_t, uint32_t> kmer_row, code_row;
    uint32_t slots{0};
    uint32_t fresh()
    {
        if (slots >= max_slots)
            throw budget_exceeded{};
        return slots++;
    }
    void fill(uint32_t, uint64_t) {}
    void note(std::vector<uint64_t> const & in, std::vector<uint64_t> & list, std::unordered_map<uint64_t, uint32_t> & row) const
    {
        for (uint64_t v : in)
            if (row.emplace(v, static_cast<uint32_t>(list.size())).second)
            {
                if (list.size() >= max_probes)
                    throw budget_exceeded{};
                list.push_back(v);
            }
    }
    void step(std::vector<uint64_t> const & in, std::vector<uint32_t> const &, std::vector<uint8_t> & res)
    {
        note(in, kmers, kmer_row);
        std::fill(res.begin(), res.end(), uint8_t{1});
    }
    void step_dgram(std::vector<uint64_t> const & in, std::vector<uint32_t> const &, std::vector<uint8_t> & res)
    {
        note(in, codes, code_row);
        std::fill(res.begin(), res.end(), uint8_t{1});
    }
    void or_pairs(std::vector<uint32_t> const &, std::vector<uint32_t> const &) {}
};

// Path bitvectors in host words over membership rows that are already here: the reference's own arithmetic
// (`path_ &= ibf_->query(kmer)`, `path_ |= other.path_`, `path_.none()`), no device round trip per rank
struct host_paths
{
    using row_map = std::unordered_map<uint64_t, uint32_t>;
    uint64_t W;
    row_map const & kmer_row; // k-mer → row of kmer_rows (W words each)
    row_map const & code_row;
    uint64_t const * kmer_rows;
    uint64_t const * code_rows;
    std::vector<uint64_t> words;
    uint32_t slots{0};
    host_paths(uint64_t w, row_map const & km, row_map const & cm, uint64_t const * kr, uint64_t const * cr)
        : W(w), kmer_row(km), code_row(cm), kmer_rows(kr), code_rows(cr)
    {
        words.reserve(64 * W);
    }
    uint32_t fresh()
    {
        words.resize(words.size() + W, 0);
        return slots++;
    }
    void fill(uint32_t slot, uint64_t bins)
    {
        for (uint64_t w = 0; w < W; ++w)
            words[slot * W + w] = bins >= 64 * (w + 1) ? ~0ull : (bins > 64 * w ? (1ull << (bins - 64 * w)) - 1 : 0);
    }
    void and_rows(std::vector<uint64_t> const & in, std::vector<uint32_t> const & sl, std::vector<uint8_t> & res,
                  std::unordered_map<uint64_t, uint32_t> const & row_of, uint64_t const * rows)
    {
        for (size_t i = 0; i < in.size(); ++i)
        {
            uint64_t const * row = rows + static_cast<size_t>(row_of.at(in[i])) * W;
            uint64_t * path = words.data() + static_cast<size_t>(sl[i]) * W;
            uint64_t any = 0;
            for (uint64_t w = 0; w < W; ++w)
                any |= (path[w] &= row[w]);
            res[i] = any != 0;
        }
    }
    void step(std::vector<uint64_t> const & in, std::vector<uint32_t> const & sl, std::vector<uint8_t> & res)
    {
        and_rows(in, sl, res, kmer_row, kmer_rows);
    }
    void step_dgram(std::vector<uint64_t> const & in, std::vector<uint32_t> const & sl, std::vector<uint8_t> & res)
    {
        and_rows(in, sl, res, code_row, code_rows);
    }
    void or_pairs(std::vector<uint32_t> const & dst, std::vector<uint32_t> const & src)
    {
        for (size_t i = 0; i < dst.size(); ++i)
            for (uint64_t w = 0; w < W; ++w)
                words[static_cast<size_t>(dst[i]) * W + w] |= words[static_cast<size_t>(src[i]) * W + w];
    }
};

// The walk itself, over any store of path bitvectors (`Paths`): HBM-resident slots driven rank by rank (device_paths),
// a dry run that only records which k-mers the walk can ask for (recording_paths), or host words over membership rows
// fetched in one batch (host_paths).  Returns the slot that holds path_matrix.
template <class Paths>
uint32_t walk(kgraph const & g, DeviceIndex & index, collect_stats * stats, Paths & paths, uint64_t bins)
{
    size_t const n_nodes = g.size();
    uint8_t const k = index.k();
    unsigned const bits = index.is_dna() ? 2 : 5;
    // create_selection_bitmask (otf_collector.h:168-182): the low (k-1) symbols
    uint64_t submask = 0;
    for (unsigned i = 0; i + 1 < k; ++i)
        submask = (submask << bits) | ((1ull << bits) - 1);

    uint32_t const match_slot = paths.fresh(); // path_matrix = bitvector(bins): all clear (fresh slots are zero)
    uint32_t const start_slot = paths.fresh();
    paths.fill(start_slot, bins); // hit_vector = bitvector(bins, true)

    std::vector<bucket> table(n_nodes);
    std::vector<int> node_of_rank(n_nodes, -1);
    for (size_t id = 0; id < n_nodes; ++id)
        node_of_rank[g.rank[id]] = static_cast<int>(id);

    std::vector<uint32_t> or_dst, or_src;
    auto push = [&](int node, item it, bool may_hand_over) {
        bucket & b = table[g.rank[node]];
        uint64_t const key = it.kmer & submask;
        uint32_t const found = b.by_key.find(key);
        if (found != rh_slots::npos)
        {
            // absorb (otf_collector.h:204-208): only the path is merged, kmer_/shift_count_ of the first arrival stay
            or_dst.push_back(b.items[found].slot);
            or_src.push_back(it.slot);
            return;
        }
        if (!may_hand_over)
        {
            uint32_t copy = paths.fresh(); // zero → OR = copy
            or_dst.push_back(copy);
            or_src.push_back(it.slot);
            it.slot = copy;
        }
        b.by_key.insert_new(key, static_cast<uint32_t>(b.items.size()));
        b.items.push_back(it);
    };

    push(0, item{0, 0, start_slot}, true);

    // OTFCollector::dibf_ / has_dibf_: without `-g` the reference hands a default-constructed DGramIndex to the collector,
    // whose gap bounds are both 0 (include/query.h:258-264, dGramIndex.h:67-68)
    DGramIndex const * const dibf = index.dgram();
    uint64_t const dgram_min = dibf ? dibf->min_gap() : 0, dgram_max = dibf ? dibf->max_gap() : 0;
    std::vector<uint64_t> dgram_codes;
    std::vector<uint32_t> dgram_slots, dgram_item;

    std::vector<uint64_t> probe_kmers;
    std::vector<uint32_t> probe_slots, probe_item;
    std::vector<uint8_t> alive;

    for (size_t r = 0; r < n_nodes; ++r)
    {
        bucket & b = table[r];
        if (b.items.empty())
            continue;
        b.sort_into_pop_order();
        int const node = node_of_rank[r];
        int const sym = g.symbol[node];
        or_dst.clear();
        or_src.clear();
        if (stats)
        {
            ++stats->ranks;
            stats->items += b.items.size();
        }
        if (sym == Match)
        {
            for (item const & it : b.items)
            {
                or_dst.push_back(match_slot); // path_matrix |= top.path_
                or_src.push_back(it.slot);
            }
        }
        else if (sym == Ghost || sym == '$')
        {
            for (item const & it : b.items)
                push(g.first[node], it, true);
        }
        else if (sym == Split)
        {
            for (item const & it : b.items)
            {
                push(g.first[node], it, false);
                push(g.second[node], it, false);
            }
        }
        else if (sym == Gap)
        {
            // gap_procedure (otf_collector.h:290-312)
            uint64_t const gap = g.gap[node];
            for (item const & it : b.items)
            {
                item nx{0, 0, it.slot};
                if (!(it.shift_count < 3 || gap < dgram_min || gap > dgram_max))
                {
                    // the last three residues as update_kmer packed them (5 bits each), weighted like the index side's
                    // a1 a2 a3 — the packed codes, not aa_to_num: restated as written
                    uint64_t const a1 = (it.kmer & 0b111110000000000) >> 10, a2 = (it.kmer & 0b1111100000) >> 5, a3 = it.kmer & 0b11111;
                    nx.kmer = gap * 64000000ULL + a1 * 3200000ULL + a2 * 160000ULL + a3 * 8000ULL;
                    nx.gapped = true;
                }
                push(g.first[node], nx, true);
            }
        }
        else
        {
            // update_gapped (otf_collector.h:216-245) for items that have just jumped a Gap node, update_path (:247-279)
            // for the rest
            probe_kmers.clear();
            probe_slots.clear();
            probe_item.clear();
            dgram_codes.clear();
            dgram_slots.clear();
            dgram_item.clear();
            for (size_t i = 0; i < b.items.size(); ++i)
            {
                item & it = b.items[i];
                if (it.gapped)
                {
                    if (it.shift_count == 0)
                    {
                        it.kmer += 400ULL * dgram_aa_to_num(sym);
                        it.res1 = sym;
                        ++it.shift_count;
                    }
                    else if (it.shift_count == 1)
                    {
                        it.kmer += 20ULL * dgram_aa_to_num(sym);
                        it.res2 = sym;
                        ++it.shift_count;
                    }
                    else if (it.shift_count == 2)
                    {
                        if (dibf) // `if(has_dibf_) path_ &= dibf_->query(dgram)`
                        {
                            dgram_codes.push_back(it.kmer + dgram_aa_to_num(sym));
                            dgram_slots.push_back(it.slot);
                            dgram_item.push_back(static_cast<uint32_t>(i));
                        }
                        it.kmer = 0; // the three residues after the gap start the next k-mer
                        index.update_kmer(it.res1, it.kmer);
                        index.update_kmer(it.res2, it.kmer);
                        index.update_kmer(sym, it.kmer);
                        ++it.shift_count;
                        it.gapped = false;
                        it.res1 = it.res2 = 0;
                    }
                    continue;
                }
                if (it.shift_count + 1 < k)
                {
                    index.update_kmer(sym, it.kmer);
                    ++it.shift_count;
                    continue;
                }
                if (it.shift_count > k) // k < 3 after a gap: none of update_path's three cases (:251,:256,:268) applies
                    continue;
                uint64_t const probe = index.update_kmer(sym, it.kmer);
                if (it.shift_count + 1 == k)
                    ++it.shift_count;
                probe_kmers.push_back(probe);
                probe_slots.push_back(it.slot);
                probe_item.push_back(static_cast<uint32_t>(i));
            }
            alive.assign(b.items.size(), 1);
            if (!probe_kmers.empty())
            {
                std::vector<uint8_t> res(probe_kmers.size());
                paths.step(probe_kmers, probe_slots, res);
                for (size_t j = 0; j < res.size(); ++j)
                    alive[probe_item[j]] = res[j];
                if (stats)
                    stats->probes += probe_kmers.size();
            }
            if (!dgram_codes.empty())
            {
                std::vector<uint8_t> res(dgram_codes.size());
                paths.step_dgram(dgram_codes, dgram_slots, res);
                for (size_t j = 0; j < res.size(); ++j)
                    alive[dgram_item[j]] = res[j];
                if (stats)
                    stats->probes += dgram_codes.size();
            }
            for (size_t i = 0; i < b.items.size(); ++i)
                if (alive[i]) // `if(top.path_.none()) break;` drops dead paths at once
                    push(g.first[node], b.items[i], true);
        }
        if (!or_dst.empty())
            paths.or_pairs(or_dst, or_src);
        b.items.clear();
        b.by_key.clear();
    }
    return match_slot;
}

bitvector collect(kgraph const & g, DeviceIndex & index, collect_stats * stats, shard_hooks const * hooks)
{
    // a column shard of a flat index walks the same graph over its own bins only: paths are shard_words() wide and the
    // pruning decision is made global through hooks->any_alive
    bool const sharded = !index.is_hibf() && index.shard_count() > 1;
    uint64_t const word_begin = sharded ? index.shard_word_begin() : 0;
    uint64_t const W = index.path_words();
    uint64_t const bins = sharded ? std::min<uint64_t>(W * 64, index.bin_count() > word_begin * 64 ? index.bin_count() - word_begin * 64 : 0)
                                  : index.bin_count();
    if (sharded && !(hooks && (hooks->any_alive || hooks->comm)))
        throw error("a column-sharded index needs shard_hooks::any_alive or ::comm (the pruning decision spans all shards)");
    trx_comm_t * const comm = sharded && hooks ? hooks->comm : nullptr;

    // One device round trip for the whole query where the k-graph is small enough to probe speculatively: a dry run of the
    // walk lists every k-mer it can ask for, ONE membership batch brings their rows, and the walk proper runs on host
    // words (the reference's own arithmetic).  Pruning then saves no probes — it never saved device time on these graphs,
    // where a rank is one launch and one synchronisation (6.5 us flat, 19 us HIBF per rank) — and the candidate set and the
    // statistics are those of the pruned walk.  Column shards and wide graphs keep their paths in HBM.
    char const * const spec_env = getenv("TRX_COLLECT_SPECULATIVE"); // =off: A/B and tests
    bool const speculative_off = spec_env && std::string(spec_env) == "off";
    if (!sharded && !speculative_off)
    {
        recording_paths rec;
        bool fits = true;
        try
        {
            walk(g, index, nullptr, rec, bins);
        }
        catch (budget_exceeded const &)
        {
            fits = false;
        }
        if (fits)
        {
            std::vector<uint64_t> kmer_rows(rec.kmers.size() * W), code_rows(rec.codes.size() * W);
            if (!rec.kmers.empty())
            {
                if (index.is_hibf())
                    check(trx_hibf_membership(index.hibf(), rec.kmers.data(), rec.kmers.size(), kmer_rows.data(), nullptr));
                else
                    check(trx_bulk_contains(index.flat(), rec.kmers.data(), rec.kmers.size(), kmer_rows.data()));
            }
            if (!rec.codes.empty())
                check(trx_bulk_contains(index.dgram()->flat(), rec.codes.data(), rec.codes.size(), code_rows.data()));
            host_paths hp(W, rec.kmer_row, rec.code_row, kmer_rows.data(), code_rows.data());
            uint32_t const match_slot = walk(g, index, stats, hp, bins);
            bitvector out(index.bin_words(), 0);
            std::copy(hp.words.begin() + static_cast<size_t>(match_slot) * W, hp.words.begin() + static_cast<size_t>(match_slot + 1) * W,
                      out.begin());
            return out;
        }
    }
    prepare_pool(index, W);
    device_paths paths(index, hooks, comm, sharded, W);
    uint32_t const match_slot = walk(g, index, stats, paths, bins);
    slot_pool & pool = paths.pool;
    bitvector out(index.bin_words(), 0);
    check(trx_paths_read(pool.paths(), match_slot, 1, out.data() + word_begin));
    if (comm) // every rank contributes its column block (equal sizes: the shard count divides bin_words) and gets all
    {
        bitvector mine(out.begin() + word_begin, out.begin() + word_begin + W);
        check(trx_comm_allgather(comm, mine.data(), out.data(), W * 8));
    }
    return out;
}

namespace
{
kgraph graph_of(std::string const & regex, DeviceIndex & index, bool augment);
}

// process_query for a list of patterns (`-f`): every pattern's k-graph gets its dry run, the k-mers of ALL of them go to the
// device in one membership batch (in pieces of at most kBatchRowBytes of rows), and each walk then runs on host words.
// Patterns whose graph is over the dry run's budget, and column shards, take process_query one at a time.
std::vector<batch_answer> process_query_batch(std::vector<std::string> const & regexes, DeviceIndex & index, bool augment)
{
    constexpr size_t kChunk = 256; // patterns per membership batch: at most 256 x 1024 k-mers' rows on the host at a time
    std::vector<batch_answer> out(regexes.size());
    bool const sharded = !index.is_hibf() && index.shard_count() > 1;
    char const * const spec_env = getenv("TRX_COLLECT_SPECULATIVE");
    bool const one_by_one = sharded || (spec_env && std::string(spec_env) == "off");
    uint64_t const W = index.path_words();
    uint64_t const bins = index.bin_count();
    auto answer_alone = [&](size_t i) {
        try
        {
            out[i].bins = process_query(regexes[i], index, &out[i].stats, nullptr, augment);
            out[i].ok = true;
        }
        catch (std::exception const & e)
        {
            out[i].error = e.what();
        }
    };
    if (one_by_one)
    {
        for (size_t i = 0; i < regexes.size(); ++i)
            answer_alone(i);
        return out;
    }
    // host threads: k-graphs, dry runs and walks of different patterns have nothing in common but read-only tables
    unsigned const workers = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    auto parallel = [&](size_t n, auto && body) {
        std::atomic<size_t> next{0};
        auto work = [&] {
            for (size_t i = next++; i < n; i = next++)
                body(i);
        };
        unsigned const w = static_cast<unsigned>(std::min<size_t>(workers, n));
        if (w <= 1)
        {
            work();
            return;
        }
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < w; ++t)
            pool.emplace_back(work);
        for (std::thread & t : pool)
            t.join();
    };
    enum class fate : uint8_t { walk, alone, failed };
    bool const trace = getenv("TRX_COLLECT_TRACE") != nullptr; // where a batch spends its time, on stderr
    auto const now = [] { return std::chrono::steady_clock::now(); };
    auto const ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    double ms_dry = 0, ms_merge = 0, ms_device = 0, ms_walk = 0, ms_alone = 0;
    std::atomic<uint64_t> us_graph{0}, us_dry{0};
    for (size_t begin = 0; begin < regexes.size(); begin += kChunk)
    {
        auto const t0 = now();
        size_t const n = std::min(kChunk, regexes.size() - begin);
        std::vector<kgraph> graphs(n);
        // four times a single query's budget — a batch has the host threads to walk wider graphs, but the dry run of a wide
        // graph visits far more than its pruned walk does.  Measured on 1000 C3 patterns with -a: 1024 k-mers per pattern
        // 160 ms, 2048 176, 4096 144, 8192 145, 32 768 181
        static size_t const batch_probes = [] {
            char const * e = getenv("TRX_COLLECT_BATCH_PROBES"); // A/B of the dry run's budget inside a batch
            size_t const v = e ? strtoull(e, nullptr, 10) : 0;
            return v ? v : size_t{4096};
        }();
        std::vector<recording_paths> recs(n, recording_paths(batch_probes, 1u << 17));
        std::vector<fate> what(n, fate::walk);
        parallel(n, [&](size_t j) {
            try
            {
                auto const g0 = now();
                graphs[j] = graph_of(regexes[begin + j], index, augment);
                auto const g1 = now();
                walk(graphs[j], index, nullptr, recs[j], bins);
                if (trace)
                {
                    us_graph += static_cast<uint64_t>(ms(g0, g1) * 1000);
                    us_dry += static_cast<uint64_t>(ms(g1, now()) * 1000);
                }
            }
            catch (budget_exceeded const &)
            {
                what[j] = fate::alone;
            }
            catch (std::exception const & e)
            {
                what[j] = fate::failed;
                out[begin + j].error = e.what();
            }
        });
        // the union of the chunk's k-mers, in pattern order (the result does not depend on the order; the batch does not
        // depend on the thread count)
        auto const t1 = now();
        // groups of patterns whose k-mers' rows fit kGroupRowBytes: one membership batch and one round of walks per group
        constexpr uint64_t kGroupRowBytes = 256ull << 20;
        size_t group_begin = 0;
        while (group_begin < n)
        {
            auto const u0 = now();
            // the union of the group's k-mers, in pattern order (the result does not depend on the order; the batch does
            // not depend on the thread count)
            host_paths::row_map kmer_row, code_row;
            std::vector<uint64_t> kmers, codes;
            size_t group_end = group_begin;
            for (; group_end < n; ++group_end)
            {
                size_t const j = group_end;
                if (what[j] != fate::walk)
                    continue;
                if (group_end > group_begin && (kmers.size() + codes.size() + recs[j].kmers.size() + recs[j].codes.size()) * W * 8 > kGroupRowBytes)
                    break;
                for (uint64_t v : recs[j].kmers)
                    if (kmer_row.emplace(v, static_cast<uint32_t>(kmers.size())).second)
                        kmers.push_back(v);
                for (uint64_t v : recs[j].codes)
                    if (code_row.emplace(v, static_cast<uint32_t>(codes.size())).second)
                        codes.push_back(v);
            }
            auto const u1 = now();
            std::unique_ptr<uint64_t[]> kmer_rows(new uint64_t[kmers.size() * W + 1]), code_rows(new uint64_t[codes.size() * W + 1]);
            if (!kmers.empty())
            {
                if (index.is_hibf())
                    check(trx_hibf_membership(index.hibf(), kmers.data(), kmers.size(), kmer_rows.get(), nullptr));
                else
                    check(trx_bulk_contains(index.flat(), kmers.data(), kmers.size(), kmer_rows.get()));
            }
            if (!codes.empty())
                check(trx_bulk_contains(index.dgram()->flat(), codes.data(), codes.size(), code_rows.get()));
            auto const u2 = now();
            size_t const gb = group_begin;
            parallel(group_end - group_begin, [&](size_t jj) {
                size_t const j = gb + jj;
                if (what[j] != fate::walk)
                    return;
                batch_answer & a = out[begin + j];
                try
                {
                    host_paths hp(W, kmer_row, code_row, kmer_rows.get(), code_rows.get());
                    uint32_t const match_slot = walk(graphs[j], index, &a.stats, hp, bins);
                    a.bins.assign(hp.words.begin() + static_cast<size_t>(match_slot) * W,
                                  hp.words.begin() + static_cast<size_t>(match_slot + 1) * W);
                    a.ok = true;
                }
                catch (std::exception const & e)
                {
                    a.error = e.what();
                }
            });
            ms_merge += ms(u0, u1);
            ms_device += ms(u1, u2);
            ms_walk += ms(u2, now());
            group_begin = group_end;
        }
        auto const t4 = now();
        for (size_t j = 0; j < n; ++j) // graphs over the dry run's budget: the device-resident walk, one at a time
            if (what[j] == fate::alone)
                answer_alone(begin + j);
        ms_dry += ms(t0, t1);
        ms_alone += ms(t4, now());
    }
    if (trace)
        fprintf(stderr, "process_query_batch: %zu patterns, %u host threads: graphs + dry runs %.1f ms (thread time: graphs %.1f, dry runs "
                        "%.1f), union of k-mers %.1f, membership %.1f, walks %.1f, one at a time %.1f\n", regexes.size(), workers, ms_dry,
                us_graph.load() / 1000.0, us_dry.load() / 1000.0, ms_merge, ms_device, ms_walk, ms_alone);
    return out;
}

namespace
{
kgraph graph_of(std::string const & regex, DeviceIndex & index, bool augment)
{
    std::string rx = regex;
    if (!index.is_dna())
    {
        if (index.reduction() > 0)
            rx = reduce_query_alphabet(rx, index.redmap());
        rx = trim_regex(rx);
    }
    std::string postfix = translate(rx);
    kgraph g = construct_kgraph(postfix, index.k(), index.reduction() != TRX_RED_BASE);
    if (augment)
        g.augment();
    return g;
}
} // namespace

bitvector process_query_with_rows(std::string const & regex, DeviceIndex & index, rows_fn const & rows, collect_stats * stats,
                                  bool augment)
{
    if (index.dgram())
        throw error("process_query_with_rows: DGram filters are not supported");
    kgraph g = graph_of(regex, index, augment);
    uint64_t const W = index.bin_words(), bins = index.bin_count();
    recording_paths rec(size_t{1} << 24, size_t{1} << 24); // the caller asked for this route: no budget
    walk(g, index, nullptr, rec, bins);
    std::vector<uint64_t> kmer_rows(rec.kmers.size() * W + 1, 0), no_codes(1, 0);
    if (!rec.kmers.empty())
        rows(rec.kmers.data(), rec.kmers.size(), kmer_rows.data());
    host_paths hp(W, rec.kmer_row, rec.code_row, kmer_rows.data(), no_codes.data());
    uint32_t const match_slot = walk(g, index, stats, hp, bins);
    return bitvector(hp.words.begin() + static_cast<size_t>(match_slot) * W, hp.words.begin() + static_cast<size_t>(match_slot + 1) * W);
}

bitvector process_query(std::string const & regex, DeviceIndex & index, collect_stats * stats, shard_hooks const * hooks,
                        bool augment)
{
    // preprocess_query (include/query.h:80-94): peptides reduce (if the index is reduced) and trim, DNA only translates
    std::string rx = regex;
    if (!index.is_dna())
    {
        if (index.reduction() > 0)
            rx = reduce_query_alphabet(rx, index.redmap());
        rx = trim_regex(rx);
    }
    std::string postfix = translate(rx);
    kgraph g = construct_kgraph(postfix, index.k(), index.reduction() != TRX_RED_BASE); // include/query.h:239-240
    if (augment)
        g.augment(); // include/query.h:243: `if(augment && catsites.size() > 0) collector.augment(catsites);`
    return collect(g, index, stats, hooks);
}

} // namespace trxh
