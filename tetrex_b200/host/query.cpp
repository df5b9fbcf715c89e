// query.cpp — the query side above the collector: iter_disk_search (include/query.h:97-188), run_collection
// (include/query.h:250-291), run_multiple_queries (:330-340), drive_query (src/query.cpp:477-498), drive_index
// (src/index_base.cpp:73-118) and drive_inspection (src/inspect_idx.cpp).  Same messages on the same streams
// (seqan3::debug_stream is stderr), same output lines; every scan is one trx_regex_scan call (K4).
#include <algorithm>
#include <bit>
#include <chrono>
#include <iterator>
#include <memory>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <thread>

#include "trx_host.hpp"

namespace trxh
{

namespace
{

// comp_tab (src/query.cpp:7-16) restated: IUPAC complement for letters, identity elsewhere
char complement(char c)
{
    static char const upper[] = "TVGHEFCDIJMLKNOPQYSAABWXRZ"; // complement of 'A'..'Z'
    if (c >= 'A' && c <= 'Z')
        return upper[c - 'A'];
    if (c >= 'a' && c <= 'z')
        return static_cast<char>(upper[c - 'a'] + ('a' - 'A'));
    if (c == '`')
        return '@'; // the table maps 96 to 64
    return c;
}

uint64_t popcount(bitvector const & v)
{
    uint64_t n = 0;
    for (uint64_t w : v)
        n += std::popcount(w);
    return n;
}

struct regex_handle
{
    trx_regex_t * re{nullptr};
    regex_handle(std::string const & pattern, bool posix)
    {
        check(trx_regex_compile(pattern.c_str(), posix ? 1 : 0, &re));
    }
    ~regex_handle() { trx_regex_destroy(re); }
};

constexpr uint64_t kHitRoom = 1u << 20;

std::vector<trx_hit> scan(DeviceIndex::text_store const & store, regex_handle & rx, bitvector const & cand, bool reverse,
                          uint8_t const * map)
{
    // room for kHitRoom hits up front — untouched memory costs nothing, and the device side keeps its buffers — so that a
    // second scan is only ever needed beyond that
    uint64_t room = kHitRoom;
    for (;;)
    {
        std::unique_ptr<trx_hit[]> buf(new trx_hit[room]);
        uint64_t n = 0;
        int rc = trx_regex_scan(store.text, rx.re, cand.data(), reverse ? 1 : 0, map, buf.get(), room, &n, nullptr);
        if (rc == TRX_ECAPACITY)
        {
            room = n + n / 8 + 16;
            continue;
        }
        check(rc);
        return std::vector<trx_hit>(buf.get(), buf.get() + n);
    }
}

} // namespace

namespace
{
uint64_t search_impl(bitvector const & hits_in, std::string const & query_in, DeviceIndex & index, std::ostream * out,
                     std::ostream * rev_out);
}

uint64_t iter_disk_search(bitvector const & hits_in, std::string const & query_in, DeviceIndex & index, std::string const & dest)
{
    std::ofstream file_out;
    std::ostream * out = &std::cout;
    if (dest != "-")
    {
        file_out.open(dest);
        if (!file_out)
            throw error("Failed to open output file: " + dest);
        out = &file_out;
    }
    // reverse hits always go to stdout, whatever -o says (std::osyncstream(std::cout), src/query.cpp:186)
    return search_impl(hits_in, query_in, index, out, &std::cout);
}

uint64_t iter_disk_search(bitvector const & hits_in, std::string const & query_in, DeviceIndex & index, std::ostream & out)
{
    return search_impl(hits_in, query_in, index, &out, &out);
}

namespace
{

uint64_t search_impl(bitvector const & hits_in, std::string const & query_in, DeviceIndex & index, std::ostream * out,
                     std::ostream * rev_out)
{
    std::vector<std::string> const & libs = index.acid_libs();
    // compute_set_bins (src/query.cpp:40-76): a one-bin library is always scanned
    bitvector cand((libs.size() + 63) / 64, 0);
    if (libs.size() == 1)
        cand[0] = 1;
    else
        for (size_t w = 0; w < cand.size() && w < hits_in.size(); ++w)
            cand[w] = hits_in[w];
    if (index.shard_count() > 1) // a shard verifies its own bins only (its text store holds nothing else)
        for (size_t w = 0; w < cand.size(); ++w)
            if (w < index.shard_word_begin() || w >= index.shard_word_begin() + index.shard_words())
                cand[w] = 0;
    if (libs.size() % 64) // bits past the last bin never name a file
        cand.back() &= (1ull << (libs.size() % 64)) - 1;
    // nothing to verify here — no candidate among this shard's bins, or a shard that owns no bins at all (more shards
    // than 64-bin words): the reference's loop over set bins simply does not execute (include/query.h:120-139)
    if (popcount(cand) == 0)
        return 0;
    DeviceIndex::text_store const & store = index.text(cand);

    std::string query = query_in;
    bool const dna = index.is_dna();
    if (!dna && index.reduction() > 0)
        query = reduce_query_alphabet(query, index.redmap()); // include/query.h:144
    regex_handle rx(query, /*posix=*/!dna);                    // RE2::POSIX for peptides, default options for DNA

    uint8_t const * map = (!dna && index.reduction() > 0) ? index.redmap() : nullptr;
    std::vector<trx_hit> fwd, rev;
    if (!dna)
        fwd = scan(store, rx, cand, false, map);
    else
    {
        // both strands in flight at once (trx_regex_scan_begin/_end): the second scan's kernels follow the first one's
        // without a host round trip in between
        int t_fwd = -1, t_rev = -1;
        uint64_t const cap = kHitRoom;
        check(trx_regex_scan_begin(store.text, rx.re, cand.data(), 0, map, cap, &t_fwd));
        int const rc_begin = trx_regex_scan_begin(store.text, rx.re, cand.data(), 1, nullptr, cap, &t_rev);
        // a ticket that is not ended keeps its slot busy on the persistent text handle: whatever happens below, both
        // are ended before this scope is left
        struct ticket_guard
        {
            trx_text_t * text;
            int * ticket;
            ~ticket_guard()
            {
                if (*ticket >= 0)
                    trx_regex_scan_end(text, *ticket, nullptr, nullptr, nullptr);
            }
        } guard_fwd{store.text, &t_fwd}, guard_rev{store.text, &t_rev};
        auto finish = [&](int & ticket, bool reverse, uint8_t const * m) {
            std::unique_ptr<trx_hit[]> hits(new trx_hit[cap]);
            uint64_t n = 0;
            int const rc = trx_regex_scan_end(store.text, ticket, hits.get(), &n, nullptr);
            ticket = -1;
            if (rc == TRX_ECAPACITY)
                return scan(store, rx, cand, reverse, m); // more hits than the speculative buffer: once more, with room
            check(rc);
            return std::vector<trx_hit>(hits.get(), hits.get() + n);
        };
        fwd = finish(t_fwd, false, map);
        if (rc_begin != TRX_OK)
            check(rc_begin);
        rev = finish(t_rev, true, nullptr);
    }

    uint64_t lines = 0;
    std::string line, match;
    auto emit_forward = [&](trx_hit const & h) {
        char const * rec = store.residues.data() + store.offsets[h.record];
        match.assign(rec + h.start, h.end - h.start);
        if (map) // verify_reduced_fasta_hit prints the REDUCED residues (src/query.cpp:284-287)
            for (char & c : match)
                c = static_cast<char>(map[static_cast<uint8_t>(c)]);
        *out << libs[h.bin] << '\t' << store.names[h.record] << '\t' << match << '\t' << h.start << ',' << h.end << '\n';
        ++lines;
    };
    auto emit_reverse = [&](trx_hit const & h) {
        // coordinates are in the reverse-complemented record (src/query.cpp:167-191)
        uint64_t const len = store.offsets[h.record + 1] - store.offsets[h.record];
        char const * rec = store.residues.data() + store.offsets[h.record];
        match.clear();
        for (uint32_t p = h.start; p < h.end; ++p)
            match.push_back(complement(rec[len - 1 - p]));
        *rev_out << libs[h.bin] << '\t' << store.names[h.record] << '\t' << match << '\t' << "REVERSE STRAND HIT" << std::endl;
        ++lines;
    };
    // per bin: all forward lines, then all reverse lines (include/query.h:132-134); bins ascending (-t 1 order)
    size_t i = 0, j = 0;
    while (i < fwd.size() || j < rev.size())
    {
        uint32_t bin = UINT32_MAX;
        if (i < fwd.size())
            bin = fwd[i].bin;
        if (j < rev.size())
            bin = std::min(bin, rev[j].bin);
        for (; i < fwd.size() && fwd[i].bin == bin; ++i)
            emit_forward(fwd[i]);
        if (j < rev.size() && rev[j].bin == bin)
        {
            out->flush();
            for (; j < rev.size() && rev[j].bin == bin; ++j)
                emit_reverse(rev[j]);
        }
    }
    out->flush();
    return lines;
}

double now()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// candidate words as the verify step sees them: real bins only, this shard's bins only
bitvector verify_candidates(bitvector const & hits_in, DeviceIndex & index)
{
    std::vector<std::string> const & libs = index.acid_libs();
    bitvector cand((libs.size() + 63) / 64, 0);
    if (libs.size() == 1)
        cand[0] = 1;
    else
        for (size_t w = 0; w < cand.size() && w < hits_in.size(); ++w)
            cand[w] = hits_in[w];
    if (index.shard_count() > 1)
        for (size_t w = 0; w < cand.size(); ++w)
            if (w < index.shard_word_begin() || w >= index.shard_word_begin() + index.shard_words())
                cand[w] = 0;
    if (libs.size() % 64)
        cand.back() &= (1ull << (libs.size() % 64)) - 1;
    return cand;
}

// K4 for a batch of patterns in shared passes over the text (trx_regex_scan_multi); hits grouped per pattern
std::vector<std::vector<trx_hit>> scan_batch(DeviceIndex::text_store const & store, std::vector<regex_handle *> const & rx,
                                             std::vector<uint64_t> const & cand_flat, uint64_t stride)
{
    std::vector<trx_regex_t *> handles;
    for (regex_handle * r : rx)
        handles.push_back(r->re);
    std::vector<trx_hit> hits(1 << 16);
    std::vector<uint32_t> pats(hits.size());
    std::vector<uint64_t> per(rx.size());
    for (;;)
    {
        uint64_t n = 0;
        int const rc = trx_regex_scan_multi(store.text, handles.data(), static_cast<uint32_t>(handles.size()), cand_flat.data(),
                                            stride, hits.data(), pats.data(), hits.size(), &n, per.data(), nullptr);
        if (rc == TRX_ECAPACITY)
        {
            hits.resize(std::max<uint64_t>(2 * hits.size(), n + n / 4 + 16));
            pats.resize(hits.size());
            continue;
        }
        check(rc);
        std::vector<std::vector<trx_hit>> out(rx.size());
        uint64_t o = 0;
        for (size_t i = 0; i < rx.size(); ++i)
        {
            out[i].assign(hits.begin() + o, hits.begin() + o + per[i]);
            o += per[i];
        }
        return out;
    }
}

// iter_disk_search_set + verify_fasta_set (include/query.h:191-224, src/query.cpp:318-339): a record of a candidate bin is
// reported when EVERY query matches somewhere in it; the line names the queries in the order RE2::Set::Match returns
// them — the order in which their first matches complete in the record, ties in descending order of "(" + query + ")"
// (RE2 sorts the set's patterns and reads the match list of its final DFA state backwards).  The reference compiles the
// queries as given (no alphabet reduction, forward strand only) and writes to stdout whatever -o says.
uint64_t iter_disk_search_set(bitvector const & hits_in, std::vector<std::string> const & queries, DeviceIndex & index,
                              std::ostream & out)
{
    bitvector const cand = verify_candidates(hits_in, index);
    if (popcount(cand) == 0 || queries.empty())
        return 0;
    DeviceIndex::text_store const & store = index.text(cand);
    std::vector<std::unique_ptr<regex_handle>> owned;
    std::vector<regex_handle *> rx;
    for (std::string const & q : queries)
    {
        owned.push_back(std::make_unique<regex_handle>(q, /*posix=*/false)); // RE2::DefaultOptions; existence is all that counts
        rx.push_back(owned.back().get());
    }
    std::vector<std::vector<trx_hit>> const per = scan_batch(store, rx, cand, 0);
    // records in which every query has a match: intersect the (sorted) record lists
    std::vector<uint32_t> records;
    for (size_t i = 0; i < per.size(); ++i)
    {
        std::vector<uint32_t> mine;
        for (trx_hit const & h : per[i])
            if (mine.empty() || mine.back() != h.record)
                mine.push_back(h.record);
        if (i == 0)
            records = mine;
        else
        {
            std::vector<uint32_t> both;
            std::set_intersection(records.begin(), records.end(), mine.begin(), mine.end(), std::back_inserter(both));
            records.swap(both);
        }
    }
    std::vector<std::string> const & libs = index.acid_libs();
    uint64_t lines = 0;
    for (uint32_t r : records)
    {
        auto const * rec = reinterpret_cast<uint8_t const *>(store.residues.data() + store.offsets[r]);
        uint64_t const len = store.offsets[r + 1] - store.offsets[r];
        std::vector<std::pair<int64_t, size_t>> order;
        for (size_t i = 0; i < queries.size(); ++i)
            order.emplace_back(trx_regex_first_match_end(rx[i]->re, rec, len), i);
        std::stable_sort(order.begin(), order.end(), [&](auto const & a, auto const & b) {
            if (a.first != b.first)
                return a.first < b.first;
            return "(" + queries[a.second] + ")" > "(" + queries[b.second] + ")";
        });
        out << libs[store.rec_bin[r]] << '\t' << store.names[r] << "\tN --> ";
        for (auto const & o : order)
            out << queries[o.second] << " --> ";
        out << "C" << std::endl;
        ++lines;
    }
    return lines;
}

// run_conjunction (include/query.h:293-326): the candidate bins of a conjunction are the AND of its queries' filters
void run_conjunction(query_arguments & args, DeviceIndex & index)
{
    double const t1 = now();
    std::vector<std::string> queries; // split_string(input_regex, ':') (src/query.cpp:365-371, :463-466)
    {
        std::stringstream ss(args.regex);
        std::string token;
        while (std::getline(ss, token, ':'))
            queries.push_back(token);
    }
    bitvector hit_vector(index.bin_words(), 0);
    for (uint64_t b = 0; b < index.bin_count(); ++b)
        hit_vector[b / 64] |= 1ull << (b % 64);
    if (index.bin_count() > 1)
    {
        for (std::string const & q : queries)
        {
            bitvector got = process_query(q, index, nullptr, nullptr, args.augment);
            for (size_t w = 0; w < hit_vector.size(); ++w)
                hit_vector[w] &= got[w];
        }
    }
    else
        std::cerr << "[WARNING] Index contains only 1 bin. Unable to accelerate search using the TetRex algorithm. "
                     "Performing Linear Scan"
                  << std::endl;
    if (args.verbose)
        std::cerr << "Narrowed Search to " << popcount(hit_vector) << " possible bins" << std::endl;
    if (popcount(hit_vector) != 0)
    {
        try
        {
            iter_disk_search_set(hit_vector, queries, index, std::cout);
        }
        catch (std::exception const & e)
        {
            std::cerr << e.what() << '\n';
        }
    }
    if (args.verbose)
        std::cerr << "Query Time: " << (now() - t1) << std::endl;
}

// run_multiple_queries (include/query.h:330-340) for a whole file of motifs: the filter phase runs motif by motif (every
// motif has its own k-graph and candidate bins), the verify phase ONCE for all of them — pattern i over its own candidate
// bins, every tile of the union read once per 128 patterns.  Same files (<id>.tsv), same lines, same stderr fields as the
// loop over run_collection; "Query Time" is the motif's filter time plus its share of the shared verify.
void run_motif_batch(std::vector<std::pair<std::string, std::string>> const & motifs, DeviceIndex & index, bool augment)
{
    size_t const n = motifs.size();
    uint64_t const words = (index.acid_libs().size() + 63) / 64;
    std::vector<uint64_t> cands(n * words, 0);
    std::vector<double> filter_s(n, 0.0);
    std::vector<uint64_t> bin_count(n, 0);
    std::vector<std::unique_ptr<regex_handle>> owned(n);
    std::vector<size_t> live; // motifs that reach the verify phase
    bitvector uni(words, 0);
    // the filter phase of all motifs at once: one membership batch for their k-mers (process_query_batch)
    double const t_filter = now();
    std::vector<std::string> patterns;
    for (auto const & m : motifs)
        patterns.push_back(m.second);
    std::vector<batch_answer> answers = process_query_batch(patterns, index, augment);
    double const filter_each = (now() - t_filter) / static_cast<double>(std::max<size_t>(n, 1));
    for (size_t i = 0; i < n; ++i)
    {
        double const t1 = now();
        try
        {
            bitvector hit_vector(index.bin_words(), 0);
            for (uint64_t b = 0; b < index.bin_count(); ++b)
                hit_vector[b / 64] |= 1ull << (b % 64);
            if (!answers[i].ok)
                throw error(answers[i].error);
            bitvector const & got = answers[i].bins;
            for (size_t w = 0; w < hit_vector.size(); ++w)
                hit_vector[w] &= got[w];
            bin_count[i] = popcount(hit_vector);
            if (bin_count[i])
            {
                bitvector const c = verify_candidates(hit_vector, index);
                std::copy(c.begin(), c.end(), cands.begin() + i * words);
                owned[i] = std::make_unique<regex_handle>(motifs[i].second, /*posix=*/true);
                live.push_back(i);
                for (size_t w = 0; w < words; ++w)
                    uni[w] |= c[w];
            }
        }
        catch (std::exception const & e)
        {
            std::cerr << motifs[i].first << "\t" << e.what() << '\n';
            bin_count[i] = UINT64_MAX;
        }
        filter_s[i] = filter_each + (now() - t1);
    }
    double verify_s = 0.0;
    std::vector<std::vector<trx_hit>> per;
    DeviceIndex::text_store const * store = nullptr;
    if (!live.empty() && popcount(uni))
    {
        double const t1 = now();
        store = &index.text(uni);
        std::vector<regex_handle *> rx;
        std::vector<uint64_t> flat(live.size() * words);
        for (size_t k = 0; k < live.size(); ++k)
        {
            rx.push_back(owned[live[k]].get());
            std::copy(cands.begin() + live[k] * words, cands.begin() + (live[k] + 1) * words, flat.begin() + k * words);
        }
        per = scan_batch(*store, rx, flat, words);
        verify_s = now() - t1;
    }
    std::vector<std::string> const & libs = index.acid_libs();
    size_t k = 0;
    std::string match;
    for (size_t i = 0; i < n; ++i)
    {
        if (bin_count[i] == UINT64_MAX)
            continue;
        std::cerr << motifs[i].first << "\t" << "Bin Count: " << bin_count[i] << "\t";
        if (bin_count[i])
        {
            std::ofstream out(motifs[i].first + ".tsv");
            if (!out)
                std::cerr << "Failed to open output file: " << motifs[i].first << ".tsv" << '\n';
            else if (k < per.size())
                for (trx_hit const & h : per[k])
                {
                    char const * rec = store->residues.data() + store->offsets[h.record];
                    match.assign(rec + h.start, h.end - h.start);
                    out << libs[h.bin] << '\t' << store->names[h.record] << '\t' << match << '\t' << h.start << ',' << h.end << '\n';
                }
            ++k;
        }
        std::cerr << "Query Time: " << (filter_s[i] + verify_s / std::max<size_t>(live.size(), 1)) << std::endl;
    }
}

// run_collection (include/query.h:250-291)
void run_collection(query_arguments & args, bool read_file, DeviceIndex & index)
{
    bool const verbose = args.verbose && !read_file;
    double const t1 = now();
    bitvector hit_vector(index.bin_words(), 0);
    for (uint64_t b = 0; b < index.bin_count(); ++b)
        hit_vector[b / 64] |= 1ull << (b % 64);
    if (index.bin_count() > 1)
    {
        bitvector got = process_query(args.regex, index, nullptr, nullptr, args.augment);
        for (size_t w = 0; w < hit_vector.size(); ++w)
            hit_vector[w] &= got[w];
    }
    else
        std::cerr << "[WARNING] Index contains only 1 bin. Unable to accelerate search using the TetRex algorithm. "
                     "Performing Linear Scan"
                  << std::endl;
    if (verbose)
        std::cerr << "Narrowed Search to " << popcount(hit_vector) << " possible bins" << std::endl;
    if (read_file)
        std::cerr << "Bin Count: " << popcount(hit_vector) << "\t";
    if (popcount(hit_vector) != 0)
    {
        try
        {
            iter_disk_search(hit_vector, args.regex, index, args.ofile);
        }
        catch (std::exception const & e)
        {
            std::cerr << e.what() << '\n';
        }
    }
    double const t2 = now();
    if (verbose || read_file)
        std::cerr << "Query Time: " << (t2 - t1) << std::endl;
}

} // namespace

namespace
{

// `tetrex query --gpus N`: the single-query path on N devices of this node, one host thread and one shard per device
// (SURVEY.md §5/§8(e)).  Flat IBF: device i holds the 64-bin column blocks i of N of every row, the pruning flags are
// all-reduced and the candidate words all-gathered over NCCL inside the CUDA library (trx_comm_create_all); HIBF: every
// device holds a replica.  Every device reads, uploads and verifies only the FASTA files of its own bins; the threads'
// outputs are concatenated in shard order, which is the reference's -t 1 order.
void run_collection_multi_gpu(query_arguments & args, unsigned n)
{
    double const t1 = now();
    std::vector<trx_ctx_t *> ctxs(n, nullptr);
    std::vector<trx_comm_t *> comms(n, nullptr);
    auto cleanup = [&] {
        for (trx_comm_t * c : comms)
            if (c)
                trx_comm_destroy(c);
        for (trx_ctx_t * c : ctxs)
            if (c)
                trx_ctx_destroy(c);
    };
    try
    {
        for (unsigned i = 0; i < n; ++i)
            check(trx_ctx_create(static_cast<int>(i), nullptr, &ctxs[i]));
        check(trx_comm_create_all(ctxs.data(), static_cast<int>(n), comms.data()));
        std::vector<std::string> outputs(n), errors(n);
        std::vector<uint64_t> candidates(n, 0);
        std::vector<std::thread> threads;
        for (unsigned i = 0; i < n; ++i)
            threads.emplace_back([&, i] {
                try
                {
                    DeviceIndex index(ctxs[i]);
                    index.set_shard(i, n);
                    index.set_threads(args.t);
                    index.load(args.idx);
                    shard_hooks hooks;
                    hooks.comm = comms[i];
                    bitvector hit_vector(index.bin_words(), 0);
                    for (uint64_t b = 0; b < index.bin_count(); ++b)
                        hit_vector[b / 64] |= 1ull << (b % 64);
                    if (index.bin_count() > 1)
                    {
                        bitvector got = process_query(args.regex, index, nullptr, index.is_hibf() ? nullptr : &hooks, args.augment);
                        for (size_t w = 0; w < hit_vector.size(); ++w)
                            hit_vector[w] &= got[w];
                    }
                    candidates[i] = popcount(hit_vector);
                    if (candidates[i])
                    {
                        std::ostringstream os;
                        iter_disk_search(hit_vector, args.regex, index, os);
                        outputs[i] = os.str();
                    }
                }
                catch (std::exception const & e)
                {
                    errors[i] = e.what();
                }
            });
        for (std::thread & t : threads)
            t.join();
        for (unsigned i = 0; i < n; ++i)
            if (!errors[i].empty())
                throw error("device " + std::to_string(i) + ": " + errors[i]);
        if (args.verbose)
            std::cerr << "Narrowed Search to " << candidates[0] << " possible bins" << std::endl;
        std::ofstream file_out;
        std::ostream * out = &std::cout;
        if (args.ofile != "-")
        {
            file_out.open(args.ofile);
            if (!file_out)
                throw error("Failed to open output file: " + args.ofile);
            out = &file_out;
        }
        for (unsigned i = 0; i < n; ++i)
            *out << outputs[i];
        if (args.verbose)
            std::cerr << "Query Time: " << (now() - t1) << std::endl;
    }
    catch (...)
    {
        cleanup();
        throw;
    }
    cleanup();
}

} // namespace

void drive_query(query_arguments & args)
{
    if (args.regex == "-")
        std::cin >> args.regex; // src/query.cpp:479
    if (char const * e = getenv("TETREX_GPUS"); e && args.gpus <= 1)
        args.gpus = static_cast<unsigned>(std::max(1, atoi(e)));
    if (args.draw)
        throw error("-d (graph drawing) is outside this build: see DESIGN.md \"out of scope\"");
    {
        // load_params comes first in the reference too (src/query.cpp:480-481): an unreadable file is reported before
        // anything else happens
        uint8_t k;
        std::string molecule;
        bool is_hibf;
        read_index_params(args.idx, k, molecule, is_hibf);
    }
    if (args.gpus > 1 && args.read_file == "-" && args.conjunction == "-" && args.dibf == "-")
    {
        run_collection_multi_gpu(args, args.gpus);
        return;
    }
    trx_ctx_t * ctx = nullptr;
    check(trx_ctx_create(0, nullptr, &ctx));
    try
    {
        DeviceIndex index(ctx);
        index.set_threads(args.t);
        index.load(args.idx);
        // run_collection loads the DGram index named by -g (include/query.h:258-264); run_conjunction only would if
        // has_dibf were already set, which nothing does (:300-301) — so `-c` walks with the gap bounds 0..0 and no filter
        DGramIndex dibf(ctx);
        if (args.dibf != "-" && args.conjunction == "-")
        {
            dibf.load(args.dibf);
            index.attach_dgram(&dibf);
        }
        if (args.read_file != "-")
        {
            // read_regex_from_file + run_multiple_queries (src/query.cpp:372-395, include/query.h:330-340):
            // "id<TAB>motif" per line, matches of each go to "<id>.tsv"
            std::ifstream in(args.read_file);
            if (!in.is_open())
                throw error("Could not open file: " + args.read_file);
            std::string line;
            std::vector<std::pair<std::string, std::string>> motifs; // (id, motif)
            while (std::getline(in, line))
            {
                if (line.empty())
                    continue;
                std::istringstream iss(line);
                std::string id, motif;
                if (!std::getline(iss, id, '\t') || !std::getline(iss, motif, '\t'))
                    continue;
                motifs.emplace_back(id, motif);
            }
            // Peptide indexes without alphabet reduction verify the whole file of motifs in shared passes over the text
            // (K4 multi-pattern scan); everything else (DNA: two strands; reduced alphabets: mapped text) one by one.
            if (!index.is_dna() && index.reduction() == 0 && index.bin_count() > 1 && motifs.size() > 1)
                run_motif_batch(motifs, index, args.augment);
            else
                for (auto const & [id, motif] : motifs)
                {
                    args.regex = motif;
                    args.ofile = id + ".tsv";
                    std::cerr << id << "\t";
                    run_collection(args, true, index);
                }
        }
        else if (args.conjunction != "-")
            run_conjunction(args, index); // the reference wires -c up for peptide HIBFs only (src/query.cpp:463-466); here: any index
        else
            run_collection(args, false, index);
    }
    catch (...)
    {
        trx_ctx_destroy(ctx);
        throw;
    }
    trx_ctx_destroy(ctx);
}

void drive_index(index_arguments const & args)
{
    trx_ctx_t * ctx = nullptr;
    check(trx_ctx_create(0, nullptr, &ctx));
    try
    {
        DeviceIndex index(ctx);
        index.build(args);
        std::cerr << "Writing to disk... ";
        index.store(args.ofile + ".ibf");
        std::cerr << "DONE" << std::endl;
    }
    catch (...)
    {
        trx_ctx_destroy(ctx);
        throw;
    }
    trx_ctx_destroy(ctx);
}

void drive_inspection(std::string const & idx_path)
{
    // src/inspect_idx.cpp — reads the file only, no device needed
    std::cerr << "Reading Index from Disk... ";
    double const t1 = now();
    index_image img = read_index(idx_path);
    std::cerr << "DONE in " << now() - t1 << "s" << std::endl;
    bool const dna = img.molecule == "na";
    if (!img.is_hibf)
    {
        std::cout << "INDEX TYPE: IBF" << std::endl;
        std::cout << "BIN COUNT (BFs): " << img.flat.bins << std::endl;
        std::cout << "BIN SIZE (bits): " << img.flat.bin_size << std::endl;
    }
    else
    {
        std::cout << "INDEX TYPE: HIBF" << std::endl;
        std::cout << "FALSE POSITIVE RATE: " << std::fixed << std::setprecision(2) << img.fpr << std::endl;
    }
    std::cout << "HASH COUNT (hash functions): " << unsigned(img.hash_count) << std::endl;
    std::cout << "KMER LENGTH (" << (dna ? "bases" : "residues") << "): " << unsigned(img.k) << std::endl;
    // the reference streams the uint8_t reduction_ as a raw character here; the number is printed instead
    std::cout << "MOLECULE TYPE (alphabet): " << (dna ? "Nucleic Acid" : "Amino Acid") << " [REDUCTION=" << unsigned(img.reduction)
              << "]" << std::endl;
    std::cout << "ACID LIBRARY (filepaths):" << std::endl;
    for (auto const & path : img.acid_libs)
        std::cout << "\t- " << path << std::endl;
    std::cerr << "DONE" << std::endl;
}

} // namespace trxh
