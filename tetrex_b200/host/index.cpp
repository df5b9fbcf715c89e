// index.cpp — DeviceIndex: TetrexIndex<flavor, molecule> (include/index_base.h) with the bit matrix in HBM.
//
// build(): create_{ibf,hibf}_{dna,aa}_index + populate_index (src/index_base.cpp:25-118, include/index_ibf.h:
// 88-131, include/index_hibf.h:86-126).  The reference buffers every k-mer of every bin on the host
// (user_bin_data_) and inserts them one by one; here the FASTA bytes go to the device once and K1
// (trx_emplace_records) derives k-mers, hashes and sets bits in one pass — nothing is buffered per k-mer.
// store()/load(): store_ibf / load_ibf through the cereal-compatible file of ibf_file.cpp.
#include <exception>
#include <thread>
#include <mutex>
#include <atomic>
#include <algorithm>
#include <cmath>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <numeric>

#include "trx_host.hpp"

namespace trxh
{

void check(int rc)
{
    if (rc != TRX_OK)
        throw error(std::string(trx_last_error()));
}

DeviceIndex::~DeviceIndex()
{
    if (text_ && text_->text)
        trx_text_destroy(text_->text);
    if (scratch_.paths)
        trx_paths_destroy(scratch_.paths);
    if (scratch_.d_kmers)
        trx_dev_free(ctx_, scratch_.d_kmers);
    if (scratch_.d_hits)
        trx_dev_free(ctx_, scratch_.d_hits);
    if (flat_)
        trx_ibf_destroy(flat_);
    if (hibf_)
        trx_hibf_destroy(hibf_);
}

void DeviceIndex::init_maps()
{
    trx_peptide_maps(img_.reduction, aamap_, redmap_);
}

void DeviceIndex::describe_only(std::string const & molecule, uint8_t k, uint8_t reduction, uint64_t bins)
{
    if (molecule != "na" && molecule != "aa")
        throw error("molecule must be \"na\" or \"aa\"");
    img_ = index_image{};
    img_.molecule = molecule;
    img_.k = k;
    img_.reduction = reduction;
    img_.bin_count = bins;
    init_maps();
}

uint64_t DeviceIndex::update_kmer(int symbol, uint64_t & kmer) const
{
    unsigned const k = img_.k;
    if (is_dna())
    {
        // nucleotide_decomposer.h:113-121
        uint64_t const mask = k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1);
        uint64_t const fb = (static_cast<uint64_t>(symbol) >> 1) & 3;
        uint64_t const forward = ((kmer << 2) & mask) | fb;
        kmer = forward;
        // revComplement (:44-79): reversed order of the k 2-bit codes, each ^ 2 …
        uint64_t rc = 0, x = forward;
        for (unsigned i = 0; i < k; ++i)
        {
            rc = (rc << 2) | ((x & 3) ^ 2);
            x >>= 2;
        }
        // … except that `kmer2 >>= 4` (:60) is an arithmetic shift: with bit 63 set (k = 32) the two leading
        // bases are lost and the low nibble of the result is 0
        if (forward >> 63)
            rc &= ~0xFull;
        return forward <= rc ? forward : rc;
    }
    // peptide_decomposer.h:293-299
    uint64_t const mask = k >= 32 ? ~0ull : ((1ull << (5 * k)) - 1);
    uint64_t const forward = ((kmer << 5) & mask) | aamap_[static_cast<uint8_t>(symbol)];
    kmer = forward;
    return forward;
}

namespace
{

struct library
{
    std::string residues;           // all records of all bins, bins ascending
    std::vector<uint64_t> offsets;  // n_records + 1
    std::vector<uint32_t> rec_bin;  // user bin of each record
    std::vector<std::string> names; // ">" + name
    std::vector<uint64_t> bin_emits; // k-mers each bin hands to emplace (duplicates counted, like user_bin_data_)
    uint64_t indexed{};             // records of length >= k
};

// bins that are not wanted are skipped (they keep their ids and simply have no records): a verify shard, or a
// one-shot query that only needs its candidate bins
// `threads` (the command line's -t; the reference spends it on OpenMP over bins, include/query.h:113): the files are
// parsed by that many host threads — inflate + kseq is the cost of a one-shot query — and joined in bin order.
// 0 or 1 = as many as the machine has, at most 16.
library read_library(std::vector<std::string> const & files, int molecule, unsigned k, bool announce_short,
                     std::vector<bool> const * want = nullptr, unsigned threads = 0)
{
    library lib;
    lib.offsets.push_back(0);
    lib.bin_emits.assign(files.size(), 0);
    std::vector<size_t> todo;
    for (size_t b = 0; b < files.size(); ++b)
        if (!want || (*want)[b])
            todo.push_back(b);
    std::vector<std::vector<fasta_record>> parsed(todo.size());
    {
        unsigned workers = threads > 1 ? threads : std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
        workers = static_cast<unsigned>(std::min<size_t>(workers, todo.size()));
        std::atomic<size_t> next{0};
        std::exception_ptr failure;
        std::mutex failure_lock;
        auto work = [&] {
            for (size_t i = next++; i < todo.size(); i = next++)
            {
                try
                {
                    parsed[i] = read_fasta(files[todo[i]]);
                }
                catch (...)
                {
                    std::lock_guard<std::mutex> hold(failure_lock);
                    if (!failure)
                        failure = std::current_exception();
                    next = todo.size();
                }
            }
        };
        if (workers <= 1)
            work();
        else
        {
            std::vector<std::thread> pool;
            for (unsigned w = 0; w < workers; ++w)
                pool.emplace_back(work);
            for (std::thread & t : pool)
                t.join();
        }
        if (failure)
            std::rethrow_exception(failure);
    }
    for (size_t i = 0; i < todo.size(); ++i)
    {
        size_t const b = todo[i];
        for (fasta_record & rec : parsed[i])
        {
            if (rec.seq.size() < k)
            {
                if (announce_short)
                    std::cerr << "RECORD TOO SHORT " << rec.comment << std::endl; // include/index_ibf.h:116
            }
            else
            {
                ++lib.indexed;
                lib.bin_emits[b] += trx_emit_count(molecule, k, rec.seq.size());
            }
            lib.residues += rec.seq;
            lib.offsets.push_back(lib.residues.size());
            lib.rec_bin.push_back(static_cast<uint32_t>(b));
            lib.names.push_back(">" + rec.name);
        }
        std::vector<fasta_record>().swap(parsed[i]);
    }
    return lib;
}

struct device_library
{
    trx_ctx_t * ctx;
    void *residues{nullptr}, *offsets{nullptr}, *rec_bin{nullptr};
    device_library(trx_ctx_t * c, library const & lib) : ctx(c)
    {
        check(trx_dev_alloc(ctx, lib.residues.size() + 64, &residues));
        check(trx_dev_alloc(ctx, lib.offsets.size() * 8, &offsets));
        check(trx_dev_alloc(ctx, std::max<size_t>(lib.rec_bin.size(), 1) * 4, &rec_bin));
        check(trx_copy_h2d(ctx, residues, lib.residues.data(), lib.residues.size()));
    }
    ~device_library()
    {
        trx_dev_free(ctx, residues);
        trx_dev_free(ctx, offsets);
        trx_dev_free(ctx, rec_bin);
    }
};

// One technical IBF built on the device from records [r0, r1) of the library (a contiguous run: bins are read in
// ascending order); bin_of_record[i] = technical bin of record r0 + i.
ibf_image build_one(trx_ctx_t * ctx, device_library & dev, library const & lib, uint64_t r0, uint64_t r1,
                    std::vector<uint32_t> const & bin_of_record, uint64_t bins, uint64_t bin_size, unsigned hash_funs,
                    int molecule, unsigned k, int reduction, trx_ibf_t ** keep)
{
    trx_ibf_t * ibf = nullptr;
    check(trx_ibf_create(ctx, bins, bin_size, hash_funs, 0, 1, &ibf));
    try
    {
        uint64_t const base = lib.offsets[r0];
        std::vector<uint64_t> rebased(r1 - r0 + 1);
        for (uint64_t r = r0; r <= r1; ++r)
            rebased[r - r0] = lib.offsets[r] - base;
        check(trx_copy_h2d(ctx, dev.offsets, rebased.data(), rebased.size() * 8));
        check(trx_copy_h2d(ctx, dev.rec_bin, bin_of_record.data(), (r1 - r0) * 4));
        check(trx_emplace_records_dev(ibf, molecule, k, reduction, static_cast<uint8_t *>(dev.residues) + base,
                                      rebased.back(), static_cast<uint64_t *>(dev.offsets),
                                      static_cast<uint32_t *>(dev.rec_bin), r1 - r0));
        uint64_t geo[10];
        check(trx_ibf_geometry(ibf, geo));
        ibf_image f;
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
        if (keep)
            *keep = ibf;
        else
            trx_ibf_destroy(ibf);
        return f;
    }
    catch (...)
    {
        trx_ibf_destroy(ibf);
        throw;
    }
}

} // namespace

void DeviceIndex::build(index_arguments const & args)
{
    // drive_index (src/index_base.cpp:73-118): .lst files expand to their lines, other paths are made absolute
    std::vector<std::string> files;
    for (std::string const & f : args.acid_libs)
    {
        std::filesystem::path p{f};
        if (p.extension() == ".lst")
        {
            std::ifstream fin{p};
            if (!fin.good())
                throw error("Could not open file " + p.string() + " for reading.");
            std::string line;
            while (std::getline(fin, line))
                files.push_back(line);
        }
        else
            files.push_back(std::filesystem::absolute(p).string());
    }
    if (files.empty())
        throw error("no input libraries");
    static std::map<std::string, uint8_t> const reductions{{"murphy", 1}, {"li", 2}, {"None", 0}};
    auto red = reductions.find(args.reduction);
    if (red == reductions.end())
        throw error("unknown reduction '" + args.reduction + "' (murphy, li)");

    img_ = index_image{};
    img_.k = args.k;
    img_.molecule = args.dna ? "na" : "aa";
    img_.is_hibf = !args.idx;
    img_.acid_libs = files;
    img_.reduction = red->second;
    img_.bin_count = files.size();
    img_.bin_size_member = 0; // the reference never sets the member it serialises
    img_.fpr = args.fpr;
    img_.hash_count = args.hash_count;
    img_.bin_names = files;
    init_maps();

    int const molecule = args.dna ? TRX_MOL_NA : TRX_MOL_AA;
    library lib = read_library(files, molecule, args.k, true, nullptr, args.t);
    seq_count_ = lib.indexed;
    device_library dev(ctx_, lib);

    if (!img_.is_hibf)
    {
        // IBFIndex::init_ibf (include/index_ibf.h:88-99): bin_size from the largest bin's emit count
        uint64_t const largest = *std::max_element(lib.bin_emits.begin(), lib.bin_emits.end());
        uint64_t const bin_size = trx_compute_bitcount(largest, args.fpr);
        if (bin_size == 0)
            throw error("The bin size must be > 0 (no record of length >= k in any bin)");
        img_.flat = build_one(ctx_, dev, lib, 0, lib.rec_bin.size(), lib.rec_bin, files.size(), bin_size, args.hash_count,
                              molecule, args.k, img_.reduction, &flat_);
    }
    else if (!(getenv("TRX_HIBF_LAYOUT") && std::string(getenv("TRX_HIBF_LAYOUT")) == "simple"))
    {
        // HIBFIndex::populate_index (include/index_hibf.h:86-130): every k-mer of every bin, in decomposition order, goes to
        // seqan::hibf — sketches, layout, hierarchical build — restated in hibf_layout.cpp / hibf_build.cpp.  The k-mers
        // are decomposed on the device (ordered emission), the sets and the layout are host work like in the reference,
        // every bit is set by K1.
        uint64_t const n_records = lib.rec_bin.size();
        std::vector<uint64_t> out_offsets(n_records + 1, 0);
        uint64_t total = 0;
        for (uint64_t r = 0; r < n_records; ++r)
            total += trx_emit_count(molecule, args.k, lib.offsets[r + 1] - lib.offsets[r]);
        std::vector<uint64_t> values(total);
        check(trx_decompose_records(ctx_, molecule, args.k, img_.reduction, reinterpret_cast<uint8_t const *>(lib.residues.data()),
                                    lib.offsets.data(), n_records, values.data(), values.size(), out_offsets.data()));
        std::vector<uint64_t> bin_offsets(files.size() + 1, total);
        {
            uint64_t r = 0;
            for (uint64_t b = 0; b < files.size(); ++b)
            {
                while (r < n_records && lib.rec_bin[r] < b)
                    ++r;
                bin_offsets[b] = r < n_records ? out_offsets[r] : total;
            }
        }
        // compute_sketches on the device: K1's decomposition with a HyperLogLog sink (N3)
        std::vector<uint8_t> registers(files.size() << 12);
        {
            void * d_regs = nullptr;
            check(trx_dev_alloc(ctx_, registers.size(), &d_regs));
            check(trx_copy_h2d(ctx_, dev.offsets, lib.offsets.data(), lib.offsets.size() * 8));
            check(trx_copy_h2d(ctx_, dev.rec_bin, lib.rec_bin.data(), lib.rec_bin.size() * 4));
            int const rc = trx_hll_sketch_records_dev(ctx_, molecule, args.k, img_.reduction, static_cast<uint8_t *>(dev.residues),
                                                      lib.residues.size(), static_cast<uint64_t *>(dev.offsets),
                                                      static_cast<uint32_t *>(dev.rec_bin), n_records, files.size(), 12,
                                                      static_cast<uint8_t *>(d_regs));
            int const rc2 = rc == TRX_OK ? trx_copy_d2h(ctx_, registers.data(), d_regs, registers.size()) : rc;
            trx_dev_free(ctx_, d_regs);
            check(rc2);
        }
        img_.hibf = build_hibf_like_seqan(ctx_, values, bin_offsets, args.hash_count, args.fpr, args.t, &registers);
        upload();
    }
    else
    {
        // TRX_HIBF_LAYOUT=simple: the plain two-level layout of round 1 (no sketches, no split bins) — T = 64*ceil(sqrt(n)/64)
        // technical bins on top; with n <= T every user bin gets its own technical bin, otherwise consecutive user bins share
        // a merged bin that owns a lower-level IBF.  A valid seqan archive whose false positives differ from the reference's.
        uint64_t const n = files.size();
        uint64_t const T = 64 * static_cast<uint64_t>(std::ceil(std::sqrt(static_cast<double>(n)) / 64.0));
        hibf_image & h = img_.hibf;
        h.number_of_user_bins = n;
        uint64_t const merged = UINT64_MAX;
        if (n <= T)
        {
            uint64_t const largest = *std::max_element(lib.bin_emits.begin(), lib.bin_emits.end());
            uint64_t const bin_size = std::max<uint64_t>(trx_compute_bitcount(largest, args.fpr), 64);
            h.ibf_vector.push_back(build_one(ctx_, dev, lib, 0, lib.rec_bin.size(), lib.rec_bin, n, bin_size, args.hash_count,
                                             molecule, args.k, img_.reduction, nullptr));
            h.next_ibf_id.emplace_back(n, 0);
            std::vector<uint64_t> ub(n);
            std::iota(ub.begin(), ub.end(), 0);
            h.ibf_bin_to_user_bin_id.push_back(ub);
            h.prev_ibf_id.emplace_back(0, 0);
        }
        else
        {
            uint64_t const groups = std::min<uint64_t>(T, n);
            auto group_of = [&](uint64_t b) { return b * groups / n; };
            std::vector<uint64_t> group_emits(groups, 0);
            for (uint64_t b = 0; b < n; ++b)
                group_emits[group_of(b)] += lib.bin_emits[b];
            uint64_t const top_size =
                std::max<uint64_t>(trx_compute_bitcount(*std::max_element(group_emits.begin(), group_emits.end()), args.fpr), 64);
            std::vector<uint32_t> map(lib.rec_bin.size());
            for (size_t r = 0; r < map.size(); ++r)
                map[r] = static_cast<uint32_t>(group_of(lib.rec_bin[r]));
            h.ibf_vector.push_back(build_one(ctx_, dev, lib, 0, map.size(), map, groups, top_size, args.hash_count, molecule,
                                             args.k, img_.reduction, nullptr));
            h.next_ibf_id.emplace_back();
            h.ibf_bin_to_user_bin_id.emplace_back(groups, merged);
            h.prev_ibf_id.emplace_back(0, 0);
            for (uint64_t g = 0; g < groups; ++g)
            {
                uint64_t first = n, last = 0, largest = 0;
                for (uint64_t b = 0; b < n; ++b)
                    if (group_of(b) == g)
                    {
                        first = std::min(first, b);
                        last = b;
                        largest = std::max(largest, lib.bin_emits[b]);
                    }
                uint64_t const members = last - first + 1;
                uint64_t const r0 = std::lower_bound(lib.rec_bin.begin(), lib.rec_bin.end(), first) - lib.rec_bin.begin();
                uint64_t const r1 = std::upper_bound(lib.rec_bin.begin(), lib.rec_bin.end(), last) - lib.rec_bin.begin();
                std::vector<uint32_t> sub(r1 - r0);
                for (uint64_t r = r0; r < r1; ++r)
                    sub[r - r0] = static_cast<uint32_t>(lib.rec_bin[r] - first);
                uint64_t const size = std::max<uint64_t>(trx_compute_bitcount(largest, args.fpr), 64);
                h.ibf_vector.push_back(build_one(ctx_, dev, lib, r0, r1, sub, members, size, args.hash_count, molecule, args.k,
                                                 img_.reduction, nullptr));
                uint64_t const id = h.ibf_vector.size() - 1;
                h.next_ibf_id[0].push_back(id);
                h.next_ibf_id.emplace_back(members, id);
                std::vector<uint64_t> ub(members);
                std::iota(ub.begin(), ub.end(), first);
                h.ibf_bin_to_user_bin_id.push_back(ub);
                h.prev_ibf_id.emplace_back(0, g);
            }
        }
        upload();
    }
    // text for verify is what was just read
    text_ = std::make_unique<text_store>();
    text_->names = std::move(lib.names);
    text_->offsets = std::move(lib.offsets);
    text_->rec_bin = std::move(lib.rec_bin);
    text_->residues = std::move(lib.residues);
    text_->loaded.assign(files.size(), true);
    std::cerr << "Indexed " << seq_count_ << " sequences across " << img_.bin_count << " bins." << std::endl;
    if (img_.bin_count == 1)
        std::cerr << "[WARNING] The indexed reference library was not split into bins. The TetRex runtime will be "
                     "significantly slower."
                  << std::endl;
}

void DeviceIndex::upload()
{
    if (!img_.is_hibf)
    {
        ibf_image const & f = img_.flat;
        if (flat_)
            trx_ibf_destroy(flat_), flat_ = nullptr;
        if (shard_count_ > 1 && f.bin_words % shard_count_ != 0)
            throw error("cannot split " + std::to_string(f.bin_words) + " 64-bin columns over " + std::to_string(shard_count_) +
                        " shards (the shard count must divide the column count)");
        check(trx_ibf_create(ctx_, f.bins, f.bin_size, static_cast<uint32_t>(f.hash_funs), shard_index_, shard_count_, &flat_));
        check(trx_ibf_upload(flat_, f.words.data(), f.words.size()));
        return;
    }
    hibf_image const & h = img_.hibf;
    size_t const n = h.ibf_vector.size();
    std::vector<uint64_t> bins(n), bin_size(n), hash_funs(n), table_len(n);
    std::vector<uint64_t const *> words(n), next(n), user(n);
    for (size_t i = 0; i < n; ++i)
    {
        bins[i] = h.ibf_vector[i].bins;
        bin_size[i] = h.ibf_vector[i].bin_size;
        hash_funs[i] = h.ibf_vector[i].hash_funs;
        words[i] = h.ibf_vector[i].words.data();
        if (h.next_ibf_id[i].size() != h.ibf_bin_to_user_bin_id[i].size())
            throw archive_error("HIBF tables of different length");
        next[i] = h.next_ibf_id[i].data();
        user[i] = h.ibf_bin_to_user_bin_id[i].data();
        table_len[i] = h.next_ibf_id[i].size();
    }
    trx_hibf_desc desc{};
    desc.n_ibfs = n;
    desc.n_user_bins = h.number_of_user_bins;
    desc.bins = bins.data();
    desc.bin_size = bin_size.data();
    desc.hash_funs = hash_funs.data();
    desc.words = words.data();
    desc.next_ibf = next.data();
    desc.user_bin = user.data();
    desc.table_len = table_len.data();
    if (hibf_)
        trx_hibf_destroy(hibf_), hibf_ = nullptr;
    check(trx_hibf_upload(ctx_, &desc, &hibf_));
}

void DeviceIndex::store(std::string const & path) const
{
    write_index(img_, path);
}

void DeviceIndex::load(std::string const & path)
{
    img_ = read_index(path);
    init_maps();
    upload();
}

void DeviceIndex::set_shard(uint32_t index, uint32_t count)
{
    if (count == 0 || index >= count)
        throw error("shard " + std::to_string(index) + " of " + std::to_string(count) + " does not exist");
    if (flat_ || hibf_ || text_)
        throw error("set_shard must be called before load()");
    shard_index_ = index;
    shard_count_ = count;
}

uint64_t DeviceIndex::shard_word_begin() const
{
    // balanced contiguous split of the 64-bin words (equal parts whenever count divides, which a flat IBF requires)
    uint64_t const words = bin_words(), base = words / shard_count_, extra = words % shard_count_;
    return shard_index_ * base + std::min<uint64_t>(shard_index_, extra);
}

uint64_t DeviceIndex::shard_words() const
{
    uint64_t const words = bin_words(), base = words / shard_count_, extra = words % shard_count_;
    return base + (shard_index_ < extra ? 1 : 0);
}

DeviceIndex::text_store const & DeviceIndex::text(bitvector const & cand)
{
    // iter_disk_search opens acid_libs_[bin] of the CANDIDATE bins only, per query (include/query.h:126-131,171-177):
    // a moved file of a bin that is not a candidate is not an error, and a one-shot `tetrex query` must not pay for
    // parsing and uploading the whole collection.  The store keeps what it has loaded; it grows to the union when a
    // later query needs more, and takes every bin of the shard once a query asks for more than a quarter of them (a
    // resident process ends up with the whole text in HBM, where K4 wants it).
    size_t const n_bins = img_.acid_libs.size();
    size_t const first_bin = shard_count_ > 1 ? shard_word_begin() * 64 : 0;
    size_t const last_bin = std::min<size_t>(n_bins, shard_count_ > 1 ? (shard_word_begin() + shard_words()) * 64 : n_bins);
    auto wanted = [&](size_t b) { return b / 64 < cand.size() && ((cand[b / 64] >> (b % 64)) & 1); };
    bool missing = !text_;
    size_t n_wanted = 0;
    for (size_t b = first_bin; b < last_bin; ++b)
        if (wanted(b))
        {
            ++n_wanted;
            if (text_ && !text_->loaded[b])
                missing = true;
        }
    if (missing)
    {
        std::vector<bool> load(n_bins, false);
        bool const everything = preload_all_ || n_wanted * 4 > last_bin - first_bin;
        for (size_t b = first_bin; b < last_bin; ++b)
            load[b] = everything || wanted(b) || (text_ && text_->loaded[b]);
        for (size_t b = first_bin; b < last_bin; ++b)
            if (load[b] && (everything ? wanted(b) : true) && !std::filesystem::exists(img_.acid_libs[b]))
                throw error("File not found. Did you move/rename an indexed file?");
        if (everything) // bins nobody asked for: a missing file only means that bin has no text yet
            for (size_t b = first_bin; b < last_bin; ++b)
                if (load[b] && !wanted(b) && !std::filesystem::exists(img_.acid_libs[b]))
                    load[b] = false;
        library lib = read_library(img_.acid_libs, is_dna() ? TRX_MOL_NA : TRX_MOL_AA, img_.k, false, &load, threads_);
        if (text_ && text_->text)
            trx_text_destroy(text_->text);
        text_ = std::make_unique<text_store>();
        text_->names = std::move(lib.names);
        text_->offsets = std::move(lib.offsets);
        text_->rec_bin = std::move(lib.rec_bin);
        text_->residues = std::move(lib.residues);
        text_->loaded = std::move(load);
    }
    if (!text_->text)
        check(trx_text_upload(ctx_, reinterpret_cast<uint8_t const *>(text_->residues.data()), text_->offsets.data(),
                              text_->rec_bin.data(), text_->rec_bin.size(), img_.acid_libs.size(), &text_->text));
    return *text_;
}

} // namespace trxh
