// dgram.cpp — DGramIndex (include/dGramIndex.h, src/dGramIndex.cpp): the gapped 3+3-mer filter `tetrex query -a -g` probes
// when it jumps over a high-complexity stretch of a pattern.  The reference buffers every code of every bin in host
// vectors, sizes the filter by the largest buffer and emplaces the buffers one code at a time; here the residues go to
// the device once and two launches do the same: count per bin, then emplace (trx_dgram_count_dev / trx_dgram_emplace_dev).
#include <algorithm>
#include <filesystem>
#include <fstream>

#include "trx_host.hpp"

namespace trxh
{

DGramIndex::~DGramIndex()
{
    if (flat_)
        trx_ibf_destroy(flat_);
}

void DGramIndex::build(dindex_arguments const & args)
{
    if (args.max > TRX_DGRAM_MAX_GAP)
        throw error("upper gap bound " + std::to_string(args.max) + " exceeds " + std::to_string(TRX_DGRAM_MAX_GAP));
    img_ = dgram_image{};
    img_.min_gap = args.min;
    img_.max_gap = args.max;
    img_.pad = args.pad;
    img_.hash_count = args.hash_count;
    img_.fpr = args.fpr;
    img_.bins = args.acid_libs;
    img_.bin_count = img_.bins.size();
    // hits_(bc_, true) (:98): seqan::hibf::bit_vector(n, true) fills whole words, so the bits past the size are set too and
    // the archive carries them (tests/golden/tiny.dibf ends in ff..ff for two bins)
    img_.hits_bits = img_.bin_count;
    img_.hits_words.assign((img_.bin_count + 63) / 64, ~0ull);
    if (img_.bin_count == 0)
        throw error("The reference interleaved_bloom_filter cannot be built without bins (bin count 0)");

    // populate() (:103-111): add_fasta per bin — kseq over gzopen, every record through process_sequence
    std::string residues;
    std::vector<uint64_t> offsets{0};
    std::vector<uint32_t> rec_bin;
    for (size_t b = 0; b < img_.bins.size(); ++b)
        for (fasta_record & rec : read_fasta(img_.bins[b]))
        {
            residues += rec.seq;
            offsets.push_back(residues.size());
            rec_bin.push_back(static_cast<uint32_t>(b));
        }
    void *d_res = nullptr, *d_off = nullptr, *d_bin = nullptr, *d_cnt = nullptr;
    if (flat_)
        trx_ibf_destroy(flat_);
    flat_ = nullptr;
    try
    {
        check(trx_dev_alloc(ctx_, residues.size() + 64, &d_res));
        check(trx_dev_alloc(ctx_, offsets.size() * 8, &d_off));
        check(trx_dev_alloc(ctx_, std::max<size_t>(rec_bin.size(), 1) * 4, &d_bin));
        check(trx_dev_alloc(ctx_, img_.bin_count * 8, &d_cnt));
        check(trx_copy_h2d(ctx_, d_res, residues.data(), residues.size()));
        check(trx_copy_h2d(ctx_, d_off, offsets.data(), offsets.size() * 8));
        check(trx_copy_h2d(ctx_, d_bin, rec_bin.data(), rec_bin.size() * 4));
        uint32_t const lo = static_cast<uint32_t>(std::min<size_t>(args.min, UINT32_MAX)), hi = static_cast<uint32_t>(args.max);
        check(trx_dgram_count_dev(ctx_, static_cast<uint8_t *>(d_res), residues.size(), static_cast<uint64_t *>(d_off),
                                  static_cast<uint32_t *>(d_bin), rec_bin.size(), lo, hi, img_.bin_count,
                                  static_cast<uint64_t *>(d_cnt)));
        codes_per_bin_.assign(img_.bin_count, 0);
        check(trx_copy_d2h(ctx_, codes_per_bin_.data(), d_cnt, img_.bin_count * 8));
        // init_ibf (:144-154): find_largest_bin → compute_bitcount (float log like IBFIndex::compute_bitcount) → emplace
        uint64_t max_bin = 0;
        for (uint64_t c : codes_per_bin_)
            max_bin = std::max(max_bin, c);
        uint64_t const bin_size = trx_compute_bitcount(max_bin, args.fpr);
        if (bin_size == 0) // seqan::hibf::interleaved_bloom_filter's constructor rejects it
            throw error("The bin size must be > 0 (no gapped k-mer in any bin)");
        if (args.hash_count == 0 || args.hash_count > 5)
            throw error("The number of hash functions must be > 0 and <= 5.");
        check(trx_ibf_create(ctx_, img_.bin_count, bin_size, args.hash_count, 0, 1, &flat_));
        check(trx_dgram_emplace_dev(flat_, static_cast<uint8_t *>(d_res), residues.size(), static_cast<uint64_t *>(d_off),
                                    static_cast<uint32_t *>(d_bin), rec_bin.size(), lo, hi));
        uint64_t geo[10];
        check(trx_ibf_geometry(flat_, geo));
        ibf_image & f = img_.ibf;
        f.bins = geo[0];
        f.technical_bins = geo[1];
        f.bin_size = geo[2];
        f.hash_shift = geo[3];
        f.bin_words = geo[4];
        f.hash_funs = geo[5];
        f.words.resize(f.bin_words * f.bin_size);
        check(trx_ibf_download(flat_, f.words.data(), f.words.size()));
        f.occupancy.assign(f.technical_bins, 0);
        f.track_occupancy = false;
    }
    catch (...)
    {
        trx_dev_free(ctx_, d_res);
        trx_dev_free(ctx_, d_off);
        trx_dev_free(ctx_, d_bin);
        trx_dev_free(ctx_, d_cnt);
        throw;
    }
    trx_dev_free(ctx_, d_res);
    trx_dev_free(ctx_, d_off);
    trx_dev_free(ctx_, d_bin);
    trx_dev_free(ctx_, d_cnt);
}

void DGramIndex::store(std::string const & path) const
{
    write_dindex(img_, path);
}

void DGramIndex::load(std::string const & path)
{
    img_ = read_dindex(path);
    if (flat_)
        trx_ibf_destroy(flat_);
    flat_ = nullptr;
    ibf_image const & f = img_.ibf;
    check(trx_ibf_create(ctx_, f.bins, f.bin_size, static_cast<unsigned>(f.hash_funs), 0, 1, &flat_));
    check(trx_ibf_upload(flat_, f.words.data(), f.words.size()));
}

void DeviceIndex::attach_dgram(DGramIndex const * d)
{
    // `path_ &= dibf_->query(dgram)` (include/otf_collector.h:235) ANDs two bit vectors of bc_ and getBinCount() bits: in
    // the reference a mismatch is undefined behaviour, here it is an error
    if (d && d->bin_count() != bin_count())
        throw error("the DGram index has " + std::to_string(d->bin_count()) + " bins, the index " + std::to_string(bin_count()) +
                    ": they must be built from the same files");
    if (d && !is_hibf() && shard_count() > 1)
        throw error("a DGram index cannot be attached to a column shard of a flat index");
    dgram_ = d;
}

// drive_dindex (src/dGramIndex.cpp:19-38): `.lst` arguments are lists of files, one per line; everything else is one bin
void drive_dindex(dindex_arguments const & cmd_args)
{
    dindex_arguments args = cmd_args;
    args.acid_libs.clear();
    for (std::string const & file : cmd_args.acid_libs)
    {
        if (std::filesystem::path(file).extension() == ".lst")
        {
            std::ifstream fin{file};
            if (!fin.good() || !fin.is_open())
                throw error("Could not open file " + file + " for reading.");
            std::string line;
            while (std::getline(fin, line))
                args.acid_libs.emplace_back(line);
        }
        else
            args.acid_libs.push_back(std::filesystem::absolute(file).string());
    }
    trx_ctx_t * ctx = nullptr;
    check(trx_ctx_create(0, nullptr, &ctx));
    try
    {
        DGramIndex dindex(ctx);
        dindex.build(args);
        dindex.store(args.ofile);
    }
    catch (...)
    {
        trx_ctx_destroy(ctx);
        throw;
    }
    trx_ctx_destroy(ctx);
}

} // namespace trxh
