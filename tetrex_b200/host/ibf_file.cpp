// ibf_file.cpp — reader/writer for the reference's on-disk index: a cereal 1.3.2 BinaryOutputArchive of
// TetrexIndex (include/index_base.h:160-195).  Layout: native little-endian, no padding, no
// class-version tags; string / vector<T> = uint64 size + elements (arithmetic T as one blob); std::array = raw
// blob; bool = 1 byte.  Field order follows the serialize() members cited next to each block.
#include <cstring>
#include <fstream>

#include "trx_host.hpp"

namespace trxh
{

namespace
{

struct writer
{
    std::ofstream os;
    explicit writer(std::string const & path) : os(path, std::ios::binary)
    {
        if (!os)
            throw error("Failed to open output file: " + path);
    }
    template <class T>
    void pod(T const & v)
    {
        os.write(reinterpret_cast<char const *>(&v), sizeof(T));
    }
    void str(std::string const & s)
    {
        pod<uint64_t>(s.size());
        os.write(s.data(), static_cast<std::streamsize>(s.size()));
    }
    void strs(std::vector<std::string> const & v)
    {
        pod<uint64_t>(v.size());
        for (auto const & s : v)
            str(s);
    }
    void u64s(std::vector<uint64_t> const & v)
    {
        pod<uint64_t>(v.size());
        os.write(reinterpret_cast<char const *>(v.data()), static_cast<std::streamsize>(v.size() * 8));
    }
};

struct reader
{
    std::ifstream is;
    std::string path;
    explicit reader(std::string const & p) : is(p, std::ios::binary), path(p)
    {
        if (!is)
            throw archive_error("Failed to read from input stream: cannot open " + p);
    }
    void raw(void * dst, size_t n)
    {
        is.read(static_cast<char *>(dst), static_cast<std::streamsize>(n));
        if (static_cast<size_t>(is.gcount()) != n)
            throw archive_error("Failed to read " + std::to_string(n) + " bytes from input stream! Read " +
                                std::to_string(is.gcount()));
    }
    template <class T>
    T pod()
    {
        T v{};
        raw(&v, sizeof(T));
        return v;
    }
    uint64_t size(uint64_t elem_bytes)
    {
        uint64_t n = pod<uint64_t>();
        if (n > (1ull << 40) / (elem_bytes ? elem_bytes : 1))
            throw archive_error("implausible container size in " + path + " (not a TetRex index?)");
        return n;
    }
    std::string str()
    {
        std::string s(size(1), '\0');
        raw(s.data(), s.size());
        return s;
    }
    std::vector<std::string> strs()
    {
        std::vector<std::string> v(size(8));
        for (auto & s : v)
            s = str();
        return v;
    }
    std::vector<uint64_t> u64s()
    {
        std::vector<uint64_t> v(size(8));
        raw(v.data(), v.size() * 8);
        return v;
    }
};

// seqan::hibf::interleaved_bloom_filter::serialize (lib/hibf/include/hibf/interleaved_bloom_filter.hpp:463-477)
void write_ibf(writer & w, ibf_image const & f)
{
    w.pod<uint32_t>(1); // version
    w.pod(f.bins);
    w.pod(f.technical_bins);
    w.pod(f.bin_size);
    w.pod(f.hash_shift);
    w.pod(f.bin_words);
    w.pod(f.hash_funs);
    // bit_vector::save (lib/hibf/include/hibf/misc/bit_vector.hpp:905-921): size in BITS, then the words
    w.pod<uint64_t>(f.technical_bins * f.bin_size);
    w.os.write(reinterpret_cast<char const *>(f.words.data()), static_cast<std::streamsize>(f.words.size() * 8));
    w.u64s(f.occupancy);
    w.pod<uint8_t>(f.track_occupancy ? 1 : 0);
}

ibf_image read_ibf(reader & r)
{
    ibf_image f;
    uint32_t version = r.pod<uint32_t>();
    if (version != 1)
        throw archive_error("unsupported interleaved_bloom_filter version " + std::to_string(version));
    f.bins = r.pod<uint64_t>();
    f.technical_bins = r.pod<uint64_t>();
    f.bin_size = r.pod<uint64_t>();
    f.hash_shift = r.pod<uint64_t>();
    f.bin_words = r.pod<uint64_t>();
    f.hash_funs = r.pod<uint64_t>();
    uint64_t bits = r.pod<uint64_t>();
    if (bits != f.technical_bins * f.bin_size || f.bin_words * 64 != f.technical_bins || bits > (1ull << 44))
        throw archive_error("inconsistent interleaved_bloom_filter geometry in " + r.path);
    f.words.resize((bits + 63) / 64);
    r.raw(f.words.data(), f.words.size() * 8);
    f.occupancy = r.u64s();
    f.track_occupancy = r.pod<uint8_t>() != 0;
    return f;
}

} // namespace

// DGramIndex::serialize (include/dGramIndex.h:305-309)
void write_dindex(dgram_image const & img, std::string const & path)
{
    writer w(path);
    w.pod(img.min_gap);
    w.pod(img.max_gap);
    w.pod(img.pad);
    w.pod(img.hash_count);
    w.pod(img.fpr);
    w.strs(img.bins);
    write_ibf(w, img.ibf);
    w.pod(img.bin_count);
    w.pod(img.hits_bits); // bit_vector::save: size in bits, then its words
    w.os.write(reinterpret_cast<char const *>(img.hits_words.data()), static_cast<std::streamsize>(img.hits_words.size() * 8));
    if (!w.os)
        throw error("Failed to write " + path);
}

dgram_image read_dindex(std::string const & path)
{
    reader r(path);
    dgram_image img;
    img.min_gap = r.pod<uint64_t>();
    img.max_gap = r.pod<uint64_t>();
    img.pad = r.pod<uint64_t>();
    img.hash_count = r.pod<uint8_t>();
    img.fpr = r.pod<float>();
    img.bins = r.strs();
    img.ibf = read_ibf(r);
    img.bin_count = r.pod<uint64_t>();
    img.hits_bits = r.pod<uint64_t>();
    if (img.hits_bits > (1ull << 40))
        throw archive_error("implausible container size in " + path + " (not a TetRex DGram index?)");
    img.hits_words.resize((img.hits_bits + 63) / 64);
    r.raw(img.hits_words.data(), img.hits_words.size() * 8);
    return img;
}

void write_index(index_image const & img, std::string const & path)
{
    writer w(path);
    // TetrexIndex::serialize (include/index_base.h:160-165)
    w.pod(img.k);
    w.str(img.molecule);
    w.pod<uint8_t>(img.is_hibf ? 1 : 0);
    w.strs(img.acid_libs);
    w.pod(img.reduction);
    if (!img.is_hibf)
    {
        // IBFIndex::serialize (include/index_ibf.h:152-156)
        w.pod(img.bin_count);
        w.pod(img.bin_size_member);
        w.pod(img.hash_count);
        w.strs(img.bin_names);
        write_ibf(w, img.flat);
    }
    else
    {
        // HIBFIndex::serialize (include/index_hibf.h:154-158)
        w.pod(img.bin_count);
        w.pod(img.fpr);
        w.pod(img.hash_count);
        w.strs(img.bin_names);
        // hierarchical_interleaved_bloom_filter::serialize (…/hierarchical_interleaved_bloom_filter.hpp:267-277)
        hibf_image const & h = img.hibf;
        w.pod<uint32_t>(1);
        w.pod(h.number_of_user_bins);
        w.pod<uint64_t>(h.ibf_vector.size());
        for (auto const & f : h.ibf_vector)
            write_ibf(w, f);
        w.pod<uint64_t>(h.next_ibf_id.size());
        for (auto const & v : h.next_ibf_id)
            w.u64s(v);
        w.pod<uint64_t>(h.prev_ibf_id.size());
        for (auto const & p : h.prev_ibf_id)
        {
            w.pod(p.first);
            w.pod(p.second);
        }
        w.pod<uint64_t>(h.ibf_bin_to_user_bin_id.size());
        for (auto const & v : h.ibf_bin_to_user_bin_id)
            w.u64s(v);
    }
    // MoleculeDecomposer::serialize (include/molecule_decomposer.h:118-122) then the concrete decomposer
    bool const dna = img.molecule == "na";
    w.pod(img.k);
    w.pod<uint8_t>(dna ? 2 : 5);          // lshift_
    w.pod<uint8_t>(dna ? 0b11 : 0b11111); // rmask_
    if (dna)
    {
        // NucleotideDecomposer::serialize (include/nucleotide_decomposer.h:123-127)
        w.pod(img.k);
        w.pod(img.reduction);
        w.pod<uint8_t>(static_cast<uint8_t>(2 * img.k - 2));
        w.pod<uint64_t>(img.k >= 32 ? ~0ull : ((1ull << (2 * img.k)) - 1));
    }
    else
    {
        // PeptideDecomposer::serialize (include/peptide_decomposer.h:301-305)
        uint8_t aamap[256], redmap[256];
        trx_peptide_maps(img.reduction, aamap, redmap);
        w.pod(img.k);
        w.pod(img.reduction);
        w.pod<uint8_t>(img.reduction == TRX_RED_BASE ? 20 : 10); // alphabet_size_
        w.pod<uint64_t>(img.k >= 32 ? ~0ull : ((1ull << (5 * img.k)) - 1));
        w.os.write(reinterpret_cast<char const *>(aamap), 256);
        w.os.write(reinterpret_cast<char const *>(redmap), 256);
    }
    if (!w.os)
        throw error("write error on " + path);
}

index_image read_index(std::string const & path)
{
    reader r(path);
    index_image img;
    img.k = r.pod<uint8_t>();
    img.molecule = r.str();
    img.is_hibf = r.pod<uint8_t>() != 0;
    if (img.molecule != "na" && img.molecule != "aa")
        throw archive_error(path + " is not a TetRex index (molecule tag '" + img.molecule + "')");
    img.acid_libs = r.strs();
    img.reduction = r.pod<uint8_t>();
    if (!img.is_hibf)
    {
        img.bin_count = r.pod<uint64_t>();
        img.bin_size_member = r.pod<uint64_t>();
        img.hash_count = r.pod<uint8_t>();
        img.bin_names = r.strs();
        img.flat = read_ibf(r);
    }
    else
    {
        img.bin_count = r.pod<uint64_t>();
        img.fpr = r.pod<float>();
        img.hash_count = r.pod<uint8_t>();
        img.bin_names = r.strs();
        hibf_image & h = img.hibf;
        uint32_t version = r.pod<uint32_t>();
        if (version != 1)
            throw archive_error("unsupported hierarchical_interleaved_bloom_filter version " + std::to_string(version));
        h.number_of_user_bins = r.pod<uint64_t>();
        uint64_t n = r.size(64);
        for (uint64_t i = 0; i < n; ++i)
            h.ibf_vector.push_back(read_ibf(r));
        n = r.size(8);
        for (uint64_t i = 0; i < n; ++i)
            h.next_ibf_id.push_back(r.u64s());
        n = r.size(16);
        for (uint64_t i = 0; i < n; ++i)
        {
            uint64_t a = r.pod<uint64_t>();
            uint64_t b = r.pod<uint64_t>();
            h.prev_ibf_id.emplace_back(a, b);
        }
        n = r.size(8);
        for (uint64_t i = 0; i < n; ++i)
            h.ibf_bin_to_user_bin_id.push_back(r.u64s());
    }
    // decomposer tail: validated, not stored (it is a pure function of k / molecule / reduction)
    uint8_t ksize = r.pod<uint8_t>();
    (void)r.pod<uint8_t>();
    (void)r.pod<uint8_t>();
    if (ksize != img.k)
        throw archive_error("decomposer k does not match the index k in " + path);
    return img;
}

void read_index_params(std::string const & path, uint8_t & k, std::string & molecule, bool & is_hibf)
{
    reader r(path);
    k = r.pod<uint8_t>();
    molecule = r.str();
    is_hibf = r.pod<uint8_t>() != 0;
}

} // namespace trxh
