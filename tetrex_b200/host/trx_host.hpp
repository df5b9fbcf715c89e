// trx_host.hpp — host C++ layer of tetrex_b200: the reference's C++ API for the index-probe + verify path
// (TetrexIndex, process_query, iter_disk_search, drive_index, drive_query — /root/reference include/index_base.h,
// include/query.h) re-implemented on top of the C ABI in include/tetrex_b200.h.  Every hot loop is a C-ABI call
// into libtetrex_cuda.so; this layer only parses files and regexes, plans batches and formats output.
#pragma once

#include <cstdint>
#include <functional>
#include <iosfwd>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/tetrex_b200.h"
#include "rh_order.hpp"

namespace trxh
{

struct error : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

// thrown for unreadable / truncated index files; plays the role of cereal::Exception (src/main.cpp:54-58)
struct archive_error : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

void check(int rc); // throws trxh::error with trx_last_error() when rc != TRX_OK

// ------------------------------------------------------------------------------------------ arguments
// include/arg_parse.h:10-70 — same names and defaults; molecule / ibf are derived AFTER parsing here (the
// reference freezes them before parsing, SURVEY.md F3, so its -n / -i flags are dead).
struct index_arguments
{
    uint8_t k = 6;
    float fpr = 0.05f;
    uint8_t hash_count = 3;
    uint8_t t = 1;
    bool dna = false;  // -n
    bool idx = false;  // -i : flat IBF instead of HIBF
    std::string reduction = "None";
    std::string ofile;
    std::vector<std::string> acid_libs;
};

// include/arg_parse.h:95-123 (`tetrex dindex`): -l / -u bound the gap; hash count, fpr and pad have no option letter
struct dindex_arguments
{
    size_t min = 3;
    size_t max = 21;
    size_t pad = 1;
    uint8_t hash_count = 3;
    float fpr = 0.05f;
    std::string ofile;
    std::vector<std::string> acid_libs;
};

struct query_arguments
{
    uint8_t t = 1;
    bool verbose = false;
    bool draw = false;
    bool augment = false;
    std::string idx;
    std::string regex;
    unsigned gpus = 1; // --gpus N / TETREX_GPUS: one shard of the index per device of this node (not a reference flag)
    std::string read_file = "-";
    std::string conjunction = "-";
    std::string dibf = "-";
    std::string ofile = "-";
};

// --------------------------------------------------------------------------------------------- index file
// One technical IBF as serialised by seqan::hibf::interleaved_bloom_filter.
struct ibf_image
{
    uint64_t bins{}, technical_bins{}, bin_size{}, hash_shift{}, bin_words{}, hash_funs{};
    std::vector<uint64_t> words;     // ceil(technical_bins*bin_size/64)
    std::vector<uint64_t> occupancy; // technical_bins zeros on every TetRex path
    bool track_occupancy{false};
};

struct hibf_image
{
    uint64_t number_of_user_bins{};
    std::vector<ibf_image> ibf_vector;
    std::vector<std::vector<uint64_t>> next_ibf_id;
    std::vector<std::pair<uint64_t, uint64_t>> prev_ibf_id;
    std::vector<std::vector<uint64_t>> ibf_bin_to_user_bin_id;
};

// TetrexIndex as it lies on disk (include/index_base.h:160-165 + index_ibf.h:152-156 / index_hibf.h:154-158
// + the decomposer tail, molecule_decomposer.h:118-122)
struct index_image
{
    uint8_t k{};
    std::string molecule;          // "na" / "aa"
    bool is_hibf{};
    std::vector<std::string> acid_libs;
    uint8_t reduction{};           // 0 none, 1 murphy, 2 li
    uint64_t bin_count{};
    uint64_t bin_size_member{};    // flat only; the reference always writes 0 here
    float fpr{0.05f};              // HIBF only
    uint8_t hash_count{};
    std::vector<std::string> bin_names; // tech_bins_ / user_bins_
    ibf_image flat;
    hibf_image hibf;
};

// DGramIndex as it lies on disk (include/dGramIndex.h:305-309): archive(min_gap_, max_gap_, pad_, hc_, fpr_, bins_, dibf_,
// bc_, hits_)
struct dgram_image
{
    uint64_t min_gap{}, max_gap{}, pad{};
    uint8_t hash_count{};
    float fpr{};
    std::vector<std::string> bins; // one FASTA file per bin
    ibf_image ibf;
    uint64_t bin_count{1};
    uint64_t hits_bits{};             // hits_: bitvector(bc_, true) as it was when the index was saved
    std::vector<uint64_t> hits_words;
};
void write_dindex(dgram_image const & img, std::string const & path);
dgram_image read_dindex(std::string const & path);

// seqan::hibf::hierarchical_interleaved_bloom_filter{config} (lib/hibf/src/hierarchical_interleaved_bloom_filter.cpp:212-245)
// for user bins given as k-mer lists in HIBFIndex::emplace order (user bin b = values[offsets[b], offsets[b+1])): the
// reference's sketches, layout and hierarchical build (hibf_layout.cpp, hibf_build.cpp) with K1 setting the bits.
// sketch_registers (optional): the user bins' HyperLogLog sketches (12 bits, one after the other) when they were computed
// on the device (trx_hll_sketch_records_dev); otherwise they are computed here from the lists.
hibf_image build_hibf_like_seqan(trx_ctx_t * ctx, std::vector<uint64_t> const & values, std::vector<uint64_t> const & offsets,
                                 size_t hash_count, float fpr, size_t threads, std::vector<uint8_t> const * sketch_registers = nullptr);

void write_index(index_image const & img, std::string const & path);
index_image read_index(std::string const & path);
// index_params / load_params (include/index_base.h:5-16,197-202): only the three leading fields
void read_index_params(std::string const & path, uint8_t & k, std::string & molecule, bool & is_hibf);

// ------------------------------------------------------------------------------------------------- FASTA
struct fasta_record
{
    std::string name;    // header up to the first whitespace (kseq name.s)
    std::string comment; // rest of the header line
    std::string seq;     // concatenated sequence lines
};
// kseq_read over gzopen (include/kseq.h): plain or gzip, FASTA or FASTQ. Throws trxh::error if unopenable.
std::vector<fasta_record> read_fasta(std::string const & path);

// ------------------------------------------------------------------------------------- regex front end
// translate() / PostfixConverter (include/utils.h:96-468, src/utils.cpp:3-15): infix → postfix with '-' as the
// explicit concatenation operator; '.' and [^..] expand over the 20 amino acids.
std::string translate(std::string const & pattern);
// trimRegEx / reduce_query_alphabet / preprocess_query (src/query.cpp:78-155, include/query.h:80-94)
std::string trim_regex(std::string const & regex);
std::string reduce_query_alphabet(std::string const & regex, uint8_t const redmap[256]);

// --------------------------------------------------------------------------------------------- k-graph
enum node_symbol : int { Match = 256, Ghost = 257, Split = 258, Gap = 259 };

// Linearised acyclic NFA of construct_kgraph (src/construct_nfa.cpp:265-335): node ids, symbols and the
// two-slot arc map exactly as the reference numbers them, plus the collector's ranks
// (include/otf_collector.h:328-339).
// Catsite (include/construction_tools.h:151-185): a concatenation whose right operand has so many paths that walking it
// would explode; `-a` replaces the operand by Gap nodes, one per distinct length.
struct catsite
{
    int cleavage_site{-1};  // node before the high-complexity subgraph
    int cleavage_start{-1}; // its entry node
    int cleavage_end{-1};   // its exit node
    int downstream{-1};     // where the exit node leads (complete())
    rh_slots gaps;          // lengths of the paths through it, in the reference's robin_hood iteration order
};

struct kgraph
{
    std::vector<int> symbol;
    std::vector<int> first, second; // arc_map slots; -1 = none
    std::vector<int> rank;
    std::vector<size_t> gap;        // gap_map: length a Gap node stands for
    std::vector<std::pair<int, int>> arcs; // in creation order (source, target)
    std::vector<catsite> catsites;  // detect_bad_graphs (src/construction_tools.cpp:161-180), in detection order
    // OTFCollector::augment (include/otf_collector.h:466-493): merge neighbouring catsites, bridge each with Gap nodes,
    // rank again.  No-op without catsites.
    void augment();
    int add_node(int sym);
    void add_arc(int source, int target); // update_arc_map (src/construction_tools.cpp:136-158)
    void compute_ranks();
    size_t size() const { return symbol.size(); }
};
// reduced = true → construct_reduced_kgraph (src/construct_reduced_nfa.cpp), used when the index alphabet is reduced
kgraph construct_kgraph(std::string const & postfix, uint8_t k, bool reduced = false);

// ----------------------------------------------------------------------------------------- DGram index
// DGramIndex (include/dGramIndex.h:64-310) with its filter in HBM: gapped 3+3-mers "abc, gap g, def" of every record, one
// bin per FASTA file.  build() = constructor + populate(): two launches over the residues (count per bin → size the
// filter like compute_bitcount → emplace); the collector probes it from gapped items when it is attached to an index.
class DGramIndex
{
public:
    explicit DGramIndex(trx_ctx_t * ctx) : ctx_(ctx) {}
    ~DGramIndex();
    DGramIndex(DGramIndex const &) = delete;
    DGramIndex & operator=(DGramIndex const &) = delete;
    void build(dindex_arguments const & args);
    void store(std::string const & path) const; // save_dindex (:312-318)
    void load(std::string const & path);        // load_dindex (:320-326) + upload
    uint64_t min_gap() const { return img_.min_gap; }
    uint64_t max_gap() const { return img_.max_gap; }
    uint64_t bin_count() const { return img_.bin_count; }
    dgram_image const & image() const { return img_; }
    std::vector<uint64_t> const & codes_per_bin() const { return codes_per_bin_; } // dgram_buffer_[b].size(), after build()
    trx_ibf_t * flat() const { return flat_; }

private:
    trx_ctx_t * ctx_;
    dgram_image img_;
    trx_ibf_t * flat_{nullptr};
    std::vector<uint64_t> codes_per_bin_;
};
void drive_dindex(dindex_arguments const & args); // src/dGramIndex.cpp:19-38

// ------------------------------------------------------------------------------------------------ index
// TetrexIndex<flavor, molecule> with its bit matrix in HBM.
class DeviceIndex
{
public:
    explicit DeviceIndex(trx_ctx_t * ctx) : ctx_(ctx) {}
    // One process per GPU: this handle is shard `index` of `count`.  Call before load().
    //  * flat IBF: the handle holds the 64-bin column block [word_begin, word_begin + words) of EVERY row
    //    (trx_ibf_create(..., shard_index, shard_count); count must divide bin_words), path bitvectors are that wide;
    //  * HIBF: the (small) filter is replicated, every shard computes the whole candidate vector;
    //  * verify text: only the FASTA files of the shard's bins are read and uploaded, bins keep their global ids.
    void set_shard(uint32_t index, uint32_t count);
    uint32_t shard_index() const { return shard_index_; }
    uint32_t shard_count() const { return shard_count_; }
    uint64_t shard_word_begin() const;  // first 64-bin word this shard owns
    uint64_t shard_words() const;       // number of 64-bin words this shard owns
    uint64_t path_words() const { return is_hibf() ? bin_words() : shard_words(); } // width of a collector path slot
    ~DeviceIndex();
    DeviceIndex(DeviceIndex const &) = delete;
    DeviceIndex & operator=(DeviceIndex const &) = delete;

    // create_{ibf,hibf}_{dna,aa}_index + populate_index (src/index_base.cpp:25-70): reads the FASTA bins,
    // sizes the filter with the reference's rule and sets every bit on the GPU (K1).
    void build(index_arguments const & args);
    void store(std::string const & path) const; // store_ibf
    void load(std::string const & path);        // load_ibf + upload to the device
    // No filter at all: only what the regex front end and the k-graph walk need to know about an index (molecule "na" /
    // "aa", k, reduction, number of bins).  For collect_with_rows, whose caller brings the membership rows; needs no device.
    void describe_only(std::string const & molecule, uint8_t k, uint8_t reduction, uint64_t bins);

    uint8_t k() const { return img_.k; }
    bool is_dna() const { return img_.molecule == "na"; }
    bool is_hibf() const { return img_.is_hibf; }
    uint8_t reduction() const { return img_.reduction; }
    uint64_t bin_count() const { return img_.bin_count; }
    uint64_t bin_words() const { return (img_.bin_count + 63) / 64; }
    std::vector<std::string> const & acid_libs() const { return img_.acid_libs; }
    index_image const & image() const { return img_; }
    trx_ctx_t * ctx() const { return ctx_; }
    trx_ibf_t * flat() const { return flat_; }
    trx_hibf_t * hibf() const { return hibf_; }
    uint8_t const * aamap() const { return aamap_; }
    uint8_t const * redmap() const { return redmap_; }
    uint64_t sequences_indexed() const { return seq_count_; }

    // TetrexIndex::update_kmer (nucleotide_decomposer.h:113-121 / peptide_decomposer.h:293-299): rolls `symbol`
    // into `kmer` (forward value) and returns the value that is probed.
    uint64_t update_kmer(int symbol, uint64_t & kmer) const;

    // residue store for verify: built on first use from acid_libs (the files iter_disk_search would re-read)
    struct text_store
    {
        trx_text_t * text{nullptr};
        std::vector<std::string> names;      // per record: ">" + name
        std::vector<uint64_t> offsets;       // n_records + 1
        std::vector<uint32_t> rec_bin;
        std::string residues;
        std::vector<bool> loaded;            // per bin: its FASTA file has been read into this store
    };
    // the store with (at least) the candidate bins' records loaded and uploaded; `cand` = ceil(bins/64) words
    text_store const & text(std::vector<uint64_t> const & cand);
    // resident processes (servers, benchmarks) load every bin on first use instead of growing query by query
    void preload_all_text(bool on) { preload_all_ = on; }
    void set_threads(unsigned n) { threads_ = n; } // -t: host threads that parse the bins' FASTA files

    // Device scratch of the collector, kept between queries: the path-slot pool (cudaMalloc + cudaFree of up to
    // 128 MB per query cost 0.6 ms — more than the whole walk) and the k-mer / hit-vector buffers of the HIBF route.
    // `used` slots of the previous query are zeroed when the next one starts.
    struct collector_scratch
    {
        trx_paths_t * paths{nullptr};
        uint64_t words{}, capacity{}, used{};
        void * d_kmers{nullptr};
        void * d_hits{nullptr};
        uint64_t kmers_bytes{}, hits_bytes{};
    };
    collector_scratch & scratch() { return scratch_; }

    // `-g`: the DGram index the collector's gapped items probe (OTFCollector::dibf_ / has_dibf_).  Not owned.
    void attach_dgram(DGramIndex const * d);
    DGramIndex const * dgram() const { return dgram_; }

private:
    void upload();
    void init_maps();
    trx_ctx_t * ctx_;
    uint32_t shard_index_{0}, shard_count_{1};
    index_image img_;
    trx_ibf_t * flat_{nullptr};
    trx_hibf_t * hibf_{nullptr};
    uint8_t aamap_[256]{}, redmap_[256]{};
    uint64_t seq_count_{};
    bool preload_all_{false};
    unsigned threads_{0};
    std::unique_ptr<text_store> text_;
    collector_scratch scratch_;
    DGramIndex const * dgram_{nullptr};
};

// ------------------------------------------------------------------------------------------ query path
using bitvector = std::vector<uint64_t>; // ceil(bins/64) words, bit b = bin b (seqan::hibf::bit_vector words)

// Hooks for the column-sharded collector.  any_alive: called once per probe step with one byte per probed item
// (1 = some bin of THIS shard is still set on the item's path); it must OR the bytes over all shards in place
// (an all-reduce(MAX) of n bytes) so that every shard prunes the same items — `path_.none()` is a statement about
// ALL bins (include/otf_collector.h:383).  Unset = single device.
struct shard_hooks
{
    std::function<void(uint8_t * alive, size_t n)> any_alive;
    // preferred: a communicator of the CUDA library (NCCL over NVLink).  The alive bytes are then reduced ON THE DEVICE,
    // on the stream of the probe kernel that wrote them (trx_paths_step_sharded) — no callback, no host hop — and the
    // final candidate words are all-gathered the same way, so every rank returns the whole vector.
    trx_comm_t * comm{nullptr};
};

struct collect_stats
{
    uint64_t probes{};   // k-mers probed on the device
    uint64_t ranks{};    // level-synchronous rounds
    uint64_t items{};    // collector items processed
};

// process_query (include/query.h:226-247): preprocess, k-graph, OTFCollector::collect with every probe, path
// AND, none() test and path OR executed on the device.
// A sharded flat index returns the whole-width vector with only its own words filled in (the caller ORs / gathers).
// augment: collector.augment(catsites) before the walk (`-a`); a DGram index attached to `index` (`-g`) is probed by the
// gapped items (include/otf_collector.h:216-245,290-312).
bitvector process_query(std::string const & regex, DeviceIndex & index, collect_stats * stats = nullptr,
                        shard_hooks const * hooks = nullptr, bool augment = false);
// the collector alone on a ready k-graph (OTFCollector::collect, include/otf_collector.h:341-393)
bitvector collect(kgraph const & g, DeviceIndex & index, collect_stats * stats = nullptr, shard_hooks const * hooks = nullptr);

// process_query for a whole motif file (`tetrex query -f`): one membership batch for the k-mers of all patterns.  Per
// pattern: the candidate bins (bin_words() words), or the text of the exception process_query would have thrown.
struct batch_answer
{
    bool ok{false};
    bitvector bins;
    collect_stats stats{};
    std::string error;
};
std::vector<batch_answer> process_query_batch(std::vector<std::string> const & regexes, DeviceIndex & index, bool augment = false);

// process_query over membership rows the CALLER supplies (another filter, a cache, a remote shard): the k-graph is walked
// once without pruning to list the k-mers it can ask for, `rows(kmers, n, out)` fills n x bin_words() words
// (out[i * bin_words + w] = IBFIndex::query(kmers[i]) word w), and the walk proper runs on them.  No device call is made
// — which is also what lets the collector's rules be checked against the reference without a GPU (tests/test_host_cpu.py).
// `index` only has to be described (DeviceIndex::describe_only); DGram filters are not supported here.
using rows_fn = std::function<void(uint64_t const * kmers, size_t n, uint64_t * out_words)>;
bitvector process_query_with_rows(std::string const & regex, DeviceIndex & index, rows_fn const & rows, collect_stats * stats = nullptr,
                                  bool augment = false);

// iter_disk_search (include/query.h:97-188): verifies the candidate bins on the device (K4) and writes the
// reference's lines "bin_path\t>name\tmatch\tstart,end\n" (+ "…\tREVERSE STRAND HIT" for DNA) to `dest`
// ("-" = stdout).  Returns the number of lines written.
uint64_t iter_disk_search(bitvector const & hits, std::string const & query, DeviceIndex & index, std::string const & dest);
// the same into a stream (forward and reverse-strand lines alike); a sharded index prints the lines of its own bins —
// shards own ascending bin ranges, so concatenating the shards' outputs in shard order is the -t 1 print order
uint64_t iter_disk_search(bitvector const & hits, std::string const & query, DeviceIndex & index, std::ostream & out);

void drive_index(index_arguments const & args);
void drive_query(query_arguments & args);
void drive_inspection(std::string const & idx_path);

} // namespace trxh
