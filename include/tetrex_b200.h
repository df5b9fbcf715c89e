/* tetrex_b200.h — C ABI of libtetrex_cuda.so: the B200 (sm_100a) implementation of TetRex's
 * index-probe + verify hot path.
 *
 * The reference (remyschwab/TetRex @ 6e856c0) has no FFI; its boundary for this path is the C++
 * API in its include/ headers.  Every entry point below names the reference function whose inner loop it
 * replaces.  The host C++ layer (tetrex_b200/host/) implements the reference's own types
 * (TetrexIndex, process_query, iter_disk_search, drive_index, drive_query) on top of these calls;
 * Python (the tetrex_b200 package, ctypes) binds the same symbols for tests and bench.py.
 *
 * Conventions
 *   - every call returns int: 0 = TRX_OK, negative = TRX_E*; text via trx_last_error() (thread local)
 *   - opaque handles, created/destroyed in pairs; a handle belongs to one context = one CUDA
 *     device + one stream; one context per host thread (mirrors hibf's "one agent per thread")
 *   - plain entry points take HOST buffers, copy in/out and return after the result is in the
 *     caller's buffer; *_dev entry points take DEVICE pointers, only enqueue work on the context's
 *     stream and return immediately (trx_ctx_sync waits)
 *   - there is no CPU fallback anywhere: without a usable sm_100 device trx_ctx_create fails
 *   - bit matrix layout = seqan::hibf::interleaved_bloom_filter's (lib/hibf/include/hibf/
 *     interleaved_bloom_filter.hpp:463-477): bit (row r, bin b) is bit b%64 (LSB first) of
 *     uint64 word r*bin_words + b/64
 */
#ifndef TETREX_B200_H
#define TETREX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TRX_ABI_VERSION 1

enum {
    TRX_OK = 0,
    TRX_EINVAL = -1,     /* bad argument (std::logic_error / std::invalid_argument in the reference) */
    TRX_ECUDA = -2,      /* CUDA runtime error, text in trx_last_error() */
    TRX_ENODEV = -3,     /* no sm_100 device: the product never falls back to the CPU */
    TRX_ENOMEM = -4,
    TRX_ECAPACITY = -5,  /* caller-provided output buffer too small; required size is reported */
    TRX_EREGEX = -6,     /* pattern not supported / malformed / matches the empty string */
    TRX_ERANGE = -7      /* index out of range (bin id >= bins, slot >= capacity, ...) */
};

enum { TRX_MOL_NA = 0, TRX_MOL_AA = 1 };                      /* TetrexIndex::molecule_ "na" / "aa" */
enum { TRX_RED_BASE = 0, TRX_RED_MURPHY = 1, TRX_RED_LI = 2 }; /* TetrexIndex::reduction_ */

typedef struct trx_ctx trx_ctx_t;
typedef struct trx_ibf trx_ibf_t;
typedef struct trx_hibf trx_hibf_t;
typedef struct trx_paths trx_paths_t;
typedef struct trx_text trx_text_t;
typedef struct trx_regex trx_regex_t;
typedef struct trx_comm trx_comm_t;

/* ------------------------------------------------------------------------------------------ context */
const char *trx_last_error(void);
int trx_abi_version(void);
/* device: CUDA ordinal.  stream: a cudaStream_t owned by the caller (e.g. torch's current stream) or
 * NULL to let the context create its own non-blocking stream. */
int trx_ctx_create(int device, void *stream, trx_ctx_t **out);
void trx_ctx_destroy(trx_ctx_t *ctx);
int trx_ctx_sync(trx_ctx_t *ctx);
void *trx_ctx_stream(trx_ctx_t *ctx);
/* number of kernels this context has launched since creation (bench.py's gpu_launches) */
uint64_t trx_ctx_launches(trx_ctx_t *ctx);
/* CUDA device ordinal of the context (callers that open further contexts on the same device, one per host thread) */
int trx_ctx_device(const trx_ctx_t *ctx);
/* raw device memory for callers that have no CUDA runtime of their own (the C++ host layer) */
int trx_dev_alloc(trx_ctx_t *ctx, uint64_t bytes, void **out);
int trx_dev_free(trx_ctx_t *ctx, void *ptr);
int trx_dev_memset(trx_ctx_t *ctx, void *dst, int value, uint64_t bytes);
int trx_copy_h2d(trx_ctx_t *ctx, void *dst_dev, const void *src_host, uint64_t bytes); /* synchronous */
int trx_copy_d2h(trx_ctx_t *ctx, void *dst_host, const void *src_dev, uint64_t bytes); /* synchronous */
/* CUDA-event timing on the context's stream (so callers without torch can time kernels) */
int trx_timer_start(trx_ctx_t *ctx);
int trx_timer_stop_ms(trx_ctx_t *ctx, float *ms); /* records, synchronises, returns elapsed ms */

/* ------------------------------------------------------------------- a1/a2: k-mer decomposition */
/* Host-side helpers (no device work), so the caller can size buffers like IBFIndex::find_largest_bin
 * (include/index_ibf.h:81-86) without hashing: emits per record = L+1 (DNA, the wrap-around quirk of
 * nucleotide_decomposer.h:104-111) or L-k+1 (peptide, peptide_decomposer.h:280-291); 0 if L < k. */
uint64_t trx_emit_count(int molecule, uint32_t k, uint64_t record_len);
/* PeptideDecomposer::create_residue_maps / create_r2r_maps (peptide_decomposer.h:59-213) */
void trx_peptide_maps(int reduction, uint8_t aamap[256], uint8_t redmap[256]);
/* IBFIndex::compute_bitcount (include/index_ibf.h:133-139): float log, double arithmetic, ceil */
uint64_t trx_compute_bitcount(uint64_t max_bin_emits, float fpr);

/* Emits the packed k-mers of n_records records exactly as {Nucleotide,Peptide}Decomposer::
 * decompose_record would hand them to TetrexIndex::emplace, in order.  residues: concatenated
 * record bytes; rec_offsets: n_records+1 byte offsets; out_offsets: n_records+1 emit offsets
 * (exclusive prefix sum of trx_emit_count); out_values: out_offsets[n_records] values. */
int trx_decompose_records(trx_ctx_t *ctx, int molecule, uint32_t k, int reduction, const uint8_t *residues,
                          const uint64_t *rec_offsets, uint64_t n_records, uint64_t *out_values,
                          uint64_t out_capacity, uint64_t *out_offsets);

/* ----------------------------------------------------------------------- a3/a4/a5: flat IBF */
/* seqan::hibf::interleaved_bloom_filter ctor (lib/hibf/src/interleaved_bloom_filter.cpp:32-53).
 * The matrix lives in HBM, zeroed.  shard_index/shard_count partition the bin_words columns of
 * every row in contiguous 64-bin blocks: this handle then owns words
 * [shard_index*bin_words/shard_count, (shard_index+1)*bin_words/shard_count) of each row
 * (shard_count must divide bin_words; 0/1 = whole matrix).  Row indices are shard independent. */
int trx_ibf_create(trx_ctx_t *ctx, uint64_t bins, uint64_t bin_size, uint32_t hash_funs, uint32_t shard_index,
                   uint32_t shard_count, trx_ibf_t **out);
void trx_ibf_destroy(trx_ibf_t *ibf);
/* out[0..10) = bins, technical_bins, bin_size, hash_shift, bin_words, hash_funs, n_words (whole
 * matrix), shard_word_begin, shard_words (per row), shard_n_words */
int trx_ibf_geometry(const trx_ibf_t *ibf, uint64_t out[10]);
void *trx_ibf_device_words(trx_ibf_t *ibf);
int trx_ibf_clear(trx_ibf_t *ibf);
/* words = the whole matrix in file layout (n_words); a shard keeps only its columns. */
int trx_ibf_upload(trx_ibf_t *ibf, const uint64_t *words, uint64_t n_words);
/* inverse; a shard writes only its columns into `words` (caller zero-fills / merges shards). */
int trx_ibf_download(trx_ibf_t *ibf, uint64_t *words, uint64_t n_words);

/* K1a — interleaved_bloom_filter::emplace (…cpp:119-140) for n (value, bin) pairs.  A shard ignores
 * pairs whose bin it does not own. */
int trx_emplace_values(trx_ibf_t *ibf, const uint64_t *values, const uint32_t *bins, uint64_t n);
int trx_emplace_values_dev(trx_ibf_t *ibf, const uint64_t *d_values, const uint32_t *d_bins, uint64_t n);
/* K1b — IBFIndex::populate_index's inner loops fused (include/index_ibf.h:101-131: decompose_record
 * → emplace → hash_and_fit → set bit) straight from residues; nothing is buffered per k-mer.
 * rec_bin[i] = user bin of record i.  Records shorter than k are skipped like the reference does. */
int trx_emplace_records(trx_ibf_t *ibf, int molecule, uint32_t k, int reduction, const uint8_t *residues,
                        const uint64_t *rec_offsets, const uint32_t *rec_bin, uint64_t n_records);
int trx_emplace_records_dev(trx_ibf_t *ibf, int molecule, uint32_t k, int reduction, const uint8_t *d_residues,
                            uint64_t n_residues, const uint64_t *d_rec_offsets, const uint32_t *d_rec_bin,
                            uint64_t n_records);

/* K1b for a SLAB of neighbouring bins: the same fused loop, with the promise-free hint that the records of this call
 * belong to bins [first_bin, first_bin + n_bins) — what IBFIndex::populate_index's outer loop over files
 * (include/index_ibf.h:105-112: one FASTA file = one bin) gives for free when files are streamed in order.  When the matrix
 * is far larger than L2 the bits of those bins are collected in a dense copy of their word columns (L2-sliced passes)
 * and folded into the interleaved matrix in one strided pass, instead of one DRAM read-modify-write per bit.  Records
 * whose bin lies outside the range are still emplaced correctly (straight into the matrix).  trx_emplace_records (host
 * buffers) derives the range from rec_bin itself. */
int trx_emplace_records_slab_dev(trx_ibf_t *ibf, int molecule, uint32_t k, int reduction, const uint8_t *d_residues,
                                 uint64_t n_residues, const uint64_t *d_rec_offsets, const uint32_t *d_rec_bin,
                                 uint64_t n_records, uint64_t first_bin, uint64_t n_bins);

/* N2 — DGramIndex, the gapped 3+3-mer filter behind `tetrex dindex` / `tetrex query -g` (include/dGramIndex.h).
 * DGramIndex::process_sequence (:194-243) run for every record at once: position i >= 2 with i + min_gap + 3 < L, the
 * three residues ending at i, a gap of min_gap..max_gap, the three residues from j = i + gap + 1 (stop at the first gap
 * with j + 2 >= L); all six must be upper-case letters of "ACDEFGHIKLMNPQRSTVWY"; records shorter than min_gap + 7 give
 * nothing; code = gap*20^6 + a1*20^5 + a2*20^4 + a3*20^3 + b1*400 + b2*20 + b3.
 *  trx_dgram_count   : per_bin[b] = how many codes bin b receives, duplicates included — what dgram_buffer_[b].size()
 *                      is after populate() (:103-111); its maximum sizes the filter (find_largest_bin / compute_bitcount,
 *                      :128-142).  per_bin[0..n_bins) is overwritten.
 *  trx_dgram_emplace : init_ibf's loop (:149-153): every code is emplaced into its record's bin.
 * max_gap is limited to TRX_DGRAM_MAX_GAP (the reference's default upper bound is 21). */
#define TRX_DGRAM_MAX_GAP 4096u
int trx_dgram_count(trx_ctx_t *ctx, const uint8_t *residues, const uint64_t *rec_offsets, const uint32_t *rec_bin,
                    uint64_t n_records, uint32_t min_gap, uint32_t max_gap, uint64_t n_bins, uint64_t *per_bin);
int trx_dgram_count_dev(trx_ctx_t *ctx, const uint8_t *d_residues, uint64_t n_residues, const uint64_t *d_rec_offsets,
                        const uint32_t *d_rec_bin, uint64_t n_records, uint32_t min_gap, uint32_t max_gap,
                        uint64_t n_bins, uint64_t *d_per_bin);
int trx_dgram_emplace(trx_ibf_t *ibf, const uint8_t *residues, const uint64_t *rec_offsets, const uint32_t *rec_bin,
                      uint64_t n_records, uint32_t min_gap, uint32_t max_gap);
int trx_dgram_emplace_dev(trx_ibf_t *ibf, const uint8_t *d_residues, uint64_t n_residues, const uint64_t *d_rec_offsets,
                          const uint32_t *d_rec_bin, uint64_t n_records, uint32_t min_gap, uint32_t max_gap);

/* N3 — seqan::hibf::sketch::compute_sketches (lib/hibf/src/sketch/compute_sketches.cpp:27-43) with hyperloglog::add
 * (lib/hibf/src/sketch/hyperloglog.cpp:54-70) fused into K1's decomposition: the HyperLogLog sketch of every user bin,
 * straight from the residues, over exactly the k-mers HIBFIndex::emplace would buffer (include/index_hibf.h:81-84,110-114).
 * registers: n_bins << sketch_bits bytes (bin b's sketch at b << sketch_bits), seqan's register layout; sketch_bits 5..20
 * (the reference uses 12).  The layout algorithm that consumes them runs on the host (tetrex_b200/host/hibf_layout.cpp). */
int trx_hll_sketch_records(trx_ctx_t *ctx, int molecule, uint32_t k, int reduction, const uint8_t *residues,
                           const uint64_t *rec_offsets, const uint32_t *rec_bin, uint64_t n_records, uint64_t n_bins,
                           uint32_t sketch_bits, uint8_t *registers);
int trx_hll_sketch_records_dev(trx_ctx_t *ctx, int molecule, uint32_t k, int reduction, const uint8_t *d_residues,
                               uint64_t n_residues, const uint64_t *d_rec_offsets, const uint32_t *d_rec_bin,
                               uint64_t n_records, uint64_t n_bins, uint32_t sketch_bits, uint8_t *d_registers);

/* K2 — containment_agent_type::bulk_contains (…cpp:203-290) / IBFIndex::query (index_ibf.h:146-150)
 * for a batch: out_words[i*shard_words + w] = AND_h row_h(kmers[i])[w]. */
int trx_bulk_contains(trx_ibf_t *ibf, const uint64_t *kmers, uint64_t n, uint64_t *out_words);
int trx_bulk_contains_dev(trx_ibf_t *ibf, const uint64_t *d_kmers, uint64_t n, uint64_t *d_out_words);

/* K2 without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments and result as trx_bulk_contains (host k-mers in, host hit vectors out, returns when they are there), but the
 * device staging of k-mers and hit vectors is a fixed file-scope `static __device__` array of TRX_NOALLOC_SCRATCH_BYTES
 * instead of the context's grow-on-demand buffers; the call makes no cudaMalloc / cudaFree / allocating library call.
 *   - limit: n*8 (rounded up to 256) + n*shard_words*8 bytes must fit the array; beyond that TRX_ECAPACITY with the
 *     numbers in trx_last_error().  trx_bulk_contains_noalloc_max_kmers(ibf) = the largest n for this filter
 *     (C3, 16 words per k-mer: 30 840; C4 whole, 128 words: 4 064).
 *   - every GPU that loads the library holds the array until the process exits.
 *   - ALL calls on one device share the array, and it keeps the last call's contents (nothing needs clearing: every byte a
 *     call reads was written by that call).  Calls that overlap on the GPU — two contexts / streams / host threads on the
 *     same device — RACE on it: serialise them.  The call itself synchronises its stream before it returns. */
#define TRX_NOALLOC_SCRATCH_BYTES (4u << 20)
int trx_bulk_contains_noalloc(trx_ibf_t *ibf, const uint64_t *kmers, uint64_t n, uint64_t *out_words);
uint64_t trx_bulk_contains_noalloc_max_kmers(const trx_ibf_t *ibf);

/* ------------------------------------------------------ a6: collector path algebra in HBM */
/* A pool of `capacity` path bitvectors ("slots") of words_per_slot uint64 each — the path_ member
 * of every live CollectorsItem (include/otf_collector.h:19-37) — resident in HBM. */
int trx_paths_create(trx_ctx_t *ctx, uint64_t words_per_slot, uint64_t capacity, trx_paths_t **out);
void trx_paths_destroy(trx_paths_t *p);
void *trx_paths_device_words(trx_paths_t *p);
/* slots [first, first+n): bits [0,bins) set, padding clear = bitvector(bins, true)
 * (otf_collector.h:344); bins == 0 → all clear = bitvector(bins) (path_matrix, :343). */
int trx_paths_fill(trx_paths_t *p, uint64_t first, uint64_t n, uint64_t bins);
/* update_path's probe + `path_ &= hits` + `path_.none()` (otf_collector.h:247-279,383) for n items
 * in one launch: slot[slots[i]] &= bulk_contains(kmers[i]); alive[i] = any bit left.  Each slot may
 * appear once per call.  Host arrays in, alive (n bytes) out. */
int trx_paths_step(trx_ibf_t *ibf, trx_paths_t *p, const uint64_t *kmers, const uint32_t *slots, uint64_t n,
                   uint8_t *alive);
/* the same launch on device arrays: enqueues on the context's stream and returns (alive bytes land in d_alive). */
int trx_paths_step_dev(trx_ibf_t *ibf, trx_paths_t *p, const uint64_t *d_kmers, const uint32_t *d_slots, uint64_t n,
                       uint8_t *d_alive);
/* trx_paths_step for a 64-bin column shard: `path_.none()` is a statement about ALL bins, so the
 * alive bytes are OR-ed over the shards (trx_comm_any_alive_dev, on the device, same stream) before they are copied
 * out; every rank must make the same call with the same k-mers and slots. */
int trx_paths_step_sharded(trx_ibf_t *ibf, trx_paths_t *p, trx_comm_t *comm, const uint64_t *kmers, const uint32_t *slots,
                           uint64_t n, uint8_t *alive);
/* same, against precomputed hit vectors resident on the device (HIBF route: trx_hibf_membership_dev
 * wrote d_hits[hit_index[i]*words_per_slot ...]) */
int trx_paths_and_hits(trx_paths_t *p, const uint64_t *d_hits, const uint32_t *hit_index, const uint32_t *slots,
                       uint64_t n, uint8_t *alive);
/* push/absorb/split_procedure and `path_matrix |= top.path_` (otf_collector.h:190-208,283-288,363):
 * slot[dst[i]] |= slot[src[i]] for n pairs, any multiplicity of dst (atomic OR). */
int trx_paths_or(trx_paths_t *p, const uint32_t *dst, const uint32_t *src, uint64_t n);
int trx_paths_read(trx_paths_t *p, uint64_t first, uint64_t n, uint64_t *out_words);

/* -------------------------------------------------------------------------- a8: HIBF probe */
/* Flattened seqan::hibf::hierarchical_interleaved_bloom_filter (…/hierarchical_interleaved_bloom_
 * filter.hpp:209-238).  Per IBF i: geometry, matrix words, and per technical bin next_ibf_id and
 * ibf_bin_to_user_bin_id (UINT64_MAX = merged, UINT64_MAX-1 = deleted). */
typedef struct trx_hibf_desc {
    uint64_t n_ibfs;
    uint64_t n_user_bins;
    const uint64_t *bins;            /* [n_ibfs] ibf.bin_count() */
    const uint64_t *bin_size;        /* [n_ibfs] */
    const uint64_t *hash_funs;       /* [n_ibfs] */
    const uint64_t *const *words;    /* [n_ibfs] → technical_bins*bin_size/64 words each */
    const uint64_t *const *next_ibf; /* [n_ibfs] → table_len[i] entries */
    const uint64_t *const *user_bin; /* [n_ibfs] → table_len[i] entries */
    const uint64_t *table_len;       /* [n_ibfs] */
} trx_hibf_desc;
int trx_hibf_upload(trx_ctx_t *ctx, const trx_hibf_desc *desc, trx_hibf_t **out);
void trx_hibf_destroy(trx_hibf_t *h);
/* K3 — HIBFIndex::query = membership_for({kmer}, 1) → populate_bitvector (include/index_hibf.h:
 * 132-147; …hpp:321-355) as a level-synchronous frontier.  out_words: n × ceil(n_user_bins/64).
 * stats (nullable, host): [0] = IBFs visited in total, [1] = levels run. */
int trx_hibf_membership(trx_hibf_t *h, const uint64_t *kmers, uint64_t n, uint64_t *out_words, uint64_t *stats);
int trx_hibf_membership_dev(trx_hibf_t *h, const uint64_t *d_kmers, uint64_t n, uint64_t *d_out_words,
                            uint64_t *stats);
/* The same in seqan's own result form: membership_for returns the ids of the user bins that hold the value
 * (…hpp:321-355,424-439); HIBFIndex::query only then turns them into a bitvector (include/index_hibf.h:132-147).
 * counts[i] = number of user bins of k-mer i; ids = the ids of k-mer 0, then of k-mer 1, …, ascending inside a k-mer, as
 * id_bytes = 2 (number_of_user_bins <= 65536) or 4 byte integers; cap = room in ids (entries), *n_ids = entries written.
 * The rows are compacted on the device, so a host caller pays for the ids it gets, not for ceil(bins/64) words per k-mer.
 * TRX_ECAPACITY with *n_ids = a lower bound of what is needed when cap is too small. */
int trx_hibf_membership_ids(trx_hibf_t *h, const uint64_t *kmers, uint64_t n, uint32_t *counts, void *ids, uint32_t id_bytes,
                            uint64_t cap, uint64_t *n_ids, uint64_t *stats);
uint64_t trx_hibf_user_bin_words(const trx_hibf_t *h);

/* ------------------------------------------------------------------------ a9/a10: verify */
/* Residue store in HBM: what iter_disk_search re-reads from disk on every query
 * (include/query.h:97-188).  Records are grouped by bin, bins ascending, file order inside a bin:
 * rec_bin must be non-decreasing. */
int trx_text_upload(trx_ctx_t *ctx, const uint8_t *residues, const uint64_t *rec_offsets, const uint32_t *rec_bin,
                    uint64_t n_records, uint64_t n_bins, trx_text_t **out);
void trx_text_destroy(trx_text_t *t);

/* Compiles the verify regex as iter_disk_search does ("(" + query + ")"): posix_longest = 1 →
 * RE2::POSIX, leftmost-longest (peptide, query.h:146-149); 0 → default RE2, leftmost-first (DNA,
 * query.h:101-104).  Supported syntax = what TetRex's own lexer accepts (include/utils.h:96-468):
 * literals, '.', [..] / [^..] without ranges, ( ) | * + ? {m} {m,n}, backslash escapes.  Anything
 * else, more than TRX_REGEX_MAX_POSITIONS symbol positions after {m,n} expansion, or a pattern that
 * matches the empty string (the reference spins forever on those) → TRX_EREGEX. */
#define TRX_REGEX_MAX_POSITIONS 250
int trx_regex_compile(const char *pattern, int posix_longest, trx_regex_t **out);
void trx_regex_destroy(trx_regex_t *re);
int trx_regex_positions(const trx_regex_t *re);
/* Host only (output formatting of the conjunction path, src/query.cpp:318-339): end (exclusive) of the earliest-ENDING
 * match of re in text[0, len), -1 if none.  RE2::Set::Match lists the rules of a set in the order in which their first
 * matches complete; verify_fasta_set prints them in that order. */
int64_t trx_regex_first_match_end(const trx_regex_t *re, const uint8_t *text, uint64_t len);
/* Host only (tests): the determinised matcher phase 2 uses for POSIX leftmost-longest patterns whose subset automaton stays
 * small (FindAndConsume's anchored step, lib/re2/re2/re2.h:668-669).  info[0..3) = states, byte classes, first accepting
 * state; returns 0 if the pattern has no such matcher (leftmost-first, or too many subsets).  trx_regex_dfa_match_end:
 * end (exclusive) of the longest match that STARTS at text[start], -1 if none, -2 if there is no DFA.
 * trx_regex_dfa_starts: the reverse searcher phase 1 uses for patterns without a selective necessary factor (bounded match
 * length, no anchors), run over text[0, len) as one record: starts[p] = 1 iff some match starts at p; returns the longest
 * possible match length, 0 if the pattern has no such searcher. */
int trx_regex_dfa_info(trx_regex_t *re, uint32_t info[3]);
uint32_t trx_regex_dfa_starts(trx_regex_t *re, const uint8_t *text, uint64_t len, uint8_t *starts);
int64_t trx_regex_dfa_match_end(trx_regex_t *re, const uint8_t *text, uint64_t len, uint64_t start);

typedef struct trx_hit {
    uint32_t bin;
    uint32_t record; /* global record index as uploaded */
    uint32_t start;  /* byte offset in the record (reverse strand: in the reverse-complemented record) */
    uint32_t end;    /* exclusive */
} trx_hit;

/* K4 — verify_fasta_hit / verify_reduced_fasta_hit / reverse_verify_fasta_hit (src/query.cpp:167-315):
 * every record of every bin whose bit is set in candidate_bin_words (ceil(n_bins/64) words, host) is
 * scanned with non-overlapping FindAndConsume semantics.  map256 (nullable, host): byte map applied
 * to the text before matching (redmap_ of a reduced index).  reverse_strand: scan comp_tab-
 * complemented, reversed records (src/query.cpp:7-16,167-191).  hits come back sorted by
 * (bin, record, start) = the reference's -t 1 print order.  If more than cap hits exist, *n_hits holds the
 * total, the call returns TRX_ECAPACITY and the contents of hits are unspecified (call again with room).
 * stats (nullable, 2 entries): [0] = residues scanned, [1] = literals (fixed window) or atoms (gapped factor) of the
 * SIMD prefilter (0 = the exact automaton kernel ran as phase 1). */
int trx_regex_scan(trx_text_t *text, trx_regex_t *re, const uint64_t *candidate_bin_words, int reverse_strand,
                   const uint8_t *map256, trx_hit *hits, uint64_t cap, uint64_t *n_hits, uint64_t *stats);

/* The same scan in two halves, for batches of queries (run_multiple_queries, include/query.h:330-340, loops over a
 * file of regexes): _begin plans and enqueues everything on the context's stream and returns a ticket at once, _end
 * waits for that scan and delivers its hits exactly as trx_regex_scan does (which is _begin + _end).  Up to
 * TRX_SCAN_SLOTS scans of one text may be in flight, each on a stream of its own (NOT the context's stream: order
 * other work against a scan through _end); their kernels run back to back on the device while the host is already
 * planning the next one — the ~50 us of per-call latency (plan, launches, read-back, sync) is paid once per
 * batch instead of once per regex.  A further _begin with every slot taken returns TRX_ECAPACITY.  The regex handle and
 * the candidate / map arrays are only read during _begin.  Tickets may be ended in any order.  _end with hits == NULL
 * and n_hits == NULL abandons the scan: it waits for it and frees the slot without delivering anything. */
#define TRX_SCAN_SLOTS 4
int trx_regex_scan_begin(trx_text_t *text, trx_regex_t *re, const uint64_t *candidate_bin_words, int reverse_strand,
                         const uint8_t *map256, uint64_t cap, int *ticket);
int trx_regex_scan_end(trx_text_t *text, int ticket, trx_hit *hits, uint64_t *n_hits, uint64_t *stats);

/* K4 for a BATCH of patterns in one pass over the text — run_multiple_queries (include/query.h:330-340: a file of motifs,
 * each verified against its own candidate bins) and the Set-based verify of the conjunction path (include/query.h:191-224,
 * src/query.cpp:318-339) without re-reading the text per pattern.  Pattern i is scanned over the bins set in
 * candidate_bin_words[i * cand_stride_words ...] (cand_stride_words = 0: all patterns share one bitmap); forward strand,
 * no byte map.  hits come back grouped by pattern, each group in trx_regex_scan's order; hit_pattern[j] = pattern of
 * hits[j]; per_pattern_hits (nullable) = group sizes.  More than cap hits: TRX_ECAPACITY, *n_hits = how many there are
 * (nothing is delivered).
 * Every tile of the union of candidate bins is read ONCE per pass of up to 128 patterns.  First stage of a pass: the
 * patterns' class windows, bit-parallel over the patterns (trx_regex_window_plan shows a pattern's window); patterns
 * without one share a pass that evaluates their necessary-factor plans on occurrence bitmasks of the batch's alphabet;
 * patterns with neither, or with more than 32 768 candidate records, are scanned one at a time inside the same call
 * (same results).  The hits of all patterns stay in one device list until the call sorts them there — (pattern, record,
 * start) — and copies them out once; passes are enqueued back to back, nothing is read back in between.
 * TRX_MULTI_WINDOW=off / TRX_REGEX_MULTI=off switch the first stage / the shared passes off (A/B, tests).
 * stats (nullable, 4 entries): residues read by the shared passes, patterns answered by them, patterns answered one at a
 * time, shared passes launched. */
int trx_regex_scan_multi(trx_text_t *text, trx_regex_t *const *regexes, uint32_t n_patterns,
                         const uint64_t *candidate_bin_words, uint64_t cand_stride_words, trx_hit *hits,
                         uint32_t *hit_pattern, uint64_t cap, uint64_t *n_hits, uint64_t *per_pattern_hits, uint64_t *stats);

/* Host only, no device needed: the phase-1 prefilter trx_regex_scan would pick for this regex on a text with the
 * given relative byte frequencies (what RE2 does internally with its prefilter/memchr accelerations,
 * lib/re2/re2/prog.h ConfigurePrefixAccel — here the plan is visible so that tests can check it is a NECESSARY
 * condition of every match).  out[0] = 0 none (exact automaton), 1 fixed window, 2 gapped atoms.
 * window: out[1] = literals in bytes 0..3, out[2] = in bytes 4..7, out[3] = address of the match's first byte minus
 *         address of window byte 0, out[8+j] = byte offset of literal j in the window, out[16+j] = literal byte.
 * atoms:  out[1] = distinct literal bytes, out[2] = terms, out[5] = atoms, out[3], out[4] = the match's first byte
 *         lies at an address in [anchor + out[3], anchor + out[4]]; term i: byte out[8+i] at anchor + d for a d in
 *         [out[24+i], out[40+i]], belonging to atom out[56+i] — an atom holds if ANY of its terms does (a class of up
 *         to 3 bytes), the SIMD stage passes where ALL atoms hold.  out[6] = confirm checks, evaluated only at the
 *         anchors that pass: check c holds if the byte at anchor + d is in its class for a d in
 *         [out[72+10c], out[73+10c]] (signed), class = 256-bit mask in out[74+10c .. 82+10c) (bit b of word w = byte
 *         32w + b). */
int trx_regex_filter_plan(const trx_regex_t *re, const double *byte_freq, int reverse_strand, const uint8_t *map256,
                          int32_t out[136]);

/* Host only: the class window trx_regex_scan_multi's bit-parallel first stage (regex_window_multi) would use for the
 * pattern on a text with these byte frequencies: up to 6 CONSECUTIVE text bytes every match contains, each from a byte
 * class.  out[0] = 1 usable / 0 none (the pattern then takes the term-list stage), out[1] = length, out[2], out[3] = the
 * window's first byte lies at [match start + out[2], match start + out[3]]; classes[8k .. 8k+8) = 256-bit class of window
 * byte k.  out[4] = confirm checks evaluated where the window holds (other necessary positions of the pattern): check c
 * holds if the byte at window byte 0 + d is in its class for a d in [out[5+10c], out[6+10c]] (signed), class = 256-bit
 * mask in out[7+10c .. 15+10c).  pass[0] = the fraction of text positions expected to pass the window, pass[1] = window
 * and checks.  classes and pass may be NULL. */
int trx_regex_window_plan(const trx_regex_t *re, const double *byte_freq, int32_t out[65], uint32_t classes[48], double pass[2]);

/* ------------------------------------------------------------------- multi-GPU: collectives (8(e)) */
/* One communicator per context = per GPU (NCCL over NVLink/NVSwitch, bound at run time: no link dependency, single-GPU
 * use never loads it).  The reference has nothing of the kind (single process, OpenMP over bins, include/query.h:113);
 * what it replaces is the implicit "all bins are in this address space" of OTFCollector::collect.
 *   one process per GPU:  rank 0 calls trx_comm_unique_id, the caller ships the 128 bytes to every rank by any
 *                         channel (torch.distributed, a file, MPI), every rank calls trx_comm_create;
 *   one process, N GPUs:  trx_comm_create_all(ctxs, n, comms) (ncclCommInitAll).
 * All _dev calls enqueue on the context's stream (ordered after the kernels already queued there) and return. */
#define TRX_COMM_ID_BYTES 128
int trx_comm_unique_id(uint8_t id[TRX_COMM_ID_BYTES]);
int trx_comm_create(trx_ctx_t *ctx, int rank, int world, const uint8_t id[TRX_COMM_ID_BYTES], trx_comm_t **out);
int trx_comm_create_all(trx_ctx_t *const *ctxs, int n, trx_comm_t **out);
void trx_comm_destroy(trx_comm_t *comm);
int trx_comm_rank(const trx_comm_t *comm);
int trx_comm_world(const trx_comm_t *comm);
uint64_t trx_comm_collectives(const trx_comm_t *comm); /* collectives issued so far (bench: what crossed NVLink) */
int trx_comm_version(void);                            /* NCCL version code, -1 if NCCL cannot be loaded */
/* in-place OR of n alive bytes over all ranks (all-reduce MAX): otf_collector.h:383's none() made global */
int trx_comm_any_alive_dev(trx_comm_t *comm, uint8_t *d_alive, uint64_t n);
/* d_recv[r*bytes_per_rank ...] = rank r's d_send: final candidate words, hit counts, hit tuples */
int trx_comm_allgather_dev(trx_comm_t *comm, const void *d_send, void *d_recv, uint64_t bytes_per_rank);
int trx_comm_bcast_dev(trx_comm_t *comm, void *d_buf, uint64_t bytes, int root);
/* host buffers (small control messages): stages, gathers, reads back, synchronises */
int trx_comm_allgather(trx_comm_t *comm, const void *send, void *recv, uint64_t bytes_per_rank);

#ifdef __cplusplus
}
#endif
#endif /* TETREX_B200_H */
