/* tetrex_host.h — C facade of the host layer (tetrex_b200/host/capi.cpp → tetrex_b200/lib/libtrxhost.so).
 *
 * include/tetrex_b200.h is the boundary of the CUDA library; this header is the boundary of the C++ layer that mirrors
 * the reference's query API on top of it, for drivers that are not C++ (one process per GPU under torchrun:
 * tetrex_b200/sharded.py).  Reference interfaces, relative to /root/reference:
 *   load_ibf                      include/index_base.h:189-195
 *   process_query                 include/query.h:226-247   (preprocess, construct_kgraph, OTFCollector::collect)
 *   iter_disk_search              include/query.h:97-188    (verify_fasta_hit & co., src/query.cpp:167-315)
 *   DGramIndex / drive_dindex     include/dGramIndex.h:64-326, src/dGramIndex.cpp:19-38
 * Every function returns 0, or -1 with the message in trxh_last_error() (thread-local): the messages are the ones the
 * C++ layer throws (trxh::error / trxh::archive_error), i.e. what src/main.cpp prints for std::runtime_error /
 * cereal::Exception.  There is no CPU fallback underneath: the handle lives on the device of the trx_ctx_t. */
#ifndef TETREX_HOST_H
#define TETREX_HOST_H

#include <stdint.h>

#include "tetrex_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

const char *trxh_last_error(void);

/* Loads shard `shard_index` of `shard_count` of an index file onto ctx's device (1 shard = the whole index).  Flat IBF:
 * the 64-bin column block of every row (shard_count must divide bin_words); HIBF: a replica; in both cases verify reads
 * only the FASTA files of the shard's bins. */
int trxh_open(trx_ctx_t *ctx, const char *path, uint32_t shard_index, uint32_t shard_count, void **out);
void trxh_close(void *handle);
/* out[0..8) = bin_count, bin_words, is_hibf, is_dna, k, shard_word_begin, shard_words, reduction */
int trxh_info(void *handle, uint64_t out[8]);

/* A communicator of the CUDA library (trx_comm_create, include/tetrex_b200.h) for this handle: a column-sharded
 * trxh_process_query then needs no callback — alive bytes are all-reduced on the device, on the probe kernel's stream, and
 * the candidate words are all-gathered, so out_words is complete on every rank.  NULL detaches.  Not owned. */
int trxh_attach_comm(void *handle, trx_comm_t *comm);
/* resident processes (servers, benchmarks): load every bin of the shard into HBM on the first verify; the default reads
 * the candidate bins' FASTA files only, like iter_disk_search (include/query.h:126-131) */
int trxh_preload_text(void *handle, int on);

/* `tetrex query -a` (include/arg_parse.h:65, include/query.h:243): trxh_process_query runs OTFCollector::augment
 * (include/otf_collector.h:466-493) on the k-graph before the walk — high-complexity stretches become Gap nodes. */
int trxh_set_augment(void *handle, int on);
/* `tetrex query -g FILE` (include/arg_parse.h:68, include/query.h:258-264): load_dindex (include/dGramIndex.h:320-326) onto
 * the handle's device; gapped collector items probe it (include/otf_collector.h:216-245).  The DGram index must have the
 * index's bin count.  NULL or "" detaches. */
int trxh_attach_dgram(void *handle, const char *path);
/* `tetrex dindex` (src/dGramIndex.cpp:19-38, include/dGramIndex.h:90-154): DGramIndex(min, max, 1, hash_count, fpr, files),
 * populate() and save_dindex(out_path) with the counting and the emplacing done on ctx's device (trx_dgram_count_dev /
 * trx_dgram_emplace_dev).  File names are stored as given.  per_bin (nullable): dgram_buffer_[b].size() per file. */
int trxh_dindex_build(trx_ctx_t *ctx, const char *const *files, uint64_t n_files, uint64_t min_gap, uint64_t max_gap,
                      uint32_t hash_count, float fpr, const char *out_path, uint64_t *per_bin);

/* ORs n bytes over all shards in place (all-reduce MAX): the collector's `path_.none()` (include/otf_collector.h:383) is
 * a statement about all bins.  Called once per probing rank of the k-graph, on the calling thread. */
typedef void (*trxh_alive_fn)(uint8_t *alive, uint64_t n, void *user);

/* process_query: candidate bins of `regex` as bin_words words (bit b = bin b, ANDed with the real bins like
 * run_collection does).  A column shard fills only its own words and REQUIRES `alive`; otherwise alive may be NULL.
 * stats (nullable, 3 entries): k-mers probed, ranks visited, collector items. */
int trxh_process_query(void *handle, const char *regex, trxh_alive_fn alive, void *user, uint64_t *out_words, uint64_t *stats);

/* process_query for n patterns at once — the filter phase of `tetrex query -f FILE` (src/query.cpp:463-466 loops over the
 * motifs and calls run_collection for each): the k-mers of ALL patterns' k-graphs go to the device in one membership
 * batch, every walk then runs on host words.  out_words: n x bin_words; status[i] = 1 answered / 0 the pattern was
 * refused (what trxh_process_query reports as an error); stats (nullable): n x 3. */
int trxh_process_query_batch(void *handle, const char *const *regexes, uint64_t n, uint64_t *out_words, uint8_t *status,
                             uint64_t *stats);

/* process_query (include/query.h:226-247) over membership rows the CALLER supplies — another filter implementation, a cache,
 * a remote shard.  The k-graph is walked once without pruning to list the k-mers it can ask for, rows() is called once
 * with all of them (out_words[i * ceil(bins/64) + w] = word w of IBFIndex::query(kmers[i])), then the walk proper runs on
 * those rows with the reference's rules.  No index file, no device.  molecule: 0 DNA, 1 peptide; reduction 0 / 1 murphy /
 * 2 li; out_words: ceil(bins/64) words; stats (nullable): probes, ranks, items of the pruned walk. */
typedef void (*trxh_rows_fn)(const uint64_t *kmers, uint64_t n, uint64_t *out_words, void *user);
int trxh_process_query_with_rows(int molecule, uint32_t k, int reduction, uint64_t bins, const char *regex, int augment,
                                 trxh_rows_fn rows, void *user, uint64_t *out_words, uint64_t *stats);

/* iter_disk_search over the shard's own bins: the reference's output lines, in its -t 1 order, as one malloc'ed buffer
 * (trxh_free).  cand_words: bin_words words (bits of other shards' bins are ignored). */
int trxh_verify(void *handle, const uint64_t *cand_words, const char *regex, char **out_text, uint64_t *out_len, uint64_t *n_lines);
void trxh_free(void *p);

/* Host only (no device): the order in which the collector drains one rank whose items arrived with these keys
 * (`kmer_ & submask_`), i.e. robin_hood::unordered_map slot order as OTFCollector::collect sees it through
 * `comp_table_[i].begin()` (include/otf_collector.h:355-357, include/robin_hood.h).  Duplicate keys are absorbed.
 * out[j] = index into keys of the j-th item popped; returns the number of items. */
uint64_t trxh_rank_pop_order(const uint64_t *keys, uint64_t n, uint32_t *out);

/* Host only (no device): the layout seqan::hibf computes for user bins given as k-mer lists — HyperLogLog sketches
 * (lib/hibf/src/sketch/hyperloglog.cpp), k-mer count estimates, compute_layout (every file of lib/hibf/src/layout,
 * lib/hibf/src/sketch/toolbox.cpp) as HIBFIndex::populate_index triggers it (include/index_hibf.h:124-129).  Dump (uint64
 * words): tmax, top_level_max_bin_id, #max_bins, per max bin {len, indices..., id}, #user_bins, per user bin {len, prev...,
 * storage_TB_id, number_of_technical_bins, idx}, then the n_bins estimates.  Returns the number of words (may exceed cap),
 * 0 on error. */
uint64_t trxh_hibf_layout(const uint64_t *values, const uint64_t *bin_offsets, uint64_t n_bins, uint32_t hash_funs, float fpr,
                          uint64_t *out, uint64_t cap);

/* Host only: iteration order of seqan::hibf's k-mer sets (robin_hood::unordered_flat_set<uint64_t>, lib/hibf/include/hibf/
 * contrib/robin_hood.hpp) after a script of op[i] = 0: insert(keys[i]) / 1: clear().  insert_into_ibf cuts a split user
 * bin into technical bins along that order (lib/hibf/src/build/insert_into_ibf.cpp:28-45), so it is part of the index. */
uint64_t trxh_flat_set_script(const uint64_t *keys, const uint8_t *op, uint64_t n, uint64_t *out);

/* diagnostic: print a native backtrace on SIGSEGV */
void trxh_debug_backtrace_on_segv(void);

#ifdef __cplusplus
}
#endif
#endif /* TETREX_HOST_H */
