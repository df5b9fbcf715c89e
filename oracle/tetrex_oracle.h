/* tetrex_oracle.h — CPU restatement of TetRex's index-probe arithmetic.  TEST INFRASTRUCTURE ONLY:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * liboracle_tetrex.so.  The product (tetrex_b200/) never links or calls it.
 *
 * Parity status: PINNED — tests/test_oracle_vs_ref.py checks every function below against the
 * unmodified reference compiled into oracle/_ref/libtetrex_ref.so (oracle/Makefile, target `ref`),
 * and tests/golden/ holds vectors generated from that library (tests/golden/make_golden.py). */
#ifndef TETREX_ORACLE_H
#define TETREX_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* include/nucleotide_decomposer.h:81-111, include/peptide_decomposer.h:261-291.
 * molecule 0 = DNA, 1 = peptide; reduction 0 Base, 1 Murphy, 2 Li.  Returns the number of emits
 * (written up to cap). */
uint64_t orc_decompose(int molecule, int k, int reduction, const uint8_t *seq, uint64_t len, uint64_t *out, uint64_t cap);
/* update_kmer (nucleotide_decomposer.h:113-121, peptide_decomposer.h:293-299): rolls `symbol` into
 * *kmer (forward value) and returns the value that is probed (canonical for DNA). */
uint64_t orc_update_kmer(int molecule, int k, int reduction, int symbol, uint64_t *kmer);
void orc_peptide_maps(int reduction, uint8_t aamap[256], uint8_t redmap[256]);
/* include/index_ibf.h:133-139 */
uint64_t orc_compute_bitcount(uint64_t n, float fpr);

/* lib/hibf/include/hibf/interleaved_bloom_filter.hpp:177-187 without the final * technical_bins */
uint64_t orc_hash_row(uint64_t value, int seed_index, uint64_t bin_size);
/* lib/hibf/src/interleaved_bloom_filter.cpp:119-140; words = bin_words*bin_size uint64 */
void orc_emplace(uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, const uint64_t *values,
                 const uint32_t *bin_ids, uint64_t n);
/* …cpp:203-290; out = n × bin_words; threads > 1 → OpenMP over the batch */
void orc_bulk_contains(const uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, const uint64_t *values,
                       uint64_t n, uint64_t *out, int threads);
uint64_t orc_bulk_contains_checksum(const uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs,
                                    const uint64_t *values, uint64_t n, int threads);

/* lib/hibf/include/hibf/hierarchical_interleaved_bloom_filter.hpp:321-355 with one value, threshold 1,
 * followed by HIBFIndex::populate_bitvector (include/index_hibf.h:132-147).
 * Per IBF i: bins[i], bin_size[i], hash_funs[i], words[i], next_ibf[i], user_bin[i] (table_len[i] entries).
 * out = n × ceil(n_user_bins/64). */
void orc_hibf_membership(uint64_t n_ibfs, uint64_t n_user_bins, const uint64_t *bins, const uint64_t *bin_size,
                         const uint64_t *hash_funs, const uint64_t *const *words, const uint64_t *const *next_ibf,
                         const uint64_t *const *user_bin, const uint64_t *table_len, const uint64_t *values, uint64_t n,
                         uint64_t *out, int threads);

/* include/index_ibf.h:88-131 for in-memory records: decompose every record, emplace into rec_bin[r] */
void orc_emplace_records(uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, int molecule, int k,
                         int reduction, const uint8_t *text, const uint64_t *rec_offsets, const uint32_t *rec_bin,
                         uint64_t n_records, int threads);

#ifdef __cplusplus
}
#endif
#endif
