/* tetrex_oracle.c — CPU restatement (plain C) of the reference's index-probe arithmetic.
 * TEST INFRASTRUCTURE ONLY, see tetrex_oracle.h.  Each function cites the reference lines it follows;
 * it is written scalar and sequential on purpose, so that it reads like the reference's own loops. */
#include "tetrex_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

static const uint64_t k_seeds[5] = {13572355802537770549ULL, 13043817825332782213ULL, 10650232656628343401ULL,
                                    16499269484942379435ULL, 4893150838803335377ULL};

/* ------------------------------------------------------------------------------------- a2 maps */
void orc_peptide_maps(int reduction, uint8_t aamap[256], uint8_t redmap[256])
{
    /* include/peptide_decomposer.h:59-151: code tables, one assignment per letter */
    static const char *letters = "ABCDEFGHIJKLMNOPQRSTUVWXYZ";
    /*                                   A  B  C  D  E  F  G  H  I  J  K  L  M   N   O   P   Q   R   S   T   U   V   W   X   Y   Z */
    static const uint8_t base_code[26] = {0, 2, 1, 2, 3, 4, 5, 6, 7, 9, 8, 9, 10, 11, 20, 12, 13, 14, 15, 16, 20, 17, 18, 20, 19, 3};
    static const uint8_t murphy_code[26] = {0, 1, 2, 1, 1, 3, 4, 5, 6, 6, 7, 6, 6, 1, 7, 8, 1, 7, 9, 9, 2, 6, 3, 9, 3, 1};
    static const uint8_t li_code[26] = {0, 1, 2, 1, 1, 3, 4, 5, 6, 7, 8, 7, 7, 5, 8, 9, 1, 8, 0, 0, 2, 6, 3, 0, 3, 1};
    /* include/peptide_decomposer.h:153-213: residue → representative letter */
    static const char *murphy_red = "ABCBBFGHIIKIIBKPBKSSCIFSFB";
    static const char *li_red = "ABCBBFGHIJKJJHKPBKAACIFAFB";
    memset(aamap, 0, 256);
    memset(redmap, 0, 256);
    const uint8_t *code = reduction == 1 ? murphy_code : reduction == 2 ? li_code : base_code;
    const char *red = reduction == 1 ? murphy_red : li_red; /* Base falls into the Li branch (:184-212) */
    for (int i = 0; i < 26; ++i)
    {
        aamap[(uint8_t)letters[i]] = code[i];
        redmap[(uint8_t)letters[i]] = (uint8_t)red[i];
    }
}

/* ---------------------------------------------------------------------------------- a1 helpers */
static uint64_t dna_mask(int k)
{
    return k >= 32 ? ~0ULL : ((1ULL << (2 * k)) - 1); /* nucleotide_decomposer.h:36 */
}

static uint64_t dna_revcomp(uint64_t kmer, int k)
{
    /* nucleotide_decomposer.h:44-79 computes this with two nibble LUT shuffles and a byte reversal;
     * stated directly: reverse the order of the k 2-bit codes and complement each (code ^ 2). */
    uint64_t rc = 0;
    uint64_t const top_bit = kmer >> 63;
    for (int i = 0; i < k; ++i)
    {
        rc = (rc << 2) | ((kmer & 3) ^ 2);
        kmer >>= 2;
    }
    /* :60 `kmer2 >>= 4` is an ARITHMETIC shift of a signed 64-bit vector lane: when bit 63 is set (k = 32 and
     * the first base is T/G) the top byte's shuffle index gets bit 7 and pshufb returns 0 for it, so the two
     * leading bases contribute nothing: the lowest nibble of the result is 0. */
    if (top_bit)
        rc &= ~0xFULL;
    return rc;
}

uint64_t orc_decompose(int molecule, int k, int reduction, const uint8_t *seq, uint64_t len, uint64_t *out, uint64_t cap)
{
    uint64_t n = 0;
    if (k <= 0 || len < (uint64_t)k)
        return 0; /* include/index_ibf.h:114-118 skips such records */
#define EMIT(v)                                                                                                        \
    do                                                                                                                 \
    {                                                                                                                  \
        if (n < cap)                                                                                                   \
            out[n] = (v);                                                                                              \
        ++n;                                                                                                           \
    } while (0)
    if (molecule == 0)
    {
        /* nucleotide_decomposer.h:81-111 */
        uint64_t const mask = dna_mask(k);
        int const left_shift = 2 * k - 2;
        uint64_t fwd = 0;
        for (int i = 0; i < k; ++i)
            fwd = (fwd << 2) | ((seq[i] >> 1) & 3);
        uint64_t rev = dna_revcomp(fwd, k);
        EMIT(fwd <= rev ? fwd : rev);
        for (uint64_t i = 0; i < len; ++i) /* the WHOLE record, from 0 (:110) */
        {
            uint64_t fb = (seq[i] >> 1) & 3;
            /* :95-96 — `auto fb = (base>>1)&3` is an int, so `(fb^0b10)<<left_shift_` is a 32-bit shift: for
             * k >= 16 (left_shift_ >= 30) it overflows / is undefined in C++ and, as compiled (x86 shl masks the
             * count to 5 bits), yields (fb^2) << (left_shift_ & 31), sign-extended when it is ORed into the uint64
             * store.  The index side of the reference therefore does NOT hold true reverse complements for k >= 16;
             * reproduced here because index bits must match. */
            uint64_t cb = (uint64_t)(int64_t)(int32_t)((uint32_t)(fb ^ 2) << (left_shift & 31));
            fwd = ((fwd << 2) & mask) | fb;
            rev = ((rev >> 2) & mask) | cb;
            EMIT(fwd <= rev ? fwd : rev);
        }
    }
    else
    {
        /* peptide_decomposer.h:261-291 */
        uint8_t aamap[256], redmap[256];
        orc_peptide_maps(reduction, aamap, redmap);
        uint64_t const mask = k >= 32 ? ~0ULL : ((1ULL << (5 * k)) - 1);
        uint64_t h = 0;
        for (int i = 0; i < k; ++i)
            h = (h << 5) | aamap[seq[i]];
        EMIT(h);
        for (uint64_t i = (uint64_t)k; i < len; ++i)
        {
            h = ((h << 5) & mask) | aamap[seq[i]];
            EMIT(h);
        }
    }
#undef EMIT
    return n;
}

uint64_t orc_update_kmer(int molecule, int k, int reduction, int symbol, uint64_t *kmer)
{
    if (molecule == 0)
    {
        /* nucleotide_decomposer.h:113-121 */
        uint64_t fb = ((uint64_t)symbol >> 1) & 3;
        uint64_t forward = ((*kmer << 2) & dna_mask(k)) | fb;
        *kmer = forward;
        uint64_t reverse = dna_revcomp(forward, k);
        return forward <= reverse ? forward : reverse;
    }
    /* peptide_decomposer.h:293-299 */
    uint8_t aamap[256], redmap[256];
    orc_peptide_maps(reduction, aamap, redmap);
    uint64_t const mask = k >= 32 ? ~0ULL : ((1ULL << (5 * k)) - 1);
    uint64_t forward = ((*kmer << 5) & mask) | aamap[(uint8_t)symbol];
    *kmer = forward;
    return forward;
}

uint64_t orc_compute_bitcount(uint64_t n, float fpr)
{
    /* include/index_ibf.h:133-139: std::log(float) is logf */
    double numerator = -(double)n * (double)logf(fpr);
    double denominator = pow(log(2.0), 2);
    return (uint64_t)ceil(numerator / denominator);
}

/* ------------------------------------------------------------------------------------ a4 / a5 */
uint64_t orc_hash_row(uint64_t h, int seed_index, uint64_t bin_size)
{
    /* interleaved_bloom_filter.hpp:177-187; hash_shift = countl_zero(bin_size) (…cpp:48) */
    int const hash_shift = __builtin_clzll(bin_size);
    h *= k_seeds[seed_index];
    h ^= h >> hash_shift;
    h *= 11400714819323198485ULL;
    return (uint64_t)(((__uint128_t)h * (__uint128_t)bin_size) >> 64);
}

void orc_emplace(uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, const uint64_t *values,
                 const uint32_t *bin_ids, uint64_t n)
{
    uint64_t const technical_bins = ((bins + 63) / 64) * 64;
    for (uint64_t i = 0; i < n; ++i)
        for (int f = 0; f < hash_funs; ++f)
        {
            uint64_t idx = orc_hash_row(values[i], f, bin_size) * technical_bins + bin_ids[i];
            words[idx >> 6] |= 1ULL << (idx & 63);
        }
}

void orc_bulk_contains(const uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, const uint64_t *values,
                       uint64_t n, uint64_t *out, int threads)
{
    uint64_t const bin_words = (bins + 63) / 64;
    if (threads < 1)
        threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i)
    {
        uint64_t *res = out + (uint64_t)i * bin_words;
        const uint64_t *row = words + orc_hash_row(values[i], 0, bin_size) * bin_words;
        memcpy(res, row, bin_words * 8);
        for (int f = 1; f < hash_funs; ++f)
        {
            row = words + orc_hash_row(values[i], f, bin_size) * bin_words;
            for (uint64_t w = 0; w < bin_words; ++w)
                res[w] &= row[w];
        }
    }
}

uint64_t orc_bulk_contains_checksum(const uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs,
                                    const uint64_t *values, uint64_t n, int threads)
{
    uint64_t const bin_words = (bins + 63) / 64;
    uint64_t total = 0;
    if (threads < 1)
        threads = 1;
#pragma omp parallel num_threads(threads) reduction(+ : total)
    {
        uint64_t *res = (uint64_t *)malloc(bin_words * 8);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < (int64_t)n; ++i)
        {
            orc_bulk_contains(words, bins, bin_size, hash_funs, values + i, 1, res, 1);
            for (uint64_t w = 0; w < bin_words; ++w)
                total += res[w] * (w + 1);
        }
        free(res);
    }
    return total;
}

/* ------------------------------------------------------------------------------------------ a8 */
typedef struct
{
    uint64_t n_ibfs;
    const uint64_t *bins, *bin_size, *hash_funs, *table_len;
    const uint64_t *const *words, *const *next_ibf, *const *user_bin;
} hibf_view;

static void membership_impl(const hibf_view *h, uint64_t value, uint64_t ibf_idx, uint64_t *out_bits, uint64_t *scratch,
                            int depth)
{
    /* hierarchical_interleaved_bloom_filter.hpp:321-355 with values = {value}, threshold = 1 */
    uint64_t const bins = h->bins[ibf_idx];
    uint64_t const bin_words = (bins + 63) / 64;
    uint64_t *res = (uint64_t *)malloc(bin_words * 8);
    orc_bulk_contains(h->words[ibf_idx], bins, h->bin_size[ibf_idx], (int)h->hash_funs[ibf_idx], &value, 1, res, 1);
    uint64_t const n_tb = h->table_len[ibf_idx] < bins ? h->table_len[ibf_idx] : bins; /* result.size() = bin_count */
    unsigned sum = 0;
    for (uint64_t bin = 0; bin < n_tb; ++bin)
    {
        sum += (res[bin >> 6] >> (bin & 63)) & 1; /* bulk_count of one value */
        uint64_t const user_bin_id = h->user_bin[ibf_idx][bin];
        if (user_bin_id == UINT64_MAX - 1) /* bin_kind::deleted */
            continue;
        if (user_bin_id == UINT64_MAX) /* bin_kind::merged */
        {
            if (sum >= 1 && depth < 64)
                membership_impl(h, value, h->next_ibf[ibf_idx][bin], out_bits, scratch, depth + 1);
            sum = 0;
        }
        else if (bin + 1 == n_tb || user_bin_id != h->user_bin[ibf_idx][bin + 1])
        {
            if (sum >= 1)
                out_bits[user_bin_id >> 6] |= 1ULL << (user_bin_id & 63); /* populate_bitvector */
            sum = 0;
        }
    }
    free(res);
}

void orc_hibf_membership(uint64_t n_ibfs, uint64_t n_user_bins, const uint64_t *bins, const uint64_t *bin_size,
                         const uint64_t *hash_funs, const uint64_t *const *words, const uint64_t *const *next_ibf,
                         const uint64_t *const *user_bin, const uint64_t *table_len, const uint64_t *values, uint64_t n,
                         uint64_t *out, int threads)
{
    hibf_view h = {n_ibfs, bins, bin_size, hash_funs, table_len, words, next_ibf, user_bin};
    uint64_t const out_words = (n_user_bins + 63) / 64;
    if (threads < 1)
        threads = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i)
    {
        uint64_t *dst = out + (uint64_t)i * out_words;
        memset(dst, 0, out_words * 8);
        membership_impl(&h, values[i], 0, dst, NULL, 0);
    }
}

/* IBFIndex::populate_index + init_ibf (include/index_ibf.h:88-131) for records held in memory: every k-mer
 * decompose_record emits for record r is emplaced into technical bin rec_bin[r].  OpenMP over records with
 * atomic ORs (bit sets commute, so the result equals the reference's serial loop). */
void orc_emplace_records(uint64_t *words, uint64_t bins, uint64_t bin_size, int hash_funs, int molecule, int k,
                         int reduction, const uint8_t *text, const uint64_t *rec_offsets, const uint32_t *rec_bin,
                         uint64_t n_records, int threads)
{
    uint64_t const technical_bins = ((bins + 63) / 64) * 64;
    if (threads < 1)
        threads = 1;
#pragma omp parallel num_threads(threads)
    {
        uint64_t cap = 1 << 16;
        uint64_t *buf = (uint64_t *)malloc(cap * 8);
#pragma omp for schedule(dynamic, 64)
        for (int64_t r = 0; r < (int64_t)n_records; ++r)
        {
            uint64_t len = rec_offsets[r + 1] - rec_offsets[r];
            if (len + 2 > cap)
            {
                cap = len + 2;
                buf = (uint64_t *)realloc(buf, cap * 8);
            }
            uint64_t n = orc_decompose(molecule, k, reduction, text + rec_offsets[r], len, buf, cap);
            for (uint64_t i = 0; i < n; ++i)
                for (int f = 0; f < hash_funs; ++f)
                {
                    uint64_t idx = orc_hash_row(buf[i], f, bin_size) * technical_bins + rec_bin[r];
                    uint64_t bit = 1ULL << (idx & 63);
#pragma omp atomic
                    words[idx >> 6] |= bit;
                }
        }
        free(buf);
    }
}
