/* oracle/fm_oracle.h -- TEST INFRASTRUCTURE ONLY (CPU checker), never shipped or timed as
 * the product.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.
 *
 * Plain-C restatement of the reference's FM-index exact-match COUNT over a reference-built
 * .awfmi file.  Parity status: PINNED -- checked in tests/test_oracle.py against the
 * reference's golden vectors (tests/test_count_kmers.py:22-44 of the reference) and
 * against the reference's own compiled code (oracle/_ref, see oracle/Makefile).
 */
#ifndef FM_ORACLE_H
#define FM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ora_index ora_index;

/* Reads header, BWT blocks, prefix sums and seed table of an .awfmi v8 DNA index
 * (layout: AvxWindowFmIndex/src/AwFmFile.c:20-171).  Returns NULL on failure and stores a
 * negative code in *err (-1 open, -2 format, -3 version/alphabet, -4 alloc, -5 short read). */
ora_index *ora_index_load(const char *path, int *err);
void ora_index_free(ora_index *ix);
uint64_t ora_bwt_length(const ora_index *ix);
uint32_t ora_seed_length(const ora_index *ix);
void ora_prefix_sums(const ora_index *ix, uint64_t out[6]);

/* counts[q] = number of occurrences of seq[idx[q] : idx[q]+len[q]] in the indexed text,
 * five-letter alphabet (a c g t/u, anything else = x), truncated to u32 like the
 * reference (AwFmParallelSearch.c:136-157).  len[q]==0 or idx+len>seq_len -> returns -1.
 * *lf_steps (optional) accumulates the number of backward-extension steps executed. */
int ora_count(const ora_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
              const uint64_t *len, uint64_t n, uint32_t *counts, uint64_t *lf_steps);

/* Same, for the reverse complement of each k-mer exactly as newmap/search.py:677-697
 * builds it: bytes ACGTacgt are complemented, every other byte passes through unchanged,
 * then the string is reversed. */
int ora_count_rc(const ora_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                 const uint64_t *len, uint64_t n, uint32_t *counts, uint64_t *lf_steps);

/* search.py:769-871 in closed form; mask[i]=1 marks an ambiguous start position. */
void ora_upper_bounds(const uint8_t *mask, uint64_t num_kmers, uint64_t seg_len, uint32_t max_k,
                      uint32_t *ub);

/* search.py:383-548 (range mode) replayed per position over one segment buffer.
 * out[i] (u32) = minimum unique length or 0.  zero_info = {position, length, round} of the
 * zero-count query the reference would have reported first (position -1 if none).
 * Returns -1 on invalid arguments (incl. 0 < initial_len < min_k, where the reference
 * does not terminate). */
int ora_min_unique_range(const ora_index *ix, const uint8_t *seg, uint64_t seg_len,
                         uint64_t num_kmers, uint32_t min_k, uint32_t max_k, uint32_t initial_len,
                         int use_rc, int nthreads, uint32_t *out, int64_t zero_info[3],
                         uint64_t *lf_steps, uint64_t *queries);

/* search.py:551-644 (explicit k list) replayed per position. */
int ora_min_unique_list(const ora_index *ix, const uint8_t *seg, uint64_t seg_len,
                        uint64_t num_kmers, const uint32_t *klist, uint32_t nk, int use_rc,
                        int nthreads, uint32_t *out, int64_t zero_info[3], uint64_t *lf_steps,
                        uint64_t *queries);

#ifdef __cplusplus
}
#endif
#endif
