/* oracle/fm_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU checker); see fm_oracle.h.
 *
 * Scalar restatement of the reference's exact-match count.  Every function cites the
 * reference lines it follows (paths relative to /root/reference/AvxWindowFmIndex/src).
 * Written independently: plain loops and __builtin_popcountll, no SIMD; OpenMP only across independent positions.
 */
#include "fm_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_POSITIONS 256u /* AwFmIndex.h: AW_FM_POSITIONS_PER_FM_BLOCK */
#define BLOCK_BYTES 160u     /* struct AwFmNucleotideBlock, AwFmIndex.h:60-63 */

struct ora_index {
  uint64_t bwt_length;
  uint64_t n_blocks;
  uint32_t seed_len;
  uint8_t *blocks;     /* n_blocks * 160 B, file order */
  uint64_t prefix[6];  /* AwFmCreate.c:299-305 */
  uint64_t *seed;      /* 4^seed_len pairs {start,end}, AwFmKmerTable.c:24-35 */
  uint64_t seed_entries;
};

/* AwFmLetter.c:9-20 : case-fold by OR 0x20; a0 c1 g2 t/u3, everything else 4.
 * ('$' maps to 5 in the reference and then indexes past prefixSums -- undefined there;
 * the oracle folds it into 4 and the tests never feed it.) */
static inline unsigned letter_of(uint8_t ch) {
  switch (ch | 0x20) {
    case 'a': return 0;
    case 'c': return 1;
    case 'g': return 2;
    case 't': return 3;
    case 'u': return 3;
    default: return 4;
  }
}

/* AwFmLetter.c:91-118 (nucleotide branch): ambiguous = not a,c,g,t,u after tolower. */
static inline int is_ambiguous(uint8_t ch) { return letter_of(ch) == 4; }

static uint64_t rd_u64(const uint8_t *p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
}

/* Occ(letter, pos): occurrences of letter in BWT[0..pos] inclusive.
 * AwFmSearch.c:44-95 (block/local split), AwFmOccurrence.c:12-34 (plane logic),
 * AwFmSimdConfig.c:81-100 (mask = bits 0..local inclusive). */
static uint64_t occ(const ora_index *ix, unsigned letter, uint64_t pos) {
  const uint64_t b = pos / BLOCK_POSITIONS;
  const unsigned local = (unsigned)(pos % BLOCK_POSITIONS);
  const uint8_t *blk = ix->blocks + b * BLOCK_BYTES;
  uint64_t total = rd_u64(blk + 96 + 8 * letter);
  for (unsigned w = 0; w < 4; ++w) {
    if (w * 64 > local) break;
    const uint64_t p0 = rd_u64(blk + 0 + 8 * w);
    const uint64_t p1 = rd_u64(blk + 32 + 8 * w);
    const uint64_t p2 = rd_u64(blk + 64 + 8 * w);
    uint64_t v;
    switch (letter) {
      case 0: v = p2 & p1; break;        /* A 0b110 */
      case 1: v = p2 & p0; break;        /* C 0b101 */
      case 2: v = p1 & p0; break;        /* G 0b011 */
      case 3: v = ~p2 & ~p1 & p0; break; /* T 0b001 */
      default: v = ~p2 & ~p0 & p1; break; /* x 0b010 */
    }
    if (local / 64 == w) {
      const unsigned keep = local % 64; /* bits 0..keep */
      v &= (keep == 63) ? ~0ULL : ((1ULL << (keep + 1)) - 1);
    }
    total += (uint64_t)__builtin_popcountll(v);
  }
  return total;
}

/* One backward-extension step, AwFmSearch.c:44-95. */
static inline void lf_step(const ora_index *ix, uint64_t *start, uint64_t *end, unsigned letter) {
  const uint64_t c = ix->prefix[letter];
  const uint64_t ns = c + occ(ix, letter, *start - 1);
  const uint64_t ne = c + occ(ix, letter, *end) - 1;
  *start = ns;
  *end = ne;
}

typedef uint8_t (*fetch_fn)(const uint8_t *seq, uint64_t seq_len, uint64_t idx, uint64_t len, uint64_t j);

/* j-th character (0 = leftmost) of the forward k-mer seq[idx:idx+len]. */
static uint8_t fetch_fwd(const uint8_t *seq, uint64_t seq_len, uint64_t idx, uint64_t len, uint64_t j) {
  (void)seq_len; (void)len;
  return seq[idx + j];
}

/* j-th character of the k-mer newmap builds for the other strand (search.py:22,682-689):
 * translate ACGTacgt -> TGCAtgca (all other bytes unchanged), then reverse. */
static uint8_t fetch_rc(const uint8_t *seq, uint64_t seq_len, uint64_t idx, uint64_t len, uint64_t j) {
  (void)seq_len;
  uint8_t ch = seq[idx + len - 1 - j];
  switch (ch) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'T': return 'A';
    case 'a': return 't';
    case 'c': return 'g';
    case 'g': return 'c';
    case 't': return 'a';
    default: return ch;
  }
}

/* AwFmParallelSearch.c:182-255 for one query: seed (table or from-scratch over the last
 * min(len,seedLen) characters), then extend leftwards while the range is valid. */
static uint32_t count_one(const ora_index *ix, fetch_fn fetch, const uint8_t *seq, uint64_t seq_len,
                          uint64_t idx, uint64_t len, uint64_t *steps) {
  const uint64_t sl = ix->seed_len;
  uint64_t start, end;
  uint64_t next; /* number of characters (from the right) already consumed */
  int use_table = len >= sl;
  if (use_table) { /* AwFmKmerTable.c:8-21 */
    for (uint64_t j = len - sl; j < len; ++j)
      if (is_ambiguous(fetch(seq, seq_len, idx, len, j))) { use_table = 0; break; }
  }
  if (use_table) { /* AwFmKmerTable.c:24-35 */
    uint64_t t = 0;
    for (uint64_t j = len - sl; j < len; ++j) t = t * 4 + letter_of(fetch(seq, seq_len, idx, len, j));
    start = ix->seed[2 * t];
    end = ix->seed[2 * t + 1];
    next = sl;
  } else { /* AwFmParallelSearch.c:195-209 + AwFmSearch.c:424-435 */
    const uint64_t sub = len < sl ? len : sl;
    unsigned l = letter_of(fetch(seq, seq_len, idx, len, len - 1));
    start = ix->prefix[l];
    end = ix->prefix[l + 1] - 1;
    next = 1;
    while (next < sub && start <= end) {
      lf_step(ix, &start, &end, letter_of(fetch(seq, seq_len, idx, len, len - 1 - next)));
      ++next;
      if (steps) ++*steps;
    }
    next = sub; /* the extension loop below resumes at seedLen+1 regardless (…:229-233) */
  }
  /* AwFmParallelSearch.c:224-255: letter index seedLen+1 .. len from the right. */
  if (next < sl) next = sl;
  while (next < len && start <= end) {
    lf_step(ix, &start, &end, letter_of(fetch(seq, seq_len, idx, len, len - 1 - next)));
    ++next;
    if (steps) ++*steps;
  }
  /* AwFmIndexStruct.c:127-130, truncated to u32 at AwFmParallelSearch.c:156 */
  return (start <= end) ? (uint32_t)(end - start + 1) : 0u;
}

static int count_many(const ora_index *ix, fetch_fn fetch, const uint8_t *seq, uint64_t seq_len,
                      const uint64_t *idx, const uint64_t *len, uint64_t n, uint32_t *counts,
                      uint64_t *lf_steps) {
  for (uint64_t q = 0; q < n; ++q) {
    if (len[q] == 0 || idx[q] + len[q] > seq_len) return -1;
  }
  for (uint64_t q = 0; q < n; ++q)
    counts[q] = count_one(ix, fetch, seq, seq_len, idx[q], len[q], lf_steps);
  return 0;
}

int ora_count(const ora_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
              const uint64_t *len, uint64_t n, uint32_t *counts, uint64_t *lf_steps) {
  return count_many(ix, fetch_fwd, seq, seq_len, idx, len, n, counts, lf_steps);
}

int ora_count_rc(const ora_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                 const uint64_t *len, uint64_t n, uint32_t *counts, uint64_t *lf_steps) {
  return count_many(ix, fetch_rc, seq, seq_len, idx, len, n, counts, lf_steps);
}

uint64_t ora_bwt_length(const ora_index *ix) { return ix->bwt_length; }
uint32_t ora_seed_length(const ora_index *ix) { return ix->seed_len; }
void ora_prefix_sums(const ora_index *ix, uint64_t out[6]) { memcpy(out, ix->prefix, sizeof ix->prefix); }

void ora_index_free(ora_index *ix) {
  if (!ix) return;
  free(ix->blocks);
  free(ix->seed);
  free(ix);
}

/* AwFmFile.c:174-300: "AwFmIndex\n" | u32 version(8) | u32 flags | u8 saRatio | u8 seedLen |
 * u8 alphabet(2 = DNA) | u8 storeSeq | u64 bwtLength | blocks | prefixSums[6] | seed table. */
ora_index *ora_index_load(const char *path, int *err) {
  int e = 0;
  ora_index *ix = NULL;
  FILE *f = fopen(path, "rb");
  if (!f) { e = -1; goto fail; }
  uint8_t hdr[30];
  if (fread(hdr, 1, 30, f) != 30) { e = -5; goto fail; }
  if (memcmp(hdr, "AwFmIndex\n", 10) != 0) { e = -2; goto fail; }
  uint32_t version;
  memcpy(&version, hdr + 10, 4);
  if (version != 8 || hdr[20] != 2) { e = -3; goto fail; }
  ix = calloc(1, sizeof *ix);
  if (!ix) { e = -4; goto fail; }
  ix->seed_len = hdr[19];
  ix->bwt_length = rd_u64(hdr + 22);
  if (ix->bwt_length == 0 || ix->seed_len > 16) { e = -2; goto fail; }
  ix->n_blocks = 1 + (ix->bwt_length - 1) / BLOCK_POSITIONS; /* AwFmIndexStruct.c:102-104 */
  ix->seed_entries = 1ULL << (2 * ix->seed_len);
  ix->blocks = malloc(ix->n_blocks * BLOCK_BYTES);
  ix->seed = malloc(ix->seed_entries * 16);
  if (!ix->blocks || !ix->seed) { e = -4; goto fail; }
  if (fread(ix->blocks, BLOCK_BYTES, ix->n_blocks, f) != ix->n_blocks) { e = -5; goto fail; }
  if (fread(ix->prefix, 8, 6, f) != 6) { e = -5; goto fail; }
  if (fread(ix->seed, 16, ix->seed_entries, f) != ix->seed_entries) { e = -5; goto fail; }
  fclose(f);
  if (err) *err = 0;
  return ix;
fail:
  if (f) fclose(f);
  ora_index_free(ix);
  if (err) *err = e;
  return NULL;
}

/* ------------------------------------------------------------------------------------------
 * Per-position restatement of newmap/search.py (paths relative to /root/reference).
 * The reference iterates "for each round, for all unfinished positions"; positions never
 * interact, so the same trajectory is replayed here one position at a time.  The only
 * cross-position effect -- the RuntimeError raised in the first round that contains a zero
 * count, naming the lowest such position (search.py:699-722) -- is reproduced by keeping the
 * lexicographically smallest (round, position) zero event.
 * ---------------------------------------------------------------------------------------- */
#include <omp.h>

/* search.py:23,744-766 : a start position is searchable iff its byte is one of b'ACGTactg'. */
static inline int start_allowed(uint8_t ch) {
  switch (ch) {
    case 'A': case 'C': case 'G': case 'T': case 'a': case 'c': case 'g': case 't': return 1;
    default: return 0;
  }
}

/* search.py:769-871 restated as a closed form (proved equivalent case by case in DESIGN.md,
 * pinned by the 16 vectors of tests/test_upper_search_bound_truncation.py):
 *   ub[i] = min(max_k, (first ambiguous START position j>i with j<num_kmers) - i)
 *   or, when no ambiguous start position follows i, min(max_k, seg_len - i).
 * `mask` has num_kmers entries (1 = ambiguous).  Entries at ambiguous positions are left at
 * max_k like the reference leaves them. */
void ora_upper_bounds(const uint8_t *mask, uint64_t num_kmers, uint64_t seg_len, uint32_t max_k,
                      uint32_t *ub) {
  uint64_t next_amb = UINT64_MAX; /* none seen to the right yet */
  for (uint64_t r = num_kmers; r-- > 0;) {
    if (mask[r]) {
      next_amb = r;
      ub[r] = max_k;
    } else {
      const uint64_t room = (next_amb == UINT64_MAX) ? seg_len - r : next_amb - r;
      ub[r] = room < max_k ? (uint32_t)room : max_k;
    }
  }
}

typedef struct {
  uint64_t round, pos, len;
  int seen;
} zero_event;

static inline void note_zero(zero_event *z, uint64_t round, uint64_t pos, uint64_t len) {
  if (!z->seen || round < z->round || (round == z->round && pos < z->pos)) {
    z->seen = 1; z->round = round; z->pos = pos; z->len = len;
  }
}

/* search.py:647-697 for one (position, length): forward count, plus the count of the
 * complemented-reversed k-mer unless --norc; summed in uint32. */
static inline uint32_t both_strands(const ora_index *ix, const uint8_t *seg, uint64_t seg_len,
                                    uint64_t i, uint64_t k, int use_rc, uint64_t *steps) {
  uint32_t c = count_one(ix, fetch_fwd, seg, seg_len, i, k, steps);
  if (use_rc) c += count_one(ix, fetch_rc, seg, seg_len, i, k, steps);
  return c;
}

int ora_min_unique_range(const ora_index *ix, const uint8_t *seg, uint64_t seg_len,
                         uint64_t num_kmers, uint32_t min_k, uint32_t max_k, uint32_t initial_len,
                         int use_rc, int nthreads, uint32_t *out, int64_t zero_info[3],
                         uint64_t *lf_steps, uint64_t *queries) {
  if (min_k == 0 || min_k > max_k || num_kmers > seg_len) return -1;
  if (initial_len && initial_len < min_k) return -1; /* the reference never terminates here */
  uint8_t *mask = malloc(num_kmers ? num_kmers : 1);
  uint32_t *ub = malloc((num_kmers ? num_kmers : 1) * sizeof *ub);
  if (!mask || !ub) { free(mask); free(ub); return -4; }
  for (uint64_t i = 0; i < num_kmers; ++i) mask[i] = !start_allowed(seg[i]);
  ora_upper_bounds(mask, num_kmers, seg_len, max_k, ub);
  zero_event zall = {0, 0, 0, 0};
  uint64_t steps_all = 0, queries_all = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel num_threads(nthreads)
  {
    zero_event z = {0, 0, 0, 0};
    uint64_t steps = 0, nq = 0;
#pragma omp for schedule(dynamic, 4096)
    for (uint64_t i = 0; i < num_kmers; ++i) {
      out[i] = 0;
      if (mask[i] || ub[i] < min_k) continue; /* search.py:399,437 */
      uint32_t lo = min_k, hi = ub[i], best = 0;
      uint32_t q = (lo + hi) / 2; /* search.py:424-426 */
      if (initial_len && q > initial_len) q = initial_len; /* :429-433 */
      for (uint64_t round = 0;; ++round) {
        const uint32_t c = both_strands(ix, seg, seg_len, i, q, use_rc, &steps);
        ++nq;
        if (c == 0) note_zero(&z, round, i, q);
        if (c == 1) { /* search.py:489-508,524-527 */
          if (best == 0 || q < best) best = q;
          if (q == lo) break;
          hi = q - 1;
        } else if (c > 1) { /* :513-517,532-535 */
          if (q == hi) break;
          lo = q + 1;
        }
        /* c == 0: neither rule fires in the reference; it has already raised. */
        if (c == 0) break;
        q = (lo + hi) / 2; /* :540-542 */
      }
      out[i] = best;
    }
#pragma omp critical
    {
      if (z.seen) note_zero(&zall, z.round, z.pos, z.len);
      steps_all += steps;
      queries_all += nq;
    }
  }
  if (zero_info) {
    zero_info[0] = zall.seen ? (int64_t)zall.pos : -1;
    zero_info[1] = (int64_t)zall.len;
    zero_info[2] = (int64_t)zall.round;
  }
  if (lf_steps) *lf_steps += steps_all;
  if (queries) *queries += queries_all;
  free(mask);
  free(ub);
  return 0;
}

/* search.py:551-644.  k-mers are cut at the end of the buffer, the cut length is what is
 * counted, the nominal k is what is recorded (:590,:600,:634); an upper-case 'N' anywhere in
 * the (cut) k-mer retires the position (:593-596); a total of exactly 1 finishes it (:639). */
int ora_min_unique_list(const ora_index *ix, const uint8_t *seg, uint64_t seg_len,
                        uint64_t num_kmers, const uint32_t *klist, uint32_t nk, int use_rc,
                        int nthreads, uint32_t *out, int64_t zero_info[3], uint64_t *lf_steps,
                        uint64_t *queries) {
  if (nk == 0 || num_kmers > seg_len) return -1;
  for (uint32_t j = 0; j < nk; ++j) if (klist[j] == 0) return -1;
  zero_event zall = {0, 0, 0, 0};
  uint64_t steps_all = 0, queries_all = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel num_threads(nthreads)
  {
    zero_event z = {0, 0, 0, 0};
    uint64_t steps = 0, nq = 0;
#pragma omp for schedule(dynamic, 4096)
    for (uint64_t i = 0; i < num_kmers; ++i) {
      out[i] = 0;
      if (!start_allowed(seg[i])) continue;
      for (uint32_t j = 0; j < nk; ++j) {
        const uint64_t k = klist[j];
        const uint64_t cut = (i + k <= seg_len) ? k : seg_len - i;
        if (memchr(seg + i, 'N', cut)) break;
        const uint32_t c = both_strands(ix, seg, seg_len, i, cut, use_rc, &steps);
        ++nq;
        if (c == 0) { note_zero(&z, j, i, cut); break; }
        if (c == 1) { out[i] = (uint32_t)k; break; }
      }
    }
#pragma omp critical
    {
      if (z.seen) note_zero(&zall, z.round, z.pos, z.len);
      steps_all += steps;
      queries_all += nq;
    }
  }
  if (zero_info) {
    zero_info[0] = zall.seen ? (int64_t)zall.pos : -1;
    zero_info[1] = (int64_t)zall.len;
    zero_info[2] = (int64_t)zall.round;
  }
  if (lf_steps) *lf_steps += steps_all;
  if (queries) *queries += queries_all;
  return 0;
}

/* Several sequences x several indexes (search.py:461,489-517 range; :575,:627-639 list; :656-697 the count):
 * the count of a (position, length) is the uint32 sum over every index and every sequence of both_strands(); a
 * length is unique when the sum equals the number of SEQUENCES.  Start mask, 'N' rule and upper bounds come from
 * segs[0] only (:391-394, :557-560).  A sum in (0, n_seq) matches neither rule of the range loop (:489-535): the
 * reference would repeat the query for ever; reported as return code 1 (outputs for other positions are valid). */
static uint32_t multi_count(const ora_index *const *ixs, uint32_t n_ix, const uint8_t *const *segs, uint32_t n_seq,
                            uint64_t seg_len, uint64_t i, uint64_t k, int use_rc, uint64_t *steps) {
  uint32_t c = 0;
  for (uint32_t x = 0; x < n_ix; ++x)
    for (uint32_t s = 0; s < n_seq; ++s) c += both_strands(ixs[x], segs[s], seg_len, i, k, use_rc, steps);
  return c;
}

int ora_min_unique_multi(const ora_index *const *ixs, uint32_t n_ix, const uint8_t *const *segs, uint32_t n_seq,
                         uint64_t seg_len, uint64_t num_kmers, int range_mode, const uint32_t *klist, uint32_t nk,
                         uint32_t initial_len, int use_rc, uint32_t *out, int64_t zero_info[3], uint64_t *lf_steps,
                         uint64_t *queries) {
  if (n_ix == 0 || n_seq == 0 || nk == 0 || num_kmers > seg_len) return -1;
  const uint8_t *seg = segs[0];
  uint32_t min_k = klist[0], max_k = klist[0];
  for (uint32_t j = 0; j < nk; ++j) {
    if (klist[j] == 0) return -1;
    if (klist[j] < min_k) min_k = klist[j];
    if (klist[j] > max_k) max_k = klist[j];
  }
  if (range_mode && initial_len && initial_len < min_k) return -1;
  uint8_t *mask = calloc(num_kmers ? num_kmers : 1, 1);
  uint32_t *ub = malloc((num_kmers ? num_kmers : 1) * sizeof *ub);
  if (!mask || !ub) { free(mask); free(ub); return -4; }
  for (uint64_t i = 0; i < num_kmers; ++i) mask[i] = !start_allowed(seg[i]);
  if (range_mode) ora_upper_bounds(mask, num_kmers, seg_len, max_k, ub);
  zero_event z = {0, 0, 0, 0};
  uint64_t steps = 0, nq = 0;
  int shorted = 0;
  for (uint64_t i = 0; i < num_kmers; ++i) {
    out[i] = 0;
    if (mask[i]) continue;
    uint32_t best = 0;
    if (range_mode) {
      if (ub[i] < min_k) continue;
      uint32_t lo = min_k, hi = ub[i];
      uint32_t q = (lo + hi) / 2;
      if (initial_len && q > initial_len) q = initial_len;
      for (uint64_t round = 0;; ++round) {
        const uint32_t c = multi_count(ixs, n_ix, segs, n_seq, seg_len, i, q, use_rc, &steps);
        ++nq;
        if (c == 0) { note_zero(&z, round, i, q); break; }
        if (c == n_seq) {
          if (best == 0 || q < best) best = q;
          if (q == lo) break;
          hi = q - 1;
        } else if (c > n_seq) {
          if (q == hi) break;
          lo = q + 1;
        } else { shorted = 1; break; }
        q = (lo + hi) / 2;
      }
    } else {
      for (uint32_t j = 0; j < nk; ++j) {
        const uint64_t k = klist[j];
        const uint64_t cut = (i + k <= seg_len) ? k : seg_len - i;
        if (memchr(seg + i, 'N', cut)) break;
        const uint32_t c = multi_count(ixs, n_ix, segs, n_seq, seg_len, i, cut, use_rc, &steps);
        ++nq;
        if (c == 0) { note_zero(&z, j, i, cut); break; }
        if (c == n_seq && best == 0) best = (uint32_t)k; /* :627-636 */
        if (c == 1) break;                               /* :639 */
      }
    }
    out[i] = best;
  }
  if (zero_info) {
    zero_info[0] = z.seen ? (int64_t)z.pos : -1;
    zero_info[1] = (int64_t)z.len;
    zero_info[2] = (int64_t)z.round;
  }
  if (lf_steps) *lf_steps += steps;
  if (queries) *queries += nq;
  free(mask);
  free(ub);
  return shorted;
}
