// nm_core.cuh -- per-thread bodies of every kernel on the `newmap search` hot path.
//
// Each kernel in nm_kernels.cu is "one thread = one call of a function in this header", so the
// same bodies can be compiled by a host compiler and driven serially by the CPU-side logic tests
// (tests/hostemu; test infrastructure, never loaded by the product).  Everything here is
// integer arithmetic: u64/u32 suffix-array pointers, popcount rank, u8/16/32 lengths.
//
// What is restated (paths relative to hoffmangroup/newmap, AwFm = AvxWindowFmIndex/src):
//   letter coding                AwFm/AwFmLetter.c:9-20
//   block layout / rank          AwFm/AwFmIndex.h:60-63, AwFmOccurrence.c:12-34, AwFmSimdConfig.c:81-100
//   backward-extension step      AwFm/AwFmSearch.c:44-95
//   seed table lookup            AwFm/AwFmKmerTable.c:8-35
//   batched count                AwFm/AwFmParallelSearch.c:136-255
//   strand handling, zero check  newmap/search.py:647-724
//   range / list search          newmap/search.py:383-548, :551-644, bounds :744-871
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define NM_HD __host__ __device__ __forceinline__
#else
#define NM_HD inline
#endif

namespace nm {

// ---------------------------------------------------------------------------------------------
// small portability layer (device intrinsics <-> host builtins)
// ---------------------------------------------------------------------------------------------
NM_HD int popc64(uint64_t v) {
#if defined(__CUDA_ARCH__)
  return __popcll(v);
#else
  return __builtin_popcountll(v);
#endif
}
NM_HD uint64_t ld64(const uint64_t *p) {
#if defined(__CUDA_ARCH__)
  return __ldg((const unsigned long long *)p);
#else
  return *p;
#endif
}
NM_HD uint32_t ld32(const uint32_t *p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
NM_HD uint8_t ld8(const uint8_t *p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
struct u32x4 {
  uint32_t x, y, z, w;
};
NM_HD u32x4 ld128(const uint32_t *p) {
#if defined(__CUDA_ARCH__)
  const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
  return u32x4{v.x, v.y, v.z, v.w};
#else
  return u32x4{p[0], p[1], p[2], p[3]};
#endif
}
NM_HD void atomic_add64(unsigned long long *p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicAdd(p, v);
#else
  *p += v;
#endif
}
NM_HD void atomic_min64(unsigned long long *p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicMin(p, v);
#else
  if (v < *p) *p = v;
#endif
}
NM_HD void atomic_or64(unsigned long long *p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicOr(p, v);
#else
  *p |= v;
#endif
}

// ---------------------------------------------------------------------------------------------
// constants
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kNativeBlockBytes = 160;   // AwFmIndex.h:60-63: 3 x 256-bit planes + 8 x u64
constexpr uint32_t kRank32Words = 8;          // derived layout: 4 x u32 counts + 2 x u64 planes per 64 positions
constexpr uint32_t kPoison = 0xFFFFFFFFu;     // count word of a rank32 block that holds an x or $ symbol

// segment codes (one byte per base, written by prep_thread)
constexpr uint8_t kLetterMask = 0x07;  // 0..3 = a c g t/u, 4 = anything else ("x")
constexpr uint8_t kIrregular = 0x08;   // letter x, or U/u (complemented differently by search.py:22): generic kernel only
constexpr uint8_t kBarrier = 0x10;     // no queried k-mer may contain this position
constexpr uint8_t kNoStart = 0x20;     // search.py:23,744-766: start byte not in b'ACGTactg'

constexpr uint32_t kNone = 0u;               // G: no unique suffix within bounds
constexpr uint32_t kInf = 0xFFFFFFFFu;       // R: other strand still present at the longest length
constexpr uint32_t kPending = 0xFFFFFFFEu;   // G/R: to be computed by the generic kernel
constexpr uint32_t kSeedInvalid = 0xFFFFFFFFu;
constexpr uint32_t kNoBarrier = 0xFFFFFFFFu;
constexpr uint32_t kTile = 64;               // nb[] tile (positions per thread of nb_tile_thread)

// counters[] slots
enum { CTR_LF = 0, CTR_QUERIES = 1, CTR_GENERIC = 2, CTR_ANOMALY = 3, CTR_ZERO_KEY = 4, CTR_REF_LF = 5, CTR_N = 8 };

struct SeedEntry32 {
  uint32_t start, len;
};

struct DevIndex {
  const uint8_t *native;      // (n_blocks + 1) * 160 B, file order, one zero-plane pad block
  const uint32_t *rank32;     // (n64 + 1) * 8 u32, or nullptr
  const SeedEntry32 *seed32;  // 4^seed_len entries, or nullptr
  const uint64_t *seed64;     // 4^seed_len {start,end} pairs exactly as in the file
  uint64_t C[6];              // prefixSums, AwFmCreate.c:299-305
  uint64_t bwt_len;
  uint32_t seed_len;
};

// AwFmLetter.c:9-20: case-fold with |0x20; a0 c1 g2 t3 u3, everything else 4
NM_HD unsigned letter_of(uint8_t ch) {
  const uint8_t c = ch | 0x20;
  return c == 'a' ? 0u : c == 'c' ? 1u : c == 'g' ? 2u : (c == 't' || c == 'u') ? 3u : 4u;
}
// search.py:23: ALLOWED_BASES = b'ACGTactg'
NM_HD bool start_allowed(uint8_t ch) {
  const uint8_t c = ch | 0x20;
  return c == 'a' || c == 'c' || c == 'g' || c == 't';
}
// search.py:22: bytes.maketrans(b'ACGTacgt', b'TGCAtgca'); every other byte passes through
NM_HD uint8_t complement_byte(uint8_t ch) {
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

// ---------------------------------------------------------------------------------------------
// rank on the unmodified .awfmi block (exclusive: occurrences of `letter` in BWT[0, p))
// AwFmOccurrence.c:12-34 plane logic with codes A=110 C=101 G=011 T=001 x=010 (AwFmLetter.c:37-40)
// ---------------------------------------------------------------------------------------------
NM_HD uint64_t occ_native(const uint8_t *blocks, unsigned letter, uint64_t p) {
  const uint64_t *b = reinterpret_cast<const uint64_t *>(blocks + (p >> 8) * kNativeBlockBytes);
  const unsigned off = (unsigned)(p & 255);
  uint64_t total = ld64(b + 12 + letter);
  const unsigned wlast = off >> 6;
  for (unsigned w = 0; w <= wlast; ++w) {
    const uint64_t mask = (w < wlast) ? ~0ull : ((1ull << (off & 63)) - 1ull);
    if (!mask) break;
    const uint64_t p0 = ld64(b + w), p1 = ld64(b + 4 + w), p2 = ld64(b + 8 + w);
    uint64_t v;
    switch (letter) {
      case 0: v = p2 & p1; break;
      case 1: v = p2 & p0; break;
      case 2: v = p1 & p0; break;
      case 3: v = ~p2 & ~p1 & p0; break;
      default: v = ~p2 & ~p0 & p1; break;
    }
    total += (uint64_t)popc64(v & mask);
  }
  return total;
}

struct Range64 {
  uint64_t lo, hi;  // half-open SA interval; the reference's inclusive {start,end} = {lo, hi-1}
};

// AwFmSearch.c:44-95: start' = C[l] + Occ(l,start-1), end' = C[l] + Occ(l,end) - 1
NM_HD void lf_native(const DevIndex &ix, Range64 &r, unsigned letter) {
  const uint64_t c = ix.C[letter];
  r.lo = c + occ_native(ix.native, letter, r.lo);
  r.hi = c + occ_native(ix.native, letter, r.hi);
}

// ---------------------------------------------------------------------------------------------
// rank on the derived 32-byte layout (a,c,g,t only; bwt_len < 2^32-1)
// block = { u32 count[4] | u64 plane0 | u64 plane1 }, code = plane1<<1 | plane0, 64 positions.
// A block that holds an x or $ symbol carries kPoison in all four counts.
// ---------------------------------------------------------------------------------------------
NM_HD uint32_t occ_rank32(const uint32_t *rank32, unsigned letter, uint32_t p, bool &poison) {
  const uint32_t *w = rank32 + (size_t)(p >> 6) * kRank32Words;
  const uint32_t c = ld32(w + letter);
  const u32x4 pl = ld128(w + 4);
  const uint64_t p0 = (uint64_t)pl.x | ((uint64_t)pl.y << 32);
  const uint64_t p1 = (uint64_t)pl.z | ((uint64_t)pl.w << 32);
  const uint64_t m = ((letter & 1u) ? p0 : ~p0) & ((letter & 2u) ? p1 : ~p1);
  const uint64_t mask = (1ull << (p & 63u)) - 1ull;
  poison = poison || (c == kPoison);
  return c + (uint32_t)popc64(m & mask);
}

// ---------------------------------------------------------------------------------------------
// exact-match count of one k-mer with the reference's own strategy (AwFmParallelSearch.c:182-255):
// seed table when len >= seedLen and the last seedLen characters are unambiguous, otherwise the
// last min(len,seedLen) characters from scratch; then leftwards while the range is valid.
// RC = false: k-mer is seq[idx, idx+len).  RC = true: the k-mer search.py:682-689 builds for the
// other strand (complement ACGTacgt, pass everything else through, reverse).
// ---------------------------------------------------------------------------------------------
template <bool RC>
NM_HD uint8_t kmer_char(const uint8_t *seq, uint64_t idx, uint64_t len, uint64_t j) {
  return RC ? complement_byte(ld8(seq + idx + len - 1 - j)) : ld8(seq + idx + j);
}

template <bool RC>
NM_HD uint32_t count_one(const DevIndex &ix, const uint8_t *seq, uint64_t idx, uint64_t len, uint32_t &steps) {
  const uint64_t sl = ix.seed_len;
  Range64 r;
  bool use_table = len >= sl;
  uint64_t key = 0;
  if (use_table) {  // AwFmKmerTable.c:8-21 and :24-35 (leftmost seed character most significant)
    for (uint64_t j = len - sl; j < len; ++j) {
      const unsigned l = letter_of(kmer_char<RC>(seq, idx, len, j));
      if (l == 4u) { use_table = false; break; }
      key = key * 4u + l;
    }
  }
  uint64_t next;  // characters consumed from the right
  if (use_table) {
    const uint64_t s = ld64(ix.seed64 + 2 * key), e = ld64(ix.seed64 + 2 * key + 1);
    r.lo = s;
    r.hi = (s <= e) ? e + 1 : s;
    next = sl;
  } else {  // AwFmParallelSearch.c:195-209, AwFmSearch.c:424-435
    const uint64_t sub = len < sl ? len : sl;
    const unsigned l = letter_of(kmer_char<RC>(seq, idx, len, len - 1));
    r.lo = ix.C[l];
    r.hi = ix.C[l + 1];
    next = 1;
    while (next < sub && r.lo < r.hi) {
      lf_native(ix, r, letter_of(kmer_char<RC>(seq, idx, len, len - 1 - next)));
      ++next;
      ++steps;
    }
    next = sub;
  }
  if (next < sl) next = sl;  // the extension loop starts at letter seedLen+1 from the right (:229-233)
  while (next < len && r.lo < r.hi) {
    lf_native(ix, r, letter_of(kmer_char<RC>(seq, idx, len, len - 1 - next)));
    ++next;
    ++steps;
  }
  return (r.lo < r.hi) ? (uint32_t)(r.hi - r.lo) : 0u;  // AwFmIndexStruct.c:127-130, u32 at AwFmParallelSearch.c:156
}

// search.py:647-697 for one (position, length): forward count plus, unless --norc, the other strand
NM_HD uint32_t count_both(const DevIndex &ix, const uint8_t *seg, uint64_t i, uint64_t k, bool use_rc,
                          uint32_t &steps) {
  uint32_t c = count_one<false>(ix, seg, i, k, steps);
  if (use_rc) c += count_one<true>(ix, seg, i, k, steps);
  return c;
}

// ---------------------------------------------------------------------------------------------
// segment description shared by the search kernels
// ---------------------------------------------------------------------------------------------
struct SegParams {
  const uint8_t *seg;    // raw bytes, seg_len
  uint8_t *codes;        // seg_len
  uint32_t *seedkey;     // seg_len: packed seed_len letters starting at p, or kSeedInvalid
  uint32_t *nb;          // seg_len: position of the first barrier >= p inside p's tile, or kNoBarrier
  uint32_t *tile_first;  // ceil(seg_len / kTile): first barrier of the tile, or kNoBarrier
  uint32_t *G;           // seg_len + 1: length of the shortest unique suffix ending at e (forward strand)
  uint32_t *R;           // seg_len: shortest length from i at which the other strand has no match
  void *out;             // num_kmers entries of out_width bytes
  const uint32_t *klist; // list mode: nk lengths in the order given
  unsigned long long *counters;  // CTR_N slots
  uint32_t seg_len, num_kmers;
  uint32_t kmin, kmax, nk, initial;
  int mode;       // 0 range, 1 list
  int use_rc;
  int out_width;  // 1, 2, 4
  int fast;       // rank32 + seed table usable for this search (kmin >= seed_len, bwt_len < 2^32-1)
  int want_ref_lf;
};

NM_HD void store_out(const SegParams &P, uint32_t i, uint32_t v) {
  if (P.out_width == 1) reinterpret_cast<uint8_t *>(P.out)[i] = (uint8_t)v;
  else if (P.out_width == 2) reinterpret_cast<uint16_t *>(P.out)[i] = (uint16_t)v;
  else reinterpret_cast<uint32_t *>(P.out)[i] = v;
}

// ---------------------------------------------------------------------------------------------
// (T) the reference's own trajectory, one position per thread.  search.py:383-548 / :551-644
// replayed per position (positions never interact); the zero-count event the reference would
// report first is the smallest (round, position) -- kept with atomicMin, its length in G[pos].
// ---------------------------------------------------------------------------------------------
NM_HD void note_zero(const SegParams &P, uint32_t round, uint32_t pos, uint32_t len) {
  P.G[pos] = len;  // a position sees at most one zero event (the search of that position stops there)
  atomic_min64(&P.counters[CTR_ZERO_KEY], ((unsigned long long)round << 32) | pos);
}

// closed form of search.py:769-871 (see oracle/fm_oracle.c ora_upper_bounds): distance to the next
// ambiguous START position (first num_kmers bytes only), else to the end of the buffer, capped at max_k
NM_HD uint32_t next_barrier(const SegParams &P, uint32_t i) {
  uint32_t v = P.nb[i];
  if (v != kNoBarrier) return v;
  const uint32_t ntiles = (P.seg_len + kTile - 1) / kTile;
  for (uint32_t t = i / kTile + 1; t < ntiles && (uint64_t)t * kTile < (uint64_t)i + P.kmax; ++t) {
    v = P.tile_first[t];
    if (v != kNoBarrier) return v;
  }
  return kNoBarrier;
}

NM_HD uint32_t length_limit(const SegParams &P, uint32_t i) {
  uint32_t lim = P.seg_len - i;
  if (lim > P.kmax) lim = P.kmax;
  const uint32_t b = next_barrier(P, i);
  if (b != kNoBarrier && b - i < lim) lim = b - i;
  return lim;
}

NM_HD void trajectory_thread(const DevIndex &ix, const SegParams &P, uint32_t i, uint32_t &steps, uint32_t &nq) {
  uint32_t best = 0;
  const uint8_t ch = ld8(P.seg + i);
  if (start_allowed(ch)) {
    if (P.mode == 0) {
      const uint32_t ub = length_limit(P, i);
      if (ub >= P.kmin) {                       // search.py:437
        uint32_t lo = P.kmin, hi = ub;
        uint32_t q = (lo + hi) / 2;             // :424-426
        if (P.initial && q > P.initial) q = P.initial;  // :429-433
        for (uint32_t round = 0;; ++round) {
          const uint32_t c = count_both(ix, P.seg, i, q, P.use_rc != 0, steps);
          ++nq;
          if (c == 0) { note_zero(P, round, i, q); break; }
          if (c == 1) {                         // :489-508, :524-527
            if (best == 0 || q < best) best = q;
            if (q == lo) break;
            hi = q - 1;
          } else {                              // :513-517, :532-535
            if (q == hi) break;
            lo = q + 1;
          }
          q = (lo + hi) / 2;                    // :540-542
        }
      }
    } else {
      const uint32_t b = next_barrier(P, i);    // list mode: barriers are upper-case 'N' (:593-596)
      for (uint32_t j = 0; j < P.nk; ++j) {
        const uint32_t k = P.klist[j];
        const uint32_t cut = (k <= P.seg_len - i) ? k : P.seg_len - i;  // :590,:600
        if (b != kNoBarrier && b - i < cut) break;
        const uint32_t c = count_both(ix, P.seg, i, cut, P.use_rc != 0, steps);
        ++nq;
        if (c == 0) { note_zero(P, j, i, cut); break; }
        if (c == 1) { best = k; break; }        // :627-639 (nominal k recorded)
      }
    }
  }
  store_out(P, i, best);
}

// ---------------------------------------------------------------------------------------------
// (P) prep: raw bytes -> codes + packed seed keys.  One thread per position.
// range mode: barrier = ambiguous start position among the first num_kmers bytes (search.py:744-871;
//             the look-ahead bytes are NOT masked -- SURVEY.md §8 A7 quirk); list mode: barrier = 'N'.
// ---------------------------------------------------------------------------------------------
NM_HD uint8_t code_of(const SegParams &P, uint32_t p) {
  const uint8_t ch = ld8(P.seg + p);
  const unsigned l = letter_of(ch);
  uint8_t c = (uint8_t)l;
  const bool allowed = start_allowed(ch);
  if (l == 4u || !allowed) c |= kIrregular;  // x, or U/u (letter t but not complemented by search.py:22)
  if (!allowed) c |= kNoStart;
  const bool barrier = (P.mode == 0) ? (!allowed && p < P.num_kmers) : (ch == 'N');
  if (barrier) c |= kBarrier;
  return c;
}

NM_HD void prep_thread(const DevIndex &ix, const SegParams &P, uint32_t p) {
  const uint8_t c0 = code_of(P, p);
  P.codes[p] = c0;
  uint32_t key = 0;
  const uint32_t sl = ix.seed_len;
  bool ok = (uint64_t)p + sl <= P.seg_len;
  if (ok) {
    for (uint32_t j = 0; j < sl; ++j) {
      const uint8_t c = (j == 0) ? c0 : code_of(P, p + j);
      if (c & (kIrregular | kBarrier)) { ok = false; break; }
      key = (key << 2) | (c & 3u);
    }
  }
  P.seedkey[p] = ok ? key : kSeedInvalid;
}

// one thread per tile of kTile positions: nb[p] = first barrier at or after p inside the tile
NM_HD void nb_tile_thread(const SegParams &P, uint32_t t) {
  const uint32_t begin = t * kTile;
  uint32_t end = begin + kTile;
  if (end > P.seg_len) end = P.seg_len;
  uint32_t next = kNoBarrier;
  for (uint32_t p = end; p-- > begin;) {
    if (P.codes[p] & kBarrier) next = p;
    P.nb[p] = next;
  }
  P.tile_first[t] = next;
}

// seed key of the other strand's k-mer starting (in search order) at i: letters 3-c[i], 3-c[i+1], ...
// with c[i] LEAST significant (it is the last character of the reverse-complemented k-mer)
NM_HD uint32_t rc_seed_key(uint32_t fwdkey, uint32_t sl) {
  uint32_t r = 0;
  for (uint32_t j = 0; j < sl; ++j) {
    r = (r << 2) | (3u - (fwdkey & 3u));
    fwdkey >>= 2;
  }
  return r;
}

// ---------------------------------------------------------------------------------------------
// (F) trajectory-free search, fast kernel: one thread per position p runs two chains, interleaved:
//   forward strand, leftwards from end e = p:   G[p] = min{ l : occ(seg[p-l, p)) == 1 }
//   other strand,  rightwards from start i = p: R[p] = min{ l : occ(rc(seg[p, p+l))) == 0 }
// both seeded from the seed table at depth seed_len (so values <= seed_len read as seed_len; the
// caller guarantees kmin >= seed_len) and extended on the rank32 layout.  Anything irregular is
// marked kPending for the generic kernel.  Why this gives the reference's bytes: DESIGN.md §3.
// ---------------------------------------------------------------------------------------------
struct Chain32 {
  uint32_t lo, hi;  // half-open
  uint32_t d;       // characters consumed
  bool active;
};

NM_HD void free_fast_thread(const DevIndex &ix, const SegParams &P, uint32_t p, uint32_t &steps) {
  const uint32_t sl = ix.seed_len;
  uint32_t g = kNone, rr = kInf;
  Chain32 f{0, 0, sl, false}, r{0, 0, sl, false};
  bool poison = false;

  // ---- forward strand, end e = p
  if (p >= 1 && !(P.codes[p - 1] & kBarrier)) {
    uint32_t key = kSeedInvalid;
    if (p >= sl && !(P.mode == 1 && p == P.seg_len)) key = P.seedkey[p - sl];
    if (key == kSeedInvalid) {
      g = kPending;
    } else {
      const SeedEntry32 s = ix.seed32[key];
      f.lo = s.start;
      f.hi = s.start + s.len;
      if (s.len == 0) atomic_or64(&P.counters[CTR_ANOMALY], 1ull);
      else if (s.len == 1) g = sl;
      else f.active = true;
    }
  }
  // ---- other strand, start i = p
  const bool want_r = P.use_rc && p < P.num_kmers && !(P.codes[p] & kNoStart);
  if (want_r) {
    const uint32_t key = P.seedkey[p];
    if (key == kSeedInvalid) {  // includes p + seed_len > seg_len, i.e. every truncated list-mode length < seed_len
      rr = kPending;
    } else {
      const SeedEntry32 s = ix.seed32[rc_seed_key(key, sl)];
      r.lo = s.start;
      r.hi = s.start + s.len;
      if (s.len == 0) rr = sl;
      else r.active = true;
    }
  }

  while (f.active || r.active) {
    unsigned fl = 0, rl = 0;
    if (f.active) {  // next character to the left: seg[p-1-d]
      if (f.d >= P.kmax || f.d >= p) { f.active = false; g = kNone; }
      else {
        const uint8_t c = P.codes[p - 1 - f.d];
        if (c & kBarrier) { f.active = false; g = kNone; }
        else if (c & kIrregular) { f.active = false; g = kPending; }
        else fl = c & 3u;
      }
    }
    if (r.active) {  // next character to the right: seg[p+d], complemented
      if (r.d >= P.kmax || p + r.d >= P.seg_len) { r.active = false; rr = kInf; }
      else {
        const uint8_t c = P.codes[p + r.d];
        if (c & kBarrier) { r.active = false; rr = kInf; }
        else if (c & kIrregular) { r.active = false; rr = kPending; }
        else rl = 3u - (c & 3u);
      }
    }
    uint32_t flo = 0, fhi = 0, rlo = 0, rhi = 0;
    if (f.active) {
      flo = occ_rank32(ix.rank32, fl, f.lo, poison);
      fhi = occ_rank32(ix.rank32, fl, f.hi, poison);
    }
    if (r.active) {
      rlo = occ_rank32(ix.rank32, rl, r.lo, poison);
      rhi = occ_rank32(ix.rank32, rl, r.hi, poison);
    }
    if (poison) {  // a block with an x/$ symbol was touched: both searches go to the generic kernel
      if (f.active) g = kPending;
      if (r.active) rr = kPending;
      break;
    }
    if (f.active) {
      const uint32_t c = (uint32_t)ix.C[fl];
      f.lo = c + flo;
      f.hi = c + fhi;
      ++f.d;
      ++steps;
      const uint32_t n = f.hi - f.lo;
      if (n <= 1) {
        f.active = false;
        if (n == 1) g = f.d;
        else atomic_or64(&P.counters[CTR_ANOMALY], 1ull);  // a substring of the segment is absent from the index
      }
    }
    if (r.active) {
      const uint32_t c = (uint32_t)ix.C[rl];
      r.lo = c + rlo;
      r.hi = c + rhi;
      ++r.d;
      ++steps;
      if (r.hi == r.lo) { r.active = false; rr = r.d; }
    }
  }
  P.G[p] = g;
  if (p < P.seg_len) P.R[p] = (p < P.num_kmers) ? rr : kInf;
}

// ---------------------------------------------------------------------------------------------
// (G) trajectory-free search, generic kernel: the same two quantities computed exactly (from
// depth 1, no seed table) on the unmodified .awfmi blocks with the five-letter alphabet and u64
// pointers.  Runs for entries marked kPending, or for every position when `all` is set
// (kmin < seed_len, or bwt_len >= 2^32-1).
// ---------------------------------------------------------------------------------------------
NM_HD unsigned rc_letter_of_code(uint8_t c) {
  const unsigned l = c & kLetterMask;
  if (c & kIrregular) return l == 3u ? 3u : 4u;  // U/u keeps letter t (search.py:22 does not complement it)
  return 3u - l;
}

NM_HD void free_generic_thread(const DevIndex &ix, const SegParams &P, uint32_t p, bool all, uint32_t &steps,
                               uint32_t &nsearch) {
  // ---- forward strand, end e = p
  if (all || P.G[p] == kPending) {
    uint32_t g = kNone;
    if (p >= 1) {
      Range64 r{0, 0};
      for (uint32_t d = 0; d < P.kmax && d < p; ++d) {
        const uint8_t c = P.codes[p - 1 - d];
        if (c & kBarrier) break;
        const unsigned l = c & kLetterMask;
        if (d == 0) { r.lo = ix.C[l]; r.hi = ix.C[l + 1]; }
        else { lf_native(ix, r, l); ++steps; }
        const uint64_t n = r.hi > r.lo ? r.hi - r.lo : 0;
        if (n == 1) { g = d + 1; break; }
        if (n == 0) { atomic_or64(&P.counters[CTR_ANOMALY], 1ull); break; }
      }
      ++nsearch;
    }
    P.G[p] = g;
  }
  // ---- other strand, start i = p
  if (p < P.seg_len && (all || P.R[p] == kPending)) {
    uint32_t rr = kInf;
    if (P.use_rc && p < P.num_kmers && !(P.codes[p] & kNoStart)) {
      Range64 r{0, 0};
      for (uint32_t d = 0; d < P.kmax && p + d < P.seg_len; ++d) {
        const uint8_t c = P.codes[p + d];
        if (c & kBarrier) break;
        const unsigned l = rc_letter_of_code(c);
        if (d == 0) { r.lo = ix.C[l]; r.hi = ix.C[l + 1]; }
        else { lf_native(ix, r, l); ++steps; }
        if (r.hi <= r.lo) { rr = d + 1; break; }
      }
      ++nsearch;
    }
    P.R[p] = rr;
  }
}

// ---------------------------------------------------------------------------------------------
// (R) resolve: out[i] from G, R and the length limit.  One thread per start position.
//   unique(i,k)  <=>  occ(seg[i,i+k)) == 1 and other strand absent  <=>  G[i+k] in [1,k] and R[i] <= k
// range mode: smallest k in [kmin, lim(i)] (binary search: the predicate is monotone in k);
// list mode: first k of the list, with the reference's 'N' rule and end-of-buffer truncation.
// ---------------------------------------------------------------------------------------------
NM_HD bool fwd_unique(const SegParams &P, uint32_t i, uint32_t e) {
  const uint32_t g = P.G[e];
  return g != kNone && g <= e - i;
}

// LF steps the reference's binary search would execute for position i whose answer is `ans`
// (0 = none) with other-strand threshold r (seeded searches; x-containing k-mers approximated)
NM_HD uint64_t ref_trajectory_steps(const SegParams &P, uint32_t sl, uint32_t lim, uint32_t ans, uint32_t r) {
  uint64_t total = 0;
  uint32_t lo = P.kmin, hi = lim;
  uint32_t q = (lo + hi) / 2;
  if (P.initial && q > P.initial) q = P.initial;
  for (;;) {
    total += q > sl ? q - sl : 0;                        // forward strand: always the full length
    if (P.use_rc) {
      const uint32_t upto = (r != kInf && r < q) ? r : q;  // other strand: until the range empties
      total += upto > sl ? upto - sl : 0;
    }
    const bool one = ans != 0 && q >= ans;
    if (one) { if (q == lo) break; hi = q - 1; }
    else { if (q == hi) break; lo = q + 1; }
    q = (lo + hi) / 2;
  }
  return total;
}

NM_HD void resolve_thread(const DevIndex &ix, const SegParams &P, uint32_t i, uint64_t &ref_lf) {
  uint32_t ans = 0;
  if (!(P.codes[i] & kNoStart)) {
    const uint32_t r = P.use_rc ? P.R[i] : 0u;
    if (P.mode == 0) {
      const uint32_t lim = length_limit(P, i);
      if (lim >= P.kmin) {
        if (r != kInf) {
          const uint32_t k0 = r > P.kmin ? r : P.kmin;
          if (k0 <= lim && fwd_unique(P, i, i + lim)) {
            uint32_t lo = i + k0, hi = i + lim;
            while (lo < hi) {
              const uint32_t mid = lo + (hi - lo) / 2;
              if (fwd_unique(P, i, mid)) hi = mid; else lo = mid + 1;
            }
            ans = lo - i;
          }
        }
        if (P.want_ref_lf) ref_lf += ref_trajectory_steps(P, ix.seed_len, lim, ans, P.use_rc ? r : 0u);
      }
    } else {
      const uint32_t b = next_barrier(P, i);
      for (uint32_t j = 0; j < P.nk; ++j) {
        const uint32_t k = P.klist[j];
        const uint32_t cut = (k <= P.seg_len - i) ? k : P.seg_len - i;
        if (b != kNoBarrier && b - i < cut) break;
        if ((r != kInf && r <= cut) && fwd_unique(P, i, i + cut)) { ans = k; break; }
      }
    }
  }
  store_out(P, i, ans);
}

// ---------------------------------------------------------------------------------------------
// index transcode: unmodified .awfmi blocks -> rank32 layout; seed table -> {u32 start, u32 len}
// one thread per 64 BWT positions (a quarter of a native block)
// ---------------------------------------------------------------------------------------------
NM_HD void transcode_thread(const uint8_t *native, uint64_t n_blocks, uint64_t bwt_len, uint32_t *rank32,
                            uint64_t b64) {
  const uint64_t nb = b64 >> 2;
  const unsigned w = (unsigned)(b64 & 3);
  uint32_t *o = rank32 + b64 * kRank32Words;
  if (nb >= n_blocks) {  // pad block(s): only ever read with an empty mask
    const uint64_t *last = reinterpret_cast<const uint64_t *>(native + n_blocks * kNativeBlockBytes);  // pad block counters = totals
    for (int l = 0; l < 4; ++l) o[l] = (uint32_t)last[12 + l];
    o[4] = o[5] = o[6] = o[7] = 0;
    return;
  }
  const uint64_t *b = reinterpret_cast<const uint64_t *>(native + nb * kNativeBlockBytes);
  uint64_t cnt[4] = {b[12], b[13], b[14], b[15]};
  for (unsigned v = 0; v < w; ++v) {
    const uint64_t p0 = b[v], p1 = b[4 + v], p2 = b[8 + v];
    cnt[0] += popc64(p2 & p1);
    cnt[1] += popc64(p2 & p0);
    cnt[2] += popc64(p1 & p0);
    cnt[3] += popc64(~p2 & ~p1 & p0);
  }
  const uint64_t p0 = b[w], p1 = b[4 + w], p2 = b[8 + w];
  const uint64_t mc = p2 & p0, mg = p1 & p0, mt = ~p2 & ~p1 & p0, ma = p2 & p1;
  uint64_t valid = ~0ull;  // positions of this word that exist in the BWT
  const uint64_t first = b64 * 64;
  if (first + 64 > bwt_len) valid = (first >= bwt_len) ? 0ull : ((1ull << (bwt_len - first)) - 1ull);
  const uint64_t other = ~(ma | mc | mg | mt) & valid;  // x or $ inside the BWT
  const uint64_t plane0 = mc | mt, plane1 = mg | mt;
  if (other) {
    o[0] = o[1] = o[2] = o[3] = kPoison;
  } else {
    for (int l = 0; l < 4; ++l) o[l] = (uint32_t)cnt[l];
  }
  o[4] = (uint32_t)plane0;
  o[5] = (uint32_t)(plane0 >> 32);
  o[6] = (uint32_t)plane1;
  o[7] = (uint32_t)(plane1 >> 32);
}

NM_HD void seed_transcode_thread(const uint64_t *seed64, SeedEntry32 *seed32, uint64_t t) {
  const uint64_t s = seed64[2 * t], e = seed64[2 * t + 1];
  SeedEntry32 o;
  o.start = (uint32_t)s;
  o.len = (s <= e) ? (uint32_t)(e - s + 1) : 0u;
  seed32[t] = o;
}

// FNV-1a-like fold used by nm_workspace_checksum (order independent: sum of per-position hashes)
NM_HD uint64_t checksum_term(uint64_t i, uint32_t v) {
  uint64_t h = (i + 1) * 0x9E3779B97F4A7C15ull ^ ((uint64_t)v * 0xC2B2AE3D27D4EB4Full);
  h ^= h >> 29;
  h *= 0xBF58476D1CE4E5B9ull;
  h ^= h >> 32;
  return h;
}

}  // namespace nm
