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
NM_HD int ctz64(uint64_t v) {  // v != 0
#if defined(__CUDA_ARCH__)
  return __ffsll((long long)v) - 1;
#else
  return __builtin_ctzll(v);
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
NM_HD void atomic_max64(unsigned long long *p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicMax(p, v);
#else
  if (v > *p) *p = v;
#endif
}
NM_HD void atomic_or32(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  atomicOr(p, v);
#else
  *p |= v;
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
constexpr uint8_t kSeedBad = 0x40;     // the window of P.sl letters starting here holds an irregular letter, a barrier or
                                       // the segment end: seedkey[p] is not a seed-table key

constexpr uint32_t kNone = 0u;               // G: no unique suffix within bounds
constexpr uint32_t kInf = 0xFFFFFFFFu;       // R: other strand still present at the longest length
constexpr uint32_t kPending = 0xFFFFFFFEu;   // G/R: to be computed by the generic kernel
constexpr uint32_t kDeep = 0xFFFFFFFDu;      // G/R: undecided at the depth cap of the first pass
constexpr uint32_t kNoBarrier = 0xFFFFFFFFu;
constexpr uint32_t kCertOwnChains = 1;      // chains of their own a certificate thread may run for positions no bound settles
constexpr uint32_t kTile = 64;               // nb[] tile (positions per thread of nb_tile_thread)

// counters[] slots
enum { CTR_LF = 0, CTR_QUERIES = 1, CTR_GENERIC = 2, CTR_ANOMALY = 3, CTR_ZERO_KEY = 4, CTR_REF_LF = 5,
       CTR_LF_PASS1 = 6,  // LF steps of the first (depth-capped, every position) pass of the trajectory-free search
       CTR_SEEDS = 7,     // seed-table lookups of the trajectory-free search (one 32-byte sector each)
       CTR_SEEDS_PASS1 = 8,  // the part of CTR_SEEDS looked up by the first pass
       CTR_SHORT = 9,     // multi search: positions whose summed count fell below the number of sequences (not 0)
       CTR_UNIQUE = 10, CTR_AMBIG = 11, CTR_MINLEN = 12, CTR_MAXLEN = 13,  // NM_OPT_OUT_STATS: reductions over the output
       CTR_N = 16 };

struct alignas(8) SeedEntry32 {  // one 64-bit load
  uint32_t start, len;
};

// entry of the derived table: the SA interval of the d-mer AND of its reverse complement, so that one 16-byte
// access seeds the forward-strand chain that ends where the d-mer ends and the other-strand chain that starts
// where it starts
struct alignas(16) SeedEntryPair {
  SeedEntry32 fwd, rc;
};

struct DevIndex {
  const uint8_t *native;      // (n_blocks + 1) * 160 B, file order, one zero-plane pad block
  const uint32_t *rank32;     // (n64 + 1) * 8 u32, or nullptr
  const SeedEntry32 *seed32;  // 4^seed_len entries, or nullptr
  const uint64_t *seed64;     // 4^seed_len {start,end} pairs exactly as in the file
  const SeedEntryPair *seedx; // derived on the device at load time: 4^seedx_len entries of a DEEPER seed table
  uint32_t seedx_len;         // (seed_len < seedx_len <= 16), or nullptr / 0
  const uint32_t *c32;        // prefixSums as u32[6] in device memory (rank32 path), or nullptr
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
// count[l] = prefixSums[l] + occurrences of l before the block: the rank of a position inside the block IS the new
// suffix-array pointer of the backward-extension step (AwFmSearch.c:44-95 adds the prefix sum after the rank; here it
// is folded in when the layout is derived, one add and one table lookup less on every dependent step).
// A block that holds an x or $ symbol carries kPoison in all four counts.  One block = one 32-byte
// DRAM sector, fetched with a single 256-bit load (LDG.E.256 on sm_100a).
// ---------------------------------------------------------------------------------------------
struct Rank32Block {
  uint32_t c[4];
  uint64_t p0, p1;
};

NM_HD Rank32Block ld_rank32(const uint32_t *rank32, uint32_t blk) {
  const uint32_t *w = rank32 + (size_t)blk * kRank32Words;
  Rank32Block r;
#if defined(__CUDA_ARCH__)
  uint32_t v4, v5, v6, v7;
  asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.c[0]), "=r"(r.c[1]), "=r"(r.c[2]), "=r"(r.c[3]), "=r"(v4), "=r"(v5), "=r"(v6), "=r"(v7)
               : "l"(w));
  r.p0 = (uint64_t)v4 | ((uint64_t)v5 << 32);
  r.p1 = (uint64_t)v6 | ((uint64_t)v7 << 32);
#else
  for (int l = 0; l < 4; ++l) r.c[l] = w[l];
  r.p0 = (uint64_t)w[4] | ((uint64_t)w[5] << 32);
  r.p1 = (uint64_t)w[6] | ((uint64_t)w[7] << 32);
#endif
  return r;
}

#if defined(NM_BRANCHY_SEL)   // (round-1 form, kept for A/B runs: tools/sweep_opts.py with NEWMAP_B200_LIB)
NM_HD uint32_t sel4(unsigned letter, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return letter == 0 ? a : letter == 1 ? b : letter == 2 ? c : d;
}
NM_HD uint32_t occ_in_block(const Rank32Block &b, unsigned letter, uint32_t off) {
  const uint64_t m = ((letter & 1u) ? b.p0 : ~b.p0) & ((letter & 2u) ? b.p1 : ~b.p1);
  return sel4(letter, b.c[0], b.c[1], b.c[2], b.c[3]) + (uint32_t)popc64(m & ((1ull << off) - 1ull));
}
#else
NM_HD uint32_t sel4(unsigned letter, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  // branch-free (the lanes of a warp hold different letters: a branch per letter would run four paths one after the other)
  const uint32_t m1 = 0u - (letter & 1u), m2 = 0u - ((letter >> 1) & 1u);
  const uint32_t lo = a ^ ((a ^ b) & m1), hi = c ^ ((c ^ d) & m1);
  return lo ^ ((lo ^ hi) & m2);
}

NM_HD uint32_t occ_in_block(const Rank32Block &b, unsigned letter, uint32_t off) {
  // positions whose 2-bit code equals `letter`: plane0 must equal bit 0 of the letter, plane1 bit 1 (x ^ ~M = x where the
  // bit is set, ~x where it is not)
  const uint64_t n1 = (uint64_t)(letter & 1u) - 1ull, n2 = (uint64_t)((letter >> 1) & 1u) - 1ull;   // ~M1, ~M2
  const uint64_t m = (b.p0 ^ n1) & (b.p1 ^ n2);
  return sel4(letter, b.c[0], b.c[1], b.c[2], b.c[3]) + (uint32_t)popc64(m & ((1ull << off) - 1ull));
}
#endif

// one backward-extension step of the half-open interval [lo,hi) by `letter` (0..4) with u32 pointers:
// rank32 blocks for a,c,g,t (one sector when both bounds share a block), the unmodified .awfmi block
// for x or when a poisoned block is met.  C32 = prefix sums as u32[6].
NM_HD void lf_fast(const DevIndex &ix, const uint32_t *C32, unsigned letter, uint32_t &lo, uint32_t &hi) {
  if (letter < 4u) {
    const uint32_t blo = lo >> 6, bhi = hi >> 6;
    // both loads are issued before either is used: a chain is a sequence of DEPENDENT steps, and a step whose bounds
    // lie in two blocks must cost one memory latency, not two
    const Rank32Block a = ld_rank32(ix.rank32, blo);
    Rank32Block b = a;
    if (bhi != blo) b = ld_rank32(ix.rank32, bhi);
    const uint32_t nlo = occ_in_block(a, letter, lo & 63u);
    const uint32_t nhi = occ_in_block(b, letter, hi & 63u);
    const bool poison = a.c[0] == kPoison || b.c[0] == kPoison;
    if (!poison) {      // (the block counts include the prefix sum of the letter)
      lo = nlo;
      hi = nhi;
      return;
    }
  }
  const uint32_t c = C32[letter];
  const uint32_t nlo = (uint32_t)occ_native(ix.native, letter, lo);
  const uint32_t nhi = (uint32_t)occ_native(ix.native, letter, hi);
  lo = c + nlo;
  hi = c + nhi;
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
  uint32_t *pk;          // seg_len/16 + 2 words: the letters packed 2 bits each, position 16w + j in bits 31-2j..30-2j of
                         // word w (an earlier position is more significant, like the digits of a seed-table key)
  const SeedEntry32 *seedtab;  // seed table the chains start from: 4^sl entries of the file's table ...
  const SeedEntryPair *seedpair;  // ... or of the derived both-strand table (exactly one of the two is set)
  uint32_t sl;           // its depth; kmin >= sl
  uint64_t *barmask;     // ceil(seg_len / kTile): bit j of word t = position t * kTile + j is a barrier
  uint32_t *tile_first;  // ceil(seg_len / kTile): first barrier of the tile, or kNoBarrier
  uint32_t *G;           // seg_len + 1: length of the shortest unique suffix ending at e (forward strand)
  uint32_t *R;           // seg_len: shortest length from i at which the other strand has no match
  uint32_t *deepmap;     // (seg_len + 1) bits, cleared by prep: bit p = the first pass left G[p] or R[p] kDeep.  Deep entries are
                         // 1-2 % of a segment: passes 2 and 3 look here before they read G / R at all (nullptr: they scan)
  void *out;             // num_kmers entries of out_width bytes
  const uint32_t *klist; // list mode: nk lengths in the order given
  unsigned long long *counters;  // CTR_N slots
  uint32_t seg_len, num_kmers;
  uint32_t kmin, kmax, nk, initial;
  uint32_t keff;         // shortest length the trajectory-free passes work with: max(kmin, sl).  When kmin < sl (a range that
                         // starts below the depth of the seed table) the lengths kmin .. sl-1 are settled by resolve_thread
                         // with exact counts, for the few positions the table does not rule out (`sub`)
  uint8_t *sub;          // seg_len (only when keff > kmin): 1 = the sl-1 letters starting here may be unique (subdepth_candidate)
  int mode;       // 0 range, 1 list
  int use_rc;
  int out_width;  // 1, 2, 4
  int fast;       // rank32 + seed table usable for this search (kmin >= seed_len, bwt_len < 2^32-1)
  int want_ref_lf;
  uint32_t dcap;   // first-pass depth cap of the trajectory-free search (characters, > seed_len)
  uint32_t gsize;  // positions per certificate group of the deep pass
  uint32_t own_chains;  // chains of their own a certificate thread may run for positions no bound settles
  uint32_t cert_runs;   // certificate pass: runs of consecutive deep entries a thread takes per group and strand
  uint32_t stall;       // first pass: a chain whose interval kept its size for this many steps in a row is left deep (0 = off)
  uint32_t stride;      // sparse first pass: one seed window per `stride` positions is looked up unconditionally (1 = every window)
};

NM_HD uint32_t load_out(const SegParams &P, uint32_t i) {
  if (P.out_width == 1) return reinterpret_cast<const uint8_t *>(P.out)[i];
  if (P.out_width == 2) return reinterpret_cast<const uint16_t *>(P.out)[i];
  return reinterpret_cast<const uint32_t *>(P.out)[i];
}
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
  const uint64_t m = P.barmask[i / kTile] >> (i % kTile);
  if (m) return i + (uint32_t)ctz64(m);
  uint32_t v;
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
// (M) several sequences x several indexes at once (search.py:461,489-517 range mode, :575,:627-639 list mode,
// :656-697 the count): the count of a (position, length) is summed over every index and every sequence, both strands
// unless --norc, in uint32; a length is unique when the sum equals the number of SEQUENCES.  Start mask, barriers and
// length limits come from the first sequence only (search.py:391-394, :557-560).  One position per thread on the
// reference's own query trajectory.  A sum below the number of sequences that is not 0 matches no rule of the
// reference's range loop (it would repeat the same query for ever): the position is left at what it has found and
// counted in CTR_SHORT for the caller to refuse.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kMaxMulti = 8;
struct MultiSet {
  DevIndex ix[kMaxMulti];
  const uint8_t *seg[kMaxMulti];  // seg[0] == P.seg; all of P.seg_len bytes
  uint32_t n_ix, n_seq;
};

NM_HD uint32_t count_multi(const MultiSet &M, uint64_t i, uint64_t k, bool use_rc, uint32_t &steps) {
  uint32_t c = 0;
  for (uint32_t x = 0; x < M.n_ix; ++x)
    for (uint32_t s = 0; s < M.n_seq; ++s) c += count_both(M.ix[x], M.seg[s], i, k, use_rc, steps);
  return c;
}

NM_HD void trajectory_multi_thread(const MultiSet &M, const SegParams &P, uint32_t i, uint32_t &steps, uint32_t &nq,
                                   uint32_t &nshort) {
  uint32_t best = 0;
  const uint32_t S = M.n_seq;
  const uint8_t ch = ld8(P.seg + i);
  if (start_allowed(ch)) {
    if (P.mode == 0) {
      const uint32_t ub = length_limit(P, i);
      if (ub >= P.kmin) {
        uint32_t lo = P.kmin, hi = ub;
        uint32_t q = (lo + hi) / 2;
        if (P.initial && q > P.initial) q = P.initial;
        for (uint32_t round = 0;; ++round) {
          const uint32_t c = count_multi(M, i, q, P.use_rc != 0, steps);
          ++nq;
          if (c == 0) { note_zero(P, round, i, q); break; }
          if (c == S) {                         // search.py:489-508, :524-527
            if (best == 0 || q < best) best = q;
            if (q == lo) break;
            hi = q - 1;
          } else if (c > S) {                   // :513-517, :532-535
            if (q == hi) break;
            lo = q + 1;
          } else {
            ++nshort;
            break;
          }
          q = (lo + hi) / 2;
        }
      }
    } else {
      const uint32_t b = next_barrier(P, i);
      for (uint32_t j = 0; j < P.nk; ++j) {
        const uint32_t k = P.klist[j];
        const uint32_t cut = (k <= P.seg_len - i) ? k : P.seg_len - i;
        if (b != kNoBarrier && b - i < cut) break;
        const uint32_t c = count_multi(M, i, cut, P.use_rc != 0, steps);
        ++nq;
        if (c == 0) { note_zero(P, j, i, cut); break; }
        if (c == S && best == 0) best = k;      // :627-636: first length whose sum equals the number of sequences
        if (c == 1) break;                      // :639: only a sum of exactly 1 retires the position
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
  (void)ix;
  const uint8_t c0 = code_of(P, p);
  const uint32_t sl = P.sl;
  bool ok = (uint64_t)p + sl <= P.seg_len;
  if (ok) {
    for (uint32_t j = 0; j < sl; ++j) {
      const uint8_t c = (j == 0) ? c0 : code_of(P, p + j);
      if (c & (kIrregular | kBarrier)) { ok = false; break; }
    }
  }
  P.codes[p] = ok ? c0 : (uint8_t)(c0 | kSeedBad);
}

// one thread per word of pk: the letters of 16 positions, 2 bits each, the earliest position in the top bits
// (positions beyond the segment read as letter 0; a window that reaches them carries kSeedBad)
NM_HD void pack_thread(const SegParams &P, uint32_t w) {
  uint32_t v = 0;
  for (uint32_t j = 0; j < 16u; ++j) {
    const uint64_t p = (uint64_t)w * 16u + j;
    const uint32_t l = p < P.seg_len ? (uint32_t)(P.codes[p] & 3u) : 0u;
    v = (v << 2) | l;
  }
  P.pk[w] = v;
}

// one thread per tile of kTile positions: barrier bits of the tile, first barrier of the tile
NM_HD void nb_tile_thread(const SegParams &P, uint32_t t) {
  const uint32_t begin = t * kTile;
  uint32_t end = begin + kTile;
  if (end > P.seg_len) end = P.seg_len;
  uint64_t m = 0;
  for (uint32_t p = begin; p < end; ++p)
    if (P.codes[p] & kBarrier) m |= 1ull << (p - begin);
  P.barmask[t] = m;
  P.tile_first[t] = m ? begin + (uint32_t)ctz64(m) : kNoBarrier;
}

// seed-table key of the window of P.sl letters starting at q (q + P.sl <= seg_len): two words of pk and a funnel shift
NM_HD uint32_t window_key(const SegParams &P, uint32_t q) {
  const uint32_t w = q >> 4, r = 2u * (q & 15u);
  const uint32_t hi = ld32(P.pk + w), lo = ld32(P.pk + w + 1);
#if defined(__CUDA_ARCH__)
  const uint32_t x = __funnelshift_l(lo, hi, r);
#else
  const uint32_t x = r ? (hi << r) | (lo >> (32u - r)) : hi;
#endif
  return x >> (32u - 2u * P.sl);
}

// seed key of the other strand's k-mer starting (in search order) at i: letters 3-c[i], 3-c[i+1], ...
// with c[i] LEAST significant (it is the last character of the reverse-complemented k-mer)
NM_HD uint32_t rc_seed_key(uint32_t fwdkey, uint32_t sl) {
  uint32_t x = ~fwdkey;  // 3 - c == ~c on two bits
#if defined(__CUDA_ARCH__)
  x = __brev(x);
#else
  x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
  x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
  x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
  x = ((x >> 8) & 0x00FF00FFu) | ((x & 0x00FF00FFu) << 8);
  x = (x >> 16) | (x << 16);
#endif
  x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);  // bit reversal also swapped the two bits of a letter
  return x >> (32u - 2u * sl);                              // the sl letters now sit at the top
}

// ---------------------------------------------------------------------------------------------
// (F) trajectory-free search on the rank32 layout.  Two quantities per position p:
//   forward strand, leftwards from end e = p:   G[p] = min{ l : occ(seg[p-l, p)) == 1 }   (kNone if none)
//   other strand,  rightwards from start i = p: R[p] = min{ l : occ(rc(seg[p, p+l))) == 0 } (kInf if none)
// each found by ONE backward search ("chain") seeded from the seed table at depth seed_len (values
// below seed_len then read as seed_len; the caller guarantees kmin >= seed_len) or, when the seed
// window holds a barrier / irregular letter / the segment edge, from depth 1.
// Three passes keep long repeats cheap:
//   pass 1 (free_fast_thread)   every position, chains capped at depth P.dcap -> value or kDeep
//   pass 2 (free_cert_thread)   one thread per group of P.gsize positions: ONE chain from the group's
//                               outermost kDeep position, run P.gsize deeper than kmax.  A string that
//                               CONTAINS seg[e'-L, e') (resp. seg[i', i'+L)) and still occurs twice
//                               (resp. still occurs on the other strand) proves the same of the
//                               contained string, so one chain settles every kDeep position of a
//                               group that lies in a long exact repeat
//   pass 3 (free_rest_thread)   positions still kDeep: their own chain, uncapped
// Why the resolved bytes equal the reference's: DESIGN.md section 3.
// ---------------------------------------------------------------------------------------------
struct ChainResult {
  uint32_t d_hit;   // depth at which the chain's stop condition was met (count == 1 / count == 0), 0 = never
  uint32_t d_stop;  // depth explored; the count at every depth <= d_stop (and < d_hit) did not meet the condition
  bool wall;        // stopped because the next character is a barrier or beyond the segment
};

// A chain = one backward search, kept as {half-open SA interval, depth}.  kind 0: forward strand, pattern
// seg[p-d, p) extended leftwards from end p, stops when count == 1.  kind 1: other strand, pattern
// rc(seg[p, p+d)) extended rightwards from start p, stops when count == 0.
struct ChainState {
  uint32_t lo, hi, d;
};

// where the chain starts: the seed table at depth P.sl when the window of P.sl letters is clean, else depth 1
// (the caller has checked fwd_wanted / rc_wanted)
NM_HD bool chain_seeded(const SegParams &P, uint32_t kind, uint32_t p) {
  if (kind == 0u) return p >= P.sl && !(P.mode == 1 && p == P.seg_len) && !(P.codes[p - P.sl] & kSeedBad);
  return !(P.codes[p] & kSeedBad);
}
NM_HD uint32_t chain_seed_key(const SegParams &P, uint32_t kind, uint32_t p) {
  return kind == 0u ? window_key(P, p - P.sl) : rc_seed_key(window_key(P, p), P.sl);
}
// seed-table entry of the chain (kind, p): the window of P.sl letters ending at p (forward strand) or starting at p
NM_HD SeedEntry32 seed_entry(const SegParams &P, uint32_t kind, uint32_t p) {
  const uint32_t key = window_key(P, kind == 0u ? p - P.sl : p);
  if (P.seedpair) {
#if defined(__CUDA_ARCH__)
    const uint4 e = __ldg(reinterpret_cast<const uint4 *>(P.seedpair + key));
    return kind == 0u ? SeedEntry32{e.x, e.y} : SeedEntry32{e.z, e.w};
#else
    return kind == 0u ? P.seedpair[key].fwd : P.seedpair[key].rc;
#endif
  }
  return P.seedtab[kind == 0u ? key : rc_seed_key(key, P.sl)];
}
NM_HD unsigned rc_letter_of_code(uint8_t c) {
  const unsigned l = c & kLetterMask;
  if (c & kIrregular) return l == 3u ? 3u : 4u;  // U/u keeps letter t (search.py:22 does not complement it)
  return 3u - l;
}
NM_HD ChainState chain_from_seed(const SeedEntry32 s, uint32_t sl) { return ChainState{s.start, s.start + s.len, sl}; }
NM_HD ChainState chain_from_letter(const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t p) {
  const unsigned l = kind == 0u ? (unsigned)(P.codes[p - 1] & kLetterMask) : rc_letter_of_code(P.codes[p]);
  return ChainState{C32[l], C32[l + 1], 1u};
}
NM_HD ChainState chain_init(const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t p, uint32_t &seeds) {
  if (chain_seeded(P, kind, p)) {
    ++seeds;
    return chain_from_seed(seed_entry(P, kind, p), P.sl);
  }
  return chain_from_letter(C32, P, kind, p);
}

// the chain's own stop condition at its present depth (no memory access)
NM_HD bool chain_finished(const SegParams &P, uint32_t kind, const ChainState &s, ChainResult &r) {
  const uint32_t n = s.hi - s.lo;
  if (kind == 0u ? n > 1u : n != 0u) return false;
  r.d_hit = (kind == 0u && n == 0u) ? 0u : s.d;
  if (kind == 0u && n == 0u) atomic_or64(&P.counters[CTR_ANOMALY], 1ull);  // a substring of the segment is absent from the index
  r.d_stop = s.d;
  r.wall = false;
  return true;
}

// One iteration of a chain: either it ends here (returns true, r filled) or it takes one backward-extension
// step.  Written once for both kinds so that the lanes of a warp that hold chains of either kind stay converged.
NM_HD bool chain_iterate(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t p,
                         ChainState &s, uint32_t maxdepth, uint32_t &steps, ChainResult &r) {
  if (chain_finished(P, kind, s, r)) return true;
  r.d_hit = 0;
  r.d_stop = s.d;
  const bool edge = kind == 0u ? s.d >= p : p + s.d >= P.seg_len;
  if (edge) { r.wall = true; return true; }
  const uint8_t c = P.codes[kind == 0u ? p - 1u - s.d : p + s.d];
  if (c & kBarrier) { r.wall = true; return true; }
  r.wall = false;
  if (s.d >= maxdepth) return true;
  lf_fast(ix, C32, kind == 0u ? (unsigned)(c & kLetterMask) : rc_letter_of_code(c), s.lo, s.hi);
  ++s.d;
  ++steps;
  return false;
}

// First pass only.  A chain whose interval has kept its size for P.stall steps in a row is almost surely inside a
// repeat (two random 16-letter matches part ways within a step or two; the copies of a repeat stay together) and
// would run on to the depth cap only to be marked deep there: it is marked deep now.  Deep entries are settled
// exactly by passes 2 and 3 whatever made them deep, so this moves work, never a result.
struct StallState {
  uint32_t n, run;
};
NM_HD bool chain_stalled(const SegParams &P, StallState &st, const ChainState &s) {
  const uint32_t n = s.hi - s.lo;
  st.run = (n == st.n) ? st.run + 1u : 0u;
  st.n = n;
  return P.stall && st.run >= P.stall;
}
NM_HD ChainResult run_chain(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t p,
                            uint32_t maxdepth, uint32_t &steps, uint32_t &seeds) {
  ChainState s = chain_init(C32, P, kind, p, seeds);
  ChainResult r{0, 0, false};
  while (!chain_iterate(ix, C32, P, kind, p, s, maxdepth, steps, r)) {}
  return r;
}
NM_HD ChainResult run_fwd_chain(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t e,
                                uint32_t maxdepth, uint32_t &steps, uint32_t &seeds) {
  return run_chain(ix, C32, P, 0u, e, maxdepth, steps, seeds);
}
NM_HD ChainResult run_rc_chain(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t i,
                               uint32_t maxdepth, uint32_t &steps, uint32_t &seeds) {
  return run_chain(ix, C32, P, 1u, i, maxdepth, steps, seeds);
}

// value stored in G (none = kNone) / R (none = kInf) for a finished chain of a pass capped at `cap`
NM_HD uint32_t chain_value(const SegParams &P, const ChainResult &r, uint32_t cap, uint32_t none) {
  if (r.d_hit) return r.d_hit;
  if (!r.wall && r.d_stop < P.kmax && r.d_stop >= cap) return kDeep;
  return none;
}

// the first pass's chain for (kind, p): value for G / R
NM_HD uint32_t run_chain_pass1(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t p,
                               uint32_t cap, uint32_t none, uint32_t &steps, uint32_t &seeds) {
  ChainState s = chain_init(C32, P, kind, p, seeds);
  StallState st{s.hi - s.lo, 0u};
  ChainResult r{0, 0, false};
  while (!chain_iterate(ix, C32, P, kind, p, s, cap, steps, r))
    if (chain_stalled(P, st, s)) return kDeep;
  return chain_value(P, r, cap, none);
}

NM_HD bool fwd_wanted(const SegParams &P, uint32_t e) { return e >= 1 && !(P.codes[e - 1] & kBarrier); }
NM_HD bool rc_wanted(const SegParams &P, uint32_t i) {
  return P.use_rc && i < P.num_kmers && !(P.codes[i] & kNoStart);
}

NM_HD void mark_deep(const SegParams &P, uint32_t p) {
  if (P.deepmap) atomic_or32(&P.deepmap[p >> 5], 1u << (p & 31u));
}
NM_HD bool maybe_deep(const SegParams &P, uint32_t p) {
  return !P.deepmap || ((ld32(P.deepmap + (p >> 5)) >> (p & 31u)) & 1u);
}
// any position of [lo, hi] marked?  (whole words are looked at: a neighbour's bit only costs the scan it would have saved)
NM_HD bool maybe_deep_range(const SegParams &P, uint32_t lo, uint32_t hi) {
  if (!P.deepmap) return true;
  for (uint32_t w = lo >> 5; w <= (hi >> 5); ++w)
    if (ld32(P.deepmap + w)) return true;
  return false;
}

// pass 1 and pass 3 share this: `cap` = P.dcap in pass 1, kmax in pass 3 (which only redoes kDeep entries)
// (parts: bit 0 = the forward-strand chain ending at p, bit 1 = the other-strand chain starting at p)
NM_HD void free_chain_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t p, bool first,
                             uint32_t &steps, uint32_t &seeds, uint32_t parts = 3u) {
  const uint32_t cap = first ? (P.dcap < P.kmax ? P.dcap : P.kmax) : P.kmax;
  if (!first && !maybe_deep(P, p)) return;
  if ((parts & 1u) && (first || P.G[p] == kDeep)) {
    uint32_t g = kNone;
    if (fwd_wanted(P, p)) {
      g = first ? run_chain_pass1(ix, C32, P, 0u, p, cap, kNone, steps, seeds)
                : chain_value(P, run_fwd_chain(ix, C32, P, p, cap, steps, seeds), cap, kNone);
    }
    P.G[p] = g;
    if (g == kDeep) mark_deep(P, p);
  }
  if ((parts & 2u) && p < P.seg_len && (first || P.R[p] == kDeep)) {
    uint32_t rr = kInf;
    if (rc_wanted(P, p)) {
      rr = first ? run_chain_pass1(ix, C32, P, 1u, p, cap, kInf, steps, seeds)
                 : chain_value(P, run_rc_chain(ix, C32, P, p, cap, steps, seeds), cap, kInf);
    }
    P.R[p] = rr;
    if (rr == kDeep) mark_deep(P, p);
  }
}

// ---------------------------------------------------------------------------------------------
// Sparse first pass.  The resolver only ever asks "G[e] <= t" and "R[i] <= t" with t >= kmin (range mode: every
// queried length is >= kmin; list mode: the same except where the segment end cuts a k-mer short, see below), so
// an UPPER BOUND b <= kmin is as good as the exact value.  Bounds come for free from a neighbour:
//   G[e+1] <= G[e] + 1   (a unique string stays unique when it is extended; needs position e not to be a barrier)
//   R[i-1] <= R[i] + 1   (an absent string stays absent when it is extended; needs start i-1 to be wanted)
// With a seed table of depth sl < kmin a window that is already unique (forward strand) / absent (other strand)
// at depth sl therefore settles kmin - sl further ends to its right / starts to its left without touching
// memory.  The positions are cut into cells of `stride` = min(kmin - sl + 1, cap) consecutive ends / starts; a
// cell is walked in the direction the bounds spread -- ends upwards, starts downwards: the chain of the first
// position runs (one table lookup, usually decided there), its value spreads as far as it stays <= kmin, the next
// unsettled position runs its own chain, and so on.  A cell inside a repeat degenerates to one chain per position,
// which is what the dense pass does everywhere.
// Exact values are kept where a bound would not do: list mode compares against the CUT length where the segment
// end truncates a k-mer (search.py:590,600), which may be below kmin -- the end e = seg_len and the starts
// i > seg_len - kmin get chains of their own there.
// Stored bounds are valid upper bounds <= kmin, so the later passes (which only rewrite kDeep entries and use
// G[e-1] + 1 / R[i+1] + 1 as upper bounds themselves) need no change.
// ---------------------------------------------------------------------------------------------
// Lengths below the depth of the table (a search range that starts at kmin < sl, e.g. 12:255 on the 16-letter table).
// The count is non-increasing in the length, so a position can only have an answer below sl when it is unique at sl
// itself; and the 4 table entries that share the first sl-1 letters of the window (keys 4x .. 4x+3: one aligned 64-byte
// group, which the memory system fetches whole anyway) give a LOWER bound on the two-strand count of those sl-1
// letters: the occurrences followed (forward strand) / preceded (other strand) by a regular letter.  A bound >= 2
// rules the position out; the rest (a few per cent in a genome of 4^(sl-1) letters or more) gets exact counts.
NM_HD bool subdepth_candidate(const SegParams &P, uint32_t q) {
  const uint32_t key = window_key(P, q) & ~3u;
  uint32_t lb = 0;
  for (uint32_t c = 0; c < 4u; ++c) {
    const SeedEntryPair e = P.seedpair[key + c];
    lb += e.fwd.len + (P.use_rc ? e.rc.len : 0u);
    if (lb > 1u) return false;
  }
  return true;
}

constexpr uint32_t kMaxStride = 8;
// stride chosen when none is asked for: as wide as the slack between the table depth and the shortest length allows
NM_HD uint32_t sparse_auto_stride(uint32_t kmin, uint32_t sl, bool pair_table, uint32_t asked) {
  if (!pair_table || kmin <= sl) return 1u;
  uint32_t s = kmin - sl + 1u;
  if (s > kMaxStride) s = kMaxStride;
  if (asked && asked < s) s = asked;
  return s;
}
NM_HD bool sparse_numeric(const SegParams &P, uint32_t v) { return v >= 1u && v <= P.kmax; }
NM_HD bool sparse_exact_only_end(const SegParams &P, uint32_t e) { return P.mode == 1 && e == P.seg_len; }
NM_HD bool sparse_exact_only_start(const SegParams &P, uint32_t i) { return P.mode == 1 && P.seg_len - i < P.keff; }

// ends are grouped [0, head), [head, head + S), ... with head = sl + S - 1: the window of sl letters that starts at
// q = c * S + (S - 1) ends at the first end of cell c and starts at the last start of cell c (starts [c*S, c*S + S))
NM_HD uint32_t sparse_fwd_cell_last(const SegParams &P, uint32_t e) {
  const uint32_t S = P.stride, head = P.sl + S - 1u;
  uint32_t last = e < head ? head - 1u : head + (e - head) / S * S + (S - 1u);
  return last > P.seg_len ? P.seg_len : last;
}
NM_HD uint32_t sparse_rc_cell_first(const SegParams &P, uint32_t i) { return i / P.stride * P.stride; }

// next end in [cur, last] that needs a chain (the unwanted ones before it are settled as "none"); false = cell done
NM_HD bool sparse_next_fwd(const SegParams &P, uint32_t &cur, uint32_t last) {
  for (; cur <= last; ++cur) {
    if (fwd_wanted(P, cur)) return true;
    P.G[cur] = kNone;
  }
  return false;
}
// next start in [first, end1) from the right that needs a chain: on return true it is end1 - 1
NM_HD bool sparse_next_rc(const SegParams &P, uint32_t first, uint32_t &end1) {
  for (; end1 > first; --end1) {
    if (rc_wanted(P, end1 - 1u)) return true;
    P.R[end1 - 1u] = kInf;
  }
  return false;
}
// G[cur] = v has just been settled: spread v + 1, v + 2, ... over the ends that follow; returns the first end not settled
NM_HD uint32_t sparse_spread_fwd(const SegParams &P, uint32_t cur, uint32_t last, uint32_t v) {
  uint32_t e = cur + 1u;
  if (!sparse_numeric(P, v)) return e;
  for (uint32_t b = v + 1u; e <= last && b <= P.keff && fwd_wanted(P, e) && !sparse_exact_only_end(P, e); ++e, ++b) P.G[e] = b;
  return e;
}
// R[end1 - 1] = v has just been settled: spread over the starts before it; returns the new end1
NM_HD uint32_t sparse_spread_rc(const SegParams &P, uint32_t first, uint32_t end1, uint32_t v) {
  uint32_t n = end1 - 1u;   // starts [first, n) remain
  if (!sparse_numeric(P, v)) return n;
  for (uint32_t b = v + 1u; n > first && b <= P.keff && rc_wanted(P, n - 1u) && !sparse_exact_only_start(P, n - 1u); --n, ++b)
    P.R[n - 1u] = b;
  return n;
}

// One cell side walked serially (the readable definition, run by the host-side logic tests; k_pass1_sparse keeps many
// of these in flight per block and takes the first lookups of all cells of a block together).
// kind 0: the ends [first_pos, its cell's last]; kind 1: the starts [its cell's first, first_pos].
NM_HD void sparse_side_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t kind, uint32_t first_pos,
                              uint32_t &steps, uint32_t &seeds) {
  const uint32_t cap = P.dcap < P.kmax ? P.dcap : P.kmax;
  if (kind == 0u) {
    uint32_t cur = first_pos;
    const uint32_t last = sparse_fwd_cell_last(P, first_pos);
    while (sparse_next_fwd(P, cur, last)) {
      const uint32_t v = run_chain_pass1(ix, C32, P, 0u, cur, cap, kNone, steps, seeds);
      P.G[cur] = v;
      if (v == kDeep) mark_deep(P, cur);
      cur = sparse_spread_fwd(P, cur, last, v);
    }
  } else {
    uint32_t end1 = first_pos + 1u;
    const uint32_t first = sparse_rc_cell_first(P, first_pos);
    while (sparse_next_rc(P, first, end1)) {
      const uint32_t v = run_chain_pass1(ix, C32, P, 1u, end1 - 1u, cap, kInf, steps, seeds);
      P.R[end1 - 1u] = v;
      if (v == kDeep) mark_deep(P, end1 - 1u);
      end1 = sparse_spread_rc(P, first, end1, v);
    }
  }
}
// every cell side of a segment, serially: the whole sparse first pass (host-side logic tests)
NM_HD void sparse_pass1_serial(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t &steps, uint32_t &seeds) {
  const uint32_t S = P.stride, head = P.sl + S - 1u;
  sparse_side_thread(ix, C32, P, 0u, 0u, steps, seeds);                                   // the head cell: ends [0, head)
  for (uint64_t e = head; e <= P.seg_len; e += S) sparse_side_thread(ix, C32, P, 0u, (uint32_t)e, steps, seeds);
  for (uint64_t i = 0; i < P.seg_len; i += S) {                                           // starts [i, i + S)
    const uint64_t top = i + S - 1u < P.seg_len ? i + S - 1u : P.seg_len - 1u;
    sparse_side_thread(ix, C32, P, 1u, (uint32_t)top, steps, seeds);
  }
}

// What the chain r run from the end e0 says about a deep end e <= e0 by itself and by containment; true when G[e]
// was settled (written), false when it stays deep
NM_HD bool cert_contain_fwd(const SegParams &P, const ChainResult &r, uint32_t e0, uint32_t e) {
  const uint32_t dt = e0 - e;
  uint32_t L = P.kmax;                             // deepest length e itself may be asked about
  if (r.wall && r.d_stop > dt && r.d_stop - dt < L) L = r.d_stop - dt;
  if (dt == 0) {
    P.G[e] = (r.d_hit && r.d_hit <= L) ? r.d_hit : kNone;
    return true;
  }
  const bool beyond_wall = r.wall && r.d_stop <= dt;  // e lies beyond the wall e0's chain ran into
  if (!beyond_wall && L + dt <= r.d_stop && (r.d_hit == 0 || r.d_hit > L + dt)) {
    P.G[e] = kNone;                                // seg[e-L, e0) occurs twice, so does seg[e-L, e)
    return true;
  }
  return false;
}
// mirror image: the chain r run from the start i0 and a deep start i >= i0
NM_HD bool cert_contain_rc(const SegParams &P, const ChainResult &r, uint32_t i0, uint32_t i) {
  const uint32_t dt = i - i0;
  uint32_t L = P.kmax;
  if (r.wall && r.d_stop > dt && r.d_stop - dt < L) L = r.d_stop - dt;
  if (dt == 0) {
    P.R[i] = (r.d_hit && r.d_hit <= L) ? r.d_hit : kInf;
    return true;
  }
  const bool beyond_wall = r.wall && r.d_stop <= dt;
  if (!beyond_wall && L + dt <= r.d_stop && (r.d_hit == 0 || r.d_hit > L + dt)) {
    P.R[i] = kInf;                                 // rc(seg[i0, i+L)) still occurs, so does rc(seg[i, i+L))
    return true;
  }
  return false;
}

// pass 2: group t covers ends e in (t*gs, (t+1)*gs] and starts i in [t*gs, (t+1)*gs)
// (parts: bit 0 = the forward-strand half, bit 1 = the other-strand half; the kernel gives each half its own thread
// so that the two long chains of a group run side by side)
//
// One run [a, b] of consecutive deep ends, forward strand: ONE chain from its LARGEST end b, run (b - a) deeper than
// max_k; it contains seg[e-L, e) for every smaller e of the run.
NM_HD void cert_run_fwd(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t a, uint32_t b, uint32_t &budget,
                        uint32_t &steps, uint32_t &seeds) {
  const ChainResult r = run_fwd_chain(ix, C32, P, b, P.kmax + (b - a), steps, seeds);
  // nu: the deepest explored length at which seg[b-nu, b) still occurs at least twice
  const uint32_t nu = r.d_hit ? r.d_hit - 1u : r.d_stop;
  uint32_t e = a;
  for (;;) {
    // settle ends from the left until one needs a chain of its own.  (That chain is run AFTER this loop: the lanes of a
    // warp meet such an end in different iterations, and a chain inside the loop would run for one lane at a time.)
    uint32_t need = 0;
    for (; e <= b; ++e) {
      if (P.G[e] != kDeep) continue;
      const uint32_t dt = b - e;
      if (cert_contain_fwd(P, r, b, e)) continue;
      const bool beyond_wall = r.wall && r.d_stop <= dt;  // e lies beyond the wall b's chain ran into
      // seg[e-l, e) lies inside seg[b-nu, b) for l <= nu - dt, so G[e] >= nu - dt + 1; and a unique suffix of
      // length g ending at e-1 makes the suffix of length g+1 ending at e unique, so G[e] <= G[e-1] + 1.  Where the
      // two bounds meet (a run of positions since the last difference between two copies of a repeat: G grows by
      // one per position) the value is settled without a chain of its own.
      // (e-1 may belong to the group on the left, whose thread may be writing it right now: what is read is either
      // kDeep or a final value, both are safe; only the number of chains run depends on the timing)
      if (!beyond_wall && nu >= dt) {
        const uint32_t left = P.G[e - 1];
        if (left != kNone && left != kDeep && left != kPending && left + 1u == nu - dt + 1u) {
          P.G[e] = left + 1u;
          continue;
        }
      }
      if (!need) need = e;                           // the first such end gets a chain (if the budget allows) ...
      if (budget) break;                             // ... and the ends after it may then interpolate from its value
    }
    if (!need || !budget) break;                     // (without budget the loop above ran to the end: the rest stays deep)
    --budget;
    P.G[need] = chain_value(P, run_fwd_chain(ix, C32, P, need, P.kmax, steps, seeds), P.kmax, kNone);
    e = need + 1;
  }
}
// mirror image, other strand: ONE chain from the SMALLEST start a of the run; it contains seg[i, i+L) for every larger i
NM_HD void cert_run_rc(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t a, uint32_t b, uint32_t &budget,
                       uint32_t &steps, uint32_t &seeds) {
  const ChainResult r = run_rc_chain(ix, C32, P, a, P.kmax + (b - a), steps, seeds);
  const uint32_t nu = r.d_hit ? r.d_hit - 1u : r.d_stop;  // rc(seg[a, a+nu)) still occurs
  uint32_t n = b - a + 1u;                           // starts a .. a+n-1 still to look at, from the right: R[i] <= R[i+1] + 1
  for (;;) {
    uint32_t need = kNoBarrier;
    for (; n > 0u; --n) {
      const uint32_t i = a + n - 1u;
      if (P.R[i] != kDeep) continue;
      const uint32_t dt = i - a;
      if (cert_contain_rc(P, r, a, i)) continue;
      const bool beyond_wall = r.wall && r.d_stop <= dt;
      if (!beyond_wall && nu >= dt && i + 1 < P.seg_len) {
        const uint32_t right = P.R[i + 1];
        if (right != kInf && right != kDeep && right != kPending && right + 1u == nu - dt + 1u) {
          P.R[i] = right + 1u;
          continue;
        }
      }
      if (need == kNoBarrier) need = i;
      if (budget) break;
    }
    if (need == kNoBarrier || !budget) break;
    --budget;
    P.R[need] = chain_value(P, run_rc_chain(ix, C32, P, need, P.kmax, steps, seeds), P.kmax, kInf);
    n = need - a;                                    // continue with the starts to the left of it
  }
}

// A group's deep entries fall into runs of consecutive positions: inside one repeat copy the runs are separated by
// the dcap positions after each difference between the copies (those were settled by the first pass), so a group no
// wider than dcap holds one run and a wider one holds a few.  Each run gets its own anchor chain; at most kCertRuns
// runs per group and strand are taken, the rest stays deep for the third pass.
constexpr uint32_t kCertRuns = 1;  // default of P.cert_runs (profiles/r2_sweep_i_cert_runs.jsonl; 4 with groups of 20 in round 1)
NM_HD void free_cert_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t t, uint32_t &steps,
                            uint32_t &seeds, uint32_t parts = 3u) {
  const uint32_t gs = P.gsize;
  if (parts & 1u) {  // ---- forward strand
    const uint32_t first = t * gs + 1;
    uint32_t last = first + gs - 1;
    if (last > P.seg_len) last = P.seg_len;
    if (first > last || !maybe_deep_range(P, first, last)) parts &= ~1u;
  }
  if (parts & 1u) {
    const uint32_t first = t * gs + 1;
    uint32_t last = first + gs - 1;
    if (last > P.seg_len) last = P.seg_len;
    // (the r-th runs of the lanes of a warp are taken in the same iteration of this loop, so their chains run side by
    // side; with the chain inside the scanning loop each lane would reach it in another iteration, one after the other)
    uint32_t budget = P.own_chains, e = first;
    for (uint32_t r = 0; r < P.cert_runs; ++r) {
      while (e <= last && P.G[e] != kDeep) ++e;
      if (e > last) break;
      uint32_t b = e;
      while (b < last && P.G[b + 1] == kDeep) ++b;
      cert_run_fwd(ix, C32, P, e, b, budget, steps, seeds);
      e = b + 1;
    }
  }
  if (P.use_rc && (parts & 2u)) {  // ---- other strand
    const uint32_t first = t * gs;
    uint32_t end = first + gs;
    if (end > P.num_kmers) end = P.num_kmers;
    if (first >= end || !maybe_deep_range(P, first, end - 1u)) return;
    uint32_t budget = P.own_chains, i = first;
    for (uint32_t r = 0; r < P.cert_runs; ++r) {
      while (i < end && P.R[i] != kDeep) ++i;
      if (i >= end) break;
      uint32_t b = i;
      while (b + 1 < end && P.R[b + 1] == kDeep) ++b;
      cert_run_rc(ix, C32, P, i, b, budget, steps, seeds);
      i = b + 1;
    }
  }
}

struct CoarseGroup {  // one (group, strand) of the coarse round: result of its anchor chain
  uint32_t d_hit, d_stop, wall;
};

// Coarse round: one anchor chain per group of P.gsize (256) positions and strand, run max_k + P.gsize - 1 deep, and
// the containment rule alone.  It settles the inside of long exact repeats (tandem arrays, recent duplications), where
// every group of the fine round would otherwise pay a chain of max_k steps.  The anchors are fixed positions -- the
// LAST end of the group for the forward strand, the FIRST start for the other strand -- so the chains need nothing
// from the first pass and run BESIDE it (another stream): their latency, max_k + 255 dependent steps, hides behind
// it, and outside repeats they end after a step or two.  What the round cannot settle stays deep.
// group t: ends e in (t*gs, (t+1)*gs], starts i in [t*gs, (t+1)*gs); part 0 forward strand, part 1 other strand.
NM_HD uint32_t coarse_anchor(const SegParams &P, uint32_t t, uint32_t part) {
  if (part == 0u) {
    const uint64_t e0 = ((uint64_t)t + 1u) * P.gsize;
    return e0 > P.seg_len ? P.seg_len : (uint32_t)e0;
  }
  return t * P.gsize;
}
NM_HD CoarseGroup coarse_anchor_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t t, uint32_t part,
                                       uint32_t &steps, uint32_t &seeds) {
  const uint32_t a = coarse_anchor(P, t, part);
  const bool wanted = part == 0u ? fwd_wanted(P, a) : rc_wanted(P, a);
  if (!wanted) return CoarseGroup{0u, 0u, 1u};   // a wall at depth 0: says nothing about any other position
  const ChainResult r = run_chain(ix, C32, P, part, a, P.kmax + P.gsize - 1u, steps, seeds);
  return CoarseGroup{r.d_hit, r.d_stop, r.wall ? 1u : 0u};
}
// position p: the end p (forward strand) and the start p (other strand), if still deep, against their group's anchor
NM_HD void coarse_settle_thread(const SegParams &P, const CoarseGroup *groups, uint32_t p) {
  if (p >= 1u && p <= P.seg_len && P.G[p] == kDeep) {
    const uint32_t t = (p - 1u) / P.gsize;
    const CoarseGroup g = groups[2u * t];
    cert_contain_fwd(P, ChainResult{g.d_hit, g.d_stop, g.wall != 0u}, coarse_anchor(P, t, 0u), p);
  }
  if (P.use_rc && p < P.num_kmers && P.R[p] == kDeep) {
    const uint32_t t = p / P.gsize;
    const CoarseGroup g = groups[2u * t + 1u];
    cert_contain_rc(P, ChainResult{g.d_hit, g.d_stop, g.wall != 0u}, coarse_anchor(P, t, 1u), p);
  }
}

// ---------------------------------------------------------------------------------------------
// (G) trajectory-free search, generic kernel: the same two quantities computed exactly (from
// depth 1, no seed table) on the unmodified .awfmi blocks with the five-letter alphabet and u64
// pointers.  Runs for entries marked kPending, or for every position when `all` is set
// (kmin < seed_len, or bwt_len >= 2^32-1).
// ---------------------------------------------------------------------------------------------
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
      const uint32_t kq = P.keff;                        // == kmin unless the range starts below the table depth
      if (lim >= kq) {
        if (r != kInf) {
          const uint32_t k0 = r > kq ? r : kq;
          if (k0 <= lim) {
            if (fwd_unique(P, i, i + k0)) {              // most positions: unique at the shortest length they may have
              ans = k0;
            } else if (k0 < lim && fwd_unique(P, i, i + lim)) {
              uint32_t lo = i + k0 + 1, hi = i + lim;
              while (lo < hi) {
                const uint32_t mid = lo + (hi - lo) / 2;
                if (fwd_unique(P, i, mid)) hi = mid; else lo = mid + 1;
              }
              ans = lo - i;
            }
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

// (V) strict mode (NM_OPT_STRICT): is every k-mer the reference could query present in the index?
// The reference raises on the first queried k-mer whose two-strand count is 0 (search.py:699-722) -- possible only
// when the sequence does not match the index.  The trajectory-free passes notice an absent string only when one of
// their chains runs into it, and a chain stops as soon as its interval has size 1, so a difference inside a
// near-identical repeat can go unseen.  Every queried k-mer is a prefix of S[i, i + lim(i)); all of those for the
// starts of a block [t*B, (t+1)*B) lie inside S[t*B, (t+1)*B - 1 + kmax), cut at barriers (no queried k-mer contains
// one): ONE backward search per barrier-free stretch of that string proves them all present, at (B + kmax) / B
// steps per position.  An empty interval raises CTR_ANOMALY, upon which nm_workspace_sync replays the segment on the
// reference's own trajectory, which reports the reference's error -- or none, when the reference's trajectory never
// queries the absent string; either way the bytes and the exception are the reference's.
NM_HD void presence_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t t, uint32_t B, uint32_t &steps) {
  const uint64_t first = (uint64_t)t * B;
  if (first >= P.num_kmers) return;
  uint64_t last_start = first + B - 1u;
  if (last_start >= P.num_kmers) last_start = P.num_kmers - 1u;
  uint64_t end = last_start + P.kmax;                       // one past the last letter any window of the block reaches
  if (end > P.seg_len) end = P.seg_len;
  bool have = false;
  uint32_t lo = 0, hi = 0;
  for (uint64_t p = end; p-- > first;) {
    const uint8_t c = P.codes[p];
    if (c & kBarrier) { have = false; continue; }
    const unsigned letter = c & kLetterMask;
    if (!have) {
      lo = C32[letter];
      hi = C32[letter + 1u];
      have = true;
    } else {
      lf_fast(ix, C32, letter, lo, hi);
      ++steps;
    }
    if (lo >= hi) {
      atomic_or64(&P.counters[CTR_ANOMALY], 1ull);
      return;
    }
  }
}

// (S) lengths kmin .. keff-1, below the depth of the table the chains started from (range mode, keff > kmin): after
// resolve_thread.  Possible only where the position is unique at keff -- or cannot reach keff at all -- and where the
// table does not rule it out (P.sub, subdepth_candidate): exact two-strand counts from the longest such length
// downwards (the count is non-increasing in the length).  A clean window is counted on the rank32 layout from the
// file's own seed table (its last seed_len letters looked up, the rest one step each); anything else -- an
// irregular letter nearby, a barrier or the segment end before keff -- with the reference's own count (count_both).
NM_HD uint32_t count_clean_fast(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t i, uint32_t k,
                                uint32_t &steps) {
  // S[i, i+k), k >= seed_len, inside a clean window of P.sl >= k letters starting at i: 2-bit letters from pk
  const uint32_t fsl = ix.seed_len;
  const uint32_t w = window_key(P, i);                       // P.sl letters, the first one most significant
  auto letter_at = [&](uint32_t j) -> unsigned { return (w >> (2u * (P.sl - 1u - j))) & 3u; };   // letter i + j
  uint32_t total = 0;
  {  // forward strand: the last fsl letters from the table, then leftwards
    uint32_t key = 0;
    for (uint32_t j = k - fsl; j < k; ++j) key = (key << 2) | letter_at(j);
    const SeedEntry32 e = ix.seed32[key];
    uint32_t lo = e.start, hi = e.start + e.len;
    for (uint32_t j = k - fsl; j-- > 0u && lo < hi;) { lf_fast(ix, C32, letter_at(j), lo, hi); ++steps; }
    total += hi > lo ? hi - lo : 0u;
  }
  if (P.use_rc) {  // other strand: the pattern is comp(S[i+k-1]) .. comp(S[i]); its last fsl letters are comp of S[i+fsl-1] .. S[i]
    uint32_t key = 0;
    for (uint32_t j = fsl; j-- > 0u;) key = (key << 2) | (3u - letter_at(j));
    const SeedEntry32 e = ix.seed32[key];
    uint32_t lo = e.start, hi = e.start + e.len;
    for (uint32_t j = fsl; j < k && lo < hi; ++j) { lf_fast(ix, C32, 3u - letter_at(j), lo, hi); ++steps; }
    total += hi > lo ? hi - lo : 0u;
  }
  return total;
}

NM_HD void subdepth_thread(const DevIndex &ix, const uint32_t *C32, const SegParams &P, uint32_t i, uint32_t &steps) {
  const uint8_t c = P.codes[i];
  if (c & kNoStart) return;
  const uint32_t kq = P.keff;
  const bool clean = !(c & kSeedBad);                      // then no barrier within P.sl = kq letters either: lim >= kq
  uint32_t ans;
  uint32_t lim = kq;
  if (clean) {
    if (!P.sub[i]) return;
    ans = load_out(P, i);
    if (ans != kq) return;                                 // not unique at kq: not unique at anything shorter
  } else {
    lim = length_limit(P, i);
    if (lim < P.kmin) return;
    ans = load_out(P, i);
    if (lim >= kq && ans != kq) return;
  }
  uint32_t k = lim < kq ? lim : kq - 1u;
  if (clean) {
    if (count_clean_fast(ix, C32, P, i, k, steps) != 1u) return;
    while (k > P.kmin && count_clean_fast(ix, C32, P, i, k - 1u, steps) == 1u) --k;
  } else {
    if (count_both(ix, P.seg, i, k, P.use_rc != 0, steps) != 1u) return;
    while (k > P.kmin && count_both(ix, P.seg, i, k - 1u, P.use_rc != 0, steps) == 1u) --k;
  }
  store_out(P, i, k);
}

// ---------------------------------------------------------------------------------------------
// index transcode: unmodified .awfmi blocks -> rank32 layout; seed table -> {u32 start, u32 len}
// one thread per 64 BWT positions (a quarter of a native block)
// ---------------------------------------------------------------------------------------------
NM_HD void transcode_thread(const uint8_t *native, uint64_t n_blocks, uint64_t bwt_len, uint32_t *rank32,
                            uint64_t b64, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
  const uint32_t C[4] = {c0, c1, c2, c3};   // prefixSums of a, c, g, t: folded into the block counts
  const uint64_t nb = b64 >> 2;
  const unsigned w = (unsigned)(b64 & 3);
  uint32_t *o = rank32 + b64 * kRank32Words;
  if (nb >= n_blocks) {  // pad block(s): only ever read with an empty mask
    const uint64_t *last = reinterpret_cast<const uint64_t *>(native + n_blocks * kNativeBlockBytes);  // pad block counters = totals
    for (int l = 0; l < 4; ++l) o[l] = C[l] + (uint32_t)last[12 + l];
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
    for (int l = 0; l < 4; ++l) o[l] = C[l] + (uint32_t)cnt[l];
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

// Deeper seed table, derived on the device at load time (one thread per entry of the level above): the
// SA interval of the (d+1)-mer c.w is one backward-extension step (AwFmSearch.c:44-95) of the interval of
// the d-mer w, and its key is c * 4^d + key(w) because the leftmost character is the most significant
// digit (AwFmKmerTable.c:24-35).  Both rank blocks are loaded once for the four letters.
NM_HD void seed_extend_thread(const DevIndex &ix, const uint32_t *C32, const SeedEntry32 *src, uint32_t d,
                              SeedEntry32 *dst, uint64_t key) {
  const SeedEntry32 s = src[key];
  const uint64_t stride = 1ull << (2 * d);
  if (s.len == 0) {
    for (unsigned c = 0; c < 4u; ++c) dst[c * stride + key] = SeedEntry32{0u, 0u};
    return;
  }
  const uint32_t lo = s.start, hi = s.start + s.len;
  const Rank32Block a = ld_rank32(ix.rank32, lo >> 6);
  const Rank32Block b = (hi >> 6) == (lo >> 6) ? a : ld_rank32(ix.rank32, hi >> 6);
  const bool poison = a.c[0] == kPoison || b.c[0] == kPoison;
  for (unsigned c = 0; c < 4u; ++c) {
    uint32_t nlo, nhi;
    if (!poison) {
      nlo = occ_in_block(a, c, lo & 63u);
      nhi = occ_in_block(b, c, hi & 63u);
    } else {
      nlo = C32[c] + (uint32_t)occ_native(ix.native, c, lo);
      nhi = C32[c] + (uint32_t)occ_native(ix.native, c, hi);
    }
    dst[c * stride + key] = SeedEntry32{nhi > nlo ? nlo : 0u, nhi > nlo ? nhi - nlo : 0u};
  }
}

// both-strand table from the finished one-strand table of depth d: pair[key] = {t[key], t[rc(key)]}
NM_HD void seed_pair_thread(const SeedEntry32 *t, uint32_t d, SeedEntryPair *pair, uint64_t key) {
  SeedEntryPair e;
  e.fwd = t[key];
  e.rc = t[rc_seed_key((uint32_t)key, d)];
  pair[key] = e;
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
