// nm_kernels.cu -- CUDA kernels (sm_100a) of the `newmap search` hot path.
//
// Pointer-chasing rank work: no tensor cores.  What matters here is (i) enough independent
// searches in flight to cover HBM latency (one or two dependent chains per thread, blocks sized
// for full occupancy, grids >> 148 SMs), (ii) one 32-byte sector per rank query on the rank32
// layout, (iii) coalesced streaming of the segment arrays.  Per-thread bodies live in nm_core.cuh.
#include "nm_kernels.cuh"

namespace nm {

namespace {

constexpr int kBlock = 256;
// resident blocks per SM the register allocation aims at (tuning experiments build variants with other values)
#ifndef NM_CHAIN_MINB
#define NM_CHAIN_MINB 8
#endif
#ifndef NM_P1_MINB
#define NM_P1_MINB 6
#endif
#ifndef NM_CERT_MINB
#define NM_CERT_MINB 16
#endif

inline unsigned grid_for(uint64_t n, int block = kBlock) { return (unsigned)((n + block - 1) / block); }

__device__ __forceinline__ void flush_counter(unsigned long long *ctr, uint32_t v) {
  // warp-aggregate, one atomic per warp
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(ctr, (unsigned long long)v);
}
__device__ __forceinline__ void flush_counter64(unsigned long long *ctr, unsigned long long v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(ctr, v);
}

__global__ void __launch_bounds__(kBlock) k_transcode(const uint8_t *native, uint64_t n_blocks, uint64_t bwt_len,
                                                      uint32_t *rank32, uint64_t n64_total, uint4 C) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n64_total) transcode_thread(native, n_blocks, bwt_len, rank32, t, C.x, C.y, C.z, C.w);
}

__global__ void __launch_bounds__(kBlock) k_seed_transcode(const uint64_t *seed64, SeedEntry32 *seed32, uint64_t n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) seed_transcode_thread(seed64, seed32, t);
}

// one level of the deeper seed table: thread = one entry of depth d, writes its four children of depth d + 1
__global__ void __launch_bounds__(kBlock) k_seed_extend(DevIndex ix, const SeedEntry32 *src, uint32_t d, SeedEntry32 *dst) {
  __shared__ uint32_t sC[8];
  if (threadIdx.x < 6) sC[threadIdx.x] = ix.c32[threadIdx.x];
  __syncthreads();
  const uint64_t key = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (key < (1ull << (2 * d))) seed_extend_thread(ix, sC, src, d, dst, key);
}

__global__ void __launch_bounds__(kBlock) k_seed_pair(const SeedEntry32 *t, uint32_t d, SeedEntryPair *pair) {
  const uint64_t key = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (key < (1ull << (2 * d))) seed_pair_thread(t, d, pair, key);
}

template <bool RC>
__global__ void __launch_bounds__(kBlock) k_count(DevIndex ix, const uint8_t *seq, const uint64_t *idx,
                                                  const uint64_t *len, uint64_t n, uint32_t *counts,
                                                  unsigned long long *counters) {
  const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0;
  if (q < n) counts[q] = count_one<RC>(ix, seq, idx[q], len[q], steps);
  flush_counter(&counters[CTR_LF], steps);
}

// ---------------------------------------------------------------------------------------------
// prep: raw bytes -> codes (+ kSeedBad), letters packed 2 bits each, barrier bits and first barrier per tile.
// Writes what prep_thread + pack_thread + nb_tile_thread (nm_core.cuh: the readable definition, run by the
// host-side logic tests) write.  One block = kPrepTile positions; its bytes (+ the 15 the last seed windows reach
// into) are brought into shared memory by ONE bulk asynchronous copy (cp.async.bulk, the 1-D TMA path: no
// per-thread loads of the raw sequence at all), then one thread = 16 positions: 16 codes -> one 16-byte store,
// their letters -> one word of pk, their barrier bits -> a quarter of a barmask word (combined with shuffles).
// ---------------------------------------------------------------------------------------------
constexpr int kPrepV = 16;                      // positions per thread
constexpr int kPrepTile = kBlock * kPrepV;      // 4096 positions per block
constexpr int kPrepHalo = 32;                   // bytes after the tile the last windows may read (sl - 1 <= 15)
static_assert(kTile == 64 && kPrepV * 4 == (int)kTile, "prep kernel assumes 4 threads x 16 positions per tile");

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(kBlock) k_prep(DevIndex ix, SegParams P) {
  (void)ix;
  // the counters of the search that follows start here (no separate copy: a 128-byte host-to-device copy would queue
  // behind the 10 MB segment uploads of the other workspaces on the copy engine)
  if (blockIdx.x == 0 && threadIdx.x < CTR_N)
    P.counters[threadIdx.x] = (threadIdx.x == CTR_ZERO_KEY || threadIdx.x == CTR_MINLEN) ? ~0ull : 0ull;
  // ... and so does the map of positions the first pass leaves deep: every block clears the words of its own tile
  // (the grid covers seg_len + 32 positions, so the word of the end position seg_len is someone's)
  if (P.deepmap && threadIdx.x < kPrepTile / 32) {
    const uint64_t w = (uint64_t)blockIdx.x * (kPrepTile / 32) + threadIdx.x;
    if (w <= ((uint64_t)P.seg_len >> 5)) P.deepmap[w] = 0u;
  }
  // raw bytes of the tile from the 16-byte-aligned address at or below its first byte: head (<= 15) + tile + halo
  __shared__ alignas(128) uint8_t s_raw[kPrepTile + kPrepHalo + 16 + 16];
  __shared__ alignas(8) unsigned long long s_bar;
  __shared__ uint8_t s_code[kPrepTile + kPrepHalo];
  __shared__ uint8_t s_lut[kBlock];
  static_assert(kBlock == 256, "one table entry per thread");
  const uint64_t b0 = (uint64_t)blockIdx.x * kPrepTile;                 // first position of the tile
  const uint8_t *src = P.seg + b0;
  const uint32_t head = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u);
  const bool have = b0 < P.seg_len;                                     // (the last block may only have pk padding to write)
  uint64_t want = have ? (uint64_t)P.seg_len - b0 : 0;                  // bytes of the segment from the tile start on
  if (want > kPrepTile + kPrepHalo) want = kPrepTile + kPrepHalo;
  // (the buffers behind P.seg are padded: reading up to the next multiple of 16 past the segment end is in bounds)
  const uint32_t bytes = (uint32_t)((head + want + 15u) & ~15ull);
  if (have && threadIdx.x == 0) {
    const uint32_t bar = smem_u32(&s_bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(s_raw)), "l"(src - head), "r"(bytes), "r"(bar) : "memory");
  }
  // codes of the tile + halo, each computed once, through a 256-entry table (code_of, nm_core.cuh, evaluated once per
  // byte VALUE and block instead of once per byte): bits 0-6 = letter | irregular | no-start, bit 7 = the byte is 'N'
  // (the barrier of list mode; range mode: a no-start byte among the first num_kmers positions)
  {
    const uint8_t ch = (uint8_t)threadIdx.x;
    const unsigned l = letter_of(ch);
    const bool allowed = start_allowed(ch);
    uint8_t c = (uint8_t)l;
    if (l == 4u || !allowed) c |= kIrregular;
    if (!allowed) c |= kNoStart;
    if (ch == 'N') c |= 0x80;
    s_lut[threadIdx.x] = c;
  }
  __syncthreads();   // the barrier is initialised before anyone waits on it; the table is complete
  if (have) {
    const uint32_t bar = smem_u32(&s_bar);
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}"
                   : "=r"(done) : "r"(bar) : "memory");
    }
  }
  for (uint32_t j = threadIdx.x; j < kPrepTile + kPrepHalo; j += kBlock) {
    const uint64_t p = b0 + j;
    uint32_t c = kIrregular;                                            // beyond the segment: invalidates every window that reaches it
    if (p < P.seg_len) {
      const uint32_t t = s_lut[s_raw[head + j]];
      const bool barrier = (P.mode == 0) ? ((t & kNoStart) && p < P.num_kmers) : (t & 0x80u) != 0u;
      c = (t & 0x7Fu) | (barrier ? (uint32_t)kBarrier : 0u);
    }
    s_code[j] = (uint8_t)c;
  }
  __syncthreads();
  const uint32_t j0 = threadIdx.x * kPrepV;
  const uint64_t p0 = b0 + j0;
  const uint32_t sl = P.sl;
  // `bad` = characters still to shift out before the window is free of irregular / barrier / out-of-range characters
  uint32_t bad = 0;
  for (uint32_t j = 0; j + 1 < sl; ++j) bad = (s_code[j0 + j] & (kIrregular | kBarrier)) ? sl : (bad ? bad - 1 : 0);
  uint32_t cw[4] = {0, 0, 0, 0}, pkw = 0, barbits = 0;
#pragma unroll
  for (int i = 0; i < kPrepV; ++i) {
    bad = (s_code[j0 + sl - 1 + i] & (kIrregular | kBarrier)) ? sl : (bad ? bad - 1 : 0);
    uint32_t c = s_code[j0 + i];
    pkw = (pkw << 2) | ((p0 + i < P.seg_len) ? (c & 3u) : 0u);
    if (c & kBarrier) barbits |= 1u << i;
    if (bad) c |= kSeedBad;
    cw[i >> 2] |= c << (8 * (i & 3));
  }
  if (p0 < P.seg_len) {
    if (p0 + kPrepV <= P.seg_len) {
      *reinterpret_cast<uint4 *>(P.codes + p0) = make_uint4(cw[0], cw[1], cw[2], cw[3]);
    } else {
#pragma unroll
      for (int i = 0; i < kPrepV; ++i)
        if (p0 + i < P.seg_len) P.codes[p0 + i] = (uint8_t)(cw[i >> 2] >> (8 * (i & 3)));
    }
  }
  if (p0 < (uint64_t)P.seg_len + 32u) P.pk[p0 >> 4] = pkw;              // pk has seg_len / 16 + 2 words
  // barrier bits of the 64-position tile: 4 lanes x 16 bits
  unsigned long long m = (unsigned long long)barbits << (16 * (threadIdx.x & 3));
  m |= __shfl_xor_sync(0xffffffffu, m, 1);
  m |= __shfl_xor_sync(0xffffffffu, m, 2);
  if ((threadIdx.x & 3) == 0 && p0 < P.seg_len) {
    P.barmask[p0 / kTile] = m;
    P.tile_first[p0 / kTile] = m ? (uint32_t)p0 + (uint32_t)(__ffsll((long long)m) - 1) : kNoBarrier;
  }
}

__global__ void __launch_bounds__(kBlock) k_trajectory(DevIndex ix, SegParams P) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0, nq = 0;
  if (i < P.num_kmers) trajectory_thread(ix, P, i, steps, nq);
  flush_counter(&P.counters[CTR_LF], steps);
  flush_counter(&P.counters[CTR_QUERIES], nq);
}

__global__ void __launch_bounds__(kBlock) k_trajectory_multi(const MultiSet *M, SegParams P) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0, nq = 0, nshort = 0;
  if (i < P.num_kmers) trajectory_multi_thread(*M, P, i, steps, nq, nshort);
  flush_counter(&P.counters[CTR_LF], steps);
  flush_counter(&P.counters[CTR_QUERIES], nq);
  flush_counter(&P.counters[CTR_SHORT], nshort);
}

// prefix sums as u32[6] staged in shared memory (dynamic indexing of kernel parameters would spill to local memory)
__device__ __forceinline__ const uint32_t *stage_c32(const DevIndex &ix, uint32_t *sC) {
  if (threadIdx.x < 6) sC[threadIdx.x] = ix.c32[threadIdx.x];
  __syncthreads();
  return sC;
}

// pass 1 (first = true): every position, depth-capped; pass 3 (first = false): positions still kDeep
__global__ void __launch_bounds__(kBlock, NM_CHAIN_MINB) k_free_chain(DevIndex ix, SegParams P, bool first) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  // (giving the two chains of a position to different warps in pass 3 was measured slower: 0.134 against 0.110 ms)
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0, seeds = 0;
  if (p <= P.seg_len) free_chain_thread(ix, C32, P, p, first, steps, seeds);
  flush_counter(&P.counters[first ? CTR_LF_PASS1 : CTR_LF], steps);  // nm_lib.cu adds the first pass into lf_steps
  flush_counter(&P.counters[first ? CTR_SEEDS_PASS1 : CTR_SEEDS], seeds);
}

// ---------------------------------------------------------------------------------------------
// pass 1, sparse (nm_core.cuh "Sparse first pass"; sparse_side_thread is the serial definition of what one cell
// side does).  One block = kSpCells * kBlock cells of P.stride positions; a cell has a forward side (its ends,
// walked upwards) and an other-strand side (its starts, walked downwards), and ONE window of the both-strand table
// -- the one that ends at the cell's first end and starts at its last start -- opens both.
//   phase A  every thread issues the table lookups of its cells back to back (independent random sectors in
//            flight, no divergence) and walks each side as far as it goes without touching the index again: with a
//            deep table most cells are finished here, one 16-byte access for `stride` positions and both strands;
//   phase B  the sides that are not finished are queued in shared memory; the lanes pop them and run them to their
//            end, one memory access per loop iteration whatever the side is doing -- a table lookup for the next
//            unsettled position or one backward-extension step of its chain -- with the addresses computed first
//            and the 32-byte loads issued in one place, so lanes in different states share an iteration.
// ---------------------------------------------------------------------------------------------
constexpr int kSpCells = 2;                              // cells per thread
constexpr int kSpItems = 2 * kSpCells * kBlock + 8;      // sides a block can queue (+ the head cell of block 0)

__global__ void __launch_bounds__(kBlock, NM_P1_MINB) k_pass1_sparse(DevIndex ix, SegParams P) {
  __shared__ uint32_t sC[8];
  __shared__ uint4 s_items[kSpItems];    // {position, lo, hi, depth | kind << 31 | needs-seed << 30}
  __shared__ uint32_t s_limit[kSpItems]; // last end of the cell (forward side) / first start of the cell (other strand)
  __shared__ uint32_t s_count, s_next;
  if (threadIdx.x == 0) s_count = s_next = 0;
  const uint32_t *C32 = stage_c32(ix, sC);  // __syncthreads inside
  const uint32_t cap = P.dcap < P.kmax ? P.dcap : P.kmax;
  const uint32_t S = P.stride, sl = P.sl;
  const uint32_t *pair_words = reinterpret_cast<const uint32_t *>(P.seedpair);
  uint32_t steps = 0, seeds = 0;

  auto push = [&](uint32_t kind, uint32_t pos, uint32_t limit, const ChainState &s, bool need_seed) {
    const uint32_t at = atomicAdd(&s_count, 1u);
    s_items[at] = make_uint4(pos, s.lo, s.hi, s.d | (kind << 31) | (need_seed ? 1u << 30 : 0u));
    s_limit[at] = limit;
  };
  // the value v of `pos` is known: store it, spread the bounds it gives, find the side's next position that needs a
  // chain.  true = the side goes on at `pos` (updated); false = the side is finished
  auto settle = [&](uint32_t kind, uint32_t &pos, uint32_t limit, uint32_t v) -> bool {
    if (v == kDeep) mark_deep(P, pos);
    if (kind == 0u) {
      P.G[pos] = v;
      uint32_t cur = sparse_spread_fwd(P, pos, limit, v);
      if (!sparse_next_fwd(P, cur, limit)) return false;
      pos = cur;
    } else {
      P.R[pos] = v;
      uint32_t end1 = sparse_spread_rc(P, limit, pos + 1u, v);
      if (!sparse_next_rc(P, limit, end1)) return false;
      pos = end1 - 1u;
    }
    return true;
  };
  // entry of the both-strand table out of the 32-byte sector that holds it (two entries per sector)
  auto entry_of = [&](const Rank32Block &sec, uint32_t key, uint32_t kind) -> SeedEntry32 {
    if (key & 1u) return kind == 0u ? SeedEntry32{(uint32_t)sec.p0, (uint32_t)(sec.p0 >> 32)}
                                    : SeedEntry32{(uint32_t)sec.p1, (uint32_t)(sec.p1 >> 32)};
    return kind == 0u ? SeedEntry32{sec.c[0], sec.c[1]} : SeedEntry32{sec.c[2], sec.c[3]};
  };

  // ---- phase A
  {
    uint64_t q[kSpCells];
    bool win_ok[kSpCells];
    uint32_t key[kSpCells];
    uint4 ent[kSpCells];
#pragma unroll
    for (int j = 0; j < kSpCells; ++j) {
      const uint64_t c = ((uint64_t)blockIdx.x * kSpCells + j) * kBlock + threadIdx.x;
      q[j] = c * S + (S - 1u);            // window of the cell: ends at its first end, starts at its last start
      win_ok[j] = q[j] + sl <= P.seg_len && !(ld8(P.codes + q[j]) & kSeedBad);
    }
#pragma unroll
    for (int j = 0; j < kSpCells; ++j) key[j] = win_ok[j] ? window_key(P, (uint32_t)q[j]) : 0u;
#pragma unroll
    for (int j = 0; j < kSpCells; ++j) {
      ent[j] = make_uint4(0u, 0u, 0u, 0u);
      if (win_ok[j]) {
        if (P.sub) {
          // a range that starts below the table depth (stride 1, q = the start position): the four entries that share
          // the window's first sl-1 letters, one aligned 64-byte group (subdepth_candidate, nm_core.cuh)
          const uint4 *grp = reinterpret_cast<const uint4 *>(P.seedpair + (key[j] & ~3u));
          const uint4 g0 = __ldg(grp), g1 = __ldg(grp + 1), g2 = __ldg(grp + 2), g3 = __ldg(grp + 3);
          const uint32_t k3 = key[j] & 3u;
          ent[j] = k3 == 0u ? g0 : k3 == 1u ? g1 : k3 == 2u ? g2 : g3;
          uint32_t lb = g0.y + g1.y + g2.y + g3.y;
          if (P.use_rc) lb += g0.w + g1.w + g2.w + g3.w;
          P.sub[q[j]] = lb <= 1u ? 1 : 0;
        } else {
          ent[j] = __ldg(reinterpret_cast<const uint4 *>(P.seedpair + key[j]));
        }
        ++seeds;
      } else if (P.sub && q[j] < P.seg_len) {
        P.sub[q[j]] = 0;
      }
    }
#pragma unroll
    for (int j = 0; j < kSpCells; ++j) {
      // forward side: ends [q + sl, q + sl + S - 1]
      const uint64_t e0 = q[j] + sl;
      if (e0 <= P.seg_len) {
        uint32_t cur = (uint32_t)e0;
        const uint32_t last = e0 + S - 1u > P.seg_len ? P.seg_len : (uint32_t)(e0 + S - 1u);
        if (sparse_next_fwd(P, cur, last)) {
          if (cur == (uint32_t)e0 && win_ok[j] && !sparse_exact_only_end(P, cur)) {      // chain_seeded(P, 0, cur)
            const ChainState s = chain_from_seed(SeedEntry32{ent[j].x, ent[j].y}, sl);
            ChainResult r{0, 0, false};
            if (!chain_finished(P, 0u, s, r)) push(0u, cur, last, s, false);
            else if (settle(0u, cur, last, chain_value(P, r, cap, kNone))) push(0u, cur, last, ChainState{0, 0, 0}, true);
          } else {
            push(0u, cur, last, ChainState{0, 0, 0}, true);
          }
        }
      }
      // other-strand side: starts [q - (S - 1), q]
      const uint64_t first64 = q[j] - (S - 1u);
      if (first64 < P.seg_len) {
        const uint32_t first = (uint32_t)first64;
        uint32_t end1 = q[j] < P.seg_len ? (uint32_t)q[j] + 1u : P.seg_len;
        if (sparse_next_rc(P, first, end1)) {
          uint32_t pos = end1 - 1u;
          if (pos == (uint32_t)q[j] && win_ok[j]) {                                      // chain_seeded(P, 1, pos)
            const ChainState s = chain_from_seed(SeedEntry32{ent[j].z, ent[j].w}, sl);
            ChainResult r{0, 0, false};
            if (!chain_finished(P, 1u, s, r)) push(1u, pos, first, s, false);
            else if (settle(1u, pos, first, chain_value(P, r, cap, kInf))) push(1u, pos, first, ChainState{0, 0, 0}, true);
          } else {
            push(1u, pos, first, ChainState{0, 0, 0}, true);
          }
        }
      }
    }
    // the head cell: the ends [0, sl + S - 1) no window of the grid ends at
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      uint32_t cur = 0;
      const uint32_t last = sparse_fwd_cell_last(P, 0u);
      if (sparse_next_fwd(P, cur, last)) push(0u, cur, last, ChainState{0, 0, 0}, true);
    }
  }
  __syncthreads();

  // ---- phase B
  const uint32_t count = s_count;
  bool have = false, need_seed = false;
  uint32_t kind = 0, pos = 0, limit = 0;
  ChainState s{0, 0, 0};
  StallState stall{0, 0};
  for (;;) {
    if (!have) {
      const uint32_t at = atomicAdd(&s_next, 1u);
      if (at >= count) break;
      const uint4 it = s_items[at];
      pos = it.x;
      limit = s_limit[at];
      s = ChainState{it.y, it.z, it.w & 0x3FFFFFFFu};
      kind = it.w >> 31;
      need_seed = (it.w >> 30) & 1u;
      stall = StallState{s.hi - s.lo, 0u};
      have = true;
    }
    // -- what this iteration does: 0 nothing to load, 1 one step on rank32, 2 table lookup, 3 one step with the letter x
    uint32_t action = 0, key = 0;
    unsigned letter = 0;
    bool done = false;
    ChainResult r{0, 0, false};
    if (need_seed) {
      if (chain_seeded(P, kind, pos)) {
        key = window_key(P, kind == 0u ? pos - sl : pos);
        action = 2u;
      } else {
        s = chain_from_letter(C32, P, kind, pos);
        stall = StallState{s.hi - s.lo, 0u};
        need_seed = false;
      }
    }
    if (!need_seed) {
      if (chain_finished(P, kind, s, r)) {
        done = true;
      } else {
        r.d_hit = 0;
        r.d_stop = s.d;
        const bool edge = kind == 0u ? s.d >= pos : pos + s.d >= P.seg_len;
        if (edge) {
          r.wall = true;
          done = true;
        } else {
          const uint8_t c = ld8(P.codes + (kind == 0u ? pos - 1u - s.d : pos + s.d));
          if (c & kBarrier) {
            r.wall = true;
            done = true;
          } else if (s.d >= cap) {
            done = true;
          } else {
            letter = kind == 0u ? (unsigned)(c & kLetterMask) : rc_letter_of_code(c);
            action = letter < 4u ? 1u : 3u;
          }
        }
      }
    }
    // -- the loads of the iteration, in one place
    Rank32Block A{{0, 0, 0, 0}, 0, 0}, B{{0, 0, 0, 0}, 0, 0};
    if (action == 1u || action == 2u) {
      const uint32_t *base = action == 1u ? ix.rank32 : pair_words;
      const uint32_t ia = action == 1u ? s.lo >> 6 : key >> 1;
      const uint32_t ib = action == 1u ? s.hi >> 6 : ia;
      A = ld_rank32(base, ia);
      B = A;
      if (ib != ia) B = ld_rank32(base, ib);
    }
    // -- what came back
    bool stalled = false;
    if (action == 2u) {
      ++seeds;
      s = chain_from_seed(entry_of(A, key, kind), sl);
      stall = StallState{s.hi - s.lo, 0u};
      need_seed = false;
      done = chain_finished(P, kind, s, r);
    } else if (action == 1u) {
      uint32_t nlo = occ_in_block(A, letter, s.lo & 63u), nhi = occ_in_block(B, letter, s.hi & 63u);
      if (A.c[0] == kPoison || B.c[0] == kPoison) {      // a block with an x or $ symbol: the unmodified .awfmi block
        nlo = C32[letter] + (uint32_t)occ_native(ix.native, letter, s.lo);
        nhi = C32[letter] + (uint32_t)occ_native(ix.native, letter, s.hi);
      }
      s.lo = nlo;   // (the block counts include the prefix sum of the letter)
      s.hi = nhi;
      ++s.d;
      ++steps;
      stalled = chain_stalled(P, stall, s);
    } else if (action == 3u) {
      lf_fast(ix, C32, letter, s.lo, s.hi);
      ++s.d;
      ++steps;
      stalled = chain_stalled(P, stall, s);
    }
    if (done || stalled) {
      const uint32_t v = stalled ? kDeep : chain_value(P, r, cap, kind == 0u ? kNone : kInf);
      if (settle(kind, pos, limit, v)) need_seed = true;
      else have = false;
    }
  }
  flush_counter(&P.counters[CTR_LF_PASS1], steps);
  flush_counter(&P.counters[CTR_SEEDS_PASS1], seeds);
}

// pass 2: two threads per group of P.gsize positions (forward-strand chain, other-strand chain)
__global__ void __launch_bounds__(64, NM_CERT_MINB) k_free_cert(DevIndex ix, SegParams P, uint32_t ngroups) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  // a warp = 32 consecutive groups of ONE strand (even warps: forward, odd warps: other strand): the two halves of
  // free_cert_thread are different code, lanes of both kinds in one warp would run them one after the other
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t warp = t >> 5, group = (warp >> 1) * 32u + (t & 31u);
  uint32_t steps = 0, seeds = 0;
  if (group < ngroups) free_cert_thread(ix, C32, P, group, steps, seeds, 1u << (warp & 1u));
  flush_counter(&P.counters[CTR_LF], steps);
  flush_counter(&P.counters[CTR_SEEDS], seeds);
}

// coarse round (coarse_anchor_thread / coarse_settle_thread, nm_core.cuh): the anchor chains run on the workspace's
// second stream beside the first pass (blocks of 32 threads: the few long chains hold few thread slots), the settle
// kernel after both
__global__ void __launch_bounds__(32) k_coarse_anchor(DevIndex ix, SegParams P, uint32_t ngroups, CoarseGroup *groups) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  // a warp = 32 consecutive groups of one strand, like k_free_cert
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t warp = j >> 5, t = (warp >> 1) * 32u + (j & 31u), part = warp & 1u;
  uint32_t steps = 0, seeds = 0;
  if (t < ngroups && (part == 0u || P.use_rc)) groups[2u * t + part] = coarse_anchor_thread(ix, C32, P, t, part, steps, seeds);
  flush_counter(&P.counters[CTR_LF], steps);
  flush_counter(&P.counters[CTR_SEEDS], seeds);
}
__global__ void __launch_bounds__(kBlock) k_coarse_settle(SegParams P, const CoarseGroup *groups) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p <= P.seg_len) coarse_settle_thread(P, groups, p);
}

__global__ void __launch_bounds__(kBlock) k_free_generic(DevIndex ix, SegParams P, bool all) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0, ns = 0;
  if (p <= P.seg_len) free_generic_thread(ix, P, p, all, steps, ns);
  flush_counter(&P.counters[CTR_LF], steps);
  flush_counter(&P.counters[CTR_GENERIC], ns);
}

__global__ void __launch_bounds__(kBlock) k_resolve(DevIndex ix, SegParams P) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t ref_lf = 0;
  if (i < P.num_kmers) resolve_thread(ix, P, i, ref_lf);
  if (P.want_ref_lf) flush_counter64(&P.counters[CTR_REF_LF], ref_lf);
}

// strict mode (presence_thread, nm_core.cuh): one long chain per block of B start positions, beside the search
__global__ void __launch_bounds__(64) k_presence(DevIndex ix, SegParams P, uint32_t B, uint32_t nblocks) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0;
  if (t < nblocks) presence_thread(ix, C32, P, t, B, steps);
  flush_counter(&P.counters[CTR_LF], steps);
}

// lengths below the depth of the table the chains started from (subdepth_thread, nm_core.cuh): after k_resolve
__global__ void __launch_bounds__(kBlock) k_subdepth(DevIndex ix, SegParams P) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0;
  if (i < P.num_kmers) subdepth_thread(ix, C32, P, i, steps);
  flush_counter(&P.counters[CTR_LF], steps);
}

// NM_OPT_OUT_STATS: what `newmap search --verbose` reports per sequence (search.py:318-347,885-902), reduced on the device
__global__ void __launch_bounds__(kBlock) k_out_stats(SegParams P) {
  uint32_t uniq = 0, amb = 0, mn = 0xFFFFFFFFu, mx = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.num_kmers; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t v = load_out(P, (uint32_t)i);
    if (v) { ++uniq; mn = min(mn, v); mx = max(mx, v); }
    if (!start_allowed(ld8(P.seg + i))) ++amb;
  }
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  flush_counter(&P.counters[CTR_UNIQUE], uniq);
  flush_counter(&P.counters[CTR_AMBIG], amb);
  if ((threadIdx.x & 31) == 0) {
    if (mx) atomicMax(&P.counters[CTR_MAXLEN], (unsigned long long)mx);
    if (mn != 0xFFFFFFFFu) atomicMin(&P.counters[CTR_MINLEN], (unsigned long long)mn);
  }
}

__global__ void __launch_bounds__(kBlock) k_checksum(const void *out, int width, uint64_t n, unsigned long long *sum) {
  unsigned long long acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t v = width == 1 ? ((const uint8_t *)out)[i] : width == 2 ? ((const uint16_t *)out)[i] : ((const uint32_t *)out)[i];
    acc += checksum_term(i, v);
  }
  flush_counter64(sum, acc);
}

// random-access probe.  DEP = true: every thread walks MLP chains whose next offset depends on the data just
// loaded (loads of one chain are serialised exactly like the LF steps of one search).  DEP = false: the
// offsets depend only on (thread, step), so all loads of a thread are independent -- the most the memory
// system can do on random granules.  n_granules is a power of two; 32-bit hashing keeps the ALU work small.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
template <int WORDS, int MLP, bool DEP>
__global__ void __launch_bounds__(kBlock) k_probe(const uint32_t *buf, uint32_t granule_mask, uint64_t chains,
                                                  uint32_t steps, unsigned long long *sink) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= chains) return;
  uint32_t x[MLP];
#pragma unroll
  for (int m = 0; m < MLP; ++m) x[m] = mix32((uint32_t)(t * MLP + m) * 0x9E3779B9u + 0x1234567u);
  uint32_t acc = 0;
  for (uint32_t s = 0; s < steps; ++s) {
    uint32_t v[MLP];
#pragma unroll
    for (int m = 0; m < MLP; ++m) {
      x[m] = mix32(x[m] + s);
      const uint4 *p = reinterpret_cast<const uint4 *>(buf + (size_t)(x[m] & granule_mask) * WORDS);
      uint32_t a = 0;
#pragma unroll
      for (int j = 0; j < WORDS / 4; ++j) {
        const uint4 q = __ldg(p + j);
        a += q.x ^ q.y ^ q.z ^ q.w;
      }
      v[m] = a;
    }
#pragma unroll
    for (int m = 0; m < MLP; ++m) {
      acc += v[m];
      if (DEP) x[m] += v[m];  // data dependence
    }
  }
  if (acc == 0x7fffffffu) atomicAdd(sink, 1ull);
}

}  // namespace

void launch_transcode(const uint8_t *native, uint64_t n_blocks, uint64_t bwt_len, uint32_t *rank32,
                      uint64_t n64_total, const uint64_t *prefix_sums, cudaStream_t st) {
  const uint4 C = make_uint4((uint32_t)prefix_sums[0], (uint32_t)prefix_sums[1], (uint32_t)prefix_sums[2], (uint32_t)prefix_sums[3]);
  k_transcode<<<grid_for(n64_total), kBlock, 0, st>>>(native, n_blocks, bwt_len, rank32, n64_total, C);
}
void launch_seed_transcode(const uint64_t *seed64, SeedEntry32 *seed32, uint64_t entries, cudaStream_t st) {
  k_seed_transcode<<<grid_for(entries), kBlock, 0, st>>>(seed64, seed32, entries);
}
void launch_seed_extend(const DevIndex &ix, const SeedEntry32 *src, uint32_t d, SeedEntry32 *dst, cudaStream_t st) {
  k_seed_extend<<<grid_for(1ull << (2 * d)), kBlock, 0, st>>>(ix, src, d, dst);
}
void launch_seed_pair(const SeedEntry32 *t, uint32_t d, SeedEntryPair *pair, cudaStream_t st) {
  k_seed_pair<<<grid_for(1ull << (2 * d)), kBlock, 0, st>>>(t, d, pair);
}
void launch_count(const DevIndex &ix, const uint8_t *seq, const uint64_t *idx, const uint64_t *len, uint64_t n,
                  bool rc, uint32_t *counts, unsigned long long *counters, cudaStream_t st) {
  if (n == 0) return;
  if (rc) k_count<true><<<grid_for(n), kBlock, 0, st>>>(ix, seq, idx, len, n, counts, counters);
  else k_count<false><<<grid_for(n), kBlock, 0, st>>>(ix, seq, idx, len, n, counts, counters);
}
void launch_prep(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  if (P.seg_len == 0) return;
  k_prep<<<grid_for((uint64_t)P.seg_len + 32u, kPrepTile), kBlock, 0, st>>>(ix, P);
}
void launch_trajectory(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  k_trajectory<<<grid_for(P.num_kmers), kBlock, 0, st>>>(ix, P);
}
void launch_trajectory_multi(const MultiSet *M, const SegParams &P, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  k_trajectory_multi<<<grid_for(P.num_kmers), kBlock, 0, st>>>(M, P);
}
void launch_free_chain(const DevIndex &ix, const SegParams &P, bool first, cudaStream_t st) {
  k_free_chain<<<grid_for((uint64_t)P.seg_len + 1), kBlock, 0, st>>>(ix, P, first);
}
void launch_free_pass1(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  // cells: the forward sides cover the ends up to seg_len, the other-strand sides the starts below seg_len
  const uint64_t cells = ((uint64_t)P.seg_len + P.stride) / P.stride + 1u;
  const unsigned g = grid_for(cells, kSpCells * kBlock);
  k_pass1_sparse<<<g ? g : 1u, kBlock, 0, st>>>(ix, P);
}
void launch_free_cert(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  const uint32_t ngroups = (uint32_t)(((uint64_t)P.seg_len + 1 + P.gsize - 1) / P.gsize);
  k_free_cert<<<grid_for(2ull * (((uint64_t)ngroups + 31) / 32 * 32), 64), 64, 0, st>>>(ix, P, ngroups);
}
void launch_coarse_anchor(const DevIndex &ix, const SegParams &P, CoarseGroup *groups, cudaStream_t st) {
  const uint32_t ngroups = (uint32_t)(((uint64_t)P.seg_len + 1 + P.gsize - 1) / P.gsize);
  k_coarse_anchor<<<grid_for(2ull * (((uint64_t)ngroups + 31) / 32 * 32), 32), 32, 0, st>>>(ix, P, ngroups, groups);
}
void launch_coarse_settle(const SegParams &P, const CoarseGroup *groups, cudaStream_t st) {
  k_coarse_settle<<<grid_for((uint64_t)P.seg_len + 1), kBlock, 0, st>>>(P, groups);
}
void launch_free_generic(const DevIndex &ix, const SegParams &P, bool all, cudaStream_t st) {
  k_free_generic<<<grid_for((uint64_t)P.seg_len + 1), kBlock, 0, st>>>(ix, P, all);
}
void launch_resolve(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  k_resolve<<<grid_for(P.num_kmers), kBlock, 0, st>>>(ix, P);
}
void launch_presence(const DevIndex &ix, const SegParams &P, uint32_t B, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  const uint32_t nblocks = (uint32_t)(((uint64_t)P.num_kmers + B - 1) / B);
  k_presence<<<grid_for(nblocks, 64), 64, 0, st>>>(ix, P, B, nblocks);
}
void launch_subdepth(const DevIndex &ix, const SegParams &P, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  k_subdepth<<<grid_for(P.num_kmers), kBlock, 0, st>>>(ix, P);
}
void launch_out_stats(const SegParams &P, cudaStream_t st) {
  if (P.num_kmers == 0) return;
  unsigned g = grid_for(P.num_kmers);
  if (g > 148 * 8) g = 148 * 8;
  k_out_stats<<<g, kBlock, 0, st>>>(P);
}
void launch_checksum(const void *out, int out_width, uint64_t n, unsigned long long *sum, cudaStream_t st) {
  if (n == 0) return;
  unsigned g = grid_for(n);
  if (g > 148 * 8) g = 148 * 8;
  k_checksum<<<g, kBlock, 0, st>>>(out, out_width, n, sum);
}
template <int WORDS, bool DEP>
static void launch_probe_w(const uint32_t *b, uint32_t mask, uint64_t chains, uint32_t steps, uint32_t mlp,
                           unsigned long long *sink, cudaStream_t st) {
  if (mlp >= 8) k_probe<WORDS, 8, DEP><<<grid_for(chains), kBlock, 0, st>>>(b, mask, chains, steps, sink);
  else if (mlp >= 4) k_probe<WORDS, 4, DEP><<<grid_for(chains), kBlock, 0, st>>>(b, mask, chains, steps, sink);
  else if (mlp >= 2) k_probe<WORDS, 2, DEP><<<grid_for(chains), kBlock, 0, st>>>(b, mask, chains, steps, sink);
  else k_probe<WORDS, 1, DEP><<<grid_for(chains), kBlock, 0, st>>>(b, mask, chains, steps, sink);
}
void launch_probe(const uint8_t *buf, uint64_t n_granules_pow2, uint32_t granule, uint64_t chains, uint32_t steps,
                  uint32_t mlp, bool dependent, unsigned long long *sink, cudaStream_t st) {
  const uint32_t *b = reinterpret_cast<const uint32_t *>(buf);
  const uint32_t mask = (uint32_t)(n_granules_pow2 - 1);
  if (dependent) {
    if (granule == 32) launch_probe_w<8, true>(b, mask, chains, steps, mlp, sink, st);
    else if (granule == 64) launch_probe_w<16, true>(b, mask, chains, steps, mlp, sink, st);
    else launch_probe_w<32, true>(b, mask, chains, steps, mlp, sink, st);
  } else {
    if (granule == 32) launch_probe_w<8, false>(b, mask, chains, steps, mlp, sink, st);
    else if (granule == 64) launch_probe_w<16, false>(b, mask, chains, steps, mlp, sink, st);
    else launch_probe_w<32, false>(b, mask, chains, steps, mlp, sink, st);
  }
}

}  // namespace nm
