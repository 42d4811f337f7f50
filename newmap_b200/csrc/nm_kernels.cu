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
                                                      uint32_t *rank32, uint64_t n64_total) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n64_total) transcode_thread(native, n_blocks, bwt_len, rank32, t);
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

// prep, fused and vectorised: one thread = 8 consecutive positions, 8 threads = one kTile (64) tile.
// Writes what prep_thread + nb_tile_thread (nm_core.cuh, kept as the readable definition and for the
// host-side logic tests) write: codes, packed seed keys (rolling window), next barrier inside the tile,
// first barrier of the tile.  8-byte / 32-byte stores, tile scan with shuffles.
constexpr int kPrepV = 8;
static_assert(kTile == 64 && kPrepV * 8 == (int)kTile, "prep kernel assumes 8 threads x 8 positions per tile");

__global__ void __launch_bounds__(kBlock) k_prep(DevIndex ix, SegParams P) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t p0 = (uint64_t)t * kPrepV;
  const uint32_t sl = P.sl;
  const bool live = p0 < P.seg_len;
  uint32_t first_barrier = kNoBarrier;
  unsigned long long cw = 0;  // the thread's 8 codes, one per byte (no dynamically indexed arrays: no local memory)
  // the block's codes, each computed once, + the sl - 1 codes after them the last windows reach into
  __shared__ unsigned long long s_code[kBlock + 2];
  {
    const uint64_t b0 = (uint64_t)blockIdx.x * blockDim.x * kPrepV;
    if (live) {
#pragma unroll
      for (int j = 0; j < kPrepV; ++j) {
        const uint8_t cj = (p0 + j < P.seg_len) ? code_of(P, (uint32_t)(p0 + j)) : (uint8_t)kIrregular;
        cw |= (unsigned long long)cj << (8 * j);
      }
    } else {
      cw = ~0ull;  // beyond the segment: irregular, invalidates every window that reaches it
    }
    s_code[threadIdx.x] = cw;
    if (threadIdx.x < 16) {
      const uint64_t q = b0 + (uint64_t)kBlock * kPrepV + threadIdx.x;
      reinterpret_cast<uint8_t *>(s_code + kBlock)[threadIdx.x] =
          (q < P.seg_len) ? code_of(P, (uint32_t)q) : (uint8_t)kIrregular;
    }
    __syncthreads();
    if (!live) cw = 0;
  }
  if (live) {
    // rolling seed key; `bad` = characters still to shift out before the window is free of irregular /
    // barrier / out-of-range characters
    const unsigned long long w1 = s_code[threadIdx.x + 1], w2 = s_code[threadIdx.x + 2];
    auto code_at = [&](uint32_t j) -> uint32_t {  // code of position p0 + j, j < 24
      const unsigned long long w = j < 8u ? cw : (j < 16u ? w1 : w2);
      return (uint32_t)(w >> (8u * (j & 7u))) & 0xFFu;
    };
    uint32_t key = 0, bad = 0;
    const uint32_t keymask = (sl >= 16) ? 0xFFFFFFFFu : ((1u << (2 * sl)) - 1u);
    auto push = [&](uint32_t j) {
      const uint32_t cj = code_at(j);
      key = ((key << 2) | (cj & 3u)) & keymask;
      bad = (cj & (kIrregular | kBarrier)) ? sl : (bad ? bad - 1 : 0);
    };
    for (uint32_t j = 0; j + 1 < sl; ++j) push(j);
    uint32_t keys[kPrepV];
    unsigned long long badbits = 0;
#pragma unroll
    for (int i = 0; i < kPrepV; ++i) {
      push(sl - 1 + i);
      keys[i] = bad ? 0u : key;
      if (bad) badbits |= (unsigned long long)kSeedBad << (8 * i);
    }
    cw |= badbits;
    if (p0 + kPrepV <= P.seg_len) {
      *reinterpret_cast<uint2 *>(P.codes + p0) = make_uint2((uint32_t)cw, (uint32_t)(cw >> 32));
      uint4 *kd = reinterpret_cast<uint4 *>(P.seedkey + p0);
      kd[0] = make_uint4(keys[0], keys[1], keys[2], keys[3]);
      kd[1] = make_uint4(keys[4], keys[5], keys[6], keys[7]);
    } else {
#pragma unroll
      for (int j = 0; j < kPrepV; ++j)
        if (p0 + j < P.seg_len) { P.codes[p0 + j] = (uint8_t)(cw >> (8 * j)); P.seedkey[p0 + j] = keys[j]; }
    }
#pragma unroll
    for (int j = kPrepV - 1; j >= 0; --j)
      if (p0 + j < P.seg_len && ((cw >> (8 * j)) & kBarrier)) first_barrier = (uint32_t)(p0 + j);
  }
  // suffix-min over the 8 lanes of the tile: nearest barrier in the lanes to the right of this one
  uint32_t right = kNoBarrier, incl = first_barrier;
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
    const uint32_t v = __shfl_down_sync(0xffffffffu, incl, o, 8);
    if ((threadIdx.x & 7) + o < 8) incl = min(incl, v);
  }
  {
    const uint32_t v = __shfl_down_sync(0xffffffffu, incl, 1, 8);
    if ((threadIdx.x & 7) + 1 < 8) right = v;
  }
  if (!live) return;
  uint32_t next = right, nbv[kPrepV];
#pragma unroll
  for (int j = kPrepV - 1; j >= 0; --j) {
    if (p0 + j < P.seg_len && ((cw >> (8 * j)) & kBarrier)) next = (uint32_t)(p0 + j);
    nbv[j] = next;
  }
  if (p0 + kPrepV <= P.seg_len) {
    uint4 *nd = reinterpret_cast<uint4 *>(P.nb + p0);
    nd[0] = make_uint4(nbv[0], nbv[1], nbv[2], nbv[3]);
    nd[1] = make_uint4(nbv[4], nbv[5], nbv[6], nbv[7]);
  } else {
#pragma unroll
    for (int j = 0; j < kPrepV; ++j)
      if (p0 + j < P.seg_len) P.nb[p0 + j] = nbv[j];
  }
  if ((threadIdx.x & 7) == 0) P.tile_first[p0 / kTile] = incl;
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
__global__ void __launch_bounds__(kBlock, 8) k_free_chain(DevIndex ix, SegParams P, bool first) {
  __shared__ uint32_t sC[8];
  const uint32_t *C32 = stage_c32(ix, sC);
  // (giving the two chains of a position to different warps in pass 3 was measured slower: 0.134 against 0.110 ms)
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t steps = 0, seeds = 0;
  if (p <= P.seg_len) free_chain_thread(ix, C32, P, p, first, steps, seeds);
  flush_counter(&P.counters[first ? CTR_LF_PASS1 : CTR_LF], steps);  // nm_lib.cu adds the first pass into lf_steps
  flush_counter(&P.counters[first ? CTR_SEEDS_PASS1 : CTR_SEEDS], seeds);
}

// pass 1, tiled: one block = kP1Tile consecutive positions = 2 * kP1Tile chains (forward strand ending at p, other
// strand starting at p).  With a deep seed table most chains are over after their table lookup, so:
//   phase A  every thread issues its four seed lookups back to back (four independent random sectors in flight per
//            thread, no divergence), settles the chains that are already decided and queues the rest in shared memory;
//   phase B  the lanes pop queued chains and extend them one step per iteration, taking the next chain as soon as
//            theirs ends, so the warps stay full while the DRAM latency of each step is covered by the other lanes.
// Same values as free_chain_thread(first = true) (nm_core.cuh) position by position.
constexpr int kP1Win = 2;  // seed windows per thread (3 or 4: shorter kernel when timed alone, slower job: 8.62 -> 7.98 G positions/s)
constexpr int kP1Tile = kP1Win * kBlock;

__global__ void __launch_bounds__(kBlock, 8) k_pass1_tiled(DevIndex ix, SegParams P) {
  __shared__ uint32_t sC[8];
  __shared__ uint4 s_items[2 * kP1Tile + 16];  // {p, lo, hi, d | kind << 31}; + the seedless ends of block 0
  __shared__ uint32_t s_count, s_next;
  if (threadIdx.x == 0) s_count = s_next = 0;
  const uint32_t *C32 = stage_c32(ix, sC);  // __syncthreads inside
  const uint32_t cap = P.dcap < P.kmax ? P.dcap : P.kmax;
  const uint32_t t0 = blockIdx.x * kP1Tile;
  uint32_t steps = 0, seeds = 0;

  // ---- phase A.  Thread work = two seed windows q (letters q .. q+sl-1); a window seeds the other-strand chain that
  // STARTS at q and the forward-strand chain that ENDS at q + sl, so with the both-strand table one 16-byte entry
  // serves two chains.  Code bytes first, then keys, then entries: the loads of each group are in flight together.
  SeedEntry32 ent[2 * kP1Win];
  uint32_t st[2 * kP1Win];  // 0 outside the segment, 1 not wanted, 2 starts from a letter, 3 starts from the seed table
  {
    uint32_t cself[kP1Win], clast[kP1Win];  // codes[q], codes[q + sl - 1]
#pragma unroll
    for (int j = 0; j < kP1Win; ++j) {
      const uint32_t q = t0 + j * kBlock + threadIdx.x;
      const uint64_t e = (uint64_t)q + P.sl;
      cself[j] = q < P.seg_len ? (uint32_t)ld8(P.codes + q) : (uint32_t)(kNoStart | kSeedBad);
      clast[j] = e <= P.seg_len ? (uint32_t)ld8(P.codes + e - 1) : (uint32_t)kBarrier;
    }
    uint32_t key[kP1Win];
#pragma unroll
    for (int j = 0; j < kP1Win; ++j) {
      const uint32_t q = t0 + j * kBlock + threadIdx.x;
      const uint64_t e = (uint64_t)q + P.sl;
      const bool window_ok = !(cself[j] & kSeedBad);
      uint32_t v0 = 0, v1 = 0;
      if (e <= P.seg_len)  // fwd_wanted / chain_seeded(kind 0) for end e, on the bytes already loaded
        v0 = !(clast[j] & kBarrier) ? ((window_ok && !(P.mode == 1 && e == P.seg_len)) ? 3u : 2u) : 1u;
      if (q < P.seg_len)   // rc_wanted / chain_seeded(kind 1) for start q
        v1 = (P.use_rc && q < P.num_kmers && !(cself[j] & kNoStart)) ? (window_ok ? 3u : 2u) : 1u;
      st[2 * j] = v0;
      st[2 * j + 1] = v1;
      key[j] = (v0 == 3u || v1 == 3u) ? ld32(P.seedkey + q) : 0u;
    }
#pragma unroll
    for (int j = 0; j < kP1Win; ++j) {
      ent[2 * j] = ent[2 * j + 1] = SeedEntry32{0u, 0u};
      if (st[2 * j] == 3u || st[2 * j + 1] == 3u) {
        if (P.seedpair) {
          const uint4 e = __ldg(reinterpret_cast<const uint4 *>(P.seedpair + key[j]));
          ent[2 * j] = SeedEntry32{e.x, e.y};
          ent[2 * j + 1] = SeedEntry32{e.z, e.w};
          ++seeds;
        } else {
          if (st[2 * j] == 3u) {
            const uint2 e = __ldg(reinterpret_cast<const uint2 *>(P.seedtab + key[j]));
            ent[2 * j] = SeedEntry32{e.x, e.y};
            ++seeds;
          }
          if (st[2 * j + 1] == 3u) {
            const uint2 e = __ldg(reinterpret_cast<const uint2 *>(P.seedtab + rc_seed_key(key[j], P.sl)));
            ent[2 * j + 1] = SeedEntry32{e.x, e.y};
            ++seeds;
          }
        }
      }
    }
  }
#pragma unroll
  for (int slot = 0; slot < 2 * kP1Win; ++slot) {
    const uint32_t kind = slot & 1;
    const uint32_t p = t0 + (slot >> 1) * kBlock + threadIdx.x + (kind == 0u ? P.sl : 0u);
    if (st[slot] == 0) continue;
    uint32_t *dst = kind == 0u ? P.G + p : P.R + p;
    const uint32_t none = kind == 0u ? kNone : kInf;
    if (st[slot] == 1) { *dst = none; continue; }
    const ChainState s = st[slot] == 3 ? chain_from_seed(ent[slot], P.sl) : chain_from_letter(C32, P, kind, p);
    ChainResult r{0, 0, false};
    if (chain_finished(P, kind, s, r)) {
      *dst = chain_value(P, r, cap, none);
    } else {
      const uint32_t at = atomicAdd(&s_count, 1u);
      s_items[at] = make_uint4(p, s.lo, s.hi, s.d | (kind << 31));
    }
  }
  // forward-strand chains that end before the first whole window (e < sl) have no seed: the first block takes them
  if (blockIdx.x == 0 && threadIdx.x < P.sl && threadIdx.x <= P.seg_len) {
    const uint32_t e = threadIdx.x;
    if (fwd_wanted(P, e)) {
      const ChainState s = chain_from_letter(C32, P, 0u, e);
      ChainResult r{0, 0, false};
      if (chain_finished(P, 0u, s, r)) {
        P.G[e] = chain_value(P, r, cap, kNone);
      } else {
        const uint32_t at = atomicAdd(&s_count, 1u);
        s_items[at] = make_uint4(e, s.lo, s.hi, s.d);
      }
    } else {
      P.G[e] = kNone;
    }
  }
  __syncthreads();

  // ---- phase B
  const uint32_t count = s_count;
  bool have = false;
  uint32_t p = 0, kind = 0;
  ChainState s{0, 0, 0};
  StallState stall{0, 0};
  for (;;) {
    if (!have) {
      const uint32_t at = atomicAdd(&s_next, 1u);
      if (at >= count) break;
      const uint4 it = s_items[at];
      p = it.x;
      s = ChainState{it.y, it.z, it.w & 0x7FFFFFFFu};
      kind = it.w >> 31;
      stall = StallState{s.hi - s.lo, 0u};
      have = true;
    }
    ChainResult r{0, 0, false};
    if (chain_iterate(ix, C32, P, kind, p, s, cap, steps, r)) {
      *(kind == 0u ? P.G + p : P.R + p) = chain_value(P, r, cap, kind == 0u ? kNone : kInf);
      have = false;
    } else if (chain_stalled(P, stall, s)) {      // run_chain_pass1 (nm_core.cuh)
      *(kind == 0u ? P.G + p : P.R + p) = kDeep;
      have = false;
    }
  }
  flush_counter(&P.counters[CTR_LF_PASS1], steps);
  flush_counter(&P.counters[CTR_SEEDS_PASS1], seeds);
}

// pass 2: two threads per group of P.gsize positions (forward-strand chain, other-strand chain)
__global__ void __launch_bounds__(64, 16) k_free_cert(DevIndex ix, SegParams P, uint32_t ngroups) {
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
                      uint64_t n64_total, cudaStream_t st) {
  k_transcode<<<grid_for(n64_total), kBlock, 0, st>>>(native, n_blocks, bwt_len, rank32, n64_total);
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
  k_prep<<<grid_for(((uint64_t)P.seg_len + kPrepV - 1) / kPrepV), kBlock, 0, st>>>(ix, P);
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
  const unsigned g = grid_for((uint64_t)P.seg_len, kP1Tile);
  k_pass1_tiled<<<g ? g : 1u, kBlock, 0, st>>>(ix, P);
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
