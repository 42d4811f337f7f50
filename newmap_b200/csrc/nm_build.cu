// nm_build.cu -- GPU construction of a reference-format .awfmi (v8, DNA) index.
//
// the reference builds its index on the CPU (FastaVector -> sanitize ->
// libdivsufsort -> BWT blocks -> prefix sums -> recursive seed table -> bit-packed sampled suffix
// array -> file; AwFmCreate.c:128-389, AwFmSuffixArray.c:52-96, AwFmFile.c:20-171), which takes
// minutes per Gbp.  This builds the same FILE, byte for byte, on the device: prefix-doubling suffix
// sort (radix sort of (rank, rank+h) pairs, only the still-ambiguous suffixes are re-sorted each
// round), BWT bit-planes + counters, level-by-level seed table with the reference's own
// backward-step arithmetic, sampled suffix array packed exactly like the reference's in-place
// packer.  The suffix array of a text that ends in a unique '$' is unique, so any correct sorter
// gives divsufsort's array.  Byte identity with the reference builder is a GPU test
// (tests/test_gpu_build.py).  The search path never depends on this file.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/newmap_b200.h"
#include "nm_core.cuh"

namespace {

constexpr int kB = 256;
inline unsigned grid_for(uint64_t n, int b = kB) { return (unsigned)((n + b - 1) / b); }

struct Err {
  std::string msg;
};
thread_local std::string g_build_error;

#define BUILD_CUDA(call)                                                       \
  do {                                                                         \
    cudaError_t e__ = (call);                                                  \
    if (e__ != cudaSuccess) {                                                  \
      g_build_error = std::string(#call) + ": " + cudaGetErrorString(e__);     \
      return NM_ECUDA;                                                         \
    }                                                                          \
  } while (0)

// text codes in suffix-sort order: '$' < a < c < g < t < x  (bytes 0x24 < 0x61 < 0x63 < 0x67 < 0x74 < 0x78)
__device__ __forceinline__ uint8_t code_of_raw(uint8_t ch) {  // AwFmLetter.c:23-34 sanitize, then order code
  const uint8_t c = ch | 0x20;
  return c == 'a' ? 1 : c == 'c' ? 2 : c == 'g' ? 3 : c == 't' ? 4 : 5;
}

// codes[i] for i < n0 from the FastaVector sequence buffer (incl. its '\0' separators -> x), codes[n0] = '$'
__global__ void k_codes(const uint8_t *raw, uint64_t n0, uint8_t *codes) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n0) codes[i] = code_of_raw(raw[i]);
  else if (i == n0) codes[i] = 0;
}

// first 21 characters, 3 bits each, most significant first (beyond the text: 0)
__global__ void k_init_keys(const uint8_t *codes, uint64_t n, uint64_t *keys, uint32_t *vals) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = 0;
  for (int j = 0; j < 21; ++j) {
    const uint64_t p = i + j;
    k = (k << 3) | (p < n ? codes[p] : 0);
  }
  keys[i] = k;
  vals[i] = (uint32_t)i;
}

// head[j] = j if element j starts a new key group else 0 (max-scanned into the group's first position)
__global__ void k_heads(const uint64_t *keys, uint64_t m, uint32_t *head) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  head[j] = (j == 0 || keys[j] != keys[j - 1]) ? (uint32_t)j : 0u;
}

// after the first sort: ISA[SA[j]] = first position of j's group; flag non-singletons
__global__ void k_first_ranks(const uint64_t *keys, const uint32_t *sa, const uint32_t *rank, uint64_t n,
                              uint32_t *isa, uint8_t *active) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  isa[sa[j]] = rank[j];
  const bool head = (j == 0 || keys[j] != keys[j - 1]);
  const bool next_head = (j + 1 == n || keys[j + 1] != keys[j]);
  active[j] = !(head && next_head);
}

// doubling round: key = (current rank, rank of the suffix h further on + 1)
__global__ void k_round_keys(const uint32_t *suf, const uint32_t *isa, uint64_t m, uint64_t n, uint64_t h,
                             uint64_t *keys) {
  const uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  const uint64_t s = suf[a];
  const uint64_t lo = (s + h < n) ? (uint64_t)isa[s + h] + 1 : 0;
  keys[a] = ((uint64_t)isa[s] << 32) | lo;
}

// gstart[a] = a where the HIGH half (old group) changes, else 0
__global__ void k_group_starts(const uint64_t *keys, uint64_t m, uint32_t *gstart) {
  const uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  gstart[a] = (a == 0 || (keys[a] >> 32) != (keys[a - 1] >> 32)) ? (uint32_t)a : 0u;
}

// pos[a] = SA position of active element a after this round; newhead[a] = pos if it starts a sub-group
__global__ void k_positions(const uint64_t *keys, const uint32_t *gstart, uint64_t m, uint32_t *pos, uint32_t *newhead) {
  const uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  const uint32_t p = (uint32_t)(keys[a] >> 32) + ((uint32_t)a - gstart[a]);
  pos[a] = p;
  newhead[a] = (a == 0 || keys[a] != keys[a - 1]) ? p : 0u;
}

__global__ void k_apply_round(const uint64_t *keys, const uint32_t *suf, const uint32_t *pos, const uint32_t *newrank,
                              uint64_t m, uint32_t *sa, uint32_t *isa, uint8_t *active) {
  const uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  sa[pos[a]] = suf[a];
  isa[suf[a]] = newrank[a];
  const bool head = (a == 0 || keys[a] != keys[a - 1]);
  const bool next_head = (a + 1 == m || keys[a + 1] != keys[a]);
  active[a] = !(head && next_head);
}

// ---- BWT blocks: one thread per 64 BWT positions writes its three plane words and its letter counts
// letter index of the reference: a0 c1 g2 t3 x4 $5 (AwFmLetter.c:9-20); plane codes AwFmLetter.c:37-40
__global__ void k_bwt_words(const uint8_t *codes, const uint32_t *sa, uint64_t n, uint8_t *blocks, uint32_t *wordcnt) {
  const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nwords = ((n + 255) / 256) * 4;
  if (w >= nwords) return;
  uint64_t p0 = 0, p1 = 0, p2 = 0;
  uint32_t cnt[6] = {0, 0, 0, 0, 0, 0};
  const uint8_t vec_of_letter[6] = {6, 5, 3, 1, 2, 4};
  for (int b = 0; b < 64; ++b) {
    const uint64_t j = w * 64 + b;
    if (j >= n) break;
    const uint32_t s = sa[j];
    const unsigned letter = (s == 0) ? 5u : (unsigned)(codes[s - 1] == 0 ? 5 : codes[s - 1] - 1);
    const unsigned v = vec_of_letter[letter];
    p0 |= (uint64_t)(v & 1u) << b;
    p1 |= (uint64_t)((v >> 1) & 1u) << b;
    p2 |= (uint64_t)((v >> 2) & 1u) << b;
    ++cnt[letter];
  }
  uint64_t *blk = reinterpret_cast<uint64_t *>(blocks + (w >> 2) * 160);
  blk[(w & 3)] = p0;
  blk[4 + (w & 3)] = p1;
  blk[8 + (w & 3)] = p2;
  for (int l = 0; l < 6; ++l) wordcnt[l * nwords + w] = cnt[l];
}

// per-block letter counts (sum of 4 words) as u64 for the scan
__global__ void k_block_counts(const uint32_t *wordcnt, uint64_t nblocks, uint64_t *blkcnt) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nblocks * 6) return;
  const uint64_t l = t / nblocks, b = t % nblocks;
  const uint32_t *w = wordcnt + l * nblocks * 4 + b * 4;
  blkcnt[t] = (uint64_t)w[0] + w[1] + w[2] + w[3];
}

__global__ void k_fill_counters(const uint64_t *base, uint64_t nblocks, uint8_t *blocks) {
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblocks) return;
  uint64_t *blk = reinterpret_cast<uint64_t *>(blocks + b * 160);
  for (int l = 0; l < 6; ++l) blk[12 + l] = base[(uint64_t)l * nblocks + b];
  blk[18] = 0;
  blk[19] = 0;
}

// ---- seed table, one level per launch: entry(l . P') = backward step of entry(P') by l
// (AwFmCreate.c:354-389; index = l * 4^len(P') + index(P'); ranges inclusive {start,end} as in the file)
__global__ void k_seed_level(nm::DevIndex ix, const uint64_t *prev, uint64_t prev_n, uint64_t *next) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= prev_n * 4) return;
  const unsigned l = (unsigned)(t / prev_n);
  const uint64_t parent = t % prev_n;
  const uint64_t start = prev[2 * parent], end = prev[2 * parent + 1];
  const uint64_t c = l == 0 ? ix.C[0] : l == 1 ? ix.C[1] : l == 2 ? ix.C[2] : ix.C[3];
  next[2 * t] = c + nm::occ_native(ix.native, l, start);            // C + Occ(l, start-1) inclusive
  next[2 * t + 1] = c + nm::occ_native(ix.native, l, end + 1) - 1;  // C + Occ(l, end) - 1
}

// ---- sampled suffix array, packed like AwFmSuffixArray.c:52-96 (values LSB-first, `width` bits each)
__global__ void k_pack_sa(const uint32_t *sa, uint64_t nsamples, uint32_t ratio, uint32_t width, uint64_t packed_bytes,
                          uint8_t *out) {
  // one thread per output byte: gather the bits of the (at most 2 for width >= 8, else more) samples covering it
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= packed_bytes) return;
  const uint64_t bit0 = b * 8;
  uint32_t v = 0;
  uint64_t i = bit0 / width;
  for (; i < nsamples && i * width < bit0 + 8; ++i) {
    const uint64_t val = sa[i * ratio];
    const int64_t shift = (int64_t)(i * width) - (int64_t)bit0;  // position of the value's bit 0 in this byte
    const uint64_t piece = shift >= 0 ? (val << shift) : (val >> (-shift));
    v |= (uint32_t)(piece & 0xFF);
  }
  out[b] = (uint8_t)v;
}

}  // namespace

namespace {

struct DevBuf {
  void *p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
  template <typename T> T *as() { return reinterpret_cast<T *>(p); }
};

bool write_all(FILE *f, const void *p, size_t n) { return fwrite(p, 1, n, f) == n; }

// device buffer -> file through a bounded pinned staging buffer
int dump_device(FILE *f, const void *dev, uint64_t bytes, uint8_t *stage, size_t stage_bytes) {
  uint64_t done = 0;
  while (done < bytes) {
    const size_t want = (size_t)((bytes - done < stage_bytes) ? bytes - done : stage_bytes);
    BUILD_CUDA(cudaMemcpy(stage, (const uint8_t *)dev + done, want, cudaMemcpyDeviceToHost));
    if (!write_all(f, stage, want)) return NM_EOPEN;
    done += want;
  }
  return NM_OK;
}

}  // namespace

extern "C" {

const char *nm_build_last_error(void) { return g_build_error.c_str(); }

int nm_index_build(const uint8_t *sequence, uint64_t sequence_len, const uint8_t *header, uint64_t header_len,
                   const uint64_t *metadata, uint64_t metadata_count, uint32_t sa_ratio, uint32_t seed_len,
                   int device, const char *out_path, double *seconds_sort) {
  if (!sequence || !out_path || sequence_len == 0 || sa_ratio == 0 || sa_ratio > 255 || seed_len == 0 || seed_len > 15)
    return NM_EARG;
  const uint64_t n = sequence_len + 1;  // + '$'  (AwFmCreate.c:160-179)
  if (n >= 0xFFFFFFF0ull) return NM_EUNSUPPORTED;
  BUILD_CUDA(cudaSetDevice(device));
  cudaEvent_t ev0, ev1;
  BUILD_CUDA(cudaEventCreate(&ev0));
  BUILD_CUDA(cudaEventCreate(&ev1));

  DevBuf d_codes, d_isa, d_keys[2], d_vals[2], d_aux0, d_aux1, d_aux2, d_flag, d_tmp, d_count;
  BUILD_CUDA(d_codes.alloc(n + 32));
  {
    DevBuf d_raw;
    BUILD_CUDA(d_raw.alloc(sequence_len));
    BUILD_CUDA(cudaMemcpy(d_raw.p, sequence, sequence_len, cudaMemcpyHostToDevice));
    k_codes<<<grid_for(n), kB>>>(d_raw.as<uint8_t>(), sequence_len, d_codes.as<uint8_t>());
    BUILD_CUDA(cudaDeviceSynchronize());
  }
  BUILD_CUDA(cudaEventRecord(ev0));
  BUILD_CUDA(d_isa.alloc(n * 4));
  for (int i = 0; i < 2; ++i) {
    BUILD_CUDA(d_keys[i].alloc(n * 8));
    BUILD_CUDA(d_vals[i].alloc(n * 4));
  }
  BUILD_CUDA(d_aux0.alloc(n * 4));
  BUILD_CUDA(d_aux1.alloc(n * 4));
  BUILD_CUDA(d_aux2.alloc(n * 4));
  BUILD_CUDA(d_flag.alloc(n));
  BUILD_CUDA(d_count.alloc(8));

  // temp storage large enough for every CUB call below
  size_t tmp_bytes = 0, need = 0;
  {
    cub::DoubleBuffer<uint64_t> k(d_keys[0].as<uint64_t>(), d_keys[1].as<uint64_t>());
    cub::DoubleBuffer<uint32_t> v(d_vals[0].as<uint32_t>(), d_vals[1].as<uint32_t>());
    cub::DeviceRadixSort::SortPairs(nullptr, need, k, v, (int64_t)n, 0, 64);
    tmp_bytes = need;
    cub::DeviceScan::InclusiveScan(nullptr, need, d_aux0.as<uint32_t>(), d_aux1.as<uint32_t>(), cub::Max(), (int64_t)n);
    if (need > tmp_bytes) tmp_bytes = need;
    cub::DeviceSelect::Flagged(nullptr, need, d_vals[0].as<uint32_t>(), d_flag.as<uint8_t>(), d_vals[1].as<uint32_t>(),
                               d_count.as<uint64_t>(), (int64_t)n);
    if (need > tmp_bytes) tmp_bytes = need;
    cub::DeviceScan::ExclusiveSum(nullptr, need, d_keys[0].as<uint64_t>(), d_keys[1].as<uint64_t>(), (int64_t)(n / 256 + 2));
    if (need > tmp_bytes) tmp_bytes = need;
  }
  BUILD_CUDA(d_tmp.alloc(tmp_bytes));

  // ---- suffix array by prefix doubling -----------------------------------------------------------
  uint32_t *sa = nullptr;  // final suffix array lives in one of d_vals[] after the first sort; we copy it out
  DevBuf d_sa;
  BUILD_CUDA(d_sa.alloc(n * 4));
  sa = d_sa.as<uint32_t>();
  {
    cub::DoubleBuffer<uint64_t> k(d_keys[0].as<uint64_t>(), d_keys[1].as<uint64_t>());
    cub::DoubleBuffer<uint32_t> v(d_vals[0].as<uint32_t>(), d_vals[1].as<uint32_t>());
    k_init_keys<<<grid_for(n), kB>>>(d_codes.as<uint8_t>(), n, k.Current(), v.Current());
    size_t tb = tmp_bytes;
    BUILD_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tb, k, v, (int64_t)n, 0, 63));
    BUILD_CUDA(cudaMemcpy(sa, v.Current(), n * 4, cudaMemcpyDeviceToDevice));
    k_heads<<<grid_for(n), kB>>>(k.Current(), n, d_aux0.as<uint32_t>());
    tb = tmp_bytes;
    BUILD_CUDA(cub::DeviceScan::InclusiveScan(d_tmp.p, tb, d_aux0.as<uint32_t>(), d_aux1.as<uint32_t>(), cub::Max(), (int64_t)n));
    k_first_ranks<<<grid_for(n), kB>>>(k.Current(), sa, d_aux1.as<uint32_t>(), n, d_isa.as<uint32_t>(), d_flag.as<uint8_t>());
    // active suffixes, in suffix-array order
    tb = tmp_bytes;
    BUILD_CUDA(cub::DeviceSelect::Flagged(d_tmp.p, tb, sa, d_flag.as<uint8_t>(), v.Alternate(), d_count.as<uint64_t>(), (int64_t)n));
    uint64_t m = 0;
    BUILD_CUDA(cudaMemcpy(&m, d_count.p, 8, cudaMemcpyDeviceToHost));
    v.selector ^= 1;  // Current() = active suffixes
    uint64_t h = 21;
    int rounds = 0;
    while (m > 0) {
      if (++rounds > 64) { g_build_error = "suffix sort did not converge"; return NM_EFORMAT; }
      k_round_keys<<<grid_for(m), kB>>>(v.Current(), d_isa.as<uint32_t>(), m, n, h, k.Current());
      tb = tmp_bytes;
      BUILD_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tb, k, v, (int64_t)m, 0, 64));
      k_group_starts<<<grid_for(m), kB>>>(k.Current(), m, d_aux0.as<uint32_t>());
      tb = tmp_bytes;
      BUILD_CUDA(cub::DeviceScan::InclusiveScan(d_tmp.p, tb, d_aux0.as<uint32_t>(), d_aux1.as<uint32_t>(), cub::Max(), (int64_t)m));
      k_positions<<<grid_for(m), kB>>>(k.Current(), d_aux1.as<uint32_t>(), m, d_aux0.as<uint32_t>(), d_aux2.as<uint32_t>());
      tb = tmp_bytes;
      BUILD_CUDA(cub::DeviceScan::InclusiveScan(d_tmp.p, tb, d_aux2.as<uint32_t>(), d_aux1.as<uint32_t>(), cub::Max(), (int64_t)m));
      k_apply_round<<<grid_for(m), kB>>>(k.Current(), v.Current(), d_aux0.as<uint32_t>(), d_aux1.as<uint32_t>(), m, sa,
                                         d_isa.as<uint32_t>(), d_flag.as<uint8_t>());
      tb = tmp_bytes;
      BUILD_CUDA(cub::DeviceSelect::Flagged(d_tmp.p, tb, v.Current(), d_flag.as<uint8_t>(), v.Alternate(), d_count.as<uint64_t>(), (int64_t)m));
      BUILD_CUDA(cudaMemcpy(&m, d_count.p, 8, cudaMemcpyDeviceToHost));
      v.selector ^= 1;
      h *= 2;
    }
  }
  BUILD_CUDA(cudaEventRecord(ev1));
  BUILD_CUDA(cudaEventSynchronize(ev1));
  if (seconds_sort) {
    float ms = 0;
    cudaEventElapsedTime(&ms, ev0, ev1);
    *seconds_sort = ms * 1e-3;
  }
  // the big sort buffers are no longer needed
  for (int i = 0; i < 2; ++i) { cudaFree(d_keys[i].p); d_keys[i].p = nullptr; cudaFree(d_vals[i].p); d_vals[i].p = nullptr; }
  cudaFree(d_isa.p); d_isa.p = nullptr;
  cudaFree(d_aux0.p); d_aux0.p = nullptr; cudaFree(d_aux1.p); d_aux1.p = nullptr; cudaFree(d_aux2.p); d_aux2.p = nullptr;

  // ---- BWT blocks + counters + prefix sums ---------------------------------------------------------
  const uint64_t nblocks = 1 + (n - 1) / 256, nwords = nblocks * 4;
  DevBuf d_blocks, d_wordcnt, d_blkcnt, d_base;
  BUILD_CUDA(d_blocks.alloc((nblocks + 1) * 160));
  BUILD_CUDA(cudaMemset(d_blocks.p, 0, (nblocks + 1) * 160));
  BUILD_CUDA(d_wordcnt.alloc(nwords * 6 * 4));
  BUILD_CUDA(d_blkcnt.alloc(nblocks * 6 * 8));
  BUILD_CUDA(d_base.alloc(nblocks * 6 * 8));
  k_bwt_words<<<grid_for(nwords), kB>>>(d_codes.as<uint8_t>(), sa, n, d_blocks.as<uint8_t>(), d_wordcnt.as<uint32_t>());
  k_block_counts<<<grid_for(nblocks * 6), kB>>>(d_wordcnt.as<uint32_t>(), nblocks, d_blkcnt.as<uint64_t>());
  uint64_t totals[6];
  for (int l = 0; l < 6; ++l) {
    size_t tb = tmp_bytes;
    BUILD_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tb, d_blkcnt.as<uint64_t>() + (uint64_t)l * nblocks,
                                             d_base.as<uint64_t>() + (uint64_t)l * nblocks, (int64_t)nblocks));
    uint64_t last_base = 0, last_cnt = 0;
    BUILD_CUDA(cudaMemcpy(&last_base, d_base.as<uint64_t>() + (uint64_t)l * nblocks + nblocks - 1, 8, cudaMemcpyDeviceToHost));
    BUILD_CUDA(cudaMemcpy(&last_cnt, d_blkcnt.as<uint64_t>() + (uint64_t)l * nblocks + nblocks - 1, 8, cudaMemcpyDeviceToHost));
    totals[l] = last_base + last_cnt;
  }
  k_fill_counters<<<grid_for(nblocks), kB>>>(d_base.as<uint64_t>(), nblocks, d_blocks.as<uint8_t>());
  uint64_t prefix[6];  // AwFmCreate.c:299-305
  prefix[0] = 1;
  for (int l = 1; l < 6; ++l) prefix[l] = prefix[l - 1] + totals[l - 1];
  if (prefix[5] != n) { g_build_error = "prefix sums do not add up"; return NM_EFORMAT; }
  {  // pad block so that a rank at position 256*nblocks is defined (never written to the file)
    uint64_t pad[20];
    memset(pad, 0, sizeof pad);
    for (int l = 0; l < 6; ++l) pad[12 + l] = totals[l];
    BUILD_CUDA(cudaMemcpy(d_blocks.as<uint8_t>() + nblocks * 160, pad, 160, cudaMemcpyHostToDevice));
  }

  // ---- seed table, level by level -------------------------------------------------------------------
  nm::DevIndex ix{};
  ix.native = d_blocks.as<uint8_t>();
  for (int l = 0; l < 6; ++l) ix.C[l] = prefix[l];
  ix.bwt_len = n;
  ix.seed_len = seed_len;
  const uint64_t entries = 1ull << (2 * seed_len);
  DevBuf d_seedA, d_seedB;
  BUILD_CUDA(d_seedA.alloc(entries * 16));
  BUILD_CUDA(d_seedB.alloc(entries * 16 / 4 + 64));
  uint64_t *cur = (seed_len % 2) ? d_seedA.as<uint64_t>() : d_seedB.as<uint64_t>();   // so that the last level lands in A
  uint64_t *nxt = (seed_len % 2) ? d_seedB.as<uint64_t>() : d_seedA.as<uint64_t>();
  {
    uint64_t lvl1[8];
    for (int l = 0; l < 4; ++l) { lvl1[2 * l] = prefix[l]; lvl1[2 * l + 1] = prefix[l + 1] - 1; }
    BUILD_CUDA(cudaMemcpy(cur, lvl1, sizeof lvl1, cudaMemcpyHostToDevice));
  }
  uint64_t cur_n = 4;
  for (uint32_t len = 1; len < seed_len; ++len) {
    k_seed_level<<<grid_for(cur_n * 4), kB>>>(ix, cur, cur_n, nxt);
    uint64_t *t = cur; cur = nxt; nxt = t;
    cur_n *= 4;
  }
  BUILD_CUDA(cudaDeviceSynchronize());
  const uint64_t *seed_final = cur;

  // ---- sampled suffix array ---------------------------------------------------------------------------
  const uint32_t width = 64 - (uint32_t)__builtin_clzll(n - 1);        // AwFmSuffixArray.c:12-19
  const uint64_t nsamples = (n + sa_ratio - 1) / sa_ratio;
  const uint64_t total_bits = nsamples * width;
  const uint64_t packed_bytes = (total_bits + 7) / 8;
  const uint64_t sa_bytes = packed_bytes + 8;                          // + AW_FM_SUFFIX_ARRAY_END_PADDING_BYTES
  DevBuf d_packed;
  BUILD_CUDA(d_packed.alloc(sa_bytes));
  k_pack_sa<<<grid_for(packed_bytes), kB>>>(sa, nsamples, sa_ratio, width, packed_bytes, d_packed.as<uint8_t>());
  {
    // the reference packs IN PLACE over its u64 suffix array and then realloc()s: the 8 padding bytes are
    // whatever u64 suffix-array bytes lay at those offsets (never reached by the packer)
    uint8_t pad[8];
    for (int b = 0; b < 8; ++b) {
      const uint64_t off = packed_bytes + b, idx = off / 8;
      uint32_t v = 0;
      if (idx < n) BUILD_CUDA(cudaMemcpy(&v, sa + idx, 4, cudaMemcpyDeviceToHost));
      const uint64_t v64 = v;
      pad[b] = (uint8_t)(v64 >> (8 * (off % 8)));
    }
    BUILD_CUDA(cudaMemcpy(d_packed.as<uint8_t>() + packed_bytes, pad, 8, cudaMemcpyHostToDevice));
  }
  BUILD_CUDA(cudaDeviceSynchronize());
  BUILD_CUDA(cudaGetLastError());

  // ---- write the file (AwFmFile.c:20-171) ---------------------------------------------------------------
  FILE *f = fopen(out_path, "w+b");
  if (!f) return NM_EOPEN;
  const size_t stage_bytes = 64u << 20;
  uint8_t *stage = nullptr;
  if (cudaHostAlloc((void **)&stage, stage_bytes, cudaHostAllocDefault) != cudaSuccess) { fclose(f); return NM_EALLOC; }
  int rc = NM_OK;
  const uint32_t version = 8, flags = 1;
  const uint8_t cfg[4] = {(uint8_t)sa_ratio, (uint8_t)seed_len, 2, 0};
  bool ok = write_all(f, "AwFmIndex\n", 10) && write_all(f, &version, 4) && write_all(f, &flags, 4) &&
            write_all(f, cfg, 4) && write_all(f, &n, 8);
  if (ok) rc = dump_device(f, d_blocks.p, nblocks * 160, stage, stage_bytes);
  ok = ok && rc == NM_OK && write_all(f, prefix, 48);
  if (ok) rc = dump_device(f, seed_final, entries * 16, stage, stage_bytes);
  if (ok && rc == NM_OK) rc = dump_device(f, d_packed.p, sa_bytes, stage, stage_bytes);
  ok = ok && rc == NM_OK && write_all(f, &header_len, 8) && write_all(f, &metadata_count, 8) &&
       (header_len == 0 || write_all(f, header, header_len)) &&
       (metadata_count == 0 || write_all(f, metadata, metadata_count * 16));
  cudaFreeHost(stage);
  if (fclose(f) != 0) ok = false;
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  if (rc != NM_OK) return rc;
  return ok ? NM_OK : NM_EOPEN;
}

}  // extern "C"
