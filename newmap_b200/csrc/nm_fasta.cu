// nm_fasta.cu -- FASTA ingest on the device (C-ABI nm_fasta_*, include/newmap_b200.h).
//
// What the reference does on the host, line by line in Python (newmap/fasta.py:20-190: find header lines,
// right-strip and concatenate the other lines, cut windows), is done here at memory speed: the raw text goes to
// HBM through pinned staging while it is being read, one pass locates the header lines, a stream compaction
// (per-tile counts -> exclusive scan -> scatter) removes newlines and header lines, and the records come out as
// contiguous runs of bases that the search workspaces read in place.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/newmap_b200.h"
#include "nm_fasta.cuh"

namespace {

constexpr uint32_t kFaTile = 4096;          // bytes per compaction tile (one block of 256 threads x 16 bytes)
constexpr uint32_t kFaThreads = 256;
constexpr uint32_t kFaPerThread = kFaTile / kFaThreads;
constexpr uint64_t kMaxHeaders = 1ull << 22;
constexpr uint64_t kStageBytes = 2ull << 20;   // pinned staging chunk (two per reader thread; pinning costs time per byte)

int fa_fail(cudaError_t e, const char *what) {
  nm_internal_set_cuda_error(what, cudaGetErrorString(e));
  return NM_ECUDA;
}
#define FA_CUDA(call)                                   \
  do {                                                  \
    cudaError_t e__ = (call);                           \
    if (e__ != cudaSuccess) return fa_fail(e__, #call); \
  } while (0)

// a header line starts where a '>' or ';' follows a newline (or opens the text)
__global__ void __launch_bounds__(256) k_fa_headers(const uint8_t *raw, uint64_t n, unsigned long long *count,
                                                    uint64_t *list, uint64_t cap) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint8_t c = raw[i];
    if ((c == '>' || c == ';') && (i == 0 || raw[i - 1] == '\n')) {
      const unsigned long long at = atomicAdd(count, 1ull);
      if (at < cap) list[at] = i;
    }
  }
}

// end of each header line: position of its '\n' (or n)
__global__ void __launch_bounds__(256) k_fa_header_ends(const uint8_t *raw, uint64_t n, const uint64_t *starts,
                                                        uint64_t nh, uint64_t *ends) {
  const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= nh) return;
  uint64_t i = starts[h];
  while (i < n && raw[i] != '\n') ++i;
  ends[h] = i;
}

// index of the last header that starts at or before i, or nh if none
__device__ __forceinline__ uint64_t header_at_or_before(const uint64_t *starts, uint64_t nh, uint64_t i) {
  uint64_t lo = 0, hi = nh;  // first header with start > i
  while (lo < hi) {
    const uint64_t mid = (lo + hi) >> 1;
    if (starts[mid] <= i) lo = mid + 1; else hi = mid;
  }
  return lo ? lo - 1 : nh;
}

struct TileHeaders {  // the headers that can overlap one tile: [first, last)
  uint64_t first, last;
};
__device__ __forceinline__ TileHeaders tile_headers(const uint64_t *starts, const uint64_t *ends, uint64_t nh,
                                                    uint64_t t0, uint64_t t1) {
  TileHeaders th{0, 0};
  if (nh == 0) return th;
  const uint64_t h = header_at_or_before(starts, nh, t0);  // may reach into the tile from the left
  uint64_t first = (h == nh) ? 0 : (ends[h] >= t0 ? h : h + 1);
  uint64_t last = first;
  while (last < nh && starts[last] < t1) ++last;
  th.first = first;
  th.last = last;
  return th;
}
__device__ __forceinline__ bool in_header(const uint64_t *starts, const uint64_t *ends, const TileHeaders &th, uint64_t i) {
  for (uint64_t h = th.first; h < th.last; ++h)
    if (i >= starts[h] && i <= ends[h]) return true;
  return false;
}
__device__ __forceinline__ bool is_odd_space(uint8_t c) {
  return c == ' ' || c == '\t' || c == '\r' || c == 0x0b || c == 0x0c;
}

// pass 1 of the compaction: kept bytes per tile; pass 2: scatter.  A byte is kept unless it is '\n' or part of a header line.
template <bool SCATTER>
__global__ void __launch_bounds__(kFaThreads) k_fa_compact(const uint8_t *raw, uint64_t n, const uint64_t *starts,
                                                           const uint64_t *ends, uint64_t nh, unsigned long long *tile_count,
                                                           const unsigned long long *tile_base, uint8_t *seq,
                                                           unsigned int *odd) {
  __shared__ TileHeaders s_th;
  __shared__ uint32_t s_warp[kFaThreads / 32];
  const uint64_t t0 = (uint64_t)blockIdx.x * kFaTile;
  const uint64_t t1 = t0 + kFaTile < n ? t0 + kFaTile : n;
  if (threadIdx.x == 0) s_th = tile_headers(starts, ends, nh, t0, t1);
  __syncthreads();
  const TileHeaders th = s_th;
  const uint64_t b0 = t0 + (uint64_t)threadIdx.x * kFaPerThread;
  uint8_t c[kFaPerThread];
  uint32_t keep = 0;  // bit j: byte b0 + j is kept
  bool any_odd = false;
#pragma unroll
  for (uint32_t j = 0; j < kFaPerThread; ++j) {
    const uint64_t i = b0 + j;
    c[j] = i < t1 ? raw[i] : (uint8_t)'\n';
    if (i < t1 && c[j] != '\n' && !(th.first < th.last && in_header(starts, ends, th, i))) {
      keep |= 1u << j;
      any_odd = any_odd || is_odd_space(c[j]);
    }
  }
  const uint32_t mine = __popc(keep);
  // block-wide exclusive scan of `mine`
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((threadIdx.x & 31) >= o) incl += v;
  }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();
  uint32_t before = incl - mine, total = 0;
#pragma unroll
  for (uint32_t w = 0; w < kFaThreads / 32; ++w) {
    if (w < (threadIdx.x >> 5)) before += s_warp[w];
    total += s_warp[w];
  }
  if (!SCATTER) {
    if (threadIdx.x == 0) tile_count[blockIdx.x] = total;
    if (any_odd) atomicOr(odd, 1u);
  } else {
    uint8_t *dst = seq + tile_base[blockIdx.x] + before;
#pragma unroll
    for (uint32_t j = 0; j < kFaPerThread; ++j)
      if (keep & (1u << j)) *dst++ = c[j];
  }
}

// bases kept before raw position x (x is a line start): tile_base of its tile + the kept bytes of the tile before x
__global__ void __launch_bounds__(256) k_fa_seqpos(const uint8_t *raw, uint64_t n, const uint64_t *starts,
                                                   const uint64_t *ends, uint64_t nh, const unsigned long long *tile_base,
                                                   uint64_t *seqpos) {
  const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= nh) return;
  const uint64_t x = starts[h];
  const uint64_t t0 = x / kFaTile * kFaTile;
  const uint64_t t1 = t0 + kFaTile < n ? t0 + kFaTile : n;
  const TileHeaders th = tile_headers(starts, ends, nh, t0, t1);
  uint64_t kept = 0;
  for (uint64_t i = t0; i < x; ++i)
    if (raw[i] != '\n' && !in_header(starts, ends, th, i)) ++kept;
  seqpos[h] = tile_base[x / kFaTile] + kept;
}

// header lines packed one after the other for a single device->host copy
__global__ void __launch_bounds__(256) k_fa_gather_headers(const uint8_t *raw, const uint64_t *starts, const uint64_t *ends,
                                                           const uint64_t *packed_off, uint64_t nh, uint8_t *out) {
  const uint64_t h = blockIdx.x;
  if (h >= nh) return;
  const uint64_t s = starts[h], len = ends[h] - s;
  for (uint64_t j = threadIdx.x; j < len; j += blockDim.x) out[packed_off[h] + j] = raw[s + j];
}

}  // namespace

struct nm_fasta {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint8_t *d_raw = nullptr;
  uint64_t raw_cap = 0, raw_len = 0;
  uint8_t *d_seq = nullptr;
  uint64_t seq_len = 0;
  bool finished = false;
  uint32_t odd = 0;
  float ms_parse = 0;
  std::vector<nm_fasta_record> recs;
  std::string headers;
  struct Stage {
    uint8_t *h = nullptr;
    cudaEvent_t done = nullptr;
    bool busy = false;
  } stage[2];
  int next_stage = 0;
};

const uint8_t *nm_fasta_device_sequence(const nm_fasta *fa, uint64_t *length, int *device) {
  if (!fa || !fa->finished) return nullptr;
  if (length) *length = fa->seq_len;
  if (device) *device = fa->device;
  return fa->d_seq;
}

namespace {

int fa_reserve(nm_fasta *fa, uint64_t need) {
  if (need <= fa->raw_cap) return NM_OK;
  uint64_t cap = fa->raw_cap ? fa->raw_cap * 2 : (64ull << 20);
  if (cap < need) cap = need;
  uint8_t *p = nullptr;
  if (cudaMalloc((void **)&p, cap + 256) != cudaSuccess) {
    cudaGetLastError();
    return NM_EALLOC;
  }
  if (fa->raw_len) FA_CUDA(cudaMemcpyAsync(p, fa->d_raw, fa->raw_len, cudaMemcpyDeviceToDevice, fa->stream));
  FA_CUDA(cudaStreamSynchronize(fa->stream));
  cudaFree(fa->d_raw);
  fa->d_raw = p;
  fa->raw_cap = cap;
  return NM_OK;
}

template <typename T>
struct DevBuf {
  T *p = nullptr;
  ~DevBuf() { cudaFree(p); }
  cudaError_t alloc(uint64_t n) { return cudaMalloc((void **)&p, (n ? n : 1) * sizeof(T)); }
};

}  // namespace

extern "C" {

int nm_fasta_open(int device, uint64_t size_hint, nm_fasta **out) {
  if (!out) return NM_EARG;
  *out = nullptr;
  int ndev = 0;
  FA_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return NM_EARG;
  FA_CUDA(cudaSetDevice(device));
  nm_fasta *fa = new (std::nothrow) nm_fasta();
  if (!fa) return NM_EALLOC;
  fa->device = device;
  if (cudaStreamCreateWithFlags(&fa->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete fa;
    return NM_ECUDA;
  }
  if (size_hint) {
    const int rc = fa_reserve(fa, size_hint);
    if (rc != NM_OK) {
      nm_fasta_close(fa);
      return rc;
    }
  }
  *out = fa;
  return NM_OK;
}

void nm_fasta_close(nm_fasta *fa) {
  if (!fa) return;
  cudaSetDevice(fa->device);
  if (fa->stream) cudaStreamSynchronize(fa->stream);
  for (auto &s : fa->stage) {
    if (s.h) cudaFreeHost(s.h);
    if (s.done) cudaEventDestroy(s.done);
  }
  cudaFree(fa->d_raw);
  cudaFree(fa->d_seq);
  if (fa->stream) cudaStreamDestroy(fa->stream);
  delete fa;
}

int nm_fasta_feed(nm_fasta *fa, const void *data, uint64_t n) {
  if (!fa || fa->finished || (n && !data)) return NM_EARG;
  FA_CUDA(cudaSetDevice(fa->device));
  int rc = fa_reserve(fa, fa->raw_len + n);
  if (rc != NM_OK) return rc;
  const uint8_t *src = (const uint8_t *)data;
  while (n) {
    nm_fasta::Stage &s = fa->stage[fa->next_stage];
    fa->next_stage ^= 1;
    if (!s.h) {
      FA_CUDA(cudaHostAlloc((void **)&s.h, kStageBytes, cudaHostAllocDefault));
      FA_CUDA(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    }
    if (s.busy) FA_CUDA(cudaEventSynchronize(s.done));
    const uint64_t m = n < kStageBytes ? n : kStageBytes;
    memcpy(s.h, src, m);
    FA_CUDA(cudaMemcpyAsync(fa->d_raw + fa->raw_len, s.h, m, cudaMemcpyHostToDevice, fa->stream));
    FA_CUDA(cudaEventRecord(s.done, fa->stream));
    s.busy = true;
    fa->raw_len += m;
    src += m;
    n -= m;
  }
  return NM_OK;
}

int nm_fasta_feed_file(nm_fasta *fa, const char *path, uint32_t threads) {
  return nm_fasta_feed_file_range(fa, path, 0, ~0ull, threads);
}

int nm_fasta_feed_file_range(nm_fasta *fa, const char *path, uint64_t offset, uint64_t length, uint32_t threads) {
  if (!fa || fa->finished || !path) return NM_EARG;
  FA_CUDA(cudaSetDevice(fa->device));
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return NM_EOPEN;
  struct stat st;
  if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {
    close(fd);
    return NM_EOPEN;
  }
  const uint64_t file_size = (uint64_t)st.st_size;
  if (offset > file_size) offset = file_size;
  const uint64_t size = length < file_size - offset ? length : file_size - offset;   // bytes [offset, offset + size)
  int rc = fa_reserve(fa, fa->raw_len + size);
  if (rc != NM_OK) {
    close(fd);
    return rc;
  }
  FA_CUDA(cudaStreamSynchronize(fa->stream));
  if (threads == 0) threads = 4;
  const uint64_t nchunks = (size + kStageBytes - 1) / kStageBytes;
  if (threads > nchunks) threads = (uint32_t)(nchunks ? nchunks : 1);
  std::atomic<uint64_t> next{0};
  std::atomic<int> status{NM_OK};
  uint8_t *base = fa->d_raw + fa->raw_len;
  const int device = fa->device;
  auto worker = [&]() {
    // two pinned buffers per reader: the file read of one chunk overlaps the host->device copy of the previous one
    uint8_t *h[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t s = nullptr;
    bool busy[2] = {false, false};
    bool ok = cudaSetDevice(device) == cudaSuccess && cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess;
    for (int b = 0; ok && b < 2; ++b)
      ok = cudaHostAlloc((void **)&h[b], kStageBytes, cudaHostAllocDefault) == cudaSuccess &&
           cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming) == cudaSuccess;
    int b = 0;
    while (ok && status.load() == NM_OK) {
      const uint64_t c = next.fetch_add(1);
      if (c >= nchunks) break;
      const uint64_t off = c * kStageBytes, len = size - off < kStageBytes ? size - off : kStageBytes;
      if (busy[b] && cudaEventSynchronize(ev[b]) != cudaSuccess) { ok = false; break; }
      uint64_t got = 0;
      while (got < len) {
        const ssize_t r = pread(fd, h[b] + got, len - got, (off_t)(offset + off + got));
        if (r <= 0) break;
        got += (uint64_t)r;
      }
      if (got != len) { status.store(NM_ESHORT); break; }
      if (cudaMemcpyAsync(base + off, h[b], len, cudaMemcpyHostToDevice, s) != cudaSuccess ||
          cudaEventRecord(ev[b], s) != cudaSuccess) { ok = false; break; }
      busy[b] = true;
      b ^= 1;
    }
    if (s) cudaStreamSynchronize(s);
    if (!ok) status.store(NM_ECUDA);
    for (int k = 0; k < 2; ++k) {
      if (h[k]) cudaFreeHost(h[k]);
      if (ev[k]) cudaEventDestroy(ev[k]);
    }
    if (s) cudaStreamDestroy(s);
  };
  std::vector<std::thread> pool;
  for (uint32_t t = 1; t < threads; ++t) pool.emplace_back(worker);
  worker();
  for (auto &t : pool) t.join();
  close(fd);
  if (status.load() != NM_OK) return status.load();
  fa->raw_len += size;
  return NM_OK;
}

}  // extern "C"

extern "C" {

int nm_fasta_finish(nm_fasta *fa, nm_fasta_info *info) {
  if (!fa || fa->finished) return NM_EARG;
  FA_CUDA(cudaSetDevice(fa->device));
  cudaStream_t st = fa->stream;
  for (auto &s : fa->stage)
    if (s.busy) FA_CUDA(cudaEventSynchronize(s.done));
  FA_CUDA(cudaStreamSynchronize(st));
  const uint64_t n = fa->raw_len;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  FA_CUDA(cudaEventCreate(&e0));
  FA_CUDA(cudaEventCreate(&e1));
  struct EvGuard {
    cudaEvent_t a, b;
    ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); }
  } evg{e0, e1};
  FA_CUDA(cudaEventRecord(e0, st));
  fa->recs.clear();
  fa->headers.clear();
  fa->seq_len = 0;
  if (n) {
    // ---- header lines
    const uint64_t cap = std::min<uint64_t>(kMaxHeaders, n / 2 + 1);
    DevBuf<unsigned long long> d_count;
    DevBuf<uint64_t> d_list;
    DevBuf<unsigned int> d_odd;
    if (d_count.alloc(1) != cudaSuccess || d_list.alloc(cap) != cudaSuccess || d_odd.alloc(1) != cudaSuccess) return NM_EALLOC;
    FA_CUDA(cudaMemsetAsync(d_count.p, 0, 8, st));
    FA_CUDA(cudaMemsetAsync(d_odd.p, 0, 4, st));
    k_fa_headers<<<148 * 16, 256, 0, st>>>(fa->d_raw, n, d_count.p, d_list.p, cap);
    unsigned long long nh = 0;
    FA_CUDA(cudaMemcpyAsync(&nh, d_count.p, 8, cudaMemcpyDeviceToHost, st));
    FA_CUDA(cudaStreamSynchronize(st));
    if (nh > cap) return NM_EUNSUPPORTED;
    std::vector<uint64_t> starts(nh), ends(nh), seqpos(nh);
    DevBuf<uint64_t> d_ends, d_seqpos, d_poff;
    if (d_ends.alloc(nh) != cudaSuccess || d_seqpos.alloc(nh) != cudaSuccess || d_poff.alloc(nh) != cudaSuccess) return NM_EALLOC;
    if (nh) {
      FA_CUDA(cudaMemcpyAsync(starts.data(), d_list.p, nh * 8, cudaMemcpyDeviceToHost, st));
      FA_CUDA(cudaStreamSynchronize(st));
      std::sort(starts.begin(), starts.end());
      FA_CUDA(cudaMemcpyAsync(d_list.p, starts.data(), nh * 8, cudaMemcpyHostToDevice, st));
      k_fa_header_ends<<<(unsigned)((nh + 255) / 256), 256, 0, st>>>(fa->d_raw, n, d_list.p, nh, d_ends.p);
      FA_CUDA(cudaMemcpyAsync(ends.data(), d_ends.p, nh * 8, cudaMemcpyDeviceToHost, st));
    }
    // ---- compaction: counts per tile, exclusive scan, scatter
    const uint64_t ntiles = (n + kFaTile - 1) / kFaTile;
    DevBuf<unsigned long long> d_tcount, d_tbase;
    if (d_tcount.alloc(ntiles + 1) != cudaSuccess || d_tbase.alloc(ntiles + 1) != cudaSuccess) return NM_EALLOC;
    FA_CUDA(cudaMemsetAsync(d_tcount.p + ntiles, 0, 8, st));
    k_fa_compact<false><<<(unsigned)ntiles, kFaThreads, 0, st>>>(fa->d_raw, n, d_list.p, d_ends.p, nh, d_tcount.p, nullptr,
                                                                nullptr, d_odd.p);
    size_t tmp_bytes = 0;
    FA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_tcount.p, d_tbase.p, (int64_t)(ntiles + 1), st));
    DevBuf<uint8_t> d_tmp;
    if (d_tmp.alloc(tmp_bytes) != cudaSuccess) return NM_EALLOC;
    FA_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp_bytes, d_tcount.p, d_tbase.p, (int64_t)(ntiles + 1), st));
    unsigned long long total = 0;
    FA_CUDA(cudaMemcpyAsync(&total, d_tbase.p + ntiles, 8, cudaMemcpyDeviceToHost, st));
    FA_CUDA(cudaMemcpyAsync(&fa->odd, d_odd.p, 4, cudaMemcpyDeviceToHost, st));
    FA_CUDA(cudaStreamSynchronize(st));
    cudaFree(fa->d_seq);
    fa->d_seq = nullptr;
    if (cudaMalloc((void **)&fa->d_seq, total + 256) != cudaSuccess) {
      cudaGetLastError();
      return NM_EALLOC;
    }
    FA_CUDA(cudaMemsetAsync(fa->d_seq + total, 0, 256, st));
    k_fa_compact<true><<<(unsigned)ntiles, kFaThreads, 0, st>>>(fa->d_raw, n, d_list.p, d_ends.p, nh, nullptr, d_tbase.p,
                                                               fa->d_seq, nullptr);
    fa->seq_len = total;
    // ---- records: bases kept before each header line; header text packed for one copy
    if (nh) {
      k_fa_seqpos<<<(unsigned)((nh + 255) / 256), 256, 0, st>>>(fa->d_raw, n, d_list.p, d_ends.p, nh, d_tbase.p, d_seqpos.p);
      FA_CUDA(cudaMemcpyAsync(seqpos.data(), d_seqpos.p, nh * 8, cudaMemcpyDeviceToHost, st));
      FA_CUDA(cudaStreamSynchronize(st));
      std::vector<uint64_t> poff(nh);
      uint64_t packed = 0;
      for (uint64_t h = 0; h < nh; ++h) {
        poff[h] = packed;
        packed += ends[h] - starts[h];
      }
      DevBuf<uint8_t> d_hdr;
      if (d_hdr.alloc(packed) != cudaSuccess) return NM_EALLOC;
      FA_CUDA(cudaMemcpyAsync(d_poff.p, poff.data(), nh * 8, cudaMemcpyHostToDevice, st));
      k_fa_gather_headers<<<(unsigned)nh, 256, 0, st>>>(fa->d_raw, d_list.p, d_ends.p, d_poff.p, nh, d_hdr.p);
      fa->headers.resize(packed);
      if (packed) FA_CUDA(cudaMemcpyAsync(&fa->headers[0], d_hdr.p, packed, cudaMemcpyDeviceToHost, st));
      FA_CUDA(cudaStreamSynchronize(st));
      if (seqpos[0] > 0) fa->recs.push_back(nm_fasta_record{0, seqpos[0], 0, 0});  // bases before the first header line
      for (uint64_t h = 0; h < nh; ++h) {
        const uint64_t end = h + 1 < nh ? seqpos[h + 1] : (uint64_t)total;
        fa->recs.push_back(nm_fasta_record{seqpos[h], end - seqpos[h], poff[h], ends[h] - starts[h]});
      }
    } else if (total) {
      fa->recs.push_back(nm_fasta_record{0, (uint64_t)total, 0, 0});
    }
    FA_CUDA(cudaGetLastError());
  }
  FA_CUDA(cudaEventRecord(e1, st));
  FA_CUDA(cudaStreamSynchronize(st));
  FA_CUDA(cudaEventElapsedTime(&fa->ms_parse, e0, e1));
  cudaFree(fa->d_raw);  // the text itself is no longer needed
  fa->d_raw = nullptr;
  fa->raw_cap = 0;
  fa->finished = true;
  if (info) {
    memset(info, 0, sizeof *info);
    info->raw_bytes = n;
    info->sequence_bytes = fa->seq_len;
    info->records = fa->recs.size();
    info->odd_whitespace = fa->odd ? 1u : 0u;
    info->ms_parse = fa->ms_parse;
  }
  return NM_OK;
}

int nm_fasta_records(const nm_fasta *fa, nm_fasta_record *out, uint64_t capacity) {
  if (!fa || !fa->finished || (!out && capacity)) return NM_EARG;
  const uint64_t m = std::min<uint64_t>(capacity, fa->recs.size());
  if (m) memcpy(out, fa->recs.data(), m * sizeof(nm_fasta_record));
  return NM_OK;
}

int nm_fasta_headers(const nm_fasta *fa, char *out, uint64_t capacity, uint64_t *needed) {
  if (!fa || !fa->finished) return NM_EARG;
  if (needed) *needed = fa->headers.size();
  if (out && capacity >= fa->headers.size() && !fa->headers.empty()) memcpy(out, fa->headers.data(), fa->headers.size());
  return NM_OK;
}

int nm_fasta_read(const nm_fasta *fa, uint64_t seq_offset, uint64_t n, uint8_t *out) {
  if (!fa || !fa->finished || (n && !out) || seq_offset > fa->seq_len || n > fa->seq_len - seq_offset) return NM_EARG;
  FA_CUDA(cudaSetDevice(fa->device));
  if (n) FA_CUDA(cudaMemcpy(out, fa->d_seq + seq_offset, n, cudaMemcpyDeviceToHost));
  return NM_OK;
}

}  // extern "C"
