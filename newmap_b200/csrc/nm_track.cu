// nm_track.cu -- `newmap track` on the device (C-ABI nm_track_*, include/newmap_b200.h).
//
// newmap/track.py:22-62 marks +1 where 0 < unique[j] <= k, -1 k places later, and takes the cumulative sum: the
// value at i is the number of marked j in (i-k, i].  Here: an inclusive prefix sum of the marks (cub::DeviceScan)
// and one subtraction per position; streaming, HBM-bound.  The BED runs (track.py:65-97) are the positions where
// (count > 0) changes (cub::DeviceSelect); the WIG text (track.py:100-121) is a length scan plus a scatter from
// the k+1 strings a count can print as.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/newmap_b200.h"
#include "nm_fasta.cuh"

namespace {

int tr_fail(cudaError_t e, const char *what) {
  nm_internal_set_cuda_error(what, cudaGetErrorString(e));
  return NM_ECUDA;
}
#define TR_CUDA(call)                                   \
  do {                                                  \
    cudaError_t e__ = (call);                           \
    if (e__ != cudaSuccess) return tr_fail(e__, #call); \
  } while (0)

template <typename T>
__global__ void __launch_bounds__(256) k_tr_marks(const T *unique, uint64_t n, uint32_t k, uint32_t *marks) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t u = (uint32_t)unique[i];
    marks[i] = (u != 0u && u <= k) ? 1u : 0u;   // track.py:45-48
  }
}

// counts[i] = prefix[i] - prefix[i-k]  (prefix inclusive; track.py:51-55)
__global__ void __launch_bounds__(256) k_tr_window(const uint32_t *prefix, uint64_t n, uint32_t k, uint32_t *counts) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) counts[i] = prefix[i] - (i >= k ? prefix[i - k] : 0u);
}

// flag the first position of every maximal run of equal (count > 0)  (track.py:70-75)
__global__ void __launch_bounds__(256) k_tr_change(const uint32_t *counts, uint64_t n, uint8_t *flag) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (i == 0 || (counts[i] > 0u) != (counts[i - 1] > 0u)) ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_tr_run_values(const uint32_t *counts, const uint64_t *starts, uint64_t nruns,
                                                       uint8_t *values) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < nruns) values[r] = counts[starts[r]] > 0u ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_tr_text_len(const uint32_t *counts, uint64_t m, const uint32_t *offsets,
                                                     uint32_t *len) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) {
    const uint32_t c = counts[i];
    len[i] = offsets[c + 1] - offsets[c] + 1u;  // + '\n'
  }
}

__global__ void __launch_bounds__(256) k_tr_text_put(const uint32_t *counts, uint64_t m, const uint32_t *offsets,
                                                     const char *strings, const unsigned long long *at, char *out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t c = counts[i];
  const uint32_t a = offsets[c], b = offsets[c + 1];
  char *dst = out + at[i];
  for (uint32_t j = a; j < b; ++j) *dst++ = strings[j];
  *dst = '\n';
}

__device__ __forceinline__ uint32_t dec_digits(uint64_t v) {
  uint32_t d = 1;
  while (v >= 10) { v /= 10; ++d; }
  return d;
}
__device__ __forceinline__ char *put_dec(char *dst, uint64_t v, uint32_t digits) {
  for (uint32_t j = digits; j-- > 0;) { dst[j] = (char)('0' + v % 10); v /= 10; }
  return dst + digits;
}
// one BED line per run: "<chr>\t<start>\t<end>\tk<k>\t<0|1>\t.\n"  (track.py:86-97); fixed = bytes of a line without the two numbers
__global__ void __launch_bounds__(256) k_tr_bed_len(const uint64_t *starts, uint64_t r0, uint64_t m, uint64_t nruns,
                                                    uint64_t n, uint32_t fixed, uint32_t *len) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint64_t r = r0 + i, s = starts[r], e = r + 1 < nruns ? starts[r + 1] : n;
  len[i] = fixed + dec_digits(s) + dec_digits(e);
}
__global__ void __launch_bounds__(256) k_tr_bed_put(const uint64_t *starts, const uint32_t *counts, uint64_t r0, uint64_t m,
                                                    uint64_t nruns, uint64_t n, const char *chr, uint32_t chr_len,
                                                    const char *ktag, uint32_t ktag_len, const unsigned long long *at,
                                                    char *out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint64_t r = r0 + i, s = starts[r], e = r + 1 < nruns ? starts[r + 1] : n;
  char *dst = out + at[i];
  for (uint32_t j = 0; j < chr_len; ++j) *dst++ = chr[j];
  *dst++ = '\t';
  dst = put_dec(dst, s, dec_digits(s));
  *dst++ = '\t';
  dst = put_dec(dst, e, dec_digits(e));
  *dst++ = '\t';
  for (uint32_t j = 0; j < ktag_len; ++j) *dst++ = ktag[j];
  *dst++ = '\t';
  *dst++ = counts[s] > 0u ? '1' : '0';
  *dst++ = '\t';
  *dst++ = '.';
  *dst = '\n';
}

}  // namespace

struct nm_track {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint64_t cap = 0;       // positions the buffers hold
  uint64_t n = 0;
  uint32_t k = 0;
  void *d_unique = nullptr;
  uint32_t *d_a = nullptr, *d_counts = nullptr;  // marks / prefix / text lengths; window counts
  void *d_tmp = nullptr;
  size_t tmp_bytes = 0;
  uint64_t *d_runs = nullptr;  // starts of the single-read runs of the last nm_track_runs
  uint64_t nruns = 0;
  uint8_t *h_stage = nullptr;  // pinned staging for uploads and downloads
  uint64_t stage_bytes = 0;
};

namespace {

int tr_reserve(nm_track *tr, uint64_t n) {
  if (n <= tr->cap) return NM_OK;
  cudaFree(tr->d_unique);
  cudaFree(tr->d_a);
  cudaFree(tr->d_counts);
  tr->d_unique = nullptr;
  tr->d_a = tr->d_counts = nullptr;
  tr->cap = 0;
  const uint64_t cap = n + n / 8 + 1024;
  if (cudaMalloc(&tr->d_unique, cap * 4) != cudaSuccess || cudaMalloc((void **)&tr->d_a, cap * 4) != cudaSuccess ||
      cudaMalloc((void **)&tr->d_counts, cap * 4) != cudaSuccess) {
    cudaGetLastError();
    return NM_EALLOC;
  }
  tr->cap = cap;
  return NM_OK;
}
int tr_tmp(nm_track *tr, size_t bytes) {
  if (bytes <= tr->tmp_bytes) return NM_OK;
  cudaFree(tr->d_tmp);
  tr->d_tmp = nullptr;
  tr->tmp_bytes = 0;
  if (cudaMalloc(&tr->d_tmp, bytes) != cudaSuccess) {
    cudaGetLastError();
    return NM_EALLOC;
  }
  tr->tmp_bytes = bytes;
  return NM_OK;
}
int tr_stage(nm_track *tr, uint64_t bytes) {
  if (bytes <= tr->stage_bytes) return NM_OK;
  if (tr->h_stage) cudaFreeHost(tr->h_stage);
  tr->h_stage = nullptr;
  tr->stage_bytes = 0;
  if (cudaHostAlloc((void **)&tr->h_stage, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return NM_EALLOC;
  }
  tr->stage_bytes = bytes;
  return NM_OK;
}
inline unsigned blocks(uint64_t n) { return (unsigned)((n + 255) / 256); }
constexpr uint64_t kTrStage = 64ull << 20;

}  // namespace

extern "C" {

int nm_track_open(int device, nm_track **out) {
  if (!out) return NM_EARG;
  *out = nullptr;
  int ndev = 0;
  TR_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return NM_EARG;
  TR_CUDA(cudaSetDevice(device));
  nm_track *tr = new (std::nothrow) nm_track();
  if (!tr) return NM_EALLOC;
  tr->device = device;
  if (cudaStreamCreateWithFlags(&tr->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete tr;
    return NM_ECUDA;
  }
  *out = tr;
  return NM_OK;
}

void nm_track_close(nm_track *tr) {
  if (!tr) return;
  cudaSetDevice(tr->device);
  if (tr->stream) cudaStreamSynchronize(tr->stream);
  cudaFree(tr->d_unique);
  cudaFree(tr->d_a);
  cudaFree(tr->d_counts);
  cudaFree(tr->d_tmp);
  cudaFree(tr->d_runs);
  if (tr->h_stage) cudaFreeHost(tr->h_stage);
  if (tr->stream) cudaStreamDestroy(tr->stream);
  delete tr;
}

int nm_track_compute(nm_track *tr, const void *unique, uint64_t n, int width, uint32_t kmer_length) {
  if (!tr || (n && !unique) || (width != 1 && width != 2 && width != 4) || kmer_length == 0) return NM_EARG;
  if (n >= 0xFFFFFFFFull) return NM_EUNSUPPORTED;  // u32 prefix sums: one sequence per call
  TR_CUDA(cudaSetDevice(tr->device));
  int rc = tr_reserve(tr, n);
  if (rc != NM_OK) return rc;
  rc = tr_stage(tr, kTrStage);
  if (rc != NM_OK) return rc;
  cudaStream_t st = tr->stream;
  tr->n = n;
  tr->k = kmer_length;
  tr->nruns = 0;
  if (n == 0) return NM_OK;
  // upload through pinned staging (the caller's buffer is pageable: an np.fromfile array or a memory map)
  const uint64_t bytes = n * (uint64_t)width;
  for (uint64_t off = 0; off < bytes; off += kTrStage) {
    const uint64_t m = bytes - off < kTrStage ? bytes - off : kTrStage;
    memcpy(tr->h_stage, (const uint8_t *)unique + off, m);
    TR_CUDA(cudaMemcpyAsync((uint8_t *)tr->d_unique + off, tr->h_stage, m, cudaMemcpyHostToDevice, st));
    TR_CUDA(cudaStreamSynchronize(st));
  }
  if (width == 1) k_tr_marks<uint8_t><<<blocks(n), 256, 0, st>>>((const uint8_t *)tr->d_unique, n, kmer_length, tr->d_a);
  else if (width == 2) k_tr_marks<uint16_t><<<blocks(n), 256, 0, st>>>((const uint16_t *)tr->d_unique, n, kmer_length, tr->d_a);
  else k_tr_marks<uint32_t><<<blocks(n), 256, 0, st>>>((const uint32_t *)tr->d_unique, n, kmer_length, tr->d_a);
  size_t need = 0;
  TR_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, tr->d_a, tr->d_a, (int64_t)n, st));
  rc = tr_tmp(tr, need);
  if (rc != NM_OK) return rc;
  TR_CUDA(cub::DeviceScan::InclusiveSum(tr->d_tmp, need, tr->d_a, tr->d_a, (int64_t)n, st));
  k_tr_window<<<blocks(n), 256, 0, st>>>(tr->d_a, n, kmer_length, tr->d_counts);
  TR_CUDA(cudaGetLastError());
  TR_CUDA(cudaStreamSynchronize(st));
  return NM_OK;
}

int nm_track_counts(nm_track *tr, uint64_t first, uint64_t m, uint32_t *out) {
  if (!tr || first > tr->n || m > tr->n - first || (m && !out)) return NM_EARG;
  TR_CUDA(cudaSetDevice(tr->device));
  if (m) TR_CUDA(cudaMemcpy(out, tr->d_counts + first, m * 4, cudaMemcpyDeviceToHost));
  return NM_OK;
}

int nm_track_runs(nm_track *tr, uint64_t *starts, uint8_t *values, uint64_t capacity, uint64_t *nruns) {
  if (!tr || !nruns) return NM_EARG;
  TR_CUDA(cudaSetDevice(tr->device));
  cudaStream_t st = tr->stream;
  const uint64_t n = tr->n;
  *nruns = 0;
  if (n == 0) return NM_OK;
  // flags in the (now free) byte buffer of the input, selected positions in d_a viewed as u64 pairs
  uint8_t *flag = (uint8_t *)tr->d_unique;
  k_tr_change<<<blocks(n), 256, 0, st>>>(tr->d_counts, n, flag);
  uint64_t *sel = nullptr;
  unsigned long long *d_num = nullptr;
  struct Guard {
    uint64_t *a;
    unsigned long long *b;
    ~Guard() { cudaFree(a); cudaFree(b); }
  } guard{nullptr, nullptr};
  if (cudaMalloc((void **)&sel, (n + 1) * 8) != cudaSuccess || cudaMalloc((void **)&d_num, 8 + n + 16) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(sel);
    cudaFree(d_num);
    return NM_EALLOC;
  }
  guard.a = sel;
  guard.b = d_num;
  cub::CountingInputIterator<uint64_t> iota(0);
  size_t need = 0;
  TR_CUDA(cub::DeviceSelect::Flagged(nullptr, need, iota, flag, sel, d_num, (int64_t)n, st));
  int rc = tr_tmp(tr, need);
  if (rc != NM_OK) return rc;
  TR_CUDA(cub::DeviceSelect::Flagged(tr->d_tmp, need, iota, flag, sel, d_num, (int64_t)n, st));
  unsigned long long num = 0;
  TR_CUDA(cudaMemcpyAsync(&num, d_num, 8, cudaMemcpyDeviceToHost, st));
  TR_CUDA(cudaStreamSynchronize(st));
  *nruns = num;
  cudaFree(tr->d_runs);               // the run starts stay on the device for nm_track_bed_text
  tr->d_runs = sel;
  tr->nruns = num;
  guard.a = nullptr;
  const uint64_t m = num < capacity ? num : capacity;
  if (m && starts) TR_CUDA(cudaMemcpy(starts, sel, m * 8, cudaMemcpyDeviceToHost));
  if (m && values) {
    uint8_t *d_val = (uint8_t *)(d_num + 1);
    k_tr_run_values<<<blocks(m), 256, 0, st>>>(tr->d_counts, sel, m, d_val);
    TR_CUDA(cudaMemcpyAsync(values, d_val, m, cudaMemcpyDeviceToHost, st));
    TR_CUDA(cudaStreamSynchronize(st));
  }
  return NM_OK;
}

int nm_track_text(nm_track *tr, const char *strings, const uint32_t *offsets, uint64_t first, uint64_t m, char *out,
                  uint64_t capacity, uint64_t *written) {
  if (!tr || !strings || !offsets || !written || first > tr->n || m > tr->n - first || (m && !out)) return NM_EARG;
  TR_CUDA(cudaSetDevice(tr->device));
  cudaStream_t st = tr->stream;
  *written = 0;
  if (m == 0) return NM_OK;
  if (m >= 0x7FFFFFFFull) return NM_EARG;
  const uint32_t k = tr->k;
  const uint32_t total_chars = offsets[k + 1];
  // the k+1 strings and their offsets
  char *d_str = nullptr;
  uint32_t *d_off = nullptr;
  unsigned long long *d_at = nullptr;
  char *d_text = nullptr;
  struct Guard {
    void *p[4];
    ~Guard() { for (void *q : p) cudaFree(q); }
  } guard{{nullptr, nullptr, nullptr, nullptr}};
  uint32_t longest = 0;
  for (uint32_t c = 0; c <= k; ++c) longest = offsets[c + 1] - offsets[c] > longest ? offsets[c + 1] - offsets[c] : longest;
  const uint64_t text_cap = m * (uint64_t)(longest + 1);
  if (cudaMalloc((void **)&d_str, total_chars + 1) != cudaSuccess) return NM_EALLOC;
  guard.p[0] = d_str;
  if (cudaMalloc((void **)&d_off, (k + 2) * 4) != cudaSuccess) return NM_EALLOC;
  guard.p[1] = d_off;
  if (cudaMalloc((void **)&d_at, (m + 1) * 8) != cudaSuccess) return NM_EALLOC;
  guard.p[2] = d_at;
  if (cudaMalloc((void **)&d_text, text_cap) != cudaSuccess) return NM_EALLOC;
  guard.p[3] = d_text;
  TR_CUDA(cudaMemcpyAsync(d_str, strings, total_chars, cudaMemcpyHostToDevice, st));
  TR_CUDA(cudaMemcpyAsync(d_off, offsets, (k + 2) * 4, cudaMemcpyHostToDevice, st));
  uint32_t *len = tr->d_a;  // the prefix sums are no longer needed
  k_tr_text_len<<<blocks(m), 256, 0, st>>>(tr->d_counts + first, m, d_off, len);
  TR_CUDA(cudaMemsetAsync(len + m, 0, 4, st));
  size_t need = 0;
  TR_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, len, d_at, (int64_t)(m + 1), st));
  int rc = tr_tmp(tr, need);
  if (rc != NM_OK) return rc;
  TR_CUDA(cub::DeviceScan::ExclusiveSum(tr->d_tmp, need, len, d_at, (int64_t)(m + 1), st));
  unsigned long long total = 0;
  TR_CUDA(cudaMemcpyAsync(&total, d_at + m, 8, cudaMemcpyDeviceToHost, st));
  k_tr_text_put<<<blocks(m), 256, 0, st>>>(tr->d_counts + first, m, d_off, d_str, d_at, d_text);
  TR_CUDA(cudaStreamSynchronize(st));
  if (total > capacity) return NM_EARG;
  TR_CUDA(cudaMemcpy(out, d_text, total, cudaMemcpyDeviceToHost));
  *written = total;
  return NM_OK;
}

int nm_track_bed_text(nm_track *tr, const char *chr_name, uint64_t first_run, uint64_t m, char *out, uint64_t capacity,
                      uint64_t *written) {
  if (!tr || !chr_name || !written || first_run > tr->nruns || m > tr->nruns - first_run || (m && !out)) return NM_EARG;
  TR_CUDA(cudaSetDevice(tr->device));
  cudaStream_t st = tr->stream;
  *written = 0;
  if (m == 0) return NM_OK;
  if (m >= 0x7FFFFFFFull) return NM_EARG;
  char ktag[16];
  const uint32_t ktag_len = (uint32_t)snprintf(ktag, sizeof ktag, "k%u", tr->k);
  const uint32_t chr_len = (uint32_t)strlen(chr_name);
  const uint32_t fixed = chr_len + 1 + 1 + 1 + ktag_len + 1 + 1 + 1 + 1 + 1;  // tabs, k tag, value, '.', newline
  char *d_names = nullptr, *d_text = nullptr;
  unsigned long long *d_at = nullptr;
  struct Guard {
    void *p[3];
    ~Guard() { for (void *q : p) cudaFree(q); }
  } guard{{nullptr, nullptr, nullptr}};
  const uint64_t text_cap = m * (uint64_t)(fixed + 40);
  if (cudaMalloc((void **)&d_names, chr_len + ktag_len + 2) != cudaSuccess) return NM_EALLOC;
  guard.p[0] = d_names;
  if (cudaMalloc((void **)&d_at, (m + 1) * 8) != cudaSuccess) return NM_EALLOC;
  guard.p[1] = d_at;
  if (cudaMalloc((void **)&d_text, text_cap) != cudaSuccess) return NM_EALLOC;
  guard.p[2] = d_text;
  TR_CUDA(cudaMemcpyAsync(d_names, chr_name, chr_len, cudaMemcpyHostToDevice, st));
  TR_CUDA(cudaMemcpyAsync(d_names + chr_len, ktag, ktag_len, cudaMemcpyHostToDevice, st));
  if (m + 1 > tr->cap) return NM_EARG;  // runs <= positions
  uint32_t *len = tr->d_a;
  k_tr_bed_len<<<blocks(m), 256, 0, st>>>(tr->d_runs, first_run, m, tr->nruns, tr->n, fixed, len);
  TR_CUDA(cudaMemsetAsync(len + m, 0, 4, st));
  size_t need = 0;
  TR_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, len, d_at, (int64_t)(m + 1), st));
  int rc = tr_tmp(tr, need);
  if (rc != NM_OK) return rc;
  TR_CUDA(cub::DeviceScan::ExclusiveSum(tr->d_tmp, need, len, d_at, (int64_t)(m + 1), st));
  unsigned long long total = 0;
  TR_CUDA(cudaMemcpyAsync(&total, d_at + m, 8, cudaMemcpyDeviceToHost, st));
  k_tr_bed_put<<<blocks(m), 256, 0, st>>>(tr->d_runs, tr->d_counts, first_run, m, tr->nruns, tr->n, d_names, chr_len,
                                        d_names + chr_len, ktag_len, d_at, d_text);
  TR_CUDA(cudaStreamSynchronize(st));
  if (total > capacity) return NM_EARG;
  TR_CUDA(cudaMemcpy(out, d_text, total, cudaMemcpyDeviceToHost));
  *written = total;
  return NM_OK;
}

}  // extern "C"
