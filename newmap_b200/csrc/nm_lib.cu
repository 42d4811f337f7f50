// nm_lib.cu -- host side of libnewmap_b200.so: .awfmi loader, device-resident index, segment
// workspaces and the C-ABI of include/newmap_b200.h.  No CPU fallback anywhere: every entry point
// that computes launches the kernels of nm_kernels.cu or fails.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/newmap_b200.h"
#include "nm_fasta.cuh"
#include "nm_kernels.cuh"

namespace {

thread_local std::string g_cuda_error;

int cuda_fail(cudaError_t e, const char *what) {
  g_cuda_error = std::string(what) + ": " + cudaGetErrorString(e);
  return NM_ECUDA;
}
#define NM_CUDA(call)                                  \
  do {                                                 \
    cudaError_t e__ = (call);                          \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
  } while (0)


}  // namespace

struct nm_index {
  nm::DevIndex dev{};
  int device = 0;
  uint64_t n_blocks = 0;
  uint64_t seed_entries = 0;
  uint32_t sa_ratio = 0;
  bool has_rank32 = false;
  uint64_t device_bytes = 0;
  void *d_native = nullptr, *d_rank32 = nullptr, *d_seed32 = nullptr, *d_seed64 = nullptr, *d_c32 = nullptr;
  void *d_seedx = nullptr;             // deeper seed table derived on the device (set_seed_depth)
  uint64_t seedx_bytes = 0;
  nm_workspace *default_ws = nullptr;  // used by nm_min_unique
  std::mutex mu;                       // serialises calls that share default_ws / count scratch
};

struct nm_workspace {
  nm_index *ix = nullptr;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  uint64_t max_seg_len = 0;
  // device buffers
  uint8_t *d_seg = nullptr, *d_codes = nullptr;
  const uint8_t *seg_ext = nullptr;  // segment attached from a parsed FASTA (nm_workspace_attach) instead of uploaded
  uint32_t *d_pk = nullptr, *d_tile_first = nullptr, *d_G = nullptr, *d_R = nullptr, *d_deepmap = nullptr;
  uint64_t *d_barmask = nullptr;
  uint8_t *d_sub = nullptr;  // candidates for lengths below the table depth (allocated by the first search that needs it)
  void *d_out = nullptr;
  uint32_t *d_klist = nullptr;
  uint32_t klist_cap = 0;
  unsigned long long *d_counters = nullptr;
  unsigned long long *h_counters = nullptr;  // pinned
  cudaEvent_t ev[10]{};
  uint32_t launches = 0;  // kernels launched by the last search (run_search)
  // last call
  nm::SegParams P{};
  uint64_t seg_len = 0;
  bool have_search = false;
  int algo_req = 0, algo_used = 0;
  bool timed_h2d = false, timed_d2h = false;
  void *pending_out = nullptr;
  uint32_t opt_ref_lf = 0, opt_dcap = 0, opt_gsize = 0, opt_pass1 = 0, opt_gsize2 = ~0u, opt_own_chains = ~0u, opt_stall = ~0u, opt_cert_runs = 0, opt_stride = 0, opt_out_stats = 0, opt_subdepth = 0, opt_strict = 0;
  // several sequences x several indexes (nm_workspace_multi); cleared by the next upload / attach
  nm::CoarseGroup *d_cgroups = nullptr;  // coarse certificate round: 2 per group of >= kMinGsize2 positions
  uint64_t cgroup_cap = 0;
  cudaStream_t aux_stream = nullptr;     // the coarse round's anchor chains run here, beside the first pass
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_strict = nullptr;
  nm::MultiSet *d_multi = nullptr;
  uint8_t *d_extra[nm::kMaxMulti]{};  // extra sequences' segments, max_seg_len + 16 bytes each, allocated on demand
  uint32_t multi_ix = 0, multi_seq = 0;
};

namespace {

// AwFmFile.c:20-171 (writer) / :174-300 (reader): "AwFmIndex\n" | u32 version | u32 featureFlags |
// u8 suffixArrayCompressionRatio | u8 kmerLengthInSeedTable | u8 alphabetType | u8 storeOriginalSequence |
// u64 bwtLength | blocks | u64 prefixSums[6] | seed table | ...
constexpr size_t kHeaderBytes = 30;

// Deeper seed table: level d+1 from level d with one backward-extension step per entry, starting from the
// index's own table.  depth 0 (or <= seed_len) drops it.  Needs the rank32 layout.
int set_seed_depth(nm_index *ix, uint32_t depth) {
  NM_CUDA(cudaSetDevice(ix->device));
  NM_CUDA(cudaDeviceSynchronize());
  if (ix->d_seedx) {
    cudaFree(ix->d_seedx);
    ix->d_seedx = nullptr;
    ix->device_bytes -= ix->seedx_bytes;
    ix->seedx_bytes = 0;
    ix->dev.seedx = nullptr;
    ix->dev.seedx_len = 0;
  }
  const uint32_t sl = ix->dev.seed_len;
  if (depth <= sl) return NM_OK;
  if (!ix->has_rank32 || depth > 16) return NM_EUNSUPPORTED;
  void *cur = ix->d_seed32;
  for (uint32_t d = sl; d < depth; ++d) {
    void *next = nullptr;
    const size_t bytes = (size_t)sizeof(nm::SeedEntry32) << (2 * (d + 1));
    if (cudaMalloc(&next, bytes) != cudaSuccess) {
      cudaGetLastError();
      if (cur != ix->d_seed32) cudaFree(cur);
      return NM_EALLOC;
    }
    nm::launch_seed_extend(ix->dev, (const nm::SeedEntry32 *)cur, d, (nm::SeedEntry32 *)next, 0);
    const cudaError_t e = cudaDeviceSynchronize();
    if (cur != ix->d_seed32) cudaFree(cur);
    cur = next;
    if (e != cudaSuccess) {
      cudaFree(cur);
      return cuda_fail(e, "k_seed_extend");
    }
  }
  // both strands in one entry: pair[key] = {t[key], t[reverse complement of key]}
  void *pair = nullptr;
  const size_t pair_bytes = (size_t)sizeof(nm::SeedEntryPair) << (2 * depth);
  if (cudaMalloc(&pair, pair_bytes) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(cur);
    return NM_EALLOC;
  }
  nm::launch_seed_pair((const nm::SeedEntry32 *)cur, depth, (nm::SeedEntryPair *)pair, 0);
  const cudaError_t e = cudaDeviceSynchronize();
  cudaFree(cur);
  if (e != cudaSuccess) {
    cudaFree(pair);
    return cuda_fail(e, "k_seed_pair");
  }
  ix->d_seedx = pair;
  ix->seedx_bytes = pair_bytes;
  ix->device_bytes += ix->seedx_bytes;
  ix->dev.seedx = (const nm::SeedEntryPair *)pair;
  ix->dev.seedx_len = depth;
  return NM_OK;
}

// Depth chosen at load time: the smallest d with 4^d >= 4 * bwtLength (beyond that almost every d-mer of the
// text is already unique, so a chain is one table lookup plus a step or two), at most 16 (u32 keys), and the
// table within what the free device memory allows (see below).  NEWMAP_B200_SEED_DEPTH overrides
// (0 = none).  Purely a speed matter: results do not depend on it.
uint32_t auto_seed_depth(const nm_index *ix) {
  if (const char *v = getenv("NEWMAP_B200_SEED_DEPTH")) return (uint32_t)strtoul(v, nullptr, 10);
  uint32_t d = ix->dev.seed_len;
  while (d < 16 && (1ull << (2 * d)) < 4 * ix->dev.bwt_len) ++d;
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return 0;
  // peak while building: the one-strand table (8 B x 4^d) next to the both-strand one (16 B x 4^d); leave 40 % free
  while (d > ix->dev.seed_len && ((24ull << (2 * d)) > free_b / 5 * 3)) --d;
  return d;
}

int load_index(const char *path, int device, nm_index **out) {
  int ndev = 0;
  NM_CUDA(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) return cuda_fail(cudaErrorNoDevice, "cudaGetDeviceCount");
  if (device < 0 || device >= ndev) return NM_EARG;
  NM_CUDA(cudaSetDevice(device));

  FILE *f = fopen(path, "rb");
  if (!f) return NM_EOPEN;
  struct Closer {
    FILE *f;
    ~Closer() { if (f) fclose(f); }
  } closer{f};
  uint8_t hdr[kHeaderBytes];
  if (fread(hdr, 1, kHeaderBytes, f) != kHeaderBytes) return NM_ESHORT;
  if (memcmp(hdr, "AwFmIndex\n", 10) != 0) return NM_EFORMAT;
  uint32_t version;
  memcpy(&version, hdr + 10, 4);
  if (version != 8 || hdr[20] != 2) return NM_EVERSION;  // AwFmIndexStruct.c version, AwFmAlphabetDna = 2
  uint64_t bwt_len;
  memcpy(&bwt_len, hdr + 22, 8);
  const uint32_t seed_len = hdr[19];
  if (bwt_len == 0 || seed_len == 0 || seed_len > 15) return NM_EFORMAT;

  nm_index *ix = new (std::nothrow) nm_index();
  if (!ix) return NM_EALLOC;
  struct Guard {
    nm_index *ix;
    ~Guard() { if (ix) nm_index_close(ix); }
  } guard{ix};
  ix->device = device;
  ix->sa_ratio = hdr[18];
  ix->n_blocks = 1 + (bwt_len - 1) / 256;  // AwFmIndexStruct.c:102-104
  ix->seed_entries = 1ull << (2 * seed_len);
  const uint64_t native_bytes = (ix->n_blocks + 1) * nm::kNativeBlockBytes;
  const uint64_t seed_bytes = ix->seed_entries * 16;

  // stage through pinned host memory in chunks so the file streams straight into HBM
  const size_t chunk = 64u << 20;
  uint8_t *stage = nullptr;
  if (cudaHostAlloc((void **)&stage, chunk, cudaHostAllocDefault) != cudaSuccess) return NM_EALLOC;
  struct StageFree {
    uint8_t *p;
    ~StageFree() { cudaFreeHost(p); }
  } stage_free{stage};

  if (cudaMalloc(&ix->d_native, native_bytes) != cudaSuccess) return NM_EALLOC;
  if (cudaMalloc(&ix->d_seed64, seed_bytes) != cudaSuccess) return NM_EALLOC;
  ix->device_bytes = native_bytes + seed_bytes;

  auto stream_in = [&](void *dst, uint64_t bytes) -> int {
    uint64_t done = 0;
    while (done < bytes) {
      const size_t want = (size_t)((bytes - done < chunk) ? bytes - done : chunk);
      if (fread(stage, 1, want, f) != want) return NM_ESHORT;
      if (cudaMemcpy((uint8_t *)dst + done, stage, want, cudaMemcpyHostToDevice) != cudaSuccess) return NM_ECUDA;
      done += want;
    }
    return NM_OK;
  };
  int rc = stream_in(ix->d_native, ix->n_blocks * nm::kNativeBlockBytes);
  if (rc != NM_OK) return rc;
  uint64_t prefix[6];
  if (fread(prefix, 8, 6, f) != 6) return NM_ESHORT;
  if (prefix[5] != bwt_len || prefix[0] != 1) return NM_EFORMAT;
  for (int l = 0; l < 5; ++l)
    if (prefix[l] > prefix[l + 1]) return NM_EFORMAT;
  rc = stream_in(ix->d_seed64, seed_bytes);
  if (rc != NM_OK) return rc;

  // pad block after the last one: zero planes, counters = totals, so that a rank at p == 256*n_blocks is defined
  uint64_t pad[20];
  memset(pad, 0, sizeof pad);
  for (int l = 0; l < 5; ++l) pad[12 + l] = prefix[l + 1] - prefix[l];
  pad[12 + 5] = 1;
  NM_CUDA(cudaMemcpy((uint8_t *)ix->d_native + ix->n_blocks * nm::kNativeBlockBytes, pad, nm::kNativeBlockBytes,
                     cudaMemcpyHostToDevice));

  ix->dev.native = (const uint8_t *)ix->d_native;
  ix->dev.seed64 = (const uint64_t *)ix->d_seed64;
  for (int l = 0; l < 6; ++l) ix->dev.C[l] = prefix[l];
  ix->dev.bwt_len = bwt_len;
  ix->dev.seed_len = seed_len;

  // derived layouts (device-side transcode of the unchanged file contents)
  if (bwt_len < 0xFFFFFFFEull) {
    const uint64_t n64 = (ix->n_blocks + 1) * 4;
    if (cudaMalloc(&ix->d_rank32, n64 * nm::kRank32Words * 4) != cudaSuccess) return NM_EALLOC;
    if (cudaMalloc(&ix->d_seed32, ix->seed_entries * sizeof(nm::SeedEntry32)) != cudaSuccess) return NM_EALLOC;
    ix->device_bytes += n64 * nm::kRank32Words * 4 + ix->seed_entries * sizeof(nm::SeedEntry32);
    nm::launch_transcode(ix->dev.native, ix->n_blocks, bwt_len, (uint32_t *)ix->d_rank32, n64, prefix, 0);
    nm::launch_seed_transcode(ix->dev.seed64, (nm::SeedEntry32 *)ix->d_seed32, ix->seed_entries, 0);
    NM_CUDA(cudaGetLastError());
    NM_CUDA(cudaDeviceSynchronize());
    uint32_t c32[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int l = 0; l < 6; ++l) c32[l] = (uint32_t)prefix[l];
    if (cudaMalloc(&ix->d_c32, sizeof c32) != cudaSuccess) return NM_EALLOC;
    NM_CUDA(cudaMemcpy(ix->d_c32, c32, sizeof c32, cudaMemcpyHostToDevice));
    ix->dev.c32 = (const uint32_t *)ix->d_c32;
    ix->dev.rank32 = (const uint32_t *)ix->d_rank32;
    ix->dev.seed32 = (const nm::SeedEntry32 *)ix->d_seed32;
    ix->has_rank32 = true;
    const int rc = set_seed_depth(ix, auto_seed_depth(ix));
    if (rc != NM_OK && rc != NM_EALLOC && rc != NM_EUNSUPPORTED) return rc;  // without the table the searches still run
  }
  guard.ix = nullptr;
  *out = ix;
  return NM_OK;
}

constexpr uint32_t kMinGsize2 = 32;  // smallest group of the coarse certificate round (sizes its buffers)

int ws_alloc(nm_workspace *ws) {
  const uint64_t n = ws->max_seg_len;
  const uint64_t ntiles = (n + nm::kTile - 1) / nm::kTile + 1;
  NM_CUDA(cudaMalloc((void **)&ws->d_seg, n + 16));
  NM_CUDA(cudaMalloc((void **)&ws->d_codes, n + 16));
  NM_CUDA(cudaMalloc((void **)&ws->d_pk, (n / 16 + 4) * 4));
  NM_CUDA(cudaMalloc((void **)&ws->d_barmask, ntiles * 8));
  NM_CUDA(cudaMalloc((void **)&ws->d_deepmap, (n / 32 + 4) * 4));
  NM_CUDA(cudaMalloc((void **)&ws->d_tile_first, ntiles * 4));
  NM_CUDA(cudaMalloc((void **)&ws->d_G, (n + 2) * 4));
  NM_CUDA(cudaMalloc((void **)&ws->d_R, (n + 1) * 4));
  NM_CUDA(cudaMalloc(&ws->d_out, (n + 1) * 4));
  ws->cgroup_cap = 2 * ((n + 1) / kMinGsize2 + 2);
  NM_CUDA(cudaMalloc((void **)&ws->d_cgroups, ws->cgroup_cap * sizeof(nm::CoarseGroup)));
  NM_CUDA(cudaStreamCreateWithFlags(&ws->aux_stream, cudaStreamNonBlocking));
  NM_CUDA(cudaEventCreateWithFlags(&ws->ev_fork, cudaEventDisableTiming));
  NM_CUDA(cudaEventCreateWithFlags(&ws->ev_join, cudaEventDisableTiming));
  NM_CUDA(cudaEventCreateWithFlags(&ws->ev_strict, cudaEventDisableTiming));
  NM_CUDA(cudaMalloc((void **)&ws->d_counters, nm::CTR_N * 8));
  NM_CUDA(cudaHostAlloc((void **)&ws->h_counters, nm::CTR_N * 8, cudaHostAllocDefault));
  for (auto &e : ws->ev) NM_CUDA(cudaEventCreate(&e));
  return NM_OK;
}

enum { EV_H2D0 = 0, EV_H2D1, EV_PREP1, EV_PASS1, EV_CERT1, EV_SEARCH1, EV_GENERIC1, EV_RESOLVE1, EV_D2H0, EV_D2H1 };

// launches prep + the chosen algorithm on ws->stream; nothing is synchronised here
int run_search(nm_workspace *ws, int algo) {
  nm_index *ix = ws->ix;
  nm::SegParams &P = ws->P;
  cudaStream_t st = ws->stream;
  // (the counters are reset by block 0 of k_prep; an empty segment launches no kernel at all)
  if (P.seg_len == 0) {
    static const unsigned long long init_counters[nm::CTR_N] = {0, 0, 0, 0, ~0ull, 0, 0, 0, 0, 0, 0, 0, ~0ull, 0, 0, 0};
    NM_CUDA(cudaMemcpyAsync(ws->d_counters, init_counters, sizeof init_counters, cudaMemcpyHostToDevice, st));
  }
  NM_CUDA(cudaEventRecord(ws->ev[EV_H2D1], st));
  nm::launch_prep(ix->dev, P, st);
  ws->launches = P.seg_len ? 1u : 0u;
  NM_CUDA(cudaEventRecord(ws->ev[EV_PREP1], st));
  if (algo == NM_ALGO_TRAJECTORY) {
    if (ws->multi_ix) nm::launch_trajectory_multi(ws->d_multi, P, st);
    else nm::launch_trajectory(ix->dev, P, st);
    ws->launches += P.num_kmers ? 1u : 0u;
    NM_CUDA(cudaEventRecord(ws->ev[EV_PASS1], st));
    NM_CUDA(cudaEventRecord(ws->ev[EV_CERT1], st));
    NM_CUDA(cudaEventRecord(ws->ev[EV_SEARCH1], st));
    NM_CUDA(cudaEventRecord(ws->ev[EV_GENERIC1], st));
    NM_CUDA(cudaEventRecord(ws->ev[EV_RESOLVE1], st));
  } else {
    const bool fast = (algo != NM_ALGO_FREE_GENERIC) && ix->has_rank32 && P.kmin >= ix->dev.seed_len;
    P.fast = fast ? 1 : 0;
    // strict mode: the presence chains (presence_thread, nm_core.cuh) need only the codes: they run on the
    // workspace's second stream beside the passes and are joined before the counters are read
    const bool strict = ws->opt_strict && ix->has_rank32 && P.num_kmers;
    if (strict) {
      NM_CUDA(cudaEventRecord(ws->ev_fork, st));
      NM_CUDA(cudaStreamWaitEvent(ws->aux_stream, ws->ev_fork, 0));
      nm::launch_presence(ix->dev, P, ws->opt_strict, ws->aux_stream);
      NM_CUDA(cudaEventRecord(ws->ev_strict, ws->aux_stream));
      ++ws->launches;
    }
    const bool deep_passes = fast && P.dcap < P.kmax;
    // coarse round (coarse_anchor_thread, nm_core.cuh): anchor chains of max_k + gsize2 - 1 steps at fixed positions,
    // on the workspace's second stream BESIDE the first pass, settled after both.  Off by default: it removes a third
    // of a segment's LF steps (2.8 -> 2.0 per position on the 1 Gbp bench genome), but beside the first pass, which
    // keeps the memory system at its random-access ceiling, a dependent step costs about 3 us and the 455-step chains
    // end 0.6 ms after it (8.73 -> 8.12 G positions/s); run alone after it they take 0.7 ms (profiles/r1_gsize2_*).
    const uint32_t g2 = ws->opt_gsize2 == ~0u ? 0u : ws->opt_gsize2;
    const bool coarse = deep_passes && g2 > P.gsize && g2 >= kMinGsize2;
    nm::SegParams Pc = P;
    if (coarse) {
      Pc.gsize = g2;
      NM_CUDA(cudaEventRecord(ws->ev_fork, st));
      NM_CUDA(cudaStreamWaitEvent(ws->aux_stream, ws->ev_fork, 0));
      nm::launch_coarse_anchor(ix->dev, Pc, ws->d_cgroups, ws->aux_stream);
      NM_CUDA(cudaEventRecord(ws->ev_join, ws->aux_stream));
      ++ws->launches;
    }
    if (fast) {
      // tiled kernel when the chains are short (deep seed table); with the file's 12-letter table the chains run
      // 4-5 steps on average and one thread per position is faster (measured: profiles/r1_seed_depth_sweep.md)
      // the sparse kernel needs the both-strand table; one thread per position otherwise (the file's 12-letter table:
      // chains of 4-5 steps) or when asked for (NM_OPT_PASS1 = 1: the dense reference of the tests)
      const bool sparse = P.seedpair != nullptr && ws->opt_pass1 != 1u;
      if (sparse) nm::launch_free_pass1(ix->dev, P, st);
      else { P.stride = 1; nm::launch_free_chain(ix->dev, P, true, st); }
      ++ws->launches;
    }
    NM_CUDA(cudaEventRecord(ws->ev[EV_PASS1], st));
    if (deep_passes) {
      if (coarse) {
        NM_CUDA(cudaStreamWaitEvent(st, ws->ev_join, 0));
        nm::launch_coarse_settle(Pc, ws->d_cgroups, st);
        ++ws->launches;
      }
      nm::launch_free_cert(ix->dev, P, st);
      ++ws->launches;
    }
    NM_CUDA(cudaEventRecord(ws->ev[EV_CERT1], st));
    if (deep_passes) { nm::launch_free_chain(ix->dev, P, false, st); ++ws->launches; }
    NM_CUDA(cudaEventRecord(ws->ev[EV_SEARCH1], st));
    if (!fast) { nm::launch_free_generic(ix->dev, P, true, st); ++ws->launches; }
    NM_CUDA(cudaEventRecord(ws->ev[EV_GENERIC1], st));
    if (strict) NM_CUDA(cudaStreamWaitEvent(st, ws->ev_strict, 0));
    nm::launch_resolve(ix->dev, P, st);
    ws->launches += P.num_kmers ? 1u : 0u;
    if (P.keff > P.kmin) {   // a range that starts below the depth of the table in use: the lengths below it
      nm::launch_subdepth(ix->dev, P, st);
      ws->launches += P.num_kmers ? 1u : 0u;
    }
    NM_CUDA(cudaEventRecord(ws->ev[EV_RESOLVE1], st));
  }
  if (ws->opt_out_stats) {
    nm::launch_out_stats(P, st);
    ws->launches += P.num_kmers ? 1u : 0u;
  }
  NM_CUDA(cudaGetLastError());
  NM_CUDA(cudaMemcpyAsync(ws->h_counters, ws->d_counters, nm::CTR_N * 8, cudaMemcpyDeviceToHost, st));
  ws->algo_used = algo == NM_ALGO_AUTO ? NM_ALGO_AUTO : algo;
  return NM_OK;
}

}  // namespace

void nm_internal_set_cuda_error(const char *what, const char *message) {
  g_cuda_error = std::string(what) + ": " + message;
}

extern "C" {

int nm_abi_version(void) { return NM_ABI_VERSION; }

const char *nm_strerror(int code) {
  switch (code) {
    case NM_OK: return "ok";
    case NM_EOPEN: return "could not open the index file";
    case NM_EFORMAT: return "not a valid .awfmi index";
    case NM_EVERSION: return "unsupported .awfmi version or alphabet (need v8, DNA)";
    case NM_EALLOC: return "allocation failed";
    case NM_ESHORT: return "index file is truncated";
    case NM_EARG: return "invalid argument";
    case NM_ECUDA: return "CUDA error";
    case NM_EUNSUPPORTED: return "unsupported";
    case NM_ELIMIT: return "request exceeds the fixed scratch of the no-allocation entry point (NM_NOALLOC_SCRATCH_BYTES)";
    default: return "unknown error";
  }
}

const char *nm_last_cuda_error(void) { return g_cuda_error.c_str(); }


int nm_device_count(void) {
  int n = 0;
  const cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
  return n;
}

int nm_index_open(const char *path, int device, nm_index **out) {
  if (!path || !out) return NM_EARG;
  *out = nullptr;
  return load_index(path, device, out);
}

void nm_index_close(nm_index *ix) {
  if (!ix) return;
  cudaSetDevice(ix->device);
  if (ix->default_ws) nm_workspace_destroy(ix->default_ws);
  cudaFree(ix->d_native);
  cudaFree(ix->d_rank32);
  cudaFree(ix->d_seed32);
  cudaFree(ix->d_seedx);
  cudaFree(ix->d_seed64);
  cudaFree(ix->d_c32);
  delete ix;
}

int nm_index_set_seed_depth(nm_index *ix, uint32_t depth) {
  if (!ix) return NM_EARG;
  std::lock_guard<std::mutex> lock(ix->mu);
  return set_seed_depth(ix, depth);
}

int nm_index_info_get(const nm_index *ix, nm_index_info *info) {
  if (!ix || !info) return NM_EARG;
  info->bwt_length = ix->dev.bwt_len;
  info->n_blocks = ix->n_blocks;
  for (int l = 0; l < 6; ++l) info->prefix_sums[l] = ix->dev.C[l];
  info->device_bytes = ix->device_bytes;
  info->seed_length = ix->dev.seed_len;
  info->derived_seed_length = ix->dev.seedx_len;
  info->sa_compression_ratio = ix->sa_ratio;
  info->has_rank32 = ix->has_rank32 ? 1u : 0u;
  info->device = ix->device;
  return NM_OK;
}

int nm_count_from_sequence(nm_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                           const uint64_t *len, uint64_t n, int reverse_complement, uint32_t *counts,
                           uint64_t *lf_steps) {
  if (!ix || (n && (!idx || !len || !counts)) || (seq_len && !seq)) return NM_EARG;
  for (uint64_t q = 0; q < n; ++q)  // newmap-count.c:63-68,184-190
    if (len[q] == 0 || idx[q] > seq_len || len[q] > seq_len - idx[q]) return NM_EARG;
  if (n == 0) return NM_OK;
  std::lock_guard<std::mutex> lock(ix->mu);
  NM_CUDA(cudaSetDevice(ix->device));
  uint8_t *d_seq = nullptr;
  uint64_t *d_idx = nullptr, *d_len = nullptr;
  uint32_t *d_counts = nullptr;
  unsigned long long *d_ctr = nullptr;
  int rc = NM_OK;
  auto body = [&]() -> int {
    NM_CUDA(cudaMalloc((void **)&d_seq, seq_len + 16));
    NM_CUDA(cudaMalloc((void **)&d_idx, n * 8));
    NM_CUDA(cudaMalloc((void **)&d_len, n * 8));
    NM_CUDA(cudaMalloc((void **)&d_counts, n * 4));
    NM_CUDA(cudaMalloc((void **)&d_ctr, nm::CTR_N * 8));
    NM_CUDA(cudaMemset(d_ctr, 0, nm::CTR_N * 8));
    NM_CUDA(cudaMemcpy(d_seq, seq, seq_len, cudaMemcpyHostToDevice));
    NM_CUDA(cudaMemcpy(d_idx, idx, n * 8, cudaMemcpyHostToDevice));
    NM_CUDA(cudaMemcpy(d_len, len, n * 8, cudaMemcpyHostToDevice));
    nm::launch_count(ix->dev, d_seq, d_idx, d_len, n, reverse_complement != 0, d_counts, d_ctr, 0);
    NM_CUDA(cudaGetLastError());
    NM_CUDA(cudaMemcpy(counts, d_counts, n * 4, cudaMemcpyDeviceToHost));
    if (lf_steps) {
      unsigned long long v = 0;
      NM_CUDA(cudaMemcpy(&v, d_ctr + nm::CTR_LF, 8, cudaMemcpyDeviceToHost));
      *lf_steps += v;
    }
    return NM_OK;
  };
  rc = body();
  cudaFree(d_seq);
  cudaFree(d_idx);
  cudaFree(d_len);
  cudaFree(d_counts);
  cudaFree(d_ctr);
  return rc;
}

int nm_workspace_create(nm_index *ix, uint64_t max_seg_len, void *cuda_stream, nm_workspace **out) {
  if (!ix || !out || max_seg_len == 0 || max_seg_len >= 0xFFFFFF00ull) return NM_EARG;
  *out = nullptr;
  NM_CUDA(cudaSetDevice(ix->device));
  nm_workspace *ws = new (std::nothrow) nm_workspace();
  if (!ws) return NM_EALLOC;
  ws->ix = ix;
  ws->max_seg_len = max_seg_len;
  if (const char *v = getenv("NEWMAP_B200_GSIZE2")) ws->opt_gsize2 = (uint32_t)strtoul(v, nullptr, 10);  // tuning experiments
  if (const char *v = getenv("NEWMAP_B200_STALL")) ws->opt_stall = (uint32_t)strtoul(v, nullptr, 10);
  if (const char *v = getenv("NEWMAP_B200_OWN_CHAINS")) ws->opt_own_chains = (uint32_t)strtoul(v, nullptr, 10);
  if (cuda_stream) {
    ws->stream = (cudaStream_t)cuda_stream;
  } else {
    if (cudaStreamCreateWithFlags(&ws->stream, cudaStreamNonBlocking) != cudaSuccess) {
      delete ws;
      return NM_ECUDA;
    }
    ws->own_stream = true;
  }
  const int rc = ws_alloc(ws);
  if (rc != NM_OK) {
    nm_workspace_destroy(ws);
    return rc;
  }
  *out = ws;
  return NM_OK;
}

void nm_workspace_destroy(nm_workspace *ws) {
  if (!ws) return;
  cudaSetDevice(ws->ix->device);
  cudaStreamSynchronize(ws->stream);
  cudaFree(ws->d_seg);
  cudaFree(ws->d_codes);
  cudaFree(ws->d_pk);
  cudaFree(ws->d_barmask);
  cudaFree(ws->d_deepmap);
  cudaFree(ws->d_sub);
  cudaFree(ws->d_tile_first);
  cudaFree(ws->d_G);
  cudaFree(ws->d_R);
  cudaFree(ws->d_out);
  cudaFree(ws->d_klist);
  cudaFree(ws->d_multi);
  cudaFree(ws->d_cgroups);
  if (ws->aux_stream) cudaStreamDestroy(ws->aux_stream);
  if (ws->ev_fork) cudaEventDestroy(ws->ev_fork);
  if (ws->ev_join) cudaEventDestroy(ws->ev_join);
  if (ws->ev_strict) cudaEventDestroy(ws->ev_strict);
  for (auto *p : ws->d_extra) cudaFree(p);
  cudaFree(ws->d_counters);
  if (ws->h_counters) cudaFreeHost(ws->h_counters);
  for (auto &e : ws->ev)
    if (e) cudaEventDestroy(e);
  if (ws->own_stream) cudaStreamDestroy(ws->stream);
  delete ws;
}

int nm_workspace_upload(nm_workspace *ws, const uint8_t *seg, uint64_t seg_len) {
  if (!ws || (seg_len && !seg) || seg_len > ws->max_seg_len) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  NM_CUDA(cudaEventRecord(ws->ev[EV_H2D0], ws->stream));
  if (seg_len) NM_CUDA(cudaMemcpyAsync(ws->d_seg, seg, seg_len, cudaMemcpyHostToDevice, ws->stream));
  ws->seg_len = seg_len;
  ws->seg_ext = nullptr;
  ws->timed_h2d = true;
  ws->have_search = false;
  ws->multi_ix = ws->multi_seq = 0;
  return NM_OK;
}

int nm_workspace_multi(nm_workspace *ws, nm_index *const *indexes, uint32_t n_indexes, const uint8_t *const *segments,
                       uint32_t n_segments) {
  if (!ws || !indexes || n_indexes == 0 || n_indexes > nm::kMaxMulti || n_segments + 1 > nm::kMaxMulti ||
      (n_segments && !segments) || ws->seg_ext)
    return NM_EARG;
  for (uint32_t x = 0; x < n_indexes; ++x)
    if (!indexes[x] || indexes[x]->device != ws->ix->device) return NM_EARG;
  for (uint32_t j = 0; j < n_segments; ++j)
    if (!segments[j] && ws->seg_len) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  if (!ws->d_multi) NM_CUDA(cudaMalloc((void **)&ws->d_multi, sizeof(nm::MultiSet)));
  nm::MultiSet M;
  memset(&M, 0, sizeof M);
  M.n_ix = n_indexes;
  M.n_seq = n_segments + 1;
  for (uint32_t x = 0; x < n_indexes; ++x) M.ix[x] = indexes[x]->dev;
  M.seg[0] = ws->d_seg;
  for (uint32_t j = 0; j < n_segments; ++j) {
    if (!ws->d_extra[j]) NM_CUDA(cudaMalloc((void **)&ws->d_extra[j], ws->max_seg_len + 16));
    if (ws->seg_len)
      NM_CUDA(cudaMemcpyAsync(ws->d_extra[j], segments[j], ws->seg_len, cudaMemcpyHostToDevice, ws->stream));
    M.seg[j + 1] = ws->d_extra[j];
  }
  // the set is read by the kernel of the search that follows on the same stream; the copy source is this stack
  // frame, so the copy is synchronous with respect to the host
  NM_CUDA(cudaMemcpyAsync(ws->d_multi, &M, sizeof M, cudaMemcpyHostToDevice, ws->stream));
  NM_CUDA(cudaStreamSynchronize(ws->stream));
  ws->multi_ix = n_indexes;
  ws->multi_seq = n_segments + 1;
  ws->have_search = false;
  return NM_OK;
}

int nm_workspace_attach(nm_workspace *ws, const nm_fasta *fa, uint64_t seq_offset, uint64_t seg_len) {
  if (!ws || !fa) return NM_EARG;
  uint64_t total = 0;
  int device = -1;
  const uint8_t *seq = nm_fasta_device_sequence(fa, &total, &device);
  if (!seq || device != ws->ix->device || seg_len > ws->max_seg_len || seq_offset > total || seg_len > total - seq_offset)
    return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  NM_CUDA(cudaEventRecord(ws->ev[EV_H2D0], ws->stream));
  ws->seg_ext = seq + seq_offset;
  ws->multi_ix = ws->multi_seq = 0;
  ws->seg_len = seg_len;
  ws->timed_h2d = true;
  ws->have_search = false;
  return NM_OK;
}

int nm_workspace_search(nm_workspace *ws, uint64_t num_kmers, int mode, const uint32_t *klist, uint32_t nk,
                        uint32_t initial_search_length, int use_rc, int algo, int out_width) {
  if (!ws || !klist || nk == 0) return NM_EARG;
  if (num_kmers > ws->seg_len) return NM_EARG;
  if (out_width != 1 && out_width != 2 && out_width != 4) return NM_EARG;
  if (algo != NM_ALGO_AUTO && algo != NM_ALGO_TRAJECTORY && algo != NM_ALGO_FREE_GENERIC) return NM_EARG;
  uint32_t kmin, kmax;
  if (mode == NM_MODE_RANGE) {
    if (nk != 2) return NM_EARG;
    kmin = klist[0];
    kmax = klist[1];
    if (kmin == 0 || kmin > kmax) return NM_EARG;
    // the reference never terminates when the first query is clamped below the lower bound
    if (initial_search_length && initial_search_length < kmin) return NM_EARG;
  } else if (mode == NM_MODE_LIST) {
    kmin = kmax = klist[0];
    for (uint32_t j = 0; j < nk; ++j) {
      if (klist[j] == 0) return NM_EARG;
      if (klist[j] < kmin) kmin = klist[j];
      if (klist[j] > kmax) kmax = klist[j];
    }
    if (initial_search_length) return NM_EARG;  // search.py:137-140
  } else {
    return NM_EARG;
  }
  if (kmax >= 0xFFFFFFF0u) return NM_EARG;
  if ((out_width == 1 && kmax > 0xFF) || (out_width == 2 && kmax > 0xFFFF)) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  if (!ws->timed_h2d) NM_CUDA(cudaEventRecord(ws->ev[EV_H2D0], ws->stream));
  if (mode == NM_MODE_LIST) {
    if (nk > ws->klist_cap) {
      NM_CUDA(cudaStreamSynchronize(ws->stream));
      cudaFree(ws->d_klist);
      ws->d_klist = nullptr;
      NM_CUDA(cudaMalloc((void **)&ws->d_klist, (size_t)nk * 4));
      ws->klist_cap = nk;
    }
    NM_CUDA(cudaMemcpyAsync(ws->d_klist, klist, (size_t)nk * 4, cudaMemcpyHostToDevice, ws->stream));
  }
  if (ws->multi_ix) algo = NM_ALGO_TRAJECTORY;  // the summed count is searched on the reference's own trajectory
  nm::SegParams &P = ws->P;
  P.seg = ws->seg_ext ? ws->seg_ext : ws->d_seg;
  P.codes = ws->d_codes;
  P.pk = ws->d_pk;
  P.barmask = ws->d_barmask;
  P.deepmap = ws->d_deepmap;
  P.tile_first = ws->d_tile_first;
  P.G = ws->d_G;
  P.R = ws->d_R;
  P.out = ws->d_out;
  P.klist = ws->d_klist;
  P.counters = ws->d_counters;
  P.seg_len = (uint32_t)ws->seg_len;
  P.num_kmers = (uint32_t)num_kmers;
  P.kmin = kmin;
  // No k-mer is longer than the segment, so the chains never need to go deeper than that: with the bound clamped the
  // sums the certificate pass forms (kmax + group size, L + dt) stay far below 2^32 for any --search-range upper end.
  // (The nominal lengths of a k list, the output width and the reference's query trajectory use the values given.)
  P.kmax = (uint32_t)std::min<uint64_t>(kmax, std::max<uint64_t>(ws->seg_len, kmin));
  P.nk = nk;
  P.initial = initial_search_length;
  P.mode = mode;
  P.use_rc = use_rc ? 1 : 0;
  P.out_width = out_width;
  P.fast = 0;
  {
    // chains start from the deeper derived table when every queried length reaches its depth; the estimate of the
    // reference's LF steps (NM_OPT_REF_LF) needs thresholds exact down to the index's own seed length
    const nm::DevIndex &dv = ws->ix->dev;
    // ... or, for a RANGE that starts below its depth but reaches it (12:255 on the 16-letter table), when the lengths
    // below the depth are settled separately (P.keff, resolve_thread): NM_OPT_SUBDEPTH 1 switches that off
    const bool sub = dv.seedx && mode == NM_MODE_RANGE && kmin < dv.seedx_len && kmax >= dv.seedx_len &&
                     kmin >= dv.seed_len && ws->opt_subdepth == 0u && ws->opt_ref_lf != 2u && algo == NM_ALGO_AUTO;
    const bool deep = dv.seedx && (kmin >= dv.seedx_len || sub) && ws->opt_ref_lf != 2u;
    P.want_ref_lf = (ws->opt_ref_lf && !deep) ? 1 : 0;
    P.sl = deep ? dv.seedx_len : dv.seed_len;
    P.seedtab = deep ? nullptr : dv.seed32;
    P.seedpair = deep ? dv.seedx : nullptr;
    P.keff = sub ? dv.seedx_len : kmin;
    P.sub = nullptr;
    if (sub) {
      if (!ws->d_sub) NM_CUDA(cudaMalloc((void **)&ws->d_sub, ws->max_seg_len + 16));
      P.sub = ws->d_sub;
    }
    const uint32_t sl = P.sl;
    // (profiles/r2_sweep_*.jsonl: with the sparse first pass a shallow cap pays -- the chains that do not end within a few
    //  steps are inside repeats, where the certificate pass settles 20 positions with one chain)
    P.dcap = ws->opt_dcap ? ws->opt_dcap : (deep ? sl + 3 : sl + 12);
    if (P.dcap <= sl) P.dcap = sl + 1;
    P.gsize = ws->opt_gsize ? ws->opt_gsize : 40;  // profiles/r2_sweep_f_gsize.jsonl (round 1, dense first pass: 20)
    P.own_chains = ws->opt_own_chains == ~0u ? nm::kCertOwnChains : ws->opt_own_chains;
    P.cert_runs = ws->opt_cert_runs ? ws->opt_cert_runs : nm::kCertRuns;
    // only when passes 2 and 3 will run (run_search: deep_passes).  Off by default: with 2 it removes 45 % of the first
    // pass's LF steps on the 1 Gbp bench genome, but those are the cheap ones (neighbouring positions inside a repeat
    // walk the same rank blocks, L2 hits) and the later passes pay more than it saves (8.45 -> 8.03 G positions/s).
    P.stall = P.dcap < kmax ? (ws->opt_stall == ~0u ? 0u : ws->opt_stall) : 0u;
    // sparse first pass (nm_core.cuh): one window per `stride` positions, the rest settled by bounds spread from it
    P.stride = nm::sparse_auto_stride(P.keff, sl, deep, ws->opt_stride);
  }
  ws->algo_req = algo;
  ws->have_search = true;
  ws->timed_d2h = false;
  ws->pending_out = nullptr;
  return run_search(ws, algo);
}

int nm_workspace_download(nm_workspace *ws, void *out) {
  if (!ws || !ws->have_search || (!out && ws->P.num_kmers)) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  NM_CUDA(cudaEventRecord(ws->ev[EV_D2H0], ws->stream));
  if (ws->P.num_kmers)
    NM_CUDA(cudaMemcpyAsync(out, ws->d_out, (size_t)ws->P.num_kmers * ws->P.out_width, cudaMemcpyDeviceToHost,
                            ws->stream));
  NM_CUDA(cudaEventRecord(ws->ev[EV_D2H1], ws->stream));
  ws->timed_d2h = true;
  ws->pending_out = out;
  return NM_OK;
}

int nm_workspace_sync(nm_workspace *ws, nm_stats *stats) {
  if (!ws) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  NM_CUDA(cudaStreamSynchronize(ws->stream));
  if (!ws->have_search) return NM_OK;
  bool replayed = false;
  float ms[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  auto el = [&](int a, int b, float &dst) { cudaEventElapsedTime(&dst, ws->ev[a], ws->ev[b]); };
  if (ws->timed_h2d) el(EV_H2D0, EV_H2D1, ms[0]);
  el(EV_H2D1, EV_PREP1, ms[1]);
  el(EV_PREP1, EV_SEARCH1, ms[2]);
  el(EV_PREP1, EV_PASS1, ms[7]);
  el(EV_PASS1, EV_CERT1, ms[8]);
  el(EV_CERT1, EV_SEARCH1, ms[9]);
  el(EV_SEARCH1, EV_GENERIC1, ms[3]);
  el(EV_GENERIC1, EV_RESOLVE1, ms[4]);
  if (ws->timed_d2h) el(EV_D2H0, EV_D2H1, ms[5]);
  if (ws->timed_d2h) el(EV_H2D0, EV_D2H1, ms[6]); else el(EV_H2D0, EV_RESOLVE1, ms[6]);
  const uint32_t free_launches = ws->launches;
  const unsigned long long lf_pass1 = ws->h_counters[nm::CTR_LF_PASS1],
                           seeds_pass1 = ws->h_counters[nm::CTR_SEEDS_PASS1],
                           seeds = ws->h_counters[nm::CTR_SEEDS] + seeds_pass1;
  unsigned long long free_lf = ws->h_counters[nm::CTR_LF] + lf_pass1;
  if (ws->algo_req == NM_ALGO_AUTO && ws->h_counters[nm::CTR_ANOMALY]) {
    // A substring of the segment is absent from the index (sequence / index mismatch).  The
    // reference reports the first zero count on ITS query trajectory: replay that trajectory.
    void *out = ws->pending_out;
    const bool d2h = ws->timed_d2h;
    int rc = run_search(ws, NM_ALGO_TRAJECTORY);
    if (rc != NM_OK) return rc;
    if (d2h && ws->P.num_kmers)
      NM_CUDA(cudaMemcpyAsync(out, ws->d_out, (size_t)ws->P.num_kmers * ws->P.out_width, cudaMemcpyDeviceToHost,
                              ws->stream));
    NM_CUDA(cudaStreamSynchronize(ws->stream));
    replayed = true;
  }
  ws->timed_h2d = false;
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->positions = ws->P.num_kmers;
    stats->lf_steps = ws->h_counters[nm::CTR_LF] + (replayed ? free_lf : lf_pass1);
    stats->lf_steps_pass1 = lf_pass1;
    stats->seed_lookups = seeds;
    stats->seed_lookups_pass1 = seeds_pass1;
    stats->launches = ws->launches + (replayed ? free_launches : 0u);
    stats->seed_depth = ws->P.sl;
    stats->stride = ws->P.stride;
    if (ws->opt_out_stats) {
      stats->out_stats = 1;
      stats->unique_positions = ws->h_counters[nm::CTR_UNIQUE];
      stats->ambiguous_positions = ws->h_counters[nm::CTR_AMBIG];
      stats->min_unique = ws->h_counters[nm::CTR_MINLEN] == ~0ull ? 0u : (uint32_t)ws->h_counters[nm::CTR_MINLEN];
      stats->max_unique = (uint32_t)ws->h_counters[nm::CTR_MAXLEN];
    }
    stats->short_counts = (uint32_t)std::min<unsigned long long>(ws->h_counters[nm::CTR_SHORT], 0xFFFFFFFFull);
    stats->ms_pass1 = ms[7];
    stats->ms_cert = ms[8];
    stats->ms_pass3 = ms[9];
    stats->queries = ws->h_counters[nm::CTR_QUERIES];
    stats->generic_searches = ws->h_counters[nm::CTR_GENERIC];
    stats->ref_lf_steps = ws->h_counters[nm::CTR_REF_LF];
    const unsigned long long key = ws->h_counters[nm::CTR_ZERO_KEY];
    stats->zero_pos = -1;
    stats->zero_len = 0;
    stats->zero_round = -1;
    if (key != ~0ull) {
      stats->zero_pos = (int64_t)(key & 0xFFFFFFFFull);
      stats->zero_round = (int64_t)(key >> 32);
      uint32_t zl = 0;
      NM_CUDA(cudaMemcpy(&zl, ws->d_G + stats->zero_pos, 4, cudaMemcpyDeviceToHost));
      stats->zero_len = zl;
    }
    stats->algo_used = replayed ? NM_ALGO_TRAJECTORY : (uint32_t)ws->algo_req;
    stats->replayed = replayed ? 1u : 0u;
    stats->ms_h2d = ms[0];
    stats->ms_prep = ms[1];
    stats->ms_search = ms[2];
    stats->ms_generic = ms[3];
    stats->ms_resolve = ms[4];
    stats->ms_d2h = ms[5];
    stats->ms_total = ms[6];
  }
  return NM_OK;
}

int nm_workspace_option(nm_workspace *ws, int option, uint64_t value) {
  if (!ws) return NM_EARG;
  switch (option) {
    case NM_OPT_REF_LF: ws->opt_ref_lf = value ? 2u : 0u; return NM_OK;  // 2 = asked for explicitly
    case NM_OPT_DCAP: ws->opt_dcap = (uint32_t)value; return NM_OK;
    case NM_OPT_PASS1: if (value > 2) return NM_EARG; ws->opt_pass1 = (uint32_t)value; return NM_OK;
    case NM_OPT_GSIZE: if (value > 4096) return NM_EARG; ws->opt_gsize = (uint32_t)value; return NM_OK;
    case NM_OPT_OWN_CHAINS: if (value > 64) return NM_EARG; ws->opt_own_chains = (uint32_t)value; return NM_OK;
    case NM_OPT_STALL: if (value > 64) return NM_EARG; ws->opt_stall = (uint32_t)value; return NM_OK;
    case NM_OPT_CERT_RUNS: if (value > 64) return NM_EARG; ws->opt_cert_runs = (uint32_t)value; return NM_OK;
    case NM_OPT_GSIZE2: if (value > 65536) return NM_EARG; ws->opt_gsize2 = (uint32_t)value; return NM_OK;
    case NM_OPT_STRICT:
      if (value > (1u << 20)) return NM_EARG;
      ws->opt_strict = value == 1 ? 1024u : (uint32_t)value;   // 1 = the default block of 1024 start positions
      return NM_OK;
    case NM_OPT_SUBDEPTH: ws->opt_subdepth = value ? 1u : 0u; return NM_OK;
    case NM_OPT_OUT_STATS: ws->opt_out_stats = value ? 1u : 0u; return NM_OK;
    case NM_OPT_STRIDE: if (value > nm::kMaxStride) return NM_EARG; ws->opt_stride = (uint32_t)value; return NM_OK;
    default: return NM_EARG;
  }
}

int nm_workspace_checksum(nm_workspace *ws, uint64_t *sum) {
  if (!ws || !sum || !ws->have_search) return NM_EARG;
  NM_CUDA(cudaSetDevice(ws->ix->device));
  unsigned long long *d = nullptr;
  NM_CUDA(cudaMalloc((void **)&d, 8));
  cudaMemsetAsync(d, 0, 8, ws->stream);
  nm::launch_checksum(ws->d_out, ws->P.out_width, ws->P.num_kmers, d, ws->stream);
  unsigned long long v = 0;
  cudaError_t e = cudaMemcpyAsync(&v, d, 8, cudaMemcpyDeviceToHost, ws->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ws->stream);
  cudaFree(d);
  if (e != cudaSuccess) return cuda_fail(e, "checksum");
  *sum = v;
  return NM_OK;
}

int nm_min_unique(nm_index *ix, const uint8_t *seg, uint64_t seg_len, uint64_t num_kmers, int mode,
                  const uint32_t *klist, uint32_t nk, uint32_t initial_search_length, int use_rc, int algo,
                  int out_width, void *out, nm_stats *stats) {
  if (!ix) return NM_EARG;
  if (seg_len == 0 || num_kmers == 0) {
    if (stats) { memset(stats, 0, sizeof *stats); stats->zero_pos = -1; stats->zero_round = -1; }
    return num_kmers == 0 ? NM_OK : NM_EARG;
  }
  std::lock_guard<std::mutex> lock(ix->mu);
  if (!ix->default_ws || ix->default_ws->max_seg_len < seg_len) {
    if (ix->default_ws) nm_workspace_destroy(ix->default_ws);
    ix->default_ws = nullptr;
    uint64_t cap = seg_len < (1u << 20) ? (1u << 20) : seg_len + seg_len / 8;
    if (cap >= 0xFFFFFF00ull) cap = seg_len;
    const int rc = nm_workspace_create(ix, cap, nullptr, &ix->default_ws);
    if (rc != NM_OK) return rc;
  }
  nm_workspace *ws = ix->default_ws;
  ws->opt_ref_lf = stats ? 1u : 0u;  // 1 = computed when the search runs on the index's own seed table anyway
  int rc = nm_workspace_upload(ws, seg, seg_len);
  if (rc == NM_OK) rc = nm_workspace_search(ws, num_kmers, mode, klist, nk, initial_search_length, use_rc, algo, out_width);
  if (rc == NM_OK) rc = nm_workspace_download(ws, out);
  if (rc == NM_OK) rc = nm_workspace_sync(ws, stats);
  return rc;
}

void *nm_host_alloc(uint64_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}
void nm_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

int nm_probe_random_access(int device, uint64_t bytes, uint32_t granule, uint64_t chains, uint32_t steps,
                           uint32_t mlp, int dependent, double *loads_per_second, float *ms_out) {
  if ((granule != 32 && granule != 64 && granule != 128) || bytes < granule || !chains || !steps) return NM_EARG;
  if ((bytes & (bytes - 1)) || bytes / granule > 0x100000000ull) return NM_EARG;  // power of two, 32-bit granule index
  NM_CUDA(cudaSetDevice(device));
  uint8_t *buf = nullptr;
  unsigned long long *sink = nullptr;
  cudaEvent_t a = nullptr, b = nullptr;
  auto body = [&]() -> int {
    NM_CUDA(cudaMalloc((void **)&buf, bytes));
    NM_CUDA(cudaMalloc((void **)&sink, 8));
    NM_CUDA(cudaMemset(buf, 0x5a, bytes));
    NM_CUDA(cudaMemset(sink, 0, 8));
    NM_CUDA(cudaEventCreate(&a));
    NM_CUDA(cudaEventCreate(&b));
    const uint64_t ng = bytes / granule;
    nm::launch_probe(buf, ng, granule, chains, steps < 4 ? steps : 4, mlp, dependent != 0, sink, 0);  // warm-up
    NM_CUDA(cudaDeviceSynchronize());
    NM_CUDA(cudaEventRecord(a, 0));
    nm::launch_probe(buf, ng, granule, chains, steps, mlp, dependent != 0, sink, 0);
    NM_CUDA(cudaEventRecord(b, 0));
    NM_CUDA(cudaEventSynchronize(b));
    NM_CUDA(cudaGetLastError());
    float ms = 0;
    NM_CUDA(cudaEventElapsedTime(&ms, a, b));
    if (ms_out) *ms_out = ms;
    if (loads_per_second) *loads_per_second = (double)chains * steps * (mlp >= 8 ? 8 : mlp >= 4 ? 4 : mlp >= 2 ? 2 : 1) / (ms * 1e-3);
    return NM_OK;
  };
  const int rc = body();
  if (a) cudaEventDestroy(a);
  if (b) cudaEventDestroy(b);
  cudaFree(buf);
  cudaFree(sink);
  return rc;
}

// ---------------------------------------------------------------------------------------------
// nm_count_from_sequence without any allocation (include/newmap_b200.h,
// no counterpart in the reference).  The scratch nm_count_from_sequence cudaMallocs per call -- device copies of
// the sequence, the query lists, the counts and the counters -- lives in one file-scope device array.  In HOST code
// the array's name is only a stand-in for the symbol, not a device address: the address is asked for once per
// device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernel gets.
// ---------------------------------------------------------------------------------------------
}  // extern "C"

static __device__ __align__(256) unsigned char g_noalloc_scratch[NM_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;                       // one call at a time per process: every call shares the array
unsigned char *g_noalloc_ptr[kMaxDevices];     // device address of g_noalloc_scratch on each device, once known

int noalloc_scratch(int device, unsigned char **out) {
  if (device < 0 || device >= kMaxDevices) return NM_EARG;
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    NM_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // (the caller has made `device` current)
    g_noalloc_ptr[device] = static_cast<unsigned char *>(p);
  }
  *out = g_noalloc_ptr[device];
  return NM_OK;
}
}  // namespace

extern "C" {

int nm_count_from_sequence_noalloc(nm_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                                   const uint64_t *len, uint64_t n, int reverse_complement, uint32_t *counts,
                                   uint64_t *lf_steps) {
  if (!ix || (n && (!idx || !len || !counts)) || (seq_len && !seq)) return NM_EARG;
  for (uint64_t q = 0; q < n; ++q)  // newmap-count.c:63-68,184-190
    if (len[q] == 0 || idx[q] > seq_len || len[q] > seq_len - idx[q]) return NM_EARG;
  if (n == 0) return NM_OK;
  // layout: [counters, 256 B][idx, 8 B x n64][len, 8 B x n64][counts, 4 B x n64][sequence + 16]
  if (n > NM_NOALLOC_SCRATCH_BYTES || seq_len > NM_NOALLOC_SCRATCH_BYTES) return NM_ELIMIT;
  const uint64_t n64 = (n + 63) / 64 * 64;
  const uint64_t seq_bytes = (seq_len + 16 + 255) / 256 * 256;
  if (256 + 20 * n64 + seq_bytes > NM_NOALLOC_SCRATCH_BYTES) return NM_ELIMIT;
  std::lock_guard<std::mutex> all(g_noalloc_mu);
  std::lock_guard<std::mutex> lock(ix->mu);
  NM_CUDA(cudaSetDevice(ix->device));
  unsigned char *base = nullptr;
  const int rc = noalloc_scratch(ix->device, &base);
  if (rc != NM_OK) return rc;
  unsigned long long *d_ctr = reinterpret_cast<unsigned long long *>(base);
  uint64_t *d_idx = reinterpret_cast<uint64_t *>(base + 256);
  uint64_t *d_len = d_idx + n64;
  uint32_t *d_counts = reinterpret_cast<uint32_t *>(d_len + n64);
  uint8_t *d_seq = reinterpret_cast<uint8_t *>(d_counts + n64);
  static_assert(nm::CTR_N * 8 <= 256, "the counters fit the first 256 bytes of the scratch");
  NM_CUDA(cudaMemsetAsync(d_ctr, 0, 256, 0));   // the array keeps the last call's contents: the counters start at zero
  NM_CUDA(cudaMemcpyAsync(d_seq, seq, seq_len, cudaMemcpyHostToDevice, 0));
  NM_CUDA(cudaMemcpyAsync(d_idx, idx, n * 8, cudaMemcpyHostToDevice, 0));
  NM_CUDA(cudaMemcpyAsync(d_len, len, n * 8, cudaMemcpyHostToDevice, 0));
  nm::launch_count(ix->dev, d_seq, d_idx, d_len, n, reverse_complement != 0, d_counts, d_ctr, 0);
  NM_CUDA(cudaGetLastError());
  NM_CUDA(cudaMemcpy(counts, d_counts, n * 4, cudaMemcpyDeviceToHost));
  if (lf_steps) {
    unsigned long long v = 0;
    NM_CUDA(cudaMemcpy(&v, d_ctr + nm::CTR_LF, 8, cudaMemcpyDeviceToHost));
    *lf_steps += v;
  }
  return NM_OK;
}

}  // extern "C"
