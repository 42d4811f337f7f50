// nm_kernels.cuh -- launch wrappers of the kernels in nm_kernels.cu (host-side declarations)
#pragma once
#include <cuda_runtime.h>

#include "nm_core.cuh"

namespace nm {

void launch_transcode(const uint8_t *native, uint64_t n_blocks, uint64_t bwt_len, uint32_t *rank32,
                      uint64_t n64_total, const uint64_t *prefix_sums, cudaStream_t st);
void launch_seed_transcode(const uint64_t *seed64, SeedEntry32 *seed32, uint64_t entries, cudaStream_t st);
void launch_seed_extend(const DevIndex &ix, const SeedEntry32 *src, uint32_t d, SeedEntry32 *dst, cudaStream_t st);
void launch_seed_pair(const SeedEntry32 *t, uint32_t d, SeedEntryPair *pair, cudaStream_t st);
void launch_count(const DevIndex &ix, const uint8_t *seq, const uint64_t *idx, const uint64_t *len, uint64_t n,
                  bool rc, uint32_t *counts, unsigned long long *counters, cudaStream_t st);
void launch_prep(const DevIndex &ix, const SegParams &P, cudaStream_t st);
void launch_trajectory(const DevIndex &ix, const SegParams &P, cudaStream_t st);
void launch_trajectory_multi(const MultiSet *M, const SegParams &P, cudaStream_t st);
void launch_free_chain(const DevIndex &ix, const SegParams &P, bool first, cudaStream_t st);  // pass 1 / pass 3
void launch_free_pass1(const DevIndex &ix, const SegParams &P, cudaStream_t st);                 // pass 1, sparse kernel
void launch_free_cert(const DevIndex &ix, const SegParams &P, cudaStream_t st);               // pass 2
void launch_coarse_anchor(const DevIndex &ix, const SegParams &P, CoarseGroup *groups, cudaStream_t st);
void launch_coarse_settle(const SegParams &P, const CoarseGroup *groups, cudaStream_t st);
void launch_free_generic(const DevIndex &ix, const SegParams &P, bool all, cudaStream_t st);
void launch_resolve(const DevIndex &ix, const SegParams &P, cudaStream_t st);
void launch_presence(const DevIndex &ix, const SegParams &P, uint32_t B, cudaStream_t st);
void launch_subdepth(const DevIndex &ix, const SegParams &P, cudaStream_t st);
void launch_out_stats(const SegParams &P, cudaStream_t st);
void launch_checksum(const void *out, int out_width, uint64_t n, unsigned long long *sum, cudaStream_t st);
void launch_probe(const uint8_t *buf, uint64_t n_granules_pow2, uint32_t granule, uint64_t chains, uint32_t steps,
                  uint32_t mlp, bool dependent, unsigned long long *sink, cudaStream_t st);

}  // namespace nm
