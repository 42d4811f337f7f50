// tests/hostemu/hostemu.cpp -- TEST INFRASTRUCTURE ONLY.
// Compiles the per-thread kernel bodies of newmap_b200/csrc/nm_core.cuh with the host compiler
// and runs them serially (thread id = loop index) in the same order nm_lib.cu launches them, so
// the algorithmic logic of the CUDA path can be checked against the oracle in the build
// container, which has no GPU.  The product never loads this library.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../newmap_b200/csrc/nm_core.cuh"

using namespace nm;

struct EmuIndex {
  std::vector<uint8_t> native;
  std::vector<uint32_t> rank32;
  std::vector<SeedEntry32> seed32;
  std::vector<uint64_t> seed64;
  std::vector<SeedEntryPair> seedx;  // deeper derived both-strand table (emu_index_set_seed_depth)
  uint32_t c32[8] = {0};
  DevIndex dev{};
  bool has_rank32 = false;
};

extern "C" {

EmuIndex *emu_index_open(const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) return nullptr;
  uint8_t hdr[30];
  if (fread(hdr, 1, 30, f) != 30) { fclose(f); return nullptr; }
  uint64_t bwt_len;
  memcpy(&bwt_len, hdr + 22, 8);
  const uint32_t sl = hdr[19];
  EmuIndex *ix = new EmuIndex();
  const uint64_t nb = 1 + (bwt_len - 1) / 256;
  ix->native.assign((nb + 1) * 160, 0);
  if (fread(ix->native.data(), 160, nb, f) != nb) { fclose(f); delete ix; return nullptr; }
  uint64_t prefix[6];
  if (fread(prefix, 8, 6, f) != 6) { fclose(f); delete ix; return nullptr; }
  const uint64_t entries = 1ull << (2 * sl);
  ix->seed64.resize(entries * 2);
  if (fread(ix->seed64.data(), 16, entries, f) != entries) { fclose(f); delete ix; return nullptr; }
  fclose(f);
  uint64_t *pad = reinterpret_cast<uint64_t *>(ix->native.data() + nb * 160);
  for (int l = 0; l < 5; ++l) pad[12 + l] = prefix[l + 1] - prefix[l];
  pad[17] = 1;
  ix->dev.native = ix->native.data();
  ix->dev.seed64 = ix->seed64.data();
  for (int l = 0; l < 6; ++l) ix->dev.C[l] = prefix[l];
  ix->dev.bwt_len = bwt_len;
  ix->dev.seed_len = sl;
  const uint64_t n64 = (nb + 1) * 4;
  ix->rank32.resize(n64 * 8);
  for (uint64_t t = 0; t < n64; ++t)
    transcode_thread(ix->native.data(), nb, bwt_len, ix->rank32.data(), t, (uint32_t)prefix[0], (uint32_t)prefix[1], (uint32_t)prefix[2], (uint32_t)prefix[3]);
  ix->seed32.resize(entries);
  for (uint64_t t = 0; t < entries; ++t) seed_transcode_thread(ix->seed64.data(), ix->seed32.data(), t);
  for (int l = 0; l < 6; ++l) ix->c32[l] = (uint32_t)prefix[l];
  ix->dev.c32 = ix->c32;
  ix->dev.rank32 = ix->rank32.data();
  ix->dev.seed32 = ix->seed32.data();
  ix->has_rank32 = true;
  return ix;
}

void emu_index_close(EmuIndex *ix) { delete ix; }

uint32_t emu_rc_seed_key(uint32_t key, uint32_t sl) { return rc_seed_key(key, sl); }

// the level-by-level construction nm_lib.cu set_seed_depth runs on the device
int emu_index_set_seed_depth(EmuIndex *ix, uint32_t depth) {
  ix->seedx.clear();
  ix->dev.seedx = nullptr;
  ix->dev.seedx_len = 0;
  const uint32_t sl = ix->dev.seed_len;
  if (depth <= sl) return 0;
  if (depth > 16) return -8;
  std::vector<SeedEntry32> cur = ix->seed32;
  for (uint32_t d = sl; d < depth; ++d) {
    std::vector<SeedEntry32> next((size_t)1 << (2 * (d + 1)));
    for (uint64_t key = 0; key < (1ull << (2 * d)); ++key) seed_extend_thread(ix->dev, ix->c32, cur.data(), d, next.data(), key);
    cur.swap(next);
  }
  ix->seedx.resize((size_t)1 << (2 * depth));
  for (uint64_t key = 0; key < (1ull << (2 * depth)); ++key) seed_pair_thread(cur.data(), depth, ix->seedx.data(), key);
  ix->dev.seedx = ix->seedx.data();
  ix->dev.seedx_len = depth;
  return 0;
}

int emu_count(EmuIndex *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx, const uint64_t *len,
              uint64_t n, int rc, uint32_t *counts, uint64_t *lf) {
  uint32_t steps = 0;
  for (uint64_t q = 0; q < n; ++q) {
    if (len[q] == 0 || idx[q] + len[q] > seq_len) return -6;
    counts[q] = rc ? count_one<true>(ix->dev, seq, idx[q], len[q], steps) : count_one<false>(ix->dev, seq, idx[q], len[q], steps);
  }
  if (lf) *lf += steps;
  return 0;
}

// stats: [0] lf, [1] queries, [2] generic searches, [3] anomaly, [4] zero pos (-1), [5] zero len, [6] zero round, [7] ref lf,
//        [8] seed-table lookups, [9] stride of the sparse first pass
int emu_min_unique(EmuIndex *ix, const uint8_t *seg, uint32_t seg_len, uint32_t num_kmers, int mode,
                   const uint32_t *klist, uint32_t nk, uint32_t initial, int use_rc, int algo, int out_width,
                   void *out, int64_t *stats, uint32_t dcap, uint32_t gsize, uint32_t gsize2, uint32_t stall, uint32_t stride) {
  std::vector<uint8_t> codes(seg_len + 16);
  std::vector<uint32_t> pk(seg_len / 16 + 2), tf((seg_len + kTile - 1) / kTile + 1), G(seg_len + 2), R(seg_len + 1);
  std::vector<uint64_t> barmask((seg_len + kTile - 1) / kTile + 1);
  std::vector<uint32_t> deepmap(seg_len / 32 + 4, 0u);   // cleared by k_prep on the device
  unsigned long long ctr[CTR_N] = {0, 0, 0, 0, ~0ull, 0, 0, 0};
  SegParams P{};
  P.deepmap = (gsize + dcap) % 3 == 2 ? nullptr : deepmap.data();   // (without the map passes 2 and 3 scan G / R: same results)
  P.seg = seg; P.codes = codes.data(); P.pk = pk.data(); P.barmask = barmask.data(); P.tile_first = tf.data();
  P.G = G.data(); P.R = R.data(); P.out = out; P.klist = klist; P.counters = ctr;
  P.seg_len = seg_len; P.num_kmers = num_kmers; P.nk = nk; P.initial = initial; P.mode = mode;
  P.use_rc = use_rc; P.out_width = out_width;
  P.gsize = gsize ? gsize : 32;
  P.own_chains = kCertOwnChains;
  P.cert_runs = 1 + (gsize + dcap) % 4;
  if (mode == 0) { P.kmin = klist[0]; P.kmax = klist[1]; }
  else { P.kmin = P.kmax = klist[0]; for (uint32_t j = 0; j < nk; ++j) { if (klist[j] < P.kmin) P.kmin = klist[j]; if (klist[j] > P.kmax) P.kmax = klist[j]; } }
  // nm_lib.cu nm_workspace_search: the deeper table when every queried length reaches its depth (the estimate of the
  // reference's LF steps is then off, so it is only requested on the index's own table)
  // ... or (nm_lib.cu) a range that starts below its depth but reaches it, the lengths below settled by resolve_thread
  const bool sub = ix->dev.seedx && mode == 0 && P.kmin < ix->dev.seedx_len && P.kmax >= ix->dev.seedx_len &&
                   P.kmin >= ix->dev.seed_len && algo == 0 && !(stride & 0x40000000u);
  const bool strict = stride != 0xFFFFFFFFu && (stride & 0x20000000u);   // bit 29 of the stride argument: NM_OPT_STRICT
  if (stride != 0xFFFFFFFFu) stride &= ~0x60000000u;     // bit 30 of the stride argument: NM_OPT_SUBDEPTH 1
  const bool deep = ix->dev.seedx && (P.kmin >= ix->dev.seedx_len || sub);
  std::vector<uint8_t> subv(sub ? seg_len + 16 : 0);
  P.keff = sub ? ix->dev.seedx_len : P.kmin;
  P.sub = sub ? subv.data() : nullptr;
  P.sl = deep ? ix->dev.seedx_len : ix->dev.seed_len;
  P.seedtab = deep ? nullptr : ix->dev.seed32;
  P.seedpair = deep ? ix->dev.seedx : nullptr;
  P.want_ref_lf = deep ? 0 : 1;
  P.dcap = dcap ? dcap : P.sl + 12; if (P.dcap <= P.sl) P.dcap = P.sl + 1;
  P.stall = P.dcap < P.kmax ? stall : 0;
  // nm_lib.cu: 0 = the widest stride the slack allows; 0xFFFFFFFF = the dense thread-per-position first pass
  P.stride = stride == 0xFFFFFFFFu ? 1u : sparse_auto_stride(P.keff, P.sl, deep, stride);
  for (uint32_t p = 0; p < seg_len; ++p) prep_thread(ix->dev, P, p);
  for (uint32_t w = 0; w < seg_len / 16 + 2; ++w) pack_thread(P, w);
  const uint32_t ntiles = (seg_len + kTile - 1) / kTile;
  for (uint32_t t = 0; t < ntiles; ++t) nb_tile_thread(P, t);
  uint32_t steps = 0, nq = 0, ns = 0, seeds = 0;
  uint64_t ref_lf = 0;
  if (algo == 1) {
    for (uint32_t i = 0; i < num_kmers; ++i) trajectory_thread(ix->dev, P, i, steps, nq);
  } else {
    const bool fast = (algo != 2) && ix->has_rank32 && P.kmin >= ix->dev.seed_len;
    P.fast = fast;
    if (strict && num_kmers) {   // nm_lib.cu run_search: presence chains, one per block of start positions
      const uint32_t B = 64 + (gsize + dcap) % 100, nblocks = (num_kmers + B - 1) / B;
      uint32_t ps = 0;
      for (uint32_t t = 0; t < nblocks; ++t) presence_thread(ix->dev, ix->c32, P, t, B, ps);
    }
    if (fast) {
      if (sub)   // k_pass1_sparse phase A: which starts the table does not rule out for the lengths below its depth
        for (uint32_t q = 0; q < seg_len; ++q) subv[q] = !(codes[q] & kSeedBad) && subdepth_candidate(P, q);
      if (deep && stride != 0xFFFFFFFFu) sparse_pass1_serial(ix->dev, ix->c32, P, steps, seeds);   // k_pass1_sparse
      else for (uint32_t p = 0; p <= seg_len; ++p) free_chain_thread(ix->dev, ix->c32, P, p, true, steps, seeds);
      if (P.dcap < P.kmax) {
        uint32_t s2 = 0;
        if (gsize2 > P.gsize) {  // nm_lib.cu run_search: coarse round (anchor chains at fixed positions, containment only)
          SegParams Pc = P;
          Pc.gsize = gsize2;
          const uint32_t nc = (seg_len + 1 + Pc.gsize - 1) / Pc.gsize;
          std::vector<CoarseGroup> groups(2 * (size_t)nc);
          for (uint32_t t = 0; t < nc; ++t)
            for (uint32_t part = 0; part < 2; ++part)
              if (part == 0 || P.use_rc) groups[2 * t + part] = coarse_anchor_thread(ix->dev, ix->c32, Pc, t, part, s2, seeds);
          for (uint32_t p = 0; p <= seg_len; ++p) coarse_settle_thread(Pc, groups.data(), p);
        }
        const uint32_t ngroups = (seg_len + 1 + P.gsize - 1) / P.gsize;
        for (uint32_t t = 0; t < ngroups; ++t) free_cert_thread(ix->dev, ix->c32, P, t, s2, seeds);
        ns = s2;  // reported in stats[2]: LF steps of the certificate pass
        steps += s2;
        for (uint32_t p = 0; p <= seg_len; ++p) free_chain_thread(ix->dev, ix->c32, P, p, false, steps, seeds);
      }
    } else {
      for (uint32_t p = 0; p <= seg_len; ++p) free_generic_thread(ix->dev, P, p, true, steps, ns);
    }
    for (uint32_t i = 0; i < num_kmers; ++i) resolve_thread(ix->dev, P, i, ref_lf);
    if (P.keff > P.kmin) for (uint32_t i = 0; i < num_kmers; ++i) subdepth_thread(ix->dev, ix->c32, P, i, steps);
  }
  if (stats) {
    stats[0] = steps; stats[1] = nq; stats[2] = ns; stats[3] = (int64_t)ctr[CTR_ANOMALY];
    stats[4] = -1; stats[5] = 0; stats[6] = -1; stats[7] = (int64_t)ref_lf; stats[8] = seeds; stats[9] = P.stride;
    if (ctr[CTR_ZERO_KEY] != ~0ull) {
      stats[4] = (int64_t)(ctr[CTR_ZERO_KEY] & 0xFFFFFFFFull);
      stats[6] = (int64_t)(ctr[CTR_ZERO_KEY] >> 32);
      stats[5] = G[stats[4]];
    }
  }
  return 0;
}

// nm_workspace_multi + nm_workspace_search: several sequences x several indexes on the reference's trajectory.
// stats: [0] lf, [1] queries, [2] short counts, [4] zero pos (-1), [5] zero len, [6] zero round
int emu_min_unique_multi(EmuIndex *const *ixs, uint32_t n_ix, const uint8_t *const *segs, uint32_t n_seq, uint32_t seg_len,
                         uint32_t num_kmers, int mode, const uint32_t *klist, uint32_t nk, uint32_t initial, int use_rc,
                         int out_width, void *out, int64_t *stats) {
  if (n_ix == 0 || n_ix > kMaxMulti || n_seq == 0 || n_seq > kMaxMulti) return -1;
  std::vector<uint8_t> codes(seg_len + 16);
  std::vector<uint32_t> pk(seg_len / 16 + 2), tf((seg_len + kTile - 1) / kTile + 1), G(seg_len + 2), R(seg_len + 1);
  std::vector<uint64_t> barmask((seg_len + kTile - 1) / kTile + 1);
  unsigned long long ctr[CTR_N] = {0, 0, 0, 0, ~0ull, 0, 0, 0};
  SegParams P{};
  P.seg = segs[0]; P.codes = codes.data(); P.pk = pk.data(); P.barmask = barmask.data(); P.tile_first = tf.data();
  P.G = G.data(); P.R = R.data(); P.out = out; P.klist = klist; P.counters = ctr;
  P.seg_len = seg_len; P.num_kmers = num_kmers; P.nk = nk; P.initial = initial; P.mode = mode;
  P.use_rc = use_rc; P.out_width = out_width; P.gsize = 16;
  if (mode == 0) { P.kmin = klist[0]; P.kmax = klist[1]; }
  else { P.kmin = P.kmax = klist[0]; for (uint32_t j = 0; j < nk; ++j) { if (klist[j] < P.kmin) P.kmin = klist[j]; if (klist[j] > P.kmax) P.kmax = klist[j]; } }
  const DevIndex &d0 = ixs[0]->dev;
  P.sl = d0.seed_len; P.seedtab = d0.seed32; P.seedpair = nullptr; P.dcap = P.sl + 12; P.stride = 1; P.keff = P.kmin;
  for (uint32_t p = 0; p < seg_len; ++p) prep_thread(d0, P, p);
  for (uint32_t w = 0; w < seg_len / 16 + 2; ++w) pack_thread(P, w);
  const uint32_t ntiles = (seg_len + kTile - 1) / kTile;
  for (uint32_t t = 0; t < ntiles; ++t) nb_tile_thread(P, t);
  MultiSet M{};
  M.n_ix = n_ix; M.n_seq = n_seq;
  for (uint32_t x = 0; x < n_ix; ++x) M.ix[x] = ixs[x]->dev;
  for (uint32_t j = 0; j < n_seq; ++j) M.seg[j] = segs[j];
  uint32_t steps = 0, nq = 0, nshort = 0;
  for (uint32_t i = 0; i < num_kmers; ++i) trajectory_multi_thread(M, P, i, steps, nq, nshort);
  if (stats) {
    stats[0] = steps; stats[1] = nq; stats[2] = nshort; stats[3] = 0; stats[4] = -1; stats[5] = 0; stats[6] = -1; stats[7] = 0;
    if (ctr[CTR_ZERO_KEY] != ~0ull) {
      stats[4] = (int64_t)(ctr[CTR_ZERO_KEY] & 0xFFFFFFFFull);
      stats[6] = (int64_t)(ctr[CTR_ZERO_KEY] >> 32);
      stats[5] = G[stats[4]];
    }
  }
  return 0;
}

}  // extern "C"
