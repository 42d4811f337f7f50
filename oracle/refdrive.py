"""oracle/refdrive.py -- TEST / BASELINE INFRASTRUCTURE ONLY.

The reference's CPU implementation of the hot path, runnable where /root/reference is absent
(the GPU box): the reference's own compiled C (oracle/_ref/_c_newmap_count_kmers.so =
src/newmap-count.c + AvxWindowFmIndex, AVX2 + OpenMP, built by oracle/Makefile from the sources
where they lie) driven by a numpy restatement of the host loops of newmap/search.py:383-724
(binary_search / linear_search / get_kmer_counts), call for call: one forward and one
reverse-complement `count_kmers_from_sequence` per round, Python lists across the boundary, the
index re-read from disk inside every call -- exactly what `newmap search` does.  Used by
bench.py (`cpu_baseline`, `--impl reference`) and checked against the imported reference in
tests/test_oracle.py where /root/reference exists."""
from __future__ import annotations

import time

import numpy as np

from . import ora

COMPLEMENT = bytes.maketrans(b"ACGTacgt", b"TGCAtgca")
_ALLOWED = np.zeros(256, dtype=bool)
_ALLOWED[list(b"ACGTacgt")] = True


class Timer:
    def __init__(self):
        self.native = 0.0
        self.calls = 0


def _counts(index_path, seq: bytes, rcseq, idx: np.ndarray, lens: np.ndarray, nthreads: int, use_rc: bool,
            timer: Timer | None, mod=None) -> np.ndarray:
    mod = mod or ora.ref_modules()[0]
    il, ll = idx.tolist(), lens.tolist()
    t0 = time.perf_counter()
    total = np.array(mod.count_kmers_from_sequence(str(index_path), seq, il, ll, nthreads), dtype=np.uint32)
    if use_rc:
        ridx = (len(seq) - idx - lens).tolist()
        total += np.array(mod.count_kmers_from_sequence(str(index_path), rcseq, ridx, ll, nthreads), dtype=np.uint32)
    if timer is not None:
        timer.native += time.perf_counter() - t0
        timer.calls += 2 if use_rc else 1
    zero = np.nonzero(total == 0)[0]
    if zero.size:
        i, k = int(idx[zero[0]]), int(lens[zero[0]])
        raise ora.ZeroCount(ora.zero_count_message(seq[i:i + k].decode("utf-8")))
    return total


def search_segment(index_path, seg: bytes, num_kmers: int, kmer_lengths, is_range: bool, initial: int = 0,
                   use_rc: bool = True, nthreads: int = 1, timer: Timer | None = None, count_module=None) -> np.ndarray:
    """uint32 minimum unique lengths of one segment buffer with the reference's control flow.  `count_module`: what
    stands behind `from newmap._c_newmap_count_kmers import count_kmers_from_sequence` (search.py:12) -- the reference's
    own extension by default; the tests also pass the product's drop-in (newmap_b200.count) to prove the boundary."""
    buf = np.frombuffer(seg, dtype=np.uint8)
    rcseq = seg.translate(COMPLEMENT)[::-1] if use_rc else None
    amb = ~_ALLOWED[buf[:num_kmers]]
    finished = amb.copy()
    unique = np.zeros(num_kmers, dtype=np.int64)
    if is_range:
        kmin, kmax = min(kmer_lengths), max(kmer_lengths)
        upper = ora.upper_bounds(amb.astype(np.uint8), len(seg), kmax).astype(np.int64)
        lower = np.full(num_kmers, kmin, dtype=np.int64)
        query = (upper + lower) // 2
        if initial:
            query[query > initial] = initial
        finished[upper < kmin] = True
        while not finished.all():
            idx = np.nonzero(~finished)[0]
            q = query[idx]
            c = _counts(index_path, seg, rcseq, idx, q, nthreads, use_rc, timer, count_module)
            one, many = c == 1, c > 1
            u = unique[idx]
            unique[idx] = np.where(one & ((u == 0) | (q < u)), q, u)
            fin = finished[idx]
            fin = np.where(one, q == lower[idx], fin)
            fin = np.where(many, q == upper[idx], fin)
            finished[idx] = fin
            upper[idx] = np.where(one, q - 1, upper[idx])
            lower[idx] = np.where(many, q + 1, lower[idx])
            query[idx] = (upper[idx] + lower[idx]) // 2
    else:
        n_pos = np.where(buf == ord("N"))[0]
        for k in kmer_lengths:
            idx = np.nonzero(~finished)[0]
            if idx.size:
                # positions whose (end-truncated) k-mer holds an upper-case N are retired (search.py:593-596)
                nxt = np.searchsorted(n_pos, idx)
                has_n = np.zeros(idx.size, dtype=bool)
                ok = nxt < n_pos.size
                has_n[ok] = n_pos[nxt[ok]] < np.minimum(idx[ok] + k, len(seg))
                finished[idx[has_n]] = True
                idx = idx[~has_n]
            if idx.size == 0:
                break
            lens = np.minimum(k, len(seg) - idx)
            c = _counts(index_path, seg, rcseq, idx, lens, nthreads, use_rc, timer, count_module)
            u = unique[idx]
            unique[idx] = np.where((c == 1) & (u == 0), k, u)
            finished[idx] = c == 1
    return unique.astype(np.uint32)
