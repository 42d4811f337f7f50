"""Host logic of the partitioned ingest of a sharded run (newmap_b200/search.py: _line_start_at_or_after,
_partition_table, _plan_groups): every rank reads only its byte range of the FASTA (+ the stretch its last window
reaches into); from what the ranks found there -- emulated here with a host parser that lists records the way
nm_fasta_finish does -- each must lay out the same global window plan as a single reader of the whole file, own each
window exactly once, and find the window's bases at the local offset it computed."""
import os

import numpy as np
import pytest

from newmap_b200 import search as S
from newmap_b200.fasta import window_plan


def _device_like_records(text: bytes):
    """[(header line, offset into the compacted bases, bases)] + the compacted bases, as nm_fasta_finish lists them."""
    recs, bases, cur = [], bytearray(), None
    for line in text.split(b"\n"):
        if line[:1] in (b">", b";"):
            if cur is None and bases:
                recs.append([b"", 0, len(bases)])
            cur = [line, len(bases), 0]
            recs.append(cur)
        else:
            bases += line
            if cur is not None:
                cur[2] += len(line)
    if cur is None and bases:
        recs.append([b"", 0, len(bases)])
    return [tuple(r) for r in recs], bytes(bases)


def _rank_view(path, text, rank, world, seqlen):
    """What _partitioned_ingest computes before the exchange, with the host parser in place of the GPU's."""
    size = len(text)
    fd = os.open(path, os.O_RDONLY)
    try:
        lo = S._line_start_at_or_after(fd, size * rank // world, size)
        hi = S._line_start_at_or_after(fd, size * (rank + 1) // world, size)
    finally:
        os.close(fd)
    tail_len = min(size - hi, seqlen + seqlen // 4 + 65536)
    records, bases = _device_like_records(text[lo:hi + tail_len])
    tail = text[hi:hi + tail_len]
    t_recs, t_bases = _device_like_records(tail)
    n_tail_headers = sum(1 for h, _, _ in t_recs if h)
    tail_first = t_recs[0][2] if t_recs and not t_recs[0][0] else 0
    recs = records[:len(records) - n_tail_headers]
    frags = []
    for n, (header, off, length) in enumerate(recs):
        own = length - (tail_first if n == len(recs) - 1 else 0)
        frags.append((None if (not header and n == 0 and rank > 0) else S._record_id(header), off, own))
    return frags, (tail_first if recs else 0), bases, (lo, hi)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("width,batch,max_k", [(60, 1000, 41), (7, 333, 20), (80, 5000, 200), (61, 100, 12)])
def test_partitioned_plan_equals_the_whole_file_plan(tmp_path, world, width, batch, max_k):
    rng = np.random.default_rng(world * 1000 + width)
    acgt = np.frombuffer(b"ACGTNacgt", dtype=np.uint8)
    parts, whole = [], {}
    for name, n in (("chr1 desc", 23117), ("chr2", 9000), ("empty", 0), ("chr3", 41), ("chr3", 700), ("chr4", 15555)):
        seq = acgt[rng.integers(0, acgt.size, n)].tobytes()
        parts.append(b">" + name.encode() + b"\n")
        for i in range(0, n, width):
            parts.append(seq[i:i + width] + b"\n")
        whole.setdefault(name.split()[0].encode(), []).append(seq)
    text = b"".join(parts)
    path = tmp_path / "g.fa"
    path.write_bytes(text)
    seqlen, look = batch + max_k - 1, max_k - 1
    cfg = S.SearchConfig(fasta_filepaths=[path], fmindex_filepaths=[path], kmer_lengths=[max_k], is_binary_search=False)
    # the reference plan: one reader of the whole file
    recs_all, bases_all = _device_like_records(text)
    want_groups = S._plan_groups(cfg, [(S._record_id(h), n) for h, _, n in recs_all], seqlen, look)
    offs_all = [off for _, off, _ in recs_all]
    views = [_rank_view(path, text, r, world, seqlen) for r in range(world)]
    ranges = [v[3] for v in views]
    assert ranges[0][0] == 0 and ranges[-1][1] == len(text) and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    fragments = [[(h, own) for h, _, own in v[0]] for v in views]
    owners = {}
    for rank, (frags, tail_bases, bases, _) in enumerate(views):
        table, locate = S._partition_table(fragments, rank, [off for _, off, _ in frags], tail_bases)
        records = [(rec[0], rec[1]) for rec in table]
        assert [(i, n) for i, n in records if n] == [(S._record_id(h), n) for h, _, n in recs_all if n]
        groups = S._plan_groups(cfg, records, seqlen, look)
        assert [(g[0], g[2], len(g[1])) for g in groups] == [(g[0], g[2], len(g[1])) for g in want_groups]
        live_all = [k for k, (_, _, n) in enumerate(recs_all) if n]
        live = [k for k, (_, n) in enumerate(records) if n]
        for g, gw in zip(groups, want_groups):
            for (rec_no, start, length, nk, foff, number), (rec_w, start_w, length_w, nk_w, foff_w, number_w) in zip(g[1], gw[1]):
                assert (start, length, nk, foff, number) == (start_w, length_w, nk_w, foff_w, number_w)
                assert live.index(rec_no) == live_all.index(rec_w)
                owner, off, fits = locate(rec_no, start, length)
                assert fits and 0 <= owner < world
                assert owners.setdefault(number, owner) == owner          # every rank names the same owner
                if owner == rank:
                    assert bases[off:off + length] == bases_all[offs_all[rec_w] + start:offs_all[rec_w] + start + length]
    assert sorted(owners) == list(range(sum(len(g[1]) for g in want_groups)))
    assert len(set(owners.values())) > 1                                  # the windows really are spread over ranks


def test_line_start_search(tmp_path):
    p = tmp_path / "t"
    p.write_bytes(b"abc\ndefgh\n\nij")
    fd = os.open(p, os.O_RDONLY)
    try:
        got = [S._line_start_at_or_after(fd, pos, 13) for pos in range(0, 15)]
    finally:
        os.close(fd)
    assert got == [0, 4, 4, 4, 4, 10, 10, 10, 10, 10, 10, 11, 13, 13, 13]
