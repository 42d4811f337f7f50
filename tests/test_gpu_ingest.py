"""FASTA ingest on the device (nm_fasta_*, newmap_b200.search._write_unique_counts_device) against the host reader
(newmap_b200/fasta.py, which tests/test_fasta_segments.py pins to the reference's segmentation cases)."""
import gzip
import io
import os

import numpy as np
import pytest

from conftest import GOLD
from newmap_b200.fasta import _records
from oracle import ora

pytestmark = pytest.mark.gpu


def _host_records(text: bytes):
    recs, cur, parts = [], None, []
    for sid, piece, finished in _records(io.BytesIO(text)):
        if cur is None:
            cur = sid
        parts.append(piece)
        if finished:
            recs.append((sid, b"".join(parts)))
            parts, cur = [], None
    return recs


def _device_records(nm, text: bytes, chunks=None):
    with nm.Fasta(0) as fa:
        if chunks:
            for i in range(0, len(text), chunks):
                fa.feed(text[i:i + chunks])
        else:
            fa.feed(text)
        fa.finish()
        out = []
        for header, off, n in fa.records:
            tok = header.split()
            out.append((tok[0][1:] if tok else b"", fa.read(off, n)))
        return out, fa.info


@pytest.fixture(scope="module")
def nm():
    from newmap_b200 import _lib
    _lib.load()
    return _lib


def _random_fasta(rng, nrec, maxlen, width, leading=False, trailing_newline=True, blank_lines=True):
    parts = []
    if leading:
        parts.append(b"ACGTNNACGT\nAC\n")
    for r in range(nrec):
        L = int(rng.integers(0, maxlen))
        seq = bytes(np.frombuffer(b"ACGTNacgtn", dtype=np.uint8)[rng.integers(0, 10, L)])
        start = b">" if r % 5 else b";"
        parts.append(start + f"rec{r} some description {r}".encode() + b"\n")
        for i in range(0, L, width):
            parts.append(seq[i:i + width] + b"\n")
            if blank_lines and rng.integers(0, 50) == 0:
                parts.append(b"\n")
    text = b"".join(parts)
    return text if trailing_newline else text.rstrip(b"\n")


def test_device_parse_equals_host_reader(nm):
    rng = np.random.default_rng(5)
    cases = [(3, 300, 60, False, True), (40, 20000, 61, True, False), (1, 100000, 80, False, True),
             (200, 50, 7, True, True), (0, 1, 60, True, True), (5, 9000, 4096, False, False)]
    for nrec, maxlen, width, leading, trailing in cases:
        text = _random_fasta(rng, nrec, maxlen, width, leading, trailing)
        want = [(i, s) for i, s in _host_records(text)]
        for chunks in (None, 1000):
            got, info = _device_records(nm, text, chunks)
            assert info.odd_whitespace == 0 and info.raw_bytes == len(text)
            assert [g for g in got if g[1]] == [w for w in want if w[1]], (nrec, maxlen, width, leading, trailing, chunks)
    got, info = _device_records(nm, b"")
    assert got == [] and info.records == 0
    got, info = _device_records(nm, b">only header")
    assert [g for g in got if g[1]] == []
    # white space that needs real right-stripping is reported, not parsed
    for text in (b">a\nACGT \nAC\n", b">a\r\nACGT\r\nAC\r\n", b">a\nAC\tGT\n"):
        assert _device_records(nm, text)[1].odd_whitespace == 1


def test_feed_file_reads_with_threads(nm, tmp_path):
    rng = np.random.default_rng(9)
    text = _random_fasta(rng, 6, 30_000_000, 60, blank_lines=False)       # > 32 MiB chunks: several readers
    p = tmp_path / "big.fa"
    p.write_bytes(text)
    want = [(i, s) for i, s in _host_records(text) if s]
    with nm.Fasta(0, size_hint=len(text)) as fa:
        fa.feed_file(p, threads=3)
        fa.finish()
        assert fa.info.raw_bytes == len(text)
        got = [(h.split()[0][1:], n) for h, off, n in fa.records if n]
        assert got == [(i, len(s)) for i, s in want]
        for (h, off, n), (_, s) in zip([r for r in fa.records if r[2]], want):
            assert fa.read(off, min(n, 1000)) == s[:1000]
            assert fa.read(off + n - min(n, 777), min(n, 777)) == s[-min(n, 777):]


def _run_search(tmp_path, name, fasta, index, ingest, monkeypatch, **kw):
    from pathlib import Path
    from newmap_b200 import search as S
    out = tmp_path / f"{name}_{ingest}"
    out.mkdir()
    monkeypatch.setattr(S, "INGEST", ingest)
    S.write_unique_counts(S.SearchConfig(fasta_filepaths=[Path(fasta)], fmindex_filepaths=[Path(index)],
                                         output_directory=out, **kw))
    return {p.name: p.read_bytes() for p in sorted(out.iterdir())}


@pytest.mark.parametrize("batch", [15, 1000, 10_000_000])
def test_device_ingest_writes_the_same_files(nm, tmp_path, monkeypatch, batch):
    for fasta, index, kl, rng_mode in (("genome.fa", "genome_s1.awfmi", list(range(4, 11)), True),
                                       ("quirk.fa", "quirk_s6.awfmi", list(range(20, 201)), True),
                                       ("quirk.fa", "quirk_s4.awfmi", [24, 36, 50, 100], False),
                                       ("quirk_iupac.fa", "quirk_iupac_s5.awfmi", list(range(8, 41)), True)):
        kw = dict(kmer_lengths=kl, is_binary_search=rng_mode, kmer_batch_size=batch)
        a = _run_search(tmp_path, f"{fasta}{batch}{len(kl)}", GOLD / fasta, GOLD / index, "host", monkeypatch, **kw)
        b = _run_search(tmp_path, f"{fasta}{batch}{len(kl)}", GOLD / fasta, GOLD / index, "device", monkeypatch, **kw)
        assert a.keys() == b.keys() and a, (fasta, batch)
        for k in a:
            assert a[k] == b[k], (fasta, batch, k)


def test_device_ingest_gz_filters_and_sharding(nm, tmp_path, monkeypatch):
    gz = tmp_path / "quirk.fa.gz"
    with gzip.open(gz, "wb") as fh:
        fh.write((GOLD / "quirk.fa").read_bytes())
    kw = dict(kmer_lengths=list(range(20, 101)), is_binary_search=True, kmer_batch_size=700)
    a = _run_search(tmp_path, "gz", GOLD / "quirk.fa", GOLD / "quirk_s6.awfmi", "host", monkeypatch, **kw)
    b = _run_search(tmp_path, "gz", gz, GOLD / "quirk_s6.awfmi", "device", monkeypatch, **kw)
    assert a == b
    c = _run_search(tmp_path, "inc", GOLD / "quirk.fa", GOLD / "quirk_s6.awfmi", "device", monkeypatch,
                    include_sequence_ids=[b"chr2"], **kw)
    assert list(c) == ["chr2.unique.uint8"] and c["chr2.unique.uint8"] == a["chr2.unique.uint8"]
    # two ranks share the output files (host-side gather at window offsets)
    from pathlib import Path
    from newmap_b200 import search as S
    out = tmp_path / "sharded"
    out.mkdir()
    monkeypatch.setattr(S, "INGEST", "device")
    for rank in (1, 0):
        S.write_unique_counts(S.SearchConfig(fasta_filepaths=[GOLD / "quirk.fa"], fmindex_filepaths=[GOLD / "quirk_s6.awfmi"],
                                             output_directory=out, rank=rank, world_size=2, **kw))
    assert {p.name: p.read_bytes() for p in sorted(out.iterdir())} == a


@pytest.mark.parametrize("ingest", ["host", "device"])
def test_records_that_repeat_an_id_like_the_reference(nm, tmp_path, monkeypatch, ingest):
    """tests/golden/dup_*.fa (oracle/gen_dup_golden.py): an id twice in a row is APPENDED to one file
    (search.py:268-357 opens a file anew only when the id changes), an id that comes back after another one truncates
    it -- both paths write what the unmodified reference wrote, also when two ranks share the files."""
    gold = np.load(GOLD / "dup_ids.npz")
    runs = [dict(kmer_lengths=list(range(20, 61)), is_binary_search=True, kmer_batch_size=1000),
            dict(kmer_lengths=[22, 30, 45], is_binary_search=False, kmer_batch_size=700)]
    for tag in ("append", "return"):
        for n, kw in enumerate(runs):
            got = _run_search(tmp_path, f"{tag}{n}", GOLD / f"dup_{tag}.fa", GOLD / f"dup_{tag}.awfmi", ingest, monkeypatch, **kw)
            assert sorted(got) == ["a.unique.uint8", "b.unique.uint8"]
            for sid in ("a", "b"):
                want = gold[f"{tag}|run{n}|{sid}"]
                assert np.array_equal(np.frombuffer(got[f"{sid}.unique.uint8"], dtype=np.uint8), want), (tag, n, sid, ingest)
    # sharded over two ranks (files pre-filled with stale bytes longer than the result: the size is set by the owner
    # of a file's last window)
    from newmap_b200 import search as S
    out = tmp_path / f"sharded_{ingest}"
    out.mkdir()
    for sid in ("a", "b"):
        (out / f"{sid}.unique.uint8").write_bytes(b"\xff" * 9000)
    monkeypatch.setattr(S, "INGEST", ingest)
    for rank in (1, 0):
        S.write_unique_counts(S.SearchConfig(fasta_filepaths=[GOLD / "dup_append.fa"], fmindex_filepaths=[GOLD / "dup_append.awfmi"],
                                             output_directory=out, rank=rank, world_size=2, **runs[0]))
    for sid in ("a", "b"):
        assert np.array_equal(np.fromfile(out / f"{sid}.unique.uint8", dtype=np.uint8), gold[f"append|run0|{sid}"]), sid


def test_verbose_statistics_come_from_the_device(nm, tmp_path, monkeypatch, capsys):
    """--verbose no longer drops to the host reader: the per-sequence summary (search.py:885-902) is reduced on the
    device (NM_OPT_OUT_STATS) and equals the one the host path computes from the bytes."""
    logs = {}
    for ingest in ("host", "device"):
        capsys.readouterr()
        _run_search(tmp_path, "verbose", GOLD / "quirk.fa", GOLD / "quirk_s6.awfmi", ingest, monkeypatch,
                    kmer_lengths=list(range(20, 101)), is_binary_search=True, kmer_batch_size=1500, verbose=True)
        logs[ingest] = capsys.readouterr().err.splitlines()
    assert logs["host"] == logs["device"]
    assert any("unique lengths found" in t for t in logs["device"]) and any("Processing 1500 k-mers" in t for t in logs["device"])


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's index builder)")
def test_many_sequences_interleaved_writes(nm, tmp_path, monkeypatch):
    """More sequences than the device path deals round-robin at a time (INTERLEAVE = 8), of very different lengths,
    some shorter than a window: same files as the host path, which writes file after file; and with the interleaving
    off and the number of files open at a time bounded."""
    from newmap_b200 import search as S
    rng = np.random.default_rng(23)
    acgt = np.frombuffer(b"ACGTN", dtype=np.uint8)
    parts = []
    for j in range(21):
        n = int(rng.integers(5, 9000)) if j % 5 else 30
        parts.append(b">s%d\n" % j + acgt[rng.integers(0, 5 if j % 3 == 0 else 4, n)].tobytes() + b"\n")
    fa = tmp_path / "many.fa"
    fa.write_bytes(b"".join(parts))
    ixp = tmp_path / "many.awfmi"
    ora.ref_build_index(fa, ixp, 8, 5)
    kw = dict(kmer_lengths=list(range(12, 61)), is_binary_search=True, kmer_batch_size=700)
    a = _run_search(tmp_path, "many", fa, ixp, "host", monkeypatch, **kw)
    b = _run_search(tmp_path, "many", fa, ixp, "device", monkeypatch, **kw)
    assert a.keys() == b.keys() and len(a) == 21
    assert a == b
    monkeypatch.setattr(S, "INTERLEAVE", 1)
    assert _run_search(tmp_path, "many_il1", fa, ixp, "device", monkeypatch, **kw) == a
    monkeypatch.setattr(S, "INTERLEAVE", 5)
    monkeypatch.setattr(S, "SLOT_CACHE", False)
    assert _run_search(tmp_path, "many_il5", fa, ixp, "device", monkeypatch, **kw) == a
