"""Pins the CPU oracle (oracle/fm_oracle.c + oracle/ora.py) against the reference:
(1) the reference's own golden vectors (tests/test_count_kmers.py:22-44,
tests/test_unique_counts.py:17-21, tests/test_upper_search_bound_truncation.py,
tests/test_sequence_buffer_iter.py of the reference), (2) committed fixtures produced by the
unmodified reference (oracle/gen_golden.py), (3) live differential runs against the
reference where /root/reference and oracle/_ref exist (skipped elsewhere)."""
import io

import numpy as np
import pytest

from conftest import GOLD, run_klist
from oracle import ora, pyref


@pytest.fixture(scope="module")
def genome_index():
    return ora.OracleIndex(GOLD / "genome_s1.awfmi")


def test_header_and_prefix_sums(genome_index):
    # decoded dump of the reference-built 408-byte index
    assert genome_index.bwt_length == 53
    assert genome_index.seed_length == 1
    assert genome_index.prefix_sums == [1, 20, 32, 44, 51, 53]


@pytest.mark.parametrize("ix", ["genome_s1", "genome_s3"])
def test_reference_count_goldens(ix):
    o = ora.OracleIndex(GOLD / f"{ix}.awfmi")
    # reference tests/test_count_kmers.py:22-26
    assert o.count_kmers([b"AAAA", b"AT", b"TAT", b"CCC", b"NNN", b"TCGT"]).tolist() == [9, 3, 1, 8, 0, 0]
    # reference tests/test_count_kmers.py:38-44
    assert o.count(b"AAAAATTTTTATCGAATCGA", [0, 4, 9], [4, 2, 3]).tolist() == [9, 3, 1]


def test_count_fixture_vectors(manifest, golden_counts):
    for c in manifest["counts"]:
        o = ora.OracleIndex(GOLD / f"{c['index']}.awfmi")
        seq = dict(ora.fasta_records(GOLD / c["fasta"]))[c["sequence"].encode()]
        got = o.count(seq, golden_counts[c["key"] + "|idx"], golden_counts[c["key"] + "|len"])
        assert np.array_equal(got, golden_counts[c["key"] + "|count"]), c["key"]


def test_rc_count_matches_translate_reverse():
    o = ora.OracleIndex(GOLD / "quirk_iupac_s5.awfmi")
    seq = dict(ora.fasta_records(GOLD / "quirk_iupac.fa"))[b"chr1"]
    rng = np.random.default_rng(5)
    idx = rng.integers(0, len(seq) - 70, size=300)
    ln = rng.integers(1, 70, size=300)
    rcseq = seq.translate(bytes.maketrans(b"ACGTacgt", b"TGCAtgca"))[::-1]
    ridx = len(seq) - idx - ln
    assert np.array_equal(o.count(seq, idx, ln, rc=True), o.count(rcseq, ridx, ln))


def test_reference_unique_goldens():
    # reference tests/test_unique_counts.py:17-21,50-67 (k=4..10, batch 15, both modes)
    chr1 = [0, 10, 9, 8, 7, 6, 5, 4, 4, 4, 6, 5, 4, 4, 4, 0, 0, 0, 0, 0]
    chr2 = [10, 10, 9, 8, 7, 6, 5, 4, 4, 4] + [0] * 20
    for ix in ("genome_s1", "genome_s3"):
        o = ora.OracleIndex(GOLD / f"{ix}.awfmi")
        for is_range in (True, False):
            res = ora.min_unique(GOLD / "genome.fa", o, list(range(4, 11)), is_range, 15)
            assert res[b"chr1"].tolist() == chr1 and res[b"chr2"].tolist() == chr2
            assert res[b"chr1"].dtype == np.uint8


def test_search_fixture_vectors(manifest, golden_searches):
    for run in manifest["searches"]:
        o = ora.OracleIndex(GOLD / f"{run['index']}.awfmi")
        res = ora.min_unique(GOLD / run["fasta"], o, run_klist(run), run["range"], run["batch"],
                             initial_search_length=run.get("initial", 0), use_rc=not run.get("norc", False))
        assert [k.decode() for k in res] == run["sequences"]
        for sid, arr in res.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            assert arr.dtype == want.dtype and np.array_equal(arr, want), (run["id"], sid)


# ---- upper-bound truncation: the reference's 16 vectors, restated as (mask, seg_len, max_k, want)
def _ub_cases():
    F, T = False, True
    return [
        # all unambiguous, look-ahead long enough / short / none
        ([F] * 5, 5 + 3, 4, [4, 4, 4, 4, 4]),
        ([F] * 5, 5 + 1, 4, [4, 4, 4, 3, 2]),
        ([F] * 5, 5, 4, [4, 4, 3, 2, 1]),
        ([F] * 3, 3, 10, [3, 2, 1]),
        # all ambiguous: untouched
        ([T] * 4, 4 + 2, 3, [3, 3, 3, 3]),
        # gap at the start
        ([T, T, F, F, F, F], 6 + 3, 4, [4, 4, 4, 4, 4, 4]),
        ([T, T, F, F, F, F], 6, 4, [4, 4, 4, 3, 2, 1]),
        # gap in the centre
        ([F, F, F, T, T, F, F, F], 8 + 2, 3, [3, 2, 1, 3, 3, 3, 3, 3]),
        ([F, F, F, T, T, F, F, F], 8, 3, [3, 2, 1, 3, 3, 3, 2, 1]),
        ([F, F, F, F, F, T, F, F], 8, 4, [4, 4, 3, 2, 1, 4, 2, 1]),
        # gap at the end
        ([F, F, F, F, T, T], 6 + 3, 4, [4, 3, 2, 1, 4, 4]),
        ([F, F, F, F, T, T], 6, 3, [3, 3, 2, 1, 3, 3]),
        # several gaps
        ([T, F, F, T, F, T, F, F, F], 9 + 1, 3, [3, 2, 1, 3, 1, 3, 3, 3, 2]),
        ([F, T, F, T, F, T], 6 + 2, 3, [1, 3, 1, 3, 1, 3]),
        ([T, F, T, F, T, F], 6 + 2, 3, [3, 1, 3, 1, 3, 3]),
        ([T, F, T, F, T, F], 6, 3, [3, 1, 3, 1, 3, 1]),
    ]


@pytest.mark.parametrize("mask,seg_len,max_k,want", _ub_cases())
def test_upper_bounds_closed_form(mask, seg_len, max_k, want):
    assert ora.upper_bounds(np.array(mask, dtype=np.uint8), seg_len, max_k).tolist() == want


@pytest.mark.skipif(not pyref.available(), reason="needs /root/reference (build container only)")
def test_upper_bounds_against_reference_function():
    search, _ = pyref.load()
    rng = np.random.default_rng(3)
    cases = [(np.array(m, dtype=bool), s, k) for m, s, k, _ in _ub_cases()]
    for _ in range(400):
        n = int(rng.integers(1, 40))
        max_k = int(rng.integers(1, 12))
        p = rng.choice([0.0, 0.1, 0.5, 0.9, 1.0])
        cases.append((rng.random(n) < p, n + int(rng.integers(0, max_k)), max_k))
    for mask, seg_len, max_k in cases:
        ub = np.full(mask.size, max_k, dtype=np.uint32)
        search.update_upper_search_bound(ub, mask, max_k, seg_len)
        got = ora.upper_bounds(mask.astype(np.uint8), seg_len, max_k)
        keep = ~mask          # entries at ambiguous positions are never read
        assert np.array_equal(got[keep], ub[keep]), (mask, seg_len, max_k)
    # the 16 restated vectors agree with the reference function too
    for m, s, k, want in _ub_cases():
        ub = np.full(len(m), k, dtype=np.uint32)
        search.update_upper_search_bound(ub, np.array(m, dtype=bool), k, s)
        assert ub.tolist() == want


# ---- segmentation: closed form vs the reference generator
def _seg_cases():
    chr2 = (GOLD / ".." / "golden" / "quirk.fa").read_bytes()
    small = b">chr2\nCGCANCAGA\nGCANCGNCG\n"       # the reference's tests/data/chr2.fa
    genome = (GOLD / "genome.fa").read_bytes()
    out = []
    for data in (small, genome, chr2, b">a\n\n>b\nACGT\n>c\n", b"ACGT\nAC\n", b">x y z\nAC\nGT\nA\n;c\nGG\n"):
        for length, overlap in [(5, 2), (4, 0), (9, 3), (18, 17), (10, 9), (3, 2), (100, 10), (1000, 39), (1039, 39)]:
            out.append((data, length, overlap))
    return out


def test_segments_reference_expectations():
    # reference tests/test_sequence_buffer_iter.py: chr2.fa, length 5 overlap 2 -> windows every 3 bases
    segs = list(ora.segments_of(b"chr2", b"CGCANCAGAGCANCGNCG", 3, 2))
    assert [s.data for s in segs] == [b"CGCAN", b"ANCAG", b"AGAGC", b"GCANC", b"NCGNC", b"NCG"]
    assert [s.epilogue for s in segs] == [False] * 5 + [True]
    segs = list(ora.segments_of(b"chr2", b"CGCANCAGAGCANCGNCG", 16, 2))
    assert [s.data for s in segs] == [b"CGCANCAGAGCANCGNCG"] and segs[0].epilogue


@pytest.mark.skipif(not pyref.available(), reason="needs /root/reference (build container only)")
def test_segments_against_reference_generator(tmp_path):
    _, fasta = pyref.load()
    for n, (data, length, overlap) in enumerate(_seg_cases()):
        want = [(s.id, s.data, s.epilogue) for s in fasta.sequence_segments(io.BytesIO(data), length, overlap)]
        p = tmp_path / f"c{n}.fa"
        p.write_bytes(data)
        got = [(s.id, s.data, s.epilogue) for s in ora.segments(p, length - overlap, overlap)]
        assert got == want, (data[:40], length, overlap)


@pytest.mark.skipif(not pyref.available(), reason="needs /root/reference (build container only)")
def test_live_differential_against_reference(tmp_path):
    """Fresh random genome -> reference index builder -> reference search vs oracle."""
    from newmap_b200 import synth
    recs = synth.make_genome([30_000, 12_000], seed=99, batch=5000)
    fa = tmp_path / "g.fa"
    synth.write_fasta(fa, recs)
    ix = tmp_path / "g.awfmi"
    ora.ref_build_index(fa, ix, 8, 7)
    o = ora.OracleIndex(ix)
    for klist, is_range, batch in [(list(range(20, 201)), True, 5000), ([24, 36, 50, 100], False, 7000),
                                   (list(range(12, 256)), True, 9000)]:
        want = pyref.run_search(fa, ix, klist, is_range, batch, tmp_path / "out", num_threads=4)
        got = ora.min_unique(fa, o, klist, is_range, batch)
        assert set(got) == set(want)
        for sid in want:
            assert np.array_equal(got[sid], want[sid]), (sid, is_range, batch)


def test_zero_count_raises_like_reference():
    o = ora.OracleIndex(GOLD / "genome_s1.awfmi")
    with pytest.raises(ora.ZeroCount) as e:
        o.min_unique_segment(b"ACGTTTGCATGCAAGT", 16, list(range(3, 6)), True)
    assert "was not found in the index" in str(e.value)


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref")
def test_refdrive_matches_oracle_and_goldens(manifest, golden_searches):
    """The CPU-baseline driver (reference C code + restated host loop) gives the reference's bytes."""
    from oracle import refdrive
    for run in manifest["searches"]:
        if run["id"] % 4:       # a quarter of the runs is plenty: each round re-reads the index like the reference
            continue
        klist = run_klist(run)
        max_k = max(klist)
        got = {}
        for seg in ora.segments(GOLD / run["fasta"], run["batch"], max_k - 1):
            n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
            out = refdrive.search_segment(GOLD / f"{run['index']}.awfmi", seg.data, n, klist, run["range"],
                                          run.get("initial", 0), not run.get("norc", False), 2)
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            assert np.array_equal(np.concatenate(parts).astype(want.dtype), want), (run["id"], sid)
