"""newmap_b200.fasta.sequence_segments against the reference's expectations
(tests/test_sequence_buffer_iter.py of the reference), the oracle's closed form, and -- in the
build container -- the reference generator itself."""
import gzip
import io

import pytest

from conftest import GOLD
from newmap_b200 import fasta
from oracle import ora, pyref

SMALL = b">chr2\nCGCANCAGA\nGCANCGNCG\n"     # the reference's tests/data/chr2.fa


def _got(data, length, overlap):
    return [(s.id, s.data, s.epilogue) for s in fasta.sequence_segments(io.BytesIO(data), length, overlap)]


def test_reference_expectations():
    segs = _got(SMALL, 5, 2)
    assert [d for _, d, _ in segs] == [b"CGCAN", b"ANCAG", b"AGAGC", b"GCANC", b"NCGNC", b"NCG"]
    assert [e for _, _, e in segs] == [False] * 5 + [True]
    assert all(i == b"chr2" for i, _, _ in segs)
    assert _got(SMALL, 18, 2) == [(b"chr2", b"CGCANCAGAGCANCGNCG", True)]
    assert _got(SMALL, 100, 10) == [(b"chr2", b"CGCANCAGAGCANCGNCG", True)]
    # exact fill: the last whole window is the epilogue
    assert _got(b">a\nACGTAC\n", 4, 2) == [(b"a", b"ACGT", False), (b"a", b"GTAC", True)]


CASES = [SMALL, (GOLD / "genome.fa").read_bytes(), (GOLD / "quirk.fa").read_bytes(),
         b">a\n\n>b\nACGT\n>c\n", b"ACGT\nAC\n", b">x y z\nAC\nGT\nA\n;c\nGG\n", b">a\r\nAC GT \r\nTT\r\n>b\nA",
         b"", b">only\n", b">a\nACGT", b"\n\n>a\nAC\n\nGT\n"]
SHAPES = [(5, 2), (4, 0), (9, 3), (18, 17), (10, 9), (3, 2), (100, 10), (1000, 39), (1039, 39), (1, 0)]


@pytest.mark.parametrize("data", CASES)
def test_against_oracle_closed_form(tmp_path, data):
    p = tmp_path / "x.fa"
    p.write_bytes(data)
    for length, overlap in SHAPES:
        want = [(s.id, s.data, s.epilogue) for s in ora.segments(p, length - overlap, overlap)]
        assert _got(data, length, overlap) == want, (length, overlap)


@pytest.mark.skipif(not pyref.available(), reason="needs /root/reference (build container only)")
@pytest.mark.parametrize("data", CASES)
def test_against_reference_generator(data):
    _, ref_fasta = pyref.load()
    for length, overlap in SHAPES:
        want = [(s.id, s.data, s.epilogue) for s in ref_fasta.sequence_segments(io.BytesIO(data), length, overlap)]
        assert _got(data, length, overlap) == want, (length, overlap)


def test_block_boundaries_and_gzip(tmp_path, monkeypatch):
    data = (GOLD / "quirk.fa").read_bytes()
    want = _got(data, 1039, 39)
    for block in (7, 61, 64, 1000):
        monkeypatch.setattr(fasta, "_READ_BLOCK", block)
        assert _got(data, 1039, 39) == want
    gz = tmp_path / "q.fa.gz"
    with gzip.open(gz, "wb") as fh:
        fh.write(data)
    from newmap_b200.util import optional_gzip_open
    with optional_gzip_open(gz, "rb") as fh:
        assert [(s.id, s.data, s.epilogue) for s in fasta.sequence_segments(fh, 1039, 39)] == want


def test_window_plan_is_sequence_segments_in_closed_form():
    """fasta.window_plan (used by the device ingest path, where a record is one contiguous device buffer) cuts exactly
    the windows sequence_segments cuts."""
    import io
    import numpy as np
    from newmap_b200.fasta import sequence_segments, window_plan
    rng = np.random.default_rng(11)
    for _ in range(300):
        sl = int(rng.integers(2, 60))
        ov = int(rng.integers(0, sl))
        L = int(rng.integers(0, 400))
        seq = bytes(rng.integers(65, 70, L, dtype=np.uint8))
        fa = b">r\n" + b"\n".join(seq[i:i + 17] for i in range(0, L, 17)) + b"\n"
        want = [(s.data, s.epilogue) for s in sequence_segments(io.BytesIO(fa), sl, ov)]
        got = [(seq[a:a + n], ep) for a, n, ep in window_plan(L, sl, ov)]
        assert got == want, (L, sl, ov)
