"""nm_count_from_sequence_noalloc (include/newmap_b200.h): the count
entry point for callers that cannot call cudaMalloc.  Same counts, bit for bit (integer work: tolerance 0), as
nm_count_from_sequence through the C-ABI, on two calls in a row that share the library's fixed scratch array (the
second one smaller, so stale contents of the first lie right behind it), on both strands, with the LF-step counter
starting at zero each time; a request beyond the scratch is refused with a clear error and launches nothing.
Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import numpy as np
import pytest

from conftest import GOLD
from oracle import ora

pytestmark = pytest.mark.gpu


def _queries(rng, seq_len, n, max_len):
    idx = rng.integers(0, seq_len - 1, n).astype(np.uint64)
    ln = np.minimum(rng.integers(1, max_len, n).astype(np.uint64), np.uint64(seq_len) - idx)
    return idx, ln


def test_no_alloc_count_equals_the_allocating_entry_point_twice_in_a_row():
    from newmap_b200 import _lib
    seq = dict(ora.fasta_records(GOLD / "quirk_iupac.fa"))[b"chr1"]
    rng = np.random.default_rng(17)
    with _lib.Index(GOLD / "quirk_iupac_s5.awfmi", device=0) as ix:
        for call, (n, max_len) in enumerate(((5000, 60), (700, 25))):
            idx, ln = _queries(rng, len(seq), n, max_len)
            for rc in (False, True):
                want, want_steps = ix.count(seq, idx, ln, reverse_complement=rc, return_steps=True)
                got, got_steps = ix.count(seq, idx, ln, reverse_complement=rc, return_steps=True, no_alloc=True)
                assert np.array_equal(got, want), (call, rc, np.nonzero(got != want)[0][:10])
                assert got_steps == want_steps, (call, rc)
            assert want.any()
        # the reference's own golden vector through the new entry point (tests/test_count_kmers.py:38-44)
    with _lib.Index(GOLD / "genome_s1.awfmi", device=0) as ix:
        assert ix.count(b"AAAAATTTTTATCGAATCGA", [0, 4, 9], [4, 2, 3], no_alloc=True).tolist() == [9, 3, 1]
        # same argument rule as nm_count_from_sequence
        with pytest.raises(ValueError):
            ix.count(b"ACGT", [0], [0], no_alloc=True)
        with pytest.raises(IndexError):
            ix.count(b"ACGT", [3], [2], no_alloc=True)
        # beyond the fixed scratch: refused, nothing launched
        big = b"ACGT" * (_lib.NM_NOALLOC_SCRATCH_BYTES // 4)
        with pytest.raises(ValueError, match="fixed scratch"):
            ix.count(big, [0], [4], no_alloc=True)
        many = np.zeros(_lib.NM_NOALLOC_SCRATCH_BYTES // 20 + 64, dtype=np.uint64)
        with pytest.raises(ValueError, match="fixed scratch"):
            ix.count(b"ACGTACGT", many, many + 2, no_alloc=True)
        # ... and the entry point still works afterwards
        assert ix.count(b"AAAAATTTTTATCGAATCGA", [0, 4, 9], [4, 2, 3], no_alloc=True).tolist() == [9, 3, 1]
