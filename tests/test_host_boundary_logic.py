"""Host-side logic of the boundary that needs no GPU: argument checking of the count shim happens before any device
is touched and in the order the reference's "sOb" / "sy#OOb" parsing gives (src/newmap-count.c:28-60, 91-140), and the
window / file plan of the device path (newmap_b200/search.py _plan_groups) follows search.py:268-357."""
import pytest

from newmap_b200 import count as shim
from newmap_b200 import search as S
from newmap_b200.fasta import window_plan


@pytest.fixture()
def no_device(monkeypatch):
    def boom(*a, **k):
        raise AssertionError("the index must not be opened for an argument error")
    monkeypatch.setattr(shim, "open_index", boom)


def test_count_kmers_argument_errors_come_first(no_device):
    for args, exc in ((("ix", [b"AA"], 1.5), TypeError), (("ix", [b"AA"], 256), OverflowError), (("ix", [b"AA"], -1), OverflowError),
                      (("ix", (b"AA",), 1), TypeError), (("ix", ["AA"], 1), TypeError), (("ix", [b"AA", b""], 1), ValueError)):
        with pytest.raises(exc):
            shim.count_kmers(*args)


def test_count_from_sequence_argument_errors_come_first(no_device):
    seq = b"ACGTACGT"
    cases = (((seq, [0], [2], 300), OverflowError), (("ACGT", [0], [2], 300), TypeError),        # the sequence is parsed before num_threads
             ((bytearray(seq), [0], [2], 1), TypeError), ((seq, (0,), [2], 1), TypeError), ((seq, [0], (2,), 1), TypeError),
             ((seq, [0, 1], [2], 1), ValueError), ((seq, [0], [2.0], 1), TypeError), ((seq, [-1], [2], 1), IndexError),
             ((seq, [0], [-1], 1), IndexError), ((seq, [7], [2], 1), IndexError))
    for args, exc in cases:
        with pytest.raises(exc):
            shim.count_kmers_from_sequence("ix", *args)


def _cfg(**kw):
    return S.SearchConfig(fasta_filepaths=["x.fa"], fmindex_filepaths=["x.awfmi"], kmer_lengths=[20, 40], is_binary_search=False, **kw)


def test_plan_groups_appends_consecutive_ids_and_numbers_windows():
    records = [(b"a", 2500), (b"a", 1800), (b"empty", 0), (b"b", 900), (b"a", 700)]
    groups = S._plan_groups(_cfg(), records, 1000 + 39, 39)
    assert [g[0] for g in groups] == [b"a", b"b", b"a"]                 # the empty record opens no file; "a" comes back
    a = groups[0]
    assert a[2] == 2500 + 1800                                          # both records land in one file ...
    offsets = [w[4] for w in a[1]]
    assert offsets == sorted(offsets) and offsets[0] == 0
    nk = [w[3] for w in a[1]]
    assert sum(nk) == 4300 and offsets[-1] + nk[-1] == 4300             # ... back to back
    numbers = [w[5] for g in groups for w in g[1]]
    assert numbers == list(range(len(numbers)))                        # the sharding key runs over the whole FASTA
    for rec_no, start, length, k, off, _ in a[1]:
        assert (start, length) in [(s, n) for s, n, _ in window_plan(records[rec_no][1], 1039, 39)]


def test_plan_groups_include_exclude_quirk():
    records = [(b"a", 100), (b"b", 100), (b"c", 100)]
    assert [g[0] for g in S._plan_groups(_cfg(include_sequence_ids=[b"b"]), records, 139, 39)] == [b"b"]
    # search.py:270-278 tests `id not in include_ids` first: with only --exclude-sequences every sequence is skipped
    assert S._plan_groups(_cfg(exclude_sequence_ids=[b"b"]), records, 139, 39) == []
