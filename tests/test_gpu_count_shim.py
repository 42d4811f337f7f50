"""The boundary of the path: newmap_b200.count is the drop-in for the reference's extension
`newmap._c_newmap_count_kmers` (src/newmap-count.c:28-206) -- same two functions, same argument meaning, same
exceptions -- backed by the CUDA library.

* the reference's own tests/test_count_kmers.py:21-44, case for case, against the shim;
* the shim against the reference's compiled extension (oracle/_ref) call for call, results and exceptions;
* INTEGRATION.md option A on the GPU: the reference's host loops (newmap/search.py:383-724 as restated call for
  call in oracle/refdrive.py) driving the SHIM instead of the reference's extension must reproduce the files the
  unmodified reference wrote (tests/golden/searches.npz).  (The unmodified newmap/search.py itself runs on the shim
  in tests/test_reference_host_on_shim.py, where /root/reference exists.)"""
import numpy as np
import pytest

from conftest import GOLD, run_klist
from oracle import ora

pytestmark = pytest.mark.gpu

DEFAULT_COMPRESSION_RATIO, DEFAULT_SEED_LENGTH = 8, 12      # newmap/main.py:13-14


@pytest.fixture(scope="module")
def genome_index(tmp_path_factory):
    # setUp of tests/test_count_kmers.py:14-19: the index is generated from genome.fa with the CLI defaults
    from newmap_b200.index import generate_fm_index
    out = tmp_path_factory.mktemp("shim") / "genome.awfmi"
    generate_fm_index(str(GOLD / "genome.fa"), str(out), DEFAULT_COMPRESSION_RATIO, DEFAULT_SEED_LENGTH)
    return str(out)


def test_count_kmers(genome_index):                         # tests/test_count_kmers.py:21-26
    from newmap_b200.count import count_kmers
    counts = count_kmers(genome_index, [b'AAAA', b'AT', b'TAT', b'CCC', b'NNN', b'TCGT'], 1)
    assert counts == [9, 3, 1, 8, 0, 0]


def test_count_wrong_type(genome_index):                    # :28-31
    from newmap_b200.count import count_kmers
    with pytest.raises(TypeError):
        count_kmers(genome_index, ["AAAA"], 1)


def test_empty_byte_string(genome_index):                   # :33-36
    from newmap_b200.count import count_kmers
    with pytest.raises(ValueError):
        count_kmers(genome_index, [b'AAAA', b'', b'TAT'], 1)


def test_sequence_counting(genome_index):                   # :38-44
    from newmap_b200.count import count_kmers_from_sequence
    counts = count_kmers_from_sequence(genome_index, b'AAAAATTTTTATCGAATCGA', [0, 4, 9], [4, 2, 3], 1)
    assert counts == [9, 3, 1]


def _outcome(fn, *args):
    try:
        return ("ok", fn(*args))
    except Exception as e:          # noqa: BLE001 -- the exception type and text ARE the result compared
        return (type(e).__name__, str(e))


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's compiled extension)")
def test_shim_against_the_reference_extension_call_for_call():
    from newmap_b200 import count as shim
    ref = ora.ref_modules()[0]
    seq = b"AAAAATTTTTATCGAATCGA"
    for ix in (str(GOLD / "genome_s1.awfmi"), str(GOLD / "genome_s3.awfmi"), str(GOLD / "quirk_s4.awfmi")):
        # (a zero-length k-mer is answered 0 by the reference on a seed-length-1 index and is undefined behaviour -- it
        #  reads kmer[-1] -- on any other, so it is not compared live; test_zero_length_kmers_count_zero pins the shim)
        for args in ((seq, [0], [-1]), (seq, [-1], [2]), (b"AAAA", [3], [2]), ("AAAA", [0], [2]),
                     (b"AAAA", (0,), [2]), (b"AAAA", [0], (2,)), (b"AAAA", [0], [2.0]), (b"AAAA", [0, 1], [2]),
                     (bytearray(b"AAAA"), [0], [2]), (seq, [0, 4, 9], [4, 2, 3]), (seq, [0, 1, 2, 5], [7, 19, 3, 15]),
                     (b"NNNNACGTnnacgtRYKM", [0, 4, 3, 8, 14], [4, 4, 6, 6, 4]), (seq, [], [])):
            for threads in (1, 0, 255, 256, -1, 1.5, True):
                a = _outcome(ref.count_kmers_from_sequence, ix, *args, threads)
                b = _outcome(shim.count_kmers_from_sequence, ix, *args, threads)
                assert a == b, (ix, args, threads, a, b)
        for kmers in ([b"AAAA", b"AT", b"TAT", b"CCC", b"NNN", b"TCGT"], ["AAAA"], [b"AAAA", b""], (b"AA",), [],
                      [b"A"], [b"acgt", b"ACGTN", b"nnn", b"T" * 40], [b"AA", 3]):
            for threads in (1, 300, -1, 2.5, True):
                a = _outcome(ref.count_kmers, ix, kmers, threads)
                b = _outcome(shim.count_kmers, ix, kmers, threads)
                assert a == b, (ix, kmers, threads, a, b)


def test_zero_length_kmers_count_zero():
    from newmap_b200 import count as shim
    ix = str(GOLD / "genome_s1.awfmi")      # seed length 1: the one index on which the reference defines this (-> 0)
    assert shim.count_kmers_from_sequence(ix, b"AAAA", [0, 1, 4], [0, 2, 0], 1) == [0, 14, 0]
    if ora.ref_available():
        assert ora.ref_modules()[0].count_kmers_from_sequence(ix, b"AAAA", [0, 1, 4], [0, 2, 0], 1) == [0, 14, 0]


def test_reference_host_loop_on_the_shim_reproduces_reference_files(manifest, golden_searches):
    from newmap_b200 import count as shim
    from oracle import refdrive
    done = 0
    for run in manifest["searches"]:
        if run["batch"] < 500 and run["fasta"] != "genome.fa":
            continue                                    # keep the number of count calls (each a GPU round trip) modest
        klist = run_klist(run)
        max_k = max(klist)
        index = GOLD / f"{run['index']}.awfmi"
        got = {}
        for seg in ora.segments(GOLD / run["fasta"], run["batch"], max_k - 1):
            n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
            out = refdrive.search_segment(index, seg.data, n, klist, run["range"], run.get("initial", 0),
                                          not run.get("norc", False), 1, count_module=shim)
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            assert np.array_equal(arr, want), (run, sid, np.nonzero(arr != want)[0][:10])
        done += 1
    assert done >= 10
