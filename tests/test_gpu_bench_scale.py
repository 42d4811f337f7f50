"""Parity at BENCH scale against the reference itself.

The 100 Mbp bench genome (BASELINE configs[1], `c2_100m` of bench.py) is indexed twice -- by the reference's
own builder (oracle/_ref: AwFmCreate + libdivsufsort) and by the GPU builder -- and the two files must be
identical.  Whole 10 M-position segments (the CLI's default batch) then go through the C-ABI on the
REFERENCE-BUILT index, and contiguous sub-samples of >= 200 k positions of each are answered by the reference's
compiled count code under the reference's host loop (oracle/refdrive.py) on the same bytes.

A sub-sample [a, a+m) of a segment reproduces the segment's own answers when it keeps the segment's look-ahead
semantics (an ambiguous byte in the LOOK-AHEAD of a batch is not seen by the upper-bound
truncation): the tail sub-sample (its look-ahead IS the segment's) always does, an inner one does when its own
look-ahead bytes are all unambiguous -- both kinds are used."""
import os

import numpy as np
import pytest

from oracle import ora

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref")]

BATCH = 10_000_000
SAMPLE = 200_000
CORES = min(os.cpu_count() or 1, 64)


@pytest.fixture(scope="module")
def c2(tmp_path_factory):
    """(FASTA, reference-built index, GPU-built index) of the 100 Mbp bench genome."""
    import bench
    from newmap_b200 import index, synth
    d = tmp_path_factory.mktemp("c2_100m")
    fa = d / "genome.fa"
    lengths, seed = synth.config_lengths("c2_100m")
    synth.write_fasta(fa, synth.make_genome(lengths, seed))
    ref, gpu = d / "ref.awfmi", d / "gpu.awfmi"
    ora.ref_build_index(fa, ref, bench.SA_RATIO, bench.SEED_LEN)
    index.generate_fm_index(fa, gpu, bench.SA_RATIO, bench.SEED_LEN, device=0)
    return fa, ref, gpu


def test_gpu_built_index_equals_reference_built_at_100mbp(c2):
    _, ref, gpu = c2
    a, b = np.fromfile(ref, dtype=np.uint8), np.fromfile(gpu, dtype=np.uint8)
    assert a.size == b.size, (a.size, b.size)
    bad = np.nonzero(a != b)[0]
    assert bad.size == 0, (bad.size, bad[:5])


def _segments(fa, max_k):
    return list(ora.segments(fa, BATCH, max_k - 1))


def _subsamples(seg, n, max_k):
    """[(a, m)]: the tail sub-sample and one inner sub-sample whose look-ahead is unambiguous."""
    allowed = np.zeros(256, dtype=bool)
    allowed[list(b"ACGTacgt")] = True
    buf = np.frombuffer(seg.data, dtype=np.uint8)
    out = [(n - SAMPLE, SAMPLE)]
    for a in range(2_000_000, n - 2 * SAMPLE, 1_000_003):
        look = buf[a + SAMPLE:a + SAMPLE + max_k - 1]
        if allowed[look].all() and not allowed[buf[a:a + SAMPLE]].all():   # clean look-ahead, some N inside
            out.append((a, SAMPLE))
            break
    else:
        out.append((3_000_000, SAMPLE) if allowed[buf[3_000_000 + SAMPLE:3_000_000 + SAMPLE + max_k - 1]].all() else out[0])
    return out


@pytest.mark.parametrize("mode", ["range_quirk", "range_epilogue", "list"])
def test_whole_segments_against_the_reference_on_subsamples(c2, mode):
    from newmap_b200 import _lib
    from oracle import refdrive
    fa, ref, _ = c2
    is_range = mode != "list"
    klist = list(range(20, 201)) if is_range else [24, 36, 50, 100]
    max_k = max(klist)
    segs = _segments(fa, max_k)
    allowed = np.zeros(256, dtype=bool)
    allowed[list(b"ACGTacgt")] = True
    if mode == "range_quirk":
        # a segment whose look-ahead (the first bytes of the next batch) holds an ambiguous byte: SURVEY §8 A7
        pick = [s for s in segs if not s.epilogue and
                not allowed[np.frombuffer(s.data, dtype=np.uint8)[BATCH:]].all()]
        assert pick, "the bench genome places N-runs right after multiples of the batch size"
        seg = pick[0]
    elif mode == "range_epilogue":
        seg = [s for s in segs if s.epilogue][-1]
    else:
        seg = segs[4]
    n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
    assert n >= 5_000_000
    with _lib.Index(ref, device=0) as ix:
        got, st = ix.min_unique(seg.data, n, klist, is_range, algo="auto", dtype=np.uint32)
        assert st.replayed == 0 and st.zero_pos < 0
        assert ix.derived_seed_length >= 13          # the production configuration: deep both-strand table
        for a, m in _subsamples(seg, n, max_k):
            tail = a + m == n
            data = seg.data[a:] if tail else seg.data[a:a + m + max_k - 1]
            want = refdrive.search_segment(ref, data, m, klist, is_range, 0, True, CORES)
            bad = np.nonzero(got[a:a + m] != want)[0]
            assert bad.size == 0, (mode, a, m, bad[:10], got[a:a + m][bad[:10]], want[bad[:10]])
            assert want.any()
