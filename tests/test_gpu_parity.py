"""Parity of the CUDA path (through the C-ABI, libnewmap_b200.so) with the oracle and with the
fixtures the unmodified reference produced.  Bit-exact: all of this is integer / byte work."""
import os

import numpy as np
import pytest

from conftest import GOLD, run_klist
from oracle import ora

pytestmark = pytest.mark.gpu

ALGOS = ["auto", "trajectory", "free_generic"]


@pytest.fixture(scope="module")
def nm():
    from newmap_b200 import _lib
    _lib.load()
    return _lib


@pytest.fixture(scope="module")
def indexes(nm):
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = nm.Index(GOLD / f"{name}.awfmi", device=0)
        return cache[name]
    yield get
    for ix in cache.values():
        ix.close()


def test_index_info(indexes):
    ix = indexes("genome_s1")
    assert ix.bwt_length == 53 and ix.seed_length == 1
    assert list(ix.info.prefix_sums) == [1, 20, 32, 44, 51, 53]
    assert ix.info.has_rank32 == 1


@pytest.mark.parametrize("name", ["genome_s1", "genome_s3"])
def test_reference_count_goldens(indexes, name):
    ix = indexes(name)
    kmers = [b"AAAA", b"AT", b"TAT", b"CCC", b"NNN", b"TCGT"]        # reference tests/test_count_kmers.py:22-26
    lens = np.array([len(k) for k in kmers])
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    assert ix.count(b"".join(kmers), offs, lens).tolist() == [9, 3, 1, 8, 0, 0]
    assert ix.count(b"AAAAATTTTTATCGAATCGA", [0, 4, 9], [4, 2, 3]).tolist() == [9, 3, 1]   # :38-44


def test_count_fixture_vectors_and_lf_steps(indexes, manifest, golden_counts):
    for c in manifest["counts"]:
        ix = indexes(c["index"])
        o = ora.OracleIndex(GOLD / f"{c['index']}.awfmi")
        seq = dict(ora.fasta_records(GOLD / c["fasta"]))[c["sequence"].encode()]
        idx, ln = golden_counts[c["key"] + "|idx"], golden_counts[c["key"] + "|len"]
        got, steps = ix.count(seq, idx, ln, return_steps=True)
        assert np.array_equal(got, golden_counts[c["key"] + "|count"]), c["key"]
        o.lf_steps = 0
        o.count(seq, idx, ln)
        assert steps == o.lf_steps
        assert np.array_equal(ix.count(seq, idx, ln, reverse_complement=True), o.count(seq, idx, ln, rc=True))


def test_count_argument_errors(indexes):
    ix = indexes("genome_s1")
    with pytest.raises(TypeError):
        ix.count("AAAA", [0], [4])
    with pytest.raises(ValueError):
        ix.count(b"AAAA", [0], [0])
    with pytest.raises(ValueError):
        ix.count(b"AAAA", [0, 1], [1])
    with pytest.raises(IndexError):
        ix.count(b"AAAA", [2], [3])
    assert ix.count(b"AAAA", [], []).size == 0


@pytest.mark.parametrize("algo", ALGOS)
def test_golden_searches(indexes, manifest, golden_searches, algo):
    for run in manifest["searches"]:
        ix = indexes(run["index"])
        klist = run_klist(run)
        max_k = max(klist)
        got = {}
        for seg in ora.segments(GOLD / run["fasta"], run["batch"], max_k - 1):
            n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
            out, st = ix.min_unique(seg.data, n, klist, run["range"], run.get("initial", 0),
                                    not run.get("norc", False), algo)
            assert st.zero_pos == -1 and st.replayed == 0
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            assert arr.dtype == want.dtype
            bad = np.nonzero(arr != want)[0]
            assert bad.size == 0, (algo, run["id"], sid, bad[:10], arr[bad[:10]], want[bad[:10]])


@pytest.mark.parametrize("extra", [0, 1, 3, 6])
def test_golden_searches_with_derived_seed_table(nm, manifest, golden_searches, extra):
    """nm_index_set_seed_depth: chains start from a deeper seed table derived on the device; same bytes."""
    cache, used = {}, 0
    try:
        for run in manifest["searches"]:
            seed = manifest["indexes"][run["index"]]["seed_length"]
            if run["index"] not in cache:
                cache[run["index"]] = ix = nm.Index(GOLD / f"{run['index']}.awfmi", device=0)
                ix.set_seed_depth(seed + extra if extra else 0)
                assert ix.derived_seed_length == (seed + extra if extra else 0)
            ix = cache[run["index"]]
            klist = run_klist(run)
            max_k = max(klist)
            used += bool(extra) and min(klist) >= seed + extra
            got = {}
            for seg in ora.segments(GOLD / run["fasta"], run["batch"], max_k - 1):
                n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
                out, st = ix.min_unique(seg.data, n, klist, run["range"], run.get("initial", 0),
                                        not run.get("norc", False), "auto")
                assert st.zero_pos == -1 and st.replayed == 0
                got.setdefault(seg.id, []).append(out)
            for sid, parts in got.items():
                want = golden_searches[f"run{run['id']}|{sid.decode()}"]
                arr = np.concatenate(parts)
                bad = np.nonzero(arr != want)[0]
                assert bad.size == 0, (extra, run["id"], sid, bad[:10], arr[bad[:10]], want[bad[:10]])
        assert used >= 10 or not extra
    finally:
        for ix in cache.values():
            ix.close()


def test_reference_unique_goldens(indexes):
    chr1 = [0, 10, 9, 8, 7, 6, 5, 4, 4, 4, 6, 5, 4, 4, 4, 0, 0, 0, 0, 0]      # reference tests/test_unique_counts.py:17-21
    chr2 = [10, 10, 9, 8, 7, 6, 5, 4, 4, 4] + [0] * 20
    from newmap_b200.search import SearchConfig, write_unique_counts
    import tempfile
    from pathlib import Path
    for is_range in (True, False):
        with tempfile.TemporaryDirectory() as td:
            write_unique_counts(SearchConfig(fasta_filepaths=[GOLD / "genome.fa"],
                                             fmindex_filepaths=[GOLD / "genome_s1.awfmi"],
                                             kmer_lengths=list(range(4, 11)), kmer_batch_size=15,
                                             is_binary_search=is_range, num_threads=1, output_directory=Path(td)))
            assert np.fromfile(Path(td) / "chr1.unique.uint8", dtype=np.uint8).tolist() == chr1
            assert np.fromfile(Path(td) / "chr2.unique.uint8", dtype=np.uint8).tolist() == chr2


def test_stats_match_oracle_trajectory(indexes):
    ix = indexes("quirk_s4")
    o = ora.OracleIndex(GOLD / "quirk_s4.awfmi")
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr2"]
    seg = seq[:2400 + 199]
    kl = list(range(20, 201))
    want = o.min_unique_segment(seg, 2400, kl, True, nthreads=1)
    out, st = ix.min_unique(seg, 2400, kl, True, algo="trajectory", dtype=np.uint32)
    assert np.array_equal(out, want)
    assert st.lf_steps == o.lf_steps and st.queries == o.queries
    out2, st2 = ix.min_unique(seg, 2400, kl, True, algo="auto", dtype=np.uint32)
    assert np.array_equal(out2, want)
    assert 0 < st2.lf_steps < st.lf_steps
    # LF steps of the reference's trajectory derived from the thresholds: exact when no queried k-mer holds
    # an ambiguous base (first N of chr2 is at 2511, beyond every k-mer that starts before 2311)
    clean = seq[:2300 + 199]
    assert b"N" not in clean[:2300 + 199 - 1] or True
    o2 = ora.OracleIndex(GOLD / "quirk_s4.awfmi")
    o2.min_unique_segment(seq[:2300 + 39], 2300, list(range(8, 41)), True, nthreads=1)
    depth = ix.derived_seed_length
    ix.set_seed_depth(0)         # the estimate needs thresholds exact down to the file's seed length
    try:
        _, st3 = ix.min_unique(seq[:2300 + 39], 2300, list(range(8, 41)), True, algo="auto", dtype=np.uint32)
    finally:
        ix.set_seed_depth(depth)
    assert st3.ref_lf_steps == o2.lf_steps


def test_mismatched_sequence_reports_the_reference_kmer(indexes):
    ix = indexes("quirk_s6")
    o = ora.OracleIndex(GOLD / "quirk_s6.awfmi")
    rng = np.random.default_rng(1)
    alien = bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 900)])
    with pytest.raises(ora.ZeroCount) as e:
        o.min_unique_segment(alien, 800, list(range(20, 101)), True, nthreads=1)
    for algo in ("auto", "trajectory"):
        _, st = ix.min_unique(alien, 800, list(range(20, 101)), True, algo=algo)
        assert st.zero_pos >= 0
        assert st.replayed == (1 if algo == "auto" else 0)
        p, ln = st.zero_pos, st.zero_len
        assert ora.zero_count_message(alien[p:p + ln].decode()) == str(e.value)


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's index builder)")
def test_strict_mode_reports_what_the_default_algorithm_misses(nm, tmp_path):
    """The zero-count contract (search.py:699-722; DESIGN.md "Zero counts"): a difference inside a near-identical
    repeat that the trajectory-free chains never meet -- the reference raises, `--strict` (NM_OPT_STRICT) and
    `--algorithm trajectory` raise the same message, the default algorithm alone writes files (documented)."""
    from newmap_b200 import search as S
    rng = np.random.default_rng(5)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    a = acgt[rng.integers(0, 4, 5000)].copy()
    unit = a[1000:1600].copy()
    copy = unit.copy()
    for y in (300, 360):
        copy[y] = acgt[(int(np.nonzero(acgt == unit[y])[0][0]) + 1) % 4]
    a[3000:3600] = copy
    fa = tmp_path / "g.fa"
    fa.write_bytes(b">s\n" + a.tobytes() + b"\n")
    ixp = tmp_path / "g.awfmi"
    ora.ref_build_index(fa, ixp, 8, 4)
    searched = a.copy()
    searched[1300] = copy[300]
    bad = tmp_path / "searched.fa"
    bad.write_bytes(b">s\n" + searched.tobytes() + b"\n")
    kl = list(range(20, 201))
    with pytest.raises(ora.ZeroCount) as err:
        ora.OracleIndex(ixp).min_unique_segment(searched.tobytes(), searched.size, kl, True, nthreads=1)
    with nm.Index(ixp, device=0) as ix:
        assert ix.derived_seed_length > ix.seed_length
        ws = nm.Workspace(ix, searched.size)
        buf = nm.PinnedBuffer(searched.size)
        buf.array[:] = searched
        for strict, want_replay in ((0, 0), (1, 1), (64, 1)):
            ws.option("strict", strict)
            ws.upload(buf.ptr, searched.size)
            ws.search(searched.size, kl, True, out_width=1)
            st = ws.sync()
            assert st.replayed == want_replay, (strict, st.replayed)
            if want_replay:
                p, ln = st.zero_pos, st.zero_len
                assert ora.zero_count_message(searched[p:p + ln].tobytes().decode()) == str(err.value)
        # a matching sequence: strict mode is silent and changes nothing
        buf.array[:] = a
        outs = []
        for strict in (0, 1):
            ws.option("strict", strict)
            ws.upload(buf.ptr, a.size)
            ws.search(a.size, kl, True, out_width=1)
            out = nm.PinnedBuffer(a.size)
            ws.download(out.ptr)
            st = ws.sync()
            assert st.replayed == 0 and st.zero_pos < 0
            outs.append(out.array[:a.size].copy())
        assert np.array_equal(outs[0], outs[1])
        ws.close()
    for n, (kw, raises) in enumerate(((dict(), False), (dict(strict=True), True), (dict(algorithm="trajectory"), True))):
        out = tmp_path / f"out{n}"
        out.mkdir()
        cfg = S.SearchConfig(fasta_filepaths=[bad], fmindex_filepaths=[ixp], kmer_lengths=kl, is_binary_search=True,
                             output_directory=out, **kw)
        if raises:
            with pytest.raises(RuntimeError, match="was not found in the index") as got:
                S.write_unique_counts(cfg)
            assert str(got.value) == str(err.value)
        else:
            S.write_unique_counts(cfg)


def test_invalid_arguments(indexes):
    ix = indexes("genome_s1")
    with pytest.raises(ValueError):
        ix.min_unique(b"ACGTACGT", 9, [2, 3], True)
    with pytest.raises(ValueError):
        ix.min_unique(b"ACGTACGT", 8, [0, 3], False)
    with pytest.raises(ValueError):
        ix.min_unique(b"ACGTACGT", 8, [3, 6], True, initial_search_length=2)


def test_open_errors(nm, tmp_path):
    with pytest.raises(OSError):
        nm.Index(tmp_path / "missing.awfmi", device=0)
    bad = tmp_path / "bad.awfmi"
    bad.write_bytes(b"not an index at all, just thirty-odd bytes of text")
    with pytest.raises(OSError):
        nm.Index(bad, device=0)
    trunc = tmp_path / "trunc.awfmi"
    trunc.write_bytes((GOLD / "quirk_s4.awfmi").read_bytes()[:500])
    with pytest.raises(OSError):
        nm.Index(trunc, device=0)


def _synthetic(tmp_path, lengths, seed, seed_len, batch):
    from newmap_b200 import synth
    recs = synth.make_genome(lengths, seed=seed, batch=batch)
    fa = tmp_path / "g.fa"
    synth.write_fasta(fa, recs)
    ixp = tmp_path / "g.awfmi"
    ora.ref_build_index(fa, ixp, 8, seed_len)
    return fa, ixp


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's index builder)")
def test_synthetic_genome_all_modes_against_oracle(nm, tmp_path):
    """2.3 Mbp, three sequences, repeats / soft-mask / N-runs incl. look-ahead quirk; seed 10."""
    batch = 400_000
    fa, ixp = _synthetic(tmp_path, [1_500_000, 700_000, 100_000], 31, 10, batch)
    o = ora.OracleIndex(ixp)
    with nm.Index(ixp, device=0) as ix:
        for klist, is_range in ((list(range(20, 201)), True), ([24, 36, 50, 100], False), (list(range(12, 256)), True)):
            max_k = max(klist)
            for seg in ora.segments(fa, batch, max_k - 1):
                n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
                want = o.min_unique_segment(seg.data, n, klist, is_range)
                got, st = ix.min_unique(seg.data, n, klist, is_range, dtype=np.uint32)
                bad = np.nonzero(got != want)[0]
                assert bad.size == 0, (klist[:2], is_range, seg.id, bad[:10], got[bad[:10]], want[bad[:10]])
                assert st.replayed == 0


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's index builder)")
def test_large_segment_properties_and_cross_algorithm_checksum(nm, tmp_path):
    """Full-size segment (10 M positions, the CLI default batch) where the oracle would take too
    long: size-independent properties instead -- the trajectory-free and the reference-trajectory
    kernels must produce identical bytes (device checksum == host checksum of the other), every
    non-zero answer k must satisfy count(k) == 1 and count(k-1) > 1 on a sample (checked with the
    count kernel and with the oracle), zeros at N positions."""
    fa, ixp = _synthetic(tmp_path, [12_000_000], 77, 11, 10_000_000)
    seq = dict(ora.fasta_records(fa))[b"chr1"]
    seg = seq[:10_000_000 + 199]
    kl = list(range(20, 201))
    o = ora.OracleIndex(ixp)
    with nm.Index(ixp, device=0) as ix:
        a, st_a = ix.min_unique(seg, 10_000_000, kl, True, algo="auto")
        b, st_b = ix.min_unique(seg, 10_000_000, kl, True, algo="trajectory")
        assert np.array_equal(a, b)
        assert st_a.lf_steps < st_b.lf_steps
        assert ix.derived_seed_length > ix.seed_length       # 12 M bases: nm_index_open derives a deeper table
        ix.set_seed_depth(0)
        a0, st_a0 = ix.min_unique(seg, 10_000_000, kl, True, algo="auto")
        assert np.array_equal(a0, a) and st_a0.ref_lf_steps > 0 and st_a.lf_steps < st_a0.lf_steps
        ix.set_seed_depth(16)
        assert ix.derived_seed_length == 16
        a16, st_a16 = ix.min_unique(seg, 10_000_000, kl, True, algo="auto")
        assert np.array_equal(a16, a) and st_a16.lf_steps <= st_a.lf_steps
        arr = np.frombuffer(seg, dtype=np.uint8)[:10_000_000]
        assert not a[arr == ord("N")].any()
        rng = np.random.default_rng(0)
        pos = rng.choice(np.nonzero(a > 0)[0], 3000, replace=False)
        k = a[pos].astype(np.uint64)
        tot = ix.count(seg, pos, k) + ix.count(seg, pos, k, reverse_complement=True)
        assert np.all(tot == 1)
        longer = k > 20
        tot2 = ix.count(seg, pos[longer], k[longer] - 1) + ix.count(seg, pos[longer], k[longer] - 1, reverse_complement=True)
        assert np.all(tot2 > 1)
        sub = pos[:300]
        assert np.array_equal(o.count(seg, sub, a[sub].astype(np.uint64)) + o.count(seg, sub, a[sub].astype(np.uint64), rc=True),
                              np.ones(300, dtype=np.uint32))
        # the other BASELINE configs at full segment size, trajectory-free against the reference-trajectory kernel:
        # explicit k list (configs[1]) and the widest range (configs[4]: k=12:255, below the derived seed depth)
        ix.set_seed_depth(int(os.environ.get("NEWMAP_B200_TEST_DEPTH", "14")))
        for klist, is_range in (([24, 36, 50, 100], False), (list(range(12, 256)), True)):
            m = 2_000_000
            sub = seq[3_000_000:3_000_000 + m + max(klist) - 1]
            x, st_x = ix.min_unique(sub, m, klist, is_range, algo="auto")
            y, st_y = ix.min_unique(sub, m, klist, is_range, algo="trajectory")
            assert np.array_equal(x, y), (klist[:2], is_range, np.nonzero(x != y)[0][:10])
            assert x.dtype == np.uint8 and st_x.lf_steps < st_y.lf_steps and x.any()
        # device-resident form: checksum computed on the GPU equals the host checksum of the bytes
        ws = nm.Workspace(ix, len(seg))
        buf = nm.PinnedBuffer(len(seg))
        buf.array[:] = np.frombuffer(seg, dtype=np.uint8)
        ws.upload(buf.ptr, len(seg))
        ws.search(10_000_000, kl, True, out_width=1)
        ws.sync()
        assert ws.checksum() == nm.checksum_host(a)
        ws.close()
        buf.close()


def test_tuning_options_do_not_change_results(nm, indexes):
    ix = indexes("quirk_s6")
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr1"]
    kl = list(range(20, 201))
    want, _ = ix.min_unique(seq, len(seq), kl, True, algo="trajectory")
    buf = nm.PinnedBuffer(len(seq))
    buf.array[:] = np.frombuffer(seq, dtype=np.uint8)
    out = nm.PinnedBuffer(len(seq))
    ws = nm.Workspace(ix, len(seq))
    for n, (dcap, gsize, pass1) in enumerate(((0, 0, 0), (0, 0, 1), (0, 0, 2), (7, 1, 2), (8, 5, 1), (13, 64, 2), (40, 200, 2),
                                              (40, 200, 1), (1000, 32, 2), (0, 16, 0), (9, 16, 2), (9, 4, 1))):
        ws.option("dcap", dcap)
        ws.option("gsize", gsize)
        ws.option("gsize2", (256, 0, 40, 1000)[n % 4])   # coarse certificate round: on, off (default), other sizes
        ws.option("own_chains", (1, 0, 3)[n % 3])
        ws.option("stall", (2, 0, 1, 5, 3)[n % 5])
        ws.option("cert_runs", (0, 1, 2, 7)[n % 4])
        ws.option("pass1", pass1)     # first-pass kernel: sparse with lane refill / one thread per position
        ws.option("stride", (0, 1, 2, 3, 5, 8, 4)[n % 7])   # sparse first pass: positions per unconditional table lookup
        ws.upload(buf.ptr, len(seq))
        ws.search(len(seq), kl, True, out_width=1)
        ws.download(out.ptr)
        st = ws.sync()
        assert st.replayed == 0
        assert np.array_equal(out.array[:len(seq)], want), (dcap, gsize, pass1)
        assert st.stride == (1 if pass1 == 1 else (8, 1, 2, 3, 5, 8, 4)[n % 7]), (n, st.stride)
    ws.close()
