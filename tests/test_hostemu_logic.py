"""CPU-side check of the CUDA path's LOGIC: the per-thread kernel bodies (nm_core.cuh) compiled
for the host and run serially (tests/hostemu) against the oracle and the reference-made golden
fixtures.  The same comparisons run on the real kernels in tests/test_gpu_parity.py."""
import numpy as np
import pytest

from conftest import GOLD, run_klist
from hostemu import EmuIndex
from oracle import ora

ALGOS = {"auto": 0, "trajectory": 1, "free_generic": 2}


def _segments(run):
    max_k = run["k"][1] if run["range"] else max(run["k"])
    for seg in ora.segments(GOLD / run["fasta"], run["batch"], max_k - 1):
        n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
        yield seg, n


@pytest.mark.parametrize("algo", list(ALGOS))
def test_golden_searches(manifest, golden_searches, algo):
    cache = {}
    for run in manifest["searches"]:
        ix = cache.setdefault(run["index"], EmuIndex(GOLD / f"{run['index']}.awfmi"))
        klist = run_klist(run)
        got = {}
        for seg, n in _segments(run):
            out, stats = ix.min_unique_segment(seg.data, n, klist, run["range"], run.get("initial", 0),
                                               not run.get("norc", False), ALGOS[algo])
            assert stats[3] == 0 and stats[4] == -1, (run["id"], stats)
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            bad = np.nonzero(arr != want)[0]
            assert bad.size == 0, (algo, run, sid, bad[:10], arr[bad[:10]], want[bad[:10]])


def test_golden_counts(manifest, golden_counts):
    for c in manifest["counts"]:
        ix = EmuIndex(GOLD / f"{c['index']}.awfmi")
        o = ora.OracleIndex(GOLD / f"{c['index']}.awfmi")
        seq = dict(ora.fasta_records(GOLD / c["fasta"]))[c["sequence"].encode()]
        idx, ln = golden_counts[c["key"] + "|idx"], golden_counts[c["key"] + "|len"]
        got, lf = ix.count(seq, idx, ln)
        assert np.array_equal(got, golden_counts[c["key"] + "|count"])
        o.lf_steps = 0
        assert np.array_equal(ix.count(seq, idx, ln, rc=True)[0], o.count(seq, idx, ln, rc=True))


def test_trajectory_lf_steps_and_zero_event_match_oracle():
    ix = EmuIndex(GOLD / "quirk_s4.awfmi")
    o = ora.OracleIndex(GOLD / "quirk_s4.awfmi")
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr1"]
    seg = seq[:1199]
    o.lf_steps = o.queries = 0
    want = o.min_unique_segment(seg, 1000, list(range(20, 201)), True, nthreads=1)
    out, stats = ix.min_unique_segment(seg, 1000, list(range(20, 201)), True, algo=1)
    assert np.array_equal(out, want)
    assert stats[0] == o.lf_steps and stats[1] == o.queries
    # a sequence that is not in the index: same first zero-count k-mer as the oracle reports
    rng = np.random.default_rng(1)
    alien = bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 600)])
    with pytest.raises(ora.ZeroCount) as e:
        o.min_unique_segment(alien, 500, list(range(20, 101)), True, nthreads=1)
    for algo in (0, 1):
        out, stats = ix.min_unique_segment(alien, 500, list(range(20, 101)), True, algo=algo)
        if algo == 0:
            assert stats[3] == 1     # the trajectory-free search flags the anomaly; nm_lib.cu then replays
        else:
            p, ln = int(stats[4]), int(stats[5])
            assert ora.zero_count_message(alien[p:p + ln].decode()) == str(e.value)


def test_ref_lf_estimate_matches_trajectory_on_clean_sequence():
    """stats.ref_lf_steps (LF steps of the reference's trajectory, derived from the thresholds) equals
    the steps the trajectory kernel really executes when no query contains an ambiguous base."""
    ix = EmuIndex(GOLD / "quirk_s4.awfmi")
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr2"]
    seg = seq[0:2400 + 39]      # no N inside (first N of chr2 is at 2511)
    assert b"N" not in seg
    kl = list(range(8, 41))
    _, st_t = ix.min_unique_segment(seg, 2400, kl, True, algo=1)
    _, st_f = ix.min_unique_segment(seg, 2400, kl, True, algo=0)
    assert st_f[7] == st_t[0]


def test_random_small_genomes_all_algorithms_agree_with_oracle(tmp_path):
    if not ora.ref_available():
        pytest.skip("needs oracle/_ref to build fresh indexes")
    rng = np.random.default_rng(2024)
    for trial in range(6):
        n1, n2 = int(rng.integers(200, 3000)), int(rng.integers(50, 800))
        alphabet = np.frombuffer(b"ACGT" if trial % 2 else b"ACGTacgtNNn", dtype=np.uint8)
        recs = []
        for name, n in (("s1", n1), ("s2", n2)):
            a = alphabet[rng.integers(0, alphabet.size, n)].copy()
            if n > 300:   # plant repeats so that long k-mers stay non-unique
                a[n // 2:n // 2 + 120] = a[10:130]
            recs.append((name, a))
        fa = tmp_path / f"t{trial}.fa"
        with open(fa, "wb") as fh:
            for name, a in recs:
                fh.write(b">" + name.encode() + b"\n" + a.tobytes() + b"\n")
        seed = int(rng.integers(1, 7))
        ixp = tmp_path / f"t{trial}.awfmi"
        ora.ref_build_index(fa, ixp, 8, seed)
        o, e = ora.OracleIndex(ixp), EmuIndex(ixp)
        deep = EmuIndex(ixp).set_seed_depth(seed + 1 + trial % 3)
        for klist, is_range in ((list(range(3, 40)), True), (list(range(seed, 70)), True), ([5, 9, 17, 40], False),
                                ([30, 12, 12, 50], False)):
            batch = int(rng.integers(20, 500))
            max_k = max(klist)
            for seg in ora.segments(fa, batch, max_k - 1):
                n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
                want = o.min_unique_segment(seg.data, n, klist, is_range, nthreads=1)
                for algo in (0, 1, 2):
                    got, st = e.min_unique_segment(seg.data, n, klist, is_range, algo=algo)
                    assert np.array_equal(got, want), (trial, klist[:3], is_range, algo, batch)
                got, st = deep.min_unique_segment(seg.data, n, klist, is_range, algo=0)
                assert np.array_equal(got, want), (trial, klist[:3], is_range, "derived seed table", batch)


@pytest.mark.parametrize("dcap_extra,gsize,gsize2", [(1, 1, 0), (1, 5, 40), (2, 32, 0), (4, 64, 256), (12, 32, 33), (3, 200, 0),
                                                      (1000, 32, 64), (2, 16, 256), (1, 3, 1000), (4, 16, 64)])
def test_certificate_pass_is_exact_for_any_cap_and_group(manifest, golden_searches, dcap_extra, gsize, gsize2):
    """The depth cap, the certificate group size and the group size of the coarse certificate round are tuning knobs:
    results never depend on them."""
    cache = {}
    for run in manifest["searches"]:
        if run["index"].startswith("genome"):
            continue
        ix = cache.setdefault(run["index"], EmuIndex(GOLD / f"{run['index']}.awfmi"))
        seed = manifest["indexes"][run["index"]]["seed_length"]
        klist = run_klist(run)
        got = {}
        for seg, n in _segments(run):
            out, stats = ix.min_unique_segment(seg.data, n, klist, run["range"], run.get("initial", 0),
                                               not run.get("norc", False), 0, dcap=seed + dcap_extra, gsize=gsize,
                                               gsize2=gsize2, stall=(dcap_extra + gsize) % 4)
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            bad = np.nonzero(arr != want)[0]
            assert bad.size == 0, (dcap_extra, gsize, run, sid, bad[:10], arr[bad[:10]], want[bad[:10]])


@pytest.mark.parametrize("extra", [1, 2, 4, 6])
def test_derived_seed_table_never_changes_a_result(manifest, golden_searches, extra):
    """The deeper seed table derived at load time (nm_index_set_seed_depth) is a speed matter only: every golden
    search whose shortest length reaches the derived depth starts its chains from it and returns the same bytes."""
    cache, used = {}, 0
    for run in manifest["searches"]:
        seed = manifest["indexes"][run["index"]]["seed_length"]
        depth = seed + extra
        ix = cache.get(run["index"])
        if ix is None:
            ix = cache[run["index"]] = EmuIndex(GOLD / f"{run['index']}.awfmi").set_seed_depth(depth)
        klist = run_klist(run)
        used += min(klist) >= depth
        got = {}
        for seg, n in _segments(run):
            out, stats = ix.min_unique_segment(seg.data, n, klist, run["range"], run.get("initial", 0),
                                               not run.get("norc", False), 0)
            assert stats[3] == 0 and stats[4] == -1, (run["id"], stats)
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            bad = np.nonzero(arr != want)[0]
            assert bad.size == 0, (extra, run, sid, bad[:10], arr[bad[:10]], want[bad[:10]])
    assert used >= 10


def test_derived_seed_table_entries_equal_counts():
    """Every entry of the derived table is the SA interval of its d-mer: length == occurrence count."""
    ix = EmuIndex(GOLD / "quirk_s4.awfmi").set_seed_depth(7)
    o = ora.OracleIndex(GOLD / "quirk_s4.awfmi")
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr1"][:1500]
    # a 7-mer search seeded at depth 7 reports G == 7 exactly where the 7-mer is unique on the forward strand
    idx = np.arange(0, len(seq) - 7)
    cnt = o.count(seq, idx, np.full(idx.size, 7))
    out, _ = ix.min_unique_segment(seq, len(seq) - 6, [7, 7], True, use_rc=False, algo=0)
    clean = np.array([all(c in b"ACGTacgt" for c in seq[i:i + 7]) for i in idx])
    assert np.array_equal((out[:idx.size] == 7)[clean], (cnt == 1)[clean])


def test_rc_seed_key_bit_trick():
    import ctypes as C
    from hostemu import lib
    L = lib()
    L.emu_rc_seed_key.restype = C.c_uint32
    L.emu_rc_seed_key.argtypes = [C.c_uint32, C.c_uint32]
    rng = np.random.default_rng(5)
    for sl in range(1, 17):
        for _ in range(50):
            letters = rng.integers(0, 4, sl)
            key = 0
            for c in letters:
                key = (key << 2) | int(c)
            want = 0
            for j, c in enumerate(letters):   # c[i] least significant, complemented
                want |= (3 - int(c)) << (2 * j)
            assert L.emu_rc_seed_key(key, sl) == want


def test_diverged_repeats_interpolation_certificate(tmp_path):
    """Copies of a repeat that differ at a few positions (forward and reverse-complement copies, two and three
    copies): between two differences the shortest unique length grows by one per position, which the certificate
    pass settles from one chain per group plus the neighbour's value.  Every cap / group size gives the oracle's bytes."""
    if not ora.ref_available():
        pytest.skip("needs oracle/_ref to build fresh indexes")
    rng = np.random.default_rng(77)
    comp = np.zeros(256, dtype=np.uint8)
    comp[list(b"ACGT")] = list(b"TGCA")
    for trial in range(4):
        n = 6000
        a = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, n)].copy()
        unit = a[100:700].copy()

        def mutate(u, every):
            u = u.copy()
            for pos in range(int(rng.integers(5, every)), u.size, every):
                u[pos] = (b"ACGT".replace(bytes([u[pos]]), b""))[int(rng.integers(0, 3))]
            return u
        a[1500:2100] = mutate(unit, 90)                       # second copy, a difference every 90
        a[3000:3600] = mutate(unit, 37)                       # third copy, other differences
        a[4200:4800] = comp[mutate(unit, 55)][::-1]           # reverse-complement copy
        if trial % 2:
            a[5000:5005] = ord("N")
        fa = tmp_path / f"d{trial}.fa"
        fa.write_bytes(b">s\n" + a.tobytes() + b"\n")
        seed = 3 + trial
        ixp = tmp_path / f"d{trial}.awfmi"
        ora.ref_build_index(fa, ixp, 8, seed)
        o, e = ora.OracleIndex(ixp), EmuIndex(ixp)
        if trial >= 2:
            e.set_seed_depth(seed + 2)
        klist = list(range(seed + 2, 121))
        for batch in (n, 1000):
            for seg in ora.segments(fa, batch, max(klist) - 1):
                m = len(seg.data) if seg.epilogue else len(seg.data) - (max(klist) - 1)
                want = o.min_unique_segment(seg.data, m, klist, True, nthreads=1)
                for dcap_extra, gsize in ((1, 32), (3, 8), (2, 64), (6, 5), (1, 1)):
                    got, st = e.min_unique_segment(seg.data, m, klist, True, algo=0, dcap=seed + 2 + dcap_extra, gsize=gsize,
                                                   gsize2=(0, 48, 256)[(dcap_extra + gsize) % 3], stall=(1, 2, 0, 3)[gsize % 4])
                    bad = np.nonzero(got != want)[0]
                    assert bad.size == 0, (trial, batch, dcap_extra, gsize, bad[:10], got[bad[:10]], want[bad[:10]])


@pytest.mark.parametrize("stride", [0, 2, 3, 5, 8, -1])
def test_sparse_first_pass_is_exact_for_any_stride(manifest, golden_searches, stride):
    """The sparse first pass (one seed window per `stride` positions looked up unconditionally, upper bounds
    G[e+1] <= G[e]+1 / R[i-1] <= R[i]+1 spread from it, chains of their own only for what no bound settles) gives the
    reference's bytes for every stride, in range and in list mode (incl. the end-of-buffer truncation of list mode,
    where a bound would not do), and looks up fewer windows than the dense pass."""
    cache, sparse_runs, lookups = {}, 0, [0, 0]
    for run in manifest["searches"]:
        seed = manifest["indexes"][run["index"]]["seed_length"]
        klist = run_klist(run)
        depth = min(seed + 1 + run["id"] % 3, 16)
        key = (run["index"], depth)
        ix = cache.get(key)
        if ix is None:
            ix = cache[key] = EmuIndex(GOLD / f"{run['index']}.awfmi").set_seed_depth(depth)
        got = {}
        for seg, n in _segments(run):
            out, stats = ix.min_unique_segment(seg.data, n, klist, run["range"], run.get("initial", 0),
                                               not run.get("norc", False), 0, stride=stride,
                                               dcap=depth + 1 + run["id"] % 7, gsize=(20, 5, 64)[run["id"] % 3])
            assert stats[3] == 0 and stats[4] == -1, (run["id"], stats)
            if min(klist) > depth and stride >= 0:
                want_stride = min(min(klist) - depth + 1, 8, stride or 8)
                assert stats[9] == want_stride, (run["id"], stats[9], want_stride)
                sparse_runs += want_stride > 1
                if stride == 0 and want_stride > 1:
                    lookups[0] += stats[8]
                    lookups[1] += ix.min_unique_segment(seg.data, n, klist, run["range"], run.get("initial", 0),
                                                        not run.get("norc", False), 0, stride=-1)[1][8]
            got.setdefault(seg.id, []).append(out)
        for sid, parts in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            arr = np.concatenate(parts)
            bad = np.nonzero(arr != want)[0]
            assert bad.size == 0, (stride, run, sid, bad[:10], arr[bad[:10]], want[bad[:10]])
    if stride >= 0:
        assert sparse_runs >= 10
    if stride == 0:
        assert lookups[0] < 0.8 * lookups[1], lookups      # fewer table lookups than the dense pass makes


def test_range_that_starts_below_the_table_depth(manifest, golden_searches, tmp_path):
    """k = a:b with a below the depth of the derived table (BASELINE configs[4]: 12:255 on the 16-letter table): chains
    start from the derived table, lengths below its depth are settled with exact counts where the table does not rule
    them out (subdepth_candidate).  Same bytes as the reference's for every golden range run that qualifies, with the
    path on and off (bit 30 of the stride argument = NM_OPT_SUBDEPTH 1), and for random genomes against the oracle."""
    cache, used = {}, 0
    for run in manifest["searches"]:
        if not run["range"]:
            continue
        seed = manifest["indexes"][run["index"]]["seed_length"]
        klist = run_klist(run)
        for depth in {min(max(klist), min(klist) + 1 + run["id"] % 2, 11), min(max(klist), min(klist) + 3, 11)}:   # (host RAM: 16 B x 4^depth)
            if not (seed <= min(klist) < depth <= max(klist)) or depth <= seed:
                continue
            ix = cache.get((run["index"], depth))
            if ix is None:
                ix = cache[(run["index"], depth)] = EmuIndex(GOLD / f"{run['index']}.awfmi").set_seed_depth(depth)
            for off in (0, 0x40000000):
                got = {}
                for seg, n in _segments(run):
                    out, stats = ix.min_unique_segment(seg.data, n, klist, True, run.get("initial", 0),
                                                       not run.get("norc", False), 0, stride=off)
                    assert stats[3] == 0 and stats[4] == -1, (run["id"], stats)
                    got.setdefault(seg.id, []).append(out)
                for sid, parts in got.items():
                    want = golden_searches[f"run{run['id']}|{sid.decode()}"]
                    arr = np.concatenate(parts)
                    bad = np.nonzero(arr != want)[0]
                    assert bad.size == 0, (depth, off, run, sid, bad[:10], arr[bad[:10]], want[bad[:10]])
            used += 1
    assert used >= 8, used
    if not ora.ref_available():
        return
    rng = np.random.default_rng(99)
    for trial in range(4):
        n = int(rng.integers(3000, 9000))
        alphabet = np.frombuffer(b"ACGT" if trial % 2 else b"ACGTacgtN", dtype=np.uint8)
        a = alphabet[rng.integers(0, alphabet.size, n)].copy()
        a[n // 2:n // 2 + 300] = a[100:400]
        fa = tmp_path / f"s{trial}.fa"
        fa.write_bytes(b">s\n" + a.tobytes() + b"\n")
        ixp = tmp_path / f"s{trial}.awfmi"
        ora.ref_build_index(fa, ixp, 8, 3)
        o = ora.OracleIndex(ixp)
        e = EmuIndex(ixp).set_seed_depth(7 + trial % 2)         # a 6 k genome: many 5- and 6-mers are already unique
        for klist in (list(range(3, 40)), list(range(5, 12)), list(range(4, 8 + trial % 2))):
            for batch in (n, 977):
                for seg in ora.segments(fa, batch, max(klist) - 1):
                    m = len(seg.data) if seg.epilogue else len(seg.data) - (max(klist) - 1)
                    want = o.min_unique_segment(seg.data, m, klist, True, nthreads=1)
                    for norc in (False, True):
                        if norc:
                            want = o.min_unique_segment(seg.data, m, klist, True, use_rc=False, nthreads=1)
                        got, st = e.min_unique_segment(seg.data, m, klist, True, use_rc=not norc, algo=0)
                        bad = np.nonzero(got != want)[0]
                        assert bad.size == 0, (trial, klist[0], klist[-1], batch, norc, bad[:10], got[bad[:10]], want[bad[:10]])
                        assert (want[want > 0] < 7).any()           # answers below the table depth do occur here


def test_strict_mode_sees_the_mismatch_the_default_algorithm_misses(tmp_path):
    """The zero-count contract (search.py:699-722).  Two copies of a repeat differ at two sites 60 apart; the FASTA
    that is searched carries, in one copy, the other copy's letter at the first site.  Every k-mer that spans both
    sites is absent from the index: the reference raises.  The chains of the trajectory-free passes stop at the first
    site they meet (one copy left = unique), so the default algorithm sees nothing; NM_OPT_STRICT does, and the replay
    on the reference's trajectory then reports the reference's own k-mer.  On a matching FASTA strict mode is silent."""
    if not ora.ref_available():
        pytest.skip("needs oracle/_ref to build a fresh index")
    rng = np.random.default_rng(5)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    a = acgt[rng.integers(0, 4, 5000)].copy()
    unit = a[1000:1600].copy()
    copy = unit.copy()
    y1, y2 = 300, 360
    for y in (y1, y2):
        copy[y] = acgt[(int(np.nonzero(acgt == unit[y])[0][0]) + 1) % 4]
    a[3000:3600] = copy
    fa = tmp_path / "g.fa"
    fa.write_bytes(b">s\n" + a.tobytes() + b"\n")
    ixp = tmp_path / "g.awfmi"
    ora.ref_build_index(fa, ixp, 8, 4)
    searched = a.copy()
    searched[1000 + y1] = copy[y1]                      # first copy now reads like the second one at y1, not at y2
    kl = list(range(20, 201))
    o, e = ora.OracleIndex(ixp), EmuIndex(ixp).set_seed_depth(8)
    with pytest.raises(ora.ZeroCount) as err:
        o.min_unique_segment(searched.tobytes(), searched.size, kl, True, nthreads=1)
    STRICT = 0x20000000
    _, st = e.min_unique_segment(searched.tobytes(), searched.size, kl, True, algo=0)
    assert st[3] == 0                                    # the default algorithm does not notice
    _, st = e.min_unique_segment(searched.tobytes(), searched.size, kl, True, algo=0, stride=STRICT)
    assert st[3] == 1                                    # strict mode does -> nm_workspace_sync replays on the trajectory
    _, st = e.min_unique_segment(searched.tobytes(), searched.size, kl, True, algo=1)
    p, ln = int(st[4]), int(st[5])
    assert ora.zero_count_message(searched[p:p + ln].tobytes().decode()) == str(err.value)
    for dcap, gsize in ((0, 0), (9, 7), (11, 33)):       # (they also pick the block size of the presence chains here)
        want = o.min_unique_segment(a.tobytes(), a.size, kl, True, nthreads=1)
        got, st = e.min_unique_segment(a.tobytes(), a.size, kl, True, algo=0, stride=STRICT, dcap=dcap, gsize=gsize)
        assert st[3] == 0 and np.array_equal(got, want)
