"""Several FASTA files x several index files at once (search.py:260-265,461,489-517,656-697).

CPU: the oracle (oracle/fm_oracle.c ora_min_unique_multi) and the host-compiled kernel body
(trajectory_multi_thread) against goldens the unmodified reference wrote (oracle/gen_multi_golden.py).
GPU: nm_workspace_multi through the C-ABI and `newmap search` with comma-separated inputs against the same goldens."""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle import ora
import hostemu

MG = Path(__file__).resolve().parent / "golden" / "multi"
MANIFEST = json.loads((MG / "manifest.json").read_text())
GOLD = np.load(MG / "searches.npz")


def _k(run):
    return list(range(run["k"][0], run["k"][1] + 1)) if run["range"] else run["k"]


def _tuples(run):
    max_k = max(_k(run))
    for tup in zip(*[ora.segments(MG / f, run["batch"], max_k - 1) for f in run["fastas"]]):
        seg = tup[0]
        yield tup, (len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1))


@pytest.mark.parametrize("run", MANIFEST, ids=[f"run{r['id']}" for r in MANIFEST])
def test_oracle_multi_matches_reference_goldens(run):
    ixs = [ora.OracleIndex(MG / i) for i in run["indexes"]]
    got = ora.min_unique_multi([MG / f for f in run["fastas"]], ixs, _k(run), run["range"], run["batch"],
                               run.get("initial", 0), not run.get("norc", False))
    assert sorted(k.decode() for k in got) == sorted(run["sequences"])
    for sid, arr in got.items():
        assert np.array_equal(arr, GOLD[f"run{run['id']}|{sid.decode()}"]), (run, sid)


@pytest.mark.parametrize("run", MANIFEST, ids=[f"run{r['id']}" for r in MANIFEST])
def test_kernel_body_multi_matches_reference_goldens(run):
    ixs = [hostemu.EmuIndex(MG / i) for i in run["indexes"]]
    out = {}
    for tup, n in _tuples(run):
        vals, st = hostemu.min_unique_multi_segment(ixs, [t.data for t in tup], n, _k(run), run["range"],
                                                    run.get("initial", 0), not run.get("norc", False))
        assert st[2] == 0 and st[4] == -1
        out.setdefault(tup[0].id.decode(), []).append(vals)
    for sid, parts in out.items():
        gold = GOLD[f"run{run['id']}|{sid}"]
        assert np.array_equal(np.concatenate(parts).astype(gold.dtype), gold), (run, sid)


def test_multi_zero_and_short_counts():
    """A sequence that is in none of the indexes: zero count reported like search.py:699-722; a summed count between
    0 and the number of sequences is flagged (the reference's loop would not end)."""
    ct, ga = ora.OracleIndex(MG / "ct.awfmi"), ora.OracleIndex(MG / "partB.awfmi")
    segs = list(ora.segments(MG / "ct.fa", 1_000_000, 19))
    other = list(ora.segments(MG / "base.fa", 1_000_000, 19))
    # two sequences, one index that holds only the first: every sum is >= 1 but mostly < 2
    with pytest.raises(ora.ShortCount):
        ora.min_unique_multi_segment([ct], [segs[0].data, other[0].data], len(segs[0].data), list(range(12, 21)), True)
    e = [hostemu.EmuIndex(MG / "ct.awfmi")]
    _, st = hostemu.min_unique_multi_segment(e, [segs[0].data, other[0].data], len(segs[0].data), list(range(12, 21)), True)
    assert st[2] > 0
    # an index that does not hold the sequence at all: first zero count on the trajectory
    with pytest.raises(ora.ZeroCount) as ei:
        ora.min_unique_multi_segment([ga], [segs[0].data], len(segs[0].data), list(range(12, 21)), True)
    _, st = hostemu.min_unique_multi_segment([hostemu.EmuIndex(MG / "partB.awfmi")], [segs[0].data], len(segs[0].data),
                                             list(range(12, 21)), True)
    p, ln = int(st[4]), int(st[5])
    assert p >= 0 and bytes(segs[0].data[p:p + ln]).decode() in str(ei.value)


@pytest.mark.gpu
@pytest.mark.parametrize("run", MANIFEST, ids=[f"run{r['id']}" for r in MANIFEST])
def test_gpu_multi_search_matches_reference_goldens(run, tmp_path):
    from newmap_b200 import main
    k = f"{run['k'][0]}:{run['k'][1]}" if run["range"] else ",".join(map(str, run["k"]))
    argv = ["search", "--search-range", k, "--kmer-batch-size", str(run["batch"]), "--output-directory", str(tmp_path)]
    if run.get("norc"):
        argv.append("--norc")
    if run.get("initial"):
        argv += ["--initial-search-length", str(run["initial"])]
    argv += [",".join(str(MG / f) for f in run["fastas"]), ",".join(str(MG / i) for i in run["indexes"])]
    main.parse_subcommands(argv)
    for sid in run["sequences"]:
        gold = GOLD[f"run{run['id']}|{sid}"]
        got = np.fromfile(tmp_path / f"{sid}.unique.{gold.dtype.name}", dtype=gold.dtype)
        assert np.array_equal(got, gold), (run, sid)


@pytest.mark.gpu
def test_gpu_multi_errors_and_api(tmp_path):
    from newmap_b200 import _lib, search
    cfg = search.SearchConfig(fasta_filepaths=[MG / "ct.fa", MG / "base.fa"], fmindex_filepaths=[MG / "ct.awfmi", MG / "ga.awfmi"],
                              kmer_lengths=list(range(12, 21)), is_binary_search=True, output_directory=tmp_path)
    with pytest.raises(RuntimeError, match="below the number of sequences|not found in the index"):
        search.write_unique_counts(cfg)
    cfg = search.SearchConfig(fasta_filepaths=[MG / "ct.fa"], fmindex_filepaths=[MG / "partB.awfmi", MG / "partB.awfmi"],
                              kmer_lengths=list(range(12, 21)), is_binary_search=True, output_directory=tmp_path)
    with pytest.raises(RuntimeError, match="not found in the index"):
        search.write_unique_counts(cfg)
    # more sequences than indexes: search.py:664-668 indexes the index list by sequence number
    cfg = search.SearchConfig(fasta_filepaths=[MG / "ct.fa", MG / "ga.fa"], fmindex_filepaths=[MG / "ct.awfmi"],
                              kmer_lengths=list(range(12, 21)), is_binary_search=True, output_directory=tmp_path)
    with pytest.raises(IndexError):
        search.write_unique_counts(cfg)
    # get_kmer_counts sums over every index and sequence (search.py:647-697)
    cfg = search.SearchConfig(fasta_filepaths=[MG / "ct.fa", MG / "ga.fa"], fmindex_filepaths=[MG / "ct.awfmi", MG / "ga.awfmi"],
                              kmer_lengths=[16], is_binary_search=False, output_directory=tmp_path)
    a = next(ora.segments(MG / "ct.fa", 1_000_000, 15)).data
    b = next(ora.segments(MG / "ga.fa", 1_000_000, 15)).data
    idx, lens = [20, 100, 1400, 1405], [16, 12, 30, 8]
    got = search.get_kmer_counts(cfg, [a, b], idx, lens)
    want = np.zeros(len(idx), dtype=np.uint32)
    for ixp in ("ct.awfmi", "ga.awfmi"):
        o = ora.OracleIndex(MG / ixp)
        for s in (a, b):
            want += o.count(s, idx, lens) + o.count(s, idx, lens, rc=True)
    assert np.array_equal(got, want)
    # argument checks of the C-ABI entry
    ix = _lib.Index(MG / "ct.awfmi", seed_depth=0)
    ws = _lib.Workspace(ix, 1000)
    with pytest.raises(ValueError):
        ws.multi([ix] * 9, [])
    with pytest.raises(ValueError):
        ws.multi([], [])
