"""`newmap track` (SURVEY.md §8 f N2): the numpy restatement (oracle/track_oracle.py) against files the unmodified
reference wrote (tests/golden/track/, oracle/gen_track_golden.py) and against the reference itself where it exists;
the GPU path (newmap_b200/track.py -> nm_track_*) against both.  Text output: byte-exact."""
import json
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import GOLD
from oracle import track_oracle as T

TG = GOLD / "track"
CASES = json.loads((TG / "manifest.json").read_text())
DT = {".uint8": np.uint8, ".uint16": np.uint16, ".uint32": np.uint32}


def _load(name):
    return np.fromfile(TG / name, dtype=DT[Path(name).suffix]), name[:name.find(".unique")]


@pytest.mark.parametrize("case", CASES, ids=[c["bed"][:-4] for c in CASES])
def test_oracle_reproduces_reference_files(case):
    bed = b"".join(T.bed_text(*(lambda a, n: (a, case["k"], n))(*_load(f))) for f in case["files"])
    wig = b"".join(T.wig_text(*(lambda a, n: (a, case["k"], n))(*_load(f))) for f in case["files"])
    assert bed == (TG / case["bed"]).read_bytes()
    assert wig == (TG / case["wig"]).read_bytes()


@pytest.mark.skipif(not Path("/root/reference/newmap/track.py").is_file(), reason="needs the reference checkout")
def test_oracle_against_live_reference(tmp_path):
    sys.path.insert(0, "/root/reference")
    try:
        from newmap import track as ref
    finally:
        sys.path.remove("/root/reference")
    rng = np.random.default_rng(3)
    for k in (1, 7, 24, 99, 100, 101, 255):
        a = rng.integers(0, 256, 4000).astype(np.uint8)
        a[rng.random(a.size) < 0.5] = 0
        p = tmp_path / f"s{k}.unique.uint8"
        a.tofile(p)
        assert np.array_equal(ref.create_multiread_mappability_from_unique_file(p, k, np.uint8), T.multiread(a, k))
        ref.write_mappability_files([p], k, str(tmp_path / "o.bed"), str(tmp_path / "o.wig"), False)
        assert (tmp_path / "o.bed").read_bytes() == T.bed_text(a, k, f"s{k}")
        assert (tmp_path / "o.wig").read_bytes() == T.wig_text(a, k, f"s{k}")


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c["bed"][:-4] for c in CASES])
def test_gpu_track_reproduces_reference_files(case, tmp_path):
    from newmap_b200 import track
    bed, wig = tmp_path / "o.bed", tmp_path / "o.wig"
    bed.write_bytes(b"stale")                  # removed first, like track.py:150-151
    track.write_mappability_files([TG / f for f in case["files"]], case["k"], str(bed), str(wig), False)
    assert bed.read_bytes() == (TG / case["bed"]).read_bytes()
    assert wig.read_bytes() == (TG / case["wig"]).read_bytes()
    for f in case["files"]:
        a, _ = _load(f)
        got = track.create_multiread_mappability_from_unique_file(TG / f, case["k"], a.dtype)
        assert got.dtype == np.float64 and np.array_equal(got, T.multiread(a, case["k"]))


@pytest.mark.gpu
def test_gpu_track_large_random_and_cli(tmp_path, capsysbinary):
    from newmap_b200 import main, track
    rng = np.random.default_rng(8)
    for dtype, hi, k, n in ((np.uint8, 256, 24, 3_000_000), (np.uint16, 600, 150, 1_000_003), (np.uint32, 300, 100, 70_000)):
        a = rng.integers(0, hi, n).astype(dtype)
        a[rng.random(n) < 0.6] = 0
        for s in rng.integers(0, n - 5000, 40):
            a[s:s + int(rng.integers(1, 5000))] = 0
        p = tmp_path / f"chrR{k}.unique.{np.dtype(dtype).name}"
        a.tofile(p)
        with track.Track() as tr:
            tr.compute(a, k)
            assert np.array_equal(tr.counts(), T.window_counts(a, k).astype(np.uint32))
            assert np.array_equal(tr.counts(1000, 77), T.window_counts(a, k)[1000:1077].astype(np.uint32))
        bed, wig = tmp_path / f"r{k}.bed", tmp_path / f"r{k}.wig"
        main.parse_subcommands(["track", str(k), str(p), "-s", str(bed), "-m", str(wig)])
        assert bed.read_bytes() == T.bed_text(a, k, f"chrR{k}")
        assert wig.read_bytes() == T.wig_text(a, k, f"chrR{k}")
    # default read length 24 when the first positional is a file; BED to standard output by default (track.py:256-277)
    p = tmp_path / "chrR24.unique.uint8"
    capsysbinary.readouterr()
    main.parse_subcommands(["track", str(p)])
    assert capsysbinary.readouterr().out == T.bed_text(np.fromfile(p, dtype=np.uint8), 24, "chrR24")
    with pytest.raises(ValueError):
        track.write_mappability_files([p], 24, "-", "-", False)
    with pytest.raises(ValueError):
        track.write_mappability_files([tmp_path / "x.unique.uint64"], 24, "a", "", False)


@pytest.mark.gpu
def test_gpu_track_chunked_text(tmp_path, monkeypatch):
    """BED and WIG text assembled in several device chunks equals the one-chunk text."""
    from newmap_b200 import track
    rng = np.random.default_rng(9)
    n, k = 200_003, 37
    a = rng.integers(0, 80, n).astype(np.uint8)
    a[rng.random(n) < 0.5] = 0
    p = tmp_path / "chrC.unique.uint8"
    a.tofile(p)
    monkeypatch.setattr(track, "_RUN_CHUNK", 1001)
    monkeypatch.setattr(track, "_TEXT_CHUNK", 7777)
    bed, wig = tmp_path / "c.bed", tmp_path / "c.wig"
    track.write_mappability_files([p], k, str(bed), str(wig), False)
    assert bed.read_bytes() == T.bed_text(a, k, "chrC")
    assert wig.read_bytes() == T.wig_text(a, k, "chrC")
    with track.Track() as tr:
        tr.compute(a, k)
        nruns = tr.count_runs()
        starts, values = tr.runs()
        assert nruns == starts.size
        assert bytes(tr.bed_text("chrC", 0, 0)) == b""
        last = bytes(tr.bed_text("chrC", nruns - 1, 1))
        assert last == f"chrC\t{int(starts[-1])}\t{n}\tk{k}\t{int(values[-1])}\t.\n".encode()
