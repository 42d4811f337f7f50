"""INTEGRATION.md option A, proven where the reference's source exists (the build container): the UNMODIFIED
/root/reference/newmap/search.py runs with the name its line 12 imports
(`from newmap._c_newmap_count_kmers import count_kmers_from_sequence`) bound to the product's drop-in
newmap_b200.count.count_kmers_from_sequence, and must write the files the reference wrote with its own extension
(tests/golden/searches.npz).  There is no GPU here, so the resident index behind the shim is the host-compiled
kernel body of the count kernel (tests/hostemu: count_one of nm_core.cuh, the code k_count runs); the argument
checking, list conversion and both-strand call pattern are the shim's own.  The same host loop drives the real
CUDA library in tests/test_gpu_count_shim.py."""
import numpy as np
import pytest

from conftest import GOLD, run_klist
from hostemu import EmuIndex
from oracle import pyref

pytestmark = pytest.mark.skipif(not pyref.available(), reason="needs /root/reference and oracle/_ref")


class _HostResidentIndex:
    """What newmap_b200.count.open_index returns, with EmuIndex in the place of the HBM-resident index."""
    _cache: dict = {}

    def __init__(self, path):
        self.emu = self._cache.setdefault(str(path), EmuIndex(path))

    def count(self, seq, idx, length, reverse_complement=False):
        return self.emu.count(bytes(seq), idx, length, rc=reverse_complement)[0]


@pytest.fixture()
def reference_search_on_shim(monkeypatch):
    from newmap_b200 import count as shim
    search, _ = pyref.load()
    monkeypatch.setattr(shim, "open_index", lambda path, *a, **k: _HostResidentIndex(path))
    monkeypatch.setattr(search, "count_kmers_from_sequence", shim.count_kmers_from_sequence)   # search.py:12
    return search


@pytest.mark.parametrize("fasta", ["genome.fa", "quirk.fa"])
def test_unmodified_reference_search_writes_the_golden_files_through_the_shim(reference_search_on_shim, manifest,
                                                                              golden_searches, tmp_path, fasta):
    done = 0
    for run in manifest["searches"]:
        if run["fasta"] != fasta or run.get("include") or run.get("exclude"):
            continue
        if fasta == "quirk.fa" and run["index"] != "quirk_s4":
            continue
        klist = run_klist(run)
        out = tmp_path / f"run{run['id']}"
        got = pyref.run_search(GOLD / fasta, GOLD / f"{run['index']}.awfmi", klist, run["range"], run["batch"], out,
                               initial_search_length=run.get("initial", 0), use_rc=not run.get("norc", False))
        assert got
        for sid, arr in got.items():
            want = golden_searches[f"run{run['id']}|{sid.decode()}"]
            assert np.array_equal(arr, want), (run, sid)
        done += 1
    assert done >= 5


def test_zero_count_error_of_the_reference_loop_through_the_shim(reference_search_on_shim, tmp_path):
    """search.py:699-722 raises on a k-mer that is not in the index; the shim must hand it the zero to see."""
    fa = tmp_path / "alien.fa"
    rng = np.random.default_rng(3)
    fa.write_bytes(b">a\n" + bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 300)]) + b"\n")
    with pytest.raises(RuntimeError, match="was not found in the index"):
        pyref.run_search(fa, GOLD / "quirk_s4.awfmi", list(range(20, 41)), True, 1000, tmp_path / "o")
