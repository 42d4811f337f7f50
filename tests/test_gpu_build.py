"""GPU index construction (nm_index_build) writes the reference builder's file byte for byte."""
import numpy as np
import pytest

from conftest import GOLD
from oracle import ora

pytestmark = pytest.mark.gpu


def _same(a, b):
    x, y = a.read_bytes(), b.read_bytes()
    if x == y:
        return True, ""
    n = min(len(x), len(y))
    d = np.nonzero(np.frombuffer(x[:n], np.uint8) != np.frombuffer(y[:n], np.uint8))[0]
    return False, f"sizes {len(x)} vs {len(y)}, first differing byte {d[0] if d.size else n}, {d.size} differ"


@pytest.mark.parametrize("name,fasta", [("genome_s1", "genome.fa"), ("genome_s3", "genome.fa"), ("quirk_s4", "quirk.fa"),
                                        ("quirk_s6", "quirk.fa"), ("quirk_iupac_s5", "quirk_iupac.fa")])
def test_committed_reference_indexes_are_reproduced(manifest, tmp_path, name, fasta):
    from newmap_b200 import index
    meta = manifest["indexes"][name]
    out = tmp_path / f"{name}.awfmi"
    index.generate_fm_index(GOLD / fasta, out, meta["compression_ratio"], meta["seed_length"], device=0)
    ok, why = _same(out, GOLD / f"{name}.awfmi")
    assert ok, why


@pytest.mark.skipif(not ora.ref_available(), reason="needs oracle/_ref (the reference's index builder)")
@pytest.mark.parametrize("lengths,seed_len,ratio", [([300_000, 40_000, 1_000], 8, 8), ([2_000_000], 10, 1),
                                                    ([1_500_000, 500_000], 11, 32)])
def test_synthetic_genomes_match_reference_builder(tmp_path, lengths, seed_len, ratio):
    from newmap_b200 import index, synth
    fa = tmp_path / "g.fa"
    synth.write_fasta(fa, synth.make_genome(lengths, seed=len(lengths) * 17 + seed_len, batch=100_000))
    ref, got = tmp_path / "ref.awfmi", tmp_path / "gpu.awfmi"
    ora.ref_build_index(fa, ref, ratio, seed_len)
    index.generate_fm_index(fa, got, ratio, seed_len, device=0)
    ok, why = _same(got, ref)
    assert ok, why


def test_gpu_built_index_searches_like_the_reference_built_one(tmp_path):
    from newmap_b200 import _lib, index
    out = tmp_path / "q.awfmi"
    index.generate_fm_index(GOLD / "quirk.fa", out, 8, 6, device=0)
    seq = dict(ora.fasta_records(GOLD / "quirk.fa"))[b"chr1"]
    kl = list(range(20, 201))
    with _lib.Index(out, device=0) as a, _lib.Index(GOLD / "quirk_s6.awfmi", device=0) as b:
        x, _ = a.min_unique(seq, len(seq), kl, True)
        y, _ = b.min_unique(seq, len(seq), kl, True)
    assert np.array_equal(x, y)
