"""oracle/gen_multi_golden.py -- TEST INFRASTRUCTURE ONLY.

Goldens for the several-FASTA / several-index mode of `newmap search`, written by the UNMODIFIED
reference in the build container (newmap/search.py imported from /root/reference by oracle/pyref.py, its own index
builder and count extension from oracle/_ref):

    python -m oracle.gen_multi_golden        ->  tests/golden/multi/

Inputs: a small genome with N runs, soft-masked runs and repeats; its C->T and G->A conversions (the bisulfite use
the mode was written for: two FASTA files of equal lengths with their two indexes) and the genome split over two
index files (one FASTA, two indexes).
"""
from __future__ import annotations

import json
import tempfile
from pathlib import Path

import numpy as np

from . import ora, pyref
from .gen_golden import _rc, write_fasta

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden" / "multi"


def base_genome(seed: int = 21):
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)

    def rnd(n):
        return acgt[rng.integers(0, 4, size=n)]

    a = rnd(2600)
    a[:11] = ord("N")
    a[700:705] = ord("N")
    a[1001:1003] = ord("N")          # just after a batch boundary of 1000 (look-ahead quirk)
    a[300:420] |= 0x20
    a[1400:1650] = a[100:350]        # exact duplication
    a[1800:2000] = _rc(a[450:650])   # reverse-complement duplication
    d = a[2050:2250].copy()
    d[::23] = rnd(d[::23].size)
    a[2300:2500] = d                 # diverged duplication
    a[2590:] = ord("N")
    b = rnd(1500)
    b[200:420] = a[120:340]          # duplication between sequences
    b[900:903] = ord("N")
    b[1100:1300] = np.tile(rnd(5), 40)
    return [("chrA", a), ("chrB second", b)]


def convert(records, src: bytes, dst: bytes):
    t = np.arange(256, dtype=np.uint8)
    for x, y in zip(src, dst):
        t[x] = y
    return [(name, t[seq]) for name, seq in records]


def run_reference(fastas, indexes, k, is_range, batch, outdir, use_rc=True, initial=0):
    search, _ = pyref.load()
    outdir = Path(outdir)
    cfg = search.SearchConfig(
        fasta_filepaths=[Path(f) for f in fastas], fmindex_filepaths=[Path(i) for i in indexes],
        kmer_lengths=list(k), is_binary_search=is_range, use_reverse_complement=use_rc, output_directory=outdir,
        include_sequence_ids=[], exclude_sequence_ids=[], num_threads=2, kmer_batch_size=batch,
        initial_search_length=initial)
    search.write_unique_counts(cfg)
    suffix = "uint8" if max(k) <= 255 else "uint16"
    return {f.name[: -len(f".unique.{suffix}")]: np.fromfile(f, dtype=getattr(np, suffix))
            for f in sorted(outdir.glob(f"*.unique.{suffix}"))}


def main():
    assert pyref.available(), "needs /root/reference and `make -C oracle ref`"
    GOLD.mkdir(parents=True, exist_ok=True)
    g = base_genome()
    write_fasta(GOLD / "base.fa", g)
    write_fasta(GOLD / "ct.fa", convert(g, b"Cc", b"Tt"))
    write_fasta(GOLD / "ga.fa", convert(g, b"Gg", b"Aa"))
    write_fasta(GOLD / "partA.fa", g[:1])
    write_fasta(GOLD / "partB.fa", g[1:])
    for name in ("ct", "ga", "partA", "partB"):
        ora.ref_build_index(GOLD / f"{name}.fa", GOLD / f"{name}.awfmi", 8, 4)
    runs = []
    for batch in (1000, 1_000_000):
        runs.append(dict(fastas=["ct.fa", "ga.fa"], indexes=["ct.awfmi", "ga.awfmi"], k=list(range(8, 61)), range=True, batch=batch))
        runs.append(dict(fastas=["ct.fa", "ga.fa"], indexes=["ct.awfmi", "ga.awfmi"], k=[10, 16, 24, 40], range=False, batch=batch))
        runs.append(dict(fastas=["base.fa"], indexes=["partA.awfmi", "partB.awfmi"], k=list(range(6, 41)), range=True, batch=batch))
    runs.append(dict(fastas=["ct.fa", "ga.fa"], indexes=["ct.awfmi", "ga.awfmi"], k=list(range(8, 61)), range=True, batch=700, norc=True))
    runs.append(dict(fastas=["ct.fa", "ga.fa"], indexes=["ct.awfmi", "ga.awfmi"], k=list(range(5, 45)), range=True, batch=900, initial=12))
    runs.append(dict(fastas=["ct.fa", "ga.fa"], indexes=["ct.awfmi", "ga.awfmi"], k=[12, 30], range=False, batch=650, norc=True))
    runs.append(dict(fastas=["base.fa"], indexes=["partA.awfmi", "partB.awfmi"], k=[8, 14, 22], range=False, batch=800))
    runs.append(dict(fastas=["base.fa"], indexes=["partB.awfmi", "partA.awfmi"], k=list(range(250, 281)), range=True, batch=1200))
    arrays, manifest = {}, []
    for n, r in enumerate(runs):
        with tempfile.TemporaryDirectory() as td:
            res = run_reference([GOLD / f for f in r["fastas"]], [GOLD / i for i in r["indexes"]], r["k"], r["range"],
                                r["batch"], td, use_rc=not r.get("norc", False), initial=r.get("initial", 0))
        for sid, arr in res.items():
            arrays[f"run{n}|{sid}"] = arr
        r = dict(r, id=n, sequences=list(res))
        if r["range"]:
            r["k"] = [min(r["k"]), max(r["k"])]
        manifest.append(r)
        print(n, {k: (v.size, int(np.count_nonzero(v))) for k, v in res.items()})
    np.savez_compressed(GOLD / "searches.npz", **arrays)
    (GOLD / "manifest.json").write_text(json.dumps(manifest, indent=1))


if __name__ == "__main__":
    main()
