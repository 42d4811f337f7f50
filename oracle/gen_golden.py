"""oracle/gen_golden.py -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/ by running the UNMODIFIED reference in the build container:
indexes through the reference's own index builder (oracle/_ref, compiled from
/root/reference by oracle/Makefile) and expected outputs through the reference's own Python
driver (newmap/search.py imported from /root/reference by oracle/pyref.py) and its own count
extension.  Run where /root/reference exists:

    python -m oracle.gen_golden

Everything it writes is small (a few hundred kB in total) and committed, because
/root/reference does not travel to the GPU box.
"""
from __future__ import annotations

import json
import tempfile
from pathlib import Path

import numpy as np

from . import ora, pyref

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"

# the reference's own 50 bp fixture (tests/data/genome.fa), restated as data
GENOME_FA = b">chr1\nAAAAATTTTT\nATCGAATCGA\n>chr2\nAAAAAAAAAA\nGGGGGGGGGG\nCCCCCCCCCC\n"


def _rc(a: np.ndarray) -> np.ndarray:
    t = np.arange(256, dtype=np.uint8)
    for x, y in zip(b"ACGTacgt", b"TGCAtgca"):
        t[x] = y
    return t[a[::-1]]


def quirk_genome(seed: int = 7, iupac: bool = False) -> list[tuple[str, np.ndarray]]:
    """Three short sequences packed with the cases SURVEY.md §8(c) lists: N-runs at the start,
    centre and end, N-runs that begin just after a multiple of the batch sizes used below
    (look-ahead quirk), soft-masked runs, tandem repeats, forward and reverse-complement
    duplications (with and without substitutions), reverse-palindromes."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)

    def rnd(n):
        return acgt[rng.integers(0, 4, size=n)]

    a = rnd(6200)
    a[:37] = ord("N")                      # leading run
    a[1003:1009] = ord("N")                # just after batch 1000 (look-ahead of segment 0)
    a[2000:2001] = ord("N")                # exactly at a batch boundary
    a[2995:3010] = ord("N")                # straddles a batch boundary
    a[4030:4033] = ord("N")                # inside look-ahead of segment 3 for k<=40
    a[1400:1600] |= 0x20                   # soft-masked
    a[1700:1900] = np.tile(rnd(7), 29)[:200]          # tandem, unit 7
    a[2100:2160] = np.tile(rnd(2), 30)                # dinucleotide repeat
    a[2300:2700] = a[300:700]                         # forward duplication, exact
    a[3100:3400] = _rc(a[700:1000])                   # reverse-complement duplication, exact
    d = a[300:600].copy(); d[::37] = rnd(d[::37].size)
    a[4300:4600] = d                                   # diverged duplication
    half = rnd(40); a[5000:5080] = np.concatenate([half, _rc(half)])   # reverse-palindrome
    a[6150:] = ord("N")                    # trailing run
    b = rnd(3500)
    b[500:900] = a[320:720]                           # inter-sequence duplication
    b[1500:1800] = _rc(b[1000:1300])
    b[1800:2100] |= 0x20
    b[2511:2513] = ord("N")
    b[3000:3200] = np.tile(rnd(50), 4)
    c = rnd(800)
    c[100:130] = ord("N")
    c[780:] = a[100:120]
    if iupac:
        # (no U/u: the reference's index builder sorts 'u' as its own symbol but codes it as
        #  't' in the BWT, which corrupts the index -- observed here, out of scope)
        a[1250] = ord("n"); a[1260] = ord("R"); a[4100] = ord("y"); b[700] = ord("M")
        b[1007] = ord("n"); c[400] = ord("s"); c[410] = ord("K")
    return [("chr1", a), ("chr2 some description", b), ("chrM", c)]


def write_fasta(path, records, width=60):
    with open(path, "wb") as fh:
        for name, arr in records:
            fh.write(b">" + name.encode() + b"\n")
            raw = arr.tobytes()
            for i in range(0, len(raw), width):
                fh.write(raw[i:i + width] + b"\n")


def main():
    assert pyref.available(), "needs /root/reference and `make -C oracle ref`"
    GOLD.mkdir(parents=True, exist_ok=True)
    count_mod, _ = ora.ref_modules()
    manifest = {"indexes": {}, "searches": [], "counts": []}

    (GOLD / "genome.fa").write_bytes(GENOME_FA)
    write_fasta(GOLD / "quirk.fa", quirk_genome())
    write_fasta(GOLD / "quirk_iupac.fa", quirk_genome(iupac=True))
    for name, fa, ratio, seed in [("genome_s1", "genome.fa", 1, 1), ("genome_s3", "genome.fa", 8, 3),
                                  ("quirk_s4", "quirk.fa", 8, 4), ("quirk_s6", "quirk.fa", 8, 6),
                                  ("quirk_iupac_s5", "quirk_iupac.fa", 8, 5)]:
        ora.ref_build_index(GOLD / fa, GOLD / f"{name}.awfmi", ratio, seed)
        manifest["indexes"][name] = {"fasta": fa, "compression_ratio": ratio, "seed_length": seed}

    # ---- count goldens: the reference's own vectors + random (index, length) queries
    rng = np.random.default_rng(11)
    cases = {}
    for ixname, fa in [("genome_s1", "genome.fa"), ("genome_s3", "genome.fa"), ("quirk_s4", "quirk.fa"),
                       ("quirk_iupac_s5", "quirk_iupac.fa")]:
        for sid, seq in ora.fasta_records(GOLD / fa):
            n = len(seq)
            m = 400 if n > 100 else 120
            idx = rng.integers(0, n, size=m)
            ln = np.minimum(rng.integers(1, 64 if n > 100 else 12, size=m), n - idx)
            out = count_mod.count_kmers_from_sequence(str(GOLD / f"{ixname}.awfmi"), seq, idx.tolist(),
                                                      ln.tolist(), 1)
            key = f"{ixname}|{sid.decode()}"
            cases[key + "|idx"] = idx.astype(np.uint64)
            cases[key + "|len"] = ln.astype(np.uint64)
            cases[key + "|count"] = np.array(out, dtype=np.uint32)
            manifest["counts"].append({"index": ixname, "fasta": fa, "sequence": sid.decode(), "key": key})
    np.savez_compressed(GOLD / "counts.npz", **cases)

    # ---- search goldens through the reference's write_unique_counts
    runs = []
    for ix, fa in [("genome_s1", "genome.fa"), ("genome_s3", "genome.fa")]:
        for mode in ("range", "list"):
            runs.append(dict(index=ix, fasta=fa, k=list(range(4, 11)), range=(mode == "range"), batch=15))
        runs.append(dict(index=ix, fasta=fa, k=list(range(20, 201)), range=True, batch=10_000_000))
        runs.append(dict(index=ix, fasta=fa, k=list(range(2, 8)), range=True, batch=7, norc=True))
        runs.append(dict(index=ix, fasta=fa, k=[3, 5, 8], range=False, batch=4))
    for ix, fa in [("quirk_s4", "quirk.fa"), ("quirk_s6", "quirk.fa"), ("quirk_iupac_s5", "quirk_iupac.fa")]:
        for batch in (1000, 257, 1_000_000):
            runs.append(dict(index=ix, fasta=fa, k=list(range(8, 41)), range=True, batch=batch))
            runs.append(dict(index=ix, fasta=fa, k=[8, 12, 20, 33], range=False, batch=batch))
        runs.append(dict(index=ix, fasta=fa, k=list(range(20, 201)), range=True, batch=1000))
        runs.append(dict(index=ix, fasta=fa, k=list(range(12, 256)), range=True, batch=2000))
        runs.append(dict(index=ix, fasta=fa, k=list(range(3, 30)), range=True, batch=500, initial=12))
        runs.append(dict(index=ix, fasta=fa, k=list(range(10, 61)), range=True, batch=1000, norc=True))
        runs.append(dict(index=ix, fasta=fa, k=[24, 36, 50, 100], range=False, batch=1500))
        runs.append(dict(index=ix, fasta=fa, k=[16, 30], range=False, batch=3000, norc=True))
        runs.append(dict(index=ix, fasta=fa, k=list(range(250, 301)), range=True, batch=4000))   # uint16 output
    arrays = {}
    for n, r in enumerate(runs):
        with tempfile.TemporaryDirectory() as td:
            res = pyref.run_search(GOLD / r["fasta"], GOLD / f"{r['index']}.awfmi", r["k"], r["range"],
                                   r["batch"], td, initial_search_length=r.get("initial", 0),
                                   use_rc=not r.get("norc", False), num_threads=2)
        for sid, arr in res.items():
            arrays[f"run{n}|{sid.decode()}"] = arr
        r = dict(r, id=n, sequences=[s.decode() for s in res])
        if r["range"]:
            r["k"] = [min(r["k"]), max(r["k"])]     # stored as [min, max] for range runs
        manifest["searches"].append(r)
    np.savez_compressed(GOLD / "searches.npz", **arrays)
    (GOLD / "manifest.json").write_text(json.dumps(manifest, indent=1))
    print(f"wrote {len(runs)} search goldens, {len(manifest['counts'])} count goldens to {GOLD}")


if __name__ == "__main__":
    main()
