"""Deterministic synthetic genomes for parity tests and bench.py.

Not part of the search path: it only manufactures FASTA inputs of the shapes BASELINE.json
names (uniform ACGT background, soft-masked runs, tandem repeats, segmental duplications on
both strands with a few substitutions, telomere/centromere/short N-runs placed so that some
fall inside the look-ahead of a batch boundary).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.arange(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTacgt", b"TGCAtgca"):
    _COMP[_a] = _b


def revcomp(a: np.ndarray) -> np.ndarray:
    return _COMP[a[::-1]]


def make_sequence(length: int, rng: np.random.Generator, *, batch: int = 10_000_000,
                  repeat_scale: float = 1.0, n_runs: bool = True) -> np.ndarray:
    """One chromosome-like byte array of `length` bases."""
    seq = _ACGT[rng.integers(0, 4, size=length, dtype=np.uint8)]
    if length < 64:
        return seq
    # tandem repeats: one every ~1 Mbp (at least one), unit in {2,7,50,171}, 5-50 kb long
    n_tr = max(1, int(length / 1_000_000 * repeat_scale))
    for _ in range(n_tr):
        unit = int(rng.choice([2, 7, 50, 171]))
        span = int(min(rng.integers(5_000, 50_000), max(2 * unit, length // 20)))
        at = int(rng.integers(0, max(1, length - span)))
        u = _ACGT[rng.integers(0, 4, size=unit, dtype=np.uint8)]
        seq[at:at + span] = np.tile(u, span // unit + 1)[:span]
    # segmental duplications: one every ~2 Mbp, 10-200 kb, half reverse-complemented,
    # 0 / 0.1 / 1 % substitutions
    n_sd = max(1, int(length / 2_000_000 * repeat_scale))
    for j in range(n_sd):
        span = int(min(rng.integers(10_000, 200_000), max(32, length // 25)))
        src = int(rng.integers(0, max(1, length - span)))
        dst = int(rng.integers(0, max(1, length - span)))
        piece = seq[src:src + span].copy()
        if j % 2:
            piece = revcomp(piece)
        rate = (0.0, 0.001, 0.01)[j % 3]
        if rate:
            hit = np.nonzero(rng.random(span) < rate)[0]
            piece[hit] = _ACGT[rng.integers(0, 4, size=hit.size, dtype=np.uint8)]
        seq[dst:dst + span] = piece
    # soft-masked (lower-case) runs: ~5 % of bases in 1-10 kb runs
    covered, target = 0, int(0.05 * length)
    while covered < target:
        span = int(min(rng.integers(1_000, 10_000), max(8, length // 50)))
        at = int(rng.integers(0, max(1, length - span)))
        seq[at:at + span] |= 0x20
        covered += span
    if n_runs:
        tel = int(min(10_000, length // 100))
        if tel:
            seq[:tel] = ord("N")
            seq[length - tel:] = ord("N")
        cen = int(length * rng.uniform(0.01, 0.03))
        at = int(rng.integers(length // 3, max(length // 3 + 1, 2 * length // 3 - cen)))
        seq[at:at + cen] = ord("N")
        # short runs; several right after a multiple of the batch size (look-ahead quirk, §8 A7)
        for j in range(20):
            span = int(rng.integers(1, 100))
            if j < 8 and length > batch:
                base = int(rng.integers(1, length // batch + 1)) * batch
                at = base + int(rng.integers(0, 199))
            else:
                at = int(rng.integers(0, length))
            at = min(max(at, 0), max(0, length - span))
            seq[at:at + span] = ord("N")
    return seq


def make_genome(lengths, seed: int, **kw) -> list[tuple[str, np.ndarray]]:
    rng = np.random.default_rng(seed)
    return [(f"chr{i + 1}", make_sequence(int(n), rng, **kw)) for i, n in enumerate(lengths)]


def write_fasta(path, records, width: int = 60) -> None:
    path = Path(path)
    with open(path, "wb") as fh:
        for name, arr in records:
            fh.write(b">" + name.encode() + b"\n")
            n = arr.size
            full = n // width
            if full:
                body = np.empty((full, width + 1), dtype=np.uint8)
                body[:, :width] = arr[:full * width].reshape(full, width)
                body[:, width] = 10
                fh.write(body.tobytes())
            if n % width:
                fh.write(arr[full * width:].tobytes() + b"\n")


# GRCh38-like proportions for a 24-sequence genome (chr1..22, X, Y), fractions of the total.
GRCH38_FRACTIONS = np.array([248, 242, 198, 190, 181, 171, 159, 145, 138, 133, 135, 133, 114, 107,
                             102, 90, 83, 80, 58, 64, 46, 50, 156, 57], dtype=np.float64)
GRCH38_FRACTIONS /= GRCH38_FRACTIONS.sum()


def config_lengths(name: str) -> tuple[list[int], int]:
    """(sequence lengths, seed) for the BASELINE.json synthetic configs."""
    if name == "c2_100m":
        return [100_000_000], 100
    if name == "c3_1g":
        w = np.array([5, 4, 3.5, 3, 2.5, 2, 1.5, 1.0])
        return [int(x) for x in (w / w.sum() * 1_000_000_000)], 1000
    if name == "c4_3g":
        return [int(x) for x in GRCH38_FRACTIONS * 3_100_000_000], 3100
    raise KeyError(name)
