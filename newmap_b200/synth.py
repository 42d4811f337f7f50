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
                  repeat_scale: float = 1.0, n_runs: bool = True, background: np.ndarray | None = None) -> np.ndarray:
    """One chromosome-like byte array of `length` bases (on `background` when given, else on uniform random ACGT)."""
    seq = background if background is not None else _ACGT[rng.integers(0, 4, size=length, dtype=np.uint8)]
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


# interspersed repeats of a mammalian genome, coarsely: SINE-like (300 bp, 11 %), LINE-like (6 kb, truncated copies,
# 18 %), LTR-like (7 %), DNA-transposon-like (3 %), each old enough to have diverged 8-20 % from its consensus, and a
# young subfamily of each of the first two (2-3 % divergence): 43 % of the genome in all
MAMMALIAN_FAMILIES = [(300, 0.09, 0.12, False), (300, 0.02, 0.025, False), (6000, 0.15, 0.15, True),
                      (6000, 0.03, 0.03, True), (450, 0.07, 0.18, False), (1200, 0.03, 0.20, True)]


def make_genome(lengths, seed: int, families=None, **kw) -> list[tuple[str, np.ndarray]]:
    rng = np.random.default_rng(seed)
    out = []
    for i, n in enumerate(lengths):
        if families:
            # interspersed repeats first (one consensus set per genome: the same families on every chromosome), then
            # the tandem repeats, duplications, soft-masking and N-runs of make_sequence on top
            base = _ACGT[rng.integers(0, 4, size=int(n), dtype=np.uint8)]
            fam_rng = np.random.default_rng(seed + 7919)                      # same consensus sequences on every chromosome
            consensus_state = fam_rng.bit_generator.state
            add_interspersed_repeats_with_state(base, rng, families, consensus_state)
            out.append((f"chr{i + 1}", make_sequence(int(n), rng, background=base, **kw)))
        else:
            out.append((f"chr{i + 1}", make_sequence(int(n), rng, **kw)))
    return out


def add_interspersed_repeats_with_state(seq, rng, families, consensus_state) -> None:
    """Overwrites parts of `seq` with copies of repeat families, in place: the repeat spectrum of a mammalian genome
    (about half of it interspersed repeats, most of them old and diverged) rather than the few young exact repeats
    make_sequence plants.  families = [(consensus length, fraction of the sequence to cover, divergence, truncate)]:
    every copy is the family consensus with `divergence` of its bases substituted independently per copy, on either
    strand; `truncate`: copies keep a random-length 3' part only (like LINE elements).  The consensus sequences come
    from a generator in `consensus_state`, so that every chromosome of a genome carries the same families."""
    crng = np.random.default_rng()
    crng.bit_generator.state = consensus_state
    n = seq.size
    for length, fraction, divergence, truncate in families:
        consensus = _ACGT[crng.integers(0, 4, size=length, dtype=np.uint8)]
        copies = int(n * fraction / (length * (0.55 if truncate else 1.0)))
        done = 0
        while done < copies:
            m = min(copies - done, max(1, (8 << 20) // length))
            keep = rng.integers(length // 10, length + 1, size=m) if truncate else np.full(m, length)
            start = rng.integers(0, max(1, n - length), size=m)
            col = np.arange(length)[None, :]
            live = col >= (length - keep)[:, None]
            piece = np.broadcast_to(consensus, (m, length)).copy()
            hit = rng.random((m, length)) < divergence
            piece[hit] = _ACGT[rng.integers(0, 4, size=int(hit.sum()), dtype=np.uint8)]
            flip = rng.random(m) < 0.5
            piece[flip] = _COMP[piece[flip][:, ::-1]]
            live[flip] = live[flip][:, ::-1]
            pos = start[:, None] + col
            seq[pos[live]] = piece[live]
            done += m


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
    if name == "c3r_1g":     # c3_1g's shape with the interspersed-repeat spectrum of a mammalian genome (MAMMALIAN_FAMILIES)
        w = np.array([5, 4, 3.5, 3, 2.5, 2, 1.5, 1.0])
        return [int(x) for x in (w / w.sum() * 1_000_000_000)], 1001
    raise KeyError(name)
