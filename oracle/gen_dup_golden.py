"""oracle/gen_dup_golden.py -- TEST INFRASTRUCTURE ONLY.

A FASTA whose records repeat an id: twice in a row (the reference APPENDS the second record to the same file,
newmap/search.py:268-357: a file is opened anew only when the id changes) and once more after another id (the
reference truncates the file and starts again).  Writes tests/golden/dup_{append,return}.{fa,awfmi} and dup_ids.npz: the FASTAs, their indexes
by the reference's builder and the files the UNMODIFIED reference wrote for them.  Run where /root/reference exists:
    python -m oracle.gen_dup_golden"""
from __future__ import annotations

import tempfile
from pathlib import Path

import numpy as np

from . import ora, pyref

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
RUNS = [dict(k=[20, 60], range=True, batch=1000), dict(k=[22, 30, 45], range=False, batch=700)]


def main():
    rng = np.random.default_rng(11)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    pool = acgt[rng.integers(0, 4, 400)]
    out = {}
    for tag, recs in (("append", [("a first", 2500), ("a second", 1800), ("b", 900)]),
                      ("return", [("a first", 1500), ("b", 900), ("a again", 700)])):
        parts = []
        for name, n in recs:
            s = acgt[rng.integers(0, 4, n)].copy()
            s[100:400] = pool[:300]                 # shared repeat so that long k-mers stay non-unique
            s[n // 2:n // 2 + 5] = ord("N")
            parts.append(b">" + name.encode() + b"\n" + s.tobytes() + b"\n")
        fa = GOLD / f"dup_{tag}.fa"
        fa.write_bytes(b"".join(parts))
        ix = GOLD / f"dup_{tag}.awfmi"
        ora.ref_build_index(fa, ix, 8, 5)
        for n, run in enumerate(RUNS):
            kl = list(range(run["k"][0], run["k"][1] + 1)) if run["range"] else run["k"]
            with tempfile.TemporaryDirectory() as tmp:
                got = pyref.run_search(fa, ix, kl, run["range"], run["batch"], tmp)
            for sid, arr in got.items():
                out[f"{tag}|run{n}|{sid.decode()}"] = arr
    np.savez_compressed(GOLD / "dup_ids.npz", **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
