"""`newmap track` on the GPU against the numpy restatement of the reference (oracle/track_oracle.py, test
infrastructure; the timed CPU leg only) on one synthetic chromosome.  python tools/track_bench.py [positions] [k]"""
import json
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from newmap_b200 import track  # noqa: E402
from oracle import track_oracle as T  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 24
rng = np.random.default_rng(1)
a = rng.integers(20, 60, n).astype(np.uint8)
a[rng.random(n) < 0.3] = 0
for s in rng.integers(0, n - 200_000, 200):
    a[s:s + int(rng.integers(100, 200_000))] = 0
with tempfile.TemporaryDirectory() as td:
    td = Path(td)
    p = td / "chrS.unique.uint8"
    a.tofile(p)
    res = {"positions": n, "k": k}
    for label, single, multi in (("bed", str(td / "o.bed"), ""), ("bed+wig", str(td / "o2.bed"), str(td / "o.wig"))):
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            track.write_mappability_files([p], k, single, multi, False)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        res[f"gpu_{label}_s"] = round(best, 3)
        res[f"gpu_{label}_positions_per_s"] = n / best
    res["wig_bytes"] = (td / "o.wig").stat().st_size
    m = min(n, 5_000_000)          # CPU leg on a bounded sample
    t0 = time.perf_counter()
    bed = T.bed_text(a[:m], k, "chrS")
    t1 = time.perf_counter()
    wig = T.wig_text(a[:m], k, "chrS")
    t2 = time.perf_counter()
    res.update({"cpu_sample_positions": m, "cpu_bed_positions_per_s": m / (t1 - t0), "cpu_wig_positions_per_s": m / (t2 - t1)})
    track.write_mappability_files([p], k, str(td / "c.bed"), "", False)
    res["bed_prefix_equal"] = (td / "c.bed").read_bytes()[:len(bed) - 200] == bed[:len(bed) - 200]
    res["wig_prefix_equal"] = (td / "o.wig").read_bytes()[:len(wig) - 200] == wig[:len(wig) - 200]
print(json.dumps(res))
