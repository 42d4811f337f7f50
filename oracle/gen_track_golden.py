"""Writes tests/golden/track/: <id>.unique.uintN inputs and the BED / WIG files the UNMODIFIED reference
(/root/reference/newmap/track.py, imported here) produces from them.  Run where /root/reference exists:
    python oracle/gen_track_golden.py"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, "/root/reference")
from newmap import track as ref_track  # noqa: E402

out = ROOT / "tests" / "golden" / "track"
out.mkdir(parents=True, exist_ok=True)
rng = np.random.default_rng(2024)
searches = np.load(ROOT / "tests" / "golden" / "searches.npz")
cases = []


def unique_array(n, dtype, hi):
    a = rng.integers(0, hi, n).astype(dtype)
    a[rng.random(n) < 0.35] = 0                       # positions with no unique length
    s = int(rng.integers(0, max(1, n - 40)))
    a[s:s + 40] = 0                                   # an unmappable stretch longer than most read lengths
    return a


inputs = {
    "chrA.unique.uint8": unique_array(700, np.uint8, 120),
    "chr.B.v2.unique.uint8": unique_array(333, np.uint8, 255),       # a name with periods (track.py:158-160)
    "chrC.unique.uint16": unique_array(500, np.uint16, 900),
    "chrD.unique.uint32": unique_array(260, np.uint32, 70000),
    "chrQ.unique.uint8": next(searches[k] for k in searches.files if searches[k].dtype == np.uint8 and searches[k].size > 900)[:900],
}
for name, arr in inputs.items():
    arr.tofile(out / name)
for k, files in ((24, ["chrA.unique.uint8", "chr.B.v2.unique.uint8"]), (1, ["chrA.unique.uint8"]),
                 (10, ["chrC.unique.uint16"]), (100, ["chrA.unique.uint8", "chrQ.unique.uint8"]),
                 (150, ["chrC.unique.uint16", "chrD.unique.uint32"]), (1000, ["chrD.unique.uint32"]),
                 (36, ["chrQ.unique.uint8"])):
    tag = f"k{k}_" + "_".join(f.split(".unique")[0] for f in files)
    bed, wig = out / f"{tag}.bed", out / f"{tag}.wig"
    ref_track.write_mappability_files([out / f for f in files], k, str(bed), str(wig), False)
    cases.append({"k": k, "files": files, "bed": bed.name, "wig": wig.name})
(out / "manifest.json").write_text(json.dumps(cases, indent=1))
print(f"{len(cases)} cases, {sum(p.stat().st_size for p in out.iterdir())} bytes in {out}")
