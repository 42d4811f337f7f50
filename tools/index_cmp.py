"""The bench genome indexed by the reference's own builder (oracle/_ref) and by the GPU builder: are the two
.awfmi files the same bytes?  One-off evidence for profiles/ (the reference builder takes minutes per Gbp).
    python tools/index_cmp.py c3_1g [out.json] [--install]
--install: the reference-built file becomes bench.py's cached workload index (so a bench run that follows loads a
reference-built index without building it again)."""
import hashlib
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
d = bench.CACHE / f"{name}_cmp"
d.mkdir(parents=True, exist_ok=True)
fa = d / "genome.fa"
lengths, seed = synth.config_lengths(name)
synth.write_fasta(fa, synth.make_genome(lengths, seed))
res = {"workload": name, "bases": int(sum(lengths)), "sequences": len(lengths)}
for who in ("gpu", "reference"):
    t0 = time.time()
    bench.build_index(fa, d / f"{who}.awfmi", who)
    res[f"{who}_build_s"] = round(time.time() - t0, 1)


def digest(p):
    h = hashlib.sha256()
    with open(p, "rb") as fh:
        while True:
            b = fh.read(64 << 20)
            if not b:
                break
            h.update(b)
    return h.hexdigest()


a, b = d / "gpu.awfmi", d / "reference.awfmi"
res["bytes"] = [a.stat().st_size, b.stat().st_size]
res["sha256"] = [digest(a), digest(b)]
res["identical"] = res["bytes"][0] == res["bytes"][1] and res["sha256"][0] == res["sha256"][1]
if not res["identical"] and res["bytes"][0] == res["bytes"][1]:
    x, y = np.memmap(a, dtype=np.uint8, mode="r"), np.memmap(b, dtype=np.uint8, mode="r")
    first = None
    for off in range(0, x.size, 256 << 20):
        bad = np.nonzero(x[off:off + (256 << 20)] != y[off:off + (256 << 20)])[0]
        if bad.size:
            first = int(off + bad[0])
            break
    res["first_difference"] = first
print(json.dumps(res))
if len(sys.argv) > 2 and not sys.argv[2].startswith("-"):
    Path(sys.argv[2]).write_text(json.dumps(res, indent=1) + "\n")
if "--install" in sys.argv:
    w = bench.CACHE / name
    w.mkdir(parents=True, exist_ok=True)
    fa.rename(w / "genome.fa")
    b.rename(w / "genome.awfmi")
    (w / "builder.txt").write_text("the reference's own builder (oracle/_ref: AwFmCreate + libdivsufsort)")
    (w / ".done").write_text("ok")
