"""Every segment of a bench genome searched twice on the GPU -- trajectory-free (the product path) and the
reference-trajectory kernel (the reference's own sequence of count queries) -- and the device checksums compared.
    python tools/cross_check_genome.py [workload] [first_segment] [segments]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
fa, ixp = bench.prepare_workload(workload)
from newmap_b200.fasta import sequence_segments  # noqa: E402
from newmap_b200.search import get_num_kmers  # noqa: E402
with open(fa, "rb") as fh:
    every = [(s, get_num_kmers(s, bench.K_RANGE[1])) for s in sequence_segments(fh, bench.BATCH + bench.K_RANGE[1] - 1, bench.K_RANGE[1] - 1)]
total = len(every)
count = int(sys.argv[3]) if len(sys.argv) > 3 else total
segs = every[first:first + count]
ix = _lib.Index(ixp, device=0)
cap = max(len(s.data) for s, _ in segs)
ws = _lib.Workspace(ix, cap)
pb = _lib.PinnedBuffer(cap)
kl = list(range(*bench.K_RANGE)) + [bench.K_RANGE[1]]
bad, positions, t0 = [], 0, time.perf_counter()
ms = {"auto": 0.0, "trajectory": 0.0}
for j, (seg, n) in enumerate(segs):
    m = len(seg.data)
    pb.array[:m] = np.frombuffer(seg.data, dtype=np.uint8)
    sums = {}
    for algo in ("auto", "trajectory"):
        ws.upload(pb.ptr, m)
        ws.search(n, kl, True, algo=algo, out_width=1)
        st = ws.sync()
        assert st.zero_pos < 0 and not st.replayed
        ms[algo] += st.ms_search + st.ms_resolve + st.ms_prep
        sums[algo] = ws.checksum()
    positions += n
    if sums["auto"] != sums["trajectory"]:
        bad.append(first + j)
print(json.dumps({"workload": workload, "segments": len(segs), "of": total, "positions": positions, "k": list(bench.K_RANGE),
                  "segments_that_differ": bad, "ms_trajectory_free": round(ms["auto"], 1),
                  "ms_reference_trajectory": round(ms["trajectory"], 1), "wall_s": round(time.perf_counter() - t0, 1)}))
