"""Experiment: group size of the coarse certificate round (and of the fine one) vs. time of searched segments.
    python tools/exp_gsize2.py [workload] [first_segment] [segments]"""
import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
first = int(sys.argv[2]) if len(sys.argv) > 2 else 1
count = int(sys.argv[3]) if len(sys.argv) > 3 else 4
fa, ixp = bench.prepare_workload(workload)
(segs, total) = bench.load_segments(fa, first, count)
ix = _lib.Index(ixp, device=0)
cap = max(len(s.data) for s, _ in segs)
ws = _lib.Workspace(ix, cap)
pbs = []
for seg, n in segs:
    pb = _lib.PinnedBuffer(len(seg.data))
    pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
    pbs.append(pb)
kl = list(range(20, 201))
want = None
PAIRS = ((16, 0), (16, 64), (16, 128), (16, 256), (16, 512), (16, 1024), (32, 256), (32, 512), (8, 128), (8, 256))
if len(sys.argv) > 4:      # "16:0,24:0,32:0"
    PAIRS = tuple(tuple(int(x) for x in pr.split(":")) for pr in sys.argv[4].split(","))
for pair in PAIRS:
    gsize, gsize2 = pair[0], pair[1]
    dcap = pair[2] if len(pair) > 2 else 0      # "gsize:gsize2:dcap"
    ws.option("gsize", gsize)
    ws.option("gsize2", gsize2)
    ws.option("dcap", dcap)
    own = pair[3] if len(pair) > 3 else 1        # "gsize:gsize2:dcap:own_chains"
    ws.option("own_chains", own)
    runs = pair[4] if len(pair) > 4 else 0       # "gsize:gsize2:dcap:own_chains:cert_runs"
    ws.option("cert_runs", runs)
    tot = {}
    sums = []
    for (seg, n), pb in zip(segs, pbs):
        ws.upload(pb.ptr, len(seg.data))
        best = None
        for _ in range(4):
            ws.search(n, kl, True, algo="auto", out_width=1)
            st = ws.sync().as_dict()
            if best is None or st["ms_search"] < best["ms_search"]:
                best = st
        sums.append(ws.checksum())
        for k in ("ms_search", "ms_pass1", "ms_cert", "ms_pass3", "lf_steps", "lf_steps_pass1"):
            tot[k] = tot.get(k, 0) + best[k]
    want = want or sums
    row = {"gsize": gsize, "gsize2": gsize2, "dcap": dcap, "own_chains": own, "cert_runs": runs, "same": sums == want, "segments": len(segs)}
    row.update({k: (round(v / len(segs), 4) if isinstance(v, float) else v // len(segs)) for k, v in tot.items()})
    print(json.dumps(row), flush=True)
