"""Experiment: per-kernel event times of searches queued back to back on ONE stream against the same searches issued
one at a time.  `python tools/exp_inregion.py [workload] [workspaces] [shared]`: with `shared` every workspace is put on the
same torch stream; without it each has the stream the library creates for it (their kernels overlap and the event
times are then not kernel durations)."""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
nws = int(sys.argv[2]) if len(sys.argv) > 2 else 10
fa, ixp = bench.prepare_workload(workload)
(segs, total) = bench.load_segments(fa, 1, nws)
ix = _lib.Index(ixp, device=0)
shared = None
if len(sys.argv) > 3 and sys.argv[3] == "shared":
    import torch
    shared = torch.cuda.Stream()
wss, keep = [], []
for seg, n in segs:
    ws = _lib.Workspace(ix, len(seg.data), cuda_stream=shared.cuda_stream if shared else None)
    pb = _lib.PinnedBuffer(len(seg.data))
    pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
    ws.upload(pb.ptr, len(seg.data))
    ws.sync()
    wss.append((ws, n))
    keep.append(pb)
kl = list(range(20, 201))
KEYS = ("ms_prep", "ms_pass1", "ms_cert", "ms_pass3", "ms_resolve", "ms_search")


def run(depth):
    rows, t0 = [], time.perf_counter()
    for rep in range(4):
        for i in range(0, len(wss), depth):
            for ws, n in wss[i:i + depth]:
                ws.search(n, kl, True, algo="auto", out_width=1)
            for ws, n in wss[i:i + depth]:
                st = ws.sync().as_dict()
                if rep:
                    rows.append(st)
        if rep == 0:
            t0 = time.perf_counter()
    wall = (time.perf_counter() - t0) * 1e3 / (3 * len(wss))
    out = {"stream": "shared" if shared else "own", "queued": depth, "wall_ms_per_search": round(wall, 4)}
    out.update({k: round(sum(r[k] for r in rows) / len(rows), 4) for k in KEYS})
    print(json.dumps(out), flush=True)


for d in (1, 2, 5, nws):
    run(d)
