"""Experiment: effect of cudaLimitMaxL2FetchGranularity (NEWMAP_B200_L2_FETCH) on the random-sector probe and on
one searched segment.  Run once per setting:  NEWMAP_B200_L2_FETCH=32 python tools/exp_l2fetch.py [workload]"""
import json
import os
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
res = {"l2_fetch": os.environ.get("NEWMAP_B200_L2_FETCH", "default"), "workload": workload, "probe": {}}
for nbytes in (512 << 20, 2 << 30):
    for g in (32, 128):
        res["probe"][f"{nbytes >> 20}MiB_{g}B"] = _lib.probe_random_access(0, nbytes, g, 8 << 20, 32, 1, True)[0] / 1e9
fa, ixp = bench.prepare_workload(workload)
(segs, total) = bench.load_segments(fa, 1, 1)
seg, n = segs[0]
ix = _lib.Index(ixp, device=0)
ws = _lib.Workspace(ix, len(seg.data))
pb = _lib.PinnedBuffer(len(seg.data))
pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
ws.upload(pb.ptr, len(seg.data))
kl = list(range(20, 201))
best = None
for _ in range(6):
    ws.search(n, kl, True, algo="auto", out_width=1)
    st = ws.sync().as_dict()
    if best is None or st["ms_search"] < best["ms_search"]:
        best = st
res["segment"] = {k: v for k, v in best.items() if k.startswith("ms_") or k.endswith("steps")}
res["checksum"] = ws.checksum()
print(json.dumps(res), flush=True)
