"""Experiment: derived seed-table depth (and first-pass depth cap) vs. time of one searched segment.
    python tools/exp_seed_depth.py [workload] [segment_no]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
segno = int(sys.argv[2]) if len(sys.argv) > 2 else 1
fa, ixp = bench.prepare_workload(workload)
(segs, total) = bench.load_segments(fa, segno, 1)
seg, n = segs[0]
t0 = time.time()
ix = _lib.Index(ixp, device=0)
print(json.dumps({"open_s": time.time() - t0, "auto_depth": ix.derived_seed_length,
                  "device_gb": ix.info.device_bytes / 1e9}), flush=True)
ws = _lib.Workspace(ix, len(seg.data))
pb = _lib.PinnedBuffer(len(seg.data))
pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
ws.upload(pb.ptr, len(seg.data))
kl = list(range(20, 201))
want = None
for depth, dcap, pass1 in [(0, 0, 1), (0, 0, 2), (13, 0, 2), (14, 0, 2), (15, 0, 2), (16, 28, 1), (16, 28, 2), (16, 18, 2),
                           (16, 20, 2), (16, 24, 2), (16, 32, 2), (14, 20, 2), (15, 20, 2), (16, 0, 0)]:
    t0 = time.time()
    ix.set_seed_depth(depth)
    tb = time.time() - t0
    ws.option("dcap", dcap)
    ws.option("pass1", pass1)
    best = None
    for _ in range(5):
        ws.search(n, kl, True, algo="auto", out_width=1)
        st = ws.sync().as_dict()
        if best is None or st["ms_search"] < best["ms_search"]:
            best = st
    cs = ws.checksum()
    want = want or cs
    row = {"depth": depth, "dcap": dcap, "pass1": pass1, "build_s": round(tb, 3), "same": cs == want}
    row.update({k: (round(v, 4) if isinstance(v, float) else v) for k, v in best.items()
                if k.startswith("ms_p") or k in ("ms_search", "ms_cert", "ms_resolve", "lf_steps", "lf_steps_pass1",
                                                "seed_lookups", "seed_depth")})
    print(json.dumps(row), flush=True)
