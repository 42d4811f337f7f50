"""Experiment: certificate group size / depth cap vs. time of one searched segment.
    python tools/exp_gsize.py [workload] [segment_no]"""
import json
import sys
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
ix = _lib.Index(ixp, device=0)
ws = _lib.Workspace(ix, len(seg.data))
pb = _lib.PinnedBuffer(len(seg.data))
pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
ws.upload(pb.ptr, len(seg.data))
kl = list(range(20, 201))
want = None
for gsize in (8, 16, 32):
    for dcap in (20, 24, 28):
        ws.option("dcap", dcap)
        ws.option("gsize", gsize)
        best = None
        for _ in range(5):
            ws.search(n, kl, True, algo="auto", out_width=1)
            st = ws.sync().as_dict()
            if best is None or st["ms_search"] < best["ms_search"]:
                best = st
        cs = ws.checksum()
        want = want or cs
        row = {"gsize": gsize, "dcap": dcap, "same": cs == want}
        row.update({k: (round(v, 4) if isinstance(v, float) else v) for k, v in best.items()
                    if k in ("ms_search", "ms_pass1", "ms_cert", "ms_pass3", "lf_steps", "lf_steps_pass1")})
        print(json.dumps(row), flush=True)
