"""Short single-GPU program for ncu: a few searches of one 10 M-position segment of a bench
workload (default c2_100m).  `python tools/profile_one.py [workload] [repeats] [algo] [segment] [k range]`."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c2_100m"
repeats = int(sys.argv[2]) if len(sys.argv) > 2 else 3
algo = sys.argv[3] if len(sys.argv) > 3 else "auto"
segment = int(sys.argv[4]) if len(sys.argv) > 4 else 1
krange = sys.argv[5] if len(sys.argv) > 5 else "20:200"
bench.K_RANGE = tuple(int(x) for x in krange.split(":"))
fa, ixp = bench.prepare_workload(workload)
seg, n = bench.sample_segment(fa, segment)
ix = _lib.Index(ixp, device=0)
ws = _lib.Workspace(ix, len(seg.data))
pb = _lib.PinnedBuffer(len(seg.data))
pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
ws.upload(pb.ptr, len(seg.data))
kl = list(range(bench.K_RANGE[0], bench.K_RANGE[1] + 1))
for _ in range(repeats):
    ws.search(n, kl, True, algo=algo, out_width=1)
    st = ws.sync()
print({k: v for k, v in st.as_dict().items() if k.startswith("ms_") or "steps" in k or "lookups" in k or k in ("generic_searches", "stride", "launches")})
