"""Host cost of one search call sequence (attach/upload, search, download, sync) with next to no GPU work: tiny segments
on the golden index, a pool of workspaces like bench.py.  `python tools/host_overhead.py`"""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from newmap_b200 import _lib  # noqa: E402

gold = Path(__file__).resolve().parent.parent / "tests" / "golden"
ix = _lib.Index(gold / "quirk_s6.awfmi", device=0)
seq = b"".join((gold / "quirk.fa").read_bytes().split(b">")[1].split(b"\n")[1:])[40:40 + 1199]
n = 1000
kl = list(range(20, 201))
pool = [_lib.Workspace(ix, len(seq)) for _ in range(10)]
inp = _lib.PinnedBuffer(len(seq))
inp.array[:] = np.frombuffer(seq, dtype=np.uint8)
outs = [_lib.PinnedBuffer(n) for _ in pool]
res = {}
for mode in ("search_only", "upload_search_download"):
    for ws in pool:
        ws.upload(inp.ptr, len(seq))
    reps = 3000
    for timed in (False, True):
        t0 = time.perf_counter()
        busy = [False] * len(pool)
        for j in range(reps):
            w = j % len(pool)
            if busy[w]:
                pool[w].sync()
            if mode != "search_only":
                pool[w].upload(inp.ptr, len(seq))
            pool[w].search(n, kl, True, out_width=1)
            if mode != "search_only":
                pool[w].download(outs[w].ptr)
            busy[w] = True
        for ws in pool:
            ws.sync()
        dt = time.perf_counter() - t0
    res[mode + "_us_per_search"] = round(dt / reps * 1e6, 1)
print(json.dumps(res))
