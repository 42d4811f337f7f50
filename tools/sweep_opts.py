"""Tuning sweep on a bench workload: the device-resident arm of bench.py (a pool of workspaces over the parsed genome)
timed for several settings of the workspace options.  One process, index opened once.
    python tools/sweep_opts.py c3_1g 'gsize2=0,256;stall=0,2;dcap=0,22;pool=10' [segments]"""
import itertools
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib  # noqa: E402
from newmap_b200.count import open_index  # noqa: E402
from newmap_b200.fasta import window_plan  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3_1g"
spec = sys.argv[2] if len(sys.argv) > 2 else "gsize2=0,256"
S = int(sys.argv[3]) if len(sys.argv) > 3 else 100
krange = sys.argv[4] if len(sys.argv) > 4 else "20:200"
kmin, kmax = (int(x) for x in krange.split(":"))
axes = [(a.split("=")[0], [int(v) for v in a.split("=")[1].split(",")]) for a in spec.split(";") if a]
fa, ixp = bench.prepare_workload(workload)
index = open_index(ixp, 0)
fa_dev = _lib.Fasta(0, size_hint=fa.stat().st_size)
fa_dev.feed_file(fa)
fa_dev.finish()
windows = []
for _h, off, ln in fa_dev.records:
    for start, length, epi in window_plan(ln, bench.BATCH + kmax - 1, kmax - 1):
        windows.append((off + start, length, length if epi else length - (kmax - 1)))
mine = [windows[j % len(windows)] for j in range(S)]
kl = list(range(kmin, kmax + 1))
max_len = max(ln for _, ln, _ in mine)
import torch  # noqa: E402  (events only)
base_sums = None
for combo in itertools.product(*[v for _, v in axes]):
    opts = dict(zip([n for n, _ in axes], combo))
    npool = opts.pop("pool", 10)
    pool = [_lib.Workspace(index, max_len) for _ in range(npool)]
    for ws in pool:
        for k, v in opts.items():
            ws.option(k, v)

    def step(collect=None, sums=None):
        busy = [False] * npool
        for j, (off, ln, n) in enumerate(mine):
            ws = pool[j % npool]
            if busy[j % npool]:
                st = ws.sync()
                if collect is not None:
                    collect.append(st.as_dict())
                if sums is not None:
                    sums.append(ws.checksum())
            ws.attach(fa_dev, off, ln)
            ws.search(n, kl, True, out_width=1)
            busy[j % npool] = True
        for j in range(max(0, len(mine) - npool), len(mine)):
            st = pool[j % npool].sync()
            if collect is not None:
                collect.append(st.as_dict())
            if sums is not None:
                sums.append(pool[j % npool].checksum())

    sums: list = []
    step(sums=sums)
    if base_sums is None:
        base_sums = sums
    same = sums == base_sums
    step()
    torch.cuda.synchronize()
    stats: list = []
    t0 = time.perf_counter()
    for _ in range(5):
        step(stats)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    pos = sum(n for _, _, n in mine)
    iso = []
    for off, ln, n in mine[:10]:
        pool[0].attach(fa_dev, off, ln)
        pool[0].search(n, kl, True, out_width=1)
        iso.append(pool[0].sync().as_dict())
    m = lambda k: round(sum(s[k] for s in iso) / len(iso), 4)  # noqa: E731
    print(json.dumps({"workload": workload, "k": krange, **dict(zip([n for n, _ in axes], combo)), "gpos_per_s": round(pos / dt / 1e9, 3),
                      "same_bytes": same, "ms_prep": m("ms_prep"), "ms_pass1": m("ms_pass1"), "ms_cert": m("ms_cert"),
                      "ms_pass3": m("ms_pass3"), "ms_resolve": m("ms_resolve"),
                      "lf_per_pos": round(sum(s["lf_steps"] for s in iso) / sum(s["positions"] for s in iso), 3),
                      "lf1_per_pos": round(sum(s["lf_steps_pass1"] for s in iso) / sum(s["positions"] for s in iso), 3),
                      "seeds_per_pos": round(sum(s["seed_lookups"] for s in iso) / sum(s["positions"] for s in iso), 3),
                      "stride": iso[0]["stride"], "seed_depth": iso[0]["seed_depth"]}), flush=True)
    for ws in pool:
        ws.close()
