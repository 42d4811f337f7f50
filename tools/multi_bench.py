"""Several FASTA x several index mode at size: the C->T and G->A conversions of the 100 Mbp bench
genome with their two indexes (GPU builder), k=20:200, 10 M-position segments through nm_workspace_multi; the CPU leg
is the oracle (oracle/fm_oracle.c ora_min_unique_multi, one thread) on a bounded sample of the first segment.
    python tools/multi_bench.py [segments] [cpu_sample_positions]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200 import _lib, index, synth  # noqa: E402
from oracle import ora  # noqa: E402

nseg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
cpu_sample = int(sys.argv[2]) if len(sys.argv) > 2 else 20_000
d = bench.CACHE / "multi_100m"
d.mkdir(parents=True, exist_ok=True)
lengths, seed = synth.config_lengths("c2_100m")
recs = synth.make_genome(lengths, seed)
paths = {}
for name, src, dst in (("ct", b"Cc", b"Tt"), ("ga", b"Gg", b"Aa")):
    fa, ix = d / f"{name}.fa", d / f"{name}.awfmi"
    if not ix.is_file():
        t = np.arange(256, dtype=np.uint8)
        for x, y in zip(src, dst):
            t[x] = y
        synth.write_fasta(fa, [(n, t[s]) for n, s in recs])
        index.generate_fm_index(fa, ix, bench.SA_RATIO, bench.SEED_LEN, device=0)
    paths[name] = (fa, ix)
kl = list(range(20, 201))
segs = [bench.load_segments(paths[n][0], 1, nseg)[0] for n in ("ct", "ga")]
ixs = [_lib.Index(paths[n][1], device=0, seed_depth=0) for n in ("ct", "ga")]
cap = max(len(s.data) for s, _ in segs[0])
ws = _lib.Workspace(ixs[0], cap)
out = _lib.PinnedBuffer(cap)
res = {"workload": "C->T and G->A conversions of the 100 Mbp bench genome, 2 FASTA x 2 indexes, k=20:200", "segments": nseg}
tot_ms, tot_n, lf, nq = 0.0, 0, 0, 0
first_out = None
for j in range(nseg):
    (a, n), (b, _) = segs[0][j], segs[1][j]
    A, B = _lib.as_u8(a.data), _lib.as_u8(b.data)
    best = None
    for _ in range(2):
        ws.upload(A.ctypes.data, A.size)
        ws.multi(ixs, [B.ctypes.data])
        ws.search(n, kl, True, algo="trajectory", out_width=1)
        ws.download(out.ptr)
        st = ws.sync()
        assert st.zero_pos < 0 and st.short_counts == 0
        best = st.ms_search if best is None else min(best, st.ms_search)
    if first_out is None:
        first_out = out.array[:n].copy()
    tot_ms += best
    tot_n += n
    lf += st.lf_steps
    nq += st.queries
res.update({"gpu_positions_per_s": tot_n / (tot_ms * 1e-3), "gpu_ms_per_segment": tot_ms / nseg,
            "lf_steps_per_position": lf / tot_n, "queries_per_position": nq / tot_n,
            "unique_fraction": float(np.count_nonzero(first_out)) / first_out.size})
# CPU leg: the oracle on the first cpu_sample positions of segment 1 (buffer cut to sample + look-ahead)
(a, n), (b, _) = segs[0][0], segs[1][0]
m = min(cpu_sample, n)
cut = min(len(a.data), m + 199)
oix = [ora.OracleIndex(paths[nm][1]) for nm in ("ct", "ga")]
t0 = time.perf_counter()
want = ora.min_unique_multi_segment(oix, [bytes(a.data[:cut]), bytes(b.data[:cut])], m, kl, True)
dt = time.perf_counter() - t0
# barriers are ambiguous START positions among the first num_kmers bytes only (look-ahead quirk), and num_kmers is m here:
# compare the positions whose 200-byte window ends before m
safe = max(0, m - 200)
res.update({"cpu_sample_positions": m, "cpu_positions_per_s": m / dt, "cpu_threads": 1,
            "sample_equal": bool(np.array_equal(want[:safe].astype(np.uint8), first_out[:safe]))})
print(json.dumps(res))
