"""Whole `newmap search` run (newmap_b200.search.write_unique_counts: FASTA parse, H2D, kernels, D2H, file writes)
timed on a bench workload, with a host-side profile.  `python tools/cli_e2e.py [workload] [--profile]`"""
import cProfile
import io
import json
import pstats
import shutil
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from newmap_b200.search import SearchConfig, write_unique_counts  # noqa: E402
from newmap_b200 import search as S  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "c3_1g"
fa, ixp = bench.prepare_workload(workload)
out = Path("/tmp/newmap_b200_cli_out")
shutil.rmtree(out, ignore_errors=True)
out.mkdir(parents=True)
cfg = SearchConfig(fasta_filepaths=[fa], fmindex_filepaths=[ixp], kmer_lengths=list(range(20, 201)),
                   is_binary_search=True, kmer_batch_size=10_000_000, output_directory=out)
t0 = time.perf_counter()
S.open_index(ixp, None)          # index load (file read + upload + derived layouts), cached for the search below
t_open = time.perf_counter() - t0
pr = cProfile.Profile() if "--profile" in sys.argv else None
t0 = time.perf_counter()
if pr:
    pr.enable()
write_unique_counts(cfg)
if pr:
    pr.disable()
t_search = time.perf_counter() - t0
positions = sum(p.stat().st_size for p in out.glob("*.unique.uint8"))
print(json.dumps({"workload": workload, "positions": positions, "index_open_s": round(t_open, 3),
                  "search_s": round(t_search, 3), "positions_per_s_search": positions / t_search,
                  "positions_per_s_with_open": positions / (t_search + t_open)}), flush=True)
if pr:
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(18)
    print(s.getvalue()[:4000])
