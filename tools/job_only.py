"""The job of bench.py alone (write_unique_counts over the whole workload genome with RANK / WORLD_SIZE), with the
phase trace of every rank -- for finding what bounds the sharded run.  Under torchrun:
    python -m torch.distributed.run --nproc-per-node 8 ... tools/job_only.py c4_3g [repeats]"""
import json
import os
import shutil
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c4_3g"
repeats = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if rank == 0:
    bench.prepare_workload(workload, device=local)
if world > 1:
    dist.barrier()
fa, ixp = bench.prepare_workload(workload, device=local)
from newmap_b200 import search  # noqa: E402
from newmap_b200.count import open_index  # noqa: E402
t0 = time.perf_counter()
open_index(ixp, local)
t_open = time.perf_counter() - t0
out = bench.CACHE / "job_only_out"
kl = list(range(20, 201))
for rep in range(repeats):
    if rank == 0:
        shutil.rmtree(out, ignore_errors=True)
        out.mkdir(parents=True)
    if world > 1:
        dist.barrier()
    cfg = search.SearchConfig(fasta_filepaths=[fa], fmindex_filepaths=[ixp], kmer_lengths=kl, is_binary_search=True,
                              kmer_batch_size=bench.BATCH, output_directory=out, device=local, rank=rank, world_size=world)
    t0 = time.perf_counter()
    search.write_unique_counts(cfg)
    mine = time.perf_counter() - t0
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0
    print(json.dumps({"rep": rep, "rank": rank, "world": world, "index_open_s": round(t_open, 3), "mine_s": round(mine, 3),
                      "wall_s": round(wall, 3), "partitioned": search.PARTITIONED, "slot_cache": search.SLOT_CACHE,
                      "trace": {k: round(v, 3) for k, v in search.LAST_RUN_TRACE}}), flush=True)
if world > 1:
    dist.destroy_process_group()
