"""Worker of tests/test_sharded_host.py: one rank of a world_size-N gloo group running the product's
host driver (newmap_b200.search.write_unique_counts) with the GPU pieces replaced by an
oracle-backed stand-in, so the sharding / gather logic is exercised on CPU."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


class _Buf:
    def __init__(self, n):
        self.array = np.zeros(n, dtype=np.uint8)
        self.ptr = self.array          # the fake workspace takes the array itself as "pointer"

    def close(self):
        pass


class _FakeStats:
    zero_pos = -1


class _FakeWorkspace:
    def __init__(self, oracle_index):
        self.o = oracle_index

    def upload(self, buf, n):
        self.seg = bytes(buf[:n])

    def search(self, num_kmers, klist, is_range, initial, use_rc, algo, width):
        self.res = self.o.min_unique_segment(self.seg, num_kmers, klist, is_range, initial, use_rc, nthreads=1)
        self.width = width

    def download(self, out):
        dt = {1: np.uint8, 2: np.uint16, 4: np.uint32}[self.width]
        raw = self.res.astype(dt).view(np.uint8)
        out[:raw.size] = raw

    def sync(self):
        return _FakeStats()

    def close(self):
        pass


def run(rank, world, port, fasta, index, klist, is_range, batch, outdir):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from newmap_b200 import search
    from oracle import ora
    oracle_index = ora.OracleIndex(index)

    class Slot:
        def __init__(self, ix, max_len, itemsize):
            self.ws, self.inp, self.out, self.job = _FakeWorkspace(oracle_index), _Buf(max_len), _Buf(max_len * itemsize), None

        def close(self):
            pass

    search.open_index = lambda path, device=None: object()
    search._Slot = Slot
    search.INGEST = "host"       # the host reader + per-segment upload path (the device ingest path needs a GPU)
    dist.barrier()
    search.write_unique_counts(search.SearchConfig(
        fasta_filepaths=[Path(fasta)], fmindex_filepaths=[Path(index)], kmer_lengths=list(klist),
        is_binary_search=is_range, kmer_batch_size=batch, output_directory=Path(outdir), rank=rank, world_size=world))
    dist.barrier()
    dist.destroy_process_group()


def run_partitioned(rank, world, port, fasta, index, klist, is_range, batch, outdir):
    """One rank of a sharded run as a launcher starts it (RANK / WORLD_SIZE / MASTER_* in the environment), all ranks
    on GPU 0: the real device path with the partitioned ingest (tests/test_gpu_partitioned.py)."""
    import os
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                      NEWMAP_B200_DEVICE="0")
    from newmap_b200 import search
    search.INGEST = "device"
    search.write_unique_counts(search.SearchConfig(
        fasta_filepaths=[Path(fasta)], fmindex_filepaths=[Path(index)], kmer_lengths=list(klist),
        is_binary_search=is_range, kmer_batch_size=batch, output_directory=Path(outdir), device=0, rank=rank, world_size=world))
    trace = dict(search.LAST_RUN_TRACE)
    assert "exchanged" in trace and "planned" in trace, trace      # the partitioned ingest really ran
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()
