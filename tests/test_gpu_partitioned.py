"""A sharded run as a launcher starts it (RANK / WORLD_SIZE / MASTER_ADDR in the environment): every rank reads and
parses only its byte range of the FASTA (newmap_b200/search.py _partitioned_ingest), the ranks exchange header lines
and base counts (gloo), and the shared output files must equal what the unmodified reference wrote.  All ranks use
GPU 0 here (nothing in a kernel waits for another rank)."""
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import GOLD, run_klist
import _shard_worker

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_sharded_run_matches_goldens(manifest, golden_searches, tmp_path, world):
    runs = [r for r in manifest["searches"] if r["fasta"] in ("quirk.fa", "quirk_iupac.fa") and r["batch"] in (257, 1000, 1500)
            and not r.get("include") and not r.get("exclude") and not r.get("norc") and not r.get("initial")][world::5][:3]
    assert runs
    for run in runs:
        out = tmp_path / f"w{world}_r{run['id']}"
        out.mkdir()
        (out / "chr1.unique.uint8").write_bytes(b"\x07" * 100000)      # stale, too-long files must not survive
        mp.spawn(_shard_worker.run_partitioned, nprocs=world, join=True,
                 args=(world, _free_port(), str(GOLD / run["fasta"]), str(GOLD / f"{run['index']}.awfmi"),
                       run_klist(run), run["range"], run["batch"], str(out)))
        for sid in run["sequences"]:
            want = golden_searches[f"run{run['id']}|{sid}"]
            got = np.fromfile(out / f"{sid}.unique.{want.dtype.name}", dtype=want.dtype)
            assert np.array_equal(got, want), (world, run["id"], sid)
