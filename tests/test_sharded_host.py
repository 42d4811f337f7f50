"""N > 1 host path on CPU: world_size-2 and -3 gloo groups run the product's write_unique_counts with
position sharding (segments dealt to ranks, results gathered into shared files at
window * batch offsets); the files must equal the reference-made goldens byte for byte."""
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import GOLD, run_klist
import _shard_worker


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,run_id_mod", [(2, 0), (3, 1)])
def test_sharded_write_unique_counts_matches_goldens(manifest, golden_searches, tmp_path, world, run_id_mod):
    runs = [r for r in manifest["searches"] if r["index"] == "quirk_s4" and r["batch"] in (257, 1000)
            and not r.get("norc") and not r.get("initial")][run_id_mod::2][:2]
    assert runs
    for run in runs:
        out = tmp_path / f"w{world}_r{run['id']}"
        out.mkdir()
        # stale, too-long files must not survive
        (out / "chr1.unique.uint8").write_bytes(b"\x07" * 100000)
        mp.spawn(_shard_worker.run, nprocs=world, join=True,
                 args=(world, _free_port(), str(GOLD / run["fasta"]), str(GOLD / f"{run['index']}.awfmi"),
                       run_klist(run), run["range"], run["batch"], str(out)))
        for sid in run["sequences"]:
            want = golden_searches[f"run{run['id']}|{sid}"]
            got = np.fromfile(out / f"{sid}.unique.{want.dtype.name}", dtype=want.dtype)
            assert np.array_equal(got, want), (world, run["id"], sid)
