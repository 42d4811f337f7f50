"""Random-access roofline probe sweep (run on the GPU box): loads/s of dependent random loads by
granule, working-set size and chains per thread.  Writes JSON lines to stdout."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from newmap_b200 import _lib  # noqa: E402

for nbytes in (64 << 20, 512 << 20, 1 << 30, 2 << 30, 8 << 30):
    for granule in (32, 64, 128):
        for dep in (True, False):
          for mlp in (1, 4, 8):
            for chains in (1 << 20, 1 << 23):
                rate, ms = _lib.probe_random_access(0, nbytes, granule, chains, 32, mlp, dep)
                print(json.dumps({"bytes": nbytes, "granule": granule, "mlp": mlp, "threads": chains, "dependent": dep,
                                  "loads_per_s": rate, "gbs": rate * granule / 1e9, "ms": ms}), flush=True)
