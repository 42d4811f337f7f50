"""One random-access probe launch per mode, for ncu (dram sectors per load, L2 hit rate)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from newmap_b200 import _lib  # noqa: E402

for nbytes in (1 << 30, 4 << 30):
    for granule in (32, 64, 128):
        print(nbytes >> 20, granule, _lib.probe_random_access(0, nbytes, granule, 4 << 20, 16, 1, True))
