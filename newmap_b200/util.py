"""Small host helpers with the reference's names (newmap/util.py:1-23)."""
from __future__ import annotations

import gzip
import sys
from pathlib import Path

INDEX_EXTENSION = "awfmi"


def optional_gzip_open(file_path, mode: str):
    """gzip.open for *.gz, plain open otherwise (newmap/util.py:10-18)."""
    file_path = Path(file_path)
    return gzip.open(file_path, mode) if file_path.suffix == ".gz" else open(file_path, mode)


def verbose_print(verbose: bool, *args):
    if verbose:
        print(*args, file=sys.stderr)
