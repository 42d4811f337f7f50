"""oracle/gen_e2e_golden.py -- TEST INFRASTRUCTURE ONLY.

Runs the reference's own end-to-end script (tests/test_end_to_end.sh:1-22) through the UNMODIFIED reference CLI
(newmap/main.py parse_subcommands, imported from /root/reference with the reference's compiled extensions from
oracle/_ref bound as newmap._c_newmap_*) and keeps what it wrote: the index, the unique-length files, the verbose
logs and both track outputs.  tests/test_gpu_cli_e2e.py runs the same four commands through `python -m newmap_b200`
and compares bytes.  Run where /root/reference exists:
    python -m oracle.gen_e2e_golden"""
from __future__ import annotations

import importlib
import importlib.metadata
import json
import os
import shutil
import sys
import tempfile
from pathlib import Path

from . import pyref

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "tests" / "golden" / "e2e"
# the four commands of tests/test_end_to_end.sh (newmap --version aside), cwd = a directory holding data/genome.fa
COMMANDS = [
    ["index", "--compression-ratio=32", "--seed-length=1", "data/genome.fa"],
    ["search", "--verbose", "--output-directory=unique_lengths", "--search-range=4:10", "data/genome.fa"],
    ["track", "--verbose", "unique_lengths/chr1.unique.uint8", "unique_lengths/chr2.unique.uint8"],
    ["track", "--verbose", "--multi-read", "-", "8", "unique_lengths/chr1.unique.uint8", "unique_lengths/chr2.unique.uint8"],
]


def reference_main():
    pyref.load()
    real = importlib.metadata.version
    importlib.metadata.version = lambda name: "0.0" if name == "newmap" else real(name)   # not pip-installed here
    try:
        return importlib.import_module("newmap.main")
    finally:
        importlib.metadata.version = real


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--run":       # one command of the reference CLI, in this process
        ref = reference_main()
        sys.argv = ["newmap"] + sys.argv[2:]
        ref.parse_subcommands()
        return
    import subprocess
    shutil.rmtree(OUT, ignore_errors=True)
    (OUT / "unique_lengths").mkdir(parents=True)
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "data").mkdir()
        shutil.copy(ROOT / "tests" / "golden" / "genome.fa", tmp / "data" / "genome.fa")
        logs = []
        env = dict(os.environ, PYTHONPATH=str(ROOT))
        for cmd in COMMANDS:
            r = subprocess.run([sys.executable, "-m", "oracle.gen_e2e_golden", "--run"] + cmd, cwd=tmp, env=env,
                               capture_output=True, check=True)
            logs.append({"argv": cmd, "stdout": r.stdout.decode(), "stderr": r.stderr.decode()})
        shutil.copy(tmp / "genome.awfmi", OUT / "genome.awfmi")
        for f in sorted((tmp / "unique_lengths").iterdir()):
            shutil.copy(f, OUT / "unique_lengths" / f.name)
    (OUT / "commands.json").write_text(json.dumps(logs, indent=1) + "\n")
    print(f"{len(logs)} commands, {sum(p.stat().st_size for p in OUT.rglob('*') if p.is_file())} bytes in {OUT}")


if __name__ == "__main__":
    main()
