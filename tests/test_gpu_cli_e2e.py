"""The reference's end-to-end script (tests/test_end_to_end.sh:1-22: `newmap index` -> `newmap search --verbose`
-> `newmap track` twice) through `python -m newmap_b200`, compared with what the UNMODIFIED reference CLI wrote
for the same commands (tests/golden/e2e, made by oracle/gen_e2e_golden.py): index file, unique-length files and
both track outputs byte for byte; every line of our verbose log must be a line of the reference's, in order."""
import json
import os
import shutil
import subprocess
import sys
from pathlib import Path

import pytest

from conftest import GOLD, ROOT

pytestmark = pytest.mark.gpu
E2E = GOLD / "e2e"


def _is_subsequence(ours: list, theirs: list) -> tuple:
    it = iter(theirs)
    for line in ours:
        if not any(line == t for t in it):
            return False, line
    return True, ""


def test_end_to_end_script_through_the_cli(tmp_path):
    (tmp_path / "data").mkdir()
    shutil.copy(GOLD / "genome.fa", tmp_path / "data" / "genome.fa")
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    v = subprocess.run([sys.executable, "-m", "newmap_b200", "--version"], cwd=tmp_path, env=env, capture_output=True)
    assert v.returncode == 0 and v.stdout.startswith(b"newmap")
    for step in json.loads((E2E / "commands.json").read_text()):
        r = subprocess.run([sys.executable, "-m", "newmap_b200"] + step["argv"], cwd=tmp_path, env=env, capture_output=True)
        assert r.returncode == 0, (step["argv"], r.stderr.decode()[-2000:])
        assert r.stdout.decode() == step["stdout"], step["argv"]
        ok, line = _is_subsequence(r.stderr.decode().splitlines(), step["stderr"].splitlines())
        assert ok, (step["argv"], "line not in the reference's log:", line)
        if step["argv"][0] == "search":
            # the per-sequence summary of --verbose (search.py:885-902) is there in full
            for want in [t for t in step["stderr"].splitlines() if "unique length" in t or "ambiguity" in t and "skipped" in t
                         or "maximum unique" in t or "minimum unique" in t]:
                assert want in r.stderr.decode().splitlines(), want
    assert (tmp_path / "genome.awfmi").read_bytes() == (E2E / "genome.awfmi").read_bytes()
    for f in sorted((E2E / "unique_lengths").iterdir()):
        assert (tmp_path / "unique_lengths" / f.name).read_bytes() == f.read_bytes(), f.name
    assert sorted(p.name for p in (tmp_path / "unique_lengths").iterdir()) == sorted(p.name for p in (E2E / "unique_lengths").iterdir())
