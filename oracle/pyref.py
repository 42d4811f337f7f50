"""oracle/pyref.py -- TEST INFRASTRUCTURE ONLY.

Imports the UNMODIFIED Python reference (newmap/search.py, newmap/fasta.py) from
/root/reference with the reference's own C extensions (oracle/_ref) bound in place of the
``newmap._c_newmap_*`` modules it expects.  Works only where /root/reference exists (the build
container): used by oracle/gen_golden.py to produce tests/golden/ and by the differential
tests that skip themselves elsewhere.  Nothing on the GPU box imports this.
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

REFERENCE_ROOT = Path("/root/reference")


def available() -> bool:
    from . import ora
    return (REFERENCE_ROOT / "newmap" / "search.py").is_file() and ora.ref_available()


def load():
    """Returns the reference's (search_module, fasta_module)."""
    from . import ora
    if "newmap.search" in sys.modules and getattr(sys.modules["newmap"], "__graft_ref__", False):
        return sys.modules["newmap.search"], sys.modules["newmap.fasta"]
    count_mod, gen_mod = ora.ref_modules()
    # newmap/__init__.py is empty; newmap/main.py (importlib.metadata.version) is not imported.
    pkg = types.ModuleType("newmap")
    pkg.__path__ = [str(REFERENCE_ROOT / "newmap")]
    pkg.__graft_ref__ = True
    sys.modules["newmap"] = pkg
    sys.modules["newmap._c_newmap_count_kmers"] = count_mod
    sys.modules["newmap._c_newmap_generate_index"] = gen_mod
    import importlib
    search = importlib.import_module("newmap.search")
    fasta = importlib.import_module("newmap.fasta")
    return search, fasta


def run_search(fasta_path, index_path, kmer_lengths, is_range, kmer_batch_size, outdir,
               initial_search_length=0, use_rc=True, num_threads=1, include=(), exclude=()):
    """write_unique_counts of the reference; returns {id: numpy array} read back from outdir."""
    import numpy as np
    search, _ = load()
    outdir = Path(outdir)
    outdir.mkdir(parents=True, exist_ok=True)
    cfg = search.SearchConfig(
        fasta_filepaths=[Path(fasta_path)], fmindex_filepaths=[Path(index_path)],
        kmer_lengths=list(kmer_lengths), is_binary_search=is_range,
        use_reverse_complement=use_rc, output_directory=outdir,
        include_sequence_ids=list(include), exclude_sequence_ids=list(exclude),
        num_threads=num_threads, kmer_batch_size=kmer_batch_size,
        initial_search_length=initial_search_length)
    search.write_unique_counts(cfg)
    max_k = max(kmer_lengths)
    suffix = "uint8" if max_k <= 255 else ("uint16" if max_k <= 65535 else "uint32")
    res = {}
    for f in sorted(outdir.glob(f"*.unique.{suffix}")):
        res[f.name[: -len(f".unique.{suffix}")].encode()] = np.fromfile(f, dtype=getattr(np, suffix))
    return res
