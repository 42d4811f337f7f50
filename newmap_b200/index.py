"""`newmap index` on the GPU: FASTA -> reference-format .awfmi.

Same entry point name and argument order as the reference's extension
(`generate_fm_index(fasta, index, compression_ratio, seed_length)`, src/newmap-generate-index.c:11-82,
newmap/index.py:7-23).  The FASTA is read with FastaVector's rules
(AvxWindowFmIndex/lib/FastaVector/src/FastaVector.c:50-168); suffix sorting, BWT blocks, seed table
and sampled suffix array are built by nm_index_build on the device, and the file written is byte
for byte what the reference's CPU builder writes (tests/test_gpu_build.py)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

DEFAULT_COMPRESSION_RATIO = 8
DEFAULT_SEED_LENGTH = 12
_DROP_SEQ = bytes(range(32)) + bytes(range(128, 256))       # FastaVector keeps sequence chars >= ' ' (signed char)
_DROP_HDR = bytes(c for c in range(32) if c != 9) + bytes(range(128, 256))


def read_fasta_vector(path):
    """(sequence uint8 array, header bytes, metadata uint64 array of shape (n, 2)) as FastaVector builds them."""
    with open(path, "rb") as fh:
        data = fh.read()
    if b";" in data:
        raise NotImplementedError("newmap_b200.index: ';' comments in FASTA are not supported by the GPU builder")
    if data and not data.startswith(b">"):
        raise ValueError("newmap_b200.index: FASTA must start with a '>' header line")
    seq_parts, hdr_parts, meta = [], [], []
    hdr_end = seq_end = 0
    pos = 0 if data.startswith(b">") else -1
    while pos >= 0:
        eol = data.find(b"\n", pos)
        if eol < 0:
            eol = len(data)
        h = data[pos + 1:eol].translate(None, _DROP_HDR) + (b"\0" if eol < len(data) else b"")
        nxt = data.find(b"\n>", eol)
        body = data[eol + 1:(nxt + 1 if nxt >= 0 else len(data))].translate(None, _DROP_SEQ) + b"\0"
        hdr_parts.append(h)
        seq_parts.append(body)
        hdr_end += len(h)
        seq_end += len(body)
        meta.append((hdr_end, seq_end))
        pos = nxt + 1 if nxt >= 0 else -1
    seq = np.frombuffer(b"".join(seq_parts), dtype=np.uint8)
    return seq, b"".join(hdr_parts), np.array(meta, dtype=np.uint64).reshape(-1, 2)


def generate_fm_index(fasta_path, index_path, compression_ratio: int = DEFAULT_COMPRESSION_RATIO,
                      seed_length: int = DEFAULT_SEED_LENGTH, device: int | None = None) -> float:
    """Builds `index_path` from `fasta_path` on the GPU; returns the device seconds of the suffix sort."""
    if not os.path.isfile(fasta_path):
        raise FileNotFoundError(f"No such file: {fasta_path}")      # newmap-generate-index.c:32-36
    if device is None:
        device = int(os.environ.get("NEWMAP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    seq, hdr, meta = read_fasta_vector(fasta_path)
    if seq.size == 0:
        raise ValueError("empty FASTA")
    if np.any((seq | 0x20) == ord("u")) or np.any(seq == ord("$")) or np.any(seq == 0x04):
        raise ValueError("newmap_b200.index: U/u or '$' in the sequence (the reference's own index is inconsistent "
                         "for these inputs)")
    L = _lib.load()
    secs = C.c_double(0)
    hb = np.frombuffer(hdr, dtype=np.uint8) if hdr else np.zeros(1, dtype=np.uint8)
    rc = L.nm_index_build(seq.ctypes.data, seq.size, hb.ctypes.data, len(hdr), meta.ctypes.data, meta.shape[0],
                          int(compression_ratio), int(seed_length), device, os.fsencode(str(index_path)),
                          C.byref(secs))
    if rc != 0:
        raise _lib.NewmapError(rc, f"nm_index_build ({L.nm_build_last_error().decode()})")
    return secs.value


def main(args):
    generate_fm_index(args.fasta_file, args.output or os.path.splitext(os.path.basename(args.fasta_file))[0] + ".awfmi",
                      args.compression_ratio, args.seed_length)
