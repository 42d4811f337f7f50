"""oracle/ora.py -- TEST INFRASTRUCTURE ONLY (CPU checker), never the product path.

Python face of the CPU oracle:

* ctypes binding of ``liboracle.so`` (``fm_oracle.c``: scalar restatement of the reference's
  FM-index count and of newmap/search.py replayed per position);
* ``ref_modules()``: the reference's own two C extensions compiled from the sources where
  they lie under /root/reference into ``oracle/_ref`` (see ``Makefile``);
* ``segments()``: closed-form restatement of newmap/fasta.py:20-190;
* ``min_unique()``: whole-FASTA oracle (segmenter + per-position search) returning the
  arrays the reference would write to ``<id>.unique.uintN``.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference``
legs may import this module.  Parity status: PINNED (tests/test_oracle.py: reference golden
vectors + differential runs against oracle/_ref and against the imported Python reference,
fixtures under tests/golden/).
"""
from __future__ import annotations

import ctypes as C
import gzip
import importlib.machinery
import importlib.util
import os
import subprocess
from pathlib import Path
from typing import Iterator, NamedTuple

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "liboracle.so"
REF_DIR = HERE / "_ref"


def build(ref: bool = True) -> None:
    """Compile liboracle.so and, when /root/reference is present, oracle/_ref."""
    targets = ["oracle"]
    if ref and Path("/root/reference/src/newmap-count.c").is_file():
        targets.append("ref")
    subprocess.run(["make", "-s", "-C", str(HERE)] + targets, check=True)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.is_file():
            build(ref=False)
        L = C.CDLL(str(LIB_PATH))
        u8p, u32p, u64p, i64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint32, C.c_uint64, C.c_int64))
        L.ora_index_load.restype = C.c_void_p
        L.ora_index_load.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
        L.ora_index_free.argtypes = [C.c_void_p]
        L.ora_bwt_length.restype = C.c_uint64
        L.ora_bwt_length.argtypes = [C.c_void_p]
        L.ora_seed_length.restype = C.c_uint32
        L.ora_seed_length.argtypes = [C.c_void_p]
        L.ora_prefix_sums.argtypes = [C.c_void_p, u64p]
        for f in (L.ora_count, L.ora_count_rc):
            f.restype = C.c_int
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, u64p, u64p, C.c_uint64, u32p, u64p]
        L.ora_upper_bounds.argtypes = [u8p, C.c_uint64, C.c_uint64, C.c_uint32, u32p]
        L.ora_min_unique_range.restype = C.c_int
        L.ora_min_unique_range.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32,
                                           C.c_uint32, C.c_uint32, C.c_int, C.c_int, u32p, i64p, u64p, u64p]
        L.ora_min_unique_list.restype = C.c_int
        L.ora_min_unique_list.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, u32p, C.c_uint32,
                                          C.c_int, C.c_int, u32p, i64p, u64p, u64p]
        L.ora_min_unique_multi.restype = C.c_int
        L.ora_min_unique_multi.argtypes = [C.POINTER(C.c_void_p), C.c_uint32, C.POINTER(C.c_void_p), C.c_uint32,
                                           C.c_uint64, C.c_uint64, C.c_int, u32p, C.c_uint32, C.c_uint32, C.c_int,
                                           u32p, i64p, u64p, u64p]
        _lib = L
    return _lib


def _ptr(a: np.ndarray, t):
    return a.ctypes.data_as(C.POINTER(t))


class ZeroCount(RuntimeError):
    """The reference's RuntimeError for a generated k-mer absent from the index."""


def zero_count_message(kmer: str) -> str:
    # newmap/search.py:719-722
    return ("The following generated k-mer was not found in the index:\n"
            f"{kmer}\nPossibly a mismatch between the sequence and the index.")


class OracleIndex:
    """A reference-built .awfmi held in host memory by the C oracle."""

    def __init__(self, path):
        err = C.c_int(0)
        self._h = lib().ora_index_load(os.fsencode(str(path)), C.byref(err))
        if not self._h:
            raise OSError(f"oracle: could not load index {path} (code {err.value})")
        self.path = str(path)
        self.lf_steps = 0
        self.queries = 0

    def close(self):
        if self._h:
            lib().ora_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def bwt_length(self) -> int:
        return lib().ora_bwt_length(self._h)

    @property
    def seed_length(self) -> int:
        return lib().ora_seed_length(self._h)

    @property
    def prefix_sums(self) -> list[int]:
        out = (C.c_uint64 * 6)()
        lib().ora_prefix_sums(self._h, out)
        return list(out)

    def count(self, seq: bytes, idx, length, rc: bool = False) -> np.ndarray:
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        if idx.shape != length.shape:
            raise ValueError("Both lists of indices and lengths must be the same length")
        out = np.zeros(idx.size, dtype=np.uint32)
        steps = C.c_uint64(0)
        buf = np.frombuffer(seq, dtype=np.uint8)
        fn = lib().ora_count_rc if rc else lib().ora_count
        r = fn(self._h, buf.ctypes.data, buf.size, _ptr(idx, C.c_uint64), _ptr(length, C.c_uint64),
               idx.size, _ptr(out, C.c_uint32), C.byref(steps))
        if r != 0:
            raise IndexError("oracle: empty k-mer or index+length beyond the sequence")
        self.lf_steps += steps.value
        return out

    def count_kmers(self, kmers: list[bytes]) -> np.ndarray:
        lens = np.array([len(k) for k in kmers], dtype=np.uint64)
        offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint64) if len(kmers) else lens
        return self.count(b"".join(kmers), offs, lens)

    def min_unique_segment(self, seg: bytes, num_kmers: int, kmer_lengths, is_range: bool,
                           initial_search_length: int = 0, use_rc: bool = True,
                           nthreads: int = 0) -> np.ndarray:
        """uint32 array of num_kmers minimum unique lengths for one segment buffer; raises
        ZeroCount exactly where newmap/search.py:699-722 would."""
        buf = np.frombuffer(seg, dtype=np.uint8)
        out = np.zeros(max(num_kmers, 1), dtype=np.uint32)
        zero = np.full(3, -1, dtype=np.int64)
        steps, queries = C.c_uint64(0), C.c_uint64(0)
        nthreads = nthreads or (os.cpu_count() or 1)
        if is_range:
            r = lib().ora_min_unique_range(self._h, buf.ctypes.data, buf.size, num_kmers,
                                           min(kmer_lengths), max(kmer_lengths), initial_search_length,
                                           int(use_rc), nthreads, _ptr(out, C.c_uint32),
                                           _ptr(zero, C.c_int64), C.byref(steps), C.byref(queries))
        else:
            kl = np.ascontiguousarray(kmer_lengths, dtype=np.uint32)
            r = lib().ora_min_unique_list(self._h, buf.ctypes.data, buf.size, num_kmers,
                                          _ptr(kl, C.c_uint32), kl.size, int(use_rc), nthreads,
                                          _ptr(out, C.c_uint32), _ptr(zero, C.c_int64),
                                          C.byref(steps), C.byref(queries))
        if r != 0:
            raise ValueError(f"oracle: invalid search arguments (code {r})")
        self.lf_steps += steps.value
        self.queries += queries.value
        if zero[0] >= 0:
            p, ln = int(zero[0]), int(zero[1])
            raise ZeroCount(zero_count_message(bytes(seg[p:p + ln]).decode("utf-8")))
        return out[:num_kmers]


class ShortCount(RuntimeError):
    """A summed count below the number of sequences that is not 0: the reference's range loop never ends."""


def min_unique_multi_segment(indexes: list, segs: list, num_kmers: int, kmer_lengths, is_range: bool,
                             initial_search_length: int = 0, use_rc: bool = True) -> np.ndarray:
    """Several sequences x several indexes for one tuple of segment buffers (search.py:383-644 with
    count == num_sequences, :656-697)."""
    bufs = [np.frombuffer(s, dtype=np.uint8) for s in segs]
    assert all(b.size == bufs[0].size for b in bufs)
    ih = (C.c_void_p * len(indexes))(*[ix._h for ix in indexes])
    sp = (C.c_void_p * len(bufs))(*[b.ctypes.data for b in bufs])
    kl = np.ascontiguousarray([min(kmer_lengths), max(kmer_lengths)] if is_range else list(kmer_lengths), dtype=np.uint32)
    out = np.zeros(max(num_kmers, 1), dtype=np.uint32)
    zero = np.full(3, -1, dtype=np.int64)
    steps, queries = C.c_uint64(0), C.c_uint64(0)
    r = lib().ora_min_unique_multi(ih, len(indexes), sp, len(bufs), bufs[0].size, num_kmers, int(is_range),
                                   _ptr(kl, C.c_uint32), kl.size, initial_search_length, int(use_rc),
                                   _ptr(out, C.c_uint32), _ptr(zero, C.c_int64), C.byref(steps), C.byref(queries))
    if r < 0:
        raise ValueError(f"oracle: invalid search arguments (code {r})")
    if zero[0] >= 0:
        p, ln = int(zero[0]), int(zero[1])
        raise ZeroCount(zero_count_message(bytes(segs[0][p:p + ln]).decode("utf-8")))
    if r == 1:
        raise ShortCount("summed count below the number of sequences")
    return out[:num_kmers]


def min_unique_multi(fasta_paths: list, indexes: list, kmer_lengths, is_range: bool, kmer_batch_size: int,
                     initial_search_length: int = 0, use_rc: bool = True) -> dict:
    """{sequence id of the first file: array} for several FASTA / index files (search.py:222-357, zip(*sequences))."""
    max_k = max(kmer_lengths)
    dt = out_dtype(max_k)
    out: dict[bytes, list] = {}
    for tup in zip(*[segments(p, kmer_batch_size, max_k - 1) for p in fasta_paths]):
        seg = tup[0]
        n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)
        vals = min_unique_multi_segment(indexes, [t.data for t in tup], n, kmer_lengths, is_range,
                                        initial_search_length, use_rc)
        out.setdefault(seg.id, []).append(vals.astype(dt))
    return {k: np.concatenate(v) for k, v in out.items()}


def upper_bounds(mask, seg_len: int, max_k: int) -> np.ndarray:
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    ub = np.zeros(max(mask.size, 1), dtype=np.uint32)
    lib().ora_upper_bounds(_ptr(mask, C.c_uint8), mask.size, seg_len, max_k, _ptr(ub, C.c_uint32))
    return ub[:mask.size]


# --------------------------------------------------------------------------------------------
# The reference's own compiled C (oracle/_ref)
# --------------------------------------------------------------------------------------------
_ref_cache: dict = {}


def ref_available() -> bool:
    return (REF_DIR / "_c_newmap_count_kmers.so").is_file() and \
        (REF_DIR / "_c_newmap_generate_index.so").is_file()


def ref_modules():
    """(count_module, generate_module) = the reference's CPython extensions built by
    oracle/Makefile from /root/reference/src/*.c + AvxWindowFmIndex/src/*.c (unmodified)."""
    if "m" not in _ref_cache:
        mods = []
        for name in ("_c_newmap_count_kmers", "_c_newmap_generate_index"):
            path = REF_DIR / f"{name}.so"
            if not path.is_file():
                raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
            loader = importlib.machinery.ExtensionFileLoader(name, str(path))
            spec = importlib.util.spec_from_loader(name, loader)
            mod = importlib.util.module_from_spec(spec)
            loader.exec_module(mod)
            mods.append(mod)
        _ref_cache["m"] = tuple(mods)
    return _ref_cache["m"]


def ref_build_index(fasta_path, index_path, compression_ratio: int = 8, seed_length: int = 12) -> None:
    """`newmap index` through the reference's own extension (newmap/index.py:7-23)."""
    ref_modules()[1].generate_fm_index(str(fasta_path), str(index_path), compression_ratio, seed_length)


# --------------------------------------------------------------------------------------------
# Segmentation (newmap/fasta.py:20-190) in closed form
# --------------------------------------------------------------------------------------------
class Segment(NamedTuple):
    id: bytes
    data: bytes
    epilogue: bool


def _open(path):
    path = Path(path)
    return gzip.open(path, "rb") if path.suffix == ".gz" else open(path, "rb")


def fasta_records(path) -> Iterator[tuple[bytes, bytes]]:
    """(id, sequence) per record: a line starting with '>' or ';' opens a record whose id is
    the first whitespace token minus its first byte (fasta.py:60-75); data lines are
    right-stripped and concatenated (fasta.py:47,98)."""
    cur, parts = b"", []
    with _open(path) as fh:
        for line in fh:
            line = line.rstrip()
            if line.startswith((b">", b";")):
                if parts:
                    yield cur, b"".join(parts)
                cur, parts = line.split()[0][1:], []
            elif line:
                parts.append(line)
    if parts:
        yield cur, b"".join(parts)


def segments_of(seq_id: bytes, seq: bytes, batch: int, lookahead: int) -> Iterator[Segment]:
    """fasta.py:109-190 for one record: windows of batch+lookahead bytes starting at multiples
    of `batch` while a whole window fits; whatever remains (if anything beyond the overlap
    is new) is the epilogue, otherwise the last full window is."""
    total, want = len(seq), batch + lookahead
    if total == 0:
        return
    starts = []
    s = 0
    while s + want <= total:
        starts.append(s)
        s += batch
    consumed = (starts[-1] + want) if starts else 0
    if consumed < total:
        for st in starts:
            yield Segment(seq_id, seq[st:st + want], False)
        tail_start = (starts[-1] + batch) if starts else 0
        yield Segment(seq_id, seq[tail_start:], True)
    else:
        for st in starts[:-1]:
            yield Segment(seq_id, seq[st:st + want], False)
        yield Segment(seq_id, seq[starts[-1]:starts[-1] + want], True)


def segments(path, batch: int, lookahead: int) -> Iterator[Segment]:
    for sid, seq in fasta_records(path):
        yield from segments_of(sid, seq, batch, lookahead)


def out_dtype(max_k: int):
    # search.py:204-212
    return np.uint8 if max_k <= 255 else (np.uint16 if max_k <= 65535 else np.uint32)


def min_unique(fasta_path, index: OracleIndex, kmer_lengths, is_range: bool, kmer_batch_size: int,
               initial_search_length: int = 0, use_rc: bool = True, nthreads: int = 0) -> dict:
    """{sequence id: array} = contents the reference writes to <id>.unique.uintN
    (search.py:197-380, single FASTA / single index, no include/exclude filter)."""
    max_k = max(kmer_lengths)
    dt = out_dtype(max_k)
    out: dict[bytes, list] = {}
    for seg in segments(fasta_path, kmer_batch_size, max_k - 1):
        n = len(seg.data) if seg.epilogue else len(seg.data) - (max_k - 1)  # search.py:727-741
        vals = index.min_unique_segment(seg.data, n, kmer_lengths, is_range, initial_search_length,
                                        use_rc, nthreads)
        out.setdefault(seg.id, []).append(vals.astype(dt))
    return {k: np.concatenate(v) for k, v in out.items()}
