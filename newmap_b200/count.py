"""Drop-in for the reference's C extension `newmap._c_newmap_count_kmers` (src/newmap-count.c):
same two functions, same argument meaning, same exception types -- backed by the CUDA library
through ctypes, with the index kept resident in HBM between calls instead of being re-read from
disk on every call (newmap-count.c:9-17,136)."""
from __future__ import annotations

import os
import threading

from . import _lib

_cache: dict = {}
_cache_lock = threading.Lock()


def open_index(index_path, device: int | None = None, seed_depth: int | None = None) -> _lib.Index:
    """Process-wide cache of resident indexes keyed by (real path, mtime, size, device, derived seed depth asked for)."""
    if device is None:
        device = int(os.environ.get("NEWMAP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    path = os.path.realpath(os.fspath(index_path))
    try:
        st = os.stat(path)
    except OSError:
        raise OSError("Could not load reference index from file\n")     # newmap-count.c:13-16
    key = (path, st.st_mtime_ns, st.st_size, device, seed_depth)
    with _cache_lock:
        ix = _cache.get(key)
        if ix is None or ix._h is None:
            ix = _lib.Index(path, device, seed_depth)
            _cache[key] = ix
        return ix


def close_all() -> None:
    from . import search
    search.release_cached_slots()          # workspaces must go before their index
    with _cache_lock:
        for ix in _cache.values():
            ix.close()
        _cache.clear()


def _check_threads(num_threads) -> None:
    """The reference parses num_threads with the 'b' format (newmap-count.c:36,102): an integer (bool included) in
    0..255, else TypeError / OverflowError.  Accepted for interface compatibility and otherwise ignored on the GPU."""
    if not isinstance(num_threads, int):
        raise TypeError(f"'{type(num_threads).__name__}' object cannot be interpreted as an integer")
    if num_threads < 0:
        raise OverflowError("unsigned byte integer is less than minimum")
    if num_threads > 255:
        raise OverflowError("unsigned byte integer is greater than maximum")


def count_kmers(index_path: str, kmers: list, num_threads: int = 1) -> list:
    """newmap-count.c:28-89: occurrences of each byte-string k-mer in the index."""
    _check_threads(num_threads)
    if not isinstance(kmers, list):
        raise TypeError("Second argument must be a list of kmer byte strings")
    for k in kmers:
        if not isinstance(k, bytes):
            raise TypeError("All elements of the kmer list must be byte strings")
        if len(k) == 0:
            raise ValueError("All elements of the kmer list must have non-zero length")
    ix = open_index(index_path)
    lens = [len(k) for k in kmers]
    offs, acc = [], 0
    for n in lens:
        offs.append(acc)
        acc += n
    return ix.count(b"".join(kmers), offs, lens).tolist()


def count_kmers_from_sequence(index_path: str, sequence, indices: list, lengths: list, num_threads: int = 1) -> list:
    """newmap-count.c:91-206: occurrences of sequence[i:i+l] for each (i, l)."""
    # argument parsing order of "sy#OOb" (newmap-count.c:102): the sequence buffer is looked at before num_threads
    if isinstance(sequence, str):
        raise TypeError("a bytes-like object is required, not 'str'")
    if isinstance(sequence, bytearray):     # 'y#' takes read-only buffers only
        raise TypeError("argument 2 must be read-only bytes-like object, not bytearray")
    _check_threads(num_threads)
    if not isinstance(indices, list):
        raise TypeError("Third argument must be a list of integers")
    if not isinstance(lengths, list):
        raise TypeError("Fourth argument must be a list of integers")
    if len(indices) != len(lengths):
        raise ValueError("Both lists of indices and lengths must be the same length")
    for i, n in zip(indices, lengths):
        if not isinstance(i, int) or not isinstance(n, int):
            raise TypeError("All elements of the the index and length lists must be integers")
        if i < 0:
            raise IndexError("All elements of the the index list must be non-negative integers")
        # newmap-count.c:176-194: a negative length wraps in size_t and fails the bounds check (the ValueError the
        # source names is unreachable: its test looks at the index again)
        if n < 0 or i + n > len(sequence):
            raise IndexError("The sum of the index and length of each k-mer must be less than or equal to the "
                             "length of the input byte sequence")
    ix = open_index(index_path)
    # a zero-length k-mer is not an error here (unlike count_kmers): the reference answers 0 for it
    live = [q for q, n in enumerate(lengths) if n > 0]
    out = [0] * len(lengths)
    if live:
        got = ix.count(sequence, [indices[q] for q in live], [lengths[q] for q in live]).tolist()
        for q, c in zip(live, got):
            out[q] = c
    return out
