"""tests/hostemu -- TEST INFRASTRUCTURE ONLY: the per-thread kernel bodies of
newmap_b200/csrc/nm_core.cuh compiled for the host and run serially, so that the logic of the
CUDA path is checked against the oracle where there is no GPU.  Never imported by the product."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "libhostemu.so"
SRC = [HERE / "hostemu.cpp", HERE.parent.parent / "newmap_b200" / "csrc" / "nm_core.cuh"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not LIB.is_file() or any(s.stat().st_mtime > LIB.stat().st_mtime for s in SRC):
            subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-Wextra",
                            "-o", str(LIB), str(SRC[0])], check=True)
        L = C.CDLL(str(LIB))
        L.emu_index_open.restype = C.c_void_p
        L.emu_index_open.argtypes = [C.c_char_p]
        L.emu_index_close.argtypes = [C.c_void_p]
        L.emu_index_set_seed_depth.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_count.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int,
                                C.c_void_p, C.c_void_p]
        L.emu_min_unique.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_uint32,
                                     C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.emu_min_unique_multi.argtypes = [C.POINTER(C.c_void_p), C.c_uint32, C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32,
                                           C.c_uint32, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                           C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def min_unique_multi_segment(indexes: list, segs: list, num_kmers, kmer_lengths, is_range, initial=0, use_rc=True,
                             out_width=4):
    """trajectory_multi_thread over one tuple of segment buffers: (values, stats)."""
    bufs = [np.frombuffer(s, dtype=np.uint8) for s in segs]
    ih = (C.c_void_p * len(indexes))(*[ix._h for ix in indexes])
    sp = (C.c_void_p * len(bufs))(*[b.ctypes.data for b in bufs])
    kl = np.array([min(kmer_lengths), max(kmer_lengths)] if is_range else list(kmer_lengths), dtype=np.uint32)
    out = np.zeros(max(num_kmers, 1), dtype={1: np.uint8, 2: np.uint16, 4: np.uint32}[out_width])
    stats = np.zeros(8, dtype=np.int64)
    r = lib().emu_min_unique_multi(ih, len(indexes), sp, len(bufs), bufs[0].size, num_kmers, 0 if is_range else 1,
                                   kl.ctypes.data, kl.size, initial, int(use_rc), out_width, out.ctypes.data,
                                   stats.ctypes.data)
    assert r == 0
    return out[:num_kmers], stats


class EmuIndex:
    def __init__(self, path):
        self._h = lib().emu_index_open(os.fsencode(str(path)))
        if not self._h:
            raise OSError(f"hostemu: cannot load {path}")

    def set_seed_depth(self, depth: int):
        if lib().emu_index_set_seed_depth(self._h, depth):
            raise ValueError(f"hostemu: seed depth {depth} unsupported")
        return self

    def count(self, seq: bytes, idx, length, rc=False):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        out = np.zeros(idx.size, dtype=np.uint32)
        lf = C.c_uint64(0)
        buf = np.frombuffer(seq, dtype=np.uint8)
        r = lib().emu_count(self._h, buf.ctypes.data, buf.size, idx.ctypes.data, length.ctypes.data, idx.size,
                            int(rc), out.ctypes.data, C.byref(lf))
        if r:
            raise IndexError("bad query")
        return out, lf.value

    def min_unique_segment(self, seg: bytes, num_kmers, kmer_lengths, is_range, initial=0, use_rc=True, algo=0,
                           out_width=4, dcap=0, gsize=0, gsize2=0, stall=0, stride=0):
        """stride: 0 = widest the slack allows (default of the library), n = at most n, -1 = dense thread-per-position pass."""
        buf = np.frombuffer(seg, dtype=np.uint8)
        kl = np.array([min(kmer_lengths), max(kmer_lengths)] if is_range else list(kmer_lengths), dtype=np.uint32)
        out = np.zeros(max(num_kmers, 1), dtype={1: np.uint8, 2: np.uint16, 4: np.uint32}[out_width])
        stats = np.zeros(10, dtype=np.int64)
        lib().emu_min_unique(self._h, buf.ctypes.data, buf.size, num_kmers, 0 if is_range else 1, kl.ctypes.data,
                             kl.size, initial, int(use_rc), algo, out_width, out.ctypes.data, stats.ctypes.data, dcap, gsize, gsize2, stall, stride & 0xFFFFFFFF)
        return out[:num_kmers], stats
