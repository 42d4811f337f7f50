"""ctypes binding of libnewmap_b200.so (include/newmap_b200.h).

The library is hand-written CUDA for sm_100a; there is no CPU fallback.  Importing this module
never touches the GPU; opening an index does, and raises if no CUDA device is usable."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
# (NEWMAP_B200_LIB: another build of the same library, for tuning experiments)
LIB_PATH = Path(os.environ.get("NEWMAP_B200_LIB") or PKG_DIR / "libnewmap_b200.so")

NM_MODE_RANGE, NM_MODE_LIST = 0, 1
NM_ALGO_AUTO, NM_ALGO_TRAJECTORY, NM_ALGO_FREE_GENERIC = 0, 1, 2
ALGOS = {"auto": NM_ALGO_AUTO, "trajectory": NM_ALGO_TRAJECTORY, "free_generic": NM_ALGO_FREE_GENERIC}
NM_EOPEN, NM_EFORMAT, NM_EVERSION, NM_EALLOC, NM_ESHORT, NM_EARG, NM_ECUDA, NM_EUNSUPPORTED, NM_ELIMIT = range(-1, -10, -1)
NM_NOALLOC_SCRATCH_BYTES = 4 << 20


class IndexInfo(C.Structure):
    _fields_ = [("bwt_length", C.c_uint64), ("n_blocks", C.c_uint64), ("prefix_sums", C.c_uint64 * 6),
                ("device_bytes", C.c_uint64), ("seed_length", C.c_uint32), ("sa_compression_ratio", C.c_uint32),
                ("has_rank32", C.c_uint32), ("device", C.c_int32), ("derived_seed_length", C.c_uint32),
                ("reserved", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("positions", C.c_uint64), ("lf_steps", C.c_uint64), ("queries", C.c_uint64),
                ("generic_searches", C.c_uint64), ("ref_lf_steps", C.c_uint64), ("zero_pos", C.c_int64),
                ("zero_len", C.c_int64), ("zero_round", C.c_int64), ("algo_used", C.c_uint32),
                ("replayed", C.c_uint32), ("ms_h2d", C.c_float), ("ms_prep", C.c_float), ("ms_search", C.c_float),
                ("ms_generic", C.c_float), ("ms_resolve", C.c_float), ("ms_d2h", C.c_float), ("ms_total", C.c_float),
                ("ms_pass1", C.c_float), ("ms_cert", C.c_float), ("ms_pass3", C.c_float), ("launches", C.c_uint32),
                ("seed_depth", C.c_uint32), ("short_counts", C.c_uint32), ("lf_steps_pass1", C.c_uint64),
                ("seed_lookups", C.c_uint64), ("seed_lookups_pass1", C.c_uint64), ("stride", C.c_uint32),
                ("out_stats", C.c_uint32), ("unique_positions", C.c_uint64), ("ambiguous_positions", C.c_uint64),
                ("min_unique", C.c_uint32), ("max_unique", C.c_uint32)]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


class FastaInfo(C.Structure):
    _fields_ = [("raw_bytes", C.c_uint64), ("sequence_bytes", C.c_uint64), ("records", C.c_uint64),
                ("odd_whitespace", C.c_uint32), ("reserved", C.c_uint32), ("ms_parse", C.c_float),
                ("reserved2", C.c_float)]


class FastaRecord(C.Structure):
    _fields_ = [("seq_offset", C.c_uint64), ("seq_length", C.c_uint64), ("header_offset", C.c_uint64),
                ("header_length", C.c_uint64)]


# every symbol include/newmap_b200.h declares: (restype, argtypes)
_vp, _u64, _u32, _i = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
SIGNATURES = {
    "nm_abi_version": (_i, []),
    "nm_strerror": (C.c_char_p, [_i]),
    "nm_last_cuda_error": (C.c_char_p, []),
    "nm_device_count": (_i, []),
    "nm_index_open": (_i, [C.c_char_p, _i, C.POINTER(_vp)]),
    "nm_index_close": (None, [_vp]),
    "nm_index_info_get": (_i, [_vp, C.POINTER(IndexInfo)]),
    "nm_index_set_seed_depth": (_i, [_vp, _u32]),
    "nm_count_from_sequence": (_i, [_vp, _vp, _u64, _vp, _vp, _u64, _i, _vp, C.POINTER(_u64)]),
    "nm_count_from_sequence_noalloc": (_i, [_vp, _vp, _u64, _vp, _vp, _u64, _i, _vp, C.POINTER(_u64)]),
    "nm_min_unique": (_i, [_vp, _vp, _u64, _u64, _i, _vp, _u32, _u32, _i, _i, _i, _vp, C.POINTER(Stats)]),
    "nm_workspace_create": (_i, [_vp, _u64, _vp, C.POINTER(_vp)]),
    "nm_workspace_destroy": (None, [_vp]),
    "nm_workspace_upload": (_i, [_vp, _vp, _u64]),
    "nm_workspace_search": (_i, [_vp, _u64, _i, _vp, _u32, _u32, _i, _i, _i]),
    "nm_workspace_download": (_i, [_vp, _vp]),
    "nm_workspace_sync": (_i, [_vp, C.POINTER(Stats)]),
    "nm_workspace_option": (_i, [_vp, _i, _u64]),
    "nm_workspace_checksum": (_i, [_vp, C.POINTER(_u64)]),
    "nm_fasta_open": (_i, [_i, _u64, C.POINTER(_vp)]),
    "nm_fasta_close": (None, [_vp]),
    "nm_fasta_feed": (_i, [_vp, _vp, _u64]),
    "nm_fasta_feed_file": (_i, [_vp, C.c_char_p, _u32]),
    "nm_fasta_feed_file_range": (_i, [_vp, C.c_char_p, _u64, _u64, _u32]),
    "nm_fasta_finish": (_i, [_vp, C.POINTER(FastaInfo)]),
    "nm_fasta_records": (_i, [_vp, C.POINTER(FastaRecord), _u64]),
    "nm_fasta_headers": (_i, [_vp, _vp, _u64, C.POINTER(_u64)]),
    "nm_fasta_read": (_i, [_vp, _u64, _u64, _vp]),
    "nm_workspace_multi": (_i, [_vp, C.POINTER(_vp), C.c_uint32, C.POINTER(_vp), C.c_uint32]),
    "nm_workspace_attach": (_i, [_vp, _vp, _u64, _u64]),
    "nm_track_open": (_i, [_i, C.POINTER(_vp)]),
    "nm_track_close": (None, [_vp]),
    "nm_track_compute": (_i, [_vp, _vp, _u64, _i, _u32]),
    "nm_track_counts": (_i, [_vp, _u64, _u64, _vp]),
    "nm_track_runs": (_i, [_vp, _vp, _vp, _u64, C.POINTER(_u64)]),
    "nm_track_bed_text": (_i, [_vp, C.c_char_p, _u64, _u64, _vp, _u64, C.POINTER(_u64)]),
    "nm_track_text": (_i, [_vp, C.c_char_p, _vp, _u64, _u64, _vp, _u64, C.POINTER(_u64)]),
    "nm_host_alloc": (_vp, [_u64]),
    "nm_host_free": (None, [_vp]),
    "nm_index_build": (_i, [_vp, _u64, _vp, _u64, _vp, _u64, _u32, _u32, _i, C.c_char_p, C.POINTER(C.c_double)]),
    "nm_build_last_error": (C.c_char_p, []),
    "nm_probe_random_access": (_i, [_i, _u64, _u32, _u64, _u32, _u32, _i, C.POINTER(C.c_double), C.POINTER(C.c_float)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library; fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.is_file():
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                              f"g.build()'` or `make -C newmap_b200/csrc` (there is no CPU fallback)")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        if L.nm_abi_version() != 3:
            raise ImportError("libnewmap_b200.so ABI version mismatch")
        _lib = L
    return _lib


class NewmapError(RuntimeError):
    def __init__(self, code: int, what: str):
        L = load()
        msg = L.nm_strerror(code).decode()
        if code == NM_ECUDA:
            msg += f" ({L.nm_last_cuda_error().decode()})"
        super().__init__(f"{what}: {msg}")
        self.code = code


def check(code: int, what: str) -> int:
    if code < 0:
        if code in (NM_EOPEN, NM_EFORMAT, NM_EVERSION, NM_ESHORT):
            # newmap-count.c:13-16: index load failures surface as OSError
            raise OSError(f"Could not load reference index from file ({load().nm_strerror(code).decode()})")
        if code == NM_EALLOC:
            raise MemoryError(what)
        raise NewmapError(code, what)
    return code


def as_u8(buf) -> np.ndarray:
    """Zero-copy uint8 view of a bytes-like object (TypeError for str, like the 'y#' format of
    newmap-count.c:102)."""
    if isinstance(buf, str):
        raise TypeError("a bytes-like object is required, not 'str'")
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    if a.dtype != np.uint8 or not a.flags.c_contiguous:
        a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


def out_dtype(max_k: int):
    # search.py:204-212
    if max_k <= 0xFF:
        return np.dtype(np.uint8), "uint8"
    if max_k <= 0xFFFF:
        return np.dtype(np.uint16), "uint16"
    return np.dtype(np.uint32), "uint32"


class Index:
    """A reference-built .awfmi resident in the HBM of one GPU (opened once, not once per call
    like newmap-count.c:9-17,136)."""

    def __init__(self, path, device: int | None = None, seed_depth: int | None = None):
        """seed_depth: depth of the derived seed table (None = chosen by the library, 0 = none)."""
        L = load()
        if device is None:
            device = int(os.environ.get("NEWMAP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        h = _vp()
        old = os.environ.get("NEWMAP_B200_SEED_DEPTH")
        if seed_depth is not None:
            os.environ["NEWMAP_B200_SEED_DEPTH"] = str(int(seed_depth))
        try:
            check(L.nm_index_open(os.fsencode(str(path)), device, C.byref(h)), f"nm_index_open({path})")
        finally:
            if seed_depth is not None:
                if old is None:
                    os.environ.pop("NEWMAP_B200_SEED_DEPTH", None)
                else:
                    os.environ["NEWMAP_B200_SEED_DEPTH"] = old
        self._h = h
        self.path = str(path)
        self.device = device
        info = IndexInfo()
        check(L.nm_index_info_get(self._h, C.byref(info)), "nm_index_info_get")
        self.info = info

    def close(self):
        if getattr(self, "_h", None):
            load().nm_index_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def seed_length(self) -> int:
        return int(self.info.seed_length)

    @property
    def derived_seed_length(self) -> int:
        """Depth of the deeper seed table derived on the device at load time (0 = none)."""
        return int(self.info.derived_seed_length)

    def set_seed_depth(self, depth: int) -> None:
        """Rebuild the derived seed table at `depth` letters (<= seed_length drops it).  Speed only."""
        L = load()
        check(L.nm_index_set_seed_depth(self._h, int(depth)), "nm_index_set_seed_depth")
        check(L.nm_index_info_get(self._h, C.byref(self.info)), "nm_index_info_get")

    @property
    def bwt_length(self) -> int:
        return int(self.info.bwt_length)

    def count(self, seq, idx, length, reverse_complement: bool = False, return_steps: bool = False,
              no_alloc: bool = False):
        """u32 occurrence counts of seq[idx[q]:idx[q]+length[q]] (nm_count_from_sequence; no_alloc: the variant that
        calls no cudaMalloc and refuses requests beyond its fixed scratch with ValueError)."""
        s = as_u8(seq)
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        if idx.shape != length.shape:
            raise ValueError("Both lists of indices and lengths must be the same length")
        out = np.zeros(idx.size, dtype=np.uint32)
        steps = _u64(0)
        fn = load().nm_count_from_sequence_noalloc if no_alloc else load().nm_count_from_sequence
        rc = fn(self._h, s.ctypes.data, s.size, idx.ctypes.data, length.ctypes.data,
                idx.size, int(reverse_complement), out.ctypes.data, C.byref(steps))
        if rc == NM_ELIMIT:
            raise ValueError(load().nm_strerror(rc).decode())
        if rc == NM_EARG:
            if np.any(length == 0):
                raise ValueError("All elements of the kmer list must have non-zero length")
            raise IndexError("The sum of the index and length of each k-mer must be less than or equal to the "
                             "length of the input byte sequence")
        check(rc, "nm_count_from_sequence")
        return (out, steps.value) if return_steps else out

    def min_unique(self, seg, num_kmers: int, kmer_lengths, is_range: bool, initial_search_length: int = 0,
                   use_rc: bool = True, algo: str = "auto", dtype=None):
        """(array of num_kmers minimum unique lengths, Stats) for one segment buffer."""
        s = as_u8(seg)
        kl = _klist(kmer_lengths, is_range)
        dt = np.dtype(dtype) if dtype is not None else out_dtype(int(max(kmer_lengths)))[0]
        out = np.zeros(max(int(num_kmers), 1), dtype=dt)
        st = Stats()
        rc = load().nm_min_unique(self._h, s.ctypes.data, s.size, int(num_kmers),
                                  NM_MODE_RANGE if is_range else NM_MODE_LIST, kl.ctypes.data, kl.size,
                                  int(initial_search_length), int(use_rc), ALGOS[algo], dt.itemsize,
                                  out.ctypes.data, C.byref(st))
        if rc == NM_EARG:
            raise ValueError("invalid search arguments (k-mer lengths must be positive, min <= max, "
                             "initial search length inside the range, num_kmers <= segment length)")
        check(rc, "nm_min_unique")
        return out[:num_kmers], st


def _klist(kmer_lengths, is_range: bool) -> np.ndarray:
    if is_range:
        return np.array([min(kmer_lengths), max(kmer_lengths)], dtype=np.uint32)
    return np.ascontiguousarray(list(kmer_lengths), dtype=np.uint32)


class Workspace:
    """One CUDA stream + device scratch for segments up to max_seg_len (nm_workspace_*)."""

    def __init__(self, index: Index, max_seg_len: int, cuda_stream: int | None = None):
        self.index = index
        h = _vp()
        check(load().nm_workspace_create(index._h, int(max_seg_len), _vp(cuda_stream) if cuda_stream else None,
                                         C.byref(h)), "nm_workspace_create")
        self._h = h
        self._keep = None

    def close(self):
        if getattr(self, "_h", None):
            if getattr(self.index, "_h", None):      # (an index closed first took the device context of its workspaces'
                load().nm_workspace_destroy(self._h)   #  owner with it: the handle is dropped, not dereferenced)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def option(self, name: str, value: int):
        """'ref_lf' | 'dcap' | 'gsize' | 'pass1' | 'gsize2' | 'own_chains' | 'stall' | 'cert_runs' | 'stride' | 'out_stats' | 'subdepth' | 'strict'
        (NM_OPT_*): instrumentation / tuning, never changes results."""
        check(load().nm_workspace_option(self._h, {"ref_lf": 1, "dcap": 2, "gsize": 3, "pass1": 4, "gsize2": 5, "own_chains": 6,
                                                   "stall": 7, "cert_runs": 8, "stride": 9, "out_stats": 10, "subdepth": 11, "strict": 12}[name], int(value)),
              "nm_workspace_option")

    def attach(self, fasta: "Fasta", seq_offset: int, seg_len: int):
        """The next search reads its segment from the parsed FASTA on the device (no host->device copy)."""
        rc = load().nm_workspace_attach(self._h, fasta._h, int(seq_offset), int(seg_len))
        if rc == NM_EARG:
            raise ValueError("segment outside the parsed sequence, longer than the workspace, or another device")
        check(rc, "nm_workspace_attach")

    def upload(self, seg_ptr: int, seg_len: int):
        check(load().nm_workspace_upload(self._h, seg_ptr, int(seg_len)), "nm_workspace_upload")

    def multi(self, indexes: list, segment_ptrs: list):
        """Several sequences x several indexes (search.py:461,656-697): the next search sums the counts over `indexes`
        and over the uploaded segment plus the host buffers `segment_ptrs` (same length, one per further FASTA)."""
        ih = (_vp * len(indexes))(*[ix._h for ix in indexes])
        sp = (_vp * max(len(segment_ptrs), 1))(*[_vp(p) for p in segment_ptrs])
        rc = load().nm_workspace_multi(self._h, ih, len(indexes), sp, len(segment_ptrs))
        if rc == NM_EARG:
            raise ValueError("between 1 and 8 indexes on the workspace's device and at most 7 further sequences")
        check(rc, "nm_workspace_multi")

    def search(self, num_kmers: int, kmer_lengths, is_range: bool, initial_search_length: int = 0,
               use_rc: bool = True, algo: str = "auto", out_width: int = 1):
        kl = _klist(kmer_lengths, is_range)
        self._keep = kl
        rc = load().nm_workspace_search(self._h, int(num_kmers), NM_MODE_RANGE if is_range else NM_MODE_LIST,
                                        kl.ctypes.data, kl.size, int(initial_search_length), int(use_rc),
                                        ALGOS[algo], int(out_width))
        if rc == NM_EARG:
            raise ValueError("invalid search arguments")
        check(rc, "nm_workspace_search")

    def download(self, out_ptr: int):
        check(load().nm_workspace_download(self._h, out_ptr), "nm_workspace_download")

    def sync(self) -> Stats:
        st = Stats()
        check(load().nm_workspace_sync(self._h, C.byref(st)), "nm_workspace_sync")
        return st

    def checksum(self) -> int:
        v = _u64(0)
        check(load().nm_workspace_checksum(self._h, C.byref(v)), "nm_workspace_checksum")
        return v.value


class Fasta:
    """The text of a FASTA file parsed on the GPU (nm_fasta_*): records = [(header line bytes, offset, length)] into
    one device buffer of bases that workspaces read in place (Workspace.attach)."""

    def __init__(self, device: int | None = None, size_hint: int = 0):
        if device is None:
            device = int(os.environ.get("NEWMAP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        h = _vp()
        check(load().nm_fasta_open(int(device), int(size_hint), C.byref(h)), "nm_fasta_open")
        self._h = h
        self.device = device
        self.info = None
        self.records: list = []

    def feed(self, data) -> None:
        a = as_u8(data)
        check(load().nm_fasta_feed(self._h, a.ctypes.data, a.size), "nm_fasta_feed")

    def feed_file(self, path, threads: int = 0, offset: int = 0, length: int | None = None) -> None:
        """The file's text, or its bytes [offset, offset + length), read by `threads` readers into pinned staging."""
        if offset == 0 and length is None:
            check(load().nm_fasta_feed_file(self._h, os.fsencode(str(path)), int(threads)), f"nm_fasta_feed_file({path})")
        else:
            check(load().nm_fasta_feed_file_range(self._h, os.fsencode(str(path)), int(offset),
                                                  int(length) if length is not None else 0xFFFFFFFFFFFFFFFF, int(threads)),
                  f"nm_fasta_feed_file_range({path})")

    def finish(self) -> "Fasta":
        L = load()
        info = FastaInfo()
        check(L.nm_fasta_finish(self._h, C.byref(info)), "nm_fasta_finish")
        self.info = info
        n = int(info.records)
        recs = (FastaRecord * max(n, 1))()
        check(L.nm_fasta_records(self._h, recs, n), "nm_fasta_records")
        need = _u64(0)
        check(L.nm_fasta_headers(self._h, None, 0, C.byref(need)), "nm_fasta_headers")
        text = C.create_string_buffer(max(int(need.value), 1))
        check(L.nm_fasta_headers(self._h, text, need.value, None), "nm_fasta_headers")
        raw = text.raw[:need.value]
        self.records = [(raw[r.header_offset:r.header_offset + r.header_length], int(r.seq_offset), int(r.seq_length))
                        for r in recs[:n]]
        return self

    def read(self, offset: int, n: int) -> bytes:
        out = np.empty(max(int(n), 1), dtype=np.uint8)
        check(load().nm_fasta_read(self._h, int(offset), int(n), out.ctypes.data), "nm_fasta_read")
        return out[:n].tobytes()

    def close(self):
        if getattr(self, "_h", None):
            load().nm_fasta_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PinnedBuffer:
    """Page-locked host staging buffer exposed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        self.ptr = load().nm_host_alloc(int(nbytes))
        if not self.ptr:
            raise MemoryError("nm_host_alloc")
        self.nbytes = int(nbytes)
        self.array = np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr))

    def close(self):
        if getattr(self, "ptr", None):
            self.array = None
            load().nm_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def checksum_host(arr: np.ndarray) -> int:
    """numpy twin of the device checksum (nm_core.cuh checksum_term), for device-resident runs."""
    i = np.arange(1, arr.size + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = (i * np.uint64(0x9E3779B97F4A7C15)) ^ (arr.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F))
        h ^= h >> np.uint64(29)
        h *= np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(32)
        return int(h.sum(dtype=np.uint64))


def probe_random_access(device: int, nbytes: int, granule: int, chains: int, steps: int, mlp: int = 1,
                        dependent: bool = True):
    rate, ms = C.c_double(0), C.c_float(0)
    check(load().nm_probe_random_access(device, int(nbytes), granule, int(chains), steps, mlp, int(dependent), C.byref(rate), C.byref(ms)),
          "nm_probe_random_access")
    return rate.value, ms.value
