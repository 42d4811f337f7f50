"""`newmap track` on the B200 path: single- and multi-read mappability tracks from <id>.unique.uintN.

Same public surface as the reference's newmap/track.py (function names, arguments, files written, stdout handling);
the arithmetic runs on the GPU through libnewmap_b200.so (nm_track_*, newmap_b200/csrc/nm_track.cu):

  * multi-read mappability of position i = (number of j in (i-k, i] with 0 < unique[j] <= k) / k
    (track.py:22-62: +1/-1 interval marks, cumulative sum, one division at the end);
  * single-read mappability = that number > 0, written as BED runs (track.py:65-97);
  * the WIG text (track.py:100-121) prints one of only k+1 different values per line: the k+1 strings are formatted
    here, once, with the reference's own format expression, and the device assembles the text from them.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from pathlib import Path
from typing import BinaryIO

import numpy as np

from . import _lib
from .util import verbose_print

DEFAULT_MAPPABILITY_READ_LENGTH = 24
STDOUT_FILENAME = "-"
CHROMOSOME_FILENAME_DELIMITER = ".unique"
BED_FILE_LINE_FORMAT = "{}\t{}\t{}\tk{}\t{}\t.\n"
WIG_FIXED_STEP_DECLARATION_FORMAT = "fixedStep chrom={} start={} step=1 span=1\n"
MULTIREAD_MAPPABILITY_TYPE = np.float64
_TEXT_CHUNK = 32 << 20      # positions per device text chunk
_RUN_CHUNK = 8 << 20        # BED lines per device text chunk


class Track:
    """Window counts of one sequence resident on the GPU (nm_track_*)."""

    def __init__(self, device: int | None = None):
        if device is None:
            device = int(os.environ.get("NEWMAP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        h = C.c_void_p()
        _lib.check(_lib.load().nm_track_open(int(device), C.byref(h)), "nm_track_open")
        self._h = h
        self.n = 0
        self.k = 0

    def compute(self, unique: np.ndarray, kmer_length: int) -> "Track":
        if unique.dtype not in (np.uint8, np.uint16, np.uint32):
            raise ValueError(f"unsupported unique length type {unique.dtype}")
        if kmer_length <= 0:
            raise ValueError("read length must be positive")
        a = np.ascontiguousarray(unique)
        _lib.check(_lib.load().nm_track_compute(self._h, a.ctypes.data, a.size, a.dtype.itemsize, int(kmer_length)),
                   "nm_track_compute")
        self.n, self.k = int(a.size), int(kmer_length)
        return self

    def counts(self, first: int = 0, m: int | None = None) -> np.ndarray:
        m = self.n - first if m is None else m
        out = np.empty(max(m, 1), dtype=np.uint32)
        _lib.check(_lib.load().nm_track_counts(self._h, int(first), int(m), out.ctypes.data), "nm_track_counts")
        return out[:m]

    def runs(self):
        """(starts uint64, values uint8) of the maximal runs of equal single-read mappability."""
        L = _lib.load()
        n = C.c_uint64(0)
        _lib.check(L.nm_track_runs(self._h, None, None, 0, C.byref(n)), "nm_track_runs")
        starts = np.empty(max(n.value, 1), dtype=np.uint64)
        values = np.empty(max(n.value, 1), dtype=np.uint8)
        _lib.check(L.nm_track_runs(self._h, starts.ctypes.data, values.ctypes.data, n.value, C.byref(n)), "nm_track_runs")
        return starts[:n.value], values[:n.value]

    def count_runs(self) -> int:
        """Number of maximal runs of equal single-read mappability; their starts stay on the device for bed_text."""
        n = C.c_uint64(0)
        _lib.check(_lib.load().nm_track_runs(self._h, None, None, 0, C.byref(n)), "nm_track_runs")
        return int(n.value)

    def bed_text(self, chr_name: str, first_run: int, m: int) -> np.ndarray:
        """The BED lines of runs [first_run, first_run + m) as a uint8 array, assembled on the device."""
        name = chr_name.encode()
        cap = m * (len(name) + 64)
        out = np.empty(max(cap, 1), dtype=np.uint8)
        w = C.c_uint64(0)
        _lib.check(_lib.load().nm_track_bed_text(self._h, name, int(first_run), int(m), out.ctypes.data, cap, C.byref(w)),
                   "nm_track_bed_text")
        return out[:w.value]

    def text(self, strings: list, first: int, m: int) -> np.ndarray:
        """b''.join(strings[count[i]] + b'\\n' for i in range(first, first + m)) as a uint8 array, assembled on the device."""
        blob = b"".join(strings)
        off = np.zeros(len(strings) + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(s) for s in strings])
        cap = m * (max(len(s) for s in strings) + 1)
        out = np.empty(max(cap, 1), dtype=np.uint8)
        w = C.c_uint64(0)
        _lib.check(_lib.load().nm_track_text(self._h, blob, off.ctypes.data, int(first), int(m), out.ctypes.data, cap,
                                             C.byref(w)), "nm_track_text")
        return out[:w.value]

    def close(self):
        if getattr(self, "_h", None):
            _lib.load().nm_track_close(self._h)
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


def float_format(value: float, decimal_places: int) -> bytes:
    """track.py:114-121."""
    if value == 0.0:
        return b"0.0"
    return f"{value:.{decimal_places}f}".encode()


def create_multiread_mappability_from_unique_file(unique_lengths_filename: Path, kmer_length: int, data_type):
    """float64 multi-read mappability of every position (track.py:22-62), window counts from the GPU."""
    unique = np.fromfile(str(unique_lengths_filename), dtype=data_type)
    with Track() as tr:
        counts = tr.compute(unique, kmer_length).counts()
    out = counts.astype(MULTIREAD_MAPPABILITY_TYPE)
    out /= kmer_length
    return out


def write_single_read_bed(bed_file: BinaryIO, kmer_length: int, multi_read_mappability, chr_name: str):
    """track.py:65-97 on a float64 array (kept for callers of the reference's function)."""
    single = np.where(multi_read_mappability > 0.0, 1, 0)
    change = np.concatenate(([True], single[1:] != single[:-1])) if single.size else np.zeros(0, dtype=bool)
    starts = np.nonzero(change)[0]
    lengths = np.diff(np.append(starts, single.size))
    _write_bed_runs(bed_file, kmer_length, chr_name, starts, starts + lengths, single[starts])


def _write_bed_runs(bed_file: BinaryIO, kmer_length: int, chr_name: str, starts, ends, values):
    bed_file.write(b"".join(f"{chr_name}\t{int(s)}\t{int(e)}\tk{kmer_length}\t{int(v)}\t.\n".encode()
                            for s, e, v in zip(starts, ends, values)))


def write_multi_read_wig(wig_file: BinaryIO, multi_read_mappability, chr_name: str, decimal_places: int = 2):
    """track.py:100-111 on a float64 array (kept for callers of the reference's function)."""
    wig_file.write(WIG_FIXED_STEP_DECLARATION_FORMAT.format(chr_name, 1).encode())
    for chunk in np.nditer(multi_read_mappability, flags=["external_loop", "buffered"]):
        wig_file.write(b"\n".join(float_format(v, decimal_places) for v in chunk) + b"\n")


def safe_remove(filename: str):
    if filename and filename != STDOUT_FILENAME and Path(filename).exists():
        Path(filename).unlink()


def _data_type_of(unique_count_filename: Path):
    suffix = unique_count_filename.suffix
    if suffix == ".uint8":
        return np.uint8
    if suffix == ".uint16":
        return np.uint16
    if suffix == ".uint32":
        return np.uint32
    raise ValueError(f"Unknown extension on unique length file: \"{suffix}\"")


def write_mappability_files(unique_count_filenames: list, kmer_length: int, single_read_bed_filename: str,
                            multi_read_wig_filename: str, verbose: bool):
    """track.py:131-236: same files, same order, same bytes; counts, run boundaries and the WIG text come from the GPU."""
    if single_read_bed_filename == STDOUT_FILENAME and multi_read_wig_filename == STDOUT_FILENAME:
        raise ValueError("Cannot output both single-read and multi-read files to standard output")
    elif not single_read_bed_filename and not multi_read_wig_filename:
        raise ValueError("Must specify at least one output file")
    safe_remove(single_read_bed_filename)
    safe_remove(multi_read_wig_filename)
    decimal_places = int(np.ceil(np.log10(kmer_length)))
    strings = [float_format(c / kmer_length, decimal_places) for c in range(kmer_length + 1)]
    with Track() as tr:
        for unique_count_filename in unique_count_filenames:
            unique_count_filename = Path(unique_count_filename)
            file_basename = unique_count_filename.name
            chr_name = file_basename[:file_basename.find(CHROMOSOME_FILENAME_DELIMITER)]
            data_type = _data_type_of(unique_count_filename)
            verbose_print(verbose, f"Calculating mappability regions from minimum unique k-mer lengths in file: "
                                   f"{unique_count_filename}")
            unique = np.fromfile(str(unique_count_filename), dtype=data_type)
            tr.compute(unique, kmer_length)
            verbose_print(verbose, "Chromosome size:")
            verbose_print(verbose, "{}\t{}".format(chr_name, unique.shape[0]))
            if single_read_bed_filename:
                shown = "standard output" if single_read_bed_filename == STDOUT_FILENAME else single_read_bed_filename
                verbose_print(verbose, f"Appending single-read mappability regions to {shown}")
                nruns = tr.count_runs()
                fh = sys.stdout.buffer if single_read_bed_filename == STDOUT_FILENAME else open(single_read_bed_filename, "ab")
                try:
                    for first in range(0, nruns, _RUN_CHUNK):
                        fh.write(tr.bed_text(chr_name, first, min(_RUN_CHUNK, nruns - first)).data)
                finally:
                    if fh is not sys.stdout.buffer:
                        fh.close()
                    else:
                        fh.flush()
            if multi_read_wig_filename:
                shown = "standard output" if multi_read_wig_filename == STDOUT_FILENAME else multi_read_wig_filename
                verbose_print(verbose, f"Appending multi-read mappability regions to {shown}")
                fh = sys.stdout.buffer if multi_read_wig_filename == STDOUT_FILENAME else open(multi_read_wig_filename, "ab")
                try:
                    fh.write(WIG_FIXED_STEP_DECLARATION_FORMAT.format(chr_name, 1).encode())
                    for first in range(0, tr.n, _TEXT_CHUNK):
                        fh.write(tr.text(strings, first, min(_TEXT_CHUNK, tr.n - first)).data)
                finally:
                    if fh is not sys.stdout.buffer:
                        fh.close()
                    else:
                        fh.flush()


def check_unique_file_existence(filename: Path):
    if not filename.exists():
        raise FileNotFoundError(f"Unique count file does not exist: {filename}")


def main(args):
    """track.py:245-284."""
    unique_count_filenames = [Path(filename) for filename in args.unique_count_files]
    kmer_length = args.read_length
    single_read_bed_filename = args.single_read
    multi_read_wig_filename = args.multi_read
    for filename in unique_count_filenames:
        check_unique_file_existence(filename)
    if not kmer_length.isdigit():
        additional = Path(kmer_length)
        check_unique_file_existence(additional)
        unique_count_filenames.insert(0, additional)
        kmer_length = DEFAULT_MAPPABILITY_READ_LENGTH
    else:
        kmer_length = int(kmer_length)
    if not single_read_bed_filename and not multi_read_wig_filename:
        single_read_bed_filename = STDOUT_FILENAME
    write_mappability_files(unique_count_filenames, kmer_length, single_read_bed_filename, multi_read_wig_filename,
                            args.verbose)
