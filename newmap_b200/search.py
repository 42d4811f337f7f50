"""`newmap search` host driver on the B200 path.

Same public surface as the reference's newmap/search.py -- SearchConfig (fields, from_args),
write_unique_counts, binary_search, linear_search, get_kmer_counts, get_num_kmers,
get_ambiguous_sequence_mask, update_upper_search_bound -- and the same outputs
(<outdir>/<id>.unique.uint8|16|32, raw little-endian, one value per base), but the per-position
search no longer lives in numpy + a count extension: each segment buffer is handed to
libnewmap_b200.so (nm_workspace_* through ctypes), which computes the minimum unique length of
every position on the GPU.  Segments are pipelined over a few workspaces so that host->device
copies, kernels and device->host copies of consecutive segments overlap.  `num_threads` is
accepted and ignored.  There is no CPU fallback.
"""
from __future__ import annotations

import os
import sys
from collections import deque
from contextlib import ExitStack
from dataclasses import dataclass, field
from functools import partial
from math import ceil, log2
from pathlib import Path
from typing import Callable

import numpy as np

from . import _lib
from .count import open_index
from .fasta import SequenceSegment, sequence_segments, window_plan
from .util import INDEX_EXTENSION, optional_gzip_open, verbose_print

KMER_RANGE_SEPARATOR = ":"
SEQUENCE_ID_SEPARATOR = ","
INDEX_FILE_SEPARATOR = ","
FASTA_FILE_SEPARATOR = ","

COMPLEMENT_TRANSLATE_TABLE = bytes.maketrans(b"ACGTacgt", b"TGCAtgca")
ALLOWED_BASES = b"ACGTactg"          # sic (search.py:23): the set is right, the order is not
UNIQUE_COUNT_FILENAME_FORMAT = "{}.unique.{}"
PIPELINE_DEPTH = int(os.environ.get("NEWMAP_B200_PIPELINE", "3"))
DEVICE_PIPELINE_DEPTH = int(os.environ.get("NEWMAP_B200_DEVICE_PIPELINE", "4"))
WRITER_THREADS = int(os.environ.get("NEWMAP_B200_WRITERS", "8"))   # threads writing <id>.unique.uintN windows (pwrite)
READER_THREADS = int(os.environ.get("NEWMAP_B200_READERS", "0"))   # threads reading the FASTA text into pinned staging (0 = library default)
# "device": the FASTA text is parsed on the GPU (nm_fasta_*) and segments are searched where they lie;
# "host": newmap_b200/fasta.py cuts the segments and every segment is uploaded.  Default: device (with --verbose too:
# the per-sequence statistics of the verbose log are a reduction on the device, NM_OPT_OUT_STATS).
INGEST = os.environ.get("NEWMAP_B200_INGEST", "auto")
# a run sharded by its launcher (torchrun: RANK / WORLD_SIZE / MASTER_ADDR) reads, per rank, only the part of an
# uncompressed FASTA its windows lie in; 0 = every rank reads and parses the whole text
PARTITIONED = os.environ.get("NEWMAP_B200_PARTITIONED", "1") != "0"
INTERLEAVE = int(os.environ.get("NEWMAP_B200_INTERLEAVE", "8"))    # sequences whose windows are dealt round-robin (device path)
# [(phase name, seconds since the start of the run)] of the last device-ingest run of this process (bench.py reads it)
LAST_RUN_TRACE: list = []


def nil_search_log(*_: str):
    pass


@dataclass(frozen=True)
class SearchConfig:
    """Field-for-field the reference's SearchConfig (search.py:32-63); `device` and `algorithm`
    are additions with neutral defaults."""
    fasta_filepaths: list
    fmindex_filepaths: list
    kmer_lengths: list
    is_binary_search: bool
    use_reverse_complement: bool = True
    output_directory: Path = field(default_factory=Path.cwd)
    include_sequence_ids: list = field(default_factory=list)
    exclude_sequence_ids: list = field(default_factory=list)
    num_threads: int = 1                 # accepted, ignored on the GPU
    kmer_batch_size: int = 1000000
    initial_search_length: int = 0
    verbose: bool = False
    device: int | None = None            # CUDA device (default: $NEWMAP_B200_DEVICE, $LOCAL_RANK, 0)
    algorithm: str = "auto"              # "auto" | "trajectory" | "free_generic" (include/newmap_b200.h)
    strict: bool = False                 # NM_OPT_STRICT: prove every k-mer the reference could query present in the index
    rank: int = 0                        # position sharding: this process handles the segments whose running
    world_size: int = 1                  # number (over the whole FASTA) is congruent to rank modulo world_size
    log: Callable[..., None] = field(init=False, repr=False, compare=False)

    def __post_init__(self):
        object.__setattr__(self, "log", partial(verbose_print, True) if self.verbose else nil_search_log)

    @classmethod
    def from_args(cls, args):
        """argparse namespace of `newmap search` -> SearchConfig (search.py:65-184)."""
        output_directory = Path(args.output_directory) if args.output_directory else Path(".")
        output_directory.mkdir(parents=True, exist_ok=True)
        fasta_filenames = [Path(p) for p in args.fasta_file.split(FASTA_FILE_SEPARATOR)]
        if not args.index_file:
            index_filenames = [Path(Path(fasta_filenames[0]).stem + "." + INDEX_EXTENSION)]
        else:
            index_filenames = [Path(p) for p in args.index_file.split(INDEX_FILE_SEPARATOR)]
        for index_filename in index_filenames:
            if not index_filename.is_file():
                raise FileNotFoundError(f"Index file not found: {index_filename}")
        spec = args.search_range
        if KMER_RANGE_SEPARATOR in spec:
            try:
                lo, hi = map(int, spec.split(KMER_RANGE_SEPARATOR))
            except ValueError:
                raise ValueError("Could not parse k-mer search range format")
            if lo > hi:
                raise ValueError("K-mer range start length is larger than the end length")
            kmer_lengths, binary = list(range(lo, hi + 1)), True
        else:
            kmer_lengths, binary = list(map(int, spec.split(","))), False
        if args.initial_search_length and not binary:
            raise ValueError("Initial search length only valid when a range of k-mer lengths is given")
        if args.include_sequences and args.exclude_sequences:
            raise ValueError("Cannot specify both include and exclude sequences")
        include = [s.encode() for s in args.include_sequences.split(SEQUENCE_ID_SEPARATOR)] \
            if args.include_sequences else []
        exclude = [s.encode() for s in args.exclude_sequences.split(SEQUENCE_ID_SEPARATOR)] \
            if args.exclude_sequences else []
        return cls(fasta_filepaths=fasta_filenames, fmindex_filepaths=index_filenames,
                   kmer_lengths=kmer_lengths, is_binary_search=binary,
                   use_reverse_complement=not args.norc, output_directory=output_directory,
                   include_sequence_ids=include, exclude_sequence_ids=exclude, verbose=args.verbose,
                   num_threads=args.num_threads, kmer_batch_size=args.kmer_batch_size,
                   initial_search_length=args.initial_search_length,
                   device=getattr(args, "device", None), algorithm=getattr(args, "algorithm", "auto"),
                   strict=bool(getattr(args, "strict", False)),
                   rank=int(os.environ.get("RANK", "0")), world_size=int(os.environ.get("WORLD_SIZE", "1")))


# ------------------------------------------------------------------------------------------------
# helpers with the reference's names
# ------------------------------------------------------------------------------------------------
def get_num_kmers(sequence_segment: SequenceSegment, max_kmer_length: int) -> int:
    """search.py:727-741."""
    n = len(sequence_segment.data)
    return n if sequence_segment.epilogue else n - (max_kmer_length - 1)


def get_ambiguous_sequence_mask(sequence_segment: SequenceSegment, num_positions: int) -> np.ndarray:
    """True where the start byte is not one of b'ACGTacgt' (search.py:744-766)."""
    buf = np.frombuffer(sequence_segment.data, dtype=np.uint8, count=num_positions)
    lut = np.ones(256, dtype=bool)
    lut[list(ALLOWED_BASES)] = False
    return lut[buf]


def update_upper_search_bound(upper_search_bound: np.ndarray, ambiguity_mask: np.ndarray,
                              max_search_length: int, sequence_buffer_length: int) -> None:
    """In-place upper k bound per position (search.py:769-871), in closed form: distance to the
    next ambiguous start position, else to the end of the buffer, capped at max_search_length;
    entries at ambiguous positions are left untouched.  (The GPU path computes the same quantity
    in nm_core.cuh length_limit(); this host twin exists for API parity and tests.)"""
    n = ambiguity_mask.size
    assert sequence_buffer_length - n < max_search_length, \
        "Excess sequence buffer length is greater than the maximum search length"
    if n == 0:
        return
    pos = np.arange(n, dtype=np.int64)
    nxt = np.where(ambiguity_mask, pos, np.int64(sequence_buffer_length))
    nxt = np.minimum.accumulate(nxt[::-1])[::-1]          # next ambiguous position at or after i
    room = np.minimum(nxt - pos, max_search_length)
    keep = ~ambiguity_mask
    upper_search_bound[keep] = room[keep].astype(upper_search_bound.dtype)


def _raise_zero_count(segment_data: bytes, st: _lib.Stats):
    """search.py:699-722."""
    p, n = int(st.zero_pos), int(st.zero_len)
    kmer = bytes(segment_data[p:p + n]).decode("utf-8")
    raise RuntimeError(f"The following generated k-mer was not found in the index:\n{kmer}\n"
                       f"Possibly a mismatch between the sequence and the index.")


def _is_multi(config: SearchConfig) -> bool:
    return len(config.fasta_filepaths) > 1 or len(config.fmindex_filepaths) > 1


def get_kmer_counts(config: SearchConfig, sequences: list, sequence_indices: list, kmer_lengths: list) -> np.ndarray:
    """u32 counts of sequences[*][i:i+k] summed over every index, every sequence and both strands
    (search.py:647-724) on the GPU."""
    total = np.zeros(len(sequence_indices), dtype=np.uint32)
    for fmindex_filepath in config.fmindex_filepaths:
        ix = open_index(fmindex_filepath, config.device)
        for i, seq in enumerate(sequences):
            if len(sequences) > 1:
                # search.py:664-668 names the i-th INDEX file for the i-th sequence: IndexError with fewer indexes
                config.log(f"Counting k-mers from {config.fmindex_filepaths[i]} in index {fmindex_filepath}")
            total += ix.count(seq, sequence_indices, kmer_lengths)
            if config.use_reverse_complement:
                total += ix.count(seq, sequence_indices, kmer_lengths, reverse_complement=True)
    zero = np.nonzero(total == 0)[0]
    if zero.size:
        i, k = sequence_indices[int(zero[0])], kmer_lengths[int(zero[0])]
        raise RuntimeError(f"The following generated k-mer was not found in the index:\n"
                           f"{bytes(sequences[0][i:i + k]).decode('utf-8')}\n"
                           f"Possibly a mismatch between the sequence and the index.")
    return total


class _MultiSearcher:
    """Several FASTA files x several index files (search.py:260-265,461,489-517,656-697): one
    workspace, every index resident on the same GPU without derived tables (the summed count is searched on the
    reference's own query trajectory, nm_workspace_multi)."""

    def __init__(self, config: SearchConfig, max_seg_len: int):
        self.config = config
        self.indexes = [open_index(p, config.device, seed_depth=0) for p in config.fmindex_filepaths]
        self.ws = _lib.Workspace(self.indexes[0], max_seg_len)
        self.max_seg_len = max_seg_len

    def close(self):
        self.ws.close()

    def search(self, sequence_segments_: tuple, num_kmers: int, data_type) -> np.ndarray:
        config = self.config
        first = sequence_segments_[0]
        n = len(first.data)
        datas = [_lib.as_u8(s.data) for s in sequence_segments_]
        for d in datas[1:]:
            if d.size != n:
                raise ValueError("the sequences of the FASTA files given together must be equal in length "
                                 "(search.py:391-392)")
        if len(sequence_segments_) > len(self.indexes) and num_kmers and \
                not get_ambiguous_sequence_mask(first, num_kmers).all():
            # search.py:664-668 (see get_kmer_counts): the reference stops here with this error
            raise IndexError("list index out of range")
        if n > self.max_seg_len:
            self.ws.close()
            self.ws = _lib.Workspace(self.indexes[0], n)
            self.max_seg_len = n
        out = np.zeros(max(num_kmers, 1), dtype=data_type)
        self.ws.upload(datas[0].ctypes.data, n)
        self.ws.multi(self.indexes, [d.ctypes.data for d in datas[1:]])
        self.ws.search(num_kmers, config.kmer_lengths, config.is_binary_search, config.initial_search_length,
                       config.use_reverse_complement, "trajectory", np.dtype(data_type).itemsize)
        self.ws.download(out.ctypes.data)
        st = self.ws.sync()
        if st.zero_pos >= 0:
            _raise_zero_count(first.data, st)
        if st.short_counts:
            raise RuntimeError(f"{st.short_counts} position(s) had a total count below the number of sequences "
                               f"({len(sequence_segments_)}): the sequences do not all occur in the indexes given "
                               "(the reference does not terminate on such input)")
        return out[:num_kmers]


def _search_segment(config: SearchConfig, sequence_segments_: tuple, data_type) -> tuple:
    seg = sequence_segments_[0]
    max_k = max(config.kmer_lengths)
    num_kmers = get_num_kmers(seg, max_k)
    if len(sequence_segments_) > 1 or len(config.fmindex_filepaths) > 1:
        ms = _MultiSearcher(config, len(seg.data))
        try:
            out = ms.search(tuple(sequence_segments_), num_kmers, data_type)
        finally:
            ms.close()
    else:
        ix = open_index(config.fmindex_filepaths[0], config.device)
        out, st = ix.min_unique(seg.data, num_kmers, config.kmer_lengths, config.is_binary_search,
                                config.initial_search_length, config.use_reverse_complement, config.algorithm,
                                dtype=data_type)
        if st.zero_pos >= 0:
            _raise_zero_count(seg.data, st)
    ambiguous = int(get_ambiguous_sequence_mask(seg, num_kmers).sum())
    return out, ambiguous


def binary_search(config: SearchConfig, sequence_segments: tuple, min_kmer_length: int, max_kmer_length: int,
                  data_type) -> tuple:
    """(unique_lengths, ambiguous_positions_skipped) for one segment, range mode (search.py:383-548)."""
    assert min_kmer_length == min(config.kmer_lengths) and max_kmer_length == max(config.kmer_lengths)
    return _search_segment(config, sequence_segments, data_type)


def linear_search(config: SearchConfig, sequence_segments: tuple, num_kmers: int, data_type) -> tuple:
    """(unique_lengths, ambiguous_positions_skipped) for one segment, list mode (search.py:551-644)."""
    return _search_segment(config, sequence_segments, data_type)


# ------------------------------------------------------------------------------------------------
# the pipelined driver
# ------------------------------------------------------------------------------------------------
class _Slot:
    """One in-flight segment: pinned staging for input and output + a workspace (stream)."""

    def __init__(self, ix: _lib.Index, max_seg_len: int, out_itemsize: int):
        self.ws = _lib.Workspace(ix, max_seg_len)
        self.inp = _lib.PinnedBuffer(max_seg_len)
        self.out = _lib.PinnedBuffer(max_seg_len * out_itemsize)
        self.job = None

    def close(self):
        self.ws.close()
        self.inp.close()
        self.out.close()


def write_unique_counts(config: SearchConfig):
    """FASTA -> <id>.unique.uintN, one value per base (search.py:197-380)."""
    max_kmer_length = max(config.kmer_lengths)
    min_kmer_length = min(config.kmer_lengths)
    data_type, unique_count_suffix = _lib.out_dtype(max_kmer_length)
    if config.is_binary_search and max_kmer_length > min_kmer_length:
        config.log("Max {} iterations over range {}-{}".format(
            ceil(log2(max_kmer_length - min_kmer_length) + 1), min_kmer_length, max_kmer_length))

    if _is_multi(config):
        return _write_unique_counts_multi(config, data_type, unique_count_suffix)
    ix = open_index(config.fmindex_filepaths[0], config.device)
    lookahead_length = max_kmer_length - 1
    requested_sequence_length = config.kmer_batch_size + lookahead_length
    if INGEST in ("device", "auto"):
        if _write_unique_counts_device(config, ix, data_type, unique_count_suffix):
            return
    slots: list[_Slot] = []
    inflight: deque = deque()
    totals = dict(ambiguous=0, unique=0, none=0, max_len=min_kmer_length, min_len=max_kmer_length)
    sequence_processed = False
    current_sequence_id = b""
    current_file = None
    sharded = config.world_size > 1
    if not 0 <= config.rank < config.world_size:
        raise ValueError("rank must be in [0, world_size)")
    segment_number = -1          # running number over all processed segments: the sharding key
    file_offset = 0              # positions already emitted into the current file (records that follow one another
                                 # with the same id are appended to one file, search.py:268-357)
    last_window_mine = False     # this rank searched the last window emitted into the current file

    def finish(slot: _Slot):
        seg, num_kmers, fh, offset = slot.job
        st = slot.ws.sync()
        if st.zero_pos >= 0:
            _raise_zero_count(seg.data, st)
        res = slot.out.array[:num_kmers * data_type.itemsize].view(data_type)
        if config.verbose:
            ambiguous = int(get_ambiguous_sequence_mask(seg, num_kmers).sum())
            unique = int(np.count_nonzero(res))
            totals["ambiguous"] += ambiguous
            totals["unique"] += unique
            totals["none"] += num_kmers - unique - ambiguous
            if unique:
                totals["max_len"] = max(totals["max_len"], int(res.max()))
                totals["min_len"] = min(totals["min_len"], int(res[res != 0].min()))
            else:
                config.log("No unique lengths found for this sequence segment")
        # every segment lands at window_number * batch of <id>.unique.uintN (host-side gather: ranks share
        # the files, nothing is exchanged between GPUs); with one rank this is the reference's append
        os.pwrite(fh.fileno(), res.tobytes(), offset * data_type.itemsize)
        slot.job = None

    def drain():
        while inflight:
            finish(inflight.popleft())

    def close_current_file():
        # several ranks share the file and nobody truncated it at open: the rank that searched its last window sets
        # the size (the other ranks' windows all lie below it, whenever they are written)
        if current_file is not None:
            if sharded and last_window_mine:
                os.ftruncate(current_file.fileno(), file_offset * data_type.itemsize)
            current_file.close()

    with ExitStack() as stack:
        fasta_file = stack.enter_context(optional_gzip_open(config.fasta_filepaths[0], "rb"))
        stack.callback(lambda: [s.close() for s in slots])
        for segment in sequence_segments(fasta_file, requested_sequence_length, lookahead_length):
            if current_sequence_id != segment.id:
                # include / exclude filter, evaluated exactly like search.py:268-278 -- including its
                # quirk that with only --exclude-sequences given every sequence is skipped
                if config.include_sequence_ids or config.exclude_sequence_ids:
                    if (segment.id not in config.include_sequence_ids) or (segment.id in config.exclude_sequence_ids):
                        continue
                drain()
                sequence_processed = True
                print_summary_statisitcs(config, current_sequence_id, totals["unique"], totals["ambiguous"],
                                         totals["none"], totals["max_len"], totals["min_len"])
                current_sequence_id = segment.id
                path = config.output_directory / UNIQUE_COUNT_FILENAME_FORMAT.format(
                    segment.id.decode(), unique_count_suffix)
                close_current_file()
                # one rank: truncate like search.py:305.  Several ranks share the file: create without truncating
                # (another rank may already have written its windows); see close_current_file.
                if sharded:
                    fd = os.open(path, os.O_RDWR | os.O_CREAT, 0o644)
                    current_file = os.fdopen(fd, "r+b")
                else:
                    current_file = open(path, "wb")
                file_offset = 0
                config.log("Writing unique lengths for sequence ID: {}".format(segment.id.decode()))
            num_kmers = get_num_kmers(segment, max_kmer_length)
            segment_number += 1
            offset = file_offset
            file_offset += num_kmers
            last_window_mine = segment_number % config.world_size == config.rank
            if not last_window_mine:
                continue
            config.log("Processing {} k-mers".format(num_kmers))
            if len(inflight) >= PIPELINE_DEPTH:
                finish(inflight.popleft())
            slot = next((s for s in slots if s.job is None), None)
            if slot is None:
                slot = _Slot(ix, requested_sequence_length, data_type.itemsize)
                if config.strict:
                    slot.ws.option("strict", 1)
                slots.append(slot)
            n = len(segment.data)
            slot.inp.array[:n] = np.frombuffer(segment.data, dtype=np.uint8)
            slot.job = (segment, num_kmers, current_file, offset)
            slot.ws.upload(slot.inp.ptr, n)
            slot.ws.search(num_kmers, config.kmer_lengths, config.is_binary_search, config.initial_search_length,
                           config.use_reverse_complement, config.algorithm, data_type.itemsize)
            slot.ws.download(slot.out.ptr)
            inflight.append(slot)
        drain()
        close_current_file()
        current_file = None
        print_summary_statisitcs(config, current_sequence_id, totals["unique"], totals["ambiguous"],
                                 totals["none"], totals["max_len"], totals["min_len"])
    if not sequence_processed:
        if config.include_sequence_ids:
            raise ValueError(f"None of the included sequences were found: {config.include_sequence_ids}")
        elif config.exclude_sequence_ids:
            raise ValueError("The excluded sequences were too strict and nothing was processed: "
                             f"{config.exclude_sequence_ids}")


def _write_unique_counts_multi(config: SearchConfig, data_type, unique_count_suffix: str):
    """write_unique_counts for several FASTA and / or index files (search.py:222-357 with zip(*sequences)): the n-th
    segments of every file are searched together; ids, filters and output files follow the first file."""
    max_kmer_length = max(config.kmer_lengths)
    min_kmer_length = min(config.kmer_lengths)
    lookahead_length = max_kmer_length - 1
    requested_sequence_length = config.kmer_batch_size + lookahead_length
    if not 0 <= config.rank < config.world_size:
        raise ValueError("rank must be in [0, world_size)")
    sharded = config.world_size > 1
    totals = dict(ambiguous=0, unique=0, none=0, max_len=min_kmer_length, min_len=max_kmer_length)
    sequence_processed = False
    current_sequence_id = b""
    current_file = None
    segment_number = -1
    segment_in_sequence = 0
    searcher = _MultiSearcher(config, requested_sequence_length)
    with ExitStack() as stack:
        stack.callback(searcher.close)
        fasta_files = [stack.enter_context(optional_gzip_open(f, "rb")) for f in config.fasta_filepaths]
        sequences = [sequence_segments(f, requested_sequence_length, lookahead_length) for f in fasta_files]
        for segment_list in zip(*sequences):
            segment = segment_list[0]
            if current_sequence_id != segment.id:
                if config.include_sequence_ids or config.exclude_sequence_ids:
                    if (segment.id not in config.include_sequence_ids) or (segment.id in config.exclude_sequence_ids):
                        continue
                sequence_processed = True
                print_summary_statisitcs(config, current_sequence_id, totals["unique"], totals["ambiguous"],
                                         totals["none"], totals["max_len"], totals["min_len"])
                current_sequence_id = segment.id
                path = config.output_directory / UNIQUE_COUNT_FILENAME_FORMAT.format(
                    segment.id.decode(), unique_count_suffix)
                if current_file is not None:
                    current_file.close()
                if sharded:
                    fd = os.open(path, os.O_RDWR | os.O_CREAT, 0o644)
                    current_file = stack.enter_context(os.fdopen(fd, "r+b"))
                else:
                    current_file = stack.enter_context(open(path, "wb"))
                segment_in_sequence = 0
                config.log("Writing unique lengths for sequence ID: {}".format(segment.id.decode()))
            num_kmers = get_num_kmers(segment, max_kmer_length)
            segment_number += 1
            offset = segment_in_sequence * config.kmer_batch_size
            segment_in_sequence += 1
            if segment_number % config.world_size != config.rank:
                continue
            config.log("Processing {} k-mers".format(num_kmers))
            res = searcher.search(segment_list, num_kmers, data_type)
            if config.verbose:
                ambiguous = int(get_ambiguous_sequence_mask(segment, num_kmers).sum())
                unique = int(np.count_nonzero(res))
                totals["ambiguous"] += ambiguous
                totals["unique"] += unique
                totals["none"] += num_kmers - unique - ambiguous
                if unique:
                    totals["max_len"] = max(totals["max_len"], int(res.max()))
                    totals["min_len"] = min(totals["min_len"], int(res[res != 0].min()))
                else:
                    config.log("No unique lengths found for this sequence segment")
            os.pwrite(current_file.fileno(), res.tobytes(), offset * data_type.itemsize)
            if segment.epilogue:
                os.ftruncate(current_file.fileno(), (offset + num_kmers) * data_type.itemsize)
        print_summary_statisitcs(config, current_sequence_id, totals["unique"], totals["ambiguous"],
                                 totals["none"], totals["max_len"], totals["min_len"])
    if not sequence_processed:
        if config.include_sequence_ids:
            raise ValueError(f"None of the included sequences were found: {config.include_sequence_ids}")
        elif config.exclude_sequence_ids:
            raise ValueError("The excluded sequences were too strict and nothing was processed: "
                             f"{config.exclude_sequence_ids}")


def _record_id(header: bytes) -> bytes:
    tokens = header.split()
    return tokens[0][1:] if tokens else b""       # fasta.py:60-75


def _plan_groups(config: SearchConfig, records, requested_sequence_length: int, lookahead_length: int) -> list:
    """[id, [(record number, start in record, length, num_kmers, file offset, running number)], file size] per output
    file, from [(id, record length)]: consecutive records with the same id share one file and are APPENDED to it
    (search.py:268-357 opens a file anew only when the id changes); a window lands where the windows before it in that
    file end.  The include / exclude filter is evaluated exactly like search.py:268-278, quirk included."""
    groups: list = []
    previous_id = None
    segment_number = -1
    for rec_no, (seq_id, rec_len) in enumerate(records):
        if rec_len == 0:
            continue
        if previous_id != seq_id:
            if config.include_sequence_ids or config.exclude_sequence_ids:
                if (seq_id not in config.include_sequence_ids) or (seq_id in config.exclude_sequence_ids):
                    continue
            previous_id = seq_id
            groups.append([seq_id, [], 0])
        g = groups[-1]
        for start, length, epilogue in window_plan(rec_len, requested_sequence_length, lookahead_length):
            num_kmers = length if epilogue else length - lookahead_length
            segment_number += 1
            g[1].append((rec_no, start, length, num_kmers, g[2], segment_number))
            g[2] += num_kmers
    return groups


def _line_start_at_or_after(fd: int, pos: int, size: int):
    """Smallest p >= pos that is the start of a line of the file (0, the file size, or one past a newline); None when
    there is no newline within 64 MiB (a FASTA with whole sequences on one line: not partitioned)."""
    if pos <= 0:
        return 0
    if pos >= size:
        return size
    at = pos - 1
    while at < min(size, pos + (64 << 20)):
        chunk = os.pread(fd, 1 << 20, at)
        if not chunk:
            break
        k = chunk.find(b"\n")
        if k >= 0:
            return at + k + 1
        at += len(chunk)
    return size if at >= size else None


def _all_gather(obj, world_size: int):
    """One object per rank, in rank order, over torch.distributed (the process group of the caller, else a gloo group
    from the launcher's MASTER_ADDR / MASTER_PORT): the only exchange of a sharded run -- a few numbers per rank."""
    import torch.distributed as dist
    if not dist.is_initialized():
        dist.init_process_group("gloo")
    out = [None] * world_size
    dist.all_gather_object(out, obj)
    return out


def _partition_table(fragments: list, rank: int, local_offsets: list, tail_bases: int):
    """The global record table of a partitioned run and this rank's view of it.
    fragments[r] = [(id or None, bases inside r's byte range)] in file order -- None: the piece goes on with the record
    the previous rank left open; local_offsets = where this rank's pieces begin in its own sequence buffer;
    tail_bases = bases of its last piece's record that it read beyond its range.
    Returns (table [[id, length, pieces [rank, first base in the record, bases, local offset]]], locate)."""
    table: list = []
    for r, fr in enumerate(fragments):
        for header, own in fr:
            if header is None and table:
                table[-1][1] += own
                table[-1][2].append([r, table[-1][1] - own, own, None])
            else:
                table.append([header if header is not None else b"", own, [[r, 0, own, None]]])
    mine = [piece for rec in table for piece in rec[2] if piece[0] == rank]
    for piece, off in zip(mine, local_offsets):
        piece[3] = off
    last_piece = mine[-1] if mine else None

    def locate(rec_no: int, start: int, length: int):
        """(owner of the window, its local sequence offset or None, fits): owner = the rank whose range holds the
        window's first base; the bases it needs beyond its range must lie in the stretch it read after it."""
        for piece in table[rec_no][2]:
            r, first_base, own, off = piece
            if first_base <= start < first_base + own:
                if r != rank:
                    return r, None, True
                beyond = start + length - (first_base + own)
                fits = beyond <= 0 or (piece is last_piece and beyond <= tail_bases)
                return r, off + (start - first_base), fits
        return -1, None, False          # (a record without bases has no window)

    return table, locate


def _partitioned_ingest(config: SearchConfig, ix: _lib.Index, path: Path, requested_sequence_length: int,
                        lookahead_length: int, mark):
    """Sharded run, uncompressed FASTA: every rank reads ONLY ITS byte range of the text (+ the stretch its last window
    reaches into), parses it on its GPU, and the ranks exchange what they found -- header lines and base counts -- so
    that each can lay out the same global window plan and take the windows that start in its range.
    Returns (fasta, records [(id, length)], locate(record number, start) -> (owner rank, local sequence offset)) or
    None when the text is not fit for it (every rank then reads the whole file: _write_unique_counts_device)."""
    rank, world = config.rank, config.world_size
    size = path.stat().st_size
    fd = os.open(path, os.O_RDONLY)
    try:
        lo = _line_start_at_or_after(fd, size * rank // world, size)
        hi = _line_start_at_or_after(fd, size * (rank + 1) // world, size)
        ok = lo is not None and hi is not None
        fa = None
        frags, tail_first, tail_all = [], 0, 0
        if ok:
            # enough text after the range for the longest window that can start inside it (lines of >= 8 bases)
            tail_len = min(size - hi, requested_sequence_length + requested_sequence_length // 4 + 65536)
            fa = _lib.Fasta(ix.device, size_hint=hi - lo + tail_len)
            fa.feed_file(path, READER_THREADS, offset=lo, length=hi - lo + tail_len)
            mark("text_on_device")
            try:
                fa.finish()
            except RuntimeError:
                ok = False
            mark("parsed")
            ok = ok and not fa.info.odd_whitespace
        if ok:
            # the stretch after the range, looked at on the host: how many of the parsed bases and header lines are its
            tail = np.frombuffer(os.pread(fd, tail_len, hi), dtype=np.uint8) if tail_len else np.zeros(0, dtype=np.uint8)
            newline = tail == 10
            line_start = np.empty(tail.size, dtype=bool)
            if tail.size:
                line_start[0] = True
                line_start[1:] = newline[:-1]
            headers = np.nonzero(line_start & ((tail == ord(">")) | (tail == ord(";"))))[0]
            first = int(headers[0]) if headers.size else tail.size
            tail_first = first - int(newline[:first].sum())          # bases of the tail before its first header line
            recs = fa.records[:len(fa.records) - headers.size]        # the records that begin (or go on) inside the range
            for n, (header, off, length) in enumerate(recs):
                own = length - (tail_first if n == len(recs) - 1 else 0)
                # a headerless first piece goes on with the record the previous rank left open (rank 0: id b"")
                frags.append((None if (not header and n == 0 and rank > 0) else _record_id(header), off, own))
            if not recs:
                tail_first = 0
            tail_all = tail_first
    finally:
        os.close(fd)
    everyone = _all_gather((ok, [(h, own) for h, _, own in frags]), world)
    mark("exchanged")
    if not all(o for o, _ in everyone):
        if fa is not None:
            fa.close()
        return None
    table, locate = _partition_table([fr for _, fr in everyone], rank, [off for _, off, _ in frags], tail_all)
    return fa, [(rec[0], rec[1]) for rec in table], locate


def _write_unique_counts_device(config: SearchConfig, ix: _lib.Index, data_type, unique_count_suffix: str) -> bool:
    """write_unique_counts with the FASTA parsed on the GPU: the text is streamed to the device while it is read
    (reader threads, pinned staging), records become contiguous runs of bases there, segments are the windows
    fasta.window_plan gives and are searched in place; results come back through pinned buffers and are written by a
    small pool of writer threads.  A run sharded over several ranks by its launcher (RANK / WORLD_SIZE / MASTER_ADDR in
    the environment) reads only its part of the text (_partitioned_ingest).  Returns False (nothing written) when the
    text needs the host reader."""
    from concurrent.futures import ThreadPoolExecutor
    max_kmer_length = max(config.kmer_lengths)
    lookahead_length = max_kmer_length - 1
    requested_sequence_length = config.kmer_batch_size + lookahead_length
    path = Path(config.fasta_filepaths[0])
    import time
    trace = os.environ.get("NEWMAP_B200_TRACE")
    t_start = time.perf_counter()
    marks = []

    def mark(name):
        marks.append((name, time.perf_counter() - t_start))
    sharded = config.world_size > 1
    if not 0 <= config.rank < config.world_size:
        raise ValueError("rank must be in [0, world_size)")
    itemsize = data_type.itemsize
    slots: list = []
    fa = None
    # workspaces and pinned output buffers are allocated while the file is being read
    maker = ThreadPoolExecutor(max_workers=1)
    made = maker.submit(_take_slots, ix, requested_sequence_length, itemsize,
                        DEVICE_PIPELINE_DEPTH + max(2, WRITER_THREADS - 1))
    try:
        partitioned = None
        if (sharded and PARTITIONED and path.suffix != ".gz" and os.environ.get("MASTER_ADDR")
                and os.environ.get("WORLD_SIZE") == str(config.world_size) and os.environ.get("RANK") == str(config.rank)):
            partitioned = _partitioned_ingest(config, ix, path, requested_sequence_length, lookahead_length, mark)
        if partitioned is not None:
            fa, records, locate = partitioned
            groups = _plan_groups(config, records, requested_sequence_length, lookahead_length)
            placed = [[locate(rec_no, start, length) for rec_no, start, length, _, _, _ in wins] for _, wins, _ in groups]
            fits = all(f for row in placed for _, _, f in row)
            if not all(_all_gather(fits, config.world_size)):
                raise RuntimeError("partitioned ingest: a window reaches beyond the text a rank read after its range "
                                   "(lines shorter than 8 bases?); run with NEWMAP_B200_PARTITIONED=0")
            # [id, [(local sequence offset, length, num_kmers, file offset, mine)], file size]
            plan = [[sid, [(off, length, nk, foff, owner == config.rank)
                           for (_, _, length, nk, foff, _), (owner, off, _) in zip(wins, row)], total]
                    for (sid, wins, total), row in zip(groups, placed)]
            mark("planned")
        else:
            fa = _lib.Fasta(ix.device, size_hint=path.stat().st_size if path.suffix != ".gz" else 0)
            mark("fasta_open")
            if path.suffix == ".gz":
                with optional_gzip_open(path, "rb") as fh:
                    while True:
                        block = fh.read(64 << 20)
                        if not block:
                            break
                        fa.feed(block)
            else:
                fa.feed_file(path, READER_THREADS)
            mark("text_on_device")
            try:
                fa.finish()
                mark("parsed")
            except RuntimeError:
                return False            # e.g. more header lines than the device parser handles
            if fa.info.odd_whitespace:
                return False            # right-stripping beyond newlines: the host reader implements it
            offsets = [off for _, off, _ in fa.records]
            groups = _plan_groups(config, [(_record_id(h), n) for h, _, n in fa.records], requested_sequence_length,
                                  lookahead_length)
            plan = [[sid, [(offsets[rec_no] + start, length, nk, foff, number % config.world_size == config.rank)
                           for rec_no, start, length, nk, foff, number in wins], total] for sid, wins, total in groups]
        slots = made.result()
        mark("slots_ready")
        inflight: deque = deque()
        files: list = []
        current_sequence_id = None
        # --verbose: the log lines come in the reference's order (a segment is finished before the next one is
        # announced) and the statistics of search.py:318-347 come from a reduction on the device
        depth = 1 if config.verbose else DEVICE_PIPELINE_DEPTH
        min_kmer_length = min(config.kmer_lengths)
        totals = dict(ambiguous=0, unique=0, none=0, max_len=min_kmer_length, min_len=max_kmer_length)
        for s_ in slots:                       # (slots may come from an earlier run of this process: set both ways)
            s_.ws.option("out_stats", 1 if config.verbose else 0)
            s_.ws.option("strict", 1 if config.strict else 0)

        def write_out(slot, fd, nbytes, offset, truncate_to):
            os.pwrite(fd, memoryview(slot.out.array)[:nbytes], offset)
            if truncate_to is not None:
                os.ftruncate(fd, truncate_to)

        def finish(slot):
            (seq_off, seg_len, num_kmers, fd, offset, last_of_file) = slot.job
            st = slot.ws.sync()
            if st.zero_pos >= 0:
                _raise_zero_count(fa.read(seq_off, seg_len), st)
            if config.verbose:
                unique, ambiguous = int(st.unique_positions), int(st.ambiguous_positions)
                totals["ambiguous"] += ambiguous
                totals["unique"] += unique
                totals["none"] += num_kmers - unique - ambiguous
                if unique:
                    totals["max_len"] = max(totals["max_len"], int(st.max_unique))
                    totals["min_len"] = min(totals["min_len"], int(st.min_unique))
                else:
                    config.log("No unique lengths found for this sequence segment")
            slot.pending = pool.submit(write_out, slot, fd, num_kmers * itemsize, offset * itemsize,
                                       (offset + num_kmers) * itemsize if last_of_file else None)
            slot.job = None

        sequence_processed = bool(plan)
        seen_paths: set = set()

        with ThreadPoolExecutor(max_workers=max(1, WRITER_THREADS)) as pool:
            try:
                def settle_writes():
                    while inflight:
                        finish(inflight.popleft())
                    for s_ in slots:
                        if s_.pending is not None:
                            s_.pending.result()
                            s_.pending = None

                def open_group(seq_id):
                    nonlocal current_sequence_id
                    if config.verbose:
                        while inflight:
                            finish(inflight.popleft())
                        print_summary_statisitcs(config, current_sequence_id or b"", totals["unique"],
                                                 totals["ambiguous"], totals["none"], totals["max_len"],
                                                 totals["min_len"])
                    current_sequence_id = seq_id
                    out_path = config.output_directory / UNIQUE_COUNT_FILENAME_FORMAT.format(
                        seq_id.decode(), unique_count_suffix)
                    if out_path in seen_paths:
                        # the id came back after another one: the reference truncates the file and starts again
                        if sharded:
                            raise ValueError(f"sequence id {seq_id.decode()!r} appears again after another sequence: "
                                             "not supported when the run is sharded over several ranks")
                        settle_writes()
                    seen_paths.add(out_path)
                    if len(files) >= 256:           # many sequences: no more than a few hundred files open at a time
                        settle_writes()
                        for fd_ in files:
                            os.close(fd_)
                        files.clear()
                    # one rank: truncate like search.py:305.  Several ranks share the file: create without truncating
                    # (another rank may already have written its windows); the owner of the last window sets the size.
                    flags = os.O_RDWR | os.O_CREAT | (0 if sharded else os.O_TRUNC)
                    fd_ = os.open(out_path, flags, 0o644)
                    files.append(fd_)
                    config.log("Writing unique lengths for sequence ID: {}".format(seq_id.decode()))
                    return fd_

                def submit(fd_, window, last_of_file):
                    seq_off, length, num_kmers, offset, _ = window
                    config.log("Processing {} k-mers".format(num_kmers))
                    if len(inflight) >= depth:
                        finish(inflight.popleft())
                    slot = next((s for s in slots if s.job is None and (s.pending is None or s.pending.done())), None)
                    if slot is None:
                        slot = next(s for s in slots if s.job is None)
                    if slot.pending is not None:
                        slot.pending.result()
                        slot.pending = None
                    slot.job = (seq_off, length, num_kmers, fd_, offset, last_of_file)
                    slot.ws.attach(fa, seq_off, length)
                    slot.ws.search(num_kmers, config.kmer_lengths, config.is_binary_search,
                                   config.initial_search_length, config.use_reverse_complement, config.algorithm,
                                   itemsize)
                    slot.ws.download(slot.out.ptr)
                    inflight.append(slot)

                # Windows of the same sequence land in the same file, and buffered writes to one file serialise in the
                # kernel (the inode lock): the writer threads only work side by side when the windows in flight belong
                # to different files.  So the windows of up to INTERLEAVE sequences are dealt round-robin -- every
                # window knows its file offset, the order is free -- unless the log must follow the reference's order
                # (--verbose) or an id comes back after another one (the file is then truncated in between).
                ids = [seq_id for seq_id, _, _ in plan]
                group = 1 if (config.verbose or len(set(ids)) != len(ids)) else max(1, INTERLEAVE)
                for g0 in range(0, len(plan), group):
                    batch = plan[g0:g0 + group]
                    fds = [open_group(seq_id) for seq_id, _, _ in batch]
                    for n_win in range(max(len(wins) for _, wins, _ in batch)):
                        for fd_, (_, wins, _) in zip(fds, batch):
                            if n_win < len(wins) and wins[n_win][4]:
                                submit(fd_, wins[n_win], n_win == len(wins) - 1)
                mark("all_segments_launched")
                while inflight:
                    finish(inflight.popleft())
                mark("all_segments_searched")
                if config.verbose:
                    print_summary_statisitcs(config, current_sequence_id or b"", totals["unique"], totals["ambiguous"],
                                             totals["none"], totals["max_len"], totals["min_len"])
                for s_ in slots:
                    if s_.pending is not None:
                        s_.pending.result()
                mark("all_written")
            finally:
                pool.shutdown(wait=True)
                for fd in files:
                    os.close(fd)
        if not sequence_processed:
            if config.include_sequence_ids:
                raise ValueError(f"None of the included sequences were found: {config.include_sequence_ids}")
            elif config.exclude_sequence_ids:
                raise ValueError("The excluded sequences were too strict and nothing was processed: "
                                 f"{config.exclude_sequence_ids}")
        return True
    finally:
        try:
            slots = slots or made.result()
        except Exception:
            pass
        maker.shutdown(wait=True)
        _return_slots(ix, requested_sequence_length, itemsize, slots)
        if fa is not None:
            fa.close()
        mark("closed")
        LAST_RUN_TRACE[:] = marks
        if trace:
            print("[newmap_b200 ingest] " + ", ".join(f"{n} {t:.3f}s" for n, t in marks), file=sys.stderr, flush=True)


# Workspaces + pinned output buffers of the device path, kept between the runs of one process (a resident index
# stays resident; so does its scratch): 11 slots are 110 cudaMalloc calls and 110 MB of page-locked memory, which
# costs more than searching a chromosome -- and page-locking serialises between the ranks of a node.
# NEWMAP_B200_SLOT_CACHE=0 frees them after every run.
SLOT_CACHE = os.environ.get("NEWMAP_B200_SLOT_CACHE", "1") != "0"
_slot_cache: dict = {}


def _take_slots(ix: _lib.Index, max_seg_len: int, itemsize: int, count: int) -> list:
    slots = _slot_cache.pop((id(ix), ix._h.value if ix._h else 0, max_seg_len, itemsize), []) if SLOT_CACHE else []
    for s_ in slots:
        s_.job = s_.pending = None
    while len(slots) < count:
        slots.append(_DeviceSlot(ix, max_seg_len, itemsize))
    return slots


def _return_slots(ix: _lib.Index, max_seg_len: int, itemsize: int, slots: list) -> None:
    if SLOT_CACHE and ix._h:
        for old in _slot_cache.values():            # one set at a time: another index or shape replaces it
            for s_ in old:
                s_.close()
        _slot_cache.clear()
        _slot_cache[(id(ix), ix._h.value, max_seg_len, itemsize)] = slots
    else:
        for s_ in slots:
            s_.close()


def release_cached_slots() -> None:
    for old in _slot_cache.values():
        for s_ in old:
            s_.close()
    _slot_cache.clear()


import atexit  # noqa: E402
atexit.register(release_cached_slots)


class _DeviceSlot:
    """One in-flight segment of the device ingest path: a workspace (stream) + pinned staging for its output."""

    def __init__(self, ix: _lib.Index, max_seg_len: int, out_itemsize: int):
        self.ws = _lib.Workspace(ix, max_seg_len)
        self.out = _lib.PinnedBuffer(max_seg_len * out_itemsize)
        self.job = None
        self.pending = None

    def close(self):
        self.ws.close()
        self.out.close()


def print_summary_statisitcs(config, sequence_id, total_unique_lengths_count, total_ambiguous_positions,
                             total_no_unique_lengths_count, max_length_found, min_length_found):
    """search.py:885-902 (name kept, typo included)."""
    if total_unique_lengths_count:
        config.log(f"Finished writing unique lengths for sequence ID: {sequence_id.decode()}")
        config.log(f"{total_unique_lengths_count} unique lengths found")
        config.log(f"{total_ambiguous_positions} positions skipped due to ambiguity")
        config.log(f"{total_no_unique_lengths_count} positions with no unique length found")
        config.log(f"{max_length_found}-mer maximum unique length found")
        config.log(f"{min_length_found}-mer minimum unique length found")


def main(args):
    config = SearchConfig.from_args(args)
    write_unique_counts(config)
