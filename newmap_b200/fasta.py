"""FASTA -> overlapping sequence segments, with the reference's semantics (newmap/fasta.py:20-190)
and a different mechanism: the file is read in large binary blocks, newlines are removed with
bytes.translate, and windows are cut from one growing buffer per record.

Semantics kept (they are parity-relevant):
  * a line whose first byte is '>' or ';' opens a record; its id is the first whitespace-delimited
    token without that first byte (fasta.py:60-75);
  * other lines are right-stripped and concatenated (fasta.py:47,98);
  * a record of L bases is cut into windows of `sequence_length` bytes that start at multiples of
    (sequence_length - sequence_overlap_length); the last segment of a record carries
    epilogue=True and is whatever remains after the last whole window -- or that last whole window
    itself when nothing new remains (fasta.py:109-190).
"""
from __future__ import annotations

from typing import BinaryIO, Iterator

FASTA_FILE_IGNORE_DELIMITERS = (b">", b";")
_READ_BLOCK = 32 << 20
_ODD_WHITESPACE = (b" ", b"\t", b"\r", b"\x0b", b"\x0c")


class SequenceSegment:
    __slots__ = ("id", "data", "epilogue")

    def __init__(self, sequence_id: bytes, data: bytes = b"", epilogue: bool = False):
        self.id = sequence_id
        self.data = data
        self.epilogue = epilogue

    def is_empty(self) -> bool:
        return len(self.data) == 0


def _strip_lines(block: bytes) -> bytes:
    """Concatenation of the right-stripped lines of `block` (which holds whole lines only)."""
    if any(ws in block for ws in _ODD_WHITESPACE):
        return b"".join(line.rstrip() for line in block.split(b"\n"))
    return block.translate(None, b"\n")


def _records(fasta_file: BinaryIO) -> Iterator[tuple[bytes, bytes, bool]]:
    """Yields (record id, next piece of that record's bases, record_finished)."""
    current_id = b""
    carry = b""
    started = False          # any line seen for the current record (so an empty record still closes)
    while True:
        block = fasta_file.read(_READ_BLOCK)
        at_eof = not block
        data = carry + block
        if not at_eof:
            cut = data.rfind(b"\n")
            if cut < 0:
                carry = data
                continue
            data, carry = data[:cut + 1], data[cut + 1:]
        else:
            carry = b""
        pos = 0
        n = len(data)
        while pos < n:
            # next header line: at pos itself or after a newline
            if data[pos:pos + 1] in FASTA_FILE_IGNORE_DELIMITERS:
                hdr = pos
            else:
                h1 = data.find(b"\n>", pos)
                h2 = data.find(b"\n;", pos)
                cands = [h for h in (h1, h2) if h >= 0]
                hdr = min(cands) + 1 if cands else -1
            end = hdr if hdr >= 0 else n
            if end > pos:
                piece = _strip_lines(data[pos:end])
                if piece:
                    yield current_id, piece, False
                started = True
            if hdr < 0:
                break
            eol = data.find(b"\n", hdr)
            if eol < 0:
                eol = n
            line = data[hdr:eol].rstrip()
            if started:
                yield current_id, b"", True
            current_id = line.split()[0][1:] if line.split() else b""
            started = True
            pos = eol + 1
        if at_eof:
            if started:
                yield current_id, b"", True
            return


def sequence_segments(fasta_file: BinaryIO, sequence_length: int,
                      sequence_overlap_length: int = 0) -> Iterator[SequenceSegment]:
    step = sequence_length - sequence_overlap_length
    if step <= 0:
        raise ValueError("sequence_length must exceed sequence_overlap_length")
    buf = bytearray()       # bases of the current record from window start `base` onwards
    emitted_any = False     # a whole window of this record has been cut already
    held: bytes | None = None   # last whole window, kept back until we know whether it is the epilogue

    for seq_id, piece, finished in _records(fasta_file):
        if piece:
            buf += piece
            # cut every whole window that is certainly not the last one
            while len(buf) >= sequence_length:
                if held is not None:
                    yield SequenceSegment(seq_id, held)
                held = bytes(buf[:sequence_length])
                emitted_any = True
                del buf[:step]
        if finished:
            # what remains beyond the overlap of the last whole window is new data
            remaining_new = len(buf) - (sequence_overlap_length if emitted_any else 0)
            if remaining_new > 0:
                if held is not None:
                    yield SequenceSegment(seq_id, held)
                yield SequenceSegment(seq_id, bytes(buf), epilogue=True)
            elif held is not None:
                yield SequenceSegment(seq_id, held, epilogue=True)
            buf.clear()
            emitted_any = False
            held = None


def window_plan(record_length: int, sequence_length: int, sequence_overlap_length: int = 0) -> list:
    """[(start, length, epilogue)] of the segments `sequence_segments` cuts from a record of `record_length` bases --
    the same windows in closed form, for the device ingest path where a record is already one contiguous buffer."""
    step = sequence_length - sequence_overlap_length
    if step <= 0:
        raise ValueError("sequence_length must exceed sequence_overlap_length")
    L = record_length
    whole = 0 if L < sequence_length else (L - sequence_length) // step + 1
    left = L - whole * step                                 # bytes in the buffer after the last whole window was cut
    remaining_new = left - (sequence_overlap_length if whole else 0)
    plan = [(w * step, sequence_length, False) for w in range(whole)]
    if remaining_new > 0:
        plan.append((whole * step, left, True))
    elif plan:
        plan[-1] = (plan[-1][0], sequence_length, True)
    return plan
