"""`newmap index | search | track` command line, flag-for-flag the reference's (newmap/main.py:29-251), on the
B200 path.  `python -m newmap_b200 search genome.fa genome.awfmi -r 20:200 -o out/`."""
from __future__ import annotations

import sys
from argparse import ArgumentParser

from . import __version__, index, search, track
from .util import INDEX_EXTENSION

DEFAULT_KMER_BATCH_SIZE = 10000000
DEFAULT_THREAD_COUNT = 1
DEFAULT_KMER_SEARCH_RANGE = "20:200"
FASTA_FILE_METAVAR = "fasta_file"


def build_parser() -> ArgumentParser:
    parser = ArgumentParser(prog="newmap", description="Newmap (B200): shortest unique sequence length at each "
                                                       "position of a reference sequence")
    parser.add_argument("--version", action="version", version=f"%(prog)s-b200 {__version__}")
    sub = parser.add_subparsers(metavar="", required=True)
    b = sub.add_parser("index", help="Create an FM index from sequences (on the GPU, reference file format)")
    b.set_defaults(func=index.main)
    b.add_argument("fasta_file", help="Reference sequence file in FASTA format")
    b.add_argument("--output", "-i", metavar="FILE", default=None,
                   help=f"index file to write (default: basename of fasta_file with the {INDEX_EXTENSION} extension)")
    b.add_argument("--compression-ratio", "-c", type=int, default=index.DEFAULT_COMPRESSION_RATIO)
    b.add_argument("--seed-length", "-s", type=int, default=index.DEFAULT_SEED_LENGTH)
    p = sub.add_parser("search", help="Find the shortest unique sequence length at each position in sequences")
    p.set_defaults(func=search.main)
    p.add_argument("fasta_file", metavar=FASTA_FILE_METAVAR, help="(gzipped) fasta file for kmer generation")
    p.add_argument("index_file", nargs="?",
                   help=f"reference index file to count occurances in (default: basename of {FASTA_FILE_METAVAR} "
                        f"with the {INDEX_EXTENSION} extension)")
    g = p.add_argument_group("output arguments")
    g.add_argument("--search-range", "-r", metavar="RANGE", default=DEFAULT_KMER_SEARCH_RANGE,
                   help="comma separated list of increasing lengths, or an inclusive range a:b "
                        f"(default: {DEFAULT_KMER_SEARCH_RANGE})")
    g.add_argument("--output-directory", "-o", metavar="DIR", default=".")
    g.add_argument("--include-sequences", "-i", metavar="IDS")
    g.add_argument("--exclude-sequences", "-x", metavar="IDS")
    g.add_argument("--norc", action="store_true", help="do not search the reverse-complement strand")
    g.add_argument("--verbose", "-v", action="store_true")
    q = p.add_argument_group("performance arguments")
    q.add_argument("--initial-search-length", "-l", type=int, metavar="LENGTH", default=0)
    q.add_argument("--kmer-batch-size", "-s", default=DEFAULT_KMER_BATCH_SIZE, metavar="SIZE", type=int)
    q.add_argument("--num-threads", "-t", default=DEFAULT_THREAD_COUNT, metavar="NUM", type=int,
                   help="accepted for compatibility; ignored on the GPU")
    q.add_argument("--device", type=int, default=None, help="CUDA device (default: $NEWMAP_B200_DEVICE or 0)")
    q.add_argument("--algorithm", choices=["auto", "trajectory", "free_generic"], default="auto",
                   help="auto: trajectory-free kernels (same bytes as the reference on a FASTA that matches the index); "
                        "trajectory: replay the reference's binary / linear search on the GPU, query for query")
    q.add_argument("--strict", action="store_true",
                   help="with --algorithm auto: also prove that every k-mer the reference could query occurs in the index "
                        "(about one more LF step per position).  The reference stops with 'k-mer was not found in the "
                        "index' when the FASTA does not match the index; without --strict the default algorithm reports "
                        "that only when one of its own searches meets the absent k-mer, which a difference inside a "
                        "near-identical repeat can escape; with --strict (or --algorithm trajectory) the error is the "
                        "reference's in every case")
    t = sub.add_parser("track", help="Calculate single and multi-read mappability tracks from shortest unique "
                                     "sequence lengths (on the GPU)")
    t.set_defaults(func=track.main)
    t.add_argument("read_length", nargs="?", default=str(track.DEFAULT_MAPPABILITY_READ_LENGTH), metavar="read_length",
                   help=f"read length the mappability is calculated for (default {track.DEFAULT_MAPPABILITY_READ_LENGTH})")
    t.add_argument("unique_count_files", nargs="+", help="one or more unique count files")
    o = t.add_argument_group("output arguments")
    o.add_argument("--single-read", "-s", metavar="FILE",
                   help=f"single-read mappability BED file ('{track.STDOUT_FILENAME}' for standard output; the default "
                        "when --multi-read is not given)")
    o.add_argument("--multi-read", "-m", metavar="FILE",
                   help=f"multi-read mappability WIG file ('{track.STDOUT_FILENAME}' for standard output)")
    o.add_argument("--verbose", "-v", action="store_true")
    return parser


def parse_subcommands(argv=None):
    args = build_parser().parse_args(argv)
    args.func(args)


def main():
    parse_subcommands(sys.argv[1:])


if __name__ == "__main__":
    main()
