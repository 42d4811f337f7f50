/* newmap_b200.h -- C-ABI of the B200-native `newmap search` hot path.
 *
 * One shared library (newmap_b200/libnewmap_b200.so, hand-written CUDA for sm_100a) that a
 * host program binds with ctypes / dlopen.  Plain pointers and sizes only; all buffers passed in
 * are HOST buffers owned by the caller unless a function says otherwise.  Every function
 * returns an int: >= 0 success, < 0 one of the NM_E* codes (nm_strerror()).  There is no CPU
 * fallback: without a usable CUDA device nm_index_open fails with NM_ECUDA.
 *
 * Reference interfaces replaced (paths relative to hoffmangroup/newmap):
 *   src/newmap-count.c:9-17,136    createIndex / awFmReadIndexFromFile    -> nm_index_open
 *   src/newmap-count.c:202-203     awFmDeallocIndex                        -> nm_index_close
 *   src/newmap-count.c:91-206      count_kmers_from_sequence               -> nm_count_from_sequence
 *   src/newmap-count.c:28-89       count_kmers                             -> nm_count_from_sequence
 *                                                                             (k-mers concatenated by the caller)
 *   AvxWindowFmIndex/src/AwFmParallelSearch.c:136-255  awFmParallelSearchCount -> the kernels behind both
 *   newmap/search.py:383-548       binary_search (+ :647-724 get_kmer_counts,
 *                                  :744-871 ambiguity mask / upper bounds) -> nm_min_unique (mode NM_MODE_RANGE)
 *   newmap/search.py:551-644       linear_search                           -> nm_min_unique (mode NM_MODE_LIST)
 */
#ifndef NEWMAP_B200_H
#define NEWMAP_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define NM_ABI_VERSION 3

/* error codes (negative = failure, mirroring enum AwFmReturnCode's sign convention,
 * AvxWindowFmIndex/src/AwFmIndex.h:131-137) */
#define NM_OK 0
#define NM_EOPEN (-1)      /* index file could not be opened */
#define NM_EFORMAT (-2)    /* not an .awfmi file / truncated / inconsistent header */
#define NM_EVERSION (-3)   /* .awfmi version != 8 or alphabet != DNA */
#define NM_EALLOC (-4)     /* host or device allocation failed */
#define NM_ESHORT (-5)     /* short read */
#define NM_EARG (-6)       /* invalid argument (empty k-mer, index+length beyond the sequence, ...) */
#define NM_ECUDA (-7)      /* CUDA runtime error; nm_last_cuda_error() has the text */
#define NM_EUNSUPPORTED (-8)
#define NM_ELIMIT (-9)     /* request larger than the fixed scratch of nm_count_from_sequence_noalloc */

#define NM_MODE_RANGE 0    /* klist = {min_k, max_k}: minimum over every k in [min_k, max_k] */
#define NM_MODE_LIST 1     /* klist = explicit lengths, tried in the order given */

#define NM_ALGO_AUTO 0       /* trajectory-free search (result identical to the reference's) and, if it
                                meets a k-mer absent from the index, a replay on the reference's
                                trajectory so that the reported k-mer is the reference's */
#define NM_ALGO_TRAJECTORY 1 /* the reference's own query trajectory (binary / linear search of
                                search.py), one thread per position */
#define NM_ALGO_FREE_GENERIC 2 /* trajectory-free search forced onto the unmodified .awfmi block layout */

typedef struct nm_index nm_index;
typedef struct nm_workspace nm_workspace;

typedef struct nm_index_info {
  uint64_t bwt_length;
  uint64_t n_blocks;        /* 160-byte .awfmi blocks */
  uint64_t prefix_sums[6];
  uint64_t device_bytes;    /* HBM held by this index */
  uint32_t seed_length;
  uint32_t sa_compression_ratio;
  uint32_t has_rank32;      /* 1 when the 32-byte rank layout was built (bwt_length < 2^32-1) */
  int32_t device;
  uint32_t derived_seed_length; /* depth of the deeper seed table derived on the device, 0 = none */
  uint32_t reserved;
} nm_index_info;

typedef struct nm_stats {
  uint64_t positions;         /* num_kmers of the call */
  uint64_t lf_steps;          /* backward-extension steps executed on the device */
  uint64_t queries;           /* (position,length) count queries executed (trajectory algo) */
  uint64_t generic_searches;  /* searches routed to the unmodified-layout kernel (x letters, exceptions ...) */
  uint64_t ref_lf_steps;      /* LF steps the reference's trajectory would have executed (0 if not computed: it
                                 needs the search to start from the file's own seed table -- NM_OPT_REF_LF forces
                                 that; nm_min_unique fills it only when the derived table is not in use) */
  int64_t zero_pos;           /* first (round, position) whose two-strand count was 0, else -1 */
  int64_t zero_len;
  int64_t zero_round;
  uint32_t algo_used;         /* NM_ALGO_* that produced the output */
  uint32_t replayed;          /* 1 if NM_ALGO_AUTO fell back to the trajectory replay */
  float ms_h2d, ms_prep, ms_search, ms_generic, ms_resolve, ms_d2h, ms_total; /* CUDA-event times of this call */
  /* trajectory-free search, pass by pass (ms_search = ms_pass1 + ms_cert + ms_pass3) */
  float ms_pass1, ms_cert, ms_pass3;
  uint32_t launches;          /* kernels launched for this call */
  uint32_t seed_depth;        /* depth of the seed table the chains started from */
  uint32_t short_counts;      /* nm_workspace_multi searches: positions whose summed count fell below the number of
                                 sequences without being 0 (the reference's range loop never ends there) */
  uint64_t lf_steps_pass1;    /* part of lf_steps executed by the first pass (every position, depth-capped) */
  uint64_t seed_lookups;      /* seed-table lookups of the trajectory-free search */
  uint64_t seed_lookups_pass1;
  uint32_t stride;            /* sparse first pass: positions per unconditional seed-table lookup (1 = every window) */
  uint32_t out_stats;         /* 1 when the four fields below were computed (NM_OPT_OUT_STATS) */
  uint64_t unique_positions;  /* positions with a non-zero answer */
  uint64_t ambiguous_positions; /* start positions not in b'ACGTacgt' (search.py:744-766: skipped) */
  uint32_t min_unique, max_unique; /* smallest / largest non-zero answer; 0 when there is none */
} nm_stats;

int nm_abi_version(void);
const char *nm_strerror(int code);
const char *nm_last_cuda_error(void);
int nm_device_count(void);              /* >= 0, or NM_ECUDA */

/* Reads header, BWT blocks, prefix sums and seed table of a reference-built .awfmi (v8, DNA;
 * AvxWindowFmIndex/src/AwFmFile.c:174-409; the suffix array and FastaVector sections are not
 * read), uploads them unchanged to `device` and derives the 32-byte rank layout there. */
int nm_index_open(const char *path, int device, nm_index **out);
void nm_index_close(nm_index *ix);
int nm_index_info_get(const nm_index *ix, nm_index_info *info);
/* The seed table of the file (AwFmKmerTable.c:24-35, depth kmerLengthInSeedTable, 12 by default) saves the
 * first seedLen backward-extension steps of a query.  With 180 GB of HBM a DEEPER table pays: nm_index_open
 * derives one on the device from the file's table and the BWT (one extension step per entry and level; each
 * entry holds the interval of its d-mer and of the reverse complement, 16 B; depth chosen from bwtLength and
 * free memory, at most 16 = 69 GB; NEWMAP_B200_SEED_DEPTH overrides, 0 = none) and
 * searches whose shortest length reaches that depth start their chains from it.  This call rebuilds it at
 * another depth (<= seed length: drop it).  No search may be in flight.  Results never depend on it. */
int nm_index_set_seed_depth(nm_index *ix, uint32_t depth);

/* counts[q] = occurrences of seq[idx[q] : idx[q]+len[q]] in the indexed text (u32, five-letter
 * alphabet a c g t/u x exactly like AwFmParallelSearch.c:136-255).  reverse_complement != 0
 * counts the k-mer newmap builds for the other strand (search.py:22,682-689).  Returns NM_EARG
 * if any len[q]==0 or idx[q]+len[q] > seq_len (nothing is launched).  lf_steps may be NULL. */
int nm_count_from_sequence(nm_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                           const uint64_t *len, uint64_t n, int reverse_complement, uint32_t *counts,
                           uint64_t *lf_steps);

/* The same counts as nm_count_from_sequence (same arguments, same results, same NM_EARG rule) WITHOUT ANY ALLOCATION:
 * no cudaMalloc / cudaFree / allocating library routine is called.  (No counterpart in the reference: added
 * for callers that cannot allocate -- code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it.)  nm_count_from_sequence allocates five device buffers per call; here the device
 * copies of seq / idx / len, the counts and the LF-step counter live in ONE fixed array of
 * NM_NOALLOC_SCRATCH_BYTES of device memory inside the library (a file-scope `static __device__` array: every GPU
 * that loads the library holds it until the process exits).  A request needs 256 + 20 * round_up(n, 64) +
 * round_up(seq_len + 16, 256) bytes of it; a larger one is refused with NM_ELIMIT and nothing is launched.
 * Every call on a device shares that array: the calls of one process are serialised by the library and each call
 * clears what it needs to start at zero, but work that another process or stream of the caller's makes overlap with a
 * call on the same GPU through the same array would race -- do not alias it. */
#define NM_NOALLOC_SCRATCH_BYTES (4u << 20)
int nm_count_from_sequence_noalloc(nm_index *ix, const uint8_t *seq, uint64_t seq_len, const uint64_t *idx,
                                   const uint64_t *len, uint64_t n, int reverse_complement, uint32_t *counts,
                                   uint64_t *lf_steps);

/* One segment buffer of search.py (data = num_kmers start positions + look-ahead):
 * out[i], i < num_kmers, = minimum unique length or 0, written as uint8/16/32 (out_width 1/2/4).
 * stats may be NULL.  A two-strand count of 0 is NOT an error here: it is reported in
 * stats->zero_* and the caller raises (search.py:699-722). */
int nm_min_unique(nm_index *ix, const uint8_t *seg, uint64_t seg_len, uint64_t num_kmers, int mode,
                  const uint32_t *klist, uint32_t nk, uint32_t initial_search_length, int use_rc,
                  int algo, int out_width, void *out, nm_stats *stats);

/* ---- pipelined / device-resident form of nm_min_unique -------------------------------------
 * A workspace owns one CUDA stream and device scratch for segments up to max_seg_len bytes.
 * upload/search/download are asynchronous on that stream; sync waits and fills stats.  Several
 * workspaces on one index overlap copies with kernels.  `cuda_stream` may be a cudaStream_t
 * created by the caller (e.g. torch's) or NULL to let the library create one. */
int nm_workspace_create(nm_index *ix, uint64_t max_seg_len, void *cuda_stream, nm_workspace **out);
void nm_workspace_destroy(nm_workspace *ws);
int nm_workspace_upload(nm_workspace *ws, const uint8_t *seg, uint64_t seg_len);
int nm_workspace_search(nm_workspace *ws, uint64_t num_kmers, int mode, const uint32_t *klist, uint32_t nk,
                        uint32_t initial_search_length, int use_rc, int algo, int out_width);
int nm_workspace_download(nm_workspace *ws, void *out);   /* num_kmers * out_width bytes of the last search */
int nm_workspace_sync(nm_workspace *ws, nm_stats *stats);
/* Tuning / instrumentation knobs of a workspace; none of them changes a result byte. */
#define NM_OPT_REF_LF 1   /* 1: also compute stats.ref_lf_steps (LF steps of the reference's trajectory) and, for that,
                             start the chains from the file's seed table instead of the derived one; default 0 */
#define NM_OPT_DCAP 2     /* first-pass depth cap of the trajectory-free search, characters; 0 = seed depth in use + 3 (derived
                             table) or + 12 (the file's table) */
#define NM_OPT_GSIZE 3    /* positions per certificate group; 0 = 40 */
#define NM_OPT_PASS1 4    /* kernel of the first pass: 0 = chosen from the seed depth in use (default), 1 = one thread per
                             position, 2 = tiled (seed lookups batched, lanes refilled from a queue) */
#define NM_OPT_GSIZE2 5   /* positions per group of the coarse certificate round that runs before the fine one (one chain,
                             that much deeper, settles the inside of long exact repeats: fewer LF steps, longer critical path);
                             default 0 = no coarse round, as is any value below 32 */
#define NM_OPT_OWN_CHAINS 6 /* chains of its own a certificate thread may run for a position no bound settles (the rest
                             goes to the third pass); default 1 */
#define NM_OPT_STALL 7    /* first pass: a chain whose interval kept its size for this many steps in a row is left to the
                             later passes at once instead of at the depth cap (it is inside a repeat); default 0 = off */
#define NM_OPT_CERT_RUNS 8 /* certificate pass: runs of consecutive unsettled positions a thread takes per group and strand
                             (one anchor chain each; the rest goes to the third pass); 0 = default 1 */
#define NM_OPT_STRIDE 9   /* sparse first pass: at most this many positions per unconditional seed-table lookup (the rest is
                             settled by upper bounds spread from it, or by chains of their own); 0 = as many as the slack
                             between the table depth and the shortest length allows (at most 8), 1 = every window */
#define NM_OPT_OUT_STATS 10 /* 1: one more reduction over the output fills stats.unique_positions, ambiguous_positions,
                             min_unique, max_unique (what `newmap search --verbose` prints per sequence); default 0 */
#define NM_OPT_SUBDEPTH 11 /* a range that starts below the depth of the derived seed table but reaches it (12:255 on the 16-letter
                             table): 0 (default) = chains start from the derived table and the lengths below its depth are
                             settled with exact counts for the few positions the table does not rule out; 1 = the whole
                             search falls back to the file's own table */
#define NM_OPT_STRICT 12  /* strict index / sequence check for NM_ALGO_AUTO.  The reference raises on the first queried k-mer
                             whose count is 0 (search.py:699-722); the trajectory-free passes see an absent string only when
                             a chain of theirs runs into it, which a difference inside a near-identical repeat can escape.
                             > 0: one more backward search per block of `value` start positions (1 = 1024) proves every k-mer
                             the reference could query present; on failure the segment is replayed on the reference's own
                             trajectory, so bytes and error are the reference's in every case.  Costs about
                             (value + max_k) / value LF steps per position.  Default 0 (off) */
int nm_workspace_option(nm_workspace *ws, int option, uint64_t value);
/* 64-bit FNV-style checksum of the last search's output computed on the device (for
 * device-resident runs that never copy the output back). */
int nm_workspace_checksum(nm_workspace *ws, uint64_t *sum);

/* ---- several sequences x several indexes at once ----------------------------------------------
 * Replaces newmap/search.py:260-265 (zip of the n-th segments of every FASTA file), :461,:489-517 (range mode) /
 * :575,:627-639 (list mode) and :656-697 (get_kmer_counts): the count of a (position, length) is the sum over
 * every index and every sequence, both strands unless use_rc == 0, and a length is unique when the sum equals the
 * number of sequences.  Call after nm_workspace_upload (which carries the segment of the FIRST file: start mask,
 * barriers and length limits come from it alone, search.py:391-394) and before nm_workspace_search:
 *   indexes   n_indexes (1..8) open indexes on the workspace's device, in the order of --index-file;
 *   segments  n_segments (0..7) HOST buffers, the same segment of the 2nd, 3rd ... FASTA file, each of the uploaded
 *             segment's length (the reference assumes equal lengths, search.py:391).
 * The following nm_workspace_search runs on the reference's own query trajectory whatever `algo` says.  The setting
 * lasts until the next nm_workspace_upload / nm_workspace_attach.  nm_stats.short_counts > 0 after the search means
 * the reference would not have terminated (a sum below the number of sequences that is not 0). */
int nm_workspace_multi(nm_workspace *ws, nm_index *const *indexes, uint32_t n_indexes, const uint8_t *const *segments,
                       uint32_t n_segments);

/* ---- FASTA ingest on the device (replaces newmap/fasta.py:20-190 sequence_segments for the plain-text case) ------
 * The raw text of a FASTA file is copied to the GPU as it is read (pinned staging, several reader threads), and
 * parsed there: header lines (first byte '>' or ';' at a line start) are located, newlines and header lines are
 * removed by a stream compaction, and every record becomes a contiguous run of bases in device memory.  Search
 * workspaces then take their segment windows straight from that memory (nm_workspace_attach): no per-segment
 * host->device copy, no host-side parsing.  Semantics are those of the reference's reader: a record's id is the
 * first whitespace-delimited token of its header line without the first byte, its bases are the other lines
 * right-stripped and concatenated.  Right-stripping is only implemented for '\n' line ends: if a sequence line
 * holds any other white space (space, tab, CR, VT, FF) nm_fasta_finish reports odd_whitespace = 1 and the caller
 * must use its own reader (newmap_b200/fasta.py does). */
typedef struct nm_fasta nm_fasta;
typedef struct nm_fasta_info {
  uint64_t raw_bytes;       /* text bytes received */
  uint64_t sequence_bytes;  /* bases kept (all records) */
  uint64_t records;         /* records, including a leading one without header line if the text starts with bases */
  uint32_t odd_whitespace;  /* 1: white space other than '\n' inside sequence lines -- results of this parse are not usable */
  uint32_t reserved;
  float ms_parse;           /* device time of the parse */
  float reserved2;
} nm_fasta_info;
typedef struct nm_fasta_record {
  uint64_t seq_offset;      /* first base of the record in the device sequence buffer */
  uint64_t seq_length;
  uint64_t header_offset;   /* header line (without its newline) in the packed header text, nm_fasta_headers() */
  uint64_t header_length;   /* 0 for a leading record without header line */
} nm_fasta_record;
int nm_fasta_open(int device, uint64_t size_hint, nm_fasta **out);
void nm_fasta_close(nm_fasta *fa);
/* appends n bytes of text (any host memory) */
int nm_fasta_feed(nm_fasta *fa, const void *data, uint64_t n);
/* appends a whole plain-text file, read by `threads` reader threads (0 = 4) straight into pinned staging */
int nm_fasta_feed_file(nm_fasta *fa, const char *path, uint32_t threads);
/* The same for the bytes [offset, offset + length) of the file (clipped to its size): a rank of a sharded run reads only
 * the part of the text its windows lie in (newmap_b200/search.py, partitioned ingest). */
int nm_fasta_feed_file_range(nm_fasta *fa, const char *path, uint64_t offset, uint64_t length, uint32_t threads);
int nm_fasta_finish(nm_fasta *fa, nm_fasta_info *info);
int nm_fasta_records(const nm_fasta *fa, nm_fasta_record *out, uint64_t capacity);      /* fills min(records, capacity) */
int nm_fasta_headers(const nm_fasta *fa, char *out, uint64_t capacity, uint64_t *needed); /* packed header lines */
/* copies bases [seq_offset, seq_offset + n) of the device sequence buffer to host memory (error messages, tests) */
int nm_fasta_read(const nm_fasta *fa, uint64_t seq_offset, uint64_t n, uint8_t *out);
/* the workspace's next search reads its segment from the parsed sequence buffer instead of an uploaded one */
int nm_workspace_attach(nm_workspace *ws, const nm_fasta *fa, uint64_t seq_offset, uint64_t seg_len);

/* ---- `newmap track` on the device (newmap/track.py:22-111): mappability tracks from <id>.unique.uintN ------------
 * multi-read mappability of position i for read length k = (number of j in (i-k, i] with 0 < unique[j] <= k) / k
 * (track.py:22-62: +1/-1 interval marks and a cumulative sum); single-read mappability = that number > 0
 * (track.py:65-68).  The device keeps the integer window counts; the BED runs and the WIG text are produced from
 * them (run starts by stream compaction; the text by a per-position length scan and a scatter from the k+1
 * possible strings, which the caller formats once on the host so that the digits are the reference's own). */
typedef struct nm_track nm_track;
int nm_track_open(int device, nm_track **out);
void nm_track_close(nm_track *tr);
/* uploads n values of width 1/2/4 bytes (host memory) and computes the window counts for read length k */
int nm_track_compute(nm_track *tr, const void *unique, uint64_t n, int width, uint32_t kmer_length);
/* window counts (u32) of positions [first, first + m) to host memory */
int nm_track_counts(nm_track *tr, uint64_t first, uint64_t m, uint32_t *out);
/* maximal runs of equal single-read value: *nruns = their number; starts/values filled up to capacity (may be NULL) */
int nm_track_runs(nm_track *tr, uint64_t *starts, uint8_t *values, uint64_t capacity, uint64_t *nruns);
/* BED lines "<chr>\t<start>\t<end>\tk<k>\t<0|1>\t.\n" (track.py:86-97) of runs [first_run, first_run + m) of the last
 * nm_track_runs, assembled on the device; *written = bytes produced, NM_EARG if they exceed capacity */
int nm_track_bed_text(nm_track *tr, const char *chr_name, uint64_t first_run, uint64_t m, char *out, uint64_t capacity,
                      uint64_t *written);
/* text of positions [first, first + m): strings[offsets[c] .. offsets[c+1]) + '\n' for the count c of each position
 * (offsets has kmer_length + 2 entries); *written = bytes produced, NM_EARG if they exceed capacity */
int nm_track_text(nm_track *tr, const char *strings, const uint32_t *offsets, uint64_t first, uint64_t m, char *out,
                  uint64_t capacity, uint64_t *written);

/* Pinned host memory for segment / output staging (cudaHostAlloc / cudaFreeHost). */
void *nm_host_alloc(uint64_t bytes);
void nm_host_free(void *p);

/* GPU construction of a reference-format .awfmi (v8, DNA), byte-identical to what the reference's
 * CPU builder writes (AwFmCreate.c:128-389 awFmCreateIndexFromFasta + AwFmFile.c:20-171), from the
 * three FastaVector buffers of the FASTA (lib/FastaVector/src/FastaVector.c:50-168): `sequence` = all
 * sequences concatenated with a '\0' after each, `header` = header lines each followed by '\0',
 * `metadata` = metadata_count pairs {headerEndPosition, sequenceEndPosition}.  bwtLength must stay
 * below 2^32-16.  seconds_sort (optional) receives the device time of the suffix sort. */
int nm_index_build(const uint8_t *sequence, uint64_t sequence_len, const uint8_t *header, uint64_t header_len,
                   const uint64_t *metadata, uint64_t metadata_count, uint32_t sa_ratio, uint32_t seed_len,
                   int device, const char *out_path, double *seconds_sort);
const char *nm_build_last_error(void);

/* Random-access roofline probe: `chains` threads each issue, `steps` times, `mlp` (1, 2, 4 or 8) loads of
 * `granule` bytes (32, 64 or 128, granule-aligned) at pseudo-random offsets of a `bytes`-sized (power of
 * two) device buffer.  dependent != 0: the next offset of each of the mlp chains depends on the data
 * loaded (like the LF steps of a search); dependent == 0: all loads are independent (the ceiling of the
 * memory system for random granules).  Returns loads/second (CUDA events). */
int nm_probe_random_access(int device, uint64_t bytes, uint32_t granule, uint64_t chains, uint32_t steps,
                           uint32_t mlp, int dependent, double *loads_per_second, float *ms);

#ifdef __cplusplus
}
#endif
#endif
