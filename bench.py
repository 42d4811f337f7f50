#!/usr/bin/env python
"""bench.py -- genome positions/s of the `newmap search` hot path (min-unique-k, k=20:200).

    python bench.py --gpus N --steps K --warmup W [--workload c2_100m|c3_1g|c4_3g] [--impl reference]

One process per GPU (torchrun for N > 1).  The index is replicated on every GPU, positions are
sharded by 10 M-position segment (the reference's own batching unit), there is no data-path
collective: torch.distributed is used only for the barrier and the max-over-ranks of the time.

Workload: N = 1 -> c3_1g (BASELINE configs[2], 1 Gbp); N > 1 -> c4_3g (BASELINE configs[3], 3.1 Gbp).
A step = one pass of the hot path over this rank's `segments_per_step` segments (200 x 10 M positions,
taken cyclically from the genome starting at rank * segments_per_step: weak scaling).
  value : positions/s with the parsed genome already resident in HBM (CUDA events around K steps)
  e2e   : the same through the C-ABI with pinned HOST buffers, H2D + kernels + D2H every step,
          pipelined over a few workspaces (what newmap_b200.search.write_unique_counts does)
  job   : the whole genome through newmap_b200.search.write_unique_counts with RANK / WORLD_SIZE
          (FASTA file in, <id>.unique.uint8 files out), wall clock, max over ranks: STRONG scaling
          of the thing a user runs, index-open time reported separately
  roofline : dominant kernel of the search -- (LF steps x 64 B + seed lookups x 32 B) / its CUDA-event
          time measured with one search on the GPU at a time, against the measured streaming HBM peak
          (MEASURED_PEAKS.json) and against a random 32 B-sector probe measured in the same run
  parity_sample : the reference's own answers (oracle/_ref under the reference's host loop) for a
          contiguous sample of the bench genome compared byte for byte with the GPU's
  cpu_baseline / --impl reference : the reference's own AVX2/OpenMP count code (oracle/_ref) under
          the reference's host loop (oracle/refdrive.py) on a bounded sample, all host threads
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

K_RANGE = (20, 200)
BATCH = 10_000_000
SEED_LEN, SA_RATIO = 12, 8
CACHE = Path(os.environ.get("NEWMAP_B200_CACHE", "/tmp/newmap_b200_cache"))
WORKLOADS = {
    "c2_100m": "synthetic 100 Mbp single-chromosome genome (BASELINE configs[1] genome), k=20:200",
    "c3_1g": "synthetic 1 Gbp 8-chromosome genome (BASELINE configs[2]), k=20:200",
    "c4_3g": "synthetic 3.1 Gbp 24-sequence genome (BASELINE configs[3]), k=20:200",
    "c3r_1g": "synthetic 1 Gbp 8-chromosome genome with a mammalian interspersed-repeat spectrum (43 % of it copies of "
              "diverged repeat families, newmap_b200/synth.py MAMMALIAN_FAMILIES), k=20:200",
}
# the reference's index builder (libdivsufsort, one thread, ~11 B of host RAM per base) takes ~40 s per 100 Mbp:
# used up to this genome size, the GPU builder (same file byte for byte, profiles/r2_index_cmp_*.json) above it
REF_BUILDER_MAX_BP = int(float(os.environ.get("NEWMAP_B200_REF_BUILDER_MAX_BP", "1.2e9")))


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def default_workload(world: int) -> str:
    return os.environ.get("NEWMAP_B200_WORKLOAD") or ("c3_1g" if world == 1 else "c4_3g")


# --------------------------------------------------------------------------------------------
# workload: synthetic genome + .awfmi index, cached on local disk
# --------------------------------------------------------------------------------------------
def build_index(fa: Path, ix: Path, builder: str, device: int = 0) -> str:
    """Writes the .awfmi of `fa`; returns a note naming who built it."""
    t0 = time.time()
    if builder == "reference":
        from oracle import ora   # the reference's index builder lives in oracle/_ref (not on the timed path)
        if not ora.ref_available():
            raise SystemExit("bench: oracle/_ref (reference index builder) is missing; run __graft_entry__.build() "
                             "where /root/reference exists")
        ora.ref_build_index(fa, ix, SA_RATIO, SEED_LEN)
        note = "the reference's own builder (oracle/_ref: AwFmCreate + libdivsufsort)"
    elif builder == "gpu-subprocess":
        # the reference arm must not map libnewmap_b200.so: `newmap index` of the product runs in a child process
        env = dict(os.environ, NEWMAP_B200_DEVICE=str(device))
        subprocess.run([sys.executable, "-m", "newmap_b200", "index", str(fa), "--output", str(ix),
                        "--compression-ratio", str(SA_RATIO), "--seed-length", str(SEED_LEN)], check=True, env=env,
                       cwd=str(ROOT))
        note = "GPU builder (newmap_b200 index; same file as the reference builder's, profiles/r2_index_cmp_*.json)"
    elif builder == "gpu":
        from newmap_b200 import index
        index.generate_fm_index(fa, ix, SA_RATIO, SEED_LEN, device=device)
        note = "GPU builder (newmap_b200 index; same file as the reference builder's, profiles/r2_index_cmp_*.json)"
    else:
        raise ValueError(builder)
    log(f"[bench] index built by {note} in {time.time() - t0:.1f}s ({ix.stat().st_size / 1e6:.0f} MB)")
    return note


def prepare_workload(name: str, arm: str = "ours", device: int = 0) -> tuple[Path, Path]:
    """FASTA + .awfmi for the workload; built once per box.  Index construction is outside the hot path; who
    builds it: $NEWMAP_B200_INDEX_BUILDER = reference | gpu | auto (default).  auto = the reference's own builder
    for genomes up to REF_BUILDER_MAX_BP (so the search loads a REFERENCE-BUILT index), the GPU builder above."""
    from newmap_b200 import synth
    d = CACHE / name
    fa, ix, done = d / "genome.fa", d / "genome.awfmi", d / ".done"
    if done.is_file():
        return fa, ix
    d.mkdir(parents=True, exist_ok=True)
    lengths, seed = synth.config_lengths(name)
    t0 = time.time()
    synth.write_fasta(fa, synth.make_genome(lengths, seed, families=synth.MAMMALIAN_FAMILIES if name == "c3r_1g" else None))
    log(f"[bench] genome {name}: {sum(lengths)} bp written in {time.time() - t0:.1f}s")
    builder = os.environ.get("NEWMAP_B200_INDEX_BUILDER", "auto")
    if builder == "auto":
        builder = "reference" if sum(lengths) <= REF_BUILDER_MAX_BP else "gpu"
    if builder == "gpu" and arm == "reference":
        builder = "gpu-subprocess"
    note = build_index(fa, ix, builder, device)
    (d / "builder.txt").write_text(note)
    done.write_text("ok")
    return fa, ix


def base_line(workload: str, world: int, steps: int, warmup: int, k_range: str | None, segments_per_step: int) -> dict:
    """The part of the JSON line both arms print identically (metric, unit, config)."""
    note = CACHE / workload / "builder.txt"
    line = {"metric": "genome positions/sec, min-unique-k search k=20:200", "unit": "positions/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": WORKLOADS[workload], "workload_id": workload, "kmer_batch_size": BATCH,
                       "index": f".awfmi v8 (reference format), seed {SEED_LEN}, SA ratio {SA_RATIO}; built by "
                                + (note.read_text() if note.is_file() else "unknown"),
                       "parallelism": f"index replicated x{world}, segments sharded, no collective",
                       "segments_per_step_per_gpu": segments_per_step,
                       "l2": "no flush: every step streams segments_per_step x 10 MB of sequence and random-reads an "
                             "index (rank32 + seed table, 70 GB) far larger than the 126 MB L2"}}
    if world > 1 and workload == "c4_3g":
        # N = 1 runs BASELINE configs[2] (1 Gbp), N > 1 configs[3] (3.1 Gbp): a ratio of the two lines mixes a change of
        # genome with the scaling.  The single-GPU figure for THIS genome, for whoever wants the scaling alone:
        line["config"]["same_workload_on_one_gpu"] = ("15.7 G positions/s resident on c4_3g with 1 GPU "
                                                      "(profiles/r2_sweep_i_cert_runs.jsonl); the N=1 line of bench.py is c3_1g")
    if k_range:
        line["metric"] = f"genome positions/sec, min-unique-k search k={k_range}"
        line["config"]["k_range"] = k_range
        line["config"]["workload"] = line["config"]["workload"].replace("k=20:200", f"k={k_range}")
    return line


def sample_segment(fa: Path, number: int):
    """The `number`-th segment the reference's FASTA reader cuts from the genome (host reader, stops there)."""
    from newmap_b200.fasta import sequence_segments
    from newmap_b200.search import get_num_kmers
    max_k = K_RANGE[1]
    with open(fa, "rb") as fh:
        for j, s in enumerate(sequence_segments(fh, BATCH + max_k - 1, max_k - 1)):
            if j == number:
                return s, get_num_kmers(s, max_k)
    raise SystemExit(f"bench: the genome has no segment {number}")


# --------------------------------------------------------------------------------------------
# clocks: NVML from one thread of rank 0 (no forked nvidia-smi inside the timed region)
# --------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, device: int, enabled: bool = True, period: float = 0.05):
        self.device, self.enabled, self.period = device, enabled, period
        self.rows, self._stop, self._t = [], threading.Event(), None
        self.h = self.nv = None
        if enabled:
            try:
                import pynvml
                pynvml.nvmlInit()
                idx = device
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                if vis:
                    try:
                        idx = int(vis.split(",")[device])
                    except (ValueError, IndexError):
                        idx = device
                self.h, self.nv = pynvml.nvmlDeviceGetHandleByIndex(idx), pynvml
            except Exception as e:    # the line then says clocks were not sampled
                log(f"[bench] NVML unavailable: {e}")

    def _run(self):
        nv, h = self.nv, self.h
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append((sm, mx, reasons))
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.h is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=2)

    def summary(self) -> dict:
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0,
                    "how": "NVML, rank 0 only" + ("" if self.enabled else " (not this rank)")}
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                "hw_power_brake_slowdown": 0x80}
        seen = 0
        for _, _, r in self.rows:
            seen |= int(r)
        return {"sm_mhz": float(np.median([r[0] for r in self.rows])), "sm_max_mhz": float(max(r[1] for r in self.rows)),
                "reasons": sorted(n for n, b in bits.items() if seen & b), "samples": len(self.rows),
                "how": "NVML from one thread of rank 0 every 50 ms during the timed region"}


# --------------------------------------------------------------------------------------------
# CPU reference arm
# --------------------------------------------------------------------------------------------
def cpu_reference_run(fa: Path, ix: Path, steps: int, warmup: int, sample_positions: int) -> dict:
    """The reference's AVX2/OpenMP count path (oracle/_ref) under the reference's host loop on a
    bounded sample: the first `sample_positions` start positions of the second segment."""
    from oracle import ora, refdrive
    if not ora.ref_available():
        raise SystemExit("bench: oracle/_ref missing")
    seg, _ = sample_segment(fa, 1)
    max_k = K_RANGE[1]
    data = seg.data[:sample_positions + max_k - 1]
    cores = min(os.cpu_count() or 1, 255)
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    timer = refdrive.Timer()
    per_step = None
    for _ in range(warmup):
        per_step = _timed(lambda: refdrive.search_segment(ix, data, sample_positions, kl, True, 0, True, cores))
    # One step costs the same ~16 index re-reads whatever the sample size (1.7 s each for the 3.1 Gbp index): when K steps
    # would run past the budget, fewer steps are executed -- the figure is a rate, and the line says how many it is over
    asked, budget = steps, float(os.environ.get("NEWMAP_B200_REF_ARM_BUDGET_S", "420"))
    if per_step and steps * per_step > budget:
        steps = max(2, int(budget / per_step))
    t0 = time.perf_counter()
    for _ in range(steps):
        out = refdrive.search_segment(ix, data, sample_positions, kl, True, 0, True, cores, timer)
    dt = time.perf_counter() - t0
    # Every count call of the reference re-reads the whole index from disk (src/newmap-count.c:9-17,136).
    # In a real run that cost is spread over a 10 M-position batch, not over this sample: `value_amortised`
    # charges the sample only its share of it (measured with a one-query call).  `value` is what was measured.
    mod = ora.ref_modules()[0]
    reload_s = min(_timed(lambda: mod.count_kmers_from_sequence(str(ix), data[:64], [0], [20], cores)) for _ in range(3))
    share = min(1.0, sample_positions / BATCH)
    dt_amortised = max(dt - timer.calls * reload_s * (1.0 - share), 1e-9)
    return {"value": sample_positions * steps / dt, "value_amortised": sample_positions * steps / dt_amortised,
            "seconds": dt, "seconds_amortised": dt_amortised, "cores": cores, "steps_executed": steps, "steps_asked": asked,
            "index_reload_s_per_call": reload_s, "native_share": timer.native / dt if dt else None,
            "sample": f"{sample_positions} positions of segment 1 x {steps} step(s)"
                      + (f" (of {asked} asked for: {budget:.0f} s budget)" if steps != asked else "")
                      + f", k={K_RANGE[0]}:{K_RANGE[1]}, "
                      f"{timer.calls} count_kmers_from_sequence calls, each re-reading the index ({reload_s:.2f}s) as "
                      f"the reference does; value_amortised charges that re-read at the sample's share of a 10M batch",
            "out": out, "data": data}


def parity_sample(index, r: dict, sample_positions: int) -> dict:
    """The bytes the reference computed for the sample against the GPU's for the same buffer (through the C-ABI)."""
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    got, st = index.min_unique(r["data"], sample_positions, kl, True, algo="auto", dtype=np.uint32)
    want = np.asarray(r["out"], dtype=np.uint32)
    bad = np.nonzero(got != want)[0]
    res = {"positions": int(sample_positions), "mismatches": int(bad.size), "source": "oracle/_ref",
           "what": "reference's compiled count code under the reference's host loop vs nm_min_unique (algo auto) "
                   "on the same bytes of the bench genome (segment 1)",
           "unique_positions": int(np.count_nonzero(want)), "checksum_reference": int(want.astype(np.uint64).sum())}
    if bad.size:
        res["first"] = [(int(i), int(got[i]), int(want[i])) for i in bad[:5]]
    return res


def _timed(fn) -> float:
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


# --------------------------------------------------------------------------------------------
# the job a user runs: FASTA file -> <id>.unique.uint8 files, sharded over the ranks
# --------------------------------------------------------------------------------------------
def job_run(fa: Path, ix: Path, device: int, rank: int, world: int, barrier, reduce_max) -> dict:
    """newmap_b200.search.write_unique_counts over the whole workload genome with RANK / WORLD_SIZE, index already
    resident (the reference re-reads its index on every count call; here it is opened once per process): wall clock
    between barriers = first byte read to last byte written, max over ranks; best of 3."""
    import shutil
    from newmap_b200 import search
    out = CACHE / "job_out"
    best, phases = None, {}
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    for _ in range(3):
        if rank == 0:
            shutil.rmtree(out, ignore_errors=True)
            out.mkdir(parents=True)
        barrier()
        cfg = search.SearchConfig(fasta_filepaths=[fa], fmindex_filepaths=[ix], kmer_lengths=kl, is_binary_search=True,
                                  kmer_batch_size=BATCH, output_directory=out, device=device, rank=rank, world_size=world)
        t0 = time.perf_counter()
        search.write_unique_counts(cfg)
        mine = time.perf_counter() - t0
        barrier()
        wall = time.perf_counter() - t0
        dt = reduce_max(mine)
        marks = dict(getattr(search, "LAST_RUN_TRACE", []) or [])
        if best is None or dt < best:
            best = dt
            text = marks.get("text_on_device", 0.0)
            parsed = marks.get("parsed", text)
            searched = marks.get("all_segments_searched", parsed)
            written = marks.get("all_written", searched)
            phases = {"read_fasta_to_device_s": reduce_max(text), "parse_s": reduce_max(parsed - text),
                      "search_and_download_s": reduce_max(searched - parsed),
                      "write_tail_s": reduce_max(written - searched), "teardown_s": reduce_max(mine - written),
                      "wall_between_barriers_s": reduce_max(wall)}
    positions = sum(p.stat().st_size for p in out.glob("*.unique.uint8"))
    barrier()
    if rank == 0:
        shutil.rmtree(out, ignore_errors=True)
    return {"value": positions / best, "unit": "positions/s", "positions": positions, "seconds": best,
            "scaling": "strong", "phases_max_over_ranks": phases,
            "what": "write_unique_counts with RANK/WORLD_SIZE: every rank reads and parses the FASTA text on its GPU, "
                    "searches the segments whose running number is congruent to its rank, pwrites its windows of "
                    "<id>.unique.uint8 (page cache); index resident; max over ranks, best of 3"}


# --------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS),
                    help="default: c3_1g on one GPU, c4_3g on several ($NEWMAP_B200_WORKLOAD overrides)")
    ap.add_argument("--segments-per-step", type=int, default=int(os.environ.get("NEWMAP_B200_SEGMENTS_PER_STEP", "200")))
    ap.add_argument("--cpu-sample", type=int, default=int(os.environ.get("NEWMAP_B200_CPU_SAMPLE", "1000000")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-job", action="store_true")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--k-range", default=os.environ.get("NEWMAP_B200_K_RANGE") or None,
                    help="other search range than the metric's 20:200, e.g. 12:255 (BASELINE configs[4]); the JSON "
                         "line then says so in metric and config.k_range")
    ap.add_argument("--pool", type=int, default=int(os.environ.get("NEWMAP_B200_POOL", "10")),
                    help="workspaces (one CUDA stream each) the device-resident arm deals its segments over")
    ap.add_argument("--e2e-depth", type=int, default=int(os.environ.get("NEWMAP_B200_E2E_DEPTH", "8")),
                    help="workspaces (streams) the end-to-end arm pipelines over")
    args = ap.parse_args()

    global K_RANGE
    if args.k_range:
        K_RANGE = tuple(int(x) for x in args.k_range.split(":"))
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    workload = args.workload or default_workload(world)
    S = max(1, args.segments_per_step)

    # ---------------- reference arm: CPU only, rank 0 only, never maps libnewmap_b200.so
    if args.impl == "reference":
        if rank != 0:
            return
        fa, ix = prepare_workload(workload, arm="reference", device=local)
        r = cpu_reference_run(fa, ix, max(1, args.steps), min(args.warmup, 1), args.cpu_sample)
        line = base_line(workload, world, args.steps, args.warmup, args.k_range, S)
        line.update(impl="reference", value=r["value"], ms_per_step=1e3 * r["seconds"] / max(1, r["steps_executed"]),
                    steps_executed=r["steps_executed"],
                    cpu_baseline={"value": r["value"], "unit": "positions/s", "cores": r["cores"], "kind": "reference",
                                  "sample": r["sample"], "native_share": r["native_share"],
                                  "value_amortised": r["value_amortised"],
                                  "index_reload_s_per_call": r["index_reload_s_per_call"]},
                    e2e={"value": r["value"], "unit": "positions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    gpu_launches=0)
        print(json.dumps(line), flush=True)
        return

    # ---------------- our arm
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    # rank 0 prints ONE JSON line on stdout: everything else that writes to fd 1 (NCCL's version banner is printed
    # from C at its first collective, whatever NCCL_DEBUG says) is sent to stderr, the line goes to the saved fd
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()

    def reduce_max(x: float) -> float:
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    if rank == 0:
        prepare_workload(workload, device=local)
    barrier()
    fa, ix_path = prepare_workload(workload, device=local)

    from newmap_b200 import _lib
    from newmap_b200.count import open_index
    from newmap_b200.fasta import window_plan
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    max_k = K_RANGE[1]
    t0 = time.time()
    index = open_index(ix_path, local)     # the cache write_unique_counts (job) also uses: one resident copy
    index_open_s = time.time() - t0
    log(f"[bench r{rank}] index resident in {index_open_s:.1f}s: bwt {index.bwt_length}, "
        f"{index.info.device_bytes / 1e9:.2f} GB HBM (derived seed table depth {index.derived_seed_length})")

    # ---- the genome, parsed on the device once (the product's ingest path); segments = the reference's windows
    t0 = time.time()
    fa_dev = _lib.Fasta(local, size_hint=fa.stat().st_size)
    fa_dev.feed_file(fa)
    fa_dev.finish()
    windows = []
    for _hdr, rec_off, rec_len in fa_dev.records:
        for start, length, epilogue in window_plan(rec_len, BATCH + max_k - 1, max_k - 1):
            windows.append((rec_off + start, length, length if epilogue else length - (max_k - 1)))
    mine = [windows[(rank * S + j) % len(windows)] for j in range(S)]
    positions_per_step = sum(n for _, _, n in mine)
    max_len = max(ln for _, ln, _ in mine)
    log(f"[bench r{rank}] genome on the device in {time.time() - t0:.1f}s: {len(windows)} segments, "
        f"{S} per step starting at {rank * S % len(windows)}")
    stream = torch.cuda.current_stream()

    # ---- device-resident arm: a pool of workspaces (one stream each) searches this rank's windows where they lie
    pool = [_lib.Workspace(index, max_len) for _ in range(max(1, args.pool))]

    def check(st):
        if st.replayed or st.zero_pos >= 0:
            raise SystemExit("bench: a k-mer of the genome is absent from the index (index / FASTA mismatch)")

    def step_resident(collect=None, sums=None):
        busy = [False] * len(pool)
        for j, (off, ln, n) in enumerate(mine):
            w = j % len(pool)
            ws = pool[w]
            if busy[w]:
                st = ws.sync()
                check(st)
                if collect is not None:
                    collect.append(st.as_dict())
                if sums is not None:
                    sums.append(ws.checksum())
            ws.attach(fa_dev, off, ln)
            ws.search(n, kl, True, algo=args.algo, out_width=1)
            busy[w] = True
        # the last len(pool) searches, in launch order
        first = max(0, len(mine) - len(pool))
        for j in range(first, len(mine)):
            ws = pool[j % len(pool)]
            st = ws.sync()
            check(st)
            if collect is not None:
                collect.append(st.as_dict())
            if sums is not None:
                sums.append(ws.checksum())

    checks_resident: list = []
    step_resident(sums=checks_resident)
    for _ in range(max(args.warmup, 3) - 1):
        step_resident()
    stats: list = []
    barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local, enabled=(rank == 0)) as clocks:
        ev0.record(stream)
        for _ in range(args.steps):
            step_resident(stats)        # every workspace is synchronised inside
        ev1.record(stream)
        torch.cuda.synchronize()
        barrier()
    ms_value = ev0.elapsed_time(ev1)

    # ---- end-to-end arm: pinned host in, pinned host out, `depth` workspaces with their own streams
    depth = max(1, args.e2e_depth)
    host_in = _lib.PinnedBuffer(sum(ln for _, ln, _ in mine))
    in_off = []
    acc = 0
    for off, ln, _ in mine:
        host_in.array[acc:acc + ln] = np.frombuffer(fa_dev.read(off, ln), dtype=np.uint8)
        in_off.append(acc)
        acc += ln
    slots = [(pool[j], _lib.PinnedBuffer(max_len)) for j in range(min(depth, len(pool)))]
    depth = len(slots)

    def step_e2e(verify=False):
        pending = []
        oks = []
        for j, (off, ln, n) in enumerate(mine):
            ws, outb = slots[j % depth]
            if len(pending) >= depth:
                w, ob, nn, jj = pending.pop(0)
                check(w.sync())
                if verify:
                    oks.append(_lib.checksum_host(ob.array[:nn]) == checks_resident[jj])
            ws.upload(host_in.ptr + in_off[j], ln)
            ws.search(n, kl, True, algo=args.algo, out_width=1)
            ws.download(outb.ptr)
            pending.append((ws, outb, n, j))
        for w, ob, nn, jj in pending:
            check(w.sync())
            if verify:
                oks.append(_lib.checksum_host(ob.array[:nn]) == checks_resident[jj])
        return oks

    oks = step_e2e(verify=True)
    if not all(oks):
        raise SystemExit("bench: e2e output differs from the device-resident output")
    step_e2e()
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    barrier()

    # ---- the same searches one at a time (nothing else on the GPU): per-kernel CUDA-event times for the roofline.
    # In the timed regions above the segments' kernels overlap across streams, so their event times are not
    # durations of a kernel that has the GPU to itself.
    iso: list = []
    iso_positions = 0
    ws0 = pool[0]
    for _ in range(3):
        for off, ln, n in mine[:10]:
            ws0.attach(fa_dev, off, ln)
            ws0.search(n, kl, True, algo=args.algo, out_width=1)
            iso.append(ws0.sync().as_dict())
            iso_positions += n

    # ---- cross-check one segment against the reference-trajectory kernel (different algorithm, same bytes)
    off0, ln0, n0 = mine[0]
    ws0.attach(fa_dev, off0, ln0)
    ws0.search(n0, kl, True, algo="trajectory", out_width=1)
    st_traj = ws0.sync()
    if ws0.checksum() != checks_resident[0]:
        raise SystemExit("bench: trajectory-free output differs from the reference-trajectory kernel")

    # ---- the job (strong scaling of the whole run), every rank.  What the arms above hold is released first (2 GB of
    # pinned host memory, the parsed genome, the workspaces): the job is what a user runs, in a process that holds
    # the index and nothing else
    host_in.close()
    for _, outb in slots:
        outb.close()
    for ws in pool:
        ws.close()
    fa_dev.close()
    job = None
    if not args.no_job and not args.k_range:
        job = job_run(fa, ix_path, local, rank, world, barrier, reduce_max)
        job["index_open_s_max_over_ranks"] = reduce_max(index_open_s)

    # ---- max over ranks
    ms_value, ms_e2e = reduce_max(ms_value), reduce_max(ms_e2e)
    if world > 1:
        agg = torch.tensor([float(positions_per_step)], device="cuda", dtype=torch.float64)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        positions_all = agg.item()
    else:
        positions_all = float(positions_per_step)

    if rank == 0:
        nsearch = max(1, len(iso))
        lf = sum(s["lf_steps"] for s in iso)
        ms_search = sum(s["ms_search"] for s in iso)
        seeds = sum(s["seed_lookups"] for s in iso)
        lf1 = sum(s["lf_steps_pass1"] for s in iso)
        seeds1 = sum(s["seed_lookups_pass1"] for s in iso)
        kernel_ms = {k: sum(s[k] for s in iso) / nsearch for k in
                     ("ms_prep", "ms_pass1", "ms_cert", "ms_pass3", "ms_generic", "ms_resolve")}
        # the dominant kernel of the search and its algorithmic bytes: pass 1 (k_pass1_sparse / k_free_chain) = its LF
        # steps + its seed lookups; certificate pass (k_free_cert) = the LF steps and lookups that are not pass 1's,
        # shared with pass 3 in proportion to time (the counters of passes 2 and 3 are not separate)
        ms_pass1, ms_cert, ms_pass3 = (sum(s[k] for s in iso) for k in ("ms_pass1", "ms_cert", "ms_pass3"))
        deep = iso and iso[0]["seed_depth"] >= 13
        if ms_pass1 >= ms_cert:
            dom_name = ("k_pass1_sparse" if deep else "k_free_chain") + " (pass 1 of the trajectory-free search)"
            dom_ms, dom_lf, dom_seeds = ms_pass1, lf1, seeds1
        else:
            dom_name = "k_free_cert (certificate pass of the trajectory-free search: one anchor chain per run of deep positions)"
            share = ms_cert / max(ms_cert + ms_pass3, 1e-9)
            dom_ms, dom_lf, dom_seeds = ms_cert, (lf - lf1) * share, (seeds - seeds1) * share
        bytes_dom = dom_lf * 64 + dom_seeds * 32
        bytes_all = lf * 64 + seeds * 32
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"     # dram__bytes_read+write per launch from an `ncu --set full` capture
        if tf.is_file():
            t = json.loads(tf.read_text()).get(workload, {})
            traffic = t.get("k_pass1_sparse" if ms_pass1 >= ms_cert else "k_free_cert")
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.is_file():
            peaks = json.loads(pk.read_text())
        hbm_peak, peak_src = (peaks.get("hbm_gbs"), "measured (MEASURED_PEAKS.json)") if peaks.get("hbm_gbs") else (6650.0, "fallback")
        # random-sector probes on this GPU, same run (2 GiB buffer): dependent chains (what a search does) and
        # independent loads (the ceiling of the memory system for random 32 B sectors)
        probe, probe_indep = {}, {}
        for g in (32, 64, 128):
            probe[g] = max(_lib.probe_random_access(local, 2 << 30, g, 8 << 20, 32, 1, True)[0],
                           _lib.probe_random_access(local, 2 << 30, g, 2 << 20, 32, 4, True)[0])
            probe_indep[g] = max(_lib.probe_random_access(local, 2 << 30, g, 2 << 20, 32, 8, False)[0],
                                 _lib.probe_random_access(local, 2 << 30, g, 8 << 20, 32, 4, False)[0])
        achieved = bytes_dom / (dom_ms * 1e-3) / 1e9 if dom_ms else None
        # the whole step as it really runs (kernels of several segments overlapped across streams): DRAM bytes per segment
        # from the ncu capture of every kernel of a search (profiles/traffic.json) over the timed region's time per segment
        whole_step = None
        if tf.is_file():
            t = json.loads(tf.read_text()).get(workload, {})
            seg_bytes = sum(v for k, v in t.items() if k.startswith("k_") and isinstance(v, (int, float)))
            if seg_bytes and not args.k_range:
                ms_seg = ms_value / args.steps / S
                whole_step = {"dram_bytes_per_segment_ncu": seg_bytes, "ms_per_segment_pipelined": ms_seg,
                              "achieved": seg_bytes / (ms_seg * 1e-3) / 1e9, "unit": "GB/s",
                              "frac_of_streaming_peak": seg_bytes / (ms_seg * 1e-3) / 1e9 / hbm_peak,
                              "random_accesses_per_s": seg_bytes / 64 / (ms_seg * 1e-3),
                              "ratio_to_random_64B_probe": seg_bytes / 64 / (ms_seg * 1e-3) / (probe_indep[64] or 1),
                              "how": "every DRAM byte of one search (ncu, one segment: random sectors of the index AND the "
                                     "streamed codes / G / R / output) / time per segment of the timed region; the probe = "
                                     "independent random 64-byte accesses per second measured in this run (a DRAM sector miss "
                                     "fetches its 64-byte pair) -- a ratio above 1 says the step moves more bytes than purely "
                                     "random traffic could, i.e. part of it streams"}
        line = base_line(workload, world, args.steps, args.warmup, args.k_range, S)
        line.update({
            "value": positions_all * args.steps / (ms_value * 1e-3),
            "ms_per_step": ms_value / args.steps,
            "e2e": {"value": positions_all * args.steps / (ms_e2e * 1e-3), "unit": "positions/s",
                    "h2d_bytes_per_step": int(sum(ln for _, ln, _ in mine)),
                    "d2h_bytes_per_step": int(positions_per_step), "ms_per_step": ms_e2e / args.steps,
                    "timing": f"host clock bracketed by device synchronize; {depth} pipelined workspaces"},
            "gpu_launches": int(sum(s["launches"] for s in stats)),
            "clocks": clocks.summary(),
            "roofline": {"bound": "hbm", "kernel": dom_name,
                         "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": (achieved / hbm_peak) if achieved else None, "peak_source": peak_src,
                         "traffic": traffic,
                         "algorithmic_bytes_per_unit": "64 B per LF step (2 rank queries x one 32 B sector) + 32 B per "
                                                       "seed-table lookup (one sector)",
                         "algorithmic_bytes_per_launch": bytes_dom / nsearch,
                         "lf_steps_per_launch": dom_lf / nsearch, "seed_lookups_per_launch": dom_seeds / nsearch,
                         "ms_per_launch": dom_ms / nsearch,
                         "timing": "CUDA events on the workspace stream, one search at a time (no overlap)",
                         "lf_steps_note": "the certificate pass reads neighbour groups' values while they are being "
                                          "written (results do not depend on it, the number of chains run does): LF-step "
                                          "counts vary by a fraction of a percent between runs",
                         "whole_search": {"kernels": "pass 1 + k_free_cert + k_free_chain (pass 3)",
                                          "achieved": bytes_all / (ms_search * 1e-3) / 1e9 if ms_search else None,
                                          "ms_per_segment": ms_search / nsearch,
                                          "lf_steps_per_position": lf / max(1, iso_positions),
                                          "seed_lookups_per_position": seeds / max(1, iso_positions),
                                          "lf_steps_pass1_per_position": lf1 / max(1, iso_positions),
                                          "seed_lookups_pass1_per_position": seeds1 / max(1, iso_positions)},
                         "whole_step": whole_step,
                         # the same figure for each kernel of the search that walks the index (one search at a time)
                         "per_kernel": {
                             name: {"ms_per_launch": ms_k / nsearch, "algorithmic_bytes_per_launch": b_k / nsearch,
                                    "achieved": (b_k / (ms_k * 1e-3) / 1e9) if ms_k else None,
                                    "frac": (b_k / (ms_k * 1e-3) / 1e9 / hbm_peak) if ms_k else None}
                             for name, ms_k, b_k in (
                                 ("k_pass1_sparse" if deep else "k_free_chain(first)", ms_pass1, lf1 * 64 + seeds1 * 32),
                                 ("k_free_cert", ms_cert, ((lf - lf1) * 64 + (seeds - seeds1) * 32) * ms_cert / max(ms_cert + ms_pass3, 1e-9)),
                                 ("k_free_chain", ms_pass3, ((lf - lf1) * 64 + (seeds - seeds1) * 32) * ms_pass3 / max(ms_cert + ms_pass3, 1e-9)))},
                         "seed_depth": iso[0]["seed_depth"] if iso else None,
                         "ref_trajectory_lf_steps_per_position": st_traj.lf_steps / max(1, n0),
                         "random_sector_probe": {"dependent_loads_per_s": probe, "independent_loads_per_s": probe_indep,
                                                 "gbs_32B_independent": probe_indep[32] * 32 / 1e9,
                                                 "how": "2 GiB buffer; best of {8 Mi threads x 1, 2 Mi x 4} dependent chains "
                                                        "and {2 Mi x 8, 8 Mi x 4} independent loads, 32 steps"},
                         "frac_random_sector_algorithmic": ((bytes_dom / 32) / (dom_ms * 1e-3) / probe_indep[32]) if dom_ms else None},
            "kernel_ms": kernel_ms,
            "trajectory_kernel": {"positions": n0, "ms": st_traj.ms_search, "lf_steps": st_traj.lf_steps,
                                  "queries": st_traj.queries, "positions_per_s": n0 / (st_traj.ms_search * 1e-3)},
            "index_resident": {"open_s": index_open_s, "hbm_gb": index.info.device_bytes / 1e9,
                               "derived": f"rank32 layout + both-strand seed table of depth {index.derived_seed_length}, "
                                          "derived on the device at load"},
            "resident_pool": f"{len(pool)} workspaces, one stream each; {len(windows)} segments in the genome",
        })
        if job is not None:
            line["job"] = job
        if not args.no_cpu_baseline and not args.k_range:
            # the reference on a bounded sample of THIS genome: baseline figure AND byte-for-byte parity check
            sample = args.cpu_sample if world == 1 else min(args.cpu_sample, 200_000)
            r = cpu_reference_run(fa, ix_path, 1, 0, sample)
            par = parity_sample(index, r, sample)
            line["parity_sample"] = par
            if par["mismatches"]:
                raise SystemExit(f"bench: GPU output differs from the reference on the bench genome: {par}")
            if world == 1:
                line["cpu_baseline"] = {"value": r["value"], "unit": "positions/s", "cores": r["cores"],
                                        "kind": "reference", "sample": r["sample"], "native_share": r["native_share"],
                                        "value_amortised": r["value_amortised"],
                                        "index_reload_s_per_call": r["index_reload_s_per_call"]}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    barrier()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
