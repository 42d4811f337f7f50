#!/usr/bin/env python
"""bench.py -- genome positions/s of the `newmap search` hot path (min-unique-k, k=20:200).

    python bench.py --gpus N --steps K --warmup W [--workload c2_100m|c3_1g|c4_3g] [--impl reference]

One process per GPU (torchrun for N > 1).  The index is replicated on every GPU, positions are
sharded by 10 M-position segment (the reference's own batching unit), there is no data-path
collective: torch.distributed is used only for the barrier and the max-over-ranks of the time.

A step = one pass of the hot path over this rank's `segments_per_step` segments.
  value : positions/s with the segments already resident in HBM (CUDA events around K steps)
  e2e   : the same through the C-ABI with pinned HOST buffers, H2D + kernels + D2H every step,
          pipelined over a few workspaces (what newmap_b200.search.write_unique_counts does)
  roofline : dominant kernel (k_free_chain, pass 1) -- (LF steps x 64 B + seed lookups x 32 B) / its CUDA-event time
          measured with one search on the GPU at a time, against
          the measured streaming HBM peak (MEASURED_PEAKS.json) and against a random 32 B-sector
          probe measured in the same run
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
}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# --------------------------------------------------------------------------------------------
# workload: synthetic genome + reference-built index, cached on local disk
# --------------------------------------------------------------------------------------------
def prepare_workload(name: str) -> tuple[Path, Path]:
    """FASTA + .awfmi for the workload; built once per box (the index by the reference's own
    builder, oracle/_ref -- index construction is outside the hot path)."""
    from newmap_b200 import synth
    d = CACHE / name
    fa, ix, done = d / "genome.fa", d / "genome.awfmi", d / ".done"
    if done.is_file():
        return fa, ix
    d.mkdir(parents=True, exist_ok=True)
    lengths, seed = synth.config_lengths(name)
    t0 = time.time()
    synth.write_fasta(fa, synth.make_genome(lengths, seed))
    log(f"[bench] genome {name}: {sum(lengths)} bp written in {time.time() - t0:.1f}s")
    builder = os.environ.get("NEWMAP_B200_INDEX_BUILDER", "auto")
    t0 = time.time()
    built = None
    if builder in ("auto", "gpu"):
        # index construction is setup, not the timed path: use the GPU builder (writes the reference builder's
        # file byte for byte -- tests/test_gpu_build.py) when a CUDA device is there, else the reference's own
        try:
            from newmap_b200 import _lib, index
            if _lib.load().nm_device_count() > 0:
                secs = index.generate_fm_index(fa, ix, SA_RATIO, SEED_LEN,
                                               device=int(os.environ.get("LOCAL_RANK", "0")))
                built = f"GPU builder (suffix sort {secs:.1f}s on device)"
        except Exception as e:     # fall through to the reference builder
            if builder == "gpu":
                raise
            log(f"[bench] GPU index build failed ({e}); using the reference builder")
    if built is None:
        from oracle import ora   # the reference's index builder lives in oracle/_ref (not on the timed path)
        if not ora.ref_available():
            raise SystemExit("bench: oracle/_ref (reference index builder) is missing; run __graft_entry__.build() "
                             "where /root/reference exists")
        ora.ref_build_index(fa, ix, SA_RATIO, SEED_LEN)
        built = "reference builder (oracle/_ref)"
    log(f"[bench] index built by {built} in {time.time() - t0:.1f}s ({ix.stat().st_size / 1e6:.0f} MB)")
    (d / "builder.txt").write_text(built)
    done.write_text("ok")
    return fa, ix


def load_segments(fa: Path, first: int, count: int) -> list:
    """`count` segments of the genome starting at segment number `first` (cyclic)."""
    from newmap_b200.fasta import sequence_segments
    from newmap_b200.search import get_num_kmers
    max_k = K_RANGE[1]
    with open(fa, "rb") as fh:
        segs = [(s, get_num_kmers(s, max_k)) for s in sequence_segments(fh, BATCH + max_k - 1, max_k - 1)]
    total = len(segs)
    return [segs[(first + j) % total] for j in range(count)], total


# --------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.rows, self._stop, self._t = device, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.device)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self) -> dict:
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# CPU reference arm
# --------------------------------------------------------------------------------------------
def cpu_reference_run(fa: Path, ix: Path, steps: int, warmup: int, sample_positions: int) -> dict:
    """The reference's AVX2/OpenMP count path (oracle/_ref) under the reference's host loop on a
    bounded sample: the first `sample_positions` start positions of the second segment."""
    from oracle import ora, refdrive
    if not ora.ref_available():
        raise SystemExit("bench: oracle/_ref missing")
    (segs, _total) = load_segments(fa, 1, 1)
    seg, _ = segs[0]
    max_k = K_RANGE[1]
    data = seg.data[:sample_positions + max_k - 1]
    cores = min(os.cpu_count() or 1, 255)
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    timer = refdrive.Timer()
    for _ in range(warmup):
        refdrive.search_segment(ix, data, sample_positions, kl, True, 0, True, cores)
    t0 = time.perf_counter()
    for _ in range(steps):
        out = refdrive.search_segment(ix, data, sample_positions, kl, True, 0, True, cores, timer)
    dt = time.perf_counter() - t0
    # Every count call of the reference re-reads the whole index from disk (src/newmap-count.c:9-17,136).
    # In a real run that cost is spread over a 10 M-position batch, not over this sample: measure it with a
    # one-query call and charge the sample only its share, so the baseline is not understated.
    mod = ora.ref_modules()[0]
    reload_s = min(_timed(lambda: mod.count_kmers_from_sequence(str(ix), data[:64], [0], [20], cores)) for _ in range(3))
    share = min(1.0, sample_positions / BATCH)
    dt_amortised = max(dt - timer.calls * reload_s * (1.0 - share), 1e-9)
    return {"value": sample_positions * steps / dt_amortised, "value_raw_sample": sample_positions * steps / dt,
            "seconds": dt, "seconds_amortised": dt_amortised, "cores": cores,
            "index_reload_s_per_call": reload_s, "native_share": timer.native / dt if dt else None,
            "sample": f"{sample_positions} positions of segment 1 x {steps} step(s), k=20:200, "
                      f"{timer.calls} count_kmers_from_sequence calls; per-call index re-read ({reload_s:.2f}s) "
                      f"charged at the sample's share of a 10M batch",
            "checksum_input": int(np.asarray(out, dtype=np.uint64).sum())}


def file_to_file_run(fa: Path, ix: Path, device: int) -> dict:
    """newmap_b200.search.write_unique_counts over the whole workload genome, index already resident (the reference
    re-reads its index on every count call; here it is opened once per process)."""
    import shutil
    from newmap_b200.search import SearchConfig, write_unique_counts
    out = CACHE / "file_to_file_out"
    best = None
    for _ in range(2):
        shutil.rmtree(out, ignore_errors=True)
        out.mkdir(parents=True)
        cfg = SearchConfig(fasta_filepaths=[fa], fmindex_filepaths=[ix], kmer_lengths=list(range(K_RANGE[0], K_RANGE[1] + 1)),
                           is_binary_search=True, kmer_batch_size=BATCH, output_directory=out, device=device)
        dt = _timed(lambda: write_unique_counts(cfg))
        best = dt if best is None else min(best, dt)
    positions = sum(p.stat().st_size for p in out.glob("*.unique.uint8"))
    shutil.rmtree(out, ignore_errors=True)
    return {"value": positions / best, "unit": "positions/s", "positions": positions, "seconds": best,
            "what": "write_unique_counts: FASTA text read and parsed on the GPU, every segment searched, "
                    "<id>.unique.uint8 written (page cache); best of 2"}


def _timed(fn) -> float:
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


# --------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("NEWMAP_B200_WORKLOAD", "c3_1g"), choices=list(WORKLOADS))
    ap.add_argument("--segments-per-step", type=int, default=10)
    ap.add_argument("--cpu-sample", type=int, default=int(os.environ.get("NEWMAP_B200_CPU_SAMPLE", "1000000")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--k-range", default=None, help="other search range than the metric's 20:200, e.g. 12:255 "
                    "(BASELINE configs[4]); the JSON line then says so in config.k_range")
    ap.add_argument("--resident-streams", type=int, default=int(os.environ.get("NEWMAP_B200_RESIDENT_STREAMS", "0")),
                    help="CUDA streams the device-resident arm deals its workspaces over (0 = every workspace its own stream)")
    ap.add_argument("--e2e-depth", type=int, default=int(os.environ.get("NEWMAP_B200_E2E_DEPTH", "4")),
                    help="workspaces (streams) the end-to-end arm pipelines over")
    args = ap.parse_args()

    global K_RANGE
    if args.k_range:
        K_RANGE = tuple(int(x) for x in args.k_range.split(":"))
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    base = {"metric": "genome positions/sec, min-unique-k search k=20:200", "unit": "positions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload], "workload_id": args.workload,
                       "kmer_batch_size": BATCH, "index": f".awfmi v8 (reference format), seed {SEED_LEN}, SA ratio {SA_RATIO}",
                       "parallelism": f"index replicated x{world}, segments sharded, no collective"}}
    if args.k_range:
        base["metric"] = f"genome positions/sec, min-unique-k search k={args.k_range}"
        base["config"]["k_range"] = args.k_range
        base["config"]["workload"] = base["config"]["workload"].replace("k=20:200", f"k={args.k_range}")

    # ---------------- reference arm: CPU only, rank 0 only
    if args.impl == "reference":
        if rank != 0:
            return
        fa, ix = prepare_workload(args.workload)
        r = cpu_reference_run(fa, ix, max(1, args.steps), min(args.warmup, 1), args.cpu_sample)
        line = dict(base, impl="reference", value=r["value"], n_gpus=world,
                    ms_per_step=1e3 * r["seconds_amortised"] / max(1, args.steps),
                    cpu_baseline={"value": r["value"], "unit": "positions/s", "cores": r["cores"], "kind": "reference",
                                  "sample": r["sample"], "native_share": r["native_share"],
                                  "value_raw_sample": r["value_raw_sample"],
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

    if rank == 0:
        fa, ix_path = prepare_workload(args.workload)
    barrier()
    fa, ix_path = prepare_workload(args.workload)

    from newmap_b200 import _lib
    builder_note = (CACHE / args.workload / "builder.txt")
    base["config"]["index"] += "; built by " + (builder_note.read_text() if builder_note.is_file() else "unknown")
    kl = list(range(K_RANGE[0], K_RANGE[1] + 1))
    S = args.segments_per_step
    (segs, total_segments) = load_segments(fa, rank * S, S)
    t0 = time.time()
    index = _lib.Index(ix_path, device=local)
    log(f"[bench r{rank}] index resident in {time.time() - t0:.1f}s: bwt {index.bwt_length}, "
        f"{index.info.device_bytes / 1e9:.2f} GB HBM (derived seed table depth {index.derived_seed_length}); "
        f"{S} of {total_segments} segments per step")
    base["config"]["index"] += (f"; derived on the device at load: rank32 layout, seed table of depth "
                                f"{index.derived_seed_length} ({index.info.device_bytes / 1e9:.1f} GB HBM in all)")
    positions_per_step = sum(n for _, n in segs)
    max_len = max(len(s.data) for s, _ in segs)
    stream = torch.cuda.current_stream()

    # ---- device-resident arm: one workspace per segment.  --resident-streams 0 (default): every workspace on the
    # stream the library creates for it (kernels of different segments overlap); N >= 1: dealt over N torch streams
    # (1 = the kernels of a step strictly one after the other)
    resident = []
    pinned_in = []
    side = [torch.cuda.Stream() for _ in range(args.resident_streams)]
    for j, (seg, n) in enumerate(segs):
        ws = _lib.Workspace(index, len(seg.data), cuda_stream=side[j % len(side)].cuda_stream if side else None)
        pb = _lib.PinnedBuffer(len(seg.data))
        pb.array[:] = np.frombuffer(seg.data, dtype=np.uint8)
        ws.upload(pb.ptr, len(seg.data))
        resident.append((ws, n))
        pinned_in.append(pb)
    torch.cuda.synchronize()

    def step_resident(collect=None):
        for ws, n in resident:
            ws.search(n, kl, True, algo=args.algo, out_width=1)
        for ws, n in resident:
            st = ws.sync()
            if st.replayed or st.zero_pos >= 0:
                raise SystemExit("bench: a k-mer of the genome is absent from the index (index / FASTA mismatch)")
            if collect is not None:
                collect.append(st.as_dict())

    for _ in range(max(args.warmup, 3)):
        step_resident()
    checks_resident = [ws.checksum() for ws, _ in resident]
    stats: list = []
    barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        ev0.record(stream)
        for s_ in side:            # fork: the side streams start after ev0 ...
            s_.wait_event(ev0)
        for _ in range(args.steps):
            step_resident(stats)
        for s_ in side:            # ... and ev1 is recorded after all of them (step_resident has synchronised each)
            stream.wait_stream(s_)
        ev1.record(stream)
        torch.cuda.synchronize()
        barrier()
    ms_value = ev0.elapsed_time(ev1)

    # ---- end-to-end arm: pinned host in, pinned host out, 3 workspaces with their own streams
    depth = max(1, args.e2e_depth)
    slots = [(_lib.Workspace(index, max_len), _lib.PinnedBuffer(max_len)) for _ in range(depth)]

    def step_e2e(verify=False):
        pending = []
        oks = []
        for j, (pb, (seg, n)) in enumerate(zip(pinned_in, segs)):
            ws, outb = slots[j % depth]
            if len(pending) >= depth:
                w, ob, nn, jj = pending.pop(0)
                w.sync()
                if verify:
                    oks.append(_lib.checksum_host(ob.array[:nn]) == checks_resident[jj])
            ws.upload(pb.ptr, len(seg.data))
            ws.search(n, kl, True, algo=args.algo, out_width=1)
            ws.download(outb.ptr)
            pending.append((ws, outb, n, j))
        for w, ob, nn, jj in pending:
            w.sync()
            if verify:
                oks.append(_lib.checksum_host(ob.array[:nn]) == checks_resident[jj])
        return oks

    oks = step_e2e(verify=True)
    if not all(oks):
        raise SystemExit("bench: e2e output differs from the device-resident output")
    for _ in range(2):
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
    for _ in range(args.steps):
        for ws, n in resident:
            ws.search(n, kl, True, algo=args.algo, out_width=1)
            iso.append(ws.sync().as_dict())

    # ---- cross-check one segment against the reference-trajectory kernel (different algorithm, same bytes)
    ws0, n0 = resident[0]
    ws0.search(n0, kl, True, algo="trajectory", out_width=1)
    st_traj = ws0.sync()
    if ws0.checksum() != checks_resident[0]:
        raise SystemExit("bench: trajectory-free output differs from the reference-trajectory kernel")

    # ---- max over ranks
    if world > 1:
        t = torch.tensor([ms_value, ms_e2e], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_value, ms_e2e = t.tolist()
        agg = torch.tensor([float(positions_per_step)], device="cuda", dtype=torch.float64)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        positions_all = agg.item()
    else:
        positions_all = float(positions_per_step)

    if rank == 0:
        lf = sum(s["lf_steps"] for s in iso)
        ms_search = sum(s["ms_search"] for s in iso)
        launches = len(iso)
        lf1 = sum(s["lf_steps_pass1"] for s in iso)
        seeds = sum(s["seed_lookups"] for s in iso)
        seeds1 = sum(s["seed_lookups_pass1"] for s in iso)
        ms_pass1 = sum(s["ms_pass1"] for s in iso)
        bytes1 = lf1 * 64 + seeds1 * 32
        bytes_all = lf * 64 + seeds * 32
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"     # dram__bytes_read+write per launch from an `ncu --set full` capture
        if tf.is_file():
            traffic = json.loads(tf.read_text()).get(args.workload, {}).get("k_free_chain_pass1")
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
        achieved = bytes1 / (ms_pass1 * 1e-3) / 1e9 if ms_pass1 else None
        line = dict(base)
        line.update({
            "value": positions_all * args.steps / (ms_value * 1e-3),
            "ms_per_step": ms_value / args.steps,
            "e2e": {"value": positions_all * args.steps / (ms_e2e * 1e-3), "unit": "positions/s",
                    "h2d_bytes_per_step": int(sum(len(s.data) for s, _ in segs)),
                    "d2h_bytes_per_step": int(positions_per_step), "ms_per_step": ms_e2e / args.steps,
                    "timing": f"host clock bracketed by device synchronize; {depth} pipelined workspaces"},
            "gpu_launches": int(sum(s["launches"] for s in stats)),
            "clocks": clocks.summary(),
            "roofline": {"bound": "hbm", "kernel": ("k_pass1_tiled" if iso and iso[0]["seed_depth"] >= 13 else "k_free_chain") +
                                   " (pass 1 of the trajectory-free search: every position, depth-capped)",
                         "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": (achieved / hbm_peak) if achieved else None, "peak_source": peak_src,
                         "traffic": traffic,
                         "algorithmic_bytes_per_unit": "64 B per LF step (2 rank queries x one 32 B sector) + 32 B per "
                                                       "seed-table lookup (one sector)",
                         "algorithmic_bytes_per_launch": bytes1 / max(1, launches),
                         "lf_steps_per_launch": lf1 / max(1, launches), "seed_lookups_per_launch": seeds1 / max(1, launches),
                         "ms_per_launch": ms_pass1 / max(1, launches),
                         "timing": "CUDA events on the workspace stream, one search at a time (no overlap)",
                         "whole_search": {"kernels": "pass 1 + k_free_cert + k_free_chain (pass 3)",
                                          "achieved": bytes_all / (ms_search * 1e-3) / 1e9 if ms_search else None,
                                          "ms_per_segment": ms_search / max(1, launches),
                                          "lf_steps_per_position": lf / max(1, positions_per_step * args.steps),
                                          "seed_lookups_per_position": seeds / max(1, positions_per_step * args.steps)},
                         "seed_depth": iso[0]["seed_depth"] if iso else None,
                         "ref_trajectory_lf_steps_per_position": st_traj.lf_steps / max(1, n0),
                         "random_sector_probe": {"dependent_loads_per_s": probe, "independent_loads_per_s": probe_indep,
                                                 "gbs_32B_independent": probe_indep[32] * 32 / 1e9,
                                                 "how": "2 GiB buffer; best of {8 Mi threads x 1, 2 Mi x 4} dependent chains "
                                                        "and {2 Mi x 8, 8 Mi x 4} independent loads, 32 steps"},
                         "frac_random_sector": ((bytes1 / 32) / (ms_pass1 * 1e-3) / probe_indep[32]) if ms_pass1 else None},
            "kernel_ms": {k: sum(s[k] for s in iso) / max(1, launches) for k in
                          ("ms_prep", "ms_pass1", "ms_cert", "ms_pass3", "ms_generic", "ms_resolve")},
            "trajectory_kernel": {"positions": n0, "ms": st_traj.ms_search, "lf_steps": st_traj.lf_steps,
                                  "queries": st_traj.queries, "positions_per_s": n0 / (st_traj.ms_search * 1e-3)},
        })
        line["config"]["l2"] = "index (rank32 + seed table) larger than L2 except for c2_100m (rank32 50 MB); no flush"
        line["config"]["segments_per_step_per_gpu"] = S
        line["config"]["resident_streams"] = args.resident_streams or "one per workspace (%d)" % len(resident)
        if world == 1:
            # the whole `newmap search` run a user makes: FASTA file in, <id>.unique.uint8 files out (device ingest path)
            line["file_to_file"] = file_to_file_run(fa, ix_path, local)
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_run(fa, ix_path, 1, 0, args.cpu_sample)
            line["cpu_baseline"] = {"value": r["value"], "unit": "positions/s", "cores": r["cores"],
                                    "kind": "reference", "sample": r["sample"], "native_share": r["native_share"],
                                    "value_raw_sample": r["value_raw_sample"],
                                    "index_reload_s_per_call": r["index_reload_s_per_call"]}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    barrier()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
