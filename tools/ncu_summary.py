"""Summaries for profiles/: (a) `ncu_summary.py launches <csv>` -- per-kernel totals and shares of an ncu
gpu__time_duration launch list; (b) `ncu_summary.py full <ncu-rep>` -- one markdown row per captured launch."""
import csv
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__sectors_read.sum",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"]


def short(name):
    name = name.split("(")[0]
    return name.split("::")[-1]


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows:
        if r is hdr or r[ik] == "Kernel Name":
            continue
        try:
            v = float(r[iv].replace(",", ""))
        except ValueError:
            continue
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iu].strip().replace("usecond", "us").replace("nsecond", "ns").replace("msecond", "ms").replace("second", "s"), 1e-6)
        tot[short(r[ik])] += v * scale
        cnt[short(r[ik])] += 1
    total = sum(tot.values())
    print("| kernel | launches | total ms | mean ms | share |\n|---|---|---|---|---|")
    for k in sorted(tot, key=lambda k: -tot[k]):
        print(f"| {k} | {cnt[k]} | {tot[k]:.3f} | {tot[k] / cnt[k]:.4f} | {100 * tot[k] / total:.1f} % |")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    keys = [k for k in KEYS if k in hdr]
    print("| kernel | " + " | ".join(keys) + " |\n|---|" + "---|" * len(keys))
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        print(f"| {short(d['Kernel Name'])} | " + " | ".join(f"{d[k]} {u[k]}".strip() for k in keys) + " |")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
