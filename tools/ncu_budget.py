"""Per-kernel DRAM / instruction budget of one search from an `ncu --metrics ... --csv` log (last launch of each kernel):
    python tools/ncu_budget.py gpurun_out/budget.csv > profiles/budget.md"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ik, im, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
per = collections.OrderedDict()
for r in rows:
    if r is hdr or r[ik] == "Kernel Name":
        continue
    per.setdefault((r[iid], r[ik].split("(")[0].split("::")[-1]), {})[r[im]] = float(r[iv].replace(",", ""))
last = {}
for (_, name), m in per.items():
    last[name] = m
print("| kernel | time us | DRAM sectors read M | written M | warp instructions M | L2 hit % | warps active % |")
print("|---|---|---|---|---|---|---|")
tot_s = tot_t = 0.0
for name in ("k_prep", "k_pass1_sparse", "k_free_cert", "k_free_chain", "k_resolve", "k_subdepth", "k_presence"):
    m = last.get(name)
    if not m:
        continue
    rd, wr = m.get("dram__sectors_read.sum", 0), m.get("dram__sectors_write.sum", 0)
    tot_s += rd + wr
    tot_t += m["gpu__time_duration.sum"]
    print(f"| {name} | {m['gpu__time_duration.sum'] / 1e3:.1f} | {rd / 1e6:.2f} | {wr / 1e6:.2f} | "
          f"{m.get('smsp__inst_executed.sum', 0) / 1e6:.1f} | {m.get('lts__t_sector_hit_rate.pct', float('nan')):.1f} | "
          f"{m.get('sm__warps_active.avg.pct_of_peak_sustained_active', float('nan')):.1f} |")
print(f"\ntotal {tot_t / 1e3:.0f} us stand-alone; {tot_s / 1e6:.1f} M DRAM sectors = {tot_s / 2e6:.1f} M 64-byte accesses "
      f"-> {tot_s / 2 / 48e9 * 1e3:.3f} ms at the measured 48 G random accesses/s")
