#!/usr/bin/env python
"""Turns gpurun_out/ ncu artefacts into the small text summaries kept under profiles/.

    python scripts/ncu_summary.py launches gpurun_out/launches.csv profiles/NAME.md "command line"
    python scripts/ncu_summary.py report   gpurun_out/X.ncu-rep    profiles/NAME.txt "title"
"""
import collections
import csv
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "sm__cycles_active.avg", "smsp__cycles_active.avg", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "sm__maximum_warps_per_active_cycle_pct"]


def launches(src, dst, cmd):
    lines = [l for l in open(src) if not l.startswith("==")]
    agg = collections.defaultdict(list)
    for row in csv.DictReader(lines):
        try:
            agg[row["Kernel Name"]].append(float(row["Metric Value"].replace(",", "")))
        except (ValueError, KeyError):
            continue
    tot = sum(sum(v) for v in agg.values())
    ours = [k for k in agg if "b200attn" in k]
    tot_ours = sum(sum(agg[k]) for k in ours) or 1.0
    with open(dst, "w") as f:
        f.write(f"# ncu launch list\n\nCommand: `{cmd}`\n(cold-cache, serialised launches: compare shares, not absolutes; "
                "`native::`/`at_cuda_detail::` kernels are torch building the synthetic inputs)\n\n")
        f.write("| kernel | launches | mean ns | share of all | share of b200attn kernels |\n|---|---|---|---|---|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            so = f"{sum(v) / tot_ours:.3f}" if k in ours else "-"
            f.write(f"| `{k[:100]}` | {len(v)} | {sum(v) / len(v):.0f} | {sum(v) / tot:.3f} | {so} |\n")


def report(src, dst, title):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write(f"{title}\nsource: {src} (scratch; ncu --set full --clock-control none --import-source on)\n")
        for r in rows[2:]:
            f.write(f"\n== {r[hdr.index('Kernel Name')][:110]}  (ID {r[0]})\n")
            for k in hdr:
                if k in KEEP or "pcsamp_warps_issue_stalled" in k and not k.endswith("_not_issued") or \
                        k.startswith("smsp__average_warps_issue_stalled") or k.startswith("sm__inst_executed_pipe_") and k.endswith(".sum"):
                    v = r[hdr.index(k)]
                    if v in ("", "0", "n/a"):
                        continue
                    f.write(f"{k:86s} {v:>18s} {units[hdr.index(k)]}\n")


if __name__ == "__main__":
    {"launches": launches, "report": report}[sys.argv[1]](*sys.argv[2:5])
