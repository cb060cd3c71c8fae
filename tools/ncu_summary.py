#!/usr/bin/env python
"""Summarise gpurun_out/launches.csv (ncu launch list) and a full .ncu-rep into markdown."""
import collections
import csv
import subprocess
import sys


def launches(path):
    with open(path) as fh:
        lines = [l for l in fh if not l.startswith("==")]
    agg = collections.defaultdict(list)
    for row in csv.DictReader(lines):
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except Exception:
            continue
        unit = row["Metric Unit"]
        v = v / 1000 if unit == "ns" else v * 1000 if unit == "ms" else v * 1e6 if unit == "s" else v
        agg[row["Kernel Name"].split("(")[0][:48]].append(v)
    tot = sum(sum(v) for v in agg.values())
    print("| kernel | launches | mean µs | share of GPU time |\n|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("| `%s` | %d | %.2f | %.1f %% |" % (k, len(v), sum(v) / len(v), 100 * sum(v) / tot))


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr, units = r[0], r[1]
    ki = hdr.index("Kernel Name")
    print("\n| metric | unit | " + " | ".join("launch %d" % i for i in range(len(r) - 2)) + " |")
    print("|---|---|" + "---|" * (len(r) - 2))
    print("| kernel | | " + " | ".join("`%s`" % row[ki].split("(")[0][:40] for row in r[2:]) + " |")
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print("| %s | %s | %s |" % (w, units[i], " | ".join(row[i] for row in r[2:])))


if __name__ == "__main__":
    launches(sys.argv[1])
    if len(sys.argv) > 2:
        full(sys.argv[2])
