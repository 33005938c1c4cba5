#!/usr/bin/env python3
"""Summarise ncu output for profiles/: (1) a launch list (`--metrics gpu__time_duration.sum --csv`) into per-kernel
count / total / share; (2) a `--set full` report (.ncu-rep, via `ncu -i ... --page raw --csv`) into one line per
captured launch with the counters the roofline discussion uses."""
import csv
import subprocess
import sys
from collections import OrderedDict


def launch_list(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.reader(lines)
    hdr = next(rd)
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rd:
        if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        v_us = v / 1e3 if unit.startswith("ns") else v * (1e3 if unit.startswith("ms") else 1.0)
        rows.append((r[ix["Kernel Name"]].split("(")[0], v_us))
    agg = OrderedDict()
    for k, v in rows:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(rows)} launches, {tot / 1e3:.3f} ms total device time (cold-cache, serialised under ncu: compare shares)")
    print("kernel,launches,total_us,avg_us,share")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k},{n},{t:.1f},{t / n:.2f},{t / tot:.4f}")


def full_report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(out.splitlines()))
    hdr = rd[0]
    ix = {h: i for i, h in enumerate(hdr)}
    want = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active"]
    want = [w for w in want if w in ix]
    print(",".join(want))
    print(",".join(rd[1][ix[w]] for w in want))
    for r in rd[2:]:
        print(",".join(r[ix[w]].split("(")[0] if w == "Kernel Name" else r[ix[w]] for w in want))


if __name__ == "__main__":
    {"launches": launch_list, "full": full_report}[sys.argv[1]](sys.argv[2])
