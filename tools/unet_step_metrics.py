#!/usr/bin/env python3
"""Per-kernel summary of ONE denoise step from an ncu metrics CSV, and the DRAM traffic per GEMM launch that
bench.py reports as roofline.traffic.

  ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,\\
      sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none --csv \\
      --log-file unet1_metrics.csv python tools/profile_step.py unet1
  python tools/unet_step_metrics.py unet1_metrics.csv profiles/r01_unet_step_ncu_metrics.csv profiles/roofline_traffic.json
"""
import collections
import csv
import json
import sys

BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
USEC = {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}
TENSOR = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"


def main(src, out_csv, out_json):
    lines = [l for l in open(src) if not l.startswith("==")]
    rd = csv.reader(lines)
    hdr = next(rd)
    ix = {h: i for i, h in enumerate(hdr)}
    per = collections.defaultdict(dict)
    for r in rd:
        if len(r) < len(hdr):
            continue
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("b200sd::", "").split("<")[0]
        per[(r[ix["ID"]], name)][r[ix["Metric Name"]]] = (float(r[ix["Metric Value"]].replace(",", "")), r[ix["Metric Unit"]])
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
    for (_, k), m in per.items():
        us = m["gpu__time_duration.sum"][0] * USEC.get(m["gpu__time_duration.sum"][1], 1)
        a = agg[k]
        a[0] += 1
        a[1] += us
        a[2] += m["dram__bytes_read.sum"][0] * BYTES.get(m["dram__bytes_read.sum"][1], 1)
        a[3] += m["dram__bytes_write.sum"][0] * BYTES.get(m["dram__bytes_write.sum"][1], 1)
        a[4] += m[TENSOR][0] * us
    tot = sum(a[1] for a in agg.values())
    rows = ["kernel,launches,total_us,share,dram_read_MB_per_launch,dram_write_MB_per_launch,tensor_active_pct_time_weighted"]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        rows.append(f"{k},{a[0]},{a[1]:.1f},{a[1] / tot:.4f},{a[2] / a[0] / 1e6:.3f},{a[3] / a[0] / 1e6:.3f},{a[4] / a[1]:.1f}")
    with open(out_csv, "w") as f:
        f.write("# one eager denoise step (SD1.5, batch 8 -> 16 samples); ncu per-launch metrics, cold-cache and serialised: "
                "compare shares, not absolutes\n" + "\n".join(rows) + "\n")
    g = agg["gemm_conv_kernel"]
    json.dump({"gemm_conv_kernel_dram_bytes_per_launch": (g[2] + g[3]) / g[0], "launches": g[0],
               "source": f"{out_csv} (ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over the GEMM launches "
                         "of one denoise step)"}, open(out_json, "w"), indent=1)
    print("\n".join(rows[:9]))


if __name__ == "__main__":
    main(*sys.argv[1:4])
