#!/usr/bin/env python3
"""Print a compact, fixed set of counters from an .ncu-rep (one column per captured launch)."""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum.per_cycle_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__lts2xbar_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warp_latency_per_inst_issued.ratio"]
STALL = "smsp__average_warps_issue_stalled_"


def main(path, pattern=""):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(out.splitlines()))
    hdr, units, data = rd[0], rd[1], rd[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in data if pattern in r[ix["Kernel Name"]]]
    print("kernel:", [r[ix["Kernel Name"]].split("(")[0][-28:] for r in data])
    for k in KEYS:
        if k in ix:
            print(f"{k:82s} {units[ix[k]]:10s}", [r[ix[k]][:10] for r in data])
    for h in hdr:
        if h.startswith(STALL):
            vals = [r[ix[h]] for r in data]
            try:
                if max(float(v) for v in vals) >= 0.3:
                    print(f"stall {h[len(STALL):-len('_per_issue_active.ratio')]:30s}", [v[:6] for v in vals])
            except ValueError:
                pass


if __name__ == "__main__":
    main(*sys.argv[1:])
