#!/usr/bin/env python3
"""Read the source page of an .ncu-rep (`ncu -i rep --page source --csv`) and print, per kernel, the SASS lines with
the most stall samples plus the stall-reason totals.  usage: ncu_src.py report.ncu-rep [top=40]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None:
        if cur["hdr"] is None:
            cur["hdr"] = row
        elif len(row) >= len(cur["hdr"]) - 2:
            cur["rows"].append(row)
for b in blocks:
    h = b["hdr"]
    ci = {n: i for i, n in enumerate(h)}
    samp = ci["# Samples"]
    stall_cols = [(n, i) for n, i in ci.items() if n.startswith("stall_") and "Not Issued" not in n]
    tot = sum(int(r[samp] or 0) for r in b["rows"])
    print(f"== {b['name'][:90]}  samples {tot}, {len(b['rows'])} SASS lines")
    sums = sorted(((sum(int(r[i] or 0) for r in b["rows"]), n) for n, i in stall_cols), reverse=True)
    print("   stall totals:", ", ".join(f"{n[6:]} {100 * v / max(1, tot):.1f}%" for v, n in sums[:8]))
    ex = ci["Instructions Executed"]
    print("   warp instructions executed:", sum(int(r[ex] or 0) for r in b["rows"]))
    rows = sorted(b["rows"], key=lambda r: -int(r[samp] or 0))[:top]
    for r in rows:
        reasons = sorted(((int(r[i] or 0), n[6:]) for n, i in stall_cols), reverse=True)[:2]
        print(f"   {100 * int(r[samp] or 0) / max(1, tot):5.1f}%  {r[ci['Address']][-5:]}  {r[ci['Source']][:70]:70s} "
              + " ".join(f"{n}:{v}" for v, n in reasons if v))
