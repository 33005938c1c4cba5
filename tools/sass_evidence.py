#!/usr/bin/env python3
"""Per-kernel SASS evidence of the built library: counts of the tcgen05 / TMEM / TMA mnemonics
(UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor loads / stores) and of the legacy
tensor path (HMMA) per kernel of libb200sd.so, from `cuobjdump -sass`.  Writes profiles/<tag>_sass_summary.txt and one
grep extract per tensor-core kernel."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
so = os.path.join(ROOT, "k8s-nvidia-gpus_b200", "libb200sd.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA(?:\.2CTA)?|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|HMMA|MUFU\.EX2|REDG|ATOMG|ATOMS)\b")
kern, counts, lines = None, collections.OrderedDict(), collections.defaultdict(list)
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern).replace("b200sd::", "")
        counts[kern] = collections.Counter()
        continue
    if kern:
        for mm in pat.findall(ln):
            counts[kern][mm] += 1
            if mm.startswith(("UTC", "LDTM", "STTM", "UTMA")):
                lines[kern].append(ln.strip())
out = [f"SASS mnemonic counts per kernel of {os.path.relpath(so, ROOT)} (cuobjdump -sass; CUDA 12.9, sm_100a)", ""]
for k, c in counts.items():
    out.append(f"{k}\n    " + ("  ".join(f"{n}={v}" for n, v in sorted(c.items())) or "(none of the listed mnemonics)"))
path = os.path.join(ROOT, "profiles", f"{tag}_sass_summary.txt")
open(path, "w").write("\n".join(out) + "\n")
keep = ("void gemm_conv_kernel<false, false, true, true>", "void gemm_conv_kernel<false, true, false, false>",
        "attention_kernel", "void xattn_kernel<5, 3>")
for k, ls in lines.items():
    if k not in keep:
        continue
    short = re.sub(r"[^A-Za-z0-9_]+", "_", k)[:60].strip("_")
    open(os.path.join(ROOT, "profiles", f"{tag}_sass_{short}.txt"), "w").write(
        f"{k}: tcgen05 / TMEM / TMA instructions (cuobjdump -sass | grep)\n" + "\n".join(ls) + "\n")
print(path)
