#!/usr/bin/env python3
"""Per-kernel GPU durations of ONE graph-replayed denoise step (SD1.5, batch 8) from CUPTI activity records
(torch.profiler): hardware timestamps, no instrumentation gaps.  Cross-check for bench.py's CUDA-event breakdown."""
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import b200sd  # noqa: E402

pipe = b200sd.StableDiffusionPipeline.from_pretrained("sd15", scheduler="ddim")
job = pipe.job(8, 64, 64, 7.5, True)
sched = pipe.make_schedule(25)
job.step_table[: sched.num_evals].copy_(sched.table)
job.init_latents.normal_()
job.reset()
for _ in range(3):
    job.step_plan.launch()
torch.cuda.synchronize()
what = sys.argv[1] if len(sys.argv) > 1 else "unet"
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    if what == "unet":
        for _ in range(3):
            job.step_plan.launch()
    else:
        job.vae_plan.run()
    torch.cuda.synchronize()
agg = collections.OrderedDict()
spans = []
for e in prof.events():
    if e.device_type.name != "CUDA":
        continue
    spans.append((e.time_range.start, e.time_range.end))
    name = e.name.split("(")[0].replace("void ", "").replace("b200sd::", "")[:40]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += e.device_time_total if hasattr(e, "device_time_total") else e.cuda_time_total
tot = sum(a[1] for a in agg.values())
print(f"{what}: {tot / 1e3:.3f} ms of kernel time ({'3 graph-replayed denoise steps' if what == 'unet' else '1 VAE decode of 8 images'})")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:42s} {n:5d} launches {t / 1e3:9.3f} ms  avg {t / n:8.1f} us")
spans.sort()
if spans:
    busy = sum(b - a for a, b in spans)
    wall = spans[-1][1] - spans[0][0]
    gaps = sorted(max(0.0, spans[i + 1][0] - spans[i][1]) for i in range(len(spans) - 1))
    print(f"span first start -> last end: {wall / 1e3:.3f} ms; kernels busy {busy / 1e3:.3f} ms; idle between kernels "
          f"{(wall - busy) / 1e3:.3f} ms over {len(gaps)} boundaries (median gap {gaps[len(gaps) // 2]:.2f} us, "
          f"p90 {gaps[int(len(gaps) * 0.9)]:.2f} us)")
