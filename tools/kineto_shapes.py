#!/usr/bin/env python3
"""Per-shape GPU time of the GEMM / attention launches of one graph-replayed denoise step (CUPTI timestamps matched to
the plan's call list by launch order)."""
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import b200sd  # noqa: E402

# usage: kineto_shapes.py [model=sd15] [batch=8] [latent_side=64]   (e.g. sdxl 2 128)
model = sys.argv[1] if len(sys.argv) > 1 else "sd15"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
side = int(sys.argv[3]) if len(sys.argv) > 3 else 64
pipe = b200sd.StableDiffusionPipeline.from_pretrained(model, scheduler="euler" if "xl" in model else "ddim")
job = pipe.job(batch, side, side, 7.5, True, zero_negative=pipe.is_xl)
sched = pipe.make_schedule(25)
job.step_table[: sched.num_evals].copy_(sched.table)
job.init_latents.normal_()
job.reset()
for _ in range(3):
    job.step_plan.launch()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    job.step_plan.launch()
    torch.cuda.synchronize()
evs = sorted([e for e in prof.events() if e.device_type.name == "CUDA" and "emset" not in e.name],
             key=lambda e: e.time_range.start)
calls = [c for c in job.step_plan.calls if c.name != "b200sd_memset_zero"]
assert len(evs) == len(calls), (len(evs), len(calls))
agg = collections.OrderedDict()
frontier = evs[0].time_range.start  # kernels may become resident under their predecessor's tail: count each from the end of what ran before it
for e, c in zip(evs, calls):
    p = c.params
    if c.name == "b200sd_gemm":
        taps = p.kh * p.kw if p.kh else p.ksize ** 2
        key = (f"gemm M={p.nimg * p.h * p.w_} K={taps * (p.c0 + p.c1)} N={p.n} k{p.kh or p.ksize}x{p.kw or p.ksize}"
               f"{'s2' if p.stride == 2 else ''} act{p.act}{' res' if p.residual else ''}{' gn' if p.gn_stats else ''}")
        fl = 2.0 * p.nimg * p.h * p.w_ * p.n * taps * (p.c0 + p.c1)
    elif c.name == "b200sd_attention":
        key = f"attn nq={p.nq} nk={p.nk} d={p.d}"
        fl = 4.0 * p.batch * p.heads * p.nq * p.nk * p.d
    else:
        key, fl = c.name.replace("b200sd_", ""), 0.0
    a = agg.setdefault(key, [0, 0.0, 0.0])
    a[0] += 1
    a[1] += max(0.0, e.time_range.end - max(e.time_range.start, frontier))
    frontier = max(frontier, e.time_range.end)
    a[2] += fl
tot = sum(a[1] for a in agg.values())
print(f"one denoise step: {tot / 1e3:.3f} ms kernel time")
for k, (n, t, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    print(f"{k:52s} x{n:3d} {t / 1e3:7.3f} ms {100 * t / tot:5.1f}%  {fl / t / 1e6 if fl else 0:7.0f} TF/s")
