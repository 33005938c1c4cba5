#!/usr/bin/env python3
"""Per-shape GPU time of the GEMM / attention launches of one graph-replayed denoise step (CUPTI timestamps matched to
the plan's call list by launch order)."""
import collections
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import b200sd  # noqa: E402
import bench  # noqa: E402

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
        rows = p.nimg * p.h * p.w_
        n_out = p.n // 2 if p.act == b200sd.lib.ACT_GEGLU else p.n
        # algorithmic bytes: every input pixel row once (stride 2 reads a 4x larger input), the weights once, the output
        # once (4 bytes when fp32), the residual once
        by = 2.0 * rows * (p.c0 + p.c1) * (4 if p.stride == 2 else 1) + 2.0 * p.n * taps * (p.c0 + p.c1) \
            + rows * n_out * (4.0 if p.out_dtype == b200sd.lib.F32 else 2.0) + (2.0 * rows * n_out if p.residual else 0.0)
    elif c.name == "b200sd_attention":
        key = f"attn nq={p.nq} nk={p.nk} d={p.d}"
        fl = 4.0 * p.batch * p.heads * p.nq * p.nk * p.d
        by = 2.0 * p.batch * p.heads * p.d * (2 * p.nq + 2 * p.nk)
    else:
        key, fl = c.name.replace("b200sd_", ""), 0.0
        by = bench.algorithmic_bytes(c)
    a = agg.setdefault(key, [0, 0.0, 0.0, 0.0])
    a[3] += by
    a[0] += 1
    a[1] += max(0.0, e.time_range.end - max(e.time_range.start, frontier))
    frontier = max(frontier, e.time_range.end)
    a[2] += fl
tot = sum(a[1] for a in agg.values())
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
TF, BW = peaks["bf16_tflops_sustained"] * 1e12, peaks["hbm_gbs"] * 1e9
print(f"one denoise step: {tot / 1e3:.3f} ms kernel time; per-shape roof = max(flops / {TF / 1e12:.1f} TFLOP/s, "
      f"algorithmic bytes / {BW / 1e9:.1f} GB/s) from MEASURED_PEAKS.json (bytes: inputs, weights, output, residual once)")
roof_tot = {}
for k, (n, t, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    roof = max(fl / TF, by / BW) * 1e6  # us
    fam = k.split()[0]
    r = roof_tot.setdefault(fam, [0.0, 0.0])
    r[0] += roof
    r[1] += t
for k, (n, t, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    roof = max(fl / TF, by / BW) * 1e6
    bound = "tensor" if fl / TF >= by / BW else "hbm"
    print(f"{k:52s} x{n:3d} {t / 1e3:7.3f} ms {100 * t / tot:5.1f}%  {fl / t / 1e6 if fl else 0:7.0f} TF/s "
          f"{by / t / 1e3 if by else 0:7.0f} GB/s  {fl / by if by and fl else 0:6.0f} flop/B  {bound:6s} {roof / t if t else 0:5.2f} of roof")
for fam, (roof, t) in sorted(roof_tot.items(), key=lambda kv: -kv[1][1]):
    print(f"family {fam:28s} {t / 1e3:7.3f} ms   roof {roof / 1e3:7.3f} ms   {roof / t if t else 0:5.2f} of the per-shape roofline")
print(f"whole step: roof {sum(r[0] for r in roof_tot.values()) / 1e3:.3f} ms / {tot / 1e3:.3f} ms = "
      f"{sum(r[0] for r in roof_tot.values()) / tot:.2f}")
