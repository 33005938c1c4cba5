#!/usr/bin/env python3
"""bench.py — images/sec of the SD-class batch text-to-image hot path on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (fp32 oracle) on the host cores

One bench "step" = one full pass of the hot path over one batch of synthetic prompts: CLIP text encoding, cross-
attention K/V projection, 25 DDIM denoise steps (CFG 7.5, UNet on the duplicated batch) and VAE decode to uint8 images,
SD 1.5 architecture at 512x512, batch 8 per GPU (BASELINE.json configs[1]; configs[2] when N > 1: prompts sharded
data-parallel, one NCCL gather of the uint8 images per batch).  Weights are random-init (no checkpoints offline).

Prints ONE JSON line (rank 0).  `value`: inputs resident in HBM.  `e2e`: through the public pipeline call with host
prompts / host latents / host images, copies inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

UNET_GF_PER_SAMPLE = 803.27  # SURVEY.md §8(d), algorithmic 2*MAC of the SD1.5 UNet forward at 64x64 latent
VAE_GF_PER_IMAGE = 2514.5
CLIP_GF_PER_SEQ = 13.30


def synthetic_prompts(n, offset=0):
    subj = ["panda", "astronaut", "robot", "dragon", "cat", "horse", "castle", "cabin"]
    return [f"a photo of synthetic subject {i + offset:03d}, a {subj[(i + offset) % len(subj)]} riding a motorbike, "
            f"variant {(i + offset) % 7}" for i in range(n)]


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"tf_burst": d.get("bf16_tflops"), "tf_sustained": d.get("bf16_tflops_sustained"),
                "hbm_gbs": d.get("hbm_gbs"), "source": "measured"}
    return {"tf_burst": 1590.0, "tf_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for nme, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        try:
            os.unlink(self.path)
        except OSError:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------- reference arm
def build_oracle(cfg):
    """fp32 CPU oracle modules with their own default (random) init — same distributions as the GPU arm's weights."""
    from oracle.clip import CLIPTextModel
    from oracle.unet import UNet2DConditionModel
    from oracle.vae import AutoencoderKLDecoder
    torch.manual_seed(0)
    return (UNet2DConditionModel(cfg["unet"]).eval(), AutoencoderKLDecoder(cfg["vae"]).eval(),
            CLIPTextModel(cfg["text_encoders"][0]).eval())


@torch.no_grad()
def cpu_sample(cfg, unet, vae, clip, steps, warmup, inference_steps, h, w):
    """Bounded sample of the reference CPU path for ONE image: `steps` timed CFG-paired UNet evaluations (2 samples),
    plus one VAE decode and one 2-sequence CLIP pass; images/sec = 1 / (inference_steps * t_unet + t_vae + t_clip)."""
    ids = torch.randint(0, cfg["text_encoders"][0]["vocab_size"], (2, 77))
    t0 = time.perf_counter()
    ctx = clip(ids)
    t_clip = time.perf_counter() - t0
    x = torch.randn(2, 4, h, w)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        unet(x, torch.tensor([961.0]), ctx)
        times.append(time.perf_counter() - t0)
    t_unet = sum(times[warmup:]) / max(1, steps)
    t0 = time.perf_counter()
    vae.decode(torch.randn(1, 4, h, w))
    t_vae = time.perf_counter() - t0
    ips = 1.0 / (inference_steps * t_unet + t_vae + t_clip)
    return {"ips": ips, "t_unet_pair_s": t_unet, "t_vae_s": t_vae, "t_clip_s": t_clip}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import importlib
    Wt = importlib.import_module("k8s-nvidia-gpus_b200.weights")
    cfg = Wt.load_config(args.model)
    torch.set_num_threads(os.cpu_count() or 1)
    unet, vae, clip = build_oracle(cfg)
    f = 2 ** (len(cfg["vae"]["block_out_channels"]) - 1)
    h, w = args.height // f, args.width // f
    r = cpu_sample(cfg, unet, vae, clip, args.steps, args.warmup, args.inference_steps, h, w)
    pb, mode = resolve_batch(args, args.gpus)
    sample = (f"1 image: {args.steps} timed + {args.warmup} warm-up CFG-paired UNet evaluations (2 samples, "
              f"{h}x{w} latent), 1 VAE decode, 1 CLIP pass of 2 sequences; images/sec = 1/({args.inference_steps}"
              f"*t_unet + t_vae + t_clip)")
    line = {
        "impl": "reference", "metric": "images/sec", "value": r["ips"], "unit": "images/s", "n_gpus": args.gpus,
        # same unit as the CUDA arm: ms for one bench step = one whole batch of images (extrapolated from the sample)
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * pb * args.gpus / r["ips"],
        "sample_ms_per_unet_pair": r["t_unet_pair_s"] * 1e3,
        "higher_is_better": True, "scaling": mode, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, cfg, args.gpus),
        "cpu_baseline": {"value": r["ips"], "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": r["ips"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "detail": {k: round(v, 4) for k, v in r.items()},
    }
    emit(json.dumps(line))
    return 0


def resolve_batch(args, n_gpus):
    """Per-GPU batch and scaling mode.  One GPU: BASELINE.json configs[1] (batch 8).  Several GPUs: configs[2] as
    written - a FIXED 64-prompt batch sharded over the ranks (32 / 16 / 8 per GPU at N = 2 / 4 / 8; strong scaling) -
    unless --weak keeps --batch prompts per GPU.  --global-batch overrides the total."""
    if args.global_batch is None and (n_gpus == 1 or args.weak):
        return args.batch, "weak"
    g = args.global_batch if args.global_batch is not None else 64
    if g % n_gpus != 0:
        raise SystemExit(f"bench.py: global batch {g} is not divisible by {n_gpus} GPUs")
    return g // n_gpus, "strong"


def workload_config(args, cfg, n_gpus):
    pb, mode = resolve_batch(args, n_gpus)
    return {"workload": f"{cfg['name']} architecture (random-init), {args.width}x{args.height}, "
                        f"{args.inference_steps} {args.scheduler.upper()} steps, CFG {args.guidance_scale}, "
                        + (f"batch {pb} per GPU" if mode == "weak" else f"{pb * n_gpus}-prompt batch sharded over {n_gpus} GPU(s)"),
            "global_batch": pb * n_gpus, "per_gpu_batch": pb, "scheduler": args.scheduler,
            "inference_steps": args.inference_steps, "guidance_scale": args.guidance_scale,
            "resolution": [args.width, args.height], "parallelism": f"dp{n_gpus} (prompt sharding, image gather)",
            "l2": "inputs larger than L2: each step streams ~2 GB of weights plus activations through a 126 MB L2; "
                  "no explicit flush"}


# ---------------------------------------------------------------------------------------------------- CUDA arm
def flops_of_call(call):
    p = call.params
    if call.name == "b200sd_gemm":
        m = p.nimg * p.h * p.w_
        taps = p.kh * p.kw if p.kh else p.ksize * p.ksize  # MACs actually executed (folded up-convs: 4 x 2x2 taps)
        return 2.0 * m * p.n * taps * (p.c0 + p.c1), "gemm"
    if call.name == "b200sd_attention":
        return 4.0 * p.batch * p.heads * p.nq * p.nk * p.d, "attention"
    return 0.0, call.name.replace("b200sd_", "")


def profile_plan(plan, blocker=None):
    """Per-launch device time with CUDA events on the launching stream.  A long-running `blocker` plan is enqueued
    first so the host runs ahead of the GPU and launch latency does not leak into the per-kernel intervals; the
    event-to-event gap of an EMPTY interval (measured the same way, right after) is subtracted from every interval."""
    from b200sd import ops
    s = ops.current_stream()
    n = len(plan.calls)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    cal = [torch.cuda.Event(enable_timing=True) for _ in range(65)]
    if blocker is not None:
        for _ in range(8):  # ~0.2 s of queued GPU work: the host stays ahead for the whole instrumented pass
            blocker.run()
    ev[0].record()
    for i, c in enumerate(plan.calls):
        c(s)
        ev[i + 1].record()
    for e in cal:
        e.record()
    torch.cuda.synchronize()
    gaps = sorted(cal[i].elapsed_time(cal[i + 1]) for i in range(64))
    gap = gaps[len(gaps) // 2]
    agg = {}
    for i, c in enumerate(plan.calls):
        fl, kind = flops_of_call(c)
        a = agg.setdefault(kind, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
        a["launches"] += 1
        a["ms"] += max(0.0, ev[i].elapsed_time(ev[i + 1]) - gap)
        a["flops"] += fl
        a["bytes"] += algorithmic_bytes(c)
    agg["_event_gap_ms"] = gap
    return agg


def algorithmic_bytes(call):
    """Algorithmic HBM bytes of the bandwidth-bound kernels (16-bit, each element read / written once
    per pass)."""
    p = call.params
    if call.name == "b200sd_groupnorm_stats":
        return 2.0 * p.nimg * p.hw * (p.c0 + p.c1)
    if call.name == "b200sd_groupnorm_apply":
        return 4.0 * p.nimg * p.hw * (p.c0 + p.c1)
    if call.name == "b200sd_layernorm":
        return 4.0 * p.rows * p.c
    return 0.0


def exclusive_us(spans):
    """spans: (start, end) of kernels sorted by start.  Returns each kernel's time counted from the end of everything
    launched before it: a kernel launched with programmatic dependent launch becomes resident (and its record starts)
    while its predecessor is still running and then waits for it, so the overlapped head belongs to the predecessor.
    Equals end - start when nothing overlaps; the values add up to the busy time of the stream."""
    out, frontier = [], None
    for start, end in spans:
        if frontier is None:
            frontier = start
        out.append(max(0.0, end - max(start, frontier)))
        frontier = max(frontier, end)
    return out


def cupti_profile(plan, replays=3):
    """Per-kernel device durations of `replays` graph replays of `plan` from CUPTI activity records (torch.profiler:
    hardware start / end timestamps of every kernel, no serialisation between launches, no replay - unlike ncu), matched
    to the plan's call list by launch order; overlapped heads (programmatic dependent launch) are not counted twice.  Returns {kind: {launches, ms, flops, bytes}} averaged per replay, plus
    the busy / span totals, or None when the profiler is unavailable."""
    try:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(replays):
                plan.launch()
            torch.cuda.synchronize()
        evs = sorted([e for e in prof.events() if e.device_type.name == "CUDA" and "emset" not in e.name.lower()
                      and "emcpy" not in e.name.lower()], key=lambda e: e.time_range.start)
    except Exception as exc:  # noqa: BLE001
        return {"error": repr(exc)[:200]}
    calls = [c for c in plan.calls if c.name != "b200sd_memset_zero"]
    if len(evs) != replays * len(calls):
        return {"error": f"{len(evs)} kernel records for {replays} x {len(calls)} calls"}
    agg = {}
    excl = exclusive_us([(e.time_range.start, e.time_range.end) for e in evs])
    for i, e in enumerate(evs):
        c = calls[i % len(calls)]
        fl, kind = flops_of_call(c)
        a = agg.setdefault(kind, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
        a["launches"] += 1
        a["ms"] += excl[i] * 1e-3  # see exclusive_us
        a["flops"] += fl
        a["bytes"] += algorithmic_bytes(c)
    for a in agg.values():
        for k in a:
            a[k] /= replays
    busy = sum(a["ms"] for a in agg.values())
    span = (evs[-1].time_range.end - evs[0].time_range.start) * 1e-3 / replays
    return {"kinds": agg, "busy_ms": busy, "span_ms": span}


def run_cuda(args):
    import importlib
    import torch.distributed as dist
    import b200sd
    P = importlib.import_module("k8s-nvidia-gpus_b200.pipeline")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL_DEBUG is left as the caller set it: claim_stdout() already keeps NCCL's banner off the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    b200sd.lib.load()
    info = b200sd.lib.device_info()
    pipe = P.StableDiffusionPipeline.from_pretrained(args.model, device=f"cuda:{local}", scheduler=args.scheduler)
    cfg = pipe.cfg
    f = pipe.vae_scale_factor
    h, w = args.height // f, args.width // f
    B, scaling = resolve_batch(args, world)
    sched = pipe.make_schedule(args.inference_steps)
    evals = sched.num_evals
    job = pipe.job(B, h, w, args.guidance_scale, True, zero_negative=pipe.is_xl)

    # synthetic inputs for this rank's shard (global image index = rank * B + i; seed = 1234 + index)
    prompts = synthetic_prompts(B, rank * B)
    gens = [torch.Generator().manual_seed(1234 + rank * B + i) for i in range(B)]
    ids = pipe.tokenize(prompts, [""] * B, True)
    lat = pipe.prepare_latents(B, h, w, gens, sched.init_noise_sigma)
    job.ids_dev.copy_(ids)
    job.eos_dev.copy_(pipe.eos_rows(ids, job.n_text))
    job.init_latents.copy_(lat.permute(0, 2, 3, 1).reshape(B, h * w, 4))
    job.step_table[:evals].copy_(sched.table)
    gathered = None
    if world > 1:
        gathered = torch.empty((world,) + tuple(job.images.shape), dtype=torch.uint8, device=job.images.device)

    def one_batch_resident():
        job.reset()
        job.run(evals)
        if world > 1:
            dist.all_gather_into_tensor(gathered, job.images)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_batch_resident()
    sampler = ClockSampler(local)
    sync_all()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for _ in range(args.steps):
        one_batch_resident()
    e1.record()
    sync_all()
    ms = torch.tensor([e0.elapsed_time(e1)], device=job.images.device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = ms.item()

    # UNet ms per denoise step (graph replay of one CFG-paired evaluation of the local batch + scheduler update)
    job.reset()
    torch.cuda.synchronize()
    u0, u1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    u0.record()
    for _ in range(evals):
        job.step_plan.launch()
    u1.record()
    torch.cuda.synchronize()
    unet_ms = u0.elapsed_time(u1) / evals
    # per-kernel CUPTI durations of the same graph-replayed loop (rank 0), while the clock sampler still runs
    cupti = cupti_vae = None
    if rank == 0 and not args.no_cupti:
        job.reset()
        cupti = cupti_profile(job.step_plan, replays=evals)
        cupti_vae = cupti_profile(job.vae_plan, replays=2)
    clocks = sampler.stop() if rank == 0 else None

    # end to end through the public call: host prompts / CPU-generator latents in, host uint8 images out
    def one_batch_e2e():
        g = [torch.Generator().manual_seed(1234 + rank * B + i) for i in range(B)]
        out = pipe(prompts, num_inference_steps=args.inference_steps, guidance_scale=args.guidance_scale,
                   width=args.width, height=args.height, generator=g, output_type="u8")
        if world > 1:
            dist.all_gather_into_tensor(gathered, job.images)
        return out

    one_batch_e2e()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_batch_e2e()
    sync_all()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=job.images.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_ips = args.steps * B * world / e2e_s.item()
    # the tensors pipe() copies every call: token ids, EOS rows, initial latents, the whole step table
    h2d = sum(t.numel() * t.element_size() for t in (job.ids_host, job.eos_host, job.lat_host, job.table_host))
    d2h = job.img_host.numel()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # per-kernel breakdown of one denoise step (eager launches with events; VAE plan as the run-ahead blocker)
    job.reset()
    prof = profile_plan(job.step_plan, blocker=job.vae_plan)
    vprof = profile_plan(job.vae_plan, blocker=job.vae_plan)
    event_gap_ms = prof.pop("_event_gap_ms")
    vprof.pop("_event_gap_ms")
    peaks = measured_peaks()
    gm = prof["gemm"]
    # Dominant kernel: gemm_conv_kernel.  Its time comes from CUPTI activity records of the graph-replayed denoise loop
    # (hardware timestamps of every kernel of `evals` back-to-back replays, taken right after the timed region on the
    # same warm, power-capped GPU; `cupti` below).  The CUDA-event figures are kept beside it: events between launches
    # serialise the step more than the graph replay does, so there the kernel's time is its SHARE of the intervals x
    # the event-timed graph step.
    raw_sum_ms = sum(a["ms"] for a in prof.values())
    gemm_share = gm["ms"] / raw_sum_ms
    gemm_ms_event_share = gemm_share * unet_ms
    cupti_ok = isinstance(cupti, dict) and "kinds" in cupti and "gemm" in cupti["kinds"]
    if cupti_ok:
        gemm_ms_in_step = cupti["kinds"]["gemm"]["ms"]
        method = (f"CUPTI activity records (torch.profiler) of {evals} graph replays of the denoise step, kernel "
                  "durations matched to the plan's calls by launch order, each counted from the end of the kernels launched "
                  "before it (programmatic dependent launch lets a GEMM's set-up start under its predecessor's tail); "
                  "achieved = sum 2MNK / sum of those durations")
    else:
        gemm_ms_in_step = gemm_ms_event_share
        method = ("share of the step from per-launch CUDA-event intervals (eager pass) x the CUDA-event-timed graph "
                  "step of the timed region (CUPTI unavailable: %s)" % (cupti or {}).get("error"))
    achieved_tf = gm["flops"] / (gemm_ms_in_step * 1e-3) / 1e12
    traffic, traffic_source = None, None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get("gemm_conv_kernel_dram_bytes_per_launch")
            traffic_source = ("not measured by this run: ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                              f"mean over one denoise step, from profiles/roofline_traffic.json ({tj.get('source', 'earlier ncu session')})")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "gemm_conv_kernel (tcgen05 implicit-GEMM conv/linear)",
                "achieved": achieved_tf, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": achieved_tf / peaks["tf_sustained"], "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": f"{peaks['source']} (sustained cuBLAS bf16; burst {peaks['tf_burst']})",
                "launches_per_unet_step": gm["launches"], "flops_per_launch": gm["flops"] / gm["launches"],
                "avg_launch_ms": gemm_ms_in_step / gm["launches"],
                "method": method,
                "share_of_unet_step": (gemm_ms_in_step / cupti["busy_ms"]) if cupti_ok else gemm_share,
                "event_share_of_unet_step": gemm_share,
                "event_share_tflops": gm["flops"] / (gemm_ms_event_share * 1e-3) / 1e12,
                "eager_interval_avg_launch_ms": gm["ms"] / gm["launches"],
                "eager_interval_tflops": gm["flops"] / (gm["ms"] * 1e-3) / 1e12}
    def cupti_table(cp):
        if not (isinstance(cp, dict) and "kinds" in cp):
            return cp
        t = {"busy_ms": round(cp["busy_ms"], 4), "span_ms": round(cp["span_ms"], 4)}
        for k, a in sorted(cp["kinds"].items(), key=lambda kv: -kv[1]["ms"]):
            t[k] = {"launches": int(round(a["launches"])), "ms": round(a["ms"], 4),
                    "tflops": round(a["flops"] / (a["ms"] * 1e-3) / 1e12, 1) if a["flops"] else None,
                    "gbs": round(a["bytes"] / (a["ms"] * 1e-3) / 1e9, 1) if a["bytes"] else None,
                    "frac_of_hbm_peak": round(a["bytes"] / (a["ms"] * 1e-3) / 1e9 / peaks["hbm_gbs"], 3)
                    if a["bytes"] else None}
        return t
    cupti_breakdown, cupti_vae_breakdown = cupti_table(cupti), cupti_table(cupti_vae)
    step_flops = sum(a["flops"] for a in prof.values())
    def table(pr):
        return {k: {"launches": a["launches"], "ms": round(a["ms"], 4),
                    "tflops": round(a["flops"] / (a["ms"] * 1e-3) / 1e12, 1) if a["flops"] else None,
                    "gbs": round(a["bytes"] / (a["ms"] * 1e-3) / 1e9, 1) if a["bytes"] and a["ms"] > 0 else None}
                for k, a in sorted(pr.items(), key=lambda kv: -kv[1]["ms"])}
    breakdown, vae_breakdown = table(prof), table(vprof)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        unet, vae, clip = build_oracle(cfg)
        r = cpu_sample(cfg, unet, vae, clip, 2, 1, args.inference_steps, h, w)
        cpu = {"value": r["ips"], "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
               "sample": f"1 image: 2 timed + 1 warm-up CFG-paired UNet evaluations, 1 VAE decode, 1 CLIP pass; "
                         f"images/sec = 1/({args.inference_steps}*t_unet + t_vae + t_clip)",
               "detail": {k: round(v, 4) for k, v in r.items()}}

    ips = args.steps * B * world / (total_ms * 1e-3)
    img_tf = (evals * 2 * UNET_GF_PER_SAMPLE + VAE_GF_PER_IMAGE + 2 * CLIP_GF_PER_SEQ) / 1e3
    line = {
        "metric": "images/sec", "value": ips, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "fp16" if b200sd.lib.act_dtype() == torch.float16 else "bf16",
        "data": "synthetic", "config": workload_config(args, cfg, world),
        "unet_ms_per_step": unet_ms,
        "unet_tflops": 2 * B * UNET_GF_PER_SAMPLE / 1e3 / (unet_ms * 1e-3) if cfg["name"] == "sd15" else None,
        "unet_step_algorithmic_tflop": step_flops / 1e12,
        "image_algorithmic_tflop": img_tf if cfg["name"] == "sd15" else None,
        "e2e": {"value": e2e_ips, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": args.steps * job.launches(evals),
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "unet_step_cupti": cupti_breakdown, "vae_decode_cupti": cupti_vae_breakdown, "unet_step_breakdown": breakdown, "vae_breakdown": vae_breakdown,
        "breakdown_note": f"CUDA-event intervals of an eager pass minus the empty-interval gap ({event_gap_ms * 1e3:.1f} us)",
        "device": {"sm_count": info["sm_count"], "cc": list(info["cc"])},
    }
    emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_sweep(args):
    """BASELINE.json configs[4]: total batch swept over --sweep (e.g. 1,2,4,...,128) on the GPUs of this launch.  The
    prompts of each level are sharded over the ranks like the product does (contiguous slices, remainder to the low
    ranks; levels smaller than the GPU count leave GPUs idle - reported as is).  One JSON line per level."""
    import importlib
    import torch.distributed as dist
    import b200sd
    P = importlib.import_module("k8s-nvidia-gpus_b200.pipeline")
    D = b200sd.dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    # the CPU baseline runs on rank 0 BEFORE the process group exists: the other ranks then sleep in the rendezvous
    # instead of spinning in an NCCL barrier on the cores the baseline is using
    cpu = None
    Wt = importlib.import_module("k8s-nvidia-gpus_b200.weights")
    cfg0 = Wt.load_config(args.model)
    f0 = 2 ** (len(cfg0["vae"]["block_out_channels"]) - 1)
    if rank == 0 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        unet, vae, clip = build_oracle(cfg0)
        r = cpu_sample(cfg0, unet, vae, clip, 2, 1, args.inference_steps, args.height // f0, args.width // f0)
        cpu = {"value": r["ips"], "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
               "sample": "1 image: 2 timed + 1 warm-up CFG-paired UNet evaluations, 1 VAE decode, 1 CLIP pass"}
        del unet, vae, clip
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(minutes=30))
    pipe = P.StableDiffusionPipeline.from_pretrained(args.model, device=f"cuda:{local}", scheduler=args.scheduler)
    cfg = pipe.cfg
    f = pipe.vae_scale_factor
    h, w = args.height // f, args.width // f
    sched = pipe.make_schedule(args.inference_steps)
    evals = sched.num_evals
    dev = torch.device("cuda", local)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lines = []
    for total in [int(x) for x in args.sweep.split(",")]:
        b0, b1 = D.shard_range(total, world, rank)
        B = b1 - b0
        counts = D.shard_sizes(total, world)
        job = None
        if B > 0:
            job = pipe.job(B, h, w, args.guidance_scale, True, zero_negative=pipe.is_xl)
            prompts = synthetic_prompts(B, b0)
            gens = [torch.Generator().manual_seed(1234 + b0 + i) for i in range(B)]
            ids = pipe.tokenize(prompts, [""] * B, True)
            lat = pipe.prepare_latents(B, h, w, gens, sched.init_noise_sigma)
            job.ids_dev.copy_(ids)
            job.eos_dev.copy_(pipe.eos_rows(ids, job.n_text))
            job.init_latents.copy_(lat.permute(0, 2, 3, 1).reshape(B, h * w, 4))
            job.step_table[:evals].copy_(sched.table)
        empty = torch.zeros((0, h * f, w * f, 3), dtype=torch.uint8, device=dev)

        def one_batch():
            if job is not None:
                job.reset()
                job.run(evals)
            if world > 1:
                D.gather_images(job.images if job is not None else empty, counts)

        for _ in range(max(1, args.warmup)):
            one_batch()
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            one_batch()
        e1.record()
        sync_all()
        t = torch.tensor([e0.elapsed_time(e1), 0.0], device=dev)
        if job is not None:
            job.reset()
            u0, u1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            u0.record()
            for _ in range(evals):
                job.step_plan.launch()
            u1.record()
            torch.cuda.synchronize()
            t[1] = u0.elapsed_time(u1) / evals
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            ips = args.steps * total / (t[0].item() * 1e-3)
            line = {"metric": "images/sec", "sweep_total_batch": total, "value": ips, "unit": "images/s",
                    "n_gpus": world, "gpus_busy": sum(1 for c in counts if c > 0), "per_gpu_batches": counts,
                    "steps": args.steps, "warmup": max(1, args.warmup), "ms_per_step": t[0].item() / args.steps,
                    "latency_ms_per_batch": t[0].item() / args.steps, "unet_ms_per_step_max": t[1].item(),
                    "split_k_policy": os.environ.get("B200SD_SPLITK", "shape"),
                    "vs_cpu_reference": (ips / cpu["value"]) if cpu else None, "cpu_baseline": cpu,
                    "config": {"workload": f"{cfg['name']} architecture (random-init), {args.width}x{args.height}, "
                                           f"{args.inference_steps} {args.scheduler.upper()} steps, CFG "
                                           f"{args.guidance_scale}, total batch {total} sharded over {world} GPU(s)"}}
            lines.append(line)
            emit(json.dumps(line))
        pipe._jobs.clear()   # the next level builds its own plans; give the activations back first
        job = None
        torch.cuda.empty_cache()
    if rank == 0 and args.sweep_out:
        with open(args.sweep_out, "w") as fo:
            json.dump(lines, fo, indent=1)
    if world > 1:
        dist.destroy_process_group()
    return 0


_RESULT_FD = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: keep a private duplicate of fd 1 for it and point fd 1 at stderr, so that
    nothing printed by libraries underneath (NCCL's version banner, C-level printf, warnings) can land there."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(text: str):
    data = (text + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(text + "\n")
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main(argv=None):
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="sd15")
    ap.add_argument("--batch", type=int, default=8, help="images per GPU per bench step (one GPU, or with --weak)")
    ap.add_argument("--global-batch", type=int, default=None,
                    help="total images per bench step, sharded over the GPUs (default with --gpus N > 1: 64 = configs[2])")
    ap.add_argument("--weak", action="store_true", help="several GPUs: keep --batch images per GPU (weak scaling)")
    ap.add_argument("--inference-steps", type=int, default=25)
    ap.add_argument("--guidance-scale", type=float, default=7.5)
    ap.add_argument("--width", type=int, default=512)
    ap.add_argument("--height", type=int, default=512)
    ap.add_argument("--scheduler", default="ddim")
    ap.add_argument("--sweep", default=None, help="batch-size sweep (configs[4]): comma-separated TOTAL batch sizes, one "
                                                   "JSON line each (e.g. 1,2,4,8,16,32,64,128)")
    ap.add_argument("--sweep-out", default=None, help="also write the sweep lines to this JSON file")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cupti", action="store_true", help="skip the CUPTI per-kernel pass (roofline falls back to "
                                                            "the CUDA-event share method)")
    args = ap.parse_args(argv)
    if args.impl == "reference":
        return run_reference(args)
    if args.sweep:
        return run_sweep(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and args.gpus > 1:
        print(f"bench.py: --gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})", file=sys.stderr)
        return 2
    return run_cuda(args)


if __name__ == "__main__":
    raise SystemExit(main())
