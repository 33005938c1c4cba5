#!/usr/bin/env python3
"""Batch image generation — drop-in for the reference client (boundary B1).

Positional arguments, flags, output file names and stdout lines are those of the reference's
scripts/batch_generate.py (argv :46-52, files `{out_dir}/{prefix}_{idx:02d}.png` :18-19, messages :22, :28, :41, :57;
wire body `{"prompt", "steps"}` with a 600 s timeout :20-24):

    batch_generate.py PROMPT COUNT PREFIX [OUT_DIR] [--steps N] [--url URL] [--delay S]

Deliberate differences: `traceback` is imported (the reference references it without importing it, so any failed
request there dies with NameError instead of continuing, :30-35); the `--steps` help text states the real default.

Additions (supersets, all optional): `--local` generates in-process on the local B200(s) through
b200sd.StableDiffusionPipeline instead of HTTP — batched (`--batch-size`), seeded per image (`--seed` + index),
with `--guidance-scale/--width/--height/--scheduler/--model`, `--prompts-file` (one prompt per line, cycled) and PNG
encoding on `--png-threads` threads overlapped with the next batch, and sharded over all GPUs of the node when started
under `torchrun --nproc-per-node N -- batch_generate.py ...` (one process per GPU, same files for any N; the `--` keeps
torchrun's own parser away from `--local`); `--concurrency N` keeps N HTTP requests in flight
so that b200sd's batching server can coalesce them (default 1: the reference's sequential loop).
"""
import argparse
import os
import sys
import time
import traceback
from pathlib import Path

DEFAULT_URL = "http://192.168.0.176:30800/generate"
HTTP_TIMEOUT_S = 600


def _target(out_dir: Path, prefix: str, idx: int) -> Path:
    return out_dir / f"{prefix}_{idx:02d}.png"


def generate(prompt: str, steps: int, url: str, out_dir: Path, prefix: str, count: int, delay: float,
             concurrency: int = 1) -> None:
    """HTTP mode: `count` POSTs, one PNG per response; failures are reported and the loop goes on.  concurrency = 1 is
    the reference's strictly sequential loop; more keeps that many requests in flight so that a batching server
    (b200sd's) can coalesce them."""
    import requests

    out_dir.mkdir(parents=True, exist_ok=True)
    http = requests.Session()

    def one(idx: int) -> None:
        target = _target(out_dir, prefix, idx)
        print(f"[*] Generating {target.name} -> {target}")
        try:
            reply = http.post(url, json={"prompt": prompt, "steps": steps}, timeout=HTTP_TIMEOUT_S)
            reply.raise_for_status()
            target.write_bytes(reply.content)
            print(f"    ✔ done in {reply.headers.get('X-Gen-Time', '?')}s")
        except requests.exceptions.RequestException as exc:
            print(f"    ✖ Request failed for {target.name}: {exc}")
            traceback.print_exc()
        except Exception as exc:  # noqa: BLE001 — keep going, like the reference intends
            print(f"    ✖ Unexpected error for {target.name}: {exc}")
            traceback.print_exc()

    if concurrency > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(concurrency) as pool:
            list(pool.map(one, range(1, count + 1)))
    else:
        for idx in range(1, count + 1):
            one(idx)
            if delay > 0 and idx != count:
                time.sleep(delay)
    print("[*] Generation loop finished (with possible errors).")


def generate_local(prompts, steps, out_dir: Path, prefix: str, count: int, *, batch_size: int, seed: int,
                   guidance_scale: float, width: int, height: int, scheduler, model: str, pipe=None,
                   png_threads: int = 4) -> None:
    """Local mode: the same files, produced by b200sd on this machine's GPU in batches of `batch_size`."""
    import torch

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return generate_local_sharded(prompts, steps, out_dir, prefix, count, batch_size=batch_size, seed=seed,
                                      guidance_scale=guidance_scale, width=width, height=height, scheduler=scheduler,
                                      model=model, pipe=pipe, png_threads=png_threads)
    if pipe is None:
        import b200sd
        pipe = b200sd.StableDiffusionPipeline.from_pretrained(model, scheduler=scheduler)
    out_dir.mkdir(parents=True, exist_ok=True)
    # PNG encoding (zlib, releases the GIL) runs on a small thread pool while the GPU works on the next batch
    from concurrent.futures import ThreadPoolExecutor
    from PIL import Image
    pool = ThreadPoolExecutor(max_workers=png_threads)
    pending = []

    def save(arr, path):
        Image.fromarray(arr).save(path, format="PNG")

    for start in range(1, count + 1, batch_size):
        idxs = list(range(start, min(count, start + batch_size - 1) + 1))
        batch_prompts = [prompts[(i - 1) % len(prompts)] for i in idxs]
        gens = [torch.Generator().manual_seed(seed + i - 1) for i in idxs]
        for i in idxs:
            print(f"[*] Generating {_target(out_dir, prefix, i).name} -> {_target(out_dir, prefix, i)}")
        t0 = time.time()
        try:
            images = pipe(batch_prompts, num_inference_steps=steps, guidance_scale=guidance_scale, width=width,
                          height=height, generator=gens, output_type="u8").images
            pending += [pool.submit(save, img, _target(out_dir, prefix, i)) for i, img in zip(idxs, images)]
            dt = time.time() - t0
            for _ in idxs:
                print(f"    ✔ done in {dt / len(idxs):.2f}s")
        except Exception as exc:  # noqa: BLE001
            print(f"    ✖ Unexpected error for {_target(out_dir, prefix, idxs[0]).name}: {exc}")
            traceback.print_exc()
    for f in pending:
        try:
            f.result()
        except Exception as exc:  # noqa: BLE001
            print(f"    ✖ Unexpected error while writing a PNG: {exc}")
    pool.shutdown()
    print("[*] Generation loop finished (with possible errors).")


def generate_local_sharded(prompts, steps, out_dir: Path, prefix: str, count: int, *, batch_size: int, seed: int,
                           guidance_scale: float, width: int, height: int, scheduler, model: str, pipe=None,
                           png_threads: int = 4, backend: str | None = None) -> None:
    """Local mode under `torchrun --nproc-per-node N`: one process per GPU, the `count` images sharded over the ranks
    by b200sd.dist.generate_sharded (contiguous slices, image i always seeded with seed + i - 1, so the files are
    byte-identical for any N), one gather of the uint8 images, rank 0 writes the same files as the single-GPU mode."""
    import torch
    import torch.distributed as dist
    import b200sd
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    own_group = not dist.is_initialized()
    if pipe is None:
        torch.cuda.set_device(local)
        if own_group:
            dist.init_process_group(backend or "nccl", device_id=torch.device("cuda", local))
        pipe = b200sd.StableDiffusionPipeline.from_pretrained(model, device=f"cuda:{local}", scheduler=scheduler)
    elif own_group:
        dist.init_process_group(backend or "gloo")
    all_prompts = [prompts[i % len(prompts)] for i in range(count)]
    t0 = time.time()
    images = b200sd.dist.generate_sharded(pipe, all_prompts, base_seed=seed, max_batch=batch_size,
                                          num_inference_steps=steps, guidance_scale=guidance_scale, width=width,
                                          height=height)
    dt = time.time() - t0
    if rank == 0:
        from concurrent.futures import ThreadPoolExecutor
        from PIL import Image
        out_dir.mkdir(parents=True, exist_ok=True)
        arr = images.cpu().numpy()
        with ThreadPoolExecutor(max_workers=png_threads) as pool:
            futs = []
            for i in range(count):
                target = _target(out_dir, prefix, i + 1)
                print(f"[*] Generating {target.name} -> {target}")
                futs.append(pool.submit(lambda a, t: Image.fromarray(a).save(t, format="PNG"), arr[i], target))
                print(f"    ✔ done in {dt / max(1, count):.2f}s")
            for f in futs:
                f.result()
        print("[*] Generation loop finished (with possible errors).")
    if own_group:
        dist.barrier()
        dist.destroy_process_group()


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description="Batch-generate images via the SD15 API")
    ap.add_argument("prompt", help="prompt to send to the API")
    ap.add_argument("count", type=int, help="number of images to generate")
    ap.add_argument("prefix", help="output filename prefix, e.g. piggy")
    ap.add_argument("out_dir", nargs="?", default="outputs", help="directory to save images (default: outputs)")
    ap.add_argument("--steps", type=int, default=40, help="diffusion steps per image (default: 40)")
    ap.add_argument("--url", default=DEFAULT_URL, help=f"API endpoint (default: {DEFAULT_URL})")
    ap.add_argument("--delay", type=float, default=0, help="seconds to sleep between requests")
    ap.add_argument("--concurrency", type=int, default=1,
                    help="requests kept in flight (default 1 = the reference's sequential loop; an addition)")
    loc = ap.add_argument_group("local generation (additions to the reference CLI)")
    loc.add_argument("--local", action="store_true", help="generate in-process on the local GPU instead of HTTP")
    loc.add_argument("--batch-size", type=int, default=8)
    loc.add_argument("--seed", type=int, default=1234, help="image i uses seed + i")
    loc.add_argument("--guidance-scale", type=float, default=7.5)
    loc.add_argument("--width", type=int, default=512)
    loc.add_argument("--height", type=int, default=512)
    loc.add_argument("--scheduler", default=None, choices=[None, "ddim", "pndm", "euler"])
    loc.add_argument("--model", default="sd15")
    loc.add_argument("--png-threads", type=int, default=4, help="PNG encoder threads overlapping the GPU work")
    loc.add_argument("--prompts-file", default=None, help="one prompt per line; overrides PROMPT, cycled over COUNT")
    return ap


def main(argv: list) -> int:
    args = build_parser().parse_args(argv)
    out_dir = Path(args.out_dir)
    if args.local:
        prompts = [args.prompt]
        if args.prompts_file:
            prompts = [l.strip() for l in Path(args.prompts_file).read_text().splitlines() if l.strip()] or prompts
        generate_local(prompts, args.steps, out_dir, args.prefix, args.count, batch_size=args.batch_size,
                       seed=args.seed, guidance_scale=args.guidance_scale, width=args.width, height=args.height,
                       scheduler=args.scheduler, model=args.model, png_threads=args.png_threads)
    else:
        generate(args.prompt, args.steps, args.url, out_dir, args.prefix, args.count, args.delay, args.concurrency)
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"All done. Images saved under {out_dir.resolve()}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main(sys.argv[1:]))
