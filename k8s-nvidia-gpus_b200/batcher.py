"""Server-side dynamic batching.

The reference service answers every `/generate` with its own `pipe(prompt, ...)` call, straight from Starlette's
thread pool and without any guard around the stateful pipeline (cluster-config/apps/sd15-api/configmap.yaml:86-121);
its client sends `count` requests one after another (scripts/batch_generate.py:17-24).  One image per call cannot
feed a B200 - the batch-8 step is 3x cheaper per image than the batch-1 step - so concurrent requests with the same
shape (steps, guidance, width, height) are coalesced here into ONE batched `pipe([...], generator=[...])` call:

  * a request's image does not depend on what it is batched with: every image draws its own `[1, 4, h, w]` latents
    from its own CPU generator (the reference's CPU branch, configmap.yaml:92) and the kernels are batch-invariant
    (tests/test_sd15_gpu.py), so seed -> image is the same as for a lone request;
  * batches are padded up to a fixed set of sizes (default 1, 2, 4, 8) so that only a handful of CUDA graphs / buffer
    sets ever exist per image shape; the padding rows repeat the last request and are dropped;
  * exactly one worker thread talks to the pipeline, which also fixes the reference's unguarded concurrent `pipe()`.

Host-side logic only (threads, queues); the GPU work is whatever `pipe` does.
"""
from __future__ import annotations

import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field

import torch


@dataclass
class _Req:
    prompt: str
    key: tuple  # (steps, guidance_scale, width, height)
    generator: torch.Generator | None
    future: Future = field(default_factory=Future)
    t_enqueue: float = field(default_factory=time.monotonic)


class DynamicBatcher:
    def __init__(self, pipe_getter, max_batch: int = 8, max_wait_ms: float = 10.0, buckets=(1, 2, 4, 8),
                 output_type: str = "pil"):
        """pipe_getter: callable returning the pipeline (built lazily, on the worker thread's first batch).
        max_wait_ms: how long the first request of a batch waits for company before the batch is launched."""
        self._get_pipe = pipe_getter
        self.buckets = tuple(sorted(b for b in buckets if b <= max_batch)) or (1,)
        self.max_batch = self.buckets[-1]
        self.max_wait = max_wait_ms / 1e3
        self.output_type = output_type
        self._cv = threading.Condition()
        self._queue: list[_Req] = []
        self._stop = False
        self.stats = {"requests": 0, "batches": 0, "images": 0, "padded": 0, "largest": 0}
        self._worker = threading.Thread(target=self._run, name="b200sd-batcher", daemon=True)
        self._worker.start()

    # ------------------------------------------------------------------ client side
    def submit(self, prompt: str, steps: int, guidance_scale: float, width: int, height: int,
               seed: int | None = None) -> Future:
        gen = torch.Generator(device="cpu").manual_seed(seed) if seed is not None else None
        req = _Req(prompt, (int(steps), float(guidance_scale), int(width), int(height)), gen)
        with self._cv:
            if self._stop:
                raise RuntimeError("batcher is shut down")
            self._queue.append(req)
            self.stats["requests"] += 1
            self._cv.notify_all()
        return req.future

    def generate(self, *a, **kw):
        """Blocking form of submit(): returns the image or raises the pipeline's exception."""
        return self.submit(*a, **kw).result()

    def shutdown(self):
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        self._worker.join(timeout=5)

    # ------------------------------------------------------------------ worker
    def _take_batch(self):
        """Oldest request first; its shape key selects the batch; waits up to max_wait for it to fill."""
        with self._cv:
            while not self._queue and not self._stop:
                self._cv.wait()
            if self._stop and not self._queue:
                return None
            head = self._queue[0]
            deadline = head.t_enqueue + self.max_wait
            while True:
                same = [r for r in self._queue if r.key == head.key]
                left = deadline - time.monotonic()
                if len(same) >= self.max_batch or left <= 0 or self._stop:
                    break
                self._cv.wait(timeout=left)
            batch = same[: self.max_batch]
            taken = set(map(id, batch))
            self._queue = [r for r in self._queue if id(r) not in taken]
            return batch

    def _run(self):
        while True:
            batch = self._take_batch()
            if batch is None:
                return
            n = len(batch)
            size = next(b for b in self.buckets if b >= n)
            steps, gs, width, height = batch[0].key
            prompts = [r.prompt for r in batch] + [batch[-1].prompt] * (size - n)
            # an unseeded request gets a fresh generator (the reference leaves `generator=None`: a random image)
            gens = [r.generator if r.generator is not None else torch.Generator(device="cpu").manual_seed(
                torch.seed() & 0x7FFFFFFF) for r in batch]
            gens += [torch.Generator(device="cpu").manual_seed(0) for _ in range(size - n)]
            try:
                pipe = self._get_pipe()
                out = pipe(prompts, num_inference_steps=steps, guidance_scale=gs, width=width, height=height,
                           generator=gens, output_type=self.output_type)
                images = out.images
                for i, r in enumerate(batch):
                    r.future.set_result(images[i])
            except BaseException as e:  # noqa: BLE001 - every waiter must learn about the failure
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)
            st = self.stats
            st["batches"] += 1
            st["images"] += n
            st["padded"] += size - n
            st["largest"] = max(st["largest"], n)
