"""HTTP front end with the reference service's contract (drop-in boundary B2).

Same surface as the FastAPI app embedded in cluster-config/apps/sd15-api/configmap.yaml:16-121:
  POST /generate  {prompt, steps?=30, guidance_scale?=7.5, seed?, width?=512, height?=512} -> image/png,
                  header X-Gen-Time "<seconds>.2fs"; 400 "prompt is required" for a blank prompt; falsy optional
                  fields fall back to the defaults (configmap.yaml:86-121)
  GET  /healthz   {"ok": true}                                   (configmap.yaml:60-62; the Deployment's readinessProbe)
  GET  /last      last PNG or 404                                (configmap.yaml:80-84)
  GET  /          HTML preview of the last image                 (configmap.yaml:64-78)
Differences, all deliberate: the pipeline is b200sd's; concurrent requests of the same shape are coalesced into one
batched pipeline call by a single worker thread (batcher.py; B200SD_MAX_BATCH, default 8, and B200SD_BATCH_WAIT_MS,
default 10 - the reference calls its stateful pipeline from Starlette's thread pool unguarded, one image per call,
configmap.yaml:105-112); PNG encoding stays on the request threads, off the GPU path; the seed drives a CPU
generator — the reference's CPU branch (configmap.yaml:92) — so a seed reproduces the oracle's image whether or not
the request was batched.

    uvicorn app:app --host 0.0.0.0 --port 8000      # from the repo root, exactly the command of deployment.yaml:42
                                                    # (env MODEL_ID as in deployment.yaml:45-46)
"""
from __future__ import annotations

import base64
import io
import logging
import os
import time
from threading import Lock

from fastapi import FastAPI, HTTPException
from fastapi.responses import HTMLResponse, Response
from pydantic import BaseModel

logger = logging.getLogger("sd15-api")

DEFAULT_STEPS, DEFAULT_GUIDANCE, DEFAULT_SIZE = 30, 7.5, 512
MIN_SIZE, MAX_SIZE, MAX_STEPS = 8, int(os.getenv("B200SD_MAX_SIZE", "2048")), 1000


class GenReq(BaseModel):
    prompt: str
    steps: int | None = DEFAULT_STEPS
    guidance_scale: float | None = DEFAULT_GUIDANCE
    seed: int | None = None
    width: int | None = DEFAULT_SIZE
    height: int | None = DEFAULT_SIZE


def create_app(pipe=None, pipe_factory=None, max_batch: int | None = None, max_wait_ms: float | None = None) -> FastAPI:
    """pipe: an object with the StableDiffusionPipeline call surface; pipe_factory: built lazily on first use."""
    from .batcher import DynamicBatcher
    app = FastAPI(title="SD15 Minimal API (b200sd)")
    state = {"pipe": pipe, "last": None}
    last_lock, pipe_lock = Lock(), Lock()
    max_batch = int(os.getenv("B200SD_MAX_BATCH", "8")) if max_batch is None else max_batch
    max_wait_ms = float(os.getenv("B200SD_BATCH_WAIT_MS", "10")) if max_wait_ms is None else max_wait_ms

    def get_pipe():
        if state["pipe"] is None:
            with pipe_lock:
                if state["pipe"] is None:
                    state["pipe"] = (pipe_factory or _default_factory)()
        return state["pipe"]

    batcher = DynamicBatcher(get_pipe, max_batch=max_batch, max_wait_ms=max_wait_ms)
    app.state.batcher = batcher

    @app.get("/healthz")
    def healthz():
        return {"ok": True}

    @app.get("/", response_class=HTMLResponse)
    def index():
        if state["last"] is None:
            return "<h1>SD1.5 GPU API</h1><p>No image generated yet. POST /generate to create one.</p>"
        b64 = base64.b64encode(state["last"]).decode("ascii")
        return ("<html><head><title>SD1.5 Demo</title></head><body style='background:#0b0b0f;color:#f0f0f0;"
                "font-family:sans-serif;'><h1>Latest image</h1>"
                f"<img src='data:image/png;base64,{b64}' alt='latest image' style='max-width:90vw;height:auto;'/>"
                "<p>POST <code>/generate</code> with a prompt to update this preview.</p></body></html>")

    @app.get("/last")
    def last_image():
        if state["last"] is None:
            raise HTTPException(404, "No image generated yet")
        return Response(content=state["last"], media_type="image/png")

    @app.post("/generate")
    def generate(req: GenReq):
        if not req.prompt or not req.prompt.strip():
            raise HTTPException(400, "prompt is required")
        steps = req.steps or DEFAULT_STEPS
        gs = req.guidance_scale or DEFAULT_GUIDANCE
        width, height = req.width or DEFAULT_SIZE, req.height or DEFAULT_SIZE
        # client-controlled shapes are bounded BEFORE they reach the GPU (each new size builds plans and a CUDA graph):
        # upstream's own rule (multiples of 8) plus a size / step range; errors are 400s, not 500s
        if width % 8 or height % 8 or not (MIN_SIZE <= width <= MAX_SIZE) or not (MIN_SIZE <= height <= MAX_SIZE):
            raise HTTPException(400, f"width and height must be multiples of 8 in [{MIN_SIZE}, {MAX_SIZE}]")
        if not (1 <= steps <= MAX_STEPS):
            raise HTTPException(400, f"steps must be in [1, {MAX_STEPS}]")
        if not (0.0 <= gs <= 100.0):
            raise HTTPException(400, "guidance_scale must be in [0, 100]")
        t0 = time.time()
        logger.info("Generating prompt='%s' steps=%s guidance=%.2f seed=%s size=%sx%s", req.prompt, steps, gs,
                    req.seed if req.seed is not None else "auto", width, height)
        # joins the batch of same-shaped requests in flight; pipeline errors surface as 500 like the reference's
        image = batcher.generate(req.prompt, steps, gs, width, height, req.seed)
        buf = io.BytesIO()
        image.save(buf, format="PNG")
        png = buf.getvalue()
        latency = time.time() - t0
        logger.info("Completed generation in %.2fs", latency)
        with last_lock:
            state["last"] = png
        return Response(content=png, media_type="image/png", headers={"X-Gen-Time": f"{latency:.2f}s"})

    return app


def _default_factory():
    from .pipeline import StableDiffusionPipeline
    # MODEL_ID: a checkpoint directory in diffusers layout (what the reference's HF cache holds) or an architecture
    # name; a hub id / missing path raises (no silent random weights - see from_pretrained)
    model_id = os.getenv("MODEL_ID", "/models/stable-diffusion-v1-5")
    pipe = StableDiffusionPipeline.from_pretrained(model_id)
    pipe.enable_attention_slicing()
    pipe.enable_vae_slicing()
    logger.info("Loaded %s on CUDA (b200sd)", model_id)
    return pipe


logging.basicConfig(level=logging.INFO)
app = create_app()
