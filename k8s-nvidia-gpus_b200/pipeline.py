"""`StableDiffusionPipeline`-compatible front end of the B200 hot path (drop-in boundary B3).

Mirrors exactly the surface the reference server uses (cluster-config/apps/sd15-api/configmap.yaml:41-47, 105-112):

    pipe = StableDiffusionPipeline.from_pretrained(MODEL_ID, torch_dtype=...)
    pipe.enable_attention_slicing(); pipe.enable_vae_slicing(); pipe.vae.to("cpu"); pipe = pipe.to("cuda")
    image = pipe(prompt, num_inference_steps=..., guidance_scale=..., width=..., height=..., generator=gen).images[0]

plus the batch form `pipe([prompt, ...], generator=[gen, ...])` that BASELINE.json's batch configs name.  The three
memory toggles are accepted and ignored (identity maths).  Everything between tokenisation and the
uint8 image runs as b200sd CUDA kernels; there is no CPU or library fallback.
"""
from __future__ import annotations

import collections
import logging
import os

import numpy as np
import torch

from . import lib as L
from . import ops
from . import schedulers as S
from . import checkpoint as Ck
from . import weights as Wt
from .clip import CLIPTextEngine
from .tokenizer import CLIPTokenizer
from .unet import UNetEngine
from .vae import VAEDecoderEngine

MAX_EVALS = 1024
MAX_JOBS = int(os.environ.get("B200SD_MAX_JOBS", "4"))  # prepared (batch, size) shapes kept per pipeline (LRU)
logger = logging.getLogger("b200sd")


def sinusoid(values: torch.Tensor, dim: int) -> torch.Tensor:
    """diffusers `Timesteps(dim, flip_sin_to_cos=True, downscale_freq_shift=0)` on the host (fp32): [len(values), dim]
    = [cos | sin](v * exp(-ln(10000) * i / (dim / 2))).  Used for SDXL's six size / crop ids (StableDiffusionXLPipeline
    ._get_add_time_ids + UNet2DConditionModel.add_time_proj), which depend on the image size only."""
    half = dim // 2
    freqs = torch.exp(-np.log(10000.0) * torch.arange(half, dtype=torch.float32) / half)
    a = values.to(torch.float32)[:, None] * freqs[None]
    return torch.cat([torch.cos(a), torch.sin(a)], dim=-1)


class PipelineOutput:
    def __init__(self, images, nsfw_content_detected=None, latents=None):
        self.images = images
        self.nsfw_content_detected = nsfw_content_detected
        self.latents = latents

    def __getitem__(self, i):
        return (self.images, self.nsfw_content_detected)[i]


class _VaeHandle:
    """`pipe.vae.to("cpu")` (reference configmap.yaml:44-45, VAE_CPU toggle) is accepted and ignored: the decoder
    always runs on the GPU."""

    def __init__(self, engine):
        self.engine = engine

    def to(self, *_a, **_k):
        return self


class Job:
    """Everything prepared for one (batch, latent size, guidance, scheduler kind) shape: device buffers + plans.
    `run(num_evals)` enqueues text encoder -> K/V projection -> denoise steps -> VAE decode on the current stream,
    reading token ids / initial latents from `ids_dev` / `init_latents` (already resident in HBM)."""

    def __init__(self, pipe, batch, h, w, guidance_scale, cfg_on, zero_negative=False):
        self.pipe, self.batch, self.h, self.w = pipe, batch, h, w
        self.cfg_on = cfg_on
        dev = pipe.device
        nenc = len(pipe.text_encoders)
        self.n_text = batch * (2 if cfg_on else 1)
        # token ids of every text encoder back to back ([encoder][n_text * 77]); EOS row per sequence for the pooled one
        self.ids_dev = torch.zeros(nenc * self.n_text * 77, dtype=torch.int32, device=dev)
        self.eos_dev = torch.zeros(self.n_text, dtype=torch.int32, device=dev)
        self.init_latents = torch.zeros(batch, h * w, 4, dtype=torch.float32, device=dev)
        self.latents = torch.zeros(batch, h * w, 4, dtype=torch.float32, device=dev)
        self.history = torch.zeros(4, batch, h * w, 4, dtype=torch.float32, device=dev)
        self.cur_sample = torch.zeros(batch, h * w, 4, dtype=torch.float32, device=dev)
        self.step_table = torch.zeros(MAX_EVALS, 16, dtype=torch.float32, device=dev)
        self.step_counter = torch.zeros(2, dtype=torch.int32, device=dev)
        # pinned staging for the end-to-end path (host -> device inputs, device -> host images)
        self.ids_host = torch.zeros(nenc * self.n_text * 77, dtype=torch.int32).pin_memory()
        self.eos_host = torch.zeros(self.n_text, dtype=torch.int32).pin_memory()
        self.lat_host = torch.zeros(batch, h * w, 4, dtype=torch.float32).pin_memory()
        self.table_host = torch.zeros(MAX_EVALS, 16, dtype=torch.float32).pin_memory()
        self.text_plans, ctxs, pooled = [], [], None
        for e, enc in enumerate(pipe.text_encoders):
            ids = self.ids_dev[e * self.n_text * 77:(e + 1) * self.n_text * 77]
            plan, ctx, pl = enc.build_plan(self.n_text, ids, self.eos_dev)
            self.text_plans.append(plan)
            ctxs.append(ctx)
            pooled = pl if pl is not None else pooled
        self.ctx, self.pooled = ctxs, pooled
        # SDXL, negative_prompt=None (force_zeros_for_empty_prompt): the unconditional half of every text output is 0
        self.zero_plan = None
        if zero_negative and cfg_on:
            from .engine import Plan
            self.zero_plan = Plan()
            for t in ctxs + ([pooled] if pooled is not None else []):
                self.zero_plan.calls.append(ops.memset_zero(t, batch * (t.numel() // self.n_text) * t.element_size()))
        time_emb = None
        if pipe.cfg["unet"].get("addition_embed_type") == "text_time":
            f = pipe.vae_scale_factor
            # add_time_ids = original_size + crops_coords_top_left + target_size, same for both CFG halves
            ids6 = torch.tensor([h * f, w * f, 0, 0, h * f, w * f], dtype=torch.float32)
            row = sinusoid(ids6, pipe.cfg["unet"]["addition_time_embed_dim"]).reshape(1, -1)
            time_emb = row.repeat(self.n_text, 1).to(ops.act_dtype()).to(dev).contiguous()
        self.kv_plan, self.kv, self.aug_emb = pipe.unet.build_kv_plan(self.n_text, ctxs, self.pooled, time_emb)
        self.step_plan, self.dbg = pipe.unet.build_step_plan(
            batch, h, w, self.latents, self.step_table, self.step_counter, self.kv, guidance_scale=guidance_scale,
            cfg_on=cfg_on, history=self.history, cur_sample=self.cur_sample, aug_emb=self.aug_emb)
        self.vae_plan, self.images = pipe.vae_engine.build_plan(batch, h, w, self.latents)
        f = pipe.vae_scale_factor
        self.img_host = torch.zeros(batch, h * f, w * f, 3, dtype=torch.uint8).pin_memory()
        # one eager pass validates every launch configuration before capture
        self.run_text()
        self.vae_plan.run()
        if pipe.use_cuda_graph:
            self.step_plan.capture()
        else:
            self.step_plan.run()
        torch.cuda.synchronize()

    def run_text(self):
        for pl in self.text_plans:
            pl.run()
        if self.zero_plan is not None:
            self.zero_plan.run()
        self.kv_plan.run()

    def reset(self):
        """Start a new generation from `init_latents` (device -> device)."""
        self.latents.copy_(self.init_latents)
        self.step_counter.zero_()
        self.history.zero_()

    def run(self, num_evals, decode=True, on_step=None):
        self.run_text()
        for i in range(num_evals):
            self.step_plan.launch()
            if on_step is not None:
                on_step(i)
        if decode:
            self.vae_plan.run()

    def launches(self, num_evals, decode=True):
        n = sum(len(pl) for pl in self.text_plans) + len(self.kv_plan) + num_evals * len(self.step_plan)
        n += len(self.zero_plan) if self.zero_plan is not None else 0
        return n + (len(self.vae_plan) if decode else 0)


class StableDiffusionPipeline:
    def __init__(self, cfg: dict, unet_sd: dict, vae_sd: dict, text_sd, tokenizer: CLIPTokenizer | None = None,
                 device: str = "cuda", scheduler: str | None = None, vae_dtype: str | None = None):
        """text_sd: one state dict, or a list with one per entry of cfg["text_encoders"] (SDXL: CLIP-L, OpenCLIP-bigG
        with text_projection)."""
        if not torch.cuda.is_available():
            raise RuntimeError("b200sd needs a CUDA device (sm_100a); there is no CPU path")
        text_sds = list(text_sd) if isinstance(text_sd, (list, tuple)) else [text_sd]
        if len(text_sds) != len(cfg["text_encoders"]) or len(text_sds) > 2:
            raise ValueError("one state dict per text encoder (at most two)")
        self.cfg = cfg
        dev = torch.device(device)
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        torch.cuda.set_device(self.device)
        self.unet = UNetEngine(cfg["unet"], unet_sd, self.device)
        # 16-bit type of the VAE decoder ("fp16" | "bf16"): configs/sdxl.json asks for bf16 (fp32 exponent range: real
        # SDXL VAE activations overflow fp16, upstream force_upcasts that decoder to fp32); the UNet / text encoders
        # stay fp16, whose 11-bit significand the CFG-amplified latents need
        self.vae_engine = VAEDecoderEngine(cfg["vae"], vae_sd, self.device, variant=vae_dtype)
        self.vae = _VaeHandle(self.vae_engine)
        self.text_encoders = [CLIPTextEngine(c, sd, self.device) for c, sd in zip(cfg["text_encoders"], text_sds)]
        self.text_encoder = self.text_encoders[0]
        self.tokenizer = tokenizer or CLIPTokenizer.from_pretrained_or_synthetic(cfg["text_encoders"][0]["vocab_size"])
        self.tokenizers = [self.tokenizer]
        if len(self.text_encoders) == 2:
            # SDXL's tokenizer_2: same vocabulary and merges, but pads with "!" (id 0) instead of <|endoftext|>
            import copy
            tok2 = copy.copy(self.tokenizer)
            tok2.pad_token_id = 0
            self.tokenizers.append(tok2)
        self.is_xl = cfg["unet"].get("addition_embed_type") == "text_time"
        self.scheduler_kind = (scheduler or cfg["scheduler"]["type"]).lower()
        self.vae_scale_factor = self.vae_engine.scale_factor
        self.use_cuda_graph = os.environ.get("B200SD_NO_GRAPH", "0") != "1"
        self._jobs = collections.OrderedDict()
        self.last_timings = {}
        self.model_id = cfg.get("name", "?")

    # ------------------------------------------------------------------ construction (reference configmap.yaml:41)
    @classmethod
    def from_pretrained(cls, model_id: str = "sd15", torch_dtype=None, device: str = "cuda", seed: int = 0,
                        scheduler: str | None = None, random_init: bool | None = None, **_ignored):
        """model_id: a checkpoint directory in diffusers layout (model_index.json, */config.json, *.safetensors,
        tokenizer/, scheduler/ - loaded as is), a config name under configs/ ("sd15", "tiny", "sdxl"), a config .json
        path.  Architecture names give random-init weights (seeded, diffusers key layout; a WARNING is logged).  Anything
        else - a hub-style id such as "runwayml/stable-diffusion-v1-5", a mistyped or unmounted path - raises
        FileNotFoundError unless random_init=True / B200SD_RANDOM_WEIGHTS=1 maps it to the sd15 / sdxl architecture with
        random weights; `torch_dtype` is accepted for signature
        compatibility (compute is 16-bit operands with fp32 accumulation and statistics; lib.act_dtype())."""
        if Ck.is_checkpoint_dir(model_id):
            # a real checkpoint in diffusers layout: configs, safetensors weights, scheduler
            # default (PNDM for SD1.5, Euler for SDXL) and the tokenizer's vocab.json / merges.txt
            cfg, unet_sd, vae_sd, text_sds, tok_dir = Ck.load_diffusers_dir(model_id)
            tok = CLIPTokenizer.from_files(tok_dir) if tok_dir else None
            pipe = cls(cfg, unet_sd, vae_sd, text_sds, tokenizer=tok, device=device, scheduler=scheduler)
            pipe.model_id = model_id
            return pipe
        name = model_id
        explicit = os.path.isfile(model_id) or os.path.exists(os.path.join(Wt.CONFIG_DIR, f"{model_id}.json"))
        if not explicit:
            # neither a checkpoint directory nor an architecture name: a hub id, or a mistyped / unmounted path.  Never
            # answer requests with noise silently - random weights need an explicit opt-in.
            if random_init is None:
                random_init = os.environ.get("B200SD_RANDOM_WEIGHTS", "0") == "1"
            if not random_init:
                raise FileNotFoundError(
                    f"{model_id!r} is neither a checkpoint directory in diffusers layout (model_index.json) nor an "
                    f"architecture name under configs/ ({', '.join(sorted(Wt.config_names()))}); hub ids cannot be "
                    "downloaded here.  Pass random_init=True (or B200SD_RANDOM_WEIGHTS=1) to run that architecture "
                    "with random-init weights.")
            name = "sdxl" if "xl" in model_id.lower() else "sd15"
        logger.warning("b200sd: %r -> architecture %r with RANDOM-INIT weights (seed %d); images are noise",
                       model_id, name, seed)
        cfg = Wt.load_config(name)
        kw = dict(round_dtype=ops.act_dtype(), device=device, store_dtype=ops.act_dtype())  # drawn on the GPU
        unet_sd = Wt.random_state_dict(Wt.unet_schema(cfg["unet"]), seed, **kw)
        vkw = dict(kw)
        if cfg["vae"].get("dtype") == "bf16":
            vkw.update(round_dtype=torch.bfloat16, store_dtype=torch.bfloat16)
        vae_sd = Wt.random_state_dict(Wt.vae_decoder_schema(cfg["vae"]), seed + 1, **vkw)
        text_sd = [Wt.random_state_dict(Wt.clip_text_schema(c), seed + 2 + i, **kw)
                   for i, c in enumerate(cfg["text_encoders"])]
        pipe = cls(cfg, unet_sd, vae_sd, text_sd, device=device, scheduler=scheduler)
        pipe.model_id = model_id
        return pipe

    # memory toggles of the reference (configmap.yaml:42-43, 46-47): identity maths, ignored on 180 GB parts
    def enable_attention_slicing(self, *_a, **_k):
        return None

    def enable_vae_slicing(self):
        return None

    def to(self, *_a, **_k):
        return self

    # ------------------------------------------------------------------ stages
    def job(self, batch, h, w, guidance_scale, cfg_on, zero_negative=False) -> Job:
        """The prepared plans / buffers / captured graph for one SHAPE.  The guidance scale is not part of the key: it
        travels in the device step table (B200SD_ST_GUIDANCE).  At most MAX_JOBS shapes are kept (least recently used
        is dropped, its buffers go back to the allocator) so that client-chosen sizes cannot exhaust HBM."""
        key = (batch, h, w, cfg_on, bool(zero_negative))
        job = self._jobs.get(key)
        if job is None:
            while len(self._jobs) >= max(1, MAX_JOBS):
                self._jobs.popitem(last=False)
                torch.cuda.empty_cache()
            job = self._jobs[key] = Job(self, batch, h, w, guidance_scale, cfg_on, zero_negative)
        self._jobs.move_to_end(key)
        return job

    def make_schedule(self, num_inference_steps, scheduler=None):
        kind = (scheduler or self.scheduler_kind).lower()
        return S.make_schedule(kind, num_inference_steps,
                               **{k: v for k, v in self.cfg["scheduler"].items() if k != "type"})

    def tokenize(self, prompts, negative, cfg_on):
        """int32 [n_encoders * n_text * 77], per encoder in encode_prompt order: [negative...; positive...]."""
        texts = (list(negative) if cfg_on else []) + list(prompts)
        return torch.cat([torch.tensor(tok(texts), dtype=torch.int32).reshape(-1) for tok in self.tokenizers])

    def eos_rows(self, ids: torch.Tensor, n_text: int) -> torch.Tensor:
        """Flat row index (n * 77 + position) of the first EOS token of every sequence of the LAST encoder - where
        CLIPTextModelWithProjection pools (transformers modeling_clip.py: first position equal to eos_token_id)."""
        last = ids.reshape(len(self.tokenizers), n_text, 77)[-1]
        pos = (last == self.tokenizers[-1].eos_token_id).to(torch.int32).argmax(dim=-1)
        return (torch.arange(n_text, dtype=torch.int32) * 77 + pos.to(torch.int32)).to(torch.int32)

    def prepare_latents(self, batch, h, w, generator, init_sigma):
        """randn_tensor semantics: one draw of the full batch for a single generator, one [1,4,h,w] draw per image for
        a list; the draw happens on the generator's own device (CPU generator = the reference's CPU branch,
        configmap.yaml:92) and is returned NCHW fp32."""
        shape = (batch, 4, h, w)
        if isinstance(generator, (list, tuple)):
            if len(generator) != batch:
                raise ValueError(f"You have passed a list of generators of length {len(generator)}, but requested an "
                                 f"effective batch size of {batch}.")
            lat = torch.cat([torch.randn((1,) + shape[1:], generator=g, device=g.device, dtype=torch.float32).cpu()
                             for g in generator], dim=0)
        elif generator is not None:
            lat = torch.randn(shape, generator=generator, device=generator.device, dtype=torch.float32).cpu()
        else:
            lat = torch.randn(shape, dtype=torch.float32)
        return lat * init_sigma

    # ------------------------------------------------------------------ the call (reference configmap.yaml:105-112)
    @torch.no_grad()
    def __call__(self, prompt, num_inference_steps: int = 50, guidance_scale: float = 7.5, width: int | None = None,
                 height: int | None = None, generator=None, latents: torch.Tensor | None = None,
                 negative_prompt=None, output_type: str = "pil", scheduler: str | None = None,
                 callback_on_step_end=None, **_ignored):
        torch.cuda.set_device(self.device)  # calls may come from any thread (the server's batching worker)
        f = self.vae_scale_factor
        height = height or self.cfg["unet"]["sample_size"] * f
        width = width or self.cfg["unet"]["sample_size"] * f
        if height % 8 != 0 or width % 8 != 0:
            raise ValueError(f"`height` and `width` have to be divisible by 8 but are {height} and {width}.")
        if isinstance(prompt, str):
            prompts = [prompt]
        elif isinstance(prompt, (list, tuple)):
            prompts = list(prompt)
        else:
            raise ValueError(f"`prompt` has to be of type `str` or `list` but is {type(prompt)}")
        batch = len(prompts)
        if negative_prompt is None:
            negative = [""] * batch
        elif isinstance(negative_prompt, str):
            negative = [negative_prompt] * batch
        else:
            negative = list(negative_prompt)
            if len(negative) != batch:
                raise ValueError("`negative_prompt` batch size does not match `prompt`")
        cfg_on = guidance_scale > 1.0
        h, w = height // f, width // f
        sched = self.make_schedule(num_inference_steps, scheduler)
        if sched.num_evals > MAX_EVALS:
            raise ValueError("too many inference steps")
        # StableDiffusionXLPipeline (force_zeros_for_empty_prompt): no negative prompt -> zero embeddings, not ""
        zero_negative = self.is_xl and negative_prompt is None
        job = self.job(batch, h, w, guidance_scale, cfg_on, zero_negative)

        # host side: tokenise, draw latents (NCHW, upstream RNG order) -> NHWC pinned staging
        ids = self.tokenize(prompts, negative, cfg_on)
        job.ids_host.copy_(ids)
        job.eos_host.copy_(self.eos_rows(ids, job.n_text))
        if latents is None:
            lat = self.prepare_latents(batch, h, w, generator, sched.init_noise_sigma)
        else:
            lat = latents.to(torch.float32).cpu() * sched.init_noise_sigma
        job.lat_host.copy_(lat.permute(0, 2, 3, 1).reshape(batch, h * w, 4))
        job.table_host[: sched.num_evals].copy_(sched.table)
        job.table_host[: sched.num_evals, L.ST_GUIDANCE] = float(guidance_scale)

        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        job.ids_dev.copy_(job.ids_host, non_blocking=True)
        job.eos_dev.copy_(job.eos_host, non_blocking=True)
        job.init_latents.copy_(job.lat_host, non_blocking=True)
        job.step_table.copy_(job.table_host, non_blocking=True)
        job.reset()
        on_step = None
        if callback_on_step_end is not None:
            def on_step(i):
                cb = job.latents.view(batch, h, w, 4).permute(0, 3, 1, 2).contiguous().cpu()
                callback_on_step_end(self, i, sched.timesteps[i], {"latents": cb})
        decode = output_type != "latent"
        job.run(sched.num_evals, decode=decode, on_step=on_step)
        if not decode:
            out = job.latents.view(batch, h, w, 4).permute(0, 3, 1, 2).contiguous().cpu()
            return PipelineOutput(images=out, latents=out)
        job.img_host.copy_(job.images, non_blocking=True)
        ev[1].record()
        ev[1].synchronize()
        img = job.img_host.numpy().copy()
        self.last_timings = {"gpu_ms": ev[0].elapsed_time(ev[1]), "unet_evals": sched.num_evals,
                             "launches": job.launches(sched.num_evals)}
        if output_type == "u8":
            images = img
        elif output_type == "np":
            images = img.astype(np.float32) / 255.0
        elif output_type == "pil":
            from PIL import Image
            images = [Image.fromarray(a) for a in img]
        else:
            raise ValueError(f"unknown output_type {output_type!r}")
        return PipelineOutput(images=images, nsfw_content_detected=None)


class StableDiffusionXLPipeline(StableDiffusionPipeline):
    """`diffusers.StableDiffusionXLPipeline` surface (BASELINE.json configs[3]): two text encoders (penultimate hidden
    states concatenated, pooled + projected output of the second), `text_time` size conditioning with
    original_size = target_size = (height, width) and no crop, Euler-discrete by default, and zero (not "")
    unconditional embeddings when no negative prompt is given."""

    @classmethod
    def from_pretrained(cls, model_id: str = "sdxl", **kw):
        return super().from_pretrained(model_id, **kw)


def make_generator(seed: int, device: str = "cpu") -> torch.Generator:
    return torch.Generator(device=device).manual_seed(seed)
