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

import os
from types import SimpleNamespace

import numpy as np
import torch

from . import ops
from . import schedulers as S
from . import weights as Wt
from .clip import CLIPTextEngine
from .tokenizer import CLIPTokenizer
from .unet import UNetEngine
from .vae import VAEDecoderEngine

MAX_EVALS = 1024


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


class StableDiffusionPipeline:
    def __init__(self, cfg: dict, unet_sd: dict, vae_sd: dict, text_sd: dict, tokenizer: CLIPTokenizer | None = None,
                 device: str = "cuda", scheduler: str | None = None):
        if not torch.cuda.is_available():
            raise RuntimeError("b200sd needs a CUDA device (sm_100a); there is no CPU path")
        if len(cfg["text_encoders"]) != 1 or cfg["unet"].get("addition_embed_type"):
            raise NotImplementedError("dual text encoders / text_time conditioning (SDXL) are not wired yet")
        self.cfg = cfg
        dev = torch.device(device)
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        torch.cuda.set_device(self.device)
        self.unet = UNetEngine(cfg["unet"], unet_sd, self.device)
        self.vae_engine = VAEDecoderEngine(cfg["vae"], vae_sd, self.device)
        self.vae = _VaeHandle(self.vae_engine)
        self.text_encoder = CLIPTextEngine(cfg["text_encoders"][0], text_sd, self.device)
        self.tokenizer = tokenizer or CLIPTokenizer.from_pretrained_or_synthetic(cfg["text_encoders"][0]["vocab_size"])
        self.scheduler_kind = (scheduler or cfg["scheduler"]["type"]).lower()
        self.vae_scale_factor = self.vae_engine.scale_factor
        self.use_cuda_graph = os.environ.get("B200SD_NO_GRAPH", "0") != "1"
        self._text_cache = {}
        self._step_cache = {}
        self._vae_cache = {}
        self.step_table = torch.zeros(MAX_EVALS, 16, dtype=torch.float32, device=self.device)
        self.step_counter = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.last_timings = {}

    # ------------------------------------------------------------------ construction (reference configmap.yaml:41)
    @classmethod
    def from_pretrained(cls, model_id: str = "sd15", torch_dtype=None, device: str = "cuda", seed: int = 0,
                        scheduler: str | None = None, **_ignored):
        """model_id: a config name under configs/ ("sd15", "tiny"), a config .json path, or a hub-style id such as
        "runwayml/stable-diffusion-v1-5" (mapped to the sd15 architecture).  No checkpoints exist offline, so
        weights are random-init (seeded) in the diffusers key layout; `torch_dtype` is accepted for signature
        compatibility (compute is bf16 operands / fp32 accumulation and statistics)."""
        name = model_id
        if not os.path.exists(model_id) and not os.path.exists(os.path.join(Wt.CONFIG_DIR, f"{model_id}.json")):
            low = model_id.lower()
            name = "sdxl" if "xl" in low else "sd15"
        cfg = Wt.load_config(name)
        unet_sd = Wt.random_state_dict(Wt.unet_schema(cfg["unet"]), seed)
        vae_sd = Wt.random_state_dict(Wt.vae_decoder_schema(cfg["vae"]), seed + 1)
        text_sd = Wt.random_state_dict(Wt.clip_text_schema(cfg["text_encoders"][0]), seed + 2)
        return cls(cfg, unet_sd, vae_sd, text_sd, device=device, scheduler=scheduler)

    # memory toggles of the reference (configmap.yaml:42-43, 46-47): identity maths, ignored on 180 GB parts
    def enable_attention_slicing(self, *_a, **_k):
        return None

    def enable_vae_slicing(self):
        return None

    def to(self, *_a, **_k):
        return self

    # ------------------------------------------------------------------ stages
    def _encode(self, prompts, negative):
        """[uncond; cond] bf16 [2B, 77, C] — StableDiffusionPipeline.encode_prompt ordering."""
        texts = list(negative) + list(prompts)
        ids = torch.tensor(self.tokenizer(texts), dtype=torch.int32)
        n = len(texts)
        if n not in self._text_cache:
            ids_dev = torch.zeros(n * 77, dtype=torch.int32, device=self.device)
            plan, out = self.text_encoder.build_plan(n, ids_dev)
            kv_plan, kv = self.unet.build_kv_plan(n, out)
            self._text_cache[n] = (ids_dev, plan, out, kv_plan, kv)
        ids_dev, plan, out, kv_plan, kv = self._text_cache[n]
        ids_dev.copy_(ids.reshape(-1), non_blocking=True)
        plan.run()
        kv_plan.run()
        return ids, out, kv

    def _step_plan(self, batch, h, w, kv, guidance_scale, cfg_on, n_text):
        key = (batch, h, w, float(guidance_scale), cfg_on, n_text)
        if key not in self._step_cache:
            latents = torch.zeros(batch, h * w, 4, dtype=torch.float32, device=self.device)
            history = torch.zeros(4, batch, h * w, 4, dtype=torch.float32, device=self.device)
            cur = torch.zeros(batch, h * w, 4, dtype=torch.float32, device=self.device)
            plan, dbg = self.unet.build_step_plan(batch, h, w, latents, self.step_table, self.step_counter, kv,
                                                  guidance_scale=guidance_scale, cfg_on=cfg_on, history=history,
                                                  cur_sample=cur)
            if self.use_cuda_graph:
                saved = (latents.clone(), self.step_counter.clone())
                plan.capture()
                latents.copy_(saved[0])
                self.step_counter.copy_(saved[1])
            self._step_cache[key] = (plan, latents, history, cur, dbg)
        return self._step_cache[key]

    def _vae_plan(self, batch, h, w, latents):
        key = (batch, h, w, latents.data_ptr())
        if key not in self._vae_cache:
            self._vae_cache[key] = self.vae_engine.build_plan(batch, h, w, latents)
        return self._vae_cache[key]

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
        kind = (scheduler or self.scheduler_kind).lower()
        sched = S.make_schedule(kind, num_inference_steps, **{k: v for k, v in self.cfg["scheduler"].items()
                                                              if k != "type"})
        if sched.num_evals > MAX_EVALS:
            raise ValueError("too many inference steps")

        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        _ids, _ctx, kv = self._encode(prompts, negative if cfg_on else [])
        n_text = len(prompts) + (len(negative) if cfg_on else 0)
        ev[1].record()

        plan, lat_dev, history, cur, _dbg = self._step_plan(batch, h, w, kv, guidance_scale, cfg_on, n_text)
        if latents is None:
            latents = self.prepare_latents(batch, h, w, generator, sched.init_noise_sigma)
        else:
            latents = latents.to(torch.float32).cpu() * sched.init_noise_sigma
        # NCHW (upstream layout, RNG order) -> NHWC device latents
        lat_dev.copy_(latents.permute(0, 2, 3, 1).reshape(batch, h * w, 4).contiguous(), non_blocking=True)
        self.step_table[: sched.num_evals].copy_(sched.table, non_blocking=True)
        self.step_counter.zero_()
        if kind == "pndm":
            history.zero_()
        for i, t in enumerate(sched.timesteps):
            plan.launch()
            if callback_on_step_end is not None:
                cb_lat = lat_dev.view(batch, h, w, 4).permute(0, 3, 1, 2).contiguous().cpu()
                callback_on_step_end(self, i, t, {"latents": cb_lat})
        ev[2].record()
        if output_type == "latent":
            out = lat_dev.view(batch, h, w, 4).permute(0, 3, 1, 2).contiguous().cpu()
            return PipelineOutput(images=out, latents=out)
        vplan, img_dev = self._vae_plan(batch, h, w, lat_dev)
        vplan.run()
        ev[3].record()
        img = img_dev.cpu().numpy()  # [B, H, W, 3] uint8 (device -> host copy synchronises)
        self.last_timings = {"text_ms": ev[0].elapsed_time(ev[1]), "denoise_ms": ev[1].elapsed_time(ev[2]),
                             "vae_ms": ev[2].elapsed_time(ev[3]), "unet_evals": sched.num_evals,
                             "launches_per_step": plan.kernel_launches(), "vae_launches": vplan.kernel_launches()}
        final_lat = None
        if output_type == "u8":
            images = img
        elif output_type == "np":
            images = img.astype(np.float32) / 255.0
        elif output_type == "pil":
            from PIL import Image
            images = [Image.fromarray(a) for a in img]
        else:
            raise ValueError(f"unknown output_type {output_type!r}")
        return PipelineOutput(images=images, nsfw_content_detected=None, latents=final_lat)


def make_generator(seed: int, device: str = "cpu") -> torch.Generator:
    return torch.Generator(device=device).manual_seed(seed)
