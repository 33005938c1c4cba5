"""Oracle text-to-image pipeline (fp32 CPU) — restates diffusers==0.30.2
`pipelines/stable_diffusion/pipeline_stable_diffusion.py::StableDiffusionPipeline.__call__` as invoked by the
reference at configmap.yaml:105-112 on its CPU branch (TORCH_DTYPE float32, no autocast, CPU generator:
configmap.yaml:31, 92, 103), with `safety_checker=None` (the checker is outside the hot path).

    tokenise -> CLIP text ([negative; positive]) -> randn latents * init_noise_sigma -> for t: UNet on the
    duplicated batch, CFG combine, scheduler.step -> VAE decode(latents / scaling_factor) -> clamp(x/2+.5) ->
    NHWC -> (x*255).round().astype(uint8)

TEST INFRASTRUCTURE — see oracle/__init__.py.  Also the timed CPU baseline of bench.py.
"""
from __future__ import annotations

import numpy as np
import torch

from .clip import CLIPTextModel
from .schedulers import make_scheduler
from .unet import UNet2DConditionModel
from .vae import AutoencoderKLDecoder


class OraclePipeline:
    def __init__(self, cfg: dict, unet_sd: dict, vae_sd: dict, text_sd, tokenize, scheduler: str | None = None,
                 eos_token_id: int | None = None):
        """tokenize: callable(list[str]) -> list[list[int]] (oracle.tokenizer.OracleTokenizer, or golden ids); for two
        text encoders (SDXL) a list of two callables (the second pads with id 0) and text_sd a list of two dicts."""
        self.cfg = cfg
        self.unet = UNet2DConditionModel(cfg["unet"]).eval()
        self.vae = AutoencoderKLDecoder(cfg["vae"]).eval()
        text_sds = list(text_sd) if isinstance(text_sd, (list, tuple)) else [text_sd]
        self.text_encoders = [CLIPTextModel(c).eval() for c in cfg["text_encoders"]]
        self.text_encoder = self.text_encoders[0]
        self.unet.load_state_dict(unet_sd, strict=True)
        self.vae.load_state_dict(vae_sd, strict=True)
        for enc, sd in zip(self.text_encoders, text_sds):
            enc.load_state_dict(sd, strict=True)
        self.tokenizers = list(tokenize) if isinstance(tokenize, (list, tuple)) else [tokenize]
        self.tokenize = self.tokenizers[0]
        self.eos_token_id = eos_token_id
        self.is_xl = cfg["unet"].get("addition_embed_type") == "text_time"
        self.scheduler_kind = scheduler or cfg["scheduler"]["type"]

    def _encode_xl(self, texts):
        """diffusers 0.30.2 pipeline_stable_diffusion_xl.py::encode_prompt for one list of texts: penultimate hidden
        states of both encoders concatenated on the channel axis, pooled text_embeds of the second."""
        hs, pooled = [], None
        for enc, tok in zip(self.text_encoders, self.tokenizers):
            h, p = enc.encode_xl(torch.tensor(tok(texts), dtype=torch.long), self.eos_token_id)
            hs.append(h)
            pooled = p if p is not None else pooled
        return torch.cat(hs, dim=-1), pooled

    @torch.no_grad()
    def __call__(self, prompt, num_inference_steps=50, guidance_scale=7.5, width=None, height=None, generator=None,
                 latents=None, negative_prompt=None, scheduler=None, decode=True, max_evals=None):
        f = self.vae.scale_factor
        height = height or self.cfg["unet"]["sample_size"] * f
        width = width or self.cfg["unet"]["sample_size"] * f
        prompts = [prompt] if isinstance(prompt, str) else list(prompt)
        b = len(prompts)
        neg = [negative_prompt or ""] * b if not isinstance(negative_prompt, (list, tuple)) else list(negative_prompt)
        do_cfg = guidance_scale > 1.0
        added = None
        if self.is_xl:
            emb, pooled = self._encode_xl(prompts)
            if do_cfg:
                if negative_prompt is None:  # force_zeros_for_empty_prompt (SDXL-base model_index.json)
                    neg_emb, neg_pooled = torch.zeros_like(emb), torch.zeros_like(pooled)
                else:
                    neg_emb, neg_pooled = self._encode_xl(neg)
                emb, pooled = torch.cat([neg_emb, emb]), torch.cat([neg_pooled, pooled])
            # _get_add_time_ids: original_size + crops_coords_top_left + target_size, identical for both halves
            tid = torch.tensor([[height, width, 0, 0, height, width]], dtype=torch.float32).repeat(emb.shape[0], 1)
            added = {"text_embeds": pooled, "time_ids": tid}
        else:
            ids = torch.tensor(self.tokenize(prompts), dtype=torch.long)
            emb = self.text_encoder(ids)
            if do_cfg:
                neg_emb = self.text_encoder(torch.tensor(self.tokenize(neg), dtype=torch.long))
                emb = torch.cat([neg_emb, emb])
        sched = make_scheduler(self.cfg["scheduler"], scheduler or self.scheduler_kind)
        sched.set_timesteps(num_inference_steps)
        shape = (b, 4, height // f, width // f)
        if latents is None:
            if isinstance(generator, (list, tuple)):
                latents = torch.cat([torch.randn((1,) + shape[1:], generator=g, dtype=torch.float32) for g in generator])
            else:
                latents = torch.randn(shape, generator=generator, dtype=torch.float32)
        latents = latents * sched.init_noise_sigma
        per_step = []
        eps_per_step = []
        for i, t in enumerate(sched.timesteps):
            if max_evals is not None and i >= max_evals:
                break
            x_in = torch.cat([latents] * 2) if do_cfg else latents
            x_in = sched.scale_model_input(x_in, t)
            eps = self.unet(x_in, t.to(torch.float32), emb, added)
            eps_per_step.append(eps.clone())
            if do_cfg:
                e_u, e_c = eps.chunk(2)
                eps = e_u + guidance_scale * (e_c - e_u)
            latents = sched.step(eps, t, latents)
            per_step.append(latents.clone())
        out = {"latents": latents, "per_step": per_step, "eps": eps_per_step, "text": emb, "added": added,
               "timesteps": [int(t) for t in sched.timesteps]}
        if decode:
            image = self.vae.decode(latents / self.cfg["vae"]["scaling_factor"])
            out["image_f32"] = image
            img = (image / 2 + 0.5).clamp(0, 1).permute(0, 2, 3, 1).float().numpy()
            out["images"] = (img * 255).round().astype("uint8")
        return out
