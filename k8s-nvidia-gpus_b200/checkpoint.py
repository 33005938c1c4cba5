"""Real-weight loading: a diffusers-layout checkpoint directory -> (config, state dicts, tokenizer)
(SURVEY.md §8(f) rank 3, A.6).

What `StableDiffusionPipeline.from_pretrained(MODEL_ID)` finds in the reference's model cache
(cluster-config/apps/sd15-api/configmap.yaml:41, deployment.yaml:45-46; HF hub snapshot layout):

    model_index.json
    unet/config.json            unet/diffusion_pytorch_model[.fp16].safetensors
    vae/config.json             vae/diffusion_pytorch_model[.fp16].safetensors
    text_encoder/config.json    text_encoder/model[.fp16].safetensors        (+ text_encoder_2/ for SDXL)
    tokenizer/vocab.json        tokenizer/merges.txt
    scheduler/scheduler_config.json

The safetensors container is read directly (8-byte little-endian header length, JSON header, raw little-endian
tensors) - no dependency beyond torch - and validated against weights.py's schema: every key and shape must match.
"""
from __future__ import annotations

import json
import os
import struct

import numpy as np
import torch

from . import weights as Wt

_DTYPES = {"F32": (torch.float32, 4), "F16": (torch.float16, 2), "BF16": (torch.bfloat16, 2), "F64": (torch.float64, 8),
           "I64": (torch.int64, 8), "I32": (torch.int32, 4), "U8": (torch.uint8, 1), "BOOL": (torch.bool, 1)}


def read_safetensors(path: str, keep=None) -> dict:
    """{name: tensor} (dtype as stored).  keep: optional predicate on the tensor name (skipped tensors are not read)."""
    out = {}
    with open(path, "rb") as f:
        (hlen,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(hlen))
        base = 8 + hlen
        mm = np.memmap(path, dtype=np.uint8, mode="r")
        for name, info in header.items():
            if name == "__metadata__" or (keep is not None and not keep(name)):
                continue
            if info["dtype"] not in _DTYPES:
                raise ValueError(f"{path}: tensor {name} has unsupported dtype {info['dtype']}")
            dt, size = _DTYPES[info["dtype"]]
            b, e = info["data_offsets"]
            n = int(np.prod(info["shape"])) if info["shape"] else 1
            if e - b != n * size:
                raise ValueError(f"{path}: tensor {name} has {e - b} bytes, expected {n * size}")
            raw = torch.from_numpy(np.array(mm[base + b:base + e]))  # private copy: the file can go away
            out[name] = raw.view(dt).reshape(info["shape"])
    return out


def _find(dirname: str, stems) -> str:
    for stem in stems:
        for suffix in (".safetensors", ".fp16.safetensors"):
            p = os.path.join(dirname, stem + suffix)
            if os.path.exists(p):
                return p
    raise FileNotFoundError(f"no safetensors weights ({', '.join(stems)}) under {dirname}")


def _json(path: str) -> dict:
    with open(path) as f:
        return json.load(f)


# ------------------------------------------------------------------------------------------------ config translation
def unet_config(c: dict) -> dict:
    """diffusers UNet2DConditionModel config.json -> configs/*.json "unet" entry."""
    ch = list(c["block_out_channels"])
    n = len(ch)
    ahd = c.get("num_attention_heads") or c["attention_head_dim"]  # diffusers' misnomer: this IS the head count
    heads = list(ahd) if isinstance(ahd, (list, tuple)) else [ahd] * n
    depth = c.get("transformer_layers_per_block", 1)
    depth = list(depth) if isinstance(depth, (list, tuple)) else [depth] * n
    for key, want in (("act_fn", "silu"), ("flip_sin_to_cos", True), ("freq_shift", 0), ("center_input_sample", False),
                      ("upcast_attention", False), ("resnet_time_scale_shift", "default"), ("class_embed_type", None),
                      ("only_cross_attention", False), ("dual_cross_attention", False)):
        if c.get(key, want) != want:
            raise NotImplementedError(f"UNet config {key}={c[key]!r} is outside the supported path (expects {want!r})")
    return {"in_channels": c["in_channels"], "out_channels": c["out_channels"], "sample_size": c["sample_size"],
            "block_out_channels": ch, "layers_per_block": c["layers_per_block"],
            "down_block_types": list(c["down_block_types"]), "up_block_types": list(c["up_block_types"]),
            "transformer_layers_per_block": depth, "num_attention_heads": heads,
            "cross_attention_dim": c["cross_attention_dim"], "norm_num_groups": c["norm_num_groups"],
            "norm_eps": c["norm_eps"], "use_linear_projection": bool(c.get("use_linear_projection", False)),
            "addition_embed_type": c.get("addition_embed_type"),
            "addition_time_embed_dim": c.get("addition_time_embed_dim"),
            "projection_class_embeddings_input_dim": c.get("projection_class_embeddings_input_dim")}


def vae_config(c: dict) -> dict:
    return {"latent_channels": c["latent_channels"], "out_channels": c["out_channels"],
            "block_out_channels": list(c["block_out_channels"]), "layers_per_block": c["layers_per_block"],
            "norm_num_groups": c["norm_num_groups"], "scaling_factor": c.get("scaling_factor", 0.18215),
            # upstream runs a force_upcast VAE (SDXL) in fp32 because its activations overflow fp16: bf16 image here
            "dtype": "bf16" if c.get("force_upcast", False) else "fp16"}


def text_config(c: dict, penultimate: bool) -> dict:
    out = {"vocab_size": c["vocab_size"], "hidden_size": c["hidden_size"], "intermediate_size": c["intermediate_size"],
           "num_layers": c["num_hidden_layers"], "num_heads": c["num_attention_heads"],
           "max_position": c["max_position_embeddings"], "act": c["hidden_act"], "eps": c.get("layer_norm_eps", 1e-5),
           "output": "penultimate" if penultimate else "last_hidden_state"}
    if "WithProjection" in "".join(c.get("architectures", [])):
        out["projection_dim"] = c["projection_dim"]
    return out


_SCHED = {"PNDMScheduler": "pndm", "DDIMScheduler": "ddim", "EulerDiscreteScheduler": "euler"}


def scheduler_config(c: dict) -> dict:
    cls = c.get("_class_name", "PNDMScheduler")
    if cls not in _SCHED:
        raise NotImplementedError(f"scheduler {cls} (implemented: {sorted(_SCHED)})")
    if c.get("beta_schedule", "scaled_linear") != "scaled_linear":
        raise NotImplementedError(f"beta_schedule {c['beta_schedule']!r}")
    if c.get("prediction_type", "epsilon") != "epsilon":
        raise NotImplementedError(f"prediction_type {c['prediction_type']!r}")
    # what make_schedule() hard-codes; a checkpoint that asks for anything else must not load silently
    want = {"timestep_spacing": "leading", "set_alpha_to_one": False, "clip_sample": False, "thresholding": False,
            "use_karras_sigmas": False, "rescale_betas_zero_snr": False, "interpolation_type": "linear"}
    if cls == "PNDMScheduler":
        want["skip_prk_steps"] = True
    if cls == "DDIMScheduler" and "clip_sample" not in c:
        raise NotImplementedError("DDIMScheduler without clip_sample=false (upstream default is true: x0 clipping)")
    for k, v in want.items():
        if k in c and c[k] != v:
            raise NotImplementedError(f"scheduler_config {k}={c[k]!r} is not implemented (only {v!r})")
    return {"type": _SCHED[cls], "steps_offset": c.get("steps_offset", 0), "beta_start": c["beta_start"],
            "beta_end": c["beta_end"], "num_train_timesteps": c.get("num_train_timesteps", 1000)}


# ------------------------------------------------------------------------------------------------ state dicts
_VAE_LEGACY = {"query": "to_q", "key": "to_k", "value": "to_v", "proj_attn": "to_out.0"}


def vae_decoder_state_dict(sd: dict) -> dict:
    """Decoder half of an AutoencoderKL checkpoint; pre-0.18 attention names (query/key/value/proj_attn, as in the
    original runwayml/stable-diffusion-v1-5 files) are mapped the way diffusers' loader does."""
    out = {}
    for k, v in sd.items():
        if not (k.startswith("decoder.") or k.startswith("post_quant_conv.")):
            continue
        parts = k.split(".")
        if "attentions" in parts and parts[-2] in _VAE_LEGACY:
            parts[-2:-1] = _VAE_LEGACY[parts[-2]].split(".")
            k = ".".join(parts)
        if "attentions" in parts and v.dim() == 4:  # 1x1-conv form of the attention projections in LDM-era files
            v = v[:, :, 0, 0]
        out[k] = v
    return out


def text_state_dict(sd: dict) -> dict:
    return {k: v for k, v in sd.items() if not k.endswith("position_ids")}


def check_against_schema(name: str, sd: dict, schema) -> dict:
    missing = [k for k in schema if k not in sd]
    extra = [k for k in sd if k not in schema]
    bad = [f"{k}: {tuple(sd[k].shape)} != {tuple(schema[k][1])}" for k in schema
           if k in sd and tuple(sd[k].shape) != tuple(schema[k][1])]
    if missing or extra or bad:
        raise ValueError(f"{name} checkpoint does not match the architecture: {len(missing)} missing "
                         f"(e.g. {missing[:3]}), {len(extra)} unexpected (e.g. {extra[:3]}), {len(bad)} shape "
                         f"mismatches (e.g. {bad[:3]})")
    return {k: sd[k] for k in schema}


def is_checkpoint_dir(path: str) -> bool:
    return os.path.isdir(path) and os.path.exists(os.path.join(path, "model_index.json"))


def load_diffusers_dir(path: str):
    """-> (cfg, unet_sd, vae_sd, [text_sd, ...], tokenizer_dir or None).  Tensors stay in their stored dtype on the
    CPU; engine.DeviceWeights converts per use."""
    index = _json(os.path.join(path, "model_index.json"))
    xl = "XL" in index.get("_class_name", "")
    ucfg = unet_config(_json(os.path.join(path, "unet", "config.json")))
    vcfg = vae_config(_json(os.path.join(path, "vae", "config.json")))
    enc_dirs = ["text_encoder"] + (["text_encoder_2"] if os.path.isdir(os.path.join(path, "text_encoder_2")) else [])
    tcfgs = [text_config(_json(os.path.join(path, d, "config.json")), penultimate=xl) for d in enc_dirs]
    scfg = scheduler_config(_json(os.path.join(path, "scheduler", "scheduler_config.json")))
    cfg = {"name": os.path.basename(os.path.normpath(path)), "unet": ucfg, "vae": vcfg, "text_encoders": tcfgs,
           "scheduler": scfg}
    unet_sd = check_against_schema("unet", read_safetensors(_find(os.path.join(path, "unet"),
                                                                 ["diffusion_pytorch_model"])), Wt.unet_schema(ucfg))
    vae_raw = read_safetensors(_find(os.path.join(path, "vae"), ["diffusion_pytorch_model"]),
                               keep=lambda k: k.startswith("decoder.") or k.startswith("post_quant_conv."))
    vae_sd = check_against_schema("vae", vae_decoder_state_dict(vae_raw), Wt.vae_decoder_schema(vcfg))
    text_sds = [check_against_schema(d, text_state_dict(read_safetensors(_find(os.path.join(path, d), ["model"]))),
                                     Wt.clip_text_schema(c)) for d, c in zip(enc_dirs, tcfgs)]
    tok_dir = os.path.join(path, "tokenizer")
    if not (os.path.exists(os.path.join(tok_dir, "vocab.json")) and os.path.exists(os.path.join(tok_dir, "merges.txt"))):
        tok_dir = None
    return cfg, unet_sd, vae_sd, text_sds, tok_dir
