"""Test helper: write a checkpoint directory in diffusers / HF-hub layout (config.json files in UPSTREAM key names,
weights with the upstream `safetensors` library) from this repo's config + state dicts."""
import json
import os

import torch


def _dump(path, obj):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        json.dump(obj, f)


def write_diffusers_dir(root, cfg, unet_sd, vae_sd, text_sds, tokenizer_mod, dtype=torch.float16, legacy_vae=True,
                        scheduler_class="PNDMScheduler"):
    from safetensors.torch import save_file
    u, v = cfg["unet"], cfg["vae"]
    xl = u.get("addition_embed_type") == "text_time"
    _dump(os.path.join(root, "model_index.json"),
          {"_class_name": "StableDiffusionXLPipeline" if xl else "StableDiffusionPipeline", "_diffusers_version": "0.30.2"})
    heads = u["num_attention_heads"]
    _dump(os.path.join(root, "unet", "config.json"), {
        "_class_name": "UNet2DConditionModel", "in_channels": u["in_channels"], "out_channels": u["out_channels"],
        "sample_size": u["sample_size"], "block_out_channels": u["block_out_channels"],
        "layers_per_block": u["layers_per_block"], "down_block_types": u["down_block_types"],
        "up_block_types": u["up_block_types"], "attention_head_dim": heads if xl else heads[0],
        "transformer_layers_per_block": u["transformer_layers_per_block"] if xl else 1,
        "cross_attention_dim": u["cross_attention_dim"], "norm_num_groups": u["norm_num_groups"],
        "norm_eps": u["norm_eps"], "use_linear_projection": u["use_linear_projection"], "act_fn": "silu",
        "flip_sin_to_cos": True, "freq_shift": 0, "addition_embed_type": u.get("addition_embed_type"),
        "addition_time_embed_dim": u.get("addition_time_embed_dim"),
        "projection_class_embeddings_input_dim": u.get("projection_class_embeddings_input_dim")})
    _dump(os.path.join(root, "vae", "config.json"), {
        "_class_name": "AutoencoderKL", "latent_channels": v["latent_channels"], "out_channels": v["out_channels"],
        "in_channels": 3, "block_out_channels": v["block_out_channels"], "layers_per_block": v["layers_per_block"],
        "norm_num_groups": v["norm_num_groups"], "scaling_factor": v["scaling_factor"],
        "force_upcast": v.get("dtype") == "bf16"})
    _dump(os.path.join(root, "scheduler", "scheduler_config.json"), {
        "_class_name": scheduler_class, "beta_start": cfg["scheduler"]["beta_start"], "beta_end": cfg["scheduler"]["beta_end"],
        "beta_schedule": "scaled_linear", "num_train_timesteps": 1000, "steps_offset": 1, "skip_prk_steps": True,
        "set_alpha_to_one": False, "prediction_type": "epsilon"})
    cast = lambda sd: {k: t.to(dtype).contiguous() for k, t in sd.items()}  # noqa: E731
    save_file(cast(unet_sd), os.path.join(root, "unet", "diffusion_pytorch_model.safetensors"))
    vae_out = {}
    ren = {"to_q": "query", "to_k": "key", "to_v": "value", "to_out.0": "proj_attn"}
    for k, t in vae_sd.items():
        if legacy_vae and ".attentions." in k:
            for new, old in ren.items():
                k = k.replace("." + new + ".", "." + old + ".")
        vae_out[k] = t
    vae_out["encoder.conv_in.weight"] = torch.zeros(8, 3, 3, 3)  # the encoder half is present in real files: ignored
    vae_out["quant_conv.weight"] = torch.zeros(8, 8, 1, 1)
    os.makedirs(os.path.join(root, "vae"), exist_ok=True)
    save_file(cast(vae_out), os.path.join(root, "vae", "diffusion_pytorch_model.fp16.safetensors"))
    for i, (tc, sd) in enumerate(zip(cfg["text_encoders"], text_sds)):
        d = os.path.join(root, "text_encoder" if i == 0 else "text_encoder_2")
        proj = tc.get("projection_dim")
        _dump(os.path.join(d, "config.json"), {
            "architectures": ["CLIPTextModelWithProjection" if proj else "CLIPTextModel"], "vocab_size": tc["vocab_size"],
            "hidden_size": tc["hidden_size"], "intermediate_size": tc["intermediate_size"],
            "num_hidden_layers": tc["num_layers"], "num_attention_heads": tc["num_heads"],
            "max_position_embeddings": tc["max_position"], "hidden_act": tc["act"], "layer_norm_eps": tc["eps"],
            "projection_dim": proj or 512})
        out = cast(sd)
        out["text_model.embeddings.position_ids"] = torch.arange(77)[None]  # old transformers checkpoints carry it
        save_file(out, os.path.join(d, "model.safetensors"))
    os.makedirs(os.path.join(root, "tokenizer"), exist_ok=True)
    tokenizer_mod.write_vocab_files(os.path.join(root, "tokenizer"), cfg["text_encoders"][0]["vocab_size"])
    return root
