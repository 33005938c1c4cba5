"""Weight schema (diffusers / transformers state-dict key layout) and random initialisation.

`from_pretrained` in the reference (configmap.yaml:41) loads real checkpoints; offline there are none, so the
benchmarks and parity tests use random-init weights of the named architecture.  The schema below is what a real
safetensors checkpoint in diffusers layout provides, so such a checkpoint loads unchanged.
"""
from __future__ import annotations

import json
import math
import os
from collections import OrderedDict

import torch

CONFIG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs")


def config_names():
    """Architecture names available under configs/."""
    return [f[:-5] for f in os.listdir(CONFIG_DIR) if f.endswith(".json")]


def load_config(name_or_path: str) -> dict:
    path = name_or_path if os.path.exists(name_or_path) else os.path.join(CONFIG_DIR, f"{name_or_path}.json")
    with open(path) as f:
        return json.load(f)


# ------------------------------------------------------------------------------------------------ shape enumerators
def _lin(d, name, cin, cout, bias=True):
    d[name + ".weight"] = ("linear", (cout, cin))
    if bias:
        d[name + ".bias"] = ("bias", (cout,), cin)


def _conv(d, name, cin, cout, k):
    d[name + ".weight"] = ("conv", (cout, cin, k, k))
    d[name + ".bias"] = ("bias", (cout,), cin * k * k)


def _norm(d, name, c):
    d[name + ".weight"] = ("ones", (c,))
    d[name + ".bias"] = ("zeros", (c,))


def _resnet(d, p, cin, cout, temb):
    _norm(d, p + ".norm1", cin)
    _conv(d, p + ".conv1", cin, cout, 3)
    if temb:
        _lin(d, p + ".time_emb_proj", temb, cout)
    _norm(d, p + ".norm2", cout)
    _conv(d, p + ".conv2", cout, cout, 3)
    if cin != cout:
        _conv(d, p + ".conv_shortcut", cin, cout, 1)


def _transformer(d, p, c, ctx, depth, linear_proj):
    _norm(d, p + ".norm", c)
    if linear_proj:
        _lin(d, p + ".proj_in", c, c)
    else:
        _conv(d, p + ".proj_in", c, c, 1)
    for i in range(depth):
        b = f"{p}.transformer_blocks.{i}"
        _norm(d, b + ".norm1", c)
        for n, kdim in (("attn1", c), ("attn2", ctx)):
            _lin(d, f"{b}.{n}.to_q", c, c, bias=False)
            _lin(d, f"{b}.{n}.to_k", kdim, c, bias=False)
            _lin(d, f"{b}.{n}.to_v", kdim, c, bias=False)
            _lin(d, f"{b}.{n}.to_out.0", c, c)
        _norm(d, b + ".norm2", c)
        _norm(d, b + ".norm3", c)
        _lin(d, b + ".ff.net.0.proj", c, 8 * c)
        _lin(d, b + ".ff.net.2", 4 * c, c)
    if linear_proj:
        _lin(d, p + ".proj_out", c, c)
    else:
        _conv(d, p + ".proj_out", c, c, 1)


def unet_schema(cfg: dict) -> OrderedDict:
    d = OrderedDict()
    ch = cfg["block_out_channels"]
    temb = ch[0] * 4
    ctx, lin = cfg["cross_attention_dim"], cfg["use_linear_projection"]
    heads, depth, layers = cfg["num_attention_heads"], cfg["transformer_layers_per_block"], cfg["layers_per_block"]
    _conv(d, "conv_in", cfg["in_channels"], ch[0], 3)
    _lin(d, "time_embedding.linear_1", ch[0], temb)
    _lin(d, "time_embedding.linear_2", temb, temb)
    if cfg.get("addition_embed_type") == "text_time":
        _lin(d, "add_embedding.linear_1", cfg["projection_class_embeddings_input_dim"], temb)
        _lin(d, "add_embedding.linear_2", temb, temb)
    cout = ch[0]
    for i, t in enumerate(cfg["down_block_types"]):
        cin, cout = cout, ch[i]
        for j in range(layers):
            _resnet(d, f"down_blocks.{i}.resnets.{j}", cin if j == 0 else cout, cout, temb)
            if t.startswith("CrossAttn"):
                _transformer(d, f"down_blocks.{i}.attentions.{j}", cout, ctx, depth[i], lin)
        if i < len(ch) - 1:
            _conv(d, f"down_blocks.{i}.downsamplers.0.conv", cout, cout, 3)
    _resnet(d, "mid_block.resnets.0", ch[-1], ch[-1], temb)
    _transformer(d, "mid_block.attentions.0", ch[-1], ctx, depth[-1], lin)
    _resnet(d, "mid_block.resnets.1", ch[-1], ch[-1], temb)
    rch, rdepth = ch[::-1], depth[::-1]
    cout = rch[0]
    for i, t in enumerate(cfg["up_block_types"]):
        cprev, cout = cout, rch[i]
        cin = rch[min(i + 1, len(ch) - 1)]
        for j in range(layers + 1):
            skip = cin if j == layers else cout
            rin = cprev if j == 0 else cout
            _resnet(d, f"up_blocks.{i}.resnets.{j}", rin + skip, cout, temb)
            if t.startswith("CrossAttn"):
                _transformer(d, f"up_blocks.{i}.attentions.{j}", cout, ctx, rdepth[i], lin)
        if i < len(ch) - 1:
            _conv(d, f"up_blocks.{i}.upsamplers.0.conv", cout, cout, 3)
    _norm(d, "conv_norm_out", ch[0])
    _conv(d, "conv_out", ch[0], cfg["out_channels"], 3)
    return d


def vae_decoder_schema(cfg: dict) -> OrderedDict:
    d = OrderedDict()
    ch = cfg["block_out_channels"]
    lc = cfg["latent_channels"]
    _conv(d, "post_quant_conv", lc, lc, 1)
    _conv(d, "decoder.conv_in", lc, ch[-1], 3)
    _resnet(d, "decoder.mid_block.resnets.0", ch[-1], ch[-1], 0)
    a = "decoder.mid_block.attentions.0"
    _norm(d, a + ".group_norm", ch[-1])
    for n in ("to_q", "to_k", "to_v", "to_out.0"):
        _lin(d, f"{a}.{n}", ch[-1], ch[-1])
    _resnet(d, "decoder.mid_block.resnets.1", ch[-1], ch[-1], 0)
    rch = ch[::-1]
    cout = rch[0]
    for i in range(len(ch)):
        cin, cout = cout, rch[i]
        for j in range(cfg["layers_per_block"] + 1):
            _resnet(d, f"decoder.up_blocks.{i}.resnets.{j}", cin if j == 0 else cout, cout, 0)
        if i < len(ch) - 1:
            _conv(d, f"decoder.up_blocks.{i}.upsamplers.0.conv", cout, cout, 3)
    _norm(d, "decoder.conv_norm_out", ch[0])
    _conv(d, "decoder.conv_out", ch[0], cfg["out_channels"], 3)
    return d


def clip_text_schema(cfg: dict) -> OrderedDict:
    d = OrderedDict()
    c = cfg["hidden_size"]
    d["text_model.embeddings.token_embedding.weight"] = ("embed", (cfg["vocab_size"], c))
    d["text_model.embeddings.position_embedding.weight"] = ("embed", (cfg["max_position"], c))
    for l in range(cfg["num_layers"]):
        p = f"text_model.encoder.layers.{l}"
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            _lin(d, f"{p}.self_attn.{n}", c, c)
        _norm(d, p + ".layer_norm1", c)
        _lin(d, p + ".mlp.fc1", c, cfg["intermediate_size"])
        _lin(d, p + ".mlp.fc2", cfg["intermediate_size"], c)
        _norm(d, p + ".layer_norm2", c)
    _norm(d, "text_model.final_layer_norm", c)
    if cfg.get("projection_dim"):
        d["text_projection.weight"] = ("linear", (cfg["projection_dim"], c))
    return d


def num_params(schema: OrderedDict) -> int:
    return sum(math.prod(v[1]) for v in schema.values())


def random_state_dict(schema: OrderedDict, seed: int = 0, round_dtype=torch.float16, device="cpu",
                      store_dtype=torch.float32) -> OrderedDict:
    """torch default-init equivalents (kaiming-uniform(a=sqrt 5) weights, U(+-1/sqrt(fan_in)) biases, N(0,1)*0.02
    embeddings, norm affine 1/0).  Values are rounded to `round_dtype`-representable numbers (fp16 = the library's
    default 16-bit type; None keeps full fp32) so the fp32 CPU oracle and the 16-bit GPU path hold identical weights.
    `device="cuda"` draws directly on the GPU (benchmarks: 860 M parameters in a second instead of a minute) and
    `store_dtype=torch.float16` halves the staging memory; the CPU stream (device="cpu") is the one parity tests use."""
    g = torch.Generator(device=device).manual_seed(seed)
    sd = OrderedDict()
    for name, spec in schema.items():
        kind, shape = spec[0], spec[1]
        if kind in ("linear", "conv"):
            fan_in = math.prod(shape[1:])
            bound = 1.0 / math.sqrt(fan_in)
            t = (torch.rand(shape, generator=g, device=device) * 2 - 1) * bound
        elif kind == "bias":
            bound = 1.0 / math.sqrt(spec[2])
            t = (torch.rand(shape, generator=g, device=device) * 2 - 1) * bound
        elif kind == "embed":
            t = torch.randn(shape, generator=g, device=device) * 0.02
        elif kind == "ones":
            t = torch.ones(shape, device=device)
        else:
            t = torch.zeros(shape, device=device)
        if round_dtype is not None:
            t = t.to(round_dtype)
        sd[name] = t.to(store_dtype)
    return sd
