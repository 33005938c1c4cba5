"""Oracle CLIP text encoder (fp32) — restates transformers==4.44.2 `models/clip/modeling_clip.py`
(CLIPTextModel: CLIPTextEmbeddings, CLIPEncoderLayer pre-LN with causal mask, final_layer_norm) as pinned at
reference configmap.yaml:12 and used by `pipe(...)` (configmap.yaml:105) to encode the prompt and the empty negative
prompt.  Pinned against the installed transformers.CLIPTextModel in tests (tests/golden/clip_tiny.pt).
TEST INFRASTRUCTURE — see oracle/__init__.py."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def quick_gelu(x):
    return x * torch.sigmoid(1.702 * x)


class _Embeddings(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.token_embedding = nn.Embedding(cfg["vocab_size"], cfg["hidden_size"])
        self.position_embedding = nn.Embedding(cfg["max_position"], cfg["hidden_size"])

    def forward(self, ids):
        pos = torch.arange(ids.shape[1])[None]
        return self.token_embedding(ids) + self.position_embedding(pos)


class _SelfAttn(nn.Module):
    def __init__(self, c, heads):
        super().__init__()
        self.heads = heads
        self.q_proj, self.k_proj = nn.Linear(c, c), nn.Linear(c, c)
        self.v_proj, self.out_proj = nn.Linear(c, c), nn.Linear(c, c)

    def forward(self, x):
        b, n, c = x.shape
        h, d = self.heads, c // self.heads
        q = self.q_proj(x).view(b, n, h, d).transpose(1, 2) * d ** -0.5
        k = self.k_proj(x).view(b, n, h, d).transpose(1, 2)
        v = self.v_proj(x).view(b, n, h, d).transpose(1, 2)
        s = q @ k.transpose(-1, -2)
        mask = torch.full((n, n), float("-inf")).triu(1)
        s = torch.softmax(s + mask, dim=-1)
        return self.out_proj((s @ v).transpose(1, 2).reshape(b, n, c))


class _MLP(nn.Module):
    def __init__(self, c, inter, act):
        super().__init__()
        self.fc1, self.fc2 = nn.Linear(c, inter), nn.Linear(inter, c)
        self.act = act

    def forward(self, x):
        x = self.fc1(x)
        x = quick_gelu(x) if self.act == "quick_gelu" else F.gelu(x)
        return self.fc2(x)


class _Layer(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        c = cfg["hidden_size"]
        self.layer_norm1 = nn.LayerNorm(c, eps=cfg["eps"])
        self.self_attn = _SelfAttn(c, cfg["num_heads"])
        self.layer_norm2 = nn.LayerNorm(c, eps=cfg["eps"])
        self.mlp = _MLP(c, cfg["intermediate_size"], cfg["act"])

    def forward(self, x):
        x = x + self.self_attn(self.layer_norm1(x))
        return x + self.mlp(self.layer_norm2(x))


class _Encoder(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.layers = nn.ModuleList([_Layer(cfg) for _ in range(cfg["num_layers"])])


class _TextTransformer(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.embeddings = _Embeddings(cfg)
        self.encoder = _Encoder(cfg)
        self.final_layer_norm = nn.LayerNorm(cfg["hidden_size"], eps=cfg["eps"])


class CLIPTextModel(nn.Module):
    """CLIPTextModel, or CLIPTextModelWithProjection when cfg has "projection_dim" (SDXL's second encoder: state-dict
    key `text_projection.weight`, no bias)."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.text_model = _TextTransformer(cfg)
        if cfg.get("projection_dim"):
            self.text_projection = nn.Linear(cfg["hidden_size"], cfg["projection_dim"], bias=False)

    def encode_xl(self, ids, eos_token_id):
        """StableDiffusionXLPipeline.encode_prompt's use of one encoder: (hidden_states[-2], text_embeds or None).
        text_embeds = text_projection(pooled), pooled = final_layer_norm(last hidden state) at the first EOS token
        (transformers 4.44.2 modeling_clip.py CLIPTextTransformer.forward, eos_token_id != 2 branch; with CLIP's
        vocabulary the legacy argmax branch picks the same position because EOS is the largest id)."""
        tm = self.text_model
        x = tm.embeddings(ids)
        hidden = [x]
        for layer in tm.encoder.layers:
            x = layer(x)
            hidden.append(x)
        embeds = None
        if self.cfg.get("projection_dim"):
            last = tm.final_layer_norm(x)
            pos = (ids == eos_token_id).int().argmax(dim=-1)
            embeds = self.text_projection(last[torch.arange(ids.shape[0]), pos])
        return hidden[-2], embeds

    def forward(self, ids, output="last_hidden_state"):
        """ids int64 [B,77] -> [B,77,C].  "last_hidden_state" (SD1.5: post final LN) or "penultimate" (SDXL:
        hidden_states[-2], no final LN)."""
        tm = self.text_model
        x = tm.embeddings(ids)
        hidden = [x]
        for layer in tm.encoder.layers:
            x = layer(x)
            hidden.append(x)
        if output == "penultimate":
            return hidden[-2]
        return tm.final_layer_norm(x)
