"""Oracle UNet2DConditionModel (fp32, NCHW) — restates diffusers==0.30.2
`models/unets/unet_2d_condition.py`, `unet_2d_blocks.py`, `resnet.py`, `attention.py`, `transformer_2d.py`,
`embeddings.py` as pinned by the reference (configmap.yaml:13) and called at configmap.yaml:105.
Module/parameter names follow the diffusers state-dict schema so real checkpoints load unchanged.
TEST INFRASTRUCTURE — see oracle/__init__.py."""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


def timestep_embedding(t: torch.Tensor, dim: int, flip_sin_to_cos: bool = True, shift: float = 0.0) -> torch.Tensor:
    """embeddings.get_timestep_embedding: exp(-ln(1e4) * i / (half - shift)); [sin|cos], flipped to [cos|sin]."""
    half = dim // 2
    exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32) / (half - shift)
    emb = t[:, None].float() * torch.exp(exponent)[None, :]
    emb = torch.cat([torch.sin(emb), torch.cos(emb)], dim=-1)
    if flip_sin_to_cos:
        emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
    return emb


class TimestepEmbedding(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.linear_1 = nn.Linear(cin, cout)
        self.linear_2 = nn.Linear(cout, cout)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin, cout, temb_ch, groups, eps):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_ch, cout) if temb_ch else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.time_emb_proj is not None:
            h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class Attention(nn.Module):
    """attention_processor.Attention with AttnProcessor math (SlicedAttnProcessor, enabled by the reference at
    configmap.yaml:42, computes the same product slice by slice)."""

    def __init__(self, dim, heads, ctx_dim=None, qkv_bias=False):
        super().__init__()
        self.heads = heads
        self.to_q = nn.Linear(dim, dim, bias=qkv_bias)
        self.to_k = nn.Linear(ctx_dim or dim, dim, bias=qkv_bias)
        self.to_v = nn.Linear(ctx_dim or dim, dim, bias=qkv_bias)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim), nn.Identity()])

    def forward(self, x, ctx=None):
        ctx = x if ctx is None else ctx
        b, n, c = x.shape
        h = self.heads
        q = self.to_q(x).view(b, n, h, c // h).transpose(1, 2)
        k = self.to_k(ctx).view(b, ctx.shape[1], h, c // h).transpose(1, 2)
        v = self.to_v(ctx).view(b, ctx.shape[1], h, c // h).transpose(1, 2)
        s = torch.softmax(q @ k.transpose(-1, -2) * (c // h) ** -0.5, dim=-1)
        o = (s @ v).transpose(1, 2).reshape(b, n, c)
        return self.to_out[0](o)


class GEGLU(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.proj = nn.Linear(cin, cout * 2)

    def forward(self, x):
        u, g = self.proj(x).chunk(2, dim=-1)
        return u * F.gelu(g)


class FeedForward(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Identity(), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        return self.net[2](self.net[0](x))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, heads, ctx_dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, ctx_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, x, ctx):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x), ctx)
        return x + self.ff(self.norm3(x))


class Transformer2DModel(nn.Module):
    def __init__(self, dim, heads, ctx_dim, depth, groups, linear_proj):
        super().__init__()
        self.linear_proj = linear_proj
        self.norm = nn.GroupNorm(groups, dim, eps=1e-6)
        self.proj_in = nn.Linear(dim, dim) if linear_proj else nn.Conv2d(dim, dim, 1)
        self.transformer_blocks = nn.ModuleList([BasicTransformerBlock(dim, heads, ctx_dim) for _ in range(depth)])
        self.proj_out = nn.Linear(dim, dim) if linear_proj else nn.Conv2d(dim, dim, 1)

    def forward(self, x, ctx):
        b, c, h, w = x.shape
        r = x
        x = self.norm(x)
        if self.linear_proj:
            x = self.proj_in(x.permute(0, 2, 3, 1).reshape(b, h * w, c))
        else:
            x = self.proj_in(x).permute(0, 2, 3, 1).reshape(b, h * w, c)
        for blk in self.transformer_blocks:
            x = blk(x, ctx)
        if self.linear_proj:
            x = self.proj_out(x).reshape(b, h, w, c).permute(0, 3, 1, 2)
        else:
            x = self.proj_out(x.reshape(b, h, w, c).permute(0, 3, 1, 2))
        return x + r


class Downsample2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class DownBlock(nn.Module):
    def __init__(self, cin, cout, temb, layers, groups, eps, attn, heads, ctx, depth, linproj, down):
        super().__init__()
        self.resnets = nn.ModuleList(
            [ResnetBlock2D(cin if i == 0 else cout, cout, temb, groups, eps) for i in range(layers)])
        if attn:
            self.attentions = nn.ModuleList(
                [Transformer2DModel(cout, heads, ctx, depth, groups, linproj) for _ in range(layers)])
        self.has_attn = attn
        if down:
            self.downsamplers = nn.ModuleList([Downsample2D(cout)])
        self.has_down = down

    def forward(self, x, temb, ctx):
        outs = []
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.has_attn:
                x = self.attentions[i](x, ctx)
            outs.append(x)
        if self.has_down:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class UpBlock(nn.Module):
    def __init__(self, cin, cout, cprev, temb, layers, groups, eps, attn, heads, ctx, depth, linproj, up):
        super().__init__()
        res = []
        for i in range(layers):
            skip = cin if i == layers - 1 else cout
            rin = cprev if i == 0 else cout
            res.append(ResnetBlock2D(rin + skip, cout, temb, groups, eps))
        self.resnets = nn.ModuleList(res)
        if attn:
            self.attentions = nn.ModuleList(
                [Transformer2DModel(cout, heads, ctx, depth, groups, linproj) for _ in range(layers)])
        self.has_attn = attn
        if up:
            self.upsamplers = nn.ModuleList([Upsample2D(cout)])
        self.has_up = up

    def forward(self, x, skips, temb, ctx):
        for i, res in enumerate(self.resnets):
            x = torch.cat([x, skips.pop()], dim=1)  # running tensor first, skip second
            x = res(x, temb)
            if self.has_attn:
                x = self.attentions[i](x, ctx)
        if self.has_up:
            x = self.upsamplers[0](x)
        return x


class MidBlock(nn.Module):
    def __init__(self, c, temb, groups, eps, heads, ctx, depth, linproj):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, temb, groups, eps) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DModel(c, heads, ctx, depth, groups, linproj)])

    def forward(self, x, temb, ctx):
        x = self.resnets[0](x, temb)
        x = self.attentions[0](x, ctx)
        return self.resnets[1](x, temb)


class UNet2DConditionModel(nn.Module):
    def __init__(self, cfg: dict):
        super().__init__()
        self.cfg = cfg
        ch = cfg["block_out_channels"]
        temb = ch[0] * 4
        g, eps = cfg["norm_num_groups"], cfg["norm_eps"]
        ctx, lin = cfg["cross_attention_dim"], cfg["use_linear_projection"]
        heads, depth, layers = cfg["num_attention_heads"], cfg["transformer_layers_per_block"], cfg["layers_per_block"]
        self.conv_in = nn.Conv2d(cfg["in_channels"], ch[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(ch[0], temb)
        if cfg.get("addition_embed_type") == "text_time":
            self.add_embedding = TimestepEmbedding(cfg["projection_class_embeddings_input_dim"], temb)
        downs = []
        cout = ch[0]
        for i, t in enumerate(cfg["down_block_types"]):
            cin, cout = cout, ch[i]
            downs.append(DownBlock(cin, cout, temb, layers, g, eps, t.startswith("CrossAttn"), heads[i], ctx, depth[i],
                                   lin, down=i < len(ch) - 1))
        self.down_blocks = nn.ModuleList(downs)
        self.mid_block = MidBlock(ch[-1], temb, g, eps, heads[-1], ctx, depth[-1], lin)
        ups = []
        rch, rheads, rdepth = ch[::-1], heads[::-1], depth[::-1]
        cout = rch[0]
        for i, t in enumerate(cfg["up_block_types"]):
            cprev, cout = cout, rch[i]
            cin = rch[min(i + 1, len(ch) - 1)]
            ups.append(UpBlock(cin, cout, cprev, temb, layers + 1, g, eps, t.startswith("CrossAttn"), rheads[i], ctx,
                               rdepth[i], lin, up=i < len(ch) - 1))
        self.up_blocks = nn.ModuleList(ups)
        self.conv_norm_out = nn.GroupNorm(g, ch[0], eps=eps)
        self.conv_out = nn.Conv2d(ch[0], cfg["out_channels"], 3, padding=1)

    def forward(self, x, t, ctx, added_cond=None):
        """x [B,4,h,w], t scalar or [B], ctx [B,77,D] -> eps [B,4,h,w]"""
        if not torch.is_tensor(t):
            t = torch.tensor([t], dtype=torch.float32)
        t = t.reshape(-1).expand(x.shape[0])
        ch0 = self.cfg["block_out_channels"][0]
        emb = self.time_embedding(timestep_embedding(t, ch0))
        if self.cfg.get("addition_embed_type") == "text_time":
            tid = timestep_embedding(added_cond["time_ids"].flatten(), self.cfg["addition_time_embed_dim"])
            tid = tid.reshape(x.shape[0], -1)
            emb = emb + self.add_embedding(torch.cat([added_cond["text_embeds"], tid], dim=-1))
        h = self.conv_in(x)
        skips = [h]
        for blk in self.down_blocks:
            h, outs = blk(h, emb, ctx)
            skips += outs
        h = self.mid_block(h, emb, ctx)
        for blk in self.up_blocks:
            h = blk(h, skips, emb, ctx)
        return self.conv_out(F.silu(self.conv_norm_out(h)))
