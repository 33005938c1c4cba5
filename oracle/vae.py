"""Oracle AutoencoderKL.decode (fp32, NCHW) — restates diffusers==0.30.2 `models/autoencoders/autoencoder_kl.py`,
`vae.py` (Decoder), `unet_2d_blocks.py` (UNetMidBlock2D / UpDecoderBlock2D) as pinned at reference
configmap.yaml:13; reached through `pipe(...)` at configmap.yaml:105 (`enable_vae_slicing()` at :43 only batches the
same math one image at a time).  TEST INFRASTRUCTURE — see oracle/__init__.py."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet import ResnetBlock2D, Upsample2D


class VaeAttention(nn.Module):
    """Single-head spatial self-attention of the VAE mid block (Attention with group_norm, residual_connection)."""

    def __init__(self, c, groups):
        super().__init__()
        self.group_norm = nn.GroupNorm(groups, c, eps=1e-6)
        self.to_q = nn.Linear(c, c)
        self.to_k = nn.Linear(c, c)
        self.to_v = nn.Linear(c, c)
        self.to_out = nn.ModuleList([nn.Linear(c, c), nn.Identity()])

    def forward(self, x):
        b, c, h, w = x.shape
        r = x
        t = self.group_norm(x).view(b, c, h * w).transpose(1, 2)
        q, k, v = self.to_q(t), self.to_k(t), self.to_v(t)
        s = torch.softmax(q @ k.transpose(-1, -2) * c ** -0.5, dim=-1)
        o = self.to_out[0](s @ v)
        return o.transpose(1, 2).reshape(b, c, h, w) + r


class VaeMidBlock(nn.Module):
    def __init__(self, c, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(c, c, None, groups, 1e-6) for _ in range(2)])
        self.attentions = nn.ModuleList([VaeAttention(c, groups)])

    def forward(self, x):
        return self.resnets[1](self.attentions[0](self.resnets[0](x)))


class UpDecoderBlock(nn.Module):
    def __init__(self, cin, cout, layers, groups, up):
        super().__init__()
        self.resnets = nn.ModuleList(
            [ResnetBlock2D(cin if i == 0 else cout, cout, None, groups, 1e-6) for i in range(layers)])
        self.has_up = up
        if up:
            self.upsamplers = nn.ModuleList([Upsample2D(cout)])

    def forward(self, x):
        for r in self.resnets:
            x = r(x)
        return self.upsamplers[0](x) if self.has_up else x


class Decoder(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        ch = cfg["block_out_channels"]
        g = cfg["norm_num_groups"]
        self.conv_in = nn.Conv2d(cfg["latent_channels"], ch[-1], 3, padding=1)
        self.mid_block = VaeMidBlock(ch[-1], g)
        rch = ch[::-1]
        ups, cout = [], rch[0]
        for i in range(len(ch)):
            cin, cout = cout, rch[i]
            ups.append(UpDecoderBlock(cin, cout, cfg["layers_per_block"] + 1, g, up=i < len(ch) - 1))
        self.up_blocks = nn.ModuleList(ups)
        self.conv_norm_out = nn.GroupNorm(g, ch[0], eps=1e-6)
        self.conv_out = nn.Conv2d(ch[0], cfg["out_channels"], 3, padding=1)

    def forward(self, z):
        x = self.mid_block(self.conv_in(z))
        for b in self.up_blocks:
            x = b(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))


class AutoencoderKLDecoder(nn.Module):
    """Decoder half of AutoencoderKL (text-to-image never runs the encoder)."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.post_quant_conv = nn.Conv2d(cfg["latent_channels"], cfg["latent_channels"], 1)
        self.decoder = Decoder(cfg)

    @property
    def scale_factor(self):
        return 2 ** (len(self.cfg["block_out_channels"]) - 1)

    def decode(self, z):
        return self.decoder(self.post_quant_conv(z))
