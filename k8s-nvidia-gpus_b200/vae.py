"""AutoencoderKL.decode + image post-processing as a plan of b200sd kernels.

Replaces `vae.decode(latents / scaling_factor)` and `image_processor.postprocess` (clamp(x/2+.5) -> NHWC -> *255 ->
round -> uint8) at the end of `pipe(...)` (reference configmap.yaml:105-112).
The reference's enable_vae_slicing() (configmap.yaml:43) decodes one image at a time to fit a 6 GB card; with 180 GB
of HBM the decoder runs batched.

The single-head mid-block attention has head dim 512 (> the flash kernel's 160 limit and > one SM's TMEM/smem budget
for a fused tile), so it runs unfused on the same GEMM kernel: S = QK^T (alpha = 1/sqrt(C)) -> row softmax ->
O = P V with V^T produced directly by a GEMM with swapped operand roles (V^T = W_v X^T); V's bias is added after the
product (rows of P sum to 1).  The fp32 logits are produced in blocks of query rows (at most ATTN_LOGIT_BYTES at a
time, reused for every block and image), so no [hw, hw] matrix is ever materialised: at 1024x1024 (16 384 tokens) the
full matrix would be 1 GiB per image.

The 16-bit type of the decoder is a per-engine choice (`variant`): "bf16" binds the plan to the bfloat16 image of the
library - the SDXL VAE needs fp32's exponent range (upstream `force_upcast`s it to fp32), while the UNet stays fp16.
"""
from __future__ import annotations

import os

import torch

from . import lib as L
from . import ops
from .engine import Act, Builder, DeviceWeights


ATTN_LOGIT_BYTES = 64 << 20


class VAEDecoderEngine:
    def __init__(self, cfg: dict, state_dict: dict, device="cuda", variant: str | None = None):
        self.cfg = cfg
        self.device = torch.device(device)
        self.W = DeviceWeights(state_dict, self.device)
        self.variant = variant or cfg.get("dtype", "fp16")

    @property
    def scale_factor(self):
        return 2 ** (len(self.cfg["block_out_channels"]) - 1)

    def _mid_attention(self, bld: Builder, p: str, x: Act, groups: int):
        W = self.W
        c, hw, n = x.c, x.hw, x.nimg
        gn = bld.groupnorm(W, p + ".group_norm", x, groups, 1e-6, silu=False)
        assert c % 64 == 0 and hw % 64 == 0, "unfused VAE attention needs C and HW multiples of 64"
        q = bld.linear(W, p + ".to_q", gn, c)
        k = bld.linear(W, p + ".to_k", gn, c)   # dense [hw, c] rows: consumed as the B ("weight") operand of QK^T
        # V^T[c_out, token] = W_v[c_out, :] . X[token, :]  — weights play the A role, activations the B role
        wv = W.act16(p + ".to_v.weight")
        vt = bld.alloc((n, c, hw))
        rb = max(128, min(hw, (ATTN_LOGIT_BYTES // (4 * hw)) // 128 * 128))  # query rows per logit block
        s = bld.alloc((rb, hw), torch.float32)
        pm = bld.alloc((rb, hw))
        o = bld.alloc((n, x.h, x.w, c))
        q2, k2, gn2, o2 = q.t.view(n, hw, c), k.t.view(n, hw, c), gn.t.view(n, hw, c), o.view(n, hw, c)
        for i in range(n):
            bld.add(ops.gemm(wv, gn2[i], vt[i], nimg=1, h=1, w_=c, c0=c, n=hw, ksize=1))
            for r0 in range(0, hw, rb):
                r = min(rb, hw - r0)
                bld.add(ops.gemm(q2[i, r0:r0 + r], k2[i], s, nimg=1, h=1, w_=r, c0=c, n=hw, ksize=1, alpha=c ** -0.5))
                bld.add(ops.softmax_rows(s, pm, rows=r, cols=hw))
                bld.add(ops.gemm(pm, vt[i], o2[i, r0:r0 + r], nimg=1, h=1, w_=r, c0=hw, n=c, ksize=1,
                                 bias=W.vec(p + ".to_v.bias")))
        bld.free(gn, q, k, vt, s, pm)
        oa = Act(o, n, x.h, x.w, c)
        out = bld.linear(W, p + ".to_out.0", oa, c, residual=x)
        bld.free(oa)
        return out

    def build_plan(self, batch: int, h: int, w: int, latents: torch.Tensor):
        """latents f32 NHWC [batch, h*w, 4] -> (plan, uint8 images [batch, h*f, w*f, 3])."""
        with L.use(self.variant):  # every call of this plan binds to the engine's image of the library
            return self._build_plan(batch, h, w, latents)

    def _build_plan(self, batch: int, h: int, w: int, latents: torch.Tensor):
        cfg, W = self.cfg, self.W
        ch = cfg["block_out_channels"]
        g = cfg["norm_num_groups"]
        layers = cfg["layers_per_block"] + 1
        lc = cfg["latent_channels"]
        assert lc == 4
        # Statistics ride in the conv epilogues for K >= 1536 only.  Measured: lowering the bar to K >= 1024 (the
        # 128-channel convs at 512x512, N = 128) removes 1.4 ms of statistics passes but costs those epilogue-bound
        # convs 6 ms (decode 21.6 -> 26.3 ms for 8 images).  B200SD_VAE_GN_MIN_K for A/B runs.
        bld = Builder(self.device, fuse_gn_stats=True, gn_fuse_min_k=int(os.environ.get("B200SD_VAE_GN_MIN_K", 24 * 64)))
        z = bld.persistent((batch, h, w, 8))
        bld.add(ops.latent_to_act(latents, z, b=batch, hw=h * w, scale=1.0 / cfg["scaling_factor"]))
        # post_quant_conv 1x1 (4 -> 4) written into a zero-padded 8-channel buffer (TMA wants 16-byte pixels)
        z2 = bld.persistent((batch, h, w, 8))
        bld.conv(W, "post_quant_conv", Act(z, batch, h, w, 8), lc, 1, out=z2, ldc=8, wsplit=(lc,))
        x = bld.conv(W, "decoder.conv_in", Act(z2, batch, h, w, 8), ch[-1], 3, wsplit=(lc,), gn_groups=g)
        y = bld.resnet(W, "decoder.mid_block.resnets.0", x, ch[-1], g, 1e-6)
        bld.free(x)
        y2 = self._mid_attention(bld, "decoder.mid_block.attentions.0", y, g)
        bld.free(y)
        x = bld.resnet(W, "decoder.mid_block.resnets.1", y2, ch[-1], g, 1e-6)
        bld.free(y2)
        rch = ch[::-1]
        for i in range(len(ch)):
            for j in range(layers):
                y = bld.resnet(W, f"decoder.up_blocks.{i}.resnets.{j}", x, rch[i], g, 1e-6)
                bld.free(x)
                x = y
            if i < len(ch) - 1:
                up = bld.upsample_conv(W, f"decoder.up_blocks.{i}.upsamplers.0.conv", x, rch[i], gn_groups=g)
                bld.free(x)
                x = up
        gn = bld.groupnorm(W, "decoder.conv_norm_out", x, g, 1e-6, silu=True)
        bld.free(x)
        H, Wd = gn.h, gn.w
        rgb = bld.alloc((batch, H, Wd, 4), torch.float32)
        bld.conv(W, "decoder.conv_out", gn, cfg["out_channels"], 3, out=rgb, ldc=4)
        bld.free(gn)
        img = bld.persistent((batch, H, Wd, 3), dtype=torch.uint8, zero=False)
        bld.add(ops.denorm_u8(rgb, img, npix=batch * H * Wd, ldx=4))
        self.last_rgb = rgb
        return bld.finish(), img
