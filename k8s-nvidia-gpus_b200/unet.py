"""UNet2DConditionModel denoise step as a flat plan of b200sd kernels.

Replaces `unet(latent_model_input, t, encoder_hidden_states)` + CFG combine + `scheduler.step` inside the denoising
loop of `pipe(...)` (reference configmap.yaml:105-112).  One plan = one denoise step
for a fixed (batch, latent size); it reads the step table / step counter on the device, so the captured CUDA graph is
replayed once per step with no host interaction.

Data flow per step (NHWC bf16, CFG batch = [uncond; cond]):
  latents f32 -> prep (scale, bf16, duplicate) -> conv_in -> down blocks -> mid -> up blocks (skip concat is virtual:
  GroupNorm and the convs read two sources) -> GN+SiLU -> conv_out (f32) -> CFG + scheduler update of the latents.
Cross-attention K/V depend only on the prompt: they are projected once per prompt batch by `kv_plan`.
"""
from __future__ import annotations

import os

import torch

from . import lib as L
from . import ops
from .engine import Act, Builder, DeviceWeights


def _resnet_names(cfg):
    """(name, cout) of every ResnetBlock2D in execution order (for the fused time-embedding projection)."""
    ch = cfg["block_out_channels"]
    layers = cfg["layers_per_block"]
    out = []
    for i in range(len(ch)):
        out += [(f"down_blocks.{i}.resnets.{j}", ch[i]) for j in range(layers)]
    out += [("mid_block.resnets.0", ch[-1]), ("mid_block.resnets.1", ch[-1])]
    rch = ch[::-1]
    for i in range(len(ch)):
        out += [(f"up_blocks.{i}.resnets.{j}", rch[i]) for j in range(layers + 1)]
    return out


def _attn_blocks(cfg):
    """(transformer prefix, channels, heads, depth) of every Transformer2DModel in execution order."""
    ch, heads, depth = cfg["block_out_channels"], cfg["num_attention_heads"], cfg["transformer_layers_per_block"]
    layers = cfg["layers_per_block"]
    out = []
    for i, t in enumerate(cfg["down_block_types"]):
        if t.startswith("CrossAttn"):
            out += [(f"down_blocks.{i}.attentions.{j}", ch[i], heads[i], depth[i]) for j in range(layers)]
    out.append(("mid_block.attentions.0", ch[-1], heads[-1], depth[-1]))
    rch, rh, rd = ch[::-1], heads[::-1], depth[::-1]
    for i, t in enumerate(cfg["up_block_types"]):
        if t.startswith("CrossAttn"):
            out += [(f"up_blocks.{i}.attentions.{j}", rch[i], rh[i], rd[i]) for j in range(layers + 1)]
    return out


class UNetEngine:
    def __init__(self, cfg: dict, state_dict: dict, device="cuda"):
        self.cfg = cfg
        self.device = torch.device(device)
        self.W = DeviceWeights(state_dict, self.device)
        self.seq = 77
        # LayerNorm folded into the neighbouring GEMMs (b200sd.h ln_*; no layernorm kernel, no normalised tensor) is
        # built and parity-tested but OFF by default: measured on the SD1.5 batch-8 step it LOSES 0.8 ms (18.71 -> 19.49
        # ms) - the 0.63 ms of layernorm kernels go, but the short-K linears around them are epilogue-bound and pay more
        # for the extra epilogue work (GEGLU 65536x320x2560: 0.67 -> 1.24 ms; out-projections + 30 % for the row sums).
        # B200SD_UNET_FOLD_LN=1 turns it on.
        self.fold_ln = os.environ.get("B200SD_UNET_FOLD_LN", "0") == "1"

    # ------------------------------------------------------------------ prompt-dependent precompute
    def build_kv_plan(self, nctx: int, ctx, text_embeds: torch.Tensor | None = None,
                      time_ids_emb: torch.Tensor | None = None):
        """ctx: 16-bit [nctx, 77, cross_dim], or one such tensor per text encoder (widths adding up to cross_dim) ->
        per attention block 16-bit [nctx, 77, 2C] (K | V), computed once per
        prompt batch.  SDXL (`addition_embed_type == "text_time"`): also aug_emb fp32 [nctx, 4*C0] =
        add_embedding(concat(text_embeds [nctx, P], time_ids_emb [nctx, 6*D])) - the concat is virtual (two A sources
        of one GEMM).  Returns (plan, kv dict, aug_emb or None)."""
        bld = Builder(self.device)
        W = self.W
        ctxs = list(ctx) if isinstance(ctx, (list, tuple)) else [ctx]
        if len(ctxs) > 2:
            raise ValueError("at most two text encoders")
        seq = ctxs[0].shape[1]
        dims = tuple(int(t.shape[2]) for t in ctxs)
        if sum(dims) != self.cfg["cross_attention_dim"]:
            raise ValueError(f"text encoder widths {dims} do not add up to cross_attention_dim")
        self.seq = seq
        kv = {}
        for p, c, heads, depth in _attn_blocks(self.cfg):
            for l in range(depth):
                b = f"{p}.transformer_blocks.{l}.attn2"
                # the channel concat of the encoders' outputs is virtual: two A sources walked along K
                w = W.cat_conv([b + ".to_k", b + ".to_v"], dims)
                out = bld.persistent((nctx, 1, seq, 2 * c), zero=False)
                bld.add(ops.gemm(ctxs[0], w, out, nimg=1, h=1, w_=nctx * seq, c0=dims[0],
                                 a1=ctxs[1] if len(ctxs) > 1 else None, c1=dims[1] if len(ctxs) > 1 else 0, n=2 * c,
                                 ksize=1, weights_static=True))
                kv[b] = out
        aug = None
        if self.cfg.get("addition_embed_type") == "text_time":
            if text_embeds is None or time_ids_emb is None:
                raise ValueError("text_time conditioning needs text_embeds and the time-id embedding")
            temb_dim = self.cfg["block_out_channels"][0] * 4
            c0, c1 = text_embeds.shape[1], time_ids_emb.shape[1]
            if c0 + c1 != self.cfg["projection_class_embeddings_input_dim"]:
                raise ValueError(f"text_embeds ({c0}) + time ids ({c1}) != projection_class_embeddings_input_dim")
            h1 = bld.alloc((nctx, 1, 1, temb_dim))
            bld.add(ops.gemm(text_embeds, W.conv("add_embedding.linear_1", (c0, c1)), h1, nimg=1, h=1, w_=nctx, c0=c0,
                             a1=time_ids_emb, c1=c1, n=temb_dim, ksize=1, bias=W.vec("add_embedding.linear_1.bias"),
                             act=L.ACT_SILU, weights_static=True))
            aug = bld.persistent((nctx, temb_dim), dtype=torch.float32, zero=False)
            bld.add(ops.gemm(h1, W.conv("add_embedding.linear_2"), aug, nimg=1, h=1, w_=nctx, c0=temb_dim, n=temb_dim,
                             ksize=1, bias=W.vec("add_embedding.linear_2.bias"), weights_static=True))
        return bld.finish(), kv, aug

    # ------------------------------------------------------------------ one denoise step
    def build_step_plan(self, batch: int, h: int, w: int, latents, step_table, step_counter, kv: dict, *,
                        guidance_scale: float, cfg_on: bool = True, history=None, cur_sample=None, aug_emb=None):
        """latents f32 [batch, h*w, 4] (updated in place).  Returns (plan, debug dict)."""
        cfg = self.cfg
        W = self.W
        # GroupNorm statistics in the conv epilogues (engine.Builder): worth ~1 % of the timed step here (A/B in one
        # process pair on the same box: 19.38 vs 19.59 ms); B200SD_UNET_FUSE_GN=0 turns it off for such comparisons
        bld = Builder(self.device, fuse_gn_stats=os.environ.get("B200SD_UNET_FUSE_GN", "1") == "1")
        n = batch * (2 if cfg_on else 1)
        ch = cfg["block_out_channels"]
        g, eps = cfg["norm_num_groups"], cfg["norm_eps"]
        temb_dim = ch[0] * 4
        layers = cfg["layers_per_block"]
        heads, depth, lin = cfg["num_attention_heads"], cfg["transformer_layers_per_block"], cfg["use_linear_projection"]
        dbg = {}

        # --- input + time embedding
        xin = bld.persistent((n, h, w, 8))
        bld.add(ops.prep_unet_input(latents, xin, step_table, step_counter, b=batch, hw=h * w, cfg=cfg_on))
        tsin = bld.alloc((n, 1, 1, ch[0]))
        bld.add(ops.timestep_embedding(tsin, step_table, step_counter, rows=n, dim=ch[0]))
        t1 = bld.linear(W, "time_embedding.linear_1", Act(tsin, n, 1, 1, ch[0]), temb_dim, act=L.ACT_SILU)
        # silu(emb) is all the resnets need (emb itself is never used elsewhere); SDXL adds aug_emb before the SiLU
        emb = bld.linear(W, "time_embedding.linear_2", t1, temb_dim, act=L.ACT_SILU, rowbias=aug_emb,
                         ld_rowbias=temb_dim if aug_emb is not None else 0, per_image_rows=aug_emb is not None)
        res = _resnet_names(cfg)
        names = [r[0] + ".time_emb_proj" for r in res]
        offs, o = {}, 0
        for nme, c in res:
            offs[nme] = o
            o += c
        ntot = o
        tproj = bld.linear_w(emb, W.cat_linear(names, "temb"), W.cat_bias(names, "temb"), ntot, out_dtype=torch.float32)
        tflat = tproj.t.view(n, ntot)

        def temb_for(name):
            return tflat[:, offs[name]:]

        def resnet(name, x, cout):
            return bld.resnet(W, name, x, cout, g, eps, temb=temb_for(name), ld_temb=ntot)

        def transformer(p, x: Act, nheads, ndepth):
            c = x.c
            d = c // nheads
            gn = bld.groupnorm(W, p + ".norm", x, g, 1e-6, silu=False)
            # Every LayerNorm of the block is folded into the layers around it (engine.Builder.ln_linear): its producer
            # emits per-row statistics, its consumer runs on the raw residual stream with gamma-scaled weights.
            fold = self.fold_ln
            t = bld.conv(W, p + ".proj_in", gn, c, 1, rowstats=fold)
            bld.free(gn)
            for l in range(ndepth):
                b = f"{p}.transformer_blocks.{l}"
                qkv_names = [b + ".attn1.to_q", b + ".attn1.to_k", b + ".attn1.to_v"]
                if fold:
                    qkv = bld.ln_linear(W, b + ".norm1", qkv_names, t, 3 * c, bias=False)
                else:
                    ln = bld.layernorm(W, b + ".norm1", t)
                    qkv = bld.linear_w(ln, W.cat_linear(qkv_names), None, 3 * c)
                    bld.free(ln)
                q2 = qkv.t.view(-1, 3 * c)
                a = bld.attention(q2, q2[:, c:], q2[:, 2 * c:], batch=n, heads=nheads, nq=x.hw, nk=x.hw, d=d, ldq=3 * c,
                                  ldk=3 * c, ldv=3 * c, like=t)
                bld.free(qkv)
                t2 = bld.linear(W, b + ".attn1.to_out.0", a, c, residual=t, rowstats=fold)
                bld.free(a, t)
                if fold:
                    q = bld.ln_linear(W, b + ".norm2", [b + ".attn2.to_q"], t2, c, bias=False)
                else:
                    ln = bld.layernorm(W, b + ".norm2", t2)
                    q = bld.linear(W, b + ".attn2.to_q", ln, c, bias=False)
                    bld.free(ln)
                kvt = kv[b + ".attn2"].view(-1, 2 * c)
                a = bld.attention(q.t, kvt, kvt[:, c:], batch=n, heads=nheads, nq=x.hw, nk=self.seq, d=d, ldq=c,
                                  ldk=2 * c, ldv=2 * c, like=t2)
                bld.free(q)
                t3 = bld.linear(W, b + ".attn2.to_out.0", a, c, residual=t2, rowstats=fold)
                bld.free(a, t2)
                perm = ops.geglu_perm(8 * c)
                if fold:
                    hg = bld.ln_linear(W, b + ".norm3", [b + ".ff.net.0.proj"], t3, 8 * c, key="geglu", perm=perm,
                                       act=L.ACT_GEGLU)
                else:
                    ln = bld.layernorm(W, b + ".norm3", t3)
                    hg = bld.linear_w(ln, W.cat_linear([b + ".ff.net.0.proj"], "geglu", perm),
                                      W.cat_bias([b + ".ff.net.0.proj"], "geglu", perm), 8 * c, act=L.ACT_GEGLU)
                    bld.free(ln)
                t = bld.linear(W, b + ".ff.net.2", hg, c, residual=t3, rowstats=fold and l + 1 < ndepth)
                bld.free(hg, t3)
            out = bld.conv(W, p + ".proj_out", t, c, 1, residual=x)
            bld.free(t)
            return out

        # --- down path
        x = bld.conv(W, "conv_in", Act(xin, n, h, w, 8), ch[0], 3, wsplit=(cfg["in_channels"],), gn_groups=g)
        skips = [x]
        for i, btype in enumerate(cfg["down_block_types"]):
            for j in range(layers):
                y = resnet(f"down_blocks.{i}.resnets.{j}", x, ch[i])
                if btype.startswith("CrossAttn"):
                    y2 = transformer(f"down_blocks.{i}.attentions.{j}", y, heads[i], depth[i])
                    bld.free(y)
                    y = y2
                x = y
                skips.append(x)
            if i < len(ch) - 1:
                x = bld.conv_s2(W, f"down_blocks.{i}.downsamplers.0.conv", x, ch[i], gn_groups=g)
                skips.append(x)
        # --- mid
        y = resnet("mid_block.resnets.0", x, ch[-1])
        y2 = transformer("mid_block.attentions.0", y, heads[-1], depth[-1])
        bld.free(y)
        x = resnet("mid_block.resnets.1", y2, ch[-1])
        bld.free(y2)
        # --- up path (skip concat never materialised)
        rch, rh, rd = ch[::-1], heads[::-1], depth[::-1]
        for i, btype in enumerate(cfg["up_block_types"]):
            for j in range(layers + 1):
                skip = skips.pop()
                y = resnet(f"up_blocks.{i}.resnets.{j}", (x, skip), rch[i])
                bld.free(x, skip)
                if btype.startswith("CrossAttn"):
                    y2 = transformer(f"up_blocks.{i}.attentions.{j}", y, rh[i], rd[i])
                    bld.free(y)
                    y = y2
                x = y
            if i < len(ch) - 1:
                up = bld.upsample_conv(W, f"up_blocks.{i}.upsamplers.0.conv", x, rch[i])
                bld.free(x)
                x = up
        # --- head
        gn = bld.groupnorm(W, "conv_norm_out", x, g, eps, silu=True)
        bld.free(x)
        eps_out = bld.persistent((n, h, w, cfg["out_channels"]), dtype=torch.float32, zero=False)
        bld.conv(W, "conv_out", gn, cfg["out_channels"], 3, out=eps_out)
        bld.free(gn)
        dbg["eps"] = eps_out
        dbg["unet_calls"] = len(bld.plan.calls)
        bld.add(ops.cfg_scheduler_step(eps_out, latents, history, cur_sample, step_table, step_counter, b=batch,
                                       hw=h * w, ld_eps=cfg["out_channels"], cfg=cfg_on,
                                       guidance_scale=guidance_scale))
        return bld.finish(), dbg
