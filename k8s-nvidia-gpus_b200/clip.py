"""CLIP text encoder as a plan of b200sd kernels.

Replaces `text_encoder(input_ids)[0]` inside StableDiffusionPipeline.encode_prompt (reference configmap.yaml:105;
SURVEY.md §8(a) a7, A.3): token + position embedding, pre-LN layers with causal self-attention (flash kernel,
`causal=1`) and quick_gelu / gelu MLP (fused into the fc1 GEMM epilogue), final LayerNorm.

SDXL (StableDiffusionXLPipeline.encode_prompt): both encoders hand over `hidden_states[-2]` (the input of
the last layer, no final LayerNorm); the second one (CLIPTextModelWithProjection) additionally returns
`text_embeds` = text_projection(final_layer_norm(last hidden state)[EOS token]).
"""
from __future__ import annotations

import torch

from . import lib as L
from . import ops
from .engine import Act, Builder, DeviceWeights


class CLIPTextEngine:
    def __init__(self, cfg: dict, state_dict: dict, device="cuda"):
        self.cfg = cfg
        self.device = torch.device(device)
        self.W = DeviceWeights(state_dict, self.device)

    def build_plan(self, nseq: int, ids: torch.Tensor, eos_rows: torch.Tensor | None = None):
        """ids: int32 [nseq * 77] on device (contents may change between runs).  Returns (plan, out [nseq, 77, C],
        pooled) — out is last_hidden_state, or hidden_states[-2] when cfg["output"] == "penultimate" (SDXL); pooled is
        None unless cfg has "projection_dim": then [nseq, projection_dim] text_embeds, which needs `eos_rows`
        (int32 [nseq] on device: flat row index n * 77 + position of sequence n's EOS token)."""
        cfg, W = self.cfg, self.W
        c, heads, seq = cfg["hidden_size"], cfg["num_heads"], cfg["max_position"]
        d = c // heads
        act = {"quick_gelu": L.ACT_QUICK_GELU, "gelu": L.ACT_GELU}.get(cfg["act"])
        if act is None:
            raise ValueError("CLIP MLP activation %r (quick_gelu and gelu are implemented)" % cfg["act"])
        penult = cfg.get("output") == "penultimate"
        proj = cfg.get("projection_dim")
        if proj and eos_rows is None:
            raise ValueError("a text encoder with a projection needs eos_rows")
        bld = Builder(self.device)
        x = Act(bld.alloc((nseq, 1, seq, c)), nseq, 1, seq, c)
        bld.add(ops.clip_embed(ids, W.act16("text_model.embeddings.token_embedding.weight"),
                               W.act16("text_model.embeddings.position_embedding.weight"), x.t, rows=nseq * seq,
                               seq=seq, c=c, vocab=cfg["vocab_size"]))
        nl = cfg["num_layers"] - (1 if penult and not proj else 0)  # the pooled output needs every layer
        out = None
        for l in range(nl):
            keep_x = penult and l == cfg["num_layers"] - 1
            if keep_x:
                out = x.t.view(nseq, seq, c)  # hidden_states[-2]: what the last layer reads; never recycled below
            p = f"text_model.encoder.layers.{l}"
            ln = bld.layernorm(W, p + ".layer_norm1", x, cfg["eps"])
            names = [p + ".self_attn.q_proj", p + ".self_attn.k_proj", p + ".self_attn.v_proj"]
            qkv = bld.linear_w(ln, W.cat_linear(names), W.cat_bias(names), 3 * c)
            bld.free(ln)
            q2 = qkv.t.view(-1, 3 * c)
            a = bld.attention(q2, q2[:, c:], q2[:, 2 * c:], batch=nseq, heads=heads, nq=seq, nk=seq, d=d, ldq=3 * c,
                              ldk=3 * c, ldv=3 * c, causal=True, like=x)
            bld.free(qkv)
            x2 = bld.linear(W, p + ".self_attn.out_proj", a, c, residual=x)
            bld.free(a)
            if not keep_x:
                bld.free(x)
            ln = bld.layernorm(W, p + ".layer_norm2", x2, cfg["eps"])
            hmid = bld.linear(W, p + ".mlp.fc1", ln, cfg["intermediate_size"], act=act)
            bld.free(ln)
            x = bld.linear(W, p + ".mlp.fc2", hmid, c, residual=x2)
            bld.free(hmid, x2)
        fin_w, fin_b = W.vec("text_model.final_layer_norm.weight"), W.vec("text_model.final_layer_norm.bias")
        pooled = None
        if not penult:
            out = bld.persistent((nseq, seq, c), zero=False)
            bld.add(ops.layernorm(x.t, out, fin_w, fin_b, rows=nseq * seq, c=c, eps=cfg["eps"]))
        elif not proj:
            out = x.t.view(nseq, seq, c)  # L - 1 layers were run: x is hidden_states[-2]
        if proj:
            # pooled = final LN of the EOS row only: gather the rows (embedding-gather kernel with a zero position
            # table), normalise them, project (no bias)
            eos = Act(bld.alloc((nseq, 1, 1, c)), nseq, 1, 1, c)
            zero_pos = bld.persistent((1, c))
            bld.add(ops.clip_embed(eos_rows, x.t.view(nseq * seq, c), zero_pos, eos.t, rows=nseq, seq=1, c=c,
                                   vocab=nseq * seq))
            ln = Act(bld.alloc((nseq, 1, 1, c)), nseq, 1, 1, c)
            bld.add(ops.layernorm(eos.t, ln.t, fin_w, fin_b, rows=nseq, c=c, eps=cfg["eps"]))
            pooled = bld.persistent((nseq, proj), zero=False)
            bld.add(ops.gemm(ln.t, W.cat_linear(["text_projection"]), pooled, nimg=1, h=1, w_=nseq, c0=c, n=proj,
                             ksize=1, weights_static=True))
        return bld.finish(), out, pooled
