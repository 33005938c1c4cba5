"""CLIP text encoder as a plan of b200sd kernels.

Replaces `text_encoder(input_ids)[0]` inside StableDiffusionPipeline.encode_prompt (reference configmap.yaml:105;
SURVEY.md §8(a) a7, A.3): token + position embedding, 12 pre-LN layers with causal self-attention (flash kernel,
`causal=1`) and quick_gelu MLP (fused into the fc1 GEMM epilogue), final LayerNorm.
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

    def build_plan(self, nseq: int, ids: torch.Tensor):
        """ids: int32 [nseq * 77] on device (contents may change between runs).  Returns (plan, out bf16
        [nseq, 77, C]) — last_hidden_state, or hidden_states[-2] when cfg["output"] == "penultimate" (SDXL)."""
        cfg, W = self.cfg, self.W
        c, heads, seq = cfg["hidden_size"], cfg["num_heads"], cfg["max_position"]
        d = c // heads
        act = L.ACT_QUICK_GELU if cfg["act"] == "quick_gelu" else None
        if act is None:
            raise NotImplementedError("CLIP MLP activation %r (only quick_gelu is implemented)" % cfg["act"])
        bld = Builder(self.device)
        x = Act(bld.alloc((nseq, 1, seq, c)), nseq, 1, seq, c)
        bld.add(ops.clip_embed(ids, W.act16("text_model.embeddings.token_embedding.weight"),
                               W.act16("text_model.embeddings.position_embedding.weight"), x.t, rows=nseq * seq,
                               seq=seq, c=c, vocab=cfg["vocab_size"]))
        nl = cfg["num_layers"] - (1 if cfg.get("output") == "penultimate" else 0)
        for l in range(nl):
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
            bld.free(a, x)
            ln = bld.layernorm(W, p + ".layer_norm2", x2, cfg["eps"])
            hmid = bld.linear(W, p + ".mlp.fc1", ln, cfg["intermediate_size"], act=act)
            bld.free(ln)
            x = bld.linear(W, p + ".mlp.fc2", hmid, c, residual=x2)
            bld.free(hmid, x2)
        out = bld.persistent((nseq, seq, c), zero=False)
        if cfg.get("output") == "penultimate":
            raise NotImplementedError("penultimate hidden-state output (SDXL) is not wired yet")
        bld.add(ops.layernorm(x.t, out, W.vec("text_model.final_layer_norm.weight"),
                              W.vec("text_model.final_layer_norm.bias"), rows=nseq * seq, c=c, eps=cfg["eps"]))
        return bld.finish(), out
