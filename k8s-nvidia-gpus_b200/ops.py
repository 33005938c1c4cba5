"""Tensor-level wrappers over the C ABI: weight packing + `Call` builders.

Every function here returns a prepared `lib.Call` (or runs it immediately when `stream` is given); device memory is
owned by torch tensors passed in by the caller — the library never allocates.
"""
from __future__ import annotations

import ctypes as C

import os

import torch

from . import lib as L


def act_dtype():
    return L.act_dtype()


def _ptr(t):
    return None if t is None else t.data_ptr()


def ceil_to(v: int, m: int) -> int:
    return (v + m - 1) // m * m


# ------------------------------------------------------------------------------------------------ weight packing
def pack_conv_weight(w: torch.Tensor, split=None) -> torch.Tensor:
    """[Cout, Cin, kh, kw] (upstream layout) -> act16 [Cout, kh*kw, c0p + c1p], channels zero-padded to
    64 per source so each 64-wide K block of the implicit GEMM belongs to exactly one source / filter tap.  Runs on
    whatever device `w` lives on."""
    cout, cin, kh, kw = w.shape
    split = split or (cin,)
    assert sum(split) == cin
    dt = act_dtype()
    wt = w.permute(0, 2, 3, 1).reshape(cout, kh * kw, cin)
    if len(split) == 1 and cin % 64 == 0:
        return wt.to(dt).contiguous()
    parts, off = [], 0
    for c in split:
        cp = ceil_to(c, 64)
        blk = torch.zeros(cout, kh * kw, cp, dtype=dt, device=w.device)
        blk[:, :, :c] = wt[:, :, off:off + c].to(dt)
        parts.append(blk)
        off += c
    return torch.cat(parts, dim=2).contiguous()


def pack_linear_weight(w: torch.Tensor, split=None) -> torch.Tensor:
    """[N, K] -> act16 [N, 1, Kp]."""
    return pack_conv_weight(w[:, :, None, None], split)


def pack_upconv_weights(w: torch.Tensor) -> list:
    """Upsample2D = nearest 2x + conv 3x3 (pad 1) folded into four 2x2 convolutions on the low-resolution input, one
    per output phase (py, px) = (row parity, column parity): output (2y+py, 2x+px) reads input rows y-1+py .. y+py and
    columns x-1+px .. x+px.  The 3x3 taps that land on the same input pixel are pre-summed (in fp32, then rounded once):
        py = 0: rows {y-1: k0,      y: k1 + k2}        py = 1: rows {y: k0 + k1,      y+1: k2}      (same along x)
    Exact in real arithmetic; 16 instead of 36 MACs per input pixel.  Returns [phase py*2+px] packed [Cout, 4, Cp]."""
    w = w.float()
    rows = {0: [w[:, :, 0:1], w[:, :, 1:2] + w[:, :, 2:3]], 1: [w[:, :, 0:1] + w[:, :, 1:2], w[:, :, 2:3]]}
    out = []
    for py in (0, 1):
        wy = torch.cat(rows[py], dim=2)  # [Cout, Cin, 2, 3]
        cols = {0: [wy[..., 0:1], wy[..., 1:2] + wy[..., 2:3]], 1: [wy[..., 0:1] + wy[..., 1:2], wy[..., 2:3]]}
        for px in (0, 1):
            out.append(pack_conv_weight(torch.cat(cols[px], dim=3)))  # [Cout, Cin, 2, 2] -> [Cout, 4, Cp]
    return out


def geglu_perm(n2: int) -> torch.Tensor:
    """Row permutation for the GEGLU input projection [2F, K]: packed 64-row groups = [32 value rows | 32 gate rows]."""
    f = n2 // 2
    assert f % 32 == 0
    idx = torch.arange(f).view(-1, 32)
    return torch.cat([idx, idx + f], dim=1).reshape(-1)


# ------------------------------------------------------------------------------------------------ call builders
_WEIGHTS_STATIC = os.environ.get("B200SD_WEIGHTS_STATIC", "1") != "0"  # A/B switch: early weight fetch (b200sd.h)
def gemm(a0, w, out, *, nimg, h, w_, c0, n, ksize=1, a1=None, c1=0, ld0=None, ld1=None, bias=None, rowbias=None,
         ld_rowbias=0, residual=None, ldr=0, ldc=None, act=L.ACT_NONE, out_dtype=None, alpha=1.0, bn_hint=0,
         gn_stats=None, gn_groups=0, gn_rows_per_img=0, kh=0, kw=0, off_h=0, off_w=0, stride=0, out_strides=None,
         split_k=0, workspace=None, ln_out_stats=None, ln_in_stats=None, ln_colsum=None, ln_eps=1e-5,
         weights_static=False):
    """weights_static: `w` holds model weights (nothing in flight writes it) - see b200sd.h.
    out_strides: (pixel, row, image) strides of the output in elements when its pixels are interleaved into a larger
    tensor (pass the strided VIEW as `out`: its data_ptr is the first pixel)."""
    p = L.GemmParams()
    p.a0, p.a1, p.w, p.out = _ptr(a0), _ptr(a1), _ptr(w), _ptr(out)
    p.bias, p.rowbias, p.residual = _ptr(bias), _ptr(rowbias), _ptr(residual)
    p.nimg, p.h, p.w_, p.c0, p.c1 = nimg, h, w_, c0, c1
    p.ld0 = ld0 if ld0 is not None else c0
    p.ld1 = ld1 if ld1 is not None else c1
    p.n, p.ksize = n, ksize
    n_out = n // 2 if act == L.ACT_GEGLU else n
    p.ldc = ldc if ldc is not None else n_out
    p.ldr = ldr if ldr else n_out
    p.ld_rowbias = ld_rowbias
    p.act = act
    if out_dtype is None:
        out_dtype = L.F32 if out.dtype == torch.float32 else L.ACT16
    p.out_dtype = out_dtype
    p.bn_hint = bn_hint
    p.alpha = alpha
    p.gn_stats, p.gn_groups, p.gn_rows_per_img = _ptr(gn_stats), gn_groups, gn_rows_per_img
    p.kh, p.kw, p.off_h, p.off_w, p.stride = kh, kw, off_h, off_w, stride
    if out_strides is not None:
        p.out_pix_stride, p.out_row_stride, p.out_img_stride = out_strides
    p.split_k = split_k
    p.weights_static = 1 if weights_static and _WEIGHTS_STATIC else 0
    p.ln_out_stats, p.ln_in_stats, p.ln_colsum, p.ln_eps = _ptr(ln_out_stats), _ptr(ln_in_stats), _ptr(ln_colsum), ln_eps
    if workspace is not None:
        p.workspace, p.workspace_bytes = _ptr(workspace), workspace.numel() * workspace.element_size()
    return L.Call("b200sd_gemm", p, keep=(a0, a1, w, out, bias, rowbias, residual, gn_stats, workspace, ln_out_stats,
                                          ln_in_stats, ln_colsum))


def gemm_workspace_bytes(call) -> int:
    """Split-K workspace bytes a prepared gemm Call needs (0 without split_k)."""
    return int(L.load().b200sd_gemm_workspace_bytes(C.byref(call.params)))


def attention(q, k, v, out, *, batch, heads, nq, nk, d, ldq, ldk, ldv, ldo, scale, causal=False):
    p = L.AttnParams()
    p.q, p.k, p.v, p.out = _ptr(q), _ptr(k), _ptr(v), _ptr(out)
    p.batch, p.heads, p.nq, p.nk, p.d = batch, heads, nq, nk, d
    p.ldq, p.ldk, p.ldv, p.ldo = ldq, ldk, ldv, ldo
    p.causal = int(causal)
    p.scale = scale
    return L.Call("b200sd_attention", p, keep=(q, k, v, out))


def groupnorm(x0, out, stats, gamma, beta, *, nimg, hw, c0, groups, eps, silu, x1=None, c1=0):
    p = L.GroupNormParams()
    p.x0, p.x1, p.stats, p.gamma, p.beta, p.out = _ptr(x0), _ptr(x1), _ptr(stats), _ptr(gamma), _ptr(beta), _ptr(out)
    p.nimg, p.hw, p.c0, p.c1, p.groups, p.silu, p.eps = nimg, hw, c0, c1, groups, int(silu), eps
    keep = (x0, x1, stats, gamma, beta, out)
    return (L.Call("b200sd_groupnorm_stats", p, keep=keep), L.Call("b200sd_groupnorm_apply", p, keep=keep))


def groupnorm_stats_buffer(nimg: int, groups: int, device="cuda") -> torch.Tensor:
    """Zeroed statistics scratch for `groupnorm` / fused GEMM statistics: int32 [nimg, groups, GN_STATS_WORDS]
    (64-bit fixed-point sums; size = b200sd_groupnorm_workspace_bytes)."""
    return torch.zeros(nimg, groups, L.GN_STATS_WORDS, dtype=torch.int32, device=device)


def groupnorm_stats_values(buf: torch.Tensor) -> torch.Tensor:
    """Decode a statistics scratch (see groupnorm_stats_buffer) to float64 [nimg, groups, 2] = (sum, sum of squares):
    each is two int64 words (H, L) with value H * 128 + L * 2^-24.  For tests and tools."""
    w = buf.contiguous().view(torch.int64).reshape(buf.shape[0], buf.shape[1], 2, 2).to(torch.float64)
    return w[..., 0] * 128.0 + w[..., 1] * 2.0 ** -24


def groupnorm_workspace_bytes(*, nimg, groups) -> int:
    p = L.GroupNormParams()
    p.nimg, p.groups = nimg, groups
    return int(L.load().b200sd_groupnorm_workspace_bytes(C.byref(p)))


def layernorm(x, out, gamma, beta, *, rows, c, eps):
    p = L.LayerNormParams()
    p.x, p.gamma, p.beta, p.out, p.rows, p.c, p.eps = _ptr(x), _ptr(gamma), _ptr(beta), _ptr(out), rows, c, eps
    return L.Call("b200sd_layernorm", p, keep=(x, gamma, beta, out))


def prep_unet_input(latents, out, step_table, step_counter, *, b, hw, cfg):
    p = L.PrepParams()
    p.latents, p.out, p.step_table, p.step_counter = _ptr(latents), _ptr(out), _ptr(step_table), _ptr(step_counter)
    p.b, p.hw, p.cfg = b, hw, int(cfg)
    return L.Call("b200sd_prep_unet_input", p, keep=(latents, out, step_table, step_counter))


def timestep_embedding(out, step_table, step_counter, *, rows, dim):
    p = L.TembParams()
    p.out, p.step_table, p.step_counter, p.rows, p.dim = _ptr(out), _ptr(step_table), _ptr(step_counter), rows, dim
    return L.Call("b200sd_timestep_embedding", p, keep=(out, step_table, step_counter))


def cfg_scheduler_step(eps, latents, history, cur_sample, step_table, step_counter, *, b, hw, ld_eps, cfg,
                       guidance_scale):
    p = L.SchedParams()
    p.eps, p.latents, p.history, p.cur_sample = _ptr(eps), _ptr(latents), _ptr(history), _ptr(cur_sample)
    p.step_table, p.step_counter = _ptr(step_table), _ptr(step_counter)
    p.b, p.hw, p.ld_eps, p.cfg, p.guidance_scale = b, hw, ld_eps, int(cfg), guidance_scale
    return L.Call("b200sd_cfg_scheduler_step", p, keep=(eps, latents, history, cur_sample, step_table, step_counter))


def clip_embed(ids, tok, pos, out, *, rows, seq, c, vocab):
    p = L.EmbedParams()
    p.ids, p.tok, p.pos, p.out, p.rows, p.seq, p.c, p.vocab = _ptr(ids), _ptr(tok), _ptr(pos), _ptr(out), rows, seq, c, vocab
    return L.Call("b200sd_clip_embed", p, keep=(ids, tok, pos, out))


def denorm_u8(x, out, *, npix, ldx):
    p = L.DenormParams()
    p.x, p.out, p.npix, p.ldx = _ptr(x), _ptr(out), npix, ldx
    return L.Call("b200sd_denorm_u8", p, keep=(x, out))


def latent_to_act(latents, out, *, b, hw, scale):
    p = L.LatentCastParams()
    p.latents, p.out, p.b, p.hw, p.scale = _ptr(latents), _ptr(out), b, hw, scale
    return L.Call("b200sd_latent_to_act", p, keep=(latents, out))


def softmax_rows(x, out, *, rows, cols):
    p = L.SoftmaxParams()
    p.x, p.out, p.rows, p.cols = _ptr(x), _ptr(out), rows, cols
    return L.Call("b200sd_softmax_rows", p, keep=(x, out))


def memset_zero(t, nbytes=None):
    p = L.MemsetParams()
    p.ptr, p.bytes = _ptr(t), (t.numel() * t.element_size() if nbytes is None else nbytes)
    return L.Call("b200sd_memset_zero", p, keep=(t,))


def current_stream() -> int:
    return torch.cuda.current_stream().cuda_stream
