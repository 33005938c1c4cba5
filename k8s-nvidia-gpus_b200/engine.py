"""Host-side plan builder: turns a model description + weights into a flat list of prepared C-ABI calls.

A `Plan` is built once per (architecture, batch, resolution); running it enqueues every kernel on the current CUDA
stream with no allocation and no host synchronisation, so a plan can be captured into a CUDA graph and replayed.
Activations are NHWC bf16 (`Act`), which makes a feature map [N, H, W, C] and its token view [N, H*W, C] the same
memory.
"""
from __future__ import annotations

import os

import torch

from . import lib as L
from . import ops


class Act:
    """NHWC 16-bit activation: tensor [nimg, h, w, c] (contiguous).  `stats` = (statistics scratch slice, groups) when
    the kernel that produced it also accumulated its GroupNorm statistics (fused into the GEMM epilogue)."""
    __slots__ = ("t", "nimg", "h", "w", "c", "stats", "rowstats")

    def __init__(self, t, nimg, h, w, c, stats=None, rowstats=None):
        self.t, self.nimg, self.h, self.w, self.c, self.stats = t, nimg, h, w, c, stats
        self.rowstats = rowstats  # per-row (sum, sum of squares) slots filled by the producing GEMM (folded LayerNorm)

    @property
    def hw(self):
        return self.h * self.w

    @property
    def rows(self):
        return self.nimg * self.h * self.w


class Plan:
    def __init__(self):
        self.calls = []
        self.keep = []
        self.graph = None

    def run(self, stream: int | None = None):
        s = ops.current_stream() if stream is None else stream
        for c in self.calls:
            c(s)

    def capture(self):
        """Capture the plan into a CUDA graph (after one eager warm-up run on the side stream)."""
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self.run()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.run()
        self.graph = g
        return g

    def launch(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.run()

    def __len__(self):
        return len(self.calls)

    def kernel_launches(self) -> int:
        return len(self.calls)


class DeviceWeights:
    """Packs a (diffusers-layout) CPU state dict into the kernels' layouts on first use and caches the result."""

    def __init__(self, sd: dict, device, prefix: str = ""):
        self.sd, self.device, self.prefix = sd, device, prefix
        self.cache = {}

    def raw(self, name):
        return self.sd[self.prefix + name].float()

    def has(self, name):
        return (self.prefix + name) in self.sd

    def _get(self, key, fn):
        if key not in self.cache:
            self.cache[key] = fn().to(self.device)
        return self.cache[key]

    def conv(self, name, split=None):
        def f():
            w = self.raw(name + ".weight")
            if w.dim() == 2:
                w = w[:, :, None, None]
            return ops.pack_conv_weight(w, split)
        return self._get(("conv", name, split), f)

    def cat_linear(self, names, key=None, perm=None):
        """Row-concatenated linear weights (fused QKV / KV / all time-embedding projections)."""
        def f():
            w = torch.cat([self.raw(n + ".weight") for n in names], dim=0)
            if perm is not None:
                w = w[perm]
            return ops.pack_linear_weight(w)
        return self._get(("catlin", tuple(names), key), f)

    def cat_conv(self, names, split):
        """Row-concatenated linear weights whose K axis is split over several A sources (virtual concat)."""
        def f():
            w = torch.cat([self.raw(n + ".weight") for n in names], dim=0)
            return ops.pack_conv_weight(w[:, :, None, None], tuple(split))
        return self._get(("catconv", tuple(names), tuple(split)), f)

    def cat_bias(self, names, key=None, perm=None):
        def f():
            b = torch.cat([self.raw(n + ".bias") for n in names], dim=0)
            if perm is not None:
                b = b[perm]
            return b.contiguous()
        return self._get(("catbias", tuple(names), key), f)

    def upconv(self, name):
        """The four phase weights of a folded nearest-2x + conv3x3 (ops.pack_upconv_weights)."""
        def f():
            return torch.stack(ops.pack_upconv_weights(self.raw(name + ".weight")))
        return self._get(("upconv", name), f)

    def ln_linear(self, ln_name, names, key=None, perm=None, bias=True):
        """A linear layer (row-concatenated `names`) with the preceding LayerNorm `ln_name` folded in (b200sd.h, ln_*):
        returns (packed W' = gamma (.) W, bias' = b + W beta (fp32), colsum[n] = sum_k W'[n, k] of the ROUNDED weights)."""
        def f():
            w = torch.cat([self.raw(n + ".weight") for n in names], dim=0)
            b = (torch.cat([self.raw(n + ".bias") for n in names], dim=0) if bias
                 else torch.zeros(w.shape[0], dtype=torch.float32, device=w.device))
            gamma, beta = self.raw(ln_name + ".weight"), self.raw(ln_name + ".bias")
            b = b + w @ beta
            wq = (w * gamma[None, :]).to(ops.act_dtype())
            if perm is not None:
                wq, b = wq[perm], b[perm]
            return {"w": ops.pack_linear_weight(wq.float()), "b": b.contiguous(),
                    "s": wq.float().sum(dim=1).contiguous()}
        key = ("lnlin", ln_name, tuple(names), key)
        if key not in self.cache:
            self.cache[key] = {k: v.to(self.device) for k, v in f().items()}
        return self.cache[key]

    def vec(self, name):
        return self._get(("vec", name), lambda: self.raw(name).contiguous())

    def act16(self, name):
        return self._get(("act16", name), lambda: self.raw(name).to(ops.act_dtype()).contiguous())


class Builder:
    def __init__(self, device, stats_words: int = 1 << 22, fuse_gn_stats: bool = False, gn_fuse_min_k: int = 24 * 64):
        """fuse_gn_stats: let long-reduction 3x3 convs accumulate the GroupNorm statistics of their output in the
        epilogue.  Measured: the VAE decoder gains 9 % (its statistics passes stream 0.5 GB tensors), the UNet step
        about 1 % (43 of 61 passes go, the conv epilogues get longer)."""
        self.device = device
        self.fuse_gn_stats = fuse_gn_stats
        # shortest reduction (elements of K) whose conv epilogue also takes the GroupNorm statistics of its output
        self.gn_fuse_min_k = gn_fuse_min_k
        self.plan = Plan()
        self.pool = {}
        # GroupNorm statistics scratch of the whole plan (4-byte words; one memset per run)
        # (handed out in chunks of stats_words; every used chunk is zeroed by one memset at the start of the plan)
        self.stats_words = stats_words
        self.stat_chunks = []   # [tensor, words used]
        # split-K policy (b200sd.h, split_k): "shape" (default) cuts the long reductions of the 8x8-level convolutions
        # into 3 slices - a function of the per-image shape only, so an image's bytes do not depend on the batch it
        # rides in; "auto" sizes the split from the tile count of the actual launch (best at batch 1-2, gives up that
        # bitwise batch invariance); "0" turns it off.
        self.split_policy = os.environ.get("B200SD_SPLITK", "shape")
        self.split_calls = []

    # ------------------------------------------------------------------ memory
    def alloc(self, shape, dtype=None):
        dtype = dtype or ops.act_dtype()
        n = 1
        for s in shape:
            n *= s
        key = (n * torch.empty((), dtype=dtype).element_size(), dtype)
        lst = self.pool.get(key)
        if lst:
            return lst.pop().view(shape)
        t = torch.empty(shape, dtype=dtype, device=self.device)
        self.plan.keep.append(t)
        return t

    def persistent(self, shape, dtype=None, zero=True):
        dtype = dtype or ops.act_dtype()
        t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.device)
        self.plan.keep.append(t)
        return t

    def free(self, *acts):
        for a in acts:
            t = a.t if isinstance(a, Act) else a
            if t is None:
                continue
            key = (t.numel() * t.element_size(), t.dtype)
            self.pool.setdefault(key, []).append(t.reshape(-1))

    def add(self, call):
        self.plan.calls.append(call)

    def stats_slice(self, nimg, groups):
        """Zero-at-plan-start statistics slots: nimg * groups of them (GroupNorm: (image, group); folded LayerNorm:
        (row, 1))."""
        nst = nimg * groups * L.GN_STATS_WORDS
        if not self.stat_chunks or self.stat_chunks[-1][1] + nst > self.stat_chunks[-1][0].numel():
            t = torch.zeros(max(self.stats_words, nst), dtype=torch.int32, device=self.device)
            self.plan.keep.append(t)
            self.stat_chunks.append([t, 0])
        ch = self.stat_chunks[-1]
        st = ch[0][ch[1]:ch[1] + nst]
        ch[1] += nst
        return st

    @staticmethod
    def _pow2_ceil(v):
        return 1 << max(0, (v - 1).bit_length())

    def split_k_for(self, rows_per_img: int, m_rows: int, n: int, nkb: int) -> int:
        """Number of K slices for a GEMM of m_rows x n with nkb 64-deep K blocks (see split_policy)."""
        if self.split_policy == "shape":
            return 3 if (rows_per_img <= 64 and nkb >= 96) else 1
        if self.split_policy == "auto":
            tiles = -(-m_rows // 128) * -(-n // 256)
            if tiles >= 90 or nkb < 32:
                return 1
            return max(1, min(8, nkb // 16, 148 // tiles))
        return 1

    def add_gemm(self, call, split_k: int):
        """Queue a gemm call; split-K calls get the plan's shared workspace patched in by finish()."""
        if split_k > 1:
            self.split_calls.append(call)
        self.add(call)

    def finish(self) -> Plan:
        if self.split_calls:
            # one zero-initialised workspace shared by every split-K call of the plan (stream order serialises them,
            # and each call leaves the counters at zero)
            need = max(ops.gemm_workspace_bytes(c) for c in self.split_calls)
            ws = torch.zeros((need + 255) // 256 * 64, dtype=torch.int32, device=self.device)
            self.plan.keep.append(ws)
            for c in self.split_calls:
                c.params.workspace, c.params.workspace_bytes = ws.data_ptr(), ws.numel() * 4
            self.split_calls = []
        for t, used in self.stat_chunks:
            if used:
                self.plan.calls.insert(0, ops.memset_zero(t, used * 4))
        return self.plan

    # ------------------------------------------------------------------ layers
    def conv(self, W: DeviceWeights, name, x, cout, ksize, *, bias=True, rowbias=None, ld_rowbias=0, residual=None,
             act=L.ACT_NONE, out_dtype=None, out=None, ldc=None, wsplit=None, gn_groups=None, rowstats=False):
        """x: Act or (Act, Act) virtual concat.  Returns Act [nimg, h, w, cout].  gn_groups: also accumulate the
        GroupNorm statistics of the output in the epilogue (long reductions only - a 3x3 conv hides it completely;
        tiles must hold >= 32 pixels of an image, see b200sd.h)."""
        xs = x if isinstance(x, tuple) else (x,)
        a0 = xs[0]
        st = None
        if (self.fuse_gn_stats and gn_groups and ksize == 3 and out is None and ldc is None and out_dtype in (None, ops.act_dtype())
                and cout % (2 * gn_groups) == 0 and 9 * sum(a.c for a in xs) >= self.gn_fuse_min_k  # long reductions only
                and min(128, self._pow2_ceil(a0.w)) * self._pow2_ceil(a0.h) >= 32):
            st = self.stats_slice(a0.nimg, gn_groups)
        split = wsplit or tuple(a.c for a in xs)  # wsplit: real weight channels when the activation is zero padded
        w = W.conv(name, split)
        b = W.vec(name + ".bias") if bias else None
        if out is None:
            out = self.alloc((a0.nimg, a0.h, a0.w, cout), out_dtype)
        a1 = xs[1] if len(xs) > 1 else None
        nkb = ksize * ksize * sum(-(-a.c // 64) for a in xs)
        sk = self.split_k_for(a0.hw, a0.rows, cout, nkb) if out_dtype in (None, ops.act_dtype()) else 1
        self.add_gemm(ops.gemm(a0.t, w, out, nimg=a0.nimg, h=a0.h, w_=a0.w, c0=a0.c, n=cout, ksize=ksize,
                               a1=None if a1 is None else a1.t, c1=0 if a1 is None else a1.c, bias=b, rowbias=rowbias,
                               ld_rowbias=ld_rowbias, residual=None if residual is None else residual.t,
                               ldr=0 if residual is None else residual.c, ldc=ldc, act=act, gn_stats=st,
                               gn_groups=gn_groups if st is not None else 0, split_k=sk, weights_static=True,
                               ln_out_stats=(rs := self.stats_slice(a0.rows, 1) if rowstats else None)), sk)
        return Act(out, a0.nimg, a0.h, a0.w, cout if ldc is None else ldc,
                   stats=None if st is None else (st, gn_groups), rowstats=rs)

    def linear_w(self, x: Act, w, b, n, *, residual=None, act=L.ACT_NONE, out_dtype=None, rowbias=None,
                 ld_rowbias=0, per_image_rows=False, rowstats=False, ln=None, ln_eps=1e-5):
        """Linear layer over the rows of x with an already packed weight.  rowstats: also emit the per-row statistics of
        the output (it feeds a folded LayerNorm).  ln: colsum vector - x is the RAW input of a LayerNorm folded into
        this layer (w, b from DeviceWeights.ln_linear; x.rowstats from its producer)."""
        assert ln is None or x.rowstats is not None, "folded LayerNorm: the input carries no row statistics"
        n_out = n // 2 if act == L.ACT_GEGLU else n
        out = self.alloc((x.nimg, x.h, x.w, n_out), out_dtype)
        if per_image_rows:
            geo = dict(nimg=x.rows, h=1, w_=1)
        else:
            geo = dict(nimg=1, h=1, w_=x.rows)
        sk = self.split_k_for(1 << 30, x.rows, n, -(-x.c // 64))  # (only the "auto" policy splits linear layers)
        self.add_gemm(ops.gemm(x.t, w, out, c0=x.c, n=n, ksize=1, bias=b,
                               residual=None if residual is None else residual.t,
                               ldr=0 if residual is None else residual.c, act=act, rowbias=rowbias,
                               ld_rowbias=ld_rowbias, split_k=sk, weights_static=True,
                               ln_out_stats=(rs := self.stats_slice(x.rows, 1) if rowstats else None),
                               ln_in_stats=x.rowstats if ln is not None else None, ln_colsum=ln, ln_eps=ln_eps,
                               **geo), sk)
        return Act(out, x.nimg, x.h, x.w, n_out, rowstats=rs)

    def linear(self, W: DeviceWeights, name, x: Act, n, *, bias=True, **kw):
        return self.linear_w(x, W.conv(name), W.vec(name + ".bias") if bias else None, n, **kw)

    def ln_linear(self, W: DeviceWeights, ln_name, names, x: Act, n, *, bias=True, key=None, perm=None, eps=1e-5, **kw):
        """LayerNorm(ln_name) -> Linear(names...) without a normalised tensor: see b200sd.h (ln_*)."""
        f = W.ln_linear(ln_name, names, key=key, perm=perm, bias=bias)
        return self.linear_w(x, f["w"], f["b"], n, ln=f["s"], ln_eps=eps, **kw)

    def groupnorm(self, W: DeviceWeights, name, x, groups, eps, silu):
        xs = x if isinstance(x, tuple) else (x,)
        a0 = xs[0]
        a1 = xs[1] if len(xs) > 1 else None
        C = sum(a.c for a in xs)
        out = self.alloc((a0.nimg, a0.h, a0.w, C))
        fused = a1 is None and a0.stats is not None and a0.stats[1] == groups
        st = a0.stats[0] if fused else self.stats_slice(a0.nimg, groups)
        s_call, a_call = ops.groupnorm(a0.t, out, st, W.vec(name + ".weight"), W.vec(name + ".bias"), nimg=a0.nimg,
                                       hw=a0.hw, c0=a0.c, groups=groups, eps=eps, silu=silu,
                                       x1=None if a1 is None else a1.t, c1=0 if a1 is None else a1.c)
        if not fused:  # otherwise the producing GEMM's epilogue has already accumulated the statistics
            self.add(s_call)
        self.add(a_call)
        return Act(out, a0.nimg, a0.h, a0.w, C)

    def layernorm(self, W: DeviceWeights, name, x: Act, eps=1e-5):
        out = self.alloc((x.nimg, x.h, x.w, x.c))
        self.add(ops.layernorm(x.t, out, W.vec(name + ".weight"), W.vec(name + ".bias"), rows=x.rows, c=x.c, eps=eps))
        return Act(out, x.nimg, x.h, x.w, x.c)

    def attention(self, q_t, k_t, v_t, *, batch, heads, nq, nk, d, ldq, ldk, ldv, causal=False, like: Act = None):
        C = heads * d
        out = self.alloc((like.nimg, like.h, like.w, C))
        self.add(ops.attention(q_t, k_t, v_t, out, batch=batch, heads=heads, nq=nq, nk=nk, d=d, ldq=ldq, ldk=ldk,
                               ldv=ldv, ldo=C, scale=d ** -0.5, causal=causal))
        return Act(out, like.nimg, like.h, like.w, C)

    def upsample_conv(self, W: DeviceWeights, name, x: Act, cout, gn_groups=None):
        """Upsample2D: nearest 2x + conv 3x3, as four 2x2 convolutions on the low-resolution input whose outputs
        interleave into [nimg, 2h, 2w, cout] (no upsampled tensor is materialised; 2.25x fewer MACs)."""
        out = self.alloc((x.nimg, 2 * x.h, 2 * x.w, cout))
        st = None
        if (self.fuse_gn_stats and gn_groups and cout % (2 * gn_groups) == 0 and 4 * x.c >= 16 * 64
                and min(128, self._pow2_ceil(x.w)) * self._pow2_ceil(x.h) >= 32):
            st = self.stats_slice(x.nimg, gn_groups)
        w4 = W.upconv(name)
        b = W.vec(name + ".bias")
        strides = (2 * cout, 4 * x.w * cout, 4 * x.h * x.w * cout)
        for py in (0, 1):
            for px in (0, 1):
                self.add(ops.gemm(x.t, w4[py * 2 + px], out[:, py::2, px::2, :], nimg=x.nimg, h=x.h, w_=x.w, c0=x.c,
                                  n=cout, ksize=1, kh=2, kw=2, off_h=py - 1, off_w=px - 1, bias=b, ldc=cout,
                                  out_strides=strides, gn_stats=st, gn_groups=gn_groups if st is not None else 0,
                                  weights_static=True))
        return Act(out, x.nimg, 2 * x.h, 2 * x.w, cout, stats=None if st is None else (st, gn_groups))

    def conv_s2(self, W: DeviceWeights, name, x: Act, cout, gn_groups=None):
        """Downsample2D: 3x3 stride-2 pad-1 conv as an implicit GEMM whose A tiles are fetched with TMA element
        strides (every other input pixel) - no patch matrix."""
        ho, wo = x.h // 2, x.w // 2
        assert x.h % 2 == 0 and x.w % 2 == 0
        out = self.alloc((x.nimg, ho, wo, cout))
        st = None
        if (self.fuse_gn_stats and gn_groups and cout % (2 * gn_groups) == 0 and 9 * x.c >= self.gn_fuse_min_k
                and min(128, self._pow2_ceil(wo)) * self._pow2_ceil(ho) >= 32):
            st = self.stats_slice(x.nimg, gn_groups)
        sk = self.split_k_for(ho * wo, x.nimg * ho * wo, cout, 9 * -(-x.c // 64))
        self.add_gemm(ops.gemm(x.t, W.conv(name), out, nimg=x.nimg, h=ho, w_=wo, c0=x.c, n=cout, ksize=3, kh=3, kw=3,
                               off_h=-1, off_w=-1, stride=2, bias=W.vec(name + ".bias"), gn_stats=st,
                               gn_groups=gn_groups if st is not None else 0, split_k=sk, weights_static=True), sk)
        return Act(out, x.nimg, ho, wo, cout, stats=None if st is None else (st, gn_groups))

    # ------------------------------------------------------------------ composite blocks (UNet + VAE)
    def resnet(self, W, p, x, cout, groups, eps, temb=None, ld_temb=0):
        """ResnetBlock2D: x Act or (Act, Act).  temb: fp32 tensor view [nimg, cout] (row stride ld_temb) or None.
        Both 3x3 convs also accumulate the GroupNorm statistics of their output (norm2 / the next block's norm)."""
        xs = x if isinstance(x, tuple) else (x,)
        cin = sum(a.c for a in xs)
        h = self.groupnorm(W, p + ".norm1", x, groups, eps, silu=True)
        h1 = self.conv(W, p + ".conv1", h, cout, 3, rowbias=temb, ld_rowbias=ld_temb, gn_groups=groups)
        self.free(h)
        h2 = self.groupnorm(W, p + ".norm2", h1, groups, eps, silu=True)
        self.free(h1)
        if cin != cout:
            sc = self.conv(W, p + ".conv_shortcut", x, cout, 1)
        else:
            assert len(xs) == 1
            sc = xs[0]
        out = self.conv(W, p + ".conv2", h2, cout, 3, residual=sc, gn_groups=groups)
        self.free(h2)
        if cin != cout:
            self.free(sc)
        return out
