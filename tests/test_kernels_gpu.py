"""GPU parity for the bandwidth-bound kernels and flash attention (C ABI) vs fp32 torch restatements."""
import math

import pytest
import torch
import torch.nn.functional as F

from test_gemm_gpu import ACT, _rand, _report

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nimg,hw,c0,c1,groups,silu", [(2, 64, 320, 0, 32, True), (3, 256, 1280, 640, 32, True),
                                                       (2, 1024, 640, 320, 32, False), (1, 4096, 128, 0, 32, True),
                                                       (2, 64, 1280, 1280, 32, True), (2, 100, 32, 0, 8, True)])
def test_groupnorm(cuda_lib, b200sd, nimg, hw, c0, c1, groups, silu):
    torch.manual_seed(0)
    ops = b200sd.ops
    C = c0 + c1
    x0 = (_rand(nimg, hw, c0).float() * 2 + 0.7).to(ACT())
    x1 = (_rand(nimg, hw, c1).float() * 0.5 - 1.1).to(ACT()) if c1 else None
    gamma = torch.randn(C, device="cuda")
    beta = torch.randn(C, device="cuda")
    stats = ops.groupnorm_stats_buffer(nimg, groups)
    out = torch.empty(nimg, hw, C, device="cuda", dtype=ACT())
    st, ap = ops.groupnorm(x0, out, stats, gamma, beta, nimg=nimg, hw=hw, c0=c0, groups=groups, eps=1e-5, silu=silu,
                           x1=x1, c1=c1)
    s = ops.current_stream()
    st(s)
    ap(s)
    torch.cuda.synchronize()
    x = x0.float() if x1 is None else torch.cat([x0.float(), x1.float()], dim=-1)
    ref = F.group_norm(x.permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    if silu:
        ref = F.silu(ref)
    _report(f"groupnorm {nimg}x{hw}x{C}", out, ref, 1e-2)


def test_groupnorm_is_deterministic_and_handles_large_means(cuda_lib, b200sd):
    """(1) The statistics are accumulated as integers: two runs on the same input give bit-identical scratch and output
    (fp32 atomics did not).  (2) mean >> std (the E[x^2] - mean^2 regime): exact totals + double finalisation keep
    the result on torch.group_norm."""
    torch.manual_seed(3)
    ops = b200sd.ops
    nimg, hw, C, groups = 4, 4096, 320, 32
    x = (20.0 + 0.25 * torch.randn(nimg, hw, C, device="cuda")).to(ACT())
    gamma, beta = torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
    outs, sts = [], []
    for _ in range(3):
        stats = ops.groupnorm_stats_buffer(nimg, groups)
        out = torch.empty(nimg, hw, C, device="cuda", dtype=ACT())
        st, ap = ops.groupnorm(x, out, stats, gamma, beta, nimg=nimg, hw=hw, c0=C, groups=groups, eps=1e-5, silu=False)
        st(ops.current_stream())
        ap(ops.current_stream())
        torch.cuda.synchronize()
        outs.append(out)
        sts.append(stats)
    assert torch.equal(sts[0], sts[1]) and torch.equal(sts[0], sts[2])
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    g = x.double().reshape(nimg, hw, groups, C // groups)
    want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1)
    got = ops.groupnorm_stats_values(sts[0])
    assert ((got - want).abs() <= 1e-6 * want.abs() + 1e-3).all()
    ref = F.group_norm(x.float().permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    _report("groupnorm mean >> std", outs[0], ref, 1e-2)
    assert ops.groupnorm_workspace_bytes(nimg=nimg, groups=groups) == stats.numel() * 4


@pytest.mark.parametrize("rows,c", [(77, 768), (4096, 320), (1000, 640), (300, 1280), (64, 32), (10, 2048)])
def test_layernorm(cuda_lib, b200sd, rows, c):
    torch.manual_seed(0)
    ops = b200sd.ops
    x = (_rand(rows, c).float() * 3 + 1).to(ACT())
    g, b = torch.randn(c, device="cuda"), torch.randn(c, device="cuda")
    out = torch.empty_like(x)
    ops.layernorm(x, out, g, b, rows=rows, c=c, eps=1e-5)(ops.current_stream())
    torch.cuda.synchronize()
    _report(f"layernorm {rows}x{c}", out, F.layer_norm(x.float(), (c,), g, b, 1e-5), 1e-2)


def test_softmax_rows(cuda_lib, b200sd):
    ops = b200sd.ops
    x = torch.randn(50, 4096, device="cuda") * 4
    out = torch.empty(50, 4096, device="cuda", dtype=ACT())
    ops.softmax_rows(x, out, rows=50, cols=4096)(ops.current_stream())
    torch.cuda.synchronize()
    _report("softmax_rows", out, torch.softmax(x, dim=-1), 1e-2)


@pytest.mark.parametrize("batch,heads,nq,nk,d,causal", [
    (1, 1, 128, 128, 64, False),     # single tile, aligned head dim
    (2, 3, 256, 384, 64, False),     # multiple kv tiles (online softmax + O rescale)
    (2, 8, 200, 77, 40, False),      # cross-attention shape: ragged nq/nk, d=40
    (2, 8, 1024, 1024, 80, False),   # two d atoms
    (1, 8, 256, 256, 160, False),    # three d atoms
    (2, 8, 64, 64, 160, False),      # 8x8 level
    (3, 12, 77, 77, 64, True),       # CLIP causal
    (1, 2, 300, 300, 64, True),      # causal across tiles
    (1, 8, 4096, 4096, 40, False),   # SD1.5 64x64 level
    # short-KV (nk <= 128, not causal): the persistent cross-attention kernel (attention_cross.cu)
    (4, 8, 4096, 77, 40, False),     # SD1.5 64x64 cross-attention: 3 pipelines, many items per CTA
    (2, 8, 1024, 77, 80, False),     # two d atoms, 2 pipelines
    (2, 8, 256, 77, 160, False),     # three d atoms, 1 pipeline
    (16, 8, 64, 77, 160, False),     # 8x8 level: one query tile per (batch, head) - a new K/V run every item
    (2, 10, 300, 77, 64, False),     # SDXL head dim, ragged queries
    (3, 4, 50, 5, 40, False),        # a handful of keys
    (1, 2, 130, 128, 64, False),     # the most keys this kernel takes (8 score chunks)
    (2, 3, 257, 100, 48, False),
])
def test_attention(cuda_lib, b200sd, batch, heads, nq, nk, d, causal):
    torch.manual_seed(0)
    ops = b200sd.ops
    C = heads * d
    # q from a fused QKV-like buffer (ld = 3C), k/v from a KV buffer (ld = 2C)
    qkv = _rand(batch, nq, 3 * C)
    kv = _rand(batch, nk, 2 * C)
    q = qkv[..., :C]
    k, v = kv[..., :C], kv[..., C:]
    out = torch.full((batch, nq, C), float("nan"), device="cuda", dtype=ACT())
    scale = d ** -0.5
    ops.attention(q, k, v, out, batch=batch, heads=heads, nq=nq, nk=nk, d=d, ldq=3 * C, ldk=2 * C, ldv=2 * C, ldo=C,
                  scale=scale, causal=causal)(ops.current_stream())
    torch.cuda.synchronize()
    qf = q.float().view(batch, nq, heads, d).transpose(1, 2)
    kf = k.float().view(batch, nk, heads, d).transpose(1, 2)
    vf = v.float().view(batch, nk, heads, d).transpose(1, 2)
    s = qf @ kf.transpose(-1, -2) * scale
    if causal:
        mask = torch.ones(nq, nk, device="cuda", dtype=torch.bool).tril()
        s = s.masked_fill(~mask, float("-inf"))
    ref = (torch.softmax(s, dim=-1) @ vf).transpose(1, 2).reshape(batch, nq, C)
    _report(f"attention b{batch} h{heads} {nq}x{nk} d{d} causal={causal}", out, ref, 2e-2)


@pytest.mark.parametrize("d,nk", [(40, 1024), (64, 1000), (80, 448), (160, 200), (40, 64), (64, 130)])
def test_attention_reference_moves_and_accumulator_merge(cuda_lib, b200sd, d, nk):
    """Scores built so that the softmax reference has to move late and differently in the even- and odd-tile
    accumulators (spikes in single key tiles, a first tile far below zero), and so that the two partial results are
    merged with very different weights."""
    torch.manual_seed(4)
    ops = b200sd.ops
    batch, heads, nq = 2, 3, 300
    C = heads * d
    q = _rand(batch, nq, C)
    k = _rand(batch, nk, C)
    v = _rand(batch, nk, C)
    kf32 = k.float().view(batch, nk, heads, d)
    qf32 = q.float().view(batch, nq, heads, d)
    # keys of tile 0: strongly anti-aligned with every query of head 0 (first-tile undershoot); one key of an odd tile
    # and one of a late even tile aligned with the mean query (late overshoot, different per accumulator)
    qm = qf32.mean(dim=1)                                 # [batch, heads, d]
    kf32[:, : min(64, nk), 0, :] = -6.0 * qm[:, None, 0, :] / qm[:, None, 0, :].norm(dim=-1, keepdim=True) - 2.0
    if nk > 64 + 7:
        kf32[:, 64 + 7, 1, :] = 40.0 * qm[:, 1, :] / qm[:, 1, :].norm(dim=-1, keepdim=True)
    if nk > 5 * 64:
        kf32[:, 4 * 64 + 3, 1, :] = 25.0 * qm[:, 1, :] / qm[:, 1, :].norm(dim=-1, keepdim=True)
        kf32[:, 5 * 64 - 1, 2, :] = 60.0
    k = kf32.reshape(batch, nk, C).to(ACT()).contiguous()
    out = torch.full((batch, nq, C), float("nan"), device="cuda", dtype=ACT())
    scale = 1.0  # unscaled scores: spread of several hundred in the log2 domain
    ops.attention(q, k, v, out, batch=batch, heads=heads, nq=nq, nk=nk, d=d, ldq=C, ldk=C, ldv=C, ldo=C,
                  scale=scale)(ops.current_stream())
    torch.cuda.synchronize()
    qf = q.float().view(batch, nq, heads, d).transpose(1, 2)
    kf = k.float().view(batch, nk, heads, d).transpose(1, 2)
    vf = v.float().view(batch, nk, heads, d).transpose(1, 2)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * scale, dim=-1) @ vf).transpose(1, 2).reshape(batch, nq, C)
    _report(f"attention spikes d{d} nk{nk}", out, ref, 2e-2)


def _table(rows):
    t = torch.zeros(len(rows), 16)
    for i, r in enumerate(rows):
        for k, v in r.items():
            t[i, k] = v
    return t.cuda()


def test_prep_temb_sched_ddim_euler(cuda_lib, b200sd):
    torch.manual_seed(0)
    ops, L = b200sd.ops, b200sd.lib
    b, hw = 3, 64
    lat = torch.randn(b, hw, 4, device="cuda")
    table = _table([
        {L.ST_TIMESTEP: 961.0, L.ST_IN_SCALE: 0.5, L.ST_KIND: 0, L.ST_C0: 0.9, L.ST_C1: 0.4, L.ST_C2: 0.6, L.ST_C3: 0.8},
        {L.ST_TIMESTEP: 41.0, L.ST_IN_SCALE: 1.0, L.ST_KIND: 1, L.ST_C0: -0.7, L.ST_C1: 3.0},
    ])
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    s = ops.current_stream()
    xin = torch.empty(2 * b, hw, 8, device="cuda", dtype=ACT())
    ops.prep_unet_input(lat, xin, table, counter, b=b, hw=hw, cfg=True)(s)
    temb = torch.empty(2 * b, 320, device="cuda", dtype=ACT())
    ops.timestep_embedding(temb, table, counter, rows=2 * b, dim=320)(s)
    torch.cuda.synchronize()
    ref_in = torch.cat([lat * 0.5, torch.zeros_like(lat)], dim=-1)
    _report("prep", xin, torch.cat([ref_in, ref_in]), 1e-2)
    i = torch.arange(160, device="cuda", dtype=torch.float32)
    ang = 961.0 * torch.exp(-math.log(10000.0) * i / 160)
    ref_t = torch.cat([torch.cos(ang), torch.sin(ang)]).expand(2 * b, 320)
    _report("temb", temb, ref_t, 1e-2)
    # SURVEY App. C anchors (t=961, dim 320)
    assert abs(temb[0, 0].item() - 0.94689709) < 4e-3 and abs(temb[0, 160].item() + 0.32153678) < 4e-3

    eps = torch.randn(2 * b, hw, 4, device="cuda")
    gs = 7.5
    e = eps[:b] + gs * (eps[b:] - eps[:b])
    lat0 = lat.clone()
    ops.cfg_scheduler_step(eps, lat, None, None, table, counter, b=b, hw=hw, ld_eps=4, cfg=True, guidance_scale=gs)(s)
    torch.cuda.synchronize()
    x0 = (lat0 - 0.9 * e) / 0.4
    _report("ddim", lat, 0.6 * x0 + 0.8 * e, 1e-5)
    assert counter.tolist() == [1, 0]
    lat1 = lat.clone()
    ops.cfg_scheduler_step(eps, lat, None, None, table, counter, b=b, hw=hw, ld_eps=4, cfg=True, guidance_scale=gs)(s)
    torch.cuda.synchronize()
    _report("euler", lat, lat1 + e * -0.7, 1e-5)
    assert counter.tolist() == [2, 0]


def test_layout_kernels(cuda_lib, b200sd):
    torch.manual_seed(0)
    ops = b200sd.ops
    s = ops.current_stream()
    ids = torch.randint(0, 1000, (2 * 77,), device="cuda", dtype=torch.int32)
    tok, pos = _rand(1000, 64), _rand(77, 64)
    emb = torch.empty(2 * 77, 64, device="cuda", dtype=ACT())
    ops.clip_embed(ids, tok, pos, emb, rows=2 * 77, seq=77, c=64, vocab=1000)(s)
    torch.cuda.synchronize()
    ref = tok.float()[ids.long()] + pos.float().repeat(2, 1)
    _report("clip_embed", emb, ref, 1e-2)

    img = torch.randn(1000, 4, device="cuda") * 1.5
    u8 = torch.empty(1000, 3, device="cuda", dtype=torch.uint8)
    ops.denorm_u8(img, u8, npix=1000, ldx=4)(s)
    torch.cuda.synchronize()
    ref = ((img[:, :3] / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)
    assert torch.equal(u8, ref)

    lat = torch.randn(2, 64, 4, device="cuda")
    o = torch.empty(2, 64, 8, device="cuda", dtype=ACT())
    ops.latent_to_act(lat, o, b=2, hw=64, scale=1 / 0.18215)(s)
    torch.cuda.synchronize()
    _report("latent_to_bf16", o, torch.cat([lat / 0.18215, torch.zeros_like(lat)], -1), 1e-2)


@pytest.mark.parametrize("n", [2, 4, 25])
def test_cfg_sched_kernel_pndm_follows_oracle_scheduler(cuda_lib, b200sd, n):
    """cfg_sched_kernel's PNDM / PLMS path driven by the product's own step table - first step (mode 1: keep the
    sample), repeated second step (mode 2: average with the stored eps, restore the sample, no history push), then the
    2-, 3- and 4-term linear multistep combines over the device history ring - against oracle.schedulers.PNDMScheduler
    fed the same guided eps (fp32, 2e-5)."""
    from oracle.schedulers import PNDMScheduler
    ops, S = b200sd.ops, b200sd.schedulers
    torch.manual_seed(n)
    b, hw, gs = 2, 96, 7.5
    sched = S.make_schedule("pndm", n, steps_offset=1)
    o = PNDMScheduler(steps_offset=1)
    o.set_timesteps(n)
    assert sched.num_evals == len(o.timesteps) == n + 1
    table = torch.zeros(sched.num_evals, 16)
    table.copy_(sched.table)
    table = table.cuda()
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    lat = torch.randn(b, hw, 4, device="cuda")
    hist = torch.zeros(4, b, hw, 4, device="cuda")
    cur = torch.zeros(b, hw, 4, device="cuda")
    x_o = lat.cpu().clone()
    s = ops.current_stream()
    for i, t in enumerate(o.timesteps):
        eps = torch.randn(2 * b, hw, 4, device="cuda") * 0.7
        ops.cfg_scheduler_step(eps, lat, hist, cur, table, counter, b=b, hw=hw, ld_eps=4, cfg=True,
                               guidance_scale=gs)(s)
        torch.cuda.synchronize()
        e = eps.cpu()
        x_o = o.step(e[:b] + gs * (e[b:] - e[:b]), t, x_o)
        _report(f"pndm n={n} eval {i}", lat, x_o.cuda(), 2e-5)
        assert counter.tolist() == [i + 1, 0]
