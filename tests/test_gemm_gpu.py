"""GPU parity: tcgen05 implicit-GEMM conv / linear kernel (through the C ABI) vs an fp32 torch restatement of the
same op on identical 16-bit-rounded inputs.  Tolerance: 16-bit output rounding (<= 2^-8 relative) + fp32 accumulation order,
i.e. |err| <= 1e-2 * max|ref| elementwise-max — stated per test."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def ACT():
    """The library's 16-bit activation dtype (fp16 by default; bf16 build optional)."""
    import b200sd
    return b200sd.lib.act_dtype()


def _report(name, got, ref, tol):
    got = got.float()
    ref = ref.float()
    err = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-12
    mx = err.max().item()
    if not (mx <= tol * scale) or torch.isnan(got).any():
        idx = torch.nonzero(err > tol * scale)
        print(f"\n[{name}] FAIL max_err={mx:.4g} scale={scale:.4g} bad={idx.shape[0]}/{err.numel()} nan={torch.isnan(got).sum().item()}")
        print("  first bad idx:", idx[:8].tolist())
        flat = err.reshape(-1, err.shape[-1])
        print("  bad rows (first 16):", torch.nonzero(flat.max(dim=1).values > tol * scale)[:16].flatten().tolist())
        print("  bad cols (first 16):", torch.nonzero(flat.max(dim=0).values > tol * scale)[:16].flatten().tolist())
        print("  got[0,:8]", got.reshape(-1, got.shape[-1])[0, :8].tolist())
        print("  ref[0,:8]", ref.reshape(-1, ref.shape[-1])[0, :8].tolist())
    assert mx <= tol * scale, f"{name}: max err {mx} > {tol} * {scale}"
    assert not torch.isnan(got).any()


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(ACT())


@pytest.mark.parametrize("M,K,N,bn", [(128, 64, 128, 128), (300, 320, 200, 0), (16, 1280, 1280, 0),
                                      (1232, 768, 640, 0), (4096, 320, 960, 0), (513, 72, 40, 64),
                                      (700, 320, 320, 0), (2048, 640, 200, 128),
                                      # forced tile modes: 1000 + bn = single-CTA tiles, 2000 + bn = CTA pairs
                                      (300, 320, 200, 2128), (700, 320, 320, 2192), (4096, 1280, 320, 2256),
                                      (65536, 320, 960, 2256), (65536, 320, 960, 1256), (1000, 64, 64, 2064),
                                      # wide tiles (256 < bn <= 512): one TMEM accumulator, two MMAs / B loads a step
                                      (4096, 1280, 1280, 1320), (4096, 320, 1000, 512), (700, 192, 328, 384),
                                      (40000, 320, 960, 1320)])
def test_linear(cuda_lib, b200sd, M, K, N, bn):
    torch.manual_seed(0)
    ops = b200sd.ops
    x = _rand(M, K)
    w = _rand(N, K, scale=K ** -0.5)
    bias = torch.randn(N, device="cuda")
    wp = ops.pack_linear_weight(w.float().cpu()).cuda()
    out = torch.full((M, N), float("nan"), device="cuda", dtype=ACT())
    ops.gemm(x, wp, out, nimg=1, h=1, w_=M, c0=K, n=N, bias=bias, bn_hint=bn)(ops.current_stream())
    torch.cuda.synchronize()
    ref = x.float() @ w.float().t() + bias
    _report(f"linear {M}x{K}x{N}", out, ref, 1e-2)


def test_linear_epilogues(cuda_lib, b200sd):
    torch.manual_seed(1)
    ops, L = b200sd.ops, b200sd.lib
    M, K, N = 520, 256, 384
    x = _rand(M, K)
    w = _rand(N, K, scale=K ** -0.5)
    bias = torch.randn(N, device="cuda")
    res = _rand(M, N)
    wp = ops.pack_linear_weight(w.float().cpu()).cuda()
    acc = x.float() @ w.float().t()
    s = ops.current_stream()
    # residual + bias, 16-bit out
    out = torch.empty(M, N, device="cuda", dtype=ACT())
    ops.gemm(x, wp, out, nimg=1, h=1, w_=M, c0=K, n=N, bias=bias, residual=res)(s)
    torch.cuda.synchronize()
    _report("bias+residual", out, acc + bias + res.float(), 1e-2)
    # silu, fp32 out, alpha
    out32 = torch.empty(M, N, device="cuda", dtype=torch.float32)
    ops.gemm(x, wp, out32, nimg=1, h=1, w_=M, c0=K, n=N, bias=bias, act=L.ACT_SILU, alpha=0.5)(s)
    torch.cuda.synchronize()
    _report("silu fp32", out32, F.silu(0.5 * acc + bias), 2e-5 * 50)
    # quick gelu
    ops.gemm(x, wp, out, nimg=1, h=1, w_=M, c0=K, n=N, bias=bias, act=L.ACT_QUICK_GELU)(s)
    torch.cuda.synchronize()
    z = acc + bias
    _report("quick_gelu", out, z * torch.sigmoid(1.702 * z), 1e-2)
    # exact (erf) gelu, fp32 out so the activation itself is checked to ~1e-6
    ops.gemm(x, wp, out32, nimg=1, h=1, w_=M, c0=K, n=N, bias=bias, act=L.ACT_GELU)(s)
    torch.cuda.synchronize()
    _report("gelu fp32", out32, F.gelu(z), 1e-5)
    # GEGLU: weight [2F, K] -> packed; out [M, F]
    Fh = 256
    w2 = _rand(2 * Fh, K, scale=K ** -0.5)
    b2 = torch.randn(2 * Fh, device="cuda")
    perm = ops.geglu_perm(2 * Fh)
    wp2 = ops.pack_linear_weight(w2.float().cpu()[perm]).cuda()
    b2p = b2[perm.cuda()].contiguous()
    outg = torch.empty(M, Fh, device="cuda", dtype=ACT())
    ops.gemm(x, wp2, outg, nimg=1, h=1, w_=M, c0=K, n=2 * Fh, bias=b2p, act=L.ACT_GEGLU)(s)
    torch.cuda.synchronize()
    z = x.float() @ w2.float().t() + b2
    _report("geglu", outg, z[:, :Fh] * F.gelu(z[:, Fh:]), 1e-2)


@pytest.mark.parametrize("nimg,H,W,cin,cout", [(2, 16, 16, 64, 128), (3, 8, 8, 128, 64), (1, 64, 64, 320, 320),
                                               (2, 32, 32, 8, 320), (2, 64, 64, 320, 4), (1, 24, 20, 32, 96),
                                               (1, 256, 256, 128, 128)])
def test_conv3x3(cuda_lib, b200sd, nimg, H, W, cin, cout):
    torch.manual_seed(2)
    ops = b200sd.ops
    x = _rand(nimg, H, W, cin)                       # NHWC
    w = _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    bias = torch.randn(cout, device="cuda")
    wp = ops.pack_conv_weight(w.float().cpu()).cuda()
    out = torch.full((nimg, H, W, cout), float("nan"), device="cuda", dtype=torch.float32)
    ops.gemm(x, wp, out, nimg=nimg, h=H, w_=W, c0=cin, n=cout, ksize=3, bias=bias)(ops.current_stream())
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=1).permute(0, 2, 3, 1)
    _report(f"conv3x3 {nimg}x{H}x{W} {cin}->{cout}", out, ref, 2e-3)


def test_wide_tiles_conv_residual_geglu_f32(cuda_lib, b200sd):
    """Wide-tile plan through the other epilogues: 3x3 conv + residual + fused GroupNorm statistics, GEGLU, fp32 out."""
    torch.manual_seed(8)
    ops, L = b200sd.ops, b200sd.lib
    s = ops.current_stream()
    nimg, H, W, cin, cout = 16, 16, 16, 128, 640
    x, w = _rand(nimg, H, W, cin), _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    bias, res = torch.randn(cout, device="cuda"), _rand(nimg, H, W, cout)
    out = torch.full((nimg, H, W, cout), float("nan"), device="cuda", dtype=ACT())
    stats = ops.groupnorm_stats_buffer(nimg, 32)
    ops.gemm(x, ops.pack_conv_weight(w.float().cpu()).cuda(), out, nimg=nimg, h=H, w_=W, c0=cin, n=cout, ksize=3,
             bias=bias, residual=res, gn_stats=stats, gn_groups=32, bn_hint=320)(s)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=1).permute(0, 2, 3, 1) + res.float()
    _report("wide conv3x3 + residual", out, ref, 1e-2)
    g = out.float().reshape(nimg, H * W, 32, cout // 32)
    want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1)
    assert (ops.groupnorm_stats_values(stats) - want).abs().max().item() <= 2e-4 * want.abs().max().item() + 1e-2
    M, K, Fh = 4096, 320, 512
    xl = _rand(M, K)
    w2, b2 = _rand(2 * Fh, K, scale=K ** -0.5), torch.randn(2 * Fh, device="cuda")
    perm = ops.geglu_perm(2 * Fh)
    outg = torch.empty(M, Fh, device="cuda", dtype=ACT())
    ops.gemm(xl, ops.pack_linear_weight(w2.float().cpu()[perm]).cuda(), outg, nimg=1, h=1, w_=M, c0=K, n=2 * Fh,
             bias=b2[perm.cuda()].contiguous(), act=L.ACT_GEGLU, bn_hint=512)(s)
    torch.cuda.synchronize()
    z = xl.float() @ w2.float().t() + b2
    _report("wide geglu", outg, z[:, :Fh] * F.gelu(z[:, Fh:]), 1e-2)
    wl = _rand(352, K, scale=K ** -0.5)
    out32 = torch.empty(M, 352, device="cuda", dtype=torch.float32)
    ops.gemm(xl, ops.pack_linear_weight(wl.float().cpu()).cuda(), out32, nimg=1, h=1, w_=M, c0=K, n=352,
             bn_hint=384)(s)
    torch.cuda.synchronize()
    _report("wide fp32 out", out32, xl.float() @ wl.float().t(), 2e-5 * 50)


@pytest.mark.parametrize("mode", [1000, 2000])
def test_conv3x3_forced_tile_modes(cuda_lib, b200sd, mode):
    """Same conv through single-CTA tiles and through CTA pairs (cta_group::2), odd number of M tiles included."""
    torch.manual_seed(5)
    ops = b200sd.ops
    for nimg, H, W, cin, cout, bn in [(3, 16, 16, 128, 320, 192), (16, 64, 64, 320, 320, 192), (5, 8, 8, 256, 128, 128)]:
        x = _rand(nimg, H, W, cin)
        w = _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
        bias = torch.randn(cout, device="cuda")
        res = _rand(nimg, H, W, cout)
        out = torch.full((nimg, H, W, cout), float("nan"), device="cuda", dtype=ACT())
        ops.gemm(x, ops.pack_conv_weight(w.float().cpu()).cuda(), out, nimg=nimg, h=H, w_=W, c0=cin, n=cout, ksize=3,
                 bias=bias, residual=res, bn_hint=mode + bn)(ops.current_stream())
        torch.cuda.synchronize()
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=1).permute(0, 2, 3, 1) + res.float()
        _report(f"conv3x3 mode {mode} {nimg}x{H}x{W} {cin}->{cout}", out, ref, 1e-2)


def test_conv3x3_concat_temb_residual(cuda_lib, b200sd):
    """ResnetBlock2D pieces in one call: two-source (skip concat) conv + bias + per-image time-embedding row bias,
    and the 1x1 shortcut over the same virtual concat feeding a residual epilogue."""
    torch.manual_seed(3)
    ops = b200sd.ops
    nimg, H, W, c0, c1, cout = 4, 8, 8, 128, 64, 128
    a0, a1 = _rand(nimg, H, W, c0), _rand(nimg, H, W, c1)
    cin = c0 + c1
    w = _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    wsc = _rand(cout, cin, 1, 1, scale=cin ** -0.5)
    bias = torch.randn(cout, device="cuda")
    temb = torch.randn(nimg, 512, device="cuda")     # row-bias slice at column offset 128, ld 512
    s = ops.current_stream()
    sc = torch.empty(nimg, H, W, cout, device="cuda", dtype=ACT())
    ops.gemm(a0, ops.pack_conv_weight(wsc.float().cpu(), (c0, c1)).cuda(), sc, nimg=nimg, h=H, w_=W, c0=c0, a1=a1,
             c1=c1, n=cout, ksize=1)(s)
    out = torch.empty(nimg, H, W, cout, device="cuda", dtype=ACT())
    ops.gemm(a0, ops.pack_conv_weight(w.float().cpu(), (c0, c1)).cuda(), out, nimg=nimg, h=H, w_=W, c0=c0, a1=a1,
             c1=c1, n=cout, ksize=3, bias=bias, rowbias=temb[:, 128:], ld_rowbias=512, residual=sc)(s)
    torch.cuda.synchronize()
    xin = torch.cat([a0, a1], dim=-1).float().permute(0, 3, 1, 2)
    ref_sc = F.conv2d(xin, wsc.float()).permute(0, 2, 3, 1)
    _report("shortcut 1x1 concat", sc, ref_sc, 1e-2)
    ref = F.conv2d(xin, w.float(), bias, padding=1).permute(0, 2, 3, 1) + temb[:, None, None, 128:256] + sc.float()
    _report("conv3x3 concat+temb+residual", out, ref, 1e-2)


@pytest.mark.parametrize("nimg,H,W,cin,cout,groups,flat", [
    (2, 16, 16, 64, 128, 32, False), (3, 8, 8, 128, 320, 32, False), (1, 24, 20, 32, 96, 8, False),
    (4, 64, 64, 64, 320, 32, False), (5, 8, 8, 64, 640, 32, True), (2, 32, 32, 64, 64, 8, True)])
def test_fused_groupnorm_statistics(cuda_lib, b200sd, nimg, H, W, cin, cout, groups, flat):
    """The epilogue's (image, group) sums and sums of squares of the conv output (bias + time-embedding row bias +
    residual included) against torch on the fp32 result; `flat` = the im2col / linear geometry with rows per image."""
    torch.manual_seed(6)
    ops = b200sd.ops
    ks = 1 if flat else 3
    x = _rand(nimg, H, W, cin)
    w = _rand(cout, cin, ks, ks, scale=(ks * ks * cin) ** -0.5)
    bias = torch.randn(cout, device="cuda")
    res = _rand(nimg, H, W, cout)
    temb = torch.randn(nimg, cout, device="cuda")
    out = torch.empty(nimg, H, W, cout, device="cuda", dtype=ACT())
    stats = ops.groupnorm_stats_buffer(nimg, groups)
    wp = ops.pack_conv_weight(w.float().cpu()).cuda()
    if flat:
        geo = dict(nimg=1, h=1, w_=nimg * H * W, gn_rows_per_img=H * W)
        kw = {}
    else:
        geo = dict(nimg=nimg, h=H, w_=W)
        kw = dict(rowbias=temb, ld_rowbias=cout)
    ops.gemm(x, wp, out, c0=cin, n=cout, ksize=ks, bias=bias, residual=res, gn_stats=stats, gn_groups=groups,
             **geo, **kw)(ops.current_stream())
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=ks // 2).permute(0, 2, 3, 1) + res.float()
    if not flat:
        ref = ref + temb[:, None, None, :]
    _report("conv with fused stats", out, ref, 1e-2)
    g = ref.reshape(nimg, H * W, groups, cout // groups)
    want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1)
    err = (ops.groupnorm_stats_values(stats) - want).abs().max().item()
    scale = want.abs().max().item()
    print("fused GN stats max err", err, "scale", scale)
    assert err <= 2e-4 * scale + 1e-2


@pytest.mark.parametrize("nimg,H,W,cin,cout", [(2, 16, 16, 32, 64), (3, 64, 64, 320, 320), (2, 8, 8, 1280, 1280),
                                               (1, 24, 40, 64, 96)])
def test_conv3x3_stride2_tma_element_strides(cuda_lib, b200sd, nimg, H, W, cin, cout):
    """Downsample2D (conv 3x3, stride 2, padding 1) as an implicit GEMM: the A tiles are fetched with TMA element
    strides (every other input pixel), no patch matrix - against F.conv2d(stride=2, padding=1)."""
    torch.manual_seed(21)
    ops = b200sd.ops
    x = _rand(nimg, H, W, cin)
    w = _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    bias = torch.randn(cout, device="cuda")
    out = torch.full((nimg, H // 2, W // 2, cout), float("nan"), device="cuda", dtype=ACT())
    ops.gemm(x, ops.pack_conv_weight(w.float().cpu()).cuda(), out, nimg=nimg, h=H // 2, w_=W // 2, c0=cin, n=cout,
             ksize=3, kh=3, kw=3, off_h=-1, off_w=-1, stride=2, bias=bias)(ops.current_stream())
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, stride=2, padding=1).permute(0, 2, 3, 1)
    _report("conv3x3 stride 2", out, ref, 1e-2)


@pytest.mark.parametrize("nimg,H,W,cin,cout,groups", [(2, 8, 8, 64, 64, 0), (2, 32, 32, 320, 640, 32),
                                                      (1, 12, 20, 128, 128, 0), (3, 16, 16, 512, 256, 32)])
def test_upsample_conv_folded_into_four_phase_convs(cuda_lib, b200sd, nimg, H, W, cin, cout, groups):
    """Upsample2D (nearest 2x, then conv 3x3 pad 1) as four 2x2 convolutions on the low-resolution input with
    pre-summed weights, each writing one output phase through a strided output tensor map - against
    F.conv2d(F.interpolate(x, scale_factor=2, mode='nearest')).  With fused GroupNorm statistics of the full output."""
    torch.manual_seed(22)
    ops = b200sd.ops
    x = _rand(nimg, H, W, cin)
    w = _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    bias = torch.randn(cout, device="cuda")
    out = torch.full((nimg, 2 * H, 2 * W, cout), float("nan"), device="cuda", dtype=ACT())
    w4 = [t.cuda() for t in ops.pack_upconv_weights(w.float().cpu())]
    stats = ops.groupnorm_stats_buffer(nimg, groups) if groups else None
    strides = (2 * cout, 4 * W * cout, 4 * H * W * cout)
    s = ops.current_stream()
    for py in (0, 1):
        for px in (0, 1):
            ops.gemm(x, w4[py * 2 + px], out[:, py::2, px::2, :], nimg=nimg, h=H, w_=W, c0=cin, n=cout, ksize=1, kh=2,
                     kw=2, off_h=py - 1, off_w=px - 1, bias=bias, ldc=cout, out_strides=strides, gn_stats=stats,
                     gn_groups=groups)(s)
    torch.cuda.synchronize()
    up = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2, mode="nearest")
    ref = F.conv2d(up, w.float(), bias, padding=1).permute(0, 2, 3, 1)
    _report("upsample + conv3x3 (folded)", out, ref, 1e-2)
    if groups:
        g = out.float().reshape(nimg, 4 * H * W, groups, cout // groups)
        want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1).double()
        got = ops.groupnorm_stats_values(stats)
        assert (got - want).abs().max().item() <= 2e-4 * want.abs().max().item() + 1e-2


@pytest.mark.parametrize("split", [2, 3, 5])
def test_split_k_conv_and_linear(cuda_lib, b200sd, split):
    """Split-K: the reduction cut into slices on different thread blocks, slices added in index order by whichever
    block arrives last.  8x8-level conv (bias + time-embedding row bias + residual + fused GroupNorm statistics), a
    GEGLU linear and an fp32-output linear - each against torch, and bit-identical across repeated runs on a SHARED
    workspace (the counters re-arm themselves)."""
    torch.manual_seed(30 + split)
    ops, L = b200sd.ops, b200sd.lib
    s = ops.current_stream()
    nimg, H, W, cin, cout, groups = 6, 8, 8, 640, 320, 32
    x, w = _rand(nimg, H, W, cin), _rand(cout, cin, 3, 3, scale=(9 * cin) ** -0.5)
    bias, res, temb = torch.randn(cout, device="cuda"), _rand(nimg, H, W, cout), torch.randn(nimg, cout, device="cuda")
    wp = ops.pack_conv_weight(w.float().cpu()).cuda()
    M, K, Fh = 300, 1280, 256
    xl = _rand(M, K)
    w2, b2 = _rand(2 * Fh, K, scale=K ** -0.5), torch.randn(2 * Fh, device="cuda")
    perm = ops.geglu_perm(2 * Fh)
    w2p = ops.pack_linear_weight(w2.float().cpu()[perm]).cuda()
    b2p = b2[perm.cuda()].contiguous()
    wl = _rand(200, K, scale=K ** -0.5)
    wlp = ops.pack_linear_weight(wl.float().cpu()).cuda()

    def calls():
        out = torch.full((nimg, H, W, cout), float("nan"), device="cuda", dtype=ACT())
        stats = ops.groupnorm_stats_buffer(nimg, groups)
        outg = torch.full((M, Fh), float("nan"), device="cuda", dtype=ACT())
        out32 = torch.full((M, 200), float("nan"), device="cuda", dtype=torch.float32)
        cs = [ops.gemm(x, wp, out, nimg=nimg, h=H, w_=W, c0=cin, n=cout, ksize=3, bias=bias, residual=res, rowbias=temb,
                       ld_rowbias=cout, gn_stats=stats, gn_groups=groups, split_k=split),
              ops.gemm(xl, w2p, outg, nimg=1, h=1, w_=M, c0=K, n=2 * Fh, bias=b2p, act=L.ACT_GEGLU, split_k=split),
              ops.gemm(xl, wlp, out32, nimg=1, h=1, w_=M, c0=K, n=200, split_k=split)]
        return cs, (out, stats, outg, out32)

    cs, _ = calls()
    need = max(ops.gemm_workspace_bytes(c) for c in cs)
    assert need > 0
    ws = torch.zeros((need + 255) // 256 * 64, dtype=torch.int32, device="cuda")  # zeroed once, shared by every call
    results = []
    for _ in range(3):
        cs, outs = calls()
        for c in cs:
            c.params.workspace, c.params.workspace_bytes = ws.data_ptr(), ws.numel() * 4
            c(s)
        torch.cuda.synchronize()
        results.append(outs)
    out, stats, outg, out32 = results[0]
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=1).permute(0, 2, 3, 1) + res.float() \
        + temb[:, None, None, :]
    _report(f"split-{split} conv3x3 + temb + residual", out, ref, 1e-2)
    g = out.float().reshape(nimg, H * W, groups, cout // groups)
    want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1).double()
    assert (ops.groupnorm_stats_values(stats) - want).abs().max().item() <= 2e-4 * want.abs().max().item() + 1e-2
    z = xl.float() @ w2.float().t() + b2
    _report(f"split-{split} geglu", outg, z[:, :Fh] * F.gelu(z[:, Fh:]), 1e-2)
    _report(f"split-{split} fp32 out", out32, xl.float() @ wl.float().t(), 1e-3)
    for other in results[1:]:
        assert all(torch.equal(a, b) for a, b in zip(results[0], other))   # order-independent slice reduction
    # without a workspace the call is rejected, nothing is launched
    bad = ops.gemm(xl, wlp, out32, nimg=1, h=1, w_=M, c0=K, n=200, split_k=split)
    with pytest.raises(Exception, match="workspace"):
        bad(s)


@pytest.mark.parametrize("M,C,N,geglu", [(4096, 320, 960, False), (1000, 640, 640, False), (4096, 320, 2560, True),
                                         (300, 1280, 10240, True)])
def test_layernorm_folded_into_producer_and_consumer(cuda_lib, b200sd, M, C, N, geglu):
    """nn.LayerNorm between two linear layers without a normalised tensor: the producer (linear + residual) adds every
    output row's sum / sum of squares to integer slots, the consumer runs on the raw rows with gamma-scaled weights and
    applies rstd_m * acc - rstd_m * mean_m * colsum_n + (b + W beta) in its epilogue - against
    F.linear(F.layer_norm(y)) (and GEGLU of it) on the same 16-bit-rounded y."""
    torch.manual_seed(40)
    ops, L = b200sd.ops, b200sd.lib
    s = ops.current_stream()
    x0, res = _rand(M, C), _rand(M, C, scale=2.0)
    w0, b0 = _rand(C, C, scale=C ** -0.5), torch.randn(C, device="cuda")
    y = torch.empty(M, C, device="cuda", dtype=ACT())
    rows = ops.groupnorm_stats_buffer(M, 1)
    ops.gemm(x0, ops.pack_linear_weight(w0.float().cpu()).cuda(), y, nimg=1, h=1, w_=M, c0=C, n=C, bias=b0, residual=res,
             ln_out_stats=rows)(s)
    torch.cuda.synchronize()
    yf = y.float()
    want = torch.stack([yf.sum(dim=1), (yf * yf).sum(dim=1)], dim=-1).double()[:, None, :]
    got = ops.groupnorm_stats_values(rows)
    assert ((got - want).abs() <= 1e-5 * want.abs() + 1e-3).all()
    # consumer
    gamma, beta = torch.randn(C, device="cuda") * 0.5 + 1.0, torch.randn(C, device="cuda") * 0.3
    w1, b1 = _rand(N, C, scale=C ** -0.5).float(), torch.randn(N, device="cuda")
    wq = (w1 * gamma[None, :]).to(ACT())
    bq = b1 + w1 @ beta
    colsum = wq.float().sum(dim=1).contiguous()
    perm = ops.geglu_perm(N).cuda() if geglu else torch.arange(N, device="cuda")
    out = torch.empty(M, N // 2 if geglu else N, device="cuda", dtype=ACT())
    ops.gemm(y, ops.pack_linear_weight(wq.float().cpu()[perm.cpu()]).cuda(), out, nimg=1, h=1, w_=M, c0=C, n=N,
             bias=bq[perm].contiguous(), act=L.ACT_GEGLU if geglu else L.ACT_NONE, ln_in_stats=rows,
             ln_colsum=colsum[perm].contiguous(), ln_eps=1e-5)(s)
    torch.cuda.synchronize()
    z = F.linear(F.layer_norm(yf, (C,), gamma, beta, 1e-5), w1, b1)
    ref = z[:, :N // 2] * F.gelu(z[:, N // 2:]) if geglu else z
    _report(f"LN folded {M}x{C}x{N} geglu={geglu}", out, ref, 1e-2)


def test_gemm_rejects_bad_arguments(cuda_lib, b200sd):
    ops = b200sd.ops
    x = _rand(64, 60)
    w = _rand(64, 64)
    out = torch.empty(64, 64, device="cuda", dtype=ACT())
    with pytest.raises(b200sd.lib.B200SDError, match="multiples of 8"):
        ops.gemm(x, w, out, nimg=1, h=1, w_=64, c0=60, n=64)(ops.current_stream())
    with pytest.raises(b200sd.lib.B200SDError, match="ksize"):
        ops.gemm(x, w, out, nimg=1, h=1, w_=64, c0=64, n=64, ksize=5)(ops.current_stream())
