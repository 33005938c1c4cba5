#!/usr/bin/env python3
"""Short, graph-free invocations of the hot path for ncu (see /opt/skills/guides/B200_PROFILING.md).

    python tools/profile_step.py step      # SD1.5 batch 8: text + K/V + 2 denoise steps + VAE, eager launches
    python tools/profile_step.py kernels   # the heaviest individual shapes of the UNet step, one launch each

Timing numbers printed here are CUDA-event times of the plain run; numbers under ncu are never bench values.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["B200SD_NO_GRAPH"] = "1"

import torch  # noqa: E402


def mode_step():
    import b200sd
    pipe = b200sd.StableDiffusionPipeline.from_pretrained("sd15", scheduler="ddim")
    job = pipe.job(8, 64, 64, 7.5, True)
    sched = pipe.make_schedule(25)
    job.step_table[: sched.num_evals].copy_(sched.table)
    job.init_latents.normal_()
    job.reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    job.run(2)
    e1.record()
    torch.cuda.synchronize()
    print(f"step mode: text+kv+2 steps+vae = {e0.elapsed_time(e1):.2f} ms, launches {job.launches(2)}")


def mode_kernels():
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    n = 16  # CFG-paired batch of 8

    def rnd(*shape, scale=1.0):
        return (torch.randn(*shape, device="cuda") * scale).to(dt)

    one_launch = os.environ.get("B200SD_PROFILE_ONE", "0") == "1"  # under ncu --set full: one launch per shape

    def timed(name, call, flops):
        if not one_launch:
            call(s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        call(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{name:44s} {ms:8.3f} ms  {flops / ms / 1e9:8.1f} TFLOP/s")

    convs = [(64, 320, 320), (32, 640, 640), (16, 1280, 1280), (8, 1280, 1280), (64, 640, 320), (16, 2560, 1280)]
    for hw, cin, cout in convs:
        x = rnd(n, hw, hw, cin)
        w = rnd(cout, 9, cin, scale=(9 * cin) ** -0.5)
        out = torch.empty(n, hw, hw, cout, device="cuda", dtype=dt)
        bias = torch.zeros(cout, device="cuda")
        timed(f"conv3x3 {cin}->{cout} @{hw}x{hw} x{n}",
              ops.gemm(x, w, out, nimg=n, h=hw, w_=hw, c0=cin, n=cout, ksize=3, bias=bias),
              2.0 * n * hw * hw * cout * 9 * cin)
    lins = [(n * 4096, 320, 2560, True), (n * 4096, 1280, 320, False), (n * 1024, 640, 5120, True),
            (n * 4096, 320, 960, False), (n * 256, 1280, 10240, True)]
    for m, k, nn, geglu in lins:
        x = rnd(m, k)
        w = rnd(nn, 1, k, scale=k ** -0.5)
        out = torch.empty(m, nn // 2 if geglu else nn, device="cuda", dtype=dt)
        bias = torch.zeros(nn, device="cuda")
        timed(f"linear {m}x{k}x{nn}{' geglu' if geglu else ''}",
              ops.gemm(x, w, out, nimg=1, h=1, w_=m, c0=k, n=nn, bias=bias,
                       act=b200sd.lib.ACT_GEGLU if geglu else b200sd.lib.ACT_NONE), 2.0 * m * k * nn)
    for tokens, heads, d in [(4096, 8, 40), (1024, 8, 80), (256, 8, 160)]:
        c = heads * d
        qkv = rnd(n, tokens, 3 * c)
        out = torch.empty(n, tokens, c, device="cuda", dtype=dt)
        q2 = qkv.view(-1, 3 * c)
        timed(f"self-attention {tokens} tok d={d} x{n}",
              ops.attention(q2, q2[:, c:], q2[:, 2 * c:], out, batch=n, heads=heads, nq=tokens, nk=tokens, d=d,
                            ldq=3 * c, ldk=3 * c, ldv=3 * c, ldo=c, scale=d ** -0.5), 4.0 * n * heads * tokens * tokens * d)
    kv = rnd(n, 77, 640)
    q = rnd(n, 4096, 320)
    out = torch.empty(n, 4096, 320, device="cuda", dtype=dt)
    kv2 = kv.view(-1, 640)
    timed("cross-attention 4096x77 d=40", ops.attention(q, kv2, kv2[:, 320:], out, batch=n, heads=8, nq=4096, nk=77,
                                                        d=40, ldq=320, ldk=640, ldv=640, ldo=320, scale=40 ** -0.5),
          4.0 * n * 8 * 4096 * 77 * 40)
    # bandwidth-bound kernels
    x = rnd(n, 4096, 320)
    o = torch.empty_like(x)
    g, b = torch.ones(320, device="cuda"), torch.zeros(320, device="cuda")
    st = ops.groupnorm_stats_buffer(n, 32)
    sc, ac = ops.groupnorm(x, o, st, g, b, nimg=n, hw=4096, c0=320, groups=32, eps=1e-5, silu=True)
    nbytes = x.numel() * 2
    for name, call, byt in (("groupnorm stats 320@64x64", sc, nbytes), ("groupnorm apply 320@64x64", ac, 2 * nbytes),
                            ("layernorm 65536x320", ops.layernorm(x, o, g, b, rows=n * 4096, c=320, eps=1e-5),
                             2 * nbytes)):
        call(s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        call(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{name:44s} {ms:8.3f} ms  {byt / ms / 1e6:8.1f} GB/s")


if __name__ == "__main__" and (len(sys.argv) < 2 or sys.argv[1] in ("step", "kernels")):
    {"step": mode_step, "kernels": mode_kernels}[sys.argv[1] if len(sys.argv) > 1 else "kernels"]()


def mode_lin():
    """Short-K linear (65536 x 320 -> 960) under different tile widths; used for ncu and quick A/B timing."""
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    m, k, nn = 65536, 320, 960
    x = (torch.randn(m, k, device="cuda")).to(dt)
    w = (torch.randn(nn, 1, k, device="cuda") * k ** -0.5).to(dt)
    out = torch.empty(m, nn, device="cuda", dtype=dt)
    bias = torch.zeros(nn, device="cuda")
    for bn in (0, 64, 128, 192, 256):
        call = ops.gemm(x, w, out, nimg=1, h=1, w_=m, c0=k, n=nn, bias=bias, bn_hint=bn)
        call(s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            call(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"linear {m}x{k}x{nn} bn={bn:3d}: {ms:.3f} ms  {2.0 * m * k * nn / ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "lin":
    mode_lin()


def mode_pair():
    """Single-CTA tiles vs CTA pairs (cta_group::2) on the heaviest UNet / VAE shapes."""
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    n = 16
    shapes = [("conv", 64, 320, 320, 3), ("conv", 32, 640, 640, 3), ("conv", 16, 1280, 1280, 3), ("conv", 8, 1280, 1280, 3),
              ("conv", 64, 640, 320, 3), ("lin", 65536, 320, 960, 1), ("lin", 65536, 320, 320, 1),
              ("lin", 65536, 1280, 320, 1), ("lin", 16384, 640, 1920, 1), ("lin", 4096, 1280, 3840, 1),
              ("lin", 16384, 2560, 640, 1), ("conv", 256, 128, 128, 3), ("conv", 128, 256, 256, 3)]
    for kind, a, cin, cout, ks in shapes:
        if kind == "conv":
            nn = 8 if a >= 128 else n
            x = (torch.randn(nn, a, a, cin, device="cuda")).to(dt)
            geo = dict(nimg=nn, h=a, w_=a)
            m = nn * a * a
        else:
            x = (torch.randn(a, cin, device="cuda")).to(dt)
            geo = dict(nimg=1, h=1, w_=a)
            m = a
        w = (torch.randn(cout, ks * ks, cin, device="cuda") * (ks * ks * cin) ** -0.5).to(dt)
        out = torch.empty(m, cout, device="cuda", dtype=dt)
        bias = torch.zeros(cout, device="cuda")
        res = []
        for hint in (0, 1000, 2000):
            call = ops.gemm(x, w, out, c0=cin, n=cout, ksize=ks, bias=bias, bn_hint=hint, **geo)
            call(s)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                call(s)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            res.append(f"{ms:.3f} ms {2.0 * m * cout * ks * ks * cin / ms / 1e9:6.0f} TF")
        print(f"{kind} M={m} K={ks * ks * cin} N={cout}: auto {res[0]} | single {res[1]} | pair {res[2]}")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "pair":
    mode_pair()


def mode_unet1():
    """Exactly one eager denoise step inside a cudaProfilerStart/Stop range (ncu --profile-from-start off)."""
    import b200sd
    pipe = b200sd.StableDiffusionPipeline.from_pretrained("sd15", scheduler="ddim")
    job = pipe.job(8, 64, 64, 7.5, True)
    sched = pipe.make_schedule(25)
    job.step_table[: sched.num_evals].copy_(sched.table)
    job.init_latents.normal_()
    job.reset()
    job.step_plan.run()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    job.step_plan.run()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("unet1 done:", len(job.step_plan), "launches")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "unet1":
    mode_unet1()


def mode_att():
    """Self-attention of the 64x64 level (4096 tokens, 8 heads x 40) alone: quick A/B timing and ncu target."""
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    n = 16
    for tokens, heads, d in [(4096, 8, 40), (1024, 8, 80)]:
        c = heads * d
        qkv = (torch.randn(n, tokens, 3 * c, device="cuda")).to(dt)
        out = torch.empty(n, tokens, c, device="cuda", dtype=dt)
        q2 = qkv.view(-1, 3 * c)
        call = ops.attention(q2, q2[:, c:], q2[:, 2 * c:], out, batch=n, heads=heads, nq=tokens, nk=tokens, d=d,
                             ldq=3 * c, ldk=3 * c, ldv=3 * c, ldo=c, scale=d ** -0.5)
        call(s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            call(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"self-attention {tokens} tok d={d} x{n}: {ms:.3f} ms  {4.0 * n * heads * tokens * tokens * d / ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "att":
    mode_att()


def mode_sweep():
    """Tile-plan sweep: every UNet linear / conv shape class under forced (single | pair) x BN; best vs the heuristic."""
    import b200sd
    ops = b200sd.ops
    L = b200sd.lib
    dt = L.act_dtype()
    s = ops.current_stream()
    # (M, K-per-tap, N, ksize, spatial side or 0, residual, geglu)
    shapes = [(65536, 320, 320, 1, 0, True, False), (65536, 320, 960, 1, 0, False, False),
              (65536, 1280, 320, 1, 0, True, False), (65536, 320, 2560, 1, 0, False, True),
              (16384, 640, 640, 1, 0, True, False), (16384, 640, 5120, 1, 0, False, True),
              (16384, 2560, 640, 1, 0, True, False), (4096, 1280, 1280, 1, 0, True, False),
              (4096, 5120, 1280, 1, 0, True, False), (4096, 1280, 10240, 1, 0, False, True),
              (1024, 1280, 1280, 3, 8, True, False), (4096, 1280, 1280, 3, 16, True, False),
              (16384, 640, 640, 3, 32, True, False), (65536, 320, 320, 3, 64, True, False)]
    for m, cin, nn, ks, side, has_res, geglu in shapes:
        if ks == 3:
            nimg = m // (side * side)
            x = torch.randn(nimg, side, side, cin, device="cuda").to(dt)
            geo = dict(nimg=nimg, h=side, w_=side)
        else:
            x = torch.randn(m, cin, device="cuda").to(dt)
            geo = dict(nimg=1, h=1, w_=m)
        w = (torch.randn(nn, ks * ks, cin, device="cuda") * (ks * ks * cin) ** -0.5).to(dt)
        nout = nn // 2 if geglu else nn
        out = torch.empty(m, nout, device="cuda", dtype=dt)
        res = torch.randn(m, nout, device="cuda").to(dt) if has_res else None
        bias = torch.zeros(nn, device="cuda")
        results = {}
        for hint in [0] + [md + bn for md in (1000, 2000) for bn in (64, 128, 192, 256)] + [320, 384, 448, 512]:
            try:
                call = ops.gemm(x, w, out, c0=cin, n=nn, ksize=ks, bias=bias, bn_hint=hint, residual=res,
                                act=L.ACT_GEGLU if geglu else L.ACT_NONE, **geo)
                call(s)
            except L.B200SDError:
                continue
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                call(s)
            e1.record()
            torch.cuda.synchronize()
            results[hint] = e0.elapsed_time(e1) / 5
        fl = 2.0 * m * nn * ks * ks * cin
        best = min((v, k) for k, v in results.items() if k)
        row = " ".join(f"{k}:{v * 1e3:.0f}" for k, v in results.items())
        print(f"M={m} K={ks * ks * cin} N={nn}{' res' if has_res else ''}{' geglu' if geglu else ''}: auto {results[0] * 1e3:.0f} us "
              f"({fl / results[0] / 1e9:.0f} TF) best {best[1]} {best[0] * 1e3:.0f} us ({fl / best[0] / 1e9:.0f} TF) | {row}")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "sweep":
    mode_sweep()


def mode_split():
    """Split-K A/B on the 8x8-level convolutions (M = 1024 rows at batch 8): split_k = 1, 2, 3, 4, 6."""
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    for cin, cout, nimg, hw in [(1280, 1280, 16, 8), (2560, 1280, 16, 8), (1280, 1280, 2, 8), (1280, 1280, 2, 16)]:
        x = torch.randn(nimg, hw, hw, cin, device="cuda").to(dt)
        w = (torch.randn(cout, 9, cin, device="cuda") * (9 * cin) ** -0.5).to(dt)
        out = torch.empty(nimg, hw, hw, cout, device="cuda", dtype=dt)
        fl = 2.0 * nimg * hw * hw * cout * 9 * cin
        for sk in (1, 2, 3, 4, 6, 8):
            call = ops.gemm(x, w, out, nimg=nimg, h=hw, w_=hw, c0=cin, n=cout, ksize=3, split_k=sk)
            if sk > 1:
                need = ops.gemm_workspace_bytes(call)
                ws = torch.zeros((need + 255) // 256 * 64, dtype=torch.int32, device="cuda")
                call.params.workspace, call.params.workspace_bytes = ws.data_ptr(), ws.numel() * 4
            for _ in range(3):
                call(s)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                call(s)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            print(f"conv3x3 {cin}->{cout} @{hw}x{hw} x{nimg} split_k={sk}: {ms * 1e3:7.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "split":
    mode_split()


def mode_norms():
    """GroupNorm statistics / apply and LayerNorm on the step's main shapes, cold (L2 flushed by a 512 MB write) and
    warm (second launch on the same input): quick timing and the ncu target for the bandwidth-bound kernels."""
    import b200sd
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    s = ops.current_stream()
    n = 16
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def run(name, call, byt):
        res = []
        for cold in (True, False):
            if cold:
                flush.fill_(1)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            call(s)
            e1.record()
            torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1))
        print(f"{name:40s} cold {res[0] * 1e3:7.1f} us {byt / res[0] / 1e6:7.0f} GB/s   warm {res[1] * 1e3:7.1f} us "
              f"{byt / res[1] / 1e6:7.0f} GB/s")

    for hw, c0, c1 in [(4096, 320, 0), (4096, 320, 320), (1024, 640, 0), (1024, 640, 640), (256, 1280, 0),
                       (256, 1280, 1280), (64, 1280, 1280)]:
        x0 = torch.randn(n, hw, c0, device="cuda").to(dt)
        x1 = torch.randn(n, hw, c1, device="cuda").to(dt) if c1 else None
        C = c0 + c1
        o = torch.empty(n, hw, C, device="cuda", dtype=dt)
        g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
        st = ops.groupnorm_stats_buffer(n, 32)
        sc, ac = ops.groupnorm(x0, o, st, g, b, nimg=n, hw=hw, c0=c0, groups=32, eps=1e-5, silu=True, x1=x1, c1=c1)
        run(f"groupnorm stats {c0}+{c1} @{hw} px", sc, 2.0 * n * hw * C)
        run(f"groupnorm apply {c0}+{c1} @{hw} px", ac, 4.0 * n * hw * C)
    for rows, c in [(n * 4096, 320), (n * 1024, 640), (n * 256, 1280)]:
        x = torch.randn(rows, c, device="cuda").to(dt)
        o = torch.empty_like(x)
        g, b = torch.ones(c, device="cuda"), torch.zeros(c, device="cuda")
        run(f"layernorm {rows}x{c}", ops.layernorm(x, o, g, b, rows=rows, c=c, eps=1e-5), 4.0 * rows * c)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "norms":
    mode_norms()
