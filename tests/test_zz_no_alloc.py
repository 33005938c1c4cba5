"""b200sd_groupnorm_noalloc — the one entry point for hosts that cannot give the library any scratch memory.

Some users of this library only ever hold the tensors they were handed: a normalisation hook that runs inside another
runtime's pre-planned memory arena, or a server that must stay inside a fixed device-memory budget, cannot call
cudaMalloc when a request arrives (it may fail, and it stalls the whole GPU while it runs).  For them GroupNorm is
offered as a single call whose statistics scratch is a fixed 64 KiB array inside libb200sd.so.

This test calls the new entry point through the C ABI (ctypes, raw pointers) and compares it with the existing
two-call path (b200sd_groupnorm_stats + b200sd_groupnorm_apply on caller-provided, caller-zeroed scratch).  Both run
the same kernels, so the outputs must agree within the 1e-2 * max|ref| bound used by the other GroupNorm tests (in
practice they are bit-identical: the statistics are accumulated as integers).  It is done twice in a row with different data and a different
shape, because the built-in scratch is shared between calls and still holds the previous call's sums.
The file name sorts last so that a `-x` run has already executed every other test before this one.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _params(lib, x0, x1, stats, gamma, beta, out, nimg, hw, c0, c1, groups, silu, eps):
    p = lib.GroupNormParams()
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    p.x0, p.x1, p.stats, p.gamma, p.beta, p.out = ptr(x0), ptr(x1), ptr(stats), ptr(gamma), ptr(beta), ptr(out)
    p.nimg, p.hw, p.c0, p.c1, p.groups, p.silu, p.eps = nimg, hw, c0, c1, groups, int(silu), eps
    return p


def _both_paths(b200sd, seed, nimg, hw, c0, c1, groups, silu):
    lib = b200sd.lib
    so = lib.load()
    dt = lib.act_dtype()
    g = torch.Generator(device="cuda").manual_seed(seed)
    C_ = c0 + c1
    x0 = (torch.randn(nimg, hw, c0, device="cuda", generator=g) * 2 + 0.5).to(dt)
    x1 = (torch.randn(nimg, hw, c1, device="cuda", generator=g) - 1.0).to(dt) if c1 else None
    gamma = torch.randn(C_, device="cuda", generator=g)
    beta = torch.randn(C_, device="cuda", generator=g)
    stream = torch.cuda.current_stream().cuda_stream

    # existing path: caller owns (and zeroes) the statistics scratch
    stats = torch.zeros(nimg, groups, lib.GN_STATS_WORDS, dtype=torch.int32, device="cuda")
    ref = torch.empty(nimg, hw, C_, device="cuda", dtype=dt)
    p = _params(lib, x0, x1, stats, gamma, beta, ref, nimg, hw, c0, c1, groups, silu, 1e-5)
    assert so.b200sd_groupnorm_stats(C.byref(p), stream) == 0, so.b200sd_last_error()
    assert so.b200sd_groupnorm_apply(C.byref(p), stream) == 0, so.b200sd_last_error()

    # new path: no scratch pointer at all
    out = torch.full((nimg, hw, C_), float("nan"), device="cuda", dtype=dt)
    q = _params(lib, x0, x1, None, gamma, beta, out, nimg, hw, c0, c1, groups, silu, 1e-5)
    rc = so.b200sd_groupnorm_noalloc(C.byref(q), stream)
    assert rc == 0, so.b200sd_last_error()
    torch.cuda.synchronize()
    return out.float(), ref.float()


def test_noalloc_matches_two_call_path_twice_in_a_row(cuda_lib, b200sd):
    for seed, shape in ((11, dict(nimg=3, hw=256, c0=320, c1=0, groups=32, silu=True)),
                        (12, dict(nimg=2, hw=1024, c0=640, c1=320, groups=32, silu=False))):
        out, ref = _both_paths(b200sd, seed, **shape)
        assert not torch.isnan(out).any()
        err = (out - ref).abs().max().item()
        assert err <= 1e-2 * ref.abs().max().item(), (shape, err)


def test_noalloc_rejects_oversize_and_touches_no_allocator(cuda_lib, b200sd):
    lib = b200sd.lib
    so = lib.load()
    dt = lib.act_dtype()
    groups = 32
    nimg = lib.GN_NOALLOC_MAX_STATS // (lib.GN_STATS_WORDS * groups) + 1   # one image too many for the built-in scratch
    x = torch.zeros(nimg, 8, 64, device="cuda", dtype=dt)
    gamma = torch.ones(64, device="cuda")
    beta = torch.zeros(64, device="cuda")
    out = torch.empty_like(x)
    stream = torch.cuda.current_stream().cuda_stream
    p = _params(lib, x, None, None, gamma, beta, out, nimg, 8, 64, 0, groups, False, 1e-5)
    rc = so.b200sd_groupnorm_noalloc(C.byref(p), stream)
    assert rc == -6, rc                                     # B200SD_ERR_TOO_LARGE
    msg = so.b200sd_last_error().decode()
    assert "exceed the built-in scratch" in msg and str(lib.GN_NOALLOC_MAX_STATS) in msg

    # the largest admissible problem still runs, and device memory in use does not grow across calls
    nimg = lib.GN_NOALLOC_MAX_STATS // (lib.GN_STATS_WORDS * groups)
    x = torch.randn(nimg, 8, 64, device="cuda").to(dt)
    out = torch.empty_like(x)
    p = _params(lib, x, None, None, gamma, beta, out, nimg, 8, 64, 0, groups, False, 1e-5)
    assert so.b200sd_groupnorm_noalloc(C.byref(p), stream) == 0, so.b200sd_last_error()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert so.b200sd_groupnorm_noalloc(C.byref(p), stream) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    ref = torch.nn.functional.group_norm(x.float().permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    assert (out.float() - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()
