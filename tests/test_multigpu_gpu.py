"""Sharding invariance on hardware: the drop-in CLI in local mode, started as one process
per GPU under torchrun (NCCL), writes byte-identical images to a single-process run of the same command - image i is
always seeded with seed + i - 1 and nothing in the CUDA path depends on the batch or rank an image is computed in.
Needs two GPUs (`gpurun --gpus 2`); skipped on a one-GPU box."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "scripts", "batch_generate.py")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _load(out_dir, count):
    from PIL import Image
    return [np.asarray(Image.open(os.path.join(out_dir, f"img_{i:02d}.png"))) for i in range(1, count + 1)]


@pytest.mark.parametrize("model,count,size,batch", [("tiny", 7, 64, 2), ("sd15", 12, 512, 8)])
def test_cli_local_is_byte_identical_across_world_sizes(cuda_lib, tmp_path, model, count, size, batch):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    args = ["a photo of a panda riding a motorbike", str(count), "img"]
    flags = ["--local", "--model", model, "--steps", "3", "--width", str(size), "--height", str(size), "--batch-size",
             str(batch), "--scheduler", "ddim", "--seed", "77"]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    one, two = str(tmp_path / "w1"), str(tmp_path / "w2")
    r = subprocess.run([sys.executable, CLI, *args, one, *flags], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "--", CLI, *args, two, *flags],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("[*] Generating") == count and "All done." in r.stdout
    a, b = _load(one, count), _load(two, count)
    assert a[0].shape == (size, size, 3)
    for i, (x, y) in enumerate(zip(a, b)):
        assert np.array_equal(x, y), f"image {i + 1} differs between 1 and 2 GPUs"
    assert not np.array_equal(a[0], a[1])  # different seeds give different images


def test_second_device_in_one_process(cuda_lib, b200sd):
    """Per-device kernel attributes and caches: the first GEMM / attention calls on cuda:1 in a process that has already
    used cuda:0 (cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device property: a process-wide 'already set'
    flag would launch with the 48 KB default there and fail)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.nn.functional as F
    ops = b200sd.ops
    dt = b200sd.lib.act_dtype()
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        with torch.cuda.device(dev):
            g = torch.Generator(device=dev).manual_seed(3)
            x = torch.randn(2, 16, 16, 64, device=dev, generator=g).to(dt)
            w = (torch.randn(128, 64, 3, 3, device=dev, generator=g) * 0.04).to(dt)
            out = torch.empty(2, 16, 16, 128, device=dev, dtype=dt)
            ops.gemm(x, ops.pack_conv_weight(w.float().cpu()).to(dev), out, nimg=2, h=16, w_=16, c0=64, n=128,
                     ksize=3)(torch.cuda.current_stream().cuda_stream)
            qkv = torch.randn(2, 300, 3 * 128, device=dev, generator=g).to(dt)
            att = torch.empty(2, 300, 128, device=dev, dtype=dt)
            q2 = qkv.view(-1, 384)
            for nk, kk, vv in ((300, q2[:, 128:], q2[:, 256:]), (77, q2[:, 128:], q2[:, 256:])):
                ops.attention(q2, kk, vv, att, batch=2, heads=2, nq=300, nk=nk, d=64, ldq=384, ldk=384, ldv=384,
                              ldo=128, scale=0.125)(torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), padding=1).permute(0, 2, 3, 1)
            assert (out.float() - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()
            assert torch.isfinite(att.float()).all()
            outs.append(out.cpu())
    assert torch.equal(outs[0], outs[1])
