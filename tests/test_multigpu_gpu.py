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
