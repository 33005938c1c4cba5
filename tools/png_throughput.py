#!/usr/bin/env python3
"""Host side of SURVEY §8(f) row 4: how many 512x512 RGB PNGs per second the encoder threads deliver (PIL / zlib releases
the GIL), next to the GPU's images/s.  usage: png_throughput.py [threads=4] [images=64] [size=512]"""
import io
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
from PIL import Image

threads = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
size = int(sys.argv[3]) if len(sys.argv) > 3 else 512
rng = np.random.default_rng(0)
# smooth synthetic "photo" (pure noise is the worst case for zlib, a flat image the best): low-frequency field + noise
base = rng.normal(size=(size // 8, size // 8, 3))
img = np.kron(base, np.ones((8, 8, 1))) * 40 + 128 + rng.normal(size=(size, size, 3)) * 6
imgs = [np.clip(img + i, 0, 255).astype(np.uint8) for i in range(n)]


def enc(a):
    buf = io.BytesIO()
    Image.fromarray(a).save(buf, format="PNG")
    return len(buf.getvalue())


for t in (1, threads):
    t0 = time.perf_counter()
    with ThreadPoolExecutor(t) as ex:
        sizes = list(ex.map(enc, imgs))
    dt = time.perf_counter() - t0
    print(f"{t} thread(s): {n / dt:6.1f} PNG/s ({size}x{size} RGB, mean {sum(sizes) / n / 1e3:.0f} kB)")
