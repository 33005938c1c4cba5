"""In-tree build of libb200sd.so (hand-written sm_100a CUDA + the C ABI declared in include/b200sd.h).

nvcc cross-compiles without a GPU; the resulting .so is git-ignored but travels with the repo snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = PKG_DIR / "libb200sd.so"
BUILD_DIR = PKG_DIR / "build"
# The 16-bit element type is a compile-time property of a library image.  Two images are built from the same sources:
# fp16 (default: the reference's own CUDA dtype) and bfloat16 (fp32 exponent range: what the SDXL VAE decoder needs,
# upstream runs it in fp32 because real weights overflow fp16).  A plan picks its image (lib.use(variant)).
VARIANTS = {"fp16": ([], LIB_PATH), "bf16": (["-DB200SD_ACT_BF16"], PKG_DIR / "libb200sd_bf16.so")}

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]
# extra nvcc flags for experiments (e.g. B200SD_NVCC_EXTRA="-DB200SD_ACT_BF16"); part of the stamp digest
NVCC_FLAGS += os.environ.get("B200SD_NVCC_EXTRA", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _digest(paths, extra=()) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS + list(extra)).encode())
    return h.hexdigest()


def sources():
    return sorted(CSRC.glob("*.cu"))


def build(force: bool = False, verbose: bool = False, variant: str = "fp16") -> Path:
    extra, lib_path = VARIANTS[variant]
    srcs = sources()
    deps = srcs + sorted(CSRC.glob("*.cuh")) + [PKG_DIR.parent / "include" / "b200sd.h"]
    obj_dir = BUILD_DIR if variant == "fp16" else BUILD_DIR / variant
    stamp = obj_dir / "stamp.txt"
    digest = _digest(deps, extra)
    if not force and lib_path.exists() and stamp.exists() and stamp.read_text() == digest:
        return lib_path
    obj_dir.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = obj_dir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    # static cudart (nvcc default): the .so is self-contained and shares the primary context with torch
    cmd = [nvcc, "-shared", "-o", str(lib_path), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(digest)
    return lib_path


def build_all(force: bool = False, verbose: bool = False):
    return [build(force, verbose, v) for v in VARIANTS]


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
