"""ctypes binding of libb200sd.so (the C ABI in include/b200sd.h).

This is the only way the Python host reaches the GPU kernels.  There is deliberately no fallback: if the shared
library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "libb200sd.so"
# one library image per 16-bit element type (build.py VARIANTS); "fp16" is the default image
_LIB_PATHS = {"fp16": _LIB_PATH, "bf16": Path(__file__).resolve().parent / "libb200sd_bf16.so"}
_libs = {}
_current = "fp16"

ACT_NONE, ACT_SILU, ACT_QUICK_GELU, ACT_GEGLU, ACT_GELU = 0, 1, 2, 3, 4
ACT16, F32 = 0, 1
BF16 = ACT16  # historical alias
GN_NOALLOC_MAX_STATS = 16384  # B200SD_GN_NOALLOC_MAX_STATS (4-byte words)
GN_STATS_WORDS = 8  # B200SD_GN_STATS_WORDS: 4-byte words of statistics scratch per (image, group)
ST_TIMESTEP, ST_IN_SCALE, ST_KIND, ST_C0, ST_C1, ST_C2, ST_C3, ST_W0, ST_PNDM_MODE, ST_PUSH, ST_GUIDANCE, ST_WIDTH = (
    0, 1, 2, 3, 4, 5, 6, 7, 11, 12, 13, 16)

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class GemmParams(C.Structure):
    _fields_ = [("a0", vp), ("a1", vp), ("w", vp), ("out", vp), ("bias", vp), ("rowbias", vp), ("residual", vp),
                ("nimg", i32), ("h", i32), ("w_", i32), ("c0", i32), ("c1", i32), ("ld0", i32), ("ld1", i32),
                ("n", i32), ("ksize", i32), ("ldc", i32), ("ldr", i32), ("ld_rowbias", i32), ("act", i32),
                ("out_dtype", i32), ("bn_hint", i32), ("alpha", f32), ("gn_stats", vp), ("gn_groups", i32),
                ("gn_rows_per_img", i32), ("kh", i32), ("kw", i32), ("off_h", i32), ("off_w", i32), ("stride", i32),
                ("out_pix_stride", i32), ("out_row_stride", i32), ("out_img_stride", i32), ("split_k", i32),
                ("workspace", vp), ("workspace_bytes", i64), ("ln_out_stats", vp), ("ln_in_stats", vp),
                ("ln_colsum", vp), ("ln_eps", f32), ("weights_static", i32)]


class AttnParams(C.Structure):
    _fields_ = [("q", vp), ("k", vp), ("v", vp), ("out", vp),
                ("batch", i32), ("heads", i32), ("nq", i32), ("nk", i32), ("d", i32),
                ("ldq", i32), ("ldk", i32), ("ldv", i32), ("ldo", i32), ("causal", i32), ("scale", f32)]


class GroupNormParams(C.Structure):
    _fields_ = [("x0", vp), ("x1", vp), ("stats", vp), ("gamma", vp), ("beta", vp), ("out", vp),
                ("nimg", i32), ("hw", i32), ("c0", i32), ("c1", i32), ("groups", i32), ("silu", i32), ("eps", f32)]


class LayerNormParams(C.Structure):
    _fields_ = [("x", vp), ("gamma", vp), ("beta", vp), ("out", vp), ("rows", i32), ("c", i32), ("eps", f32)]


class PrepParams(C.Structure):
    _fields_ = [("latents", vp), ("out", vp), ("step_table", vp), ("step_counter", vp),
                ("b", i32), ("hw", i32), ("cfg", i32)]


class TembParams(C.Structure):
    _fields_ = [("out", vp), ("step_table", vp), ("step_counter", vp), ("rows", i32), ("dim", i32)]


class SchedParams(C.Structure):
    _fields_ = [("eps", vp), ("latents", vp), ("history", vp), ("cur_sample", vp), ("step_table", vp),
                ("step_counter", vp), ("b", i32), ("hw", i32), ("ld_eps", i32), ("cfg", i32),
                ("guidance_scale", f32)]


class EmbedParams(C.Structure):
    _fields_ = [("ids", vp), ("tok", vp), ("pos", vp), ("out", vp),
                ("rows", i32), ("seq", i32), ("c", i32), ("vocab", i32)]


class DenormParams(C.Structure):
    _fields_ = [("x", vp), ("out", vp), ("npix", i64), ("ldx", i32)]


class LatentCastParams(C.Structure):
    _fields_ = [("latents", vp), ("out", vp), ("b", i32), ("hw", i32), ("scale", f32)]


class SoftmaxParams(C.Structure):
    _fields_ = [("x", vp), ("out", vp), ("rows", i32), ("cols", i32)]


class MemsetParams(C.Structure):
    _fields_ = [("ptr", vp), ("bytes", i64)]


# declaration order of include/b200sd.h (b200sd_struct_sizes reports sizeof() in this order)
STRUCTS = [GemmParams, AttnParams, GroupNormParams, LayerNormParams, PrepParams, TembParams, SchedParams,
           EmbedParams, DenormParams, LatentCastParams, SoftmaxParams, MemsetParams]

# entry point -> params struct (None = special signature); must list every symbol include/b200sd.h declares
ENTRY_POINTS = {
    "b200sd_version": None,
    "b200sd_act_dtype": None,
    "b200sd_last_error": None,
    "b200sd_device_info": None,
    "b200sd_struct_sizes": None,
    "b200sd_gemm": GemmParams,
    "b200sd_gemm_workspace_bytes": None,
    "b200sd_attention": AttnParams,
    "b200sd_groupnorm_stats": GroupNormParams,
    "b200sd_groupnorm_apply": GroupNormParams,
    "b200sd_groupnorm_noalloc": GroupNormParams,
    "b200sd_groupnorm_workspace_bytes": None,
    "b200sd_layernorm": LayerNormParams,
    "b200sd_prep_unet_input": PrepParams,
    "b200sd_timestep_embedding": TembParams,
    "b200sd_cfg_scheduler_step": SchedParams,
    "b200sd_clip_embed": EmbedParams,
    "b200sd_denorm_u8": DenormParams,
    "b200sd_latent_to_act": LatentCastParams,
    "b200sd_softmax_rows": SoftmaxParams,
    "b200sd_memset_zero": MemsetParams,
}


class B200SDError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


class use:
    """`with lib.use("bf16"):` - calls prepared inside the block bind to that image of the library (a plan is built
    once, so this is a per-plan choice of the 16-bit type: the SDXL VAE decoder runs bf16 beside an fp16 UNet)."""

    def __init__(self, variant: str):
        if variant not in _LIB_PATHS:
            raise ValueError(f"unknown library variant {variant!r} ({sorted(_LIB_PATHS)})")
        self.variant = variant

    def __enter__(self):
        global _current
        self.prev, _current = _current, self.variant
        return self

    def __exit__(self, *exc):
        global _current
        _current = self.prev
        return False


def load(variant: str | None = None):
    """Load the library image (once per variant; default: the current one).  Raises if it has not been built — there
    is no other execution path."""
    variant = variant or _current
    if variant in _libs:
        return _libs[variant]
    path = _LIB_PATHS[variant]
    if variant == "fp16" and os.environ.get("B200SD_LIB"):
        path = Path(os.environ["B200SD_LIB"]).resolve()  # A/B measurements against another build of the library
    if not path.exists():
        raise B200SDError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). b200sd has no CPU or library fallback.")
    lib = C.CDLL(str(path))
    lib.b200sd_version.restype = C.c_int
    lib.b200sd_act_dtype.restype = C.c_int
    lib.b200sd_last_error.restype = C.c_char_p
    lib.b200sd_device_info.argtypes = [C.POINTER(i32)]
    lib.b200sd_struct_sizes.argtypes = [C.POINTER(i32), i32]
    for name, st in ENTRY_POINTS.items():
        if st is None:
            continue
        fn = getattr(lib, name)
        fn.argtypes = [C.POINTER(st), vp]
        fn.restype = C.c_int
    lib.b200sd_gemm_workspace_bytes.argtypes = [C.POINTER(GemmParams)]
    lib.b200sd_gemm_workspace_bytes.restype = C.c_size_t
    lib.b200sd_groupnorm_workspace_bytes.argtypes = [C.POINTER(GroupNormParams)]
    lib.b200sd_groupnorm_workspace_bytes.restype = C.c_size_t
    _libs[variant] = lib
    return lib


def act_dtype():
    """torch dtype of the library's 16-bit activation / weight element type (fp16 by default, bf16 build optional)."""
    import torch
    return torch.float16 if load().b200sd_act_dtype() == 1 else torch.bfloat16


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = load().b200sd_last_error().decode(errors="replace")
        raise B200SDError(f"{what or 'b200sd call'} failed ({code}): {msg}")


def struct_sizes():
    lib = load()
    buf = (i32 * 32)()
    n = lib.b200sd_struct_sizes(buf, 32)
    return [buf[i] for i in range(n)]


def device_info():
    lib = load()
    buf = (i32 * 4)()
    check(lib.b200sd_device_info(buf), "device_info")
    return {"sm_count": buf[0], "cc": (buf[1], buf[2]), "max_smem_optin": buf[3]}


class Call:
    """One prepared kernel invocation: (entry point, params struct).  Calling it enqueues the kernel on `stream`."""

    __slots__ = ("fn", "params", "name", "keep")

    def __init__(self, name: str, params, keep=()):
        self.fn = getattr(load(), name)
        self.params = params
        self.name = name
        self.keep = keep  # tensors referenced by raw pointer — keep them alive

    def __call__(self, stream: int) -> None:
        rc = self.fn(C.byref(self.params), stream)
        if rc != 0:
            check(rc, self.name)
