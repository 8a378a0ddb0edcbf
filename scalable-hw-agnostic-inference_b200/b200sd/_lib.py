"""ctypes binding of libb200sd.so (include/b200sd.h).  There is no CPU fallback: if the library is missing
the import of any compute op fails loudly."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libb200sd.so"
ABI_VERSION = 8

BF16, F32, F16 = 1, 2, 3

c_i32, c_i64, c_f32, c_vp = C.c_int32, C.c_int64, C.c_float, C.c_void_p


class GemmParams(C.Structure):
    _fields_ = [("a", c_vp), ("a_dtype", c_i32), ("n_img", c_i32), ("img_h", c_i32), ("img_w", c_i32),
                ("cin", c_i32), ("taps", c_i32), ("w", c_vp), ("w_dtype", c_i32), ("n_out", c_i32), ("bias", c_vp),
                ("bias_along_m", c_i32), ("row_bias", c_vp), ("row_bias_stride", c_i64), ("residual", c_vp),
                ("residual_dtype", c_i32), ("ld_res", c_i64), ("out", c_vp), ("out_dtype", c_i32), ("ld_out", c_i64),
                ("geglu", c_i32), ("block_n", c_i32), ("workspace", c_vp), ("workspace_bytes", c_i64),
                ("split_k", c_i32), ("act", c_i32), ("tile_sums", c_vp), ("rows_per_image", c_i32),
                ("conv_stride", c_i32), ("conv_pad_mode", c_i32)]


class AttentionParams(C.Structure):
    _fields_ = [("q", c_vp), ("k", c_vp), ("v", c_vp), ("ld_q", c_i64), ("ld_k", c_i64), ("ld_v", c_i64),
                ("out", c_vp), ("ld_out", c_i64), ("dtype", c_i32), ("batch", c_i32), ("heads", c_i32),
                ("head_dim", c_i32), ("sq", c_i32), ("skv", c_i32), ("scale", c_f32), ("causal", c_i32)]


class EmbedParams(C.Structure):
    _fields_ = [("ids", c_vp), ("token_table", c_vp), ("position_table", c_vp), ("out", c_vp), ("rows", c_i64),
                ("seq_len", c_i32), ("dim", c_i32), ("vocab", c_i32)]


class GroupNormParams(C.Structure):
    _fields_ = [("x1", c_vp), ("c1", c_i32), ("x2", c_vp), ("c2", c_i32), ("in_dtype", c_i32), ("n_img", c_i32),
                ("hw", c_i32), ("groups", c_i32), ("eps", c_f32), ("gamma", c_vp), ("beta", c_vp), ("silu", c_i32),
                ("y", c_vp), ("raw", c_vp), ("out_dtype", c_i32), ("workspace", c_vp), ("workspace_bytes", c_i64),
                ("sums", c_vp), ("stat_hw", c_i64), ("sums_out", c_vp)]


class GnTileSumsParams(C.Structure):
    _fields_ = [("tile_sums1", c_vp), ("parts1", c_i32), ("c1", c_i32), ("tile_sums2", c_vp), ("parts2", c_i32),
                ("c2", c_i32), ("n_img", c_i32), ("groups", c_i32), ("sums", c_vp)]


class LayerNormParams(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("rows", c_i64), ("cols", c_i32), ("eps", c_f32), ("gamma", c_vp),
                ("beta", c_vp), ("y", c_vp), ("out_dtype", c_i32)]


class Im2colParams(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("n_img", c_i32), ("n_src", c_i32), ("in_h", c_i32),
                ("in_w", c_i32), ("cin", c_i32), ("stride", c_i32), ("out", c_vp), ("out_dtype", c_i32),
                ("k_pad", c_i32), ("pad_lo", c_i32)]


class UpsampleParams(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("n_img", c_i32), ("h", c_i32), ("w", c_i32), ("c", c_i32),
                ("out", c_vp), ("out_dtype", c_i32)]


class CfgSchedParams(C.Structure):
    _fields_ = [("eps2", c_vp), ("eps_dtype", c_i32), ("x", c_vp), ("x0_prev", c_vp), ("coef", c_vp),
                ("step_index", c_vp), ("n", c_i64)]


class SmallLinearParams(C.Structure):
    _fields_ = [("x", c_vp), ("rows", c_i32), ("k", c_i32), ("n", c_i32), ("sinusoid_dim", c_i32),
                ("silu_in", c_i32), ("silu_out", c_i32), ("w", c_vp), ("w_dtype", c_i32), ("b", c_vp), ("y", c_vp)]


class SoftmaxParams(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("ld_in", c_i64), ("y", c_vp), ("out_dtype", c_i32),
                ("ld_out", c_i64), ("rows", c_i64), ("cols", c_i32), ("scale", c_f32)]


class DenormU8Params(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("out", c_vp), ("n", c_i64)]


class CastParams(C.Structure):
    _fields_ = [("x", c_vp), ("in_dtype", c_i32), ("y", c_vp), ("out_dtype", c_i32), ("n", c_i64)]


# every symbol include/b200sd.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "b200sd_abi_version": (c_i32, []),
    "b200sd_last_error": (C.c_char_p, []),
    "b200sd_check_device": (c_i32, [c_i32]),
    "b200sd_launch_count": (c_i64, []),
    "b200sd_gemm": (c_i32, [C.POINTER(GemmParams), c_vp]),
    "b200sd_gemm_pick_block_n": (c_i32, [c_i32, c_i32]),
    "b200sd_gemm_workspace_bytes": (c_i64, [C.POINTER(GemmParams)]),
    "b200sd_gemm_tile_sum_parts": (c_i32, [C.POINTER(GemmParams)]),
    "b200sd_gn_tile_sums": (c_i32, [C.POINTER(GnTileSumsParams), c_vp]),
    "b200sd_attention": (c_i32, [C.POINTER(AttentionParams), c_vp]),
    "b200sd_attention512": (c_i32, [C.POINTER(AttentionParams), c_vp]),
    "b200sd_groupnorm_workspace_bytes": (c_i64, [C.POINTER(GroupNormParams)]),
    "b200sd_groupnorm": (c_i32, [C.POINTER(GroupNormParams), c_vp]),
    "b200sd_groupnorm_static_scratch": (c_i32, [C.POINTER(GroupNormParams), c_vp]),
    "b200sd_layernorm": (c_i32, [C.POINTER(LayerNormParams), c_vp]),
    "b200sd_im2col3x3": (c_i32, [C.POINTER(Im2colParams), c_vp]),
    "b200sd_upsample2x": (c_i32, [C.POINTER(UpsampleParams), c_vp]),
    "b200sd_cfg_sched_step": (c_i32, [C.POINTER(CfgSchedParams), c_vp]),
    "b200sd_small_linear": (c_i32, [C.POINTER(SmallLinearParams), c_vp]),
    "b200sd_softmax_rows": (c_i32, [C.POINTER(SoftmaxParams), c_vp]),
    "b200sd_denorm_u8": (c_i32, [C.POINTER(DenormU8Params), c_vp]),
    "b200sd_cast": (c_i32, [C.POINTER(CastParams), c_vp]),
    "b200sd_embed_tokens": (c_i32, [C.POINTER(EmbedParams), c_vp]),
    "b200sd_umma_selftest": (c_i32, [c_vp, c_i32, c_vp, c_i32, c_vp, c_i32, c_i32, c_i32, c_vp]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen libb200sd.so and bind every declared symbol; raises if the library or a symbol is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python scalable-hw-agnostic-inference_b200/build.py` "
            "(or __graft_entry__.build()).  b200sd has no CPU / PyTorch fallback.")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype, fn.argtypes = res, args
    if lib.b200sd_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libb200sd ABI version {lib.b200sd_abi_version()} != {ABI_VERSION} (stale build?)")
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = load().b200sd_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"b200sd {what} failed ({status}): {msg}")
