"""torch.library ops `b200sd::*` over the C ABI of libb200sd.so (include/b200sd.h).

Every op is an out-variant: the caller (torch) owns all device memory, the library only launches kernels
on torch's current CUDA stream.  CUDA only -- there is deliberately no CPU implementation.
Output element types are taken from the output tensors (bf16 / fp16 / fp32).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L

_DT = {torch.bfloat16: L.BF16, torch.float32: L.F32, torch.float16: L.F16}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"b200sd: unsupported dtype {t.dtype}") from None


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("b200sd ops are CUDA-only (no CPU fallback); got a CPU tensor")
    _check_current(t)
    return t.data_ptr()


def _stream() -> int:
    """torch's current stream on the CURRENT device.  Every launch in libb200sd goes to the current device, so callers
    that own a replica on another GPU enter `torch.cuda.device(replica_device)` first (serving.Engine does)."""
    return torch.cuda.current_stream().cuda_stream


def _check_current(t: torch.Tensor) -> None:
    if t.is_cuda and t.device.index != torch.cuda.current_device():
        raise RuntimeError(f"b200sd: tensor lives on {t.device} but the current CUDA device is "
                           f"cuda:{torch.cuda.current_device()}; wrap the call in torch.cuda.device({t.device.index})")


def launch_count() -> int:
    return int(L.load().b200sd_launch_count())


def check_device(dev: int = 0) -> None:
    L.check(L.load().b200sd_check_device(dev), "check_device")


def pick_block_n(n_out: int, geglu: bool) -> int:
    return int(L.load().b200sd_gemm_pick_block_n(n_out, int(geglu)))


# ------------------------------------------------------------------------------------------------ profiling
# bench.py brackets every launch with CUDA events on the launching stream to get per-kernel-class time and the
# algorithmic work (FLOPs for tensor-core kernels, bytes for HBM-bound ones) each launch did.
_prof = None


def profile_begin() -> None:
    global _prof
    _prof = []


def profile_end(by_shape: bool = False):
    """-> {kind: {"launches", "ms", "work"}}; work = algorithmic FLOPs (gemm, attention) or bytes (the rest).
    by_shape=True keys the result by "kind shape" instead."""
    global _prof
    recs, _prof = _prof or [], None
    torch.cuda.synchronize()
    out = {}
    for kind, work, s, e, tag in recs:
        d = out.setdefault(f"{kind} {tag}" if by_shape else kind, {"launches": 0, "ms": 0.0, "work": 0.0})
        d["launches"] += 1
        d["ms"] += s.elapsed_time(e)
        d["work"] += work
    return out


_taglog = None  # when a list: (kind, tag, work) of every launch in order (tools/ncu_by_shape.py joins it with ncu)


def _run(kind: str, work: float, fn, p, tag: str = "") -> None:
    if _taglog is not None:
        _taglog.append((kind, tag, work))
    if _prof is None:
        L.check(fn(C.byref(p), _stream()), kind)
        return
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    L.check(fn(C.byref(p), _stream()), kind)
    e.record()
    _prof.append((kind, work, s, e, tag))


def _nbytes(*ts) -> float:
    return float(sum(t.numel() * t.element_size() for t in ts if t is not None))


# ------------------------------------------------------------------------------------------------ impls
def _gemm_params(a, n_img, img_h, img_w, cin, taps, w, bias, bias_along_m, row_bias, row_bias_stride, residual, out,
                 geglu, block_n, workspace=None, split_k=0, act=0, tile_sums=None, rows_per_image=0, conv_stride=1,
                 conv_pad_mode=0):
    p = L.GemmParams()
    p.a, p.a_dtype, p.n_img, p.img_h, p.img_w, p.cin, p.taps = _ptr(a), _dt(a), n_img, img_h, img_w, cin, taps
    p.w, p.w_dtype, p.n_out = _ptr(w), _dt(w), w.shape[0]
    assert w.shape[1] == taps * cin, (tuple(w.shape), taps, cin)
    p.bias, p.bias_along_m = _ptr(bias), int(bias_along_m)
    p.row_bias, p.row_bias_stride = _ptr(row_bias), row_bias_stride
    if residual is not None:
        p.residual, p.residual_dtype, p.ld_res = _ptr(residual), _dt(residual), residual.stride(-2)
    p.out, p.out_dtype, p.ld_out = _ptr(out), _dt(out), out.stride(-2)
    p.geglu, p.block_n, p.split_k, p.act = int(geglu), block_n, split_k, act
    p.conv_stride, p.conv_pad_mode = conv_stride, conv_pad_mode
    if workspace is not None:
        p.workspace, p.workspace_bytes = _ptr(workspace), workspace.numel() * workspace.element_size()
    if tile_sums is not None:
        assert tile_sums.dtype == torch.float32 and tile_sums.is_contiguous()
        p.tile_sums, p.rows_per_image = _ptr(tile_sums), rows_per_image
    return p


def gemm_tile_sum_parts(*args, rows_per_image=0, conv_stride=1, conv_pad_mode=0) -> int:
    """slices per image of the column-sum table this problem's epilogue can emit for the next GroupNorm (0 = it cannot);
    same arguments as `gemm`"""
    p = _gemm_params(*args)
    p.rows_per_image, p.conv_stride, p.conv_pad_mode = rows_per_image, conv_stride, conv_pad_mode
    return int(L.load().b200sd_gemm_tile_sum_parts(C.byref(p)))


def _gemm(a, n_img, img_h, img_w, cin, taps, w, bias, bias_along_m, row_bias, row_bias_stride, residual, out,
          geglu, block_n, workspace=None, split_k=0, act=0, tile_sums=None, rows_per_image=0, conv_stride=1,
          conv_pad_mode=0):
    p = _gemm_params(a, n_img, img_h, img_w, cin, taps, w, bias, bias_along_m, row_bias, row_bias_stride, residual, out,
                     geglu, block_n, workspace, split_k, act, tile_sums, rows_per_image, conv_stride, conv_pad_mode)
    m = out.numel() // out.shape[-1] if conv_stride == 2 else n_img * img_h * img_w
    _run("gemm", 2.0 * m * w.shape[0] * taps * cin, L.load().b200sd_gemm, p,
         f"M{m} N{w.shape[0]} K{taps * cin} taps{taps} img{n_img}x{img_h}x{img_w}{' geglu' if geglu else ''}")


def _attention(q, k, v, out, batch, heads, sq, skv, scale, causal=False):
    assert q.dtype == k.dtype == v.dtype == out.dtype
    p = L.AttentionParams()
    p.q, p.k, p.v, p.out = _ptr(q), _ptr(k), _ptr(v), _ptr(out)
    p.ld_q, p.ld_k, p.ld_v, p.ld_out = q.stride(-2), k.stride(-2), v.stride(-2), out.stride(-2)
    p.dtype, p.batch, p.heads, p.head_dim, p.sq, p.skv, p.scale = _dt(q), batch, heads, 64, sq, skv, scale
    p.causal = int(causal)
    _run("attention", 4.0 * batch * heads * sq * skv * 64, L.load().b200sd_attention, p,
         f"b{batch} h{heads} sq{sq} skv{skv}")


def _attention512(q, k, v, out, batch, sq, skv, scale):
    """one head of dimension 512 (the VAE mid block); q / k / v / out [batch * s, 512] views, any leading dimension"""
    assert q.dtype == k.dtype == v.dtype == out.dtype and q.shape[-1] == 512
    p = L.AttentionParams()
    p.q, p.k, p.v, p.out = _ptr(q), _ptr(k), _ptr(v), _ptr(out)
    p.ld_q, p.ld_k, p.ld_v, p.ld_out = q.stride(-2), k.stride(-2), v.stride(-2), out.stride(-2)
    p.dtype, p.batch, p.heads, p.head_dim, p.sq, p.skv, p.scale = _dt(q), batch, 1, 512, sq, skv, scale
    p.causal = 0
    _run("attention", 4.0 * batch * sq * skv * 512, L.load().b200sd_attention512, p, f"b{batch} h1 d512 sq{sq} skv{skv}")


def groupnorm_workspace_bytes(n_img: int, hw: int, groups: int = 32) -> int:
    p = L.GroupNormParams()
    p.n_img, p.hw, p.groups = n_img, hw, groups
    return int(L.load().b200sd_groupnorm_workspace_bytes(C.byref(p)))


def _embed_tokens(ids, token_table, position_table, out):
    assert ids.dtype == torch.int32 and token_table.dtype == position_table.dtype == out.dtype == torch.float32
    p = L.EmbedParams()
    p.ids, p.token_table, p.position_table, p.out = _ptr(ids), _ptr(token_table), _ptr(position_table), _ptr(out)
    p.rows, p.seq_len, p.dim, p.vocab = ids.numel(), ids.shape[-1], token_table.shape[1], token_table.shape[0]
    _run("embed_tokens", _nbytes(out) * 2, L.load().b200sd_embed_tokens, p, f"rows{p.rows} dim{p.dim}")


def _gn_tile_sums(t1, t2, n_img, groups, sums):
    """[n_img, parts, C, 2] column-sum tables of one or two producers -> per-(image, group) sums [n_img, groups, 2]"""
    p = L.GnTileSumsParams()
    p.tile_sums1, p.parts1, p.c1 = _ptr(t1), t1.shape[1], t1.shape[2]
    if t2 is not None:
        p.tile_sums2, p.parts2, p.c2 = _ptr(t2), t2.shape[1], t2.shape[2]
    p.n_img, p.groups, p.sums = n_img, groups, _ptr(sums)
    assert t1.shape[0] == n_img and (t2 is None or t2.shape[0] == n_img) and sums.numel() == n_img * groups * 2
    _run("gn_tile_sums", _nbytes(t1, t2), L.load().b200sd_gn_tile_sums, p, f"n{n_img} parts{p.parts1}+{p.parts2} C{p.c1}+{p.c2}")


def _groupnorm(x1, x2, n_img, hw, groups, eps, gamma, beta, silu, y, raw, workspace, sums=None, stat_hw=0, sums_out=None):
    if x2 is not None:
        assert x2.dtype == x1.dtype
    if raw is not None:
        assert raw.dtype == y.dtype
    p = L.GroupNormParams()
    p.x1, p.c1 = _ptr(x1), x1.shape[-1]
    p.x2, p.c2 = (_ptr(x2), x2.shape[-1]) if x2 is not None else (None, 0)
    p.in_dtype, p.n_img, p.hw, p.groups, p.eps = _dt(x1), n_img, hw, groups, eps
    p.gamma, p.beta, p.silu = _ptr(gamma), _ptr(beta), int(silu)
    p.y, p.raw, p.out_dtype = _ptr(y), _ptr(raw), (_dt(y) if y is not None else L.F16)
    if workspace is not None:
        p.workspace, p.workspace_bytes = _ptr(workspace), workspace.numel() * workspace.element_size()
    if sums is not None:
        assert sums.dtype == torch.float32 and sums.numel() == n_img * groups * 2
        p.sums, p.stat_hw = _ptr(sums), stat_hw
    if sums_out is not None:
        assert sums_out.dtype == torch.float32 and sums_out.numel() == n_img * groups * 2
        p.sums_out = _ptr(sums_out)
    # algorithmic bytes: inputs read once for statistics and once to normalise, outputs written once (with the
    # statistics supplied, or only the statistics wanted, one read of the inputs)
    reads = 1 if (sums is not None or sums_out is not None) else 2
    work = reads * _nbytes(x1, x2) + (0.0 if sums_out is not None else _nbytes(y, raw))
    kind = "groupnorm" if sums is None and sums_out is None else ("groupnorm_apply" if sums is not None else "groupnorm_stats")
    _run(kind, work, L.load().b200sd_groupnorm, p, f"n{n_img} hw{hw} C{p.c1 + p.c2} {x1.dtype}")


def _layernorm(x, eps, gamma, beta, y):
    p = L.LayerNormParams()
    p.x, p.in_dtype, p.rows, p.cols, p.eps = _ptr(x), _dt(x), x.numel() // x.shape[-1], x.shape[-1], eps
    p.gamma, p.beta, p.y, p.out_dtype = _ptr(gamma), _ptr(beta), _ptr(y), _dt(y)
    _run("layernorm", _nbytes(x, y), L.load().b200sd_layernorm, p, f"rows{p.rows} cols{p.cols}")


def _im2col3x3(x, n_img, n_src, in_h, in_w, cin, stride, out, pad_lo=1):
    p = L.Im2colParams()
    p.x, p.in_dtype, p.n_img, p.n_src, p.in_h, p.in_w, p.cin, p.stride = (_ptr(x), _dt(x), n_img, n_src, in_h, in_w,
                                                                         cin, stride)
    p.out, p.out_dtype, p.k_pad, p.pad_lo = _ptr(out), _dt(out), out.shape[-1], pad_lo
    _run("im2col3x3", _nbytes(x, out), L.load().b200sd_im2col3x3, p)


def _upsample2x(x, n_img, h, w, c, out):
    p = L.UpsampleParams()
    p.x, p.in_dtype, p.n_img, p.h, p.w, p.c, p.out, p.out_dtype = _ptr(x), _dt(x), n_img, h, w, c, _ptr(out), _dt(out)
    _run("upsample2x", _nbytes(x, out), L.load().b200sd_upsample2x, p)


def _cfg_sched_step(eps2, x, x0_prev, coef, step_index):
    p = L.CfgSchedParams()
    p.eps2, p.eps_dtype, p.x, p.x0_prev = _ptr(eps2), _dt(eps2), _ptr(x), _ptr(x0_prev)
    p.coef, p.step_index, p.n = _ptr(coef), _ptr(step_index), x.numel()
    assert eps2.numel() == 2 * x.numel() and x.dtype == torch.float32 and coef.dtype == torch.float32
    assert step_index is None or step_index.dtype == torch.int32
    _run("cfg_sched_step", _nbytes(eps2, x, x, x0_prev, x0_prev), L.load().b200sd_cfg_sched_step, p)


def _small_linear(x, rows, k, sinusoid_dim, silu_in, silu_out, w, b, y):
    p = L.SmallLinearParams()
    p.x, p.rows, p.k, p.n = _ptr(x), rows, k, w.shape[0]
    p.sinusoid_dim, p.silu_in, p.silu_out = sinusoid_dim, int(silu_in), int(silu_out)
    p.w, p.w_dtype, p.b, p.y = _ptr(w), _dt(w), _ptr(b), _ptr(y)
    assert x.dtype == torch.float32 and y.dtype == torch.float32 and w.shape[1] == k
    _run("small_linear", _nbytes(w, y), L.load().b200sd_small_linear, p)


def _softmax_rows(x, y, rows, cols, scale):
    p = L.SoftmaxParams()
    p.x, p.in_dtype, p.ld_in = _ptr(x), _dt(x), x.stride(-2)
    p.y, p.out_dtype, p.ld_out = _ptr(y), _dt(y), y.stride(-2)
    p.rows, p.cols, p.scale = rows, cols, scale
    _run("softmax_rows", float(rows) * cols * (x.element_size() + 2), L.load().b200sd_softmax_rows, p)


def _denorm_u8(x, out):
    p = L.DenormU8Params()
    p.x, p.in_dtype, p.out, p.n = _ptr(x), _dt(x), _ptr(out), x.numel()
    assert out.dtype == torch.uint8 and out.numel() == x.numel()
    _run("denorm_u8", _nbytes(x, out), L.load().b200sd_denorm_u8, p)


def _cast(x, y):
    p = L.CastParams()
    p.x, p.in_dtype, p.y, p.out_dtype, p.n = _ptr(x), _dt(x), _ptr(y), _dt(y), x.numel()
    _run("cast", _nbytes(x, y), L.load().b200sd_cast, p)


def umma_selftest(a: torch.Tensor, b: torch.Tensor, b_mn_major: bool) -> torch.Tensor:
    """D[128,n] = A[128,k] @ B^T (B [n,k]) or A @ B (B [k,n], MN-major); pins the UMMA descriptor encodings."""
    k = a.shape[1]
    n = b.shape[1] if b_mn_major else b.shape[0]
    d = torch.empty(128, n, device=a.device, dtype=torch.float32)
    L.check(L.load().b200sd_umma_selftest(_ptr(a), _dt(a), _ptr(b), _dt(b), _ptr(d), n, k, int(b_mn_major),
                                          _stream()), "umma_selftest")
    return d


# ------------------------------------------------------------------------------------------------ registration
_SCHEMAS = {
    "gemm": ("(Tensor a, int n_img, int img_h, int img_w, int cin, int taps, Tensor w, Tensor? bias, bool bias_along_m, "
             "Tensor? row_bias, int row_bias_stride, Tensor? residual, Tensor(a!) out, bool geglu, int block_n, "
             "Tensor(b!)? workspace=None, int split_k=0, int act=0, Tensor(c!)? tile_sums=None, int rows_per_image=0, "
             "int conv_stride=1, int conv_pad_mode=0) -> ()",
             _gemm),
    "attention": ("(Tensor q, Tensor k, Tensor v, Tensor(a!) out, int batch, int heads, int sq, int skv, float scale, "
                  "bool causal=False) -> ()", _attention),
    "attention512": ("(Tensor q, Tensor k, Tensor v, Tensor(a!) out, int batch, int sq, int skv, float scale) -> ()",
                     _attention512),
    "embed_tokens": ("(Tensor ids, Tensor token_table, Tensor position_table, Tensor(a!) out) -> ()", _embed_tokens),
    "groupnorm": ("(Tensor x1, Tensor? x2, int n_img, int hw, int groups, float eps, Tensor? gamma, Tensor? beta, bool silu, "
                  "Tensor(a!)? y, Tensor(b!)? raw, Tensor(c!)? workspace, Tensor? sums=None, int stat_hw=0, "
                  "Tensor(d!)? sums_out=None) -> ()", _groupnorm),
    "gn_tile_sums": ("(Tensor t1, Tensor? t2, int n_img, int groups, Tensor(a!) sums) -> ()", _gn_tile_sums),
    "layernorm": ("(Tensor x, float eps, Tensor gamma, Tensor beta, Tensor(a!) y) -> ()", _layernorm),
    "im2col3x3": ("(Tensor x, int n_img, int n_src, int in_h, int in_w, int cin, int stride, Tensor(a!) out, int pad_lo=1) -> ()",
                  _im2col3x3),
    "upsample2x": ("(Tensor x, int n_img, int h, int w, int c, Tensor(a!) out) -> ()", _upsample2x),
    "cfg_sched_step": ("(Tensor eps2, Tensor(a!) x, Tensor(b!)? x0_prev, Tensor coef, Tensor? step_index) -> ()",
                       _cfg_sched_step),
    "small_linear": ("(Tensor x, int rows, int k, int sinusoid_dim, bool silu_in, bool silu_out, Tensor w, Tensor? b, "
                     "Tensor(a!) y) -> ()", _small_linear),
    "softmax_rows": ("(Tensor x, Tensor(a!) y, int rows, int cols, float scale) -> ()", _softmax_rows),
    "denorm_u8": ("(Tensor x, Tensor(a!) out) -> ()", _denorm_u8),
    "cast": ("(Tensor x, Tensor(a!) y) -> ()", _cast),
}

_library = None


def _register() -> None:
    global _library
    if _library is not None:
        return
    lib = torch.library.Library("b200sd", "DEF")
    for name, (schema, fn) in _SCHEMAS.items():
        lib.define(name + schema)
        lib.impl(name, fn, "CUDA")
        # out-variants: shape inference has nothing to compute
        lib.impl(name, (lambda *a, **k: None), "Meta")
    _library = lib


_register()
gemm = torch.ops.b200sd.gemm
attention = torch.ops.b200sd.attention
attention512 = torch.ops.b200sd.attention512
groupnorm = torch.ops.b200sd.groupnorm
gn_tile_sums = torch.ops.b200sd.gn_tile_sums
layernorm = torch.ops.b200sd.layernorm
im2col3x3 = torch.ops.b200sd.im2col3x3
upsample2x = torch.ops.b200sd.upsample2x
cfg_sched_step = torch.ops.b200sd.cfg_sched_step
small_linear = torch.ops.b200sd.small_linear
softmax_rows = torch.ops.b200sd.softmax_rows
denorm_u8 = torch.ops.b200sd.denorm_u8
cast = torch.ops.b200sd.cast
embed_tokens = torch.ops.b200sd.embed_tokens
