"""GPU: `b200sd_groupnorm_static_scratch`, the GroupNorm entry point for hosts that cannot give the library any
scratch memory (an FFI front end that owns no CUDA runtime of its own, or a server whose device memory is all
inside a pre-sized tensor arena).  It keeps its split statistics in an array compiled into libb200sd.so.

The test drives the entry point straight through the C ABI (ctypes structs from b200sd/_lib.py) and compares it with
the ordinary entry point, `b200sd_groupnorm`, which gets a caller-owned workspace.  Two calls are made back to back on
different inputs and shapes, so a result that leaned on what the previous call left in the shared array would show.
The file name sorts last on purpose: a `pytest -x` run reaches it only after every other test has run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _params(L, x1, x2, gamma, beta, y, raw, workspace, silu, eps=1e-5, groups=32):
    n, hw, c1 = x1.shape
    p = L.GroupNormParams()
    p.x1, p.c1 = x1.data_ptr(), c1
    p.x2, p.c2 = (x2.data_ptr(), x2.shape[-1]) if x2 is not None else (None, 0)
    p.in_dtype = {torch.float32: L.F32, torch.bfloat16: L.BF16, torch.float16: L.F16}[x1.dtype]
    p.n_img, p.hw, p.groups, p.eps = n, hw, groups, eps
    p.gamma, p.beta, p.silu = gamma.data_ptr(), beta.data_ptr(), int(silu)
    p.y, p.raw = y.data_ptr(), (raw.data_ptr() if raw is not None else None)
    p.out_dtype = {torch.bfloat16: L.BF16, torch.float16: L.F16}[y.dtype]
    if workspace is not None:
        p.workspace, p.workspace_bytes = workspace.data_ptr(), workspace.numel() * 4
    return p


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def test_static_scratch_groupnorm_matches_workspace_groupnorm(cuda, lib_built):
    from b200sd import _lib as L
    lib = L.load()
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator("cpu").manual_seed(11)
    # two consecutive calls: (fp32 trunk, concat of two inputs, SiLU) then (fp16 input, single tensor, no SiLU)
    cases = [(torch.float32, 2, 48 * 48, 640, 320, True), (torch.float16, 3, 24 * 24, 320, 0, False)]
    for dtype, n, hw, c1, c2, silu in cases:
        x1 = (torch.randn(n, hw, c1, generator=g) * 1.7 + 0.4).to(cuda).to(dtype)
        x2 = (torch.randn(n, hw, c2, generator=g) * 0.6 - 0.3).to(cuda).to(dtype) if c2 else None
        c = c1 + c2
        gamma = (1 + 0.1 * torch.randn(c, generator=g)).to(cuda)
        beta = (0.1 * torch.randn(c, generator=g)).to(cuda)
        # existing entry point, caller-owned workspace
        q = L.GroupNormParams()
        q.n_img, q.hw, q.groups = n, hw, 32
        ws = torch.empty(max(int(lib.b200sd_groupnorm_workspace_bytes(C.byref(q))) // 4, 1), device=cuda)
        y_ref = torch.empty(n, hw, c, device=cuda, dtype=torch.float16)
        raw_ref = torch.empty_like(y_ref)
        p = _params(L, x1, x2, gamma, beta, y_ref, raw_ref, ws, silu)
        L.check(lib.b200sd_groupnorm(C.byref(p), stream), "groupnorm")
        # new entry point: no workspace pointer at all
        y = torch.full_like(y_ref, float("nan"))
        raw = torch.full_like(y_ref, float("nan"))
        p2 = _params(L, x1, x2, gamma, beta, y, raw, None, silu)
        L.check(lib.b200sd_groupnorm_static_scratch(C.byref(p2), stream), "groupnorm_static_scratch")
        torch.cuda.synchronize()
        assert torch.isfinite(y.float()).all()
        # same kernels, same split count: one fp16 rounding is the tolerance the GroupNorm tests use (6e-4)
        assert _rel(y, y_ref) < 6e-4 and torch.equal(raw, raw_ref)
        # and both agree with torch's GroupNorm
        cat = torch.cat([x1, x2], -1) if c2 else x1
        ref = torch.nn.functional.group_norm(cat.float().permute(0, 2, 1), 32, gamma, beta, 1e-5)
        ref = (torch.nn.functional.silu(ref) if silu else ref).permute(0, 2, 1)
        assert _rel(y, ref) < 6e-4


def test_static_scratch_groupnorm_large_batch_uses_fewer_splits(cuda, lib_built):
    """CFG batch 32 at 96x96 (the benchmarked shape): the ordinary split count would need 75776 floats, more than the
    built-in 65536, so the entry point must run with fewer splits instead of refusing -- and still match."""
    from b200sd import _lib as L
    lib = L.load()
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator("cpu").manual_seed(5)
    n, hw, c = 32, 96 * 96, 320
    x = (torch.randn(n, hw, c, generator=g) * 1.3 + 0.2).to(cuda).half()
    gamma, beta = (1 + 0.1 * torch.randn(c, generator=g)).to(cuda), (0.1 * torch.randn(c, generator=g)).to(cuda)
    q = L.GroupNormParams()
    q.n_img, q.hw, q.groups = n, hw, 32
    need = int(lib.b200sd_groupnorm_workspace_bytes(C.byref(q)))
    assert need // 4 > 65536, "this shape is meant to exceed the built-in array at the ordinary split count"
    ws = torch.empty(need // 4, device=cuda)
    y_ref, y = torch.empty_like(x), torch.full_like(x, float("nan"))
    L.check(lib.b200sd_groupnorm(C.byref(_params(L, x, None, gamma, beta, y_ref, None, ws, True)), stream), "groupnorm")
    for _ in range(2):  # two calls in a row on the shared array
        L.check(lib.b200sd_groupnorm_static_scratch(C.byref(_params(L, x, None, gamma, beta, y, None, None, True)), stream),
                "groupnorm_static_scratch")
        torch.cuda.synchronize()
        assert _rel(y, y_ref) < 6e-4
        y.fill_(float("nan"))


def test_static_scratch_groupnorm_rejects_oversize(cuda, lib_built):
    """beyond the built-in 256 KiB the call must fail with a clear message, not write out of bounds"""
    from b200sd import _lib as L
    lib = L.load()
    n, hw, c = 1200, 4, 64  # even ONE split per image does not fit: 1200 images x 32 groups x 2 floats = 76800 > 65536
    x = torch.zeros(n, hw, c, device=cuda, dtype=torch.float16)
    y = torch.empty_like(x)
    gamma, beta = torch.ones(c, device=cuda), torch.zeros(c, device=cuda)
    p = _params(L, x, None, gamma, beta, y, None, None, False)
    rc = lib.b200sd_groupnorm_static_scratch(C.byref(p), torch.cuda.current_stream().cuda_stream)
    msg = lib.b200sd_last_error().decode()
    assert rc == -4 and "built-in array holds 65536" in msg and "b200sd_groupnorm" in msg
