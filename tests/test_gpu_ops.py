"""GPU parity tests of every libb200sd kernel, called through the C ABI (ctypes -> torch.library ops), against
fp32 references of the same op on the same 16-bit-rounded inputs.  Every test that takes `K` runs twice: with fp16
and with bf16 tensor-core operands.  Tolerances are stated per test."""
import math
import os
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent

pytestmark = pytest.mark.gpu
BF, FH = torch.bfloat16, torch.float16


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="module", params=[FH, BF], ids=["fp16", "bf16"])
def K(request, cuda, lib_built):
    import b200sd  # noqa: F401  (loads the library)
    from b200sd.layers import Kernels
    return Kernels(cuda, torch.float32, request.param)


def rnd(*shape, dev, seed=0, scale=1.0, dtype=torch.float32):
    g = torch.Generator("cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dev).to(dtype)


def out_tol(dt):  # one rounding of the output to dt
    return 4e-3 if dt == BF else 6e-4


# ------------------------------------------------------------------------------------------------ UMMA descriptors
@pytest.mark.parametrize("n,k", [(64, 64), (128, 128), (256, 64), (160, 192)])
@pytest.mark.parametrize("adt,bdt", [(BF, BF), (FH, FH)])
def test_umma_selftest_k_major(cuda, lib_built, n, k, adt, bdt):
    from b200sd import ops
    a, b = rnd(128, k, dev=cuda, seed=1, dtype=adt), rnd(n, k, dev=cuda, seed=2, dtype=bdt)
    d = ops.umma_selftest(a, b, False)
    ref = a.float() @ b.float().t()
    assert rel(d, ref) < 1e-5, "K-major SWIZZLE_128B descriptor / idesc formats / TMEM layout"


@pytest.mark.parametrize("n,k", [(64, 64), (64, 128), (128, 64), (256, 128)])
@pytest.mark.parametrize("dt", [BF, FH])
def test_umma_selftest_mn_major_b(cuda, lib_built, n, k, dt):
    from b200sd import ops
    a, b = rnd(128, k, dev=cuda, seed=3, dtype=dt), rnd(k, n, dev=cuda, seed=4, dtype=dt)
    d = ops.umma_selftest(a, b, True)
    ref = a.float() @ b.float()
    assert rel(d, ref) < 1e-5, "MN-major B operand descriptor (used for V in attention)"


# ------------------------------------------------------------------------------------------------ GEMM / conv
@pytest.mark.parametrize("m,n,k,block_n", [
    (128, 64, 64, 0), (300, 320, 320, 0), (1000, 640, 1024, 128), (257, 1280, 384, 256), (512, 960, 320, 160),
    (77 * 2, 256, 1024, 0), (4096, 2560, 320, 256), (130, 4, 128, 0), (64, 520, 72, 0)])
def test_linear_plain(K, cuda, m, n, k, block_n):
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    out = K.linear(a, w, bias, out_dtype=torch.float32, block_n=block_n)
    ref = a.float() @ w.float().t() + bias
    assert rel(out, ref) < 2e-5  # fp32 accumulate of identical 16-bit operands
    out_h = K.linear(a, w, bias, out_dtype=dt, block_n=block_n)
    assert rel(out_h, ref) < out_tol(dt)


@pytest.mark.parametrize("m,n,k,split", [(300, 1280, 2048, 0), (300, 1280, 2048, 4), (130, 320, 1024, 8), (1152, 640, 5120, 3),
                                         (64, 64, 512, 2)])
def test_linear_split_k_is_exact_and_deterministic(K, cuda, m, n, k, split):
    """few output tiles -> K is split across CTAs through the workspace; partials are added in split order"""
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias, res = rnd(n, dev=cuda, seed=3), rnd(m, n, dev=cuda, seed=4)
    ref = a.float() @ w.float().t() + bias + res
    outs = [K.linear(a, w, bias, residual=res, out_dtype=torch.float32, split_k=split) for _ in range(3)]
    assert rel(outs[0], ref) < 2e-5
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    unsplit = K.linear(a, w, bias, residual=res, out_dtype=torch.float32, split_k=1)
    assert rel(outs[0], unsplit) < 5e-6  # same products, different fp32 summation order
    assert int(K.gemm_ws.view(torch.int32)[:1024].abs().max()) == 0  # tile counters are left at zero


def test_conv3x3_split_k_low_res_levels(K, cuda):
    """the 24x24 / 12x12 UNet levels at CFG batch 2: 1280 channels, automatic split"""
    from b200sd.layers import pack_conv3x3
    dt = K.op_dtype
    for n, h, cin, cout in [(2, 12, 1280, 1280), (2, 24, 640, 1280)]:
        x = rnd(n, h, h, cin, dev=cuda, seed=1, dtype=dt)
        wt = rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5)
        bias = rnd(cout, dev=cuda, seed=3)
        out = K.conv3x3(x, pack_conv3x3(wt, dt), bias, out_dtype=torch.float32)
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dt).float(), bias, padding=1).permute(0, 2, 3, 1)
        assert rel(out, ref) < 2e-5


def test_linear_mixed_operand_formats_rejected(cuda, lib_built):
    """fp16 x bf16 in one tcgen05.mma traps on B200 (measured: illegal instruction), so the ABI refuses it up front"""
    from b200sd.layers import Kernels
    Kh = Kernels(cuda, torch.float32, FH)
    a, w = rnd(300, 256, dev=cuda, seed=1, dtype=FH), rnd(320, 256, dev=cuda, seed=2, scale=1 / 16, dtype=BF)
    with pytest.raises(RuntimeError, match="must be equal"):
        Kh.linear(a, w, None, out_dtype=torch.float32)
    torch.cuda.synchronize()  # the context is still healthy


@pytest.mark.parametrize("res_dtype", [torch.float32, BF, FH])
def test_linear_residual_inplace_and_bias_along_m(K, cuda, res_dtype):
    dt = K.op_dtype
    m, n, k = 700, 640, 640
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    res = rnd(m, n, dev=cuda, seed=4, dtype=res_dtype)
    ref = a.float() @ w.float().t() + bias + res.float()
    t = res.clone()
    K.linear(a, w, bias, residual=t, out=t)  # aliasing out == residual
    # 16-bit in place: the chunk is rounded to the residual's format and added to it in L2 by a TMA reduce -- two roundings
    assert rel(t, ref) < (2e-5 if res_dtype == torch.float32 else 1.5 * out_tol(res_dtype))
    bm = rnd(m, dev=cuda, seed=5)
    out = K.linear(a, w, bm, out_dtype=torch.float32, bias_along_m=True)
    assert rel(out, a.float() @ w.float().t() + bm[:, None]) < 2e-5


@pytest.mark.parametrize("c", [64, 320])
def test_linear_geglu(K, cuda, c):
    from b200sd import ops
    from b200sd.layers import pack_geglu
    dt = K.op_dtype
    m = 600
    a = rnd(m, c, dev=cuda, seed=1, dtype=dt)
    w, b = rnd(8 * c, c, dev=cuda, seed=2, scale=c ** -0.5), rnd(8 * c, dev=cuda, seed=3, scale=0.1)
    bn = ops.pick_block_n(8 * c, True)
    wp, bp = pack_geglu(w, b, bn, dt)
    out = K.linear(a, wp, bp, out_dtype=dt, geglu=True, block_n=bn)
    p = a.float() @ w.to(dt).float().t() + b
    val, gate = p.chunk(2, dim=-1)
    ref = val * F.gelu(gate)
    assert out.shape == (m, 4 * c) and rel(out, ref) < out_tol(dt) * 1.25


@pytest.mark.parametrize("n,h,w,cin,cout", [
    (2, 96, 96, 64, 64), (2, 48, 48, 128, 160), (4, 24, 24, 64, 128), (6, 12, 12, 128, 256), (3, 12, 12, 64, 64),
    (2, 16, 16, 192, 320), (1, 8, 8, 64, 4), (2, 20, 36, 72, 96), (1, 40, 200, 64, 3), (4, 2, 2, 128, 128)])
def test_conv3x3_implicit_gemm(K, cuda, n, h, w, cin, cout):
    from b200sd.layers import pack_conv3x3
    dt = K.op_dtype
    x = rnd(n, h, w, cin, dev=cuda, seed=1, dtype=dt)
    wt = rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5)
    bias = rnd(cout, dev=cuda, seed=3)
    out = K.conv3x3(x, pack_conv3x3(wt, dt), bias, out_dtype=torch.float32)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dt).float(), bias, padding=1).permute(0, 2, 3, 1)
    assert out.shape == ref.shape and rel(out, ref) < 2e-5


@pytest.mark.parametrize("pad_mode", [0, 1])
@pytest.mark.parametrize("n,h,w,cin,cout", [
    (2, 96, 96, 64, 64), (2, 48, 48, 128, 160), (4, 24, 24, 64, 128), (32, 24, 24, 128, 128), (3, 10, 14, 64, 96),
    (2, 9, 9, 64, 64), (1, 2, 2, 64, 64), (2, 192, 192, 64, 64), (6, 48, 48, 320, 320)])
def test_conv3x3_stride2_without_im2col(K, cuda, n, h, w, cin, cout, pad_mode):
    """Downsample2D through the conv kernel's stride-2 form (TMA box with element strides 2 in x and y):
    pad_mode 0 = Conv2d(3, stride=2, padding=1) of the UNet; pad_mode 1 = F.pad(x, (0,1,0,1)) + Conv2d(3, stride=2,
    padding=0) of the VAE encoder.  Even, odd and tiny sizes, multi-image pixel boxes (24x24 at 32 images), and the
    producer-side GroupNorm column sums."""
    from b200sd import ops
    from b200sd.layers import pack_conv3x3, tile_sums_of
    dt = K.op_dtype
    x = rnd(n, h, w, cin, dev=cuda, seed=1, dtype=dt)
    wt = rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5)
    bias = rnd(cout, dev=cuda, seed=3)
    xin = x.float().permute(0, 3, 1, 2)
    if pad_mode == 1:
        ref = F.conv2d(F.pad(xin, (0, 1, 0, 1)), wt.to(dt).float(), bias, stride=2)
    else:
        ref = F.conv2d(xin, wt.to(dt).float(), bias, stride=2, padding=1)
    ref = ref.permute(0, 2, 3, 1)
    out = K.conv3x3(x, pack_conv3x3(wt, dt), bias, out_dtype=torch.float32, stride=2, pad_mode=pad_mode, gn_next=True)
    assert out.shape == ref.shape and rel(out, ref) < 2e-5
    ts = tile_sums_of(out)
    if ts is not None and cout % 32 == 0:   # the column sums describe the stored values
        sums = torch.empty(n, 32, 2, device=cuda)
        ops.gn_tile_sums(ts, None, n, 32, sums)
        g = out.view(n, -1, 32, cout // 32)
        want = torch.stack([g.sum(dim=(1, 3)), (g * g).sum(dim=(1, 3))], dim=-1)
        assert rel(sums, want) < 1e-4


def test_conv3x3_temb_rowbias_and_residual(K, cuda):
    from b200sd.layers import pack_conv3x3
    dt = K.op_dtype
    n, h, w, cin, cout = 4, 24, 24, 128, 128
    x = rnd(n, h, w, cin, dev=cuda, seed=1, dtype=dt)
    wt = rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5)
    table = rnd(n, 3 * cout, dev=cuda, seed=3)
    res = rnd(n, h, w, cout, dev=cuda, seed=4)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dt).float(), None, padding=1).permute(0, 2, 3, 1)
    out = K.conv3x3(x, pack_conv3x3(wt, dt), None, row_bias=table[:, cout:2 * cout], residual=res,
                    out_dtype=torch.float32)
    assert rel(out, ref + table[:, None, None, cout:2 * cout] + res) < 2e-5
    out1 = K.conv3x3(x, pack_conv3x3(wt, dt), None, row_bias=table[:1, :cout], out_dtype=torch.float32)  # broadcast
    assert rel(out1, ref + table[0, :cout]) < 2e-5


# ------------------------------------------------------------------------------------------------ norms
@pytest.mark.parametrize("dtype", [torch.float32, BF, FH])
@pytest.mark.parametrize("c1,c2,hw,silu", [(320, 0, 96 * 96, True), (1280, 640, 24 * 24, True), (640, 320, 48 * 48, True),
                                           (128, 0, 64 * 64, False), (64, 64, 16 * 16, True), (1280, 1280, 144, True),
                                           (128, 0, 4, True)])
def test_groupnorm(K, cuda, dtype, c1, c2, hw, silu):
    n = 2
    x1 = rnd(n, hw, 1, c1, dev=cuda, seed=1, dtype=dtype) * 1.5 + 0.3
    x2 = (rnd(n, hw, 1, c2, dev=cuda, seed=2, dtype=dtype) * 0.7 - 0.2) if c2 else None
    c = c1 + c2
    gamma, beta = rnd(c, dev=cuda, seed=3) * 0.1 + 1, rnd(c, dev=cuda, seed=4) * 0.1
    y, raw = K.groupnorm(x1, x2, gamma, beta, 1e-5, silu, want_raw=True)
    cat = torch.cat([x1, x2], -1) if c2 else x1
    ref = F.group_norm(cat.float().permute(0, 3, 1, 2), 32, gamma, beta, 1e-5)
    ref = (F.silu(ref) if silu else ref).permute(0, 2, 3, 1)
    assert y.dtype == K.op_dtype and rel(y, ref) < out_tol(K.op_dtype)
    assert torch.equal(raw, cat.to(K.op_dtype))


@pytest.mark.parametrize("dtype", [torch.float32, BF])
@pytest.mark.parametrize("cols", [64, 320, 640, 1280])
def test_layernorm(K, cuda, dtype, cols):
    x = rnd(1000, cols, dev=cuda, seed=1, dtype=dtype) * 2 + 0.5
    gamma, beta = rnd(cols, dev=cuda, seed=3) * 0.1 + 1, rnd(cols, dev=cuda, seed=4) * 0.1
    y = K.layernorm(x, gamma, beta)
    assert rel(y, F.layer_norm(x.float(), (cols,), gamma, beta, 1e-5)) < out_tol(K.op_dtype)


# ------------------------------------------------------------------------------------------------ data movement
def test_im2col_conv_in_and_stride2(K, cuda):
    dt = K.op_dtype
    lat = rnd(2, 16, 16, 4, dev=cuda, seed=1)
    cols = K.im2col(lat, 4, 1, 64)  # CFG pair shares the latent: n_img = 2 * n_src
    unf = F.unfold(lat.permute(0, 3, 1, 2), 3, padding=1)  # [2, 4*9, 256] with (c, tap) order
    ref = unf.view(2, 4, 9, 256).permute(0, 3, 2, 1).reshape(2, 16, 16, 36)
    assert torch.equal(cols[:2, ..., :36], ref.to(dt)) and torch.equal(cols[2:, ..., :36], ref.to(dt))
    assert float(cols[..., 36:].abs().max()) == 0.0
    x = rnd(2, 24, 24, 64, dev=cuda, seed=2)
    c2 = K.im2col(x, 2, 2, 9 * 64)
    unf = F.unfold(x.permute(0, 3, 1, 2), 3, padding=1, stride=2)
    ref = unf.view(2, 64, 9, 144).permute(0, 3, 2, 1).reshape(2, 12, 12, 576)
    assert torch.equal(c2, ref.to(dt))
    # the VAE encoder's Downsample2D(padding=0): pad (0,1,0,1), then 3x3 stride 2 without padding
    c3 = K.im2col(x, 2, 2, 9 * 64, pad_lo=0)
    unf = F.unfold(F.pad(x.permute(0, 3, 1, 2), (0, 1, 0, 1)), 3, padding=0, stride=2)
    ref = unf.view(2, 64, 9, 144).permute(0, 3, 2, 1).reshape(2, 12, 12, 576)
    assert c3.shape == ref.shape and torch.equal(c3, ref.to(dt))
    odd = rnd(1, 9, 9, 8, dev=cuda, seed=3)   # odd sizes: (9 + 1 - 3) // 2 + 1 = 4
    c4 = K.im2col(odd, 1, 2, 72, pad_lo=0)
    unf = F.unfold(F.pad(odd.permute(0, 3, 1, 2), (0, 1, 0, 1)), 3, padding=0, stride=2)
    assert torch.equal(c4, unf.view(1, 8, 9, 16).permute(0, 3, 2, 1).reshape(1, 4, 4, 72).to(dt))


def test_upsample2x(K, cuda):
    x = rnd(2, 12, 12, 64, dev=cuda, seed=1)
    up = K.upsample2x(x)
    ref = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1)
    assert torch.equal(up, ref.to(K.op_dtype))


def test_cast_and_denorm(cuda, lib_built):
    from b200sd import ops
    x = rnd(4096, dev=cuda, seed=1) * 1.2
    for dt in (BF, FH):
        y = torch.empty(4096, device=cuda, dtype=dt)
        ops.cast(x, y)
        assert torch.equal(y, x.to(dt))
        z = torch.empty(4096, device=cuda, dtype=torch.float32)
        ops.cast(y, z)
        assert torch.equal(z, y.float())
    big = torch.full((8,), 1e6, device=cuda)
    h = torch.empty(8, device=cuda, dtype=FH)
    ops.cast(big, h)
    assert float(h[0]) == 65504.0  # fp16 stores saturate instead of producing inf
    u8 = torch.empty(4096, device=cuda, dtype=torch.uint8)
    ops.denorm_u8(x, u8)
    ref = ((x / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)
    assert int((u8.int() - ref.int()).abs().max()) <= 1 and float((u8 != ref).float().mean()) < 1e-3


# ------------------------------------------------------------------------------------------------ K4 vs the oracle
@pytest.mark.parametrize("kind", ["ddim", "dpm++2m", "euler"])
@pytest.mark.parametrize("pred", ["v_prediction", "epsilon"])
def test_cfg_sched_step_matches_oracle(cuda, lib_built, kind, pred):
    from b200sd import B200Scheduler
    from oracle import sd21 as O
    steps, B = 10, 2
    sch = B200Scheduler(kind, pred, cuda)
    sch.set_timesteps(steps)
    osch = O.SCHEDULERS[kind](pred)
    osch.set_timesteps(steps)
    assert sch._ts == osch.timesteps and math.isclose(sch.init_noise_sigma, osch.init_noise_sigma, rel_tol=1e-12)
    x_ref = rnd(B, 16, 16, 4, dev="cpu", seed=1) * osch.init_noise_sigma
    x = x_ref.clone().to(cuda)
    x0p = torch.zeros_like(x) if sch.needs_x0_prev else None
    x0p_ref = None
    idx = torch.zeros(1, dtype=torch.int32, device=cuda)
    for i in range(steps):
        eps2 = rnd(2 * B, 16, 16, 4, dev="cpu", seed=10 + i)
        m = O.cfg_combine(eps2, 7.5)
        x_ref, x0p_ref = O.linear_step(x_ref, m, x0p_ref, osch.coef(i))
        idx.fill_(i)
        sch.step_cfg(eps2.to(cuda), idx, x, x0p, 7.5)
        assert rel(x.cpu(), x_ref) < 2e-6, (kind, pred, i)
    # 16-bit model output
    for dt in (BF, FH):
        eps2 = rnd(2 * B, 16, 16, 4, dev=cuda, seed=99, dtype=dt)
        x1 = x.clone()
        idx.fill_(3)
        sch.step_cfg(eps2, idx, x1, x0p.clone() if x0p is not None else None, 7.5)
        c = osch.coef(3)
        m = O.cfg_combine(eps2.float().cpu(), 7.5)
        ref = c.cx * x.cpu() + c.cm * m + (c.cp * x0p.cpu() if x0p is not None else 0)
        assert rel(x1.cpu(), ref) < 2e-6


def test_small_linear_time_embedding(K, cuda):
    from b200sd import ops
    from oracle import sd21 as O
    dt = K.op_dtype
    t = torch.tensor([961.0, 1.0, 481.0], device=cuda)
    w1, b1 = rnd(1280, 320, dev=cuda, seed=1, scale=320 ** -0.5, dtype=dt), rnd(1280, dev=cuda, seed=2, scale=0.05)
    y = torch.empty(3, 1280, device=cuda)
    ops.small_linear(t, 3, 320, 320, False, True, w1, b1, y)
    ref = F.silu(O.timestep_embedding(t, 320) @ w1.float().t() + b1)
    assert rel(y, ref) < 1e-4
    w2 = rnd(2000, 1280, dev=cuda, seed=3, scale=1280 ** -0.5, dtype=dt)
    y2 = torch.empty(3, 2000, device=cuda)
    ops.small_linear(y, 3, 1280, 0, True, False, w2, None, y2)
    assert rel(y2, F.silu(y) @ w2.float().t()) < 1e-4


@pytest.mark.parametrize("dtype", [torch.float32, BF])
def test_softmax_rows(K, cuda, dtype):
    from b200sd import ops
    x = rnd(300, 2304, dev=cuda, seed=1, scale=3.0, dtype=dtype)
    y = torch.empty(300, 2304, device=cuda, dtype=K.op_dtype)
    ops.softmax_rows(x, y, 300, 2304, 0.25)
    assert rel(y, torch.softmax(x.float() * 0.25, -1)) < out_tol(K.op_dtype)


# ------------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("batch,heads,sq,skv", [(2, 2, 256, 256), (1, 5, 1024, 1024), (2, 3, 576, 576), (3, 2, 144, 144),
                                                (2, 4, 256, 77), (1, 2, 9216, 77), (1, 1, 64, 64), (2, 1, 2304, 2304),
                                                (4, 2, 16, 16), (2, 2, 4, 77)])
def test_flash_attention(K, cuda, batch, heads, sq, skv):
    dt = K.op_dtype
    c = heads * 64
    self_attn = sq == skv
    if self_attn:
        qkv = rnd(batch * sq, 3 * c, dev=cuda, seed=1, dtype=dt)
        q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
    else:
        q = rnd(batch * sq, c, dev=cuda, seed=1, dtype=dt)
        kv = rnd(batch * skv, 2 * c, dev=cuda, seed=2, dtype=dt)
        k, v = kv[:, :c], kv[:, c:]
    out = K.attention(q, k, v, batch, heads, sq, skv, 0.125)
    qf = q.float().view(batch, sq, heads, 64).transpose(1, 2)
    kf = k.float().view(batch, skv, heads, 64).transpose(1, 2)
    vf = v.float().view(batch, skv, heads, 64).transpose(1, 2)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * 0.125, -1) @ vf).transpose(1, 2).reshape(batch * sq, c)
    assert out.dtype == dt and rel(out, ref) < 2 * out_tol(dt)  # 16-bit P and 16-bit output


def test_flash_attention_large_logits_rescale(K, cuda):
    """running max grows by far more than 2^8 between kv tiles -> the lazy O rescale path runs"""
    dt = K.op_dtype
    batch, heads, s = 1, 2, 512
    c = heads * 64
    q = rnd(s, c, dev=cuda, seed=1, dtype=dt)
    k = rnd(s, c, dev=cuda, seed=2, dtype=dt)
    k[256:] *= 8.0
    v = rnd(s, c, dev=cuda, seed=3, dtype=dt)
    out = K.attention(q, k, v, batch, heads, s, s, 0.5)
    qf, kf, vf = (t.float().view(1, s, heads, 64).transpose(1, 2) for t in (q, k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * 0.5, -1) @ vf).transpose(1, 2).reshape(s, c)
    assert rel(out, ref) < 2 * out_tol(dt)


@pytest.mark.parametrize("env", [{"B200SD_ATTN_LAYOUT": "1"}, {"B200SD_ATTN_PAIR_BAR": "0"},
                                 {"B200SD_ATTN_LAYOUT": "1", "B200SD_ATTN_EXP_EMU": "0", "B200SD_ATTN_STAGGER": "0"}])
def test_flash_attention_other_kernel_switches(cuda, lib_built, env):
    """the attention kernel's non-default variants (quad-row softmax layout, tile-wide row-max barrier, no polynomial
    exponentials, no stagger) read their switches once per process, so they are exercised in a child process: self-,
    cross- (ragged kv tile, persistent CTAs) and causal attention, fp16 and bf16, against explicit softmax"""
    import subprocess
    import sys
    code = r"""
import sys, torch
sys.path[:0] = [%r, %r]
import b200sd
from b200sd.layers import Kernels
dev = torch.device('cuda:0')
worst = 0.0
for dt in (torch.float16, torch.bfloat16):
    K = Kernels(dev, torch.float32, dt)
    for batch, heads, sq, skv, causal in [(2, 2, 576, 576, False), (1, 5, 1024, 1024, False), (3, 4, 320, 77, False),
                                          (2, 2, 300, 300, True), (32, 20, 144, 77, False), (1, 1, 4, 4, False)]:
        g = torch.Generator('cpu').manual_seed(sq + skv)
        c = heads * 64
        q = torch.randn(batch * sq, c, generator=g).to(dev).to(dt)
        k = torch.randn(batch * skv, c, generator=g).to(dev).to(dt)
        v = torch.randn(batch * skv, c, generator=g).to(dev).to(dt)
        out = K.attention(q, k, v, batch, heads, sq, skv, 0.125, causal=causal)
        qf, kf, vf = (t.float().view(batch, -1, heads, 64).transpose(1, 2) for t in (q, k, v))
        sc = qf @ kf.transpose(-1, -2) * 0.125
        if causal:
            sc = sc.masked_fill(torch.ones(sq, skv, device=dev).triu(1).bool(), float('-inf'))
        ref = (torch.softmax(sc, -1) @ vf).transpose(1, 2).reshape(batch * sq, c)
        err = ((out.float() - ref).norm() / ref.norm()).item()
        tol = 2 * (6e-4 if dt == torch.float16 else 4e-3)
        assert err < tol, (dt, batch, heads, sq, skv, causal, err)
        worst = max(worst, err)
print('ok', worst)
""" % (str(ROOT), str(ROOT / "scalable-hw-agnostic-inference_b200"))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip().startswith("ok"), r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("batch,s", [(1, 128), (2, 256), (1, 1024), (3, 320), (2, 77), (1, 2304)])
def test_flash_attention_head_dim_512(K, cuda, batch, s):
    """the VAE mid block's single 512-wide head: full, ragged and multi-tile sequences, strided Q|K|V views of one fused
    projection, against explicit softmax(q k^T / sqrt(512)) v"""
    from b200sd import ops
    dt = K.op_dtype
    c = 512
    qkv = rnd(batch * s, 3 * c, dev=cuda, seed=s, dtype=dt)
    q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
    guard = torch.full((batch * s + 2, c), 7.0, device=cuda, dtype=dt)
    out = guard[1:-1]
    ops.attention512(q, k, v, out, batch, s, s, c ** -0.5)
    qf, kf, vf = (t.float().view(batch, s, c) for t in (q, k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * c ** -0.5, -1) @ vf).reshape(batch * s, c)
    assert rel(out, ref) < 2 * out_tol(dt)
    assert float(guard[0].float().min()) == 7.0 and float(guard[-1].float().max()) == 7.0   # nothing outside the output


def test_flash_attention_head_dim_512_large_logits(K, cuda):
    """running max grows by far more than 2^8 between kv tiles -> the lazy O rescale over all 256 O columns runs"""
    from b200sd import ops
    dt = K.op_dtype
    s, c = 512, 512
    q = rnd(s, c, dev=cuda, seed=1, dtype=dt)
    k = rnd(s, c, dev=cuda, seed=2, dtype=dt)
    k[256:] *= 6.0
    v = rnd(s, c, dev=cuda, seed=3, dtype=dt)
    out = torch.empty(s, c, device=cuda, dtype=dt)
    ops.attention512(q, k, v, out, 1, s, s, 0.25)
    ref = torch.softmax(q.float() @ k.float().t() * 0.25, -1) @ v.float()
    assert rel(out, ref) < 2 * out_tol(dt)


# ------------------------------------------------------------------------------------------------ text-encoder ops
@pytest.mark.parametrize("batch,heads,s", [(3, 2, 77), (2, 16, 77), (1, 2, 33), (2, 2, 300), (1, 1, 640)])
def test_flash_attention_causal(K, cuda, batch, heads, s):
    """causal flag (text encoder): key j visible to query i iff j <= i, across one and several kv tiles"""
    dt = K.op_dtype
    c = heads * 64
    qkv = rnd(batch * s, 3 * c, dev=cuda, seed=4, dtype=dt)
    q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
    out = K.attention(q, k, v, batch, heads, s, s, 0.125, causal=True)
    qf, kf, vf = (t.float().view(batch, s, heads, 64).transpose(1, 2) for t in (q, k, v))
    mask = torch.full((s, s), float("-inf"), device=cuda).triu(1)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * 0.125 + mask, -1) @ vf).transpose(1, 2).reshape(batch * s, c)
    assert rel(out, ref) < 2 * out_tol(dt)


def test_attention_causal_needs_square(K, cuda):
    dt = K.op_dtype
    q, kv = rnd(64, 64, dev=cuda, seed=1, dtype=dt), rnd(77, 128, dev=cuda, seed=2, dtype=dt)
    with pytest.raises(RuntimeError, match="causal"):
        K.attention(q, kv[:, :64], kv[:, 64:], 1, 1, 64, 77, 0.125, causal=True)


@pytest.mark.parametrize("m,n,k,res,out32", [(154, 4096, 1024, False, False), (154, 512, 128, False, True),
                                             (300, 320, 320, True, True), (231, 520, 64, False, False)])
def test_linear_gelu_epilogue(K, cuda, m, n, k, res, out32):
    """act = GELU(erf) on (a w^T + bias), before the residual: TMA-store, staged and row-wise epilogues"""
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    r = rnd(m, n, dev=cuda, seed=4) if res else None
    out = K.linear(a, w, bias, residual=r, out_dtype=torch.float32 if out32 else dt, gelu=True)
    ref = F.gelu(a.float() @ w.float().t() + bias)
    if res:
        ref = ref + r
    assert rel(out, ref) < (2e-5 if out32 else out_tol(dt))


def test_embed_tokens(cuda, lib_built):
    from b200sd import ops
    g = torch.Generator("cpu").manual_seed(3)
    tok, pos = torch.randn(1000, 128, generator=g).to(cuda), torch.randn(77, 128, generator=g).to(cuda)
    ids = torch.randint(0, 1000, (5, 77), generator=g).to(cuda).to(torch.int32)
    out = torch.empty(5 * 77, 128, device=cuda)
    ops.embed_tokens(ids, tok, pos, out)
    ref = tok[ids.long()] + pos[None]
    assert torch.equal(out.view(5, 77, 128), ref)  # one fp32 add per element: bit-exact


@pytest.mark.parametrize("m,n,k", [(700, 640, 640), (4096, 320, 320), (333, 520, 128)])
def test_linear_inplace_fp32_residual_by_tma_reduce_equals_the_read_path(K, cuda, m, n, k):
    """out == residual (fp32): the chunk is added in L2 by a TMA reduce; the same fp32 sum as reading the residual"""
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    res = rnd(m, n, dev=cuda, seed=4)
    t = res.clone()
    K.linear(a, w, bias, residual=t, out=t)                      # reduce-add path
    sep = K.linear(a, w, bias, residual=res, out_dtype=torch.float32)  # residual read by the epilogue
    assert torch.equal(t, sep)
    assert rel(t, a.float() @ w.float().t() + bias + res) < 2e-5


@pytest.mark.parametrize("m,n,k", [(600, 320, 1280), (257, 640, 256)])
def test_linear_fp32_residual_with_16bit_output(K, cuda, m, n, k):
    """the feed-forward's second linear: fp32 trunk residual, 16-bit result for the next GEMM (staged epilogue)"""
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    res = rnd(m, n, dev=cuda, seed=4)
    out = K.linear(a, w, bias, residual=res, out_dtype=dt)
    assert out.dtype == dt and rel(out, a.float() @ w.float().t() + bias + res) < out_tol(dt)


# ------------------------------------------------------------------------------------------------ guard bands
def _guarded(m, n, dtype, dev, sentinel=512.0):
    """an [m, n] view in the middle of a larger buffer filled with a sentinel (4 guard rows, 16/48 guard columns)"""
    buf = torch.full((m + 8, n + 64), sentinel, device=dev, dtype=dtype)
    return buf, buf[4:4 + m, 16:16 + n]


def _guards_intact(buf, m, n, sentinel=512.0):
    mask = torch.ones_like(buf, dtype=torch.bool)
    mask[4:4 + m, 16:16 + n] = False
    return bool((buf[mask].float() == sentinel).all())


@pytest.mark.parametrize("m,n,k", [(300, 320, 320), (1000, 520, 128), (129, 96, 64), (20001, 1280, 1280)])
@pytest.mark.parametrize("mode", ["plain16", "plain32", "residual32", "inplace32", "mixed", "gelu16"])
def test_linear_epilogues_write_only_their_output(K, cuda, m, n, k, mode):
    """every epilogue path (TMA store, TMA reduce, staged, row-wise; single CTA and CTA pair) against ragged M and N with
    a sentinel band around the output: nothing outside [m, n] may change, and the inside must be right"""
    dt = K.op_dtype
    a, w = rnd(m, k, dev=cuda, seed=1, dtype=dt), rnd(n, k, dev=cuda, seed=2, scale=k ** -0.5, dtype=dt)
    bias = rnd(n, dev=cuda, seed=3)
    res = rnd(m, n, dev=cuda, seed=4)
    ref = a.float() @ w.float().t() + bias
    odt = dt if mode in ("plain16", "mixed", "gelu16") else torch.float32
    buf, out = _guarded(m, n, odt, cuda)
    if mode == "inplace32":
        out.copy_(res)
        K.linear(a, w, bias, residual=out, out=out)
        ref = ref + res
    elif mode in ("residual32", "mixed"):
        K.linear(a, w, bias, residual=res, out=out)
        ref = ref + res
    elif mode == "gelu16":
        K.linear(a, w, bias, out=out, gelu=True)
        ref = F.gelu(ref)
    else:
        K.linear(a, w, bias, out=out)
    assert _guards_intact(buf, m, n)
    assert rel(out, ref) < (2e-5 if odt == torch.float32 else out_tol(dt))


@pytest.mark.parametrize("n_img,h,w,cin,cout", [(3, 12, 12, 64, 320), (2, 24, 24, 128, 640), (5, 6, 6, 64, 96)])
def test_conv_epilogue_writes_only_its_output(K, cuda, n_img, h, w, cin, cout):
    from b200sd import ops
    from b200sd.layers import pack_conv3x3
    dt = K.op_dtype
    x = rnd(n_img, h, w, cin, dev=cuda, seed=1, dtype=dt)
    wt = rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5)
    bias = rnd(cout, dev=cuda, seed=3)
    m = n_img * h * w
    buf, out = _guarded(m, cout, torch.float32, cuda)
    ops.gemm(x, n_img, h, w, cin, 9, pack_conv3x3(wt, dt), bias, False, None, 0, None, out, False, 0, K.gemm_ws, 0)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dt).float(), bias, padding=1).permute(0, 2, 3, 1).reshape(m, cout)
    assert _guards_intact(buf, m, cout)
    assert rel(out, ref) < 2e-5


@pytest.mark.parametrize("batch,heads,sq,skv", [(2, 2, 300, 300), (1, 3, 130, 77), (3, 1, 576, 576)])
def test_attention_writes_only_its_output(K, cuda, batch, heads, sq, skv):
    from b200sd import ops
    dt = K.op_dtype
    c = heads * 64
    q = rnd(batch * sq, c, dev=cuda, seed=1, dtype=dt)
    kv = rnd(batch * skv, 2 * c, dev=cuda, seed=2, dtype=dt)
    k, v = kv[:, :c], kv[:, c:]
    buf, out = _guarded(batch * sq, c, dt, cuda)
    ops.attention(q, k, v, out, batch, heads, sq, skv, 0.125)
    qf = q.float().view(batch, sq, heads, 64).transpose(1, 2)
    kf = k.float().view(batch, skv, heads, 64).transpose(1, 2)
    vf = v.float().view(batch, skv, heads, 64).transpose(1, 2)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * 0.125, -1) @ vf).transpose(1, 2).reshape(batch * sq, c)
    assert _guards_intact(buf, batch * sq, c)
    assert rel(out, ref) < 2 * out_tol(dt)


# ------------------------------------------------------------------------------------------------ GroupNorm statistics from the producer
def _sums_from_table(ts, groups):
    """[n, parts, C, 2] column sums -> per-(image, group) (sum, sum of squares), in float64"""
    n, parts, c, _ = ts.shape
    return ts.double().sum(1).view(n, groups, c // groups, 2).sum(2)


@pytest.mark.parametrize("n,h,w,cin,cout,mode", [(3, 24, 24, 64, 128, "rowbias16"), (2, 48, 48, 64, 320, "residual32"),
                                                 (2, 20, 36, 72, 96, "plain32"), (4, 96, 96, 64, 64, "rowbias16"),
                                                 (32, 48, 48, 128, 640, "residual32")])
def test_conv_epilogue_column_sums(K, cuda, n, h, w, cin, cout, mode):
    """the conv epilogue's by-product for the next GroupNorm: per 32-row slice of every tile, per output column, the sum and
    the sum of squares of what it stores -- checked against sums over the stored tensor, and the output itself must be
    the same bits as without the by-product"""
    from b200sd.layers import pack_conv3x3, tile_sums_of
    dt = K.op_dtype
    x = rnd(n, h, w, cin, dev=cuda, seed=1, dtype=dt)
    wp = pack_conv3x3(rnd(cout, cin, 3, 3, dev=cuda, seed=2, scale=(9 * cin) ** -0.5), dt)
    kw = {}
    if mode == "rowbias16":
        kw = dict(row_bias=rnd(n, cout, dev=cuda, seed=3), out_dtype=dt)
        bias = None
    elif mode == "residual32":
        kw = dict(residual=rnd(n, h, w, cout, dev=cuda, seed=4))
        bias = rnd(cout, dev=cuda, seed=3)
    else:
        kw = dict(out_dtype=torch.float32)
        bias = rnd(cout, dev=cuda, seed=3)
    out = K.conv3x3(x, wp, bias, gn_next=True, **kw)
    ts = tile_sums_of(out)
    assert ts is not None and ts.shape[0] == n and ts.shape[2:] == (cout, 2)
    plain = K.conv3x3(x, wp, bias, **kw)
    assert torch.equal(out, plain)
    got = _sums_from_table(ts, 32)
    o = out.double().view(n, h * w, 32, cout // 32)
    ref = torch.stack([o.sum((1, 3)), (o * o).sum((1, 3))], -1)
    # the table sums the fp32 values before their rounding to the output dtype
    assert rel(got, ref) < (1e-5 if out.dtype == torch.float32 else out_tol(dt))


@pytest.mark.parametrize("hw,c,inplace", [(2304, 320, True), (9216, 64, False), (256, 128, True)])
def test_linear_epilogue_column_sums(K, cuda, hw, c, inplace):
    """[M, K] problems (proj_out onto the trunk, the im2col'd convs): rows_per_image tells the epilogue where images end"""
    from b200sd.layers import tile_sums_of
    dt = K.op_dtype
    n = 3
    a = rnd(n * hw, c, dev=cuda, seed=1, dtype=dt)
    w, bias = rnd(c, c, dev=cuda, seed=2, scale=c ** -0.5, dtype=dt), rnd(c, dev=cuda, seed=3)
    res = rnd(n * hw, c, dev=cuda, seed=4)
    if inplace:
        t = res.clone()
        out = K.linear(a, w, bias, residual=t, out=t, gn_rows_per_image=hw)
        t2 = res.clone()
        K.linear(a, w, bias, residual=t2, out=t2)  # the TMA-reduce path
        assert torch.equal(out, t2)
    else:
        out = K.linear(a, w, bias, gn_rows_per_image=hw)
    ts = tile_sums_of(out)
    assert ts is not None and ts.shape == (n, hw // 128 * 4, c, 2)
    o = out.double().view(n, hw, 32, c // 32)
    ref = torch.stack([o.sum((1, 3)), (o * o).sum((1, 3))], -1)
    assert rel(_sums_from_table(ts, 32), ref) < 1e-5
    assert tile_sums_of(K.linear(a, w, bias, gn_rows_per_image=hw - 64)) is None  # tiles would span images: not offered


@pytest.mark.parametrize("c1,c2,h", [(320, 0, 48), (640, 320, 24), (128, 64, 96)])
def test_groupnorm_from_producer_column_sums(K, cuda, c1, c2, h):
    """GroupNorm whose statistics come from the producers' epilogues (two producers for the up blocks' channel concat, the
    groups straddling the seam) against torch's GroupNorm on the same tensors, and against the two-pass kernel"""
    from b200sd.layers import pack_conv3x3, tile_sums_of
    dt = K.op_dtype
    n = 2
    mk = lambda c, seed: K.conv3x3(rnd(n, h, h, 64, dev=cuda, seed=seed, dtype=dt),
                                   pack_conv3x3(rnd(c, 64, 3, 3, dev=cuda, seed=seed + 1, scale=1 / 24), dt),
                                   rnd(c, dev=cuda, seed=seed + 2) * 0.5, gn_next=True)
    x1 = mk(c1, 10)
    x2 = mk(c2, 20) if c2 else None
    assert tile_sums_of(x1) is not None and (x2 is None or tile_sums_of(x2) is not None)
    c = c1 + c2
    gamma, beta = rnd(c, dev=cuda, seed=3) * 0.1 + 1, rnd(c, dev=cuda, seed=4) * 0.1
    y, raw = K.groupnorm(x1, x2, gamma, beta, 1e-5, True, want_raw=True)
    cat = torch.cat([x1, x2], -1) if c2 else x1
    ref = F.silu(F.group_norm(cat.float().permute(0, 3, 1, 2), 32, gamma, beta, 1e-5)).permute(0, 2, 3, 1)
    assert rel(y, ref) < out_tol(dt) and torch.equal(raw, cat.to(dt))
    K.fuse_gn_stats = False
    try:
        y2 = K.groupnorm(x1, x2, gamma, beta, 1e-5, True)
    finally:
        K.fuse_gn_stats = True
    assert rel(y, y2) < out_tol(dt)


def test_groupnorm_split_phases(K, cuda):
    """statistics-only and apply-only calls (the two halves a multi-GPU row split all-reduces between): together they equal
    the one-call GroupNorm; with the image cut in two row bands and the sums added, they equal GroupNorm of the whole"""
    from b200sd import ops
    dt = K.op_dtype
    n, h, w, c = 2, 32, 48, 320
    x = rnd(n, h, w, c, dev=cuda, seed=1) * 1.5 + 0.3
    gamma, beta = rnd(c, dev=cuda, seed=3) * 0.1 + 1, rnd(c, dev=cuda, seed=4) * 0.1
    ref = K.groupnorm(x, None, gamma, beta, 1e-5, True)
    sums = torch.empty(n, 32, 2, device=cuda)
    ops.groupnorm(x, None, n, h * w, 32, 1e-5, None, None, False, None, None, K.gn_workspace(n, h * w, 32), None, 0, sums)
    y = torch.empty_like(ref)
    ops.groupnorm(x, None, n, h * w, 32, 1e-5, gamma, beta, True, y, None, None, sums, 0, None)
    assert rel(y, ref) < 0.25 * out_tol(dt)  # same sums up to their summation order (a few flipped 16-bit roundings)
    top, bot = x[:, :20].contiguous(), x[:, 20:].contiguous()
    s1, s2 = torch.empty_like(sums), torch.empty_like(sums)
    ops.groupnorm(top, None, n, 20 * w, 32, 1e-5, None, None, False, None, None, K.gn_workspace(n, 20 * w, 32), None, 0, s1)
    ops.groupnorm(bot, None, n, 12 * w, 32, 1e-5, None, None, False, None, None, K.gn_workspace(n, 12 * w, 32), None, 0, s2)
    tot = s1 + s2
    yt, yb = torch.empty(n, 20, w, c, device=cuda, dtype=dt), torch.empty(n, 12, w, c, device=cuda, dtype=dt)
    ops.groupnorm(top, None, n, 20 * w, 32, 1e-5, gamma, beta, True, yt, None, None, tot, h * w, None)
    ops.groupnorm(bot, None, n, 12 * w, 32, 1e-5, gamma, beta, True, yb, None, None, tot, h * w, None)
    assert rel(torch.cat([yt, yb], 1), ref) < out_tol(dt)
