"""Which roundings dominate the bf16-vs-fp32 UNet error?  Oracle on GPU fp32 with emulated operand rounding."""
import sys, torch, torch.nn.functional as F
sys.path.insert(0, '.')
from oracle import sd21 as O
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False
dev = 'cuda'
hw = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cfg = O.sd21_unet_config(sample_size=hw)
u = O.seeded_init_(O.UNet2DCondition(cfg), 0).eval().to(dev)
lat, ctx = O.seeded_inputs(1, hw, 1024)
x2 = torch.cat([lat, lat]).to(dev); ctx = ctx.to(dev)
with torch.no_grad():
    ref = u(x2, 961.0, ctx)
_conv, _lin, _mm = F.conv2d, F.linear, torch.Tensor.__matmul__
def run(act_dt, w_dt, mm_dt):
    r = lambda t, dt: t if dt is None else t.to(dt).float()
    def conv(x, w, b=None, *a, **k): return _conv(r(x, act_dt), r(w, w_dt), b, *a, **k)
    def lin(x, w, b=None): return _lin(r(x, act_dt), r(w, w_dt), b)
    def mm(a, b): return _mm(r(a, mm_dt), r(b, mm_dt))
    F.conv2d, F.linear, torch.Tensor.__matmul__ = conv, lin, mm
    try:
        with torch.no_grad():
            out = u(x2, 961.0, ctx)
    finally:
        F.conv2d, F.linear, torch.Tensor.__matmul__ = _conv, _lin, _mm
    return O.rel_l2(out, ref)
bf, fh = torch.bfloat16, torch.float16
for name, cfgs in [("weights bf16 only", (None, bf, None)), ("acts bf16 only", (bf, None, None)),
                   ("attn qk/pv operands bf16 only", (None, None, bf)), ("all bf16", (bf, bf, bf)),
                   ("all fp16", (fh, fh, fh)), ("acts fp16 + weights bf16 + attn fp16", (fh, bf, fh)),
                   ("acts bf16 + weights fp16", (bf, fh, bf))]:
    print(f"{name:45s} rel L2 {run(*cfgs):.3e}", flush=True)
