"""TEST INFRASTRUCTURE (not collected by pytest): how much of the UNet's single-step error against the fp32 oracle is
inherent to 16-bit tensor-core operands?  Rounds operands INSIDE the fp32 oracle (no kernels involved) and prints the
rel-L2 against the unrounded oracle -- the numbers quoted in DESIGN.md section 2.  CPU only, a few minutes at 512^2:

    python tests/precision_probe.py [latent_size=64]
"""
import sys
from pathlib import Path

import torch
import torch.nn as nn

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import sd21 as O  # noqa: E402

hw = int(sys.argv[1]) if len(sys.argv) > 1 else 64
torch.set_num_threads(torch.get_num_threads())


def rnd(t, dt):
    return t.to(dt).float()


def run(weights=None, inputs=None, attn=None):
    """weights / inputs / attn: dtype the conv+linear weights, their inputs, the attention operands are rounded to"""
    unet = O.seeded_init_(O.UNet2DCondition(O.sd21_unet_config(sample_size=hw)), 0).eval()
    hooks = []
    for m in unet.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            if weights is not None:
                m.weight.data = rnd(m.weight.data, weights)
            if inputs is not None:
                hooks.append(m.register_forward_pre_hook(lambda _m, a: (rnd(a[0], inputs),) + tuple(a[1:])))
    if attn is not None:
        orig = O.Attention.forward

        def fwd(self, x, ctx=None):
            ctx_ = x if ctx is None else ctx
            b, s, c = x.shape
            d = c // self.heads
            q = rnd(self.to_q(x), attn).view(b, s, self.heads, d).transpose(1, 2)
            k = rnd(self.to_k(ctx_), attn).view(b, ctx_.shape[1], self.heads, d).transpose(1, 2)
            v = rnd(self.to_v(ctx_), attn).view(b, ctx_.shape[1], self.heads, d).transpose(1, 2)
            p = rnd(torch.softmax(q @ k.transpose(-1, -2) * (d ** -0.5), dim=-1), attn)
            return self.to_out[0]((p @ v).transpose(1, 2).reshape(b, s, c))

        O.Attention.forward = fwd
    lat, ctx = O.seeded_inputs(1, hw, 1024)
    with torch.no_grad():
        y = unet(torch.cat([lat, lat]), 961.0, ctx)
    for h in hooks:
        h.remove()
    if attn is not None:
        O.Attention.forward = orig
    return y


ref = run()
bf, fh = torch.bfloat16, torch.float16
for name, kw in [("weights -> bf16 only", dict(weights=bf)), ("conv/linear inputs -> bf16 only", dict(inputs=bf)),
                 ("attention QK/PV operands -> bf16 only", dict(attn=bf)), ("all bf16", dict(weights=bf, inputs=bf, attn=bf)),
                 ("all fp16", dict(weights=fh, inputs=fh, attn=fh))]:
    print(f"{name:42s} rel-L2 {O.rel_l2(run(**kw), ref):.3e}", flush=True)
