"""CPU: pins the oracle (oracle/sd21.py).  The reference has no tests or golden vectors for this path
(parity unpinned), so the pins are: exact parameter counts, diffusers key names, closed-form scheduler
identities, and self-generated regression vectors under tests/golden (made by tests/golden/make_golden.py)."""
import math
from pathlib import Path

import pytest
import torch

from oracle import sd21 as O

GOLD = Path(__file__).parent / "golden"


def test_unet_param_count_matches_sd21():
    u = O.UNet2DCondition(O.sd21_unet_config())
    assert sum(p.numel() for p in u.parameters()) == 865_910_724


def test_vae_decoder_param_count_matches_sd21():
    v = O.VAEDecoder()
    # AutoencoderKL total 83,653,863 = encoder 34,163,592 + quant_conv 72 + post_quant_conv 20 + decoder 49,490,179
    assert sum(p.numel() for p in v.parameters()) == 49_490_179 + 20


def test_vae_encoder_param_count_and_keys_match_sd21():
    """AutoencoderKL.encode restatement: the published sizes (encoder 34,163,592, quant_conv 72; with the decoder half
    the AutoencoderKL total 83,653,863) and diffusers' key names"""
    e = O.VAEEncoder()
    assert sum(p.numel() for p in e.encoder.parameters()) == 34_163_592
    assert sum(p.numel() for p in e.quant_conv.parameters()) == 72
    v = O.VAEDecoder()
    assert sum(p.numel() for m in (e, v) for p in m.parameters()) == 83_653_863
    keys = set(e.state_dict().keys())
    for k in ["encoder.conv_in.weight", "encoder.down_blocks.0.resnets.1.conv2.bias",
              "encoder.down_blocks.1.resnets.0.conv_shortcut.weight", "encoder.down_blocks.2.downsamplers.0.conv.weight",
              "encoder.mid_block.attentions.0.to_q.bias", "encoder.mid_block.attentions.0.group_norm.weight",
              "encoder.mid_block.resnets.1.norm2.weight", "encoder.conv_norm_out.bias", "encoder.conv_out.weight",
              "quant_conv.bias"]:
        assert k in keys, k
    assert "encoder.down_blocks.3.downsamplers.0.conv.weight" not in keys   # the last block does not downsample


def test_vae_encoder_downsample_padding_and_gaussian_sample():
    """Downsample2D(padding=0) pads (0,1,0,1): an impulse in the LAST row / column must reach the output, one in a
    virtual row -1 cannot; the distribution's sample follows mean + exp(0.5 * clamp(logvar)) * noise"""
    d = O.EncoderDownsample2D(1)
    with torch.no_grad():
        d.conv.weight.fill_(1.0)
        d.conv.bias.zero_()
        x = torch.zeros(1, 1, 6, 6)
        x[0, 0, 5, 5] = 1.0
        y = d(x)
        assert y.shape == (1, 1, 3, 3) and float(y[0, 0, 2, 2]) == 1.0 and float(y.sum()) == 1.0
        x = torch.zeros(1, 1, 6, 6)
        x[0, 0, 0, 0] = 1.0
        assert float(d(x)[0, 0, 0, 0]) == 1.0 and float(d(x).sum()) == 1.0   # window 0 starts AT row 0: nothing before it
    m = torch.tensor([1.0, -2.0, 0.0, 40.0]).view(1, 4, 1, 1)   # mean (1, -2), logvar (0, 40 -> clamped to 20)
    n = torch.tensor([0.5, 1.0]).view(1, 2, 1, 1)
    z = O.gaussian_sample(m, n)
    assert torch.allclose(z.flatten(), torch.tensor([1.0 + 0.5, -2.0 + float(torch.exp(torch.tensor(10.0)))]))
    assert torch.equal(O.gaussian_sample(m), m[:, :2])


def test_state_dict_keys_follow_diffusers_scheme():
    keys = set(O.UNet2DCondition(O.tiny_unet_config()).state_dict().keys())
    for k in ["conv_in.weight", "time_embedding.linear_1.weight", "time_embedding.linear_2.bias",
              "down_blocks.0.resnets.1.time_emb_proj.weight", "down_blocks.0.attentions.0.proj_in.weight",
              "down_blocks.0.attentions.0.transformer_blocks.0.attn1.to_q.weight",
              "down_blocks.0.attentions.0.transformer_blocks.0.attn2.to_out.0.bias",
              "down_blocks.1.attentions.1.transformer_blocks.0.ff.net.0.proj.weight",
              "down_blocks.1.attentions.1.transformer_blocks.0.ff.net.2.bias",
              "down_blocks.2.downsamplers.0.conv.weight", "mid_block.attentions.0.norm.weight",
              "mid_block.resnets.1.conv2.bias", "up_blocks.0.resnets.2.conv_shortcut.weight",
              "up_blocks.0.upsamplers.0.conv.weight", "up_blocks.3.attentions.2.proj_out.bias",
              "conv_norm_out.weight", "conv_out.bias"]:
        assert k in keys, k
    assert not any("to_q.bias" in k for k in keys)
    assert "down_blocks.3.attentions.0.norm.weight" not in keys  # DownBlock2D has no attention
    vkeys = set(O.VAEDecoder(O.tiny_vae_config()).state_dict().keys())
    for k in ["post_quant_conv.weight", "decoder.conv_in.bias", "decoder.mid_block.attentions.0.group_norm.weight",
              "decoder.mid_block.attentions.0.to_q.bias", "decoder.mid_block.attentions.0.to_out.0.weight",
              "decoder.up_blocks.2.resnets.0.conv_shortcut.weight", "decoder.up_blocks.0.upsamplers.0.conv.weight",
              "decoder.conv_norm_out.bias", "decoder.conv_out.weight"]:
        assert k in vkeys, k
    assert "decoder.up_blocks.3.upsamplers.0.conv.weight" not in vkeys


def test_up_block_input_channels():
    u = O.UNet2DCondition(O.sd21_unet_config())
    cins = [[r.conv1.weight.shape[1] for r in b.resnets] for b in u.up_blocks]
    assert cins == [[2560, 2560, 2560], [2560, 2560, 1920], [1920, 1280, 960], [960, 640, 640]]


def test_ddim_timesteps_leading():
    s = O.DDIM()
    s.set_timesteps(25)
    assert s.timesteps[:3] == [961, 921, 881] and s.timesteps[-1] == 1 and len(s.timesteps) == 25
    s.set_timesteps(20)
    assert s.timesteps[0] == 951 and s.timesteps[-1] == 1


@pytest.mark.parametrize("kind", ["ddim", "dpm++2m", "euler"])
@pytest.mark.parametrize("pred", ["v_prediction", "epsilon"])
def test_linear_form_equals_direct_step(kind, pred):
    s = O.SCHEDULERS[kind](pred)
    s.set_timesteps(12)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 4, 8, 8, generator=g, dtype=torch.float64) * s.init_noise_sigma
    x0_prev = None
    for i in range(12):
        m = torch.randn(2, 4, 8, 8, generator=g, dtype=torch.float64)
        direct, x0_d = s.step(m, i, x, x0_prev)
        lin, x0_l = O.linear_step(x, m, x0_prev, s.coef(i))
        assert torch.allclose(direct, lin, rtol=1e-9, atol=1e-9), (kind, pred, i)
        assert torch.allclose(x0_d, x0_l, rtol=1e-9, atol=1e-9)
        x, x0_prev = direct, x0_d


def test_ddim_identities():
    s = O.DDIM("v_prediction")
    s.set_timesteps(10)
    e = O.DDIM("epsilon")
    e.set_timesteps(10)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(64, generator=g, dtype=torch.float64)
    eps = torch.randn(64, generator=g, dtype=torch.float64)
    for i in range(10):
        a = s.alphas_cumprod[s.timesteps[i]]
        x0 = (x - (1 - a).sqrt() * eps) / a.sqrt()
        v = a.sqrt() * eps - (1 - a).sqrt() * x0          # v-parameterisation of the same (x0, eps)
        out_v, x0_v = s.step(v, i, x)
        out_e, x0_e = e.step(eps, i, x)
        assert torch.allclose(out_v, out_e, atol=1e-10) and torch.allclose(x0_v, x0_e, atol=1e-10)
        assert torch.allclose(x0_v, x0, atol=1e-10)
    # last step of the 'leading' schedule lands on alpha_bar[0] (set_alpha_to_one = False)
    c = e.coef(9)
    a, a0 = float(e.alphas_cumprod[1]), float(e.alphas_cumprod[0])
    assert math.isclose(c.cx, math.sqrt(a0 / a), rel_tol=1e-12)


def test_dpm_last_step_returns_x0_and_euler_reaches_sigma_zero():
    d = O.DPMSolverPP2M("epsilon")
    d.set_timesteps(8)
    c = d.coef(7)
    assert c.cx == c.ax and c.cm == c.am and c.cp == 0.0
    e = O.Euler("epsilon")
    e.set_timesteps(8)
    c = e.coef(7)  # sigma_next = 0 -> x' = x0
    assert math.isclose(c.cx, c.ax, abs_tol=1e-12) and math.isclose(c.cm, c.am, rel_tol=1e-12)
    assert math.isclose(e.init_noise_sigma, math.sqrt(e.sigmas[0] ** 2 + 1))


def test_cfg_combine():
    eps2 = torch.arange(16.0).view(4, 4)
    m = O.cfg_combine(eps2, 7.5)
    assert torch.equal(m, eps2[:2] + 7.5 * (eps2[2:] - eps2[:2]))


def test_timestep_embedding_layout():
    e = O.timestep_embedding(torch.tensor([0.0, 961.0]), 320)
    assert e.shape == (2, 320)
    assert torch.allclose(e[0, :160], torch.ones(160)) and torch.allclose(e[0, 160:], torch.zeros(160))  # cos first
    assert math.isclose(float(e[1, 0]), math.cos(961.0), abs_tol=1e-4)


def test_seeded_init_is_deterministic_and_o1():
    a = O.seeded_init_(O.UNet2DCondition(O.tiny_unet_config()), 0)
    b = O.seeded_init_(O.UNet2DCondition(O.tiny_unet_config()), 0)
    for (k, p), (_, q) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(p, q), k
    lat, ctx = O.seeded_inputs(1, 16, 128)
    with torch.no_grad():
        out = a(torch.cat([lat, lat]), 500.0, ctx)
    assert 0.1 < float(out.std()) < 10.0


def test_golden_vectors():
    """regression pin of the oracle itself (self-generated; see tests/golden/make_golden.py)"""
    from safetensors.torch import load_file
    gold = load_file(str(GOLD / "oracle_tiny.safetensors"))
    unet = O.seeded_init_(O.UNet2DCondition(O.tiny_unet_config()), 0)
    vae = O.seeded_init_(O.VAEDecoder(O.tiny_vae_config()), 1)
    lat, ctx = O.seeded_inputs(1, 16, 128)
    with torch.no_grad():
        eps = unet(torch.cat([lat, lat]), 961.0, ctx)
        final = O.denoise_loop(unet, lat, ctx, 4, 7.5, "ddim")
        img = O.decode_to_image(vae, final)
    assert O.rel_l2(eps, gold["unet_eps_t961"]) < 1e-5
    assert O.rel_l2(final, gold["ddim4_final_latents"]) < 1e-4
    assert O.rel_l2(img, gold["ddim4_image01"]) < 1e-4
    assert torch.equal(O.to_uint8(torch.tensor([0.0, 0.5, 1.0, 0.49803922])), torch.tensor([0, 128, 255, 127], dtype=torch.uint8))


# ------------------------------------------------------------------------------------------------ text encoder (pinned)
def test_clip_text_oracle_matches_transformers_golden():
    """oracle/clip_text.py against the output of the real transformers.CLIPTextModel (tests/golden/make_clip_golden.py)"""
    from safetensors.torch import load_file
    from oracle import clip_text as C
    g = load_file(str(Path(__file__).parent / "golden" / "clip_tiny.safetensors"))
    cfg = C.tiny_clip_config()
    sd = C.seeded_state_dict(cfg, int(g["weights_seed"][0]))
    out = C.clip_text_forward(sd, cfg, g["input_ids"])
    ref = g["last_hidden_state"]
    assert out.shape == ref.shape == (3, 77, 128)
    assert ((out - ref).norm() / ref.norm()).item() < 1e-5  # fp32 vs fp32, different op order only


def test_clip_text_oracle_against_live_transformers_if_present():
    """same check against transformers imported now (skipped where the package is absent)"""
    transformers = pytest.importorskip("transformers")
    from oracle import clip_text as C
    cfg = C.tiny_clip_config()
    hf = transformers.CLIPTextConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                                     intermediate_size=cfg.intermediate_size, num_hidden_layers=cfg.num_hidden_layers,
                                     num_attention_heads=cfg.num_attention_heads, max_position_embeddings=77,
                                     hidden_act="gelu", layer_norm_eps=cfg.layer_norm_eps, projection_dim=64)
    model = transformers.CLIPTextModel(hf).eval()
    sd = C.seeded_state_dict(cfg, 3)
    model.load_state_dict(sd, strict=False)
    ids = C.seeded_ids(cfg, 2, seed=9)[:, :33]  # a shorter sequence than the position table
    with torch.no_grad():
        ref = model(input_ids=ids)[0]
    out = C.clip_text_forward(sd, cfg, ids)
    assert ((out - ref).norm() / ref.norm()).item() < 1e-5


def test_clip_text_param_count_is_sd21s():
    from oracle import clip_text as C
    n = sum(torch.Size(s).numel() for s in C.param_shapes(C.CLIPTextConfig()).values())
    assert n == 340_387_840
