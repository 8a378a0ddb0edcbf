"""Generates tests/golden/clip_tiny.safetensors from the REAL reference implementation of the text encoder:
`transformers.CLIPTextModel` (the class diffusers instantiates for `pipe.text_encoder`; transformers 5.5.0 is in this
image, there is no network to fetch the SD2.1 checkpoint, so the weights are seeded).  Run here (CPU):
    python tests/golden/make_clip_golden.py
The fixture holds the seeded weights' seed, the token ids and the model's last_hidden_state for a 2-layer, 128-wide,
2-head configuration; tests/test_oracle.py checks oracle/clip_text.py against it, and the GPU tests check the CUDA path
against the oracle on the same inputs."""
import sys
from pathlib import Path

import torch
from safetensors.torch import save_file

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import clip_text as C  # noqa: E402


def main():
    import transformers
    from transformers import CLIPTextConfig, CLIPTextModel
    cfg = C.tiny_clip_config()
    hf_cfg = CLIPTextConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                            intermediate_size=cfg.intermediate_size, num_hidden_layers=cfg.num_hidden_layers,
                            num_attention_heads=cfg.num_attention_heads,
                            max_position_embeddings=cfg.max_position_embeddings, hidden_act="gelu",
                            layer_norm_eps=cfg.layer_norm_eps, projection_dim=64)
    model = CLIPTextModel(hf_cfg).eval()
    sd = C.seeded_state_dict(cfg, seed=0)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("position_ids" in k for k in missing), missing  # buffers only
    ids = C.seeded_ids(cfg, batch=3, seed=5)
    ids[1, 40:] = 0  # a padded prompt: pad id 0 to 77 tokens, as the SD2.1 tokenizer does
    with torch.no_grad():
        out = model(input_ids=ids)[0]
    save_file({"input_ids": ids, "last_hidden_state": out.contiguous(),
               "weights_seed": torch.tensor([0]), "transformers_version": torch.tensor(
                   [int(x) for x in transformers.__version__.split(".")[:2]])},
              str(Path(__file__).with_name("clip_tiny.safetensors")))
    print("wrote clip_tiny.safetensors", tuple(out.shape), "transformers", transformers.__version__)


if __name__ == "__main__":
    main()
