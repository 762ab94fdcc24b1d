"""Synthetic (random-init) T3 checkpoints and prompts of the benchmark shapes — there is no
network for real weights.  Key layout = t3_cfg.safetensors; values N(0, 0.02²)
(initializer_range, t3-model/config.json:11), RMSNorm weights 1."""
from __future__ import annotations

import torch

BF16 = torch.bfloat16


def random_state_dict(n_layers: int = 30, *, device="cuda", seed: int = 0, text_vocab: int = 704,
                      head_scale: float = 1.0, std: float = 0.02) -> dict:
    g = torch.Generator(device=device).manual_seed(seed)
    n = lambda *s, sc=std: (torch.randn(*s, generator=g, device=device) * sc).to(BF16)  # noqa: E731
    one = lambda: torch.ones(1024, dtype=BF16, device=device)  # noqa: E731
    sd = {}
    for i in range(n_layers):
        p = f"tfmr.layers.{i}."
        for k in "qkvo":
            sd[p + f"self_attn.{k}_proj.weight"] = n(1024, 1024)
        sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"] = n(4096, 1024), n(4096, 1024)
        sd[p + "mlp.down_proj.weight"] = n(1024, 4096)
        sd[p + "input_layernorm.weight"], sd[p + "post_attention_layernorm.weight"] = one(), one()
    sd["tfmr.norm.weight"] = one()
    sd["speech_emb.weight"], sd["speech_head.weight"] = n(8194, 1024), n(8194, 1024, sc=std * head_scale)
    sd["text_emb.weight"] = n(text_vocab, 1024)
    sd["text_pos_emb.emb.weight"], sd["speech_pos_emb.emb.weight"] = n(2050, 1024), n(4100, 1024)
    return sd


def random_prompt(n_text: int, seed: int, text_vocab: int = 704) -> list[int]:
    """[START]=255, n_text-2 ids in [3, vocab) avoiding 255 and the markers 695..697, [STOP]=0."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, min(text_vocab, 695), (n_text - 2,), generator=g)
    ids[ids == 255] = 254
    return [255] + ids.tolist() + [0]


def random_cond_emb(seed: int, pinned: bool = False) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(34, 1024, generator=g) * 0.02
    return t.pin_memory() if pinned else t
