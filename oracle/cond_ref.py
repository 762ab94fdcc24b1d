"""CPU oracle for the conditioning-prefix producer (T3CondEnc + Perceiver) —
TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates, in plain torch fp32 (reference file:line):
  models/t3/modules/cond_enc.py:84-85     cond_spkr = spkr_enc(speaker_emb.view(256))
  models/t3/modules/cond_enc.py:96-101    cond_prompt_speech_emb -> perceiver
  models/t3/modules/cond_enc.py:104-110   emotion_adv_fc(emotion_adv)
  models/t3/modules/cond_enc.py:117-123   cat(spkr, clap(empty), prompt, emotion) -> [34, 1024]
  models/t3/modules/perceiver.py:150-169  AttentionBlock2: LayerNorm both inputs with ONE norm,
                                          q/k/v Linear, 4-head SDPA (d = 256, scale 1/16), proj_out, + x1
  models/t3/modules/perceiver.py:203-214  Perceiver: attn(query, h) then attn(pre_att, pre_att), same block
  tts.py:238                              prompt rows = speech_emb(tokens) + speech_pos_emb(arange(n))
  tts.py:249-257                          update_exaggeration: row 33 = emotion_adv_fc(exaggeration)

PARITY STATUS: PINNED.  The reference's own modules (cond_enc.py, perceiver.py, t3_config.py) import
and run in this container; oracle/make_golden.py runs them on seeded weights/inputs and commits the
output as tests/golden/cond_enc.npz, which tests/test_oracle_pinning.py holds this file to.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

N_QUERY, DIM, HEADS, SPK = 32, 1024, 4, 256

KEYS = {  # name -> shape, as T3CondEnc(T3Config()).state_dict()
    "spkr_enc.weight": (DIM, SPK), "spkr_enc.bias": (DIM,), "emotion_adv_fc.weight": (DIM, 1),
    "perceiver.pre_attention_query": (1, N_QUERY, DIM),
    "perceiver.attn.norm.weight": (DIM,), "perceiver.attn.norm.bias": (DIM,),
    "perceiver.attn.to_q.weight": (DIM, DIM), "perceiver.attn.to_q.bias": (DIM,),
    "perceiver.attn.to_k.weight": (DIM, DIM), "perceiver.attn.to_k.bias": (DIM,),
    "perceiver.attn.to_v.weight": (DIM, DIM), "perceiver.attn.to_v.bias": (DIM,),
    "perceiver.attn.proj_out.weight": (DIM, DIM), "perceiver.attn.proj_out.bias": (DIM,),
}


def random_state_dict(seed: int) -> dict:
    """Seeded fp32 weights in a fixed order (so a fixture only needs the seed).  Scales are chosen so
    that every stage matters: ~unit-variance activations, O(1) attention logits."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shape in KEYS.items():
        if k.endswith("norm.weight"):
            sd[k] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif k.endswith(".bias"):
            sd[k] = 0.05 * torch.randn(shape, generator=g)
        elif k == "perceiver.pre_attention_query":
            v = math.sqrt(3.0) * math.sqrt(2.0 / (N_QUERY + N_QUERY))          # perceiver.py:196-199
            sd[k] = (torch.rand(shape, generator=g) * 2 - 1) * v
        else:
            sd[k] = torch.randn(shape, generator=g) / math.sqrt(shape[-1])
    return sd


def random_inputs(seed: int, n_prompt: int = 150):
    g = torch.Generator().manual_seed(seed)
    speech_emb = torch.randn(8194, DIM, generator=g) * 0.5
    speech_pos = torch.randn(n_prompt + 8, DIM, generator=g) * 0.5
    speaker = torch.randn(SPK, generator=g)
    tokens = torch.randint(0, 6561, (n_prompt,), generator=g)
    return speech_emb, speech_pos, speaker, tokens


def _attn_block(sd: dict, x1: torch.Tensor, x2: torch.Tensor) -> torch.Tensor:
    p = "perceiver.attn."
    ln = lambda x: F.layer_norm(x, (DIM,), sd[p + "norm.weight"], sd[p + "norm.bias"], 1e-5)  # noqa: E731
    q = F.linear(ln(x1), sd[p + "to_q.weight"], sd[p + "to_q.bias"])
    k = F.linear(ln(x2), sd[p + "to_k.weight"], sd[p + "to_k.bias"])
    v = F.linear(ln(x2), sd[p + "to_v.weight"], sd[p + "to_v.bias"])
    split = lambda t: t.view(t.shape[0], HEADS, DIM // HEADS).permute(1, 0, 2)  # noqa: E731
    s = torch.einsum("hqd,hkd->hqk", split(q), split(k)) * (DIM // HEADS) ** -0.5
    o = torch.einsum("hqk,hkd->hqd", torch.softmax(s, dim=-1), split(v))
    o = o.permute(1, 0, 2).reshape(x1.shape[0], DIM)
    return x1 + F.linear(o, sd[p + "proj_out.weight"], sd[p + "proj_out.bias"])


def cond_encode(sd: dict, speaker_emb: torch.Tensor, prompt_tokens: torch.Tensor, speech_emb: torch.Tensor,
                speech_pos: torch.Tensor, emotion: float = 0.5) -> torch.Tensor:
    """fp32 [34, 1024] conditioning prefix."""
    n = prompt_tokens.shape[-1]
    h = speech_emb[prompt_tokens.view(-1)] + speech_pos[:n]
    spk = F.linear(speaker_emb.view(SPK), sd["spkr_enc.weight"], sd["spkr_enc.bias"]).unsqueeze(0)
    query = sd["perceiver.pre_attention_query"][0]
    pre = _attn_block(sd, query, h)
    body = _attn_block(sd, pre, pre)
    emo = F.linear(torch.full((1, 1), float(emotion)), sd["emotion_adv_fc.weight"])
    return torch.cat((spk, body, emo), dim=0)


def emotion_row(sd: dict, emotion: float) -> torch.Tensor:
    return F.linear(torch.full((1, 1), float(emotion)), sd["emotion_adv_fc.weight"])[0]
