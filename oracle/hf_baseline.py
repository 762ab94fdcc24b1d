"""CPU baseline: the original-Chatterbox-style T3 inference restated on HF transformers'
LlamaModel — TEST/BENCH INFRASTRUCTURE, NOT PRODUCT CODE (only bench.py's cpu_baseline /
--impl reference legs call this).

The original source (resemble-ai/chatterbox @ 70f5987d, an empty submodule in the reference:
.SUBMODULES.json:9-15) is absent, so this follows SURVEY.md App. D: batch rows = [cond; uncond]
for every pair, HF ``past_key_values``, per-step CFG merge (t3.py:732) and a per-step Python
sampler (repetition penalty → temperature → min-p → top-p → multinomial).  `speech_pos` toggles
the original's learned speech positional embedding during decode (the vLLM reference omits it,
t3.py:766).  "parity unpinned": this is a timed baseline, not a correctness oracle."""
from __future__ import annotations

import os
import time

import torch


def build_model(n_layers: int = 30, dtype=torch.float32, seed: int = 0):
    from transformers import LlamaConfig, LlamaModel
    cfg = LlamaConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=n_layers,
                      num_attention_heads=16, num_key_value_heads=16, head_dim=64, rms_norm_eps=1e-5,
                      rope_theta=500000.0, vocab_size=8, max_position_embeddings=131072,
                      attention_bias=False, mlp_bias=False, attn_implementation="sdpa",
                      rope_scaling={"factor": 8.0, "high_freq_factor": 4.0, "low_freq_factor": 1.0,
                                    "original_max_position_embeddings": 8192, "rope_type": "llama3"})
    torch.manual_seed(seed)
    m = LlamaModel(cfg).to(dtype).eval()
    g = torch.Generator().manual_seed(seed)
    emb = (torch.randn(8194, 1024, generator=g) * 0.02).to(dtype)
    head = (torch.randn(8194, 1024, generator=g) * 0.02).to(dtype)
    pos = (torch.randn(4100, 1024, generator=g) * 0.02).to(dtype)
    return m, emb, head, pos


def _sample(logits, seen, temperature, top_p, min_p, rep):
    l = logits.float()
    l = torch.where(seen, torch.where(l > 0, l / rep, l * rep), l)
    if temperature < 1e-5:
        return l.argmax(-1)
    l = l / temperature
    p = torch.softmax(l, -1)
    l = l.masked_fill(p < p.amax(-1, keepdim=True) * min_p, float("-inf"))
    ls, idx = l.sort(-1, descending=False)
    cs = torch.softmax(ls, -1).cumsum(-1)
    mask = cs <= 1 - top_p
    mask[:, -1] = False
    l = l.scatter(-1, idx, ls.masked_fill(mask, float("-inf")))
    return torch.multinomial(torch.softmax(l, -1), 1)[:, 0]


@torch.no_grad()
def run(pairs: int, prefix_len: int, n_steps: int, *, n_layers: int = 30, dtype=torch.float32,
        cfg_scale: float = 0.5, temperature: float = 0.5, top_p: float = 0.5, min_p: float = 0.0,
        rep: float = 2.0, speech_pos: bool = False, threads: int | None = None, model=None, warmup: int = 0,
        device=None):
    """Prefill `pairs` CFG pairs with random prefix embeddings of `prefix_len` rows, then time
    `n_steps` decode steps.  Returns dict(prefill_s, decode_s, steps, tokens, cores)."""
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    m, emb, head, pos = model or build_model(n_layers, dtype)
    g = torch.Generator().manual_seed(1)
    x = (torch.randn(pairs, prefix_len, 1024, generator=g) * 0.02).to(dtype)
    cuda = device is not None and torch.device(device).type == "cuda"   # bench.py's gpu_standin leg: same loop on the GPU
    sync = torch.cuda.synchronize if cuda else (lambda: None)
    if cuda:
        x = x.to(device)
    xu = x.clone()
    xu[:, 34:-1] = 0                                         # uncond: text rows zeroed (t3.py:564)
    t0 = time.perf_counter()
    out = m(inputs_embeds=torch.cat((x, xu), 0), use_cache=True)
    past = out.past_key_values
    h = out.last_hidden_state[:, -1]
    sync()
    t_prefill = time.perf_counter() - t0
    seen = torch.zeros(pairs, 8194, dtype=torch.bool, device=x.device)
    t_dec = 0.0
    for s in range(warmup + n_steps):
        sync()
        t0 = time.perf_counter()
        lg = h @ head.T
        l = lg[:pairs] + cfg_scale * (lg[:pairs] - lg[pairs:])
        tok = _sample(l, seen, temperature, top_p, min_p, rep)
        seen[torch.arange(pairs, device=x.device), tok] = True
        e = emb[tok] + (pos[s + 1] if speech_pos else 0)
        out = m(inputs_embeds=torch.cat((e, e), 0)[:, None], past_key_values=past, use_cache=True)
        past, h = out.past_key_values, out.last_hidden_state[:, -1]
        if s >= warmup:
            sync()
            t_dec += time.perf_counter() - t0
    return dict(prefill_s=t_prefill, decode_s=t_dec, steps=n_steps, tokens=pairs * n_steps, cores=threads)
