"""How many FREE-RUNNING greedy steps does the engine emit exactly the bf16 oracle's tokens?  (north_star: "identical
greedy-decoded token sequences over the first N steps".)  Sharp model (speech head x 16), seeds = different weights +
prompts; prints per-seed horizons (first step at which the free-running streams differ, or the run length) for a
4-layer model over 32 seeds and the 30-layer model over 4 seeds, the minimum of each, and — at every divergence — how far
below the oracle's top logit the engine's token sat, in bf16 ulps of that logit: DESIGN §4's account of N."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from oracle import sampler_ref
from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights


def horizon(n_layers, seed, n_steps):
    cfg = T3ModelConfig(n_layers=n_layers)
    w = T3Weights.random(cfg, seed=seed, head_scale=16.0, norm_jitter=0.1)
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=1, max_model_len=256)
    g = torch.Generator().manual_seed(1000 + seed)
    text = [255] + torch.randint(3, 600, (10 + seed % 30,), generator=g).tolist() + [0]
    cond = torch.randn(34, 1024, generator=g) * 0.02
    sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n_steps, ignore_eos=True)
    got = eng.generate([Request(text, cond, sp, uid=7)])[0]
    eng.close()
    o = T3Oracle(w, "bf16")
    pair, seen = o.new_pair(), np.zeros(8194, bool)
    lg = o.prefill_pair(pair, cond, text, 0.5).numpy()
    for s in range(n_steps):
        t = sampler_ref.sample_row(lg, seen, repetition_penalty=2.0, temperature=0.0)
        if t + 1000 != got[s]:
            l = sampler_ref.apply_repetition_penalty(lg, seen, 2.0)
            gaps.append(round(float((l[t] - l[got[s] - 1000]) / (2.0 ** -7 * abs(l[t]))), 3))   # in bf16 ulps of the top logit
            return s
        seen[t] = True
        if s + 1 < n_steps:
            lg = o.decode_pairs([pair], [t], 0.5)[0].numpy()
    return n_steps


gaps = []
res = {"steps_run": 48, "layers4": [horizon(4, s, 48) for s in range(32)], "layers30": [horizon(30, s, 48) for s in range(4)]}
res["min_layers4"], res["min_layers30"] = min(res["layers4"]), min(res["layers30"])
res["gap_at_divergence_in_bf16_ulps_of_top_logit"] = gaps          # every divergence is a near-tie of the oracle's own logits
res["max_gap_ulps"] = max(gaps) if gaps else 0.0
print(json.dumps(res))
