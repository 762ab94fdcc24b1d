#!/usr/bin/env python3
"""Prefill determinism probe: runs the same prefill several times and compares the final hidden states bitwise."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

layers = int(os.environ.get("L", "30")); B = int(os.environ.get("B", "64")); reps = int(os.environ.get("REPS", "4"))
eng = T3Engine(random_state_dict(layers, device="cuda", head_scale=16.0), device="cuda:0", max_pairs=B, max_model_len=256)
sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=2, ignore_eos=True)
reqs = [Request(random_prompt(165, seed=i), random_cond_emb(i), sp, uid=1 + i) for i in range(B)]
hs, toks = [], []
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    for k in range(reps):
        eng._init_sampler_state(reqs, P)
        eng.prefill(reqs, P)
        torch.cuda.synchronize()
        hs.append(eng.buffer("h", (2 * B, 1024), torch.bfloat16).clone())
        toks.append(eng.sb.last_tok[:B].clone())
    eng._release(held)
res = {"env": {k: v for k, v in os.environ.items() if k.startswith("CBX_")}, "L": layers, "B": B}
for k in range(1, reps):
    d = (hs[k].float() - hs[0].float()).abs()
    res[f"rep{k}"] = {"rows_differing": int((d.amax(1) > 0).sum()), "max_abs": float(d.max()), "tokens_differing": int((toks[k] != toks[0]).sum())}
print(json.dumps(res), flush=True)
