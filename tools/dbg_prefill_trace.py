#!/usr/bin/env python3
"""Prefill determinism probe with the per-layer trace: first layer / rows / columns where two runs differ."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

layers = int(os.environ.get("L", "30")); B = int(os.environ.get("B", "16")); reps = int(os.environ.get("REPS", "6"))
eng = T3Engine(random_state_dict(layers, device="cuda", head_scale=16.0), device="cuda:0", max_pairs=B, max_model_len=256)
sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=2, ignore_eos=True)
reqs = [Request(random_prompt(165, seed=i), random_cond_emb(i), sp, uid=1 + i) for i in range(B)]
T = 2 * B * 200
assert T <= eng.max_rows
traces = []
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    for k in range(reps):
        tr = torch.zeros(layers, T, 1024, dtype=torch.bfloat16, device="cuda")
        eng.set_trace(tr)
        eng._init_sampler_state(reqs, P)
        eng.prefill(reqs, P)
        torch.cuda.synchronize()
        eng.set_trace(None)
        traces.append(tr)
    eng._release(held)
print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("CBX_")}, "L": layers, "B": B, "T": T}))
for k in range(1, reps):
    ne = (traces[k] != traces[0])
    per_layer = ne.flatten(1).sum(1).tolist()
    first = next((i for i, c in enumerate(per_layer) if c), None)
    info = {"rep": k, "first_layer": first, "per_layer_counts_head": per_layer[:6]}
    if first is not None:
        idx = ne[first].nonzero()
        rows, cols = idx[:, 0], idx[:, 1]
        info.update({"n": int(idx.shape[0]), "rows": sorted(set(rows.tolist()))[:12], "n_rows": len(set(rows.tolist())),
                     "col_min": int(cols.min()), "col_max": int(cols.max()), "n_cols": len(set(cols.tolist())),
                     "max_abs": float((traces[k][first].float() - traces[0][first].float()).abs().max())})
    print(json.dumps(info), flush=True)
