#!/usr/bin/env python3
"""Standalone determinism probe of cbx_paged_attn_prefill: identical inputs, repeated launches, bitwise compare."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops
torch.manual_seed(0)
S, P = int(os.environ.get("S", "32")), 200
nblk = -(-P // 16)
T = S * P
kv = (torch.randn(( (S * nblk + 1) * 65536) // 2, device="cuda") * 0.5).to(torch.bfloat16).view(torch.uint8)
bt = torch.zeros(S, 64, dtype=torch.int32, device="cuda")
perm = torch.randperm(S * nblk, device="cuda").to(torch.int32) + 1
bt[:, :nblk] = perm.view(S, nblk)
qkv = (torch.randn(T, 3072, device="cuda")).to(torch.bfloat16)
items = ops.prefill_items([i * P for i in range(S)], [P] * S, list(range(S))).cuda()
outs = []
for k in range(int(os.environ.get("REPS", "30"))):
    outs.append(ops.paged_attn_prefill(qkv, kv, bt, items).clone())
torch.cuda.synchronize()
bad = []
for k in range(1, len(outs)):
    ne = (outs[k] != outs[0])
    if ne.any():
        rows = sorted(set(ne.nonzero()[:, 0].tolist()))
        bad.append({"rep": k, "n": int(ne.sum()), "rows": rows[:20], "n_rows": len(rows),
                    "max_abs": float((outs[k].float() - outs[0].float()).abs().max())})
print(json.dumps({"S": S, "T": T, "reps": len(outs), "differing_reps": len(bad), "examples": bad[:4]}))
