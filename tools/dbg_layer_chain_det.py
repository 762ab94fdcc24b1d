#!/usr/bin/env python3
"""One layer's [RMSNorm -> QKV GEMM + RoPE + KV append -> prefill attention -> O GEMM] as a PDL chain on fixed
inputs, repeated; reports which buffer (q, KV pages, attention output, residual) first differs bitwise."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops
from chatterbox_vllm_b200.engine import llama3_cos_sin
torch.manual_seed(0)
S, P = int(os.environ.get("S", "32")), 200
nblk = -(-P // 16)
T = S * P
dev = "cuda"
res0 = (torch.randn(T, 1024, device=dev) * 1.5).to(torch.bfloat16)
ln = (1 + 0.1 * torch.randn(1024, device=dev)).to(torch.bfloat16)
Wqkv = (torch.randn(3072, 1024, device=dev) * 0.03).to(torch.bfloat16)
Wo = (torch.randn(1024, 1024, device=dev) * 0.03).to(torch.bfloat16)
cs = llama3_cos_sin(2048).to(dev)
bt = torch.zeros(S, 64, dtype=torch.int32, device=dev)
perm = torch.randperm(S * nblk, device=dev).to(torch.int32) + 1
bt[:, :nblk] = perm.view(S, nblk)
pos = torch.arange(P, dtype=torch.int32, device=dev).repeat(S)
rseq = torch.arange(S, dtype=torch.int32, device=dev).repeat_interleave(P)
slot = ops.slot_mapping(bt, rseq, pos)
items = ops.prefill_items([i * P for i in range(S)], [P] * S, list(range(S))).cuda()
kv = torch.zeros((S * nblk + 1) * 65536, dtype=torch.uint8, device=dev)
qkv = torch.zeros(T, 3072, dtype=torch.bfloat16, device=dev)
h = torch.empty(T, 1024, dtype=torch.bfloat16, device=dev)
attn = torch.zeros(T, 1024, dtype=torch.bfloat16, device=dev)
reps = int(os.environ.get("REPS", "300"))
inner = int(os.environ.get("INNER", "4"))
ref = None
events = []
for k in range(reps):
    res = res0.clone()
    for _ in range(inner):          # a few "layers" back to back so the chain overlaps as in the engine
        ops.layer_qkv(res, ln, Wqkv, cs, pos, slot, kv, qkv=qkv, h=h)
        ops.paged_attn_prefill(qkv, kv, bt, items, out=attn)
        ops.linear_residual(attn, Wo, res)
    cur = (qkv[:, :1024].clone(), kv.clone(), attn.clone(), res.clone())
    if ref is None:
        ref = cur
        continue
    names = ["q", "kv", "attn", "res"]
    bad = [n for n, a, b in zip(names, cur, ref) if not torch.equal(a, b)]
    if bad:
        ne = (cur[2] != ref[2])
        rows = sorted(set(ne.nonzero()[:, 0].tolist()))
        events.append({"rep": k, "differs": bad, "attn_rows": rows[:8], "attn_n_rows": len(rows)})
print(json.dumps({"S": S, "T": T, "reps": reps, "inner": inner, "n_events": len(events), "events": events[:6]}))
