"""Decode-step timing probe: python tools/probe_step.py B [ctx] [steps] [graph]  (env CBX_PDL=0 disables PDL)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

B = int(sys.argv[1]); ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 300
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 64
graph = (sys.argv[4] != "0") if len(sys.argv) > 4 else True
dev = torch.device("cuda:0")
eng = T3Engine(random_state_dict(30, device=dev, seed=0), device=dev, max_pairs=B, max_model_len=ctx + steps + 64,
               max_rows=16384, use_graph=graph)
sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=steps + 40, ignore_eos=True)
reqs = [Request(random_prompt(ctx - 35, seed=i), random_cond_emb(i), sp, uid=1 + i) for i in range(B)]
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    eng._init_sampler_state(reqs, P)
    eng.prefill(reqs, P)
    for i in range(8):
        eng.decode(B, 1, ctx + 64)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(eng.stream)
    eng.decode(B, steps, ctx + 64)
    e1.record(eng.stream)
    torch.cuda.synchronize()
print(f"B={B} ctx={ctx} graph={graph} PDL={os.environ.get('CBX_PDL', '1')}: {e0.elapsed_time(e1) / steps * 1e3:.1f} us/step")
