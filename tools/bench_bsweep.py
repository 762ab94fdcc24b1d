"""Decode-step latency against the number of CFG pairs (B = 1, 2, 3, 4, 8, 16, 32; T = 2B rows), to evidence the
GEMV -> tcgen05 switch at T > 4: run once as is (GEMV kernels at T <= 4, tcgen05 above) and once with
CBX_GEMV_MAX_T=0 (tcgen05 GEMMs at every T).  Prints one JSON object: {B: p50 step ms}.  M1-shaped pairs (prefix 200),
30 layers, CUDA-graph replay, context ~330."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

dev = torch.device("cuda:0")
sd = random_state_dict(30, device=dev, seed=0)
out = {}
for B in [int(x) for x in os.environ.get("PAIRS", "1,2,3,4,8,16,32").split(",")]:
    eng = T3Engine(sd, device=dev, max_pairs=B, max_model_len=520)
    sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=300, ignore_eos=True)
    reqs = [Request(random_prompt(165, seed=i), random_cond_emb(i, pinned=True), sp, uid=1 + i) for i in range(B)]
    with torch.cuda.stream(eng.stream):
        P, held = eng._admit(reqs)
        eng._init_sampler_state(reqs, P)
        eng.prefill(reqs, P)
        eng.decode(B, 100, 320)                               # to context ~300, warm (graph captured)
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(101)]
        evs[0].record(eng.stream)
        for i in range(100):
            eng.decode(B, 1, 384)
            evs[i + 1].record(eng.stream)
        torch.cuda.synchronize()
        eng._release(held)
    d = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(100))
    out[B] = round(d[50], 4)
    eng.close()
print(json.dumps({"gemv_max_t": os.environ.get("CBX_GEMV_MAX_T", "8 (default)"), "p50_step_ms_by_pairs": out}))
