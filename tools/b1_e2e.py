"""Wall clock of ONE request through T3Engine.generate (1 CFG pair, prompt 75, 500 tokens, 30 layers, graph replay):
the single-request latency a caller sees, graph captures included.  Prints tokens/s of three runs (the first pays the
library's lazy initialisation)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

dev = torch.device("cuda:0")
sd = random_state_dict(30, device=dev, seed=0)
out = []
for run in range(3):
    eng = T3Engine(sd, device=dev, max_pairs=1, max_model_len=640)      # a fresh engine: no graphs yet
    sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=500, ignore_eos=True)
    req = Request(random_prompt(40, seed=run), random_cond_emb(run, pinned=True), sp, uid=1 + run)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    toks = eng.generate([req])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    assert len(toks[0]) == 500
    out.append(round(500 / dt, 1))
    eng.close()
print(json.dumps({"graph_buckets_only": os.environ.get("CBX_GRAPH_BUCKETS", "0"), "tokens_per_s_first_request_of_a_fresh_engine": out}))
