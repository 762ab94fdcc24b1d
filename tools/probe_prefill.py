"""Where does T3Engine.prefill spend its time?  python tools/probe_prefill.py [pairs] [prefix]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
PREFIX = int(sys.argv[2]) if len(sys.argv) > 2 else 200
dev = torch.device("cuda:0")
eng = T3Engine(random_state_dict(30, device=dev, seed=0), device=dev, max_pairs=B, max_model_len=PREFIX + 64, max_rows=16384)
sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=32, ignore_eos=True)
reqs = [Request(random_prompt(PREFIX - 35, seed=i), random_cond_emb(i, pinned=True), sp, uid=1 + i) for i in range(B)]
fw = eng.forward
gpu_ms = []


def timed_forward(T, mk):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    fw(T, mk)
    e1.record(torch.cuda.current_stream())
    gpu_ms.append((T, e0, e1))


eng.forward = timed_forward
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    for it in range(3):
        gpu_ms.clear()
        eng._init_sampler_state(reqs, P)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.prefill(reqs, P)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        f = [(T, a.elapsed_time(b)) for T, a, b in gpu_ms]
        rows = sum(T for T, _ in f)
        print(f"iter {it}: host-side call {1e3 * (t1 - t0):.1f} ms, to sync {1e3 * (t2 - t0):.1f} ms; forward GPU "
              f"{sum(m for _, m in f):.1f} ms over {len(f)} chunks {[(T, round(m, 1)) for T, m in f]}; "
              f"{rows * 1.0068e9 / (sum(m for _, m in f) / 1e3) / 1e12:.0f} TFLOP/s (GEMM flops only)")
