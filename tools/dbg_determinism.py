#!/usr/bin/env python3
"""Runs the same job repeatedly and reports where the token streams first differ (debugging aid)."""
import os, sys, json, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

def job(layers, B, steps, temp, graph, head_scale=16.0, reps=3, n_text=165):
    eng = T3Engine(random_state_dict(layers, device="cuda", head_scale=head_scale), device="cuda:0", max_pairs=B,
                   max_model_len=n_text + 40 + steps, use_graph=graph)
    sp = SamplingParams(temperature=temp, top_p=0.5 if temp else 1.0, repetition_penalty=2.0, max_tokens=steps, ignore_eos=True)
    reqs = [Request(random_prompt(n_text, seed=i), random_cond_emb(i), sp, uid=1 + i) for i in range(B)]
    outs = [eng.generate(reqs) for _ in range(reps)]
    diffs = []
    for k in range(1, reps):
        first = None
        nreq = 0
        for r in range(B):
            for s in range(steps):
                if outs[0][r][s] != outs[k][r][s]:
                    nreq += 1
                    if first is None or s < first[1]:
                        first = (r, s)
                    break
        diffs.append({"rep": k, "requests_differing": nreq, "first": first})
    eng.close()
    return diffs

if __name__ == "__main__":
    for layers, B, steps, temp, graph in [(30, 64, 96, 0.0, True), (30, 64, 96, 0.0, False), (2, 64, 200, 0.0, True),
                                          (30, 64, 96, 0.5, True), (30, 1, 200, 0.0, True), (30, 8, 150, 0.0, True)]:
        print(json.dumps({"layers": layers, "B": B, "steps": steps, "temp": temp, "graph": graph,
                          "diffs": job(layers, B, steps, temp, graph)}), flush=True)
