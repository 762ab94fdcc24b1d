#!/usr/bin/env python3
"""BASELINE.json configs 2-5 at FULL size through the public engine call (T3Engine.generate: host
prompts + cond_emb in, host token lists out, prefill and continuous admission included), 30 layers,
random-init weights (sharp variant: speech_head x 16, so greedy argmaxes are well-conditioned).  One JSON object per config on stdout; with --out also written to a file.

These are the parity-test configurations of the bench contract, not bench lines: each run checks the
size-independent properties the domain offers —
  * every request emits exactly its max_tokens (ignore_eos) and every KV block returns to the pool;
  * the same job run twice gives the same tokens (Philox streams keyed by request uid, fixed kernel order);
  * a few probed requests, re-run ALONE through the engine's GEMV regime (different kernels, different
    accumulation order, different slots and pages), agree with their tokens from the big batch over the
    first greedy steps (until bf16 accumulation-order noise flips a near-tie; reported, not gated) — the full-size stand-in
    for the per-token oracle comparison of tests/test_gpu_engine.py, which the CPU oracle cannot do at
    30 layers x hundreds of sequences in minutes.

    python tools/run_configs.py [--configs 2,3,4,5] [--out gpurun_out/configs.json] [--weights fp8_e4m3]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="2,3,4,5")
    ap.add_argument("--out", default=None)
    ap.add_argument("--weights", default="bf16")
    ap.add_argument("--layers", type=int, default=30)
    args = ap.parse_args()
    import torch
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    results = []

    def sampled(max_tokens):
        return SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=max_tokens, ignore_eos=True)

    def greedy(max_tokens):
        return SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=max_tokens, ignore_eos=True)

    def run(name, eng, reqs, probes, what):
        free0 = eng.pool.num_free()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = eng.generate(reqs)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        n_tok = sum(len(o) for o in out)
        ok_len = all(len(o) == r.params.max_tokens for o, r in zip(out, reqs))
        ok_pool = eng.pool.num_free() == free0
        again = eng.generate(reqs)
        same = again == out
        # probes: greedy requests re-run alone (T = 2 rows: GEMV regime) must reproduce their first tokens
        agree = []
        for i in probes:
            alone = eng.generate([reqs[i]])[0]
            k = 0
            while k < len(alone) and alone[k] == out[i][k]:
                k += 1
            agree.append(k)
        res = {"config": name, "what": what, "weights": eng.weight_dtype, "requests": len(reqs), "pair_slots": eng.max_pairs,
               "speech_tokens": n_tok, "seconds": dt, "speech_tokens_per_s": n_tok / dt,
               "all_lengths_exact": ok_len, "pool_restored": ok_pool, "rerun_identical": same,
               "probe_common_prefix_vs_alone": agree, "kv_pool_gb": eng.num_blocks * 65536 * eng.n_layers / 1e9,
               "gpu_launches": eng.gpu_launches}
        assert ok_len and ok_pool and same, res
        print(json.dumps(res), flush=True)
        results.append(res)

    wanted = [int(c) for c in args.configs.split(",")]
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "prompt_lengths.json")))

    if 2 in wanted:      # 64 prompts x CFG pair = 128 seqs, prefix ~200, 500 speech tokens
        eng = T3Engine(random_state_dict(args.layers, device=dev, head_scale=16.0), device=dev, max_pairs=64, max_model_len=720,
                       weight_dtype=args.weights)
        reqs = [Request(random_prompt(165, seed=i), random_cond_emb(i), sampled(500) if i % 8 else greedy(500), uid=1 + i)
                for i in range(64)]
        run("cfg2", eng, reqs, [0, 56], "64 CFG pairs, prefix 200, 500 speech tokens each")
        eng.close(); del eng; torch.cuda.empty_cache()

    if 3 in wanted:      # benchmark-text-1 shape: README chunker's 154 chunks (+ drawn sentences up to 400), 256 pair slots
        n_text = list(golden["n_text"])
        g = torch.Generator().manual_seed(3)
        while len(n_text) < 400:
            n_text.append(int(golden["n_text"][int(torch.randint(0, len(golden["n_text"]), (1,), generator=g))]))
        eng = T3Engine(random_state_dict(args.layers, device=dev, head_scale=16.0), device=dev, max_pairs=256, max_model_len=1000,
                       weight_dtype=args.weights)
        reqs = [Request(random_prompt(n, seed=1000 + i), random_cond_emb(i % 7), sampled(390) if i % 16 else greedy(390),
                        uid=5000 + i) for i, n in enumerate(n_text)]
        run("cfg3", eng, reqs, [0, 160, 384],
            f"400 prompts (README chunker lengths of docs/benchmark-text-1.txt: n_text {min(n_text)}..{max(n_text)}), "
            "256 pair slots with continuous admission, 390 tokens each")
        eng.close(); del eng; torch.cuda.empty_cache()

    if 4 in wanted:      # multilingual sizes, prefill + decode mix: 128 slots, 320 requests of unequal lengths keep admission busy
        eng = T3Engine(random_state_dict(args.layers, device=dev, text_vocab=2454, head_scale=16.0), device=dev, max_pairs=128,
                       max_model_len=1000, weight_dtype=args.weights)
        reqs = []
        for i in range(320):
            ids = random_prompt(60 + (i * 37) % 240, seed=2000 + i, text_vocab=2454)
            ids[1] = 700 + i % 20                                   # a [lang] id from the multilingual range first
            mt = 128 + 64 * (i % 5)                                 # unequal lengths: slots free up at different steps
            reqs.append(Request(ids, random_cond_emb(100 + i % 9), sampled(mt) if i % 16 else greedy(mt), uid=9000 + i,
                                penalize_prompt_ids=True))
        run("cfg4", eng, reqs, [0, 16, 304], "multilingual text table (2454 ids), 320 requests through 128 pair slots, "
            "128-384 tokens each: prefill of new pairs interleaved with decode of the running ones")
        eng.close(); del eng; torch.cuda.empty_cache()

    if 5 in wanted:      # long-context stress: max_model_len 1000, 512 pairs, KV pool >= 95 % occupied
        P, gen = 200, 800
        per_seq = -(-(P + gen) // 16)
        num_blocks = int(2 * 512 * per_seq / 0.955) + 1
        free_b, _ = torch.cuda.mem_get_info()
        need = num_blocks * 65536 * args.layers + 6e9
        if free_b < need:
            print(json.dumps({"config": "cfg5", "skipped": f"needs {need / 1e9:.0f} GB, {free_b / 1e9:.0f} GB free"}), flush=True)
        else:
            eng = T3Engine(random_state_dict(args.layers, device=dev, head_scale=16.0), device=dev, max_pairs=512, max_model_len=1000,
                           num_blocks=num_blocks, weight_dtype=args.weights)
            reqs = [Request(random_prompt(165, seed=3000 + i), random_cond_emb(200 + i % 5), sampled(gen) if i % 64 else greedy(gen),
                            uid=20000 + i) for i in range(512)]
            occ = 2 * 512 * per_seq / (num_blocks - 1)
            run("cfg5", eng, reqs, [0, 448], f"512 CFG pairs (1024 seqs), prefix 200 -> context 1000, KV pool {occ:.1%} occupied")
            eng.close(); del eng; torch.cuda.empty_cache()

    if args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
