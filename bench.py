#!/usr/bin/env python3
"""bench.py — T3 speech tokens/s at 128 CFG pairs per GPU (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W          (torchrun for N > 1, one rank per GPU)
    python bench.py --impl reference ...                   (reference arm: HF-transformers CPU path)

A "step" is one decode step of the whole batch (128 CFG pairs = 256 sequences per GPU): embed →
30 Llama layers over the paged KV cache → head → fused CFG sampler; it emits one speech token per
pair.  Workload = SURVEY §8(d) "M1": 128 pairs, conditioning prefix P = 200, 500 speech tokens,
bf16, random-init weights, synthetic prompts.  The K timed steps are placed so that their mean
context equals the workload's mean context (P + 250 = 450) whenever K < 500; with the default
K/W the timed region IS the decode phase of the 500-token job.
Inputs are larger than L2 (KV alone is ~14 GB per step), so no L2 flush is needed between steps.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "t3_speech_tokens_per_sec_128_cfg_pairs"
PAIRS, PREFIX, GEN = 128, 200, 500
W_BYTES = 1_006_757_888 + 16_781_312          # backbone + speech head, bf16 (SURVEY §8 d)
KV_TOKEN_BYTES = 122_880                      # per token per sequence, all 30 layers


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms from the positioning steps (the same decode kernels, run
    just before the timed ones) to the end of the timed region, so that even a --steps 20 run carries >= 5 samples
    under load."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(index), "-lms", "50"], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.th.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        load = [v for v in sm if v > 0.5 * max(sm)] if sm else []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": statistics.median(load) if load else None,
                "sm_max_mhz": float(self.rows[0][1]) if self.rows else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_baseline_sample(pairs=8, prefix=450, steps=4, warmup=1, model=None):
    """The original-Chatterbox-style HF-transformers CPU path (oracle/hf_baseline.py) on a bounded
    sample of the workload: `pairs` CFG pairs at the workload's mean context, fp32, all host cores.
    8 pairs per step is the batch at which the host does best (tokens/s at 2 / 8 / 16 pairs on 8 cores:
    16 / 27 / 17 — beyond that HF's eager KV-cache attention and memory traffic swamp the GEMM gain)."""
    from oracle import hf_baseline
    r = hf_baseline.run(pairs, prefix, steps, warmup=warmup, model=model)
    return {"value": r["tokens"] / r["decode_s"], "unit": "speech_tokens/s", "cores": r["cores"], "kind": "port",
            "sample": f"{pairs} CFG pairs x {steps} decode steps at context {prefix} (workload mean), fp32 HF "
                      f"LlamaModel + python sampler; prefill {r['prefill_s']:.1f}s excluded",
            "ms_per_step": 1e3 * r["decode_s"] / steps}


KNOBS = ("CBX_SKIP", "CBX_MEGA", "CBX_GEMV_STREAM", "CBX_BLOCK_STOP", "CBX_BLOCK_NOACC", "CBX_BLOCK_FUSED", "CBX_MLP_FUSED", "CBX_GEMM_FOLD", "CBX_GRAPH_BUCKETS", "CBX_GEMM_SHALLOW", "CBX_DUAL_MIN_ROWS", "CBX_DUAL_ATTN_CTAS", "CBX_DUAL_CHAIN", "CBX_PDL",
         "CBX_ATTN_BALANCED", "CBX_ATTN_SMALL", "CBX_O_UNSPLIT", "CBX_SPLIT_MAX_ROWS", "CBX_L2_PREFETCH", "CBX_GEMM_MC",
         "CBX_ATTN_PLAN_EACH", "CBX_PREFILL_ROWWISE", "CBX_BENCH_LAYERS", "CBX_GEMV_MAX_T")


def refuse_knobs():
    """The library honours profiling / experiment switches (CBX_SKIP drops kernels: results are garbage).  A bench line
    measured under one of them is not a bench line: refuse, unless the caller says it is a tuning run."""
    on = [k for k in KNOBS if os.environ.get(k) not in (None, "")]
    if on and os.environ.get("CBX_BENCH_ALLOW_KNOBS") != "1":
        sys.exit(f"bench.py: refusing to run with {on} set (experiment / profiling switches change what is computed); "
                 "set CBX_BENCH_ALLOW_KNOBS=1 for a tuning run — its line is marked and is not a bench value")
    return on


def oracle_weights(sd: dict, n_layers: int):
    """The engine's checkpoint as the oracle's weight object (CPU copies; q/k/v and gate/up stacked)."""
    from oracle.t3_ref import T3ModelConfig, T3Weights
    c = lambda k: sd[k].detach().to("cpu")  # noqa: E731
    w = T3Weights(T3ModelConfig(n_layers=n_layers, text_vocab=sd["text_emb.weight"].shape[0]))
    for i in range(n_layers):
        p = f"tfmr.layers.{i}."
        w.wqkv.append(torch_cat([c(p + f"self_attn.{n}_proj.weight") for n in "qkv"]))
        w.wo.append(c(p + "self_attn.o_proj.weight"))
        w.wgu.append(torch_cat([c(p + "mlp.gate_proj.weight"), c(p + "mlp.up_proj.weight")]))
        w.wd.append(c(p + "mlp.down_proj.weight"))
        w.ln1.append(c(p + "input_layernorm.weight"))
        w.ln2.append(c(p + "post_attention_layernorm.weight"))
    w.norm, w.speech_emb, w.speech_head = c("tfmr.norm.weight"), c("speech_emb.weight"), c("speech_head.weight")
    w.text_emb, w.text_pos, w.speech_pos = c("text_emb.weight"), c("text_pos_emb.emb.weight"), c("speech_pos_emb.emb.weight")
    return w


def torch_cat(ts):
    import torch
    return torch.cat(ts, 0)


def oracle_probe(sd, n_layers, req, tokens, cfg_scale, seed, n=8):
    """The e2e arm's first pair against the CPU oracle: teacher-forced with the engine's first n tokens, every token
    must lie in the oracle's kept set after penalty / temperature / top-p (or within 8 bf16 ulps of the top logit of its
    boundary), and with the same Philox counters most must be the very token the oracle draws.  A bench run whose
    kernels computed something else (a skipped kernel, a stale buffer) fails here instead of printing a fast number."""
    import numpy as np
    from oracle import sampler_ref
    from oracle.t3_ref import T3Oracle
    o = T3Oracle(oracle_weights(sd, n_layers), "bf16")
    pair = o.new_pair()
    seen = np.zeros(8194, bool)
    p = req.params
    lg = o.prefill_pair(pair, req.cond_emb.float().cpu(), req.text_ids, cfg_scale).numpy()
    same = 0
    for s_ in range(n):
        tok = tokens[s_] - 1000
        l = sampler_ref.apply_repetition_penalty(lg, seen, p.repetition_penalty)
        keep, _ = sampler_ref.kept_set(l, p.temperature, p.min_p, p.top_p, p.top_k)
        floor = l[keep].min() - 8 * 2.0 ** -7 * abs(float(l.max()))
        assert l[tok] >= floor, f"bench: e2e token {s_} of the probed pair ({tok}) is outside the oracle's kept set"
        want = sampler_ref.sample_row(lg, seen, repetition_penalty=p.repetition_penalty, temperature=p.temperature,
                                      min_p=p.min_p, top_p=p.top_p, top_k=p.top_k, seed=p.seed if p.seed is not None else seed,
                                      uid=req.uid, step=s_)
        same += int(want == tok)
        seen[tok] = True
        lg = o.decode_pairs([pair], [tok], cfg_scale)[0].numpy()
    assert same >= n - 3, f"bench: only {same}/{n} of the probed pair's tokens are the oracle's draws"
    return {"pair": 0, "steps": n, "in_kept_set": n, "identical_draws": same}


def gpu_standin(dev, pairs=128, prefix=418, steps=64, warmup=4):
    """A reference-class number on the same GPU: the reference's own GPU path (vllm==0.10.0 +
    torch 2.7.1) cannot be installed here, so this is the stock-library stand-in for it — HF transformers' LlamaModel
    (bf16, SDPA attention, its own KV cache) driven as Chatterbox drives T3: batch rows [cond; uncond] for every pair,
    per-step CFG merge and a torch sampler.  None of this repo's kernels are on that path.  Timed like tts.py:407-420
    times LLM.generate: wall clock around the decode loop, at the workload's mean context (prefix 418 + 64 steps)."""
    import torch
    from oracle import hf_baseline
    m, emb, head, pos = hf_baseline.build_model(30, torch.bfloat16)
    m = m.to(dev)
    model = (m, emb.to(dev), head.to(dev), pos.to(dev))
    torch.cuda.synchronize()
    r = hf_baseline.run(pairs, prefix, steps, dtype=torch.bfloat16, model=model, warmup=warmup, device=dev)
    del model, m
    torch.cuda.empty_cache()
    return {"value": r["tokens"] / r["decode_s"], "unit": "speech_tokens/s", "ms_per_step": 1e3 * r["decode_s"] / steps,
            "what": f"HF transformers LlamaModel bf16 + SDPA + HF KV cache on this GPU, eager, {pairs} CFG pairs "
                    f"({2 * pairs} rows), context {prefix}..{prefix + steps}, torch sampler; stand-in for the reference's "
                    "vLLM-0.10 path, which cannot be installed offline"}


def vllm_standin(e2e_value):
    """The serving stack the reference sits on, stock, on this GPU: tools/vllm_standin.py in its own process (vLLM starts
    an engine-core subprocess and wants half the GPU's memory: this process has released its engines by now).  Compared
    with the E2E value — both include the prefill and are wall-clock around a generate call."""
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    try:
        r = subprocess.run([sys.executable, os.path.join(here, "tools", "vllm_standin.py")], capture_output=True, text=True,
                           timeout=1200)
        out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {str(e)[:200]}"}
    if "value" in out and e2e_value:
        out["this_e2e_over_standin"] = e2e_value / out["value"]
    return out


def run_reference(args, rank, world):
    """Reference arm: the reference's CPU implementation of the path (no GPU code involved)."""
    if rank != 0:
        return
    from oracle import hf_baseline
    model = hf_baseline.build_model(30)
    pairs = 8
    r = hf_baseline.run(pairs, 450, args.steps, warmup=args.warmup, model=model)
    val = r["tokens"] / r["decode_s"]
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "speech_tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * r["decode_s"] / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{PAIRS} CFG pairs x prefix {PREFIX} x {GEN} speech tokens (M1); each step = "
                                   f"one decode step of a {pairs}-pair sample at context 450"},
            "cpu_baseline": {"value": val, "unit": "speech_tokens/s", "cores": r["cores"], "kind": "port",
                             "sample": f"{pairs} CFG pairs per step, context 450, fp32 HF LlamaModel (original-"
                                       "chatterbox-style path restated; source absent from the reference tree)"},
            "e2e": {"value": val, "unit": "speech_tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_cfg3(args, rank, world, local):
    """BASELINE config 3 through the PRODUCT's multi-GPU path (sharding.generate_sharded): 400 prompts whose text lengths
    are drawn (seed 3, with replacement) from the reference chunker's 154 chunks of docs/benchmark-text-1.txt
    (tests/golden/prompt_lengths.json: min 17, mean 152, max 484 text tokens), 390 speech tokens each (the README-derived
    mean, ignore_eos), 256 pair slots per GPU, STRONG scaling: the same 400 prompts are split over the ranks
    (longest-first round-robin), streams gathered to rank 0 over NCCL.  value = 400 x 390 tokens / wall time of the
    whole job (prefill, admission, decode, gather), max over ranks."""
    import random
    import torch
    import torch.distributed as dist
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    from chatterbox_vllm_b200.sharding import generate_sharded
    from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    lens = json.load(open(os.path.join(ROOT, "tests", "golden", "prompt_lengths.json")))["n_text"]
    rng = random.Random(3)
    n_prompts, gen = 400, 390
    n_text = [rng.choice(lens) for _ in range(n_prompts)]
    eng = T3Engine(random_state_dict(30, device=dev, seed=0), device=dev, max_pairs=256, max_model_len=1000, max_rows=17408,
                   base_seed=1234)
    sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=gen, ignore_eos=True)
    mk = lambda: [Request(random_prompt(n, seed=i), random_cond_emb(i, pinned=True), sp, uid=1 + i)   # noqa: E731
                  for i, n in enumerate(n_text)]
    warm = mk()[:2 * world]
    for r in warm:
        r.params = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=40, ignore_eos=True)
    generate_sharded(eng, warm)                                   # warm-up: graphs, tensor maps, NCCL
    times = []
    for _ in range(max(1, min(args.steps or 2, 3))):
        reqs = mk()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        out = generate_sharded(eng, reqs)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        times.append(float(dt.item()))
        if rank == 0:
            assert len(out) == n_prompts and all(len(t) == gen for t in out), "cfg3: wrong token counts"
    if rank == 0:
        best = min(times)
        print(json.dumps({"metric": "t3_speech_tokens_per_sec_cfg3_400_prompts", "value": n_prompts * gen / best,
                          "unit": "speech_tokens/s", "n_gpus": world, "scaling": "strong", "seconds": best, "runs": times,
                          "higher_is_better": True, "dtype": "bf16", "data": "synthetic", "preemptions": eng.preemptions,
                          "config": {"workload": f"cfg3: {n_prompts} prompts (benchmark-text-1 chunk lengths, mean text "
                                                 f"{sum(n_text) / n_prompts:.0f} tokens) x {gen} speech tokens, 256 pair slots "
                                                 f"per GPU, split over {world} rank(s) by sharding.generate_sharded; prefill, "
                                                 "admission, decode and the NCCL gather inside the timed region"}}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs", type=int, default=PAIRS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-b1", action="store_true")
    ap.add_argument("--no-probe", action="store_true", help="skip the CPU-oracle check of the e2e arm's first pair")
    ap.add_argument("--no-standin", action="store_true", help="skip the stock-library GPU stand-in leg")
    ap.add_argument("--vllm-standin", action="store_true",
                    help="also run stock vLLM (this image's 0.22) on a stock Llama of the T3 backbone's dimensions in a "
                         "subprocess (tools/vllm_standin.py; its engine takes ~2.5 min to start, so it is not in the default run)")
    ap.add_argument("--no-fp8", action="store_true", help="skip the separately reported FP8-e4m3 weight leg")
    ap.add_argument("--weights", default="bf16", choices=["bf16", "fp8_e4m3"],
                    help="profiling aid: run the WHOLE bench on FP8 weights (the line then says so; the default line is bf16)")
    ap.add_argument("--config", default="m1", choices=["m1", "cfg3"],
                    help="m1 (default): the BASELINE metric's workload, weak scaling; cfg3: 400 benchmark-text prompts, strong "
                         "scaling through the product's sharded generate (its own JSON line, not the driver's metric)")
    ap.add_argument("--max-rows", type=int, default=17408, help="row capacity of the engine's step buffers (prefill chunk size)")
    ap.add_argument("--prefix", type=int, default=None,
                    help="profiling aid: prompt rows per sequence (default 200, the M1 workload); with --ctx0 it puts the first "
                         "timed step at that context after the warm-up steps alone, so a profiler has few launches to skip")
    ap.add_argument("--ctx0", type=int, default=None,
                    help="profiling aid: context length at the first timed step (default: centred on the workload mean)")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        args.steps, args.warmup = args.steps or 4, 1 if args.warmup is None else args.warmup
        return run_reference(args, rank, world)
    if args.config == "cfg3":
        refuse_knobs()
        return run_cfg3(args, rank, world, local)
    args.steps = args.steps or 480
    args.warmup = max(3, 20 if args.warmup is None else args.warmup)
    knobs = refuse_knobs()
    if args.prefix is not None:
        global PREFIX
        PREFIX = args.prefix
        knobs = knobs + [f"--prefix {args.prefix}"]        # not the M1 workload: the line is marked as a tuning run

    import torch
    import torch.distributed as dist
    from chatterbox_vllm_b200 import ops
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on STDOUT at communicator creation when NCCL_DEBUG asks for it; the contract is
        # ONE JSON line there, so stdout points at stderr until the communicator exists
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    B, K, W = args.pairs, args.steps, args.warmup
    L0 = max(PREFIX + 1 + W, PREFIX + GEN // 2 - K // 2)        # context at the first timed step
    if args.ctx0 is not None:
        L0 = max(PREFIX + 1 + W, args.ctx0)
    pre = L0 - PREFIX - 1                                       # untimed decode steps (>= W warm-up)
    max_tokens = 1 + pre + K
    max_len = PREFIX + max_tokens + 8
    n_layers = int(os.environ.get("CBX_BENCH_LAYERS", "30"))    # tuning aid only: the bench line is for 30
    sd = random_state_dict(n_layers, device=dev, seed=0)
    eng = T3Engine(sd, device=dev, max_pairs=B, max_model_len=max(max_len, PREFIX + GEN + 8),
                   max_rows=args.max_rows, base_seed=1234 + rank, weight_dtype=args.weights)
    w_bytes = W_BYTES if args.weights == "bf16" else eng.projection_weight_bytes()
    sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=max_tokens, ignore_eos=True)
    # global prompt index keeps the Philox streams independent of how requests are sharded
    reqs = [Request(random_prompt(PREFIX - 35, seed=rank * B + i), random_cond_emb(rank * B + i, pinned=True), sp,
                    uid=1 + rank * B + i) for i in range(B)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident arm: prefill + untimed steps, then EXACTLY K timed decode steps
    with torch.cuda.stream(eng.stream):
        P, held = eng._admit(reqs)
        eng._init_sampler_state(reqs, P)
        eng.prefill(reqs, P)                                     # warm (tensor maps, function attributes)
        # ---- prefill leg (the tensor-pipe side of the path): the same prefixes once more, timed
        eng._init_sampler_state(reqs, P)
        torch.cuda.synchronize()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record(eng.stream)
        eng.prefill(reqs, P)
        p1.record(eng.stream)
        torch.cuda.synchronize()
        prefill_ms = p0.elapsed_time(p1)
        clocks = ClockSampler(local)                            # from the positioning steps on: the same kernels, same load
        done = 0
        while done < pre:
            n = min(32, pre - done)
            eng.decode(B, n, PREFIX + 1 + done + n)
            done += n
        barrier()
        l0 = eng.gpu_launches
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(eng.stream)
        done = 0
        while done < K:
            n = min(32, K - done)
            eng.decode(B, n, L0 + done + n)
            done += n
        ev1.record(eng.stream)
        barrier()
        ms = ev0.elapsed_time(ev1)
        clk = clocks.stop()
        launches = eng.gpu_launches - l0
        n_out = eng.sb.n_out[:B].cpu()
        assert int(n_out.min()) == max_tokens and int(eng.sb.n_finished.item()) == B, "bench: wrong token count"
        toks = eng.sb.out_tokens[:B, :max_tokens].contiguous()
        if world > 1:                                           # gather emitted token streams (the only collective)
            allt = torch.empty(world * B, max_tokens, dtype=toks.dtype, device=dev)
            dist.all_gather_into_tensor(allt, toks)
        # ---- roofline leg: the dominant kernel (paged decode attention) timed alone on its stream
        mean_ctx = L0 + K // 2
        kvl = torch.full((2 * B,), mean_ctx, dtype=torch.int32, device=dev)
        rows = torch.arange(2 * B, dtype=torch.int32, device=dev)
        qkv = eng.buffer("qkv", (2 * B, 3072), torch.bfloat16)
        layer_bytes = eng.num_blocks * ops.PAGE_BYTES
        reps = 3
        for _ in range(2):                                      # the page-range plan is built once per step, as in the engine
            ops.paged_attn_decode(qkv, eng.kv_cache[:layer_bytes], eng.block_table, rows, kvl, mean_ctx, plan="make")
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(eng.stream)
        for _ in range(reps):
            for li in range(eng.n_layers):                      # a different layer's pages each launch: cold in L2
                ops.paged_attn_decode(qkv, eng.kv_cache[li * layer_bytes:(li + 1) * layer_bytes], eng.block_table,
                                      rows, kvl, mean_ctx, plan="reuse")
        a1.record(eng.stream)
        torch.cuda.synchronize()
        attn_ms = a0.elapsed_time(a1) / (reps * eng.n_layers)
        eng._release(held)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * B * K / (ms / 1e3)

    # ---- end-to-end arm: public engine API, host buffers in, host token lists out
    e2e = None
    if not args.no_e2e:
        gen_e2e = GEN                                           # the workload's 500 tokens per pair, whatever --steps says
        sp2 = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=gen_e2e, ignore_eos=True)
        for r in reqs:
            r.params = sp2
        barrier()
        t0 = time.perf_counter()
        out = eng.generate(reqs)
        torch.cuda.synchronize()
        if world > 1:
            mine = torch.tensor(out, dtype=torch.int32, device=dev)
            allt = torch.empty(world * B, gen_e2e, dtype=torch.int32, device=dev)
            dist.all_gather_into_tensor(allt, mine)
            torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        h2d = B * 34 * 1024 * 4 + sum(2 * (35 + len(r.text_ids)) * (16 + 4 + 4 + 4 + 8) for r in reqs) + 2 * B * eng.bt_stride * 4
        d2h = B * eng.max_model_len * 4 + B * 4 + 4 * (gen_e2e // 32 + 1)
        e2e = {"value": world * B * gen_e2e / float(dt.item()), "unit": "speech_tokens/s",
               "h2d_bytes_per_step": h2d / gen_e2e, "d2h_bytes_per_step": d2h / gen_e2e,
               "gen_tokens_per_pair": gen_e2e, "mean_context": PREFIX + gen_e2e // 2, "seconds": float(dt.item()),
               "what": f"T3Engine.generate: pinned-host prompts+cond_emb in, prefill + {gen_e2e} tokens/pair, token lists out"}
        assert all(len(o) == gen_e2e for o in out)
        if rank == 0 and args.weights == "bf16" and not args.no_probe:
            e2e["oracle_probe"] = oracle_probe(sd, n_layers, reqs[0], out[0], eng.cfg_scale, eng.base_seed)

    # ---- M2 leg: p50 decode-step latency at B = 1 pair (config 1 shape: P = 75, 500 tokens)
    def b1_leg(e, wb):
        G1 = min(GEN, e.max_model_len - 76)
        sp1 = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=G1, ignore_eos=True)
        r1 = [Request(random_prompt(40, seed=1), random_cond_emb(2, pinned=True), sp1, uid=999_983)]
        with torch.cuda.stream(e.stream):
            P1, held1 = e._admit(r1)
            e._init_sampler_state(r1, P1)
            e.prefill(r1, P1)
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(G1)]
            for i in range(20):                                  # warm-up steps (graph capture of bucket 1)
                e.decode(1, 1, P1[0] + 2 + i)
            evs[0].record(e.stream)
            for i in range(1, G1 - 20):
                e.decode(1, 1, P1[0] + 21 + i)
                evs[i].record(e.stream)
            torch.cuda.synchronize()
            e._release(held1)
        dts = sorted(evs[i - 1].elapsed_time(evs[i]) for i in range(1, G1 - 20))
        p50, p95 = dts[len(dts) // 2], dts[int(len(dts) * 0.95)]
        ctx1 = P1[0] + 20 + (G1 - 20) // 2
        bytes1 = wb + 2 * (KV_TOKEN_BYTES * (ctx1 + 1) + KV_TOKEN_BYTES + 2048)
        return {"metric": "p50_decode_step_ms_b1", "p50_ms": p50, "p95_ms": p95, "steps": len(dts),
                "workload": f"1 CFG pair (2 seqs), prefix {P1[0]}, context {P1[0] + 21}..{P1[0] + G1} (mean {ctx1})",
                "bytes_per_step": bytes1, "achieved_gbs": bytes1 / (p50 / 1e3) / 1e9,
                "tokens_per_s": 1e3 / p50}

    b1 = b1_leg(eng, w_bytes) if rank == 0 and not args.no_b1 else None

    # ---- FP8-e4m3 weight leg, reported separately (north star): same M1 batch and b = 1 shape on an engine whose
    # projections + head are e4m3 bytes with per-output-feature scales; activations / KV / sampler unchanged
    fp8 = None
    if rank == 0 and not args.no_fp8 and args.weights == "bf16":
        eng.close()
        del eng
        torch.cuda.empty_cache()
        e8 = T3Engine(random_state_dict(n_layers, device=dev, seed=0), device=dev, max_pairs=B, max_model_len=max_len,
                      max_rows=17408, base_seed=1234 + rank, weight_dtype="fp8_e4m3")
        K8 = min(K, 128)
        L8 = max(PREFIX + 1 + W, PREFIX + GEN // 2 - K8 // 2)
        pre8 = L8 - PREFIX - 1
        sp8 = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=1 + pre8 + K8, ignore_eos=True)
        for r in reqs:
            r.params = sp8
        with torch.cuda.stream(e8.stream):
            P8, held8 = e8._admit(reqs)
            e8._init_sampler_state(reqs, P8)
            e8.prefill(reqs, P8)
            e8._init_sampler_state(reqs, P8)
            torch.cuda.synchronize()
            q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            q0.record(e8.stream)
            e8.prefill(reqs, P8)
            q1.record(e8.stream)
            done = 0
            while done < pre8:
                n = min(32, pre8 - done)
                e8.decode(B, n, PREFIX + 1 + done + n)
                done += n
            torch.cuda.synchronize()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(e8.stream)
            done = 0
            while done < K8:
                n = min(32, K8 - done)
                e8.decode(B, n, L8 + done + n)
                done += n
            f1.record(e8.stream)
            torch.cuda.synchronize()
            e8._release(held8)
        ms8 = f0.elapsed_time(f1)
        w8b = e8.projection_weight_bytes()
        sb8 = w8b + 2 * B * (KV_TOKEN_BYTES * (L8 + K8 // 2 + 1) + KV_TOKEN_BYTES + 2048)
        fp8 = {"weights": "fp8_e4m3 per output channel (weight-only; bf16 activations/KV, fp32 accumulate)",
               "value": B * K8 / (ms8 / 1e3), "unit": "speech_tokens/s", "steps": K8, "ms_per_step": ms8 / K8,
               "weight_bytes_per_step": w8b, "bytes_per_step": sb8, "step_achieved_gbs": sb8 / (ms8 / K8 / 1e3) / 1e9,
               "prefill_ms": q0.elapsed_time(q1),
               "tolerance": "kernels vs the oracle on dequantised weights: the bf16 path's bars; FP8 model vs bf16 model: "
                            "head-logit rel-L2 <= 0.25 (tests/test_gpu_fp8.py)"}
        if not args.no_b1:
            fp8["b1"] = b1_leg(e8, w8b)
        e8.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm, which = peaks()
    if b1:
        b1["frac_of_hbm_peak"] = b1["achieved_gbs"] / hbm
    if fp8:
        fp8["step_frac_of_hbm_peak"] = fp8["step_achieved_gbs"] / hbm
        if "b1" in fp8:
            fp8["b1"]["frac_of_hbm_peak"] = fp8["b1"]["achieved_gbs"] / hbm
    attn_bytes = 2 * B * (mean_ctx * (KV_TOKEN_BYTES // 30) + 2048 + 2048)
    traffic = None
    try:
        # ncu --set full capture of this kernel (profiles/): measured DRAM bytes ÷ algorithmic bytes of the
        # captured launch, applied to this run's launch shape (the capture uses a shorter context)
        tr = json.load(open(os.path.join(ROOT, "profiles", "attn_traffic.json")))
        traffic = attn_bytes * tr["dram_bytes_per_launch"] / tr["algorithmic_bytes_same_launch"]
    except Exception:
        pass
    step_bytes = w_bytes + 2 * B * (KV_TOKEN_BYTES * (mean_ctx + 1) + KV_TOKEN_BYTES + 2048)
    line = {"metric": METRIC, "value": value, "unit": "speech_tokens/s", "n_gpus": world, "steps": K, "warmup": pre,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.weights == "bf16" else "bf16 activations, fp8_e4m3 weights", "data": "synthetic",
            "config": {"workload": f"M1: {B} CFG pairs ({2 * B} seqs) per GPU, prefix {PREFIX}, {GEN} speech tokens, "
                                   f"T3 Llama-520M random-init; timed steps at context {L0}..{L0 + K - 1} (mean {mean_ctx})",
                       "l2": "inputs larger than L2 (KV ~%.1f GB/step)" % (2 * B * mean_ctx * KV_TOKEN_BYTES / 1e9),
                       "sampler": "T=0.5 top_p=0.5 rep=2.0 (tts.py defaults after clamp)", "cuda_graph": True},
            "clocks": clk, "gpu_launches": int(launches), "e2e": e2e,
            "roofline": {"kernel": "paged_attn_decode_bal_kernel", "bound": "hbm",
                         "achieved": attn_bytes / (attn_ms / 1e3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": attn_bytes / (attn_ms / 1e3) / 1e9 / hbm, "traffic": traffic, "peak_source": which,
                         "us_per_launch": attn_ms * 1e3, "algorithmic_bytes_per_launch": attn_bytes, "context": mean_ctx,
                         "traffic_source": "profiles/attn_traffic.json: dram__bytes of one ncu --set full capture of this "
                                           "kernel, scaled by algorithmic bytes to this launch"},
            "step_roofline": {"bytes_per_step": step_bytes, "achieved_gbs": step_bytes / (ms / K / 1e3) / 1e9,
                              "frac": step_bytes / (ms / K / 1e3) / 1e9 / hbm}}
    # prefill: 2·ΣP rows through 30 layers (2·503 378 944 FLOP per row + causal attention) + one head row per sequence
    rows_p = 2 * sum(P)
    flop_p = rows_p * 2 * 503_378_944 + sum(2 * 4 * 1024 * 30 * p * (p + 1) // 2 for p in P) + 2 * B * 2 * 8194 * 1024
    try:
        tpeak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        tpeak = 1401.4
    line["prefill"] = {"rows": rows_p, "ms": prefill_ms, "rows_per_s": rows_p / (prefill_ms / 1e3),
                       "tflops": flop_p / (prefill_ms / 1e3) / 1e12, "peak_tflops": tpeak,
                       "frac_of_bf16_peak": flop_p / (prefill_ms / 1e3) / 1e12 / tpeak,
                       "what": "T3Engine.prefill of the M1 batch: host descriptor build + H2D + embed + 30 layers (tcgen05 GEMMs, "
                               "causal paged attention) + head + first-token sampler"}
    if b1:
        line["b1"] = b1
    if fp8:
        line["fp8w"] = fp8
    if knobs:
        line["tuning_run_not_a_bench_value"] = knobs
    if not args.no_standin and world == 1 and args.weights == "bf16":
        st = gpu_standin(dev)
        st["this_over_standin"] = value / st["value"]
        line["gpu_standin"] = st
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline_sample()
    if args.vllm_standin and world == 1:
        line["gpu_standin_vllm"] = vllm_standin(line.get("e2e", {}).get("value"))
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
