"""A reference-CLASS number on the same GPU: stock vLLM (the version in this image, 0.22) serving a stock
`LlamaForCausalLM` with the T3 backbone's dimensions and dummy weights — the serving stack the reference sits on
(vllm==0.10.0 there, not installable offline), WITHOUT the reference's T3 glue: no conditioning prefix assembly, no
CFG merge, no speech embeddings.  256 independent sequences (= 128 CFG pairs' worth of rows), 200 prompt tokens,
500 sampled tokens each with the reference's sampling parameters, timed the way tts.py:407-420 times LLM.generate (wall
clock around generate, prefill included).  Reported in pair-tokens/s: 128 x 500 / seconds, the bench's unit.

Library code only (vLLM's kernels, CUDA graphs, torch.compile): none of this repo is on that path.  Prints one JSON
line; {"unavailable": "..."} if the engine does not start on this box (no network, JIT back-ends).

    python tools/vllm_standin.py [--eager] [--pairs 128] [--gen 500]
"""
import argparse
import json
import os
import sys
import tempfile
import time

os.environ.setdefault("HF_HUB_OFFLINE", "1")
os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
os.environ.setdefault("VLLM_WORKER_MULTIPROC_METHOD", "spawn")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=128)
    ap.add_argument("--prefix", type=int, default=200)
    ap.add_argument("--gen", type=int, default=500)
    ap.add_argument("--eager", action="store_true", help="no torch.compile / CUDA graphs (if the compiled path will not start)")
    args = ap.parse_args()
    d = tempfile.mkdtemp(prefix="t3_llama_cfg_")
    cfg = {"architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_size": 1024, "intermediate_size": 4096,
           "num_hidden_layers": 30, "num_attention_heads": 16, "num_key_value_heads": 16, "head_dim": 64,
           "vocab_size": 8448, "max_position_embeddings": 4096, "rms_norm_eps": 1e-5, "rope_theta": 500000.0,
           "rope_scaling": {"factor": 8.0, "high_freq_factor": 4.0, "low_freq_factor": 1.0,
                            "original_max_position_embeddings": 8192, "rope_type": "llama3"},
           "torch_dtype": "bfloat16", "tie_word_embeddings": False, "hidden_act": "silu", "attention_bias": False,
           "mlp_bias": False, "bos_token_id": 0, "eos_token_id": 1}
    json.dump(cfg, open(os.path.join(d, "config.json"), "w"))
    try:
        import numpy as np
        import vllm
        from vllm import LLM, SamplingParams
        t0 = time.perf_counter()
        llm = LLM(model=d, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=1024,
                  gpu_memory_utilization=0.5, max_num_seqs=2 * args.pairs, enable_prefix_caching=False, seed=0,
                  generation_config="vllm", enforce_eager=args.eager, disable_log_stats=True)
        t_start = time.perf_counter() - t0
        rng = np.random.default_rng(0)
        n = 2 * args.pairs
        prompts = [{"prompt_token_ids": rng.integers(2, 8194, args.prefix).tolist()} for _ in range(n)]
        sp = lambda k: SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=k,  # noqa: E731
                                      ignore_eos=True, detokenize=False)
        llm.generate(prompts, sp(16), use_tqdm=False)                      # warm: graphs, allocator
        t0 = time.perf_counter()
        outs = llm.generate(prompts, sp(args.gen), use_tqdm=False)
        dt = time.perf_counter() - t0
        got = sum(len(o.outputs[0].token_ids) for o in outs)
        assert got == n * args.gen, (got, n * args.gen)
        print(json.dumps({"value": args.pairs * args.gen / dt, "unit": "speech_tokens/s (pair-tokens: 2 sequences each)",
                          "seconds": dt, "sequences": n, "prefix": args.prefix, "gen": args.gen,
                          "sequence_tokens_per_s": got / dt, "engine_start_s": t_start, "vllm": vllm.__version__,
                          "eager": bool(args.eager),
                          "what": "stock vLLM LlamaForCausalLM, T3 backbone dimensions, dummy weights, no CFG merge / T3 glue; "
                                  "wall clock around LLM.generate (prefill included)"}), flush=True)
    except BaseException as e:  # noqa: BLE001  (engine start failures come in many types; the caller wants one line)
        print(json.dumps({"unavailable": f"{type(e).__name__}: {str(e)[:300]}"}), flush=True)
        sys.exit(0)


if __name__ == "__main__":
    main()
