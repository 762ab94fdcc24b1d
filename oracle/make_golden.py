"""Generates tests/golden/* — run ONCE in the build container (needs /root/reference,
transformers and vLLM 0.22 importable on CPU).  The outputs are committed; nothing
at test/bench time reads /root/reference.

    PYTHONPATH=/root/repo python oracle/make_golden.py

Fixtures and what pins them:
  hf_backbone.npz      transformers.LlamaModel fp32 (2-layer T3-shaped config): prefill + decode
  rope_llama3.npz      vLLM 0.22 Llama3RotaryEmbedding._compute_inv_freq + HF rope init
  sampler_vllm.npz     vLLM 0.22 torch sampler ops on seeded logits
  block_pool_vllm.json vLLM 0.22 BlockPool on a scripted admit/finish sequence
  prompt_lengths.json  reference tokenizer.json + benchmark.py chunker on docs/benchmark-text-1.txt
  cond_enc.npz         the REFERENCE'S OWN T3CondEnc / Perceiver modules (models/t3/modules/cond_enc.py,
                       perceiver.py) run on seeded weights and inputs (oracle.cond_ref.random_*)
"""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "tests", "golden")
REF = "/root/reference"
sys.path.insert(0, os.path.join(HERE, ".."))

from oracle.t3_ref import T3ModelConfig, T3Weights  # noqa: E402


def hf_backbone():
    from transformers import LlamaConfig, LlamaModel
    cfg = T3ModelConfig(n_layers=2)
    w = T3Weights.random(cfg, seed=0, norm_jitter=0.1)
    ref_cfg = json.load(open(f"{REF}/t3-model/config.json"))
    ref_cfg.update(hidden_size=1024, num_hidden_layers=2)      # undo the CFG hack (t3.py:309)
    ref_cfg.pop("architectures"), ref_cfg.pop("torch_dtype")
    m = LlamaModel(LlamaConfig(**ref_cfg)).float().eval()
    sd = {k[5:]: v.float() for k, v in w.state_dict().items() if k.startswith("tfmr.")}
    sd["embed_tokens.weight"] = torch.zeros(ref_cfg["vocab_size"], 1024)
    m.load_state_dict(sd, strict=True)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 20, 1024, generator=g) * 0.02
    x2 = torch.randn(2, 1, 1024, generator=g) * 0.02
    with torch.no_grad():
        o1 = m(inputs_embeds=x, use_cache=True)
        o2 = m(inputs_embeds=x2, past_key_values=o1.past_key_values, use_cache=True)
    np.savez_compressed(f"{OUT}/hf_backbone.npz", x=x.numpy(), x2=x2.numpy(),
                        h_prefill=o1.last_hidden_state.numpy(), h_decode=o2.last_hidden_state.numpy())


def rope():
    from transformers.modeling_rope_utils import ROPE_INIT_FUNCTIONS
    from transformers import LlamaConfig
    ref_cfg = json.load(open(f"{REF}/t3-model/config.json"))
    ref_cfg.update(hidden_size=1024)
    ref_cfg.pop("architectures"), ref_cfg.pop("torch_dtype")
    hf_inv, _ = ROPE_INIT_FUNCTIONS["llama3"](LlamaConfig(**ref_cfg), "cpu")
    out = dict(hf_inv_freq=hf_inv.numpy())
    try:
        from vllm.model_executor.layers.rotary_embedding.llama3_rope import Llama3RotaryEmbedding
        r = Llama3RotaryEmbedding.__new__(Llama3RotaryEmbedding)
        r.rotary_dim, r.scaling_factor, r.low_freq_factor, r.high_freq_factor, r.orig_max_position = 64, 8.0, 1.0, 4.0, 8192
        out["vllm_inv_freq"] = Llama3RotaryEmbedding._compute_inv_freq(r, 500000.0).numpy()
    except Exception as e:  # pragma: no cover
        print("vllm rope unavailable:", e)
    np.savez_compressed(f"{OUT}/rope_llama3.npz", **out)


def sampler():
    from vllm._custom_ops import apply_repetition_penalties_torch
    from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch
    g = torch.Generator().manual_seed(7)
    V, R = 8194, 12
    logits = torch.randn(R, V, generator=g) * 3.0
    logits[3:6] *= 2.0
    seen = torch.rand(R, V, generator=g) < 0.05
    pen = logits.clone()
    apply_repetition_penalties_torch(pen, seen, torch.zeros_like(seen), torch.full((R,), 2.0))
    cases = [(0.5, 0.0, 0.5, 0), (0.8, 0.05, 1.0, 0), (0.5, 0.1, 0.8, 0), (1.0, 0.0, 0.9, 50), (0.3, 0.0, 1.0, 20)]
    out = dict(logits=logits.numpy(), seen=seen.numpy(), penalised=pen.numpy(),
               cases=np.array(cases, np.float64))
    for ci, (T, min_p, top_p, top_k) in enumerate(cases):
        l = pen.clone() / T
        if min_p > 0:   # builtin.py:101-112
            pr = torch.softmax(l, dim=-1)
            l = l.masked_fill(pr < pr.amax(dim=-1, keepdim=True) * min_p, -float("inf"))
        l = apply_top_k_top_p_pytorch(l, torch.full((R,), top_k) if top_k else None,
                                      torch.full((R,), top_p) if top_p < 1.0 else None)
        out[f"keep_{ci}"] = torch.isfinite(l).numpy()
        out[f"probs_{ci}"] = torch.softmax(l, dim=-1).numpy()
    np.savez_compressed(f"{OUT}/sampler_vllm.npz", **out)


def block_pool():
    from vllm.v1.core.block_pool import BlockPool
    pool = BlockPool(num_gpu_blocks=64, enable_caching=False, hash_block_size=16)
    script = [("alloc", 0, 5), ("alloc", 1, 3), ("alloc", 2, 7), ("free", 1), ("alloc", 3, 4),
              ("alloc", 0, 2), ("free", 0), ("free", 2), ("alloc", 4, 20), ("alloc", 5, 6), ("free", 3),
              ("alloc", 6, 9)]
    held, trace = {}, []
    for op in script:
        if op[0] == "alloc":
            blks = pool.get_new_blocks(op[2])
            held.setdefault(op[1], []).extend(blks)
            trace.append([b.block_id for b in blks])
        else:
            pool.free_blocks(reversed(held.pop(op[1])))
            trace.append([])
    json.dump(dict(num_blocks=64, script=script, trace=trace, free_after=pool.get_num_free_blocks()),
              open(f"{OUT}/block_pool_vllm.json", "w"))


def prompt_lengths():
    from tokenizers import Tokenizer
    tok = Tokenizer.from_file(f"{REF}/src/chatterbox_vllm/models/t3/tokenizer.json")
    spec = importlib.util.spec_from_file_location("ref_benchmark", f"{REF}/benchmark.py")
    src = open(f"{REF}/benchmark.py").read().split("if __name__")[0]
    src = "\n".join(l for l in src.splitlines() if "chatterbox_vllm" not in l and "torchaudio" not in l)
    ns = {}
    exec(compile(src, "ref_benchmark", "exec"), ns)
    spec2 = importlib.util.spec_from_file_location("ref_text_utils", f"{REF}/src/chatterbox_vllm/text_utils.py")
    tu = importlib.util.module_from_spec(spec2)
    spec2.loader.exec_module(tu)
    text = open(f"{REF}/docs/benchmark-text-1.txt").read()
    # benchmark.py:60-72: drop '#' lines, split by newline, chunk each line, flatten
    lines = [l.strip() for l in text.split("\n") if not l.startswith("#") and len(l.strip()) > 0]
    chunks = [c for line in lines for c in ns["split_text_by_sentence"](line)]

    def ids(p):   # tts.py:327 + entokenizer.py:93-95
        return tok.encode(("[START]" + tu.punc_norm(p) + "[STOP]").replace(" ", "[SPACE]")).ids

    n_text = [len(ids(c)) for c in chunks]
    demo = "You are listening to a demo of the Chatterbox TTS model running on VLLM."
    cases = ["", "hello world", "wait... what; no — yes: ok", "Already ends!", "  many   spaces here ", "“Quoted” ‘text’ …"]
    json.dump(dict(n_chunks=len(chunks), n_text=n_text, chunk_chars=[len(c) for c in chunks],
                   total_prompt_tokens=sum(34 + n + 1 for n in n_text), readme_total=28754,
                   demo_text=demo, demo_ids=ids(demo),
                   punc_norm_cases=[[c, tu.punc_norm(c)] for c in cases]),
              open(f"{OUT}/prompt_lengths.json", "w"))


def cond_enc():
    """Runs the reference's own modules (they only need torch + einops) and stores their output."""
    import types
    from oracle import cond_ref
    base = f"{REF}/src/chatterbox_vllm/models/t3/modules/"
    pkg = types.ModuleType("refmods")
    pkg.__path__ = [base]
    sys.modules["refmods"] = pkg
    mods = {}
    for name in ("t3_config", "perceiver", "cond_enc"):
        spec = importlib.util.spec_from_file_location("refmods." + name, base + name + ".py")
        m = importlib.util.module_from_spec(spec)
        sys.modules["refmods." + name] = m
        spec.loader.exec_module(m)
        mods[name] = m
    enc = mods["cond_enc"].T3CondEnc(mods["t3_config"].T3Config()).eval()
    out = {}
    for case, (wseed, iseed, n_prompt, emotion) in enumerate([(11, 12, 150, 0.5), (21, 22, 37, 0.8)]):
        sd = cond_ref.random_state_dict(wseed)
        enc.load_state_dict(sd)
        speech_emb, speech_pos, speaker, tokens = cond_ref.random_inputs(iseed, n_prompt)
        prompt_emb = speech_emb[tokens] + speech_pos[:n_prompt]                       # tts.py:238
        with torch.inference_mode():
            cond = mods["cond_enc"].T3Cond(speaker_emb=speaker.view(1, -1), cond_prompt_speech_tokens=tokens.view(1, -1),
                                           cond_prompt_speech_emb=prompt_emb, emotion_adv=emotion * torch.ones(1, 1))
            y = enc(cond)
            row = enc.emotion_adv_fc(0.3 * torch.ones(1, 1))[0]                       # tts.py:254
        assert y.shape == (34, 1024)
        out[f"meta{case}"] = np.array([wseed, iseed, n_prompt], np.int64)
        out[f"emotion{case}"] = np.array([emotion], np.float64)
        out[f"cond{case}"] = y.numpy()
        out[f"row03_{case}"] = row.numpy()
    np.savez_compressed(os.path.join(OUT, "cond_enc.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[1:]
    for fn in (hf_backbone, rope, sampler, block_pool, prompt_lengths, cond_enc):
        if only and fn.__name__ not in only:
            continue
        try:
            fn()
            print("ok  ", fn.__name__)
        except Exception as e:
            print("FAIL", fn.__name__, repr(e))
