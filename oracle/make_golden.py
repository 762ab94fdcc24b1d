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
  t3_glue.npz          the REFERENCE'S OWN models/t3/t3.py glue (T3MultiModalProcessor.apply, split_prefill_decode,
                       get_input_embeddings, compute_logits, create_triangular_matrix) executed from /root/reference with
                       the vllm imports stubbed, on oracle.glue_cases inputs
  mtl_tokenizer.json   the REFERENCE'S OWN MTLTokenizer (models/t3/mtl_tokenizer.py) on its packaged vocabulary
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


def load_reference_t3():
    """Executes the reference's models/t3/t3.py FROM /root/reference with every vllm import replaced by an
    inert stub (the file needs vLLM 0.9/0.10 internals that vLLM 0.22 no longer has; none of them takes part
    in the functions pinned here).  Its sibling modules (learned_pos_emb, t3_config, cond_enc, perceiver)
    are the reference's real files.  Returns (module, restore) — call restore() to put sys.modules back."""
    import types
    saved = dict(sys.modules)

    class _Generic:
        def __class_getitem__(cls, item):
            return cls

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        m.__path__ = []
        sys.modules[name] = m
        return m

    for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
        del sys.modules[k]
    blank = lambda n: type(n, (), {})  # noqa: E731

    class _Registry:
        @staticmethod
        def register_processor(*a, **k):
            return lambda cls: cls

    class _Captured(dict):      # MultiModalKwargs.from_items([...from_elems(...build_elems(data=[x]))]) -> {"conditionals": [x]}
        @classmethod
        def from_items(cls, items):
            return cls(conditionals=[d for it in items for d in it])

    class _Item:
        @staticmethod
        def from_elems(elems):
            return elems

    class _Field:
        def build_elems(self, modality, key, data):
            return list(data)

    stub("vllm")
    stub("vllm.config", VllmConfig=blank("VllmConfig"), ModelConfig=blank("ModelConfig"))
    stub("vllm.model_executor"), stub("vllm.model_executor.layers"), stub("vllm.model_executor.models")
    stub("vllm.model_executor.layers.logits_processor", LogitsProcessor=blank("LogitsProcessor"))
    stub("vllm.model_executor.layers.vocab_parallel_embedding", ParallelLMHead=blank("ParallelLMHead"))
    stub("vllm.model_executor.models.interfaces", MultiModalEmbeddings=object, SupportsMultiModal=blank("SupportsMultiModal"))
    stub("vllm.model_executor.models.interfaces_base", VllmModelForTextGeneration=blank("VllmModelForTextGeneration"))
    stub("vllm.model_executor.models.llama", LlamaModel=blank("LlamaModel"))
    stub("vllm.model_executor.sampling_metadata", SamplingMetadata=blank("SamplingMetadata"))
    stub("vllm.multimodal", MULTIMODAL_REGISTRY=_Registry)
    stub("vllm.multimodal.inputs", MultiModalKwargs=_Captured, MultiModalKwargsItem=_Item, MultiModalBatchedField=_Field)
    stub("vllm.multimodal.parse", MultiModalDataParser=blank("MultiModalDataParser"),
         ModalityDataItems=type("ModalityDataItems", (_Generic,), {"__init__": lambda self, data, modality: setattr(self, "data", data)}))
    stub("vllm.multimodal.processing", BaseMultiModalProcessor=type("BaseMultiModalProcessor", (_Generic,), {}),
         BaseProcessingInfo=blank("BaseProcessingInfo"), MultiModalDataDict=dict, MultiModalDataItems=dict,
         MultiModalFieldConfig=blank("MultiModalFieldConfig"), PromptUpdate=blank("PromptUpdate"),
         MultiModalInputs=dict, PlaceholderRange=lambda **kw: dict(kw))
    stub("vllm.multimodal.profiling", BaseDummyInputsBuilder=blank("BaseDummyInputsBuilder"))
    stub("vllm.sequence", IntermediateTensors=blank("IntermediateTensors"))
    base = f"{REF}/src/chatterbox_vllm"
    for name, path in (("chatterbox_vllm", base), ("chatterbox_vllm.models", base + "/models"),
                       ("chatterbox_vllm.models.t3", base + "/models/t3")):
        m = types.ModuleType(name)          # package shells: the real __init__ files import vllm
        m.__path__ = [path]
        sys.modules[name] = m
    spec = importlib.util.spec_from_file_location("chatterbox_vllm.models.t3.t3", base + "/models/t3/t3.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["chatterbox_vllm.models.t3.t3"] = mod
    spec.loader.exec_module(mod)

    def restore():
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        sys.modules.update(saved)
    return mod, restore


def reference_t3_model(ref, variant: str, cfg_scale: float, head=None, dtype=torch.float32):
    """A bare object carrying exactly the attributes the reference's methods read, with the reference's
    UNBOUND functions attached (t3.py:318-343 builds the same attributes from the checkpoint)."""
    from oracle import glue_cases as gc
    tb = gc.tables(variant)
    t3conf = ref.T3Config.multilingual() if variant == "mtl" else ref.T3Config()

    class Bare:
        split_prefill_decode = ref.T3VllmModel.split_prefill_decode
        get_input_embeddings = ref.T3VllmModel.get_input_embeddings
        compute_logits = ref.T3VllmModel.compute_logits

    m = Bare()
    m.t3conf, m.dim, m.cfg_scale = t3conf, t3conf.n_channels, cfg_scale
    m.text_emb = torch.nn.Embedding.from_pretrained(tb["text_emb"].to(dtype))
    m.speech_emb = torch.nn.Embedding.from_pretrained(tb["speech_emb"].to(dtype))
    # t3.py:371-377: the precomputed tables are LearnedPositionEmbeddings.get_fixed_embedding(arange)[0]
    tp = ref.LearnedPositionEmbeddings(t3conf.max_text_tokens + 2, m.dim)
    tp.emb.weight.data.copy_(tb["text_pos"])
    sp = ref.LearnedPositionEmbeddings(t3conf.max_speech_tokens + 4, m.dim)
    sp.emb.weight.data.copy_(tb["speech_pos"])
    with torch.no_grad():
        m.precomputed_text_pos_emb = tp.get_fixed_embedding(torch.arange(t3conf.max_text_tokens + 2))[0].to(dtype)
        m.precomputed_speech_pos_emb = sp.get_fixed_embedding(torch.arange(t3conf.max_speech_tokens + 4))[0].to(dtype)
    if head is not None:
        m.speech_head = types_ns(weight=head.to(dtype))
        m.seen_logits = []

        def logits_processor(lm_head, hidden, _meta):          # vLLM LogitsProcessor: hidden @ lm_head.weight.T (fp32 accumulate)
            lg = (hidden.float() @ lm_head.weight.float().T).to(hidden.dtype)
            m.seen_logits.append(lg)
            return lg
        m.logits_processor = logits_processor
    return m


def types_ns(**kw):
    import types
    return types.SimpleNamespace(**kw)


def t3_glue():
    """The reference's own t3.py glue (apply / split_prefill_decode / get_input_embeddings / compute_logits /
    create_triangular_matrix) run on oracle.glue_cases inputs -> tests/golden/t3_glue.npz."""
    from oracle import glue_cases as gc
    ref, restore = load_reference_t3()
    out = {}
    try:
        assert (ref.PREFILL_COND_START_TOKEN, ref.PREFILL_COND_END_TOKEN, ref.PREFILL_END_TOKEN,
                ref.CONDITIONING_SIZE, ref.SPEECH_TOKEN_OFFSET) == (695, 696, 697, 34, 1000)
        out["tri_5x9"] = ref.create_triangular_matrix(5, 9).numpy()
        out["tri_40x1024_digest"] = np.array(gc.digest(ref.create_triangular_matrix(40, 1024).numpy()))

        class BareProc:                                       # T3MultiModalProcessor.apply on a tokeniser-free shell
            apply = ref.T3MultiModalProcessor.apply

            def _to_mm_items(self, mm_data):
                return mm_data

            def _apply_hf_processor(self, prompt, mm_items, kw, tok_kw, return_mm_hashes=False):
                return list(prompt), None, None, False        # (prompt_ids, mm_kwargs, mm_hashes, is_update_applied)

        for name, variant, n_text, ids_seed, cond_seed in gc.prefix_cases():
            ids, cond = gc.text_ids(variant, n_text, ids_seed), gc.cond_emb(cond_seed)
            r = BareProc().apply(ids, {"conditionals": [cond]}, {})
            final_ids, side = r["prompt_token_ids"], r["mm_kwargs"]["conditionals"][0]
            ph = r["mm_placeholders"]["conditionals"][0]
            P = len(final_ids)
            out[f"{name}/text_ids"] = np.array(ids, np.int64)
            out[f"{name}/final_prompt_ids"] = np.array(final_ids, np.int64)
            out[f"{name}/placeholder"] = np.array([ph["offset"], ph["length"]], np.int64)
            out[f"{name}/side_digest"] = np.array(gc.digest(side.numpy()))
            out[f"{name}/side_sample"] = gc.sample(side.numpy())
            out[f"{name}/side_rowsum_tail"] = side[34:].sum(1).numpy()
            model = reference_t3_model(ref, variant, 0.5)
            ids_t = torch.tensor(final_ids)
            with torch.no_grad():
                full = model.get_input_embeddings(ids_t, [side])
            assert full.shape == (P, 2048)
            out[f"{name}/full_digest"] = np.array(gc.digest(full.numpy()))
            out[f"{name}/full_sample"] = gc.sample(full.numpy())
            if name == "en_short":
                out[f"{name}/full"] = full.numpy()
            # every chunk split of App. C: each slice goes through the reference as its own call, side rows sliced alike
            for cuts in gc.chunk_splits(P):
                edges = [0, *cuts, P]
                for a, b in zip(edges[:-1], edges[1:]):
                    with torch.no_grad():
                        e = model.get_input_embeddings(ids_t[a:b], [side[a:b]])
                    assert e.shape == (b - a, 2048), (name, cuts, a, b, e.shape)
                    key = f"{name}/chunk_{a}_{b}"
                    if key + "_digest" not in out:
                        out[key + "_digest"] = np.array(gc.digest(e.numpy()))
                        out[key + "_equals_full_rows"] = np.array(bool(torch.equal(e, full[a:b])))
            # mixed batch: decode tokens of two running requests, a whole new prefix, one more decode token
            dec = torch.tensor([1000 + 17, 1000 + 6561, 1000 + 4242])
            mixed_ids = torch.cat([dec[:2], ids_t, dec[2:]])
            with torch.no_grad():
                segs = model.split_prefill_decode(mixed_ids, [side])
                mixed = model.get_input_embeddings(mixed_ids, [side])
                pure = model.get_input_embeddings(dec, None)
            out[f"{name}/mixed_seg_lens"] = np.array([len(s[0]) for s in segs], np.int64)
            out[f"{name}/mixed_seg_prefill"] = np.array([s[1] is not None for s in segs])
            out[f"{name}/mixed_digest"] = np.array(gc.digest(mixed.numpy()))
            out[f"{name}/decode_digest"] = np.array(gc.digest(pure.numpy()))
            out[f"{name}/decode_sample"] = gc.sample(pure.numpy(), 211)
        # two requests' prefixes back to back in one batch (vLLM batches prefills), second one cut short by the budget
        (n0, v0, t0, s0, c0), (n1, _, t1, s1, c1) = gc.prefix_cases()[0], gc.prefix_cases()[1]
        model = reference_t3_model(ref, "en", 0.5)
        reqs = []
        for (t, s, c) in ((t0, s0, c0), (t1, s1, c1)):
            r = BareProc().apply(gc.text_ids("en", t, s), {"conditionals": [gc.cond_emb(c)]}, {})
            reqs.append((torch.tensor(r["prompt_token_ids"]), r["mm_kwargs"]["conditionals"][0]))
        cut = 40
        ids2 = torch.cat([reqs[0][0], reqs[1][0][:cut]])
        with torch.no_grad():
            segs = model.split_prefill_decode(ids2, [reqs[0][1], reqs[1][1][:cut]])
            both = model.get_input_embeddings(ids2, [reqs[0][1], reqs[1][1][:cut]])
        out["two_prefixes/seg_lens"] = np.array([len(s[0]) for s in segs], np.int64)
        out["two_prefixes/cut"] = np.array(cut)
        out["two_prefixes/digest"] = np.array(gc.digest(both.numpy()))
        # CHATTERBOX_STRIP_LLM_EOS (t3.py:209-221,437-447): id 2 dropped from the walk
        os.environ["CHATTERBOX_STRIP_LLM_EOS"] = "1"
        try:
            ids3 = torch.tensor([695, 2, 5, 2, 697, 1003, 2, 1004])
            segs = model.split_prefill_decode(ids3, [torch.zeros(5, 1024)])
            out["strip_eos/seg_ids"] = np.array([t for s in segs for t in s[0].tolist()] , np.int64)
            out["strip_eos/seg_lens"] = np.array([len(s[0]) for s in segs], np.int64)
        finally:
            del os.environ["CHATTERBOX_STRIP_LLM_EOS"]
        # compute_logits (t3.py:725-739) at cfg 0.0 / 0.5 / 1.3, fp32 and bf16 (the dtype the reference runs in)
        head, hid = gc.speech_head(), gc.hidden_rows(2, 5)
        for dt, tag in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
            for cfg in gc.CFG_SCALES:
                model = reference_t3_model(ref, "en", cfg, head=head, dtype=dt)
                with torch.no_grad():
                    lg = model.compute_logits(hid.to(dt), None)
                assert lg.shape == (2, 9194) and torch.isinf(lg[:, :1000]).all() and (lg[:, :1000] < 0).all()
                body = lg[:, 1000:].float().numpy()
                key = f"logits/{tag}/cfg{cfg}"
                out[key + "/merged"] = body.astype(np.float32) if dt == torch.float32 else lg[:, 1000:].view(torch.int16).numpy()
                if cfg == gc.CFG_SCALES[0]:
                    c, u = model.seen_logits
                    out[f"logits/{tag}/cond"] = c.float().numpy() if dt == torch.float32 else c.view(torch.int16).numpy()
                    out[f"logits/{tag}/uncond"] = u.float().numpy() if dt == torch.float32 else u.view(torch.int16).numpy()
    finally:
        restore()
    np.savez_compressed(os.path.join(OUT, "t3_glue.npz"), **out)
    print("t3_glue keys:", len(out), "bytes:", os.path.getsize(os.path.join(OUT, "t3_glue.npz")))


def mtl_tokenizer():
    """The reference's own MTLTokenizer (models/t3/mtl_tokenizer.py, executed from /root/reference on its packaged
    mtl_tokenizer.json) on prompts in the languages whose normalisers need no third-party package here
    -> tests/golden/mtl_tokenizer.json.  (HF_HUB_OFFLINE: its Cangjie download fails and it carries on, as it would
    without network.)"""
    import types
    os.environ["HF_HUB_OFFLINE"] = "1"
    base = f"{REF}/src/chatterbox_vllm"
    saved = dict(sys.modules)
    try:
        pkg = types.ModuleType("chatterbox_vllm")
        pkg.__path__ = [base]
        sys.modules["chatterbox_vllm"] = pkg
        spec = importlib.util.spec_from_file_location("chatterbox_vllm.text_utils", base + "/text_utils.py")
        tu = importlib.util.module_from_spec(spec)
        sys.modules["chatterbox_vllm.text_utils"] = tu
        spec.loader.exec_module(tu)
        spec = importlib.util.spec_from_file_location("ref_mtl_tokenizer", base + "/models/t3/mtl_tokenizer.py")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        tok = mod.MTLTokenizer(vocab_file=base + "/models/t3/mtl_tokenizer.json")
        cases = [("[START]Hello there, General Kenobi.[STOP]", None), ("[START]Hello there.[STOP]", "en"),
                 ("[fr][START]Bonjour le monde, ça va ?[STOP]", None), ("[de][START]Über den Wolken.[STOP]", None),
                 ("[ko][START]안녕하세요. 반갑습니다.[STOP]", None), ("[START]안녕하세요.[STOP]", "ko"),
                 ("[es][START]¿Dónde está la biblioteca?[STOP]", None), ("[xx][START]Not a language tag.[STOP]", None),
                 ("[START]ＦＵＬＬ width and ﬁ ligature[STOP]", "en"), ("[tr][START]İstanbul'da hava güzel.[STOP]", None)]
        out = [dict(text=t, language_id=l, ids=[int(i) for i in tok.encode(t, language_id=l)]) for t, l in cases]
        json.dump(dict(vocab_size=tok.get_vocab_size(), cases=out), open(f"{OUT}/mtl_tokenizer.json", "w"), ensure_ascii=False)
    finally:
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        sys.modules.update(saved)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[1:]
    for fn in (hf_backbone, rope, sampler, block_pool, prompt_lengths, cond_enc, t3_glue, mtl_tokenizer):
        if only and fn.__name__ not in only:
            continue
        try:
            fn()
            print("ok  ", fn.__name__)
        except Exception as e:
            import traceback
            traceback.print_exc()
            print("FAIL", fn.__name__, repr(e))
