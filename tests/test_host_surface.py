"""CPU tests of the reference-facing host surface (chatterbox_vllm_b200.tts / text_utils /
tokenizer) against the oracle's restatement of the reference rules (oracle/prompt_ref.py) and the
committed golden fixtures.  The engine is replaced by a recording fake — no GPU compute."""
import json
import os
import random

import pytest
import torch

from chatterbox_vllm_b200 import text_utils, tts as tts_mod
from oracle import prompt_ref

REF_TOK = "/root/reference/src/chatterbox_vllm/models/t3/tokenizer.json"   # absent on the GPU box: those tests skip


def test_punc_norm_matches_oracle_rules(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "prompt_lengths.json")))
    for raw, want in g["punc_norm_cases"]:
        assert text_utils.punc_norm(raw) == want
    rng = random.Random(0)
    alphabet = list("abc XYZ.,;:!?-…—–“”‘’ \t\n") + ["...", " - ", " ,"]
    for _ in range(2000):
        s = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 30)))
        assert text_utils.punc_norm(s) == prompt_ref.punc_norm(s), repr(s)


def test_precap_and_postprocess_match_oracle(monkeypatch):
    rng = random.Random(1)
    for _ in range(200):
        prompts = [prompt_ref.wrap_prompt("x" * rng.randint(0, 700)) for _ in range(rng.randint(1, 5))]
        mt, mml = rng.choice([100, 1000, 1200]), rng.choice([300, 1000])
        assert tts_mod.pre_cap_tokens(prompts, mt, mml) == prompt_ref.pre_cap_tokens(prompts, mt, mml)
    monkeypatch.setenv("CHATTERBOX_TOKENS_PER_CHAR", "3.0")
    monkeypatch.setenv("CHATTERBOX_PRE_GUARD_MARGIN", "4")
    p = [prompt_ref.wrap_prompt("y" * 50)]
    assert tts_mod.pre_cap_tokens(p, 1000, 1000) == prompt_ref.pre_cap_tokens(p, 1000, 1000, tpc=3.0, margin=4)
    for _ in range(300):
        n = rng.randint(0, 40)
        toks = [rng.choice([6561, 6562, 7000, 8193] + list(range(0, 6561, 97))) + 1000 for _ in range(n)]
        k = rng.randint(0, 3)
        assert tts_mod.postprocess_tokens(list(toks), k).tolist() == prompt_ref.postprocess_tokens(list(toks), k).tolist()


class _FakeTok:
    def encode(self, s):
        return [255] + [3 + (ord(c) % 600) for c in s[7:-6]] + [0]


class _FakeEngine:
    def __init__(self):
        self.calls = []

    def generate(self, reqs):
        self.calls.append(reqs)
        return [[1005, 1006, 1007, 7562] for _ in reqs]

    def close(self):
        self.closed = True


def _mk(**kw):
    fc = torch.arange(1024, dtype=torch.float32).view(1024, 1) / 1024
    return tts_mod.ChatterboxTTS("cuda", 1000, _FakeEngine(), _FakeTok(), emotion_adv_fc=fc,
                                 default_cond_emb=torch.zeros(34, 1024), **kw)


def test_generate_surface_defaults_clamps_and_request_shape(monkeypatch):
    t = _mk()
    out = t.generate(["hello there", "second prompt."], exaggeration=0.5)
    assert [o.tolist() for o in out] == [[5], [5]]         # stop cut + 2-token tail crop + de-offset
    reqs = t.t3.calls[0]
    assert len(reqs) == 2 and reqs[0].text_ids[0] == 255 and reqs[0].text_ids[-1] == 0
    sp = reqs[0].params
    # tts.py:350-356 clamp (language_id defaults to 'en'), repetition_penalty 2.0, min_p swallowed
    assert (sp.temperature, sp.top_p, sp.repetition_penalty, sp.min_p) == (0.5, 0.5, 2.0, 0.0)
    wrapped = t.wrap_prompts(["hello there", "second prompt."], "en")
    assert wrapped[0] == "[START]Hello there.[STOP]"
    assert sp.max_tokens == prompt_ref.pre_cap_tokens(wrapped, 1000, 1000)
    # kwargs that reach SamplingParams in the reference
    t.generate("x", min_p=0.1, top_k=50, seed=7, skip_special_tokens=False)
    sp = t.t3.calls[-1][0].params
    # min_p is swallowed by generate_with_conds' named parameter, exactly as in the reference (tts.py:307,408-418)
    assert (sp.min_p, sp.top_k, sp.seed) == (0.0, 50, 7)
    t.generate_tokens("x", torch.zeros(34, 1024), min_p=0.1)
    assert t.t3.calls[-1][0].params.min_p == 0.1
    monkeypatch.setenv("CHATTERBOX_DETERMINISTIC", "1")
    t.generate("x")
    sp = t.t3.calls[-1][0].params
    assert (sp.temperature, sp.top_p) == (0.0, 1.0)
    with pytest.raises(ValueError, match="Unsupported language_id"):
        t.generate("x", language_id="fr")
    with pytest.raises(TypeError):
        t.generate("x", logprobs=5)
    assert t.get_supported_languages() == {"en": "English"} and t.sr == 24000
    t.shutdown()
    assert t.t3 is None


def test_update_exaggeration_and_vocoder_injection():
    seen = {}

    def voc(tokens, ref, n_timesteps):
        seen["args"] = (tokens.tolist(), ref, n_timesteps)
        return torch.zeros(1, 10)

    t = _mk(vocoder=voc, default_s3gen_ref={"k": 1})
    t.target_device = "cpu"
    cond = torch.ones(34, 1024)
    assert t.update_exaggeration(cond, 0.5) is cond
    new = t.update_exaggeration(cond, 0.8)
    assert torch.equal(new[:33], cond[:33]) and torch.allclose(new[33], 0.8 * torch.arange(1024) / 1024)
    wavs = t.generate("hi", exaggeration=0.8, diffusion_steps=5)
    assert wavs[0].shape == (1, 10) and seen["args"] == ([5], {"k": 1}, 5)
    assert torch.allclose(t.t3.calls[-1][0].cond_emb[33], 0.8 * torch.arange(1024) / 1024)


@pytest.mark.skipif(not os.path.exists(REF_TOK), reason="reference tokenizer.json not present")
def test_entokenizer_matches_survey_fixture(golden_dir):
    from chatterbox_vllm_b200.tokenizer import EnTokenizer
    g = json.load(open(os.path.join(golden_dir, "prompt_lengths.json")))
    tok = EnTokenizer(REF_TOK)
    s = "[START]" + text_utils.punc_norm("You are listening to a demo of the Chatterbox TTS model running on VLLM.") + "[STOP]"
    assert tok.encode(s) == g["demo_ids"]
    assert tok.decode(tok.encode("[START]Hello there.[STOP]")) == "[START]Hello there."


def test_guard_cap_tokens_follows_reference_formula(monkeypatch):
    """tts.py:466-486: est = clamp(ceil(2.2 * non-space chars outside [tags]), 64, 1200); cap = ceil(1.6 * est)."""
    from chatterbox_vllm_b200.tts import guard_cap_tokens
    assert guard_cap_tokens("[START]Hello world.[STOP]") == 103            # 11 chars -> est 64 -> ceil(102.4)
    long = "[START]" + "a" * 400 + "[STOP]"
    assert guard_cap_tokens(long) == 1410                                  # 400 * 2.2 = 880.0000000000001 -> 881 -> ceil(1409.6), as the reference's float math
    assert guard_cap_tokens("[START]" + "a" * 4000 + "[STOP]") == 1920     # est clamps at 1200
    monkeypatch.setenv("CHATTERBOX_TOKENS_GUARD_MULT", "2.0")
    assert guard_cap_tokens("[fr][START]x y[STOP]") == 128
