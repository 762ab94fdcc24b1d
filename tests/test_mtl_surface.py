"""CPU tests of the multilingual surface (chatterbox_vllm_b200.mtl_tts.ChatterboxMultilingualTTS, the MTL tokenizer,
ChatterboxTTS.from_local(variant=...) with a built-in voice, and the `chatterbox_vllm` import alias) against the
reference's rules (mtl_tts.py:317-520, tts.py:91-206) — the engine is a recording fake, no GPU compute — and the
reference's OWN MTLTokenizer output (tests/golden/mtl_tokenizer.json, oracle/make_golden.py::mtl_tokenizer)."""
import json
import os

import pytest
import torch

from chatterbox_vllm_b200 import tts as tts_mod
from chatterbox_vllm_b200.mtl_tts import ChatterboxMultilingualTTS
from chatterbox_vllm_b200.text_utils import SUPPORTED_LANGUAGES
from oracle import prompt_ref

REF_MTL_TOK = "/root/reference/src/chatterbox_vllm/models/t3/mtl_tokenizer.json"   # absent on the GPU box: that test skips


class _Tok:
    def __init__(self):
        self.seen = []

    def encode(self, s):
        self.seen.append(s)
        return [708, 255] + [3 + (ord(c) % 2400) for c in s if c.isalpha()][:20] + [0]


class _Eng:
    def __init__(self):
        self.calls = []

    def generate(self, reqs):
        self.calls.append(reqs)
        return [[1005, 1006, 1007, 7562] for _ in reqs]

    def close(self):
        pass


def _mk(**kw):
    fc = torch.arange(1024, dtype=torch.float32).view(1024, 1) / 1024
    return ChatterboxMultilingualTTS("cuda", 1000, _Eng(), _Tok(), emotion_adv_fc=fc, default_cond_emb=torch.zeros(34, 1024),
                                     variant="multilingual", **kw)


def test_multilingual_generate_defaults_language_lists_and_kwargs(monkeypatch):
    t = _mk()
    assert ChatterboxMultilingualTTS.get_supported_languages() == SUPPORTED_LANGUAGES and len(SUPPORTED_LANGUAGES) == 23
    out = t.generate(["bonjour le monde", "hello there"], language_id=["fr", "en"])
    assert [o.tolist() for o in out] == [[5], [5]]
    reqs = t.t3.calls[0]
    # mtl_tts.py:404-409: per-prompt "[lang][START]…[STOP]"
    assert t.tokenizer.seen == ["[fr][START]Bonjour le monde.[STOP]", "[en][START]Hello there.[STOP]"]
    sp = reqs[0].params
    # defaults 0.4 / 0.4 (mtl_tts.py:322-327) stay under the 0.5 clamp; repetition penalty 2.0; prompt ids penalised
    assert (sp.temperature, sp.top_p, sp.repetition_penalty) == (0.4, 0.4, 2.0) and reqs[0].penalize_prompt_ids
    assert sp.max_tokens == prompt_ref.pre_cap_tokens(t.tokenizer.seen, 1000, 1000)
    # no language at all: no tag, no clamp (mtl_tts.py:391-392,442-448)
    t.tokenizer.seen.clear()
    t.generate("plain prompt", temperature=0.9, top_p=0.95)
    assert t.tokenizer.seen == ["[START]Plain prompt.[STOP]"]
    sp = t.t3.calls[-1][0].params
    assert (sp.temperature, sp.top_p) == (0.9, 0.95)
    # ANY prompt with a language clamps the whole batch; a None entry gets no tag
    t.tokenizer.seen.clear()
    t.generate(["a", "b"], language_id=["de", None], temperature=0.9, top_p=0.95)
    assert t.tokenizer.seen == ["[de][START]A.[STOP]", "[START]B.[STOP]"]
    assert all((r.params.temperature, r.params.top_p) == (0.5, 0.5) for r in t.t3.calls[-1])
    # one code for every prompt
    t.tokenizer.seen.clear()
    t.generate(["a", "b"], language_id="KO")
    assert t.tokenizer.seen == ["[ko][START]A.[STOP]", "[ko][START]B.[STOP]"]
    # kwargs reach the sampling parameters — min_p INCLUDED, unlike tts.py's generate_with_conds (mtl_tts.py:508-518)
    t.generate("x", language_id="es", min_p=0.07, top_k=40, seed=3)
    sp = t.t3.calls[-1][0].params
    assert (sp.min_p, sp.top_k, sp.seed) == (0.07, 40, 3)
    t.generate_with_conds("x", {}, torch.zeros(34, 1024), language_id="es", min_p=0.11)
    assert t.t3.calls[-1][0].params.min_p == 0.11
    monkeypatch.setenv("CHATTERBOX_DETERMINISTIC", "1")
    t.generate("x", language_id="it")
    assert (t.t3.calls[-1][0].params.temperature, t.t3.calls[-1][0].params.top_p) == (0.0, 1.0)


def test_multilingual_validation_errors():
    t = _mk()
    with pytest.raises(ValueError, match="Unsupported language_id 'xx'"):
        t.generate("x", language_id="xx")
    with pytest.raises(ValueError, match="Unsupported language_id 'klingon'"):
        t.generate(["a", "b"], language_id=["fr", "klingon"])
    with pytest.raises(ValueError, match=r"Number of language_ids \(1\) must match number of prompts \(2\)"):
        t.generate(["a", "b"], language_id=["fr"])
    with pytest.raises(TypeError):
        t.generate("x", logprobs=3)


def test_multilingual_vocoder_and_exaggeration():
    seen = {}

    def voc(tokens, ref, n_timesteps):
        seen["args"] = (tokens.tolist(), ref, n_timesteps)
        return torch.zeros(1, 4)

    t = _mk(vocoder=voc, default_s3gen_ref={"k": 2})
    t.target_device = "cpu"
    w = t.generate("hola", language_id="es", exaggeration=0.7, diffusion_steps=4)
    assert w[0].shape == (1, 4) and seen["args"] == ([5], {"k": 2}, 4)
    assert torch.allclose(t.t3.calls[-1][0].cond_emb[33], 0.7 * torch.arange(1024) / 1024)


def test_english_class_on_the_multilingual_variant_prefixes_angle_tag():
    """tts.py:330-333: the English class with variant='multilingual' prefixes '<xx>'; language ids come from the 23."""
    fc = torch.arange(1024, dtype=torch.float32).view(1024, 1) / 1024
    t = tts_mod.ChatterboxTTS("cuda", 1000, _Eng(), _Tok(), emotion_adv_fc=fc, default_cond_emb=torch.zeros(34, 1024),
                              variant="multilingual")
    assert t.get_supported_languages() == SUPPORTED_LANGUAGES
    t.generate("salut", language_id="fr")
    assert t.tokenizer.seen == ["<fr>[START]Salut.[STOP]"] and t.t3.calls[-1][0].penalize_prompt_ids


def test_builtin_voice_from_conds_pt_is_used_when_no_audio_prompt(tmp_path, monkeypatch):
    """tts.py:151,217-220 / ADVICE r1: from_local loads conds.pt; generate() without audio_prompt_path runs its speaker
    embedding + prompt tokens through the conditioning encoder and hands its 'gen' dict to the vocoder."""
    from safetensors.torch import save_file
    sd = {"w": torch.zeros(1)}
    save_file(sd, str(tmp_path / "t3_cfg.safetensors"))
    save_file(sd, str(tmp_path / "t3_23lang.safetensors"))
    save_file(sd, str(tmp_path / "t3_mtl23ls_v2.safetensors"))
    spk, toks = torch.randn(1, 256), torch.randint(0, 6561, (1, 150))
    torch.save({"t3": {"speaker_emb": spk, "cond_prompt_speech_tokens": toks, "emotion_adv": 0.5 * torch.ones(1, 1, 1)},
                "gen": {"prompt_token": torch.arange(4)}}, str(tmp_path / "conds.pt"))
    built = {}

    def fake_from_state_dict(cls, state_dict, tokenizer, target_device="cuda", max_model_len=1000, max_batch_size=10,
                             variant="english", **kwargs):
        built.update(variant=variant, keys=sorted(state_dict), tok=tokenizer, max_batch_size=max_batch_size)
        fc = torch.ones(1024, 1)
        return cls(target_device, max_model_len, _Eng(), tokenizer, emotion_adv_fc=fc, variant=variant,
                   default_conds=kwargs.get("default_conds"), vocoder=kwargs.get("vocoder"))

    monkeypatch.setattr(tts_mod.ChatterboxTTS, "from_state_dict", classmethod(fake_from_state_dict))
    seen = {}

    class Enc:                                                  # stands in for cond_enc.T3CondEncoder (GPU kernels)
        def __call__(self, speaker_emb, prompt_tokens, emotion):
            seen["enc"] = (speaker_emb.shape, prompt_tokens.shape, emotion)
            return torch.full((34, 1024), 0.25)

    for cls, variant, kw in ((tts_mod.ChatterboxTTS, "english", {}), (tts_mod.ChatterboxTTS, "multilingual", {"variant": "multilingual"}),
                             (ChatterboxMultilingualTTS, "multilingual", {})):
        voc = lambda tokens, ref, n_timesteps: ref["prompt_token"]           # noqa: E731
        t = cls.from_local(tmp_path, target_device="cpu", max_batch_size=4, tokenizer=_Tok(), vocoder=voc, **kw)
        assert built["variant"] == variant and built["max_batch_size"] == 4 and isinstance(built["tok"], _Tok)
        t.cond_encoder = Enc()
        ref, cond = t.get_audio_conditionals(None)
        assert seen["enc"] == (torch.Size([256]), torch.Size([150]), 0.5) and torch.equal(cond, torch.full((34, 1024), 0.25))
        out = t.generate("hello", language_id="en")
        assert torch.equal(out[0], torch.arange(4)) and torch.equal(t.t3.calls[-1][0].cond_emb, cond)
    with pytest.raises(ValueError, match="variant 'klingon'"):
        tts_mod.ChatterboxTTS.from_local(tmp_path, variant="klingon", tokenizer=_Tok())


def test_reference_import_paths_resolve():
    from chatterbox_vllm.tts import ChatterboxTTS                    # example-tts.py:1-8
    from chatterbox_vllm.mtl_tts import ChatterboxMultilingualTTS as M, SUPPORTED_LANGUAGES as L
    from chatterbox_vllm.models.t3 import SPEECH_TOKEN_OFFSET, register_with_vllm
    from chatterbox_vllm.models.t3.entokenizer import EnTokenizer
    from chatterbox_vllm.models.t3.mtl_tokenizer import MTLTokenizer
    from chatterbox_vllm.text_utils import punc_norm
    assert ChatterboxTTS is tts_mod.ChatterboxTTS and M is ChatterboxMultilingualTTS and L is SUPPORTED_LANGUAGES
    assert SPEECH_TOKEN_OFFSET == 1000 and callable(register_with_vllm) and punc_norm("a") == "A."
    assert EnTokenizer.__name__ == "EnTokenizer" and MTLTokenizer.__name__ == "MTLTokenizer"


@pytest.mark.skipif(not os.path.exists(REF_MTL_TOK), reason="reference mtl_tokenizer.json not present")
def test_mtl_tokenizer_matches_the_reference_tokenizer(golden_dir):
    from chatterbox_vllm_b200.tokenizer import MTLTokenizer, korean_normalize
    g = json.load(open(os.path.join(golden_dir, "mtl_tokenizer.json")))
    tok = MTLTokenizer(REF_MTL_TOK)
    assert tok.vocab_size == g["vocab_size"]
    for c in g["cases"]:
        assert tok.encode(c["text"], c["language_id"]) == c["ids"], c["text"]
    # tts.py:330-333's "<xx>" prefix is parsed like "[xx]"
    assert tok.encode("<fr>[START]Bonjour.[STOP]") == tok.encode("[fr][START]Bonjour.[STOP]")
    assert tok.prepare("[START]Hi There[STOP]", "en") == "[en][START]hi[SPACE]there[STOP]"
    assert korean_normalize("한") == "\u1112\u1161\u11ab"          # syllable -> initial / medial / final Jamo
    assert tok.decode(tok.encode("[START]hello there.[STOP]")) == "[START]hello there."
