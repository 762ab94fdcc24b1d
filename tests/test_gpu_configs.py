"""BASELINE.json's own configurations on the GPU at the real depth (30 layers), shortened to <= 48 decode steps:
config 2 (64 CFG pairs, prefix 200), the metric's M1 (128 pairs, prefix 200), config 4 (multilingual vocabulary, a
prefill + decode mix through recycled slots) and config 5 (max_model_len 1000, 512 pairs, exaggeration 0.8, pool at
>= 95 % occupancy, long contexts).  In every run three probed pairs — first, middle and last slot — are held to the
CPU oracle, teacher-forced with the engine's own tokens, with the 30-layer bars of test_gpu_engine.py (every token
the oracle's argmax or within 8 bf16 ulps of its top logit, >= 80 % exact agreement: random-init logits are near
flat); the other pairs are checked through size-independent properties (exact counts, pool restored)."""
import numpy as np
import pytest
import torch

from oracle.t3_ref import T3ModelConfig, T3Weights

try:
    from test_gpu_engine import assert_stream_parity
except ImportError:
    from tests.test_gpu_engine import assert_stream_parity

pytestmark = pytest.mark.gpu
STEPS = 24


@pytest.fixture(scope="module")
def w30():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return {"en": T3Weights.random(T3ModelConfig(n_layers=30), seed=0, head_scale=16.0, norm_jitter=0.1),
            "mtl": T3Weights.random(T3ModelConfig.multilingual(n_layers=30), seed=1, head_scale=16.0, norm_jitter=0.1)}


def _reqs(n, n_text, max_tokens, *, vocab=704, seed=0, lang_tag=None, penalize=False, cond_row33=None):
    from chatterbox_vllm_b200.engine import Request, SamplingParams
    g = torch.Generator().manual_seed(seed)
    out = []
    for i in range(n):
        nt = n_text[i % len(n_text)] if isinstance(n_text, (list, tuple)) else n_text
        body = torch.randint(3, vocab, (nt - 2,), generator=g).tolist()
        body = [t if t not in (255, 695, 696, 697) else 3 for t in body]
        text = ([lang_tag, 255] + body[:nt - 3] + [0]) if lang_tag else ([255] + body + [0])
        cond = torch.randn(34, 1024, generator=g) * 0.02
        if cond_row33 is not None:
            cond[33] = cond_row33
        mt = max_tokens[i % len(max_tokens)] if isinstance(max_tokens, (list, tuple)) else max_tokens
        out.append(Request(text, cond, SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=mt, ignore_eos=True),
                           uid=1000 + i, penalize_prompt_ids=penalize))
    return out


def _probe(eng, w, reqs, got, idx):
    for i in idx:
        n = reqs[i].params.max_tokens
        assert len(got[i]) == n
        assert_stream_parity([got[i]], w, [reqs[i]], eng.cfg_scale, n, first_exact=0, min_agree=0.8, ulps=8.0,
                             temperature=0.0, repetition_penalty=2.0)


@pytest.mark.parametrize("pairs", [64, 128], ids=["config2_64pairs", "M1_128pairs"])
def test_baseline_decode_configs_prefix200(w30, pairs):
    from chatterbox_vllm_b200.engine import T3Engine
    w = w30["en"]
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=pairs, max_model_len=200 + STEPS + 8, max_rows=16384)
    reqs = _reqs(pairs, 165, STEPS, seed=pairs)                       # n_text 165 -> P = 200 (SURVEY §8 d)
    got = eng.generate(reqs, poll_every=16)
    assert all(len(t) == STEPS for t in got) and eng.pool.num_free() == eng.num_blocks - 1
    _probe(eng, w, reqs, got, (0, pairs // 2, pairs - 1))
    eng.close()


def test_baseline_config4_multilingual_prefill_decode_mix(w30):
    """128 multilingual requests (text ids up to 2453, language tag first, prompt ids >= 1000 in the penalty mask)
    through 96 slots with ragged lengths: new pairs are prefilled while the others decode."""
    from chatterbox_vllm_b200.engine import T3Engine
    w = w30["mtl"]
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=96, max_model_len=256, max_rows=16384)
    reqs = _reqs(128, [60, 150, 35, 110], [STEPS, 16, 20, 12], vocab=2454, seed=4, lang_tag=708, penalize=True)
    got = eng.generate(reqs, poll_every=8)
    assert [len(t) for t in got] == [r.params.max_tokens for r in reqs] and eng.pool.num_free() == eng.num_blocks - 1
    # the oracle's sampler sees the prompt-penalty mask the engine builds for multilingual requests
    for i in (0, 64, 127):
        n = reqs[i].params.max_tokens
        seen0 = [t - 1000 for t in reqs[i].text_ids if t >= 1000]
        assert seen0, "multilingual prompts carry ids >= 1000"
        assert_stream_parity([got[i]], w, [reqs[i]], eng.cfg_scale, n, first_exact=0, min_agree=0.8, ulps=8.0,
                             prompt_seen=[seen0], temperature=0.0, repetition_penalty=2.0)
    eng.close()


def test_baseline_config5_long_context_high_occupancy(w30):
    """512 pairs, max_model_len 1000, exaggeration 0.8 (row 33 of the conditioning replaced), KV pool sized so that
    >= 95 % of it is occupied at the end; the probed pairs carry ~900-token prefixes (57+ pages per sequence)."""
    from chatterbox_vllm_b200.engine import T3Engine
    w = w30["en"]
    g = torch.Generator().manual_seed(9)
    row33 = 0.8 * torch.randn(1024, generator=g) * 0.02               # emotion_adv_fc(0.8): Linear(1 -> 1024), no bias
    n_text = [60] * 512
    for i in (0, 255, 511):
        n_text[i] = 860 + 7 * (i % 5)                                 # P ~ 900
    reqs = _reqs(512, n_text, STEPS, seed=5, cond_row33=row33)
    blocks = 2 * sum(-(-(34 + n + 1 + STEPS) // 16) for n in n_text)
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=512, max_model_len=1000, num_blocks=int(blocks / 0.955) + 1,
                   max_rows=16384)
    peak = {"used": 0}
    orig = eng.decode

    def spy(*a):
        peak["used"] = max(peak["used"], eng.num_blocks - 1 - eng.pool.num_free())
        return orig(*a)
    eng.decode = spy
    got = eng.generate(reqs, poll_every=16)
    assert all(len(t) == STEPS for t in got) and eng.pool.num_free() == eng.num_blocks - 1
    assert peak["used"] >= 0.95 * (eng.num_blocks - 1), f"pool occupancy peaked at {peak['used'] / (eng.num_blocks - 1):.3f}"
    _probe(eng, w, reqs, got, (0, 255, 511))
    eng.close()
