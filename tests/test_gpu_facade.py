"""T3VllmModel (the vLLM-facing model class) on the GPU against the CPU oracle: embeddings bit-exact,
prefill / decode logits within the bf16 tolerance of the engine tests, chunked prefill, two requests
decoding in one batch, block recycling."""
import numpy as np
import pytest
import torch

from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights

pytestmark = pytest.mark.gpu

HEAD_MAX_ABS = 3e-2      # as tests/test_gpu_engine.py (relative to the logit scale, see there)


@pytest.fixture(scope="module")
def setup():
    from chatterbox_vllm_b200.models.t3 import T3VllmModel, build_prefill_inputs
    cfg = T3ModelConfig(n_layers=3)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    m = T3VllmModel(vllm_config=None, prefix="", max_num_seqs=4, max_model_len=256)
    loaded = m.load_weights(w.state_dict().items())
    assert "tfmr.layers.0.self_attn.q_proj.weight" in loaded and "speech_head.weight" in loaded
    return w, m, build_prefill_inputs


def _req(seed, n_text):
    g = torch.Generator().manual_seed(seed)
    text = [255] + torch.randint(3, 600, (n_text - 2,), generator=g).tolist() + [0]
    return text, torch.randn(34, 1024, generator=g) * 0.02


def _close(got, want, what):
    got, want = got.float().cpu(), want.float()
    scale = max(1.0, float(want.abs().max()))
    err = float((got - want).abs().max())
    assert err <= HEAD_MAX_ABS * scale * 4, f"{what}: max-abs {err:.4f} at logit scale {scale:.2f}"


def test_facade_prefill_decode_matches_oracle(setup):
    w, m, build = setup
    o = T3Oracle(w, "bf16")
    text, cond = _req(1, 19)
    ids, side = build(text, cond)
    P = len(ids)
    emb = m.get_input_embeddings(torch.tensor(ids), m.get_multimodal_embeddings(conditionals=[[side]]))
    c, u = o.prefill_embeds(cond, torch.tensor(text))
    assert torch.equal(emb.float().cpu(), torch.cat((c, u), 1)), "input embeddings must be bit-exact"
    hs = m.forward(None, torch.arange(P), None, emb, seq_ids=torch.zeros(P, dtype=torch.int32))
    assert hs.shape == (P, 2048)
    lg = m.compute_logits(hs[-1:], None)
    assert lg.shape == (1, 1000 + 8194) and torch.isinf(lg[0, :1000]).all() and (lg[0, :1000] < 0).all()
    pair = o.new_pair()
    want = o.prefill_pair(pair, cond, text, m.cfg_scale)
    _close(lg[0, 1000:], want, "prefill logits")
    tok = int(lg[0].argmax())
    assert tok >= 1000
    for step in range(6):                                     # decode: ids live in the +1000 space
        want = o.decode_pairs([pair], [tok - 1000], m.cfg_scale)[0]
        emb = m.get_input_embeddings(torch.tensor([tok]), [])
        hs = m.forward(None, torch.tensor([P + step]), None, emb, seq_ids=torch.tensor([0]))
        lg = m.compute_logits(hs)
        _close(lg[0, 1000:], want, f"decode step {step}")
        tok = 1000 + int(want.argmax())                       # teacher-force the oracle's token
    m.free_sequences([0])
    assert m.engine.pool.num_free() == m.engine.num_blocks - 1


def test_facade_chunked_prefill_and_batched_decode(setup):
    w, m, build = setup
    o = T3Oracle(w, "bf16")
    reqs = [_req(5, 30), _req(6, 11)]
    pairs, toks, lens = [], [], []
    for slot, (text, cond) in enumerate(reqs):
        ids, side = build(text, cond)
        P = len(ids)
        ids_t = torch.tensor(ids)
        cut = 20 if slot == 0 else 40                         # inside the cond rows / inside the text
        last = None
        for a, b in ((0, cut), (cut, P)):
            emb = m.get_input_embeddings(ids_t[a:b], [side[a:b]])
            last = m.forward(None, torch.arange(a, b), None, emb, seq_ids=torch.full((b - a,), slot, dtype=torch.int32))
        lg = m.compute_logits(last[-1:])
        pair = o.new_pair()
        want = o.prefill_pair(pair, cond, text, m.cfg_scale)
        _close(lg[0, 1000:], want, f"chunked prefill logits of request {slot}")
        pairs.append(pair), toks.append(int(want.argmax())), lens.append(P)
    for step in range(4):                                     # both requests in one decode batch (T = 4 rows)
        want = o.decode_pairs(pairs, toks, m.cfg_scale)
        emb = m.get_input_embeddings(torch.tensor([1000 + t for t in toks]), [])
        hs = m.forward(None, torch.tensor([lens[0] + step, lens[1] + step]), None, emb, seq_ids=torch.tensor([0, 1]))
        lg = m.compute_logits(hs)
        for r in range(2):
            _close(lg[r, 1000:], want[r], f"batched decode step {step} request {r}")
        toks = [int(x.argmax()) for x in want]
    m.free_sequences([0, 1])
    assert m.engine.pool.num_free() == m.engine.num_blocks - 1


def test_facade_registration_is_guarded():
    from chatterbox_vllm_b200.models import t3
    assert t3.register_with_vllm() in (True, False)           # never raises, with or without vLLM
