"""Host logic of the vLLM-facing T3 model class (prompt processor, prefill/decode segmentation, the
eight prefix-slice embedding cases) against the CPU oracle.  No GPU, no vLLM."""
import numpy as np
import pytest
import torch

from chatterbox_vllm_b200.models.t3 import (CONDITIONING_SIZE, EmbeddingTables, T3MultiModalProcessor,
                                            assemble_input_embeddings, build_prefill_inputs, split_prefill_decode)
from oracle import prompt_ref
from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights


class _Tok:
    def __init__(self):
        self.seen = []

    def encode(self, s):
        self.seen.append(s)
        return [255] + [3 + (ord(c) % 600) for c in s[:20]] + [0]


def test_processor_apply_matches_reference_layout():
    tok = _Tok()
    proc = T3MultiModalProcessor(tok)
    cond = torch.randn(34, 1024)
    out = proc.apply("[START]Hello.[STOP]", {"conditionals": [cond]}, {})
    ids = tok.encode("[START]Hello.[STOP]")
    assert out["prompt_token_ids"] == prompt_ref.final_prompt_ids(ids)
    side = out["mm_kwargs"]["conditionals"][0]
    assert np.array_equal(side.numpy(), prompt_ref.side_tensor(cond.numpy(), len(ids)))
    P = CONDITIONING_SIZE + len(ids) + 1
    assert out["mm_placeholders"]["conditionals"] == [{"offset": 0, "length": P, "is_embed": None}]
    assert out["type"] == "multimodal" and len(out["mm_hashes"]["conditionals"]) == 1
    # the language tag is pushed into the prompt string once (t3.py:138-143)
    proc.apply("[START]Bonjour.[STOP]", {"conditionals": cond}, {}, {"language_id": "FR"})
    assert tok.seen[-1] == "[fr][START]Bonjour.[STOP]"
    proc.apply("[fr][START]Bonjour.[STOP]", {"conditionals": cond}, {}, {"language_id": "fr"})
    assert tok.seen[-1] == "[fr][START]Bonjour.[STOP]"
    with pytest.raises(AssertionError):
        proc.apply("x", {}, {})
    with pytest.raises(AssertionError):
        proc.apply("x", {"conditionals": [torch.zeros(33, 1024)]}, {})


def test_split_prefill_decode_uses_markers_only():
    # decode tokens, a whole prefill block with a multilingual text id >= 1000, decode tokens again
    blk = [695] + [255] * 32 + [696, 255, 1500, 7, 0, 697]
    ids = torch.tensor([1007, 2001] + blk + [1100])
    side = torch.arange(len(blk) * 4, dtype=torch.float32).view(len(blk), 4)
    segs = split_prefill_decode(ids, [side])
    assert [s[0].tolist() for s in segs] == [[1007, 2001], blk, [1100]]
    # reference behaviour: tokens ahead of the first marker are "undecided" and take side rows as
    # prefill would (t3.py:465-468); the block then gets what is left, clamped (t3.py:424-428)
    assert segs[0][1] is not None and torch.equal(segs[0][1], side[:2])
    assert torch.equal(segs[1][1], side[2:])
    assert segs[2][1] is None
    assert split_prefill_decode(torch.tensor([], dtype=torch.long), [side]) == []


@pytest.fixture(scope="module")
def model():
    cfg = T3ModelConfig(n_layers=1)
    w = T3Weights.random(cfg, seed=3)
    sd = w.state_dict()
    bf = lambda k: sd[k].to(torch.bfloat16)  # noqa: E731
    tb = EmbeddingTables(bf("text_emb.weight"), bf("text_pos_emb.emb.weight"), bf("speech_emb.weight"),
                         bf("speech_pos_emb.emb.weight"))
    return w, tb


def _full(w, cond, text):
    c, u = T3Oracle(w, "bf16").prefill_embeds(cond, torch.tensor(text))
    return torch.cat((c, u), dim=1)


def test_full_block_and_decode_embeddings_bit_exact(model):
    w, tb = model
    text = [255, 17, 300, 42, 9, 0]
    cond = torch.randn(34, 1024) * 0.02
    ids, side = build_prefill_inputs(text, cond)
    got = assemble_input_embeddings(torch.tensor(ids), [side], tb)
    assert got.dtype == torch.bfloat16 and got.shape == (len(ids), 2048)
    assert torch.equal(got.float(), _full(w, cond, text))
    dec = assemble_input_embeddings(torch.tensor([1000 + 5, 1000 + 6561]), [], tb)
    e = tb.speech_emb[torch.tensor([5, 6561])]
    assert torch.equal(dec, torch.cat([e, e], dim=1))


def test_any_chunking_of_the_prefix_reproduces_the_full_block(model):
    """vLLM's chunked prefill may cut a prefix anywhere; every slice the reference recognises must
    give the same rows as the one-pass layout."""
    w, tb = model
    text = [255] + list(range(20, 51)) + [0]                # 33 text ids
    cond = torch.randn(34, 1024) * 0.02
    ids, side = build_prefill_inputs(text, cond)
    ids_t, P = torch.tensor(ids), len(ids)
    want = _full(w, cond, text)

    def run(cuts):
        parts, a = [], 0
        for b in list(cuts) + [P]:
            parts.append(assemble_input_embeddings(ids_t[a:b], [side[a:b]], tb))
            a = b
        return torch.cat(parts, 0).float()

    for k in (1, 5, 33, 34, 35, 50, P - 2, P - 1):          # two chunks: head + tail
        assert torch.equal(run([k]), want), f"split at {k}"
    for a, b in ((34, 40), (36, 50), (40, P - 1), (20, 45)):  # three chunks: head, middle, tail (middle is text-only or spans 696)
        got = run([a, b])
        if a >= 34:
            assert torch.equal(got, want), f"splits at {a},{b}"
        else:
            # reference quirk: a middle slice that still holds cond rows is not recognisable; the reference reads
            # them as text rows at positions trunc(rowsum - 1) (t3.py:661-685; pinned against the reference's own
            # output in tests/test_reference_glue.py) — the other chunks are still exact
            assert torch.equal(got[:a], want[:a]) and torch.equal(got[b:], want[b:])
            assert not torch.equal(got[a:b], want[a:b])


def test_mixed_batch_of_prefill_and_decode_rows(model):
    w, tb = model
    text = [255, 77, 0]
    cond = torch.randn(34, 1024) * 0.02
    ids, side = build_prefill_inputs(text, cond)
    batch = torch.tensor(ids + [1000 + 12, 1000 + 99])
    got = assemble_input_embeddings(batch, [side], tb)
    assert torch.equal(got[:len(ids)].float(), _full(w, cond, text))
    e = tb.speech_emb[torch.tensor([12, 99])]
    assert torch.equal(got[len(ids):], torch.cat([e, e], dim=1))
