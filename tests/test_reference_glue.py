"""Pins oracle/t3_ref.py, oracle/prompt_ref.py AND the product's host glue
(chatterbox_vllm_b200/models/t3/processor.py) to tests/golden/t3_glue.npz — the output of the
REFERENCE'S OWN models/t3/t3.py functions (T3MultiModalProcessor.apply, split_prefill_decode,
get_input_embeddings, compute_logits, create_triangular_matrix), executed from /root/reference by
oracle/make_golden.py::t3_glue on the seeded inputs of oracle/glue_cases.py.  Gathers and single
fp32 / bf16 adds: everything here is held bit-exact.  CPU only; nothing reads /root/reference."""
import os

import numpy as np
import pytest
import torch

from chatterbox_vllm_b200.models.t3 import processor as proc
from oracle import glue_cases as gc
from oracle import prompt_ref
from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights, cfg_merge


@pytest.fixture(scope="module")
def G(golden_dir):
    return np.load(os.path.join(golden_dir, "t3_glue.npz"))


def _tables(variant):
    tb = gc.tables(variant)
    return tb, proc.EmbeddingTables(tb["text_emb"], tb["text_pos"], tb["speech_emb"], tb["speech_pos"])


def _same(a, G, key, stride=4099):
    a = a.numpy() if torch.is_tensor(a) else np.asarray(a)
    if key + "_sample" in G.files:                      # readable failure before the digest
        np.testing.assert_array_equal(gc.sample(a, stride), G[key + "_sample"], err_msg=key)
    assert gc.digest(a) == str(G[key + "_digest"]), key


def test_triangular_code_and_constants(G):
    np.testing.assert_array_equal(proc.create_triangular_matrix(5, 9).numpy(), G["tri_5x9"])
    assert gc.digest(proc.create_triangular_matrix(40, 1024).numpy()) == str(G["tri_40x1024_digest"])
    assert (proc.PREFILL_COND_START_TOKEN, proc.PREFILL_COND_END_TOKEN, proc.PREFILL_END_TOKEN, proc.CONDITIONING_SIZE,
            proc.SPEECH_TOKEN_OFFSET) == (695, 696, 697, 34, 1000)


@pytest.mark.parametrize("case", gc.prefix_cases(), ids=lambda c: c[0])
def test_prompt_layout_matches_reference_apply(G, case):
    name, variant, n_text, ids_seed, cond_seed = case
    ids, cond = gc.text_ids(variant, n_text, ids_seed), gc.cond_emb(cond_seed)
    np.testing.assert_array_equal(np.array(ids), G[f"{name}/text_ids"])
    want_ids = G[f"{name}/final_prompt_ids"].tolist()
    # the oracle restatement
    assert prompt_ref.final_prompt_ids(ids) == want_ids
    _same(prompt_ref.side_tensor(cond.numpy(), len(ids)), G, f"{name}/side")
    # the product: the function and the processor class with the reference's apply() signature
    f_ids, side = proc.build_prefill_inputs(ids, cond)
    assert f_ids == want_ids
    _same(side, G, f"{name}/side")
    np.testing.assert_array_equal(side[34:].sum(1).numpy(), G[f"{name}/side_rowsum_tail"])
    r = proc.T3MultiModalProcessor(tokenizer=None).apply(ids, {"conditionals": [cond]}, {})
    assert r["prompt_token_ids"] == want_ids
    _same(r["mm_kwargs"]["conditionals"][0], G, f"{name}/side")
    ph = r["mm_placeholders"]["conditionals"][0]
    assert [ph["offset"], ph["length"]] == G[f"{name}/placeholder"].tolist()


@pytest.mark.parametrize("case", gc.prefix_cases(), ids=lambda c: c[0])
def test_input_embeddings_match_reference_full_block_chunks_and_mixed(G, case):
    name, variant, n_text, ids_seed, cond_seed = case
    ids, cond = gc.text_ids(variant, n_text, ids_seed), gc.cond_emb(cond_seed)
    tb, tables = _tables(variant)
    f_ids, side = proc.build_prefill_inputs(ids, cond)
    ids_t, P = torch.tensor(f_ids), len(f_ids)
    full = proc.assemble_input_embeddings(ids_t, [side], tables)
    _same(full, G, f"{name}/full")
    if f"{name}/full" in G.files:
        np.testing.assert_array_equal(full.numpy(), G[f"{name}/full"])
    # the oracle's prefill_embeds (what the CUDA prefill_embed kernel is held to on the GPU) = the two halves
    w = T3Weights(T3ModelConfig.multilingual() if variant == "mtl" else T3ModelConfig())
    w.text_emb, w.text_pos, w.speech_emb, w.speech_pos = tb["text_emb"], tb["text_pos"], tb["speech_emb"], tb["speech_pos"]
    c, u = _oracle_prefill(w, cond, ids)
    assert gc.digest(torch.cat([c, u], dim=1).numpy()) == str(G[f"{name}/full_digest"])
    # every slice of every chunk split, each as its own call like a chunked-prefill scheduler makes
    n_chunks = 0
    for cuts in gc.chunk_splits(P):
        edges = [0, *cuts, P]
        for a, b in zip(edges[:-1], edges[1:]):
            e = proc.assemble_input_embeddings(ids_t[a:b], [side[a:b]], tables)
            assert e.shape == (b - a, 2048)
            assert gc.digest(e.numpy()) == str(G[f"{name}/chunk_{a}_{b}_digest"]), (name, a, b)
            n_chunks += 1
    assert n_chunks >= 20
    # mixed batch: decode rows, a whole prefix, a decode row — segmentation and embeddings
    dec = torch.tensor([1000 + 17, 1000 + 6561, 1000 + 4242])
    mixed_ids = torch.cat([dec[:2], ids_t, dec[2:]])
    segs = proc.split_prefill_decode(mixed_ids, [side])
    assert [len(s[0]) for s in segs] == G[f"{name}/mixed_seg_lens"].tolist()
    assert [s[1] is not None for s in segs] == G[f"{name}/mixed_seg_prefill"].tolist()
    _same(proc.assemble_input_embeddings(mixed_ids, [side], tables), G, f"{name}/mixed")
    pure = proc.assemble_input_embeddings(dec, None, tables)
    _same(pure, G, f"{name}/decode", 211)
    o = _oracle(w)
    e = o.decode_embed(dec - 1000)
    assert gc.digest(torch.cat([e, e], dim=1).numpy()) == str(G[f"{name}/decode_digest"])


def _oracle(w):
    o = T3Oracle.__new__(T3Oracle)
    o.w, o.cfg, o.mode = w, w.cfg, "fp32"
    return o


def _oracle_prefill(w, cond, ids):
    return _oracle(w).prefill_embeds(cond, torch.tensor(ids))


def test_two_prefixes_in_one_batch_and_strip_eos(G, monkeypatch):
    (n0, _, t0, s0, c0), (n1, _, t1, s1, c1) = gc.prefix_cases()[0], gc.prefix_cases()[1]
    _, tables = _tables("en")
    reqs = [proc.build_prefill_inputs(gc.text_ids("en", t, s), gc.cond_emb(c)) for t, s, c in ((t0, s0, c0), (t1, s1, c1))]
    cut = int(G["two_prefixes/cut"])
    ids2 = torch.tensor(reqs[0][0] + reqs[1][0][:cut])
    mme = [reqs[0][1], reqs[1][1][:cut]]
    assert [len(s[0]) for s in proc.split_prefill_decode(ids2, mme)] == G["two_prefixes/seg_lens"].tolist()
    assert gc.digest(proc.assemble_input_embeddings(ids2, mme, tables).numpy()) == str(G["two_prefixes/digest"])
    monkeypatch.setenv("CHATTERBOX_STRIP_LLM_EOS", "1")
    segs = proc.split_prefill_decode(torch.tensor([695, 2, 5, 2, 697, 1003, 2, 1004]), [torch.zeros(5, 1024)])
    assert [t for s in segs for t in s[0].tolist()] == G["strip_eos/seg_ids"].tolist()
    assert [len(s[0]) for s in segs] == G["strip_eos/seg_lens"].tolist()


@pytest.mark.parametrize("cfg", gc.CFG_SCALES)
def test_cfg_merge_matches_reference_compute_logits(G, cfg):
    """t3.py:732 at cfg 0.0 / 0.5 / 1.3: fp32 and bf16 (the dtype the reference runs in), bit for bit, from the
    cond / uncond logits the reference's own head produced."""
    c32, u32 = torch.from_numpy(G["logits/f32/cond"]), torch.from_numpy(G["logits/f32/uncond"])
    np.testing.assert_array_equal(cfg_merge(c32, u32, cfg, "fp32").numpy(), G[f"logits/f32/cfg{cfg}/merged"])
    bits = lambda k: torch.from_numpy(G[k]).view(torch.bfloat16).float()  # noqa: E731
    c16, u16 = bits("logits/bf16/cond"), bits("logits/bf16/uncond")
    got = cfg_merge(c16, u16, cfg, "bf16")
    assert torch.equal(got, bits(f"logits/bf16/cfg{cfg}/merged"))
    if cfg == 0.0:
        assert torch.equal(got, c16)


def test_oracle_head_matches_reference_logits_processor(G):
    """The head GEMM: fp32 accumulate, one rounding to the model dtype (1 bf16 ulp = summation order)."""
    w = T3Weights(T3ModelConfig())
    w.speech_head = gc.speech_head()
    hid = gc.hidden_rows(2, 5)
    o = _oracle(w)
    lg = o.head(hid[:, :1024])
    np.testing.assert_allclose(lg.numpy(), G["logits/f32/cond"], rtol=0, atol=2e-4)
    o.mode = "bf16"
    w.speech_head = gc.speech_head().to(torch.bfloat16)
    lg16 = o.head(hid[:, 1024:].to(torch.bfloat16).float())
    want = torch.from_numpy(G["logits/bf16/uncond"]).view(torch.bfloat16).float()
    assert ((lg16 - want).abs() <= 2.0 ** -7 * want.abs() + 1e-6).all()
