"""Pins the CPU oracle against the committed golden fixtures (oracle/make_golden.py):
HF LlamaModel, vLLM 0.22 sampler/rope/block-pool code, and the reference's own
tokenizer + README token totals.  CPU only."""
import json
import os

import numpy as np
import torch

from oracle import paging_ref, prompt_ref, sampler_ref
from oracle.t3_ref import SeqState, T3ModelConfig, T3Oracle, T3Weights, llama3_inv_freq


def test_backbone_matches_hf_llama(golden_dir):
    g = np.load(os.path.join(golden_dir, "hf_backbone.npz"))
    cfg = T3ModelConfig(n_layers=2)
    o = T3Oracle(T3Weights.random(cfg, seed=0, norm_jitter=0.1), mode="fp32")
    a, b = SeqState(cfg), SeqState(cfg)
    x, x2 = torch.from_numpy(g["x"]), torch.from_numpy(g["x2"])
    h = o.forward_rows(torch.cat((x[0], x[1]), 0), [(a, 20), (b, 20)])
    assert np.abs(h.numpy() - g["h_prefill"].reshape(40, 1024)).max() < 2e-5
    h2 = o.forward_rows(torch.cat((x2[0], x2[1]), 0), [(a, 1), (b, 1)])
    assert np.abs(h2.numpy() - g["h_decode"].reshape(2, 1024)).max() < 2e-5
    assert a.len == 21 and b.len == 21


def test_rope_inv_freq_matches_vllm_and_hf(golden_dir):
    g = np.load(os.path.join(golden_dir, "rope_llama3.npz"))
    mine = llama3_inv_freq(T3ModelConfig()).numpy()
    np.testing.assert_allclose(mine, g["hf_inv_freq"], rtol=1e-6)
    np.testing.assert_array_equal(mine, g["vllm_inv_freq"])


def test_sampler_chain_matches_vllm(golden_dir):
    g = np.load(os.path.join(golden_dir, "sampler_vllm.npz"))
    logits, seen = g["logits"], g["seen"]
    for r in range(logits.shape[0]):
        pen = sampler_ref.apply_repetition_penalty(logits[r], seen[r], 2.0)
        np.testing.assert_array_equal(pen, g["penalised"][r])
        for ci, (T, min_p, top_p, top_k) in enumerate(g["cases"]):
            keep, _, margin = sampler_ref.kept_set(pen, T, min_p, top_p, int(top_k), return_margin=True)
            clear = margin > 1e-5          # skip fp-ambiguous top-p boundary tokens
            assert (keep == g[f"keep_{ci}"][r])[clear].all(), (r, ci)
            assert (keep != g[f"keep_{ci}"][r]).sum() <= 1
            if (keep == g[f"keep_{ci}"][r]).all():
                p = sampler_ref.final_probs(logits[r], seen[r], repetition_penalty=2.0, temperature=T,
                                            min_p=min_p, top_p=top_p, top_k=int(top_k))
                np.testing.assert_allclose(p, g[f"probs_{ci}"][r], atol=2e-6)


def test_greedy_is_argmax_of_penalised_logits():
    rng = np.random.default_rng(0)
    l = rng.standard_normal(8194).astype(np.float32)
    seen = np.zeros(8194, bool)
    top = int(np.argmax(l))
    assert sampler_ref.sample_row(l, seen, temperature=0.0) == top
    seen[top] = True
    l2 = l.copy(); l2[top] = abs(l2[top])
    second = int(np.argsort(-l2)[1])
    got = sampler_ref.sample_row(l2, seen, temperature=0.0, repetition_penalty=1e6)
    assert got == second


def test_philox_known_answer():
    # Random123 known-answer vectors for philox4x32-10
    out = sampler_ref.philox4x32_10((0, 0, 0, 0), (0, 0))
    assert [int(x) for x in out] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    out = sampler_ref.philox4x32_10((0xffffffff,) * 4, (0xffffffff, 0xffffffff))
    assert [int(x) for x in out] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    out = sampler_ref.philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0))
    assert [int(x) for x in out] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_race_sampling_distribution():
    rng = np.random.default_rng(3)
    l = (rng.standard_normal(8194) * 4).astype(np.float32)
    seen = np.zeros(8194, bool)
    kw = dict(temperature=0.7, top_p=0.9, min_p=0.02)
    p = sampler_ref.final_probs(l, seen, **kw)
    ids = np.nonzero(p)[0]
    n = 4000
    cnt = np.zeros(8194)
    for s in range(n):
        cnt[sampler_ref.sample_row(l, seen, seed=11, uid=5, step=s, **kw)] += 1
    assert cnt[p == 0].sum() == 0
    big = ids[p[ids] * n >= 5]
    chi2 = (((cnt[big] - p[big] * n) ** 2) / (p[big] * n)).sum()
    assert chi2 < 3.0 * len(big) + 20, (chi2, len(big))


def test_block_pool_matches_vllm(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "block_pool_vllm.json")))
    pool = paging_ref.BlockPool(g["num_blocks"])
    held = {}
    for op, want in zip(g["script"], g["trace"]):
        if op[0] == "alloc":
            got = pool.get_new_blocks(op[2])
            held.setdefault(op[1], []).extend(got)
            assert got == want
        else:
            pool.free_blocks(list(reversed(held.pop(op[1]))))
    assert pool.num_free() == g["free_after"]


def test_slot_mapping_formula():
    pool = paging_ref.BlockPool(40)
    bt = paging_ref.BlockTable(pool, max_seqs=3, max_model_len=100)
    bt.ensure(0, 33), bt.ensure(1, 16), bt.ensure(2, 1), bt.ensure(0, 49)
    assert bt.table[0, :4].tolist() == [1, 2, 3, 6] and bt.table[1, 0] == 4 and bt.table[2, 0] == 5
    s = paging_ref.slot_mapping(bt.table, [0, 0, 1, 2, 0, 1], [0, 17, 15, 0, 48, -1])
    assert s.tolist() == [16, 2 * 16 + 1, 4 * 16 + 15, 5 * 16, 6 * 16, -1] and s.dtype == np.int64
    bt.release(0)
    assert list(pool.free)[-4:] == [6, 3, 2, 1]


def test_prompt_layout_pins(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "prompt_lengths.json")))
    assert g["n_chunks"] == 154
    assert sum(len(prompt_ref.final_prompt_ids([255] + [3] * (n - 2) + [0])) for n in g["n_text"]) == g["readme_total"] == 28754
    ids = prompt_ref.final_prompt_ids(g["demo_ids"])
    assert len(ids) == 86 and ids[0] == 695 and ids[1:33] == [255] * 32 and ids[33] == 696 and ids[-1] == 697
    for raw, want in g["punc_norm_cases"]:
        assert prompt_ref.punc_norm(raw) == want
    st = prompt_ref.side_tensor(np.ones((34, 1024), np.float32), 51)
    assert st.shape == (86, 1024) and st[34].sum() == 1 and st[84].sum() == 51 and st[85].sum() == 0


def test_precap_and_postprocess():
    w = [prompt_ref.wrap_prompt("Hi."), prompt_ref.wrap_prompt("x" * 300)]
    assert prompt_ref.pre_cap_tokens(w, 1000, 1000) == int(np.ceil(64 * 1.6)) + 16
    assert prompt_ref.pre_cap_tokens(w[1:], 1000, 1200) == min(1000, int(np.ceil(min(1200, int(np.ceil(301 * 2.2))) * 1.6)) + 16)
    toks = [1005, 1006, 7561, 1007, 8000, 1008, 1009, 7562, 1010]
    assert prompt_ref.postprocess_tokens(toks).tolist() == [7]          # crop 2 before EOS, SOS-slice, <6561
    assert prompt_ref.postprocess_tokens([1001, 1002, 1003]).tolist() == [1, 2, 3]
    assert prompt_ref.clamp_sampling(0.8, 0.8, "en", False) == (0.5, 0.5)
    assert prompt_ref.clamp_sampling(0.8, 0.8, "en", True) == (0.0, 1.0)


def test_cond_encoder_oracle_matches_reference_modules():
    """tests/golden/cond_enc.npz was produced by the reference's OWN T3CondEnc / Perceiver modules
    (oracle/make_golden.py::cond_enc); the restatement must reproduce it from the seeds alone."""
    import os
    import numpy as np
    import torch
    from oracle import cond_ref
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cond_enc.npz"))
    for c in (0, 1):
        wseed, iseed, n = g[f"meta{c}"].tolist()
        sd = cond_ref.random_state_dict(wseed)
        speech_emb, speech_pos, spk, tok = cond_ref.random_inputs(iseed, n)
        y = cond_ref.cond_encode(sd, spk, tok, speech_emb, speech_pos, float(g[f"emotion{c}"][0]))
        want = torch.from_numpy(g[f"cond{c}"])
        assert y.shape == (34, 1024)
        assert float((y - want).abs().max()) <= 2e-6 * float(want.abs().max()), "oracle drifted from the reference modules"
        row = cond_ref.emotion_row(sd, 0.3)
        assert torch.allclose(row, torch.from_numpy(g[f"row03_{c}"]), rtol=0, atol=1e-7)
