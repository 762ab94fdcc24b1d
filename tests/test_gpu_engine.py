"""GPU parity tests of the fused sampler, the prefill embedding assembly and the whole T3
step (prefill + CFG-paired decode) against the CPU oracle, through the C-ABI."""
import numpy as np
import pytest
import torch

from oracle import sampler_ref
from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights, cfg_merge

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200 import ops as _ops
    _ops._lib.load()
    return _ops


def _sampler_case(ops, B, seed, *, temperature, top_p=1.0, min_p=0.0, top_k=0, rep=2.0, cfg=0.5, scale=3.0,
                  quantize=False):
    g = torch.Generator().manual_seed(seed)
    lc = torch.randn(B, 8194, generator=g) * scale
    lu = lc + torch.randn(B, 8194, generator=g) * 0.5
    if quantize:                                         # many exact ties, as real bf16 logits have
        lc, lu = (lc * 2).round() / 2, (lu * 2).round() / 2
    logits = torch.zeros(2 * B, 8256, dtype=BF16)
    logits[:B, :8194], logits[B:, :8194] = lc.to(BF16), lu.to(BF16)
    seen = torch.rand(B, 8194, generator=g) < 0.03
    sb = ops.SamplerBuffers(B, 16, "cuda", cfg_scale=cfg)
    words = np.zeros((B, 257), np.uint32)
    for b in range(B):
        for i in np.nonzero(seen[b].numpy())[0]:
            words[b, i >> 5] |= np.uint32(1 << (i & 31))
    sb.seen.copy_(torch.from_numpy(words.view(np.int32)))
    sb.temperature.fill_(temperature), sb.top_p.fill_(top_p), sb.min_p.fill_(min_p), sb.rep_pen.fill_(rep)
    sb.top_k.fill_(top_k), sb.max_out.fill_(16)
    sb.seed.copy_(torch.arange(B, dtype=torch.int64) * 7919 + 13)
    sb.uid.copy_(torch.arange(B, dtype=torch.int64) + (1 << 33))
    sb.n_out.fill_(3)
    sb.seq_len.fill_(50)
    merged = cfg_merge(logits[:B, :8194].float(), logits[B:, :8194].float(), cfg, "bf16").numpy()
    return logits.cuda(), sb, merged, seen.numpy()


def test_sampler_greedy_and_state_updates(ops):
    B = 16
    logits, sb, merged, seen = _sampler_case(ops, B, 1, temperature=0.0, quantize=True)
    sb.finished[5] = 1                                    # finished pairs are left untouched
    ops.cfg_sample(logits, sb, B)
    tok = sb.last_tok.cpu().numpy()
    for b in range(B):
        if b == 5:
            continue
        want = sampler_ref.sample_row(merged[b], seen[b], repetition_penalty=2.0, temperature=0.0)
        assert tok[b] == want, (b, tok[b], want)
        assert sb.out_tokens[b, 3].item() == want and sb.n_out[b].item() == 4
        assert (sb.seen[b, want >> 5].item() >> (want & 31)) & 1
    assert sb.n_out[5].item() == 3 and sb.seq_len[5].item() == 50 and sb.seq_len[B + 5].item() == 50
    assert sb.seq_len[0].item() == 51 and sb.seq_len[B].item() == 51
    assert sb.n_finished.item() == 0


def test_sampler_stop_and_max_tokens(ops):
    B = 4
    logits, sb, merged, seen = _sampler_case(ops, B, 2, temperature=0.0, rep=1.0)
    logits[0, 6562] = 100.0                              # pair 0 emits the stop token
    logits[1, 6562] = 100.0
    sb.ignore_eos[1] = 1                                 # pair 1 ignores it
    sb.max_out[2] = 4                                    # pair 2 hits max_tokens (n_out 3 -> 4)
    ops.cfg_sample(logits, sb, B)
    assert sb.last_tok[:2].tolist() == [6562, 6562]
    assert sb.finished.tolist() == [1, 0, 1, 0] and sb.n_finished.item() == 2


@pytest.mark.parametrize("kw", [dict(temperature=0.5, top_p=0.5), dict(temperature=0.8, min_p=0.05),
                                dict(temperature=0.5, top_p=0.8, min_p=0.1), dict(temperature=1.0, top_p=0.9, top_k=50),
                                dict(temperature=0.3, top_k=20), dict(temperature=0.5, top_p=0.5, quantize=True),
                                dict(temperature=1.0, top_p=0.95, scale=1.0)])
def test_sampler_random_matches_oracle_stream(ops, kw):
    """Same Philox counters + exponential race => the same token as the oracle.  fp32 exp/log
    differ by an ulp between CUDA and numpy, so a tiny fraction of near-ties may flip: >= 97 %
    of rows must agree exactly and every token must lie in the oracle's kept set."""
    B = 128
    quant = kw.pop("quantize", False)
    scale = kw.pop("scale", 3.0)
    logits, sb, merged, seen = _sampler_case(ops, B, 3, quantize=quant, scale=scale, **kw)
    ops.cfg_sample(logits, sb, B)
    tok = sb.last_tok.cpu().numpy()
    agree = 0
    for b in range(B):
        okw = dict(repetition_penalty=2.0, temperature=kw["temperature"], top_p=kw.get("top_p", 1.0),
                   min_p=kw.get("min_p", 0.0), top_k=kw.get("top_k", 0))
        want = sampler_ref.sample_row(merged[b], seen[b], seed=b * 7919 + 13, uid=b + (1 << 33), step=3, **okw)
        agree += int(tok[b] == want)
        l = sampler_ref.apply_repetition_penalty(merged[b], seen[b], 2.0)
        keep, _, margin = sampler_ref.kept_set(l, okw["temperature"], okw["min_p"], okw["top_p"], okw["top_k"],
                                               return_margin=True)
        assert keep[tok[b]] or margin[tok[b]] < 1e-4, f"row {b}: token {tok[b]} outside the kept set"
    assert agree >= 0.97 * B, f"only {agree}/{B} rows match the oracle stream"


@pytest.mark.parametrize("kw", [dict(temperature=0.5, top_p=0.5, scale=0.65), dict(temperature=0.5, top_p=0.5, scale=10.0),
                                dict(temperature=0.3, top_p=0.1), dict(temperature=1.0, top_p=0.9, quantize=True),
                                dict(temperature=1.0, top_p=0.999, scale=1.0), dict(temperature=0.7, top_p=0.8, min_p=0.02),
                                dict(temperature=0.5, top_p=0.5, top_k=40, quantize=True)])
def test_sampler_bucket_pivot_equals_radix_pivot(ops, kw):
    """The top-p pivot found by the bucket histogram + exact ranking is the key the MSB-first radix passes find: the
    same launch with the radix form (cbx_debug_sampler_radix_pivot) must emit the same token in every row.  Rows 0-5
    are degenerate on purpose: constant logits and two-valued logits put thousands of ids into one bucket (the radix
    fallback inside bucket_select), a single dominant id leaves one member, -inf-like tails leave zero weights."""
    from chatterbox_vllm_b200._lib import load
    B = 128
    quant = kw.pop("quantize", False)
    scale = kw.pop("scale", 3.0)
    logits, sb, _, _ = _sampler_case(ops, B, 11, quantize=quant, scale=scale, **kw)
    g = torch.Generator().manual_seed(5)
    two = torch.where(torch.rand(8194, generator=g) < 0.3, 2.0, 1.5).to(BF16).cuda()
    for half in (0, B):
        logits[half + 0, :8194] = 1.25
        logits[half + 1, :8194] = two
        logits[half + 2, :8194] = -3.0
        logits[half + 2, 777] = 30.0
        logits[half + 3, :4000] = -120.0
        logits[half + 4, :8194] = two * 40
        logits[half + 5, :8194] = 0.0
    toks = []
    try:
        for radix in (0, 1):
            load().cbx_debug_sampler_radix_pivot(radix)
            sb.n_out.fill_(3), sb.finished.zero_(), sb.n_finished.zero_()
            seen0 = sb.seen.clone()                    # the bitmap gains the emitted token: restored for the second form
            ops.cfg_sample(logits, sb, B)
            toks.append(sb.last_tok.cpu().numpy().copy())
            sb.seen.copy_(seen0)
    finally:
        load().cbx_debug_sampler_radix_pivot(0)
    assert (toks[0] == toks[1]).all(), np.nonzero(toks[0] != toks[1])[0][:10]
    assert ((toks[0] >= 0) & (toks[0] < 8194)).all()


def test_sampler_distribution_chi2(ops):
    """4096 independent draws (different uid per row) of one distribution."""
    B = 4096
    g = torch.Generator().manual_seed(11)
    row = (torch.randn(8194, generator=g) * 3).to(BF16)
    logits = torch.zeros(2 * B, 8256, dtype=BF16)
    logits[:, :8194] = row
    sb = ops.SamplerBuffers(B, 4, "cuda", cfg_scale=0.5)
    sb.temperature.fill_(0.7), sb.top_p.fill_(0.9), sb.min_p.fill_(0.02), sb.max_out.fill_(4)
    sb.uid.copy_(torch.arange(B, dtype=torch.int64))
    ops.cfg_sample(logits.cuda(), sb, B)
    cnt = np.bincount(sb.last_tok.cpu().numpy(), minlength=8194).astype(float)
    p = sampler_ref.final_probs(row.float().numpy(), np.zeros(8194, bool), temperature=0.7, top_p=0.9, min_p=0.02)
    assert cnt[p == 0].sum() <= 2
    big = np.nonzero(p * B >= 5)[0]
    chi2 = (((cnt[big] - p[big] * B) ** 2) / (p[big] * B)).sum()
    assert chi2 < 2.0 * len(big) + 30, (chi2, len(big))


def test_prefill_embed_matches_oracle(ops):
    cfg = T3ModelConfig(n_layers=1)
    w = T3Weights.random(cfg, seed=4)
    o = T3Oracle(w, "bf16")
    g = torch.Generator().manual_seed(1)
    cond = torch.randn(2, 34, 1024, generator=g) * 0.02
    text = [255, 17, 400, 2, 33, 0]
    desc = []
    for half in range(2):
        desc += [(0, 1, i, 0) for i in range(34)]
        desc += [(1, t, i, 0) if half == 0 else (2, 0, 0, 0) for i, t in enumerate(text)]
        desc += [(3, 6561, 0, 0)]
    d = lambda t: t.cuda()  # noqa: E731
    got = ops.prefill_embed(d(torch.tensor(desc, dtype=torch.int32)), d(cond), d(w.text_emb), d(w.text_pos),
                            d(w.speech_emb), d(w.speech_pos))
    c, u = o.prefill_embeds(cond[1], torch.tensor(text))
    assert torch.equal(got.cpu().float(), torch.cat((c, u), 0)), "prefill embeddings must be bit-exact"


# ------------------------------------------------------------------ whole step
def _mk_engine(n_layers, max_pairs, **kw):
    from chatterbox_vllm_b200.engine import T3Engine
    cfg = T3ModelConfig(n_layers=n_layers)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    return cfg, w, T3Engine(w.state_dict(), device="cuda:0", max_pairs=max_pairs, max_model_len=256, **kw)


def _requests(n, n_steps, seed=2, **sp):
    from chatterbox_vllm_b200.engine import Request, SamplingParams
    g = torch.Generator().manual_seed(seed)
    reqs = []
    for i in range(n):
        n_text = [12, 40, 7, 25, 18, 33][i % 6]
        text = [255] + torch.randint(3, 600, (n_text - 2,), generator=g).tolist() + [0]
        cond = torch.randn(34, 1024, generator=g) * 0.02
        reqs.append(Request(text, cond, SamplingParams(max_tokens=n_steps, ignore_eos=True, **sp), uid=100 + i))
    return reqs


def _oracle_stream(w, reqs, cfg_scale, n_steps, teacher=None, prompt_seen=None, **sp):
    """Free-running (or teacher-forced) oracle decode; returns tokens and merged logits per step.
    prompt_seen[r]: speech ids already in request r's repetition-penalty mask (the multilingual prompt quirk)."""
    o = T3Oracle(w, "bf16")
    toks, logs = [], []
    for ri, r in enumerate(reqs):
        pair = o.new_pair()
        seen = np.zeros(8194, bool)
        if prompt_seen is not None:
            seen[list(prompt_seen[ri])] = True
        lg = o.prefill_pair(pair, r.cond_emb, r.text_ids, cfg_scale).numpy()
        tr, lr = [], []
        for s in range(n_steps):
            lr.append(lg)
            t = sampler_ref.sample_row(lg, seen, seed=r.params.seed or 0, uid=r.uid, step=s, **sp)
            tr.append(t)
            fed = t if teacher is None else teacher[ri][s] - 1000
            seen[fed] = True
            if s + 1 < n_steps:
                lg = o.decode_pairs([pair], [fed], cfg_scale)[0].numpy()
        toks.append([t + 1000 for t in tr])
        logs.append(lr)
    return toks, logs

def assert_stream_parity(got, w, reqs, cfg_scale, n_steps, *, first_exact=8, min_agree=0.9, ulps=4.0, prompt_seen=None, **sp):
    """Greedy parity under bf16 accumulation-order noise: the oracle is teacher-forced with the
    engine's tokens; at every step the engine's token must be the oracle's argmax or lie within
    `ulps` bf16 ulps (2^-7 relative) of the oracle's top penalised logit (within ONE ulp during the
    first `first_exact` steps, where little rounding noise has accumulated), and >= min_agree of all
    steps must agree exactly."""
    want, logs = _oracle_stream(w, reqs, cfg_scale, n_steps, teacher=got, prompt_seen=prompt_seen, **sp)
    agree = total = 0
    for r in range(len(reqs)):
        seen = np.zeros(8194, bool)
        if prompt_seen is not None:
            seen[list(prompt_seen[r])] = True
        for s in range(n_steps):
            total += 1
            if got[r][s] == want[r][s]:
                agree += 1
            else:
                l = sampler_ref.apply_repetition_penalty(logs[r][s], seen, sp.get("repetition_penalty", 1.0))
                gap = l.max() - l[got[r][s] - 1000]
                assert s >= first_exact or gap <= 2.0 ** -7 * abs(l.max()), \
                    f"req {r}: streams differ at step {s} < {first_exact} by {gap:.4f} (top {l.max():.3f}): {got[r][:s+1]} vs {want[r][:s+1]}"
                assert gap <= ulps * 2.0 ** -7 * abs(l.max()), f"req {r} step {s}: engine token is {gap:.3f} below the oracle top logit {l.max():.3f}"
            seen[got[r][s] - 1000] = True
    assert agree >= min_agree * total, f"only {agree}/{total} steps agree"
    return agree, total


@pytest.mark.parametrize("use_graph", [False, True])
def test_engine_greedy_stream_matches_oracle_2layers(ops, use_graph):
    n_steps = 24
    cfg, w, eng = _mk_engine(2, 4, use_graph=use_graph)
    reqs = _requests(3, n_steps, temperature=0.0, repetition_penalty=2.0)
    got = eng.generate(reqs, poll_every=5)
    assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
    # engine is reusable, deterministic (same batch => same kernels => same bits), and the pool
    # returned every block
    assert eng.pool.num_free() == eng.num_blocks - 1
    assert eng.generate(reqs) == got
    # a single pair takes the GEMV path (T = 2 rows): a different accumulation order, so it is held
    # to the oracle, not to the batched run
    one = eng.generate(reqs[:1])
    assert_stream_parity(one, w, reqs[:1], eng.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
    eng.close()


@pytest.mark.parametrize("split_rows", ["512", "0"])
def test_engine_batch32_split_k_parity(ops, split_rows, monkeypatch):
    """T = 64 rows (tcgen05 regime): o_proj / down_proj run as split-K partials folded by the next
    add + RMSNorm kernel (default) or unsplit with the residual epilogue (CBX_SPLIT_MAX_ROWS=0).
    Same parity bar either way, with and without CUDA-graph replay."""
    monkeypatch.setenv("CBX_SPLIT_MAX_ROWS", split_rows)
    n_steps = 10
    for use_graph in (True, False):
        cfg, w, eng = _mk_engine(2, 32, use_graph=use_graph)
        reqs = _requests(32, n_steps, temperature=0.0, repetition_penalty=2.0)
        got = eng.generate(reqs, poll_every=4)
        assert eng.generate(reqs, poll_every=3) == got, "same batch must give the same bits"
        assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, first_exact=0, temperature=0.0,
                             repetition_penalty=2.0)
        eng.close()


@pytest.mark.parametrize("n_pairs", [1, 2])
def test_engine_persistent_decode_kernel_parity(ops, n_pairs, monkeypatch):
    """CBX_MEGA=1: the T <= 4 decode step as ONE persistent cooperative kernel (k_mega.cu: weight ring,
    tagged-word exchanges, mma.sync GEMV).  Same parity bar as the kernel chain, graph and eager,
    deterministic, and the KV pages it writes serve a later run on the default path."""
    if not ops._lib.load().cbx_has_experiments():
        pytest.skip("default build: the experiment kernels are behind `make EXPERIMENTS=1`")
    monkeypatch.setenv("CBX_MEGA", "1")
    n_steps = 40
    for use_graph in (True, False):
        cfg, w, eng = _mk_engine(3, 2, use_graph=use_graph)
        reqs = _requests(n_pairs, n_steps, temperature=0.0, repetition_penalty=2.0)
        got = eng.generate(reqs, poll_every=7)
        assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
        assert eng.generate(reqs, poll_every=4) == got, "same batch must give the same bits"
        eng.close()


def test_engine_device_handoff_equals_host_postprocessing(ops):
    """generate_device(): tokens never leave the GPU; the result must equal generate() followed by the
    reference's host-side list surgery (tts.py:448-486), with slot recycling in between."""
    from oracle import prompt_ref
    cfg, w, eng = _mk_engine(2, 3)
    # a stop id the random model emits now and then, so that streams end at different lengths
    reqs = _requests(7, 40, temperature=0.9, top_p=0.95, repetition_penalty=1.2)
    for i, r in enumerate(reqs):
        r.params.seed, r.params.ignore_eos = 11 + i, False
    host = eng.generate(reqs, poll_every=6)
    caps = [9, 1000, 1000, 20, 1000, 1000, 3]
    toks, lens = eng.generate_device(reqs, tail_crop=2, guard_caps=caps, poll_every=6)
    assert toks.is_cuda and toks.dtype == torch.int64 and lens.dtype == torch.int32
    for i in range(len(reqs)):
        want = prompt_ref.postprocess_tokens(host[i], 2)[:caps[i]]
        assert int(lens[i]) == len(want) and toks[i, :len(want)].cpu().tolist() == want.tolist(), f"request {i}"
    assert eng.pool.num_free() == eng.num_blocks - 1
    eng.close()


def test_tts_surface_device_conditionals_and_token_handoff(ops):
    """ChatterboxTTS over the real engine: conditionals encoded on the device from (speaker embedding,
    prompt tokens) by a `conditioner` standing in for the audio front-ends, exaggeration updated on
    the device, tokens handed to the vocoder as CUDA int64 without a host round trip."""
    from chatterbox_vllm_b200.tts import ChatterboxTTS
    from oracle import cond_ref, prompt_ref

    class Tok:
        def encode(self, s):
            return [255] + [3 + (ord(c) % 600) for c in s[7:-6]][:30] + [0]

    cfg = T3ModelConfig(n_layers=2)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    sd = dict(w.state_dict())
    csd = cond_ref.random_state_dict(5)
    sd.update({"cond_enc." + k: v for k, v in csd.items()})
    speech_emb, speech_pos = sd["speech_emb.weight"].float(), sd["speech_pos_emb.emb.weight"].float()
    g = torch.Generator().manual_seed(3)
    spk, ptoks = torch.randn(256, generator=g), torch.randint(0, 6561, (150,), generator=g)
    seen = {}

    def vocoder(tokens, ref, n_timesteps=10):
        seen.setdefault("tok", []).append(tokens)
        return torch.zeros(1, 8)

    tts = ChatterboxTTS.from_state_dict(sd, Tok(), "cuda", 256, 4, conditioner=lambda path: ({"ref": path}, spk, ptoks),
                                        vocoder=vocoder)
    ref, cond = tts.get_audio_conditionals("voice.wav")
    assert ref == {"ref": "voice.wav"} and cond.is_cuda and cond.shape == (34, 1024)
    want = cond_ref.cond_encode(csd, spk, ptoks, speech_emb, speech_pos, 0.5)
    assert float((cond.cpu() - want).abs().max()) <= 2e-5 * float(want.abs().max())
    c8 = tts.update_exaggeration(cond, 0.8)
    assert c8.is_cuda and torch.equal(c8[:33], cond[:33])
    assert torch.equal(c8[33].cpu(), cond_ref.emotion_row(csd, 0.8))
    wavs = tts.generate(["Hello there.", "A second, longer prompt."], audio_prompt_path="voice.wav", exaggeration=0.8,
                        max_tokens=24, temperature=0.0)
    assert len(wavs) == 2 and len(seen["tok"]) == 2
    raw = tts.generate_tokens(["Hello there.", "A second, longer prompt."], c8, max_tokens=24, temperature=0.0, raw=True)
    for t, r in zip(seen["tok"], raw):
        assert t.is_cuda and t.dtype == torch.int64
        assert t.cpu().tolist() == prompt_ref.postprocess_tokens(r, 2).tolist()
    tts.shutdown()


def test_engine_logits_tolerance_30layers(ops):
    """Full-depth (30-layer) model, oracle teacher-forced with the engine's own tokens: every
    engine token is the oracle's argmax or within 8 bf16 ulps of its top logit, and >= 80 % of
    steps agree exactly (30 layers of bf16 rounding amplify accumulation-order noise, so a
    near-tie may flip at any step, including the first)."""
    n_steps = 10
    cfg, w, eng = _mk_engine(30, 2, use_graph=True)
    reqs = _requests(2, n_steps, temperature=0.0, repetition_penalty=1.2)
    got = eng.generate(reqs)
    assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, first_exact=0, min_agree=0.8, ulps=8.0,
                         temperature=0.0, repetition_penalty=1.2)
    eng.close()


# Stated bf16 tolerance of the path (BASELINE.json north_star: "max-abs and relative error per
# layer and at the head").  Per layer, on the HF hidden state (mlp_out + residual):
#   rel-L2  = |eng - oracle|_2 / |oracle|_2      <= 3e-2 vs the bf16-emulating oracle
#                                                <= 3.5e-2 vs the fp32 oracle
# (the oracle against ITSELF with GEMMs accumulated in fp64 instead of fp32 already differs by
#  2.2e-3 after layer 1 and 7.4e-3 after layer 4 — bf16 rounding flips are amplified layer by layer;
#  the engine measures 4e-3 after layer 1 and 2.0e-2 after layer 30, i.e. it sits on that floor)
#   max-abs = max|eng - oracle| / max|oracle|    <= 6e-2
# At the head (CFG-merged inputs, bf16 logits): max-abs <= 3e-2 of the logit range.
LAYER_REL_L2_BF16, LAYER_REL_L2_FP32, LAYER_MAX_ABS, HEAD_MAX_ABS = 3e-2, 3.5e-2, 6e-2, 3e-2


def _layer_errors(tr_eng, tr_orc):
    """tr_eng bf16 [L, T, 1024] (cuda): residual stream after each layer = HF hidden state;
    tr_orc list of (mlp_out, residual) fp32 [T, 1024]."""
    out = []
    for li, (dn, res) in enumerate(tr_orc):
        want = (dn + res).to(BF16).float() if dn.dtype == torch.float32 else dn + res   # same storage rounding as the engine
        got = tr_eng[li].float().cpu()
        d = got - want
        out.append((float(d.norm() / want.norm()), float(d.abs().max() / want.abs().max())))
    return out


def _dump(name, obj):
    import json, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d = os.path.join(root, "gpurun_out")
    if os.path.isdir(d):
        json.dump(obj, open(os.path.join(d, name), "w"))


def test_engine_per_layer_error_30layers(ops):
    """Per-layer and head error of the 30-layer model for a prefill pass (tcgen05 GEMM path), a
    3-pair decode step (tcgen05 path, T = 6) and a 1-pair decode step (GEMV path, T = 2)."""
    cfg, w, eng = _mk_engine(30, 4, use_graph=False)
    report = {}
    for n_req in (3, 1):
        reqs = _requests(n_req, 4, temperature=0.0)
        P, held = eng._admit(reqs)
        eng._init_sampler_state(reqs, P)
        T = 2 * sum(P)
        tr = torch.zeros(30, T, 1024, dtype=BF16, device="cuda")
        eng.set_trace(tr)
        with torch.cuda.stream(eng.stream):
            eng.prefill(reqs, P)
        torch.cuda.synchronize()
        eng.set_trace(None)
        lg = eng.buffer("logits", (2 * n_req, 8256), BF16)[:, :8194].float().cpu()
        first = eng.sb.last_tok[:n_req].cpu().tolist()
        orcs = {m: T3Oracle(w, m) for m in ("bf16", "fp32")}
        pairs = {m: [] for m in orcs}
        r0 = 0
        for b, r in enumerate(reqs):
            for m, o in orcs.items():
                o.trace = {}
                pair = o.new_pair()
                pairs[m].append(pair)
                c, u = o.prefill_embeds(r.cond_emb, torch.tensor(r.text_ids))
                h = o.forward_rows(torch.cat((c, u), 0), [(pair[0], P[b]), (pair[1], P[b])])
                errs = _layer_errors(tr[:, r0:r0 + 2 * P[b]], o.trace["layers"])
                lim = LAYER_REL_L2_BF16 if m == "bf16" else LAYER_REL_L2_FP32
                assert max(e[0] for e in errs) <= lim, f"prefill {m} rel-L2 per layer: {errs}"
                assert max(e[1] for e in errs) <= LAYER_MAX_ABS, f"prefill {m} max-abs per layer: {errs}"
                want = o.head(torch.stack((h[P[b] - 1], h[2 * P[b] - 1])))
                herr = max(float((lg[half * n_req + b] - want[half]).abs().max() / want[half].abs().max())
                           for half in range(2))
                assert herr <= HEAD_MAX_ABS, f"prefill head {m}: {herr}"
                report[f"prefill_B{n_req}_req{b}_{m}"] = {"layers": errs, "head_max_abs": herr}
            r0 += 2 * P[b]
        # one decode step, teacher-forced with the engine's first tokens
        tr = torch.zeros(30, 2 * n_req, 1024, dtype=BF16, device="cuda")
        eng.set_trace(tr)
        with torch.cuda.stream(eng.stream):
            eng.decode(n_req, 1, max(P) + 2)
        torch.cuda.synchronize()
        eng.set_trace(None)
        lg = eng.buffer("logits", (2 * n_req, 8256), BF16)[:, :8194].float().cpu()
        for m, o in orcs.items():
            o.trace = {}
            e = o.decode_embed(torch.tensor(first))
            rows = [(p[0], 1) for p in pairs[m]] + [(p[1], 1) for p in pairs[m]]
            h = o.forward_rows(torch.cat((e, e), 0), rows)
            errs = _layer_errors(tr, o.trace["layers"])
            lim = LAYER_REL_L2_BF16 if m == "bf16" else LAYER_REL_L2_FP32
            assert max(e_[0] for e_ in errs) <= lim, f"decode {m} rel-L2 per layer: {errs}"
            assert max(e_[1] for e_ in errs) <= LAYER_MAX_ABS, f"decode {m} max-abs per layer: {errs}"
            want = o.head(h)
            herr = float(((lg - want).abs().amax(1) / want.abs().amax(1)).max())
            assert herr <= HEAD_MAX_ABS, f"decode head {m}: {herr}"
            report[f"decode_B{n_req}_{m}"] = {"layers": errs, "head_max_abs": herr}
        eng._release(held)
    _dump("layer_parity.json", report)
    eng.close()


def test_engine_prefill_logits_vs_oracle(ops):
    cfg, w, eng = _mk_engine(4, 2, use_graph=False)
    reqs = _requests(2, 2, temperature=0.0)
    P, held = eng._admit(reqs)
    eng._init_sampler_state(reqs, P)
    eng.prefill(reqs, P)
    lg = eng.buffer("logits", (4, 8256), BF16)[:, :8194].float().cpu()
    o = T3Oracle(w, "bf16")
    o32 = T3Oracle(w, "fp32")
    for b, r in enumerate(reqs):
        for orc, tol in ((o, 0.02), (o32, 0.04)):
            c, u = orc.prefill_embeds(r.cond_emb, torch.tensor(r.text_ids))
            pair = orc.new_pair()
            h = orc.forward_rows(torch.cat((c, u), 0), [(pair[0], P[b]), (pair[1], P[b])])
            want = orc.head(torch.stack((h[P[b] - 1], h[2 * P[b] - 1])))
            for half in range(2):
                err = (lg[half * 2 + b] - want[half]).abs().max().item()
                rng = want[half].abs().max().item()
                assert err <= tol * rng, f"pair {b} half {half} mode {orc.mode}: max-abs {err:.4f} vs range {rng:.2f}"
    eng._release(held)
    eng.close()


def test_engine_sampled_stream_and_waves(ops):
    """Sampling path end to end + more requests than pair slots (two static waves)."""
    n_steps = 12
    cfg, w, eng = _mk_engine(2, 2, use_graph=True)
    sp = dict(temperature=0.5, top_p=0.5, repetition_penalty=2.0)
    reqs = _requests(5, n_steps, seed=9, **sp)
    got = eng.generate(reqs)
    want, _ = _oracle_stream(w, reqs, eng.cfg_scale, n_steps, teacher=got, **sp)
    agree = sum(int(a == b) for g_, w_ in zip(got, want) for a, b in zip(g_, w_))
    assert agree >= 0.9 * 5 * n_steps, f"{agree}/{5 * n_steps}"
    assert all(len(t) == n_steps for t in got)
    eng.close()


def test_engine_continuous_admission_recycles_slots(ops):
    """7 requests with different lengths through 3 pair slots: a finished pair's slot (and KV blocks)
    is refilled while the other pairs keep decoding; every stream still matches the oracle and the
    pool gets every block back."""
    from chatterbox_vllm_b200.engine import SamplingParams
    cfg, w, eng = _mk_engine(2, 3, use_graph=True)
    lens = [5, 14, 9, 3, 11, 6, 8]
    reqs = _requests(7, 1, temperature=0.0, repetition_penalty=2.0)
    for r, n in zip(reqs, lens):
        r.params = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n, ignore_eos=True)
    got = eng.generate(reqs, poll_every=2)
    assert [len(g) for g in got] == lens
    assert eng.pool.num_free() == eng.num_blocks - 1
    for i, r in enumerate(reqs):
        assert_stream_parity([got[i]], w, [r], eng.cfg_scale, lens[i], first_exact=0, temperature=0.0,
                             repetition_penalty=2.0)
    assert eng.generate(reqs, poll_every=2) == got, "admission order is deterministic => same bits"
    # a pool too small to hold all three slots at once still completes (admission waits for blocks)
    eng.close()
    from chatterbox_vllm_b200.engine import T3Engine
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=3, max_model_len=256, num_blocks=2 * 2 * 4 + 1)
    got2 = eng.generate(reqs, poll_every=2)
    assert [len(g) for g in got2] == lens
    eng.close()


def test_engine_stops_on_eos_and_per_request_max_tokens(ops):
    from chatterbox_vllm_b200.engine import SamplingParams, T3Engine
    cfg, w, eng = _mk_engine(2, 2, use_graph=True)
    reqs = _requests(2, 30, temperature=0.0, repetition_penalty=2.0)
    free = eng.generate(reqs, poll_every=4)
    eng.close()
    stop = free[0][6]                                      # pretend this id is the stop token
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=2, max_model_len=256, stop_token=stop - 1000)
    for r, mt in zip(reqs, (30, 9)):
        r.params = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=mt, ignore_eos=False)
    got = eng.generate(reqs, poll_every=4)
    cut0 = free[0][:free[0].index(stop) + 1]
    cut1 = free[1][:9]
    if stop in cut1:
        cut1 = cut1[:cut1.index(stop) + 1]
    assert got[0] == cut0, "generation must stop right after the first stop token (it is included)"
    assert got[1] == cut1, "per-request max_tokens must be honoured"
    eng.close()


def test_engine_multilingual_shapes_and_prompt_penalty_quirk(ops):
    """Config 4 shapes: text vocab 2454 (t3_config.py:33-35), text ids up to 2453, language-tag id
    first.  With penalize_prompt_ids the reference's quirk is reproduced: prompt ids >= 1000 land in
    the repetition-penalty mask of speech id (id - 1000)."""
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    cfg = T3ModelConfig.multilingual(n_layers=2)
    w = T3Weights.random(cfg, seed=3, head_scale=16.0, norm_jitter=0.1)
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=2, max_model_len=256)
    g = torch.Generator().manual_seed(5)
    n_steps = 8
    reqs = []
    for i in range(2):
        text = [708] + [255] + torch.randint(698, 2454, (20,), generator=g).tolist() + [0]
        cond = torch.randn(34, 1024, generator=g) * 0.02
        reqs.append(Request(text, cond, SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n_steps,
                                                       ignore_eos=True), uid=50 + i, penalize_prompt_ids=True))
    got = eng.generate(reqs)
    # oracle with the same prompt-derived seen mask
    o = T3Oracle(w, "bf16")
    for ri, r in enumerate(reqs):
        pair = o.new_pair()
        seen = np.zeros(8194, bool)
        for t in r.text_ids:
            if 1000 <= t < 1000 + 8194:
                seen[t - 1000] = True
        assert seen.any()
        lg = o.prefill_pair(pair, r.cond_emb, r.text_ids, eng.cfg_scale).numpy()
        agree = 0
        for s_ in range(n_steps):
            want = sampler_ref.sample_row(lg, seen, repetition_penalty=2.0, temperature=0.0)
            tok = got[ri][s_] - 1000
            l = sampler_ref.apply_repetition_penalty(lg, seen, 2.0)
            assert tok == want or l.max() - l[tok] <= 4 * 2.0 ** -7 * abs(l.max()), (ri, s_, tok, want)
            agree += int(tok == want)
            seen[tok] = True
            lg = o.decode_pairs([pair], [tok], eng.cfg_scale)[0].numpy()
        assert agree >= n_steps - 2
    eng.close()


def test_engine_long_context_high_block_occupancy(ops):
    """Config 5 in miniature: max_model_len 1000, KV pool ~97 % occupied, contexts running up to
    ~1000 tokens, T = 64 rows.  Streams of a few probed pairs must match the oracle; every block
    returns to the pool."""
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    cfg = T3ModelConfig(n_layers=2)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    B, n_text, n_steps = 32, 900, 40
    P = 34 + n_text + 1
    per_seq = -(-(P + n_steps) // 16)
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=B, max_model_len=1000, num_blocks=int(2 * B * per_seq / 0.97) + 1)
    g = torch.Generator().manual_seed(9)
    reqs = []
    for i in range(B):
        text = [255] + torch.randint(3, 600, (n_text - 2,), generator=g).tolist() + [0]
        reqs.append(Request(text, torch.randn(34, 1024, generator=g) * 0.02,
                            SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n_steps, ignore_eos=True),
                            uid=900 + i))
    got = eng.generate(reqs, poll_every=16)
    assert all(len(t) == n_steps for t in got)
    assert eng.pool.num_free() == eng.num_blocks - 1
    probe = [0, B - 1]
    assert_stream_parity([got[i] for i in probe], w, [reqs[i] for i in probe], eng.cfg_scale, n_steps, first_exact=0,
                         min_agree=0.85, temperature=0.0, repetition_penalty=2.0)
    eng.close()


def test_engine_prefill_and_decode_are_deterministic_at_scale(ops):
    """Same job twice => same bits, at a size where rows are cut across CTAs, prefill runs in several
    chunks of tcgen05 pair GEMMs and the attention rings wrap many times (16 pairs x prefix 200, 6 layers,
    48 steps, greedy + sampled requests mixed)."""
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict
    eng = T3Engine(random_state_dict(6, device="cuda", head_scale=16.0), device="cuda:0", max_pairs=16, max_model_len=320)
    reqs = []
    for i in range(16):
        sp = SamplingParams(temperature=0.0 if i % 4 == 0 else 0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=48, ignore_eos=True)
        reqs.append(Request(random_prompt(165, seed=i), random_cond_emb(i), sp, uid=1 + i))
    first = eng.generate(reqs)
    for _ in range(3):
        assert eng.generate(reqs) == first
    eng.close()


@pytest.mark.parametrize("n_pairs", [32, 96])
def test_split_k_fold_inside_the_gemm_equals_the_separate_reduce_kernel(ops, n_pairs, monkeypatch):
    """T = 64 / 192 rows (1 / 2 row tiles), experiment builds: with CBX_GEMM_FOLD=1 the split-K o_proj / down_proj GEMMs
    fold their own partials (add + RMSNorm by the producing CTAs behind an arrival counter, k_gemm.cu EPI 5) instead of
    leaving them to rmsnorm_reduce4_kernel.  Same adds in the same order: tokens AND the last step's logits must be
    bit-identical, graph replay and eager, and the arrival counters must come back to zero (a second run gives the
    same bits).  (Measured slower than the default two-kernel form, so it is not in the default library.)"""
    if not ops._lib.load().cbx_has_experiments():
        pytest.skip("default build: the experiment kernels are behind `make EXPERIMENTS=1`")
    n_steps = 8
    out = {}
    for fold in ("1", "0"):
        monkeypatch.setenv("CBX_GEMM_FOLD", fold)
        for use_graph in (True, False):
            cfg, w, eng = _mk_engine(3, n_pairs, use_graph=use_graph)
            reqs = _requests(n_pairs, n_steps, temperature=0.0, repetition_penalty=2.0)
            l0 = eng.gpu_launches
            toks = eng.generate(reqs, poll_every=7)
            assert eng.generate(reqs, poll_every=7) == toks
            lg = eng.buffer("logits", (2 * n_pairs, 8256), BF16)[:, :8194].clone()
            out[(fold, use_graph)] = (toks, lg, eng.gpu_launches - l0)
            eng.close()
    ref_t, ref_l, ref_n = out[("0", False)]
    for key, (t, l, n) in out.items():
        assert t == ref_t, f"fold={key[0]} graph={key[1]}: tokens differ from the two-kernel form"
        assert torch.equal(l, ref_l), f"fold={key[0]} graph={key[1]}: logits differ from the two-kernel form"
    assert out[("1", False)][2] < ref_n, "the folding form did not launch fewer kernels"
