"""The engine's request scheduler on the GPU: KV blocks that follow the sequences (bit-exact against the
oracle's restatement of vLLM's block pool / block table under a scripted admit / grow / finish sequence at 25 / 50 /
95 % pool occupancy), preemption + recomputation when the pool runs dry, streaming emission, and
CHATTERBOX_CFG_SCALE through the environment (t3.py:343) at values other than the default 0.5."""
import numpy as np
import pytest
import torch

from oracle import paging_ref
from oracle.t3_ref import T3ModelConfig, T3Weights

try:
    from test_gpu_engine import _requests, assert_stream_parity
except ImportError:                                      # importmode=importlib
    from tests.test_gpu_engine import _requests, assert_stream_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def weights():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    cfg = T3ModelConfig(n_layers=2)
    return T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)


def _engine(w, max_pairs, **kw):
    from chatterbox_vllm_b200.engine import T3Engine
    kw.setdefault("max_model_len", 256)
    return T3Engine(w.state_dict(), device="cuda:0", max_pairs=max_pairs, **kw)


@pytest.mark.parametrize("occupancy", [0.25, 0.5, 0.95])
def test_engine_block_table_follows_vllm_policy_bit_exact(weights, occupancy):
    """The engine's device block table after a scripted admit / decode-window / finish / re-admit sequence equals
    oracle/paging_ref.BlockTable (vLLM block_pool.py:168-177,333-347 + single_type_kv_cache_manager.py:120) driven
    the same way: cdiv(tokens, 16) blocks per row, popped from the free queue in request order (cond row, uncond
    row), freed in reverse on finish."""
    from chatterbox_vllm_b200.engine import Request, SamplingParams
    B, look = 6, 32
    g = torch.Generator().manual_seed(int(occupancy * 100))
    n_text = [int(x) for x in torch.randint(5, 60, (10,), generator=g)]
    gen = [int(x) for x in torch.randint(40, 150, (10,), generator=g)]
    P = [34 + n + 1 for n in n_text]
    peak = 2 * sum(-(-(P[i] + gen[i]) // 16) for i in range(B))          # blocks the first wave needs at its longest
    num_blocks = int(peak / occupancy) + 2
    eng = _engine(weights, B, num_blocks=num_blocks)
    ref_pool = paging_ref.BlockPool(num_blocks)
    ref = paging_ref.BlockTable(ref_pool, 2 * B, 256)
    reqs = [Request([255] + [7] * (n_text[i] - 2) + [0], torch.zeros(34, 1024), SamplingParams(max_tokens=gen[i]), uid=i + 1)
            for i in range(10)]
    eng._reset_slots(B)
    held, length = {}, {}

    def admit(i, b):
        p, h = eng._admit([reqs[i]], [b], lookahead=look)
        held[b], length[b] = h, p[0] + 1
        for row in (b, B + b):                                         # the engine's order: a pair's cond row, then its uncond row
            ref.ensure(row, min(P[i] + gen[i], P[i] + 1 + look))

    def window(steps, slots):
        want = []
        for b in slots:
            want += [(b, length[b] + steps), (B + b, length[b] + steps)]
        assert eng._grow(want)
        for r, n in want:
            ref.ensure(r, n)
        for b in slots:
            length[b] += steps

    def finish(b):
        eng._release(held.pop(b))                                      # cond row's blocks (reversed), then the uncond row's
        ref.release(b)
        ref.release(B + b)

    def check(what):
        got = eng.block_table.cpu().numpy()
        live = sorted(list(held) + [B + b for b in held])
        np.testing.assert_array_equal(got[live], ref.table[live], err_msg=what)
        assert eng.pool.num_free() == ref_pool.num_free(), what

    for b in range(B):
        admit(b, b)
    check("first wave admitted")
    window(32, range(B))
    check("one window")
    window(17, [0, 2, 3, 5])
    finish(1)
    finish(4)
    check("two finished")
    admit(6, 1)
    admit(7, 4)
    window(32, [0, 1, 2, 3, 4, 5])
    check("slots recycled")
    for b in (0, 2):
        finish(b)
    admit(8, 0)
    window(9, [0, 1, 3, 4, 5])
    check("third wave")
    used = num_blocks - 1 - eng.pool.num_free()
    assert used > 0
    for b in list(held):
        finish(b)
    assert eng.pool.num_free() == num_blocks - 1 == ref_pool.num_free()
    eng.close()


def test_engine_preemption_recomputes_and_carries_on(weights):
    """A pool too small for every admitted pair's full length: the newest pair is preempted (blocks back to the pool,
    request to the head of the queue); once re-admitted its KV is recomputed over prefix + the tokens it had emitted
    and it carries on (vLLM v1: preemption = recompute).  The recomputed KV comes from the prefill kernels instead of
    the decode kernels (another summation order), so the streams are held to the oracle, not to the big-pool run's
    bits; every block returns, and the same schedule twice gives the same bits."""
    n_steps = 70
    reqs = _requests(6, n_steps, temperature=0.0, repetition_penalty=2.0)
    # 4 slots x 2 rows x up to ceil((75 + 70) / 16) = 10 blocks = 80 at full length; 52 admit 4 pairs, then run dry
    small = _engine(weights, 4, num_blocks=53)
    got = small.generate(reqs, poll_every=8)
    assert small.preemptions > 0, "the pool was sized to force a preemption"
    assert [len(t) for t in got] == [n_steps] * 6
    assert small.pool.num_free() == 52 and not small._rows
    assert_stream_parity(got, weights, reqs, small.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
    n0 = small.preemptions
    assert small.generate(reqs, poll_every=8) == got and small.preemptions == 2 * n0
    big = _engine(weights, 4)
    want = big.generate(reqs, poll_every=8)
    assert big.preemptions == 0
    agree = sum(a == b for g, w in zip(got, want) for a, b in zip(g, w))
    # (a greedy stream that flips one near-tie stays on its own branch, so this is a sanity bound, not a parity bar —
    # the parity bar is the oracle check above)
    assert agree >= 0.5 * 6 * n_steps, f"only {agree} of {6 * n_steps} tokens equal the un-preempted run's"
    # sampled streams: Philox is keyed by (seed, uid, step) and the step counter survives the preemption
    sreqs = _requests(6, n_steps, temperature=0.7, top_p=0.9, repetition_penalty=1.3)
    a, b = small.generate(sreqs, poll_every=8), big.generate(sreqs, poll_every=8)
    agree = sum(x == y for g, w in zip(a, b) for x, y in zip(g, w))
    assert agree >= 0.3 * 6 * n_steps, f"sampled streams: only {agree} of {6 * n_steps} tokens equal the un-preempted run's"
    small.close()
    big.close()


def test_generate_stream_hands_over_tokens_as_they_appear(weights):
    """generate_stream yields (request, new tokens, finished) every `every` steps; the concatenation per request is
    generate()'s output, first events arrive before the request is done, and slot recycling / preemption neither
    duplicate nor drop nor re-generate a token (openai_tts_server.py:392-528 is the consumer this serves)."""
    from chatterbox_vllm_b200.engine import Request, SamplingParams
    reqs = _requests(5, 40, temperature=0.0, repetition_penalty=2.0)
    for i, r in enumerate(reqs):                                      # ragged lengths: slots retire at different polls
        reqs[i] = Request(r.text_ids, r.cond_emb, SamplingParams(temperature=0.0, repetition_penalty=2.0, ignore_eos=True,
                                                                 max_tokens=12 + 9 * i), uid=r.uid)
    long = _requests(6, 70, temperature=0.0, repetition_penalty=2.0)
    for rq, slots, pool_blocks in ((reqs, 3, None), (long, 4, 53)):   # the second pool runs dry: preemption under streaming
        eng = _engine(weights, slots, num_blocks=pool_blocks)
        want = eng.generate(rq, poll_every=8)
        assert (eng.preemptions > 0) == (pool_blocks is not None), "the small pool was sized to force a preemption"
        acc = [[] for _ in rq]
        events, done = 0, set()
        for i, toks, fin in eng.generate_stream(rq, every=8):
            assert i not in done, "no event after the final one"
            acc[i] += toks
            events += 1
            if fin:
                done.add(i)
            else:
                assert 0 < len(toks) <= 9
        assert acc == want and done == set(range(len(rq)))
        assert events > len(rq), "tokens must arrive before the requests finish"
        assert_stream_parity([w[:12] for w in want], weights, rq, eng.cfg_scale, 12, temperature=0.0, repetition_penalty=2.0)
        eng.close()


@pytest.mark.parametrize("cfg", ["1.3", "0.0"])
def test_cfg_scale_from_the_environment(weights, cfg, monkeypatch):
    """CHATTERBOX_CFG_SCALE is read once at construction (t3.py:343) and reaches the fused sampler's CFG merge:
    greedy streams at cfg 1.3 / 0.0 follow the oracle run at that scale and differ from the default's."""
    n_steps = 16
    reqs = _requests(3, n_steps, temperature=0.0, repetition_penalty=2.0)
    base = _engine(weights, 4)
    assert base.cfg_scale == 0.5
    ref_tokens = base.generate(reqs)
    base.close()
    monkeypatch.setenv("CHATTERBOX_CFG_SCALE", cfg)
    eng = _engine(weights, 4)
    assert eng.cfg_scale == float(cfg)
    monkeypatch.setenv("CHATTERBOX_CFG_SCALE", "0.9")                 # read ONCE: a later change does not reach a built engine
    got = eng.generate(reqs)
    assert got != ref_tokens, "a different guidance scale must change the stream"
    assert_stream_parity(got, weights, reqs, float(cfg), n_steps, temperature=0.0, repetition_penalty=2.0)
    one = eng.generate(reqs[:1])                                     # the GEMV path (T = 2 rows) merges with the same scale
    assert_stream_parity(one, weights, reqs[:1], float(cfg), n_steps, temperature=0.0, repetition_penalty=2.0)
    eng.close()


@pytest.mark.parametrize("n_layers,seeds", [(4, range(8)), (30, range(2))], ids=["4layers", "30layers"])
def test_free_running_greedy_stream_is_the_oracles_until_a_near_tie(n_layers, seeds):
    """north_star: "identical greedy-decoded token sequences over the first N steps".  No unconditional N > 0 exists for
    two correct bf16 implementations: tools/greedy_horizon.py (profiles/r02_greedy_horizon.json) finds seeds whose very
    first step is a near-tie of the oracle's own logits.  What does hold, and is asserted here on FREE-RUNNING streams:
    the engine emits exactly the oracle's tokens up to the first step at which the oracle's top two penalised logits lie
    within 8 bf16 ulps of each other — i.e. N = the position of the first near-tie (measured: 0 .. 48+ steps)."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
    from oracle import sampler_ref
    from oracle.t3_ref import T3Oracle
    n_steps = 24
    for seed in seeds:
        w = T3Weights.random(T3ModelConfig(n_layers=n_layers), seed=seed, head_scale=16.0, norm_jitter=0.1)
        eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=1, max_model_len=256)
        g = torch.Generator().manual_seed(1000 + seed)
        text = [255] + torch.randint(3, 600, (10 + seed % 30,), generator=g).tolist() + [0]
        cond = torch.randn(34, 1024, generator=g) * 0.02
        sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n_steps, ignore_eos=True)
        got = eng.generate([Request(text, cond, sp, uid=7)])[0]
        eng.close()
        o = T3Oracle(w, "bf16")
        pair, seen = o.new_pair(), np.zeros(8194, bool)
        lg = o.prefill_pair(pair, cond, text, 0.5).numpy()
        for s in range(n_steps):
            l = sampler_ref.apply_repetition_penalty(lg, seen, 2.0)
            t = int(np.argmax(l))
            if t + 1000 != got[s]:
                gap = (l[t] - l[got[s] - 1000]) / (2.0 ** -7 * abs(l[t]))
                assert gap <= 8.0, f"seed {seed}: streams differ at step {s} although the oracle's margin is {gap:.1f} ulps"
                break
            seen[t] = True
            if s + 1 < n_steps:
                lg = o.decode_pairs([pair], [t], 0.5)[0].numpy()
