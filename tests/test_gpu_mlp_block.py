"""cbx_layer_mlp_block — the MLP half of a layer as one kernel (k_mlp.cu, T <= 4 rows) — through the C-ABI against
(a) the two kernels it replaces (layer_gate_up -> linear_residual) and (b) the op-by-op bf16 chain in fp32 torch; then
the engine's two-launch layer (env CBX_MLP_FUSED=1: attention block + MLP block, each folding the other's fixed-point
accumulator) against the oracle, against the default three-launch layer, and for what it must leave behind: empty
accumulators."""
import pytest
import torch

from oracle.t3_ref import rmsnorm
from test_gpu_engine import _mk_engine, _requests, assert_stream_parity

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
ULP = 2.0 ** -7


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200 import ops as _ops
    _ops._lib.load()
    return _ops


def _weights(seed):
    g = torch.Generator().manual_seed(seed)
    ln = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    wg = (torch.randn(4096, 1024, generator=g) * 0.03).to(BF16)
    wu = (torch.randn(4096, 1024, generator=g) * 0.03).to(BF16)
    wd = (torch.randn(1024, 4096, generator=g) * 0.02).to(BF16)
    return g, ln, wg, wu, wd


@pytest.mark.parametrize("T", [1, 2, 3, 4])
def test_fused_mlp_block_vs_chain_and_reference(ops, T):
    g, ln, wg, wu, wd = _weights(10 + T)
    res = (torch.randn(T, 1024, generator=g) * 1.5).to(BF16)
    d = lambda t: t.cuda()  # noqa: E731
    wgu = ops.pack_gate_up(d(torch.cat([wg, wu], 0).contiguous()))
    wd_d = d(wd)
    wd_t = wd_d.t().contiguous()
    act = ops.layer_gate_up(d(res), d(ln), wgu)
    chain = ops.linear_residual(act, wd_d, d(res.clone()))
    fused = ops.layer_mlp_block(d(res.clone()), d(ln), wgu, wd_t).clone()
    torch.cuda.synchronize()
    r16 = lambda x: x.to(BF16).float()  # noqa: E731
    h = rmsnorm(res.float(), ln, 1e-5, "bf16")
    gg, uu = r16(h @ wg.float().T), r16(h @ wu.float().T)
    a = r16(gg / (1 + torch.exp(-gg)) * uu)
    y = r16(a @ wd.float().T)
    want = r16(y + res.float())
    for got, what in ((fused, "fused"), (chain, "chain")):
        err = (got.float().cpu() - want).abs()
        # the activations are 1 bf16 ulp loose (SiLU through __expf), so the down-proj term gets 3 ulps + the final rounding
        tol = 3.0 * ULP * y.abs() + 1.01 * ULP * want.abs() + 2e-3
        assert (err <= tol).all(), f"T={T} {what}: max err {err.max():.3e}, {int((err > tol).sum())} elements off"
    # against the chain: same activations bit for bit, so only the down-proj summation order differs
    diff = (fused.float() - chain.float()).abs().cpu()
    assert (diff <= 1.01 * ULP * want.abs() + 1e-6).all(), f"T={T}: fused vs chain off by more than the final rounding"
    assert (diff > 0).float().mean() < 0.05, "more than 5 % of the elements round differently from the chain"
    # the accumulator is handed back empty: a second launch over the same scratch gives the same bits
    assert torch.equal(ops.layer_mlp_block(d(res.clone()), d(ln), wgu, wd_t), fused)


@pytest.mark.parametrize("n_layers", [2, 3, 5])
def test_engine_two_launch_layer_streams_and_accumulators(ops, n_layers, monkeypatch):
    """T = 2 and T = 4 rows through the engine: greedy streams against the oracle (the bar of the other engine tests),
    bit-reproducible, graph and eager agree, and the four fixed-point accumulators are zero after every run — even
    (2) and odd (3, 5) layer counts exercise both parities of the accumulator rotation and the last layer's hand-over."""
    monkeypatch.setenv("CBX_MLP_FUSED", "1")
    n_steps = 24
    streams = {}
    for use_graph in (True, False):
        cfg, w, eng = _mk_engine(n_layers, 2, use_graph=use_graph)
        for n_pairs in (1, 2):
            reqs = _requests(n_pairs, n_steps, temperature=0.0, repetition_penalty=2.0)
            got = eng.generate(reqs, poll_every=5)
            l0 = eng.gpu_launches                         # one more step over the (finished, inert) slots: count its launches
            with torch.cuda.stream(eng.stream):           # graph capture needs a real stream, not the legacy default one
                eng.decode(n_pairs, 1, 64)
            torch.cuda.synchronize()
            assert eng.gpu_launches - l0 == 2 * (n_layers - 1) + 3 + 2, "the two-launch layer did not run"
            acc = eng.buffer("block_acc", (4 * 4 * 1024,), torch.int64)
            assert int(acc.abs().max()) == 0, "a fixed-point accumulator was left non-zero after the run"
            assert eng.generate(reqs, poll_every=7) == got, "same batch must give the same bits"
            # near-ties between the top two logits of these sharp random models flip under any change of summation
            # order (the two-row run of 3 layers has one at step 2, 1.4 ulps): the 4-ulp bar applies from step 0
            assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, first_exact=0, temperature=0.0, repetition_penalty=2.0)
            streams[(use_graph, n_pairs)] = got
        eng.close()
    for n_pairs in (1, 2):
        assert streams[(True, n_pairs)] == streams[(False, n_pairs)], "graph replay and eager launches disagree"


@pytest.mark.parametrize("n_layers,n_pairs", [(3, 1), (3, 2), (4, 1), (5, 2)])
def test_engine_two_launch_layer_vs_three_launch_layer(ops, n_layers, n_pairs, monkeypatch):
    """The same decode steps on the two-launch layer (default) and on the three-launch layer it replaces (what runs
    while a trace buffer is set), teacher-forced through identical requests: the last step's logits agree within a few
    bf16 roundings almost everywhere (a stale or half-folded accumulator would show up as gross differences)."""
    monkeypatch.setenv("CBX_MLP_FUSED", "1")
    n_steps = 6
    cfg, w, eng = _mk_engine(n_layers, 2, use_graph=False)
    reqs = _requests(n_pairs, n_steps, temperature=0.0, repetition_penalty=2.0)
    a = eng.generate(reqs, poll_every=5)
    lg_a = eng.buffer("logits", (2 * n_pairs, 8256), BF16)[:, :8194].clone()
    # the trace hook writes n_layers x T rows for EVERY forward, the prefill (T = all prompt rows, < 400 here) included
    trace = torch.zeros(n_layers, 512, 1024, dtype=BF16, device="cuda")
    eng.set_trace(trace)
    b = eng.generate(reqs, poll_every=5)
    lg_b = eng.buffer("logits", (2 * n_pairs, 8256), BF16)[:, :8194].clone()
    eng.set_trace(None)
    assert float(trace.float().abs().sum()) > 0, "the traced run did not write its trace"
    acc = eng.buffer("block_acc", (4 * 4 * 1024,), torch.int64)
    assert int(acc.abs().max()) == 0
    if a == b:                                            # same tokens fed => same step inputs: compare the last step's logits
        diff = (lg_a.float() - lg_b.float()).abs()
        tol = 4.0 * ULP * lg_b.float().abs() + 0.02 * float(lg_b.float().abs().max())
        assert (diff <= tol).float().mean() > 0.99, f"max diff {diff.max():.3e} of {float(lg_b.float().abs().max()):.1f}"
    else:                                                 # a near-tie flipped a token: both streams must still meet the oracle's bar
        for got in (a, b):
            assert_stream_parity(got, w, reqs, eng.cfg_scale, n_steps, first_exact=0, temperature=0.0, repetition_penalty=2.0)
    eng.close()
