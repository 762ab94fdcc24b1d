"""Per-kernel timing (CUDA events, back-to-back launches, weights rotated over 30 copies so they
come from HBM like in the real step).  python tools/bench_kernels.py [T ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops

BF16 = torch.bfloat16
dev = torch.device("cuda:0")
L = 30


def timeit(fn, reps=5):
    """Capture L launches into a CUDA graph (no host overhead in the timed region), replay."""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for i in range(L):
            fn(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(L):
                fn(i)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * L) * 1e3


def main():
    hot = os.environ.get('CBX_HOT') == '1'     # reuse one weight copy: everything L2-resident
    Ts = [int(a) for a in sys.argv[1:]] or [2, 256]
    g = torch.Generator(device=dev).manual_seed(0)
    rn = lambda *s, sc=0.02: (torch.randn(*s, generator=g, device=dev) * sc).to(BF16)  # noqa: E731
    wqkv = [rn(3072, 1024) for _ in range(L)]
    wo = [rn(1024, 1024) for _ in range(L)]
    wgu = [rn(8192, 1024) for _ in range(L)]
    wd = [rn(1024, 4096) for _ in range(L)]
    whead = rn(8194, 1024)
    if hot:
        wqkv, wo, wgu, wd = [wqkv[0]] * L, [wo[0]] * L, [wgu[0]] * L, [wd[0]] * L
    ln = torch.ones(1024, dtype=BF16, device=dev)
    cs = rn(2048, 64, sc=1.0)
    for T in Ts:
        ctx = 300 if T <= 4 else 450
        res = rn(T, 1024, sc=1.0)
        attn_in = rn(T, 1024, sc=1.0)
        act = rn(T, 4096, sc=1.0)
        pos = torch.full((T,), ctx, dtype=torch.int32, device=dev)
        n_pages = (ctx + 16) // 16
        nb = T * n_pages + 1
        bt = (torch.arange(T * n_pages, dtype=torch.int32, device=dev) + 1).view(T, n_pages).contiguous()
        slot = (bt[:, ctx // 16].to(torch.int64) * 16 + ctx % 16).contiguous()
        n_kv = L if T <= 4 else 4
        kv = torch.zeros(n_kv * nb * 65536, dtype=torch.uint8, device=dev)
        kvl = [kv[i * nb * 65536:(i + 1) * nb * 65536] for i in range(n_kv)]
        rows = torch.arange(T, dtype=torch.int32, device=dev)
        kv_len = torch.full((T,), ctx + 1, dtype=torch.int32, device=dev)
        qkv = rn(T, 3072, sc=1.0)
        out = torch.empty(T, 1024, dtype=BF16, device=dev)
        r = {}
        h = torch.empty(T, 1024, dtype=BF16, device=dev)
        actb = torch.empty(T, 4096, dtype=BF16, device=dev)
        lgb = torch.zeros(T, 8256, dtype=BF16, device=dev)
        r["qkv"] = timeit(lambda i: ops.layer_qkv(res, ln, wqkv[i], cs, pos, slot, kvl[i % n_kv], qkv=qkv, h=h))
        r["attn"] = timeit(lambda i: ops.paged_attn_decode(qkv, kvl[i % n_kv], bt, rows, kv_len, ctx + 1, out=out))
        r["o"] = timeit(lambda i: ops.linear_residual(attn_in, wo[i], res))
        r["gate_up"] = timeit(lambda i: ops.layer_gate_up(res, ln, wgu[i], act=actb, h=h))
        r["down"] = timeit(lambda i: ops.linear_residual(act, wd[i], res))
        r["norm"] = timeit(lambda i: ops.rmsnorm(res, ln))
        r["head"] = timeit(lambda i: ops.head_logits(res, ln, whead, logits=lgb, h=h))
        B = max(T // 2, 1)
        for name, scale in (("flat", 0.3), ("sharp", 5.0)):
            logits = torch.zeros(2 * B, 8256, dtype=BF16, device=dev)
            logits[:, :8194] = rn(2 * B, 8194, sc=scale)
            sb = ops.SamplerBuffers(B, 1 << 16, dev)
            sb.temperature.fill_(0.5), sb.top_p.fill_(0.5), sb.rep_pen.fill_(2.0), sb.max_out.fill_(1 << 16)
            sb.ignore_eos.fill_(1)
            r["sample_" + name] = timeit(lambda i: ops.cfg_sample(logits, sb, B))
        print(f"T={T} ctx={ctx}: " + "  ".join(f"{k}={v:.1f}us" for k, v in r.items()), flush=True)


if __name__ == "__main__":
    main()
