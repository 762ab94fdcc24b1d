"""tcgen05 GEMM throughput at prefill sizes.  python tools/bench_gemm.py [M ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops

dev = torch.device("cuda:0")
BF16 = torch.bfloat16


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for M in [int(a) for a in sys.argv[1:]] or [16000]:
    g = torch.Generator(device=dev).manual_seed(0)
    rn = lambda *s: (torch.randn(*s, generator=g, device=dev) * 0.05).to(BF16)  # noqa: E731
    x1, x4 = rn(M, 1024), rn(M, 4096)
    res = rn(M, 1024)
    out = {}
    for name, N, K, fn in (
            ("qkv-plain", 3072, 1024, lambda W: ops.linear(x1, W)),
            ("gate_up+silu", 8192, 1024, lambda W: ops.linear(x1, W, epilogue=1)),
            ("o+res", 1024, 1024, lambda W: ops.linear_residual(x1, W, res)),
            ("down+res", 1024, 4096, lambda W: ops.linear_residual(x4, W, res)),
            ("head", 8192, 1024, lambda W: ops.linear(x1, W))):
        W = rn(N, K)
        ms = timeit(lambda: fn(W))
        out[name] = (ms, 2.0 * M * N * K / (ms / 1e3) / 1e12)
    ref = timeit(lambda: torch.matmul(x1, rn(8192, 1024).t()))
    print(f"M={M}: " + "  ".join(f"{k}={v[0] * 1e3:.0f}us/{v[1]:.0f}TF" for k, v in out.items()) +
          f"  | torch.matmul 8192x1024: {ref * 1e3:.0f}us/{2.0 * M * 8192 * 1024 / (ref / 1e3) / 1e12:.0f}TF", flush=True)
