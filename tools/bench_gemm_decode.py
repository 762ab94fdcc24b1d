"""Decode-sized (M = 256) tcgen05 GEMMs inside a CUDA graph (no host launch cost): us per launch for the
gate_up shape with and without the SiLU epilogue, and the other layer shapes.  Weights rotate over several
buffers so they stream from HBM as in the step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops

dev = torch.device("cuda:0")
BF16 = torch.bfloat16
M = int(os.environ.get("M", "256"))
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: (torch.randn(*s, generator=g, device=dev) * 0.05).to(BF16)  # noqa: E731
x1, x4, res = rn(M, 1024), rn(M, 4096), rn(M, 1024)
NW = 12


def bench(name, N, K, fn, reps=48):
    Ws = [rn(N, K) for _ in range(NW)]
    outs = fn(Ws[0], None)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for i in range(3):
            fn(Ws[i % NW], outs)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for i in range(reps):
                fn(Ws[i % NW], outs)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(5):
            gr.replay()
        e1.record(st)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / (5 * reps) * 1e3
    print(f"{name:28s} N={N:5d} K={K:5d}: {us:6.2f} us/launch  ({2.0 * M * N * K / us / 1e6:6.0f} TFLOP/s)", flush=True)


bench("gate_up + SiLU (epi 1)", 8192, 1024, lambda W, o: ops.linear(x1, W, epilogue=1, out=o))
bench("gate_up plain  (epi 0)", 8192, 1024, lambda W, o: ops.linear(x1, W, out=o))
bench("qkv plain", 3072, 1024, lambda W, o: ops.linear(x1, W, out=o))
bench("o partials x4", 1024, 1024, lambda W, o: ops.linear_partials(x1, W, 4))
bench("down partials x4", 1024, 4096, lambda W, o: ops.linear_partials(x4, W, 4))
bench("down + residual", 1024, 4096, lambda W, o: ops.linear_residual(x4, W, res))
