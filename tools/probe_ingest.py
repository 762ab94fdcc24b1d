"""How fast do bytes enter the SMs from L2?  tcgen05 GEMM (path 2, 128 x 64 tiles: 24 KB per k-step per CTA,
MMA time negligible) with n CTAs over L2-resident operands; per-CTA and aggregate ingest rates from the
slope between two K values (removes launch / prologue / epilogue)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops
os.environ["CBX_GEMM_BN"] = "64"
dev = torch.device("cuda:0")


def timeit(fn, reps=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3   # us


for M in (128, 256):
    for n in (1, 8, 37, 74, 148):
        res = []
        for K in (1024, 4096):
            x = torch.randn(M, K, device=dev).to(torch.bfloat16)
            W = torch.randn(64 * n, K, device=dev).to(torch.bfloat16)
            res.append(timeit(lambda: ops.linear(x, W, path=2)))
        ctas = n * (M // 128)
        dbytes = (4096 - 1024) // 64 * 24576
        dt = res[1] - res[0]
        print(f"M={M} CTAs={ctas:4d}: K=1024 {res[0]:6.1f} us, K=4096 {res[1]:6.1f} us -> {dbytes / dt / 1e3:6.1f} GB/s per CTA, "
              f"{ctas * dbytes / dt / 1e6:6.2f} TB/s aggregate", flush=True)
