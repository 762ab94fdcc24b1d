#!/usr/bin/env python3
"""Times cfg_sample_kernel alone (B = 1 and B = 128 pairs) under different sampler settings, to see
which stage of its serial chain the time goes to: greedy (load + argmax), temperature only
(+ softmax + exponential race), top-p (+ pivot selection), top-k + top-p.  Logits are the two shapes
the engine sees: near-flat (random-init head) and sharp (head x 16).  Each number is one CUDA graph
of 100 launches, so no host launch cost is in it.

    python tools/bench_sampler.py
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from chatterbox_vllm_b200 import ops

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    out = []
    for B in (1, 128):
        for shape, scale in (("flat", 0.65), ("sharp", 10.0)):
            g = torch.Generator().manual_seed(7)
            logits = (torch.randn(2 * B, ops.LOGITS_LD, generator=g) * scale).to(torch.bfloat16).to(dev)
            for name, kw in (("greedy", dict(temperature=0.0)),
                             ("temp", dict(temperature=0.5)),
                             ("top_p", dict(temperature=0.5, top_p=0.5)),
                             ("top_p.9", dict(temperature=0.5, top_p=0.9)),
                             ("top_k+p", dict(temperature=0.5, top_p=0.5, top_k=50)),
                             ("min_p+p", dict(temperature=0.5, top_p=0.5, min_p=0.05))):
                sb = ops.SamplerBuffers(B, 4096, dev)
                sb.temperature.fill_(kw.get("temperature", 0.0))
                sb.top_p.fill_(kw.get("top_p", 1.0))
                sb.min_p.fill_(kw.get("min_p", 0.0))
                sb.top_k.fill_(kw.get("top_k", 0))
                sb.rep_pen.fill_(2.0)
                sb.max_out.fill_(4096)
                sb.ignore_eos.fill_(1)
                sb.uid.copy_(torch.arange(B, device=dev) + 5)
                n = 100
                for _ in range(5):
                    ops.cfg_sample(logits, sb, B)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.stream(side):
                    with torch.cuda.graph(gr, stream=side):
                        for _ in range(n):
                            ops.cfg_sample(logits, sb, B)
                    sb.n_out.zero_()
                    gr.replay()
                    sb.n_out.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    gr.replay()
                    e1.record()
                torch.cuda.synchronize()
                out.append({"B": B, "logits": shape, "mode": name, "us": e0.elapsed_time(e1) * 1e3 / n})
                print(json.dumps(out[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "bench_sampler.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
