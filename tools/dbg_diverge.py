import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_engine import _mk_engine, _requests, _oracle_stream
from oracle import sampler_ref
n_steps = 24
for graph in (False, True):
    cfg, w, eng = _mk_engine(2, 4, use_graph=graph)
    reqs = _requests(3, n_steps, temperature=0.0, repetition_penalty=2.0)
    got = eng.generate(reqs, poll_every=5)
    want, logs = _oracle_stream(w, reqs, eng.cfg_scale, n_steps, teacher=got, temperature=0.0, repetition_penalty=2.0)
    for r in range(3):
        for s in range(n_steps):
            if got[r][s] != want[r][s]:
                seen = np.zeros(8194, bool)
                for t in got[r][:s]: seen[t - 1000] = True
                l = sampler_ref.apply_repetition_penalty(logs[r][s], seen, 2.0)
                print(f"graph={graph} req {r} step {s}: eng {got[r][s]} orc {want[r][s]}  l[eng]={l[got[r][s]-1000]:.4f} l[orc]={l[want[r][s]-1000]:.4f} max|l|={np.abs(l).max():.2f}")
    print("graph", graph, "done; mismatches listed above")
    eng.close()
