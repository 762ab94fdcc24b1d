"""torchrun --nproc-per-node 2 tools/shard_check.py — the product's sharded generate on two GPUs over NCCL against the
same requests on one rank's engine alone: every stream arrives, lengths and ordering are exact, each request's tokens are
(near-)identical whichever rank count ran it (uid-keyed Philox; batch composition differs, so bf16 near-ties may flip),
and both runs follow the CPU oracle (teacher-forced) on probed requests.  Prints one JSON line on rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import torch.distributed as dist
from chatterbox_vllm_b200.engine import T3Engine
from chatterbox_vllm_b200.sharding import generate_sharded, shard_indices
from oracle.t3_ref import T3ModelConfig, T3Weights
from test_gpu_engine import _requests, assert_stream_parity

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
w = T3Weights.random(T3ModelConfig(n_layers=2), seed=0, head_scale=16.0, norm_jitter=0.1)
eng = T3Engine(w.state_dict(), device=f"cuda:{local}", max_pairs=8, max_model_len=256)
n_steps = 24
for sp in (dict(temperature=0.0, repetition_penalty=2.0), dict(temperature=0.7, top_p=0.9, repetition_penalty=1.3)):
    reqs = _requests(11, n_steps, **sp)
    for r in reqs:
        r.uid = 0                                              # generate_sharded assigns global index + 1 on every rank
    out = generate_sharded(eng, reqs, all_ranks=True)
    assert [r.uid for r in reqs] == list(range(1, 12))
    assert len(out) == 11 and all(len(t) == n_steps for t in out)
    alone = eng.generate(reqs)                                 # the same requests (same uids) on one engine
    agree = sum(a == b for x, y in zip(out, alone) for a, b in zip(x, y))
    mine = shard_indices([34 + len(r.text_ids) + 1 + n_steps for r in reqs], 2, rank)
    if sp["temperature"] == 0.0:
        assert_stream_parity([out[i] for i in mine[:2]], w, [reqs[i] for i in mine[:2]], eng.cfg_scale, n_steps, **sp)
    assert agree >= 0.85 * 11 * n_steps, f"rank {rank}: only {agree}/{11 * n_steps} tokens equal the one-engine run"
    if rank == 0:
        print(json.dumps({"sampling": sp, "world": 2, "requests": 11, "tokens_equal_to_one_engine_run": agree, "of": 11 * n_steps}), flush=True)
dist.barrier()
eng.close()
dist.destroy_process_group()
