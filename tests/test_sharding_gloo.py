"""Multi-process (world_size 2, gloo, CPU) test of the request-sharding path: longest-first
round-robin partition + the single gather of token streams to rank 0."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from chatterbox_vllm_b200.sharding import gather_token_streams, shard_indices


def test_shard_indices_partition_and_balance():
    est = [50, 400, 120, 120, 7, 300, 90, 33, 260, 15]
    for world in (1, 2, 4, 8):
        parts = [shard_indices(est, world, r) for r in range(world)]
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(len(est))), "every prompt goes to exactly one rank"
        loads = [sum(est[i] for i in p) for p in parts]
        if world == 2:
            assert max(loads) - min(loads) <= max(est)
    assert shard_indices(est, 2, 0)[:2] == [1, 8]          # longest first, dealt in turn
    assert shard_indices([], 4, 3) == []


def _fake_stream(i):                                       # what a rank's engine would emit for prompt i
    g = torch.Generator().manual_seed(i)
    n = 3 + i % 5
    return (torch.randint(0, 6561, (n,), generator=g) + 1000).tolist()


def _worker(rank, world, port, est, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    idx = shard_indices(est, world, rank)
    local = [_fake_stream(i) for i in idx]
    out = gather_token_streams(local, idx, len(est))
    if rank == 0:
        q.put(out)
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


def test_gather_token_streams_world2_gloo():
    est = [50, 400, 120, 120, 7, 300, 90]                  # 7 prompts: ragged split 4 / 3
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, est, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out == [_fake_stream(i) for i in range(len(est))], "rank 0 must see every stream in global order"


def test_gather_single_process_passthrough():
    assert gather_token_streams([[1, 2], [3]], [1, 0], 2) == [[3], [1, 2]]
