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


class _FakeEngine:
    """Emits, per request, a stream that depends only on its uid — what the real engine's Philox keying guarantees."""
    device = "cpu"

    def generate(self, reqs, poll_every=32):
        return [_fake_stream(r.uid) for r in reqs]


def _req(n_text, max_tokens, uid=0):
    from chatterbox_vllm_b200.engine import Request, SamplingParams
    return Request([255] + [7] * n_text + [0], torch.zeros(1), SamplingParams(max_tokens=max_tokens), uid=uid)


def _worker_product(rank, world, port, q):
    from chatterbox_vllm_b200.sharding import generate_sharded
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    reqs = [_req(5 + 3 * i, 40 + (i * 7) % 30) for i in range(9)]          # every rank holds the same request list
    out = generate_sharded(_FakeEngine(), reqs, all_ranks=(rank == 1) or None)
    out_all = generate_sharded(_FakeEngine(), reqs, all_ranks=True)
    q.put((rank, out, out_all, [r.uid for r in reqs]))
    dist.barrier()
    dist.destroy_process_group()


def test_generate_sharded_world2_gloo_is_rank_count_independent():
    """The product path (sharding.generate_sharded, what ChatterboxTTS.generate_tokens calls under torchrun): uids are the
    global prompt index + 1 on every rank, each rank runs only its share, and the gathered streams equal the one-process
    result — rank-count independence."""
    from chatterbox_vllm_b200.sharding import generate_sharded
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_product, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict()
    for _ in range(2):
        rank, out, out_all, uids = q.get(timeout=120)
        got[rank] = (out, out_all, uids)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    single = generate_sharded(_FakeEngine(), [_req(5 + 3 * i, 40 + (i * 7) % 30) for i in range(9)])   # no process group: one rank
    assert got[0][0] == single and got[0][1] == single and got[1][1] == single
    assert got[1][0] == single                                  # rank 1 asked for all_ranks=True in its first call too
    assert got[0][2] == got[1][2] == list(range(1, 10))
