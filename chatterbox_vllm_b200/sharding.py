"""Request sharding across GPUs (SURVEY §8 e): every CFG pair is independent, so the T3 path
scales by giving each rank its own prompts — one process per GPU, no data-path collective.  The only
exchange is the gather of emitted token streams to rank 0 (NCCL over NVLink on GPUs; gloo in the
CPU tests).  The reference has no multi-GPU mode (single `LLM`, tts.py:140)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_indices(est_len: list[int], world: int, rank: int) -> list[int]:
    """Longest-first round-robin: sort prompts by estimated length (descending, stable on index),
    deal them out to ranks in turn — balances KV bytes per rank.  Returns this rank's global
    prompt indices in dealing order."""
    order = sorted(range(len(est_len)), key=lambda i: (-est_len[i], i))
    return order[rank::world]


def gather_token_streams(local: list[list[int]], local_idx: list[int], n_total: int,
                         device: torch.device | str = "cpu", all_ranks: bool = False) -> list[list[int]] | None:
    """All ranks call this with their own streams (`local[j]` belongs to global prompt
    `local_idx[j]`); rank 0 (every rank with all_ranks=True) gets the full list in global order, the
    others None.  One padded all_gather of int32 [n_local_max, max_len + 2] per rank (≤ a few MB:
    latency-bound)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        out: list = [None] * n_total
        for i, t in zip(local_idx, local):
            out[i] = list(t)
        return out
    meta = torch.tensor([len(local), max((len(t) for t in local), default=0)], dtype=torch.int64, device=device)
    metas = [torch.zeros_like(meta) for _ in range(world)]
    dist.all_gather(metas, meta)
    n_max = int(max(m[0] for m in metas))
    l_max = int(max(m[1] for m in metas))
    buf = torch.full((max(n_max, 1), l_max + 2), -1, dtype=torch.int32, device=device)   # [idx, len, tokens…]
    for j, (i, t) in enumerate(zip(local_idx, local)):
        buf[j, 0], buf[j, 1] = i, len(t)
        if t:
            buf[j, 2:2 + len(t)] = torch.tensor(t, dtype=torch.int32, device=device)
    allb = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(allb, buf)
    if dist.get_rank() != 0 and not all_ranks:
        return None
    out = [None] * n_total
    for b in allb:
        b = b.cpu()
        for row in b:
            i, n = int(row[0]), int(row[1])
            if i >= 0:
                out[i] = row[2:2 + n].tolist()
    return out


def generate_sharded(engine, reqs: list, est_len: list[int] | None = None, all_ranks: bool = False,
                     poll_every: int = 32) -> list[list[int]] | None:
    """`engine.generate` for one process per GPU (torchrun): every rank is given the SAME request list, runs its
    share — longest-first round-robin on `est_len` (default: prefix + max_tokens) — on its own engine, and the token
    streams are gathered over the default process group (NCCL on GPUs).  Requests without a uid get their global
    index + 1, so the Philox streams do not depend on the number of ranks.  With no process group (or one rank) this
    is engine.generate.  Returns the streams in request order on rank 0 (on every rank with all_ranks=True)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    for i, r in enumerate(reqs):
        if r.uid == 0:
            r.uid = i + 1
    if world == 1:
        return engine.generate(reqs, poll_every)
    if est_len is None:
        est_len = [34 + len(r.text_ids) + 1 + r.params.max_tokens for r in reqs]
    idx = shard_indices(est_len, world, dist.get_rank())
    local = engine.generate([reqs[i] for i in idx], poll_every) if idx else []
    dev = engine.device if dist.get_backend() == "nccl" else "cpu"
    return gather_token_streams(local, idx, len(reqs), device=dev, all_ranks=all_ranks)
