"""CPU oracle for paged-KV bookkeeping — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Pure-Python/numpy restatement of the block pool, block table and slot mapping
the reference inherits from vLLM (third-party, not vendored; pinned
``vllm==0.10.0`` in the reference's pyproject.toml:29 — algorithms restated from
the installed 0.22 sources, which keep the same policy):

  v1/core/block_pool.py:162-177   free queue holds blocks 0..N-1 in id order;
                                   block 0 is popped at construction as the null block
  v1/core/block_pool.py:333-347   get_new_blocks(n) = popleft n, error if n > free
  v1/core/block_pool.py:419-433   free_blocks(ordered) appends to the queue tail
  v1/core/single_type_kv_cache_manager.py:120   blocks needed = cdiv(num_tokens, block_size)
  v1/core/single_type_kv_cache_manager.py:350   a request's blocks are freed in REVERSE order
  v1/worker/block_table.py:357-377   slot = block_table[req, pos // bs]·bs + pos % bs ; pad = -1

Everything here is integer work: the CUDA/C++ side must match it BIT-EXACTLY.
PARITY STATUS: unpinned by the reference (it has no tests); the formulae are
checked against vLLM 0.22's own BlockPool in tests/test_oracle_pinning.py.
"""
from __future__ import annotations

from collections import deque

import numpy as np

BLOCK_SIZE = 16
PAD_SLOT = -1


def cdiv(a: int, b: int) -> int:
    return -(-a // b)


class BlockPool:
    def __init__(self, num_blocks: int):
        self.num_blocks = num_blocks
        self.free = deque(range(num_blocks))
        self.null_block = self.free.popleft()          # block 0

    def num_free(self) -> int:
        return len(self.free)

    def get_new_blocks(self, n: int) -> list[int]:
        if n > len(self.free):
            raise ValueError(f"Cannot get {n} free blocks from the pool")
        return [self.free.popleft() for _ in range(n)]

    def free_blocks(self, ordered: list[int]) -> None:
        self.free.extend(b for b in ordered if b != self.null_block)


class BlockTable:
    """int32 [max_seqs, max_blocks_per_seq]; unused entries are 0 (the null block),
    as in vLLM's zero-initialised block table tensor."""

    def __init__(self, pool: BlockPool, max_seqs: int, max_model_len: int, block_size: int = BLOCK_SIZE):
        self.pool, self.bs = pool, block_size
        self.table = np.zeros((max_seqs, cdiv(max_model_len, block_size)), np.int32)
        self.nblk = np.zeros(max_seqs, np.int32)

    def ensure(self, row: int, num_tokens: int) -> list[int]:
        """Grow ``row`` so it can hold ``num_tokens`` tokens."""
        need = cdiv(num_tokens, self.bs) - int(self.nblk[row])
        if need <= 0:
            return []
        new = self.pool.get_new_blocks(need)
        n0 = int(self.nblk[row])
        self.table[row, n0:n0 + need] = new
        self.nblk[row] = n0 + need
        return new

    def release(self, row: int) -> None:
        n = int(self.nblk[row])
        self.pool.free_blocks(list(reversed(self.table[row, :n].tolist())))
        self.table[row, :n] = 0
        self.nblk[row] = 0


def slot_mapping(table: np.ndarray, req: np.ndarray, pos: np.ndarray, block_size: int = BLOCK_SIZE) -> np.ndarray:
    """int64 slots for token rows (req[i], pos[i]); pos < 0 marks a padded row → -1."""
    req = np.asarray(req, np.int64)
    pos = np.asarray(pos, np.int64)
    ok = pos >= 0
    p = np.where(ok, pos, 0)
    blk = table[req, p // block_size].astype(np.int64)
    return np.where(ok, blk * block_size + p % block_size, PAD_SLOT).astype(np.int64)
