"""T3VllmModel — the reference's vLLM model class (models/t3/t3.py:302-781) over the B200 engine.

Same method names and tensor contracts as the class the reference registers as "ChatterboxT3":

    __init__(*, vllm_config, prefix)                      t3.py:305
    load_weights(Iterable[(name, Tensor)]) -> set[str]    t3.py:347
    get_multimodal_embeddings(**kw) -> list[Tensor]       t3.py:382   (vLLM >= 0.11: embed_multimodal)
    get_input_embeddings(input_ids, mm) -> [N, 2048]      t3.py:498   (vLLM >= 0.11: embed_input_ids)
    forward(input_ids, positions, intermediate_tensors, inputs_embeds) -> [N, 2048]    t3.py:742
    compute_logits(hidden_states[, sampling_metadata]) -> [B, 9194]                    t3.py:725
    get_language_model()                                  t3.py:780

The 2048-wide rows are the reference's trick for carrying the CFG pair through vLLM: columns
[0,1024) belong to the conditional sequence, [1024,2048) to the unconditional one.  ``forward``
un-stacks them into 2N rows, runs the hand-written layer kernels of libcbx_t3.so over this class's
OWN paged KV cache (64 KiB pages, both members of the pair cached and attended — the north_star
semantics) and re-stacks the normalised hidden states.

What a vLLM scheduler supplies implicitly through its attention metadata is explicit here:
``forward(..., seq_ids=IntTensor[N])`` names the request slot of every row (default: one request when
the positions are consecutive, else row i = slot i), and ``free_sequences(slots)`` returns a finished
request's blocks.  vLLM's own KV cache / attention backend are not used: their layout is not the
layout these kernels read."""
from __future__ import annotations

import os
from typing import Iterable, Optional

import torch

from ... import ops
from ...engine import BF16, T3Engine
from .processor import (CONDITIONING_SIZE, SPEECH_TOKEN_OFFSET, EmbeddingTables, assemble_input_embeddings,
                        split_prefill_decode)

BLOCK = 16


class T3VllmModel(torch.nn.Module):
    def __init__(self, *, vllm_config=None, prefix: str = "", device: str = "cuda:0", max_num_seqs: Optional[int] = None,
                 max_model_len: Optional[int] = None):
        super().__init__()
        self.vllm_config, self.prefix = vllm_config, prefix
        mc = getattr(vllm_config, "model_config", None)
        sc = getattr(vllm_config, "scheduler_config", None)
        hf = getattr(mc, "hf_config", None)
        if hf is not None and getattr(hf, "hidden_size", 1024) == 2048:
            hf.hidden_size = 1024                             # undo the reference's config hack (t3.py:307-310)
        self.is_multilingual = bool(getattr(hf, "is_multilingual", False))
        self.max_num_seqs = int(max_num_seqs or getattr(sc, "max_num_seqs", None) or 16)
        self.max_model_len = int(max_model_len or getattr(mc, "max_model_len", None) or 1000)
        self.device = torch.device(device)
        self.dim = 1024
        self.cfg_scale = float(os.environ.get("CHATTERBOX_CFG_SCALE", "0.5"))     # read once (t3.py:343)
        print("Applying CFG scale:", self.cfg_scale)
        self.engine: Optional[T3Engine] = None
        self.tables: Optional[EmbeddingTables] = None
        self.cond_enc_weights: dict = {}
        self._blocks: dict[int, list[int]] = {}               # block-table row -> blocks held

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights: Iterable[tuple[str, torch.Tensor]]) -> set[str]:
        """Checkpoint names of t3_cfg.safetensors.  q/k/v and gate/up are stacked exactly as vLLM's
        Llama loader does; ``cond_enc.*`` (conditioning encoder, outside this path) is kept aside."""
        sd, loaded = {}, set()
        for name, w in weights:
            loaded.add(name)
            if name.startswith("cond_enc."):
                self.cond_enc_weights[name] = w
            else:
                sd[name] = w
        self.engine = T3Engine(sd, device=str(self.device), max_pairs=self.max_num_seqs, max_model_len=self.max_model_len,
                               cfg_scale=self.cfg_scale, use_graph=False)
        e = self.engine
        e._B = self.max_num_seqs                              # block-table rows: slot s (cond), max_num_seqs + s (uncond)
        self.tables = EmbeddingTables(e.text_emb, e.text_pos, e.speech_emb, e.speech_pos)
        return loaded

    def get_language_model(self):
        return self.engine

    # ------------------------------------------------------------------ embeddings
    def get_multimodal_embeddings(self, **kwargs: object):
        conditionals = kwargs.get("conditionals", [])
        return [batch[0] for batch in conditionals]

    embed_multimodal = get_multimodal_embeddings              # vLLM >= 0.11 spelling

    def split_prefill_decode(self, input_ids: torch.Tensor, multimodal_embeddings):
        return split_prefill_decode(input_ids, multimodal_embeddings)

    def get_input_embeddings(self, input_ids: torch.Tensor, multimodal_embeddings=None) -> torch.Tensor:
        return assemble_input_embeddings(input_ids.to(self.device), multimodal_embeddings, self.tables)

    def embed_input_ids(self, input_ids: torch.Tensor, multimodal_embeddings=None, **kw) -> torch.Tensor:
        return self.get_input_embeddings(input_ids, multimodal_embeddings)

    # ------------------------------------------------------------------ KV bookkeeping
    def _ensure_blocks(self, row: int, n_tokens: int) -> bool:
        held = self._blocks.setdefault(row, [])
        need = -(-n_tokens // BLOCK) - len(held)
        if need <= 0:
            return False
        held.extend(self.engine.pool.alloc(need))
        return True

    def free_sequences(self, slots: Iterable[int]) -> None:
        """Return the KV blocks of finished request slots (both members of each pair)."""
        for s in slots:
            for row in (s, self.max_num_seqs + s):
                blks = self._blocks.pop(row, [])
                if blks:
                    self.engine.pool.free(list(reversed(blks)))

    # ------------------------------------------------------------------ forward
    def forward(self, input_ids: Optional[torch.Tensor], positions: torch.Tensor, intermediate_tensors=None,
                inputs_embeds: Optional[torch.Tensor] = None, *, seq_ids: Optional[torch.Tensor] = None,
                **kwargs: object) -> torch.Tensor:
        e = self.engine
        if inputs_embeds is None:
            inputs_embeds = self.get_input_embeddings(input_ids, [])
        N = inputs_embeds.shape[0]
        if 2 * N > e.max_rows:
            raise ValueError(f"{N} rows exceed the engine's row capacity {e.max_rows // 2}")
        pos_l = [int(p) for p in positions.tolist()]
        if seq_ids is None:
            # Without request ids the rows can only be ONE request's consecutive positions in slot 0 (a prefix chunk
            # or its next decode row).  Anything else is ambiguous — rows of different requests would be written into
            # one KV sequence and the logits would be silently wrong — so it is refused, not guessed.
            consecutive = all(pos_l[i + 1] == pos_l[i] + 1 for i in range(N - 1))
            others = {r % self.max_num_seqs for r in self._blocks} - {0}
            if not consecutive or others:
                raise ValueError("T3VllmModel.forward: seq_ids (the request slot of every row) is required when more than "
                                 "one request is live or the rows are not one request's consecutive positions")
            seq_l = [0] * N
        else:
            seq_l = [int(s) for s in seq_ids.tolist()]
        if N and (max(pos_l) >= self.max_model_len or max(seq_l) >= self.max_num_seqs or min(seq_l) < 0):
            raise ValueError("position or request slot out of range")
        # blocks for every (slot, position) touched, both members of the pair
        dirty = set()
        top = {}
        for s, p in zip(seq_l, pos_l):
            top[s] = max(top.get(s, 0), p + 1)
        for s, n_tok in top.items():
            for row in (s, self.max_num_seqs + s):
                if self._ensure_blocks(row, n_tok):
                    dirty.add(row)
        if dirty:
            rows = sorted(dirty)
            tab = torch.zeros(len(rows), e.bt_stride, dtype=torch.int32)
            for i, r in enumerate(rows):
                tab[i, :len(self._blocks[r])] = torch.tensor(self._blocks[r], dtype=torch.int32)
            e.block_table.index_copy_(0, torch.tensor(rows, device=self.device), tab.to(self.device))
        T = 2 * N
        dev = self.device
        pos = torch.tensor(pos_l + pos_l, dtype=torch.int32, device=dev)
        rseq = torch.tensor(seq_l + [self.max_num_seqs + s for s in seq_l], dtype=torch.int32, device=dev)
        emb = inputs_embeds.to(device=dev, dtype=BF16)
        x = e.buffer("x", (e.max_rows, self.dim), BF16)
        x[:N].copy_(emb[:, :self.dim])                        # t3.py:757,768-773: split and stack the halves
        x[N:T].copy_(emb[:, self.dim:])
        e.buffer("pos", (T,), torch.int32).copy_(pos)
        e.buffer("row_seq", (T,), torch.int32).copy_(rseq)
        e.buffer("kv_len", (T,), torch.int32).copy_(pos + 1)  # causal: row at position p sees p + 1 cached tokens
        e.buffer("slot", (T,), torch.int64).copy_(ops.slot_mapping(e.block_table, rseq, pos))
        e.forward(T, max(pos_l) + 1 if N else 1)
        h = e.buffer("h", (e.max_rows, self.dim), BF16)
        return torch.cat([h[:N], h[N:T]], dim=1)              # t3.py:776-778

    # ------------------------------------------------------------------ logits
    def compute_logits(self, hidden_states: torch.Tensor, sampling_metadata=None) -> torch.Tensor:
        """[B, 2048] -> [B, 1000 + 8194] bf16: both halves through the speech head, CFG merge with a bf16
        rounding after every op (as the reference's three torch ops), 1000 x -inf in front (t3.py:725-739)."""
        B = hidden_states.shape[0]
        hs = hidden_states.to(device=self.device, dtype=BF16)
        rows = torch.cat([hs[:, :self.dim], hs[:, self.dim:]], dim=0).contiguous()
        lg = ops.linear(rows, self.engine.speech_head, ld=8256)   # row stride a multiple of 64 elements: aligned vector stores
        cond, uncond = lg[:B], lg[B:]
        logits = cond + self.cfg_scale * (cond - uncond)
        pad = torch.full((B, SPEECH_TOKEN_OFFSET), float("-inf"), dtype=logits.dtype, device=logits.device)
        return torch.cat([pad, logits], dim=1)
