"""The per-kernel entry points of libcbx_t3.so as ``torch.library`` custom ops (namespace ``cbx``): schemas, CUDA
implementations (the ctypes calls of ``ops.py`` — the C-ABI underneath is unchanged) and fake ("meta") implementations,
so the ops are visible to the torch dispatcher, ``torch.compile`` / FakeTensor tracing and vLLM's custom-op plumbing
(north_star: "Host code is Python/PyTorch calling CUDA through a thin torch custom-op C-ABI layer"; SURVEY §8 b).

    import chatterbox_vllm_b200.torch_ops          # registers torch.ops.cbx.*
    out = torch.ops.cbx.paged_attn_decode(qkv, kv_layer, block_table, row_seq, kv_len, max_kv_len)

Ops that update a buffer in place (the residual stream, the KV pages) declare it in ``mutates_args`` and return
nothing, as torch's functionalisation requires.  There is no CPU kernel behind any of them: on a CPU tensor the
dispatcher raises NotImplementedError, it does not fall back."""
from __future__ import annotations

import torch
from torch import Tensor

from . import ops

BF16 = torch.bfloat16
_lib = torch.library


@_lib.custom_op("cbx::slot_mapping", mutates_args=(), device_types="cuda")
def slot_mapping(block_table: Tensor, req: Tensor, pos: Tensor) -> Tensor:
    """vllm block_table.py:357-377 -> int64 [n]."""
    return ops.slot_mapping(block_table, req, pos)


@slot_mapping.register_fake
def _(block_table, req, pos):
    return pos.new_empty(pos.shape, dtype=torch.int64)


@_lib.custom_op("cbx::rmsnorm", mutates_args=(), device_types="cuda")
def rmsnorm(x: Tensor, w: Tensor, eps: float = 1e-5) -> Tensor:
    return ops.rmsnorm(x, w, None, eps)


@rmsnorm.register_fake
def _(x, w, eps=1e-5):
    return torch.empty_like(x)


@_lib.custom_op("cbx::linear", mutates_args=(), device_types="cuda")
def linear(x: Tensor, W: Tensor, epilogue: int = 0, path: int = 0) -> Tensor:
    """y = x W^T (epilogue 1: SiLU(gate)·up over packed gate/up rows)."""
    return ops.linear(x, W, epilogue=epilogue, path=path)


@linear.register_fake
def _(x, W, epilogue=0, path=0):
    return x.new_empty(x.shape[0], W.shape[0] // 2 if epilogue == 1 else W.shape[0])


@_lib.custom_op("cbx::layer_qkv", mutates_args=("kv_layer",), device_types="cuda")
def layer_qkv(res: Tensor, ln_w: Tensor, wqkv: Tensor, cos_sin: Tensor, pos: Tensor, slot: Tensor, kv_layer: Tensor,
              eps: float = 1e-5) -> Tensor:
    """RMSNorm -> QKV -> RoPE; rotated q in the returned [T,3072] buffer, k / v appended to the KV pages."""
    return ops.layer_qkv(res, ln_w, wqkv, cos_sin, pos, slot, kv_layer, eps)


@layer_qkv.register_fake
def _(res, ln_w, wqkv, cos_sin, pos, slot, kv_layer, eps=1e-5):
    return res.new_empty(res.shape[0], 3072)


@_lib.custom_op("cbx::paged_attn_decode", mutates_args=(), device_types="cuda")
def paged_attn_decode(qkv: Tensor, kv_layer: Tensor, block_table: Tensor, row_seq: Tensor, kv_len: Tensor,
                      max_kv_len: int) -> Tensor:
    """vllm flash_attn.py:796-800 over the paged cache, one query row per entry -> bf16 [T,1024]."""
    return ops.paged_attn_decode(qkv, kv_layer, block_table, row_seq, kv_len, max_kv_len)


@paged_attn_decode.register_fake
def _(qkv, kv_layer, block_table, row_seq, kv_len, max_kv_len):
    return qkv.new_empty(qkv.shape[0], 1024)


@_lib.custom_op("cbx::linear_residual", mutates_args=("res",), device_types="cuda")
def linear_residual(x: Tensor, W: Tensor, res: Tensor) -> None:
    """res += x W^T in place (o_proj / down_proj with the residual epilogue)."""
    ops.linear_residual(x, W, res)


@linear_residual.register_fake
def _(x, W, res):
    return None


@_lib.custom_op("cbx::layer_gate_up", mutates_args=(), device_types="cuda")
def layer_gate_up(res: Tensor, ln_w: Tensor, wgu_packed: Tensor, eps: float = 1e-5) -> Tensor:
    return ops.layer_gate_up(res, ln_w, wgu_packed, eps)


@layer_gate_up.register_fake
def _(res, ln_w, wgu_packed, eps=1e-5):
    return res.new_empty(res.shape[0], 4096)


@_lib.custom_op("cbx::head_logits", mutates_args=(), device_types="cuda")
def head_logits(res: Tensor, norm_w: Tensor, whead: Tensor, eps: float = 1e-5) -> Tensor:
    return ops.head_logits(res, norm_w, whead, eps)


@head_logits.register_fake
def _(res, norm_w, whead, eps=1e-5):
    return res.new_empty(res.shape[0], ops.LOGITS_LD)


@_lib.custom_op("cbx::layer_attn_block", mutates_args=("res", "kv_layer"), device_types="cuda")
def layer_attn_block(res: Tensor, ln_w: Tensor, wqkv: Tensor, wo: Tensor, cos_sin: Tensor, pos: Tensor, slot: Tensor,
                     kv_layer: Tensor, block_table: Tensor, row_seq: Tensor, kv_len: Tensor, max_kv_len: int,
                     eps: float = 1e-5) -> None:
    """The attention half of a layer in one launch (T <= 4 rows): res and the KV pages are updated in place."""
    ops.layer_attn_block(res, ln_w, wqkv, wo, cos_sin, pos, slot, kv_layer, block_table, row_seq, kv_len, max_kv_len, eps)


@layer_attn_block.register_fake
def _(res, ln_w, wqkv, wo, cos_sin, pos, slot, kv_layer, block_table, row_seq, kv_len, max_kv_len, eps=1e-5):
    return None


@_lib.custom_op("cbx::layer_mlp_block", mutates_args=("res",), device_types="cuda")
def layer_mlp_block(res: Tensor, ln_w: Tensor, wgu_packed: Tensor, wd_t: Tensor, eps: float = 1e-5) -> None:
    """cbx_layer_mlp_block: the MLP half of a layer in one launch (T <= 4 rows); wd_t = Wd.t().contiguous()."""
    ops.layer_mlp_block(res, ln_w, wgu_packed, wd_t, eps)


@layer_mlp_block.register_fake
def _(res, ln_w, wgu_packed, wd_t, eps=1e-5):
    return None


OPS = ("slot_mapping", "rmsnorm", "linear", "layer_qkv", "paged_attn_decode", "linear_residual", "layer_gate_up",
       "head_logits", "layer_attn_block", "layer_mlp_block")
