"""ctypes binding of libcbx_t3.so (include/cbx_t3.h).  There is NO fallback: if the CUDA
library is missing or fails to load, importing the compute path raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcbx_t3.so")

HIDDEN, HEADS, HEAD_DIM, INTER = 1024, 16, 64, 4096
SPEECH_VOCAB, BLOCK_SIZE, LOGITS_LD, SEEN_WORDS, PAGE_BYTES = 8194, 16, 8256, 257, 65536

vp, i32p, i64p, u32p, u64p, f32p = (C.c_void_p,) * 6   # device pointers travel as integers


class SamplerState(C.Structure):
    _fields_ = [("last_tok", vp), ("out_tokens", vp), ("n_out", vp), ("max_out", vp), ("finished", vp),
                ("seq_len", vp), ("seen", vp), ("temperature", vp), ("top_p", vp), ("min_p", vp),
                ("rep_pen", vp), ("top_k", vp), ("ignore_eos", vp), ("seed", vp), ("uid", vp),
                ("n_finished", vp), ("out_stride", C.c_int), ("stop_token", C.c_int), ("cfg_scale", C.c_float)]


class Weights(C.Structure):
    _fields_ = [("n_layers", C.c_int), ("wqkv", C.POINTER(vp)), ("wo", C.POINTER(vp)), ("wgu", C.POINTER(vp)),
                ("wd", C.POINTER(vp)), ("ln1", C.POINTER(vp)), ("ln2", C.POINTER(vp)), ("norm", vp),
                ("speech_emb", vp), ("speech_head", vp), ("cos_sin", vp), ("rms_eps", C.c_float)]


class Fp8Weights(C.Structure):
    _fields_ = [("n_layers", C.c_int), ("wqkv", C.POINTER(vp)), ("wo", C.POINTER(vp)), ("wgu", C.POINTER(vp)),
                ("wd", C.POINTER(vp)), ("sqkv", C.POINTER(vp)), ("so", C.POINTER(vp)), ("sgu", C.POINTER(vp)),
                ("sd", C.POINTER(vp)), ("speech_head", vp), ("shead", vp)]


class EngineCfg(C.Structure):
    _fields_ = [("max_pairs", C.c_int), ("max_rows", C.c_int), ("max_model_len", C.c_int),
                ("num_blocks", C.c_int), ("out_stride", C.c_int)]


_PROTOS = {
    "cbx_last_error": (C.c_char_p, []),
    "cbx_abi_version": (C.c_int, []),
    "cbx_has_experiments": (C.c_int, []),
    "cbx_slot_mapping": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, vp]),
    "cbx_pool_create": (vp, [C.c_int]),
    "cbx_pool_destroy": (None, [vp]),
    "cbx_pool_num_free": (C.c_int, [vp]),
    "cbx_pool_alloc": (C.c_int, [vp, C.c_int, C.POINTER(C.c_int32)]),
    "cbx_pool_free": (C.c_int, [vp, C.POINTER(C.c_int32), C.c_int]),
    "cbx_rmsnorm": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_rmsnorm_reduce": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_linear_partials": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "cbx_linear": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "cbx_pack_gate_up": (C.c_int, [vp, vp, C.c_int, C.c_int, vp]),
    "cbx_quantize_fp8_rows": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_linear_fp8w": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "cbx_layer_qkv_fp8w": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_linear_residual_fp8w": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_linear_partials_fp8w": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "cbx_layer_gate_up_fp8w": (C.c_int, [vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_head_logits_fp8w": (C.c_int, [vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_engine_set_fp8_weights": (C.c_int, [vp, C.POINTER(Fp8Weights)]),
    "cbx_rope_kv_append": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, vp]),
    "cbx_attn_scratch_bytes": (C.c_int64, [C.c_int, C.c_int]),
    "cbx_attn_num_splits": (C.c_int, [C.c_int, C.c_int]),
    "cbx_paged_attn_decode": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_paged_attn_decode_static": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_attn_plan": (C.c_int, [vp, C.c_int, vp, vp]),
    "cbx_attn_plan_paged": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, vp]),
    "cbx_paged_attn_decode_planned": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_paged_attn_prefill": (C.c_int, [vp, vp, vp, C.c_int, vp, C.c_int, vp, vp]),
    "cbx_engine_set_prefill_items": (C.c_int, [vp, vp, C.c_int]),
    "cbx_layer_qkv": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_linear_residual": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, vp]),
    "cbx_layer_mlp_block": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_layer_mlp_block_scratch_bytes": (C.c_int64, []),
    "cbx_engine_set_down_t": (C.c_int, [vp, vp, C.c_int]),
    "cbx_layer_attn_block_scratch_bytes": (C.c_int64, []),
    "cbx_layer_attn_block_supported": (C.c_int, [C.c_int, C.c_int]),
    "cbx_layer_attn_block": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp]),
    "cbx_layer_gate_up": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_head_logits": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_float, vp]),
    "cbx_cfg_sample": (C.c_int, [vp, C.POINTER(SamplerState), C.c_int, vp]),
    "cbx_debug_sampler_radix_pivot": (C.c_int, [C.c_int]),
    "cbx_decode_prepare": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp]),
    "cbx_prefill_embed": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_int, vp]),
    "cbx_prefill_rows": (C.c_int, [vp, C.c_int, vp, vp, C.c_int, vp, vp, vp, vp, vp]),
    "cbx_cond_workspace_bytes": (C.c_int64, [C.c_int]),
    "cbx_cond_encode": (C.c_int, [vp, vp, vp, C.c_int, C.c_float, vp, vp, vp]),
    "cbx_cond_emotion_row": (C.c_int, [vp, C.c_float, vp, vp]),
    "cbx_postprocess_tokens": (C.c_int, [vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, vp, vp, vp]),
    "cbx_engine_workspace_bytes": (C.c_int64, [C.POINTER(EngineCfg)]),
    "cbx_engine_create": (vp, [C.POINTER(EngineCfg), C.POINTER(Weights), vp, vp, vp, C.POINTER(SamplerState)]),
    "cbx_engine_destroy": (None, [vp]),
    "cbx_engine_buffer": (vp, [vp, C.c_char_p]),
    "cbx_engine_set_trace": (C.c_int, [vp, vp]),
    "cbx_engine_forward": (C.c_int, [vp, C.c_int, C.c_int, vp]),
    "cbx_engine_sample": (C.c_int, [vp, C.c_int, vp]),
    "cbx_engine_sample_subset": (C.c_int, [vp, C.c_int, C.c_int, vp, vp]),
    "cbx_engine_decode": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "cbx_engine_last_launch_count": (C.c_int64, [vp]),
}
EXPORTS = tuple(_PROTOS)

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
                "Build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`make -C chatterbox_vllm_b200/csrc`.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)            # AttributeError here = ABI mismatch: fail loudly
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise RuntimeError(f"libcbx_t3 {what} failed ({rc}): {load().cbx_last_error().decode()}")
