"""T3CondEncoder — the conditioning-prefix producer on the GPU.

Replaces ``T3CondEnc`` + ``Perceiver`` + the prompt-embedding sum the reference evaluates with torch
modules and then moves to the CPU for pickling (tts.py:104-109,238-247; models/t3/modules/cond_enc.py,
perceiver.py).  Inputs are what the audio front-end (voice encoder, S3 tokenizer — out of scope)
yields: a 256-d speaker embedding and up to 150 prompt speech-token ids.  The [34, 1024] fp32 result
stays in device memory; ``emotion_row`` is the cheap path behind ``update_exaggeration``."""
from __future__ import annotations

import ctypes as C

import torch

from ._lib import check, load

F32 = torch.float32
_KEYS = (("spkr_w", "spkr_enc.weight"), ("spkr_b", "spkr_enc.bias"), ("emo_w", "emotion_adv_fc.weight"),
         ("query", "perceiver.pre_attention_query"), ("ln_w", "perceiver.attn.norm.weight"),
         ("ln_b", "perceiver.attn.norm.bias"), ("wq", "perceiver.attn.to_q.weight"), ("bq", "perceiver.attn.to_q.bias"),
         ("wk", "perceiver.attn.to_k.weight"), ("bk", "perceiver.attn.to_k.bias"), ("wv", "perceiver.attn.to_v.weight"),
         ("bv", "perceiver.attn.to_v.bias"), ("wo", "perceiver.attn.proj_out.weight"), ("bo", "perceiver.attn.proj_out.bias"))


class CondWeights(C.Structure):
    _fields_ = [(f, C.c_void_p) for f, _ in _KEYS] + [("speech_emb", C.c_void_p), ("speech_pos", C.c_void_p)]


class T3CondEncoder:
    def __init__(self, cond_enc_state: dict, speech_emb: torch.Tensor, speech_pos: torch.Tensor, device="cuda:0",
                 max_prompt: int = 150):
        """cond_enc_state: the checkpoint's ``cond_enc.*`` tensors (prefix optional); speech_emb /
        speech_pos: ``speech_emb.weight`` [8194,1024] and ``speech_pos_emb.emb.weight``, any dtype."""
        load()
        if not torch.cuda.is_available():
            raise RuntimeError("T3CondEncoder needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        sd = {k[len("cond_enc."):] if k.startswith("cond_enc.") else k: v for k, v in cond_enc_state.items()}
        self._t = {f: sd[k].to(device=self.device, dtype=F32).contiguous() for f, k in _KEYS}
        self._t["speech_emb"] = speech_emb.to(device=self.device, dtype=F32).contiguous()
        self._t["speech_pos"] = speech_pos.to(device=self.device, dtype=F32).contiguous()
        assert self._t["wq"].shape == (1024, 1024) and self._t["spkr_w"].shape == (1024, 256)
        self._w = CondWeights(**{f: t.data_ptr() for f, t in self._t.items()})
        self.max_prompt = min(max_prompt, self._t["speech_pos"].shape[0])
        self._ws = torch.empty(load().cbx_cond_workspace_bytes(self.max_prompt), dtype=torch.uint8, device=self.device)

    def __call__(self, speaker_emb: torch.Tensor, prompt_tokens: torch.Tensor, emotion_adv: float = 0.5) -> torch.Tensor:
        """fp32 CUDA [34, 1024] = speaker row ‖ 32 perceiver rows ‖ emotion row (cond_enc.py:117-123)."""
        spk = speaker_emb.to(device=self.device, dtype=F32).reshape(-1).contiguous()
        tok = prompt_tokens.to(device=self.device, dtype=torch.int32).reshape(-1).contiguous()
        if spk.numel() != 256:
            raise ValueError("speaker_emb must hold 256 values")
        if not 1 <= tok.numel() <= self.max_prompt:
            raise ValueError(f"prompt of {tok.numel()} speech tokens not in [1, {self.max_prompt}]")
        out = torch.empty(34, 1024, dtype=F32, device=self.device)
        check(load().cbx_cond_encode(C.byref(self._w), spk.data_ptr(), tok.data_ptr(), tok.numel(), float(emotion_adv),
                                     self._ws.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream),
              "cond_encode")
        return out

    def emotion_row(self, exaggeration: float) -> torch.Tensor:
        """fp32 CUDA [1024] = emotion_adv_fc(exaggeration) (tts.py:249-257)."""
        row = torch.empty(1024, dtype=F32, device=self.device)
        check(load().cbx_cond_emotion_row(C.byref(self._w), float(exaggeration), row.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "cond_emotion_row")
        return row
