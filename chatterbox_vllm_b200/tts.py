"""ChatterboxTTS — the reference's public surface for the T3 path, over the B200 engine.

Mirrors ``src/chatterbox_vllm/tts.py`` of iMash-io/chatterbox-vllm for everything up to and
including speech-token generation (tts.py:64-419) plus the token post-processing that defines the
T3→S3Gen wire format (tts.py:448-465):

    ChatterboxTTS.from_pretrained / from_local / from_state_dict
    .generate(prompts, audio_prompt_path, language_id, exaggeration, temperature, max_tokens, top_p,
              repetition_penalty, **sampling_kwargs)
    .generate_with_conds(prompts, s3gen_ref, cond_emb, ...)
    .generate_tokens(...)            (added: the measured endpoint — S3Gen is out of scope)
    .get_audio_conditionals / .update_exaggeration / .get_supported_languages / .sr / .shutdown

Differences by design:
  * ``self.t3`` is a ``T3Engine`` (libcbx_t3.so kernels), not a ``vllm.LLM``;
  * waveform synthesis (S3Gen, tts.py:425-557) is not part of this path: ``generate`` hands each
    prompt's speech tokens to an injectable ``vocoder(speech_tokens, s3gen_ref, n_timesteps=...)``
    and returns its outputs, or — when no vocoder was given — returns the token tensors;
  * conditioning from a reference WAV needs the reference's audio front-ends (VoiceEncoder,
    S3Tokenizer — out of scope): pass a precomputed ``cond_emb`` or a ``conditioner`` callable."""
from __future__ import annotations

import math
import os
import re
import time
from pathlib import Path
from typing import Any, Callable, Optional, Union

import torch

from .engine import (COND_LEN, SPEECH_TOKEN_OFFSET, START_SPEECH, STOP_SPEECH, Request, SamplingParams, T3Engine)
from .text_utils import SUPPORTED_LANGUAGES, punc_norm

S3GEN_SR = 24000                # models/s3gen/const.py (output sample rate the reference reports via .sr)
REPO_ID = "ResembleAI/chatterbox"
_TRUE = ("1", "true", "yes", "on")


def _env_flag(name: str, default: str = "0") -> bool:
    return os.environ.get(name, default).lower() in _TRUE


def drop_invalid_tokens(x: torch.Tensor) -> torch.Tensor:
    """models/s3tokenizer/__init__.py:38-78 — keep (first SOS, first EOS) exclusive when present."""
    assert x.dim() == 1
    sos = (x == START_SPEECH).nonzero()
    eos = (x == STOP_SPEECH).nonzero()
    s = int(sos[0]) + 1 if sos.numel() else 0
    e = int(eos[0]) if eos.numel() else x.numel()
    return x[s:e]


def postprocess_tokens(tok_ids: list[int], tail_crop_k: int = 2) -> torch.Tensor:
    """tts.py:448-465: cut at the first stop id (7562), crop ``tail_crop_k`` tokens before it, remove
    the +1000 offset, drop SOS/EOS framing and keep ids < 6561.  Returns int64 [n]."""
    stop = STOP_SPEECH + SPEECH_TOKEN_OFFSET
    if stop in tok_ids:
        tok_ids = tok_ids[:max(0, tok_ids.index(stop) - max(0, int(tail_crop_k)))]
    x = torch.tensor([t - SPEECH_TOKEN_OFFSET for t in tok_ids], dtype=torch.long)
    x = drop_invalid_tokens(x)
    return x[x < START_SPEECH]


def guard_cap_tokens(prompt: str) -> int:
    """tts.py:466-486: the per-prompt overrun guard applied AFTER decoding — ceil(1.6 · clamp(ceil(2.2 ·
    non-space chars of the prompt without [tags]), 64, 1200)) speech tokens at most."""
    try:
        clean = re.sub(r"\[[^\]]+\]", "", prompt)
        n_chars = sum(1 for ch in clean if not ch.isspace())
        tpc = float(os.environ.get("CHATTERBOX_TOKENS_PER_CHAR", "2.2"))
        tmin = int(os.environ.get("CHATTERBOX_TOKENS_MIN", "64"))
        tmax = int(os.environ.get("CHATTERBOX_TOKENS_MAX", "1200"))
        guard = float(os.environ.get("CHATTERBOX_TOKENS_GUARD_MULT", "1.6"))
        est = max(tmin, min(tmax, int(math.ceil(n_chars * tpc))))
        return int(math.ceil(est * guard))
    except Exception:
        return 1 << 30


def pre_cap_tokens(prompts: list[str], max_tokens: int, max_model_len: int) -> int:
    """tts.py:372-399: per-prompt estimate ceil(2.2 · non-space chars) clamped to [64, 1200], guard
    ×1.6; the batch cap is min over prompts + 16, then min with max_tokens and max_model_len.
    Env knobs as in the reference."""
    try:
        tpc = float(os.environ.get("CHATTERBOX_TOKENS_PER_CHAR", "2.2"))
        tmin = int(os.environ.get("CHATTERBOX_TOKENS_MIN", "64"))
        tmax = int(os.environ.get("CHATTERBOX_TOKENS_MAX", "1200"))
        guard = float(os.environ.get("CHATTERBOX_TOKENS_GUARD_MULT", "1.6"))
        margin = int(os.environ.get("CHATTERBOX_PRE_GUARD_MARGIN", "16"))
        caps = []
        for p in prompts:
            clean = re.sub(r"\[[^\]]+\]", "", p)
            n_chars = sum(1 for ch in clean if not ch.isspace())
            est = max(tmin, min(tmax, int(math.ceil(n_chars * tpc))))
            caps.append(int(math.ceil(est * guard)))
        cap = max(1, min(caps) + max(0, margin)) if caps else min(max_tokens, max_model_len)
        return min(cap, min(max_tokens, max_model_len))
    except Exception:
        return min(max_tokens, max_model_len)


class ChatterboxTTS:
    def __init__(self, target_device: str, max_model_len: int, t3: T3Engine, tokenizer,
                 emotion_adv_fc: torch.Tensor | None = None, default_cond_emb: torch.Tensor | None = None,
                 default_s3gen_ref: dict | None = None, vocoder: Callable | None = None,
                 conditioner: Callable | None = None, variant: str = "english", default_conds: dict | None = None):
        self.target_device, self.max_model_len = target_device, max_model_len
        self.t3, self.tokenizer, self.variant = t3, tokenizer, variant
        self.emotion_adv_fc = emotion_adv_fc          # cond_enc.emotion_adv_fc.weight [1024, 1] fp32
        self.cond_encoder = None                      # T3CondEncoder when the checkpoint carries cond_enc.*
        self.default_cond_emb, self.default_s3gen_ref = default_cond_emb, default_s3gen_ref or {}
        self.vocoder, self.conditioner = vocoder, conditioner
        # conds.pt of the checkpoint (tts.py:52-62 `Conditionals.load`): {"t3": {speaker_emb [1,256], cond_prompt_speech_tokens
        # [1,150], ...}, "gen": S3Gen reference dict} — the built-in voice used when no audio prompt is given
        self.default_conds = default_conds
        self._cond_cache: dict = {}
        self.last_t3_seconds = 0.0

    @property
    def sr(self) -> int:
        return S3GEN_SR

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_state_dict(cls, state_dict: dict, tokenizer, target_device: str = "cuda", max_model_len: int = 1000,
                        max_batch_size: int = 10, variant: str = "english", **kwargs) -> "ChatterboxTTS":
        """Build from an in-memory T3 checkpoint (key layout of t3_cfg.safetensors) — used with
        random-init weights where there is no network."""
        extra = {k: kwargs.pop(k) for k in ("default_cond_emb", "default_s3gen_ref", "vocoder", "conditioner", "default_conds")
                 if k in kwargs}
        max_pairs = kwargs.pop("max_pairs", None) or kwargs.pop("max_num_seqs", None) or max_batch_size
        eng = T3Engine(state_dict, device=target_device if target_device != "cuda" else "cuda:0",
                       max_pairs=max_pairs, max_model_len=max_model_len, **kwargs)
        fc = state_dict.get("cond_enc.emotion_adv_fc.weight")
        tts = cls(target_device, max_model_len, eng, tokenizer,
                  emotion_adv_fc=fc.float().cpu() if fc is not None else None, variant=variant, **extra)
        if "cond_enc.perceiver.attn.to_q.weight" in state_dict:   # conditioning encoder on the device (cond_enc.py, perceiver.py)
            from .cond_enc import T3CondEncoder
            tts.cond_encoder = T3CondEncoder({k: v for k, v in state_dict.items() if k.startswith("cond_enc.")},
                                             state_dict["speech_emb.weight"], state_dict["speech_pos_emb.emb.weight"],
                                             device=str(eng.device))
        return tts

    T3_WEIGHTS = {"english": "t3_cfg.safetensors", "multilingual": "t3_23lang.safetensors"}     # tts.py:104

    @classmethod
    def from_local(cls, ckpt_dir: Union[str, Path], target_device: str = "cuda", max_model_len: int = 1000,
                   compile: bool = False, max_batch_size: int = 10, variant: str = "english",
                   s3gen_use_fp16: bool = False, **kwargs) -> "ChatterboxTTS":
        """tts.py:91-158.  ``compile`` / ``s3gen_use_fp16`` are accepted for signature compatibility
        (the engine always replays CUDA graphs; S3Gen is not part of this path).  ``tokenizer=`` overrides
        the text tokenizer (default: EnTokenizer / MTLTokenizer resolved from ckpt_dir, as the reference's
        vLLM tokenizer registry would).  conds.pt, when present, becomes the built-in voice (tts.py:151)."""
        from safetensors.torch import load_file
        from .tokenizer import EnTokenizer, MTLTokenizer
        ckpt_dir = Path(ckpt_dir)
        if variant not in cls.T3_WEIGHTS:
            raise ValueError(f"variant {variant!r}: expected 'english' or 'multilingual'")
        sd = load_file(str(ckpt_dir / cls.T3_WEIGHTS[variant]))
        tok = kwargs.pop("tokenizer", None)
        if tok is None:
            tok = (EnTokenizer if variant == "english" else MTLTokenizer).from_pretrained(str(ckpt_dir))
        conds = ckpt_dir / "conds.pt"
        if conds.exists() and "default_conds" not in kwargs:
            kwargs["default_conds"] = torch.load(str(conds), weights_only=True, map_location="cpu")
        return cls.from_state_dict(sd, tok, target_device, max_model_len, max_batch_size, variant, **kwargs)

    @classmethod
    def from_pretrained(cls, repo_id: str = REPO_ID, revision: str = "1b475dffa71fb191cb6d5901215eb6f55635a9b6",
                        *args, **kwargs) -> "ChatterboxTTS":
        """tts.py:160-190 — downloads the checkpoint files from the HF Hub (needs network)."""
        from huggingface_hub import hf_hub_download
        local_path = None
        for fpath in ("t3_cfg.safetensors", "tokenizer.json", "conds.pt"):
            local_path = hf_hub_download(repo_id=repo_id, filename=fpath, revision=revision)
        return cls.from_local(Path(local_path).parent, *args, variant="english", **kwargs)

    @classmethod
    def from_pretrained_multilingual(cls, repo_id: str = REPO_ID, revision: str = "c819eeccdf99310da26bca3bc5ace120db93471a",
                                     *args, **kwargs) -> "ChatterboxTTS":
        """tts.py:192-206 — the 23-language checkpoint through the same class (variant="multilingual"); needs network.
        (The reference also symlinks the weights into ./t3-model-multilingual for vLLM; nothing reads that here.)"""
        from huggingface_hub import hf_hub_download
        local_path = None
        for fpath in ("t3_23lang.safetensors", "mtl_tokenizer.json", "conds.pt", "Cangjie5_TC.json"):
            local_path = hf_hub_download(repo_id=repo_id, filename=fpath, revision=revision)
        return cls.from_local(Path(local_path).parent, *args, variant="multilingual", **kwargs)

    def get_supported_languages(self) -> dict[str, str]:
        return SUPPORTED_LANGUAGES.copy() if self.variant == "multilingual" else {"en": "English"}

    # ------------------------------------------------------------------ conditioning
    def get_audio_conditionals(self, wav_fpath: Optional[str] = None) -> tuple[dict[str, Any], torch.Tensor]:
        """tts.py:215-247 → (s3gen_ref_dict, cond_emb [34,1024] fp32).  Cached per path like the
        reference's lru_cache(10)."""
        if wav_fpath in self._cond_cache:
            return self._cond_cache[wav_fpath]
        if wav_fpath is None and self.default_cond_emb is None and self.default_conds is not None:
            # tts.py:217-220,238-247: the built-in voice = conds.pt's speaker embedding + prompt speech tokens through
            # T3CondEnc (on the device); its "gen" dict is the vocoder's reference
            t3c = self.default_conds["t3"]
            ref, emb = self.conditionals_from_features(t3c["speaker_emb"].reshape(-1), t3c["cond_prompt_speech_tokens"].reshape(-1),
                                                       s3gen_ref=self.default_conds.get("gen", {}))
            self.default_s3gen_ref, self.default_cond_emb = ref, emb
        if wav_fpath is None and self.default_cond_emb is not None:
            out = (self.default_s3gen_ref, self.default_cond_emb)
        elif self.conditioner is not None:
            out = self.conditioner(wav_fpath)
            if len(out) == 3:                              # (s3gen_ref, speaker_emb[256], prompt speech tokens): encode here
                out = self.conditionals_from_features(out[1], out[2], s3gen_ref=out[0])
        else:
            raise NotImplementedError(
                "conditioning from audio needs the reference's VoiceEncoder/S3Tokenizer front-ends, which are "
                "outside the T3 path; pass default_cond_emb= or conditioner= at construction, or call "
                "generate_with_conds(cond_emb=...)")
        if len(self._cond_cache) >= 10:
            self._cond_cache.pop(next(iter(self._cond_cache)))
        self._cond_cache[wav_fpath] = out
        return out

    def conditionals_from_features(self, speaker_emb: torch.Tensor, prompt_tokens: torch.Tensor,
                                   s3gen_ref: Optional[dict] = None) -> tuple[Optional[dict], torch.Tensor]:
        """The T3 half of get_audio_conditionals (tts.py:238-247) on the device: voice-encoder embedding
        [256] + S3 prompt tokens [<= 150] -> cond_emb [34,1024] fp32, left in GPU memory (the reference
        moves it to the CPU so that it can be pickled to the vLLM process)."""
        if getattr(self, "cond_encoder", None) is None:
            raise RuntimeError("checkpoint has no cond_enc.* weights; cannot encode conditionals")
        return s3gen_ref, self.cond_encoder(speaker_emb, prompt_tokens, 0.5)

    def update_exaggeration(self, cond_emb: torch.Tensor, exaggeration: float) -> torch.Tensor:
        """tts.py:249-257: row 33 ← emotion_adv_fc(exaggeration) (Linear(1→1024), no bias)."""
        if exaggeration == 0.5:
            return cond_emb
        if getattr(self, "cond_encoder", None) is not None and cond_emb.is_cuda:
            new = cond_emb.clone()
            new[-1] = self.cond_encoder.emotion_row(exaggeration).to(new.dtype)
            return new
        if self.emotion_adv_fc is None:
            raise RuntimeError("checkpoint has no cond_enc.emotion_adv_fc.weight; cannot change exaggeration")
        new = cond_emb.clone()
        new[-1] = (self.emotion_adv_fc[:, 0] * float(exaggeration)).to(new.dtype)
        return new

    # ------------------------------------------------------------------ generation
    def generate(self, prompts: Union[str, list[str]], audio_prompt_path: Optional[str] = None,
                 language_id: Optional[str] = "en", exaggeration: float = 0.5, temperature: float = 0.8,
                 max_tokens=1000, top_p=0.8, repetition_penalty=2.0, *args, **kwargs) -> list[Any]:
        s3gen_ref, cond_emb = self.get_audio_conditionals(audio_prompt_path)
        return self.generate_with_conds(prompts=prompts, s3gen_ref=s3gen_ref, cond_emb=cond_emb,
                                        temperature=temperature, language_id=language_id,
                                        exaggeration=exaggeration, max_tokens=max_tokens, top_p=top_p,
                                        repetition_penalty=repetition_penalty, *args, **kwargs)

    def generate_with_conds(self, prompts: Union[str, list[str]], s3gen_ref: dict[str, Any],
                            cond_emb: torch.Tensor, language_id: Optional[str] = "en", temperature: float = 0.8,
                            exaggeration: float = 0.5, max_tokens=1000, diffusion_steps: int = 10, top_p=1.0,
                            min_p=0.05, repetition_penalty=2.0, *args, **kwargs) -> list[Any]:
        # NB the reference swallows min_p here and never forwards it (tts.py:307,408-418); kept.
        toks = self.generate_tokens(prompts, cond_emb, language_id=language_id, temperature=temperature,
                                    exaggeration=exaggeration, max_tokens=max_tokens, top_p=top_p,
                                    repetition_penalty=repetition_penalty, **kwargs)
        if self.vocoder is None:
            return toks
        return [self.vocoder(t.to(self.target_device), s3gen_ref, n_timesteps=diffusion_steps) for t in toks]

    def wrap_prompts(self, prompts: list[str], language_id: Optional[str]) -> list[str]:
        out = ["[START]" + punc_norm(p) + "[STOP]" for p in prompts]                # tts.py:327
        if self.variant == "multilingual":
            out = [f"<{language_id.lower()}>{p}" for p in out]                      # tts.py:330-333
        return out

    def _check_languages(self, language_id) -> None:                                # tts.py:316-322
        if language_id and language_id.lower() not in self.get_supported_languages():
            raise ValueError(f"Unsupported language_id '{language_id}'. Supported languages: "
                             + ", ".join(self.get_supported_languages().keys()))

    def _clamp_sampling(self, language_id) -> bool:                                 # tts.py:350-355
        return bool(language_id)

    PENALIZE_PROMPT_IDS = None          # None: by variant (the multilingual text ids reach vLLM's prompt-penalty mask)

    def generate_tokens(self, prompts: Union[str, list[str]], cond_emb: torch.Tensor,
                        language_id: Optional[str] = "en", temperature: float = 0.8, exaggeration: float = 0.5,
                        max_tokens=1000, top_p=1.0, repetition_penalty=2.0, raw: bool = False,
                        **kwargs) -> list[torch.Tensor]:
        """The T3 half of generate_with_conds (tts.py:313-419 + :448-465): returns, per prompt, the
        S3Gen-ready int64 speech tokens (or, with raw=True, the engine's id lists in vLLM id space,
        stop token included, like ``output.token_ids``)."""
        if isinstance(prompts, str):
            prompts = [prompts]
        self._check_languages(language_id)
        cond_emb = self.update_exaggeration(cond_emb, exaggeration)
        wrapped = self.wrap_prompts(prompts, language_id)
        # sampling clamps (tts.py:350-362)
        temp_use, top_p_use = temperature, top_p
        if self._clamp_sampling(language_id):
            temp_use, top_p_use = min(temperature, 0.5), min(top_p, 0.5)
        if _env_flag("CHATTERBOX_DETERMINISTIC"):
            temp_use, top_p_use = 0.0, 1.0
        cap = pre_cap_tokens(wrapped, max_tokens, self.max_model_len)
        for k in ("skip_special_tokens", "spaces_between_special_tokens", "detokenize"):   # text-only knobs
            kwargs.pop(k, None)
        sp_kw = {k: kwargs.pop(k) for k in ("min_p", "top_k", "seed", "ignore_eos") if k in kwargs}
        if kwargs:
            raise TypeError(f"unsupported SamplingParams fields for the T3 path: {sorted(kwargs)}")
        if sp_kw.get("top_k", 0) in (-1, None):
            sp_kw["top_k"] = 0
        reqs = []
        for w in wrapped:
            ids = self.tokenizer.encode(w)
            budget = self.max_model_len - (COND_LEN + len(ids) + 1)
            if budget < 1:
                raise ValueError(f"prompt of {len(ids)} text tokens does not fit max_model_len={self.max_model_len}")
            sp = SamplingParams(temperature=temp_use, top_p=top_p_use, repetition_penalty=repetition_penalty,
                                max_tokens=min(cap, budget), **sp_kw)
            pen = self.variant == "multilingual" if self.PENALIZE_PROMPT_IDS is None else self.PENALIZE_PROMPT_IDS
            reqs.append(Request(ids, cond_emb, sp, penalize_prompt_ids=pen))
        crop = int(os.environ.get("CHATTERBOX_TAIL_CROP_TOKENS", "2"))
        t0 = time.time()
        with torch.inference_mode():
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                # one process per GPU (torchrun): every rank holds the same prompts, runs its share on its own engine,
                # and all ranks get every stream back (sharding.py); the vocoder hand-off then runs on the host lists
                from .sharding import generate_sharded
                outs = generate_sharded(self.t3, reqs, all_ranks=True)
                if not raw:
                    outs = [postprocess_tokens(o, crop)[:guard_cap_tokens(w)] for o, w in zip(outs, wrapped)]
                self.last_t3_seconds = time.time() - t0
                print(f"[T3] Speech Token Generation time: {self.last_t3_seconds:.2f}s")
                return outs
            if raw or not hasattr(self.t3, "generate_device"):
                outs = self.t3.generate(reqs)
            else:
                # stop cut / tail crop / framing / filter / overrun guard run on the device; only the
                # counts come back (tts.py:448-486)
                toks, lens = self.t3.generate_device(reqs, tail_crop=crop, guard_caps=[guard_cap_tokens(w) for w in wrapped])
                outs = [toks[i, :n] for i, n in enumerate(lens.tolist())]
        self.last_t3_seconds = time.time() - t0
        print(f"[T3] Speech Token Generation time: {self.last_t3_seconds:.2f}s")
        if raw or hasattr(self.t3, "generate_device"):
            return outs
        return [postprocess_tokens(o, crop)[:guard_cap_tokens(w)] for o, w in zip(outs, wrapped)]

    def shutdown(self):
        if self.t3 is not None:
            self.t3.close()
            self.t3 = None
