"""ChatterboxMultilingualTTS — the reference's 23-language surface for the T3 path, over the B200 engine.

Mirrors ``src/chatterbox_vllm/mtl_tts.py`` (class :91-128, from_local :116-186, from_pretrained :188-268,
generate :317-372, generate_with_conds :374-520) up to the speech tokens; what differs from the English class
(``tts.ChatterboxTTS``) is exactly what differs in the reference:

  * defaults temperature 0.4 / top_p 0.4, ``language_id`` may be ``None``, one code, or one code PER PROMPT;
  * prompts are wrapped ``[lang][START]…[STOP]`` (no tag when the prompt's language is None) and tokenised by
    ``MTLTokenizer`` (lower-casing, NFKD, per-language normalisers);
  * sampling is clamped to T, top_p <= 0.5 as soon as ANY prompt names a language (mtl_tts.py:442-448);
  * every ``**kwargs`` entry goes to the sampling parameters — ``min_p`` included, which the English class's
    ``generate_with_conds`` swallows (mtl_tts.py:404,508-518 vs tts.py:307);
  * text ids above 1000 enter the repetition-penalty prompt mask: requests carry
    ``penalize_prompt_ids=True``;
  * weights file ``t3_mtl23ls_v2.safetensors``, vocabulary ``grapheme_mtl_merged_expanded_v1.json``.

Waveform synthesis (S3Gen), the voice encoder and the S3 tokenizer stay outside this path, as in ``tts.py`` here:
``vocoder=`` / ``conditioner=`` / ``default_conds`` (conds.pt) are the hand-over points."""
from __future__ import annotations

import os
from pathlib import Path
from typing import Any, Optional, Union

import torch

from .text_utils import SUPPORTED_LANGUAGES, punc_norm
from .tts import REPO_ID, ChatterboxTTS


class ChatterboxMultilingualTTS(ChatterboxTTS):
    T3_WEIGHTS_FILE = "t3_mtl23ls_v2.safetensors"                                   # mtl_tts.py:131
    PENALIZE_PROMPT_IDS = True

    @classmethod
    def get_supported_languages(cls) -> dict[str, str]:                             # mtl_tts.py:111-114 (a classmethod there too)
        return SUPPORTED_LANGUAGES.copy()

    @classmethod
    def from_local(cls, ckpt_dir: Union[str, Path], target_device: str = "cuda", max_model_len: int = 1000,
                   compile: bool = False, max_batch_size: int = 10, s3gen_use_fp16: bool = False,
                   **kwargs) -> "ChatterboxMultilingualTTS":
        """mtl_tts.py:116-186.  ``tokenizer=`` overrides MTLTokenizer.from_pretrained(ckpt_dir)."""
        from safetensors.torch import load_file
        from .tokenizer import MTLTokenizer
        ckpt_dir = Path(ckpt_dir)
        sd = load_file(str(ckpt_dir / cls.T3_WEIGHTS_FILE))
        tok = kwargs.pop("tokenizer", None) or MTLTokenizer.from_pretrained(str(ckpt_dir))
        conds = ckpt_dir / "conds.pt"
        if conds.exists() and "default_conds" not in kwargs:
            kwargs["default_conds"] = torch.load(str(conds), weights_only=True, map_location="cpu")
        return cls.from_state_dict(sd, tok, target_device, max_model_len, max_batch_size, "multilingual", **kwargs)

    @classmethod
    def from_pretrained(cls, repo_id: str = REPO_ID, revision: str = "main", *args, **kwargs) -> "ChatterboxMultilingualTTS":
        """mtl_tts.py:188-268 — snapshot of the multilingual files from the HF Hub (needs network).  The reference also
        writes ./t3-multilingual-model/config.json for vLLM; nothing reads that here."""
        from huggingface_hub import snapshot_download
        ckpt_dir = Path(snapshot_download(repo_id=repo_id, repo_type="model", revision=revision,
                                          allow_patterns=[cls.T3_WEIGHTS_FILE, "grapheme_mtl_merged_expanded_v1.json", "conds.pt",
                                                          "Cangjie5_TC.json"], token=os.getenv("HF_TOKEN")))
        return cls.from_local(ckpt_dir, *args, **kwargs)

    # ------------------------------------------------------------------ generation
    def generate(self, prompts: Union[str, list[str]], language_id: Optional[Union[str, list[str]]] = None,
                 audio_prompt_path: Optional[str] = None, exaggeration: float = 0.5, temperature: float = 0.4,
                 max_tokens=1000, top_p=0.4, repetition_penalty=2.0, *args, **kwargs) -> list[Any]:
        self._check_languages(language_id)                                          # mtl_tts.py:344-357
        s3gen_ref, cond_emb = self.get_audio_conditionals(audio_prompt_path)
        return self.generate_with_conds(prompts=prompts, language_id=language_id, s3gen_ref=s3gen_ref, cond_emb=cond_emb,
                                        temperature=temperature, exaggeration=exaggeration, max_tokens=max_tokens,
                                        top_p=top_p, repetition_penalty=repetition_penalty, *args, **kwargs)

    def generate_with_conds(self, prompts: Union[str, list[str]], s3gen_ref: dict[str, Any], cond_emb: torch.Tensor,
                            language_id: Optional[Union[str, list[str]]] = None, temperature: float = 0.4,
                            exaggeration: float = 0.5, max_tokens=1000, diffusion_steps: int = 10, top_p=0.4,
                            repetition_penalty=2.0, *args, **kwargs) -> list[Any]:
        toks = self.generate_tokens(prompts, cond_emb, language_id=language_id, temperature=temperature,
                                    exaggeration=exaggeration, max_tokens=max_tokens, top_p=top_p,
                                    repetition_penalty=repetition_penalty, **kwargs)      # min_p travels with the kwargs
        if self.vocoder is None:
            return toks
        return [self.vocoder(t.to(self.target_device), s3gen_ref, n_timesteps=diffusion_steps) for t in toks]

    # hooks of ChatterboxTTS.generate_tokens ------------------------------------------------------------------
    def _language_list(self, language_id, n: int) -> list:
        if language_id is None:
            return [None] * n
        if isinstance(language_id, str):
            return [language_id] * n
        if len(language_id) != n:                                                   # mtl_tts.py:397-398
            raise ValueError(f"Number of language_ids ({len(language_id)}) must match number of prompts ({n})")
        return list(language_id)

    def _check_languages(self, language_id) -> None:
        for lang in ([language_id] if isinstance(language_id, str) else (language_id or [])):
            if lang and lang.lower() not in SUPPORTED_LANGUAGES:
                raise ValueError(f"Unsupported language_id '{lang}'. Supported languages: " + ", ".join(SUPPORTED_LANGUAGES.keys()))

    def wrap_prompts(self, prompts: list[str], language_id) -> list[str]:           # mtl_tts.py:404-412
        return [f"[{lang.lower()}][START]{punc_norm(p)}[STOP]" if lang else f"[START]{punc_norm(p)}[STOP]"
                for p, lang in zip(prompts, self._language_list(language_id, len(prompts)))]

    def _clamp_sampling(self, language_id) -> bool:                                 # mtl_tts.py:442-448: ANY prompt with a language
        return any(language_id) if isinstance(language_id, (list, tuple)) else bool(language_id)
