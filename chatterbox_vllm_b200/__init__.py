"""chatterbox_vllm_b200 — B200-native (sm_100a) Chatterbox T3 speech-token decode path behind
the surface of iMash-io/chatterbox-vllm.  The compute path lives in libcbx_t3.so
(chatterbox_vllm_b200/csrc, C-ABI in include/cbx_t3.h); there is no CPU fallback."""
__version__ = "0.1.0"
