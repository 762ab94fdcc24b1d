/* cbx_t3.h — C-ABI of libcbx_t3.so: the B200 (sm_100a) implementation of the
 * Chatterbox T3 speech-token decode path.
 *
 * The reference (iMash-io/chatterbox-vllm) is pure Python and has no FFI; the
 * arithmetic of this path lives in vLLM / torch kernels it calls.  Each entry
 * point below therefore cites the reference CALL SITE (relative to the reference
 * repo root) or the vLLM-0.22 source ("vllm:" prefix) whose behaviour it replaces.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (cbx_last_error() gives a
 *     thread-local message); no C++ exception crosses this boundary;
 *   - all tensor pointers are DEVICE pointers owned by the caller (bf16 unless
 *     noted, row-major, innermost stride 1); nothing here allocates device memory
 *     except cbx_engine_* scratch that the caller sizes via cbx_engine_workspace_bytes;
 *   - every launch is asynchronous on the cudaStream_t passed as `void* stream`;
 *   - one engine handle per device; not re-entrant (single submitting thread, as
 *     vLLM's GPU worker is).
 */
#ifndef CBX_T3_H_
#define CBX_T3_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Fixed model geometry (t3-model/config.json; models/t3/modules/t3_config.py:1-21). */
#define CBX_HIDDEN      1024
#define CBX_HEADS       16
#define CBX_HEAD_DIM    64
#define CBX_INTER       4096
#define CBX_SPEECH_VOCAB 8194
#define CBX_BLOCK_SIZE  16          /* vllm: KV block size the reference runs with   */
#define CBX_LOGITS_LD   8256        /* padded row stride of the bf16 logits buffer   */
#define CBX_SEEN_WORDS  257         /* ceil(8194 / 32) words of seen-token bitmap    */
/* One KV page = one (layer, block): K[16 heads][16 tok][64] (16-byte chunks XOR-
 * swizzled by tok&7) followed by V[16 heads][16 tok][64]; 65536 bytes. */
#define CBX_PAGE_BYTES  65536

const char* cbx_last_error(void);
int cbx_abi_version(void);
/* 1 when the library carries the experiment kernels (`make EXPERIMENTS=1`: the persistent b = 1 step kernel selected by
 * CBX_MEGA=1, the stream-ahead GEMV selected by CBX_GEMV_STREAM=1 — both parity-green and measured slower);
 * the default build returns 0 and ignores those switches.  No counterpart in the reference. */
int cbx_has_experiments(void);

/* ---- integer bookkeeping (bit-exact) ------------------------------------------ */

/* vllm: v1/worker/block_table.py:357-377.  slot[i] = bt[req[i]][pos[i]/16]*16 + pos[i]%16,
 * or -1 when pos[i] < 0.  bt int32 [n_rows, bt_stride]; req,pos int32 [n]; slot int64 [n]. */
int cbx_slot_mapping(const int32_t* block_table, int bt_stride, const int32_t* req,
                     const int32_t* pos, int64_t* slot, int n, void* stream);

/* Host-side block pool with vLLM's policy (vllm: v1/core/block_pool.py:162-177,333-347,
 * 419-433; single_type_kv_cache_manager.py:120,350).  Pure CPU, no CUDA. */
typedef struct cbx_pool cbx_pool;
cbx_pool* cbx_pool_create(int num_blocks);                     /* block 0 reserved (null block) */
void      cbx_pool_destroy(cbx_pool*);
int       cbx_pool_num_free(const cbx_pool*);
int       cbx_pool_alloc(cbx_pool*, int n, int32_t* out_ids);  /* popleft n; <0 if n > free     */
int       cbx_pool_free(cbx_pool*, const int32_t* ids, int n); /* appended in the order given   */

/* ---- per-kernel ops (unit-testable pieces of the step) ------------------------- */

/* vllm: _C.rms_norm / fused_add_rms_norm as used by llama.py:316-333.
 * residual == NULL: out = norm(x)·w.   Else: residual = bf16(x + residual) in place,
 * out = norm(residual)·w.  x,out,residual [T,1024]; w [1024]. */
int cbx_rmsnorm(const void* x, void* residual, const void* w, void* out, int T, float eps, void* stream);

/* Split-K form of the fused add + RMSNorm: the projection that closes a residual branch left
 * `splits` fp32 partial products partials[splits][T][1024] (cbx_linear_partials); this folds them:
 * residual = bf16(bf16(sum of partials) + residual) in place, out = norm(residual)·w. */
int cbx_rmsnorm_reduce(const void* partials, int splits, void* residual, const void* w, void* out, int T,
                       float eps, void* stream);
/* partials[z][T][1024] (fp32) = x[:, Kz]·W[:, Kz]^T for the z-th of `splits` equal K slices
 * (tcgen05 GEMM, one grid.z slice per split; K % (64·splits) == 0).  Used for o_proj / down_proj
 * at decode batch sizes, where 16 column tiles alone would leave most of the 148 SMs idle. */
int cbx_linear_partials(const void* x, const void* W, void* partials, int T, int K, int splits, void* stream);

/* Per-output-feature e4m3 quantisation of a bf16 [N,K] weight: scale[n] = max|W[n,:]|/448 (1 for a zero
 * row), W8[n,k] = e4m3_rn_satfinite(W[n,k]/scale[n]).  Bit-identical to torch's
 * (W.float()/scale[:,None]).to(torch.float8_e4m3fn). */
int cbx_quantize_fp8_rows(const void* W, void* W8, float* scale, int N, int K, void* stream);
/* cbx_linear over FP8 weights: y = bf16((x · W8^T) * wscale[n]) (epilogue 1: SiLU(gate)·up of the scaled
 * pair).  GEMV regime widens the bytes in registers; tcgen05 regime widens them in shared memory
 * (converter warps write the SWIZZLE_128B bf16 operand tile).  wscale must hold round_up(N,128) floats. */
int cbx_linear_fp8w(const void* x, const void* W8, const float* wscale, void* y, int T, int N, int K, int ldy,
                    int epilogue, int path, void* stream);
/* FP8-weight forms of the fused layer pieces below (same arguments + wscale). */
int cbx_layer_qkv_fp8w(const void* res, const void* ln_w, const void* Wqkv8, const float* wscale,
                       const void* cos_sin, const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer,
                       void* h_scratch, int T, float eps, void* stream);
int cbx_linear_residual_fp8w(const void* x, const void* W8, const float* wscale, void* res, int T, int K, void* stream);
int cbx_linear_partials_fp8w(const void* x, const void* W8, const float* wscale, void* partials, int T, int K,
                             int splits, void* stream);
int cbx_layer_gate_up_fp8w(const void* res, const void* ln_w, const void* Wgu8, const float* wscale, void* act,
                           void* h_scratch, int T, float eps, void* stream);
int cbx_head_logits_fp8w(const void* res, const void* norm_w, const void* Whead8, const float* wscale, void* logits,
                         void* h_scratch, int T, float eps, void* stream);

/* y[T,N] = x[T,K] · W[N,K]^T (bf16 in, fp32 accumulate, bf16 out).  Replaces the
 * cuBLAS GEMMs behind vllm llama.py:117-120,221,225 and t3.py:729-730.
 * epilogue: 0 = store; 1 = SiLU(gate)·up with W packed by cbx_pack_gate_up (y is [T,N/2]);
 * path: 0 = auto, 1 = force SIMT GEMV (T ≤ 8), 2 = force tcgen05 GEMM.  ldy = row stride of y. */
int cbx_linear(const void* x, const void* W, void* y, int T, int N, int K, int ldy,
               int epilogue, int path, void* stream);
/* Interleaves gate/up rows in groups of 16 ([g0..15,u0..15,g16..31,...]) so both land
 * in one accumulator fragment.  src [2*I,K] (gate;up stacked, vllm llama.py:433-441). */
int cbx_pack_gate_up(const void* src, void* dst, int I, int K, void* stream);

/* vllm: _C.rotary_embedding (neox, llama.py:223) + reshape_and_cache_flash
 * (flash_attn.py:850-874).  qkv [T,3072]: q rotated in place; rotated k and v are written
 * to page slot[i] of `kv_layer` (base of this layer's pages).  cos_sin bf16 [max_pos,64]. */
int cbx_rope_kv_append(void* qkv, const int32_t* pos, const int64_t* slot, const void* cos_sin,
                       void* kv_layer, int T, void* stream);

/* vllm: flash_attn_varlen_func over the paged cache (flash_attn.py:796-800), decode form:
 * one query row per entry.  q rows are read from qkv [T,3072] (first 1024 columns);
 * row i attends to positions [0, kv_len[i]) of block-table row row_seq[i].
 * out bf16 [T,1024].  scratch: cbx_attn_scratch_bytes(T, max_kv_len). */
int64_t cbx_attn_scratch_bytes(int T, int max_kv_len);
int cbx_attn_num_splits(int T, int max_kv_len);
/* counters: int32 [T], zero before first use; the kernel leaves them zero (split-K tickets: the
 * CTA that publishes a row's last partial merges the row, so there is no second launch). */
int cbx_paged_attn_decode(const void* qkv, const void* kv_layer, const int32_t* block_table,
                          int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
                          void* out, void* scratch, int32_t* counters, int T, int max_kv_len, void* stream);
/* For 4 < T <= 4095 rows the decode kernel cuts the launch's pages, flattened in (row, page) order,
 * into 148 equal ranges (one per SM) so that ragged context lengths stay balanced.  The range plan
 * depends only on kv_len, so a caller that runs many layers over the same rows builds it once with
 * cbx_attn_plan (returns 1 when the balanced variant does not apply to T: then use
 * cbx_paged_attn_decode) and launches each layer with cbx_paged_attn_decode_planned.
 * cbx_paged_attn_decode builds the plan itself on every call. */
int cbx_attn_plan(const int32_t* kv_len, int T, void* scratch, void* stream);
/* Same, with the block table the layers will use: the plan then also holds the first three physical page ids of every
 * range, so each layer's attention launch starts its page stream one L2 round trip after its wait instead of three.
 * block_table / row_seq must be the ones later passed to cbx_paged_attn_decode_planned. */
int cbx_attn_plan_paged(const int32_t* kv_len, int T, void* scratch, void* stream, const int32_t* block_table,
                        int bt_stride, const int32_t* row_seq);
int cbx_paged_attn_decode_planned(const void* qkv, const void* kv_layer, const int32_t* block_table,
                                  int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
                                  void* out, void* scratch, int32_t* counters, int T, int max_kv_len,
                                  void* stream);

/* Allocation-free form of cbx_paged_attn_decode.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same inputs, output and arithmetic as cbx_paged_attn_decode, but the
 * scratch (split partials + page-range plan) and the ticket counters live in a file-scope device array inside
 * the library (k_attn.cu, ~1.27 MB, held by every GPU that loads the library until exit), so the caller passes
 * neither.  Limit: 5 <= T <= 4095 query rows (the balanced kernel's range; T <= 4 needs no scratch — call
 * cbx_paged_attn_decode with scratch = counters = NULL); other T returns -1 with a message.  The call itself
 * performs no cudaMalloc / cudaFree / allocating library call.  All calls on a device SHARE the array: the
 * tickets are cleared on the caller's stream at every call, but two calls that overlap on the GPU (different
 * streams, or graph branches) race — serialise them.  The first call on a device resolves the array's device
 * address (cudaGetSymbolAddress); make that call outside graph capture. */
int cbx_paged_attn_decode_static(const void* qkv, const void* kv_layer, const int32_t* block_table,
                                 int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* out,
                                 int T, int max_kv_len, void* stream);

/* Prefill form of the same attention (causal, many query rows per sequence): an item is 16
 * consecutive query rows of one sequence (query block qb = first position / 16), which attend to
 * KV pages 0..qb; each page is loaded once per item.  The rows' K/V must already be in the cache
 * (cbx_layer_qkv appends them).  items: device int32 [n_items][4] = {row0, n_q (1..16),
 * block-table row, qb}; q rows are read from qkv [T,3072], out is bf16 [T,1024]. */
int cbx_paged_attn_prefill(const void* qkv, const void* kv_layer, const int32_t* block_table, int bt_stride,
                           const int32_t* items, int n_items, void* out, void* stream);

/* ---- fused layer pieces: what cbx_engine_forward launches per layer ------------------------
 * The residual stream `res` [T,1024] is updated in place by the projections that close a
 * residual branch (vllm llama.py:316-333 fused_add_rms_norm convention, rounding points kept:
 * res = bf16(bf16(y) + res); h = bf16(bf16(res·rstd)·w)).  T <= 4 takes the weight-streaming
 * GEMV kernels (norm fused as prologue), T > 4 the tcgen05 GEMM (h_scratch [T,1024] receives the
 * normalised rows).  All launches carry programmatic stream serialisation. */
/* RMSNorm(res)·ln_w -> ·Wqkv^T -> RoPE(q,k) -> q into qkv[:, 0:1024] (qkv is [T,3072]); k, v into
 * page slot[t] of kv_layer (llama.py:221-223 + flash_attn.py:850-874). */
int cbx_layer_qkv(const void* res, const void* ln_w, const void* Wqkv, const void* cos_sin,
                  const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer, void* h_scratch,
                  int T, float eps, void* stream);
/* res += x[T,K]·W[1024,K]^T  (o_proj: K = 1024, llama.py:225; down_proj: K = 4096, llama.py:120). */
int cbx_linear_residual(const void* x, const void* W, void* res, int T, int K, void* stream);
/* act[T,4096] = SiLU(g)·u of RMSNorm(res)·ln_w·Wgu^T, Wgu packed by cbx_pack_gate_up (llama.py:117-119). */
int cbx_layer_gate_up(const void* res, const void* ln_w, const void* Wgu, void* act, void* h_scratch, int T,
                      float eps, void* stream);
/* The attention half of a layer as ONE launch for T <= 4 rows (llama.py:212-226,316-326 + flash_attn.py:796-800,
 * 850-874): res += Wo · attention(RoPE(Wqkv · RMSNorm(res)·ln_w)) with the new token's K/V appended to page slot[t]
 * — 16 clusters (one per head) of 6 CTAs; bf16 weights only.  row t attends to positions [0, kv_len[t]) of
 * block-table row row_seq[t], the new token (position kv_len[t]-1) included; kv_len[t] = 0 leaves the row's
 * attention output zero.  scratch: cbx_layer_attn_block_scratch_bytes() bytes, zero before first use (left
 * zero): the 16 heads add their O-proj partials into it as 64-bit fixed point, so the sum is order-independent.  q_out: optional [T,3072] buffer receiving the rotated q (columns 0..1023), else NULL.
 * cbx_layer_attn_block_supported(T, max_kv_len) says whether the context fits the kernel's shared-memory page
 * buffer (T = 1: 3841 tokens, 2: 1921, 3-4: 961); the engine falls back to the three-kernel chain otherwise. */
int64_t cbx_layer_attn_block_scratch_bytes(void);
int cbx_layer_attn_block_supported(int T, int max_kv_len);
int cbx_layer_attn_block(void* res, const void* ln_w, const void* Wqkv, const void* Wo, const void* cos_sin,
                         const int32_t* pos, const int64_t* slot, void* kv_layer, const int32_t* block_table,
                         int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* scratch, void* q_out,
                         int T, int max_kv_len, float eps, void* stream);
/* The MLP half of a layer as ONE launch for T <= 4 rows (llama.py:117-120,323-326): res += Wd · (SiLU(g)·u) of
 * RMSNorm(res)·ln_w · Wgu^T.  One CTA per SM owns 27-28 of the 4096 MLP features, computes their activations and
 * multiplies them by ITS rows of WdT = Wd^T ([4096,1024] row-major bf16: the transposed down-proj weights), adding the
 * fp32 partial [T,1024] into `scratch` as 64-bit fixed point with one bulk reduction (order-independent sum; four
 * copies of the accumulator, CTA b into copy b % 4, so that L2 does not serialise 148 reductions per address).  bf16
 * weights only; Wgu packed by cbx_pack_gate_up.  scratch: cbx_layer_mlp_block_scratch_bytes() bytes, zero before
 * first use (left zero).  Inside the engine the accumulator is folded by the next layer's attention block instead. */
int64_t cbx_layer_mlp_block_scratch_bytes(void);
int cbx_layer_mlp_block(void* res, const void* ln_w, const void* Wgu, const void* WdT, void* scratch, int T, float eps,
                        void* stream);
/* logits[T, CBX_LOGITS_LD] = RMSNorm(res)·norm_w · Whead[8194,1024]^T (llama.py:430 + t3.py:729-730). */
int cbx_head_logits(const void* res, const void* norm_w, const void* Whead, void* logits, void* h_scratch,
                    int T, float eps, void* stream);

/* t3.py:732 (CFG merge) + vllm v1 Sampler (sampler.py:67-300) in one kernel, one CTA per
 * pair.  logits bf16 [2B, CBX_LOGITS_LD] (rows b = cond, B+b = uncond).  Per-pair state
 * arrays are documented in cbx_sampler_state.  Emits one token per active pair. */
typedef struct cbx_sampler_state {
  int32_t*  last_tok;     /* [B] out: sampled speech id (no +1000 offset)                   */
  int32_t*  out_tokens;   /* [B, out_stride] emitted ids                                    */
  int32_t*  n_out;        /* [B] in/out                                                     */
  int32_t*  max_out;      /* [B] max_tokens per pair                                        */
  int32_t*  finished;     /* [B] in/out: 1 once stop token or max_out reached               */
  int32_t*  seq_len;      /* [2B] in/out: += 1 for rows of pairs that were active           */
  uint32_t* seen;         /* [B, CBX_SEEN_WORDS] prompt∪output bitmap (speech-id space)     */
  const float* temperature; const float* top_p; const float* min_p; const float* rep_pen; /* [B] */
  const int32_t* top_k;   /* [B] 0 = off                                                     */
  const int32_t* ignore_eos; /* [B]                                                          */
  const uint64_t* seed;   /* [B] Philox key                                                  */
  const uint64_t* uid;    /* [B] request uid (Philox counter words 2,3)                      */
  int32_t*  n_finished;   /* [1] running count of finished pairs (host polls this)           */
  int out_stride; int stop_token; float cfg_scale;
} cbx_sampler_state;
int cbx_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, void* stream);
/* Test hook: on != 0 makes every later sampler launch of this process find the top-p pivot with the MSB-first radix passes
 * (the first implementation, kept as the fallback for near-constant rows) instead of the bucket histogram + exact
 * ranking; both must emit the same tokens bit for bit (tests/test_gpu_engine.py). */
int cbx_debug_sampler_radix_pivot(int on);

/* t3.py:503-514 (decode) — x[r] = speech_emb[last_tok[r % B]] for r in [0,2B), plus the
 * per-step index work: pos[r] = seq_len[r]; slot[r] from the block table (-1 for finished
 * pairs); kv_len[r] = seq_len[r]+1 (0 for finished pairs). */
int cbx_decode_prepare(const void* speech_emb, const int32_t* last_tok, const int32_t* finished,
                       const int32_t* seq_len, const int32_t* block_table, int bt_stride,
                       void* x, int32_t* pos, int64_t* slot, int32_t* kv_len, int B, void* stream);

/* ---- conditioning-prefix producer ------------------------------------------------
 * T3CondEnc + Perceiver of the reference (models/t3/modules/cond_enc.py:57-123, perceiver.py:116-215;
 * driven from tts.py:238-257), fp32 like the reference.  All pointers are device pointers to fp32
 * tensors with the checkpoint's shapes ("cond_enc.*" of t3_cfg.safetensors). */
typedef struct {
  const float* spkr_w;      /* spkr_enc.weight                [1024,256]  */
  const float* spkr_b;      /* spkr_enc.bias                  [1024]      */
  const float* emo_w;       /* emotion_adv_fc.weight          [1024,1]    */
  const float* query;       /* perceiver.pre_attention_query  [1,32,1024] */
  const float* ln_w;        /* perceiver.attn.norm.weight     [1024]      */
  const float* ln_b;        /* perceiver.attn.norm.bias       [1024]      */
  const float* wq; const float* bq;   /* perceiver.attn.to_q       [1024,1024], [1024] */
  const float* wk; const float* bk;   /* perceiver.attn.to_k                           */
  const float* wv; const float* bv;   /* perceiver.attn.to_v                           */
  const float* wo; const float* bo;   /* perceiver.attn.proj_out                       */
  const float* speech_emb;  /* speech_emb.weight              [8194,1024] */
  const float* speech_pos;  /* speech_pos_emb.emb.weight      [>= n_prompt,1024] */
} cbx_cond_weights;
int64_t cbx_cond_workspace_bytes(int max_prompt);
/* out fp32 [34,1024] = spkr_enc(speaker_emb[256]) | perceiver(speech_emb[tok]+speech_pos[0..n)) (32 rows)
 * | emotion_adv_fc(emotion).  prompt_tokens int32 [n_prompt], 1 <= n_prompt <= 1024 (the reference uses 150). */
int cbx_cond_encode(const cbx_cond_weights* w, const float* speaker_emb, const int32_t* prompt_tokens, int n_prompt,
                    float emotion, void* workspace, float* out, void* stream);
/* row fp32 [1024] = emotion_adv_fc(emotion): what update_exaggeration (tts.py:249-257) recomputes. */
int cbx_cond_emotion_row(const cbx_cond_weights* w, float emotion, float* row, void* stream);

/* Token post-processing = the T3 -> vocoder hand-off (reference tts.py:448-486 + models/s3tokenizer/
 * __init__.py:38-78, which run as Python list surgery + a torch mask per prompt): for item i, read
 * the n_out[slots[i]] speech ids of sampler slot slots[i] (un-offset space), cut at the first
 * eos_token and crop tail_crop ids before it, keep what lies after the first sos_token and before the
 * first eos_token, drop ids >= sos_token, cap at guard_cap[i] (NULL = no cap), and write int64 ids to
 * dst[dst_rows[i]*dst_stride ...] with the count in dst_len[dst_rows[i]]. */
int cbx_postprocess_tokens(const int32_t* out_tokens, const int32_t* n_out, int out_stride, const int32_t* slots,
                           int n, int sos_token, int eos_token, int tail_crop, const int32_t* guard_cap,
                           int64_t* dst, int dst_stride, const int32_t* dst_rows, int32_t* dst_len, void* stream);

/* t3.py:548-569 (full-block prefill embeds).  desc int32 [T,4] = {kind, a, b, c}:
 * kind 0: cond_emb row  (a = request, b = row 0..33) from cond_emb fp32 [n_req,34,1024];
 * kind 1: text row       text_emb[a] + text_pos[b]; kind 2: zero row (uncond text);
 * kind 3: start-of-speech speech_emb[a] + speech_pos[b];
 * kind 4: speech_emb[a] alone — the decode embedding of an already generated token (t3.py:503-514), for a preempted
 *         request whose KV is recomputed over prefix + generated tokens (vLLM v1 preemption = recompute). */
int cbx_prefill_embed(const int32_t* desc, const float* cond_emb, const void* text_emb,
                      const void* text_pos, const void* speech_emb, const void* speech_pos,
                      void* x, int T, void* stream);

/* The rows of a prefill step from one record per sequence (T3MultiModalProcessor.apply's layout, t3.py:229-262, for both
 * members of every CFG pair): seq int32 [n_seq,8] = {first row, rows P, text tokens nt, request index into cond_emb,
 * uncond member (0/1), block-table row, offset of the sequence's ids in text_ids, offset of its history in hist}, sorted
 * by first row; rows of a sequence = 34 conditioning rows, nt text rows (zero rows for the uncond member), start-of-speech,
 * P-35-nt history rows (tokens generated before a preemption).  Writes desc [T,4] (cbx_prefill_embed's), pos, row_seq and
 * kv_len = pos + 1.  Replaces the per-row Python the reference runs in apply / get_input_embeddings for every request. */
int cbx_prefill_rows(const int32_t* seq, int n_seq, const int32_t* text_ids, const int32_t* hist, int T,
                     int32_t* desc, int32_t* pos, int32_t* row_seq, int32_t* kv_len, void* stream);

/* ---- engine: the whole T3 step behind one handle -------------------------------- */

typedef struct cbx_weights {
  int n_layers;
  const void* const* wqkv;   /* [L] -> [3072,1024]                  */
  const void* const* wo;     /* [L] -> [1024,1024]                  */
  const void* const* wgu;    /* [L] -> [8192,1024] PACKED (cbx_pack_gate_up) */
  const void* const* wd;     /* [L] -> [1024,4096]                  */
  const void* const* ln1;    /* [L] -> [1024]                       */
  const void* const* ln2;    /* [L] -> [1024]                       */
  const void* norm;          /* [1024]                              */
  const void* speech_emb;    /* [8194,1024]                         */
  const void* speech_head;   /* [8194,1024]                         */
  const void* cos_sin;       /* [max_pos,64] bf16                   */
  float rms_eps;
} cbx_weights;

typedef struct cbx_engine_cfg {
  int max_pairs;          /* B: CFG pair slots                                    */
  int max_rows;           /* row capacity of activation buffers (≥ 2B; prefill T) */
  int max_model_len;      /* tokens per sequence                                  */
  int num_blocks;         /* KV pages per layer (block 0 is the null block)       */
  int out_stride;         /* capacity of out_tokens per pair                      */
} cbx_engine_cfg;

typedef struct cbx_engine cbx_engine;
int64_t cbx_engine_workspace_bytes(const cbx_engine_cfg* cfg);
/* kv_cache: [n_layers][num_blocks] pages of CBX_PAGE_BYTES; block_table int32
 * [2B, ceil(max_model_len/16)]; workspace: cbx_engine_workspace_bytes(cfg), 1024-aligned. */
cbx_engine* cbx_engine_create(const cbx_engine_cfg* cfg, const cbx_weights* w, void* kv_cache,
                              int32_t* block_table, void* workspace, const cbx_sampler_state* st);
void cbx_engine_destroy(cbx_engine*);
/* ---- optional FP8-e4m3 weight path (north star: "an optional FP8-e4m3 per-channel weight path is
 * reported separately with its own tolerance"; SURVEY §8(d) cfg 5).  Weight-only: activations, KV
 * cache, embeddings and the sampler stay bf16 / fp32.  The reference has no counterpart (vLLM's
 * `quantization="fp8"` would be the nearest switch on `LLM(...)`, tts.py:130-140). */
typedef struct cbx_fp8_weights {
  int n_layers;
  const void* const* wqkv;   /* [L] -> [3072,1024] e4m3 bytes                 */
  const void* const* wo;     /* [L] -> [1024,1024]                            */
  const void* const* wgu;    /* [L] -> [8192,1024] quantised AFTER cbx_pack_gate_up */
  const void* const* wd;     /* [L] -> [1024,4096]                            */
  const float* const* sqkv;  /* [L] -> [3072] fp32 scale per output feature   */
  const float* const* so;    /* [L] -> [1024]                                 */
  const float* const* sgu;   /* [L] -> [8192] (packed row order)              */
  const float* const* sd;    /* [L] -> [1024]                                 */
  const void* speech_head;   /* [8194,1024] e4m3 bytes                        */
  const float* shead;        /* [8320] = 8194 scales padded with 1.0 to a multiple of 128 */
} cbx_fp8_weights;
/* Switch a created engine's projections + speech head to FP8 weights (pointers must outlive it). */
int cbx_engine_set_fp8_weights(cbx_engine*, const cbx_fp8_weights*);
/* Transposed bf16 down-proj weights per layer (wd_t[l] = Wd[l]^T, [4096,1024]) for the engine's two-launch layer of the
 * T <= 4 decode path (cbx_layer_attn_block's kernel + cbx_layer_mlp_block's kernel, each folding the other's
 * accumulator).  Without them the T <= 4 layer runs three launches.  NULL removes them.  No reference counterpart
 * (vLLM keeps one layout per weight); a layout choice of this implementation. */
int cbx_engine_set_down_t(cbx_engine*, const void* const* wd_t, int n_layers);

/* Device pointer of an internal buffer by name ("x","h","qkv","attn","act","logits","pos",
 * "slot","kv_len","row_seq","head_rows") for tests / the Python façade; NULL if unknown. */
void* cbx_engine_buffer(cbx_engine*, const char* name);

/* Parity hook (tests only): while trace != NULL, cbx_engine_forward / cbx_engine_decode copy,
 * after every layer li, the residual stream (= the HF hidden state after the layer) to
 * trace[li*T*1024] (bf16).  Set before the call, clear with NULL; do not use with graphs. */
int cbx_engine_set_trace(cbx_engine*, void* trace);

/* While n_items > 0, cbx_engine_forward uses cbx_paged_attn_prefill with these items (device
 * int32 [n_items][4]) instead of the per-row decode attention.  Cleared with n_items = 0. */
int cbx_engine_set_prefill_items(cbx_engine*, const int32_t* items, int n_items);

/* Llama forward over T rows whose embeddings are already in buffer "res" (alias "x"), with
 * pos/slot/row_seq/kv_len filled (t3.py:742-778 + vllm llama.py:394-430).  Leaves final-norm
 * hidden states in buffer "h". */
int cbx_engine_forward(cbx_engine*, int T, int max_kv_len, void* stream);
/* Head GEMM over the 2B rows listed in buffer "head_rows" (row index into "h"; cond rows
 * first) + cbx_cfg_sample (t3.py:725-739 + sampler). */
int cbx_engine_sample(cbx_engine*, int B, void* stream);
/* Admission of n_sub new pairs into a running batch of B slots (continuous batching): head GEMM
 * over head_rows[0 .. 2*n_sub) + sampler for the slots pair_map[0 .. n_sub) only (device int32). */
int cbx_engine_sample_subset(cbx_engine*, int B, int n_sub, const int32_t* pair_map, void* stream);
/* n decode steps for pair slots [0,B): forward(2B rows) → head → sample, no host sync.  The
 * step's inputs (embedding rows, positions, slots, kv lengths) were staged on the device by the
 * previous cbx_engine_sample / decode step's sampler launch.
 * use_graph != 0 replays a CUDA graph of one step (captured on first use per (B, max_kv_len
 * bucket)). */
int cbx_engine_decode(cbx_engine*, int B, int n_steps, int max_kv_len, int use_graph, void* stream);
/* kernel launches issued by the last cbx_engine_decode call (for bench.py "gpu_launches"). */
int64_t cbx_engine_last_launch_count(const cbx_engine*);

#ifdef __cplusplus
}
#endif
#endif /* CBX_T3_H_ */
