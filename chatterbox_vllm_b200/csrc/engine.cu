// engine.cu — host-side orchestration of the T3 step: the launch sequence that replaces
// T3VllmModel.forward / compute_logits (t3.py:725-778) + vLLM's LlamaModel.forward
// (vllm llama.py:394-430) + the sampler, and a CUDA-graph replay loop that replaces vLLM's
// per-step Python scheduler for the decode phase (no host sync per step).
#include <stdlib.h>
#include <string.h>

#include <map>
#include <vector>

#include "common.cuh"

namespace cbx {
struct NextStep {
  const uint4* speech_emb; uint4* res; int32_t* pos; int64_t* slot; int32_t* kv_len; int32_t* row_seq;
  const int32_t* bt; int bt_stride;
};
int64_t launches();
int init_gemm();
int init_attn();
int init_sample();
// every projection takes an optional per-output-feature scale: non-NULL = the weights are FP8-e4m3 bytes
int gemv_bf16(const void*, const void*, void*, int, int, int, int, int, cudaStream_t, const float* wscale = nullptr);
int gemv_qkv_rope(const void*, const void*, const void*, const void*, const int32_t*, const int64_t*, void*, void*, int,
                  float, cudaStream_t, const void* next_w = nullptr, size_t next_bytes = 0, const float* wscale = nullptr);
int gemv_residual(const void*, const void*, void*, int, int, cudaStream_t, const void* next_w = nullptr, size_t next_bytes = 0,
                  const float* wscale = nullptr, void* zero_ptr = nullptr, int zero_bytes = 0, const void* res_in = nullptr);
int gemv_norm_gate_up(const void*, const void*, const void*, void*, int, float, cudaStream_t, const void* next_w = nullptr,
                      size_t next_bytes = 0, const float* wscale = nullptr, const void* acc = nullptr, void* res_out = nullptr);
void attn_set_prefetch(const void*, size_t);
int attn_plan(const int32_t* kv_len, int T, void* scratch, void* stream, const int32_t* block_table, int bt_stride,
              const int32_t* row_seq, int n_ctas = 0);
int paged_attn_decode_ex(const void*, const void*, const int32_t*, int, const int32_t*, const int32_t*, void*, void*, int32_t*,
                         int, int, bool, void*, int n_ctas = 0);
int gemv_norm_head(const void*, const void*, const void*, void*, int, int, int, float, cudaStream_t, const float* wscale = nullptr);
int gemm_bf16_tc(const void*, const void*, void*, int, int, int, int, int, cudaStream_t, const float* wscale = nullptr);
int gemm_qkv_rope(const void*, const void*, void*, const void*, const int32_t*, const int64_t*, void*, int, cudaStream_t,
                  const float* wscale = nullptr);
int gemm_residual(const void*, const void*, void*, int, int, int, cudaStream_t, const float* wscale = nullptr);
int launch_decode_prepare(const void*, const int32_t*, const int32_t*, const int32_t*, const int32_t*, int, void*,
                          int32_t*, int64_t*, int32_t*, int32_t*, int32_t*, int, cudaStream_t);
int launch_cfg_sample(const void*, const cbx_sampler_state*, int, const NextStep&, cudaStream_t, const int32_t* pair_map = nullptr,
                      int n_sub = 0);
int gemm_partials(const void*, const void*, void*, int, int, int, int, cudaStream_t, const float* wscale = nullptr);
int gemm_partials_fold(const void* x, const void* W, void* part, int M, int K, int splits, cudaStream_t s, void* res,
                       const void* ln_w, void* h_out, float eps, void* counters);
bool layer_attn_block_supported(int T, int max_kv_len);
int layer_attn_block_max_len(int T);
int layer_attn_block(const void* res, const void* ln_w, const void* Wqkv, const void* Wo, const void* cos_sin, const int32_t* pos,
                     const int64_t* slot, void* kv_layer, const int32_t* bt, int bt_stride, const int32_t* row_seq,
                     const int32_t* kv_len, void* acc, void* q_out, int T, float eps, cudaStream_t s);
bool layer_mlp_block_on();
int layer_attn_block_fold(const void* res, const void* acc_in, void* res_wb, void* zero_ptr, int zero_bytes, const void* ln_w,
                          const void* Wqkv, const void* Wo, const void* cos_sin, const int32_t* pos, const int64_t* slot,
                          void* kv_layer, const int32_t* bt, int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
                          void* acc, int T, float eps, cudaStream_t s);
int layer_mlp_block(const void* x, const void* acc_in, void* res_wb, const void* ln_w, const void* Wgu, const void* WdT,
                    void* acc_out, void* zero_ptr, int zero_bytes, int T, float eps, cudaStream_t s);
struct MegaArgs {
  const bf16* ln1[32]; const bf16* wqkv[32]; const bf16* wo[32];
  const bf16* ln2[32]; const bf16* wgu[32]; const bf16* wd[32];
  const bf16* norm_w; const bf16* head_w;
  const bf16* res_in; bf16* logits; uint8_t* ws; unsigned* sync; bf16* trace;
  const bf16* cos_sin; const int32_t* pos; const int64_t* slot; const int32_t* row_seq; const int32_t* kv_len;
  const int32_t* bt; uint8_t* kv_cache; size_t layer_bytes;
  int bt_stride, T, n_layers, vocab, ldl; float eps; int flags;
};
#ifdef CBX_EXPERIMENTS   // `make EXPERIMENTS=1` links k_mega.cu (the persistent b = 1 step kernel: parity-green, slower)
size_t mega_ws_bytes();
int init_mega();
int launch_decode_mega(const MegaArgs&, cudaStream_t);
#else
inline size_t mega_ws_bytes() { return 1024; }
inline int init_mega() { return 0; }
inline int launch_decode_mega(const MegaArgs&, cudaStream_t) { return 1; }      // "not applicable": the kernel chain runs
#endif
}  // namespace cbx

// 1 when the library was built with the measured-slower experiment kernels (CBX_MEGA / CBX_GEMV_STREAM act only then)
extern "C" int cbx_has_experiments(void) {
#ifdef CBX_EXPERIMENTS
  return 1;
#else
  return 0;
#endif
}
using namespace cbx;

namespace {
__global__ void __launch_bounds__(128) gather_rows_kernel(const uint4* src,
                                                          const int32_t* rows,
                                                          uint4* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  dst[(size_t)blockIdx.x * 128 + threadIdx.x] = src[(size_t)rows[blockIdx.x] * 128 + threadIdx.x];
}
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr int kAccBytes = 4 * 1024 * 8;                // one fixed-point accumulator: [4][1024] u64
constexpr int kAccSlots = 2 + 2 * kMlpAccCopies;       // O-proj: even / odd layers; down-proj: even / odd layers x copies
}  // namespace

struct cbx_engine {
  cbx_engine_cfg cfg;
  cbx_weights w;
  std::vector<const void*> wqkv, wo, wgu, wd, ln1, ln2;
  std::vector<const void*> wd_t;   // transposed bf16 down-proj weights [4096,1024] per layer (cbx_engine_set_down_t), or empty
  // FP8-e4m3 weight path (cbx_engine_set_fp8_weights): the four projections above and the head then point at
  // e4m3 bytes and these hold their per-output-feature scales; NULL entries = bf16 weights
  std::vector<const float*> sqkv, so, sgu, sd;
  const void* head = nullptr; const float* shead = nullptr;
  bool fp8 = false;
  uint8_t* kv_cache;
  int32_t* block_table;
  int bt_stride;
  cbx_sampler_state st;
  // workspace carve-up
  bf16 *res, *h, *qkv, *attn, *act, *hs, *logits;
  int32_t *pos, *kv_len, *row_seq, *head_rows, *attn_counters;
  // 64-bit fixed-point accumulators of the fused T <= 4 kernels, [4][1024] each, all zero between steps: slots 0, 1 the
  // O-proj output of even / odd layers (k_block.cu), then the down-proj output of even / odd layers (k_mlp.cu; each in
  // kMlpAccCopies copies, common.cuh)
  unsigned long long* block_acc;
  int64_t* slot;
  void* attn_scratch;
  void* attn_scratch2;         // second plan + partial slots: the uncond half of a two-stream step (forward_dual)
  cudaStream_t s2 = nullptr;   // its stream, and the fork / join events
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaEvent_t ev_att[2][32] = {};   // "attention of layer l done" per half: the two halves' attention launches alternate
  float* part;                 // split-K partials [4][split_max_rows][1024] fp32 
  unsigned* fold_cnt;          // arrival counters of the split-K GEMMs that fold their own partials (zero between launches)
  unsigned* mega_sync;         // exit counter + epoch of the persistent decode kernel
  uint8_t* mega_ws;            // its tagged exchange buffers
  bool mega = false;           // env CBX_MEGA=1: T <= 4 decode steps run as ONE persistent kernel (k_mega.cu).  Parity-tested,
                               // but measured slower on B200 (b=1: ~1000 us/step vs ~620 us for the PDL kernel chain), so off
  int split_max_rows = 512;
  const int32_t* prefill_items = nullptr; int n_prefill_items = 0;
  bool l2_prefetch = false;  // measured: no gain at b=1 (L2->SM bandwidth ~ HBM bandwidth on B200); env CBX_L2_PREFETCH=1 enables
  int skip = 0;   // profiling aid (env CBX_SKIP bitmask: 1 qkv, 2 attn, 4 o, 8 gate_up, 16 down): results are garbage
  std::map<std::pair<int, int>, std::pair<cudaGraphExec_t, int64_t>> graphs;   // (B, kv bucket) -> exec, nodes
  int64_t last_launches;
  bf16* trace = nullptr;
};

static size_t carve(const cbx_engine_cfg* c, cbx_engine* e, uint8_t* base) {
  size_t off = 0;
  const size_t R = (size_t)c->max_rows, B2 = (size_t)2 * c->max_pairs;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 1024); return base ? base + o : nullptr; };
  void* p;
  p = take(R * 1024 * 2); if (e) e->res = (bf16*)p;
  p = take(R * 1024 * 2); if (e) e->h = (bf16*)p;
  p = take(R * 3072 * 2); if (e) e->qkv = (bf16*)p;
  p = take(R * 1024 * 2); if (e) e->attn = (bf16*)p;
  p = take(R * 4096 * 2); if (e) e->act = (bf16*)p;
  p = take(B2 * 1024 * 2); if (e) e->hs = (bf16*)p;
  p = take(B2 * CBX_LOGITS_LD * 2); if (e) e->logits = (bf16*)p;
  p = take(R * 4); if (e) e->pos = (int32_t*)p;
  p = take(R * 4); if (e) e->kv_len = (int32_t*)p;
  p = take(R * 4); if (e) e->row_seq = (int32_t*)p;
  p = take(B2 * 4); if (e) e->head_rows = (int32_t*)p;
  p = take(R * 8); if (e) e->slot = (int64_t*)p;
  p = take(R * 4); if (e) e->attn_counters = (int32_t*)p;
  p = take((R + 1024) * (CBX_HIDDEN * 4 + CBX_HEADS * 8) + 256); if (e) e->attn_scratch = p;
#ifdef CBX_EXPERIMENTS
  p = take((size_t)4 << 20); if (e) e->attn_scratch2 = p;       // >= the balanced kernel's scratch (1.3 MB)
#endif
  p = take((size_t)4 * 512 * 1024 * 4); if (e) e->part = (float*)p;
  p = take(256); if (e) e->mega_sync = (unsigned*)p;
  p = take(256); if (e) e->fold_cnt = (unsigned*)p;
  p = take(kAccBytes * kAccSlots); if (e) e->block_acc = (unsigned long long*)p;
  p = take(mega_ws_bytes()); if (e) e->mega_ws = (uint8_t*)p;
  return off;
}

extern "C" int64_t cbx_engine_workspace_bytes(const cbx_engine_cfg* cfg) { return (int64_t)carve(cfg, nullptr, nullptr); }

extern "C" cbx_engine* cbx_engine_create(const cbx_engine_cfg* cfg, const cbx_weights* w, void* kv_cache,
                                         int32_t* block_table, void* workspace, const cbx_sampler_state* st) {
  if (!cfg || !w || !kv_cache || !block_table || !workspace || !st) { set_error("cbx_engine_create: NULL argument"); return nullptr; }
  if (cfg->max_rows < 2 * cfg->max_pairs || cfg->max_pairs < 1) { set_error("cbx_engine_create: max_rows < 2*max_pairs"); return nullptr; }
  if ((reinterpret_cast<uintptr_t>(workspace) & 1023) != 0) { set_error("cbx_engine_create: workspace not 1024-aligned"); return nullptr; }
  if (init_gemm() || init_attn() || init_sample() || init_mega()) return nullptr;
  if (w->n_layers > 32) { set_error("cbx_engine_create: n_layers > 32"); return nullptr; }
  cbx_engine* e = new cbx_engine();
  e->cfg = *cfg;
  e->w = *w;
  const int L = w->n_layers;
  e->wqkv.assign(w->wqkv, w->wqkv + L); e->wo.assign(w->wo, w->wo + L); e->wgu.assign(w->wgu, w->wgu + L);
  e->wd.assign(w->wd, w->wd + L); e->ln1.assign(w->ln1, w->ln1 + L); e->ln2.assign(w->ln2, w->ln2 + L);
  e->sqkv.assign(L, nullptr); e->so.assign(L, nullptr); e->sgu.assign(L, nullptr); e->sd.assign(L, nullptr);
  e->head = w->speech_head;
  e->kv_cache = (uint8_t*)kv_cache;
  e->block_table = block_table;
  e->bt_stride = (cfg->max_model_len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  e->st = *st;
  e->last_launches = 0;
  carve(cfg, e, (uint8_t*)workspace);
  // split-K attention tickets start at zero and reset themselves
  cudaMemset(e->attn_counters, 0, (size_t)cfg->max_rows * 4);
  cudaMemset(e->mega_sync, 0, 256);
  cudaMemset(e->fold_cnt, 0, 256);
  cudaMemset(e->block_acc, 0, kAccBytes * kAccSlots);
#ifdef CBX_EXPERIMENTS
  // second stream of the two-stream decode step (forward_dual); without it the step runs on one stream
  if (cudaStreamCreateWithFlags(&e->s2, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    e->s2 = nullptr;
  }
  for (int hf = 0; hf < 2 && e->s2; ++hf)
    for (int i = 0; i < 32; ++i)
      if (cudaEventCreateWithFlags(&e->ev_att[hf][i], cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); e->s2 = nullptr; break; }
#endif
  cudaMemset(e->mega_ws, 0, mega_ws_bytes());
#ifdef CBX_EXPERIMENTS
  if (const char* v = getenv("CBX_MEGA")) e->mega = v[0] != '0';
#endif
  if (const char* v = getenv("CBX_SKIP")) e->skip = atoi(v);
  if (const char* v = getenv("CBX_L2_PREFETCH")) e->l2_prefetch = v[0] != '0';
  if (const char* v = getenv("CBX_SPLIT_MAX_ROWS")) { e->split_max_rows = atoi(v); if (e->split_max_rows > 512) e->split_max_rows = 512; }
  return e;
}

extern "C" void cbx_engine_destroy(cbx_engine* e) {
  if (!e) return;
  for (auto& g : e->graphs) cudaGraphExecDestroy(g.second.first);
  if (e->s2) cudaStreamDestroy(e->s2);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  for (auto& hf : e->ev_att) for (cudaEvent_t ev : hf) if (ev) cudaEventDestroy(ev);
  delete e;
}

extern "C" void* cbx_engine_buffer(cbx_engine* e, const char* name) {
  struct { const char* n; void* p; } tab[] = {
      {"x", e->res}, {"res", e->res}, {"h", e->h}, {"qkv", e->qkv}, {"attn", e->attn}, {"act", e->act},
      {"hs", e->hs}, {"logits", e->logits}, {"pos", e->pos}, {"slot", e->slot}, {"kv_len", e->kv_len},
      {"row_seq", e->row_seq}, {"head_rows", e->head_rows}, {"block_acc", e->block_acc}};
  for (auto& t : tab) if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

// ---- fused layer pieces (C-ABI, regime chosen by T) -----------------------------------------
static int layer_qkv(const void* res, const void* ln_w, const void* Wqkv, const float* sc, const void* cos_sin,
                     const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer, void* h_scratch, int T, float eps,
                     void* stream) {
  if (T <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (T <= 4) return gemv_qkv_rope(res, ln_w, Wqkv, cos_sin, pos, slot, qkv, kv_layer, T, eps, s, nullptr, 0, sc);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_layer_qkv: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, ln_w, h_scratch, T, eps, stream)) return -1;
  return gemm_qkv_rope(h_scratch, Wqkv, qkv, cos_sin, pos, slot, kv_layer, T, s, sc);
}
static int linear_residual(const void* x, const void* W, const float* sc, void* res, int T, int K, void* stream) {
  if (T <= 0) return 0;
  CBX_REQUIRE(K == 1024 || K == 4096, "cbx_linear_residual: K=%d unsupported", K);
  if (T <= 4) return gemv_residual(x, W, res, T, K, (cudaStream_t)stream, nullptr, 0, sc);
  return gemm_residual(x, W, res, T, 1024, K, (cudaStream_t)stream, sc);
}
static int layer_gate_up(const void* res, const void* ln_w, const void* Wgu, const float* sc, void* act, void* h_scratch,
                         int T, float eps, void* stream) {
  if (T <= 0) return 0;
  if (T <= 4) return gemv_norm_gate_up(res, ln_w, Wgu, act, T, eps, (cudaStream_t)stream, nullptr, 0, sc);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_layer_gate_up: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, ln_w, h_scratch, T, eps, stream)) return -1;
  return gemm_bf16_tc(h_scratch, Wgu, act, T, 8192, 1024, 4096, 1, (cudaStream_t)stream, sc);
}
static int head_logits(const void* res, const void* norm_w, const void* Whead, const float* sc, void* logits,
                       void* h_scratch, int T, float eps, void* stream) {
  if (T <= 0) return 0;
  if (T <= 4)
    return gemv_norm_head(res, norm_w, Whead, logits, T, CBX_SPEECH_VOCAB, CBX_LOGITS_LD, eps, (cudaStream_t)stream, sc);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_head_logits: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, norm_w, h_scratch, T, eps, stream)) return -1;
  return gemm_bf16_tc(h_scratch, Whead, logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, (cudaStream_t)stream, sc);
}
extern "C" int cbx_layer_qkv(const void* res, const void* ln_w, const void* Wqkv, const void* cos_sin,
                             const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer, void* h_scratch,
                             int T, float eps, void* stream) {
  return layer_qkv(res, ln_w, Wqkv, nullptr, cos_sin, pos, slot, qkv, kv_layer, h_scratch, T, eps, stream);
}
extern "C" int cbx_linear_residual(const void* x, const void* W, void* res, int T, int K, void* stream) {
  return linear_residual(x, W, nullptr, res, T, K, stream);
}
extern "C" int cbx_linear_partials(const void* x, const void* W, void* partials, int T, int K, int splits, void* stream) {
  if (T <= 0) return 0;
  return gemm_partials(x, W, partials, T, 1024, K, splits, (cudaStream_t)stream);
}
extern "C" int cbx_layer_gate_up(const void* res, const void* ln_w, const void* Wgu, void* act, void* h_scratch, int T,
                                 float eps, void* stream) {
  return layer_gate_up(res, ln_w, Wgu, nullptr, act, h_scratch, T, eps, stream);
}
extern "C" int cbx_head_logits(const void* res, const void* norm_w, const void* Whead, void* logits, void* h_scratch,
                               int T, float eps, void* stream) {
  return head_logits(res, norm_w, Whead, nullptr, logits, h_scratch, T, eps, stream);
}
// ---- the same pieces over FP8-e4m3 weights (W8 = e4m3 bytes, wscale = fp32 per output feature, padded to a
// multiple of 128 entries; gate/up scales in PACKED row order)
extern "C" int cbx_layer_qkv_fp8w(const void* res, const void* ln_w, const void* Wqkv8, const float* wscale,
                                  const void* cos_sin, const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer,
                                  void* h_scratch, int T, float eps, void* stream) {
  CBX_REQUIRE(wscale != nullptr, "cbx_layer_qkv_fp8w: NULL scale");
  return layer_qkv(res, ln_w, Wqkv8, wscale, cos_sin, pos, slot, qkv, kv_layer, h_scratch, T, eps, stream);
}
extern "C" int cbx_linear_residual_fp8w(const void* x, const void* W8, const float* wscale, void* res, int T, int K,
                                        void* stream) {
  CBX_REQUIRE(wscale != nullptr, "cbx_linear_residual_fp8w: NULL scale");
  return linear_residual(x, W8, wscale, res, T, K, stream);
}
extern "C" int cbx_linear_partials_fp8w(const void* x, const void* W8, const float* wscale, void* partials, int T, int K,
                                        int splits, void* stream) {
  CBX_REQUIRE(wscale != nullptr, "cbx_linear_partials_fp8w: NULL scale");
  if (T <= 0) return 0;
  return gemm_partials(x, W8, partials, T, 1024, K, splits, (cudaStream_t)stream, wscale);
}
extern "C" int cbx_layer_gate_up_fp8w(const void* res, const void* ln_w, const void* Wgu8, const float* wscale, void* act,
                                      void* h_scratch, int T, float eps, void* stream) {
  CBX_REQUIRE(wscale != nullptr, "cbx_layer_gate_up_fp8w: NULL scale");
  return layer_gate_up(res, ln_w, Wgu8, wscale, act, h_scratch, T, eps, stream);
}
extern "C" int cbx_head_logits_fp8w(const void* res, const void* norm_w, const void* Whead8, const float* wscale,
                                    void* logits, void* h_scratch, int T, float eps, void* stream) {
  CBX_REQUIRE(wscale != nullptr, "cbx_head_logits_fp8w: NULL scale");
  return head_logits(res, norm_w, Whead8, wscale, logits, h_scratch, T, eps, stream);
}
static void drop_graphs(cbx_engine* e);
// Transposed bf16 down-proj weights (WdT[l] = Wd[l]^T, [4096,1024] row-major) for the fused MLP block of the T <= 4 path
// (k_mlp.cu): a CTA that owns a slice of MLP features reads its rows of WdT contiguously.  Without them (or with
// CBX_MLP_FUSED=0) the gate_up and down-proj kernels run.  NULL / 0 layers removes them again.
extern "C" int cbx_engine_set_down_t(cbx_engine* e, const void* const* wd_t, int n_layers) {
  CBX_REQUIRE(e != nullptr, "cbx_engine_set_down_t: NULL engine");
  CBX_REQUIRE(wd_t == nullptr || n_layers == e->w.n_layers, "cbx_engine_set_down_t: %d layers, engine has %d", n_layers, e->w.n_layers);
  if (wd_t)
    for (int i = 0; i < n_layers; ++i) CBX_REQUIRE(wd_t[i] != nullptr, "cbx_engine_set_down_t: NULL pointer in layer %d", i);
  if (wd_t) e->wd_t.assign(wd_t, wd_t + n_layers); else e->wd_t.clear();
  drop_graphs(e);
  return 0;
}
// Switch a created engine to FP8-e4m3 projection weights (cbx_quantize_fp8_rows output).  Embeddings, norms,
// KV cache, activations and the sampler are unchanged (bf16 / fp32).  Graphs captured before the switch are dropped.
extern "C" int cbx_engine_set_fp8_weights(cbx_engine* e, const cbx_fp8_weights* f) {
  CBX_REQUIRE(e != nullptr && f != nullptr, "cbx_engine_set_fp8_weights: NULL argument");
  CBX_REQUIRE(f->n_layers == e->w.n_layers, "cbx_engine_set_fp8_weights: %d layers, engine has %d", f->n_layers, e->w.n_layers);
  const int L = f->n_layers;
  for (int i = 0; i < L; ++i)
    CBX_REQUIRE(f->wqkv[i] && f->wo[i] && f->wgu[i] && f->wd[i] && f->sqkv[i] && f->so[i] && f->sgu[i] && f->sd[i],
                "cbx_engine_set_fp8_weights: NULL pointer in layer %d", i);
  CBX_REQUIRE(f->speech_head && f->shead, "cbx_engine_set_fp8_weights: NULL head");
  e->wqkv.assign(f->wqkv, f->wqkv + L); e->wo.assign(f->wo, f->wo + L); e->wgu.assign(f->wgu, f->wgu + L);
  e->wd.assign(f->wd, f->wd + L);
  e->sqkv.assign(f->sqkv, f->sqkv + L); e->so.assign(f->so, f->so + L); e->sgu.assign(f->sgu, f->sgu + L);
  e->sd.assign(f->sd, f->sd + L);
  e->head = f->speech_head; e->shead = f->shead;
  e->fp8 = true;
  e->mega = false;                                     // the persistent-kernel experiment is bf16 only
  e->l2_prefetch = false;
  drop_graphs(e);
  return 0;
}

// Rows enter in buffer "res" (the residual stream starts as the embeddings, llama.py:316-320).
// Every projection that closes a residual branch adds into res in its epilogue, so a layer is
//   T <= 4 (GEMV regime, 5 kernels): [norm+QKV+RoPE+KV-append] [attention] [O+res] [norm+gate_up+SiLU] [down+res]
//   T  > 4 (tcgen05 regime, 7)     : [norm] [QKV+RoPE+KV-append] [attention] [O+res] [norm] [gate_up+SiLU] [down+res]
// Rows at or below which the weight-streaming GEMV kernels run instead of the tcgen05 GEMM (they exist for 1, 2, 4 and 8
// padded rows; the fused attention block only up to 4).  Measured (tools/bench_bsweep.py, p50 step): 3 pairs 1.157 ms on
// the GEMV kernels against 1.244 ms on the tcgen05 GEMMs, 4 pairs 1.216 against 1.244 — so the switch is at 8 rows.
// env CBX_GEMV_MAX_T=0 forces the tcgen05 path everywhere, =4 the round-1 switch point.
static int gemv_max_t() {
  static const int v = getenv("CBX_GEMV_MAX_T") ? atoi(getenv("CBX_GEMV_MAX_T")) : 8;
  return v < 0 ? 0 : (v > 8 ? 8 : v);
}

#ifdef CBX_EXPERIMENTS   // `make EXPERIMENTS=1` + env CBX_DUAL_MIN_ROWS=128: measured SLOWER (M1 3.46 -> 3.60-3.82 ms/step), see below
// ---- two-stream decode step (tcgen05 regime) -----------------------------------------------------------------------
// The cond rows [0, B) and the uncond rows [B, 2B) of a decode step do not meet before the sampler's CFG merge
// (t3.py:732), and the step is two very different kinds of work: paged attention (HBM-bound: 67 % of the M1 step at
// 95 % of the copy roofline) and a chain of six short GEMM / norm kernels per layer (latency-bound: 6 us each for
// 0.7 us of tensor work, ~12 % of the SMs' warps active).  Run as ONE stream they alternate, and the HBM idles
// through a third of the step.  Here each half runs its own layer loop on its own stream, and the attention launches
// take `dual_attn_ctas` CTAs (one per SM: the kernel's page ring fills an SM's shared memory) instead of one per SM of
// the whole GPU, so the other half's GEMM CTAs find free SMs while a half streams its KV pages.  The halves fall out of
// phase only if forced to (forward_dual chains the attention launches by events).  Weights are read twice per step
// (1 GB more, against 14 GB of KV at M1).  Same kernels, same per-row arithmetic (the parity tests pass on it).
// MEASURED on B200, M1, ms per step: one stream 3.46; two streams unchained 3.68 (104 CTAs) / 3.50 (120) / 3.53 (148);
// chained 3.82 (88 CTAs) / 3.73 (96) / 3.65 (112) / 3.63 (120) / 3.60 (132) / 3.69 (148).  The premise was wrong: the
// attention kernel's throughput is per SM (41 GB/s each: the QK / PV arithmetic on the CUDA cores keeps pace with an
// SM's 1/148 share of the HBM stream and no more), so it needs every SM to reach 95 % of the roofline and there are no
// SMs to lend.  Default off (CBX_DUAL_MIN_ROWS=0).
static int dual_attn_ctas() {
  static const int v = getenv("CBX_DUAL_ATTN_CTAS") ? atoi(getenv("CBX_DUAL_ATTN_CTAS")) : 104;
  return v;
}
static bool dual_on(const cbx_engine* e, int T) {
  static const int min_rows = getenv("CBX_DUAL_MIN_ROWS") ? atoi(getenv("CBX_DUAL_MIN_ROWS")) : 0;   // 0 = off (default)
  return e->s2 != nullptr && min_rows > 0 && T >= min_rows && T % 2 == 0 && T / 2 <= 512 && !e->trace && !e->skip && e->n_prefill_items == 0 &&
         !e->fp8 && dual_attn_ctas() > 0;
}
// One half's rows and buffers.
struct HalfCtx {
  int off, H, splits;
  cudaStream_t s;
  void* scratch;
  bf16 *res, *h, *qkv, *attn, *act;
  float* part;
  const int32_t *pos, *kv_len, *row_seq;
  const int64_t* slot;
  int32_t* counters;
  bool have_plan, pending;
};
static HalfCtx half_ctx(cbx_engine* e, int off, int H, cudaStream_t s, void* scratch) {
  HalfCtx c = {};
  const int m_tiles = (H + 127) / 128;
  c.off = off; c.H = H; c.s = s; c.scratch = scratch;
  c.splits = H <= e->split_max_rows ? (m_tiles <= 2 ? 4 : (m_tiles <= 4 ? 2 : 1)) : 1;
  c.res = e->res + (size_t)off * 1024; c.h = e->h + (size_t)off * 1024; c.qkv = e->qkv + (size_t)off * 3072;
  c.attn = e->attn + (size_t)off * 1024; c.act = e->act + (size_t)off * 4096;
  c.part = e->part + (off ? (size_t)4 * 256 * 1024 : 0);             // [splits][H][1024] per half: 4 x 256 or 2 x 512 rows
  c.pos = e->pos + off; c.kv_len = e->kv_len + off; c.row_seq = e->row_seq + off; c.slot = e->slot + off;
  c.counters = e->attn_counters + off;
  return c;
}
// [add+norm] [QKV+RoPE+append] of layer li: everything of the layer in front of its attention
static int half_pre(cbx_engine* e, HalfCtx& c, int li, uint8_t* kv) {
  const float eps = e->w.rms_eps;
  if (c.pending ? cbx_rmsnorm_reduce(c.part, c.splits, c.res, e->ln1[li], c.h, c.H, eps, c.s)
                : cbx_rmsnorm(c.res, nullptr, e->ln1[li], c.h, c.H, eps, c.s)) return -1;
  c.pending = false;
  return gemm_qkv_rope(c.h, e->wqkv[li], c.qkv, e->w.cos_sin, c.pos, c.slot, kv, c.H, c.s, nullptr);
}
static int half_attn(cbx_engine* e, HalfCtx& c, uint8_t* kv, int max_kv_len) {
  return paged_attn_decode_ex(c.qkv, kv, e->block_table, e->bt_stride, c.row_seq, c.kv_len, c.attn, c.scratch, c.counters, c.H,
                              max_kv_len, c.have_plan, c.s, dual_attn_ctas());
}
// [O] [add+norm] [gate_up+SiLU] [down]: everything of the layer behind its attention
static int half_post(cbx_engine* e, HalfCtx& c, int li) {
  const float eps = e->w.rms_eps;
  if (c.splits > 1) {
    if (gemm_partials(c.attn, e->wo[li], c.part, c.H, 1024, 1024, c.splits, c.s, nullptr)) return -1;
    if (cbx_rmsnorm_reduce(c.part, c.splits, c.res, e->ln2[li], c.h, c.H, eps, c.s)) return -1;
    if (gemm_bf16_tc(c.h, e->wgu[li], c.act, c.H, 8192, 1024, 4096, 1, c.s, nullptr)) return -1;
    if (gemm_partials(c.act, e->wd[li], c.part, c.H, 1024, 4096, c.splits, c.s, nullptr)) return -1;
    c.pending = true;
    return 0;
  }
  if (gemm_residual(c.attn, e->wo[li], c.res, c.H, 1024, 1024, c.s, nullptr)) return -1;
  if (cbx_rmsnorm(c.res, nullptr, e->ln2[li], c.h, c.H, eps, c.s)) return -1;
  if (gemm_bf16_tc(c.h, e->wgu[li], c.act, c.H, 8192, 1024, 4096, 1, c.s, nullptr)) return -1;
  return gemm_residual(c.act, e->wd[li], c.res, c.H, 1024, 4096, c.s, nullptr);
}
// The two attention launches of a layer are chained by events — cond after the uncond of the layer before, uncond after
// the cond of the same layer — so the HBM-bound kernels run back to back, never against each other, and each half's GEMM
// chain falls into the other half's attention window.  (Left to themselves the halves stay in phase: both GEMM chains
// run together, then both attention launches share the HBM, and nothing is gained — measured.)
static int forward_dual(cbx_engine* e, int T, int max_kv_len, cudaStream_t s) {
  const size_t layer_bytes = (size_t)e->cfg.num_blocks * CBX_PAGE_BYTES;
  const int H = T / 2, L = e->w.n_layers;
  static const bool chain = !(getenv("CBX_DUAL_CHAIN") && getenv("CBX_DUAL_CHAIN")[0] == '0');
  HalfCtx a = half_ctx(e, 0, H, s, e->attn_scratch), b = half_ctx(e, H, H, e->s2, e->attn_scratch2);
  int rc = 0;
  auto ok = [&](cudaError_t ce) { if (ce != cudaSuccess && rc == 0) { set_error("forward_dual: %s", cudaGetErrorString(ce)); rc = -1; } };
  ok(cudaEventRecord(e->ev_fork, s));
  ok(cudaStreamWaitEvent(e->s2, e->ev_fork, 0));
  a.have_plan = attn_plan(a.kv_len, H, a.scratch, a.s, e->block_table, e->bt_stride, a.row_seq, dual_attn_ctas()) == 0;
  b.have_plan = attn_plan(b.kv_len, H, b.scratch, b.s, e->block_table, e->bt_stride, b.row_seq, dual_attn_ctas()) == 0;
  for (int li = 0; li < L && rc == 0; ++li) {
    uint8_t* kv = e->kv_cache + (size_t)li * layer_bytes;
    if (half_pre(e, a, li, kv)) rc = -1;
    if (chain && li > 0) ok(cudaStreamWaitEvent(a.s, e->ev_att[1][li - 1], 0));
    if (half_attn(e, a, kv, max_kv_len)) rc = -1;
    if (chain) ok(cudaEventRecord(e->ev_att[0][li], a.s));
    if (half_post(e, a, li)) rc = -1;
    if (half_pre(e, b, li, kv)) rc = -1;
    if (chain) ok(cudaStreamWaitEvent(b.s, e->ev_att[0][li], 0));
    if (half_attn(e, b, kv, max_kv_len)) rc = -1;
    if (chain) ok(cudaEventRecord(e->ev_att[1][li], b.s));
    if (half_post(e, b, li)) rc = -1;
  }
  // final norm into h (the head GEMM / the sampler's row gather read it)
  for (HalfCtx* c : {&a, &b}) {
    if (rc) break;
    if (c->pending ? cbx_rmsnorm_reduce(c->part, c->splits, c->res, e->w.norm, c->h, H, e->w.rms_eps, c->s)
                   : cbx_rmsnorm(c->res, nullptr, e->w.norm, c->h, H, e->w.rms_eps, c->s)) rc = -1;
  }
  // always join: a capture must not end with the second stream still forked
  const bool joined = cudaEventRecord(e->ev_join, e->s2) == cudaSuccess && cudaStreamWaitEvent(s, e->ev_join, 0) == cudaSuccess;
  CBX_REQUIRE(joined, "forward_dual: join failed: %s", cudaGetErrorString(cudaGetLastError()));
  return rc;
}

#endif  // CBX_EXPERIMENTS

static int forward_layers(cbx_engine* e, int T, int max_kv_len, cudaStream_t s, bool final_norm) {
#ifdef CBX_EXPERIMENTS
  if (T > gemv_max_t() && dual_on(e, T)) return forward_dual(e, T, max_kv_len, s);
#endif
  const size_t layer_bytes = (size_t)e->cfg.num_blocks * CBX_PAGE_BYTES;
  const float eps = e->w.rms_eps;
  const int sk = e->skip;
  const int G = gemv_max_t();                            // rows up to which the GEMV regime runs (4)
  // split-K for the two N = 1024 projections when their 16 column tiles would leave most SMs idle
  const int m_tiles = (T + 127) / 128;
  int splits = 1;
  if (T > G && !e->trace && T <= e->split_max_rows) splits = m_tiles <= 2 ? 4 : (m_tiles <= 4 ? 2 : 1);
  bool pending = false;                                // down-proj partials not yet folded into res
  bool normed = false;                                 // ... or folded already, with h = the NEXT norm's output
  // env CBX_GEMM_FOLD=1 (`make EXPERIMENTS=1` builds only): the split-K GEMMs fold their own partials instead of leaving
  // them to a separate add+RMSNorm launch — bit-identical, measured slower (k_gemm.cu).  Read per call: the test compares
  // the two forms in one process (a captured graph keeps what it was captured with).
  const bool fold_on = getenv("CBX_GEMM_FOLD") && getenv("CBX_GEMM_FOLD")[0] == '1';
  // decode attention cuts the step's pages into equal per-CTA ranges: one plan serves all layers
  static const bool plan_each = getenv("CBX_ATTN_PLAN_EACH") != nullptr;   // debugging aid: rebuild the plan per layer
  const bool have_plan = !plan_each && e->n_prefill_items == 0 && !(sk & 2) && attn_plan(e->kv_len, T, e->attn_scratch, s, e->block_table, e->bt_stride, e->row_seq) == 0;
  const bool pf = e->l2_prefetch;
  for (int li = 0; li < e->w.n_layers; ++li) {
    uint8_t* kv = e->kv_cache + (size_t)li * layer_bytes;
    // T <= 4, bf16 weights, context within the fused block's page buffer: the attention half of the layer is ONE kernel
    // (k_block.cu): [norm+QKV+RoPE+append+attention+O+res] [norm+gate_up+SiLU] [down+res] = 3 launches per layer
    const bool fused = T <= G && !e->fp8 && !sk && !pf && e->n_prefill_items == 0 && layer_attn_block_supported(T, max_kv_len);
    if (fused && !e->trace && !e->wd_t.empty() && layer_mlp_block_on()) {
      // (env CBX_MLP_FUSED=1; measured equal to the three-launch layer below, which stays the default)
      // TWO launches per layer: [attention block] [MLP block], each closing the OTHER one's residual branch in its norm
      // prologue from a fixed-point accumulator.  Rows alternate between res (entering a layer) and h (between the two
      // blocks): the kernel that closes a branch writes the rows to the buffer its sibling CTAs are not reading.
      // Accumulators alternate with the layer's parity; each kernel clears, after its wait, the one that the NEXT
      // kernel of its kind will add into (its last reader finished two launches ago).  The last layer ends with the
      // gate_up / down kernels, which leave the residual closed in res and all four accumulators zero.
      unsigned long long* a_o = e->block_acc + (size_t)(li & 1) * (kAccBytes / 8);
      unsigned long long* a_d = e->block_acc + (size_t)(2 + (li & 1) * kMlpAccCopies) * (kAccBytes / 8);
      const unsigned long long* a_d_prev = li > 0 ? e->block_acc + (size_t)(2 + ((li - 1) & 1) * kMlpAccCopies) * (kAccBytes / 8) : nullptr;
      unsigned long long* a_o_next = e->block_acc + (size_t)((li + 1) & 1) * (kAccBytes / 8);
      // (each kernel pulling the other one's weights into L2 behind its own stream was tried: 0.517 -> 0.533 ms at b = 1)
      const bool last = li + 1 == e->w.n_layers;
      if (layer_attn_block_fold(e->res, a_d_prev, e->h, a_d, kAccBytes * kMlpAccCopies, e->ln1[li], e->wqkv[li], e->wo[li], e->w.cos_sin, e->pos,
                                e->slot, kv, e->block_table, e->bt_stride, e->row_seq, e->kv_len, a_o, T, eps, s)) return -1;
      if (!last) {
        if (layer_mlp_block(e->h, a_o, e->res, e->ln2[li], e->wgu[li], e->wd_t[li], a_d, a_o_next, kAccBytes, T, eps, s)) return -1;
      } else {
        if (gemv_norm_gate_up(e->h, e->ln2[li], e->wgu[li], e->act, T, eps, s, nullptr, 0, nullptr, a_o, e->res)) return -1;
        if (gemv_residual(e->act, e->wd[li], e->res, T, 4096, s, nullptr, 0, nullptr, e->block_acc, kAccBytes * kAccSlots, nullptr)) return -1;
      }
      continue;
    }
    if (fused) {
      // the block leaves the O-proj output in block_acc; gate_up's norm prologue closes the residual branch from it
      // (bf16(bf16(o) + res), written by its CTA 0 to buffer h — NOT in place: the other CTAs still read res); the
      // down-proj kernel adds its output to h's rows, writes res, and clears block_acc for the next layer
      if (layer_attn_block(e->res, e->ln1[li], e->wqkv[li], e->wo[li], e->w.cos_sin, e->pos, e->slot, kv, e->block_table,
                           e->bt_stride, e->row_seq, e->kv_len, e->block_acc, nullptr, T, eps, s)) return -1;
      static const bool noacc = getenv("CBX_BLOCK_NOACC") != nullptr;   // profiling aid: gate_up ignores the accumulator (garbage)
      if (gemv_norm_gate_up(e->res, e->ln2[li], e->wgu[li], e->act, T, eps, s, nullptr, 0, e->sgu[li], noacc ? nullptr : e->block_acc, e->h)) return -1;
      if (gemv_residual(e->act, e->wd[li], e->res, T, 4096, s, nullptr, 0, e->sd[li], e->block_acc, kAccBytes, noacc ? nullptr : e->h)) return -1;
      if (e->trace)
        cudaMemcpyAsync(e->trace + (size_t)li * T * CBX_HIDDEN, e->res, (size_t)T * CBX_HIDDEN * 2,
                        cudaMemcpyDeviceToDevice, s);
      continue;
    }
    if (T <= G) {
      // each kernel pulls a LATER kernel's weights into L2 (QKV -> Wo, attention -> Wgu, O -> Wd, gate_up -> next Wqkv)
      if (!(sk & 1) && gemv_qkv_rope(e->res, e->ln1[li], e->wqkv[li], e->w.cos_sin, e->pos, e->slot, e->qkv, kv, T, eps, s,
                                     pf ? e->wo[li] : nullptr, (size_t)1024 * 1024 * 2, e->sqkv[li])) return -1;
      if (pf) attn_set_prefetch(e->wgu[li], (size_t)8192 * 1024 * 2);
    } else {
      if (normed) {}                                   // the previous layer's down-proj GEMM folded AND normalised (ln1 of this layer)
      else if (pending ? cbx_rmsnorm_reduce(e->part, splits, e->res, e->ln1[li], e->h, T, eps, s)
                       : cbx_rmsnorm(e->res, nullptr, e->ln1[li], e->h, T, eps, s)) return -1;
      pending = false; normed = false;
      if (!(sk & 1) && gemm_qkv_rope(e->h, e->wqkv[li], e->qkv, e->w.cos_sin, e->pos, e->slot, kv, T, s, e->sqkv[li])) return -1;
    }
    if (e->n_prefill_items > 0) {
      if (!(sk & 2) && cbx_paged_attn_prefill(e->qkv, kv, e->block_table, e->bt_stride, e->prefill_items, e->n_prefill_items,
                                              e->attn, s)) return -1;
    } else if (!(sk & 2) && paged_attn_decode_ex(e->qkv, kv, e->block_table, e->bt_stride, e->row_seq, e->kv_len, e->attn,
                                                 e->attn_scratch, e->attn_counters, T, max_kv_len, have_plan, s)) return -1;
    if (T <= G) {
      const bool last = li + 1 == e->w.n_layers;
      if (!(sk & 4) && gemv_residual(e->attn, e->wo[li], e->res, T, 1024, s, pf ? e->wd[li] : nullptr, (size_t)1024 * 4096 * 2,
                                     e->so[li])) return -1;
      if (!(sk & 8) && gemv_norm_gate_up(e->res, e->ln2[li], e->wgu[li], e->act, T, eps, s,
                                         pf ? (last ? e->w.speech_head : e->wqkv[li + 1]) : nullptr,
                                         last ? (size_t)CBX_SPEECH_VOCAB * 1024 * 2 : (size_t)3072 * 1024 * 2, e->sgu[li])) return -1;
      if (!(sk & 16) && gemv_residual(e->act, e->wd[li], e->res, T, 4096, s, nullptr, 0, e->sd[li])) return -1;
    } else if (splits > 1) {
      static const bool o_unsplit = getenv("CBX_O_UNSPLIT") && getenv("CBX_O_UNSPLIT")[0] == '1';   // experiment
      static const bool o_split2 = getenv("CBX_O_UNSPLIT") && getenv("CBX_O_UNSPLIT")[0] == '2';    // experiment: 2 K-slices
      if (o_split2 && splits == 4) {
        if (!(sk & 4) && gemm_partials(e->attn, e->wo[li], e->part, T, 1024, 1024, 2, s, e->so[li])) return -1;
        if (cbx_rmsnorm_reduce(e->part, 2, e->res, e->ln2[li], e->h, T, eps, s)) return -1;
      } else if (o_unsplit) {
        if (!(sk & 4) && gemm_residual(e->attn, e->wo[li], e->res, T, 1024, 1024, s, e->so[li])) return -1;
        if (cbx_rmsnorm(e->res, nullptr, e->ln2[li], e->h, T, eps, s)) return -1;
      } else {
        // the split-K GEMM folds its own partials (add + RMSNorm by the producing CTAs behind an arrival counter) when
        // its grid is co-resident; rc 1 = not applicable: the two-kernel form
        int rc = fold_on && !e->fp8 && !sk ? gemm_partials_fold(e->attn, e->wo[li], e->part, T, 1024, splits, s, e->res, e->ln2[li],
                                                                 e->h, eps, e->fold_cnt) : 1;
        if (rc < 0) return -1;
        if (rc == 1) {
          if (!(sk & 4) && gemm_partials(e->attn, e->wo[li], e->part, T, 1024, 1024, splits, s, e->so[li])) return -1;
          if (cbx_rmsnorm_reduce(e->part, splits, e->res, e->ln2[li], e->h, T, eps, s)) return -1;
        }
      }
      if (!(sk & 8) && gemm_bf16_tc(e->h, e->wgu[li], e->act, T, 8192, 1024, 4096, 1, s, e->sgu[li])) return -1;
      {
        const void* ln_next = li + 1 < e->w.n_layers ? e->ln1[li + 1] : e->w.norm;   // what the fold normalises with
        int rc = fold_on && !e->fp8 && !sk ? gemm_partials_fold(e->act, e->wd[li], e->part, T, 4096, splits, s, e->res, ln_next,
                                                                 e->h, eps, e->fold_cnt) : 1;
        if (rc < 0) return -1;
        if (rc == 0) normed = true;
        else {
          if (!(sk & 16) && gemm_partials(e->act, e->wd[li], e->part, T, 1024, 4096, splits, s, e->sd[li])) return -1;
          pending = true;
        }
      }
    } else {
      if (!(sk & 4) && gemm_residual(e->attn, e->wo[li], e->res, T, 1024, 1024, s, e->so[li])) return -1;
      if (cbx_rmsnorm(e->res, nullptr, e->ln2[li], e->h, T, eps, s)) return -1;
      if (!(sk & 8) && gemm_bf16_tc(e->h, e->wgu[li], e->act, T, 8192, 1024, 4096, 1, s, e->sgu[li])) return -1;
      if (!(sk & 16) && gemm_residual(e->act, e->wd[li], e->res, T, 1024, 4096, s, e->sd[li])) return -1;
    }
    if (e->trace)
      cudaMemcpyAsync(e->trace + (size_t)li * T * CBX_HIDDEN, e->res, (size_t)T * CBX_HIDDEN * 2,
                      cudaMemcpyDeviceToDevice, s);
  }
  if (normed) return 0;                                // the last down-proj GEMM folded and applied the final norm
  if (pending) return cbx_rmsnorm_reduce(e->part, splits, e->res, e->w.norm, e->h, T, eps, s);
  if (final_norm || T > G) return cbx_rmsnorm(e->res, nullptr, e->w.norm, e->h, T, eps, s);
  return 0;
}

extern "C" int cbx_engine_forward(cbx_engine* e, int T, int max_kv_len, void* stream) {
  CBX_REQUIRE(T >= 1 && T <= e->cfg.max_rows, "cbx_engine_forward: T=%d out of range", T);
  return forward_layers(e, T, max_kv_len, (cudaStream_t)stream, true);
}

extern "C" int cbx_engine_set_prefill_items(cbx_engine* e, const int32_t* items, int n_items) {
  CBX_REQUIRE(e != nullptr && (n_items == 0 || items != nullptr), "cbx_engine_set_prefill_items: bad arguments");
  if ((n_items > 0) != (e->n_prefill_items > 0)) drop_graphs(e);
  e->prefill_items = items; e->n_prefill_items = n_items;
  return 0;
}

// A captured step bakes in everything forward_layers reads from the engine (trace pointer, split choice, weights):
// whoever changes one of those drops the captured graphs, so a later replay can never touch a freed buffer.
static void drop_graphs(cbx_engine* e) {
  for (auto& g : e->graphs) cudaGraphExecDestroy(g.second.first);
  e->graphs.clear();
}

extern "C" int cbx_engine_set_trace(cbx_engine* e, void* trace) {
  CBX_REQUIRE(e != nullptr, "cbx_engine_set_trace: NULL engine");
  if ((bf16*)trace != e->trace) drop_graphs(e);
  e->trace = (bf16*)trace;
  return 0;
}

static NextStep next_step(cbx_engine* e) {
  NextStep nx = {(const uint4*)e->w.speech_emb, (uint4*)e->res, e->pos, e->slot, e->kv_len, e->row_seq,
                 e->block_table, e->bt_stride};
  return nx;
}

extern "C" int cbx_engine_sample(cbx_engine* e, int B, void* stream) {
  CBX_REQUIRE(B >= 1 && B <= e->cfg.max_pairs, "cbx_engine_sample: B=%d out of range", B);
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t ce = launch_pdl(gather_rows_kernel, dim3(2 * B), dim3(128), 0, s, (const uint4*)e->h, (const int32_t*)e->head_rows,
                              (uint4*)e->hs);
  CBX_REQUIRE(ce == cudaSuccess, "gather_rows launch failed: %s", cudaGetErrorString(ce));
  count_launch();
  const int T = 2 * B;
  if (T <= 4) { if (gemv_bf16(e->hs, e->head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s, e->shead)) return -1; }
  else if (gemm_bf16_tc(e->hs, e->head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s, e->shead)) return -1;
  // the sampler also stages the next decode step's embeddings / positions / slots
  return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s);
}

// Admission into a running batch of B pair slots: head + sampler for the n_sub new pairs only.
// head_rows[0 .. 2*n_sub) index buffer "h" (cond rows of the new pairs, then their uncond rows);
// pair_map[i] is the slot of the i-th new pair.  Other slots' sampler state and staged inputs are untouched.
extern "C" int cbx_engine_sample_subset(cbx_engine* e, int B, int n_sub, const int32_t* pair_map, void* stream) {
  CBX_REQUIRE(B >= 1 && B <= e->cfg.max_pairs && n_sub >= 1 && n_sub <= B && pair_map, "cbx_engine_sample_subset: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int T = 2 * n_sub;
  cudaError_t ce = launch_pdl(gather_rows_kernel, dim3(T), dim3(128), 0, s, (const uint4*)e->h, (const int32_t*)e->head_rows,
                              (uint4*)e->hs);
  CBX_REQUIRE(ce == cudaSuccess, "gather_rows launch failed: %s", cudaGetErrorString(ce));
  count_launch();
  if (T <= 4) { if (gemv_bf16(e->hs, e->head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s, e->shead)) return -1; }
  else if (gemm_bf16_tc(e->hs, e->head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s, e->shead)) return -1;
  return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s, pair_map, n_sub);
}

// One decode step for pair slots [0,B): inputs were staged by the previous sampler launch.
static int decode_step(cbx_engine* e, int B, int max_kv_len, cudaStream_t s) {
  const int T = 2 * B;
  if (T <= gemv_max_t() && e->mega && e->n_prefill_items == 0 && !e->skip) {
    MegaArgs a = {};
    for (int li = 0; li < e->w.n_layers; ++li) {
      a.ln1[li] = (const bf16*)e->ln1[li]; a.wqkv[li] = (const bf16*)e->wqkv[li]; a.wo[li] = (const bf16*)e->wo[li];
      a.ln2[li] = (const bf16*)e->ln2[li]; a.wgu[li] = (const bf16*)e->wgu[li]; a.wd[li] = (const bf16*)e->wd[li];
    }
    a.norm_w = (const bf16*)e->w.norm; a.head_w = (const bf16*)e->w.speech_head;
    a.res_in = e->res; a.logits = e->logits; a.ws = e->mega_ws; a.sync = e->mega_sync;
    a.trace = e->trace; a.cos_sin = (const bf16*)e->w.cos_sin; a.pos = e->pos; a.slot = e->slot; a.row_seq = e->row_seq;
    a.kv_len = e->kv_len; a.bt = e->block_table; a.kv_cache = e->kv_cache;
    a.layer_bytes = (size_t)e->cfg.num_blocks * CBX_PAGE_BYTES;
    a.bt_stride = e->bt_stride; a.T = T; a.n_layers = e->w.n_layers; a.vocab = CBX_SPEECH_VOCAB; a.ldl = CBX_LOGITS_LD;
    a.eps = e->w.rms_eps;
    static const int mega_flags = getenv("CBX_MEGA_FLAGS") ? atoi(getenv("CBX_MEGA_FLAGS")) : 0;   // debugging aid
    a.flags = mega_flags;
    const int rc = launch_decode_mega(a, s);
    if (rc < 0) return -1;
    if (rc == 0) return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s);
  }
  if (forward_layers(e, T, max_kv_len, s, false)) return -1;
  if (T <= gemv_max_t()) {                              // final norm fused into the head GEMV
    if (gemv_norm_head(e->res, e->w.norm, e->head, e->logits, T, CBX_SPEECH_VOCAB, CBX_LOGITS_LD,
                       e->w.rms_eps, s, e->shead)) return -1;
  } else if (gemm_bf16_tc(e->h, e->head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s, e->shead)) return -1;
  return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s);
}

extern "C" int cbx_engine_decode(cbx_engine* e, int B, int n_steps, int max_kv_len, int use_graph, void* stream) {
  CBX_REQUIRE(B >= 1 && B <= e->cfg.max_pairs && n_steps >= 0, "cbx_engine_decode: bad B/n_steps");
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t l0 = launches();
  int64_t extra = 0;
  if (s == nullptr || s == cudaStreamLegacy) use_graph = 0;   // the legacy default stream cannot be captured: eager launches
  if (!use_graph) {
    for (int i = 0; i < n_steps; ++i) if (decode_step(e, B, max_kv_len, s)) return -1;
  } else {
    // What a captured step bakes in of max_kv_len: for T <= 4 whether the fused block's / the small attention kernel's
    // page buffers hold the context (so: buckets of 64 tokens); for more rows only the unbalanced attention plan — the
    // balanced kernel cuts its page ranges on the device from kv_len, so ONE graph serves every context length and a
    // long generation does not re-capture every 64 steps.
    // At T <= 4 the same holds while the fused attention block takes the step (its page buffer holds up to 1921 tokens at
    // one pair): the graph is captured for the block's longest context and serves every shorter one — a single request
    // no longer pays a capture every 64 tokens.
    static const bool balanced = !(getenv("CBX_ATTN_BALANCED") && getenv("CBX_ATTN_BALANCED")[0] == '0');
    const int T2 = 2 * B;
    const bool any_len = balanced && T2 > 4 && T2 <= 4095;
    static const bool buckets_only = getenv("CBX_GRAPH_BUCKETS") && getenv("CBX_GRAPH_BUCKETS")[0] == '1';   // A/B aid: round-1 keys
    const bool block_len = !buckets_only && T2 <= 4 && T2 <= gemv_max_t() && !e->fp8 && !e->skip && !e->l2_prefetch && !e->mega &&
                           layer_attn_block_supported(T2, max_kv_len);
    const int bucket = any_len ? e->cfg.max_model_len : block_len ? layer_attn_block_max_len(T2) : (max_kv_len + 63) / 64 * 64;
    auto key = std::make_pair(B, bucket);
    auto itr = e->graphs.find(key);
    if (itr == e->graphs.end()) {
      // warm the lazily-initialised function attributes / tensor maps outside capture
      const int64_t c0 = launches();
      CBX_REQUIRE(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess, "begin capture failed");
      const int rc = decode_step(e, B, bucket, s);
      cudaGraph_t g = nullptr;
      cudaError_t ce = cudaStreamEndCapture(s, &g);
      if (rc != 0 || ce != cudaSuccess) {
        if (g) cudaGraphDestroy(g);                    // the partial capture is not kept
        CBX_REQUIRE(rc == 0, "decode_step failed during capture: %s", cbx_last_error());
        CBX_REQUIRE(false, "end capture: %s", cudaGetErrorString(ce));
      }
      const int64_t nodes = launches() - c0;
      cudaGraphExec_t ex;
      ce = cudaGraphInstantiate(&ex, g, 0);
      cudaGraphDestroy(g);
      CBX_REQUIRE(ce == cudaSuccess, "graph instantiate: %s", cudaGetErrorString(ce));
      itr = e->graphs.emplace(key, std::make_pair(ex, nodes)).first;
      extra -= nodes;                                  // captured, not executed
    }
    for (int i = 0; i < n_steps; ++i) {
      cudaError_t ce = cudaGraphLaunch(itr->second.first, s);
      CBX_REQUIRE(ce == cudaSuccess, "graph launch: %s", cudaGetErrorString(ce));
    }
    extra += itr->second.second * n_steps;
  }
  e->last_launches = launches() - l0 + extra;
  return 0;
}

extern "C" int64_t cbx_engine_last_launch_count(const cbx_engine* e) { return e->last_launches; }
