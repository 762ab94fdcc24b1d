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
int gemv_bf16(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
int gemv_qkv_rope(const void*, const void*, const void*, const void*, const int32_t*, const int64_t*, void*, void*, int,
                  float, cudaStream_t);
int gemv_residual(const void*, const void*, void*, int, int, cudaStream_t);
int gemv_norm_gate_up(const void*, const void*, const void*, void*, int, float, cudaStream_t);
int gemv_norm_head(const void*, const void*, const void*, void*, int, int, int, float, cudaStream_t);
int gemm_bf16_tc(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
int gemm_qkv_rope(const void*, const void*, void*, const void*, const int32_t*, const int64_t*, void*, int, cudaStream_t);
int gemm_residual(const void*, const void*, void*, int, int, int, cudaStream_t);
int launch_decode_prepare(const void*, const int32_t*, const int32_t*, const int32_t*, const int32_t*, int, void*,
                          int32_t*, int64_t*, int32_t*, int32_t*, int32_t*, int, cudaStream_t);
int launch_cfg_sample(const void*, const cbx_sampler_state*, int, const NextStep&, cudaStream_t);
int paged_attn_decode_ex(const void*, const void*, const int32_t*, int, const int32_t*, const int32_t*, void*, void*,
                         int32_t*, int, int, int, void*);
}  // namespace cbx
using namespace cbx;

namespace {
__global__ void __launch_bounds__(128) gather_rows_kernel(const uint4* __restrict__ src,
                                                          const int32_t* __restrict__ rows,
                                                          uint4* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  dst[(size_t)blockIdx.x * 128 + threadIdx.x] = src[(size_t)rows[blockIdx.x] * 128 + threadIdx.x];
}
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
}  // namespace

struct cbx_engine {
  cbx_engine_cfg cfg;
  cbx_weights w;
  std::vector<const void*> wqkv, wo, wgu, wd, ln1, ln2;
  uint8_t* kv_cache;
  int32_t* block_table;
  int bt_stride;
  cbx_sampler_state st;
  // workspace carve-up
  bf16 *res, *h, *qkv, *attn, *act, *hs, *logits;
  int32_t *pos, *kv_len, *row_seq, *head_rows, *attn_counters;
  int64_t* slot;
  void* attn_scratch;
  // two-microbatch overlap (decode, T >= overlap_min_rows): the halves run on two streams so one
  // half's latency-bound GEMMs hide under the other half's HBM-bound attention
  void* attn_scratch2; int32_t* attn_counters2;
  cudaStream_t aux = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int overlap_min_rows = 64, overlap_attn_ctas = 104;
  int skip = 0;   // profiling aid (env CBX_SKIP bitmask: 1 qkv, 2 attn, 4 o, 8 gate_up, 16 down): results are garbage
  std::map<std::pair<int, int>, std::pair<cudaGraphExec_t, int64_t>> graphs;   // (B, kv bucket) -> exec, nodes
  int64_t last_launches;
  bf16* trace = nullptr;
  int trace_rows = 0;
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
  p = take(B2 * 4); if (e) e->attn_counters2 = (int32_t*)p;
  p = take((B2 + 1024) * (CBX_HIDDEN * 4 + CBX_HEADS * 8) + 256); if (e) e->attn_scratch2 = p;
  return off;
}

extern "C" int64_t cbx_engine_workspace_bytes(const cbx_engine_cfg* cfg) { return (int64_t)carve(cfg, nullptr, nullptr); }

extern "C" cbx_engine* cbx_engine_create(const cbx_engine_cfg* cfg, const cbx_weights* w, void* kv_cache,
                                         int32_t* block_table, void* workspace, const cbx_sampler_state* st) {
  if (!cfg || !w || !kv_cache || !block_table || !workspace || !st) { set_error("cbx_engine_create: NULL argument"); return nullptr; }
  if (cfg->max_rows < 2 * cfg->max_pairs || cfg->max_pairs < 1) { set_error("cbx_engine_create: max_rows < 2*max_pairs"); return nullptr; }
  if ((reinterpret_cast<uintptr_t>(workspace) & 1023) != 0) { set_error("cbx_engine_create: workspace not 1024-aligned"); return nullptr; }
  if (init_gemm() || init_attn() || init_sample()) return nullptr;
  cbx_engine* e = new cbx_engine();
  e->cfg = *cfg;
  e->w = *w;
  const int L = w->n_layers;
  e->wqkv.assign(w->wqkv, w->wqkv + L); e->wo.assign(w->wo, w->wo + L); e->wgu.assign(w->wgu, w->wgu + L);
  e->wd.assign(w->wd, w->wd + L); e->ln1.assign(w->ln1, w->ln1 + L); e->ln2.assign(w->ln2, w->ln2 + L);
  e->kv_cache = (uint8_t*)kv_cache;
  e->block_table = block_table;
  e->bt_stride = (cfg->max_model_len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  e->st = *st;
  e->last_launches = 0;
  carve(cfg, e, (uint8_t*)workspace);
  // split-K attention tickets start at zero and reset themselves
  cudaMemset(e->attn_counters, 0, (size_t)cfg->max_rows * 4);
  cudaMemset(e->attn_counters2, 0, (size_t)2 * cfg->max_pairs * 4);
  if (const char* v = getenv("CBX_SKIP")) e->skip = atoi(v);
  if (const char* v = getenv("CBX_OVERLAP_MIN_ROWS")) e->overlap_min_rows = atoi(v);      // 0 disables
  if (const char* v = getenv("CBX_OVERLAP_ATTN_CTAS")) e->overlap_attn_ctas = atoi(v);
  if (cudaStreamCreateWithFlags(&e->aux, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    set_error("cbx_engine_create: stream/event creation failed");
    delete e;
    return nullptr;
  }
  return e;
}

extern "C" void cbx_engine_destroy(cbx_engine* e) {
  if (!e) return;
  for (auto& g : e->graphs) cudaGraphExecDestroy(g.second.first);
  if (e->aux) cudaStreamDestroy(e->aux);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  delete e;
}

extern "C" void* cbx_engine_buffer(cbx_engine* e, const char* name) {
  struct { const char* n; void* p; } tab[] = {
      {"x", e->res}, {"res", e->res}, {"h", e->h}, {"qkv", e->qkv}, {"attn", e->attn}, {"act", e->act},
      {"hs", e->hs}, {"logits", e->logits}, {"pos", e->pos}, {"slot", e->slot}, {"kv_len", e->kv_len},
      {"row_seq", e->row_seq}, {"head_rows", e->head_rows}};
  for (auto& t : tab) if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

// ---- fused layer pieces (C-ABI, regime chosen by T) -----------------------------------------
extern "C" int cbx_layer_qkv(const void* res, const void* ln_w, const void* Wqkv, const void* cos_sin,
                             const int32_t* pos, const int64_t* slot, void* qkv, void* kv_layer, void* h_scratch,
                             int T, float eps, void* stream) {
  if (T <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (T <= 4) return gemv_qkv_rope(res, ln_w, Wqkv, cos_sin, pos, slot, qkv, kv_layer, T, eps, s);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_layer_qkv: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, ln_w, h_scratch, T, eps, stream)) return -1;
  return gemm_qkv_rope(h_scratch, Wqkv, qkv, cos_sin, pos, slot, kv_layer, T, s);
}
extern "C" int cbx_linear_residual(const void* x, const void* W, void* res, int T, int K, void* stream) {
  if (T <= 0) return 0;
  CBX_REQUIRE(K == 1024 || K == 4096, "cbx_linear_residual: K=%d unsupported", K);
  if (T <= 4) return gemv_residual(x, W, res, T, K, (cudaStream_t)stream);
  return gemm_residual(x, W, res, T, 1024, K, (cudaStream_t)stream);
}
extern "C" int cbx_layer_gate_up(const void* res, const void* ln_w, const void* Wgu, void* act, void* h_scratch, int T,
                                 float eps, void* stream) {
  if (T <= 0) return 0;
  if (T <= 4) return gemv_norm_gate_up(res, ln_w, Wgu, act, T, eps, (cudaStream_t)stream);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_layer_gate_up: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, ln_w, h_scratch, T, eps, stream)) return -1;
  return gemm_bf16_tc(h_scratch, Wgu, act, T, 8192, 1024, 4096, 1, (cudaStream_t)stream);
}
extern "C" int cbx_head_logits(const void* res, const void* norm_w, const void* Whead, void* logits, void* h_scratch,
                               int T, float eps, void* stream) {
  if (T <= 0) return 0;
  if (T <= 4)
    return gemv_norm_head(res, norm_w, Whead, logits, T, CBX_SPEECH_VOCAB, CBX_LOGITS_LD, eps, (cudaStream_t)stream);
  CBX_REQUIRE(h_scratch != nullptr, "cbx_head_logits: h_scratch is NULL for T > 4");
  if (cbx_rmsnorm(res, nullptr, norm_w, h_scratch, T, eps, stream)) return -1;
  return gemm_bf16_tc(h_scratch, Whead, logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, (cudaStream_t)stream);
}

// Rows enter in buffer "res" (the residual stream starts as the embeddings, llama.py:316-320).
// Every projection that closes a residual branch adds into res in its epilogue, so a layer is
//   T <= 4 (GEMV regime, 5 kernels): [norm+QKV+RoPE+KV-append] [attention] [O+res] [norm+gate_up+SiLU] [down+res]
//   T  > 4 (tcgen05 regime, 7)     : [norm] [QKV+RoPE+KV-append] [attention] [O+res] [norm] [gate_up+SiLU] [down+res]
static int forward_layers(cbx_engine* e, int r0, int T, int max_kv_len, cudaStream_t s, int attn_ctas = 148,
                          bool second = false) {
  const size_t layer_bytes = (size_t)e->cfg.num_blocks * CBX_PAGE_BYTES;
  const float eps = e->w.rms_eps;
  bf16 *res = e->res + (size_t)r0 * 1024, *h = e->h + (size_t)r0 * 1024, *qkv = e->qkv + (size_t)r0 * 3072;
  bf16 *attn = e->attn + (size_t)r0 * 1024, *act = e->act + (size_t)r0 * 4096;
  void* scratch = second ? e->attn_scratch2 : e->attn_scratch;
  int32_t* counters = second ? e->attn_counters2 : e->attn_counters;
  for (int li = 0; li < e->w.n_layers; ++li) {
    uint8_t* kv = e->kv_cache + (size_t)li * layer_bytes;
    const int sk = e->skip;
    if (!(sk & 1) && cbx_layer_qkv(res, e->ln1[li], e->wqkv[li], e->w.cos_sin, e->pos + r0, e->slot + r0, qkv, kv, h, T, eps, s)) return -1;
    if (!(sk & 2) && paged_attn_decode_ex(qkv, kv, e->block_table, e->bt_stride, e->row_seq + r0, e->kv_len + r0, attn, scratch,
                             counters, T, max_kv_len, attn_ctas, s)) return -1;
    if (!(sk & 4) && cbx_linear_residual(attn, e->wo[li], res, T, 1024, s)) return -1;
    if (!(sk & 8) && cbx_layer_gate_up(res, e->ln2[li], e->wgu[li], act, h, T, eps, s)) return -1;
    if (!(sk & 16) && cbx_linear_residual(act, e->wd[li], res, T, 4096, s)) return -1;
    if (e->trace)
      cudaMemcpyAsync(e->trace + ((size_t)li * e->trace_rows + r0) * CBX_HIDDEN, res, (size_t)T * CBX_HIDDEN * 2,
                      cudaMemcpyDeviceToDevice, s);
  }
  return 0;
}

extern "C" int cbx_engine_forward(cbx_engine* e, int T, int max_kv_len, void* stream) {
  CBX_REQUIRE(T >= 1 && T <= e->cfg.max_rows, "cbx_engine_forward: T=%d out of range", T);
  cudaStream_t s = (cudaStream_t)stream;
  e->trace_rows = T;
  if (forward_layers(e, 0, T, max_kv_len, s)) return -1;
  return cbx_rmsnorm(e->res, nullptr, e->w.norm, e->h, T, e->w.rms_eps, s);
}

extern "C" int cbx_engine_set_trace(cbx_engine* e, void* trace) {
  CBX_REQUIRE(e != nullptr, "cbx_engine_set_trace: NULL engine");
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
  if (T <= 4) { if (gemv_bf16(e->hs, e->w.speech_head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s)) return -1; }
  else if (gemm_bf16_tc(e->hs, e->w.speech_head, e->logits, T, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s)) return -1;
  // the sampler also stages the next decode step's embeddings / positions / slots
  return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s);
}

// One decode step for pair slots [0,B): inputs were staged by the previous sampler launch.
static int decode_step(cbx_engine* e, int B, int max_kv_len, cudaStream_t s) {
  const int T = 2 * B;
  e->trace_rows = T;
  if (e->overlap_min_rows > 0 && T >= e->overlap_min_rows && !e->trace) {
    // cond rows [0,B) on s, uncond rows [B,2B) on the aux stream; join before the head
    CBX_REQUIRE(cudaEventRecord(e->ev_fork, s) == cudaSuccess && cudaStreamWaitEvent(e->aux, e->ev_fork, 0) == cudaSuccess,
                "decode_step: fork failed");
    if (forward_layers(e, 0, B, max_kv_len, s, e->overlap_attn_ctas, false)) return -1;
    if (forward_layers(e, B, B, max_kv_len, e->aux, e->overlap_attn_ctas, true)) return -1;
    CBX_REQUIRE(cudaEventRecord(e->ev_join, e->aux) == cudaSuccess && cudaStreamWaitEvent(s, e->ev_join, 0) == cudaSuccess,
                "decode_step: join failed");
  } else if (forward_layers(e, 0, T, max_kv_len, s)) return -1;
  if (cbx_head_logits(e->res, e->w.norm, e->w.speech_head, e->logits, e->h, T, e->w.rms_eps, s)) return -1;
  return launch_cfg_sample(e->logits, &e->st, B, next_step(e), s);
}

extern "C" int cbx_engine_decode(cbx_engine* e, int B, int n_steps, int max_kv_len, int use_graph, void* stream) {
  CBX_REQUIRE(B >= 1 && B <= e->cfg.max_pairs && n_steps >= 0, "cbx_engine_decode: bad B/n_steps");
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t l0 = launches();
  int64_t extra = 0;
  if (!use_graph) {
    for (int i = 0; i < n_steps; ++i) if (decode_step(e, B, max_kv_len, s)) return -1;
  } else {
    const int bucket = (max_kv_len + 63) / 64 * 64;
    auto key = std::make_pair(B, bucket);
    auto itr = e->graphs.find(key);
    if (itr == e->graphs.end()) {
      // warm the lazily-initialised function attributes / tensor maps outside capture
      cudaGraph_t g;
      const int64_t c0 = launches();
      CBX_REQUIRE(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess, "begin capture failed");
      const int rc = decode_step(e, B, bucket, s);
      cudaError_t ce = cudaStreamEndCapture(s, &g);
      CBX_REQUIRE(rc == 0, "decode_step failed during capture: %s", cbx_last_error());
      CBX_REQUIRE(ce == cudaSuccess, "end capture: %s", cudaGetErrorString(ce));
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
