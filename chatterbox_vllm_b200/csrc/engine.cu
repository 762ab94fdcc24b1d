// engine.cu — host-side orchestration of the T3 step: the launch sequence that replaces
// T3VllmModel.forward / compute_logits (t3.py:725-778) + vLLM's LlamaModel.forward
// (vllm llama.py:394-430) + the sampler, and a CUDA-graph replay loop that replaces vLLM's
// per-step Python scheduler for the decode phase (no host sync per step).
#include <string.h>

#include <map>
#include <vector>

#include "common.cuh"

namespace cbx {
int64_t launches();
int init_gemm();
int init_attn();
int init_sample();
int gemv_bf16(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
int gemm_bf16_tc(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
int launch_decode_prepare(const void*, const int32_t*, const int32_t*, const int32_t*, const int32_t*, int, void*,
                          int32_t*, int64_t*, int32_t*, int32_t*, int32_t*, int, cudaStream_t);
}  // namespace cbx
using namespace cbx;

namespace {
__global__ void __launch_bounds__(128) gather_rows_kernel(const uint4* __restrict__ src,
                                                          const int32_t* __restrict__ rows,
                                                          uint4* __restrict__ dst) {
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
  bf16 *x, *res, *h, *qkv, *attn, *act, *hs, *logits;
  int32_t *pos, *kv_len, *row_seq, *head_rows;
  int64_t* slot;
  void* attn_scratch;
  std::map<std::pair<int, int>, std::pair<cudaGraphExec_t, int64_t>> graphs;   // (B, kv bucket) -> exec, nodes
  int64_t last_launches;
  bf16* trace = nullptr;
};

static size_t carve(const cbx_engine_cfg* c, cbx_engine* e, uint8_t* base) {
  size_t off = 0;
  const size_t R = (size_t)c->max_rows, B2 = (size_t)2 * c->max_pairs;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 1024); return base ? base + o : nullptr; };
  void* p;
  p = take(R * 1024 * 2); if (e) e->x = (bf16*)p;
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
  p = take((R + 1024) * (CBX_HIDDEN * 4 + CBX_HEADS * 8) + 256); if (e) e->attn_scratch = p;
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
  return e;
}

extern "C" void cbx_engine_destroy(cbx_engine* e) {
  if (!e) return;
  for (auto& g : e->graphs) cudaGraphExecDestroy(g.second.first);
  delete e;
}

extern "C" void* cbx_engine_buffer(cbx_engine* e, const char* name) {
  struct { const char* n; void* p; } tab[] = {
      {"x", e->x}, {"res", e->res}, {"h", e->h}, {"qkv", e->qkv}, {"attn", e->attn}, {"act", e->act},
      {"hs", e->hs}, {"logits", e->logits}, {"pos", e->pos}, {"slot", e->slot}, {"kv_len", e->kv_len},
      {"row_seq", e->row_seq}, {"head_rows", e->head_rows}};
  for (auto& t : tab) if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

static int linear(const void* x, const void* W, void* y, int T, int N, int K, int ldy, int epi, cudaStream_t s) {
  if (T <= 4) return gemv_bf16(x, W, y, T, N, K, ldy, epi, s);
  return gemm_bf16_tc(x, W, y, T, N, K, ldy, epi, s);
}

extern "C" int cbx_engine_forward(cbx_engine* e, int T, int max_kv_len, void* stream) {
  CBX_REQUIRE(T >= 1 && T <= e->cfg.max_rows, "cbx_engine_forward: T=%d out of range", T);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t layer_bytes = (size_t)e->cfg.num_blocks * CBX_PAGE_BYTES;
  for (int li = 0; li < e->w.n_layers; ++li) {
    if (cbx_rmsnorm(e->x, li == 0 ? nullptr : e->res, e->ln1[li], e->h, T, e->w.rms_eps, s)) return -1;
    if (li == 0) {    // residual = x for the first layer (vllm llama.py:316-320)
      if (cudaMemcpyAsync(e->res, e->x, (size_t)T * 2048, cudaMemcpyDeviceToDevice, s) != cudaSuccess) { set_error("res copy"); return -2; }
      count_launch();
    }
    if (linear(e->h, e->wqkv[li], e->qkv, T, 3072, 1024, 3072, 0, s)) return -1;
    uint8_t* kv = e->kv_cache + (size_t)li * layer_bytes;
    if (cbx_rope_kv_append(e->qkv, e->pos, e->slot, e->w.cos_sin, kv, T, s)) return -1;
    if (cbx_paged_attn_decode(e->qkv, kv, e->block_table, e->bt_stride, e->row_seq, e->kv_len, e->attn,
                              e->attn_scratch, T, max_kv_len, s)) return -1;
    if (linear(e->attn, e->wo[li], e->x, T, 1024, 1024, 1024, 0, s)) return -1;
    if (cbx_rmsnorm(e->x, e->res, e->ln2[li], e->h, T, e->w.rms_eps, s)) return -1;
    if (linear(e->h, e->wgu[li], e->act, T, 8192, 1024, 4096, 1, s)) return -1;
    if (linear(e->act, e->wd[li], e->x, T, 1024, 4096, 1024, 0, s)) return -1;
    if (e->trace) {
      const size_t n = (size_t)T * CBX_HIDDEN;
      cudaMemcpyAsync(e->trace + (size_t)(2 * li) * n, e->x, n * 2, cudaMemcpyDeviceToDevice, s);
      cudaMemcpyAsync(e->trace + (size_t)(2 * li + 1) * n, e->res, n * 2, cudaMemcpyDeviceToDevice, s);
    }
  }
  return cbx_rmsnorm(e->x, e->res, e->w.norm, e->h, T, e->w.rms_eps, s);
}

extern "C" int cbx_engine_set_trace(cbx_engine* e, void* trace) {
  CBX_REQUIRE(e != nullptr, "cbx_engine_set_trace: NULL engine");
  e->trace = (bf16*)trace;
  return 0;
}

extern "C" int cbx_engine_sample(cbx_engine* e, int B, void* stream) {
  CBX_REQUIRE(B >= 1 && B <= e->cfg.max_pairs, "cbx_engine_sample: B=%d out of range", B);
  cudaStream_t s = (cudaStream_t)stream;
  gather_rows_kernel<<<2 * B, 128, 0, s>>>((const uint4*)e->h, e->head_rows, (uint4*)e->hs);
  CBX_CHECK_LAUNCH("gather_rows");
  if (linear(e->hs, e->w.speech_head, e->logits, 2 * B, CBX_SPEECH_VOCAB, 1024, CBX_LOGITS_LD, 0, s)) return -1;
  return cbx_cfg_sample(e->logits, &e->st, B, s);
}

static int decode_step(cbx_engine* e, int B, int max_kv_len, cudaStream_t s) {
  if (launch_decode_prepare(e->w.speech_emb, e->st.last_tok, e->st.finished, e->st.seq_len, e->block_table,
                            e->bt_stride, e->x, e->pos, e->slot, e->kv_len, e->row_seq, e->head_rows, B, s)) return -1;
  if (cbx_engine_forward(e, 2 * B, max_kv_len, s)) return -1;
  return cbx_engine_sample(e, B, s);
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
