// k_cond.cu — the conditioning-prefix producer on the GPU: T3CondEnc =
// speaker Linear(256 -> 1024) ‖ Perceiver(150 prompt-token embeddings -> 32 rows) ‖ emotion Linear(1 -> 1024)
// (reference models/t3/modules/cond_enc.py:57-123, perceiver.py:116-215; tts.py:238-257).
// The reference runs these modules in fp32 (torch, TF32 off) and then ships the [34,1024] result
// through the CPU and a pickle to the vLLM process per request; here it is eight small launches
// that leave cond_emb in device memory, and a new exaggeration value costs one 1024-thread launch.
//
// All arithmetic is fp32 FMA (no tensor cores: TF32 would change the reference's numerics, and the
// whole encoder is ~1 GFLOP against ~30 MB of fp32 weights — it is bound by reading those once).
//   rows_ln      : x = speech_emb[tok] + speech_pos[i]   or the learned query   -> LayerNorm
//   linear_f32   : Y = X·W^T + b (+ residual), up to three problems per launch (q, k, v)
//   attn_f32     : 4 heads x d=256, softmax(q·k^T / 16)·v, one CTA per (query row, head)
//   edge_rows    : row 0 = spkr_enc(speaker_emb), row 33 = emotion_adv_fc(emotion)
#include "common.cuh"

#include <math.h>

using namespace cbx;

namespace {
constexpr int kDim = 1024, kQueries = 32, kHeads = 4, kHeadDim = 256, kSpk = 256;

__device__ __forceinline__ float block_sum_256(float v, float* sh) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += sh[w];
  __syncthreads();
  return t;
}

// One CTA (256 threads x 4 elements) per row.  Rows [0, n_prompt): token embedding + position;
// rows [n_prompt, n_prompt + n_plain): plain rows of `plain` (the query, or the first pass's output).
__global__ void __launch_bounds__(256) rows_ln_kernel(const int32_t* __restrict__ tok, int n_prompt,
                                                      const float* __restrict__ speech_emb, const float* __restrict__ speech_pos,
                                                      const float* __restrict__ plain, const float* __restrict__ ln_w,
                                                      const float* __restrict__ ln_b, float* __restrict__ out_prompt,
                                                      float* __restrict__ out_plain) {
  __shared__ float sh[8];
  const int r = blockIdx.x, c = threadIdx.x * 4;
  float4 x;
  float* dst;
  if (r < n_prompt) {
    const float4 e = *reinterpret_cast<const float4*>(speech_emb + (size_t)tok[r] * kDim + c);
    const float4 p = *reinterpret_cast<const float4*>(speech_pos + (size_t)r * kDim + c);
    x = make_float4(e.x + p.x, e.y + p.y, e.z + p.z, e.w + p.w);
    dst = out_prompt + (size_t)r * kDim + c;
  } else {
    x = *reinterpret_cast<const float4*>(plain + (size_t)(r - n_prompt) * kDim + c);
    dst = out_plain + (size_t)(r - n_prompt) * kDim + c;
  }
  const float mean = block_sum_256(x.x + x.y + x.z + x.w, sh) * (1.f / kDim);
  const float4 d = make_float4(x.x - mean, x.y - mean, x.z - mean, x.w - mean);
  const float var = block_sum_256(d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w, sh) * (1.f / kDim);   // biased, as nn.LayerNorm
  const float rstd = rsqrtf(var + 1e-5f);
  const float4 g = *reinterpret_cast<const float4*>(ln_w + c), b = *reinterpret_cast<const float4*>(ln_b + c);
  *reinterpret_cast<float4*>(dst) = make_float4(d.x * rstd * g.x + b.x, d.y * rstd * g.y + b.y, d.z * rstd * g.z + b.z,
                                                d.w * rstd * g.w + b.w);
}

struct LinProblem { const float* X; const float* W; const float* b; const float* res; float* Y; int M; };
struct LinBatch { LinProblem p[3]; };

// Y[M,1024] = X[M,1024]·W[1024,1024]^T + b (+ res).  32 x 32 output tile per CTA, K in slabs of 32
// through shared memory; thread (ty, tx) owns rows 4ty..4ty+3 of column tx.
__global__ void __launch_bounds__(256) linear_f32_kernel(const LinBatch batch) {
  const LinProblem pr = batch.p[blockIdx.z];
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  if (m0 >= pr.M) return;
  __shared__ float Xs[32][33], Ws[32][33];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int lr = tid >> 3, lc = (tid & 7) * 4;        // loader: row lr, 4 columns from lc
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = 0; k0 < kDim; k0 += 32) {
    float4 xv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m0 + lr < pr.M) xv = *reinterpret_cast<const float4*>(pr.X + (size_t)(m0 + lr) * kDim + k0 + lc);
    const float4 wv = *reinterpret_cast<const float4*>(pr.W + (size_t)(n0 + lr) * kDim + k0 + lc);
    Xs[lr][lc] = xv.x; Xs[lr][lc + 1] = xv.y; Xs[lr][lc + 2] = xv.z; Xs[lr][lc + 3] = xv.w;
    Ws[lr][lc] = wv.x; Ws[lr][lc + 1] = wv.y; Ws[lr][lc + 2] = wv.z; Ws[lr][lc + 3] = wv.w;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const float w = Ws[tx][k];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(Xs[ty * 4 + i][k], w, acc[i]);
    }
    __syncthreads();
  }
  const float bias = pr.b ? pr.b[n0 + tx] : 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m < pr.M) {
      float y = acc[i] + bias;
      if (pr.res) y += pr.res[(size_t)m * kDim + n0 + tx];
      pr.Y[(size_t)m * kDim + n0 + tx] = y;
    }
  }
}

// softmax(q·k^T / sqrt(256))·v for one (query row, head); n_keys <= 1024.
__global__ void __launch_bounds__(256) attn_f32_kernel(const float* __restrict__ Q, const float* __restrict__ K,
                                                       const float* __restrict__ V, int n_keys, float* __restrict__ out) {
  __shared__ float qs[kHeadDim], sc[1024], sh[8];
  const int qi = blockIdx.x, h = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  qs[tid] = Q[(size_t)qi * kDim + h * kHeadDim + tid];
  __syncthreads();
  for (int j = warp; j < n_keys; j += 8) {
    const float* kr = K + (size_t)j * kDim + h * kHeadDim;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s = fmaf(qs[lane + 32 * i], kr[lane + 32 * i], s);
    s = warp_sum(s);
    if (lane == 0) sc[j] = s * 0.0625f;
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int j = tid; j < n_keys; j += 256) mx = fmaxf(mx, sc[j]);
  mx = warp_max(mx);
  if (lane == 0) sh[warp] = mx;
  __syncthreads();
  mx = sh[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) mx = fmaxf(mx, sh[w]);
  __syncthreads();
  float sum = 0.f;
  for (int j = tid; j < n_keys; j += 256) { const float p = expf(sc[j] - mx); sc[j] = p; sum += p; }
  sum = block_sum_256(sum, sh);                        // (its barriers also publish sc[])
  const float inv = 1.f / sum;
  float o = 0.f;
  const float* vc = V + h * kHeadDim + tid;
  for (int j = 0; j < n_keys; ++j) o = fmaf(sc[j], vc[(size_t)j * kDim], o);
  out[(size_t)qi * kDim + h * kHeadDim + tid] = o * inv;
}

// block 0: row = spkr_w[1024,256]·speaker_emb + spkr_b;  block 1: row = emo_w[1024]·emotion
__global__ void __launch_bounds__(1024) edge_rows_kernel(const float* __restrict__ spkr_w, const float* __restrict__ spkr_b,
                                                         const float* __restrict__ speaker_emb, const float* __restrict__ emo_w,
                                                         float emotion, float* __restrict__ row_spk, float* __restrict__ row_emo) {
  if (blockIdx.x == 1 || row_spk == nullptr) {
    if (row_emo) row_emo[threadIdx.x] = emo_w[threadIdx.x] * emotion;
    return;
  }
  __shared__ float ve[kSpk];
  if (threadIdx.x < kSpk) ve[threadIdx.x] = speaker_emb[threadIdx.x];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int n = warp; n < kDim; n += 32) {
    const float* wr = spkr_w + (size_t)n * kSpk;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s = fmaf(wr[lane + 32 * i], ve[lane + 32 * i], s);
    s = warp_sum(s);
    if (lane == 0) row_spk[n] = s + spkr_b[n];
  }
}

int lin(const LinBatch& b, int n_prob, int max_m, cudaStream_t s) {
  linear_f32_kernel<<<dim3(kDim / 32, (max_m + 31) / 32, n_prob), 256, 0, s>>>(b);
  CBX_CHECK_LAUNCH("cond linear");
  return 0;
}
}  // namespace

extern "C" int64_t cbx_cond_workspace_bytes(int max_prompt) {
  // LN(prompt), K, V : max_prompt rows each;  LN(query), Q, A, pre, LN(pre), Q2, K2, V2, A2 : 32 rows each
  return ((int64_t)3 * max_prompt + 9 * kQueries) * kDim * 4;
}

extern "C" int cbx_cond_encode(const cbx_cond_weights* w, const float* speaker_emb, const int32_t* prompt_tokens,
                               int n_prompt, float emotion, void* workspace, float* out, void* stream) {
  CBX_REQUIRE(w && speaker_emb && prompt_tokens && workspace && out, "cbx_cond_encode: NULL argument");
  CBX_REQUIRE(n_prompt >= 1 && n_prompt <= 1024, "cbx_cond_encode: n_prompt=%d not in [1,1024]", n_prompt);
  cudaStream_t s = (cudaStream_t)stream;
  float* f = (float*)workspace;
  const size_t R = (size_t)kQueries * kDim, Pn = (size_t)n_prompt * kDim;
  float *lnp = f, *Kp = lnp + Pn, *Vp = Kp + Pn, *lnq = Vp + Pn, *Q = lnq + R, *A = Q + R, *pre = A + R, *lnpre = pre + R,
        *Q2 = lnpre + R, *K2 = Q2 + R, *V2 = K2 + R, *A2 = V2 + R;
  float* body = out + kDim;                            // rows 1..32
  // pass 1: cross-attention of the learned query over the prompt embeddings (perceiver.py:209-211)
  rows_ln_kernel<<<n_prompt + kQueries, 256, 0, s>>>(prompt_tokens, n_prompt, w->speech_emb, w->speech_pos, w->query, w->ln_w,
                                                     w->ln_b, lnp, lnq);
  CBX_CHECK_LAUNCH("cond rows_ln");
  LinBatch b1 = {{{lnq, w->wq, w->bq, nullptr, Q, kQueries}, {lnp, w->wk, w->bk, nullptr, Kp, n_prompt},
                  {lnp, w->wv, w->bv, nullptr, Vp, n_prompt}}};
  if (lin(b1, 3, n_prompt > kQueries ? n_prompt : kQueries, s)) return -2;
  attn_f32_kernel<<<dim3(kQueries, kHeads), 256, 0, s>>>(Q, Kp, Vp, n_prompt, A);
  CBX_CHECK_LAUNCH("cond attn");
  LinBatch b2 = {{{A, w->wo, w->bo, w->query, pre, kQueries}, {}, {}}};
  if (lin(b2, 1, kQueries, s)) return -2;
  // pass 2: self-attention with the SAME block (perceiver.py:212-214)
  rows_ln_kernel<<<kQueries, 256, 0, s>>>(prompt_tokens, 0, w->speech_emb, w->speech_pos, pre, w->ln_w, w->ln_b, nullptr, lnpre);
  CBX_CHECK_LAUNCH("cond rows_ln");
  LinBatch b3 = {{{lnpre, w->wq, w->bq, nullptr, Q2, kQueries}, {lnpre, w->wk, w->bk, nullptr, K2, kQueries},
                  {lnpre, w->wv, w->bv, nullptr, V2, kQueries}}};
  if (lin(b3, 3, kQueries, s)) return -2;
  attn_f32_kernel<<<dim3(kQueries, kHeads), 256, 0, s>>>(Q2, K2, V2, kQueries, A2);
  CBX_CHECK_LAUNCH("cond attn");
  LinBatch b4 = {{{A2, w->wo, w->bo, pre, body, kQueries}, {}, {}}};
  if (lin(b4, 1, kQueries, s)) return -2;
  // rows 0 and 33 (cond_enc.py:84-85,107-110)
  edge_rows_kernel<<<2, 1024, 0, s>>>(w->spkr_w, w->spkr_b, speaker_emb, w->emo_w, emotion, out, out + (size_t)(1 + kQueries) * kDim);
  CBX_CHECK_LAUNCH("cond edge rows");
  return 0;
}

extern "C" int cbx_cond_emotion_row(const cbx_cond_weights* w, float emotion, float* row, void* stream) {
  CBX_REQUIRE(w && row, "cbx_cond_emotion_row: NULL argument");
  edge_rows_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(nullptr, nullptr, nullptr, w->emo_w, emotion, nullptr, row);
  CBX_CHECK_LAUNCH("cond emotion row");
  return 0;
}
