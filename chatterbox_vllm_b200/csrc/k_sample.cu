// k_sample.cu — fused CFG merge + sampler over the 8194-token speech vocabulary.
//
// Replaces t3.py:732 (logits = cond + cfg·(cond − uncond), bf16, one rounding per op), the
// 1000-wide −inf padding of t3.py:734-738 (never materialised here; ids stay in speech space)
// and vLLM's v1 Sampler chain (sampler.py:67-300): repetition penalty over prompt∪output
// (utils.py:51-89) → greedy argmax | temperature → min-p (builtin.py:101-112) → top-k /
// top-p (topk_topp_sampler.py:328-369) → exponential race (…:395-416), plus the scheduler's
// stop / max_tokens check — with all per-request state resident on the device, so the decode
// loop never synchronises with the host.  One CTA (256 threads) per CFG pair.
//
// Specified tie rules (the reference leaves them to sort order): argmax → lowest id; top-p
// boundary ties → lowest ids kept first; top-k keeps every token tied with the k-th value.
#include "common.cuh"

using namespace cbx;
namespace cbx { int init_sample(); }

namespace {
constexpr int V = CBX_SPEECH_VOCAB;
constexpr int kThreads = 256;
constexpr int kPer = (V + kThreads - 1) / kThreads;     // 33 contiguous ids per thread
constexpr float kSamplingEps = 1e-5f;
constexpr int kSampleSmem = (2 * V + 4) * sizeof(float);

__device__ __forceinline__ uint32_t fkey(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct Red {
  float f[8];
  int i[8];
};

__device__ __forceinline__ float block_sum(float v, Red& red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red.f[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += red.f[w];
  __syncthreads();
  return s;
}
// max value, ties -> lowest index
__device__ __forceinline__ void block_argmax(float& v, int& idx, Red& red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
  if ((threadIdx.x & 31) == 0) { red.f[threadIdx.x >> 5] = v; red.i[threadIdx.x >> 5] = idx; }
  __syncthreads();
  v = red.f[0]; idx = red.i[0];
#pragma unroll
  for (int w = 1; w < 8; ++w)
    if (red.f[w] > v || (red.f[w] == v && red.i[w] < idx)) { v = red.f[w]; idx = red.i[w]; }
  __syncthreads();
}

__device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) { return __umulhi(a, b); }
// Philox-4x32-10, returns word 0
__device__ __forceinline__ uint32_t philox_word0(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}

// minimal key k in [0, 2^32) with F(k) < target, F(k) = sum of weight(i) over live ids with key_i > k
template <bool kCount>
__device__ uint32_t bisect(const float* l, const float* e, int i0, int i1, float target, Red& red) {
  uint64_t lo = 0, hi = 0xffffffffull;
  while (lo < hi) {
    const uint32_t mid = (uint32_t)((lo + hi) >> 1);
    float s = 0.f;
    for (int i = i0; i < i1; ++i) {
      const float li = l[i];
      if (fkey(li) > mid && li != -INFINITY) s += kCount ? 1.f : e[i];
    }
    s = block_sum(s, red);
    if (s < target) hi = mid; else lo = (uint64_t)mid + 1;
  }
  return (uint32_t)lo;
}

__global__ void __launch_bounds__(kThreads) cfg_sample_kernel(const bf16* __restrict__ logits,
                                                              cbx_sampler_state st, int B) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  float* l = reinterpret_cast<float*>(smem_raw);          // [V] working logits
  float* e = l + V + 2;                                   // [V] exp(l - max)
  __shared__ Red red;
  __shared__ int s_scan[kThreads];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (st.finished[b]) return;

  const bf16* lc = logits + (size_t)b * CBX_LOGITS_LD;
  const bf16* lu = logits + (size_t)(B + b) * CBX_LOGITS_LD;
  const uint32_t* seen = st.seen + (size_t)b * CBX_SEEN_WORDS;
  const float rep = st.rep_pen[b], cfg = st.cfg_scale;
  const float inv_rep = 1.0f / rep;
  // ---- CFG merge (bf16 op-by-op) + repetition penalty; strided (coalesced) pass
  for (int i = tid; i < V; i += kThreads) {
    const float c = __bfloat162float(lc[i]), u = __bfloat162float(lu[i]);
    float v = round_bf16(c + round_bf16(cfg * round_bf16(c - u)));
    if ((seen[i >> 5] >> (i & 31)) & 1u) v = v * (v > 0.f ? inv_rep : rep);
    l[i] = v;
  }
  __syncthreads();
  const int i0 = tid * kPer, i1 = min(V, i0 + kPer);
  float mx = -INFINITY; int amx = V;
  for (int i = i0; i < i1; ++i) if (l[i] > mx) { mx = l[i]; amx = i; }
  block_argmax(mx, amx, red);

  int tok = amx;
  const float temp = st.temperature[b];
  if (temp >= kSamplingEps) {
    // ---- temperature, softmax numerators
    const float mt = mx / temp;
    float z = 0.f;
    for (int i = i0; i < i1; ++i) {
      const float v = l[i] / temp;
      const float ev = expf(v - mt);
      l[i] = v; e[i] = ev; z += ev;
    }
    z = block_sum(z, red);
    // ---- min-p: drop p_i < p_max * min_p
    const float min_p = st.min_p[b];
    if (min_p > 0.f) {
      const float thr = (1.0f / z) * min_p;
      float z2 = 0.f;
      for (int i = i0; i < i1; ++i) {
        if (e[i] / z < thr) { l[i] = -INFINITY; e[i] = 0.f; }
        z2 += e[i];
      }
      z = block_sum(z2, red);
    }
    // ---- top-k: keep every token >= the k-th largest value
    const int top_k = st.top_k[b];
    if (top_k > 0 && top_k < V) {
      const uint32_t kk = bisect<true>(l, e, i0, i1, (float)top_k, red);
      float z2 = 0.f;
      for (int i = i0; i < i1; ++i) {
        if (fkey(l[i]) < kk) { l[i] = -INFINITY; e[i] = 0.f; }
        z2 += e[i];
      }
      z = block_sum(z2, red);
    }
    // ---- top-p: keep token iff the mass ranked strictly above it is < p
    const float top_p = st.top_p[b];
    if (top_p < 1.0f) {
      const float target = top_p * z;
      const uint32_t kp = bisect<false>(l, e, i0, i1, target, red);
      float g = 0.f, pv = 0.f; int nt = 0;
      for (int i = i0; i < i1; ++i) {
        const uint32_t k = fkey(l[i]);
        if (k > kp && l[i] != -INFINITY) g += e[i];
        else if (k == kp) { ++nt; pv += e[i]; }
      }
      g = block_sum(g, red);
      pv = block_sum(pv, red);                            // = (#ties)·e(v*) ; divide below
      // exclusive scan of tie counts in id order (threads own contiguous id ranges)
      s_scan[tid] = nt;
      __syncthreads();
      int before = 0, total = 0;
      for (int t = 0; t < kThreads; ++t) { const int c = s_scan[t]; if (t < tid) before += c; total += c; }
      __syncthreads();
      const float ev = total > 0 ? pv / (float)total : 0.f;
      int n_keep = total;
      if (ev > 0.f) {
        n_keep = (int)ceilf((target - g) / ev);
        if (n_keep < 1) n_keep = 1;
      }
      float z2 = 0.f;
      int rank = before;
      for (int i = i0; i < i1; ++i) {
        const uint32_t k = fkey(l[i]);
        bool drop = k < kp;
        if (k == kp) { drop = rank >= n_keep; ++rank; }
        if (drop) { l[i] = -INFINITY; e[i] = 0.f; }
        z2 += e[i];
      }
      z = block_sum(z2, red);
    }
    // ---- exponential race over survivors: argmax (e_i / z) / Exp(1)_i
    const uint64_t seed = st.seed[b], uid = st.uid[b];
    const uint32_t step = (uint32_t)st.n_out[b];
    float best = -1.f; int bi = V;
    for (int i = i0; i < i1; ++i) {
      if (e[i] > 0.f) {
        const uint32_t x = philox_word0((uint32_t)i, step, (uint32_t)uid, (uint32_t)(uid >> 32), (uint32_t)seed,
                                        (uint32_t)(seed >> 32));
        const float uu = ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-07f;   // (0,1), exact in fp32
        const float sc = (e[i] / z) / (-logf(uu));
        if (sc > best) { best = sc; bi = i; }
      }
    }
    block_argmax(best, bi, red);
    tok = bi;
  }
  // ---- emit + bookkeeping (one thread)
  if (tid == 0) {
    const int n = st.n_out[b];
    st.out_tokens[(size_t)b * st.out_stride + n] = tok;
    st.n_out[b] = n + 1;
    st.last_tok[b] = tok;
    st.seen[(size_t)b * CBX_SEEN_WORDS + (tok >> 5)] |= 1u << (tok & 31);
    st.seq_len[b] += 1;
    st.seq_len[B + b] += 1;
    if ((tok == st.stop_token && !st.ignore_eos[b]) || n + 1 >= st.max_out[b]) {
      st.finished[b] = 1;
      atomicAdd(st.n_finished, 1);
    }
  }
}
}  // namespace

namespace cbx {
int init_sample() {
  static bool done = false;
  if (done) return 0;
  cudaError_t er = cudaFuncSetAttribute(cfg_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSampleSmem);
  CBX_REQUIRE(er == cudaSuccess, "sampler smem attribute: %s", cudaGetErrorString(er));
  done = true;
  return 0;
}
}  // namespace cbx

extern "C" int cbx_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, void* stream) {
  if (B <= 0) return 0;
  const int smem = kSampleSmem;
  if (cbx::init_sample()) return -1;
  cfg_sample_kernel<<<B, kThreads, smem, (cudaStream_t)stream>>>((const bf16*)logits, *st, B);
  CBX_CHECK_LAUNCH("cfg_sample");
  return 0;
}
