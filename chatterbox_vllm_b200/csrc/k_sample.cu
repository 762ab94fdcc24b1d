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
}  // namespace
namespace cbx {
struct NextStep {             // optional fused "decode_prepare" for the following step (engine only)
  const uint4* speech_emb; uint4* res; int32_t* pos; int64_t* slot; int32_t* kv_len; int32_t* row_seq;
  const int32_t* bt; int bt_stride;
};
}  // namespace cbx
namespace {

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

// Minimal key k in [0, 2^32) with F(k) < target, where F(k) = total weight of live ids whose key
// is > k  (== the key of the pivot element at which the descending cumulative weight reaches
// target).  MSB-first 8-bit radix select: 4 histogram passes instead of a 32-step bisection.
// Weights are integers — 1 for top-k, e_i·2^40 for top-p — so the shared-memory atomics commute and
// the result is bit-reproducible.  Each thread owns a contiguous id range and run-length-merges
// equal digits before touching a bin (the top byte of a logit key is almost constant).
struct Radix {
  uint32_t lo[256], hi[256];          // 64-bit bins as two words: 32-bit shared atomics are native, 64-bit ones CAS-loop
  unsigned long long above, target;
  int digit;
};
__device__ __forceinline__ void bin_add(Radix& rx, int d, unsigned long long w) {
  const uint32_t wl = (uint32_t)w, wh = (uint32_t)(w >> 32);
  const uint32_t old = atomicAdd(&rx.lo[d], wl);
  const uint32_t carry = (old + wl < old) ? 1u : 0u;   // exactly one adder sees each wrap
  if (wh | carry) atomicAdd(&rx.hi[d], wh + carry);
}
__device__ __forceinline__ unsigned long long quant40(float e) { return __float2ull_rz(e * 1099511627776.f); }

// kCount: weights 1 and `target` given; else weights quant40(e[i]) and target = top_p · total.
template <bool kCount>
__device__ uint32_t radix_select(const float* l, const float* e, int i0, int i1, unsigned long long target_in,
                                 float top_p, Radix& rx) {
  uint32_t prefix = 0;
  if (threadIdx.x == 0) { rx.above = 0; rx.target = target_in; }
#pragma unroll 1
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    rx.lo[threadIdx.x] = 0; rx.hi[threadIdx.x] = 0;     // kThreads == 256 bins
    __syncthreads();
    int run_d = -1;
    unsigned long long run_w = 0;
    for (int i = i0; i < i1; ++i) {
      const float li = l[i];
      if (li == -INFINITY) continue;
      const uint32_t key = fkey(li);
      if (pass > 0 && (key >> (shift + 8)) != prefix) continue;
      const int d = (key >> shift) & 255;
      const unsigned long long w = kCount ? 1ull : quant40(e[i]);
      if (d != run_d) {
        if (run_w) bin_add(rx, run_d, run_w);
        run_d = d; run_w = 0;
      }
      run_w += w;
    }
    if (run_w) bin_add(rx, run_d, run_w);
    __syncthreads();
    if (threadIdx.x < 32) {
      // lane owns bins 255-8*lane .. 248-8*lane (descending digits)
      const int lane = threadIdx.x;
      unsigned long long b8[8], tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int d = 255 - 8 * lane - j;
        b8[j] = ((unsigned long long)rx.hi[d] << 32) | rx.lo[d];
        tot += b8[j];
      }
      unsigned long long incl = tot;                   // inclusive scan over lanes
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      unsigned long long target = rx.target;
      if (!kCount && pass == 0) {                      // total weight is known only now
        const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
        target = (unsigned long long)((double)top_p * (double)total);
        if (target < 1) target = 1;
      }
      unsigned long long running = rx.above + incl - tot;
      int found = -1;
      unsigned long long found_above = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (found < 0 && running < target && running + b8[j] >= target) { found = 255 - 8 * lane - j; found_above = running; }
        running += b8[j];
      }
      const unsigned ball = __ballot_sync(0xffffffffu, found >= 0);
      __syncwarp();
      if (ball == 0) {                                 // fewer live weights than the target: keep everything
        if (lane == 0) { rx.digit = 0; rx.above = running; rx.target = target; }
      } else if (lane == __ffs(ball) - 1) {
        rx.digit = found; rx.above = found_above; rx.target = target;
      }
    }
    __syncthreads();
    prefix = (prefix << 8) | (uint32_t)rx.digit;
    __syncthreads();
  }
  return prefix;
}

__global__ void __launch_bounds__(kThreads) cfg_sample_kernel(const bf16* __restrict__ logits,
                                                              cbx_sampler_state st, int B, const NextStep nx) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  float* l = reinterpret_cast<float*>(smem_raw);          // [V] working logits
  float* e = l + V + 2;                                   // [V] exp(l - max)
  __shared__ Red red;
  __shared__ Radix rx;
  __shared__ int s_scan[kThreads];
  const int b = blockIdx.x, tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();
  if (st.finished[b]) return;
  const int n_prev = st.n_out[b];
  const int len_c = st.seq_len[b], len_u = st.seq_len[B + b];

  const bf16* lc = logits + (size_t)b * CBX_LOGITS_LD;
  const bf16* lu = logits + (size_t)(B + b) * CBX_LOGITS_LD;
  const uint32_t* seen = st.seen + (size_t)b * CBX_SEEN_WORDS;
  const float rep = st.rep_pen[b], cfg = st.cfg_scale;
  const float inv_rep = 1.0f / rep;
  // ---- CFG merge (bf16 op-by-op) + repetition penalty: 128-bit loads, all issued before first use
  {
    constexpr int kChunks = V / 8;                       // 1024 full chunks + 2 tail ids
    uint4 vc[4], vu[4];
    uint32_t sb[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = tid + k * kThreads;
      vc[k] = reinterpret_cast<const uint4*>(lc)[c];
      vu[k] = reinterpret_cast<const uint4*>(lu)[c];
      sb[k] = (seen[c >> 2] >> ((c & 3) * 8)) & 0xffu;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = tid + k * kThreads;
      const uint32_t cw[4] = {vc[k].x, vc[k].y, vc[k].z, vc[k].w}, uw[4] = {vu[k].x, vu[k].y, vu[k].z, vu[k].w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float cc = (j & 1) ? bf_hi(cw[j >> 1]) : bf_lo(cw[j >> 1]);
        const float uu = (j & 1) ? bf_hi(uw[j >> 1]) : bf_lo(uw[j >> 1]);
        float v = round_bf16(cc + round_bf16(cfg * round_bf16(cc - uu)));
        if ((sb[k] >> j) & 1u) v = v * (v > 0.f ? inv_rep : rep);
        l[c * 8 + j] = v;
      }
    }
    static_assert(kChunks == 4 * kThreads, "chunk loop assumes 1024 chunks");
    if (tid < V - kChunks * 8) {
      const int i = kChunks * 8 + tid;
      const float cc = __bfloat162float(lc[i]), uu = __bfloat162float(lu[i]);
      float v = round_bf16(cc + round_bf16(cfg * round_bf16(cc - uu)));
      if ((seen[i >> 5] >> (i & 31)) & 1u) v = v * (v > 0.f ? inv_rep : rep);
      l[i] = v;
    }
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
      const uint32_t kk = radix_select<true>(l, e, i0, i1, (unsigned long long)top_k, 0.f, rx);
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
      const uint32_t kp = radix_select<false>(l, e, i0, i1, 0ull, top_p, rx);
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
    const uint32_t step = (uint32_t)n_prev;
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
  const bool fin = (tok == st.stop_token && !st.ignore_eos[b]) || n_prev + 1 >= st.max_out[b];
  if (nx.res != nullptr) {
    // ---- next decode step's inputs (t3.py:503-514 + block_table.py:357-377): both pair members are
    // fed speech_emb[tok]; position = the slot this token's K/V will occupy.
    const int half = tid >> 7, r = half * B + b;
    nx.res[(size_t)r * 128 + (tid & 127)] = nx.speech_emb[(size_t)tok * 128 + (tid & 127)];
    if ((tid & 127) == 0) {
      const int p = (half ? len_u : len_c) + 1;
      nx.pos[r] = p;
      nx.row_seq[r] = r;
      nx.kv_len[r] = fin ? 0 : p + 1;
      nx.slot[r] = fin ? (int64_t)-1
                       : (int64_t)nx.bt[(int64_t)r * nx.bt_stride + p / CBX_BLOCK_SIZE] * CBX_BLOCK_SIZE + p % CBX_BLOCK_SIZE;
    }
  }
  if (tid == 0) {
    const int n = n_prev;
    st.out_tokens[(size_t)b * st.out_stride + n] = tok;
    st.n_out[b] = n + 1;
    st.last_tok[b] = tok;
    st.seen[(size_t)b * CBX_SEEN_WORDS + (tok >> 5)] |= 1u << (tok & 31);
    st.seq_len[b] = len_c + 1;
    st.seq_len[B + b] = len_u + 1;
    if (fin) {
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

namespace cbx {
int launch_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, const NextStep& nx, cudaStream_t s) {
  if (B <= 0) return 0;
  if (init_sample()) return -1;
  cudaError_t e = launch_pdl(cfg_sample_kernel, dim3(B), dim3(kThreads), kSampleSmem, s, (const bf16*)logits, *st, B, nx);
  CBX_REQUIRE(e == cudaSuccess, "cfg_sample launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
}  // namespace cbx

extern "C" int cbx_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, void* stream) {
  return cbx::launch_cfg_sample(logits, st, B, cbx::NextStep(), (cudaStream_t)stream);
}
