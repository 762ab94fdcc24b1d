// k_sample.cu — fused CFG merge + sampler over the 8194-token speech vocabulary.
//
// Replaces t3.py:732 (logits = cond + cfg·(cond − uncond), bf16, one rounding per op), the
// 1000-wide −inf padding of t3.py:734-738 (never materialised here; ids stay in speech space)
// and vLLM's v1 Sampler chain (sampler.py:67-300): repetition penalty over prompt∪output
// (utils.py:51-89) → greedy argmax | temperature → min-p (builtin.py:101-112) → top-k /
// top-p (topk_topp_sampler.py:328-369) → exponential race (…:395-416), plus the scheduler's
// stop / max_tokens check — with all per-request state resident on the device, so the decode
// loop never synchronises with the host.  One CTA of 1024 threads per CFG pair; every
// thread keeps its 8 contiguous vocabulary ids (one 16-byte chunk of each logits row) in registers
// for the whole chain — no shared-memory copy of the row.
//
// Specified tie rules (the reference leaves them to sort order): argmax → lowest id; top-p
// boundary ties → lowest ids kept first; top-k keeps every token tied with the k-th value.
#include <stdlib.h>

#include "common.cuh"

using namespace cbx;
namespace cbx { int init_sample(); }


namespace {
constexpr int V = CBX_SPEECH_VOCAB;
constexpr int kThreads = 1024, kWarps = 32;
constexpr int NP = 10;                         // ids per thread: 8 (+2 tail ids on the last thread)
constexpr float kSamplingEps = 1e-5f;
}  // namespace
namespace cbx {
struct NextStep {             // optional fused "decode_prepare" for the following step (engine only)
  const uint4* speech_emb; uint4* res; int32_t* pos; int64_t* slot; int32_t* kv_len; int32_t* row_seq;
  const int32_t* bt; int bt_stride;
};
}  // namespace cbx
namespace {

// Loads pinned where they are written (asm volatile): nvcc otherwise sinks them below the early exit / next to their
// first use, which puts one more L2 round trip into the serial chain of this one-CTA-per-pair kernel.
__device__ __forceinline__ uint4 ld_now_u4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ld_now_u32(const void* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint64_t ld_now_u64(const void* p) {
  uint64_t v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}

__device__ __forceinline__ uint32_t fkey(float f) {      // order-preserving float -> uint32
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct Red {
  float f[kWarps], g[kWarps];
  int i[kWarps];
  uint32_t u[kWarps];
};

// Fixed-order (deterministic) block reductions: warp shuffle, then every warp re-reduces the 32
// warp partials, so all threads hold the result after one barrier pair.
__device__ __forceinline__ float block_sum(float v, Red& red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red.f[threadIdx.x >> 5] = v;
  __syncthreads();
  const float s = warp_sum(red.f[threadIdx.x & 31]);
  __syncthreads();
  return s;
}
__device__ __forceinline__ uint32_t block_min_u32(uint32_t v, Red& red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) red.u[threadIdx.x >> 5] = v;
  __syncthreads();
  v = red.u[threadIdx.x & 31];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  return v;
}
__device__ __forceinline__ void warp_argmax(float& v, int& idx) {   // ties -> lowest index
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
}
__device__ __forceinline__ void block_argmax(float& v, int& idx, Red& red) {
  warp_argmax(v, idx);
  if ((threadIdx.x & 31) == 0) { red.f[threadIdx.x >> 5] = v; red.i[threadIdx.x >> 5] = idx; }
  __syncthreads();
  v = red.f[threadIdx.x & 31]; idx = red.i[threadIdx.x & 31];
  warp_argmax(v, idx);
  __syncthreads();
}
// exclusive prefix sum of one int per thread in thread order; also returns the block total
// block_sum(a), block_sum(b) and an exclusive block scan of v behind ONE barrier pair (same summation orders as the separate
// forms, so the same bits): a block barrier of 1024 threads costs ~450-900 cycles here (tools/dbg_sampler_stamps.py).
__device__ __forceinline__ int block_sum2_excl_scan(float& a, float& b, int v, int& total, Red& red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  a = warp_sum(a);
  b = warp_sum(b);
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  // one storing lane: with lane 0 and lane 31 both storing at index w, nvcc 12.9 formed the address as base + (tid >> 3)
  // — right for lane 0 only — and lane 31 faulted with "misaligned address"
  const int warp_incl = __shfl_sync(0xffffffffu, incl, 31);
  if (lane == 0) { red.f[w] = a; red.g[w] = b; red.i[w] = warp_incl; }
  __syncthreads();
  a = warp_sum(red.f[lane]);
  b = warp_sum(red.g[lane]);
  int wt = red.i[lane], winc = wt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, winc, o);
    if (lane >= o) winc += t;
  }
  total = __shfl_sync(0xffffffffu, winc, 31);
  const int before_warp = __shfl_sync(0xffffffffu, winc - wt, w);
  __syncthreads();
  return before_warp + incl - v;
}

// Philox-4x32-10, returns word 0
__device__ __forceinline__ uint32_t philox_word0(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}

// ---- radix select ---------------------------------------------------------------------------
// Minimal key k with F(k) < target, where F(k) = total weight of live ids whose key is > k (== the
// key of the pivot at which the descending cumulative weight reaches target).  MSB-first 8-bit
// digits over the bits in which the live keys actually differ (the common prefix of min and max
// key is skipped, typically leaving 3 passes of which only the first touches every element).
// Weights are integers — 1 for top-k, e_i·2^40 for top-p — so the shared-memory atomics commute
// and the result is bit-reproducible; bins are (lo, hi) 32-bit pairs because 32-bit shared
// atomics are native while 64-bit ones are CAS loops.
struct Radix {
  uint32_t lo[256], hi[256];
  unsigned long long above, target;
  int digit;
};
__device__ __forceinline__ void bin_add(uint32_t* lo, uint32_t* hi, int d, unsigned long long w) {
  const uint32_t wl = (uint32_t)w, wh = (uint32_t)(w >> 32);
  const uint32_t old = atomicAdd(&lo[d], wl);
  const uint32_t carry = (old + wl < old) ? 1u : 0u;   // exactly one adder sees each wrap
  if (wh | carry) atomicAdd(&hi[d], wh + carry);
}
__device__ __forceinline__ void bin_add(Radix& rx, int d, unsigned long long w) { bin_add(rx.lo, rx.hi, d, w); }
__device__ __forceinline__ unsigned long long quant40(float e) { return __float2ull_rz(e * 1099511627776.f); }

template <bool kCount>
__device__ uint32_t radix_select(const float (&l)[NP], const float (&e)[NP], uint32_t kmax, unsigned long long target_in,
                                 float top_p, Radix& rx, Red& red) {
  uint32_t kmin = 0xffffffffu;
#pragma unroll
  for (int j = 0; j < NP; ++j) if (l[j] != -INFINITY) kmin = min(kmin, fkey(l[j]));
  kmin = block_min_u32(kmin, red);
  const int nbits = 32 - __clz(kmin ^ kmax);           // low bits in which live keys differ (clz(0) = 32)
  const int top = (nbits + 7) & ~7, n_pass = top >> 3;
  uint32_t prefix = top >= 32 ? 0u : (kmax >> top);
  if (threadIdx.x == 0) { rx.above = 0; rx.target = target_in; }
#pragma unroll 1
  for (int pass = 0; pass < n_pass; ++pass) {
    const int shift = top - 8 * (pass + 1);
    if (threadIdx.x < 256) { rx.lo[threadIdx.x] = 0; rx.hi[threadIdx.x] = 0; }
    __syncthreads();
    int run_d = -1;
    unsigned long long run_w = 0;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const uint32_t key = fkey(l[j]);
      const bool in = l[j] != -INFINITY && (shift + 8 >= 32 || (key >> (shift + 8)) == prefix);
      if (in) {
        const int d = (key >> shift) & 255;
        if (d != run_d) {
          if (run_w) bin_add(rx, run_d, run_w);
          run_d = d; run_w = 0;
        }
        run_w += kCount ? 1ull : quant40(e[j]);
      }
    }
    if (run_w) bin_add(rx, run_d, run_w);
    __syncthreads();
    if (threadIdx.x < 32) {
      const int lane = threadIdx.x;                    // lane owns bins 255-8*lane .. 248-8*lane
      unsigned long long b8[8], tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int d = 255 - 8 * lane - j;
        b8[j] = ((unsigned long long)rx.hi[d] << 32) | rx.lo[d];
        tot += b8[j];
      }
      unsigned long long incl = tot;
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
  return n_pass == 0 ? kmax : prefix;
}

// Top-p pivot without the radix passes (each costs 5-9 k cycles on the one SM a pair has: four block barriers, a
// 10-element loop in 1024 threads and a serial scan in warp 0; four of them were half of the launch).
//   1. one histogram over kNB LINEAR buckets of (mt - l): monotone in l, so a bucket boundary is a threshold in l; only
//      ids whose integer weight e*2^40 is non-zero take part (d < 28), and they spread over the buckets instead of
//      crowding into the few (sign, exponent) digits of a float key;
//   2. one 64-bit block scan finds the bucket in which the descending cumulative weight reaches the target;
//   3. the bucket's members (a few dozen) are gathered and ranked all-pairs on their EXACT keys: the pivot is the key
//      with  above < target <= above + weight(key)  — the same key the radix passes return, weights are integers so
//      no order of addition matters.
// More than kCap members (near-constant logits): the radix passes take over.
constexpr int kNB = kThreads, kCap = 256;
constexpr float kBucketScale = (float)kNB / 28.f;
struct Bucket {
  uint32_t lo[kNB], hi[kNB];
  unsigned long long wsum[kWarps], mw[kCap];
  uint32_t mkey[kCap];
  unsigned long long above, target;
  int bucket, cnt;
  uint32_t pivot;
};
__device__ __forceinline__ int bucket_of(float mt, float l) { return min(kNB - 1, (int)((mt - l) * kBucketScale)); }

__device__ uint32_t bucket_select(const float (&l)[NP], const float (&e)[NP], float mt, uint32_t kmax, float top_p, Bucket& bk,
                                  Radix& rx, Red& red) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  bk.lo[tid] = 0; bk.hi[tid] = 0;
  if (tid == 0) bk.cnt = 0;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    const unsigned long long w = quant40(e[j]);
    if (w) bin_add(bk.lo, bk.hi, bucket_of(mt, l[j]), w);
  }
  __syncthreads();
  const unsigned long long v = ((unsigned long long)bk.hi[tid] << 32) | bk.lo[tid];
  unsigned long long incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const unsigned long long warp_incl = __shfl_sync(0xffffffffu, incl, 31);    // lane 0 stores: see block_sum2_excl_scan
  if (lane == 0) bk.wsum[warp] = warp_incl;
  __syncthreads();
  const unsigned long long wt = bk.wsum[lane];
  unsigned long long winc = wt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, winc, o);
    if (lane >= o) winc += t;
  }
  const unsigned long long total = __shfl_sync(0xffffffffu, winc, 31);
  const unsigned long long excl = __shfl_sync(0xffffffffu, winc - wt, warp) + incl - v;
  unsigned long long target = (unsigned long long)((double)top_p * (double)total);
  if (target < 1) target = 1;
  if (excl < target && excl + v >= target) { bk.bucket = tid; bk.above = excl; bk.target = target; }   // exactly one bucket
  __syncthreads();
  const int pb = bk.bucket;
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    const unsigned long long w = quant40(e[j]);
    if (w && bucket_of(mt, l[j]) == pb) {
      const int pos = atomicAdd(&bk.cnt, 1);
      if (pos < kCap) { bk.mkey[pos] = fkey(l[j]); bk.mw[pos] = w; }
    }
  }
  __syncthreads();
  const int n = bk.cnt;
  if (n > kCap) return radix_select<false>(l, e, kmax, 0ull, top_p, rx, red);
  if (tid < n) {
    const uint32_t ki = bk.mkey[tid];
    unsigned long long a = bk.above, eq = 0;
    for (int j = 0; j < n; ++j) {
      const uint32_t kj = bk.mkey[j];
      const unsigned long long wj = bk.mw[j];
      if (kj > ki) a += wj;
      else if (kj == ki) eq += wj;
    }
    if (a < bk.target && a + eq >= bk.target) bk.pivot = ki;            // ties write the same key
  }
  __syncthreads();
  return bk.pivot;
}

__device__ __forceinline__ float cfg_pen(float cc, float uu, float cfg, bool seen, float rep, float inv_rep) {
  float v = round_bf16(cc + round_bf16(cfg * round_bf16(cc - uu)));   // t3.py:732, one bf16 rounding per op
  if (seen) v = v * (v > 0.f ? inv_rep : rep);                         // _custom_ops.py:360-373
  return v;
}

// pair_map == NULL: CTA i samples pair i from logits rows i (cond) and B+i (uncond).
// pair_map != NULL (admission of new pairs into a running batch): CTA i samples pair pair_map[i] from
// logits rows i and n_sub+i; every other pair's state is left alone.
// kRadixPivot = true: the top-p pivot comes from the radix passes instead of bucket_select — the checker build of the
// same kernel (cbx_debug_sampler_radix_pivot), against which the GPU tests require bit-identical tokens.
template <bool kRadixPivot>
__global__ void __launch_bounds__(kThreads, 1) cfg_sample_kernel(const bf16* logits,
                                                                 cbx_sampler_state st, int B, const NextStep nx,
                                                                 const int32_t* pair_map, int n_sub) {
  __shared__ Red red;
  __shared__ Radix rx;
  __shared__ Bucket bk;
  const int tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();
  const int b = pair_map ? pair_map[blockIdx.x] : (int)blockIdx.x;
  // the two logits rows and the seen byte are requested BEFORE the finished flag is looked at: one L2 round trip for the
  // whole prologue instead of two (a finished pair's loads are simply dropped)
  const bf16* lc = logits + (size_t)(pair_map ? blockIdx.x : b) * CBX_LOGITS_LD;
  const bf16* lu = logits + (size_t)(pair_map ? n_sub + blockIdx.x : B + b) * CBX_LOGITS_LD;
  const uint32_t* seen = st.seen + (size_t)b * CBX_SEEN_WORDS;
  const uint4 vc = ld_now_u4(reinterpret_cast<const uint4*>(lc) + tid), vu = ld_now_u4(reinterpret_cast<const uint4*>(lu) + tid);
  const uint32_t seen_word = ld_now_u32(seen + (tid >> 2));
  const uint64_t seed = ld_now_u64(st.seed + b), uid = ld_now_u64(st.uid + b);     // used by the race, 10 us later
  const int n_prev = (int)ld_now_u32(st.n_out + b);
  const int len_c = (int)ld_now_u32(st.seq_len + b), len_u = (int)ld_now_u32(st.seq_len + B + b);
  const float rep = __uint_as_float(ld_now_u32(st.rep_pen + b)), cfg = st.cfg_scale;
  const float temp = __uint_as_float(ld_now_u32(st.temperature + b)), min_p = __uint_as_float(ld_now_u32(st.min_p + b));
  const float top_p = __uint_as_float(ld_now_u32(st.top_p + b));
  const int top_k = (int)ld_now_u32(st.top_k + b);
  const int max_out = (int)ld_now_u32(st.max_out + b), ignore_eos = (int)ld_now_u32(st.ignore_eos + b);
  int next_block = 0;                                      // block-table entry of the NEXT position (known before sampling)
  if (nx.res != nullptr && (tid == 0 || tid == 128)) {
    const int p = (tid ? len_u : len_c) + 1;
    next_block = (int)ld_now_u32(nx.bt + (int64_t)((tid >> 7) * B + b) * nx.bt_stride + p / CBX_BLOCK_SIZE);
  }
  uint32_t tail_c = 0, tail_u = 0, tail_seen = 0;          // ids 8192, 8193 live on the last thread
  if (tid == kThreads - 1) {
    tail_c = ld_now_u32(lc + 8192), tail_u = ld_now_u32(lu + 8192);
    tail_seen = ld_now_u32(seen + CBX_SEEN_WORDS - 1);
  }
  if (ld_now_u32(st.finished + b)) return;     // volatile like the loads above, so it cannot be scheduled ahead of them
  const float inv_rep = 1.0f / rep;

  // ---- load: ids [8·tid, 8·tid+8) (+ 8192, 8193 on the last thread); CFG merge + repetition penalty
  const int id0 = tid * 8;
  float l[NP], e[NP];
  {
    const uint32_t sb = (seen_word >> ((tid & 3) * 8)) & 0xffu;
    const uint32_t cw[4] = {vc.x, vc.y, vc.z, vc.w}, uw[4] = {vu.x, vu.y, vu.z, vu.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float cc = (j & 1) ? bf_hi(cw[j >> 1]) : bf_lo(cw[j >> 1]);
      const float uu = (j & 1) ? bf_hi(uw[j >> 1]) : bf_lo(uw[j >> 1]);
      l[j] = cfg_pen(cc, uu, cfg, (sb >> j) & 1u, rep, inv_rep);
    }
    l[8] = l[9] = -INFINITY;
    if (tid == kThreads - 1) {
      l[8] = cfg_pen(bf_lo(tail_c), bf_lo(tail_u), cfg, tail_seen & 1u, rep, inv_rep);
      l[9] = cfg_pen(bf_hi(tail_c), bf_hi(tail_u), cfg, (tail_seen >> 1) & 1u, rep, inv_rep);
    }
  }
  // ---- greedy argmax on the penalised logits (ties -> lowest id)
  float mx = -INFINITY; int amx = V;
#pragma unroll
  for (int j = 0; j < NP; ++j) if (l[j] > mx) { mx = l[j]; amx = id0 + j; }
  block_argmax(mx, amx, red);
  int tok = amx;

  if (temp >= kSamplingEps) {
    // ---- temperature, softmax numerators
    const float mt = mx / temp;
    float z = 0.f;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      l[j] = l[j] / temp;
      e[j] = l[j] == -INFINITY ? 0.f : expf(l[j] - mt);
      z += e[j];
    }
    z = block_sum(z, red);
    const uint32_t kmax = fkey(mt);
    // ---- min-p: drop p_i < p_max · min_p   (builtin.py:101-112)
    if (min_p > 0.f) {
      const float thr = (1.0f / z) * min_p;
      float z2 = 0.f;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        if (e[j] / z < thr) { l[j] = -INFINITY; e[j] = 0.f; }
        z2 += e[j];
      }
      z = block_sum(z2, red);
    }
    // ---- top-k: keep every token >= the k-th largest value
    if (top_k > 0 && top_k < V) {
      const uint32_t kk = radix_select<true>(l, e, kmax, (unsigned long long)top_k, 0.f, rx, red);
      float z2 = 0.f;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        if (fkey(l[j]) < kk) { l[j] = -INFINITY; e[j] = 0.f; }
        z2 += e[j];
      }
      z = block_sum(z2, red);
    }
    // ---- top-p: keep token iff the mass ranked strictly above it is < p
    if (top_p < 1.0f) {
      const float target = top_p * z;
      const uint32_t kp = kRadixPivot ? radix_select<false>(l, e, kmax, 0ull, top_p, rx, red)
                                      : bucket_select(l, e, mt, kmax, top_p, bk, rx, red);
      float g = 0.f, pv = 0.f; int nt = 0;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const uint32_t k = fkey(l[j]);
        if (k > kp && l[j] != -INFINITY) g += e[j];
        else if (k == kp) { ++nt; pv += e[j]; }
      }
      int total;                                          // g, pv = (#ties)·e(v*), tie rank in id order
      const int before = block_sum2_excl_scan(g, pv, nt, total, red);
      const float ev = total > 0 ? pv / (float)total : 0.f;
      int n_keep = total;
      if (ev > 0.f) {
        n_keep = (int)ceilf((target - g) / ev);
        if (n_keep < 1) n_keep = 1;
      }
      float z2 = 0.f;
      int rank = before;
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const uint32_t k = fkey(l[j]);
        bool drop = k < kp;
        if (k == kp) { drop = rank >= n_keep; ++rank; }
        if (drop) { l[j] = -INFINITY; e[j] = 0.f; }
        z2 += e[j];
      }
      z = block_sum(z2, red);
    }
    // ---- exponential race over survivors: argmax (e_i / z) / Exp(1)_i   (topk_topp_sampler.py:395-416)
    const uint32_t step = (uint32_t)n_prev;
    float best = -1.f; int bi = V;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      // The kept maximum has e = 1 and 1 <= z <= 8194, so its score is >= (1/8194)/17.4 (-log u <= 25 ln 2); an element
      // with e < 1e-14 scores <= 1e-14/2.98e-8 (-log u >= 2^-25): it can never win, and skipping it keeps denormal
      // e / z (the slow path of the IEEE division, +10 us on peaked logits) out of the loop.  Same winner, bit for bit.
      if (e[j] >= 1e-14f) {
        const int i = id0 + j;
        const uint32_t x = philox_word0((uint32_t)i, step, (uint32_t)uid, (uint32_t)(uid >> 32), (uint32_t)seed,
                                        (uint32_t)(seed >> 32));
        const float uu = ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-07f;   // (0,1), exact in fp32
        const float sc = (e[j] / z) / (-logf(uu));
        if (sc > best) { best = sc; bi = i; }
      }
    }
    block_argmax(best, bi, red);
    tok = bi;
  }

  const bool fin = (tok == st.stop_token && !ignore_eos) || n_prev + 1 >= max_out;
  if (nx.res != nullptr && tid < 256) {
    // ---- next decode step's inputs (t3.py:503-514 + block_table.py:357-377): both pair members are
    // fed speech_emb[tok]; position = the slot this token's K/V will occupy.
    const int half = tid >> 7, r = half * B + b;
    nx.res[(size_t)r * 128 + (tid & 127)] = nx.speech_emb[(size_t)tok * 128 + (tid & 127)];
    if ((tid & 127) == 0) {
      const int p = (half ? len_u : len_c) + 1;
      nx.pos[r] = p;
      nx.row_seq[r] = r;
      nx.kv_len[r] = fin ? 0 : p + 1;
      nx.slot[r] = fin ? (int64_t)-1 : (int64_t)next_block * CBX_BLOCK_SIZE + p % CBX_BLOCK_SIZE;
    }
  }
  // ---- emit + bookkeeping (one thread)
  if (tid == 0) {
    st.out_tokens[(size_t)b * st.out_stride + n_prev] = tok;
    st.n_out[b] = n_prev + 1;
    st.last_tok[b] = tok;
    atomicOr(&st.seen[(size_t)b * CBX_SEEN_WORDS + (tok >> 5)], 1u << (tok & 31));   // RED.OR: no read round trip before exit
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
static int g_radix_pivot = 0;
int init_sample() { return 0; }

int launch_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, const NextStep& nx, cudaStream_t s,
                      const int32_t* pair_map, int n_sub) {
  if (B <= 0 || (pair_map && n_sub <= 0)) return 0;
  cudaError_t e = launch_pdl(g_radix_pivot ? cfg_sample_kernel<true> : cfg_sample_kernel<false>, dim3(pair_map ? n_sub : B),
                             dim3(kThreads), 0, s, (const bf16*)logits, *st, B, nx, pair_map, n_sub);
  CBX_REQUIRE(e == cudaSuccess, "cfg_sample launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
}  // namespace cbx

extern "C" int cbx_debug_sampler_radix_pivot(int on) { cbx::g_radix_pivot = on; return 0; }
extern "C" int cbx_cfg_sample(const void* logits, const cbx_sampler_state* st, int B, void* stream) {
  return cbx::launch_cfg_sample(logits, st, B, cbx::NextStep(), (cudaStream_t)stream, nullptr, 0);
}
