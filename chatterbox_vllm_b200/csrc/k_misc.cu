// k_misc.cu — small HBM-bound kernels of the T3 step + host-side block pool.
//   slot mapping, decode_prepare, prefill_embed, (add+)RMSNorm, RoPE + KV append,
//   gate/up weight packing.
#include <stdarg.h>
#include <stdlib.h>

#include <deque>

#include "common.cuh"

namespace cbx {
static thread_local char g_err[512] = "";
static int64_t g_launches = 0;
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int count_launch() { ++g_launches; return 0; }
int64_t launches() { return g_launches; }
int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      n = v < kMaxSms ? v : kMaxSms;
    else
      return kMaxSms;
  }
  return n;
}
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("CBX_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}
}  // namespace cbx
using namespace cbx;

extern "C" const char* cbx_last_error(void) { return cbx::g_err; }
extern "C" int cbx_abi_version(void) { return 1; }

// ---------------------------------------------------------------- block pool (host) ---
struct cbx_pool {
  std::deque<int32_t> free_q;
  int num_blocks;
};
extern "C" cbx_pool* cbx_pool_create(int num_blocks) {
  if (num_blocks < 1) { set_error("cbx_pool_create: num_blocks < 1"); return nullptr; }
  cbx_pool* p = new cbx_pool();
  p->num_blocks = num_blocks;
  for (int i = 1; i < num_blocks; ++i) p->free_q.push_back(i);   // block 0 popped: null block
  return p;
}
extern "C" void cbx_pool_destroy(cbx_pool* p) { delete p; }
extern "C" int cbx_pool_num_free(const cbx_pool* p) { return (int)p->free_q.size(); }
extern "C" int cbx_pool_alloc(cbx_pool* p, int n, int32_t* out) {
  CBX_REQUIRE(n >= 0 && n <= (int)p->free_q.size(), "Cannot get %d free blocks from the pool", n);
  for (int i = 0; i < n; ++i) { out[i] = p->free_q.front(); p->free_q.pop_front(); }
  return 0;
}
extern "C" int cbx_pool_free(cbx_pool* p, const int32_t* ids, int n) {
  for (int i = 0; i < n; ++i) {
    CBX_REQUIRE(ids[i] >= 0 && ids[i] < p->num_blocks, "cbx_pool_free: bad block id %d", ids[i]);
    if (ids[i] != 0) p->free_q.push_back(ids[i]);
  }
  return 0;
}

// ---------------------------------------------------------------- slot mapping --------
__global__ void slot_mapping_kernel(const int32_t* bt, int bt_stride,
                                    const int32_t* __restrict__ req, const int32_t* pos,
                                    int64_t* __restrict__ slot, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int p = pos[i];
  int64_t s = -1;
  if (p >= 0) {
    int64_t blk = bt[(int64_t)req[i] * bt_stride + p / CBX_BLOCK_SIZE];
    s = blk * CBX_BLOCK_SIZE + p % CBX_BLOCK_SIZE;
  }
  slot[i] = s;
}
extern "C" int cbx_slot_mapping(const int32_t* bt, int bt_stride, const int32_t* req, const int32_t* pos,
                                int64_t* slot, int n, void* stream) {
  if (n <= 0) return 0;
  slot_mapping_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(bt, bt_stride, req, pos, slot, n);
  CBX_CHECK_LAUNCH("slot_mapping");
  return 0;
}

// ---------------------------------------------------------------- decode prepare ------
// One CTA (128 threads) per row r in [0,2B): copies the 2 KB embedding row and thread 0
// does the row's index arithmetic.
__global__ void __launch_bounds__(128) decode_prepare_kernel(
    const uint4* __restrict__ speech_emb, const int32_t* last_tok,
    const int32_t* finished, const int32_t* seq_len,
    const int32_t* bt, int bt_stride, uint4* __restrict__ x, int32_t* __restrict__ pos,
    int64_t* __restrict__ slot, int32_t* __restrict__ kv_len, int32_t* __restrict__ row_seq,
    int32_t* __restrict__ head_rows, int B) {
  const int r = blockIdx.x, b = r % B;
  pdl_trigger();
  pdl_wait();
  const int tok = last_tok[b];
  x[(size_t)r * 128 + threadIdx.x] = speech_emb[(size_t)tok * 128 + threadIdx.x];
  if (threadIdx.x == 0) {
    const int p = seq_len[r];
    const bool live = finished[b] == 0;
    pos[r] = p;
    if (row_seq) row_seq[r] = r;
    if (head_rows) head_rows[r] = r;
    kv_len[r] = live ? p + 1 : 0;
    slot[r] = live ? (int64_t)bt[(int64_t)r * bt_stride + p / CBX_BLOCK_SIZE] * CBX_BLOCK_SIZE + p % CBX_BLOCK_SIZE
                   : (int64_t)-1;
  }
}

namespace cbx {
int launch_decode_prepare(const void* speech_emb, const int32_t* last_tok, const int32_t* finished,
                          const int32_t* seq_len, const int32_t* bt, int bt_stride, void* x, int32_t* pos,
                          int64_t* slot, int32_t* kv_len, int32_t* row_seq, int32_t* head_rows, int B,
                          cudaStream_t s) {
  cudaError_t e = launch_pdl(decode_prepare_kernel, dim3(2 * B), dim3(128), 0, s, (const uint4*)speech_emb, last_tok,
                             finished, seq_len, bt, bt_stride, (uint4*)x, pos, slot, kv_len, row_seq, head_rows, B);
  CBX_REQUIRE(e == cudaSuccess, "decode_prepare launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
}  // namespace cbx

extern "C" int cbx_decode_prepare(const void* speech_emb, const int32_t* last_tok, const int32_t* finished,
                                  const int32_t* seq_len, const int32_t* bt, int bt_stride, void* x,
                                  int32_t* pos, int64_t* slot, int32_t* kv_len, int B, void* stream) {
  CBX_REQUIRE(B > 0, "cbx_decode_prepare: B <= 0");
  // row_seq / head_rows are engine-internal identity maps; the standalone op skips them.
  return launch_decode_prepare(speech_emb, last_tok, finished, seq_len, bt, bt_stride, x, pos, slot, kv_len,
                               nullptr, nullptr, B, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- prefill embed -------
__global__ void __launch_bounds__(128) prefill_embed_kernel(
    const int4* __restrict__ desc, const float* __restrict__ cond_emb, const uint4* __restrict__ text_emb,
    const uint4* __restrict__ text_pos, const uint4* __restrict__ speech_emb,
    const uint4* __restrict__ speech_pos, uint4* __restrict__ x) {
  const int r = blockIdx.x, t = threadIdx.x;
  const int4 d = desc[r];
  uint4 o = make_uint4(0, 0, 0, 0);
  if (d.x == 0) {
    const float4* src = reinterpret_cast<const float4*>(cond_emb + ((size_t)d.y * 34 + d.z) * CBX_HIDDEN) + t * 2;
    float4 a = src[0], b = src[1];
    o = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
  } else if (d.x == 1 || d.x == 3) {
    uint4 a = (d.x == 1 ? text_emb : speech_emb)[(size_t)d.y * 128 + t];
    uint4 b = (d.x == 1 ? text_pos : speech_pos)[(size_t)d.z * 128 + t];
    o.x = pack_bf16(bf_lo(a.x) + bf_lo(b.x), bf_hi(a.x) + bf_hi(b.x));
    o.y = pack_bf16(bf_lo(a.y) + bf_lo(b.y), bf_hi(a.y) + bf_hi(b.y));
    o.z = pack_bf16(bf_lo(a.z) + bf_lo(b.z), bf_hi(a.z) + bf_hi(b.z));
    o.w = pack_bf16(bf_lo(a.w) + bf_lo(b.w), bf_hi(a.w) + bf_hi(b.w));
  } else if (d.x == 4) {                               // an already generated speech token: the decode embedding (t3.py:503-514)
    o = speech_emb[(size_t)d.y * 128 + t];
  }
  x[(size_t)r * 128 + t] = o;
}
extern "C" int cbx_prefill_embed(const int32_t* desc, const float* cond_emb, const void* text_emb,
                                 const void* text_pos, const void* speech_emb, const void* speech_pos,
                                 void* x, int T, void* stream) {
  if (T <= 0) return 0;
  prefill_embed_kernel<<<T, 128, 0, (cudaStream_t)stream>>>(
      (const int4*)desc, cond_emb, (const uint4*)text_emb, (const uint4*)text_pos, (const uint4*)speech_emb,
      (const uint4*)speech_pos, (uint4*)x);
  CBX_CHECK_LAUNCH("prefill_embed");
  return 0;
}

// ---------------------------------------------------------------- prefill row descriptors
// What T3Engine.prefill built row by row with numpy (51 200 rows at M1): the descriptor, position, block-table row and
// context length of every prefill row, from one record per SEQUENCE (a pair contributes two: cond and uncond).
// seq[s] = {first row, rows P, text tokens nt, index into cond_emb, uncond (0/1), block-table row, offset of its text ids,
// offset of its history}; a sequence's rows are [34 conditioning rows] [nt text rows] [start-of-speech] [P-35-nt history
// rows] (t3.py:229-262 layout; the uncond member's text rows are zeros, t3.py:548-569; history rows carry tokens
// generated before a preemption).  One thread per row, binary search over the sequences' first rows.
__global__ void __launch_bounds__(256) prefill_rows_kernel(const int32_t* __restrict__ seq, int n_seq,
                                                           const int32_t* __restrict__ text_ids,
                                                           const int32_t* __restrict__ hist, int T, int4* __restrict__ desc,
                                                           int32_t* __restrict__ pos, int32_t* __restrict__ row_seq,
                                                           int32_t* __restrict__ kv_len) {
  const int r = blockIdx.x * 256 + threadIdx.x;
  if (r >= T) return;
  int lo = 0, hi = n_seq - 1;
  while (lo < hi) {                                    // last sequence whose first row is <= r
    const int mid = (lo + hi + 1) >> 1;
    if (seq[mid * 8] <= r) lo = mid; else hi = mid - 1;
  }
  const int32_t* q = seq + lo * 8;
  const int i = r - q[0], nt = q[2];
  int4 d;
  if (i < 34) d = make_int4(0, q[3], i, 0);
  else if (i < 34 + nt) d = q[4] ? make_int4(2, 0, 0, 0) : make_int4(1, text_ids[q[6] + i - 34], i - 34, 0);
  else if (i == 34 + nt) d = make_int4(3, 6561, 0, 0);
  else d = make_int4(4, hist[q[7] + i - 35 - nt], 0, 0);
  desc[r] = d;
  pos[r] = i;
  row_seq[r] = q[5];
  kv_len[r] = i + 1;
}
extern "C" int cbx_prefill_rows(const int32_t* seq, int n_seq, const int32_t* text_ids, const int32_t* hist, int T,
                                int32_t* desc, int32_t* pos, int32_t* row_seq, int32_t* kv_len, void* stream) {
  if (T <= 0 || n_seq <= 0) return 0;
  CBX_REQUIRE(seq && text_ids && desc && pos && row_seq && kv_len, "cbx_prefill_rows: NULL argument");
  prefill_rows_kernel<<<(T + 255) / 256, 256, 0, (cudaStream_t)stream>>>(seq, n_seq, text_ids, hist, T, (int4*)desc, pos, row_seq, kv_len);
  CBX_CHECK_LAUNCH("prefill_rows");
  return 0;
}

// ---------------------------------------------------------------- (add +) RMSNorm -----
// One CTA of 128 threads per row; each thread owns 8 contiguous elements (one 16-byte
// load).  Rounding points follow vLLM: z = bf16(x + res); out = bf16(bf16(z·rstd)·w).
// MODE 0: out = norm(x).  MODE 1: res = bf16(x + res) in place, out = norm(res).
// MODE 2: x arrives as `splits` fp32 split-K partials [splits][T][1024] of the projection that
//         closes the residual branch: x = bf16(sum of partials), then as MODE 1.
template <int MODE>
__global__ void __launch_bounds__(128) rmsnorm_kernel(const uint4* x, uint4* __restrict__ res,
                                                      const uint4* __restrict__ w, uint4* __restrict__ out,
                                                      float eps, int splits, int T) {
  const int r = blockIdx.x, t = threadIdx.x;
  pdl_trigger();
  const uint4 wv = w[t];                               // weights: independent of the previous kernel
  pdl_wait();
  float f[8];
  uint4 q = make_uint4(0, 0, 0, 0);                   // requested with the first loads, not after the slice loop
  if (MODE >= 1) q = res[(size_t)r * 128 + t];
  if (MODE == 2) {
    const float4* p = reinterpret_cast<const float4*>(x) + ((size_t)r * 128 + t) * 2;
    float4 a = p[0], b = p[1];
    for (int z = 1; z < splits; ++z) {
      const float4 a2 = p[(size_t)z * T * 256], b2 = p[(size_t)z * T * 256 + 1];
      a.x += a2.x; a.y += a2.y; a.z += a2.z; a.w += a2.w;
      b.x += b2.x; b.y += b2.y; b.z += b2.z; b.w += b2.w;
    }
    f[0] = round_bf16(a.x); f[1] = round_bf16(a.y); f[2] = round_bf16(a.z); f[3] = round_bf16(a.w);
    f[4] = round_bf16(b.x); f[5] = round_bf16(b.y); f[6] = round_bf16(b.z); f[7] = round_bf16(b.w);
  } else {
    const uint4 v = x[(size_t)r * 128 + t];
    f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
    f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
  }
  if (MODE >= 1) {
    float g[8] = {bf_lo(q.x), bf_hi(q.x), bf_lo(q.y), bf_hi(q.y), bf_lo(q.z), bf_hi(q.z), bf_lo(q.w), bf_hi(q.w)};
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = round_bf16(f[i] + g[i]);
    res[(size_t)r * 128 + t] =
        make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  }
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
  ss = warp_sum(ss);
  __shared__ float part[4];
  if ((t & 31) == 0) part[t >> 5] = ss;
  __syncthreads();
  const float rstd = rsqrtf((part[0] + part[1] + part[2] + part[3]) * (1.0f / CBX_HIDDEN) + eps);
  float wf[8] = {bf_lo(wv.x), bf_hi(wv.x), bf_lo(wv.y), bf_hi(wv.y), bf_lo(wv.z), bf_hi(wv.z), bf_lo(wv.w), bf_hi(wv.w)};
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = round_bf16(f[i] * rstd) * wf[i];
  out[(size_t)r * 128 + t] =
      make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
// The engine's form (o_proj / down_proj leave 4 K-slices): 256 threads per row, four elements each, so a thread has just
// five loads — residual + one float4 per slice — and they are all in flight together.  The general kernel above, with
// eight elements per thread and the slice loop left to the compiler, fetched in four register-reusing waves plus the
// residual at the end: five L2 round trips in a kernel that is nothing but latency.  Same order of adds per element.
template <int S>
__global__ void __launch_bounds__(256) rmsnorm_reduce4_kernel(const float4* part, uint2* res, const uint2* __restrict__ w,
                                                              uint2* out, float eps, int T) {
  const int r = blockIdx.x, t = threadIdx.x;
  pdl_trigger();
  uint2 wv;                                            // weights: independent of the previous kernel.  ld.volatile, or ptxas
  asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(wv.x), "=r"(wv.y) : "l"(w + t));   // sinks it to its use
  pdl_wait();
  const float4* p = part + (size_t)r * 256 + t;
  const size_t zs = (size_t)T * 256;
  const uint2 q = res[(size_t)r * 256 + t];
  float4 a[S];
#pragma unroll
  for (int z = 0; z < S; ++z) a[z] = p[z * zs];
  float4 sum = a[0];
#pragma unroll
  for (int z = 1; z < S; ++z) { sum.x += a[z].x; sum.y += a[z].y; sum.z += a[z].z; sum.w += a[z].w; }
  float f[4] = {round_bf16(sum.x), round_bf16(sum.y), round_bf16(sum.z), round_bf16(sum.w)};
  const float g[4] = {bf_lo(q.x), bf_hi(q.x), bf_lo(q.y), bf_hi(q.y)};
#pragma unroll
  for (int i = 0; i < 4; ++i) f[i] = round_bf16(f[i] + g[i]);
  res[(size_t)r * 256 + t] = make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]));
  float ss = (f[0] * f[0] + f[1] * f[1]) + (f[2] * f[2] + f[3] * f[3]);
  ss = warp_sum(ss);
  __shared__ float sums[8];
  if ((t & 31) == 0) sums[t >> 5] = ss;
  __syncthreads();
  const float tot = ((sums[0] + sums[1]) + (sums[2] + sums[3])) + ((sums[4] + sums[5]) + (sums[6] + sums[7]));
  const float rstd = rsqrtf(tot * (1.0f / CBX_HIDDEN) + eps);
  const float wf[4] = {bf_lo(wv.x), bf_hi(wv.x), bf_lo(wv.y), bf_hi(wv.y)};
#pragma unroll
  for (int i = 0; i < 4; ++i) f[i] = round_bf16(f[i] * rstd) * wf[i];
  out[(size_t)r * 256 + t] = make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]));
}
extern "C" int cbx_rmsnorm(const void* x, void* residual, const void* w, void* out, int T, float eps,
                           void* stream) {
  if (T <= 0) return 0;
  cudaError_t e;
  if (residual)
    e = launch_pdl(rmsnorm_kernel<1>, dim3(T), dim3(128), 0, (cudaStream_t)stream, (const uint4*)x,
                   (uint4*)residual, (const uint4*)w, (uint4*)out, eps, 1, T);
  else
    e = launch_pdl(rmsnorm_kernel<0>, dim3(T), dim3(128), 0, (cudaStream_t)stream, (const uint4*)x,
                   (uint4*)nullptr, (const uint4*)w, (uint4*)out, eps, 1, T);
  CBX_REQUIRE(e == cudaSuccess, "rmsnorm launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
extern "C" int cbx_rmsnorm_reduce(const void* partials, int splits, void* residual, const void* w, void* out, int T,
                                  float eps, void* stream) {
  if (T <= 0) return 0;
  CBX_REQUIRE(splits >= 1 && residual != nullptr, "cbx_rmsnorm_reduce: bad arguments");
  cudaError_t e = splits == 4
      ? launch_pdl(rmsnorm_reduce4_kernel<4>, dim3(T), dim3(256), 0, (cudaStream_t)stream, (const float4*)partials,
                   (uint2*)residual, (const uint2*)w, (uint2*)out, eps, T)
      : splits == 2
      ? launch_pdl(rmsnorm_reduce4_kernel<2>, dim3(T), dim3(256), 0, (cudaStream_t)stream, (const float4*)partials,
                   (uint2*)residual, (const uint2*)w, (uint2*)out, eps, T)
      : launch_pdl(rmsnorm_kernel<2>, dim3(T), dim3(128), 0, (cudaStream_t)stream, (const uint4*)partials,
                   (uint4*)residual, (const uint4*)w, (uint4*)out, eps, splits, T);
  CBX_REQUIRE(e == cudaSuccess, "rmsnorm_reduce launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

// ---------------------------------------------------------------- RoPE + KV append ----
// One CTA of 128 threads per token row.  Threads 0..63 rotate q, 64..127 rotate k: each
// owns the 8-element chunk c (0..3) of a head's first half and its partner chunk c+4.
// Then every thread moves one 16-byte chunk of v.  K chunks land XOR-swizzled (chunk ^
// (tok & 7)) so the attention kernel's per-token 128-bit reads are bank-conflict-free.
__global__ void __launch_bounds__(128) rope_kv_append_kernel(uint4* __restrict__ qkv,
                                                             const int32_t* pos,
                                                             const int64_t* slot,
                                                             const uint4* __restrict__ cos_sin,
                                                             uint8_t* __restrict__ kv_layer) {
  const int r = blockIdx.x, t = threadIdx.x;
  uint4* row = qkv + (size_t)r * 384;                 // 3072 bf16 = 384 x 16 B
  const int which = t >> 6, item = t & 63, h = item >> 2, c = item & 3;
  const int p = pos[r];
  const int64_t s = slot[r];
  const uint4 cv = cos_sin[(size_t)p * 8 + c], sv = cos_sin[(size_t)p * 8 + 4 + c];
  uint4* src = row + which * 128 + h * 8;
  const uint4 a = src[c], b = src[c + 4];
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
  const uint32_t cw[4] = {cv.x, cv.y, cv.z, cv.w}, sw[4] = {sv.x, sv.y, sv.z, sv.w};
  uint32_t o1[4], o2[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float x1l = bf_lo(aw[i]), x1h = bf_hi(aw[i]), x2l = bf_lo(bw[i]), x2h = bf_hi(bw[i]);
    float cl = bf_lo(cw[i]), ch = bf_hi(cw[i]), sl = bf_lo(sw[i]), sh = bf_hi(sw[i]);
    o1[i] = pack_bf16(x1l * cl - x2l * sl, x1h * ch - x2h * sh);
    o2[i] = pack_bf16(x2l * cl + x1l * sl, x2h * ch + x1h * sh);
  }
  const uint4 r1 = make_uint4(o1[0], o1[1], o1[2], o1[3]), r2 = make_uint4(o2[0], o2[1], o2[2], o2[3]);
  uint8_t* page = nullptr;
  int tok = 0;
  if (s >= 0) {
    page = kv_layer + (size_t)(s / CBX_BLOCK_SIZE) * CBX_PAGE_BYTES;
    tok = (int)(s % CBX_BLOCK_SIZE);
  }
  if (which == 0) {
    src[c] = r1;
    src[c + 4] = r2;
  } else if (page) {
    uint4* kt = reinterpret_cast<uint4*>(page + h * 2048 + tok * 128);
    kt[c ^ (tok & 7)] = r1;
    kt[(c + 4) ^ (tok & 7)] = r2;
  }
  if (page) {
    const int vh = t >> 3, vc = t & 7;
    reinterpret_cast<uint4*>(page + 32768 + vh * 2048 + tok * 128)[vc] = row[256 + t];
  }
}
extern "C" int cbx_rope_kv_append(void* qkv, const int32_t* pos, const int64_t* slot, const void* cos_sin,
                                  void* kv_layer, int T, void* stream) {
  if (T <= 0) return 0;
  rope_kv_append_kernel<<<T, 128, 0, (cudaStream_t)stream>>>((uint4*)qkv, pos, slot, (const uint4*)cos_sin,
                                                              (uint8_t*)kv_layer);
  CBX_CHECK_LAUNCH("rope_kv_append");
  return 0;
}

// ---------------------------------------------------------------- gate/up packing -----
__global__ void pack_gate_up_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int I, int K8) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)2 * I * K8) return;
  const int row = (int)(i / K8), col = (int)(i % K8);
  const int which = row / I, j = row % I;               // which: 0 gate, 1 up
  const int drow = (j / 16) * 32 + which * 16 + (j % 16);
  dst[(int64_t)drow * K8 + col] = src[i];
}
extern "C" int cbx_pack_gate_up(const void* src, void* dst, int I, int K, void* stream) {
  CBX_REQUIRE(I % 16 == 0 && K % 8 == 0, "cbx_pack_gate_up: I %% 16 or K %% 8 != 0");
  const int64_t n = (int64_t)2 * I * (K / 8);
  pack_gate_up_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint4*)src,
                                                                                       (uint4*)dst, I, K / 8);
  CBX_CHECK_LAUNCH("pack_gate_up");
  return 0;
}

// ---- token post-processing: the hand-off format between T3 and the vocoder ----------------------
// Per request (reference tts.py:448-465 + s3tokenizer/__init__.py:38-78, on un-offset speech ids):
// cut at the first stop id and crop `tail_crop` tokens before it; keep what lies between the first
// SOS (exclusive) and the first EOS; drop ids >= sos (6561); cap at guard_cap (tts.py:466-486).
// One CTA per request; order-preserving compaction with a ballot scan; int64 out as S3Gen takes it.
namespace {
__global__ void __launch_bounds__(256) postprocess_tokens_kernel(const int32_t* out_tokens,
                                                                 const int32_t* n_out, int out_stride,
                                                                 const int32_t* slots, int sos, int eos,
                                                                 int tail_crop, const int32_t* __restrict__ guard_cap,
                                                                 int64_t* __restrict__ dst, int dst_stride,
                                                                 const int32_t* __restrict__ dst_rows,
                                                                 int32_t* __restrict__ dst_len) {
  __shared__ int s_stop, s_sos, s_eos, s_warp[8], s_base;
  const int item = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int32_t* src = out_tokens + (size_t)slots[item] * out_stride;
  int n = n_out[slots[item]];
  n = n < 0 ? 0 : (n > out_stride ? out_stride : n);
  if (tid == 0) { s_stop = n; s_sos = n; s_eos = n; s_base = 0; }
  __syncthreads();
  for (int i = tid; i < n; i += 256)
    if (src[i] == eos) atomicMin(&s_stop, i);
  __syncthreads();
  if (s_stop < n) {                                   // truncate at the first stop, crop the tail before it
    const int crop = tail_crop > 0 ? tail_crop : 0;
    n = s_stop - crop;
    n = n < 0 ? 0 : n;
  }
  __syncthreads();
  if (tid == 0) { s_sos = n; s_eos = n; }
  __syncthreads();
  for (int i = tid; i < n; i += 256) {
    const int t = src[i];
    if (t == sos) atomicMin(&s_sos, i);
    if (t == eos) atomicMin(&s_eos, i);
  }
  __syncthreads();
  const int start = s_sos < n ? s_sos + 1 : 0;
  const int end = s_eos < n ? s_eos : n;
  const int cap = guard_cap ? guard_cap[item] : 0x7fffffff;
  int64_t* out = dst + (size_t)dst_rows[item] * dst_stride;
  for (int i0 = start; i0 < end; i0 += 256) {         // CTA-uniform trip count
    const int i = i0 + tid;
    const int t = i < end ? src[i] : sos;
    const bool keep = i < end && t < sos;
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int off = s_base;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    off += __popc(bal & ((1u << lane) - 1));
    if (keep && off < cap && off < dst_stride) out[off] = (int64_t)t;
    __syncthreads();
    if (tid == 0) { int tot = 0; for (int w = 0; w < 8; ++w) tot += s_warp[w]; s_base += tot; }
    __syncthreads();
  }
  if (tid == 0) {
    int len = s_base;
    len = len > cap ? cap : len;
    dst_len[dst_rows[item]] = len > dst_stride ? dst_stride : len;
  }
}
}  // namespace

extern "C" int cbx_postprocess_tokens(const int32_t* out_tokens, const int32_t* n_out, int out_stride, const int32_t* slots,
                                      int n, int sos_token, int eos_token, int tail_crop, const int32_t* guard_cap,
                                      int64_t* dst, int dst_stride, const int32_t* dst_rows, int32_t* dst_len, void* stream) {
  if (n <= 0) return 0;
  CBX_REQUIRE(out_tokens && n_out && slots && dst && dst_rows && dst_len && out_stride > 0 && dst_stride > 0,
              "cbx_postprocess_tokens: bad arguments");
  postprocess_tokens_kernel<<<n, 256, 0, (cudaStream_t)stream>>>(out_tokens, n_out, out_stride, slots, sos_token, eos_token,
                                                               tail_crop, guard_cap, dst, dst_stride, dst_rows, dst_len);
  CBX_CHECK_LAUNCH("postprocess_tokens");
  return 0;
}
