// k_mega.cu — EXPERIMENT (env CBX_MEGA=1, off by default): the whole T <= 4 decode step (30 layers
// + final norm + speech head) as ONE persistent cooperative kernel, one CTA per SM.
//
// Idea: at b=1 (one CFG pair = 2 rows) a step reads 1.02 GB of weights = ~157 us of HBM time, but the
// chain of 150 dependent kernels costs 3-4.6 us each = ~620 us.  Here
//  (1) the weight stream is decoupled from the dependency chain: every CTA walks the STATIC list of
//      weight rows it will need (QKV, O, gate_up, down of each layer, then the head) through an
//      11 x 16 KB shared-memory ring of TMA bulk copies, primed at kernel start and refilled by
//      warp 15 as soon as a projection releases its groups — copies never wait for activations;
//  (2) the five all-to-all exchanges of a layer (x -> qkv -> attention -> o-proj -> gate_up -> down)
//      use no grid barrier (measured: ~2100 cycles each = release fence + atomic + poll).  Every
//      exchanged activation is a 32-bit word {bf16 value, 16-bit generation tag} (fp32 attention
//      partials: 64-bit {value, 32-bit tag}); aligned word stores are single-copy atomic, so a
//      consumer re-reads the words it needs from L2 until all carry the tag of the phase it waits
//      for — the data arrives with the signal (the "LL" idea of NCCL's low-latency protocol).
//      Write-after-read safety comes from the chain itself: a buffer written in phase p is next
//      written in phase p+2 or later, and nobody can reach phase p+2 before every CTA has finished
//      phase p+1, whose inputs are exactly the reads of phase p's buffer.
//
//   phase   waits for (tagged words)          rows per CTA         writes
//   QKV     residual row (DOWN of l-1)        <= 11 RoPE pairs     q, new k/v (tagged) + k/v -> KV page
//   ATTN    q of (row, head); the new k/v     one (row, head, split) item   fp32 partial o[64], m, l (tagged)
//   O       partials of all heads             <= 4 row pairs       residual (tagged)
//   GU      residual row                      <= 28 gate/up pairs  SiLU(g)·u (tagged)
//   DOWN    act rows                          <= 7 rows (K=4096)   residual (tagged)
//   HEAD    residual row                      <= 28 row pairs      bf16 logits (plain: next kernel reads them)
//
// The dot products run on mma.sync m16n8k16 (weights = A from the ring via ldmatrix, the <= 4
// activation rows = B, fp32 accumulate); the 16 warps split K, partial sums meet in shared memory.
// Rounding points are those of k_gemv.cu / the tcgen05 path (vLLM's: GEMM outputs, norm outputs,
// residual, rotated q/k, attention output, logits are bf16; softmax statistics fp32).
//
// RESULT (B200, b=1, ctx 300, 30 layers): parity-green (tests/test_gpu_engine.py::
// test_engine_persistent_decode_kernel_parity) but ~1000 us/step against ~620 us for the PDL kernel
// chain, so the engine keeps the chain.  Per-layer cycle stamps (-DCBX_MEGA_PROF): ~55 k cycles, ~43 k
// with the weight copies disabled.  Where it goes: every phase is a serial chain of small latencies
// that a kernel boundary also pays only once — L2 poll (~300 cycles unloaded, 2-3x that while 148
// CTAs poll), two block barriers, mbarrier waits (~250 cycles each), the K-split reduction through
// shared memory, and reading the CTA's 226 KB/layer weight slice back out of shared memory at
// 128 B/clk (~1800 cycles/layer on its own).  Five phases x ~4-5 k cycles + the attention phase's
// extra hops exceed the ~7 k cycles (3.9 us) a fused GEMV kernel of the chain costs end to end.
// Kept as a tested alternative; see DESIGN.md section 7.
// Replaces the same reference code as k_gemv.cu + k_attn.cu (vllm llama.py:117-120,212-226,316-333,
// 394-430; flash_attn.py:796-800,850-874; t3.py:729-730).
#include "common.cuh"

#include <math.h>

namespace cbx {
struct MegaArgs {
  const bf16* ln1[32]; const bf16* wqkv[32]; const bf16* wo[32];
  const bf16* ln2[32]; const bf16* wgu[32]; const bf16* wd[32];
  const bf16* norm_w; const bf16* head_w;
  const bf16* res_in; bf16* logits; uint8_t* ws; unsigned* sync; bf16* trace;
  const bf16* cos_sin; const int32_t* pos; const int64_t* slot; const int32_t* row_seq; const int32_t* kv_len;
  const int32_t* bt; uint8_t* kv_cache; size_t layer_bytes;
  int bt_stride, T, n_layers, vocab, ldl; float eps; int flags;
};
}  // namespace cbx
using namespace cbx;

namespace {
constexpr int kNW = 16;                               // warps (a 17th, dedicated producer warp would cap registers at 96)
constexpr int kConsumers = kNW * 32;
constexpr int kThreads = kConsumers;
constexpr int kRowStride = 2048 + 16;                 // one 1024-element weight row segment, padded: ldmatrix conflict-free
constexpr int kGroupBytes = 8 * kRowStride;           // ring granule: 8 row segments, one full/empty barrier pair
constexpr int kRing = 11;
constexpr int kXStride = 4096 + 8;                    // activation row stride (elements): B-fragment loads conflict-free
constexpr int kXBytes = 4 * kXStride * 2;
constexpr int kRedFloats = kNW * 68;
constexpr int kMaxSplits = 4;
constexpr int kPartWords = 66;                        // attention partial: o[64], m, l
constexpr int kWtsFloats = 4 * 16 * kMaxSplits;       // merge weights per (row, head, split)
constexpr int kSmem = kRing * kGroupBytes + kXBytes + kRedFloats * 4 + kWtsFloats * 4 + 2 * kRing * 8;
constexpr float kLog2e = 1.4426950408889634f;

// tagged exchange buffers inside MegaArgs::ws (byte offsets)
constexpr size_t kOffRes = 0;                                         // u32 [4][1024]
constexpr size_t kOffQ = kOffRes + 4 * 1024 * 4;                      // u32 [4][1024]
constexpr size_t kOffKV = kOffQ + 4 * 1024 * 4;                       // u32 [4][2][1024]   new k, v rows
constexpr size_t kOffAct = kOffKV + 4 * 2048 * 4;                     // u32 [4][4096]
constexpr size_t kOffPart = kOffAct + 4 * 4096 * 4;                   // u64 [4][16][kMaxSplits][66]
constexpr size_t kWsBytes = kOffPart + (size_t)4 * 16 * kMaxSplits * kPartWords * 8;

enum { PH_QKV = 0, PH_O = 1, PH_GU = 2, PH_DOWN = 3, PH_HEAD = 4 };
enum { TG_QKV = 0, TG_ATTN = 1, TG_O = 2, TG_GU = 3, TG_DOWN = 4 };   // tag index of a phase inside its layer

template <int PH> struct Ph;
template <> struct Ph<PH_QKV>  { static constexpr int K = 1024, NT = 2, KS = 8,  NG = 4; };
template <> struct Ph<PH_O>    { static constexpr int K = 1024, NT = 1, KS = 16, NG = 2; };
template <> struct Ph<PH_GU>   { static constexpr int K = 1024, NT = 4, KS = 4,  NG = 8; };
template <> struct Ph<PH_DOWN> { static constexpr int K = 4096, NT = 1, KS = 16, NG = 4; };
template <> struct Ph<PH_HEAD> { static constexpr int K = 1024, NT = 4, KS = 4,  NG = 8; };

// weight row of unit `u`, member ab (0 = "a" row in tile rows 0-7, 1 = "b" row in tile rows 8-15)
template <int PH>
__device__ __forceinline__ int row_of(int u, int ab) {
  if (PH == PH_QKV) return (u >> 5) * 64 + (u & 31) + 32 * ab;        // rotation partners i / i+32 of one head
  if (PH == PH_GU) return (u >> 4) * 32 + (u & 15) + 16 * ab;         // packed gate / up rows of feature u
  if (PH == PH_DOWN) return u;
  return 2 * u + ab;                                                  // O, HEAD: adjacent rows
}

__device__ __forceinline__ void bar_consumers() { __syncthreads(); }

// ---- tagged words ----------------------------------------------------------------------------
// 16-bit tags: every word of rows < T is rewritten by every launch, rows >= T are cleared at exit,
// so a stale word is at most one launch old and 160 tags per launch never alias.
__device__ __forceinline__ uint32_t tag16(uint32_t epoch, int phase_idx) {
  return (1u + (epoch * 160u + (uint32_t)phase_idx) % 65535u) << 16;
}
__device__ __forceinline__ uint32_t tag32(uint32_t epoch, int phase_idx) { return epoch * 160u + (uint32_t)phase_idx + 1u; }
__device__ __forceinline__ uint32_t tagged(float v, uint32_t tag_hi) {
  const bf16 b = __float2bfloat16_rn(v);
  return tag_hi | *reinterpret_cast<const unsigned short*>(&b);
}
__device__ __forceinline__ uint4 ld_relaxed4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool tags_ok(const uint4& v, uint32_t tag_hi) {
  return ((v.x ^ tag_hi) | (v.y ^ tag_hi) | (v.z ^ tag_hi) | (v.w ^ tag_hi)) < 0x10000u;
}
// N x 4 tagged words at p + i*stride_words: re-read until all carry the tag
template <int N>
__device__ __forceinline__ void poll_words(const uint32_t* p, int stride_words, uint32_t tag_hi, uint4 (&v)[N]) {
  bool ok;
  do {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = ld_relaxed4(p + (size_t)i * stride_words);
    ok = true;
#pragma unroll
    for (int i = 0; i < N; ++i) ok = ok && tags_ok(v[i], tag_hi);
  } while (!ok);
}
// two 4-word groups (8 consecutive elements) -> 8 packed bf16
__device__ __forceinline__ uint4 strip_tags(const uint4& lo, const uint4& hi) {
  return make_uint4((lo.x & 0xffffu) | (lo.y << 16), (lo.z & 0xffffu) | (lo.w << 16),
                    (hi.x & 0xffffu) | (hi.y << 16), (hi.z & 0xffffu) | (hi.w << 16));
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}

// ---- weight stream: the STATIC sequence of ring groups this CTA needs ---------------------------
// (layer 0: QKV 4 groups, O 2, GU 8, DOWN 4; layer 1 ...; HEAD 8).  The ring always holds the next
// kRing groups of the sequence: it is primed at kernel start and warp kNW-1 re-issues as many groups
// as a projection has just consumed, right after the consumers released them — so the copies never
// wait for activations and nobody ever waits for a free slot.
struct Units { int u0[5], n[5]; };                    // this CTA's unit range per projection (shared memory)
struct Cursor { int li, ph, gi; uint32_t pos; };      // next group to issue (warp-uniform registers)

__device__ __forceinline__ void issue_groups(const MegaArgs& a, const Units& un, Cursor& cu, int count, uint8_t* ring,
                                             uint64_t* full, uint64_t* empty, int lane) {
#pragma unroll 1
  for (int n = 0; n < count; ++n) {
    if (cu.li > a.n_layers || (cu.li == a.n_layers && cu.ph != PH_HEAD)) return;      // stream finished
    const int ph = cu.ph, gi = cu.gi;
    const int ng = ph == PH_QKV ? 4 : ph == PH_O ? 2 : ph == PH_DOWN ? 4 : 8;
    const int tile = ph == PH_DOWN ? 0 : gi >> 1, ab = ph == PH_DOWN ? 0 : gi & 1, kseg = ph == PH_DOWN ? gi : 0;
    int n_valid = un.n[ph] - tile * 8;
    n_valid = n_valid < 0 ? 0 : (n_valid > 8 ? 8 : n_valid);
    if (a.flags & 1) n_valid = 0;                      // debugging aid: no weight traffic (results are garbage)
    const uint32_t slot = cu.pos % kRing, par = (cu.pos / kRing) & 1;
    if (lane == 0) {
      mbar_wait(&empty[slot], par ^ 1);                // free by construction; kept as a guard
      if (n_valid > 0) mbar_arrive_expect_tx(&full[slot], n_valid * 2048);
      else mbar_arrive(&full[slot]);
    }
    __syncwarp();
    if (lane < n_valid) {
      const int u = un.u0[ph] + tile * 8 + lane;
      const bf16* W; int row, K = 1024;
      if (ph == PH_QKV) { W = a.wqkv[cu.li]; row = row_of<PH_QKV>(u, ab); }
      else if (ph == PH_O) { W = a.wo[cu.li]; row = row_of<PH_O>(u, ab); }
      else if (ph == PH_GU) { W = a.wgu[cu.li]; row = row_of<PH_GU>(u, ab); }
      else if (ph == PH_DOWN) { W = a.wd[cu.li]; row = u; K = 4096; }
      else { W = a.head_w; row = row_of<PH_HEAD>(u, ab); }
      bulk_g2s(ring + slot * kGroupBytes + lane * kRowStride, W + (size_t)row * K + kseg * 1024, 2048, &full[slot]);
    }
    ++cu.pos;
    if (++cu.gi == ng) {
      cu.gi = 0;
      if (ph == PH_HEAD) { cu.li = a.n_layers + 1; }
      else if (ph == PH_DOWN) { ++cu.li; cu.ph = cu.li == a.n_layers ? PH_HEAD : PH_QKV; }
      else ++cu.ph;
    }
  }
}

// ---- prologues: exchanged activations -> shared memory (bf16) ----------------------------------
// xs[t] = bf16(bf16(res[t]·rstd)·w)   (vLLM rms_norm rounding), warp t < T.  The residual row is the
// plain bf16 input of the step (layer 0) or 1024 tagged words.
__device__ __forceinline__ void norm_rows(const MegaArgs& a, bool plain, uint32_t tag_hi, const bf16* w, bf16* xs,
                                          bf16* trace_row0, int warp, int lane) {
  if (warp >= a.T) return;
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  uint4 wv[4], xv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) wv[i] = wr[lane + 32 * i];
  if (plain) {
    const uint4* xr = reinterpret_cast<const uint4*>(a.res_in + (size_t)warp * 1024);
#pragma unroll
    for (int i = 0; i < 4; ++i) xv[i] = __ldcg(xr + lane + 32 * i);
  } else {
    const uint32_t* xr = reinterpret_cast<const uint32_t*>(a.ws + kOffRes) + warp * 1024 + lane * 8;
    uint4 v[8];
    bool ok;
    do {
#pragma unroll
      for (int i = 0; i < 4; ++i) { v[2 * i] = ld_relaxed4(xr + i * 256); v[2 * i + 1] = ld_relaxed4(xr + i * 256 + 4); }
      ok = true;
#pragma unroll
      for (int i = 0; i < 8; ++i) ok = ok && tags_ok(v[i], tag_hi);
    } while (!ok);
#pragma unroll
    for (int i = 0; i < 4; ++i) xv[i] = strip_tags(v[2 * i], v[2 * i + 1]);
  }
  if (trace_row0) {
    uint4* tr = reinterpret_cast<uint4*>(trace_row0 + (size_t)warp * 1024);
#pragma unroll
    for (int i = 0; i < 4; ++i) tr[lane + 32 * i] = xv[i];
  }
  float f[4][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    unpack8(xv[i], f[i]);
#pragma unroll
    for (int e = 0; e < 8; ++e) ss += f[i][e] * f[i][e];
  }
  ss = warp_sum(ss);
  const float rstd = rsqrtf(ss * (1.0f / 1024.f) + a.eps);
  uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)warp * kXStride);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float g[8];
    unpack8(wv[i], g);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[i][e] = round_bf16(f[i][e] * rstd) * g[e];
    dst[lane + 32 * i] = make_uint4(pack_bf16(f[i][0], f[i][1]), pack_bf16(f[i][2], f[i][3]),
                                    pack_bf16(f[i][4], f[i][5]), pack_bf16(f[i][6], f[i][7]));
  }
}

// xs[t][0:4096] = act[t]: thread j owns elements [8j, 8j+8) of every row.  Spin on one 16-byte
// group per thread (148 CTAs re-reading all 32 KB per iteration would saturate L2), then read the rest.
__device__ __forceinline__ void copy_act(const MegaArgs& a, uint32_t tag_hi, bf16* xs) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(a.ws + kOffAct) + threadIdx.x * 8;
  uint4 v[8];
  do { v[0] = ld_relaxed4(src); } while (!tags_ok(v[0], tag_hi));
  bool ok;
  do {
    ok = true;
    v[1] = ld_relaxed4(src + 4);
    ok = ok && tags_ok(v[1], tag_hi);
#pragma unroll
    for (int t = 1; t < 4; ++t)
      if (t < a.T) {
        v[2 * t] = ld_relaxed4(src + t * 4096);
        v[2 * t + 1] = ld_relaxed4(src + t * 4096 + 4);
        ok = ok && tags_ok(v[2 * t], tag_hi) && tags_ok(v[2 * t + 1], tag_hi);
      }
  } while (!ok);
#pragma unroll
  for (int t = 0; t < 4; ++t)
    if (t < a.T) reinterpret_cast<uint4*>(xs + (size_t)t * kXStride)[threadIdx.x] = strip_tags(v[2 * t], v[2 * t + 1]);
}

// xs[t][h*64+d] = bf16(sum_s w_s·o_s / sum_s w_s·l_s): the attention splits of (t, h) merged.
// Every partial word is read once per CTA: the (m, l) pairs by thread (t, h), the o values by the
// thread that owns the output pair; all loads are in flight together.  NI rows x NS splits at most.
template <int NI, int NS>
__device__ __forceinline__ void merge_attn(const MegaArgs& a, uint32_t tag, const int* splits, bf16* xs, float* wts) {
  const uint64_t* part = reinterpret_cast<const uint64_t*>(a.ws + kOffPart);
  const int tid = threadIdx.x;
  const int h = tid >> 5, dp = tid & 31;              // output pair (t = it, h, 2dp) for it < T
  uint4 pv[NI][NS];
  uint4 ml[NS];
  const bool stat = tid < a.T * 16;                    // also owns the statistics of (t = tid >> 4, h = tid & 15)
  const int st_t = tid >> 4, st_h = tid & 15;
  // spin on the small (m, l) words; the o words of the same producers land within the same few hundred cycles
  if (stat) {
    bool sok;
    do {
      sok = true;
#pragma unroll
      for (int s = 0; s < NS; ++s)
        if (s < splits[st_t]) {
          ml[s] = ld_relaxed4(part + (size_t)((st_t * 16 + st_h) * kMaxSplits + s) * kPartWords + 64);
          sok = sok && ml[s].y == tag && ml[s].w == tag;
        }
    } while (!sok);
  }
  bar_consumers();
  bool ok;
  do {
    ok = true;
#pragma unroll
    for (int it = 0; it < NI; ++it)
      if (it < a.T) {
#pragma unroll
        for (int s = 0; s < NS; ++s)
          if (s < splits[it]) {
            pv[it][s] = ld_relaxed4(part + (size_t)((it * 16 + h) * kMaxSplits + s) * kPartWords + 2 * dp);
            ok = ok && pv[it][s].y == tag && pv[it][s].w == tag;
          }
      }
  } while (!ok);
  if (stat) {
    const int S = splits[st_t];
    float mx = -INFINITY;
#pragma unroll
    for (int s = 0; s < NS; ++s) if (s < S) mx = fmaxf(mx, __uint_as_float(ml[s].x));
    float wg[NS], lt = 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      wg[s] = 0.f;
      if (s < S && __uint_as_float(ml[s].x) > -INFINITY) {
        wg[s] = exp2f(__uint_as_float(ml[s].x) - mx);
        lt = fmaf(wg[s], __uint_as_float(ml[s].z), lt);
      }
    }
    const float inv = lt > 0.f ? 1.f / lt : 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) wts[tid * kMaxSplits + s] = wg[s] * inv;
  }
  bar_consumers();
#pragma unroll
  for (int it = 0; it < NI; ++it)
    if (it < a.T) {
      float o0 = 0.f, o1 = 0.f;
#pragma unroll
      for (int s = 0; s < NS; ++s)
        if (s < splits[it]) {
          const float wgt = wts[(it * 16 + h) * kMaxSplits + s];
          o0 = fmaf(wgt, __uint_as_float(pv[it][s].x), o0);
          o1 = fmaf(wgt, __uint_as_float(pv[it][s].z), o1);
        }
      *reinterpret_cast<uint32_t*>(xs + (size_t)it * kXStride + h * 64 + 2 * dp) = pack_bf16(o0, o1);
    }
}

// ---- epilogues: thread (unit u, row t) holds the two finished dot products va (a row), vb (b row)
struct EpiCtx {
  uint32_t tag_hi;            // tag of the words this phase writes
  bool first;                 // layer 0: the residual input is the plain bf16 buffer
  const float2* cs;           // QKV, shared memory [64]: cos/sin of (pos[t], i) for epilogue thread threadIdx.x
  const int64_t* slot;        // QKV, shared memory [4]: cache slot of row t (-1: do not cache)
};

// the residual words this thread will update (O: pair 2u, 2u+1; DOWN: element u), issued before the MMA
template <int PH>
__device__ __forceinline__ uint2 prefetch_old(const MegaArgs& a, const EpiCtx& c, int u, int t) {
  uint2 r = make_uint2(0, 0);
  const uint32_t* xres = reinterpret_cast<const uint32_t*>(a.ws + kOffRes);
  if (PH == PH_O) {
    if (c.first) {
      const uint32_t v = __ldcg(reinterpret_cast<const uint32_t*>(a.res_in + (size_t)t * 1024 + 2 * u));
      r = make_uint2(v & 0xffffu, v >> 16);
    } else {
      r = __ldcg(reinterpret_cast<const uint2*>(xres + t * 1024 + 2 * u));
    }
  } else if (PH == PH_DOWN) {
    r.x = __ldcg(xres + t * 1024 + u);
  }
  return r;
}

template <int PH>
__device__ __forceinline__ void epilogue(const MegaArgs& a, const EpiCtx& c, uint8_t* kv_layer, int u, int t, float va,
                                         float vb, uint2 old) {
  if (PH == PH_QKV) {
    const int head48 = u >> 5, i = u & 31, which = head48 >> 4, h = head48 & 15;
    float o1 = round_bf16(va), o2 = round_bf16(vb);
    if (which < 2) {
      const float2 cs = c.cs[threadIdx.x];
      const float x1 = o1, x2 = o2;
      o1 = x1 * cs.x - x2 * cs.y;
      o2 = x2 * cs.x + x1 * cs.y;
    }
    const uint32_t w1 = tagged(o1, c.tag_hi), w2 = tagged(o2, c.tag_hi);
    if (which == 0) {
      uint32_t* q = reinterpret_cast<uint32_t*>(a.ws + kOffQ) + t * 1024 + h * 64;
      q[i] = w1; q[i + 32] = w2;
    } else {
      uint32_t* kv = reinterpret_cast<uint32_t*>(a.ws + kOffKV) + (t * 2 + (which - 1)) * 1024 + h * 64;
      kv[i] = w1; kv[i + 32] = w2;
      const int64_t sl = c.slot[t];
      if (sl >= 0) {                                    // the cache copy is for later steps
        uint8_t* page = kv_layer + (size_t)(sl / CBX_BLOCK_SIZE) * CBX_PAGE_BYTES;
        const int tok = (int)(sl % CBX_BLOCK_SIZE);
        const unsigned short b1 = (unsigned short)w1, b2 = (unsigned short)w2;
        if (which == 1) {                               // K: 16-byte chunks XOR-swizzled by tok & 7
          uint8_t* kt = page + h * 2048 + tok * 128;
          const int c1 = (i >> 3) ^ (tok & 7), c2 = ((i + 32) >> 3) ^ (tok & 7);
          *reinterpret_cast<unsigned short*>(kt + c1 * 16 + (i & 7) * 2) = b1;
          *reinterpret_cast<unsigned short*>(kt + c2 * 16 + (i & 7) * 2) = b2;
        } else {
          unsigned short* vt = reinterpret_cast<unsigned short*>(page + 32768 + h * 2048 + tok * 128);
          vt[i] = b1; vt[i + 32] = b2;
        }
      }
    }
  } else if (PH == PH_O) {
    uint2* rp = reinterpret_cast<uint2*>(reinterpret_cast<uint32_t*>(a.ws + kOffRes) + t * 1024 + 2 * u);
    *rp = make_uint2(tagged(round_bf16(va) + __uint_as_float(old.x << 16), c.tag_hi),
                     tagged(round_bf16(vb) + __uint_as_float(old.y << 16), c.tag_hi));
  } else if (PH == PH_GU) {
    const float g = round_bf16(va), up = round_bf16(vb);
    reinterpret_cast<uint32_t*>(a.ws + kOffAct)[t * 4096 + u] = tagged(g / (1.f + __expf(-g)) * up, c.tag_hi);
  } else if (PH == PH_DOWN) {
    reinterpret_cast<uint32_t*>(a.ws + kOffRes)[t * 1024 + u] = tagged(round_bf16(va) + __uint_as_float(old.x << 16), c.tag_hi);
  } else {                                              // PH_HEAD
    *reinterpret_cast<uint32_t*>(a.logits + (size_t)t * a.ldl + 2 * u) = pack_bf16(va, vb);
  }
}

__device__ __forceinline__ void mbar_arrive_n(uint64_t* bar, uint32_t n) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n) : "memory");
}

#ifdef CBX_MEGA_PROF                                  // per-phase cycle stamps of one CTA, printed once per launch
#define PROF_CTA 3
#define STAMP(i) do { if (threadIdx.x == 0 && blockIdx.x == PROF_CTA && li == 7) stamps[i] = clock64(); } while (0)
#else
#define STAMP(i)
#endif

// ---- one projection: this CTA's <= NT tiles of 16 weight rows x K, K split over KS warps ------
template <int PH>
__device__ __forceinline__ void gemv_phase(const MegaArgs& a, const EpiCtx& c, uint8_t* kv_layer, const Units& un, Cursor& cu,
                                           uint8_t* ring, uint64_t* full, uint64_t* empty, uint32_t& cnt, const bf16* xs,
                                           float* red, int warp, int lane) {
  using P = Ph<PH>;
  const int u0 = un.u0[PH], n_units = un.n[PH];
  constexpr int KW = P::K / P::KS;                      // K elements per warp
  constexpr int NSTEP = KW / 16;
  const int tile = warp / P::KS, ks = warp % P::KS;
  // ring groups holding this warp's tile: rows 0-7 ("a"), rows 8-15 ("b"; DOWN has none)
  const int ga = PH == PH_DOWN ? (ks >> 2) : 2 * tile;
  const int gb = PH == PH_DOWN ? ga : ga + 1;
  const int koff = PH == PH_DOWN ? (ks & 3) * 256 : ks * KW;          // inside the 1024-element row segment
  const int kx = ks * KW;                                             // inside the activation row
  // epilogue thread (tile tl, unit slot eu, row et): start its residual read now
  const int tl = threadIdx.x >> 5, eu = (threadIdx.x >> 2) & 7, et = threadIdx.x & 3;
  const int ul = tl * 8 + eu;
  const bool epi = threadIdx.x < P::NT * 32 && ul < n_units && et < a.T;
  uint2 old = make_uint2(0, 0);
  if (epi && (PH == PH_O || PH == PH_DOWN)) old = prefetch_old<PH>(a, c, u0 + ul, et);

  const uint32_t pa = cnt + ga, pb = cnt + gb;
  const uint32_t slot_a = pa % kRing, slot_b = pb % kRing;
  mbar_wait(&full[slot_a], (pa / kRing) & 1);
  if (PH != PH_DOWN) mbar_wait(&full[slot_b], (pb / kRing) & 1);
  const uint32_t sa = smem_u32(ring + slot_a * kGroupBytes), sb = smem_u32(ring + slot_b * kGroupBytes);
  const int mat = lane >> 3, r8 = lane & 7;
  const uint32_t addr = ((mat & 1) ? sb : sa) + r8 * kRowStride + (koff + (mat >> 1) * 8) * 2;
  const int g = lane >> 2, cq = lane & 3;
  const uint32_t* xb = reinterpret_cast<const uint32_t*>(xs + (size_t)(g & 3) * kXStride + kx + 2 * cq);
  const uint32_t live = g < a.T ? 0xffffffffu : 0u;    // rows >= T of the B operand are zero
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
  for (int s0 = 0; s0 < NSTEP; s0 += 8) {
    constexpr int NB = NSTEP < 8 ? NSTEP : 8;
    uint32_t af[NB][4], bfr[NB][2];
#pragma unroll
    for (int s = 0; s < NB; ++s) {
      asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                   : "=r"(af[s][0]), "=r"(af[s][1]), "=r"(af[s][2]), "=r"(af[s][3]) : "r"(addr + (s0 + s) * 32));
      bfr[s][0] = xb[(s0 + s) * 8] & live;
      bfr[s][1] = xb[(s0 + s) * 8 + 4] & live;
    }
#pragma unroll
    for (int s = 0; s < NB; ++s)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(acc[s & 1][0]), "+f"(acc[s & 1][1]), "+f"(acc[s & 1][2]), "+f"(acc[s & 1][3])
                   : "r"(af[s][0]), "r"(af[s][1]), "r"(af[s][2]), "r"(af[s][3]), "r"(bfr[s][0]), "r"(bfr[s][1]));
  }
  __syncwarp();
  if (lane == 0) {                                     // 16 arrivals free a group: weight by the number of reader warps
    mbar_arrive_n(&empty[slot_a], PH == PH_DOWN ? 4 : kNW / P::KS);
    if (PH != PH_DOWN) mbar_arrive_n(&empty[slot_b], kNW / P::KS);
  }
  cnt += P::NG;
  // C fragment: [0],[1] = (row g, t = 2cq, 2cq+1); [2],[3] = (row g+8, same t)
  if (cq < 2) {
    float* mine = red + warp * 64;
    *reinterpret_cast<float2*>(mine + g * 4 + 2 * cq) = make_float2(acc[0][0] + acc[1][0], acc[0][1] + acc[1][1]);
    *reinterpret_cast<float2*>(mine + (g + 8) * 4 + 2 * cq) = make_float2(acc[0][2] + acc[1][2], acc[0][3] + acc[1][3]);
  }
  bar_consumers();
  if (warp == kNW - 1) issue_groups(a, un, cu, P::NG, ring, full, empty, lane);      // refill what was just released
  if (epi) {
    float va = 0.f, vb = 0.f;
#pragma unroll
    for (int k = 0; k < P::KS; ++k) {
      const float* pw = red + (tl * P::KS + k) * 64;
      va += pw[eu * 4 + et];
      vb += pw[(eu + 8) * 4 + et];
    }
    epilogue<PH>(a, c, kv_layer, u0 + ul, et, va, vb, old);
  }
}

// ---- attention -----------------------------------------------------------------------------------
// A CTA owns one (row, head, split) item for the whole kernel; a warp takes pages sp + S·(warp + 16n)
// of the row, one page (16 tokens) at a time, K/V rows straight from global memory.  Same lane
// mappings and fp32 online softmax as k_attn.cu.  The token written by THIS step is not read from
// the cache (that store may still be in flight): its k/v rows come from the tagged exchange buffer.
struct AttnItem {
  int row, h, sp, S, len, n_local;                   // n_local = pages of this split
  int last_j;                                        // local index of the page holding the new token, -1 if another split's
  const int32_t* btr;
};

__device__ __forceinline__ void attn_phase(const MegaArgs& a, const AttnItem& it, const uint8_t* kv_layer, uint32_t tq_hi,
                                           uint32_t tpart, float* red, float* q_s, int warp, int lane) {
  const int t16 = lane & 15, half = lane >> 4;        // QK mapping: (token, half of d)
  const int tg = lane >> 3, c = lane & 7;             // PV mapping: (4-token group, 8-dim chunk)
  const float qscale = 0.125f * kLog2e;               // 1/sqrt(64), scores in the log2 domain
  uint4 kr[4], vr[4];
  auto load_page = [&](int j) {                        // cache rows of local page j -> registers
    const int page = it.sp + j * it.S;
    const uint8_t* kt = kv_layer + (size_t)it.btr[page] * CBX_PAGE_BYTES + it.h * 2048;
    const uint8_t* vt = kt + 32768;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      kr[i] = __ldcg(reinterpret_cast<const uint4*>(kt + t16 * 128 + (((half * 4 + i) ^ (t16 & 7)) << 4)));
      vr[i] = __ldcg(reinterpret_cast<const uint4*>(vt + (tg * 4 + i) * 128 + (c << 4)));
    }
  };
  int j = warp;
  if (j < it.n_local) load_page(j);                    // in flight while q is still being produced
  // q[row][h]: polled once per CTA (16 lanes x 4 words), shared through q_s pre-scaled
  if (warp == 0 && lane < 16) {
    uint4 v;
    const uint32_t* qp = reinterpret_cast<const uint32_t*>(a.ws + kOffQ) + it.row * 1024 + it.h * 64 + lane * 4;
    do { v = ld_relaxed4(qp); } while (!tags_ok(v, tq_hi));
    *reinterpret_cast<float4*>(q_s + lane * 4) = make_float4(__uint_as_float(v.x << 16) * qscale, __uint_as_float(v.y << 16) * qscale,
                                                             __uint_as_float(v.z << 16) * qscale, __uint_as_float(v.w << 16) * qscale);
  }
  bar_consumers();
  const float4* qh = reinterpret_cast<const float4*>(q_s + half * 32);   // broadcast reads: q stays out of registers
  float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  while (j < it.n_local) {
    const int page = it.sp + j * it.S;
    const int n_valid = it.len - page * CBX_BLOCK_SIZE;
    if (j == it.last_j) {                              // patch in the new token's k / v rows (only the lanes that use them poll)
      const int tn = (it.len - 1) & 15;
      const uint32_t* nk = reinterpret_cast<const uint32_t*>(a.ws + kOffKV) + (it.row * 2) * 1024 + it.h * 64;
      const bool want_k = t16 == tn, want_v = tg == (tn >> 2);
      uint4 kn[8], vn[2];
      bool ok;
      do {
        ok = true;
        if (want_k) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { kn[i] = ld_relaxed4(nk + half * 32 + i * 4); ok = ok && tags_ok(kn[i], tq_hi); }
        }
        if (want_v) {
          vn[0] = ld_relaxed4(nk + 1024 + c * 8);
          vn[1] = ld_relaxed4(nk + 1024 + c * 8 + 4);
          ok = ok && tags_ok(vn[0], tq_hi) && tags_ok(vn[1], tq_hi);
        }
      } while (!ok);
      if (want_k) {
#pragma unroll
        for (int i = 0; i < 4; ++i) kr[i] = strip_tags(kn[2 * i], kn[2 * i + 1]);
      }
      if (want_v) {
        const uint4 vnew = strip_tags(vn[0], vn[1]);
#pragma unroll
        for (int i = 0; i < 4; ++i) if ((tn & 3) == i) vr[i] = vnew;
      }
    }
    float sc = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float f[8];
      unpack8(kr[i], f);
      const float4 qa = qh[2 * i], qb = qh[2 * i + 1];
      sc = fmaf(qa.x, f[0], sc); sc = fmaf(qa.y, f[1], sc); sc = fmaf(qa.z, f[2], sc); sc = fmaf(qa.w, f[3], sc);
      sc = fmaf(qb.x, f[4], sc); sc = fmaf(qb.y, f[5], sc); sc = fmaf(qb.z, f[6], sc); sc = fmaf(qb.w, f[7], sc);
    }
    sc += __shfl_xor_sync(0xffffffffu, sc, 16);
    sc = (t16 < n_valid) ? sc : -INFINITY;
    float mt = sc;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, o));
    const float m_new = fmaxf(m, mt);
    const float p = exp2f(sc - m_new);
    const float corr = exp2f(m - m_new);
    m = m_new;
    l = fmaf(l, corr, p);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= corr;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int tok = tg * 4 + i;
      const float pt = __shfl_sync(0xffffffffu, p, tok);
      if (tok < n_valid) {
        float f[8];
        unpack8(vr[i], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(pt, f[e], acc[e]);
      }
    }
    j += kNW;
    if (j < it.n_local) load_page(j);
  }
  // warp partial -> shared memory: o[64], m, l (l is held twice per token: both d-halves)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  const float lsum = warp_sum(l) * 0.5f;
  float* mine = red + warp * 68;
  if (lane < 8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) mine[lane * 8 + i] = acc[i];
  }
  if (lane == 0) { mine[64] = m; mine[65] = lsum; }
  bar_consumers();
  if (threadIdx.x < 64) {                              // merge the 16 warps, one thread per output dim
    const int d = threadIdx.x;
    float mx = -INFINITY;
#pragma unroll
    for (int w = 0; w < kNW; ++w) mx = fmaxf(mx, red[w * 68 + 64]);
    float o = 0.f, lt = 0.f;
    if (mx > -INFINITY) {
#pragma unroll
      for (int w = 0; w < kNW; ++w) {
        const float pm = red[w * 68 + 64];
        const float wgt = pm > -INFINITY ? exp2f(pm - mx) : 0.f;
        o = fmaf(wgt, red[w * 68 + d], o);
        lt = fmaf(wgt, red[w * 68 + 65], lt);
      }
    }
    uint2* dst = reinterpret_cast<uint2*>(a.ws + kOffPart) + (size_t)((it.row * 16 + it.h) * kMaxSplits + it.sp) * kPartWords;
    dst[d] = make_uint2(__float_as_uint(o), tpart);
    if (d == 0) {                                      // (m, l) leave as one 16-byte store
      *reinterpret_cast<uint4*>(dst + 64) = make_uint4(__float_as_uint(mx), tpart, __float_as_uint(lt), tpart);
    }
  }
}

__device__ __forceinline__ void unit_range(int n, int& u0, int& cnt) {
  u0 = (int)((int64_t)n * blockIdx.x / gridDim.x);
  cnt = (int)((int64_t)n * (blockIdx.x + 1) / gridDim.x) - u0;
}
// KV splits of a row: one page per warp per split when the grid allows it
__device__ __forceinline__ int splits_for(int len, int smax) {
  const int n_pages = (len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  const int s = (n_pages + kNW - 1) / kNW;
  return s < 1 ? 1 : (s > smax ? smax : s);
}

__global__ void __launch_bounds__(kThreads, 1) decode_mega_kernel(const __grid_constant__ MegaArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  bf16* xs = reinterpret_cast<bf16*>(smem + kRing * kGroupBytes);
  float* red = reinterpret_cast<float*>(smem + kRing * kGroupBytes + kXBytes);
  float* wts = red + kRedFloats;
  uint64_t* full = reinterpret_cast<uint64_t*>(wts + kWtsFloats);
  uint64_t* empty = full + kRing;
  __shared__ int splits[4];
  __shared__ float red_attn[kRedFloats];
  __shared__ __align__(16) float q_s[64];
  __shared__ AttnItem item;
  __shared__ float2 s_cs[64];
  __shared__ int64_t s_slot[4];
#ifdef CBX_MEGA_PROF
  __shared__ long long stamps[16];
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int smax = a.T <= 2 ? 4 : 2;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kRing; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kNW); }
    fence_barrier_init();
  }
  if (threadIdx.x < 4) splits[threadIdx.x] = threadIdx.x < a.T ? splits_for(a.kv_len[threadIdx.x], smax) : 0;
  __syncthreads();

  __shared__ Units un;
  if (threadIdx.x == 0) {
    unit_range(1536, un.u0[PH_QKV], un.n[PH_QKV]);
    unit_range(512, un.u0[PH_O], un.n[PH_O]);
    unit_range(4096, un.u0[PH_GU], un.n[PH_GU]);
    unit_range(1024, un.u0[PH_DOWN], un.n[PH_DOWN]);
    unit_range(a.vocab / 2, un.u0[PH_HEAD], un.n[PH_HEAD]);
  }
  __syncthreads();
  uint32_t cnt = 0;
  Cursor cu = {0, PH_QKV, 0, 0};
  if (warp == kNW - 1) issue_groups(a, un, cu, kRing, ring, full, empty, lane);       // prime the ring

  // ---- launch-constant state
  const uint32_t epoch = a.sync[2];
  // rows >= T of the 16-bit-tagged buffers are not written by this launch: invalidate them so that a
  // later launch with more rows cannot meet an old word whose tag has come round again
  if (a.T < 4) {
    const int per_row = 1024 + 1024 + 2048 + 4096, n = (4 - a.T) * per_row;
    for (int i = blockIdx.x * kConsumers + threadIdx.x; i < n; i += gridDim.x * kConsumers) {
      const int t = a.T + i / per_row, w = i % per_row;
      uint32_t* p = w < 1024 ? reinterpret_cast<uint32_t*>(a.ws + kOffRes) + t * 1024 + w
                  : w < 2048 ? reinterpret_cast<uint32_t*>(a.ws + kOffQ) + t * 1024 + (w - 1024)
                  : w < 4096 ? reinterpret_cast<uint32_t*>(a.ws + kOffKV) + t * 2048 + (w - 2048)
                             : reinterpret_cast<uint32_t*>(a.ws + kOffAct) + t * 4096 + (w - 4096);
      *p = 0;
    }
  }
  // attention item of this CTA
  if (threadIdx.x == 0) {
    AttnItem im = {};
    im.n_local = 0; im.last_j = -1;
    int base = 0;
    for (int t = 0; t < a.T; ++t) {
      const int n = 16 * splits[t];
      if ((int)blockIdx.x >= base && (int)blockIdx.x < base + n) {
        const int r = blockIdx.x - base;
        im.row = t; im.S = splits[t]; im.h = r / im.S; im.sp = r % im.S;
        im.len = a.kv_len[t];
        const int n_pages = (im.len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
        im.n_local = n_pages > im.sp ? (n_pages - im.sp + im.S - 1) / im.S : 0;
        const int pl = n_pages - 1;
        im.last_j = (pl >= 0 && pl % im.S == im.sp) ? pl / im.S : -1;
        im.btr = a.bt + (int64_t)a.row_seq[t] * a.bt_stride;
      }
      base += n;
    }
    item = im;
  }
  // QKV epilogue constants of thread (tile tl, unit eu, row et)
  if (threadIdx.x < 64) {
    const int tl = threadIdx.x >> 5, eu = (threadIdx.x >> 2) & 7, et = threadIdx.x & 3;
    const int ul = tl * 8 + eu;
    float2 cs = make_float2(1.f, 0.f);
    if (ul < un.n[PH_QKV] && et < a.T) {
      const int u = un.u0[PH_QKV] + ul, i = u & 31, which = (u >> 5) >> 4;
      if (which < 2) {
        const int p = a.pos[et];
        cs = make_float2(__bfloat162float(a.cos_sin[(size_t)p * 64 + i]), __bfloat162float(a.cos_sin[(size_t)p * 64 + 32 + i]));
      }
    }
    s_cs[threadIdx.x] = cs;
    if (threadIdx.x < 4) s_slot[threadIdx.x] = threadIdx.x < a.T ? a.slot[threadIdx.x] : -1;
  }
  bar_consumers();
  const bool has_item = item.S > 0;
  EpiCtx ec = {};
  ec.cs = s_cs; ec.slot = s_slot;
  const bool tracer = a.trace != nullptr && blockIdx.x == 0;

#pragma unroll 1
  for (int li = 0; li < a.n_layers; ++li) {
    uint8_t* kv_layer = a.kv_cache + (size_t)li * a.layer_bytes;
    const int p0 = li * 5;
    ec.first = li == 0;
    STAMP(0);
    norm_rows(a, li == 0, tag16(epoch, p0 - 5 + TG_DOWN), a.ln1[li], xs,
              tracer && li > 0 ? a.trace + (size_t)(li - 1) * a.T * 1024 : nullptr, warp, lane);
    bar_consumers();
    STAMP(1);
    ec.tag_hi = tag16(epoch, p0 + TG_QKV);
    gemv_phase<PH_QKV>(a, ec, kv_layer, un, cu, ring, full, empty, cnt, xs, red, warp, lane);
    STAMP(2);
    if (has_item) attn_phase(a, item, kv_layer, tag16(epoch, p0 + TG_QKV), tag32(epoch, p0 + TG_ATTN), red_attn, q_s, warp, lane);
    STAMP(3);
    if (a.T <= 2) merge_attn<2, 4>(a, tag32(epoch, p0 + TG_ATTN), splits, xs, wts);
    else merge_attn<4, 2>(a, tag32(epoch, p0 + TG_ATTN), splits, xs, wts);
    bar_consumers();
    STAMP(4);
    ec.tag_hi = tag16(epoch, p0 + TG_O);
    gemv_phase<PH_O>(a, ec, kv_layer, un, cu, ring, full, empty, cnt, xs, red, warp, lane);
    STAMP(5);
    norm_rows(a, false, tag16(epoch, p0 + TG_O), a.ln2[li], xs, nullptr, warp, lane);
    bar_consumers();
    STAMP(6);
    ec.tag_hi = tag16(epoch, p0 + TG_GU);
    gemv_phase<PH_GU>(a, ec, kv_layer, un, cu, ring, full, empty, cnt, xs, red, warp, lane);
    STAMP(7);
    copy_act(a, tag16(epoch, p0 + TG_GU), xs);
    bar_consumers();
    STAMP(8);
    ec.tag_hi = tag16(epoch, p0 + TG_DOWN);
    gemv_phase<PH_DOWN>(a, ec, kv_layer, un, cu, ring, full, empty, cnt, xs, red, warp, lane);
    STAMP(9);
  }
#ifdef CBX_MEGA_PROF
  if (threadIdx.x == 0 && blockIdx.x == PROF_CTA)
    printf("mega prof (cycles): norm %lld qkv %lld attn %lld merge %lld o %lld norm %lld gu %lld copy %lld down %lld | layer %lld\n",
           stamps[1] - stamps[0], stamps[2] - stamps[1], stamps[3] - stamps[2], stamps[4] - stamps[3], stamps[5] - stamps[4],
           stamps[6] - stamps[5], stamps[7] - stamps[6], stamps[8] - stamps[7], stamps[9] - stamps[8], stamps[9] - stamps[0]);
#endif
  norm_rows(a, false, tag16(epoch, (a.n_layers - 1) * 5 + TG_DOWN), a.norm_w, xs,
            tracer ? a.trace + (size_t)(a.n_layers - 1) * a.T * 1024 : nullptr, warp, lane);
  bar_consumers();
  gemv_phase<PH_HEAD>(a, ec, nullptr, un, cu, ring, full, empty, cnt, xs, red, warp, lane);
  // the last CTA out advances the epoch for the next launch
  if (threadIdx.x == 0) {
    const unsigned old = atomicAdd(a.sync + 1, 1u);
    if (old == gridDim.x - 1) { a.sync[1] = 0; a.sync[2] = epoch + 1; __threadfence(); }
  }
}
}  // namespace

namespace cbx {
size_t mega_ws_bytes() { return kWsBytes; }

int init_mega() {
  cudaError_t e = cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  CBX_REQUIRE(e == cudaSuccess, "mega smem attribute: %s", cudaGetErrorString(e));
  return 0;
}

// 0 = launched; 1 = shape not supported by this kernel (caller uses the per-projection kernels); < 0 error
int launch_decode_mega(const MegaArgs& a, cudaStream_t s) {
  static int n_sm = 0, coop = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, decode_mega_kernel, kThreads, kSmem);
    if (per_sm < 1) {
      fprintf(stderr, "libcbx_t3: persistent decode kernel does not fit this device (occupancy 0); using per-projection kernels\n");
      n_sm = -1;
    }
  }
  const int G = n_sm;
  auto fits = [&](int units, int cap) { return (units + G - 1) / G <= cap; };
  if (G < 16 * 4 * 2 || !coop || a.T < 1 || a.T > 4 || a.n_layers > 32 || (a.vocab & 1) || !fits(1536, 16) || !fits(512, 8) ||
      !fits(4096, 32) || !fits(1024, 8) || !fits(a.vocab / 2, 32))
    return 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(G); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, decode_mega_kernel, a);
  CBX_REQUIRE(e == cudaSuccess, "decode_mega launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
}  // namespace cbx
