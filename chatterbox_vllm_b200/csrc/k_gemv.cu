// k_gemv.cu — fused weight-streaming GEMV family for T <= 8 rows (1-4 CFG pairs: the latency / small-server decode
// regime), one kernel per projection of a Llama layer:
//
//   QKV   : RMSNorm(res)·ln1 -> x·Wqkv^T -> RoPE -> q to the qkv buffer, k/v into the KV page
//           (vllm llama.py:316-320,221-223 + reshape_and_cache_flash, flash_attn.py:850-874)
//   O     : attn·Wo^T, epilogue res = bf16(bf16(y) + res)            (llama.py:225,323-326)
//   GU    : RMSNorm(res)·ln2 -> x·Wgu^T -> SiLU(gate)·up            (llama.py:117-119)
//   DOWN  : act·Wd^T (K = 4096, split over 4 warps), epilogue res += (llama.py:120,316)
//   HEAD  : RMSNorm(res)·norm -> x·Whead^T -> bf16 logits           (llama.py:430, t3.py:729-730)
//
// HBM-bound: every weight byte is read exactly once with 128-bit L1-bypassing loads; a warp owns
// two output features (the RoPE pair i / i+32, the gate/up pair, or two adjacent rows) over a
// 1024-element K segment = 4 KB in flight per warp, all of it issued BEFORE griddepcontrol.wait
// so the weight stream of kernel n+1 overlaps the tail of kernel n (programmatic dependent
// launch).  The T x 1024 activation rows are normalised redundantly by every CTA (2-8 KB from
// L2) — cheaper than a separate launch.
//
// W8 = true: FP8-e4m3 weights with one fp32 scale per output feature (the optional weight-only
// FP8 path of the north star).  The stream halves (2 KB per warp per item); bytes are widened to f16
// with integer ops in registers (common.cuh), accumulation stays fp32, and the scale multiplies the finished
// dot product before the first bf16 rounding point, so y = bf16((x . q) * s) == x . (q * s) up to
// fp32 summation order — exactly the oracle's "dequantised weights" chain.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"

using namespace cbx;

namespace {
enum { PRO_COPY = 0, PRO_NORM = 1 };
enum { EPI_STORE = 0, EPI_SILU = 1, EPI_ROPE = 2, EPI_RES = 3 };

struct GemvArgs {
  const bf16* x;          // PRO_COPY: [T, K] input rows;  PRO_NORM: residual stream [T, 1024]
  const bf16* ln_w;       // PRO_NORM: RMSNorm weight [1024]
  const void* W;          // [N, K] bf16, or e4m3 bytes when W8
  const float* wscale;    // W8: [N] per-output-feature scale
  bf16* y;                // STORE/SILU: out [T, ldy];  RES: residual stream [T, 1024];  ROPE: qkv [T, 3072]
  const bf16* cos_sin;    // ROPE: [max_pos, 64] = cos(32) | sin(32)
  const int32_t* pos;     // ROPE: [T]
  const int64_t* slot;    // ROPE: [T] (-1 = do not cache)
  uint8_t* kv_layer;      // ROPE: this layer's pages
  int T, N, ldy;
  float eps;
  const void* next_w;     // weights of a later kernel in the chain, prefetched into L2 (may be NULL)
  size_t next_bytes;
  // after the fused attention block (k_block.cu): the O-proj output of the T rows waits in a 64-bit fixed-point
  // accumulator.  PRO_NORM with acc != NULL first closes the residual branch, x = bf16(bf16(acc * 2^-32) + x), CTA 0
  // writing the rows back to res_wb; a kernel given zero_ptr clears that many bytes after its wait (the down-proj
  // kernel hands the accumulator back empty).
  const unsigned long long* acc;
  bf16* res_wb;           // NOT the buffer x points at: the other CTAs are still reading that one
  const bf16* res_in;     // EPI_RES: the residual that is added (NULL = y, in place)
  void* zero_ptr;
  int zero_bytes;
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}

// 16 e4m3 bytes -> 16 floats = value / 256 (exact; integer widening to f16, common.cuh)
__device__ __forceinline__ void unpack16_e4m3(const uint4& v, float (&f)[16]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint32_t lo, hi;
    e4m3x4_to_f16x2_div256(w[i], lo, hi);
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    f[4 * i] = a.x; f[4 * i + 1] = a.y; f[4 * i + 2] = b.x; f[4 * i + 3] = b.y;
  }
}

template <int EPI>
__device__ __forceinline__ void item_rows(int item, int& ra, int& rb) {
  if (EPI == EPI_SILU)      { ra = (item >> 4) * 32 + (item & 15); rb = ra + 16; }   // packed gate/up
  else if (EPI == EPI_ROPE) { ra = (item >> 5) * 64 + (item & 31); rb = ra + 32; }   // rotation partners
  else                      { ra = 2 * item; rb = ra + 1; }
}

// Epilogue of one (item, row t): v0 / v1 are the finished dot products of weight rows ra / rb.
// Operands of the epilogue that do not depend on the dot products: fetched right after griddepcontrol.wait for a
// warp's first item, so their L2 round trips (pos -> cos/sin is two dependent ones) overlap the prologue and the
// FMAs instead of following them on the kernel-to-kernel critical path.
struct EpiPrefetch {
  bool valid;
  uint32_t res;          // EPI_RES: the residual pair
  float c, s;            // EPI_ROPE: cos / sin of (pos[t], i)
  int64_t slot;          // EPI_ROPE: KV slot of row t
  int pos;               // EPI_ROPE: position of row t (first-level load; cos / sin follow in phase B)
};

// phase A, right after the wait: loads whose addresses are known
template <int EPI>
__device__ __forceinline__ EpiPrefetch gemv_epilogue_prefetch(const GemvArgs& a, int item, int t) {
  EpiPrefetch pf = {true, 0u, 1.f, 0.f, -1, 0};
  if (EPI == EPI_RES) {
    pf.res = *reinterpret_cast<const uint32_t*>((a.res_in ? a.res_in : a.y) + (size_t)t * 1024 + 2 * item);
  } else if (EPI == EPI_ROPE) {
    const int which = item >> 9;                       // (item >> 5) >> 4
    if (which < 2) pf.pos = a.pos[t];
    if (which > 0) pf.slot = a.slot[t];
  }
  return pf;
}
// phase B, after the activation rows have been requested: the loads that depend on phase A (cos / sin of pos) —
// placed earlier they would stall the in-order issue of everything behind them for one L2 round trip
template <int EPI>
__device__ __forceinline__ void gemv_epilogue_prefetch2(const GemvArgs& a, int item, EpiPrefetch& pf) {
  if (EPI == EPI_ROPE && (item >> 9) < 2) {
    const int i = item & 31;
    pf.c = __bfloat162float(a.cos_sin[(size_t)pf.pos * 64 + i]);
    pf.s = __bfloat162float(a.cos_sin[(size_t)pf.pos * 64 + 32 + i]);
  }
}

template <int EPI>
__device__ __forceinline__ void gemv_epilogue(const GemvArgs& a, int item, int ra, int rb, int t, float v0, float v1,
                                              const EpiPrefetch pf = EpiPrefetch{false, 0u, 1.f, 0.f, -1, 0}) {
  if (EPI == EPI_STORE) {
    *reinterpret_cast<uint32_t*>(a.y + (size_t)t * a.ldy + ra) = pack_bf16(v0, v1);
  } else if (EPI == EPI_SILU) {
    const float g = round_bf16(v0), u = round_bf16(v1);
    a.y[(size_t)t * a.ldy + item] = __float2bfloat16_rn(g / (1.f + __expf(-g)) * u);
  } else if (EPI == EPI_RES) {
    uint32_t* rp = reinterpret_cast<uint32_t*>(a.y + (size_t)t * 1024 + ra);
    const uint32_t r = pf.valid ? pf.res : *reinterpret_cast<const uint32_t*>((a.res_in ? a.res_in : a.y) + (size_t)t * 1024 + ra);
    *rp = pack_bf16(round_bf16(v0) + bf_lo(r), round_bf16(v1) + bf_hi(r));
  } else {                                        // EPI_ROPE
    const int head48 = item >> 5, i = item & 31, which = head48 >> 4, h = head48 & 15;
    float o1 = round_bf16(v0), o2 = round_bf16(v1);
    if (which < 2) {
      float c, s;
      if (pf.valid) { c = pf.c; s = pf.s; }
      else {
        const int p = a.pos[t];
        c = __bfloat162float(a.cos_sin[(size_t)p * 64 + i]);
        s = __bfloat162float(a.cos_sin[(size_t)p * 64 + 32 + i]);
      }
      const float x1 = o1, x2 = o2;
      o1 = x1 * c - x2 * s;
      o2 = x2 * c + x1 * s;
    }
    const bf16 b1 = __float2bfloat16_rn(o1), b2 = __float2bfloat16_rn(o2);
    if (which == 0) {
      bf16* q = a.y + (size_t)t * 3072 + h * 64;
      q[i] = b1; q[i + 32] = b2;
    } else {
      const int64_t sl = pf.valid ? pf.slot : a.slot[t];
      if (sl >= 0) {
        uint8_t* page = a.kv_layer + (size_t)(sl / CBX_BLOCK_SIZE) * CBX_PAGE_BYTES;
        const int tok = (int)(sl % CBX_BLOCK_SIZE);
        if (which == 1) {                         // K: 16-byte chunks XOR-swizzled by tok & 7
          uint8_t* kt = page + h * 2048 + tok * 128;
          const int c1 = (i >> 3) ^ (tok & 7), c2 = ((i + 32) >> 3) ^ (tok & 7);
          *reinterpret_cast<bf16*>(kt + c1 * 16 + (i & 7) * 2) = b1;
          *reinterpret_cast<bf16*>(kt + c2 * 16 + (i & 7) * 2) = b2;
        } else {
          bf16* vt = reinterpret_cast<bf16*>(page + 32768 + h * 2048 + tok * 128);
          vt[i] = b1; vt[i + 32] = b2;
        }
      }
    }
  }
}

// TT = padded row count (1,2,4,8); KSEG = K / 1024 (1 or 4); NW = warps per CTA.
template <int TT, int KSEG, int PRO, int EPI, int NW, bool W8>
__global__ void __launch_bounds__(NW * 32, 1) gemv_fused_kernel(const GemvArgs a) {
  constexpr int K = KSEG * 1024;
  constexpr int ES = W8 ? 1 : 2;                      // bytes per weight element
  constexpr int NJ = W8 ? 2 : 4;                      // 128-bit loads per lane per 1024-element row segment
  constexpr int IPC = NW / KSEG;                      // items per CTA per round
  static_assert(NW % KSEG == 0, "warps must cover whole K-segment groups");
  static_assert(PRO == PRO_COPY || KSEG == 1, "norm prologue is over hidden = 1024");
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint4* xs = reinterpret_cast<uint4*>(smem_raw);     // [TT][K/8] bf16x8
  float* part = reinterpret_cast<float*>(smem_raw + TT * K * 2);   // [NW][2*TT] (KSEG > 1)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kseg = warp % KSEG;
  const int n_items = a.N / 2;
  const int T = a.T;

  pdl_trigger();
  // ---- weight prefetch of this warp's first item: independent of the previous kernel
  int item = blockIdx.x * IPC + warp / KSEG;
  const int stride = gridDim.x * IPC;
  const uint8_t* Wb = reinterpret_cast<const uint8_t*>(a.W);
  uint4 cur[2][NJ];
  if (item < n_items) {
    int ra, rb;
    item_rows<EPI>(item, ra, rb);
    const uint4* wa = reinterpret_cast<const uint4*>(Wb + ((size_t)ra * K + kseg * 1024) * ES) + lane;
    const uint4* wb = reinterpret_cast<const uint4*>(Wb + ((size_t)rb * K + kseg * 1024) * ES) + lane;
#pragma unroll
    for (int j = 0; j < NJ; ++j) { cur[0][j] = ld_stream(wa + j * 32); cur[1][j] = ld_stream(wb + j * 32); }
  }
  // RMSNorm weights of the norm prologue: also independent of the previous kernel.  Left where they are used they were
  // fetched after the row's sum of squares — one more L2 round trip on the critical path of three kernels per layer.
  uint4 lnw[4] = {};
  const bool fold_acc = PRO == PRO_NORM && a.acc != nullptr;     // CTA-uniform
  if (PRO == PRO_NORM && warp < TT && !fold_acc) {
    const uint4* wr = reinterpret_cast<const uint4*>(a.ln_w);
#pragma unroll
    for (int i = 0; i < 4; ++i) lnw[i] = ld_stream(wr + lane + 32 * i);
  }
  // accumulator form of the prologue: ALL warps share the rows (thread -> 16-byte chunk ci = tid + k * NW * 32 of the
  // TT x 128 chunks), so its norm weights are fetched in that mapping
  constexpr int kFoldIters = (TT * 128 + NW * 32 - 1) / (NW * 32);
  uint4 lnc[kFoldIters] = {};
  if (fold_acc) {
    const uint4* wr = reinterpret_cast<const uint4*>(a.ln_w);
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {
      const int ci = threadIdx.x + k * NW * 32;
      if (ci < TT * 128) lnc[k] = ld_stream(wr + (ci & 127));
    }
  }
  l2_prefetch_slice(a.next_w, a.next_bytes);
  // W8: the first item's bytes are widened to fp32 BEFORE the wait (3.5 ops per weight that would otherwise sit on
  // the kernel-to-kernel critical path; measured +0.1 us per kernel at b = 1)
  float wf[2][W8 ? 32 : 1];
  if (W8 && item < n_items) {
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      unpack16_e4m3(cur[0][j], *reinterpret_cast<float(*)[16]>(&wf[0][j * 16]));
      unpack16_e4m3(cur[1][j], *reinterpret_cast<float(*)[16]>(&wf[1][j * 16]));
    }
  }
  pdl_wait();
  if (a.zero_ptr != nullptr) {
    uint4* z = reinterpret_cast<uint4*>(a.zero_ptr);
    for (int i = blockIdx.x * (NW * 32) + threadIdx.x; i < a.zero_bytes / 16; i += gridDim.x * (NW * 32)) z[i] = make_uint4(0, 0, 0, 0);
  }
  EpiPrefetch pf = {false, 0u, 1.f, 0.f, -1, 0};
  const int item_first = item;
  const bool do_pf = (EPI == EPI_RES || EPI == EPI_ROPE) && item < n_items && kseg == 0 && lane < T;
  if (do_pf) pf = gemv_epilogue_prefetch<EPI>(a, item, lane);

  // ---- prologue: activation rows -> shared memory (bf16)
  if (PRO == PRO_NORM && fold_acc) {
    // After the fused attention block: the rows' O-proj output waits in the 64-bit fixed-point accumulator.  Close the
    // residual branch, x = bf16(bf16(acc * 2^-32) + x) (CTA 0 writes the rows back), then normalise.  Spread over all
    // warps: done by warps < TT alone, the 1024 64-bit conversions per lane-row cost 0.9 us per layer at b = 1.
    __shared__ float s_ss[TT > 4 ? TT : 4][16];
    float ssl[TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) ssl[t] = 0.f;
    uint4 xv[kFoldIters], av[kFoldIters][4];
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {             // every load of the thread in flight together
      const int ci = threadIdx.x + k * NW * 32;
      if (ci < TT * 128 && (ci >> 7) < T) {
        xv[k] = reinterpret_cast<const uint4*>(a.x)[ci];
#pragma unroll
        for (int q = 0; q < 4; ++q) av[k][q] = reinterpret_cast<const uint4*>(a.acc)[ci * 4 + q];
      }
    }
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {
      const int ci = threadIdx.x + k * NW * 32;
      if (ci < TT * 128) {
        const int t = ci >> 7;                         // warp-uniform: 32 consecutive chunks lie in one row
        uint4 o = make_uint4(0, 0, 0, 0);
        if (t < T) {
          float f[8];
          unpack8(xv[k], f);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const long long lo = (long long)(((unsigned long long)av[k][q].y << 32) | av[k][q].x);
            const long long hi = (long long)(((unsigned long long)av[k][q].w << 32) | av[k][q].z);
            f[2 * q] = round_bf16(round_bf16((float)lo * (1.f / 4294967296.f)) + f[2 * q]);
            f[2 * q + 1] = round_bf16(round_bf16((float)hi * (1.f / 4294967296.f)) + f[2 * q + 1]);
          }
          o = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
          if (blockIdx.x == 0) reinterpret_cast<uint4*>(a.res_wb)[ci] = o;
          float ss = 0.f;
#pragma unroll
          for (int e = 0; e < 8; ++e) ss += f[e] * f[e];
#pragma unroll
          for (int tt = 0; tt < TT; ++tt) if (tt == t) ssl[tt] += ss;
        }
        xs[ci] = o;                                    // the new residual row, not yet normalised
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) {
      const float v = warp_sum(ssl[t]);
      if (lane == 0) s_ss[t][warp] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {
      const int ci = threadIdx.x + k * NW * 32;
      if (ci < TT * 128 && (ci >> 7) < T) {
        const int t = ci >> 7;
        float ss = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) ss += s_ss[t][w]; // warp order: every thread (and every CTA) gets the same bits
        const float rstd = rsqrtf(ss * (1.0f / 1024.f) + a.eps);
        float f[8], w[8];
        unpack8(xs[ci], f);
        unpack8(lnc[k], w);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = round_bf16(f[e] * rstd) * w[e];
        xs[ci] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      }
    }
  } else if (PRO == PRO_NORM) {
    if (warp < TT) {
      uint4 o[4] = {};
      if (warp < T) {
        const uint4* xr = reinterpret_cast<const uint4*>(a.x + (size_t)warp * 1024);
        float f[4][8];
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          unpack8(xr[lane + 32 * i], f[i]);
#pragma unroll
          for (int e = 0; e < 8; ++e) ss += f[i][e] * f[i][e];
        }
        ss = warp_sum(ss);
        const float rstd = rsqrtf(ss * (1.0f / 1024.f) + a.eps);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float w[8];
          unpack8(lnw[i], w);
#pragma unroll
          for (int e = 0; e < 8; ++e) f[i][e] = round_bf16(f[i][e] * rstd) * w[e];
          o[i] = make_uint4(pack_bf16(f[i][0], f[i][1]), pack_bf16(f[i][2], f[i][3]), pack_bf16(f[i][4], f[i][5]),
                            pack_bf16(f[i][6], f[i][7]));
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) xs[warp * 128 + lane + 32 * i] = o[i];
    }
  } else {
    for (int i = threadIdx.x; i < TT * (K / 8); i += NW * 32) {
      const int t = i / (K / 8);
      xs[i] = t < T ? reinterpret_cast<const uint4*>(a.x)[i] : make_uint4(0, 0, 0, 0);
    }
  }
  if (do_pf) gemv_epilogue_prefetch2<EPI>(a, item_first, pf);
  __syncthreads();

  // ---- main loop over this warp's items (CTA-uniform trip count when KSEG > 1)
  for (int base = blockIdx.x * IPC; base < n_items; base += stride) {
    item = base + warp / KSEG;
    const bool valid = item < n_items;
    uint4 nxt[2][NJ];
    const int nitem = item + stride;
    if (nitem < n_items) {
      int ra, rb;
      item_rows<EPI>(nitem, ra, rb);
      const uint4* wa = reinterpret_cast<const uint4*>(Wb + ((size_t)ra * K + kseg * 1024) * ES) + lane;
      const uint4* wb = reinterpret_cast<const uint4*>(Wb + ((size_t)rb * K + kseg * 1024) * ES) + lane;
#pragma unroll
      for (int j = 0; j < NJ; ++j) { nxt[0][j] = ld_stream(wa + j * 32); nxt[1][j] = ld_stream(wb + j * 32); }
    }
    float acc[2][TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) acc[0][t] = acc[1][t] = 0.f;
    if (valid && W8) {
      // lane's 16 bytes of load j are elements [j*512 + lane*16, +16) of the segment
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
#pragma unroll
        for (int t = 0; t < TT; ++t) {
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float fx[8];
            unpack8(xs[t * (K / 8) + kseg * 128 + j * 64 + lane * 2 + c], fx);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              acc[0][t] = fmaf(wf[0][W8 ? j * 16 + c * 8 + e : 0], fx[e], acc[0][t]);
              acc[1][t] = fmaf(wf[1][W8 ? j * 16 + c * 8 + e : 0], fx[e], acc[1][t]);
            }
          }
        }
      }
    } else if (valid) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float fa[8], fb[8];
        unpack8(cur[0][j], fa);
        unpack8(cur[1][j], fb);
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float fx[8];
          unpack8(xs[t * (K / 8) + kseg * 128 + j * 32 + lane], fx);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            acc[0][t] = fmaf(fa[e], fx[e], acc[0][t]);
            acc[1][t] = fmaf(fb[e], fx[e], acc[1][t]);
          }
        }
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) { acc[0][t] = warp_sum(acc[0][t]); acc[1][t] = warp_sum(acc[1][t]); }
    if (KSEG > 1) {                                     // combine the K segments of one item
      if (lane == 0) {
#pragma unroll
        for (int t = 0; t < TT; ++t) { part[warp * 2 * TT + t] = acc[0][t]; part[warp * 2 * TT + TT + t] = acc[1][t]; }
      }
      __syncthreads();
      if (kseg == 0) {
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int q = 0; q < KSEG; ++q) { s0 += part[(warp + q) * 2 * TT + t]; s1 += part[(warp + q) * 2 * TT + TT + t]; }
          acc[0][t] = s0; acc[1][t] = s1;
        }
      }
      __syncthreads();
    }
    // ---- epilogue: lane t finishes row t
    if (valid && kseg == 0) {
      int ra, rb;
      item_rows<EPI>(item, ra, rb);
      float v0 = 0.f, v1 = 0.f;
#pragma unroll
      for (int t = 0; t < TT; ++t) if (lane == t) { v0 = acc[0][t]; v1 = acc[1][t]; }
      if (W8) { v0 *= a.wscale[ra] * 256.f; v1 *= a.wscale[rb] * 256.f; }
      if (lane < T) {
        if (item == item_first) gemv_epilogue<EPI>(a, item, ra, rb, lane, v0, v1, pf);
        else gemv_epilogue<EPI>(a, item, ra, rb, lane, v0, v1);
      }
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) { cur[0][j] = nxt[0][j]; cur[1][j] = nxt[1][j]; }
    if (W8 && nitem < n_items) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        unpack16_e4m3(cur[0][j], *reinterpret_cast<float(*)[16]>(&wf[0][j * 16]));
        unpack16_e4m3(cur[1][j], *reinterpret_cast<float(*)[16]>(&wf[1][j * 16]));
      }
    }
  }
}

#ifdef CBX_EXPERIMENTS   // `make EXPERIMENTS=1`: measured-slower variants kept for their write-ups, not shipped
// ---- stream-ahead form (experiment, off by default: see stream_mode()) ------------------------------
// The register form above holds a CTA's weights in ~100 registers x 384-512 threads, so two consecutive
// kernels of the PDL chain never share an SM: kernel n+1 starts its weight loads only when kernel n's
// CTA on that SM has exited, and every kernel of the b = 1 step pays its own HBM latency (~2 us of a
// ~4 us kernel; 27 % of the step's HBM roofline).  Here a CTA's weight rows arrive by cp.async.bulk
// (UBLKCP) into ITS shared memory, issued at entry, before griddepcontrol.wait; the kernel needs few
// registers and 16-120 KB of shared memory, so the next two or three kernels of the chain are resident
// — and streaming — while the current one computes.  A CTA owns `ipc` consecutive items (row pairs);
// after the wait a warp takes every NW-th item, reading weights with conflict-free 128-bit LDS.
template <int TT, int KSEG, int PRO, int EPI, int NW, bool W8>
__global__ void __launch_bounds__(NW * 32, NW == 8 ? 2 : 4) gemv_stream_kernel(const GemvArgs a, const int ipc) {
  constexpr int K = KSEG * 1024;
  constexpr int ES = W8 ? 1 : 2;
  constexpr int ROWB = K * ES;                         // bytes per weight row
  constexpr int NQ = ROWB / 512;                       // 128-bit loads per lane per row
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint8_t* wsm = smem_raw;                             // [ipc][2][ROWB]
  uint4* xs = reinterpret_cast<uint4*>(smem_raw + (size_t)ipc * 2 * ROWB);   // [TT][K/8] bf16x8
  uint64_t* bar = reinterpret_cast<uint64_t*>(xs + TT * (K / 8));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = a.N / 2, T = a.T;
  const int i0 = blockIdx.x * ipc;
  const int nloc = n_items - i0 < ipc ? n_items - i0 : ipc;

  pdl_trigger();
  if (warp == 0) {
    if (lane == 0) {
      mbar_init(bar, 1);
      fence_barrier_init();
      mbar_arrive_expect_tx(bar, (uint32_t)(nloc * 2 * ROWB));
    }
    __syncwarp();
    const uint8_t* Wb = reinterpret_cast<const uint8_t*>(a.W);
    for (int idx = lane; idx < 2 * nloc; idx += 32) {
      int ra, rb;
      item_rows<EPI>(i0 + (idx >> 1), ra, rb);
      bulk_g2s(wsm + (size_t)idx * ROWB, Wb + (size_t)((idx & 1) ? rb : ra) * ROWB, ROWB, bar);
    }
  }
  pdl_wait();

  // ---- prologue: activation rows -> shared memory (bf16)
  if (PRO == PRO_NORM) {
    if (warp < TT) {
      uint4 o[4] = {};
      if (warp < T) {
        const uint4* xr = reinterpret_cast<const uint4*>(a.x + (size_t)warp * 1024);
        const uint4* wr = reinterpret_cast<const uint4*>(a.ln_w);
        float f[4][8];
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          unpack8(xr[lane + 32 * i], f[i]);
#pragma unroll
          for (int e = 0; e < 8; ++e) ss += f[i][e] * f[i][e];
        }
        ss = warp_sum(ss);
        const float rstd = rsqrtf(ss * (1.0f / 1024.f) + a.eps);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float w[8];
          unpack8(wr[lane + 32 * i], w);
#pragma unroll
          for (int e = 0; e < 8; ++e) f[i][e] = round_bf16(f[i][e] * rstd) * w[e];
          o[i] = make_uint4(pack_bf16(f[i][0], f[i][1]), pack_bf16(f[i][2], f[i][3]), pack_bf16(f[i][4], f[i][5]),
                            pack_bf16(f[i][6], f[i][7]));
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) xs[warp * 128 + lane + 32 * i] = o[i];
    }
  } else {
    for (int i = threadIdx.x; i < TT * (K / 8); i += NW * 32) {
      const int t = i / (K / 8);
      xs[i] = t < T ? reinterpret_cast<const uint4*>(a.x)[i] : make_uint4(0, 0, 0, 0);
    }
  }
  __syncthreads();                                     // xs complete; the barrier's init is visible to every warp
  mbar_wait(bar, 0);                                   // the CTA's weight rows have landed

  for (int it = warp; it < nloc; it += NW) {
    const uint4* wa = reinterpret_cast<const uint4*>(wsm + (size_t)(2 * it) * ROWB);
    const uint4* wb = reinterpret_cast<const uint4*>(wsm + (size_t)(2 * it + 1) * ROWB);
    float acc[2][TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) acc[0][t] = acc[1][t] = 0.f;
#pragma unroll 4
    for (int q = 0; q < NQ; ++q) {
      const uint4 va = wa[q * 32 + lane], vb = wb[q * 32 + lane];
      if (W8) {
        float fa[16], fb[16];
        unpack16_e4m3(va, fa);
        unpack16_e4m3(vb, fb);
#pragma unroll
        for (int t = 0; t < TT; ++t) {
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float fx[8];
            unpack8(xs[t * (K / 8) + (q * 32 + lane) * 2 + c], fx);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              acc[0][t] = fmaf(fa[c * 8 + e], fx[e], acc[0][t]);
              acc[1][t] = fmaf(fb[c * 8 + e], fx[e], acc[1][t]);
            }
          }
        }
      } else {
        float fa[8], fb[8];
        unpack8(va, fa);
        unpack8(vb, fb);
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float fx[8];
          unpack8(xs[t * (K / 8) + q * 32 + lane], fx);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            acc[0][t] = fmaf(fa[e], fx[e], acc[0][t]);
            acc[1][t] = fmaf(fb[e], fx[e], acc[1][t]);
          }
        }
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) { acc[0][t] = warp_sum(acc[0][t]); acc[1][t] = warp_sum(acc[1][t]); }
    const int item = i0 + it;
    int ra, rb;
    item_rows<EPI>(item, ra, rb);
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int t = 0; t < TT; ++t) if (lane == t) { v0 = acc[0][t]; v1 = acc[1][t]; }
    if (W8) { v0 *= a.wscale[ra] * 256.f; v1 *= a.wscale[rb] * 256.f; }
    if (lane < T) gemv_epilogue<EPI>(a, item, ra, rb, lane, v0, v1);
  }
}

// env CBX_GEMV_STREAM=1 selects this form.  OFF by default: parity-green (the whole GPU suite passes on it), but
// MEASURED SLOWER on B200 — b = 1 p50 step 0.696 ms against 0.614 ms for the register form.  The weight
// latency it hides was not what bounds the step: the chain is bound by its ~2 us of dependent work per
// kernel (wait release, x rows from L2, norm, FMAs, shuffles, read-modify-write epilogue, exit flush),
// to which this form adds the mbarrier wait and a second trip of every weight byte through LDS.
inline bool stream_mode() {
  static const bool on = getenv("CBX_GEMV_STREAM") && getenv("CBX_GEMV_STREAM")[0] == '1';
  return on;
}

template <int TT, int KSEG, int PRO, int EPI, int NW, bool W8>
int launch_stream(const GemvArgs& a, cudaStream_t s) {
  constexpr int ROWB = KSEG * 1024 * (W8 ? 1 : 2);
  const int n_items = a.N / 2;
  // one CTA per SM when the items allow it; at least NW/2 items per CTA so the warps have work
  int ipc = (n_items + sm_count() - 1) / sm_count();
  if (ipc < NW / 2) ipc = NW / 2;
  const int grid = (n_items + ipc - 1) / ipc;
  const int smem = ipc * 2 * ROWB + TT * KSEG * 1024 * 2 + 16;
  static int attr_smem = 0;                            // per instantiation
  if (smem > attr_smem) {
    cudaError_t e = cudaFuncSetAttribute(gemv_stream_kernel<TT, KSEG, PRO, EPI, NW, W8>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("gemv_stream smem attribute (%d B): %s", smem, cudaGetErrorString(e)); return -2; }
    attr_smem = smem;
  }
  cudaError_t e = launch_pdl(gemv_stream_kernel<TT, KSEG, PRO, EPI, NW, W8>, dim3(grid), dim3(NW * 32), smem, s, a, ipc);
  if (e != cudaSuccess) { set_error("gemv_stream launch failed: %s", cudaGetErrorString(e)); return -2; }
  count_launch();
  return 0;
}

#endif  // CBX_EXPERIMENTS

template <int TT, int KSEG, int PRO, int EPI, int NW, bool W8>
int launch_w(const GemvArgs& a, int grid, cudaStream_t s) {
  constexpr int smem = TT * KSEG * 1024 * 2 + (KSEG > 1 ? NW * 2 * TT * 4 : 0);
  if (smem > 48 * 1024) {                              // 8 rows x K = 4096: 65 KB, above the default dynamic limit
    static bool attr = false;                          // per instantiation
    if (!attr) {
      cudaError_t e = cudaFuncSetAttribute(gemv_fused_kernel<TT, KSEG, PRO, EPI, NW, W8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) { set_error("gemv smem attribute (%d B): %s", smem, cudaGetErrorString(e)); return -2; }
      attr = true;
    }
  }
  cudaError_t e = launch_pdl(gemv_fused_kernel<TT, KSEG, PRO, EPI, NW, W8>, dim3(grid), dim3(NW * 32), smem, s, a);
  if (e != cudaSuccess) { set_error("gemv launch failed: %s", cudaGetErrorString(e)); return -2; }
  count_launch();
  return 0;
}

template <int TT, int KSEG, int PRO, int EPI, int NW>
int launch(const GemvArgs& a, int grid, cudaStream_t s) {
#ifdef CBX_EXPERIMENTS
  if (stream_mode()) {
    // stream-ahead form: 8 warps (4 for the two N = 1024 projections, whose CTAs own 4 items)
    constexpr int SW = (EPI == EPI_RES) ? 4 : 8;
    return a.wscale ? launch_stream<TT, KSEG, PRO, EPI, SW, true>(a, s) : launch_stream<TT, KSEG, PRO, EPI, SW, false>(a, s);
  }
#endif
  return a.wscale ? launch_w<TT, KSEG, PRO, EPI, NW, true>(a, grid, s) : launch_w<TT, KSEG, PRO, EPI, NW, false>(a, grid, s);
}

template <int KSEG, int PRO, int EPI, int NW>
int dispatch_t(const GemvArgs& a, int grid, cudaStream_t s) {
  if (a.T == 1) return launch<1, KSEG, PRO, EPI, NW>(a, grid, s);
  if (a.T == 2) return launch<2, KSEG, PRO, EPI, NW>(a, grid, s);
  if (a.T <= 4) return launch<4, KSEG, PRO, EPI, NW>(a, grid, s);
  return launch<8, KSEG, PRO, EPI, NW>(a, grid, s);      // 5..8 rows: 3-4 CFG pairs (engine only, CBX_GEMV_MAX_T=8)
}

// one warp per item when that needs <= one CTA per SM, else 148 CTAs looping
inline int grid_for(int n_items, int items_per_cta) {
  const int g = (n_items + items_per_cta - 1) / items_per_cta;
  return g < sm_count() ? g : sm_count();
}
}  // namespace

namespace cbx {
// Plain y = x·W^T (epilogue 0) or SiLU(gate)·up (epilogue 1).  T <= 4, K in {1024, 4096}.
int gemv_bf16(const void* x, const void* W, void* y, int T, int N, int K, int ldy, int epilogue, cudaStream_t s,
              const float* wscale) {
  CBX_REQUIRE(T >= 1 && T <= 4, "gemv: T=%d not in [1,4]", T);
  CBX_REQUIRE(K == 1024 || K == 4096, "gemv: K=%d unsupported", K);
  CBX_REQUIRE(N % 2 == 0 && (epilogue == 0 || N % 32 == 0), "gemv: N=%d unsupported", N);
  GemvArgs a = {};
  a.x = (const bf16*)x; a.W = W; a.wscale = wscale; a.y = (bf16*)y; a.T = T; a.N = N; a.ldy = ldy;
  if (K == 1024) {
    if (epilogue == 1) return dispatch_t<1, PRO_COPY, EPI_SILU, 14>(a, grid_for(N / 2, 14), s);
    if (N <= 2048) return dispatch_t<1, PRO_COPY, EPI_STORE, 4>(a, grid_for(N / 2, 4), s);
    return dispatch_t<1, PRO_COPY, EPI_STORE, 14>(a, grid_for(N / 2, 14), s);
  }
  CBX_REQUIRE(epilogue == 0, "gemv: SiLU epilogue only for K=1024");
  return dispatch_t<4, PRO_COPY, EPI_STORE, 16>(a, grid_for(N / 2, 4), s);
}

// RMSNorm(res)·ln_w -> ·Wqkv^T -> RoPE -> q into qkv[:, :1024], k/v into the page of slot[t].
int gemv_qkv_rope(const void* res, const void* ln_w, const void* Wqkv, const void* cos_sin, const int32_t* pos,
                  const int64_t* slot, void* qkv, void* kv_layer, int T, float eps, cudaStream_t s,
                  const void* next_w, size_t next_bytes, const float* wscale) {
  GemvArgs a = {};
  a.next_w = next_w; a.next_bytes = next_bytes; a.wscale = wscale;
  a.x = (const bf16*)res; a.ln_w = (const bf16*)ln_w; a.W = Wqkv; a.y = (bf16*)qkv;
  a.cos_sin = (const bf16*)cos_sin; a.pos = pos; a.slot = slot; a.kv_layer = (uint8_t*)kv_layer;
  a.T = T; a.N = 3072; a.ldy = 3072; a.eps = eps;
  return dispatch_t<1, PRO_NORM, EPI_ROPE, 12>(a, grid_for(1536, 12), s);
}
// res = bf16(bf16(x·W^T) + res), N = 1024, K in {1024, 4096}.
int gemv_residual(const void* x, const void* W, void* res, int T, int K, cudaStream_t s, const void* next_w,
                  size_t next_bytes, const float* wscale, void* zero_ptr, int zero_bytes, const void* res_in) {
  GemvArgs a = {};
  a.next_w = next_w; a.next_bytes = next_bytes; a.wscale = wscale; a.zero_ptr = zero_ptr; a.zero_bytes = zero_bytes;
  a.res_in = (const bf16*)res_in;
  a.x = (const bf16*)x; a.W = W; a.y = (bf16*)res; a.T = T; a.N = 1024; a.ldy = 1024;
  if (K == 1024) return dispatch_t<1, PRO_COPY, EPI_RES, 4>(a, grid_for(512, 4), s);
  return dispatch_t<4, PRO_COPY, EPI_RES, 16>(a, grid_for(512, 4), s);
}
// act = SiLU(g)·u of RMSNorm(res)·ln_w · Wgu^T (packed gate/up), act [T, 4096].
int gemv_norm_gate_up(const void* res, const void* ln_w, const void* Wgu, void* act, int T, float eps, cudaStream_t s,
                      const void* next_w, size_t next_bytes, const float* wscale, const void* acc, void* res_out) {
  GemvArgs a = {};
  a.next_w = next_w; a.next_bytes = next_bytes; a.wscale = wscale;
  a.acc = (const unsigned long long*)acc; a.res_wb = (bf16*)res_out;   // acc != NULL: CTA 0 writes the closed residual rows there
  a.x = (const bf16*)res; a.ln_w = (const bf16*)ln_w; a.W = Wgu; a.y = (bf16*)act;
  a.T = T; a.N = 8192; a.ldy = 4096; a.eps = eps;
  return dispatch_t<1, PRO_NORM, EPI_SILU, 14>(a, grid_for(4096, 14), s);
}
// logits = RMSNorm(res)·norm_w · Whead^T, N even.
int gemv_norm_head(const void* res, const void* norm_w, const void* Whead, void* logits, int T, int N, int ldy,
                   float eps, cudaStream_t s, const float* wscale) {
  GemvArgs a = {};
  a.wscale = wscale;
  a.x = (const bf16*)res; a.ln_w = (const bf16*)norm_w; a.W = Whead; a.y = (bf16*)logits;
  a.T = T; a.N = N; a.ldy = ldy; a.eps = eps;
  return dispatch_t<1, PRO_NORM, EPI_STORE, 14>(a, grid_for(N / 2, 14), s);
}
}  // namespace cbx
