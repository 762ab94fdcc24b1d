// k_block.cu — the attention half of a Llama layer as ONE kernel for T <= 4 rows (the 1-2 CFG-pair
// latency regime): RMSNorm(res)·ln1 -> QKV -> RoPE -> KV append -> paged attention -> O-proj -> res +=.
//
// Replaces three kernels of the GEMV chain (k_gemv.cu QKV, k_attn.cu small attention, k_gemv.cu O+res) and
// two of the layer's five all-to-all exchanges (vllm llama.py:212-226,316-326 + flash_attn.py:796-800,850-874).
// The exchange QKV -> attention is head-local, so it does not need a kernel boundary:
//
//   grid (6, 16): 16 thread-block clusters, one per HEAD, of 6 CTAs (all 16 must be co-resident: a B200 fits 22
//            clusters of 6 but only 15 of 8); 16 warps per CTA.
//   QKV      CTA r streams 32 of the head's 192 Wqkv rows (16 RoPE pairs: ranks 0-1 q, 2-3 k, 4-5 v; one pair
//            = 2 rows x 2 KB per warp) and its ~171-row slice of Wo[:, 64h:64h+64] (~21 KB) — all requested
//            BEFORE griddepcontrol.wait, like the GEMV kernels.  Each CTA normalises the T rows itself.
//   exchange rotated q/k and v (bf16-rounded) are written into ALL 6 CTAs' shared memory (DSMEM) and the
//            new token's K/V into its page; one cluster barrier.
//   attention the 6 CTAs split each row's pages (page p -> CTA p % 6; 2 x 2 KB bulk copies per page-slice,
//            requested right after the wait: old pages do not depend on this kernel's QKV); the new token
//            is taken from shared memory, not from the page.  Every CTA pushes its partial into all six
//            (DSMEM stores), one more cluster barrier, then each merges locally.
//   O-proj   CTA r multiplies the head's 64-wide output by its Wo slice and ADDS the fp32 partial [T, ~171], as
//            64-bit fixed point, into one accumulator [T,1024] shared by the 16 heads (RED.ADD.64: integer adds
//            commute, so the bits do not depend on arrival order).  No ticket, no fence, no fold in this kernel:
//            the NEXT kernel's norm prologue (k_gemv.cu, gate_up) turns the accumulator into the residual,
//            res = bf16(bf16(y) + res), and the down-proj kernel clears it (a first version folded here behind a
//            ticket: fence + atomic + fold cost 2-3 us at the tail of every launch).
//
// Rounding points are the chain's: h, qkv, rotated q/k, attention output, o-proj output and the residual are
// bf16; accumulation and softmax fp32.
#include <stdlib.h>

#include "common.cuh"

using namespace cbx;

namespace {
constexpr int kNW = 16;                    // warps per CTA
constexpr int kQkvWarps = 12;              // warps 0..11 own one Wqkv row pair each
constexpr int kCl = 6;                     // CTAs per cluster = page splits per row: 22 clusters of 6 are co-resident on a
                                           // B200 (148 SMs), but only 15 of 8 (cudaOccupancyMaxActiveClusters) — 16 are needed
constexpr int kSlices = 40;                // page-slices (one head's K+V of a page, 4 KB) a CTA can hold
constexpr int kHeads = CBX_HEADS;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kAccScale = 4294967296.f;  // 2^32: partials are far below 2^31 in magnitude, resolution 2.3e-10

// -DCBX_BLOCK_PROF: %globaltimer stamps by thread 0 of every CTA (tools/block_timeline.py reads them through
// cbx_block_prof_read): 0 entry, 1 pre-wait loads issued, 2 past the wait, 3 rows normalised, 4 QKV done, 5 exchange
// barrier passed, 6 attention loop done, 7 partial barrier passed, 8 merged, 9 O-proj + ticket, 10 end.
#ifdef CBX_BLOCK_PROF
constexpr int kProfLaunches = 128;
__device__ unsigned long long g_bprof[kProfLaunches][128][12];
int g_bprof_next = 0;
#define BPROF(k)                                                                                          \
  do {                                                                                                    \
    if (threadIdx.x == 0 && a.prof_id >= 0 && a.prof_id < kProfLaunches) {                                \
      unsigned long long t_;                                                                              \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                              \
      g_bprof[a.prof_id][blockIdx.y * kCl + blockIdx.x][k] = t_;                                            \
    }                                                                                                     \
  } while (0)
#else
#define BPROF(k)
#endif

struct BlockArgs {
  const bf16* res;           // [T,1024] residual stream (read only: the O-proj output leaves through `acc`)
  const bf16* ln_w;          // [1024]
  const bf16* Wqkv;          // [3072,1024]
  const bf16* Wo;            // [1024,1024]
  const bf16* cos_sin;       // [max_pos,64]
  const int32_t* pos;        // [T]
  const int64_t* slot;       // [T] (-1: do not cache)
  uint8_t* kv_layer;
  const int32_t* bt; int bt_stride;
  const int32_t* row_seq;    // [T]
  const int32_t* kv_len;     // [T] context length INCLUDING the new token (0 = inactive row)
  unsigned long long* acc;   // [4][1024] 64-bit fixed-point (2^-32) accumulator of the 16 heads' O-proj partials:
                             // integer adds commute, so the sum does not depend on arrival order (bit-reproducible)
  bf16* q_out;               // optional [T,3072]: rotated q (tests); may be NULL
  // FOLD form (after the fused MLP block, k_mlp.cu): the previous layer's down-proj output waits in acc_in; the norm
  // prologue closes that residual branch, x = bf16(bf16(acc_in * 2^-32) + res), CTA (0,0) writing the rows to res_wb
  // (NOT the buffer res points at: the other CTAs are still reading that one).  acc_in == NULL: the rows are copied.
  const unsigned long long* acc_in;   // [kMlpAccCopies][4][1024]: the MLP block's CTAs add into copy b % kMlpAccCopies
  bf16* res_wb;
  void* zero_ptr; int zero_bytes;   // cleared after the wait (the accumulator the MLP block of THIS layer will add into)
  int T; float eps;
  int prof_id;               // CBX_BLOCK_PROF builds only
  int stop;                  // profiling aid (env CBX_BLOCK_STOP=k): every CTA returns after phase k; results are garbage
};

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t dsmem_addr(const void* local, uint32_t rank) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
  return ra;
}
__device__ __forceinline__ float ld_dsmem(const float* local, uint32_t rank) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(dsmem_addr(local, rank)) : "memory");
  return v;
}
__device__ __forceinline__ void st_dsmem(float* local, uint32_t rank, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dsmem_addr(local, rank)), "f"(v) : "memory");
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}

template <int TT>
struct Smem {
  static constexpr int kLocal = kSlices / TT;          // page-slices per row
  static constexpr int pages = 0;
  static constexpr int xs = pages + kSlices * 4096;    // [TT][1024] bf16
  static constexpr int bars = xs + TT * 2048;          // [kSlices] mbarriers
  static constexpr int qkv = bars + kSlices * 8;       // float [TT][192]: q | k | v of the new token (bf16 values)
  static constexpr int red = qkv + TT * 192 * 4;       // float [kNW][68]: warp partials o[64], m, l
  static constexpr int parts = red + kNW * 68 * 4;     // float [kCl][TT][68]: every CTA's partial per row (pushed by the peers)
  static constexpr int attn = parts + kCl * TT * 68 * 4;   // float [TT][64]: final attention output (bf16 values)
  static constexpr int meta = attn + TT * 64 * 4;      // int [TT][2]: kv_len, local page-slices
  static constexpr int total = meta + TT * 8 + 16;
};

// Wo feature groups (4 features each, 256 in all) owned by cluster rank r: 43, 43, 43, 43, 42, 42
__host__ __device__ constexpr int wo_group0(int r) { return r * 43 - (r > 4 ? r - 4 : 0); }

template <int TT, bool FOLD>
__global__ void __launch_bounds__(kNW * 32, 1) layer_attn_block_kernel(const BlockArgs a) {
  using S = Smem<TT>;
  constexpr int kLocal = S::kLocal;
  constexpr int WPR = kNW / TT;                        // attention warps per row
  extern __shared__ __align__(1024) uint8_t smem[];
  uint4* xs = reinterpret_cast<uint4*>(smem + S::xs);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::bars);
  float* qkv_s = reinterpret_cast<float*>(smem + S::qkv);
  float* red = reinterpret_cast<float*>(smem + S::red);
  float* parts = reinterpret_cast<float*>(smem + S::parts);
  float* attn_s = reinterpret_cast<float*>(smem + S::attn);
  int* meta = reinterpret_cast<int*>(smem + S::meta);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = (int)cluster_ctarank();                // == blockIdx.x
  const int h = blockIdx.y;
  const int T = a.T;

  BPROF(0);
  pdl_trigger();
  cluster_arrive();                                    // #0: "this CTA runs" — waited for before the first DSMEM store

  // ---- before the wait: everything that does not depend on the previous kernel -----------------------------
  const int pair = r * kNW + warp;                     // 0..95: q pairs 0..31, k pairs 32..63, v pairs 64..95
  const int which = pair >> 5;                         // 0 q, 1 k, 2 v
  const int pi = pair & 31;                            // head dims (pi, pi + 32)
  uint4 wq[2][4];
  {
    const size_t ra = (size_t)which * 1024 + h * 64 + pi;
    const uint4* wa = reinterpret_cast<const uint4*>(a.Wqkv + ra * 1024) + lane;
    const uint4* wb = reinterpret_cast<const uint4*>(a.Wqkv + (ra + 32) * 1024) + lane;
#pragma unroll
    for (int j = 0; j < 4; ++j) { wq[0][j] = ld_stream(wa + j * 32); wq[1][j] = ld_stream(wb + j * 32); }
  }
  // Wo: feature group g = 4 output features; lane -> (feature g*4 + lane/8, columns 64h + 8*(lane%8) ..)
  const int g_lo = wo_group0(r), g_hi = wo_group0(r + 1);
  uint4 wo[3] = {};
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int g = g_lo + warp + kNW * j;
    if (g < g_hi) wo[j] = ld_stream(reinterpret_cast<const uint4*>(a.Wo + (size_t)(g * 4 + (lane >> 3)) * 1024 + h * 64) + (lane & 7));
  }
  uint4 lnw[4] = {};
  if (!FOLD && warp < TT) {
    const uint4* wr = reinterpret_cast<const uint4*>(a.ln_w);
#pragma unroll
    for (int i = 0; i < 4; ++i) lnw[i] = ld_stream(wr + lane + 32 * i);
  }
  // FOLD: all warps share the rows (thread -> 16-byte chunk ci = tid of the TT x 128 chunks), norm weights in that mapping
  const int ci = threadIdx.x;
  uint4 lnc = make_uint4(0, 0, 0, 0);
  if (FOLD && ci < TT * 128) lnc = ld_stream(reinterpret_cast<const uint4*>(a.ln_w) + (ci & 127));
  // page ids of this CTA's page-slices, one lane of warps 12..15 per slice: the block table is written at admission
  // and row_seq is re-staged with the same value every step — neither depends on the awaited kernels (kv_len does)
  const int sl = (warp - 12) * 32 + lane;              // slice owned by this lane (warps 12..15), else < 0 or >= kSlices
  const bool ctl = warp >= 12 && sl < kSlices;
  const int crow = ctl ? sl / kLocal : 0, cj = ctl ? sl % kLocal : 0;
  int pre_blk = 0;
  if (ctl) {                                           // a slice's barrier is initialised by the lane that issues its copies;
    mbar_init(&bars[sl], 1);                           // the consuming warps meet it only after the block barrier below
    fence_barrier_init();
    if (crow < T && r + cj * kCl < a.bt_stride) pre_blk = a.bt[(int64_t)a.row_seq[crow] * a.bt_stride + r + cj * kCl];
  }
  BPROF(1);
  pdl_wait();
  BPROF(2);
  if (a.stop == 1) return;

  // ---- after the wait ---------------------------------------------------------------------------------------
  if (ctl) {
    int len = 0, n_loc = 0;
    if (crow < T) {
      len = a.kv_len[crow];
      const int n_old = len > 0 ? len - 1 : 0;         // the new token comes from shared memory
      const int n_pages = (n_old + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
      n_loc = n_pages > r ? (n_pages - r + kCl - 1) / kCl : 0;
      if (n_loc > kLocal) n_loc = kLocal;              // host guarantees it fits; never overrun the buffer
      if (cj < n_loc) {
        uint64_t* bar = &bars[sl];
        uint8_t* dst = smem + S::pages + sl * 4096;
        const uint8_t* src = a.kv_layer + (size_t)pre_blk * CBX_PAGE_BYTES + h * 2048;
        mbar_arrive_expect_tx(bar, 4096);
        bulk_g2s(dst, src, 2048, bar);
        bulk_g2s(dst + 2048, src + 32768, 2048, bar);
      }
    }
    if (cj == 0) { meta[crow * 2] = len; meta[crow * 2 + 1] = n_loc; }
  }
  // epilogue operands that do not depend on the dot products (lane t finishes row t)
  int my_pos = 0;
  int64_t my_slot = -1;
  if (lane < T) {
    if (which < 2) my_pos = a.pos[lane];
    if (which > 0) my_slot = a.slot[lane];
  }
  if (FOLD) {
    if (a.zero_ptr != nullptr) {
      uint4* z = reinterpret_cast<uint4*>(a.zero_ptr);
      const int cta = blockIdx.y * kCl + blockIdx.x;
      for (int i = cta * (kNW * 32) + threadIdx.x; i < a.zero_bytes / 16; i += kCl * kHeads * (kNW * 32)) z[i] = make_uint4(0, 0, 0, 0);
    }
    const bool mine = ci < TT * 128 && (ci >> 7) < T;
    float f[8];
    float ss = 0.f;
    uint4 o = make_uint4(0, 0, 0, 0);
    if (mine) {
      const uint4 xv = reinterpret_cast<const uint4*>(a.res)[ci];
      unpack8(xv, f);
      if (a.acc_in != nullptr) {
        uint4 av[kMlpAccCopies][4];                    // [copy][16-byte piece of the chunk's 8 u64]: all loads in flight
#pragma unroll
        for (int c = 0; c < kMlpAccCopies; ++c)
#pragma unroll
          for (int q = 0; q < 4; ++q) av[c][q] = reinterpret_cast<const uint4*>(a.acc_in + (size_t)c * 4096)[ci * 4 + q];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          unsigned long long slo = 0ull, shi = 0ull;
#pragma unroll
          for (int c = 0; c < kMlpAccCopies; ++c) {
            slo += ((unsigned long long)av[c][q].y << 32) | av[c][q].x;
            shi += ((unsigned long long)av[c][q].w << 32) | av[c][q].z;
          }
          const long long lo = (long long)slo, hi = (long long)shi;
          f[2 * q] = round_bf16(round_bf16((float)lo * (1.f / kAccScale)) + f[2 * q]);
          f[2 * q + 1] = round_bf16(round_bf16((float)hi * (1.f / kAccScale)) + f[2 * q + 1]);
        }
      }
      o = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      if (blockIdx.x == 0 && blockIdx.y == 0 && a.res_wb != nullptr) reinterpret_cast<uint4*>(a.res_wb)[ci] = o;
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += f[e] * f[e];
    }
    ss = warp_sum(ss);                                 // 32 consecutive chunks lie in one row: warps 4t .. 4t+3 hold row t
    if (lane == 0 && ci < TT * 128) red[warp] = ss;    // `red` is free until the attention phase
    __syncthreads();
    if (ci < TT * 128) {
      if (mine) {
        const int t = ci >> 7;
        const float tot = red[4 * t] + red[4 * t + 1] + red[4 * t + 2] + red[4 * t + 3];   // fixed order: same bits in every CTA
        const float rstd = rsqrtf(tot * (1.0f / 1024.f) + a.eps);
        float w[8];
        unpack8(lnc, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = round_bf16(f[e] * rstd) * w[e];
        o = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      }
      xs[ci] = o;
    }
  } else if (warp < TT) {                              // RMSNorm of row `warp` -> xs
    uint4 o[4] = {};
    if (warp < T) {
      const uint4* xr = reinterpret_cast<const uint4*>(a.res + (size_t)warp * 1024);
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
  float rc = 1.f, rs = 0.f;
  if (which < 2 && lane < T) {
    rc = __bfloat162float(a.cos_sin[(size_t)my_pos * 64 + pi]);
    rs = __bfloat162float(a.cos_sin[(size_t)my_pos * 64 + 32 + pi]);
  }
  __syncthreads();
  BPROF(3);
  if (a.stop == 2) return;

  // ---- QKV: one row pair per warp (the GEMV kernel's summation order: the appended K/V carry the same bits) -----
  float f1, f2;                                        // lane t: the pair's values for row t, bf16-rounded
  bf16 kv_b1, kv_b2;                                   // ... as bf16: stored to q_out / the KV page at the END of the kernel
  {
    float acc[2][TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) acc[0][t] = acc[1][t] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float fa[8], fb[8];
      unpack8(wq[0][j], fa);
      unpack8(wq[1][j], fb);
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        float fx[8];
        unpack8(xs[t * 128 + j * 32 + lane], fx);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          acc[0][t] = fmaf(fa[e], fx[e], acc[0][t]);
          acc[1][t] = fmaf(fb[e], fx[e], acc[1][t]);
        }
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) { acc[0][t] = warp_sum(acc[0][t]); acc[1][t] = warp_sum(acc[1][t]); }
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int t = 0; t < TT; ++t) if (lane == t) { v0 = acc[0][t]; v1 = acc[1][t]; }
    float o1 = round_bf16(v0), o2 = round_bf16(v1);    // GEMM output rounding
    if (which < 2) {                                   // neox rotation of the pair (pi, pi + 32)
      const float x1 = o1, x2 = o2;
      o1 = x1 * rc - x2 * rs;
      o2 = x2 * rc + x1 * rs;
    }
    kv_b1 = __float2bfloat16_rn(o1); kv_b2 = __float2bfloat16_rn(o2);
    f1 = __bfloat162float(kv_b1); f2 = __bfloat162float(kv_b2);
  }
  BPROF(4);
  if (a.stop == 3) return;
  cluster_wait();                                      // #0: every CTA of the cluster has started
  {                                                    // lane = dst * TT + t writes row t's pair into CTA dst
    const int t = lane % TT, dst = lane / TT;
    const float s1 = __shfl_sync(0xffffffffu, f1, t), s2 = __shfl_sync(0xffffffffu, f2, t);
    if (dst < kCl && t < T) {
      float* p = qkv_s + t * 192 + which * 64 + pi;
      st_dsmem(p, dst, s1);
      st_dsmem(p + 32, dst, s2);
    }
  }
  cluster_arrive();                                    // #1: q / k / v of the new token are in every CTA
  cluster_wait();
  BPROF(5);
  if (a.stop == 4) return;

  // ---- attention over this CTA's page-slices: warps [t*WPR, (t+1)*WPR) serve row t --------------------------
  {
    const int t = warp / WPR, wl = warp % WPR;
    const int len = meta[t * 2], n_loc = meta[t * 2 + 1];
    const int n_old = len > 0 ? len - 1 : 0;
    const int t16 = lane & 15, half = lane >> 4;       // QK mapping: token x half of d
    const int tg = lane >> 3, c = lane & 7;            // PV mapping: token group x 16-byte chunk
    const float qscale = 0.125f * kLog2e;
    float q[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) q[i] = qkv_s[t * 192 + half * 32 + i] * qscale;
    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (r == 0 && wl == 0 && len > 0) {                // the new token, from shared memory: one "page" of one token
      float sc = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) sc = fmaf(q[i], qkv_s[t * 192 + 64 + half * 32 + i], sc);
      sc += __shfl_xor_sync(0xffffffffu, sc, 16);
      m = sc;
      l = (t16 == 0) ? 1.f : 0.f;                      // lanes 0 and 16 carry the token (l is halved below)
      if (tg == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = qkv_s[t * 192 + 128 + c * 8 + i];
      }
    }
    for (int j = wl; j < n_loc; j += WPR) {
      mbar_wait(&bars[t * kLocal + j], 0);
      const uint8_t* kt = smem + S::pages + (t * kLocal + j) * 4096;
      const uint8_t* vt = kt + 2048;
      float sc = 0.f;
      {
        const uint4* kr = reinterpret_cast<const uint4*>(kt + t16 * 128);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint4 v = kr[(half * 4 + i) ^ (t16 & 7)];
          sc = fmaf(q[i * 8 + 0], bf_lo(v.x), sc); sc = fmaf(q[i * 8 + 1], bf_hi(v.x), sc);
          sc = fmaf(q[i * 8 + 2], bf_lo(v.y), sc); sc = fmaf(q[i * 8 + 3], bf_hi(v.y), sc);
          sc = fmaf(q[i * 8 + 4], bf_lo(v.z), sc); sc = fmaf(q[i * 8 + 5], bf_hi(v.z), sc);
          sc = fmaf(q[i * 8 + 6], bf_lo(v.w), sc); sc = fmaf(q[i * 8 + 7], bf_hi(v.w), sc);
        }
      }
      sc += __shfl_xor_sync(0xffffffffu, sc, 16);
      const int n_valid = n_old - (r + j * kCl) * CBX_BLOCK_SIZE;
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
          const uint4 v = reinterpret_cast<const uint4*>(vt + tok * 128)[c];
          acc[0] = fmaf(pt, bf_lo(v.x), acc[0]); acc[1] = fmaf(pt, bf_hi(v.x), acc[1]);
          acc[2] = fmaf(pt, bf_lo(v.y), acc[2]); acc[3] = fmaf(pt, bf_hi(v.y), acc[3]);
          acc[4] = fmaf(pt, bf_lo(v.z), acc[4]); acc[5] = fmaf(pt, bf_hi(v.z), acc[5]);
          acc[6] = fmaf(pt, bf_lo(v.w), acc[6]); acc[7] = fmaf(pt, bf_hi(v.w), acc[7]);
        }
      }
    }
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
  }
  BPROF(6);
  __syncthreads();
  if (a.stop == 5) return;
  // this CTA's partial per row (its warps folded in warp order), PUSHED into every CTA of the cluster
  if (threadIdx.x < TT * 64) {
    const int t = threadIdx.x >> 6, d = threadIdx.x & 63;
    float mx = -INFINITY;
#pragma unroll
    for (int w = 0; w < WPR; ++w) mx = fmaxf(mx, red[(t * WPR + w) * 68 + 64]);
    float o = 0.f, lt = 0.f;
    if (mx > -INFINITY) {
#pragma unroll
      for (int w = 0; w < WPR; ++w) {
        const float* pw = red + (t * WPR + w) * 68;
        const float pm = pw[64];
        const float wgt = pm == -INFINITY ? 0.f : exp2f(pm - mx);
        o = fmaf(wgt, pw[d], o);
        lt = fmaf(wgt, pw[65], lt);
      }
    }
    float* dst = parts + (r * TT + t) * 68;
#pragma unroll
    for (int k = 0; k < kCl; ++k) {
      st_dsmem(dst + d, k, o);
      if (d == 0) { st_dsmem(dst + 64, k, mx); st_dsmem(dst + 65, k, lt); }
    }
  }
  cluster_arrive();                                    // #2: all six CTA partials of every row are in every CTA
  cluster_wait();                                      // (the last cluster barrier: nobody touches a peer's memory after it)
  BPROF(7);
  if (a.stop == 6) return;
  // final merge, redundantly in every CTA (each needs the head's whole output for its Wo slice), rank order
  if (threadIdx.x < TT * 64) {
    const int t = threadIdx.x >> 6, d = threadIdx.x & 63;
    float outv = 0.f;
    if (meta[t * 2] > 0) {
      float pm[kCl], mx = -INFINITY;
#pragma unroll
      for (int k = 0; k < kCl; ++k) { pm[k] = parts[(k * TT + t) * 68 + 64]; mx = fmaxf(mx, pm[k]); }
      float o = 0.f, lt = 0.f;
#pragma unroll
      for (int k = 0; k < kCl; ++k) {
        const float wgt = pm[k] == -INFINITY ? 0.f : exp2f(pm[k] - mx);
        o = fmaf(wgt, parts[(k * TT + t) * 68 + d], o);
        lt = fmaf(wgt, parts[(k * TT + t) * 68 + 65], lt);
      }
      outv = round_bf16(o / lt);                       // attention output is a bf16 tensor in the chain
    }
    attn_s[t * 64 + d] = outv;
  }
  __syncthreads();
  BPROF(8);
  if (a.stop == 7) return;

  // ---- O-proj of the head's 64 columns: up to 3 groups of 4 output features per warp ----------------------------
  {
    const int c = lane & 7;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int g = g_lo + warp + kNW * j;
      if (g < g_hi) {                                  // warp-uniform
        float w[8];
        unpack8(wo[j], w);
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float s = 0.f;
#pragma unroll
          for (int e = 0; e < 8; ++e) s = fmaf(w[e], attn_s[t * 64 + c * 8 + e], s);
          s += __shfl_xor_sync(0xffffffffu, s, 1);
          s += __shfl_xor_sync(0xffffffffu, s, 2);
          s += __shfl_xor_sync(0xffffffffu, s, 4);
          if (c == 0 && t < T)                         // RED.ADD.64: fire and forget, ordered by kernel completion
            atomicAdd(a.acc + (size_t)t * 1024 + g * 4 + (lane >> 3), (unsigned long long)__float2ll_rn(s * kAccScale));
        }
      }
    }
  }
  // ---- the new token's K / V go to their page LAST (and q to q_out): this step's attention took them from shared memory;
  // stored before the cluster barriers, their .release (a GPU-scope fence) would wait for these stores to be acknowledged
  if (lane < T) {
    if (which == 0) {
      if (a.q_out) { bf16* q = a.q_out + (size_t)lane * 3072 + h * 64; q[pi] = kv_b1; q[pi + 32] = kv_b2; }
    } else if (my_slot >= 0) {                         // append the new token's K / V to its page
      uint8_t* page = a.kv_layer + (size_t)(my_slot / CBX_BLOCK_SIZE) * CBX_PAGE_BYTES;
      const int tok = (int)(my_slot % CBX_BLOCK_SIZE);
      if (which == 1) {                                // K: 16-byte chunks XOR-swizzled by tok & 7
        uint8_t* kt = page + h * 2048 + tok * 128;
        const int c1 = (pi >> 3) ^ (tok & 7), c2 = ((pi + 32) >> 3) ^ (tok & 7);
        *reinterpret_cast<bf16*>(kt + c1 * 16 + (pi & 7) * 2) = kv_b1;
        *reinterpret_cast<bf16*>(kt + c2 * 16 + (pi & 7) * 2) = kv_b2;
      } else {
        bf16* vt = reinterpret_cast<bf16*>(page + 32768 + h * 2048 + tok * 128);
        vt[pi] = kv_b1; vt[pi + 32] = kv_b2;
      }
    }
  }
  BPROF(10);
}

bool block_on() {
  static const bool on = !(getenv("CBX_BLOCK_FUSED") && getenv("CBX_BLOCK_FUSED")[0] == '0');
  return on;
}

template <int TT, bool FOLD>
int launch_block(const BlockArgs& a, cudaStream_t s) {
  static bool attr = false;                            // per instantiation
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(layer_attn_block_kernel<TT, FOLD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Smem<TT>::total);
    if (e != cudaSuccess) { set_error("layer_attn_block smem attribute: %s", cudaGetErrorString(e)); return -2; }
    attr = true;
  }
  cudaError_t e = launch_pdl_cluster(layer_attn_block_kernel<TT, FOLD>, dim3(kCl, kHeads), dim3(kNW * 32), Smem<TT>::total, s, kCl, a);
  if (e != cudaSuccess) { set_error("layer_attn_block launch failed: %s", cudaGetErrorString(e)); return -2; }
  count_launch();
  return 0;
}

// Stand-alone form of what the engine fuses into the next two kernels: res = bf16(bf16(acc * 2^-32) + res), acc = 0.
__global__ void __launch_bounds__(256) block_finalize_kernel(unsigned long long* acc, bf16* res) {
  pdl_trigger();
  pdl_wait();
  const size_t i = (size_t)blockIdx.x * 1024 + threadIdx.x * 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float y = (float)(long long)acc[i + k] * (1.f / kAccScale);
    res[i + k] = __float2bfloat16_rn(round_bf16(y) + __bfloat162float(res[i + k]));
    acc[i + k] = 0ull;
  }
}
}  // namespace

namespace cbx {
// rows the fused block takes, and the longest context (incl. the new token) its shared-memory page buffer holds
bool layer_attn_block_supported(int T, int max_kv_len) {
  if (T < 1 || T > 4 || !block_on()) return false;
  const int TT = T == 1 ? 1 : (T == 2 ? 2 : 4);
  return max_kv_len - 1 <= (kSlices / TT) * kCl * CBX_BLOCK_SIZE;
}
// longest context (incl. the new token) the block takes for T rows; 0 when it is switched off or T is out of range
int layer_attn_block_max_len(int T) {
  if (T < 1 || T > 4 || !block_on()) return 0;
  const int TT = T == 1 ? 1 : (T == 2 ? 2 : 4);
  return (kSlices / TT) * kCl * CBX_BLOCK_SIZE + 1;
}
// acc: [4][1024] u64, zero on entry; on return (stream order) it holds the O-proj output of the T rows in 2^-32 fixed point
int layer_attn_block(const void* res, const void* ln_w, const void* Wqkv, const void* Wo, const void* cos_sin, const int32_t* pos,
                     const int64_t* slot, void* kv_layer, const int32_t* bt, int bt_stride, const int32_t* row_seq,
                     const int32_t* kv_len, void* acc, void* q_out, int T, float eps, cudaStream_t s) {
  BlockArgs a = {};
  a.res = (bf16*)res; a.ln_w = (const bf16*)ln_w; a.Wqkv = (const bf16*)Wqkv; a.Wo = (const bf16*)Wo;
  a.cos_sin = (const bf16*)cos_sin; a.pos = pos; a.slot = slot; a.kv_layer = (uint8_t*)kv_layer; a.bt = bt;
  a.bt_stride = bt_stride; a.row_seq = row_seq; a.kv_len = kv_len; a.acc = (unsigned long long*)acc;
  a.q_out = (bf16*)q_out; a.T = T; a.eps = eps;
  static const int stop = getenv("CBX_BLOCK_STOP") ? atoi(getenv("CBX_BLOCK_STOP")) : 0;
  a.stop = stop;
#ifdef CBX_BLOCK_PROF
  a.prof_id = g_bprof_next++;
#endif
  if (T == 1) return launch_block<1, false>(a, s);
  if (T == 2) return launch_block<2, false>(a, s);
  return launch_block<4, false>(a, s);
}
// FOLD form: the rows enter as res (+ acc_in, the previous layer's down-proj output in 2^-32 fixed point, or NULL); the
// closed rows are written to res_wb (!= res); zero_ptr / zero_bytes: an accumulator to clear after the wait
int layer_attn_block_fold(const void* res, const void* acc_in, void* res_wb, void* zero_ptr, int zero_bytes, const void* ln_w,
                          const void* Wqkv, const void* Wo, const void* cos_sin, const int32_t* pos, const int64_t* slot,
                          void* kv_layer, const int32_t* bt, int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
                          void* acc, int T, float eps, cudaStream_t s) {
  BlockArgs a = {};
  a.res = (bf16*)res; a.ln_w = (const bf16*)ln_w; a.Wqkv = (const bf16*)Wqkv; a.Wo = (const bf16*)Wo;
  a.cos_sin = (const bf16*)cos_sin; a.pos = pos; a.slot = slot; a.kv_layer = (uint8_t*)kv_layer; a.bt = bt;
  a.bt_stride = bt_stride; a.row_seq = row_seq; a.kv_len = kv_len; a.acc = (unsigned long long*)acc;
  a.acc_in = (const unsigned long long*)acc_in; a.res_wb = (bf16*)res_wb; a.zero_ptr = zero_ptr; a.zero_bytes = zero_bytes;
  a.T = T; a.eps = eps;
#ifdef CBX_BLOCK_PROF
  a.prof_id = g_bprof_next++;
#endif
  if (T == 1) return launch_block<1, true>(a, s);
  if (T == 2) return launch_block<2, true>(a, s);
  return launch_block<4, true>(a, s);
}
}  // namespace cbx

extern "C" int64_t cbx_layer_attn_block_scratch_bytes(void) { return (int64_t)4 * 1024 * 8; }
extern "C" int cbx_layer_attn_block_supported(int T, int max_kv_len) { return cbx::layer_attn_block_supported(T, max_kv_len) ? 1 : 0; }
extern "C" int cbx_layer_attn_block(void* res, const void* ln_w, const void* Wqkv, const void* Wo, const void* cos_sin,
                                    const int32_t* pos, const int64_t* slot, void* kv_layer, const int32_t* block_table,
                                    int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* scratch, void* q_out,
                                    int T, int max_kv_len, float eps, void* stream) {
  CBX_REQUIRE(T >= 1 && T <= 4, "cbx_layer_attn_block: T=%d not in [1,4]", T);
  CBX_REQUIRE(cbx::layer_attn_block_supported(T, max_kv_len),
              "cbx_layer_attn_block: context %d exceeds the fused block's page buffer for T=%d (or CBX_BLOCK_FUSED=0)", max_kv_len, T);
  CBX_REQUIRE(scratch != nullptr, "cbx_layer_attn_block: scratch is NULL");
  if (cbx::layer_attn_block(res, ln_w, Wqkv, Wo, cos_sin, pos, slot, kv_layer, block_table, bt_stride, row_seq, kv_len,
                            scratch, q_out, T, eps, (cudaStream_t)stream)) return -1;
  cudaError_t e = launch_pdl(block_finalize_kernel, dim3(T), dim3(256), 0, (cudaStream_t)stream, (unsigned long long*)scratch,
                             (bf16*)res);
  CBX_REQUIRE(e == cudaSuccess, "block_finalize launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

#ifdef CBX_BLOCK_PROF
// debug builds only: stamps of launches [0, n) to host (n * 128 * 12 u64); restarts the launch ids
extern "C" int cbx_block_prof_read(unsigned long long* out, int n) {
  cudaDeviceSynchronize();
  cudaError_t e = cudaMemcpyFromSymbol(out, g_bprof, (size_t)n * 128 * 12 * sizeof(unsigned long long));
  const int used = g_bprof_next;
  g_bprof_next = 0;
  return e == cudaSuccess ? used : -1;
}
#endif

// Diagnostic: how many clusters of the fused block can be co-resident (cudaOccupancyMaxActiveClusters).
extern "C" int cbx_debug_block_max_clusters(int TT, int cluster) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster, kHeads); cfg.blockDim = dim3(kNW * 32);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int n = -1;
  cudaError_t e;
  if (TT == 1) { cfg.dynamicSmemBytes = Smem<1>::total; cudaFuncSetAttribute(layer_attn_block_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<1>::total); e = cudaOccupancyMaxActiveClusters(&n, layer_attn_block_kernel<1, false>, &cfg); }
  else if (TT == 2) { cfg.dynamicSmemBytes = Smem<2>::total; cudaFuncSetAttribute(layer_attn_block_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<2>::total); e = cudaOccupancyMaxActiveClusters(&n, layer_attn_block_kernel<2, false>, &cfg); }
  else { cfg.dynamicSmemBytes = Smem<4>::total; cudaFuncSetAttribute(layer_attn_block_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<4>::total); e = cudaOccupancyMaxActiveClusters(&n, layer_attn_block_kernel<4, false>, &cfg); }
  if (e != cudaSuccess) { set_error("cudaOccupancyMaxActiveClusters: %s", cudaGetErrorString(e)); return -1; }
  return n;
}
