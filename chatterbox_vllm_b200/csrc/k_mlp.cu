// k_mlp.cu — the MLP half of a Llama layer as ONE kernel for T <= 4 rows (the 1-2 CFG-pair latency regime):
// RMSNorm(res)·ln2 -> gate_up -> SiLU(gate)·up -> down-proj, the result left in a 64-bit fixed-point accumulator.
//
// Replaces the last two kernels of the GEMV chain (k_gemv.cu gate_up, down+res; vllm llama.py:117-120,323-326) and
// the all-to-all exchange between them.  down(act) is a sum over the 4096 MLP features, so a CTA that owns a
// contiguous slice of features can finish its share of the down-projection alone:
//
//   grid      one CTA per SM (148); CTA b owns features [4096 b / G, 4096 (b+1) / G) = 27 or 28 of them; 16 warps.
//   gate_up   warp w streams the gate and the up row of features f0 + w and f0 + 16 + w (2 x 2 x 2 KB, all requested
//             BEFORE griddepcontrol.wait, like the GEMV kernels — same lane mapping, same summation order, so the
//             activations carry the bits of the chain's), SiLU(g)·u rounded to bf16 into shared memory.
//   down      the CTA's slice of the TRANSPOSED down weights (WdT[f, :] = Wd[:, f]: 28 contiguous 2 KB rows) arrives
//             by one cp.async.bulk into shared memory, also requested before the wait.  Thread i owns output features
//             2i, 2i + 1: 28 LDS.32 + 56 T FMAs.
//   reduce    the CTA's fp32 partial [T, 1024] goes to shared memory as 2^-32 fixed point and ONE bulk reduction
//             (cp.reduce.async.bulk .add.u64, SASS UBLKRED) adds it into the accumulator shared by all CTAs: integer
//             adds commute, so the sum does not depend on arrival order (bit-reproducible), and the SM issues one
//             instruction instead of 1024 T REDs.
//
// Like the fused attention block (k_block.cu) the kernel closes the PREVIOUS residual branch in its norm prologue
// (x = bf16(bf16(acc_in * 2^-32) + x), CTA 0 writing the rows back) and leaves its own branch open: the next
// layer's attention block (or cbx_acc_finalize) folds acc_out.
#include <stdlib.h>

#include "common.cuh"

using namespace cbx;

namespace {
constexpr int kNW = 16;                    // warps per CTA (4 per SM sub-partition: 128 registers per thread; 14 warps
                                           // get no more — the register file is split per sub-partition)
constexpr int kFMax = 2 * kNW;             // features a CTA can own: two rounds of one feature per warp
constexpr int kF = 4096;                   // MLP features
constexpr int kMinGrid = (kF + kFMax - 1) / kFMax;   // 128
constexpr float kAccScale = 4294967296.f;  // 2^32
constexpr int kCopies = kMlpAccCopies;     // copies of the accumulator (common.cuh: 1)

struct MlpArgs {
  const bf16* x;                     // [T,1024] residual rows, the attention branch still open when acc_in != NULL
  const unsigned long long* acc_in;  // [4][1024] O-proj output of the rows in 2^-32 fixed point, or NULL
  bf16* res_wb;                      // closed residual rows, written by CTA 0 (NOT the buffer x points at)
  const bf16* ln_w;                  // [1024]
  const bf16* Wgu;                   // [8192,1024] packed gate/up: 32-row groups of 16 gate rows then their 16 up rows
  const bf16* WdT;                   // [4096,1024] transposed down-proj weights
  unsigned long long* acc_out;       // [kCopies][4][1024] accumulator of the down-proj output (zero on entry)
  void* zero_ptr; int zero_bytes;    // cleared after the wait (an accumulator nobody reads any more)
  int T; float eps;
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}

// Weight load that stays where it is written: ptxas moves the non-coherent form (ld.global.nc, common.cuh ld_stream)
// freely across griddepcontrol.wait — it sank all sixteen of the 4-row kernel's below the wait.  A coherent load with
// a memory clobber is ordered against the wait; L1 is bypassed either way.
__device__ __forceinline__ uint4 ld_weights(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

template <int TT>
struct Smem {
  static constexpr int wd = 0;                          // [kFMax][1024] bf16: this CTA's rows of WdT
  static constexpr int xs = wd + kFMax * 2048;          // [TT][1024] bf16: normalised rows
  static constexpr int out = xs + TT * 2048;            // [TT][1024] u64: fixed-point partial (source of the bulk reduction)
  static constexpr int act = out + TT * 8192;           // float [TT][kFMax]: SiLU(g)·u (bf16 values)
  static constexpr int ss = act + TT * kFMax * 4;       // float [TT][16]: sum-of-squares partials per warp
  static constexpr int bar = ss + TT * 16 * 4;          // mbarrier of the WdT copy
  static constexpr int total = bar + 16;
};

template <int TT>
__global__ void __launch_bounds__(kNW * 32, 1) layer_mlp_block_kernel(const MlpArgs a) {
  using S = Smem<TT>;
  constexpr int kFoldIters = (TT * 128 + kNW * 32 - 1) / (kNW * 32);   // 16-byte chunks of the rows per thread
  extern __shared__ __align__(1024) uint8_t smem[];
  uint4* xs = reinterpret_cast<uint4*>(smem + S::xs);
  float* act_s = reinterpret_cast<float*>(smem + S::act);
  float* ss_s = reinterpret_cast<float*>(smem + S::ss);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + S::bar);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = a.T;
  const int f_lo = (int)((blockIdx.x * (unsigned)kF) / gridDim.x);
  const int f_hi = (int)(((blockIdx.x + 1) * (unsigned)kF) / gridDim.x);
  const int cnt = f_hi - f_lo;                         // <= kFMax: the host launches >= kMinGrid CTAs

  pdl_trigger();
  // ---- before the wait: every weight byte this CTA needs ----------------------------------------------------
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
    mbar_arrive_expect_tx(bar, (uint32_t)cnt * 2048u);
    bulk_g2s(smem + S::wd, a.WdT + (size_t)f_lo * 1024, (uint32_t)cnt * 2048u, bar);
  }
  uint4 w0[2][4], w1[2][4];                            // gate / up rows of the warp's first and second feature
  const bool has0 = warp < cnt, has1 = warp + kNW < cnt;
  if (has0) {
    const int f = f_lo + warp, ra = (f >> 4) * 32 + (f & 15);
    const uint4* wa = reinterpret_cast<const uint4*>(a.Wgu + (size_t)ra * 1024) + lane;
    const uint4* wb = reinterpret_cast<const uint4*>(a.Wgu + (size_t)(ra + 16) * 1024) + lane;
#pragma unroll
    for (int j = 0; j < 4; ++j) { w0[0][j] = ld_weights(wa + j * 32); w0[1][j] = ld_weights(wb + j * 32); }
  }
  if (has1) {
    const int f = f_lo + warp + kNW, ra = (f >> 4) * 32 + (f & 15);
    const uint4* wa = reinterpret_cast<const uint4*>(a.Wgu + (size_t)ra * 1024) + lane;
    const uint4* wb = reinterpret_cast<const uint4*>(a.Wgu + (size_t)(ra + 16) * 1024) + lane;
#pragma unroll
    for (int j = 0; j < 4; ++j) { w1[0][j] = ld_weights(wa + j * 32); w1[1][j] = ld_weights(wb + j * 32); }
  }
  // norm weights in the prologue's mapping: thread -> 16-byte chunks ci = tid + k * 512 of the TT x 128 chunks
  uint4 lnc[kFoldIters];
#pragma unroll
  for (int k = 0; k < kFoldIters; ++k) {
    const int ci = threadIdx.x + k * kNW * 32;
    lnc[k] = make_uint4(0, 0, 0, 0);
    if (ci < TT * 128) lnc[k] = ld_weights(reinterpret_cast<const uint4*>(a.ln_w) + (ci & 127));
  }
  pdl_wait();

  // ---- after the wait ---------------------------------------------------------------------------------------
  if (a.zero_ptr != nullptr) {
    uint4* z = reinterpret_cast<uint4*>(a.zero_ptr);
    for (int i = blockIdx.x * (kNW * 32) + threadIdx.x; i < a.zero_bytes / 16; i += gridDim.x * (kNW * 32)) z[i] = make_uint4(0, 0, 0, 0);
  }
  // prologue: close the attention branch, x = bf16(bf16(acc_in * 2^-32) + x), then RMSNorm — all warps share the rows
  {
    float ssl[TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) ssl[t] = 0.f;
    uint4 xv[kFoldIters], av[kFoldIters][4];
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {             // every load of the thread in flight together
      const int ci = threadIdx.x + k * kNW * 32;
      if (ci < TT * 128 && (ci >> 7) < T) {
        xv[k] = reinterpret_cast<const uint4*>(a.x)[ci];
        if (a.acc_in != nullptr) {
#pragma unroll
          for (int q = 0; q < 4; ++q) av[k][q] = reinterpret_cast<const uint4*>(a.acc_in)[ci * 4 + q];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {
      const int ci = threadIdx.x + k * kNW * 32;
      if (ci < TT * 128) {
        const int t = ci >> 7;                         // warp-uniform: 32 consecutive chunks lie in one row
        uint4 o = make_uint4(0, 0, 0, 0);
        if (t < T) {
          float f[8];
          unpack8(xv[k], f);
          if (a.acc_in != nullptr) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const long long lo = (long long)(((unsigned long long)av[k][q].y << 32) | av[k][q].x);
              const long long hi = (long long)(((unsigned long long)av[k][q].w << 32) | av[k][q].z);
              f[2 * q] = round_bf16(round_bf16((float)lo * (1.f / kAccScale)) + f[2 * q]);
              f[2 * q + 1] = round_bf16(round_bf16((float)hi * (1.f / kAccScale)) + f[2 * q + 1]);
            }
          }
          o = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
          if (blockIdx.x == 0 && a.res_wb != nullptr) reinterpret_cast<uint4*>(a.res_wb)[ci] = o;
          float ss = 0.f;
#pragma unroll
          for (int e = 0; e < 8; ++e) ss += f[e] * f[e];
#pragma unroll
          for (int tt = 0; tt < TT; ++tt) if (tt == t) ssl[tt] += ss;
        }
        xs[ci] = o;                                    // the closed residual row, not yet normalised (rows >= T: zeros)
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) {
      const float v = warp_sum(ssl[t]);
      if (lane == 0) ss_s[t * 16 + warp] = v;
    }
    for (int i = threadIdx.x; i < TT * kFMax; i += kNW * 32) act_s[i] = 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kFoldIters; ++k) {
      const int ci = threadIdx.x + k * kNW * 32;
      if (ci < TT * 128 && (ci >> 7) < T) {
        const int t = ci >> 7;
        float ss = 0.f;
#pragma unroll
        for (int w = 0; w < kNW; ++w) ss += ss_s[t * 16 + w];   // warp order: every thread (and every CTA) gets the same bits
        const float rstd = rsqrtf(ss * (1.0f / 1024.f) + a.eps);
        float f[8], w[8];
        unpack8(xs[ci], f);
        unpack8(lnc[k], w);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = round_bf16(f[e] * rstd) * w[e];
        xs[ci] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      }
    }
    __syncthreads();
  }

  // ---- gate_up: two features per warp, the GEMV kernel's summation order ------------------------------------
#pragma unroll
  for (int rnd = 0; rnd < 2; ++rnd) {
    const bool has = rnd == 0 ? has0 : has1;           // warp-uniform
    if (has) {
      float acc[2][TT];
#pragma unroll
      for (int t = 0; t < TT; ++t) acc[0][t] = acc[1][t] = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float fa[8], fb[8];
        unpack8(rnd == 0 ? w0[0][j] : w1[0][j], fa);
        unpack8(rnd == 0 ? w0[1][j] : w1[1][j], fb);
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
      if (lane < T) {
        const float g = round_bf16(v0), u = round_bf16(v1);
        act_s[lane * kFMax + rnd * kNW + warp] = round_bf16(g / (1.f + __expf(-g)) * u);
      }
    }
  }
  __syncthreads();

  // ---- down: thread i owns output features 2i, 2i + 1 over the CTA's features --------------------------------
  {
    float o0[TT], o1[TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) o0[t] = o1[t] = 0.f;
    mbar_wait(bar, 0);
    const uint32_t* wd = reinterpret_cast<const uint32_t*>(smem + S::wd) + threadIdx.x;
#pragma unroll 4
    for (int f = 0; f < cnt; ++f) {
      const uint32_t w = wd[f * 512];
      const float wl = bf_lo(w), wh = bf_hi(w);
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        const float av = act_s[t * kFMax + f];
        o0[t] = fmaf(av, wl, o0[t]);
        o1[t] = fmaf(av, wh, o1[t]);
      }
    }
    unsigned long long* out = reinterpret_cast<unsigned long long*>(smem + S::out);
#pragma unroll
    for (int t = 0; t < TT; ++t) {
      if (t < T) {
        ulonglong2 v;
        v.x = (unsigned long long)__float2ll_rn(o0[t] * kAccScale);
        v.y = (unsigned long long)__float2ll_rn(o1[t] * kAccScale);
        *reinterpret_cast<ulonglong2*>(out + t * 1024 + 2 * threadIdx.x) = v;
      }
    }
  }
  fence_proxy_async();                                 // the partial is read by the async proxy
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;"
                 ::"l"(a.acc_out + (size_t)(blockIdx.x % kCopies) * 4096), "r"(smem_u32(smem + S::out)), "r"(T * 8192) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory must outlive the read
  }
}

template <int TT>
int launch_mlp(const MlpArgs& a, cudaStream_t s) {
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(layer_mlp_block_kernel<TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<TT>::total);
    if (e != cudaSuccess) { set_error("layer_mlp_block smem attribute: %s", cudaGetErrorString(e)); return -2; }
    attr = true;
  }
  const int grid = sm_count() >= kMinGrid ? sm_count() : kMinGrid;   // a CTA's two rounds cover at most 32 features
  cudaError_t e = launch_pdl(layer_mlp_block_kernel<TT>, dim3(grid), dim3(kNW * 32), Smem<TT>::total, s, a);
  if (e != cudaSuccess) { set_error("layer_mlp_block launch failed: %s", cudaGetErrorString(e)); return -2; }
  count_launch();
  return 0;
}

// res = bf16(bf16(sum of the copies * 2^-32) + res), copies = 0: the stand-alone form of what the next kernel's prologue does
__global__ void __launch_bounds__(256) acc_finalize_kernel(unsigned long long* acc, bf16* res) {
  pdl_trigger();
  pdl_wait();
  const size_t i = (size_t)blockIdx.x * 1024 + threadIdx.x * 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    unsigned long long sum = 0ull;
#pragma unroll
    for (int c = 0; c < kCopies; ++c) { sum += acc[(size_t)c * 4096 + i + k]; acc[(size_t)c * 4096 + i + k] = 0ull; }
    const float y = (float)(long long)sum * (1.f / kAccScale);
    res[i + k] = __float2bfloat16_rn(round_bf16(y) + __bfloat162float(res[i + k]));
  }
}
}  // namespace

namespace cbx {
// env CBX_MLP_FUSED=1: the engine's T <= 4 layer becomes [attention block] [MLP block] (two launches).  OFF by default:
// parity-green, but on B200 the b = 1 step measures 0.512-0.518 ms against 0.512-0.514 ms for the three-launch layer —
// the launch it removes is paid back by the 148 same-address bulk reductions and the wider fold prologue.  Read at
// every call (a captured graph keeps what it was captured with; engines created afterwards see the new value).
bool layer_mlp_block_on() {
  const char* v = getenv("CBX_MLP_FUSED");
  return v != nullptr && v[0] == '1';
}
// x: residual rows (attention branch open if acc_in != NULL); res_wb (!= x): where CTA 0 writes the closed rows;
// acc_out: [kMlpAccCopies][4][1024] u64, zero on entry; on return (stream order) the copies add up to the down-proj output of
// the T rows in 2^-32 fixed point
int layer_mlp_block(const void* x, const void* acc_in, void* res_wb, const void* ln_w, const void* Wgu, const void* WdT,
                    void* acc_out, void* zero_ptr, int zero_bytes, int T, float eps, cudaStream_t s) {
  MlpArgs a = {};
  a.x = (const bf16*)x; a.acc_in = (const unsigned long long*)acc_in; a.res_wb = (bf16*)res_wb; a.ln_w = (const bf16*)ln_w;
  a.Wgu = (const bf16*)Wgu; a.WdT = (const bf16*)WdT; a.acc_out = (unsigned long long*)acc_out;
  a.zero_ptr = zero_ptr; a.zero_bytes = zero_bytes; a.T = T; a.eps = eps;
  if (T == 1) return launch_mlp<1>(a, s);
  if (T == 2) return launch_mlp<2>(a, s);
  return launch_mlp<4>(a, s);
}
int acc_finalize(void* acc, void* res, int T, cudaStream_t s) {
  cudaError_t e = launch_pdl(acc_finalize_kernel, dim3(T), dim3(256), 0, s, (unsigned long long*)acc, (bf16*)res);
  if (e != cudaSuccess) { set_error("acc_finalize launch failed: %s", cudaGetErrorString(e)); return -2; }
  count_launch();
  return 0;
}
}  // namespace cbx

// Stand-alone form (tests, other hosts): res <- res + down(SiLU(gate)·up of RMSNorm(res)·ln_w), through the accumulator.
// scratch: cbx_layer_mlp_block_scratch_bytes() bytes, zero on entry and zero again on return.
extern "C" int64_t cbx_layer_mlp_block_scratch_bytes(void) { return (int64_t)kCopies * 4 * 1024 * 8; }
extern "C" int cbx_layer_mlp_block(void* res, const void* ln_w, const void* Wgu, const void* WdT, void* scratch, int T,
                                   float eps, void* stream) {
  CBX_REQUIRE(T >= 1 && T <= 4, "cbx_layer_mlp_block: T=%d not in [1,4]", T);
  CBX_REQUIRE(res && ln_w && Wgu && WdT && scratch, "cbx_layer_mlp_block: NULL argument");
  if (cbx::layer_mlp_block(res, nullptr, nullptr, ln_w, Wgu, WdT, scratch, nullptr, 0, T, eps, (cudaStream_t)stream)) return -1;
  return cbx::acc_finalize(scratch, res, T, (cudaStream_t)stream) ? -1 : 0;
}
