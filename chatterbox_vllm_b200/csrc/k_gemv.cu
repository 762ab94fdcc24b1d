// k_gemv.cu — weight-streaming GEMV for T <= 4 rows (the b=1 / b=2 CFG-pair decode case).
//
// y[T,N] = x[T,K] · W[N,K]^T.  Replaces the cuBLAS GEMM vLLM would launch for
// llama.py:117-120,221,225 at tiny batch.  HBM-bound: every weight byte is read once with
// 128-bit L1-bypassing loads, two output features per warp in flight with the next K-segment
// prefetched into registers; x (T*K bf16, <= 32 KB) is staged once per CTA in shared memory.
#include "common.cuh"

using namespace cbx;

namespace {
constexpr int kWarps = 8;

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}

// TT = padded row count (1,2,4); SEGS = K / 1024; EPI: 0 store, 1 SiLU(gate)*up (packed W).
template <int TT, int SEGS, int EPI>
__global__ void __launch_bounds__(kWarps * 32, 2) gemv_kernel(const bf16* __restrict__ x,
                                                              const bf16* __restrict__ W, bf16* __restrict__ y,
                                                              int T, int N, int ldy) {
  constexpr int K = SEGS * 1024;
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint4* xs = reinterpret_cast<uint4*>(smem_raw);            // [TT][K/8]
  for (int i = threadIdx.x; i < TT * (K / 8); i += blockDim.x) {
    const int t = i / (K / 8);
    xs[i] = t < T ? reinterpret_cast<const uint4*>(x)[(size_t)t * (K / 8) + (i - t * (K / 8))] : make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_pairs = N / 2;
  for (int pr = blockIdx.x * kWarps + warp; pr < n_pairs; pr += gridDim.x * kWarps) {
    int ra, rb;
    if (EPI == 1) { ra = (pr >> 4) * 32 + (pr & 15); rb = ra + 16; }
    else          { ra = 2 * pr; rb = ra + 1; }
    const uint4* wa = reinterpret_cast<const uint4*>(W + (size_t)ra * K) + lane;
    const uint4* wb = reinterpret_cast<const uint4*>(W + (size_t)rb * K) + lane;
    float acc[2][TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) acc[0][t] = acc[1][t] = 0.f;
    uint4 cur[2][4], nxt[2][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { cur[0][j] = ld_stream(wa + j * 32); cur[1][j] = ld_stream(wb + j * 32); }
#pragma unroll
    for (int sg = 0; sg < SEGS; ++sg) {
      if (sg + 1 < SEGS) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          nxt[0][j] = ld_stream(wa + (sg + 1) * 128 + j * 32);
          nxt[1][j] = ld_stream(wb + (sg + 1) * 128 + j * 32);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float fa[8], fb[8];
        unpack8(cur[0][j], fa);
        unpack8(cur[1][j], fb);
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float fx[8];
          unpack8(xs[t * (K / 8) + sg * 128 + j * 32 + lane], fx);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            acc[0][t] = fmaf(fa[e], fx[e], acc[0][t]);
            acc[1][t] = fmaf(fb[e], fx[e], acc[1][t]);
          }
        }
      }
      if (sg + 1 < SEGS) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { cur[0][j] = nxt[0][j]; cur[1][j] = nxt[1][j]; }
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) { acc[0][t] = warp_sum(acc[0][t]); acc[1][t] = warp_sum(acc[1][t]); }
    if (lane == 0) {
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        if (t < T) {
          if (EPI == 1) {
            const float g = round_bf16(acc[0][t]), u = round_bf16(acc[1][t]);
            y[(size_t)t * ldy + pr] = __float2bfloat16_rn(g / (1.f + __expf(-g)) * u);
          } else {
            *reinterpret_cast<uint32_t*>(y + (size_t)t * ldy + ra) = pack_bf16(acc[0][t], acc[1][t]);
          }
        }
      }
    }
  }
}

template <int TT, int SEGS, int EPI>
int launch(const void* x, const void* W, void* y, int T, int N, int ldy, cudaStream_t s) {
  const int smem = TT * SEGS * 1024 * 2;
  static_assert(TT * SEGS * 1024 * 2 <= 48 * 1024, "x tile must fit the default dynamic smem limit");
  int grid = (N / 2 + kWarps - 1) / kWarps;
  if (grid > 296) grid = 296;
  gemv_kernel<TT, SEGS, EPI><<<grid, kWarps * 32, smem, s>>>((const bf16*)x, (const bf16*)W, (bf16*)y, T, N, ldy);
  CBX_CHECK_LAUNCH("gemv");
  return 0;
}

template <int SEGS, int EPI>
int dispatch_t(const void* x, const void* W, void* y, int T, int N, int ldy, cudaStream_t s) {
  if (T == 1) return launch<1, SEGS, EPI>(x, W, y, T, N, ldy, s);
  if (T == 2) return launch<2, SEGS, EPI>(x, W, y, T, N, ldy, s);
  return launch<4, SEGS, EPI>(x, W, y, T, N, ldy, s);
}
}  // namespace

namespace cbx {
// T <= 4, K in {1024, 4096}, N even (N % 32 == 0 for the SiLU epilogue).
int gemv_bf16(const void* x, const void* W, void* y, int T, int N, int K, int ldy, int epilogue,
              cudaStream_t s) {
  CBX_REQUIRE(T >= 1 && T <= 4, "gemv: T=%d not in [1,4]", T);
  CBX_REQUIRE(K == 1024 || K == 4096, "gemv: K=%d unsupported", K);
  CBX_REQUIRE(N % 2 == 0 && (epilogue == 0 || N % 32 == 0), "gemv: N=%d unsupported", N);
  if (K == 1024) {
    if (epilogue == 1) return dispatch_t<1, 1>(x, W, y, T, N, ldy, s);
    return dispatch_t<1, 0>(x, W, y, T, N, ldy, s);
  }
  CBX_REQUIRE(epilogue == 0, "gemv: SiLU epilogue only for K=1024");
  return dispatch_t<4, 0>(x, W, y, T, N, ldy, s);
}
}  // namespace cbx
