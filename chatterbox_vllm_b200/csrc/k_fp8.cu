// k_fp8.cu — weight preparation for the optional FP8-e4m3 weight path (north star: "an optional
// FP8-e4m3 per-channel weight path is reported separately with its own tolerance"; SURVEY 8(d) cfg 5).
//
// Per OUTPUT feature n (one row of a [N, K] nn.Linear weight): scale[n] = max_k |W[n,k]| / 448 (448 = the
// largest finite e4m3 value), q[n,k] = e4m3_rn_satfinite(W[n,k] / scale[n]).  The consumers (k_gemv.cu,
// k_gemm.cu W8 forms) compute bf16((x . q[n,:]) * scale[n]).  An all-zero row gets scale 1.
#include <cuda_fp8.h>

#include "common.cuh"

using namespace cbx;

namespace {
__global__ void __launch_bounds__(256) quantize_fp8_rows_kernel(const bf16* __restrict__ W, uint8_t* __restrict__ q,
                                                                float* __restrict__ scale, int K) {
  __shared__ float red[8];
  __shared__ float s_scale;
  const bf16* row = W + (size_t)blockIdx.x * K;
  float amax = 0.f;
  for (int k = threadIdx.x; k < K; k += 256) amax = fmaxf(amax, fabsf(__bfloat162float(row[k])));
  amax = warp_max(amax);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]);
    const float sc = m > 0.f ? __fdiv_rn(m, 448.f) : 1.f;
    s_scale = sc;
    scale[blockIdx.x] = sc;
  }
  __syncthreads();
  const float sc = s_scale;
  uint8_t* out = q + (size_t)blockIdx.x * K;
  for (int k = threadIdx.x; k < K; k += 256)
    out[k] = (uint8_t)__nv_cvt_float_to_fp8(__fdiv_rn(__bfloat162float(row[k]), sc), __NV_SATFINITE, __NV_E4M3);
}
}  // namespace

extern "C" int cbx_quantize_fp8_rows(const void* W, void* W8, float* scale, int N, int K, void* stream) {
  CBX_REQUIRE(W && W8 && scale && N >= 1 && K >= 1, "cbx_quantize_fp8_rows: bad arguments");
  quantize_fp8_rows_kernel<<<N, 256, 0, (cudaStream_t)stream>>>((const bf16*)W, (uint8_t*)W8, scale, K);
  CBX_CHECK_LAUNCH("quantize_fp8_rows");
  return 0;
}
