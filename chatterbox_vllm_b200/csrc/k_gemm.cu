// k_gemm.cu — bf16 GEMM on the 5th-gen tensor cores: y[M,N] = x[M,K] · W[N,K]^T.
//
// Replaces the cuBLASLt GEMMs behind vllm llama.py:117-120,221,225 and the two head GEMMs of
// t3.py:729-730 once the cond+uncond batch makes them a dense contraction (M > 4 rows).
//
// One 128 x BN output tile per CTA, BK = 64 (one 128-byte swizzle atom per row):
//   warp 0  TMA producer  : cp.async.bulk.tensor 2-D tiles of x and W (SWIZZLE_128B) into a
//                           kStages-deep shared-memory ring, mbarrier complete_tx;
//   warp 1  MMA issuer    : one elected thread issues tcgen05.mma.cta_group::1.kind::f16
//                           (M=128, N=BN, K=16) x4 per stage, accumulator in TMEM;
//                           tcgen05.commit frees the stage / signals the epilogue;
//   warp 2  TMEM allocator;
//   warps 4-7 epilogue    : tcgen05.ld 32x32b -> registers -> bf16 -> one of
//                             0 store | 1 SiLU(gate)·up | 2 RoPE + q store + KV-page append (one
//                             64-wide tile = one head) | 3 residual add in place.
// Programmatic dependent launch: the producer issues the WEIGHT tiles of the first kStages
// k-blocks before griddepcontrol.wait, so the weight stream overlaps the previous kernel's tail;
// activation tiles follow after the wait.
//
// W8 = true (weight-only FP8-e4m3, per-output-channel fp32 scale): tcgen05.mma kind::f16 takes no
// 8-bit B operand next to a bf16 A, so the weight tile travels HBM -> L2 -> shared memory as BYTES
// (half the ingest, which is what bounds the decode GEMMs) and is widened on chip: TMA drops the
// BN x 64-byte e4m3 tile into a linear staging buffer, eight converter warps (4-7, idle until the
// epilogue, plus four extra warps 8-11 the W8 form launches) widen it
// to bf16 (exact; integer + half-pipe ops only: common.cuh e4m3x4_to_bf16x2_div256 — a bf16 A next to
// an f16 B raises an illegal-instruction error on B200) straight into the SWIZZLE_128B operand layout,
// fence.proxy.async, and arrive on a per-stage "converted" barrier the MMA issuer waits on.  Weights do not depend on the previous
// kernel, so the first ring fill is loaded AND converted before griddepcontrol.wait.  The scale
// multiplies the fp32 accumulator in the epilogue, before the first bf16 rounding point.
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <tuple>

#include "common.cuh"

using namespace cbx;
namespace cbx { int init_gemm(); }

namespace {
constexpr int BM = 128, BK = 64;
constexpr int kThreads = 256;

// -DCBX_GEMM_PROF: per-CTA %globaltimer stamps of every launch (tools/gemm_timeline.py reads them through
// cbx_gemm_prof_read).  Stamps: 0 entry, 1 setup done, 2 producer past griddepcontrol.wait, 3 first stage
// landed, 4 last MMA issued, 5 accumulator complete, 6 epilogue done.
#ifdef CBX_GEMM_PROF
constexpr int kProfLaunches = 256, kProfCtas = 160;
__device__ unsigned long long g_prof[kProfLaunches][kProfCtas][8];
int g_prof_next = 0;
__device__ __forceinline__ void prof_stamp(int id, int k) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (id >= 0 && id < kProfLaunches && cta < kProfCtas) g_prof[id][cta][k] = t;
}
#define PROF(k) prof_stamp(ea.prof_id, k)
#else
#define PROF(k)
#endif

// CN = 2: the CTA pair of a cluster issues ONE tcgen05.mma.cta_group::2 (M = 256: 128 rows per CTA) per
// K = 16 step; each CTA stages its own 128 activation rows and only HALF of the BN weight rows.
template <int BN, int CN = 1, bool W8 = false>
struct Cfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN / CN * BK * 2;
  static constexpr int kB8Bytes = W8 ? BN * BK : 0;     // e4m3 staging tile (linear rows of 64 bytes)
  static constexpr int kStageBytes = kABytes + kBBytes + kB8Bytes;
  static constexpr int kStages = W8 ? (BN > 64 ? 5 : 6) : (kStageBytes > 24576 ? 6 : 8);
  static constexpr int kSmem = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  // Throughput configuration for grids of many tiles (prefill): a ring shallow enough that TWO CTAs
  // share an SM, so one CTA's epilogue (TMEM -> registers -> global) runs under the other's MMAs.
  static constexpr int kStagesShallow = kStageBytes > 24576 ? 3 : 4;
  static constexpr int kSmemShallow = kStagesShallow * kStageBytes + 1024 + 256;
};

struct EpiArgs {               // EPI 2 only
  const bf16* cos_sin; const int32_t* pos; const int64_t* slot; uint8_t* kv_layer;
  int prof_id;                 // CBX_GEMM_PROF builds only
  // EPI 5 only (split-K partials folded by the CTAs that produced them): residual stream [M,1024] (updated in place),
  // norm weights [1024], normalised rows out [M,1024], two counters per row tile (arrived, done; zero between launches)
  bf16* f_res; const bf16* f_ln; bf16* f_out; unsigned* f_cnt; float f_eps;
};

// EPI 5 tail: rows [ra, ra + 2) of the residual stream (N = 1024): res = bf16(bf16(sum of the S K-slices) + res),
// out = bf16(bf16(res * rstd) * w) — thread t owns columns 4t .. 4t+3; the adds, the warp sums and the 8-warp tree are
// rmsnorm_reduce4_kernel's (k_misc.cu), so the rows carry the same bits as the two-kernel form's.
template <int S>
__device__ __forceinline__ void fold_two_rows(const float4* part, int M, int ra, const EpiArgs& ea, float (*sums)[8]) {
  const int t = threadIdx.x;
  const uint2 wv = __ldg(reinterpret_cast<const uint2*>(ea.f_ln) + t);
  uint2* res = reinterpret_cast<uint2*>(ea.f_res);
  float4 a[2][S];
  uint2 q[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    if (ra + k < M) {
      q[k] = __ldcg(res + (size_t)(ra + k) * 256 + t);
#pragma unroll
      for (int z = 0; z < S; ++z) a[k][z] = __ldcg(part + ((size_t)z * M + ra + k) * 256 + t);
    }
  }
  float f[2][4];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    float ss = 0.f;
    if (ra + k < M) {
      float4 sum = a[k][0];
#pragma unroll
      for (int z = 1; z < S; ++z) { sum.x += a[k][z].x; sum.y += a[k][z].y; sum.z += a[k][z].z; sum.w += a[k][z].w; }
      f[k][0] = round_bf16(sum.x); f[k][1] = round_bf16(sum.y); f[k][2] = round_bf16(sum.z); f[k][3] = round_bf16(sum.w);
      const float g[4] = {bf_lo(q[k].x), bf_hi(q[k].x), bf_lo(q[k].y), bf_hi(q[k].y)};
#pragma unroll
      for (int i = 0; i < 4; ++i) f[k][i] = round_bf16(f[k][i] + g[i]);
      res[(size_t)(ra + k) * 256 + t] = make_uint2(pack_bf16(f[k][0], f[k][1]), pack_bf16(f[k][2], f[k][3]));
      ss = (f[k][0] * f[k][0] + f[k][1] * f[k][1]) + (f[k][2] * f[k][2] + f[k][3] * f[k][3]);
    }
    ss = warp_sum(ss);
    if ((t & 31) == 0) sums[k][t >> 5] = ss;
  }
  __syncthreads();
  const float wf[4] = {bf_lo(wv.x), bf_hi(wv.x), bf_lo(wv.y), bf_hi(wv.y)};
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    if (ra + k < M) {
      const float* sm = sums[k];
      const float tot = ((sm[0] + sm[1]) + (sm[2] + sm[3])) + ((sm[4] + sm[5]) + (sm[6] + sm[7]));
      const float rstd = rsqrtf(tot * (1.0f / CBX_HIDDEN) + ea.f_eps);
#pragma unroll
      for (int i = 0; i < 4; ++i) f[k][i] = round_bf16(f[k][i] * rstd) * wf[i];
      reinterpret_cast<uint2*>(ea.f_out)[(size_t)(ra + k) * 256 + t] = make_uint2(pack_bf16(f[k][0], f[k][1]), pack_bf16(f[k][2], f[k][3]));
    }
  }
  __syncthreads();                                     // sums are reused by the next pair of rows
}

// CN = 2: CTA pair along M (grid.x = row tiles, cluster dims (2, 1, 1), rank 0 = leader).  The decode GEMMs are bound by the
// bytes entering the SMs (activation re-reads are most of them); a pair computes a
// 256 x BN tile while each CTA ingests 128 activation rows + BN/2 weight rows per k-step, i.e. the
// weight bytes per output element halve.  Both producers signal the LEADER's full barrier (peer bit of
// the barrier address cleared, .cta_group::2 copy), the leader's elected thread issues the MMAs and
// multicasts its commits to both CTAs' empty / accumulator barriers, each CTA drains its own TMEM.
// 16 e4m3 bytes -> 16 bf16 (= value / 256, exact) as two 16-byte operand chunks
__device__ __forceinline__ void e4m3x16_to_bf16(const uint4& v, uint4& lo8, uint4& hi8) {
  e4m3x4_to_bf16x2_div256(v.x, lo8.x, lo8.y);
  e4m3x4_to_bf16x2_div256(v.y, lo8.z, lo8.w);
  e4m3x4_to_bf16x2_div256(v.z, hi8.x, hi8.y);
  e4m3x4_to_bf16x2_div256(v.w, hi8.z, hi8.w);
}
// r[i] *= 256 * scale[i], i < 32 (scale 16-byte aligned; the same 32 values for every row of the tile; 256 undoes
// the widening's exponent offset)
__device__ __forceinline__ void scale32(uint32_t (&r)[32], const float* __restrict__ sc) {
  const float4* sp = reinterpret_cast<const float4*>(sc);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float4 v = __ldg(sp + i);
    v.x *= 256.f; v.y *= 256.f; v.z *= 256.f; v.w *= 256.f;
    r[4 * i] = __float_as_uint(__uint_as_float(r[4 * i]) * v.x);
    r[4 * i + 1] = __float_as_uint(__uint_as_float(r[4 * i + 1]) * v.y);
    r[4 * i + 2] = __float_as_uint(__uint_as_float(r[4 * i + 2]) * v.z);
    r[4 * i + 3] = __float_as_uint(__uint_as_float(r[4 * i + 3]) * v.w);
  }
}

// W8 launches 12 warps: the extra four (8-11) are converters only.
constexpr int kThreads8 = 384;
template <int BN, int EPI, int CN, bool W8 = false>
__global__ void __launch_bounds__(W8 ? kThreads8 : kThreads, W8 ? 1 : 2) gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                              const __grid_constant__ CUtensorMap tma_b,
                                                              bf16* __restrict__ y, int M, int N, int K, int ldy,
                                                              const EpiArgs ea, const int n_stages,
                                                              const float* __restrict__ wscale) {
  using C = Cfg<BN, CN, W8>;
  static_assert(!W8 || CN == 1, "the FP8-weight form uses single-CTA tiles");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + n_stages * C::kStageBytes);
  uint64_t* empty = full + n_stages;
  uint64_t* tmem_full = empty + n_stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  uint64_t* fullb = tmem_full + 2;                     // W8: e4m3 staging tile landed
  uint64_t* conv = fullb + n_stages;                   // W8: bf16 operand tile written (the 2 warps of a converter group)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // pair mode: the two row tiles of a cluster are neighbours along grid.x (cluster dims (2, 1, 1))
  const int n0 = (CN == 1 ? blockIdx.x : blockIdx.y) * BN, m0 = (CN == 1 ? blockIdx.y : blockIdx.x) * BM;
  const int n_kb = K / BK / gridDim.z;                 // k-blocks of this CTA's K slice (split-K over blockIdx.z)
  const int kb0 = blockIdx.z * n_kb;

  uint32_t crank = 0;
  if (CN > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  constexpr uint16_t kMask = (uint16_t)((1u << CN) - 1);

  pdl_trigger();
  if (threadIdx.x == 0) PROF(0);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < n_stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    if (W8) for (int s = 0; s < n_stages; ++s) { mbar_init(&fullb[s], 1); mbar_init(&conv[s], 2); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    if (CN == 1) { tmem_alloc(tmem_slot, BN); tmem_relinquish(); }
    else { tmem_alloc_pair(tmem_slot, BN); tmem_relinquish_pair(); }
  }
  tc_fence_before();
  __syncthreads();
  if (CN > 1)      // the peer's barriers must exist before anyone signals them
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) PROF(1);

  if (warp == 0) {
    if (lane == 0) {
      auto kc = [&](int kb) { return kb0 + kb; };
      // pair: this CTA's loads complete on the leader's barrier of the same stage
      auto load = [&](uint8_t* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
        if (CN == 1) tma_load_2d(dst, map, c0, c1, bar);
        else tma_load_2d_pair(dst, map, c0, c1, bar);
      };
      const int nb = n0 + (int)crank * (BN / CN);       // this CTA's weight rows of the tile
      // weights of the first ring fill do not depend on the previous kernel
      const int n_pre = n_kb < n_stages ? n_kb : n_stages;
      if (W8) {
        // bytes of the e4m3 tile land on fullb (the converters' barrier), activations on full
        constexpr int kB8Off = C::kABytes + C::kBBytes;
        for (int kb = 0; kb < n_pre; ++kb) {
          mbar_arrive_expect_tx(&fullb[kb], C::kB8Bytes);
          tma_load_2d(smem + kb * C::kStageBytes + kB8Off, &tma_b, kc(kb) * BK, nb, &fullb[kb]);
        }
        pdl_wait();
        PROF(2);
        for (int kb = 0; kb < n_pre; ++kb) {
          mbar_arrive_expect_tx(&full[kb], C::kABytes);
          tma_load_2d(smem + kb * C::kStageBytes, &tma_a, kc(kb) * BK, m0, &full[kb]);
        }
        for (int kb = n_pre; kb < n_kb; ++kb) {
          const int s = kb % n_stages, ph = (kb / n_stages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* sa = smem + s * C::kStageBytes;
          mbar_arrive_expect_tx(&fullb[s], C::kB8Bytes);
          tma_load_2d(sa + kB8Off, &tma_b, kc(kb) * BK, nb, &fullb[s]);
          mbar_arrive_expect_tx(&full[s], C::kABytes);
          tma_load_2d(sa, &tma_a, kc(kb) * BK, m0, &full[s]);
        }
      } else {
      for (int kb = 0; kb < n_pre; ++kb) {
        if (crank == 0) mbar_arrive_expect_tx(&full[kb], CN * C::kStageBytes);
        load(smem + kb * C::kStageBytes + C::kABytes, &tma_b, kc(kb) * BK, nb, &full[kb]);
      }
      pdl_wait();
      PROF(2);
      for (int kb = 0; kb < n_pre; ++kb)
        load(smem + kb * C::kStageBytes, &tma_a, kc(kb) * BK, m0, &full[kb]);
      for (int kb = n_pre; kb < n_kb; ++kb) {
        const int s = kb % n_stages, ph = (kb / n_stages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        if (crank == 0) mbar_arrive_expect_tx(&full[s], CN * C::kStageBytes);
        uint8_t* sa = smem + s * C::kStageBytes;
        load(sa, &tma_a, kc(kb) * BK, m0, &full[s]);
        load(sa + C::kABytes, &tma_b, kc(kb) * BK, nb, &full[s]);
      }
      }
    } else {
      pdl_wait();
    }
  } else if (warp == 1) {
    pdl_wait();
    if (lane == 0 && crank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CN, BN);
      for (int kb = 0; kb < n_kb; ++kb) {
        const int s = kb % n_stages, ph = (kb / n_stages) & 1;
        mbar_wait(&full[s], ph);
        if (W8) mbar_wait(&conv[s], ph);
        if (kb == 0) PROF(3);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * C::kStageBytes);
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + C::kABytes);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {         // +32 bytes per K=16 step inside the swizzle atom
          if (CN == 1) umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          else umma_bf16_pair(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
        }
        if (CN == 1) umma_commit(&empty[s]);
        else umma_commit_pair(&empty[s], kMask);    // frees the stage in both CTAs
      }
      if (CN == 1) umma_commit(tmem_full);
      else umma_commit_pair(tmem_full, kMask);
      PROF(4);
    }
  } else if (warp < 4) {
    pdl_wait();
  } else if (W8) {
    // ---- converters (warps 4-11): e4m3 staging tile -> bf16 operand tile in the SWIZZLE_128B layout.
    // Four groups of two warps; group g owns k-blocks kb = g (mod 4), so four stages are widened
    // concurrently (one warp set walking every stage in turn is LATENCY-bound: wait -> LDS -> 14-op chain ->
    // STS -> fence -> arrive measured ~1200 cycles per stage against 256 for its MMAs).  Thread t of a group
    // takes 16-byte chunk c = i*64 + t of the tile: weight row c >> 2, bytes (c & 3)*16..+16 of its
    // 64-element k-block, i.e. bf16 chunks 2*(c & 3) and 2*(c & 3) + 1 of the 128-byte operand row,
    // stored at chunk ^ (row & 7) (TMA's 128-byte swizzle; stage bases are 1024-aligned).
    const int grp = (warp - 4) >> 1, t = threadIdx.x - 128 - grp * 64;
    constexpr int kB8Off = C::kABytes + C::kBBytes;
    const uint32_t smem0 = smem_u32(smem);
    for (int kb = grp; kb < n_kb; kb += 4) {
      const int s = kb % n_stages, ph = (kb / n_stages) & 1;
      if (kb >= n_stages) mbar_wait(&empty[s], ph ^ 1);   // the MMAs that read this operand tile are done
      mbar_wait(&fullb[s], ph);
      const uint32_t st = smem0 + s * C::kStageBytes;
#pragma unroll
      for (int i = 0; i < BN * 4 / 64; ++i) {
        const int c = i * 64 + t, row = c >> 2, ch = (c & 3) * 2;
        uint4 lo8, hi8;
        e4m3x16_to_bf16(lds128(st + kB8Off + c * 16), lo8, hi8);
        const uint32_t rowp = st + C::kABytes + row * 128;
        sts128(rowp + ((ch ^ (row & 7)) << 4), lo8);
        sts128(rowp + (((ch + 1) ^ (row & 7)) << 4), hi8);
      }
      fence_proxy_async();                              // generic-proxy writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) mbar_arrive(&conv[s]);
    }
  }
  // Epilogue.  A warp may only read the TMEM lane quarter (warp % 4); with one warp per quarter the
  // epilogue is a single warp per SM sub-partition issuing dependent math (measured: SiLU epilogue
  // +3.5 us at T = 256).  For the column-chunked epilogues the four role warps, idle by now, take every
  // other 32-column chunk of their quarter; for the RoPE / KV-append epilogue (one 64-column head at a time) every
  // other head, when the tile holds more than one (the prefill shapes: BN = 128 / 256).
  constexpr bool kAllWarps = EPI != 2 || BN > 64;
  if (warp < 8 && (warp >= 4 || kAllWarps)) {
    __syncwarp();
    const int ew = warp & 3;                           // TMEM lane quarter
    if (warp >= 4) pdl_wait();
    // EPI 2: this row's position / KV slot and its cos|sin line do not depend on the accumulator — fetch them
    // while the MMAs run (pos -> cos/sin would otherwise be two dependent L2 round trips after the last MMA)
    int pf_pos = 0;
    int64_t pf_slot = -1;
    if (EPI == 2 && m0 + ew * 32 + lane < M) {
      pf_pos = ea.pos[m0 + ew * 32 + lane];
      pf_slot = ea.slot[m0 + ew * 32 + lane];
      asm volatile("prefetch.global.L1 [%0];" ::"l"(ea.cos_sin + (size_t)pf_pos * 64));
    }
    mbar_wait(tmem_full, 0);
    if (threadIdx.x == 128) PROF(5);
    tc_fence_after();
    const int row = m0 + ew * 32 + lane;
    if (EPI == 2) {
      // ---- RoPE + q store + KV append: each 64-wide half of the tile is head (col/64) % 16 of q (0), k (1) or v (2)
#pragma unroll 1
      for (int hh = (BN > 64 && warp < 4) ? 1 : 0; hh < BN / 64; hh += (BN > 64 ? 2 : 1)) {
      uint32_t r0[32], r1[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + hh * 64, r0);
      tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + hh * 64 + 32, r1);
      tmem_ld_wait();
      if (W8) { scale32(r0, wscale + n0 + hh * 64); scale32(r1, wscale + n0 + hh * 64 + 32); }
      if (row < M) {
        const int head48 = (n0 >> 6) + hh, which = head48 >> 4, h = head48 & 15;
        uint32_t o1[16], o2[16];
        if (which < 2) {
          const int p = pf_pos;
          const uint4* cs = reinterpret_cast<const uint4*>(ea.cos_sin + (size_t)p * 64);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint4 cv = cs[c], sv = cs[4 + c];
            const uint32_t cw[4] = {cv.x, cv.y, cv.z, cv.w}, sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int e = c * 8 + i * 2;
              const float x1l = round_bf16(__uint_as_float(r0[e])), x1h = round_bf16(__uint_as_float(r0[e + 1]));
              const float x2l = round_bf16(__uint_as_float(r1[e])), x2h = round_bf16(__uint_as_float(r1[e + 1]));
              const float cl = bf_lo(cw[i]), ch = bf_hi(cw[i]), sl = bf_lo(sw[i]), sh = bf_hi(sw[i]);
              o1[c * 4 + i] = pack_bf16(x1l * cl - x2l * sl, x1h * ch - x2h * sh);
              o2[c * 4 + i] = pack_bf16(x2l * cl + x1l * sl, x2h * ch + x1h * sh);
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            o1[i] = pack_bf16(__uint_as_float(r0[2 * i]), __uint_as_float(r0[2 * i + 1]));
            o2[i] = pack_bf16(__uint_as_float(r1[2 * i]), __uint_as_float(r1[2 * i + 1]));
          }
        }
        if (which == 0) {
          uint4* dst = reinterpret_cast<uint4*>(y + (size_t)row * ldy + h * 64);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            dst[c] = make_uint4(o1[4 * c], o1[4 * c + 1], o1[4 * c + 2], o1[4 * c + 3]);
            dst[4 + c] = make_uint4(o2[4 * c], o2[4 * c + 1], o2[4 * c + 2], o2[4 * c + 3]);
          }
        } else {
          const int64_t sl = pf_slot;
          if (sl >= 0) {
            uint8_t* page = ea.kv_layer + (size_t)(sl / CBX_BLOCK_SIZE) * CBX_PAGE_BYTES;
            const int tok = (int)(sl % CBX_BLOCK_SIZE);
            const int sw = which == 1 ? (tok & 7) : 0;   // K chunks are XOR-swizzled, V chunks are not
            uint4* dst = reinterpret_cast<uint4*>(page + (which == 2 ? 32768 : 0) + h * 2048 + tok * 128);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              dst[c ^ sw] = make_uint4(o1[4 * c], o1[4 * c + 1], o1[4 * c + 2], o1[4 * c + 3]);
              dst[(4 + c) ^ sw] = make_uint4(o2[4 * c], o2[4 * c + 1], o2[4 * c + 2], o2[4 * c + 3]);
            }
          }
        }
      }
      }
    } else {
#pragma unroll 1
    for (int cb = warp < 4 ? 1 : 0; cb < BN / 32; cb += 2) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + cb * 32, r);
      tmem_ld_wait();
      const int col = n0 + cb * 32;
      if (W8 && col < N) scale32(r, wscale + col);        // wscale is padded to a multiple of 128 entries
      if (row < M && col < N) {
        if (EPI == 1) {
          // packed gate/up: columns [0,16) gate, [16,32) up of the same 16 features
          uint32_t o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float g0 = round_bf16(__uint_as_float(r[2 * i])), g1 = round_bf16(__uint_as_float(r[2 * i + 1]));
            const float u0 = round_bf16(__uint_as_float(r[16 + 2 * i])), u1 = round_bf16(__uint_as_float(r[17 + 2 * i]));
            o[i] = pack_bf16(__fdividef(g0, 1.f + __expf(-g0)) * u0, __fdividef(g1, 1.f + __expf(-g1)) * u1);
          }
          if (col + 32 <= N) {
            uint4* dst = reinterpret_cast<uint4*>(y + (size_t)row * ldy + col / 2);
            dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
            dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
          }
        } else if (EPI == 4 || EPI == 5) {
          // split-K partial: fp32 [gridDim.z][M][N] behind y (N % 32 == 0); summed by cbx_rmsnorm_reduce
          float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(y) + ((size_t)blockIdx.z * M + row) * N + col);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            dst[i] = make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]), __uint_as_float(r[4 * i + 2]),
                                 __uint_as_float(r[4 * i + 3]));
        } else if (EPI == 3) {
          // residual stream in place: res = bf16(bf16(acc) + res)   (N % 32 == 0)
          uint4* dst = reinterpret_cast<uint4*>(y + (size_t)row * ldy + col);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint4 q = dst[i];
            const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
            uint32_t ow[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              ow[j] = pack_bf16(round_bf16(__uint_as_float(r[8 * i + 2 * j])) + bf_lo(qw[j]),
                                round_bf16(__uint_as_float(r[8 * i + 2 * j + 1])) + bf_hi(qw[j]));
            dst[i] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
          }
        } else {
          if (col + 32 <= N) {
            uint4* dst = reinterpret_cast<uint4*>(y + (size_t)row * ldy + col);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              dst[i] = make_uint4(pack_bf16(__uint_as_float(r[8 * i]), __uint_as_float(r[8 * i + 1])),
                                  pack_bf16(__uint_as_float(r[8 * i + 2]), __uint_as_float(r[8 * i + 3])),
                                  pack_bf16(__uint_as_float(r[8 * i + 4]), __uint_as_float(r[8 * i + 5])),
                                  pack_bf16(__uint_as_float(r[8 * i + 6]), __uint_as_float(r[8 * i + 7])));
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col + i < N) y[(size_t)row * ldy + col + i] = __float2bfloat16_rn(__uint_as_float(r[i]));
          }
        }
      }
    }
    }
    tc_fence_before();
  }
  if (EPI == 5) __threadfence();                       // this thread's partial stores, before the CTA announces them
  __syncthreads();
  if (EPI == 5) {
    // ---- the split-K fold, by the CTAs that produced the slices.  Every CTA of a 128-row tile (column tiles x K
    // slices: 64 or 32 CTAs) announces its partials, waits until all of them have, and then adds + normalises its share
    // of the tile's rows — what a separate add+RMSNorm launch did (60 per M1 step, each a wait release + a row fetch
    // + a tail behind the GEMM's).  The host launches this form only when the whole grid is co-resident (<= one CTA
    // per SM), so the wait cannot deadlock; CTAs of the NEXT kernel cannot be ahead of ours on an SM either (they are
    // launched once every CTA of this grid has executed griddepcontrol.launch_dependents, i.e. runs).
    __shared__ float f_sums[2][8];
    const int n_cta = gridDim.x * gridDim.z, ci = blockIdx.x + gridDim.x * blockIdx.z;
    unsigned* arrive = ea.f_cnt + 2 * blockIdx.y;
    unsigned* done = arrive + 1;
    if (threadIdx.x == 0) {
      atomicAdd(arrive, 1u);
      unsigned v;
      do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(arrive) : "memory"); } while (v < (unsigned)n_cta);
    }
    __syncthreads();
    const float4* part = reinterpret_cast<const float4*>(y);
    for (int base = 2 * ci; base < BM; base += 2 * n_cta) {
      if (gridDim.z == 4) fold_two_rows<4>(part, M, m0 + base, ea, f_sums);
      else fold_two_rows<2>(part, M, m0 + base, ea, f_sums);
    }
    if (threadIdx.x == 0) {                            // the last CTA of the tile to finish hands the counters back empty
      const unsigned d = atomicAdd(done, 1u);
      if (d == (unsigned)n_cta - 1) { *arrive = 0u; *done = 0u; }
    }
  }
  if (threadIdx.x == 0) PROF(6);
  // nobody leaves while the leader's MMAs may still read this CTA's shared memory / TMEM.  An EXECUTION barrier only: no data
  // passes between the CTAs here, so the arrive is .relaxed — the default .release is a MEMBAR.ALL.GPU that made every CTA
  // wait for its epilogue's global stores to be acknowledged before it could leave (20 % of the QKV prefill GEMM's stall
  // samples, profiles/r02_ncu_prefill_qkv_stalls.txt)
  if (CN > 1)
    asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    if (CN == 1) tmem_dealloc(tmem_base, BN);
    else tmem_dealloc_pair(tmem_base, BN);
  }
}

// ---- host side: tensor maps through the driver entry point (no -lcuda link) -------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D row-major bf16 [rows, cols] tensor, box = {64 cols, box_rows}, 128-byte swizzle.
int make_map(CUtensorMap* m, const void* ptr, int rows, int cols, int box_rows) {
  typedef std::tuple<const void*, int, int, int> Key;
  static std::map<Key, CUtensorMap> cache;
  Key key(ptr, rows, cols, box_rows);
  auto itr = cache.find(key);
  if (itr != cache.end()) { *m = itr->second; return 0; }
  EncodeTiledFn enc = get_encode();
  CBX_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CBX_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d", (int)r, rows, cols);
  if (cache.size() > 4096) cache.clear();
  cache[key] = *m;
  return 0;
}

// 2-D row-major e4m3 [rows, cols] weight tensor as bytes, box = {64 bytes, box_rows}, no swizzle (the
// converter warps re-lay it out as bf16).
int make_map_u8(CUtensorMap* m, const void* ptr, int rows, int cols, int box_rows) {
  typedef std::tuple<const void*, int, int, int> Key;
  static std::map<Key, CUtensorMap> cache;
  Key key(ptr, rows, cols, box_rows);
  auto itr = cache.find(key);
  if (itr != cache.end()) { *m = itr->second; return 0; }
  EncodeTiledFn enc = get_encode();
  CBX_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CBX_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (u8) failed (%d) rows=%d cols=%d", (int)r, rows, cols);
  if (cache.size() > 4096) cache.clear();
  cache[key] = *m;
  return 0;
}

template <int BN, int EPI>
int set_attr8() {
  cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<BN, 1, true>::kSmem);
  CBX_REQUIRE(e == cudaSuccess, "gemm (fp8w) smem attribute: %s", cudaGetErrorString(e));
  return 0;
}
template <int BN, int EPI, int CN>
int set_attr() {
  cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, CN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN, CN>::kSmem);
  CBX_REQUIRE(e == cudaSuccess, "gemm smem attribute: %s", cudaGetErrorString(e));
  return 0;
}
template <int EPI>
int set_attr_pairs() { return set_attr<64, EPI, 2>() || set_attr<128, EPI, 2>() || set_attr<256, EPI, 2>(); }
}  // namespace
namespace cbx {
// Tensor map over one layer's KV pages seen as bf16 rows of 64 elements (512 rows per 64 KiB page: K
// rows then V rows), box = 256 rows, SWIZZLE_128B: the prefill attention kernel fetches a page's V
// half through it so that V lands in shared memory chunk-swizzled like K (conflict-free ldmatrix.trans).
int make_kv_rows_map(CUtensorMap* m, const void* kv_layer) {
  return make_map(m, kv_layer, 1 << 30, 64, 256);
}
int init_gemm() {
  static bool done = false;
  if (done) return 0;
  if (set_attr<64, 0, 1>() || set_attr<64, 1, 1>() || set_attr<128, 0, 1>() || set_attr<128, 1, 1>() || set_attr<64, 2, 1>() ||
      set_attr<64, 3, 1>() || set_attr<128, 3, 1>() || set_attr<64, 4, 1>() || set_attr<128, 2, 1>()) return -1;
#ifdef CBX_EXPERIMENTS
  if (set_attr<64, 5, 1>()) return -1;
#endif
  if (set_attr_pairs<0>() || set_attr_pairs<1>() || set_attr_pairs<2>() || set_attr_pairs<3>() || set_attr_pairs<4>()) return -1;
  if (set_attr8<64, 0>() || set_attr8<64, 1>() || set_attr8<64, 2>() || set_attr8<64, 3>() || set_attr8<64, 4>() ||
      set_attr8<128, 0>() || set_attr8<128, 1>() || set_attr8<128, 2>() || set_attr8<128, 3>()) return -1;
  CBX_REQUIRE(get_encode() != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  done = true;
  return 0;
}
}  // namespace cbx
namespace {
// FP8-weight form: W is [N, K] e4m3 bytes, wscale [>= round_up(N, 128)] fp32.  Deep ring, one CTA per SM.
template <int BN, int EPI>
int launch8(const void* x, const void* W, const float* wscale, void* y, int M, int N, int K, int ldy, cudaStream_t s,
            const EpiArgs& ea = EpiArgs(), int splits = 1) {
  using C = Cfg<BN, 1, true>;
  if (cbx::init_gemm()) return -1;
  const int n_tiles = (N + BN - 1) / BN, m_tiles = (M + BM - 1) / BM;
  CUtensorMap ma, mb;
  if (make_map(&ma, x, M, K, BM)) return -1;
  if (make_map_u8(&mb, W, N, K, BN)) return -1;
  cudaError_t e = launch_pdl(gemm_tc_kernel<BN, EPI, 1, true>, dim3(n_tiles, m_tiles, splits), dim3(kThreads8), C::kSmem, s, ma, mb,
                             (bf16*)y, M, N, K, ldy, ea, (int)C::kStages, wscale);
  CBX_REQUIRE(e == cudaSuccess, "gemm_tc (fp8w) launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
template <int BN, int EPI>
int launch(const void* x, const void* W, void* y, int M, int N, int K, int ldy, cudaStream_t s,
           const EpiArgs& ea = EpiArgs(), int splits = 1) {
  using C = Cfg<BN>;
  if (cbx::init_gemm()) return -1;
  const int n_tiles = (N + BN - 1) / BN, m_tiles = (M + BM - 1) / BM;
  CUtensorMap ma, mb;
  if (make_map(&ma, x, M, K, BM)) return -1;
  if (make_map(&mb, W, N, K, BN)) return -1;
  static const int shallow_env = getenv("CBX_GEMM_SHALLOW") ? atoi(getenv("CBX_GEMM_SHALLOW")) : 1;
  // shallow_env 2 (tuning aid): the shallow ring for every grid, so that the CTAs of consecutive decode GEMMs share SMs
  const bool shallow = shallow_env == 2 || (shallow_env != 0 && (int64_t)n_tiles * m_tiles * splits >= 2 * sm_count());
  const int n_stages = shallow ? C::kStagesShallow : C::kStages;
  const int smem_bytes = shallow ? C::kSmemShallow : C::kSmem;
#ifdef CBX_GEMM_PROF
  EpiArgs eap = ea; eap.prof_id = g_prof_next++;
  fprintf(stderr, "gemm_prof id=%d BN=%d EPI=%d M=%d N=%d K=%d grid=%dx%dx%d\n", eap.prof_id, BN, EPI, M, N, K, n_tiles, m_tiles, splits);
  const EpiArgs& eal = eap;
#else
  const EpiArgs& eal = ea;
#endif
  cudaError_t e = launch_pdl(gemm_tc_kernel<BN, EPI, 1>, dim3(n_tiles, m_tiles, splits), dim3(kThreads), smem_bytes, s, ma, mb,
                             (bf16*)y, M, N, K, ldy, eal, n_stages, (const float*)nullptr);
  CBX_REQUIRE(e == cudaSuccess, "gemm_tc launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
// CTA-pair form: 256 x BN tile per cluster of two CTAs along M (grid.x padded to even: rows beyond M load
// zeros and store nothing).  (An earlier experiment multicast the activation tile across two column
// tiles instead: it halved the L2 reads of A but not what each SM ingests, and measured 2-5 % slower.)
template <int BN, int EPI>
int launch_pair(const void* x, const void* W, void* y, int M, int N, int K, int ldy, cudaStream_t s,
                const EpiArgs& ea = EpiArgs(), int splits = 1) {
  using C = Cfg<BN, 2>;
  if (cbx::init_gemm()) return -1;
  const int n_tiles = (N + BN - 1) / BN, m_tiles = ((M + BM - 1) / BM + 1) & ~1;
  CUtensorMap ma, mb;
  if (make_map(&ma, x, M, K, BM)) return -1;
  if (make_map(&mb, W, N, K, BN / 2)) return -1;
  static const int shallow_env = getenv("CBX_GEMM_SHALLOW") ? atoi(getenv("CBX_GEMM_SHALLOW")) : 1;
  const bool shallow = shallow_env != 0 && (int64_t)n_tiles * m_tiles * splits >= 2 * sm_count();
  const int n_stages = shallow ? C::kStagesShallow : C::kStages;
  const int smem_bytes = shallow ? C::kSmemShallow : C::kSmem;
  cudaError_t e = launch_pdl_cluster(gemm_tc_kernel<BN, EPI, 2>, dim3(m_tiles, n_tiles, splits), dim3(kThreads), smem_bytes, s, 2,
                                     ma, mb, (bf16*)y, M, N, K, ldy, ea, n_stages, (const float*)nullptr);
  CBX_REQUIRE(e == cudaSuccess, "gemm_tc (pair) launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
// Pair tile width chosen for a problem: 0 = single-CTA tiles.  env CBX_GEMM_PAIR = 0 | 64 | 128 | 256 forces it.
// Measured on B200 (tools/bench_gemm.py, bench_gemm_decode.py): at M = 16000 the 256-wide pair runs gate_up+SiLU
// at 1235 TFLOP/s (single-CTA 128-wide tiles: 1037; cuBLAS plain GEMM of that shape: 1221), down 1116 (997),
// QKV+RoPE 831 (793); at M = 256 every pair shape is SLOWER than single-CTA tiles (gate_up 9.4 vs 8.6 us, QKV
// 8.2 vs 7.4): the cluster launch + two cluster barriers cost more than the saved ingest.  So: pairs from
// `min_rows` rows up.
int pair_width(int M, int min_rows) {
  static const int env = getenv("CBX_GEMM_PAIR") ? atoi(getenv("CBX_GEMM_PAIR")) : -1;
  if (M <= BM) return 0;                               // a lone row tile has no partner
  if (env >= 0) return env;
  return M >= min_rows ? 256 : 0;
}
template <int EPI>
int launch_pair_w(int w, const void* x, const void* W, void* y, int M, int N, int K, int ldy, cudaStream_t s,
                  const EpiArgs& ea = EpiArgs(), int splits = 1) {
  if (w == 64) return launch_pair<64, EPI>(x, W, y, M, N, K, ldy, s, ea, splits);
  if (w == 128) return launch_pair<128, EPI>(x, W, y, M, N, K, ldy, s, ea, splits);
  return launch_pair<256, EPI>(x, W, y, M, N, K, ldy, s, ea, splits);
}
}  // namespace

namespace cbx {
int gemm_bf16_tc(const void* x, const void* W, void* y, int M, int N, int K, int ldy, int epilogue,
                 cudaStream_t s, const float* wscale) {
  CBX_REQUIRE(M >= 1 && K % BK == 0 && K >= BK, "gemm_tc: bad shape M=%d K=%d", M, K);
  CBX_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0,
              "gemm_tc: operands must be 16-byte aligned");
  CBX_REQUIRE(epilogue == 0 || N % 32 == 0, "gemm_tc: SiLU epilogue needs N %% 32 == 0");
  // Narrow tiles when the grid would leave most of the 148 SMs idle.
  const int m_tiles = (M + BM - 1) / BM;
  // 64-wide tiles only when 128-wide ones would leave a third of the SMs idle: the GEMMs of this
  // path are bound by L2->SM bytes, which a 128 x 128 tile halves per output element.
  bool narrow = (int64_t)m_tiles * ((N + 127) / 128) < 100;
  if (const char* v = getenv("CBX_GEMM_BN")) narrow = atoi(v) == 64;      // tuning aid
  if (wscale) {
    if (epilogue == 1) return narrow ? launch8<64, 1>(x, W, wscale, y, M, N, K, ldy, s) : launch8<128, 1>(x, W, wscale, y, M, N, K, ldy, s);
    return narrow ? launch8<64, 0>(x, W, wscale, y, M, N, K, ldy, s) : launch8<128, 0>(x, W, wscale, y, M, N, K, ldy, s);
  }
  if (const int w = pair_width(M, 4096))
    return epilogue == 1 ? launch_pair_w<1>(w, x, W, y, M, N, K, ldy, s) : launch_pair_w<0>(w, x, W, y, M, N, K, ldy, s);
  if (epilogue == 1) return narrow ? launch<64, 1>(x, W, y, M, N, K, ldy, s) : launch<128, 1>(x, W, y, M, N, K, ldy, s);
  return narrow ? launch<64, 0>(x, W, y, M, N, K, ldy, s) : launch<128, 0>(x, W, y, M, N, K, ldy, s);
}
// EPI 2: y = qkv buffer [M, 3072]; q rotated into its first 1024 columns, k/v into the KV pages.
int gemm_qkv_rope(const void* h, const void* Wqkv, void* qkv, const void* cos_sin, const int32_t* pos,
                  const int64_t* slot, void* kv_layer, int M, cudaStream_t s, const float* wscale) {
  EpiArgs ea = {(const bf16*)cos_sin, pos, slot, (uint8_t*)kv_layer};
  if (wscale) {
    if ((M + BM - 1) / BM * 24 >= 2 * sm_count()) return launch8<128, 2>(h, Wqkv, wscale, qkv, M, 3072, 1024, 3072, s, ea);
    return launch8<64, 2>(h, Wqkv, wscale, qkv, M, 3072, 1024, 3072, s, ea);
  }
  if (const int w = pair_width(M, 8192)) return launch_pair_w<2>(w, h, Wqkv, qkv, M, 3072, 1024, 3072, s, ea);
  // prefill-sized M: 128-wide tiles (two heads per tile) halve the activation re-reads
  if ((M + BM - 1) / BM * 24 >= 2 * sm_count()) return launch<128, 2>(h, Wqkv, qkv, M, 3072, 1024, 3072, s, ea);
  return launch<64, 2>(h, Wqkv, qkv, M, 3072, 1024, 3072, s, ea);
}
// EPI 3: res[M, N] = bf16(bf16(x·W^T) + res), N % 32 == 0.
int gemm_residual(const void* x, const void* W, void* res, int M, int N, int K, cudaStream_t s, const float* wscale) {
  CBX_REQUIRE(N % 128 == 0 && K % BK == 0, "gemm_residual: bad shape N=%d K=%d", N, K);
  if (wscale) {
    const bool narrow8 = (int64_t)((M + BM - 1) / BM) * (N / 128) < sm_count();
    return narrow8 ? launch8<64, 3>(x, W, wscale, res, M, N, K, N, s) : launch8<128, 3>(x, W, wscale, res, M, N, K, N, s);
  }
  if (const int w = pair_width(M, 4096)) return launch_pair_w<3>(w, x, W, res, M, N, K, N, s);
  const bool narrow = (int64_t)((M + BM - 1) / BM) * (N / 128) < sm_count();
  return narrow ? launch<64, 3>(x, W, res, M, N, K, N, s) : launch<128, 3>(x, W, res, M, N, K, N, s);
}
// EPI 4: part[splits][M][N] fp32 = per-K-slice partial products (no rounding); N % 64 == 0.
int gemm_partials(const void* x, const void* W, void* part, int M, int N, int K, int splits, cudaStream_t s,
                  const float* wscale) {
  CBX_REQUIRE(N % 64 == 0 && splits >= 1 && K % (BK * splits) == 0, "gemm_partials: bad shape N=%d K=%d splits=%d", N, K, splits);
  if (wscale) return launch8<64, 4>(x, W, wscale, part, M, N, K, N, s, EpiArgs(), splits);
  if (const int w = pair_width(M, 4096)) return launch_pair_w<4>(w, x, W, part, M, N, K, N, s, EpiArgs(), splits);
  return launch<64, 4>(x, W, part, M, N, K, N, s, EpiArgs(), splits);
}
#ifdef CBX_EXPERIMENTS   // `make EXPERIMENTS=1` + env CBX_GEMM_FOLD=1.  Bit-identical to the two-kernel form (GPU test) and
// SLOWER: M1 3.462 -> 3.526 ms per step (154 launches instead of 214).  A programmatic-dependent-launch boundary costs
// less than the arrival counter + spin + fold serialised behind the GEMM's epilogue: the separate kernel's launch, wait
// and weight fetch overlap the GEMM's tail, the in-kernel fold cannot start before the slowest CTA of the tile.
// EPI 5: gemm_partials + the add+RMSNorm that folds them, in one launch: res[M,1024] = bf16(bf16(x·W^T) + res) summed
// over `splits` K-slices in slice order, h_out = RMSNorm(res)·ln_w (the bits of gemm_partials + cbx_rmsnorm_reduce).
// counters: 2 x (row tiles) uint32, zero on entry (left zero).  Returns 1 without launching when the form does not
// apply (the caller then runs the two kernels): N != 1024, splits not 2 or 4, or a grid that is not co-resident.
int gemm_partials_fold(const void* x, const void* W, void* part, int M, int K, int splits, cudaStream_t s, void* res,
                       const void* ln_w, void* h_out, float eps, void* counters) {
  const int m_tiles = (M + BM - 1) / BM;
  if ((splits != 2 && splits != 4) || K % (BK * splits) != 0 || pair_width(M, 4096) != 0) return 1;
  if (16 * m_tiles * splits > sm_count() || counters == nullptr) return 1;
  EpiArgs ea = {};
  ea.f_res = (bf16*)res; ea.f_ln = (const bf16*)ln_w; ea.f_out = (bf16*)h_out; ea.f_cnt = (unsigned*)counters; ea.f_eps = eps;
  return launch<64, 5>(x, W, part, M, 1024, K, 1024, s, ea, splits);
}
#else
int gemm_partials_fold(const void*, const void*, void*, int, int, int, cudaStream_t, void*, const void*, void*, float, void*) { return 1; }
#endif
int gemv_bf16(const void*, const void*, void*, int, int, int, int, int, cudaStream_t, const float* wscale = nullptr);
}  // namespace cbx

extern "C" int cbx_linear(const void* x, const void* W, void* y, int T, int N, int K, int ldy, int epilogue,
                          int path, void* stream) {
  CBX_REQUIRE(T >= 1, "cbx_linear: T < 1");
  CBX_REQUIRE(path >= 0 && path <= 2, "cbx_linear: bad path %d", path);
  const bool gemv_ok = T <= 4 && (K == 1024 || (K == 4096 && epilogue == 0));
  if (path == 1) CBX_REQUIRE(gemv_ok, "cbx_linear: GEMV path needs T <= 4 and K in {1024,4096}");
  if (path == 1 || (path == 0 && gemv_ok)) return cbx::gemv_bf16(x, W, y, T, N, K, ldy, epilogue, (cudaStream_t)stream);
  return cbx::gemm_bf16_tc(x, W, y, T, N, K, ldy, epilogue, (cudaStream_t)stream, nullptr);
}

// FP8-e4m3 weights W8 [N, K] + fp32 scale per output feature (wscale holds >= round_up(N, 128) entries).
extern "C" int cbx_linear_fp8w(const void* x, const void* W8, const float* wscale, void* y, int T, int N, int K, int ldy,
                               int epilogue, int path, void* stream) {
  CBX_REQUIRE(T >= 1 && wscale != nullptr, "cbx_linear_fp8w: T < 1 or NULL scale");
  CBX_REQUIRE(path >= 0 && path <= 2, "cbx_linear_fp8w: bad path %d", path);
  const bool gemv_ok = T <= 4 && (K == 1024 || (K == 4096 && epilogue == 0));
  if (path == 1) CBX_REQUIRE(gemv_ok, "cbx_linear_fp8w: GEMV path needs T <= 4 and K in {1024,4096}");
  if (path == 1 || (path == 0 && gemv_ok))
    return cbx::gemv_bf16(x, W8, y, T, N, K, ldy, epilogue, (cudaStream_t)stream, wscale);
  return cbx::gemm_bf16_tc(x, W8, y, T, N, K, ldy, epilogue, (cudaStream_t)stream, wscale);
}

#ifdef CBX_GEMM_PROF
// debug builds only: copy the stamps of launches [0, n) to host (n * 160 * 8 u64) and restart the launch ids
extern "C" int cbx_gemm_prof_read(unsigned long long* out, int n) {
  cudaDeviceSynchronize();
  cudaError_t e = cudaMemcpyFromSymbol(out, g_prof, (size_t)n * kProfCtas * 8 * sizeof(unsigned long long));
  const int used = g_prof_next;
  g_prof_next = 0;
  return e == cudaSuccess ? used : -1;
}
#endif
