// k_attn.cu — paged-KV decode attention (one query row per entry), sm_100a.
//
// Replaces vLLM's flash_attn_varlen_func call over the paged cache
// (vllm v1/attention/backends/flash_attn.py:796-800) for the decode shape: nkv = nh = 16,
// d = 64, so there is no GQA reuse and the op is a pure HBM stream (0.5 FMA per byte) —
// tensor cores would not help; the design goal is to keep the TMA engine saturated.
//
// Persistent kernel, one CTA per SM, 17 warps:
//   warp 16   producer: walks this CTA's work items (row, split) and issues ONE 64 KB
//             cp.async.bulk (UBLKCP) per KV page into a 3-stage shared-memory ring,
//             running ahead across item boundaries so the HBM stream never drains;
//   warps 0-15 consumers, warp = head: QK (lane = token x half-of-d, 128-bit swizzled
//             LDS), warp-shuffle online softmax in the log2 domain, PV (lane = token-group
//             x 16-byte chunk), mbarrier hand-back of the stage.
// Partial (o, m, l) per (row, split) go to scratch; the CTA that publishes a row's LAST partial
// (atomic ticket per row) merges them in the same launch — no second kernel.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"

using namespace cbx;
namespace cbx { int init_attn(); int make_kv_rows_map(CUtensorMap* m, const void* kv_layer); }

namespace {
constexpr int kStages = 3;
constexpr int kConsumerWarps = 16;
constexpr int kThreads = (kConsumerWarps + 1) * 32;
constexpr int kSmemBytes = kStages * CBX_PAGE_BYTES + 64;
constexpr float kLog2e = 1.4426950408889634f;

// balanced variant: 2 partial slots per CTA (o fp32 [1024], (m, l) per head), then the range plan
constexpr int kPlanMaxRows = 4095;              // 1024 threads x 4 rows, and entry [T] for the total
constexpr int kPlanFirst = 2 + 2 * kMaxSms;         // [kPlanFirst] = 1 when the first-page ids below are filled in
constexpr int kPrimed = 3;                      // = kStages of the balanced kernel: the first ring fill
constexpr int kPlanInts = kPlanFirst + 1 + kPrimed * kMaxSms;
constexpr int64_t kBalScratchBytes = 2 * kMaxSms * (CBX_HIDDEN * 4 + CBX_HEADS * 8) + kPlanInts * 4 + 256;

struct AttnPlan { int pps; int nsplit; };
inline AttnPlan make_plan(int T, int max_kv_len) {
  int max_pages = (max_kv_len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  if (max_pages < 1) max_pages = 1;
  int want = (sm_count() * 6 + T - 1) / T;                 // splits per row to get ~6 items per SM
  if (want < 1) want = 1;
  if (want > 32) want = 32;
  if (want > max_pages) want = max_pages;
  AttnPlan p;
  p.pps = (max_pages + want - 1) / want;
  p.nsplit = (max_pages + p.pps - 1) / p.pps;
  return p;
}

// kTicket: every CTA owns at most one (row, split) item, so the CTA that publishes a row's last
// partial merges the row in place (small batches: saves a launch on the latency-critical path).
// Otherwise partials are merged by attn_merge_kernel (large batches: keeps the page stream busy).
template <bool kTicket>
__global__ void __launch_bounds__(kThreads, 1) paged_attn_decode_kernel(
    const bf16* qkv, const uint8_t* __restrict__ kv_layer, const int32_t* bt,
    int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
    bf16* __restrict__ out, float* __restrict__ o_part, float2* __restrict__ ml_part,
    int32_t* __restrict__ counters, int T, int pps, int nsplit) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * CBX_PAGE_BYTES);
  uint64_t* empty = full + kStages;
  volatile int& s_last = *reinterpret_cast<volatile int*>(empty + kStages);   // after the 6 barriers
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  pdl_trigger();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kConsumerWarps); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();                                         // q, kv_len and this step's K/V rows come from the previous kernel

  const int n_items = T * nsplit;
  uint32_t it = 0;                                    // pages streamed so far by this CTA

  if (warp == kConsumerWarps) {
    // ------------------------------------------------------------ producer warp
    for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
      const int row = w / nsplit, split = w - row * nsplit;
      const int len = kv_len[row];
      const int n_pages = (len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
      const int p0 = split * pps, p1 = min(p0 + pps, n_pages);
      const int32_t* btr = bt + (int64_t)row_seq[row] * bt_stride;
      for (int pb = p0; pb < p1; pb += 32) {
        const int nb = min(32, p1 - pb);
        const int my_blk = lane < nb ? btr[pb + lane] : 0;
        for (int j = 0; j < nb; ++j) {
          const int blk = __shfl_sync(0xffffffffu, my_blk, j);
          if (lane == 0) {
            const uint32_t s = it % kStages, ph = (it / kStages) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&full[s], CBX_PAGE_BYTES);
            bulk_g2s(smem + s * CBX_PAGE_BYTES, kv_layer + (size_t)blk * CBX_PAGE_BYTES, CBX_PAGE_BYTES, &full[s]);
          }
          ++it;
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------- consumer warps
  const int h = warp;
  const int t16 = lane & 15, half = lane >> 4;        // QK mapping
  const int tg = lane >> 3, c = lane & 7;             // PV mapping
  const float qscale = 0.125f * kLog2e;               // 1/sqrt(64), scores kept in log2 domain

  for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
    const int row = w / nsplit, split = w - row * nsplit;
    const int len = kv_len[row];
    const int n_pages = (len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
    const int p0 = split * pps, p1 = min(p0 + pps, n_pages);
    if (p0 >= p1) {
      if (split == 0 && lane < 8)                     // row with no KV (finished pair): zeros
        reinterpret_cast<uint4*>(out + (size_t)row * CBX_HIDDEN + h * 64)[lane] = make_uint4(0, 0, 0, 0);
      continue;
    }
    // q for (row, head): this lane's half (32 elements), pre-scaled
    float q[32];
    {
      const uint4* qp = reinterpret_cast<const uint4*>(qkv + (size_t)row * 3072 + h * 64 + half * 32);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 v = qp[i];
        q[i * 8 + 0] = bf_lo(v.x) * qscale; q[i * 8 + 1] = bf_hi(v.x) * qscale;
        q[i * 8 + 2] = bf_lo(v.y) * qscale; q[i * 8 + 3] = bf_hi(v.y) * qscale;
        q[i * 8 + 4] = bf_lo(v.z) * qscale; q[i * 8 + 5] = bf_hi(v.z) * qscale;
        q[i * 8 + 6] = bf_lo(v.w) * qscale; q[i * 8 + 7] = bf_hi(v.w) * qscale;
      }
    }
    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;

    for (int pg = p0; pg < p1; ++pg, ++it) {
      const uint32_t s = it % kStages, ph = (it / kStages) & 1;
      mbar_wait(&full[s], ph);
      const uint8_t* kt = smem + s * CBX_PAGE_BYTES + h * 2048;
      const uint8_t* vt = kt + 32768;
      // ---- scores: lane = (token t16, half)
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
      const int n_valid = len - pg * CBX_BLOCK_SIZE;   // tokens of this page inside the sequence
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
      // ---- PV: lane = (token group tg, chunk c)
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
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
    }
    // ---- item epilogue: reduce over token groups, publish
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
      acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
    }
    const float lsum = warp_sum(l) * 0.5f;             // both halves hold identical p
    if (nsplit == 1) {
      if (lane < 8) {
        const float inv = 1.f / lsum;
        reinterpret_cast<uint4*>(out + (size_t)row * CBX_HIDDEN + h * 64)[lane] =
            make_uint4(pack_bf16(acc[0] * inv, acc[1] * inv), pack_bf16(acc[2] * inv, acc[3] * inv),
                       pack_bf16(acc[4] * inv, acc[5] * inv), pack_bf16(acc[6] * inv, acc[7] * inv));
      }
    } else {
      const size_t pi = (size_t)row * nsplit + split;
      if (lane < 8) {
        float4* dst = reinterpret_cast<float4*>(o_part + pi * CBX_HIDDEN + h * 64 + lane * 8);
        dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
      }
      if (lane == 0) ml_part[pi * CBX_HEADS + h] = make_float2(m, lsum);
      if (!kTicket) continue;
      // ---- ticket: the CTA that publishes the row's last partial merges all of them
      __threadfence();
      asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
      if (threadIdx.x == 0) {
        const int ns = min(nsplit, (n_pages + pps - 1) / pps);
        const int ticket = atomicAdd(&counters[row], 1);
        s_last = (ticket == ns - 1);
        if (ticket == ns - 1) counters[row] = 0;      // self-reset for the next launch
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
      if (s_last) {
        __threadfence();
        const int ns = min(nsplit, (n_pages + pps - 1) / pps);
        const int d2 = threadIdx.x * 2, hh = d2 >> 6;  // 512 threads x 2 consecutive d
        float mx = -INFINITY;
        for (int sp = 0; sp < ns; ++sp) mx = fmaxf(mx, __ldcg(&ml_part[((size_t)row * nsplit + sp) * CBX_HEADS + hh]).x);
        float ox = 0.f, oy = 0.f, lt = 0.f;
        for (int sp = 0; sp < ns; ++sp) {
          const size_t pj = (size_t)row * nsplit + sp;
          const float2 ml = __ldcg(&ml_part[pj * CBX_HEADS + hh]);
          const float wgt = exp2f(ml.x - mx);
          const float2 a = __ldcg(reinterpret_cast<const float2*>(o_part + pj * CBX_HIDDEN + d2));
          ox = fmaf(wgt, a.x, ox); oy = fmaf(wgt, a.y, oy);
          lt = fmaf(wgt, ml.y, lt);
        }
        const float inv = 1.f / lt;
        *reinterpret_cast<uint32_t*>(out + (size_t)row * CBX_HIDDEN + d2) = pack_bf16(ox * inv, oy * inv);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Balanced variant (4 < T <= kPlanMaxRows): the launch's pages, flattened in (row, page) order, are
// cut into gridDim.x equal ranges, so every CTA streams the same number of bytes whatever the
// spread of context lengths (the fixed pages-per-split plan above loses 7-14 % on ragged batches).
// A row that lies inside one range is normalised and written directly; a row cut by a range
// boundary leaves (o, m, l) partials — at most two per CTA: slot 0 for a row begun by an earlier
// CTA, slot 1 for a row continued by the next — and the CTA holding the row's last ticket merges
// them.  In (row, page) order the CTA that holds a row's head reaches it last, so merges fall at the
// end of a CTA's range rather than in the middle of its page stream.
// plan (int32): [0] total pages, [1] pages per CTA, [2 + 2c], [3 + 2c] = (row, page) where CTA c starts;
// [kPlanFirst] = 1 and [kPlanFirst + 1 + 3c + j] = physical id of the j-th page of CTA c's stream when the block table is
// given: the attention producer then issues its first ring fill straight after reading the plan, instead of after two more
// dependent L2 round trips (kv_len / row_seq of its start row, then the block-table entries) — per launch, 30 per step.
__global__ void __launch_bounds__(1024) attn_plan_kernel(const int32_t* kv_len, int T, int G,
                                                         int32_t* __restrict__ plan, const int32_t* bt, int bt_stride,
                                                         const int32_t* row_seq) {
  __shared__ int s_first[kPlanMaxRows + 1];            // s_first[r] = flattened index of row r's first page
  __shared__ int s_warp[32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  pdl_trigger();
  pdl_wait();                                          // kv_len is staged by the previous step's sampler
  kv_len = pdl_acquire(kv_len);
  int np[4], sum = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = t * 4 + i;
    np[i] = r < T ? (kv_len[r] + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE : 0;
    sum += np[i];
  }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  int run = incl - sum + (warp > 0 ? s_warp[warp - 1] : 0);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = t * 4 + i;
    if (r <= T) s_first[r] = run;
    run += np[i];
  }
  __syncthreads();
  const int total = s_first[T];
  const int q = max(1, (total + G - 1) / G);
  if (t == 0) { plan[0] = total; plan[1] = q; }
  if (t < G) {
    const int g0 = t * q;
    int row = T, page = 0;
    if (g0 < total) {
      int lo = 0, hi = T;                              // last r with s_first[r] <= g0 (it owns pages: s_first[r + 1] > g0)
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_first[mid] <= g0) lo = mid; else hi = mid;
      }
      row = lo; page = g0 - s_first[lo];
    }
    plan[2 + 2 * t] = row; plan[3 + 2 * t] = page;
    if (bt != nullptr) {                               // the first kPrimed pages of this range, in stream order
      const int n = min(kPrimed, min(q, total - g0));
      int r = row, pg = page, rr[kPrimed], pp[kPrimed], sq[kPrimed];
#pragma unroll
      for (int j = 0; j < kPrimed; ++j) {              // (row, page) of the three pages from shared memory only ...
        while (r < T && pg >= s_first[r + 1] - s_first[r]) { ++r; pg = 0; }
        rr[j] = min(r, T - 1); pp[j] = r < T ? pg : 0;
        ++pg;
      }
#pragma unroll
      for (int j = 0; j < kPrimed; ++j) sq[j] = row_seq[rr[j]];          // ... then two rounds of three loads, not six in a chain
#pragma unroll
      for (int j = 0; j < kPrimed; ++j) {
        const int blk = bt[(int64_t)sq[j] * bt_stride + pp[j]];
        if (j < n) plan[kPlanFirst + 1 + kPrimed * t + j] = blk;
      }
    }
  }
  if (t == 0) plan[kPlanFirst] = bt != nullptr ? 1 : 0;
}

__global__ void __launch_bounds__(kThreads, 1) paged_attn_decode_bal_kernel(
    const bf16* qkv, const uint8_t* __restrict__ kv_layer, const int32_t* bt,
    int bt_stride, const int32_t* row_seq, const int32_t* kv_len,
    bf16* __restrict__ out, float* __restrict__ o_part, float2* __restrict__ ml_part,
    int32_t* __restrict__ counters, const int32_t* plan, int T) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * CBX_PAGE_BYTES);
  uint64_t* empty = full + kStages;
  volatile int& s_last = *reinterpret_cast<volatile int*>(empty + kStages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  pdl_trigger();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kConsumerWarps); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();
  plan = pdl_acquire(plan);
  kv_len = pdl_acquire(kv_len);

  // everything this CTA needs from the plan in ONE round trip: ld.volatile keeps the six loads together (plain loads were
  // scheduled behind the branches that use them: flag -> branch -> ids, two more dependent trips)
  auto ld_now = [](const int32_t* p) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
  };
  const int total = ld_now(plan), per = ld_now(plan + 1);
  int row = ld_now(plan + 2 + 2 * blockIdx.x), pg = ld_now(plan + 3 + 2 * blockIdx.x);
  const int has_first = ld_now(plan + kPlanFirst);
  const int first_blk = ld_now(plan + kPlanFirst + 1 + kPrimed * blockIdx.x + min(lane, kPrimed - 1));
  const int g0 = min((int)blockIdx.x * per, total);
  int rem = min(per, total - g0);                      // pages this CTA still has to stream
  uint32_t it = 0;

  if (warp == kConsumerWarps) {
    // ------------------------------------------------------------ producer warp
    // first ring fill from the page ids the plan kernel resolved (one L2 round trip after the wait instead of three)
    static_assert(kPrimed == kStages, "the primed copies are exactly the first ring fill: no empty-barrier wait");
    uint32_t primed = 0;
    if (has_first) {
      primed = (uint32_t)max(0, min(kPrimed, rem));
      for (uint32_t j = 0; j < primed; ++j) {
        const int blk = __shfl_sync(0xffffffffu, first_blk, j);
        if (lane == 0) {
          mbar_arrive_expect_tx(&full[j], CBX_PAGE_BYTES);
          bulk_g2s(smem + j * CBX_PAGE_BYTES, kv_layer + (size_t)blk * CBX_PAGE_BYTES, CBX_PAGE_BYTES, &full[j]);
        }
      }
    }
    while (rem > 0) {
      const int n_pages = (kv_len[row] + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
      if (n_pages == 0) { ++row; continue; }
      const int p1 = min(n_pages, pg + rem);
      const int32_t* btr = bt + (int64_t)row_seq[row] * bt_stride;
      for (int pb = pg; pb < p1; pb += 32) {
        const int nb = min(32, p1 - pb);
        const int my_blk = lane < nb ? btr[pb + lane] : 0;
        for (int j = 0; j < nb; ++j) {
          const int blk = __shfl_sync(0xffffffffu, my_blk, j);
          if (lane == 0 && it >= primed) {
            const uint32_t s = it % kStages, ph = (it / kStages) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&full[s], CBX_PAGE_BYTES);
            bulk_g2s(smem + s * CBX_PAGE_BYTES, kv_layer + (size_t)blk * CBX_PAGE_BYTES, CBX_PAGE_BYTES, &full[s]);
          }
          ++it;
        }
      }
      rem -= p1 - pg;
      ++row; pg = 0;
    }
    return;
  }

  // -------------------------------------------------------------- consumer warps
  const int h = warp;
  const int t16 = lane & 15, half = lane >> 4;
  const int tg = lane >> 3, c = lane & 7;
  const float qscale = 0.125f * kLog2e;

  // rows without KV (finished pairs) belong to no range: zero them grid-wide
  for (int r = blockIdx.x * kConsumerWarps + warp; r < T; r += gridDim.x * kConsumerWarps)
    if (kv_len[r] <= 0 && lane < 8)
      for (int hh = 0; hh < CBX_HEADS; ++hh)
        reinterpret_cast<uint4*>(out + (size_t)r * CBX_HIDDEN + hh * 64)[lane] = make_uint4(0, 0, 0, 0);

  int g = g0;                                          // flattened index of (row, pg)
  while (rem > 0) {
    const int len = kv_len[row];
    const int n_pages = (len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
    if (n_pages == 0) { ++row; continue; }
    const int p0 = pg, p1 = min(n_pages, pg + rem);
    float q[32];
    {
      const uint4* qp = reinterpret_cast<const uint4*>(qkv + (size_t)row * 3072 + h * 64 + half * 32);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 v = qp[i];
        q[i * 8 + 0] = bf_lo(v.x) * qscale; q[i * 8 + 1] = bf_hi(v.x) * qscale;
        q[i * 8 + 2] = bf_lo(v.y) * qscale; q[i * 8 + 3] = bf_hi(v.y) * qscale;
        q[i * 8 + 4] = bf_lo(v.z) * qscale; q[i * 8 + 5] = bf_hi(v.z) * qscale;
        q[i * 8 + 6] = bf_lo(v.w) * qscale; q[i * 8 + 7] = bf_hi(v.w) * qscale;
      }
    }
    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;

    for (int pgi = p0; pgi < p1; ++pgi, ++it) {
      const uint32_t s = it % kStages, ph = (it / kStages) & 1;
      mbar_wait(&full[s], ph);
      const uint8_t* kt = smem + s * CBX_PAGE_BYTES + h * 2048;
      const uint8_t* vt = kt + 32768;
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
      const int n_valid = len - pgi * CBX_BLOCK_SIZE;
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
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
      acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
    }
    const float lsum = warp_sum(l) * 0.5f;
    if (p0 == 0 && p1 == n_pages) {
      if (lane < 8) {
        const float inv = 1.f / lsum;
        reinterpret_cast<uint4*>(out + (size_t)row * CBX_HIDDEN + h * 64)[lane] =
            make_uint4(pack_bf16(acc[0] * inv, acc[1] * inv), pack_bf16(acc[2] * inv, acc[3] * inv),
                       pack_bf16(acc[4] * inv, acc[5] * inv), pack_bf16(acc[6] * inv, acc[7] * inv));
      }
    } else {
      // the row's pages span CTAs c_first .. c_last; the head CTA publishes in slot 1, the others in slot 0
      const int first_page = g - p0;
      const int c_first = first_page / per, c_last = (first_page + n_pages - 1) / per;
      const size_t pi = (size_t)blockIdx.x * 2 + ((int)blockIdx.x == c_first ? 1 : 0);
      if (lane < 8) {
        float4* dst = reinterpret_cast<float4*>(o_part + pi * CBX_HIDDEN + h * 64 + lane * 8);
        dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
      }
      if (lane == 0) ml_part[pi * CBX_HEADS + h] = make_float2(m, lsum);
      __threadfence();
      asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
      if (threadIdx.x == 0) {
        const int ns = c_last - c_first + 1;
        const int ticket = atomicAdd(&counters[row], 1);
        s_last = (ticket == ns - 1);
        if (ticket == ns - 1) counters[row] = 0;      // self-reset for the next launch
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
      if (s_last) {
        __threadfence();
        const int d2 = threadIdx.x * 2, hh = d2 >> 6;  // 512 threads x 2 consecutive d
        float mx = -INFINITY;
        for (int cc = c_first; cc <= c_last; ++cc)
          mx = fmaxf(mx, __ldcg(&ml_part[((size_t)cc * 2 + (cc == c_first ? 1 : 0)) * CBX_HEADS + hh]).x);
        float ox = 0.f, oy = 0.f, lt = 0.f;
        for (int cc = c_first; cc <= c_last; ++cc) {
          const size_t pj = (size_t)cc * 2 + (cc == c_first ? 1 : 0);
          const float2 ml = __ldcg(&ml_part[pj * CBX_HEADS + hh]);
          const float wgt = exp2f(ml.x - mx);
          const float2 a = __ldcg(reinterpret_cast<const float2*>(o_part + pj * CBX_HIDDEN + d2));
          ox = fmaf(wgt, a.x, ox); oy = fmaf(wgt, a.y, oy);
          lt = fmaf(wgt, ml.y, lt);
        }
        const float inv = 1.f / lt;
        *reinterpret_cast<uint32_t*>(out + (size_t)row * CBX_HIDDEN + d2) = pack_bf16(ox * inv, oy * inv);
      }
      // the next partial of this CTA goes to the other slot, so nothing here can be overwritten early
    }
    g += p1 - p0;
    rem -= p1 - p0;
    ++row; pg = 0;
  }
}


// ---------------------------------------------------------------------------------------------
// Small-batch variant (T <= 4 rows: the 1-2 CFG-pair latency regime).  One CTA per (row, head,
// split); the S splits of a (row, head) form a thread-block cluster and merge their partial
// (m, l, o) through distributed shared memory — no global round trip, no ticket, no second
// launch.  A head's K (2 KB) and V (2 KB) of a page are contiguous, so each page costs the CTA two
// 2 KB bulk copies; all of a CTA's pages are requested up front (<= 32 pages = 128 KB in flight).
constexpr int kSmallWarps = 4;
constexpr int kSmallMaxPages = 32;                    // local pages per CTA
constexpr int kSmallSmem = kSmallMaxPages * 4096 + kSmallMaxPages * 8 + kSmallWarps * 68 * 4 + 16;

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(const float* local, uint32_t rank) {
  uint32_t ra; float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
  return v;
}

template <int S>
__global__ void __launch_bounds__(kSmallWarps * 32) paged_attn_small_kernel(
    const bf16* qkv, const uint8_t* __restrict__ kv_layer, const int32_t* bt, int bt_stride,
    const int32_t* row_seq, const int32_t* kv_len, bf16* __restrict__ out,
    const void* next_w, size_t next_bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmallMaxPages * 4096);
  float* red = reinterpret_cast<float*>(bars + kSmallMaxPages);   // [4 warps][68]: o[64], m, l
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = S > 1 ? (int)cluster_ctarank() : 0;            // == blockIdx.x (cluster spans grid.x)
  const int h = blockIdx.y, row = blockIdx.z;

  pdl_trigger();
  if (threadIdx.x < kSmallMaxPages) mbar_init(&bars[threadIdx.x], 1);
  fence_barrier_init();
  l2_prefetch_slice(next_w, next_bytes);              // a later projection's weights -> L2 while we wait
  // Page ids of this CTA's share of the row, read BEFORE the wait: the block table is written at admission and
  // row_seq is re-staged with the same value every step, so neither depends on the kernels this one waits for —
  // kv_len (how many of the ids are live) does, and is read after the wait.  Takes the row_seq -> block-table
  // round trips off the kernel-to-kernel critical path (one of the designed pre-wait loads, see the SASS audit).
  int pre_blk = 0;
  if (warp == 0 && rank + lane * S < bt_stride) pre_blk = bt[(int64_t)row_seq[row] * bt_stride + rank + lane * S];
  __syncthreads();
  pdl_wait();

  // q and kv_len are requested together (both come from the awaited kernels); nothing consumes either before both
  // loads are in flight, so the page copies start one L2 round trip after the wait, not two
  const int t16 = lane & 15, half = lane >> 4;        // QK mapping
  const int tg = lane >> 3, c = lane & 7;             // PV mapping
  const float qscale = 0.125f * kLog2e;
  uint4 qraw[4];
  {
    const uint4* qp = reinterpret_cast<const uint4*>(qkv + (size_t)row * 3072 + h * 64 + half * 32);
#pragma unroll
    for (int i = 0; i < 4; ++i) qraw[i] = qp[i];
  }
  const int len = kv_len[row];
  const int n_pages = (len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  const int n_local = n_pages > rank ? (n_pages - rank + S - 1) / S : 0;   // pages rank, rank+S, ...
  if (warp == 0 && lane < n_local) {
    const int blk = pre_blk;
    const uint8_t* src = kv_layer + (size_t)blk * CBX_PAGE_BYTES + h * 2048;
    mbar_arrive_expect_tx(&bars[lane], 4096);
    bulk_g2s(smem + lane * 4096, src, 2048, &bars[lane]);
    bulk_g2s(smem + lane * 4096 + 2048, src + 32768, 2048, &bars[lane]);
  }
  float q[32];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 v = qraw[i];
    q[i * 8 + 0] = bf_lo(v.x) * qscale; q[i * 8 + 1] = bf_hi(v.x) * qscale;
    q[i * 8 + 2] = bf_lo(v.y) * qscale; q[i * 8 + 3] = bf_hi(v.y) * qscale;
    q[i * 8 + 4] = bf_lo(v.z) * qscale; q[i * 8 + 5] = bf_hi(v.z) * qscale;
    q[i * 8 + 6] = bf_lo(v.w) * qscale; q[i * 8 + 7] = bf_hi(v.w) * qscale;
  }
  float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int j = warp; j < n_local; j += kSmallWarps) {
    mbar_wait(&bars[j], 0);
    const uint8_t* kt = smem + j * 4096;
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
    const int n_valid = len - (rank + j * S) * CBX_BLOCK_SIZE;
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
  // ---- warp partial -> shared memory
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
  __syncthreads();
  if (S > 1) cluster_sync_all();
  // ---- merge the S x 4 warp partials (cluster rank 0, one thread per output dim)
  if (rank == 0 && threadIdx.x < 64) {
    const int d = threadIdx.x;
    float mx = -INFINITY;
    for (int r = 0; r < S; ++r)
      for (int w = 0; w < kSmallWarps; ++w) mx = fmaxf(mx, S > 1 ? ld_dsmem_f32(red + w * 68 + 64, r) : red[w * 68 + 64]);
    float o = 0.f, lt = 0.f;
    if (mx > -INFINITY) {
      for (int r = 0; r < S; ++r)
        for (int w = 0; w < kSmallWarps; ++w) {
          const float* pw = red + w * 68;
          const float pm = S > 1 ? ld_dsmem_f32(pw + 64, r) : pw[64];
          if (pm == -INFINITY) continue;
          const float wgt = exp2f(pm - mx);
          o = fmaf(wgt, S > 1 ? ld_dsmem_f32(pw + d, r) : pw[d], o);
          lt = fmaf(wgt, S > 1 ? ld_dsmem_f32(pw + 65, r) : pw[65], lt);
        }
      o /= lt;
    }
    const float o2 = __shfl_down_sync(0xffffffffu, o, 1);
    if ((d & 1) == 0) *reinterpret_cast<uint32_t*>(out + (size_t)row * CBX_HIDDEN + h * 64 + d) = pack_bf16(o, o2);
  }
  if (S > 1) cluster_sync_all();                       // peers' shared memory stays alive until rank 0 has read it
}

// ---------------------------------------------------------------------------------------------
// Prefill (causal) attention over the paged cache: one item = 16 consecutive query rows of one
// sequence (a "query block" qb), which need KV pages 0..qb — each page is loaded ONCE for the 16
// queries instead of once per row.  Same persistent producer/consumer skeleton, warp = head, but the
// math runs on the tensor cores (mma.sync m16n8k16, FlashAttention-2 register pipeline): per
// (head, page) S[16x16] = Q·K^T is 8 MMAs, O[16x64] += P·V another 8, against ~1000 FMA-issue slots
// for the same work on CUDA cores (the first version of this kernel was compute-bound at 406 us per
// layer for 16 000 rows).  K fragments come from the page's pre-swizzled K rows with ldmatrix; the V
// half of a page is fetched with a 2-D TMA copy whose hardware 128-byte swizzle gives V the same
// chunk swizzle in shared memory, so ldmatrix.trans is conflict-free too.  Softmax statistics fp32;
// P is rounded to bf16 for the PV product, as flash-attention (the kernel this replaces) does.
// items int32 [n][4] = {row0 (first query row in qkv/out), n_q (1..16), block-table row, qb}.
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(kThreads, 1) paged_attn_prefill_kernel(
    const __grid_constant__ CUtensorMap vmap, const bf16* qkv, const uint8_t* __restrict__ kv_layer,
    const int32_t* bt, int bt_stride, const int4* items, int n_items, bf16* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * CBX_PAGE_BYTES);
  uint64_t* empty = full + kStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&vmap);
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kConsumerWarps); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();
  uint32_t it = 0;
  if (warp == kConsumerWarps) {                       // ---- producer: K half as stored, V half swizzled by the TMA engine
    for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
      const int4 im = items[w];
      const int32_t* btr = bt + (int64_t)im.z * bt_stride;
      for (int pb = 0; pb <= im.w; pb += 32) {
        const int nb = min(32, im.w + 1 - pb);
        const int my_blk = lane < nb ? btr[pb + lane] : 0;
        for (int j = 0; j < nb; ++j) {
          const int blk = __shfl_sync(0xffffffffu, my_blk, j);
          if (lane == 0) {
            const uint32_t s = it % kStages, ph = (it / kStages) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&full[s], CBX_PAGE_BYTES);
            bulk_g2s(smem + s * CBX_PAGE_BYTES, kv_layer + (size_t)blk * CBX_PAGE_BYTES, 32768, &full[s]);
#ifdef CBX_PF_NOSWZ
            bulk_g2s(smem + s * CBX_PAGE_BYTES + 32768, kv_layer + (size_t)blk * CBX_PAGE_BYTES + 32768, 32768, &full[s]);
#else
            tma_load_2d(smem + s * CBX_PAGE_BYTES + 32768, &vmap, 0, blk * 512 + 256, &full[s]);
#endif
          }
          ++it;
        }
      }
    }
    return;
  }
  const int h = warp, g = lane >> 2, c = lane & 3;    // MMA fragment coordinates: rows g / g+8, column pair c
  const int mat = lane >> 3, r8 = lane & 7;           // ldmatrix: this lane addresses row r8 of 8x8 matrix `mat`
  const float qscale = 0.125f * kLog2e;               // 1/sqrt(64); scores kept in the log2 domain
  for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
    const int4 im = items[w];
    const int qb = im.w;
    // Q as the A operand: 4 k-steps of 16 d; rows beyond n_q are zero
    uint32_t qa[4][4];
    {
      const uint32_t* q0 = reinterpret_cast<const uint32_t*>(qkv + (size_t)(im.x + g) * 3072 + h * 64) + c;
      const uint32_t* q1 = reinterpret_cast<const uint32_t*>(qkv + (size_t)(im.x + g + 8) * 3072 + h * 64) + c;
      const bool l0 = g < im.y, l1 = g + 8 < im.y;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        qa[ks][0] = l0 ? __ldcg(q0 + ks * 8) : 0u;
        qa[ks][1] = l1 ? __ldcg(q1 + ks * 8) : 0u;
        qa[ks][2] = l0 ? __ldcg(q0 + ks * 8 + 4) : 0u;
        qa[ks][3] = l1 ? __ldcg(q1 + ks * 8 + 4) : 0u;
      }
    }
    float m0 = -INFINITY, m1 = -INFINITY, l0s = 0.f, l1s = 0.f;
    float o[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    for (int pg = 0; pg <= qb; ++pg, ++it) {
      const uint32_t s = it % kStages, ph = (it / kStages) & 1;
      mbar_wait(&full[s], ph);
      __syncwarp();                                   // ldmatrix / mma below are .aligned: the warp must be converged after the spin
      const uint32_t kt = smem_u32(smem + s * CBX_PAGE_BYTES + h * 2048);
      const uint32_t vt = kt + 32768;
      // ---- S = Q·K^T: keys 0-7 (sc[0]) and 8-15 (sc[1])
      float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        uint32_t kb[4];
        const int key = (mat >> 1) * 8 + r8, chunk = ks * 2 + (mat & 1);
        ldsm_x4(kt + key * 128 + ((chunk ^ (key & 7)) << 4), kb);
        mma_bf16_16816(sc[0], qa[ks], kb[0], kb[1]);
        mma_bf16_16816(sc[1], qa[ks], kb[2], kb[3]);
      }
      // ---- online softmax of rows g and g+8 (this lane holds columns nt*8 + 2c, +1)
      const bool diag = pg == qb;
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int col = nt * 8 + 2 * c + e;
          sc[nt][e] = (!diag || col <= g) ? sc[nt][e] * qscale : -INFINITY;
          sc[nt][2 + e] = (!diag || col <= g + 8) ? sc[nt][2 + e] * qscale : -INFINITY;
          mx0 = fmaxf(mx0, sc[nt][e]);
          mx1 = fmaxf(mx1, sc[nt][2 + e]);
        }
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);   // key 0 of page 0 is visible to every query: finite
      const float cr0 = exp2f(m0 - mn0), cr1 = exp2f(m1 - mn1);
      m0 = mn0; m1 = mn1;
      uint32_t pa[4];
      {
        const float p00 = exp2f(sc[0][0] - mn0), p01 = exp2f(sc[0][1] - mn0), p02 = exp2f(sc[0][2] - mn1), p03 = exp2f(sc[0][3] - mn1);
        const float p10 = exp2f(sc[1][0] - mn0), p11 = exp2f(sc[1][1] - mn0), p12 = exp2f(sc[1][2] - mn1), p13 = exp2f(sc[1][3] - mn1);
        l0s = fmaf(l0s, cr0, (p00 + p01) + (p10 + p11));
        l1s = fmaf(l1s, cr1, (p02 + p03) + (p12 + p13));
        pa[0] = pack_bf16(p00, p01); pa[1] = pack_bf16(p02, p03); pa[2] = pack_bf16(p10, p11); pa[3] = pack_bf16(p12, p13);
      }
      // ---- O = O·corr + P·V
#pragma unroll
      for (int nd = 0; nd < 8; nd += 2) {
        uint32_t vb[4];
        const int key = (mat & 1) * 8 + r8, chunk = nd + (mat >> 1);
#ifdef CBX_PF_NOSWZ
        ldsm_x4_trans(vt + key * 128 + (chunk << 4), vb);
#else
        ldsm_x4_trans(vt + key * 128 + ((chunk ^ (key & 7)) << 4), vb);
#endif
        o[nd][0] *= cr0; o[nd][1] *= cr0; o[nd][2] *= cr1; o[nd][3] *= cr1;
        o[nd + 1][0] *= cr0; o[nd + 1][1] *= cr0; o[nd + 1][2] *= cr1; o[nd + 1][3] *= cr1;
        mma_bf16_16816(o[nd], pa, vb[0], vb[1]);
        mma_bf16_16816(o[nd + 1], pa, vb[2], vb[3]);
      }
      // ptxas is free to schedule the arrive ABOVE the MMAs that consume the last ldmatrix (it did: SASS LDSM,
      // SYNCS.ARRIVE, HMMA), so nothing guaranteed that the shared-memory read had been performed when the producer
      // saw the stage empty and let the TMA engine (async proxy) overwrite it: some heads of ~3 % of the rows came
      // out with a neighbouring page's K/V, differently on every launch.  The proxy fence (MEMBAR.ALL.CTA +
      // FENCE.VIEW.ASYNC) completes this thread's reads before the release.
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
    }
    l0s += __shfl_xor_sync(0xffffffffu, l0s, 1); l0s += __shfl_xor_sync(0xffffffffu, l0s, 2);
    l1s += __shfl_xor_sync(0xffffffffu, l1s, 1); l1s += __shfl_xor_sync(0xffffffffu, l1s, 2);
    const float i0 = 1.f / l0s, i1 = 1.f / l1s;
    if (g < im.y) {
      uint32_t* dst = reinterpret_cast<uint32_t*>(out + (size_t)(im.x + g) * CBX_HIDDEN + h * 64) + c;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) dst[nd * 4] = pack_bf16(o[nd][0] * i0, o[nd][1] * i0);
    }
    if (g + 8 < im.y) {
      uint32_t* dst = reinterpret_cast<uint32_t*>(out + (size_t)(im.x + g + 8) * CBX_HIDDEN + h * 64) + c;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) dst[nd * 4] = pack_bf16(o[nd][2] * i1, o[nd][3] * i1);
    }
  }
}

// Combine split partials: 256 threads per row, thread = 4 consecutive d of one head.
__global__ void __launch_bounds__(256) attn_merge_kernel(const float* o_part,
                                                         const float2* ml_part,
                                                         const int32_t* kv_len,
                                                         bf16* __restrict__ out, int pps, int nsplit) {
  const int row = blockIdx.x, t = threadIdx.x, h = t >> 4;
  pdl_trigger();
  pdl_wait();
  const int n_pages = (kv_len[row] + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
  const int ns = min(nsplit, (n_pages + pps - 1) / pps);
  if (ns == 0) return;                                 // the attention kernel already wrote zeros
  float mx = -INFINITY;
  for (int s = 0; s < ns; ++s) mx = fmaxf(mx, ml_part[((size_t)row * nsplit + s) * CBX_HEADS + h].x);
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  float lt = 0.f;
  for (int s = 0; s < ns; ++s) {
    const size_t pi = (size_t)row * nsplit + s;
    const float2 ml = ml_part[pi * CBX_HEADS + h];
    const float wgt = exp2f(ml.x - mx);
    const float4 a = reinterpret_cast<const float4*>(o_part + pi * CBX_HIDDEN)[t];
    o.x = fmaf(wgt, a.x, o.x); o.y = fmaf(wgt, a.y, o.y); o.z = fmaf(wgt, a.z, o.z); o.w = fmaf(wgt, a.w, o.w);
    lt = fmaf(wgt, ml.y, lt);
  }
  const float inv = 1.f / lt;
  reinterpret_cast<uint2*>(out + (size_t)row * CBX_HIDDEN)[t] =
      make_uint2(pack_bf16(o.x * inv, o.y * inv), pack_bf16(o.z * inv, o.w * inv));
}
}  // namespace

namespace cbx {
int init_attn() {
  static bool done = false;
  if (done) return 0;
  cudaError_t e = cudaFuncSetAttribute(paged_attn_decode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kSmemBytes);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_decode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_decode_bal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_small_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSmem);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_small_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSmem);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(paged_attn_small_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSmem);
  CBX_REQUIRE(e == cudaSuccess, "attn smem attribute: %s", cudaGetErrorString(e));
  done = true;
  return 0;
}
}  // namespace cbx

extern "C" int64_t cbx_attn_scratch_bytes(int T, int max_kv_len) {
  if (T <= 0) return 0;
  const AttnPlan p = make_plan(T, max_kv_len);
  const int64_t fixed = (int64_t)T * p.nsplit * (CBX_HIDDEN * 4 + CBX_HEADS * 8) + 256;
  return fixed > kBalScratchBytes ? fixed : kBalScratchBytes;
}
extern "C" int cbx_attn_num_splits(int T, int max_kv_len) { return T <= 0 ? 0 : make_plan(T, max_kv_len).nsplit; }

namespace cbx {
// one-shot hint consumed by the next small-batch attention launch: weights to prefetch into L2
static const void* g_next_w = nullptr;
static size_t g_next_bytes = 0;
void attn_set_prefetch(const void* w, size_t bytes) { g_next_w = w; g_next_bytes = bytes; }
int paged_attn_decode_ex(const void* qkv, const void* kv_layer, const int32_t* block_table, int bt_stride,
                         const int32_t* row_seq, const int32_t* kv_len, void* out, void* scratch, int32_t* counters,
                         int T, int max_kv_len, bool have_plan, void* stream, int n_ctas = 0);
int attn_plan(const int32_t* kv_len, int T, void* scratch, void* stream, const int32_t* block_table, int bt_stride,
              const int32_t* row_seq, int n_ctas = 0);
static bool balanced_on() {
  static const bool on = !(getenv("CBX_ATTN_BALANCED") && getenv("CBX_ATTN_BALANCED")[0] == '0');
  return on;
}
static int32_t* plan_of(void* scratch) {
  return reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(scratch) + 2 * kMaxSms * (CBX_HIDDEN * 4 + CBX_HEADS * 8));
}
}
// Page-range plan shared by every layer of a decode step (kv_len is the same for all 30 launches).
// Returns 1 when the balanced variant does not apply to T rows.
// block_table / row_seq (may be NULL): when given, each range's first pages are resolved here, once for all layers.
// n_ctas (0 = one per SM): CTAs of the attention launches that will use the plan — fewer than the SM count when the
// engine runs the two CFG halves of a step on two streams and leaves SMs to the other half's GEMMs.
int cbx::attn_plan(const int32_t* kv_len, int T, void* scratch, void* stream, const int32_t* block_table, int bt_stride,
                   const int32_t* row_seq, int n_ctas) {
  if (T <= 4 || T > kPlanMaxRows || !balanced_on() || scratch == nullptr) return 1;
  if (row_seq == nullptr) block_table = nullptr;
  const int G = n_ctas > 0 && n_ctas < sm_count() ? n_ctas : sm_count();
  cudaError_t e = launch_pdl(attn_plan_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, kv_len, T, G, plan_of(scratch),
                             block_table, bt_stride, row_seq);
  CBX_REQUIRE(e == cudaSuccess, "attn_plan launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
extern "C" int cbx_paged_attn_decode(const void* qkv, const void* kv_layer, const int32_t* block_table,
                                     int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* out,
                                     void* scratch, int32_t* counters, int T, int max_kv_len, void* stream) {
  return cbx::paged_attn_decode_ex(qkv, kv_layer, block_table, bt_stride, row_seq, kv_len, out, scratch, counters, T,
                                   max_kv_len, false, stream);
}
extern "C" int cbx_attn_plan(const int32_t* kv_len, int T, void* scratch, void* stream) {
  if (T <= 0 || cbx::init_attn()) return -1;
  return cbx::attn_plan(kv_len, T, scratch, stream, nullptr, 0, nullptr);
}
extern "C" int cbx_attn_plan_paged(const int32_t* kv_len, int T, void* scratch, void* stream, const int32_t* block_table,
                                   int bt_stride, const int32_t* row_seq) {
  if (T <= 0 || cbx::init_attn()) return -1;
  return cbx::attn_plan(kv_len, T, scratch, stream, block_table, bt_stride, row_seq);
}
extern "C" int cbx_paged_attn_decode_planned(const void* qkv, const void* kv_layer, const int32_t* block_table,
                                             int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* out,
                                             void* scratch, int32_t* counters, int T, int max_kv_len, void* stream) {
  CBX_REQUIRE(T > 4 && T <= kPlanMaxRows && cbx::balanced_on(), "cbx_paged_attn_decode_planned: no plan exists for T=%d", T);
  return cbx::paged_attn_decode_ex(qkv, kv_layer, block_table, bt_stride, row_seq, kv_len, out, scratch, counters, T,
                                   max_kv_len, true, stream);
}
// have_plan: cbx::attn_plan already ran on this stream for the same kv_len (the engine builds it once per step).
int cbx::paged_attn_decode_ex(const void* qkv, const void* kv_layer, const int32_t* block_table, int bt_stride,
                              const int32_t* row_seq, const int32_t* kv_len, void* out, void* scratch,
                              int32_t* counters, int T, int max_kv_len, bool have_plan, void* stream, int n_ctas) {
  const int max_ctas = n_ctas > 0 && n_ctas < sm_count() ? n_ctas : sm_count();
  if (T <= 0) return 0;
  if (cbx::init_attn()) return -1;
  {
    // small-batch path: clusters of S splits per (row, head)
    static const bool small_on = !(getenv("CBX_ATTN_SMALL") && getenv("CBX_ATTN_SMALL")[0] == '0');
    const int max_pages = (max_kv_len + CBX_BLOCK_SIZE - 1) / CBX_BLOCK_SIZE;
    const int S = T <= 2 ? 4 : 2;
    if (small_on && T <= 4 && max_pages <= kSmallMaxPages * S) {
      cudaStream_t st = (cudaStream_t)stream;
      dim3 grid(S, CBX_HEADS, T), block(kSmallWarps * 32);
      cudaError_t e = S == 4 ? launch_pdl_cluster(paged_attn_small_kernel<4>, grid, block, kSmallSmem, st, S, (const bf16*)qkv,
                                                  (const uint8_t*)kv_layer, block_table, bt_stride, row_seq, kv_len, (bf16*)out,
                                                  g_next_w, g_next_bytes)
                             : launch_pdl_cluster(paged_attn_small_kernel<2>, grid, block, kSmallSmem, st, S, (const bf16*)qkv,
                                                  (const uint8_t*)kv_layer, block_table, bt_stride, row_seq, kv_len, (bf16*)out,
                                                  g_next_w, g_next_bytes);
      g_next_w = nullptr; g_next_bytes = 0;
      CBX_REQUIRE(e == cudaSuccess, "paged_attn_small launch failed: %s", cudaGetErrorString(e));
      count_launch();
      return 0;
    }
  }
  if (T > 4 && T <= kPlanMaxRows && balanced_on()) {
    CBX_REQUIRE(scratch != nullptr && counters != nullptr, "cbx_paged_attn_decode: scratch / counters is NULL");
    if (!have_plan && attn_plan(kv_len, T, scratch, stream, block_table, bt_stride, row_seq, n_ctas)) return -1;
    float* o_part = reinterpret_cast<float*>(scratch);
    float2* ml_part = reinterpret_cast<float2*>(o_part + (size_t)2 * kMaxSms * CBX_HIDDEN);
    cudaError_t e = launch_pdl(paged_attn_decode_bal_kernel, dim3(max_ctas), dim3(kThreads), kSmemBytes, (cudaStream_t)stream,
                               (const bf16*)qkv, (const uint8_t*)kv_layer, block_table, bt_stride, row_seq, kv_len, (bf16*)out,
                               o_part, ml_part, counters, (const int32_t*)plan_of(scratch), T);
    CBX_REQUIRE(e == cudaSuccess, "paged_attn_decode (balanced) launch failed: %s", cudaGetErrorString(e));
    count_launch();
    return 0;
  }
  const AttnPlan p = make_plan(T, max_kv_len);
  float* o_part = reinterpret_cast<float*>(scratch);
  float2* ml_part = reinterpret_cast<float2*>(o_part + (size_t)T * p.nsplit * CBX_HIDDEN);
  CBX_REQUIRE(p.nsplit == 1 || (scratch != nullptr && counters != nullptr),
              "cbx_paged_attn_decode: scratch / counters is NULL");
  const int n_items = T * p.nsplit;
  const int grid = n_items < max_ctas ? n_items : max_ctas;
  const bool ticket = p.nsplit > 1 && n_items <= grid;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e =
      ticket ? launch_pdl(paged_attn_decode_kernel<true>, dim3(grid), dim3(kThreads), kSmemBytes, st, (const bf16*)qkv,
                          (const uint8_t*)kv_layer, block_table, bt_stride, row_seq, kv_len, (bf16*)out, o_part, ml_part,
                          counters, T, p.pps, p.nsplit)
             : launch_pdl(paged_attn_decode_kernel<false>, dim3(grid), dim3(kThreads), kSmemBytes, st, (const bf16*)qkv,
                          (const uint8_t*)kv_layer, block_table, bt_stride, row_seq, kv_len, (bf16*)out, o_part, ml_part,
                          counters, T, p.pps, p.nsplit);
  CBX_REQUIRE(e == cudaSuccess, "paged_attn_decode launch failed: %s", cudaGetErrorString(e));
  count_launch();
  if (p.nsplit > 1 && !ticket) {
    e = launch_pdl(attn_merge_kernel, dim3(T), dim3(256), 0, st, (const float*)o_part, (const float2*)ml_part, kv_len,
                   (bf16*)out, p.pps, p.nsplit);
    CBX_REQUIRE(e == cudaSuccess, "attn_merge launch failed: %s", cudaGetErrorString(e));
    count_launch();
  }
  return 0;
}

// ---- allocation-free decode attention ----------------------------------------------------------
// No counterpart in the reference; added for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The
// balanced kernel's scratch (2 partial slots per CTA + the page-range plan) and the per-row ticket
// counters live in ONE file-scope device array, kStaticScratchBytes on every GPU that loads the
// library.  The array's name is a host-side stand-in in host code: the device address is resolved
// once per device with cudaGetSymbolAddress (with that device current) and passed to the kernels.
namespace {
constexpr int kStaticMaxRows = kPlanMaxRows;                                       // 4095 query rows
constexpr size_t kStaticPartBytes = ((size_t)kBalScratchBytes + 255) / 256 * 256;  // partials + plan
constexpr size_t kStaticScratchBytes = kStaticPartBytes + (size_t)(kStaticMaxRows + 1) * 4;   // + tickets
constexpr int kStaticMaxDevices = 64;
}  // namespace
static __device__ __align__(256) unsigned char g_attn_static_scratch[kStaticScratchBytes];
static unsigned char* g_attn_static_ptr[kStaticMaxDevices];   // device address of the array, per device ordinal

extern "C" int cbx_paged_attn_decode_static(const void* qkv, const void* kv_layer, const int32_t* block_table,
                                            int bt_stride, const int32_t* row_seq, const int32_t* kv_len, void* out,
                                            int T, int max_kv_len, void* stream) {
  if (T == 0) return 0;
  CBX_REQUIRE(T > 4 && T <= kStaticMaxRows,
              "cbx_paged_attn_decode_static: T=%d outside 5..%d rows (the static scratch is sized for the balanced "
              "kernel; T <= 4 needs no scratch: call cbx_paged_attn_decode with scratch = NULL)", T, kStaticMaxRows);
  CBX_REQUIRE(cbx::balanced_on(), "cbx_paged_attn_decode_static: CBX_ATTN_BALANCED=0 disables the kernel this entry uses");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  CBX_REQUIRE(e == cudaSuccess && dev >= 0 && dev < kStaticMaxDevices, "cbx_paged_attn_decode_static: cudaGetDevice: %s",
              cudaGetErrorString(e));
  if (g_attn_static_ptr[dev] == nullptr) {            // set-up, once per device, that device current
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_attn_static_scratch);
    CBX_REQUIRE(e == cudaSuccess && p != nullptr, "cbx_paged_attn_decode_static: cudaGetSymbolAddress: %s",
                cudaGetErrorString(e));
    g_attn_static_ptr[dev] = p;
  }
  unsigned char* base = g_attn_static_ptr[dev];
  int32_t* counters = reinterpret_cast<int32_t*>(base + kStaticPartBytes);
  // every call shares the array and it keeps the last call's contents: the tickets must start at zero (a launch that
  // ran to completion leaves them zero, one that was aborted does not); partials and plan are written before being read
  e = cudaMemsetAsync(counters, 0, (size_t)T * 4, (cudaStream_t)stream);
  CBX_REQUIRE(e == cudaSuccess, "cbx_paged_attn_decode_static: clearing the tickets: %s", cudaGetErrorString(e));
  return cbx::paged_attn_decode_ex(qkv, kv_layer, block_table, bt_stride, row_seq, kv_len, out, base, counters, T,
                                   max_kv_len, false, stream);
}

// Causal attention for prefill rows whose K/V are already in the cache (vllm flash_attn.py:796-800,
// prefill form).  items: device int32 [n_items][4] = {row0, n_q, block-table row, query block}.
extern "C" int cbx_paged_attn_prefill(const void* qkv, const void* kv_layer, const int32_t* block_table, int bt_stride,
                                      const int32_t* items, int n_items, void* out, void* stream) {
  if (n_items <= 0) return 0;
  if (cbx::init_attn()) return -1;
  const int grid = n_items < sm_count() ? n_items : sm_count();
  CUtensorMap vmap;
  if (cbx::make_kv_rows_map(&vmap, kv_layer)) return -1;
  cudaError_t e = launch_pdl(paged_attn_prefill_kernel, dim3(grid), dim3(kThreads), kSmemBytes, (cudaStream_t)stream,
                             vmap, (const bf16*)qkv, (const uint8_t*)kv_layer, block_table, bt_stride, (const int4*)items,
                             n_items, (bf16*)out);
  CBX_REQUIRE(e == cudaSuccess, "paged_attn_prefill launch failed: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}
