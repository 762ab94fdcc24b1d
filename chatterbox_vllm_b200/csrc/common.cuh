// common.cuh — shared device helpers for libcbx_t3.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/cbx_t3.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libcbx_t3 is written for sm_100a only"
#endif

namespace cbx {

// ---- error plumbing ---------------------------------------------------------------
void set_error(const char* fmt, ...);
int  count_launch();                       // ++ global launch counter; returns 0
// SMs of the current device (queried once), capped at kMaxSms: the persistent kernels launch one CTA per SM and the
// attention page-range plan is laid out for at most kMaxSms ranges (B200: 148)
constexpr int kMaxSms = 148;
// Copies of the fused MLP block's down-proj accumulator (k_mlp.cu: CTA b adds into copy b % kMlpAccCopies; k_block.cu's
// FOLD prologue and acc_finalize add the copies).  L2 serialises reductions per address: with one copy the 148 CTAs' bulk
// reductions cost ~0.5 us per launch at b = 1 (measured with unread copies: step 0.512 -> 0.498 ms) — but READING four
// copies in the next kernel's prologue (70 KB per CTA through one SM's L2 port) costs more than that: 0.542 ms.  So: 1.
constexpr int kMlpAccCopies = 1;
int  sm_count();
#define CBX_CHECK_LAUNCH(name)                                                   \
  do {                                                                           \
    cudaError_t e__ = cudaGetLastError();                                        \
    if (e__ != cudaSuccess) {                                                    \
      cbx::set_error("%s launch failed: %s", name, cudaGetErrorString(e__));     \
      return -2;                                                                 \
    }                                                                            \
    cbx::count_launch();                                                         \
  } while (0)
#define CBX_REQUIRE(cond, ...)                                                   \
  do {                                                                           \
    if (!(cond)) { cbx::set_error(__VA_ARGS__); return -1; }                     \
  } while (0)

typedef __nv_bfloat16 bf16;

// ---- programmatic dependent launch (PDL) ---------------------------------------------
// Every kernel of the decode step is launched with programmatic stream serialisation: it may
// start while its predecessor is still running, prefetches what does not depend on the
// predecessor (weights), then pdl_wait()s for the predecessor's memory.  Every kernel calls
// pdl_wait() before touching activations, which keeps completion transitive along the chain.
// Both instructions are no-ops when the launch did not carry the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
bool pdl_enabled();                        // env CBX_PDL != "0"
// nvcc treats loads through `const T* __restrict__` as invariant and may hoist them ABOVE
// griddepcontrol.wait despite its memory clobber (seen in SASS: LDG.E.CONSTANT before ACQBULK), i.e.
// read what the previous kernel has not written yet.  A pointer whose first load does not depend on
// another post-wait load must be laundered after the wait: p = pdl_acquire(p).
// tools/audit_pdl_hoist.py lists every load that ended up before the wait.
// A second PDL precaution, in the cache rather than the compiler: a load through a `const T* __restrict__`
// parameter compiles to LDG.E.CONSTANT — the non-coherent read-only path, which PTX only allows for data
// that is immutable for the kernel's LIFETIME, and a PDL kernel's lifetime starts while its predecessor is
// still writing.  Rule: pointers to anything an earlier kernel of the chain writes (activations, kv_len /
// row_seq / pos / slot, partials, logits) are NOT __restrict__; only weights and tables are.
template <typename T>
__device__ __forceinline__ const T* pdl_acquire(const T* p) {
  asm volatile("" : "+l"(p)::"memory");
  return p;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster_xy(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                         int cluster_x, int cluster_y, Args... args);
// same, with a thread-block cluster of `cluster_x` CTAs along grid.x
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                      int cluster_x, Args... args) {
  return launch_pdl_cluster_xy(kern, grid, block, smem, s, cluster_x, 1, args...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster_xy(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                         int cluster_x, int cluster_y, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster_x; at[0].val.clusterDim.y = cluster_y; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- bf16 pack / unpack (bf16 -> fp32 is a 16-bit shift: one ALU op per element) ----
__device__ __forceinline__ float bf_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// streaming 128-bit global load that does not pollute L1 (weights are read once)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// L2 prefetch of this CTA's share of [base, base+bytes): thread 0 of every CTA issues bulk
// prefetches (SASS UBLKPF) for a 1/gridDim slice.  Used to pull the NEXT kernel's weights into the
// 126 MB L2 while the current kernel runs, so the weight stream does not stop at kernel boundaries.
__device__ __forceinline__ void l2_prefetch_slice(const void* base, size_t bytes) {
  if (base == nullptr || threadIdx.x != 0) return;
  const size_t n_cta = (size_t)gridDim.x * gridDim.y * gridDim.z;
  const size_t cta = blockIdx.x + (size_t)gridDim.x * (blockIdx.y + (size_t)gridDim.y * blockIdx.z);
  size_t per = ((bytes + n_cta - 1) / n_cta + 127) & ~(size_t)127;
  size_t off = cta * per;
  if (off >= bytes) return;
  if (off + per > bytes) per = (bytes - off) & ~(size_t)15;
  const char* p = reinterpret_cast<const char*>(base) + off;
  for (size_t o = 0; o < per; o += 32768) {
    const uint32_t n = (uint32_t)(per - o < 32768 ? per - o : 32768);
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p + o), "r"(n) : "memory");
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// explicit shared-space 128-bit accesses (a pointer carved out of the dynamic shared array by integer
// arithmetic is generic to the compiler: LD.E/ST.E plus a MEMBAR.ALL.CTA in front of every proxy fence)
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- mbarrier ----------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {}
}

// ---- bulk async copy global -> shared (TMA engine, 1-D; SASS: UBLKCP) ---------------
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- TMA tensor copy (2-D tile, SASS: UTMALDG) ---------------------------------------
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// ---- tcgen05 / TMEM ------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
               ::"r"(smem_u32(smem_out)), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate (kind::f16)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread complete
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(smem_u32(bar)) : "memory");
}
// ---- CTA-pair (cta_group::2) forms: one MMA spans the two CTAs of a cluster (M = 256) -------------
// issued by one warp in EACH CTA of the pair (same warp index, same smem_out offset)
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
               ::"r"(smem_u32(smem_out)), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// TMA tile into THIS CTA's shared memory; the bytes complete on the barrier at the same offset in the
// pair's leader (even) CTA: bit 24 of a shared::cluster address selects the peer inside a pair.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1) : "memory");
}
// leader CTA only: D[tmem of both CTAs] (+)= A[128 rows per CTA] * B[N/2 rows per CTA]^T
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on the barrier at this shared-memory offset in every CTA of cta_mask once the pair's MMAs complete
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(cta_mask) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (lane = TMEM lane)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp
// SmemDescriptor): start>>4 | LBO(16B, ignored for swizzled K-major) | SBO = 1024 B
// (8 rows x 128 B) | version 1 | layout SWIZZLE_128B (2).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// ---- FP8-e4m3 weight bytes -> f16 without the conversion pipe --------------------------------------
// An e4m3 byte s.eeee.mmm placed at bits 14..7 of a 16-bit lane reads as the f16 number value/256 — for
// normals, subnormals and zero alike (both formats keep 3 / 10 mantissa bits under the same hidden-bit
// rule; the exponent bias differs by 8).  The sign then sits at bit 14: adding (x & 0x4000) moves it to
// bit 15.  PRMT + shift + AND + ADD per two elements, all on the integer/FMA pipes; cvt.f16x2.e4m3x2 runs
// on the 16-lane conversion unit: the first converter built on it left the FP8 prefill 2x slower than bf16.
// Consumers fold the factor 256 into the per-output-feature scale.
__device__ __forceinline__ void e4m3x4_to_f16x2_div256(uint32_t w, uint32_t& lo, uint32_t& hi) {
  const uint32_t a = __byte_perm(w, 0, 0x4140) << 7;    // halves: byte0, byte1
  const uint32_t b = __byte_perm(w, 0, 0x4342) << 7;    // halves: byte2, byte3
  lo = a + (a & 0x40004000u);
  hi = b + (b & 0x40004000u);
}
// Same four bytes as two bf16x2 words (still value / 256, exact).  tcgen05.mma kind::f16 does NOT take a bf16 A
// next to an f16 B — the descriptor encodes the two formats separately and CUTLASS's traits only ask for two
// 16-bit types, but the B200 raises cudaErrorIllegalInstruction — so the f16 lanes are widened to fp32
// (HADD2.F32, exact, subnormals included) and the bf16 is the top half of each fp32 word (3 mantissa bits: no
// rounding).  14 integer / half-pipe ops per four weights, none on the conversion unit.
__device__ __forceinline__ void e4m3x4_to_bf16x2_div256(uint32_t w, uint32_t& lo, uint32_t& hi) {
  uint32_t h0, h1;
  e4m3x4_to_f16x2_div256(w, h0, h1);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&h0));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&h1));
  lo = __byte_perm(__float_as_uint(a.x), __float_as_uint(a.y), 0x7632);
  hi = __byte_perm(__float_as_uint(b.x), __float_as_uint(b.y), 0x7632);
}
// Instruction descriptor for kind::f16: C=f32, A=B=bf16, both K-major, M x N.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

}  // namespace cbx
