// Shared device/host helpers for the b200llm C-ABI library (sm_100a only).
//
// Everything here is hand-written inline PTX for Blackwell: mbarrier, TMA
// (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld / fences) plus the
// 128-bit vector load/store and reduction helpers used by the bandwidth kernels.
#pragma once
#include <cstdlib>

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <utility>

#include "../../include/b200llm.h"

// ----------------------------------------------------------------------------
// host-side error plumbing (b200llm.cu owns the storage)
// ----------------------------------------------------------------------------
int b200_set_error(int code, const char* fmt, ...);

#define B200_CHECK_ARG(cond, ...)                                   \
  do {                                                              \
    if (!(cond)) return b200_set_error(B200_ERR_ARG, __VA_ARGS__);  \
  } while (0)

#define B200_CHECK_CUDA(expr)                                                      \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess)                                                         \
      return b200_set_error((int)_e, "%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

#define B200_CHECK_LAUNCH() B200_CHECK_CUDA(cudaGetLastError())

static inline bool b200_aligned16(const void* p) { return (((uintptr_t)p) & 15u) == 0; }

// every launch through the library is counted so bench.py can report gpu_launches
void b200_count_launch(int n = 1);

int b200_num_sms();

#ifdef __CUDACC__
// Launch with the programmatic-stream-serialization attribute (see pdl_wait in the device section).
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                   cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
// The same for the kernels around the GEMMs (attention, norms): B200_PDL_WIDE=0 launches them the plain way (A/B runs).
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_pdl_wide(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                        cudaStream_t st, Args&&... args) {
  static const bool on = [] { const char* e = getenv("B200_PDL_WIDE"); return !(e && e[0] == '0'); }();
  if (on) return b200_launch_pdl(kernel, grid, block, smem, st, std::forward<Args>(args)...);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#endif

// Build a 2-D bf16 TMA descriptor for a row-major [rows, cols] matrix with row pitch
// `ld` elements; the box is [box_rows, box_cols] and smem is 128B-swizzled.
int b200_make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                           uint64_t ld, uint32_t box_rows, uint32_t box_cols);

// ----------------------------------------------------------------------------
// device helpers
// ----------------------------------------------------------------------------
#ifdef __CUDACC__

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier ----------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---- TMA ---------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)m) : "memory");
}
// 2-D tile load: c0 = innermost (column) coordinate, c1 = row coordinate.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"((uint64_t)m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// 2-D tile store shared -> global (bulk async-group completion): c0 = column, c1 = row coordinate
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"((uint64_t)m),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed stores have finished READING their shared-memory source (it may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// all committed stores are complete
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// 1-D bulk copy global -> shared (no descriptor); size and both addresses multiples of 16 bytes
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// explicit shared-space 128-bit accesses (pointers derived from an aligned-up dynamic smem base
// lose their address space and would otherwise compile to generic LD/ST)
__device__ __forceinline__ void sts_v4(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x for a pair of arguments on the FMA pipe instead of the MUFU (whose 16 results / clk / SM bound the attention
// kernels at head_dim 64): x = n + f with n = round(x) taken from the low mantissa bits of x + 1.5 * 2^23, |f| <= 0.5,
// 2^f by a degree-3 minimax polynomial (relative error 7.5e-5, far below the bf16 rounding the result gets), n added to
// the exponent field. Arguments below -126 (masked scores) give ~2^-126.
__device__ __forceinline__ unsigned long long pack_f2(float a, float b);
__device__ __forceinline__ void unpack_f2(unsigned long long v, float& a, float& b);
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c);
__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b);
__device__ __forceinline__ void exp2_poly_pair(float a0, float a1, float& p0, float& p1) {
  constexpr float MAGIC = 12582912.f;
  a0 = fmaxf(a0, -126.f);
  a1 = fmaxf(a1, -126.f);
  const unsigned long long x2 = pack_f2(a0, a1);
  const unsigned long long t2 = fadd2(x2, pack_f2(MAGIC, MAGIC));
  const unsigned long long r2 = fadd2(t2, pack_f2(-MAGIC, -MAGIC));
  const unsigned long long f2 = ffma2(r2, pack_f2(-1.f, -1.f), x2);
  unsigned long long q2 = ffma2(f2, pack_f2(0.055171407759189606f, 0.055171407759189606f), pack_f2(0.24261075258255005f, 0.24261075258255005f));
  q2 = ffma2(q2, f2, pack_f2(0.6932609677314758f, 0.6932609677314758f));
  q2 = ffma2(q2, f2, pack_f2(0.9999281167984009f, 0.9999281167984009f));
  float q0, q1, t0, t1;
  unpack_f2(q2, q0, q1);
  unpack_f2(t2, t0, t1);
  p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
  p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}
// Which column pairs of every 8 take the polynomial (bit i of the mask = pair i mod 8), per kernel; 0 = all on the MUFU.
// Measured (B4 T2048 h32 kv8 hd64 / B1 T4096 h32 kv8 hd128, masks 0x24, 0x4A, 0x55 in all three kernels): at head_dim 64
// every pattern is slower (forward 129 -> 137 / 137 / 145 us, backward 333 -> 345 / 347 / 353 us: those kernels are bound by
// issue slots and barrier round trips, not by the MUFU, and the polynomial costs 11 issue slots per pair against 3); at
// head_dim 128 the forward gains (203 -> 197 / 194 / 197 us), the backward does not. Hence the defaults.
#ifndef B200_ATTN_POLY_FWD
#define B200_ATTN_POLY_FWD 0x00
#endif
#ifndef B200_ATTN_POLY_FWD_HD128
#define B200_ATTN_POLY_FWD_HD128 0x4A
#endif
#ifndef B200_ATTN_POLY_DQ
#define B200_ATTN_POLY_DQ 0x00
#endif
#ifndef B200_ATTN_POLY_DKV
#define B200_ATTN_POLY_DKV 0x00
#endif

// Blackwell packed fp32 pairs (FFMA2 / FADD2 / FMUL2) and the 3-input maximum (FMNMX3): half the issue
// slots of the scalar forms in the softmax loops, which are issue-bound rather than MUFU-bound.
__device__ __forceinline__ unsigned long long pack_f2(float a, float b) {
  unsigned long long v;
  asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a), "f"(b));
  return v;
}
__device__ __forceinline__ void unpack_f2(unsigned long long v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ unsigned long long fmul2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// ---- tcgen05 -----------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// whole warp; writes the TMEM base address to *smem_slot
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_slot)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; single thread issues.
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (M rows = TMEM lanes, K bf16 elements packed two per 32-bit
// column, K-major) is read from tensor memory instead of shared memory.
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued MMAs of this thread retire
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
          smem_u32(bar))
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = lane/row)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM, same 32 lanes x 32 columns shape
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]),
      "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]),
      "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
// 32 lanes x 16 consecutive columns
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// make generic-proxy shared-memory writes visible to the async proxy (TMA / tcgen05.mma reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128B-swizzled shared-memory operand descriptor (rows at 128 B pitch, 8-row
// groups 1024 B apart). Field layout as in the PTX ISA "matrix descriptor" table:
// [0,14) addr>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) swizzle=2.
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;            // LBO (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;  // SBO = 1024 B between 8-row groups
  d |= (uint64_t)1 << 46;            // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}
// MN-major, 128B-swizzled operand: each K row holds 64 contiguous MN elements (128 B),
// 8 K-rows form one 1024 B swizzle atom (SBO), further 64-wide MN panels are `lbo_bytes` apart.
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(uint32_t m, uint32_t n, uint32_t a_mn_major,
                                                       uint32_t b_mn_major) {
  return (1u << 4)                 // D format = F32
         | (1u << 7)               // A format = BF16
         | (1u << 10)              // B format = BF16
         | (a_mn_major << 15)      // A major
         | (b_mn_major << 16)      // B major
         | ((n >> 3) << 17)        // N / 8
         | ((m >> 4) << 24);       // M / 16
}

// ---- programmatic dependent launch -------------------------------------------------------------
// A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start while its
// predecessor in the stream is still draining; everything before pdl_wait() (barrier init, TMEM
// allocation, descriptor prefetch) overlaps that tail, and nothing after it can observe stale data
// because griddepcontrol.wait returns only when the prerequisite grids have completed and flushed.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---- CTA-pair (cta_group::2) variants ----------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the same-offset mbarrier of CTA `rank` of this cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(rank)
      : "memory");
}
// TMA load issued by either CTA of a pair; the bytes are credited to the LEADER CTA's mbarrier
// (peer bit 24 of the shared::cluster address cleared).
__device__ __forceinline__ void tma_load_2d_2cta(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                                 int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"((uint64_t)m), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_slot)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
// issued by the leader CTA only: M = 256 across the pair, each CTA supplies its A rows and half of B
__device__ __forceinline__ void umma_bf16_ss_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                                  uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the same-offset mbarrier in both CTAs of the pair once prior MMAs retire
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}

// ---- counter-based dropout (K10) ----------------------------------------------------------
// One 32-bit hash of (seed, element index) per PAIR of elements; each 16-bit half is compared with
// thr16 = round(p * 65536): keep iff half >= thr16. Stateless, so the backward kernels regenerate
// exactly the mask the forward kernels used (nothing is stored).
// One 64-bit multiply folds the (pair index, seed) into 32 bits, a two-multiply 32-bit finisher mixes them: ~14
// integer instructions per pair. (The first version ran the whole splitmix64 finisher — two 64-bit multiplies and three
// 64-bit xor-shifts, ~28 instructions — per call, and the attention dK/dV kernel needs one call per ELEMENT because its
// threads walk down a column of the mask: the dropout variants of the attention kernels ran 2.5x slower than the plain
// ones.) The seed-only part is loop-invariant and hoisted by the compiler.
__host__ __device__ __forceinline__ uint32_t dropout_hash(unsigned long long pair_index, unsigned long long seed) {
  const unsigned long long sk = (seed ^ (seed >> 29)) * 0xBF58476D1CE4E5B9ull + 0x632BE59BD9B4E019ull;
  const unsigned long long z = (pair_index ^ sk) * 0x9E3779B97F4A7C15ull;
  uint32_t x = (uint32_t)(z >> 32) ^ (uint32_t)z;
  x ^= x >> 15;
  x *= 0x2C1B3C6Du;
  x ^= x >> 12;
  x *= 0x297A2D39u;
  x ^= x >> 15;
  return x;
}
// keep flags of elements 2*pair_index and 2*pair_index+1
__host__ __device__ __forceinline__ void dropout_keep2(unsigned long long pair_index, unsigned long long seed,
                                                       uint32_t thr16, bool& k0, bool& k1) {
  const uint32_t h = dropout_hash(pair_index, seed);
  k0 = (h & 0xFFFFu) >= thr16;
  k1 = (h >> 16) >= thr16;
}
__host__ __device__ __forceinline__ bool dropout_keep(unsigned long long index, unsigned long long seed, uint32_t thr16) {
  const uint32_t h = dropout_hash(index >> 1, seed);
  return ((index & 1) ? (h >> 16) : (h & 0xFFFFu)) >= thr16;
}

// ---- vector memory + math helpers ---------------------------------------------
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// 256-bit global store (sm_100: STG.E.ENL2.256): two uint4 = 16 bf16 to a 32-byte aligned address — a whole 32-byte
// sector per thread and instruction. Epilogues whose lanes each write their own row issue one sector write per lane
// either way; with 16-byte stores every sector took two instructions and arrived at L2 in halves.
__device__ __forceinline__ void st_v8(void* p, uint4 a, uint4 b) {
  asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(((unsigned long long)a.y << 32) | a.x),
               "l"(((unsigned long long)a.w << 32) | a.z), "l"(((unsigned long long)b.y << 32) | b.x),
               "l"(((unsigned long long)b.w << 32) | b.z)
               : "memory");
}
__device__ __forceinline__ uint4 ld_v4(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st_v4(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

__device__ __forceinline__ float2 bf2_to_f2(uint32_t u) {
  bf162 h = *reinterpret_cast<bf162*>(&u);
  return __bfloat1622float2(h);
}
__device__ __forceinline__ uint32_t f2_to_bf2(float a, float b) {
  bf162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_round(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}
// unpack 8 bf16 (one uint4) to 8 floats and back
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  float2 a = bf2_to_f2(v.x), b = bf2_to_f2(v.y), c = bf2_to_f2(v.z), d = bf2_to_f2(v.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 v;
  v.x = f2_to_bf2(f[0], f[1]); v.y = f2_to_bf2(f[2], f[3]);
  v.z = f2_to_bf2(f[4], f[5]); v.w = f2_to_bf2(f[6], f[7]);
  return v;
}

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
// block-wide sum; `red` is >= 33 floats of shared memory; all threads get the result
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : -INFINITY;
  t = warp_max(t);
  return t;
}

#endif  // __CUDACC__
