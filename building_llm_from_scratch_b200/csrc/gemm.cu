// K1 — bf16 GEMM on tcgen05 / TMEM fed by TMA (sm_100a).
//
//   C[M,N] = alpha * ( opA1·opB1^T + opA2·opB2^T ) (+ bias) (gelu) (+ R)
//
// One persistent CTA per SM, 256 threads, warp-specialised:
//   warp 0   : TMA producer (one lane)  — global -> 128B-swizzled smem ring (192 KB)
//   warp 1   : MMA issuer  (one lane)   — tcgen05.mma 128 x BN x 16, fp32 accumulators in TMEM
//   warp 2   : TMEM allocator / deallocator
//   warps 4-7: epilogue — tcgen05.ld TMEM -> registers -> bias/GELU/residual -> global
// The accumulator is double buffered in TMEM (2 x BN columns) so the epilogue of tile i
// overlaps the main loop of tile i+1. The second operand pair is the rank-r LoRA update
// (lora.py:26): extra K blocks accumulated into the same TMEM tile. Either operand of either
// pair may be K-major ([rows, K], K contiguous) or MN-major ([K, rows], rows contiguous);
// MN-major tiles are fetched as 64-wide panels and described to the MMA with the MN-major
// shared-memory descriptor, so backward GEMMs (dX = dY·W, dW = dY^T·X) need no transposes.
//
// Roofline: tensor pipe. Algorithmic FLOPs per launch = 2·M·N·(K1+K2).
#include <atomic>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_BYTES = BM * BK * 2;  // 16 KB
constexpr int RING_BYTES = 192 * 1024;

enum : int {
  EP_BIAS = 1,
  EP_GELU = 2,
  EP_RESIDUAL = 4,
  EP_OUT_F32 = 8,
  EP_ATOMIC = 16,
  EP_TRANS_OUT = 32,
  EP_SWIGLU_BWD = 64,   // pair kernel: the accumulator is dh; write (da | db) of SwiGLU into C, reading (a | b) from R
  EP_SWIGLU_FWD = 128,  // pair kernel: tile = 128 columns of a | the same 128 of b; write (a | b) into C and h into R
  EP_GELU_BWD = 256,    // pair kernel, staged epilogue: C = acc * gelu'(R) with R the saved pre-activation (GPT2.py:60 under autograd)
  EP_DROPOUT = 512,     // pair kernel, staged epilogue: C = R + dropout(acc + bias) (GPT2.py:85,87: x + drop_shortcut(...))
  EP_ROPE = 1024,       // pair kernel, per-thread epilogue: rotate-half RoPE on the output columns < rope_cols (q | k heads)
};

struct GemmParams {
  int M, N;
  int kb1, kb2;  // 64-wide K blocks of the main / LoRA product
  int a1_mn, b1_mn, a2_mn, b2_mn;
  int splits;
  int m_tiles, n_tiles;
  int n_fastest;   // pair kernel: walk the N tiles of an M tile-row back to back (A row stays in L2)
  // pair kernel, tail splitting: tiles >= tail_start are each cut into tail_splits K ranges so the last,
  // partly filled wave of tiles keeps every CTA pair busy; partial sums meet in tail_ws (one fp32 slot
  // per tile and split), tail_cnt counts the arrivals per (tile, CTA of the pair), zero between launches
  int tail_start, tail_splits;
  float* tail_ws;
  int* tail_cnt;
  // pair kernel, packed LoRA projections: output columns are nseg consecutive segments (q | k | v); a
  // tile starting in segment s reads its A2 K blocks at column offset s * a2_seg_stride (t_cat holds
  // one [M, r] block per segment). seg_start[i] = first column of segment i + 1; unused entries INT_MAX.
  int seg_start[3];
  int a2_seg_stride;
  int flags;
  float alpha;
  const bf16* bias;
  const bf16* R;
  long long ldr;
  void* C;
  long long ldc;
  bf16* Z;              // EP_GELU: also store the pre-activation here (bf16) for the backward pass
  long long ldz;
  // EP_DROPOUT: keep(i) is the counter-based function of (seed + *seed_dev, i = row * N + col) the stand-alone
  // dropout kernel uses, so fused and unfused runs draw the same mask
  uint32_t drop_thr16;
  float drop_inv_keep;
  unsigned long long drop_seed;
  const unsigned long long* drop_seed_dev;
  // EP_ROPE (compute_rope, Models/Llama/common_components.py:6-35, applied to the q | k heads of a packed q|k|v
  // projection): row m is position m % rope_T; tables [>= rope_T, rope_hd] bf16 whose two halves are equal
  const bf16* rope_cos;
  const bf16* rope_sin;
  int rope_T, rope_hd, rope_cols;
};

// barrier over the 4 epilogue warps (threads 128..255) of a CTA
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// erf by Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, far inside bf16's 2^-9): one reciprocal, one exponential and
// six FMAs instead of erff's ~35 instructions — the GELU epilogue of a K = 1600 GEMM is otherwise longer than its main loop.
__device__ __forceinline__ float erf_as(float x) {
  const float ax = fabsf(x);
  const float t = __fdividef(1.f, fmaf(0.3275911f, ax, 1.f));
  const float poly = t * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.061405429f, -1.453152027f), 1.421413741f), -0.284496736f), 0.254829592f);
  const float e = 1.f - poly * __expf(-ax * ax);
  return copysignf(e, x);
}
__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erf_as(x * 0.70710678118654752440f));
}
// d/dx of the erf GELU: Phi(x) + x * phi(x)
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erf_as(x * 0.70710678118654752440f));
  return cdf + x * 0.39894228040143267794f * __expf(-0.5f * x * x);
}

// Epilogue for 32 consecutive accumulator columns of one output row (one thread):
// alpha, bias, GELU, residual, then bf16 / fp32 / atomic / transposed store with tail predicates.
// Residual values for one 32-column chunk, fetched one chunk ahead of their use so the L2 latency of
// the residual read does not serialise with the TMEM load / convert / store of every chunk (short-K
// GEMMs with a residual epilogue were epilogue-bound: 23 % tensor-pipe activity at K = 512).
// rotate-half RoPE on one (first-half, second-half) pair of 32-column accumulator chunks of a head, with the
// reference's bf16 roundings (the linear's output, each product and the sum are bf16 there):
//   o1 = x1 cos - x2 sin,  o2 = x2 cos + x1 sin     (cos / sin: this row's 32 table entries for these columns)
__device__ __forceinline__ void rope_chunk_pair(uint32_t* r1, uint32_t* r2, const bf16* __restrict__ cosr,
                                                const bf16* __restrict__ sinr, float alpha) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float c[8], s[8];
    unpack8(ld_v4(cosr + q * 8), c);
    unpack8(ld_v4(sinr + q * 8), s);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x1 = bf16_round(__uint_as_float(r1[q * 8 + j]) * alpha);
      const float x2 = bf16_round(__uint_as_float(r2[q * 8 + j]) * alpha);
      r1[q * 8 + j] = __float_as_uint(bf16_round(x1 * c[j]) + bf16_round(-x2 * s[j]));
      r2[q * 8 + j] = __float_as_uint(bf16_round(x2 * c[j]) + bf16_round(x1 * s[j]));
    }
  }
}

struct ResChunk {
  uint4 v[4];
  bool vec;   // v[] holds the 32 residual values (fast path)
};
__device__ __forceinline__ void store_chunk(const GemmParams& p, long long row, bool row_ok, long long col0,
                                            const uint32_t* r, const struct ResChunk& rc);
__device__ __forceinline__ ResChunk load_residual(const GemmParams& p, long long row, bool row_ok, long long col0) {
  ResChunk rc;
  rc.vec = false;
  if ((p.flags & EP_RESIDUAL) && !(p.flags & EP_TRANS_OUT) && row_ok && col0 + 32 <= p.N && (p.ldr & 7) == 0) {
    const bf16* R = p.R + row * p.ldr + col0;
#pragma unroll
    for (int q = 0; q < 4; ++q) rc.v[q] = ld_v4(R + q * 8);
    rc.vec = true;
  }
  return rc;
}

__device__ __forceinline__ void store_chunk(const GemmParams& p, long long row, bool row_ok,
                                            long long col0, const uint32_t* r, const ResChunk& rc, const float alpha) {
  if (col0 >= p.N) return;
  const bool full_cols = col0 + 32 <= p.N;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * alpha;
  if (p.flags & EP_BIAS) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col0 + j < p.N) v[j] += __bfloat162float(p.bias[col0 + j]);
  }
  if (p.flags & EP_GELU) {
    if (p.Z != nullptr && !(p.flags & EP_TRANS_OUT)) {
      // keep the pre-activation for backward (bf16, as the separate linear would have written it) and apply the
      // GELU to that rounded value, like nn.GELU on the linear's bf16 output
      if (row_ok) {
        bf16* Zp = p.Z + row * p.ldz + col0;
        if (full_cols && (p.ldz & 15) == 0 && (((uintptr_t)p.Z) & 31) == 0) {
          st_v8(Zp, pack8(&v[0]), pack8(&v[8]));
          st_v8(Zp + 16, pack8(&v[16]), pack8(&v[24]));
        } else if (full_cols && (p.ldz & 7) == 0) {
#pragma unroll
          for (int q = 0; q < 4; ++q) st_v4(Zp + q * 8, pack8(&v[q * 8]));
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.N) Zp[j] = __float2bfloat16_rn(v[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = gelu_erf(bf16_round(v[j]));
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
    }
  }
  if (p.flags & EP_TRANS_OUT) {
    // C is [N, M]-shaped in memory: element (row, col) goes to C[col * ldc + row]
    if (row_ok) {
      if (p.flags & EP_OUT_F32) {
        float* C = (float*)p.C;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) {
            if (p.flags & EP_ATOMIC) atomicAdd(&C[(col0 + j) * p.ldc + row], v[j]);
            else C[(col0 + j) * p.ldc + row] = v[j];
          }
      } else {
        bf16* C = (bf16*)p.C;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) C[(col0 + j) * p.ldc + row] = __float2bfloat16_rn(v[j]);
      }
    }
    return;
  }
  if (!row_ok) return;
  if (p.flags & EP_RESIDUAL) {
    const bf16* R = p.R + row * p.ldr + col0;
    if (rc.vec) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float f[8];
        unpack8(rc.v[q], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[q * 8 + j] += f[j];
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) v[j] += __bfloat162float(R[j]);
    }
  }
  if (p.flags & EP_OUT_F32) {
    float* C = (float*)p.C + row * p.ldc + col0;
    if (p.flags & EP_ATOMIC) {
      if (full_cols && (p.ldc & 3) == 0 && (((uintptr_t)C) & 15) == 0) {
        // one REDG.F32x4 per four columns: a thread's 32 columns are contiguous, and the RED issue rate
        // (~1 per clock per SM) is what bounds this epilogue
#pragma unroll
        for (int q = 0; q < 8; ++q)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(C + q * 4), "f"(v[q * 4]), "f"(v[q * 4 + 1]),
                       "f"(v[q * 4 + 2]), "f"(v[q * 4 + 3])
                       : "memory");
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) atomicAdd(&C[j], v[j]);
      }
    } else if (full_cols && (p.ldc & 3) == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<float4*>(C + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) C[j] = v[j];
    }
  } else {
    bf16* C = (bf16*)p.C + row * p.ldc + col0;
    if (full_cols && (p.ldc & 15) == 0 && (((uintptr_t)p.C) & 31) == 0) {
      st_v8(C, pack8(&v[0]), pack8(&v[8]));
      st_v8(C + 16, pack8(&v[16]), pack8(&v[24]));
    } else if (full_cols && (p.ldc & 7) == 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q) st_v4(C + q * 8, pack8(&v[q * 8]));
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) C[j] = __float2bfloat16_rn(v[j]);
    }
  }
}

__device__ __forceinline__ void store_chunk(const GemmParams& p, long long row, bool row_ok, long long col0,
                                            const uint32_t* r, const ResChunk& rc) {
  store_chunk(p, row, row_ok, col0, r, rc, p.alpha);
}

template <int BN>
__global__ void __launch_bounds__(256, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                    const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                    const GemmParams p) {
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGES = RING_BYTES / STAGE_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    if (p.kb2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();   // the next kernel may start its own prologue on SMs this grid has left
  pdl_wait();                // inputs written by the previous kernel are complete and visible from here

  const int total_kb = p.kb1 + p.kb2;
  const int num_tiles = p.m_tiles * p.n_tiles;
  const int num_work = num_tiles * p.splits;

  if (warp == 0) {
    if (lane == 0) {
      // ---------------- TMA producer ----------------
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int split = w / num_tiles, tile = w % num_tiles;
        const int m_blk = tile % p.m_tiles, n_blk = tile / p.m_tiles;
        const int kb_begin = (int)((long long)split * total_kb / p.splits);
        const int kb_end = (int)((long long)(split + 1) * total_kb / p.splits);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], STAGE_BYTES);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const bool ext = kb >= p.kb1;
          const int kk = (ext ? kb - p.kb1 : kb) * BK;
          const CUtensorMap* ma = ext ? &tmA2 : &tmA1;
          const CUtensorMap* mb = ext ? &tmB2 : &tmB1;
          const int a_mn = ext ? p.a2_mn : p.a1_mn;
          const int b_mn = ext ? p.b2_mn : p.b1_mn;
          if (!a_mn) {
            tma_load_2d(sa, ma, &full[stage], kk, m_blk * BM);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)
              tma_load_2d(sa + i * 8192, ma, &full[stage], m_blk * BM + i * 64, kk);
          }
          if (!b_mn) {
            tma_load_2d(sb, mb, &full[stage], kk, n_blk * BN);
          } else {
#pragma unroll
            for (int i = 0; i < BN / 64; ++i)
              tma_load_2d(sb + i * 8192, mb, &full[stage], n_blk * BN + i * 64, kk);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ---------------- MMA issuer ----------------
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++it) {
        const int split = w / num_tiles;
        const int kb_begin = (int)((long long)split * total_kb / p.splits);
        const int kb_end = (int)((long long)(split + 1) * total_kb / p.splits);
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tempty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
          const bool ext = kb >= p.kb1;
          const int a_mn = ext ? p.a2_mn : p.a1_mn;
          const int b_mn = ext ? p.b2_mn : p.b1_mn;
          const uint32_t idesc = umma_idesc_bf16(BM, BN, (uint32_t)a_mn, (uint32_t)b_mn);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = a_mn ? umma_desc_mn128(sa + k * 2048, 8192) : umma_desc_k128(sa + k * 32);
            const uint64_t bd = b_mn ? umma_desc_mn128(sb + k * 2048, 8192) : umma_desc_k128(sb + k * 32);
            umma_bf16_ss(d_tmem, ad, bd, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tfull[as]);
      }
    }
  } else if (warp >= 4) {
    // ---------------- epilogue ----------------
    const int ew = warp - 4;  // == warp % 4: the TMEM lane quadrant this warp may read
    int it = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++it) {
      const int tile = w % num_tiles;
      const int m_blk = tile % p.m_tiles, n_blk = tile / p.m_tiles;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const long long row = (long long)m_blk * BM + ew * 32 + lane;
      const bool row_ok = row < p.M;
      ResChunk rc = load_residual(p, row, row_ok, (long long)n_blk * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * BN + c * 32);
        tmem_ld_32x32(taddr, r);
        const ResChunk cur = rc;
        if (c + 1 < BN / 32) rc = load_residual(p, row, row_ok, (long long)n_blk * BN + (c + 1) * 32);
        tmem_ld_wait();
        store_chunk(p, row, row_ok, (long long)n_blk * BN + c * 32, r, cur);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN>
int launch_bn(const CUtensorMap* maps, const GemmParams& p, int grid, cudaStream_t st) {
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGES = RING_BYTES / STAGE_BYTES;
  constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align slack*/ + (2 * STAGES + 4) * 8 + 16;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_set[dev] = true;
  }
  B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_kernel<BN>, dim3(grid), dim3(256), SMEM, st, maps[0], maps[1], maps[2],
                                  maps[3], p));
  b200_count_launch();
  return 0;
}


// ======================================================================================
// CTA-pair variant: a cluster of two CTAs on one TPC computes a 256 x 256 tile with
// tcgen05.mma.cta_group::2. Each CTA stages its own 128 A rows and HALF of the B tile, so the
// shared-memory traffic per SM (TMA writes + MMA reads) drops from 192 to 128 B/clk at full MMA
// rate — the single-CTA kernel above is capped near 2/3 of the tensor peak by exactly that.
// Only the leader CTA issues MMAs; both CTAs run a producer, an epilogue over their own 128
// TMEM lanes, and arrive remotely on the leader's accumulator-empty barrier.
// ======================================================================================
// SwiGLU backward in the epilogue of dh = dY·W3 (+ u·A3^T) (MODE 1 of the pair kernel): (a | b) come from the
// packed forward activations R [M, 2F] (b at column offset F = p.N), (da | db) go to C [M, 2F]:
//   s = sigmoid(a), da = dh * b * s * (1 + a (1 - s)), db = dh * a * s      (common_components.py:87,123 under autograd)
// The separate kernel this replaces moved 10 bytes per element of [M, F]; dh never reaches memory now.
constexpr int P_BN = 256;                       // tile N of the pair
constexpr int P_BHALF_BYTES = (P_BN / 2) * BK * 2;   // 16 KB: this CTA's half of B
constexpr int P_STAGE_BYTES = A_BYTES + P_BHALF_BYTES;
constexpr int P_STAGES = RING_BYTES / P_STAGE_BYTES; // 6
constexpr int P_SMEM = P_STAGES * P_STAGE_BYTES + 1024 + (2 * P_STAGES + 4) * 8 + 16 + 64;
// MODE 1: the epilogue's (a | b) -> (da | db) tiles pass through one 32 KB staging buffer behind the ring (measured:
// 4 ring stages + 2 or 3 buffers 260 us, 5 + 2: 243 us, 6 + 1: 233 us — the main loop needs the ring depth more)
constexpr int P_STAGES_SWIGLU = 6;
constexpr int P_IO_TILE = 128 * 64 * 2;        // one staged [128 rows x 64 columns] bf16 tile
constexpr int P_IO_BYTES = 2 * P_IO_TILE;      // staging area behind the ring: MODE 1 one (a | b) buffer, MODE 2 two R buffers
constexpr int P_SMEM_SWIGLU = P_STAGES_SWIGLU * P_STAGE_BYTES + P_IO_BYTES + 1024 + 256;

// MODE 0: 256 threads, four epilogue warps, per-thread 16-byte global accesses (one row per lane). MODE 1
// (SwiGLU-backward epilogue) and MODE 2 (residual epilogue): 384 threads — two epilogue warps per TMEM lane
// quadrant, half of each 64-column group each — and every epilogue operand / result travels by TMA through a
// swizzled staging area, because one-row-per-lane accesses cost 32 line transactions per warp instruction and
// made the load/store unit the limit once the epilogue reads as well as writes.
template <int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((MODE == 1 || MODE == 2 || MODE == 4) ? 384 : 256, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                         const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                         const __grid_constant__ CUtensorMap tmAB, const __grid_constant__ CUtensorMap tmDAB,
                         const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr bool SWIGLU = MODE == 1;
  constexpr bool STAGED = MODE == 1 || MODE == 2;   // epilogue operands / results travel by TMA through smem
  constexpr int IO_TILES = SWIGLU ? 2 : 1;       // tiles per staged group: (a | b) or the residual
  constexpr int IO_BUFS = SWIGLU ? 1 : 2;
  constexpr int BUF_BYTES = IO_TILES * P_IO_TILE;
  uint64_t* full = (uint64_t*)(smem + (STAGED ? P_STAGES_SWIGLU * P_STAGE_BYTES + P_IO_BYTES : P_STAGES * P_STAGE_BYTES));
  uint64_t* empty = full + P_STAGES;
  uint64_t* tfull = empty + P_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
  uint64_t* in_full = tempty + 3;             // staged modes: staging buffers filled by TMA
  uint64_t* out_full = in_full + IO_BUFS;     //         ... rewritten in place with the results by the 8 epilogue warps

  // MODE 4 = MODE 0's per-thread epilogue on eight epilogue warps (two per TMEM lane quadrant, half of the tile's columns
  // each): for epilogues that are arithmetic-heavy (GELU) or write two outputs
  constexpr bool WIDE = STAGED || MODE == 4;
  constexpr int STAGES = STAGED ? P_STAGES_SWIGLU : P_STAGES;
  uint8_t* sab = smem + P_STAGES_SWIGLU * P_STAGE_BYTES;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], WIDE ? 16 : 8);   // the epilogue warps of both CTAs (leader's copy is the one used)
    }
    if (STAGED) {
      for (int i = 0; i < IO_BUFS; ++i) {
        mbar_init(&in_full[i], 1);
        mbar_init(&out_full[i], 8);
      }
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_2cta(tmem_slot, 512);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    if (p.kb2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();

  const int total_kb = p.kb1 + p.kb2;
  const int num_tiles = p.m_tiles * p.n_tiles;         // m_tiles counts 256-row tiles here
  const int num_work = p.tail_start + (num_tiles - p.tail_start) * p.tail_splits;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  __shared__ int s_last;
  // work unit -> (tile, K block range); units past tail_start are the K splits of the tail tiles
  auto unit = [&](int w, int& tile, int& kb_begin, int& kb_end) {
    if (w < p.tail_start) {
      tile = w;
      kb_begin = 0;
      kb_end = total_kb;
    } else {
      const int j = w - p.tail_start;
      tile = p.tail_start + j / p.tail_splits;
      const int sp = j % p.tail_splits;
      kb_begin = (int)((long long)sp * total_kb / p.tail_splits);
      kb_end = (int)((long long)(sp + 1) * total_kb / p.tail_splits);
    }
  };

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cluster_id; w < num_work; w += num_clusters) {
        int tile, kb_begin, kb_end;
        unit(w, tile, kb_begin, kb_end);
        const int m_blk = p.n_fastest ? tile / p.n_tiles : tile % p.m_tiles;
        const int n_blk = p.n_fastest ? tile % p.n_tiles : tile / p.m_tiles;
        const int m0 = m_blk * 256 + (int)rank * 128;
        const int n0 = n_blk * P_BN + (int)rank * 128;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full[stage], 2 * P_STAGE_BYTES);
          uint8_t* sa = smem + stage * P_STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const bool ext = kb >= p.kb1;
          const int kk = (ext ? kb - p.kb1 : kb) * BK;
          const CUtensorMap* ma = ext ? &tmA2 : &tmA1;
          const CUtensorMap* mb = ext ? &tmB2 : &tmB1;
          const int a_mn = ext ? p.a2_mn : p.a1_mn;
          const int b_mn = ext ? p.b2_mn : p.b1_mn;
          if (!a_mn) {
            const int nt0 = n_blk * P_BN;       // segment of this tile selects the A2 column block
            const int seg = ext ? (nt0 >= p.seg_start[0]) + (nt0 >= p.seg_start[1]) + (nt0 >= p.seg_start[2]) : 0;
            tma_load_2d_2cta(sa, ma, &full[stage], kk + seg * p.a2_seg_stride, m0);
          } else {
            tma_load_2d_2cta(sa, ma, &full[stage], m0, kk);
            tma_load_2d_2cta(sa + 8192, ma, &full[stage], m0 + 64, kk);
          }
          if (!b_mn) {
            tma_load_2d_2cta(sb, mb, &full[stage], kk, n0);
          } else {
            tma_load_2d_2cta(sb, mb, &full[stage], n0, kk);
            tma_load_2d_2cta(sb + 8192, mb, &full[stage], n0 + 64, kk);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = cluster_id; w < num_work; w += num_clusters, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tempty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * P_BN);
        int tile, kb_begin, kb_end;
        unit(w, tile, kb_begin, kb_end);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * P_STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
          const bool ext = kb >= p.kb1;
          const int a_mn = ext ? p.a2_mn : p.a1_mn;
          const int b_mn = ext ? p.b2_mn : p.b1_mn;
          const uint32_t idesc = umma_idesc_bf16(256, P_BN, (uint32_t)a_mn, (uint32_t)b_mn);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = a_mn ? umma_desc_mn128(sa + k * 2048, 8192) : umma_desc_k128(sa + k * 32);
            const uint64_t bd = b_mn ? umma_desc_mn128(sb + k * 2048, 8192) : umma_desc_k128(sb + k * 32);
            umma_bf16_ss_2cta(d_tmem, ad, bd, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          umma_commit_2cta(&empty[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_2cta(&tfull[as]);
      }
    }
  } else if (STAGED && warp == 3) {
    // Feeds and drains the epilogue by TMA, 64 columns x 128 rows at a time, through swizzled staging buffers:
    // MODE 1 brings in the (a | b) tiles, the epilogue warps overwrite them in place with (da | db); MODE 2
    // brings in the residual tile, overwritten with the finished output tile; this thread stores the results.
    // Per-thread global accesses of one row piece each (32 rows per warp instruction) made the SM's
    // load/store unit the bottleneck of these epilogues: the SwiGLU one took 303 us against 186 us for the
    // plain GEMM, 120 us of it the stores.
    if (lane == 0) {
      tma_prefetch_desc(&tmAB);
      tma_prefetch_desc(&tmDAB);
      int n = 0;
      int q_row[IO_BUFS] = {}, q_col[IO_BUFS] = {};
      auto drain = [&](int k) {               // results of group k: staged -> global
        const int buf = k % IO_BUFS;
        mbar_wait(&out_full[buf], (uint32_t)((k / IO_BUFS) & 1));
        tma_store_2d(&tmDAB, sab + buf * BUF_BYTES, q_col[buf], q_row[buf]);
        if (IO_TILES == 2) tma_store_2d(&tmDAB, sab + buf * BUF_BYTES + P_IO_TILE, p.N + q_col[buf], q_row[buf]);
        tma_store_commit();
      };
      for (int w = cluster_id; w < num_work; w += num_clusters) {
        if (w >= p.tail_start) continue;      // K-split tail tiles use the per-thread epilogue
        int tile, kb_begin, kb_end;
        unit(w, tile, kb_begin, kb_end);
        const int m_blk = p.n_fastest ? tile / p.n_tiles : tile % p.m_tiles;
        const int n_blk = p.n_fastest ? tile % p.n_tiles : tile / p.m_tiles;
        const int row0 = m_blk * 256 + (int)rank * 128;
        for (int g = 0; g < P_BN / 64; ++g) {
          if (n_blk * P_BN + g * 64 >= p.N) break;     // N tail (N % 64 == 0): nothing to stage beyond it
          const int buf = n % IO_BUFS;
          if (n >= IO_BUFS) {
            drain(n - IO_BUFS);
            tma_store_wait_read();            // the buffer may be refilled once the stores have read it
          }
          mbar_expect_tx(&in_full[buf], BUF_BYTES);
          const int col = n_blk * P_BN + g * 64;
          tma_load_2d(sab + buf * BUF_BYTES, &tmAB, &in_full[buf], col, row0);
          if (IO_TILES == 2) tma_load_2d(sab + buf * BUF_BYTES + P_IO_TILE, &tmAB, &in_full[buf], p.N + col, row0);
          q_row[buf] = row0;
          q_col[buf] = col;
          ++n;
        }
      }
      for (int k = (n >= IO_BUFS ? n - IO_BUFS : 0); k < n; ++k) drain(k);
      tma_store_wait_all();
    }
  } else if (warp >= 4) {
    const int ew = (warp - 4) & 3;                 // TMEM lane quadrant
    const int chalf = (warp - 4) >> 2;             // WIDE: which half of the tile's column chunks
    const int c0 = WIDE ? chalf * (P_BN / 64) : 0, c1 = WIDE ? c0 + P_BN / 64 : P_BN / 32;
    auto epi_sync = [] {
      if (WIDE) asm volatile("bar.sync 1, 256;" ::: "memory");
      else asm volatile("bar.sync 1, 128;" ::: "memory");
    };
    int it = 0;
    int nst = 0;                                   // staged 64-column groups consumed so far
    for (int w = cluster_id; w < num_work; w += num_clusters, ++it) {
      int tile, kb_begin, kb_end;
      unit(w, tile, kb_begin, kb_end);
      const int m_blk = p.n_fastest ? tile / p.n_tiles : tile % p.m_tiles;
      const int n_blk = p.n_fastest ? tile % p.n_tiles : tile / p.m_tiles;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const long long row = (long long)m_blk * 256 + (long long)rank * 128 + ew * 32 + lane;
      const bool row_ok = row < p.M;
      if (w >= p.tail_start) {
        // K-split tail tile. Every split parks its raw partial sums in its own workspace slot (plain
        // coalesced 16-byte stores: slot[(column/4)][row of the tile] float4 — RED atomics issue at ~1 per
        // clock per SM, 32 K of them per CTA cost more than the tile), then bumps the (tile, CTA-of-pair)
        // counter; the split that arrives last still holds its own sums in TMEM, adds the other slots
        // to them and runs the usual epilogue. Slots need no clearing, the counter is reset here.
        const int tt = tile - p.tail_start;
        const int sp = (w - p.tail_start) % p.tail_splits;
        const int rit = (int)rank * 128 + ew * 32 + lane;
        float4* slot0 = reinterpret_cast<float4*>(p.tail_ws) + (long long)tt * p.tail_splits * (256 * (P_BN / 4));
        float4* mine = slot0 + (long long)sp * (256 * (P_BN / 4)) + rit;
        const uint32_t tacc = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN);
#pragma unroll 1
        for (int c = c0; c < c1; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tacc + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 8; ++q)
            __stcg(mine + (c * 8 + q) * 256, make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                                         __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3])));
        }
        __threadfence();
        epi_sync();
        if (threadIdx.x == 128) {
          const int old = atomicAdd(p.tail_cnt + tt * 2 + (int)rank, 1);
          s_last = (old + 1 == p.tail_splits) ? 1 : 0;
        }
        epi_sync();
        if (s_last) {
          __threadfence();
          ResChunk none;
          none.vec = false;
#pragma unroll 1
          for (int c = c0; c < c1; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(tacc + c * 32, r);
            tmem_ld_wait();
            for (int o = 0; o < p.tail_splits; ++o) {
              if (o == sp) continue;
              const float4* other = slot0 + (long long)o * (256 * (P_BN / 4)) + rit;
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 v = __ldcg(other + (c * 8 + q) * 256);
                r[4 * q] = __float_as_uint(__uint_as_float(r[4 * q]) + v.x);
                r[4 * q + 1] = __float_as_uint(__uint_as_float(r[4 * q + 1]) + v.y);
                r[4 * q + 2] = __float_as_uint(__uint_as_float(r[4 * q + 2]) + v.z);
                r[4 * q + 3] = __float_as_uint(__uint_as_float(r[4 * q + 3]) + v.w);
              }
            }
            store_chunk(p, row, row_ok, (long long)n_blk * P_BN + c * 32, r, none);
          }
          if (threadIdx.x == 128) p.tail_cnt[tt * 2 + (int)rank] = 0;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&tempty[as], 0);
        continue;
      }
      if constexpr (SWIGLU) {
        // 64 columns at a time: this warp takes the half (32 columns) given by chalf; its thread's row of a and b
        // comes out of the swizzled staging tile (16-byte piece index XOR row % 8) and da, db go back into the
        // same places; warp 3 stores the finished tiles.
        const int rloc = ew * 32 + lane;
#pragma unroll 1
        for (int g = 0; g < P_BN / 64; ++g) {
          const int n = nst++, buf = n % IO_BUFS;
          const int c = 2 * g + chalf;
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN + c * 32), r);
          mbar_wait(&in_full[buf], (uint32_t)((n / IO_BUFS) & 1));
          const uint32_t arow = smem_u32(sab) + buf * BUF_BYTES + rloc * 128;
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t off = (uint32_t)(((chalf * 4 + q) ^ (rloc & 7)) << 4);
            const float4 fa = lds_f4(arow + off), fb = lds_f4(arow + P_IO_TILE + off);
            float a[8], b[8], da[8], db[8];
            unpack8(make_uint4(__float_as_uint(fa.x), __float_as_uint(fa.y), __float_as_uint(fa.z), __float_as_uint(fa.w)), a);
            unpack8(make_uint4(__float_as_uint(fb.x), __float_as_uint(fb.y), __float_as_uint(fb.z), __float_as_uint(fb.w)), b);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float gch = __uint_as_float(r[q * 8 + j]) * p.alpha;
              const float sg = __fdividef(1.f, 1.f + __expf(-a[j]));
              da[j] = gch * b[j] * (sg * (1.f + a[j] * (1.f - sg)));
              db[j] = gch * (a[j] * sg);
            }
            const uint4 va = pack8(da), vb = pack8(db);
            sts_v4(arow + off, va.x, va.y, va.z, va.w);
            sts_v4(arow + P_IO_TILE + off, vb.x, vb.y, vb.z, vb.w);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&out_full[buf]);
        }
      } else if constexpr (MODE == 3) {
        // SwiGLU forward: the tile's first 128 columns are a = fc1 x, the last 128 the matching b = fc2 x (the
        // weights were row-interleaved that way). Both are rounded to bf16 as the separate linears would, stored
        // into the packed (a | b) buffer, and h = silu(a) * b — with the reference's bf16 roundings — goes to R.
        const long long F = p.N >> 1;
        const long long colf = (long long)n_blk * (P_BN / 2);
        bf16* Cab = (bf16*)p.C + row * p.ldc + colf;
        bf16* H = const_cast<bf16*>(p.R) + row * p.ldr + colf;
#pragma unroll 1
        for (int k = 0; k < P_BN / 64; ++k) {
          uint32_t ra[32], rb[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN + k * 32);
          tmem_ld_32x32(taddr, ra);
          tmem_ld_32x32(taddr + P_BN / 2, rb);
          tmem_ld_wait();
          if (row_ok && colf + k * 32 + 32 <= F) {
            const bool wide = ((p.ldc | p.ldr | F) & 15) == 0 && ((((uintptr_t)p.C) | ((uintptr_t)p.R)) & 31) == 0;
#pragma unroll
            for (int q2 = 0; q2 < 2; ++q2) {
              uint4 pa[2], pb[2], ph[2];
#pragma unroll
              for (int qq = 0; qq < 2; ++qq) {
                const int q = q2 * 2 + qq;
                float a[8], b[8], h[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  a[j] = bf16_round(__uint_as_float(ra[q * 8 + j]) * p.alpha);
                  b[j] = bf16_round(__uint_as_float(rb[q * 8 + j]) * p.alpha);
                  const float sg = bf16_round(__fdividef(1.f, 1.f + __expf(-a[j])));
                  h[j] = bf16_round(a[j] * sg) * b[j];
                }
                pa[qq] = pack8(a);
                pb[qq] = pack8(b);
                ph[qq] = pack8(h);
              }
              const int c = k * 32 + q2 * 16;
              if (wide) {                       // 32-byte stores: a whole sector per lane and instruction
                st_v8(Cab + c, pa[0], pa[1]);
                st_v8(Cab + F + c, pb[0], pb[1]);
                st_v8(H + c, ph[0], ph[1]);
              } else {
                st_v4(Cab + c, pa[0]); st_v4(Cab + c + 8, pa[1]);
                st_v4(Cab + F + c, pb[0]); st_v4(Cab + F + c + 8, pb[1]);
                st_v4(H + c, ph[0]); st_v4(H + c + 8, ph[1]);
              }
            }
          }
        }
      } else if constexpr (MODE == 2) {
        // staged epilogue: out = alpha * acc (+ bias) + R  |  (alpha * acc) * gelu'(R)  |  R + dropout(alpha * acc + bias),
        // written over R's tile
        const int rloc = ew * 32 + lane;
        const unsigned long long drop_seed =
            (p.flags & EP_DROPOUT) ? p.drop_seed + (p.drop_seed_dev != nullptr ? *p.drop_seed_dev : 0ull) : 0ull;
#pragma unroll 1
        for (int g = 0; g < P_BN / 64; ++g) {
          if (n_blk * P_BN + g * 64 >= p.N) break;
          const int n = nst++, buf = n % IO_BUFS;
          const int c = 2 * g + chalf;
          const long long col0 = (long long)n_blk * P_BN + c * 32;
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN + c * 32), r);
          mbar_wait(&in_full[buf], (uint32_t)((n / IO_BUFS) & 1));
          const uint32_t rrow = smem_u32(sab) + buf * BUF_BYTES + rloc * 128;
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t off = (uint32_t)(((chalf * 4 + q) ^ (rloc & 7)) << 4);
            const float4 fr = lds_f4(rrow + off);
            float res[8], v[8];
            unpack8(make_uint4(__float_as_uint(fr.x), __float_as_uint(fr.y), __float_as_uint(fr.z), __float_as_uint(fr.w)), res);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[j] = __uint_as_float(r[q * 8 + j]) * p.alpha;
              if (p.flags & EP_BIAS) v[j] += __bfloat162float(p.bias[col0 + q * 8 + j]);
            }
            if (p.flags & EP_GELU_BWD) {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] *= gelu_erf_grad(res[j]);
            } else if (p.flags & EP_DROPOUT) {
              const unsigned long long e0 = (unsigned long long)row * (unsigned long long)p.N + (unsigned long long)(col0 + q * 8);
#pragma unroll
              for (int j = 0; j < 8; j += 2) {
                bool k0, k1;
                dropout_keep2((e0 + j) >> 1, drop_seed, p.drop_thr16, k0, k1);
                // the unfused path rounds the linear's output, then the dropped value, to bf16 before the add
                v[j] = bf16_round(k0 ? bf16_round(v[j]) * p.drop_inv_keep : 0.f) + res[j];
                v[j + 1] = bf16_round(k1 ? bf16_round(v[j + 1]) * p.drop_inv_keep : 0.f) + res[j + 1];
              }
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] += res[j];
            }
            const uint4 vo = pack8(v);
            sts_v4(rrow + off, vo.x, vo.y, vo.z, vo.w);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&out_full[buf]);
        }
      } else if ((p.flags & EP_ROPE) && (long long)n_blk * P_BN < p.rope_cols) {
        // q | k tile of a packed q|k|v projection: RoPE in the epilogue (alpha is folded in here, so the stores below
        // are told alpha = 1). A head is rope_hd / 32 chunks; its halves are hc chunks apart.
        const int hc = p.rope_hd >> 6;                       // chunks per half head: 1 (hd 64) or 2 (hd 128)
        const long long t = row_ok ? row % p.rope_T : 0;
        ResChunk none;
        none.vec = false;
#pragma unroll 1
        for (int c = c0; c < c1; c += 2 * hc) {
#pragma unroll 1
          for (int k = 0; k < hc; ++k) {
            uint32_t r1[32], r2[32];
            const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN + (c + k) * 32);
            tmem_ld_32x32(taddr, r1);
            tmem_ld_32x32(taddr + hc * 32, r2);
            tmem_ld_wait();
            const long long col1 = (long long)n_blk * P_BN + (c + k) * 32;
            if (col1 < p.rope_cols)
              rope_chunk_pair(r1, r2, p.rope_cos + t * p.rope_hd + k * 32, p.rope_sin + t * p.rope_hd + k * 32, p.alpha);
            else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                r1[j] = __float_as_uint(__uint_as_float(r1[j]) * p.alpha);
                r2[j] = __float_as_uint(__uint_as_float(r2[j]) * p.alpha);
              }
            }
            store_chunk(p, row, row_ok, col1, r1, none, 1.f);
            store_chunk(p, row, row_ok, col1 + hc * 32, r2, none, 1.f);
          }
        }
      } else {
      ResChunk rc = load_residual(p, row, row_ok, (long long)n_blk * P_BN + c0 * 32);
#pragma unroll 1
      for (int c = c0; c < c1; ++c) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * P_BN + c * 32);
        tmem_ld_32x32(taddr, r);
        const ResChunk cur = rc;
        if (c + 1 < c1) rc = load_residual(p, row, row_ok, (long long)n_blk * P_BN + (c + 1) * 32);
        tmem_ld_wait();
        store_chunk(p, row, row_ok, (long long)n_blk * P_BN + c * 32, r, cur);
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty[as], 0);
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, 512);
  }
}

int launch_2cta(const CUtensorMap* maps, const GemmParams& p, int sms, cudaStream_t st) {
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM_SWIGLU));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM_SWIGLU));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));
    attr_set[dev] = true;
  }
  const int tiles = p.m_tiles * p.n_tiles;
  const int work = p.tail_start + (tiles - p.tail_start) * p.tail_splits;
  int clusters = sms / 2;
  if (work < clusters) clusters = work;
  // residual epilogue through TMA staging (MODE 2): bf16 output, whole 256-column tiles, 16-byte aligned pitches
  static const int res_tma = [] { const char* e = getenv("B200_GEMM_RES_TMA"); return e ? atoi(e) : 1; }();
  const bool fused_io = (p.flags & (EP_GELU_BWD | EP_DROPOUT)) != 0;     // these exist only in the staged epilogue
  const bool staged_res = (res_tma || fused_io) && ((p.flags & EP_RESIDUAL) || fused_io) &&
                          !(p.flags & (EP_GELU | EP_OUT_F32 | EP_TRANS_OUT | EP_ATOMIC)) && p.N % 64 == 0 && (p.ldr & 7) == 0 &&
                          (p.ldc & 7) == 0 && (((uintptr_t)p.R | (uintptr_t)p.C) & 15) == 0;
  B200_CHECK_ARG(staged_res || !fused_io, "gemm: the GELU-backward / dropout epilogues need bf16 output, N %% 64 == 0 and 16-B aligned rows");
  if (p.flags & EP_SWIGLU_BWD)
    B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_2cta_kernel<1>, dim3(2 * clusters), dim3(384), P_SMEM_SWIGLU, st, maps[0],
                                    maps[1], maps[2], maps[3], maps[4], maps[5], p));
  else if (p.flags & EP_SWIGLU_FWD)
    B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_2cta_kernel<3>, dim3(2 * clusters), dim3(256), P_SMEM, st, maps[0],
                                    maps[1], maps[2], maps[3], maps[4], maps[5], p));
  else if (staged_res)
    B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_2cta_kernel<2>, dim3(2 * clusters), dim3(384), P_SMEM_SWIGLU, st, maps[0],
                                    maps[1], maps[2], maps[3], maps[4], maps[5], p));
  else if (p.flags & EP_GELU)
    B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_2cta_kernel<4>, dim3(2 * clusters), dim3(384), P_SMEM, st, maps[0],
                                    maps[1], maps[2], maps[3], maps[4], maps[5], p));
  else
    B200_CHECK_CUDA(b200_launch_pdl(gemm_tcgen05_2cta_kernel<0>, dim3(2 * clusters), dim3(256), P_SMEM, st, maps[0],
                                    maps[1], maps[2], maps[3], maps[4], maps[5], p));
  b200_count_launch();
  return 0;
}

// ======================================================================================
// Grouped launch: up to GROUP_MAX independent skinny problems (N <= 64: the LoRA weight
// gradients dA = x^T u and dB = (dY^T t)^T, or the rank-r projections t = x A) share ONE
// persistent launch. Each of those problems alone fills at most 64-144 CTAs for ~20 us, most of it
// launch / prologue / pipeline ramp; concatenating their (tile, split) work lists keeps all SMs
// streaming and pays the fixed costs once.
// ======================================================================================
constexpr int GROUP_MAX = 16;
constexpr int G_BN = 64;
constexpr int G_B_BYTES = G_BN * BK * 2;
constexpr int G_STAGE_BYTES = A_BYTES + G_B_BYTES;
constexpr int G_STAGES = RING_BYTES / G_STAGE_BYTES;   // 8
constexpr int G_SMEM = G_STAGES * G_STAGE_BYTES + 1024 + (2 * G_STAGES + 4) * 8 + 16;

struct GroupedParams {
  int nprob;
  int work_end[GROUP_MAX];     // exclusive prefix sums of tiles * splits
  GemmParams p[GROUP_MAX];
};
struct GroupedMaps {
  CUtensorMap a[GROUP_MAX];
  CUtensorMap b[GROUP_MAX];
};

__device__ __forceinline__ int find_problem(const GroupedParams& g, int w, int& local) {
  int i = 0, begin = 0;
  while (i < g.nprob - 1 && w >= g.work_end[i]) {
    begin = g.work_end[i];
    ++i;
  }
  local = w - begin;
  return i;
}

__global__ void __launch_bounds__(256, 1)
gemm_grouped_kernel(const __grid_constant__ GroupedMaps maps, const __grid_constant__ GroupedParams g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + G_STAGES * G_STAGE_BYTES);
  uint64_t* empty = full + G_STAGES;
  uint64_t* tfull = empty + G_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * G_BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();
  const int num_work = g.work_end[g.nprob - 1];

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        int local;
        const int pi = find_problem(g, w, local);
        const GemmParams& p = g.p[pi];
        const int num_tiles = p.m_tiles * p.n_tiles;
        const int split = local / num_tiles, tile = local % num_tiles;
        const int m_blk = tile % p.m_tiles, n_blk = tile / p.m_tiles;
        const int kb_begin = (int)((long long)split * p.kb1 / p.splits);
        const int kb_end = (int)((long long)(split + 1) * p.kb1 / p.splits);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], G_STAGE_BYTES);
          uint8_t* sa = smem + stage * G_STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const int kk = kb * BK;
          if (!p.a1_mn) {
            tma_load_2d(sa, &maps.a[pi], &full[stage], kk, m_blk * BM);
          } else {
            tma_load_2d(sa, &maps.a[pi], &full[stage], m_blk * BM, kk);
            tma_load_2d(sa + 8192, &maps.a[pi], &full[stage], m_blk * BM + 64, kk);
          }
          if (!p.b1_mn) tma_load_2d(sb, &maps.b[pi], &full[stage], kk, n_blk * G_BN);
          else tma_load_2d(sb, &maps.b[pi], &full[stage], n_blk * G_BN, kk);
          if (++stage == G_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++it) {
        int local;
        const int pi = find_problem(g, w, local);
        const GemmParams& p = g.p[pi];
        const int split = local / (p.m_tiles * p.n_tiles);
        const int kb_begin = (int)((long long)split * p.kb1 / p.splits);
        const int kb_end = (int)((long long)(split + 1) * p.kb1 / p.splits);
        const int as = it & 1;
        mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * G_BN);
        const uint32_t idesc = umma_idesc_bf16(BM, G_BN, (uint32_t)p.a1_mn, (uint32_t)p.b1_mn);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * G_STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = p.a1_mn ? umma_desc_mn128(sa + k * 2048, 8192) : umma_desc_k128(sa + k * 32);
            const uint64_t bd = p.b1_mn ? umma_desc_mn128(sb + k * 2048, 8192) : umma_desc_k128(sb + k * 32);
            umma_bf16_ss(d_tmem, ad, bd, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (++stage == G_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tfull[as]);
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    int it = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++it) {
      int local;
      const int pi = find_problem(g, w, local);
      const GemmParams& p = g.p[pi];
      const int tile = local % (p.m_tiles * p.n_tiles);
      const int m_blk = tile % p.m_tiles, n_blk = tile / p.m_tiles;
      const int as = it & 1;
      mbar_wait(&tfull[as], (it >> 1) & 1);
      tc_fence_after();
      const long long row = (long long)m_blk * BM + ew * 32 + lane;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c = 0; c < G_BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * G_BN + c * 32), r);
        tmem_ld_wait();
        ResChunk none;
        none.vec = false;
        store_chunk(p, row, row_ok, (long long)n_blk * G_BN + c * 32, r, none);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * G_BN);
  }
}

// ======================================================================================
// Skinny problems on a CTA pair per 128-row tile: the two CTAs of a cluster each stream HALF of the K range into their
// own TMEM accumulator and exchange fp32 partial sums through distributed shared memory (rows 0..63 are finished by
// CTA 0, rows 64..127 by CTA 1). A skinny launch is bound by how fast one SM takes TMA data (~77 GB/s measured, L2
// hits or not): M = 8192 gives only 64 tiles, so the tile-scheduled kernel streams 768 KB per CTA on 64 of 148 SMs;
// stream-K spreads it over all SMs but pays ~5 us for the fix-up through global memory (RED, fence, counter, reload).
// Here 128 CTAs stream 384 KB each and the fix-up is a 16 KB DSMEM write plus one cluster barrier.
// Up to SP_NB problems that share the A operand (q | k | v, fc1 | fc2 projections of the same x) are ONE work item:
// A is loaded once per K block, every B has its own 64-column accumulator.
// ======================================================================================
constexpr int SP_NB = 3;
constexpr int SP_RING_BYTES = 176 * 1024;
struct PairSkinnyParams {
  int ngroups;
  int tile_end[GROUP_MAX];     // exclusive prefix sums of the groups' 128-row tiles
  int first[GROUP_MAX];        // first problem of the group
  int nb[GROUP_MAX];           // problems in the group (same A, M, K)
  GemmParams p[GROUP_MAX];
};
template <int NB>
struct PairSkinnyCfg {
  static constexpr int STAGE_BYTES = A_BYTES + NB * G_B_BYTES;
  static constexpr int STAGES = SP_RING_BYTES / STAGE_BYTES;          // 7 / 5 / 4
  static constexpr int PART_BYTES = NB * 64 * G_BN * 4;               // 64 rows x 64 columns fp32 per B
  static constexpr int SMEM = STAGES * STAGE_BYTES + PART_BYTES + 1024 + (2 * STAGES + 2) * 8 + 16;
  static constexpr uint32_t TMEM_COLS = NB == 1 ? 64 : (NB == 2 ? 128 : 256);
};

__device__ __forceinline__ void st_cluster_f4(uint32_t local_saddr, uint32_t rank, float a, float b, float c, float d) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "st.shared::cluster.v4.f32 [ra], {%2, %3, %4, %5};\n\t"
      "}" ::"r"(local_saddr),
      "r"(rank), "f"(a), "f"(b), "f"(c), "f"(d)
      : "memory");
}

template <int NB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1)
gemm_skinny_pair_kernel(const __grid_constant__ GroupedMaps maps, const __grid_constant__ PairSkinnyParams g) {
  using Cfg = PairSkinnyCfg<NB>;
  constexpr int STAGES = Cfg::STAGES, STAGE_BYTES = Cfg::STAGE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* part = smem + STAGES * STAGE_BYTES;                  // [NB][16 column quads][64 rows] float4
  uint64_t* full = (uint64_t*)(part + Cfg::PART_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint32_t* tmem_slot = (uint32_t*)(tfull + 2);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();            // also: the peer CTA is running before anything is written into its shared memory
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();

  // work item of this cluster
  const int cid = (int)(blockIdx.x >> 1);
  int gi = 0, begin = 0;
  while (gi < g.ngroups - 1 && cid >= g.tile_end[gi]) {
    begin = g.tile_end[gi];
    ++gi;
  }
  const int m_blk = cid - begin;
  const int first = g.first[gi], nb = g.nb[gi];
  const GemmParams& p0 = g.p[first];
  const int kb_begin = (int)((long long)rank * p0.kb1 / 2);
  const int kb_end = (int)((long long)(rank + 1) * p0.kb1 / 2);
  const bool has = kb_end > kb_begin;                            // K = 64: CTA 0 has nothing to add

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_expect_tx(&full[stage], (uint32_t)(A_BYTES + nb * G_B_BYTES));
        uint8_t* sa = smem + stage * STAGE_BYTES;
        const int kk = kb * BK;
        if (!p0.a1_mn) {
          tma_load_2d(sa, &maps.a[first], &full[stage], kk, m_blk * BM);
        } else {
          tma_load_2d(sa, &maps.a[first], &full[stage], m_blk * BM, kk);
          tma_load_2d(sa + 8192, &maps.a[first], &full[stage], m_blk * BM + 64, kk);
        }
        for (int i = 0; i < nb; ++i) {
          uint8_t* sb = sa + A_BYTES + i * G_B_BYTES;
          if (!g.p[first + i].b1_mn) tma_load_2d(sb, &maps.b[first + i], &full[stage], kk, 0);
          else tma_load_2d(sb, &maps.b[first + i], &full[stage], 0, kk);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        for (int i = 0; i < nb; ++i) {
          const uint32_t sb = sa + A_BYTES + i * G_B_BYTES;
          const int b_mn = g.p[first + i].b1_mn;
          const uint32_t idesc = umma_idesc_bf16(BM, G_BN, (uint32_t)p0.a1_mn, (uint32_t)b_mn);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = p0.a1_mn ? umma_desc_mn128(sa + k * 2048, 8192) : umma_desc_k128(sa + k * 32);
            const uint64_t bd = b_mn ? umma_desc_mn128(sb + k * 2048, 8192) : umma_desc_k128(sb + k * 32);
            umma_bf16_ss(tmem_base + (uint32_t)(i * G_BN), ad, bd, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (has) umma_commit(tfull);
      else mbar_arrive(tfull);
    }
    __syncwarp();
  } else if (warp >= 4) {
    // rows [64 r, 64 r + 64) of the tile are finished by CTA r: the other two epilogue warps send their partial sums
    const int ew = warp - 4;
    const uint32_t owner = (uint32_t)(ew >> 1);
    mbar_wait(tfull, 0);
    tc_fence_after();
    if (owner != rank) {
      const int rl = (ew & 1) * 32 + lane;                       // row inside the owner's 64
      for (int i = 0; i < nb; ++i) {
#pragma unroll 1
        for (int c = 0; c < G_BN / 32; ++c) {
          uint32_t r[32];
          if (has) {
            tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(i * G_BN + c * 32), r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          const uint32_t dst = smem_u32(part) + (uint32_t)(((i * 16 + c * 8) * 64 + rl) * 16);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            st_cluster_f4(dst + (uint32_t)(q * 64 * 16), owner, __uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                          __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
        }
      }
    }
  }
  __syncwarp();
  cluster_sync_all();            // the partial sums of both CTAs are in place (release / acquire at cluster scope)
  if (warp >= 4) {
    const int ew = warp - 4;
    if ((uint32_t)(ew >> 1) == rank) {
      const int rl = (ew & 1) * 32 + lane;
      const long long row = (long long)m_blk * BM + ew * 32 + lane;
      for (int i = 0; i < nb; ++i) {
        const GemmParams& p = g.p[first + i];
        const bool row_ok = row < p.M;
#pragma unroll 1
        for (int c = 0; c < G_BN / 32; ++c) {
          uint32_t r[32];
          if (has) {
            tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(i * G_BN + c * 32), r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          const uint32_t src = smem_u32(part) + (uint32_t)(((i * 16 + c * 8) * 64 + rl) * 16);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 v = lds_f4(src + (uint32_t)(q * 64 * 16));
            r[4 * q] = __float_as_uint(__uint_as_float(r[4 * q]) + v.x);
            r[4 * q + 1] = __float_as_uint(__uint_as_float(r[4 * q + 1]) + v.y);
            r[4 * q + 2] = __float_as_uint(__uint_as_float(r[4 * q + 2]) + v.z);
            r[4 * q + 3] = __float_as_uint(__uint_as_float(r[4 * q + 3]) + v.w);
          }
          ResChunk none;
          none.vec = false;
          store_chunk(p, row, row_ok, (long long)c * 32, r, none);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ======================================================================================
// Stream-K form of the grouped launch. The skinny problems are HBM-bound streams over an
// activation (t = x A, u = dY B^T, dA = x^T u, dB = t^T dY: N <= 64, K up to 128 K), so what matters is
// that all 148 SMs stream the same number of bytes. The (problem, 128-row tile, 64-wide K block)
// units of the whole group are laid end to end and cut into gridDim.x equal ranges; a CTA's range
// is a few "segments" (part of one tile's K range each).
//   segment covers the tile's whole K range, or the problem accumulates into fp32 C: usual epilogue;
//   otherwise: the raw partial sums are added (fp32 RED) into a workspace tile, a per-tile counter
//   collects the K blocks done, and the CTA that completes the count converts the workspace tile
//   (alpha, bf16 / fp32 store) and puts the zeros back, so the workspace is all-zero between launches.
// ======================================================================================
struct StreamKParams {
  int nprob;
  int unit_end[GROUP_MAX];        // exclusive prefix sums of m_tiles * kb1
  int cnt_off[GROUP_MAX];         // first tile counter of the problem
  long long ws_off[GROUP_MAX];    // first workspace float of the problem (partial tiles)
  float* ws;
  int* cnt;
  GemmParams p[GROUP_MAX];
};

struct Segment {
  int pi, tile, kb0, n;
};
__device__ __forceinline__ Segment streamk_segment(const StreamKParams& g, int u, int u_end) {
  int i = 0, begin = 0;
  while (i < g.nprob - 1 && u >= g.unit_end[i]) {
    begin = g.unit_end[i];
    ++i;
  }
  const int kb = g.p[i].kb1;
  const int local = u - begin;
  Segment s;
  s.pi = i;
  s.tile = local / kb;
  s.kb0 = local - s.tile * kb;
  s.n = min(kb - s.kb0, u_end - u);
  return s;
}


__global__ void __launch_bounds__(256, 1)
gemm_streamk_kernel(const __grid_constant__ GroupedMaps maps, const __grid_constant__ StreamKParams g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_last;
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + G_STAGES * G_STAGE_BYTES);
  uint64_t* empty = full + G_STAGES;
  uint64_t* tfull = empty + G_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * G_BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();
  const int U = g.unit_end[g.nprob - 1];
  const int u_begin = (int)((long long)blockIdx.x * U / gridDim.x);
  const int u_end = (int)((long long)(blockIdx.x + 1) * U / gridDim.x);

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int u = u_begin; u < u_end;) {
        const Segment s = streamk_segment(g, u, u_end);
        const GemmParams& p = g.p[s.pi];
        for (int kb = s.kb0; kb < s.kb0 + s.n; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], G_STAGE_BYTES);
          uint8_t* sa = smem + stage * G_STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const int kk = kb * BK;
          if (!p.a1_mn) {
            tma_load_2d(sa, &maps.a[s.pi], &full[stage], kk, s.tile * BM);
          } else {
            tma_load_2d(sa, &maps.a[s.pi], &full[stage], s.tile * BM, kk);
            tma_load_2d(sa + 8192, &maps.a[s.pi], &full[stage], s.tile * BM + 64, kk);
          }
          if (!p.b1_mn) tma_load_2d(sb, &maps.b[s.pi], &full[stage], kk, 0);
          else tma_load_2d(sb, &maps.b[s.pi], &full[stage], 0, kk);
          if (++stage == G_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        u += s.n;
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int u = u_begin; u < u_end; ++it) {
        const Segment s = streamk_segment(g, u, u_end);
        const GemmParams& p = g.p[s.pi];
        const int as = it & 1;
        mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * G_BN);
        const uint32_t idesc = umma_idesc_bf16(BM, G_BN, (uint32_t)p.a1_mn, (uint32_t)p.b1_mn);
        for (int kb = 0; kb < s.n; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * G_STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = p.a1_mn ? umma_desc_mn128(sa + k * 2048, 8192) : umma_desc_k128(sa + k * 32);
            const uint64_t bd = p.b1_mn ? umma_desc_mn128(sb + k * 2048, 8192) : umma_desc_k128(sb + k * 32);
            umma_bf16_ss(d_tmem, ad, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (++stage == G_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tfull[as]);
        u += s.n;
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    const int r_in_tile = ew * 32 + lane;
    int it = 0;
    for (int u = u_begin; u < u_end; ++it) {
      const Segment s = streamk_segment(g, u, u_end);
      const GemmParams& p = g.p[s.pi];
      u += s.n;
      const int as = it & 1;
      mbar_wait(&tfull[as], (it >> 1) & 1);
      tc_fence_after();
      const long long row = (long long)s.tile * BM + r_in_tile;
      const bool row_ok = row < p.M;
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * G_BN);
      ResChunk none;
      none.vec = false;
      if (s.n == p.kb1 || (p.flags & EP_ATOMIC)) {
#pragma unroll 1
        for (int c = 0; c < G_BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + c * 32, r);
          tmem_ld_wait();
          store_chunk(p, row, row_ok, c * 32, r, none);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[as]);
        continue;
      }
      // partial tile: workspace is [column][row-in-tile] so a warp's RED / load touches one 128 B line
      float* wt = g.ws + g.ws_off[s.pi] + (long long)s.tile * (BM * G_BN) + r_in_tile;
#pragma unroll 1
      for (int c = 0; c < G_BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) atomicAdd(wt + (c * 32 + j) * BM, __uint_as_float(r[j]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
      __threadfence();
      epi_bar_sync();
      if (threadIdx.x == 128) {
        const int old = atomicAdd(g.cnt + g.cnt_off[s.pi] + s.tile, s.n);
        s_last = (old + s.n == p.kb1) ? 1 : 0;
      }
      epi_bar_sync();
      if (s_last) {
        __threadfence();
#pragma unroll 1
        for (int c = 0; c < G_BN / 32; ++c) {
          uint32_t r[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            r[j] = __float_as_uint(__ldcg(wt + (c * 32 + j) * BM));
            __stcg(wt + (c * 32 + j) * BM, 0.f);
          }
          store_chunk(p, row, row_ok, c * 32, r, none);
        }
        if (threadIdx.x == 128) g.cnt[g.cnt_off[s.pi] + s.tile] = 0;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * G_BN);
  }
}

struct TailWs {
  void* ptr;
  size_t bytes;
};
// Tail-split scratch: one default per device (b200_gemm_set_workspace) and any number of per-stream ones
// (b200_gemm_set_stream_workspace) for hosts that run GEMMs concurrently on several streams of a device —
// two launches sharing a scratch would mix their partial tiles.
TailWs g_tail_ws[64] = {};
struct StreamWs {
  int dev;
  void* stream;
  TailWs ws;
};
std::vector<StreamWs> g_stream_ws;
std::mutex g_ws_mu;

TailWs tail_workspace_for(int dev, void* stream) {
  std::lock_guard<std::mutex> lk(g_ws_mu);
  for (const StreamWs& e : g_stream_ws)
    if (e.dev == dev && e.stream == stream) return e.ws;
  return (dev >= 0 && dev < 64) ? g_tail_ws[dev] : TailWs{nullptr, 0};
}

// which scheduling path each launch took (diagnostics for tests: b200_gemm_path_counts)
enum : int { PATH_PAIR, PATH_PAIR_TAIL_SPLIT, PATH_SINGLE_CTA, PATH_GROUPED_TILES, PATH_STREAMK, PATH_PACKED_SEGMENTS,
              PATH_SWIGLU_FWD, PATH_SWIGLU_BWD, PATH_SKINNY_PAIR, PATH_N };
std::atomic<uint64_t> g_path[PATH_N];

// Scheduling switches for A/B measurements and tests (b200_gemm_set_option): -1 = the library's own heuristic,
// 0 / 1 = off / forced. Their initial values come from the environment, read once per process.
enum : int { OPT_TAIL_SPLIT, OPT_STREAMK, OPT_N_FASTEST, OPT_SKINNY_PAIR, OPT_N };
std::atomic<int> g_opt[OPT_N];
std::once_flag g_opt_once;
int gemm_option(int which) {
  std::call_once(g_opt_once, [] {
    auto env = [](const char* name, char on, char off) {
      const char* e = getenv(name);
      return !e ? -1 : (e[0] == on ? 1 : (e[0] == off ? 0 : -1));
    };
    g_opt[OPT_TAIL_SPLIT] = env("B200_GEMM_TAIL", '1', '0');
    g_opt[OPT_STREAMK] = env("B200_GEMM_STREAMK_FORCE", '1', '0');
    g_opt[OPT_N_FASTEST] = env("B200_GEMM_ORDER", 'n', 'm');
    g_opt[OPT_SKINNY_PAIR] = env("B200_GEMM_SKINNY_PAIR", '1', '0');
  });
  return g_opt[which].load();
}

bool use_pair_kernel() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_GEMM_1CTA");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

struct Operand {
  const void* ptr;
  int64_t ld;
  int mn_major;  // 0: [rows, K] K contiguous, 1: [K, rows] rows contiguous
};

int make_operand_map(CUtensorMap* map, const Operand& o, int64_t rows, int64_t K, int box_rows) {
  if (!o.mn_major) return b200_make_tmap_2d_bf16(map, o.ptr, (uint64_t)rows, (uint64_t)K, (uint64_t)o.ld, (uint32_t)box_rows, BK);
  // stored [K, rows]: box = 64 K-rows x 64 contiguous MN elements
  return b200_make_tmap_2d_bf16(map, o.ptr, (uint64_t)K, (uint64_t)rows, (uint64_t)o.ld, BK, 64);
}

// optional epilogue extras of gemm_ex_impl
struct Extra {
  void* Z = nullptr;          // EP_GELU: pre-activation output
  int64_t ldz = 0;
  bool gelu_bwd = false;      // C = acc * gelu'(R)
  bool dropout = false;       // C = R + dropout(acc + bias)
  float drop_p = 0.f;
  uint64_t drop_seed = 0;
  const uint64_t* drop_seed_dev = nullptr;
  const void* rope_cos = nullptr;   // EP_ROPE on the output columns < rope_cols
  const void* rope_sin = nullptr;
  int64_t rope_T = 0, rope_hd = 0, rope_cols = 0;
};

struct SegInfo {
  int nseg;             // output column segments (<= 4)
  int64_t start[4];     // first column of each segment (start[0] = 0), multiples of the 256-column tile
  int64_t a2_stride;    // A2 column offset between consecutive segments
};

int gemm_ex_impl(const void* A1, int64_t lda1, int a1_mn, const void* B1, int64_t ldb1,
                 int b1_mn, int64_t K1, const void* A2, int64_t lda2, int a2_mn,
                 const void* B2, int64_t ldb2, int b2_mn, int64_t K2, const void* bias,
                 const void* R, int64_t ldr, void* C, int64_t ldc, int64_t M, int64_t N,
                 int flags, int out_f32, int accumulate, int trans_out, float alpha,
                 int splitk, void* stream, const SegInfo* seg_info, bool swiglu_bwd = false,
                 bool swiglu_fwd = false, const Extra* ex = nullptr) {
  B200_CHECK_ARG(M > 0 && N > 0 && K1 > 0 && K2 >= 0, "gemm: bad sizes M=%lld N=%lld K1=%lld K2=%lld",
                 (long long)M, (long long)N, (long long)K1, (long long)K2);
  B200_CHECK_ARG(A1 && B1 && C, "gemm: null operand");
  B200_CHECK_ARG(K2 == 0 || (A2 && B2), "gemm: K2 > 0 needs A2 and B2");
  B200_CHECK_ARG(!(flags & B200_GEMM_BIAS) || bias, "gemm: bias flag without bias pointer");
  B200_CHECK_ARG(!(flags & B200_GEMM_RESIDUAL) || R, "gemm: residual flag without R pointer");
  B200_CHECK_ARG(!(accumulate && !out_f32), "gemm: accumulate needs fp32 output");
  B200_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31) && K1 < (1ll << 31), "gemm: sizes exceed int32");

  const bool pair = use_pair_kernel() && !accumulate && !trans_out && M >= 256 && N >= 256;
  B200_CHECK_ARG(!seg_info || (pair && !a2_mn), "gemm: segmented LoRA columns need the CTA-pair kernel and a K-major A2");
  B200_CHECK_ARG(!swiglu_bwd || (pair && N % 32 == 0 && R && !out_f32 && flags == 0), "gemm: the SwiGLU-backward epilogue needs the CTA-pair kernel, N %% 32 == 0 and the packed (a | b) in R");
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  GemmParams p;
  p.M = (int)M;
  p.N = (int)N;
  p.kb1 = (int)((K1 + BK - 1) / BK);
  p.kb2 = (int)((K2 + BK - 1) / BK);
  p.a1_mn = a1_mn; p.b1_mn = b1_mn; p.a2_mn = a2_mn; p.b2_mn = b2_mn;
  p.m_tiles = pair ? (int)((M + 255) / 256) : (int)((M + BM - 1) / BM);
  p.n_tiles = (int)((N + BN - 1) / BN);
  // With M fastest every concurrent wave touches all of A; when A (M x K) is far larger than B (N x K)
  // and than the L2, walking N fastest keeps ~74/n_tiles A tile-rows plus all of B resident instead.
  p.n_fastest = 0;
  if (pair) {
    const double a_bytes = 2.0 * (double)M * (double)(K1 + K2), b_bytes = 2.0 * (double)N * (double)(K1 + K2);
    const int order = gemm_option(OPT_N_FASTEST);
    if (order >= 0) p.n_fastest = order;
    // (B up to ~3/4 of the 126 MB L2: the fc1|fc2 input-gradient GEMM has B = 67.6 MB against A = 270 MB and read
    // 1176 MB from DRAM per launch with M fastest, 3.5x its operands)
    // When neither fits (vocab projection's input gradient: A = dlogits 919 MB, B = W 525 MB) N fastest measured worse:
    // 3565 MB of DRAM reads per launch against 2479 MB with M fastest.
    else p.n_fastest = (a_bytes > 64e6 && b_bytes < 96e6 && a_bytes > 2.0 * b_bytes) ? 1 : 0;
  }
  p.alpha = alpha;
  p.bias = (const bf16*)bias;
  p.R = (const bf16*)R;
  p.ldr = ldr;
  p.C = C;
  p.ldc = ldc;
  p.flags = (flags & 7) | (out_f32 ? EP_OUT_F32 : 0) | (trans_out ? EP_TRANS_OUT : 0) | (swiglu_bwd ? EP_SWIGLU_BWD : 0) |
            (swiglu_fwd ? EP_SWIGLU_FWD : 0);
  p.Z = nullptr;
  p.ldz = 0;
  p.drop_thr16 = 0;
  p.drop_inv_keep = 1.f;
  p.drop_seed = 0;
  p.drop_seed_dev = nullptr;
  p.rope_cos = p.rope_sin = nullptr;
  p.rope_T = p.rope_hd = p.rope_cols = 0;
  if (ex && ex->rope_cos) {
    B200_CHECK_ARG(pair && ex->rope_sin && (ex->rope_hd == 64 || ex->rope_hd == 128) && ex->rope_T > 0 && ex->rope_cols > 0 &&
                       ex->rope_cols % ex->rope_hd == 0 && ex->rope_cols <= N && !out_f32 && (flags & 7) == 0 && ldc % 8 == 0,
                   "gemm: the RoPE epilogue needs the CTA-pair kernel, head_dim 64 or 128, whole heads, bf16 output, no bias / GELU / residual");
    B200_CHECK_ARG(b200_aligned16(ex->rope_cos) && b200_aligned16(ex->rope_sin), "gemm: RoPE tables must be 16-B aligned");
    p.flags |= EP_ROPE;
    p.rope_cos = (const bf16*)ex->rope_cos;
    p.rope_sin = (const bf16*)ex->rope_sin;
    p.rope_T = (int)ex->rope_T;
    p.rope_hd = (int)ex->rope_hd;
    p.rope_cols = (int)ex->rope_cols;
  }
  const bool fused_io = ex && (ex->gelu_bwd || ex->dropout);
  if (ex) {
    p.Z = (bf16*)ex->Z;
    p.ldz = ex->ldz;
    if (ex->gelu_bwd) p.flags |= EP_GELU_BWD;
    if (ex->dropout) {
      B200_CHECK_ARG(ex->drop_p >= 0.f && ex->drop_p < 1.f && N % 2 == 0, "gemm: dropout epilogue needs 0 <= p < 1 and an even N");
      p.flags |= EP_DROPOUT;
      const float t = ex->drop_p * 65536.f + 0.5f;
      p.drop_thr16 = t > 65535.f ? 65535u : (uint32_t)t;
      p.drop_inv_keep = 1.f / (1.f - ex->drop_p);
      p.drop_seed = (unsigned long long)ex->drop_seed;
      p.drop_seed_dev = (const unsigned long long*)ex->drop_seed_dev;
    }
    B200_CHECK_ARG(!fused_io || (pair && R && !out_f32 && N % 64 == 0 && ldr % 8 == 0 && ldc % 8 == 0 && b200_aligned16(R) && b200_aligned16(C)),
                   "gemm: the GELU-backward / dropout epilogues need the CTA-pair kernel (M, N >= 256), bf16 output, N %% 64 == 0 and 16-B aligned rows");
  }
  B200_CHECK_ARG(!swiglu_fwd || (pair && N % P_BN == 0 && R && !out_f32 && flags == 0),
                 "gemm: the SwiGLU-forward epilogue needs the CTA-pair kernel, N %% %d == 0 and the h buffer in R", P_BN);

  const int sms = b200_num_sms();
  const int tiles = p.m_tiles * p.n_tiles;
  const int total_kb = p.kb1 + p.kb2;
  int splits = 1;
  if (accumulate) {
    if (splitk > 0) splits = splitk;
    else if (tiles * 2 <= sms) splits = sms / tiles;
    if (splits > total_kb) splits = total_kb;
    if (splits < 1) splits = 1;
    p.flags |= EP_ATOMIC;
  }
  p.splits = splits;
  p.tail_start = tiles;
  p.tail_splits = 1;
  p.tail_ws = nullptr;
  p.tail_cnt = nullptr;
  p.seg_start[0] = p.seg_start[1] = p.seg_start[2] = 0x7fffffff;
  p.a2_seg_stride = 0;
  if (seg_info && seg_info->nseg > 1) {
    for (int i = 0; i + 1 < seg_info->nseg && i < 3; ++i) p.seg_start[i] = (int)seg_info->start[i + 1];
    p.a2_seg_stride = (int)seg_info->a2_stride;
  }
  if (pair) {
    // Tail splitting: with T tiles on G CTA pairs the last wave holds L = T mod G tiles; when it is at
    // most half full its tiles are cut into floor(G / L) K ranges (>= 4 K blocks each) so that wave
    // takes 1/s of a tile time instead of a whole one (8192 x 2048 outputs: 256 tiles on 74 pairs =
    // 3.46 waves -> 3.5 instead of 4 tile times). Needs the workspace of b200_gemm_set_workspace().
    int dev = 0;
    cudaGetDevice(&dev);
    const TailWs tw = tail_workspace_for(dev, stream);
    const int G = sms / 2;
    const int L = tiles % G;
    const int tail_opt = gemm_option(OPT_TAIL_SPLIT);
    // parking and re-reading the partial tiles costs about as much as 16 K blocks of MMA time: measured
    // 65 -> 69 us at K = 2048 (a loss) but 202 -> 189 us at K = 8192, so only deep K ranges are split
    if (tw.ptr && tiles > G && L > 0 && total_kb >= 64 && !swiglu_bwd && !swiglu_fwd && !fused_io && !(p.flags & EP_ROPE) && tail_opt != 0) {
      // s K ranges per tail tile make L * s units: the tail takes ceil(L s / G) / s tile times; every split also pays
      // the fix-up (~24 K blocks' worth). L = 38 (vocab-projection input gradient at 3584 rows: 112 tiles) gains nothing
      // from s = 2 (76 units on 74 pairs) but s = 3 gives two rounds of thirds: 2.0 -> 1.67 tile times.
      const size_t cnt_bytes = 4096;   // fixed: the counters of one launch must never be another launch's slot bytes
      const size_t slot_bytes = (size_t)256 * P_BN * sizeof(float);
      int s = 1;
      double best = 1.0;
      for (int c = 2; c <= 4 && c <= total_kb / 32; ++c) {
        if (cnt_bytes + (size_t)L * c * slot_bytes > tw.bytes || (size_t)L * 2 * sizeof(int) > cnt_bytes) break;
        const double cost = (double)((L * c + G - 1) / G) / c + 24.0 / total_kb;
        if (cost < best - 0.05) {
          best = cost;
          s = c;
        }
      }
      if (s >= 2) {
        p.tail_start = tiles - L;
        p.tail_splits = s;
        p.tail_cnt = (int*)tw.ptr;
        p.tail_ws = (float*)((char*)tw.ptr + cnt_bytes);
      }
    }
  }

  CUtensorMap maps[6];
  int rc;
  Operand oa1{A1, lda1, a1_mn}, ob1{B1, ldb1, b1_mn};
  const int b_box_rows = pair ? 128 : BN;   // the pair kernel loads half of the B tile per CTA
  if ((rc = make_operand_map(&maps[0], oa1, M, K1, BM))) return rc;
  if ((rc = make_operand_map(&maps[1], ob1, N, K1, b_box_rows))) return rc;
  if (K2 > 0) {
    Operand oa2{A2, lda2, a2_mn}, ob2{B2, ldb2, b2_mn};
    const int64_t a2_cols = seg_info ? K2 + (seg_info->nseg - 1) * seg_info->a2_stride : K2;
    if ((rc = make_operand_map(&maps[2], oa2, M, a2_cols, BM))) return rc;
    if ((rc = make_operand_map(&maps[3], ob2, N, K2, b_box_rows))) return rc;
  } else {
    maps[2] = maps[0];
    maps[3] = maps[1];
  }
  maps[4] = maps[0];
  maps[5] = maps[0];
  if (pair && !swiglu_bwd && !swiglu_fwd && ((flags & B200_GEMM_RESIDUAL) || fused_io) && !(flags & B200_GEMM_GELU) && !out_f32 &&
      N % 64 == 0 && ldr % 8 == 0 && ldc % 8 == 0 && b200_aligned16(R) && b200_aligned16(C)) {
    // residual tile in, output tile out, both by TMA (MODE 2)
    if ((rc = b200_make_tmap_2d_bf16(&maps[4], R, (uint64_t)M, (uint64_t)N, (uint64_t)ldr, 128, 64))) return rc;
    if ((rc = b200_make_tmap_2d_bf16(&maps[5], C, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 128, 64))) return rc;
  }
  if (swiglu_bwd) {   // (a | b) and (da | db), both [M, 2N], move 64 columns x 128 rows at a time through TMA
    B200_CHECK_ARG(N % P_BN == 0, "gemm: the SwiGLU-backward epilogue needs N %% %d == 0", P_BN);
    if ((rc = b200_make_tmap_2d_bf16(&maps[4], R, (uint64_t)M, (uint64_t)(2 * N), (uint64_t)ldr, 128, 64))) return rc;
    if ((rc = b200_make_tmap_2d_bf16(&maps[5], C, (uint64_t)M, (uint64_t)(2 * N), (uint64_t)ldc, 128, 64))) return rc;
  }
  const int work = tiles * splits;
  const int grid = work < sms ? work : sms;
  cudaStream_t st = (cudaStream_t)stream;
  if (pair) {
    ++g_path[PATH_PAIR];
    if (p.tail_splits > 1) ++g_path[PATH_PAIR_TAIL_SPLIT];
    if (seg_info && seg_info->nseg > 1) ++g_path[PATH_PACKED_SEGMENTS];
    if (swiglu_fwd) ++g_path[PATH_SWIGLU_FWD];
    if (swiglu_bwd) ++g_path[PATH_SWIGLU_BWD];
    return launch_2cta(maps, p, sms, st);
  }
  ++g_path[PATH_SINGLE_CTA];
  switch (BN) {
    case 64: return launch_bn<64>(maps, p, grid, st);
    case 128: return launch_bn<128>(maps, p, grid, st);
    default: return launch_bn<256>(maps, p, grid, st);
  }
}

}  // namespace

extern "C" int b200_gemm_ex(const void* A1, int64_t lda1, int a1_mn, const void* B1, int64_t ldb1,
                            int b1_mn, int64_t K1, const void* A2, int64_t lda2, int a2_mn,
                            const void* B2, int64_t ldb2, int b2_mn, int64_t K2, const void* bias,
                            const void* R, int64_t ldr, void* C, int64_t ldc, int64_t M, int64_t N,
                            int flags, int out_f32, int accumulate, int trans_out, float alpha,
                            int splitk, void* stream) {
  return gemm_ex_impl(A1, lda1, a1_mn, B1, ldb1, b1_mn, K1, A2, lda2, a2_mn, B2, ldb2, b2_mn, K2, bias, R, ldr, C,
                      ldc, M, N, flags, out_f32, accumulate, trans_out, alpha, splitk, stream, nullptr);
}

// (da | db) = SwiGLU'(a, b) * (dY·W + U·A^T): the backward of fc3's input with the SwiGLU backward in the epilogue.
extern "C" int b200_gemm_swiglu_bwd_bf16(const void* dY, int64_t lddy, const void* W, int64_t ldw, int64_t K,
                                         const void* U, int64_t ldu, const void* A, int64_t lda, int64_t r,
                                         const void* AB, int64_t ldab, void* dAB, int64_t lddab, int64_t M,
                                         int64_t F, void* stream) {
  B200_CHECK_ARG(AB && dAB && M >= 256 && F >= 256 && F % 256 == 0 && ldab >= 2 * F && lddab >= 2 * F && ldab % 8 == 0 && lddab % 8 == 0,
                 "gemm_swiglu_bwd: needs M >= 256, F a multiple of 256 and packed [M, 2F] buffers with 16-B aligned rows");
  B200_CHECK_ARG(b200_aligned16(AB) && b200_aligned16(dAB), "gemm_swiglu_bwd: pointers must be 16-B aligned");
  return gemm_ex_impl(dY, lddy, 0, W, ldw, 1, K, U, ldu, 0, A, lda, 0, r, nullptr, AB, ldab, dAB, lddab, M, F, 0, 0, 0,
                      0, 1.0f, 0, stream, nullptr, true);
}

// (a | b) = x·[W1; W2]^T (+ LoRA) and h = silu(a) * b in one launch. Wil holds the rows of W1 and W2 interleaved in
// blocks of 128 (tile = 128 columns of a, then the same 128 of b); T2 = [t1 | t2] (t_i = s x A_i), Bil [2r, 2F] is
// the matching column-interleaved block matrix (rows 0..r-1: B1 in the a columns, zeros in the b columns; rows
// r..2r-1 the other way round), so the tile's a half gets t1 B1 and its b half t2 B2.
extern "C" int b200_gemm_swiglu_fwd_bf16(const void* X, int64_t ldx, const void* Wil, int64_t ldw, int64_t K,
                                         const void* T2, int64_t ldt, const void* Bil, int64_t ldb, int64_t r2,
                                         void* AB, int64_t ldab, void* H, int64_t ldh, int64_t M, int64_t F,
                                         void* stream) {
  B200_CHECK_ARG(AB && H && M >= 256 && F >= 128 && F % 128 == 0 && ldab >= 2 * F && ldh >= F && ldab % 8 == 0 && ldh % 8 == 0,
                 "gemm_swiglu_fwd: needs M >= 256, F a multiple of 128, a packed [M, 2F] (a | b) buffer and an [M, F] h buffer");
  B200_CHECK_ARG(b200_aligned16(AB) && b200_aligned16(H), "gemm_swiglu_fwd: pointers must be 16-B aligned");
  B200_CHECK_ARG(r2 >= 0 && r2 % 8 == 0 && r2 <= 128, "gemm_swiglu_fwd: 2r=%lld must be a multiple of 8 and <= 128", (long long)r2);
  return gemm_ex_impl(X, ldx, 0, Wil, ldw, 0, K, T2, ldt, 0, Bil, ldb, 1, r2, nullptr, H, ldh, AB, ldab, M, 2 * F, 0, 0, 0,
                      0, 1.0f, 0, stream, nullptr, false, true);
}

// h = gelu(z), z = x W^T + bias (+ T Bl): the first linear of GPT-2's MLP with nn.GELU (GPT2.py:58-62) in the epilogue;
// z (bf16) is kept for backward.
extern "C" int b200_gemm_gelu_fwd_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K, const void* T,
                                       int64_t ldt, const void* Bl, int64_t ldb, int64_t r, const void* bias, void* Z,
                                       int64_t ldz, void* H, int64_t ldh, int64_t M, int64_t N, void* stream) {
  B200_CHECK_ARG(Z && H && ldz >= N && ldh >= N, "gemm_gelu_fwd: null output or pitch < N");
  Extra ex;
  ex.Z = Z;
  ex.ldz = ldz;
  return gemm_ex_impl(X, ldx, 0, W, ldw, 0, K, r > 0 ? T : nullptr, ldt, 0, r > 0 ? Bl : nullptr, ldb, 1, r > 0 ? r : 0, bias,
                      nullptr, 0, H, ldh, M, N, (bias ? B200_GEMM_BIAS : 0) | B200_GEMM_GELU, 0, 0, 0, 1.0f, 0, stream, nullptr,
                      false, false, &ex);
}

// dZ = gelu'(Z) * (dY W + U A^T): the input gradient of GPT-2's second MLP linear with the GELU backward in the epilogue
// (W [K, N] = layers.2.weight, read as an MN-major operand; U [M, r], A [N, r] the LoRA pair or NULL).
extern "C" int b200_gemm_gelu_bwd_bf16(const void* dY, int64_t lddy, const void* W, int64_t ldw, int64_t K, const void* U,
                                       int64_t ldu, const void* A, int64_t lda, int64_t r, const void* Z, int64_t ldz,
                                       void* dZ, int64_t lddz, int64_t M, int64_t N, void* stream) {
  B200_CHECK_ARG(Z && dZ && M >= 256 && N >= 256, "gemm_gelu_bwd: needs Z, dZ and M, N >= 256");
  Extra ex;
  ex.gelu_bwd = true;
  return gemm_ex_impl(dY, lddy, 0, W, ldw, 1, K, r > 0 ? U : nullptr, ldu, 0, r > 0 ? A : nullptr, lda, 0, r > 0 ? r : 0, nullptr,
                      Z, ldz, dZ, lddz, M, N, 0, 0, 0, 0, 1.0f, 0, stream, nullptr, false, false, &ex);
}

// Y = R + dropout(X W^T + bias (+ T Bl)): the block's residual add with drop_shortcut (GPT2.py:85,87) in the epilogue of
// out_proj / the second MLP linear. keep(i) is the mask b200_dropout_bf16 draws for element i = row * N + col of the
// linear's output with the same (seed, seed_dev), so backward is b200_dropout_bf16 on the incoming gradient.
extern "C" int b200_gemm_dropout_residual_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K,
                                               const void* T, int64_t ldt, const void* Bl, int64_t ldb, int64_t r,
                                               const void* bias, const void* R, int64_t ldr, void* Y, int64_t ldy,
                                               int64_t M, int64_t N, float p, uint64_t seed, const uint64_t* seed_dev,
                                               void* stream) {
  B200_CHECK_ARG(R && Y && M >= 256 && N >= 256, "gemm_dropout_residual: needs R, Y and M, N >= 256");
  Extra ex;
  ex.dropout = true;
  ex.drop_p = p;
  ex.drop_seed = seed;
  ex.drop_seed_dev = seed_dev;
  return gemm_ex_impl(X, ldx, 0, W, ldw, 0, K, r > 0 ? T : nullptr, ldt, 0, r > 0 ? Bl : nullptr, ldb, 1, r > 0 ? r : 0, bias, R,
                      ldr, Y, ldy, M, N, bias ? B200_GEMM_BIAS : 0, 0, 0, 0, 1.0f, 0, stream, nullptr, false, false, &ex);
}

// Several LoRA-wrapped linears of the same input as ONE GEMM: W is the row-concatenation of the base
// weights [sum(out_i), K], T = [t_0 | t_1 | ...] with t_i = s x A_i [M, r], Bcat = [B_0 | B_1 | ...]
// [r, sum(out_i)]; output columns of segment i get T[:, i*r:(i+1)*r] Bcat[:, segment i] added.
extern "C" int b200_gemm_lora_packed_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K,
                                          const void* T, int64_t ldt, const void* Bcat, int64_t ldb,
                                          int64_t r, int nseg, const int64_t* seg_start, void* C,
                                          int64_t ldc, int64_t M, int64_t N, void* stream) {
  B200_CHECK_ARG(nseg >= 1 && nseg <= 4 && seg_start && seg_start[0] == 0, "gemm_lora_packed: 1..4 segments, the first starting at column 0");
  B200_CHECK_ARG(r > 0 && r <= 64 && r % 8 == 0, "gemm_lora_packed: rank %lld must be a multiple of 8 and <= 64", (long long)r);
  SegInfo si;
  si.nseg = nseg;
  si.a2_stride = r;
  for (int i = 0; i < nseg; ++i) {
    B200_CHECK_ARG(seg_start[i] % P_BN == 0 && seg_start[i] < N && (i == 0 || seg_start[i] > seg_start[i - 1]),
                   "gemm_lora_packed: segment starts must be increasing multiples of %d below N", P_BN);
    si.start[i] = seg_start[i];
  }
  B200_CHECK_ARG(M >= 256 && N >= 256, "gemm_lora_packed: needs M, N >= 256");
  return gemm_ex_impl(X, ldx, 0, W, ldw, 0, K, T, ldt, 0, Bcat, ldb, 1, r, nullptr, nullptr, 0, C, ldc, M, N, 0, 0, 0,
                      0, 1.0f, 0, stream, &si);
}

// The packed q | k | v projection of an attention block (b200_gemm_lora_packed_bf16) with compute_rope
// (Models/Llama/common_components.py:6-35, called at Llama3.py:140-141) applied to the q and k heads — the first rope_cols
// output columns — in the epilogue: row m is position m % seq_len, cos / sin are [>= seq_len, head_dim] bf16 tables
// whose two column halves are equal. T / Bcat may be NULL with r = 0 (plain linears).
extern "C" int b200_gemm_qkv_rope_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K,
                                       const void* T, int64_t ldt, const void* Bcat, int64_t ldb, int64_t r,
                                       int nseg, const int64_t* seg_start, const void* cos, const void* sin,
                                       int64_t seq_len, int64_t head_dim, int64_t rope_cols, void* C, int64_t ldc,
                                       int64_t M, int64_t N, void* stream) {
  B200_CHECK_ARG(cos && sin, "gemm_qkv_rope: null RoPE tables");
  B200_CHECK_ARG(M >= 256 && N >= 256, "gemm_qkv_rope: needs M, N >= 256");
  Extra ex;
  ex.rope_cos = cos;
  ex.rope_sin = sin;
  ex.rope_T = seq_len;
  ex.rope_hd = head_dim;
  ex.rope_cols = rope_cols;
  if (r <= 0)
    return gemm_ex_impl(X, ldx, 0, W, ldw, 0, K, nullptr, 0, 0, nullptr, 0, 0, 0, nullptr, nullptr, 0, C, ldc, M, N, 0, 0, 0, 0,
                        1.0f, 0, stream, nullptr, false, false, &ex);
  B200_CHECK_ARG(nseg >= 1 && nseg <= 4 && seg_start && seg_start[0] == 0, "gemm_qkv_rope: 1..4 segments, the first starting at column 0");
  B200_CHECK_ARG(r <= 64 && r % 8 == 0, "gemm_qkv_rope: rank %lld must be a multiple of 8 and <= 64", (long long)r);
  SegInfo si;
  si.nseg = nseg;
  si.a2_stride = r;
  for (int i = 0; i < nseg; ++i) {
    B200_CHECK_ARG(seg_start[i] % P_BN == 0 && seg_start[i] < N && (i == 0 || seg_start[i] > seg_start[i - 1]),
                   "gemm_qkv_rope: segment starts must be increasing multiples of %d below N", P_BN);
    si.start[i] = seg_start[i];
  }
  return gemm_ex_impl(X, ldx, 0, W, ldw, 0, K, T, ldt, 0, Bcat, ldb, 1, r, nullptr, nullptr, 0, C, ldc, M, N, 0, 0, 0, 0, 1.0f,
                      0, stream, &si, false, false, &ex);
}

namespace {
// K splits the tile-scheduled grouped kernel gives a problem: only accumulating problems may split
int grouped_splits(const b200_gemm_problem& q, int sms, long long tiles_total) {
  if (!q.accumulate) return 1;
  const int m_tiles = (int)((q.M + BM - 1) / BM), kb = (int)((q.K + BK - 1) / BK);
  const long long target = (long long)2 * sms * m_tiles / (tiles_total > 0 ? tiles_total : 1);  // ~2 waves overall
  int splits = (int)(target / m_tiles);
  if (splits > kb / 2) splits = kb / 2;   // keep >= 2 K blocks per work item
  return splits < 1 ? 1 : splits;
}
}  // namespace

// One launch for up to B200_GEMM_GROUP_MAX skinny problems (every N <= 64, single operand pair).
extern "C" int b200_gemm_grouped_bf16(int nprob, const b200_gemm_problem* probs, void* stream) {
  B200_CHECK_ARG(nprob >= 1 && nprob <= GROUP_MAX && probs, "gemm_grouped: nprob=%d must be in [1, %d]", nprob, GROUP_MAX);
  static_assert(B200_GEMM_GROUP_MAX == GROUP_MAX, "header / kernel group size mismatch");
  GroupedParams g;
  GroupedMaps maps;
  g.nprob = nprob;
  const int sms = b200_num_sms();
  long long tiles_total = 0;
  for (int i = 0; i < nprob; ++i) {
    const b200_gemm_problem& q = probs[i];
    B200_CHECK_ARG(q.A && q.B && q.C && q.M > 0 && q.N > 0 && q.K > 0, "gemm_grouped: problem %d has a null operand or empty size", i);
    B200_CHECK_ARG(q.N <= G_BN, "gemm_grouped: problem %d has N=%lld > %d", i, (long long)q.N, G_BN);
    B200_CHECK_ARG(!(q.accumulate && !q.out_f32), "gemm_grouped: accumulate needs fp32 output (problem %d)", i);
    tiles_total += (q.M + BM - 1) / BM;
  }
  // give every problem a share of the SMs proportional to its tile count when it may split K
  int work = 0;
  for (int i = 0; i < nprob; ++i) {
    const b200_gemm_problem& q = probs[i];
    GemmParams& p = g.p[i];
    p.M = (int)q.M; p.N = (int)q.N;
    p.kb1 = (int)((q.K + BK - 1) / BK); p.kb2 = 0;
    p.a1_mn = q.a_mn; p.b1_mn = q.b_mn; p.a2_mn = p.b2_mn = 0;
    p.m_tiles = (int)((q.M + BM - 1) / BM); p.n_tiles = 1; p.n_fastest = 0;
    p.tail_start = p.m_tiles; p.tail_splits = 1; p.tail_ws = nullptr; p.tail_cnt = nullptr;
    p.alpha = q.alpha; p.bias = nullptr; p.R = nullptr; p.ldr = 0;
    p.C = q.C; p.ldc = q.ldc;
    p.flags = (q.out_f32 ? EP_OUT_F32 : 0) | (q.trans_out ? EP_TRANS_OUT : 0);
    if (q.accumulate) p.flags |= EP_ATOMIC;
    const int splits = grouped_splits(q, sms, tiles_total);
    p.splits = splits;
    work += p.m_tiles * splits;
    g.work_end[i] = work;
    Operand oa{q.A, q.lda, q.a_mn}, ob{q.B, q.ldb, q.b_mn};
    int rc;
    if ((rc = make_operand_map(&maps.a[i], oa, q.M, q.K, BM))) return rc;
    if ((rc = make_operand_map(&maps.b[i], ob, q.N, q.K, G_BN))) return rc;
  }
  for (int i = nprob; i < GROUP_MAX; ++i) g.work_end[i] = work;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_grouped_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    attr_set[dev] = true;
  }
  const int grid = work < sms ? work : sms;
  ++g_path[PATH_GROUPED_TILES];
  B200_CHECK_CUDA(b200_launch_pdl(gemm_grouped_kernel, dim3(grid), dim3(256), G_SMEM, (cudaStream_t)stream, maps, g));
  b200_count_launch();
  return 0;
}

namespace {
// workspace = [one int per 128-row tile of every non-accumulating problem | padding to 256 B |
//              one 128 x 64 fp32 tile per such tile]
void streamk_workspace_layout(int nprob, const b200_gemm_problem* probs, long long* n_cnt, long long* n_floats) {
  long long tiles = 0;
  for (int i = 0; i < nprob; ++i)
    if (!probs[i].accumulate) tiles += (probs[i].M + BM - 1) / BM;
  *n_cnt = (tiles + 63) / 64 * 64;
  *n_floats = tiles * (long long)(BM * G_BN);
}
}  // namespace

extern "C" size_t b200_gemm_grouped_workspace_bytes(int nprob, const b200_gemm_problem* probs) {
  if (nprob < 1 || !probs) return 0;
  long long n_cnt, n_floats;
  streamk_workspace_layout(nprob, probs, &n_cnt, &n_floats);
  return (size_t)(n_cnt * 4 + n_floats * 4);
}

namespace {
template <int NB>
int launch_skinny_pair(const GroupedMaps& maps, const PairSkinnyParams& g, int clusters, cudaStream_t st) {
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_skinny_pair_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, PairSkinnyCfg<NB>::SMEM));
    attr_set[dev] = true;
  }
  B200_CHECK_CUDA(b200_launch_pdl(gemm_skinny_pair_kernel<NB>, dim3(2 * clusters), dim3(256), PairSkinnyCfg<NB>::SMEM, st, maps, g));
  return 0;
}

// The group on CTA pairs (gemm_skinny_pair_kernel) when that fits: nothing accumulates into its output, K is at
// least two blocks and not so deep that 148 stream-K CTAs do better, and the tiles fill most of the machine as pairs.
// *launched = 0: not applicable, the caller goes on with stream-K / tile scheduling.
int try_skinny_pair(int nprob, const b200_gemm_problem* probs, void* stream, int* launched) {
  *launched = 0;
  if (gemm_option(OPT_SKINNY_PAIR) == 0 || gemm_option(OPT_STREAMK) == 1) return 0;     // switched off, or stream-K forced
  const int sms = b200_num_sms();
  long long tiles = 0;
  for (int i = 0; i < nprob; ++i) {
    const b200_gemm_problem& q = probs[i];
    if (!q.A || !q.B || !q.C || q.M <= 0 || q.N <= 0 || q.K <= 0 || q.N > G_BN || q.accumulate) return 0;
    if (q.K < 2 * BK || q.K > 8192) return 0;
  }
  PairSkinnyParams g;
  GroupedMaps maps;
  int ng = 0, nbmax = 1;
  for (int i = 0; i < nprob;) {
    int nb = 1;
    while (i + nb < nprob && nb < SP_NB && probs[i + nb].A == probs[i].A && probs[i + nb].lda == probs[i].lda &&
           probs[i + nb].a_mn == probs[i].a_mn && probs[i + nb].M == probs[i].M && probs[i + nb].K == probs[i].K)
      ++nb;
    g.first[ng] = i;
    g.nb[ng] = nb;
    tiles += (probs[i].M + BM - 1) / BM;
    g.tile_end[ng] = (int)tiles;
    if (nb > nbmax) nbmax = nb;
    ++ng;
    i += nb;
  }
  // pairs of CTAs: at least ~2/3 of the SMs busy (fewer tiles with a deep K range do better cut into 148 stream-K
  // ranges; with a short K range — GPT-2's M = 4096, K = 1280: 32 tiles x 20 K blocks — the pair still halves what the
  // tile-scheduled kernel streams per CTA and stream-K's fix-up costs more than the whole problem), at most a few rounds
  long long kb_max = 0;
  for (int i = 0; i < nprob; ++i) kb_max = std::max<long long>(kb_max, (probs[i].K + BK - 1) / BK);
  if ((2 * tiles < (2 * sms) / 3 && !(tiles >= 8 && kb_max <= 48)) || tiles > 4LL * sms) return 0;
  g.ngroups = ng;
  for (int i = ng; i < GROUP_MAX; ++i) {
    g.tile_end[i] = (int)tiles;
    g.first[i] = 0;
    g.nb[i] = 0;
  }
  for (int i = 0; i < nprob; ++i) {
    const b200_gemm_problem& q = probs[i];
    B200_CHECK_ARG(!(q.accumulate && !q.out_f32), "gemm_grouped: accumulate needs fp32 output (problem %d)", i);
    GemmParams& p = g.p[i];
    p.M = (int)q.M; p.N = (int)q.N;
    p.kb1 = (int)((q.K + BK - 1) / BK); p.kb2 = 0;
    p.a1_mn = q.a_mn; p.b1_mn = q.b_mn; p.a2_mn = p.b2_mn = 0;
    p.m_tiles = (int)((q.M + BM - 1) / BM); p.n_tiles = 1; p.n_fastest = 0;
    p.tail_start = p.m_tiles; p.tail_splits = 1; p.tail_ws = nullptr; p.tail_cnt = nullptr; p.splits = 1;
    p.alpha = q.alpha; p.bias = nullptr; p.R = nullptr; p.ldr = 0;
    p.C = q.C; p.ldc = q.ldc;
    p.flags = (q.out_f32 ? EP_OUT_F32 : 0) | (q.trans_out ? EP_TRANS_OUT : 0);
    Operand oa{q.A, q.lda, q.a_mn}, ob{q.B, q.ldb, q.b_mn};
    int rc;
    if ((rc = make_operand_map(&maps.a[i], oa, q.M, q.K, BM))) return rc;
    if ((rc = make_operand_map(&maps.b[i], ob, q.N, q.K, G_BN))) return rc;
  }
  int rc;
  if (nbmax == 1) rc = launch_skinny_pair<1>(maps, g, (int)tiles, (cudaStream_t)stream);
  else if (nbmax == 2) rc = launch_skinny_pair<2>(maps, g, (int)tiles, (cudaStream_t)stream);
  else rc = launch_skinny_pair<3>(maps, g, (int)tiles, (cudaStream_t)stream);
  if (rc) return rc;
  ++g_path[PATH_SKINNY_PAIR];
  b200_count_launch();
  *launched = 1;
  return 0;
}
}  // namespace

extern "C" int b200_gemm_grouped_streamk_bf16(int nprob, const b200_gemm_problem* probs, void* workspace,
                                              size_t workspace_bytes, void* stream) {
  B200_CHECK_ARG(nprob >= 1 && nprob <= GROUP_MAX && probs, "gemm_grouped_streamk: nprob=%d must be in [1, %d]", nprob, GROUP_MAX);
  {
    int launched = 0;
    const int rc = try_skinny_pair(nprob, probs, stream, &launched);
    if (rc || launched) return rc;
  }
  StreamKParams g;
  GroupedMaps maps;
  g.nprob = nprob;
  long long n_cnt, n_floats;
  streamk_workspace_layout(nprob, probs, &n_cnt, &n_floats);
  const size_t need = (size_t)(n_cnt * 4 + n_floats * 4);
  B200_CHECK_ARG(need == 0 || (workspace && workspace_bytes >= need && ((uintptr_t)workspace & 255) == 0),
                 "gemm_grouped_streamk: workspace of %zu bytes (256-B aligned, zero-filled) needed, got %zu", need, workspace_bytes);
  g.cnt = (int*)workspace;
  g.ws = (float*)((char*)workspace + n_cnt * 4);
  long long units = 0, cnt_off = 0;
  for (int i = 0; i < nprob; ++i) {
    const b200_gemm_problem& q = probs[i];
    B200_CHECK_ARG(q.A && q.B && q.C && q.M > 0 && q.N > 0 && q.K > 0, "gemm_grouped_streamk: problem %d has a null operand or empty size", i);
    B200_CHECK_ARG(q.N <= G_BN, "gemm_grouped_streamk: problem %d has N=%lld > %d", i, (long long)q.N, G_BN);
    B200_CHECK_ARG(!(q.accumulate && !q.out_f32), "gemm_grouped_streamk: accumulate needs fp32 output (problem %d)", i);
    GemmParams& p = g.p[i];
    p.M = (int)q.M; p.N = (int)q.N;
    p.kb1 = (int)((q.K + BK - 1) / BK); p.kb2 = 0;
    p.a1_mn = q.a_mn; p.b1_mn = q.b_mn; p.a2_mn = p.b2_mn = 0;
    p.m_tiles = (int)((q.M + BM - 1) / BM); p.n_tiles = 1; p.n_fastest = 0;
    p.tail_start = p.m_tiles; p.tail_splits = 1; p.tail_ws = nullptr; p.tail_cnt = nullptr; p.splits = 1;
    p.alpha = q.alpha; p.bias = nullptr; p.R = nullptr; p.ldr = 0;
    p.C = q.C; p.ldc = q.ldc;
    p.flags = (q.out_f32 ? EP_OUT_F32 : 0) | (q.trans_out ? EP_TRANS_OUT : 0) | (q.accumulate ? EP_ATOMIC : 0);
    units += (long long)p.m_tiles * p.kb1;
    B200_CHECK_ARG(units < (1ll << 31), "gemm_grouped_streamk: group too large (problem %d)", i);
    g.unit_end[i] = (int)units;
    g.cnt_off[i] = (int)cnt_off;
    g.ws_off[i] = cnt_off * (long long)(BM * G_BN);
    if (!q.accumulate) cnt_off += p.m_tiles;
    Operand oa{q.A, q.lda, q.a_mn}, ob{q.B, q.ldb, q.b_mn};
    int rc;
    if ((rc = make_operand_map(&maps.a[i], oa, q.M, q.K, BM))) return rc;
    if ((rc = make_operand_map(&maps.b[i], ob, q.N, q.K, G_BN))) return rc;
  }
  for (int i = nprob; i < GROUP_MAX; ++i) {
    g.unit_end[i] = (int)units;
    g.cnt_off[i] = 0;
    g.ws_off[i] = 0;
  }
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_streamk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    attr_set[dev] = true;
  }
  const int sms = b200_num_sms();
  {
    // Stream-K pays ~16 K blocks' worth of time for the shared-tile fix-up (RED, fence, counter, reload).
    // Replay the tile scheduler's round-robin on the host: if its busiest CTA is not clearly worse
    // than an even cut, the tiles go out whole (t = x A with K = 2048 on 64 tiles: 19 vs 23 us).
    const int sk_opt = gemm_option(OPT_STREAMK);
    long long tiles_total = 0;
    for (int i = 0; i < nprob; ++i) tiles_total += (probs[i].M + BM - 1) / BM;
    long long items = 0;
    for (int i = 0; i < nprob; ++i) items += (long long)g.p[i].m_tiles * grouped_splits(probs[i], sms, tiles_total);
    const int tgrid = (int)(items < sms ? items : sms);
    std::vector<long long> load((size_t)tgrid, 0);
    long long w = 0;
    for (int i = 0; i < nprob; ++i) {
      const int sp = grouped_splits(probs[i], sms, tiles_total);
      for (int s = 0; s < sp; ++s) {
        const long long n = (long long)(s + 1) * g.p[i].kb1 / sp - (long long)s * g.p[i].kb1 / sp;
        for (int t = 0; t < g.p[i].m_tiles; ++t, ++w) load[(size_t)(w % tgrid)] += n;
      }
    }
    long long busiest = 0;
    for (long long l : load) busiest = l > busiest ? l : busiest;
    const double even = (double)units / sms;
    const bool streamk = sk_opt == 1 || (sk_opt != 0 && (double)busiest > 1.3 * even + 16.0);
    if (!streamk) return b200_gemm_grouped_bf16(nprob, probs, stream);
  }
  // at least 4 K blocks per CTA: below that the fixed cost of a CTA outweighs its share of the stream
  long long grid = units / 4;
  if (grid > sms) grid = sms;
  if (grid < 1) grid = 1;
  ++g_path[PATH_STREAMK];
  B200_CHECK_CUDA(b200_launch_pdl(gemm_streamk_kernel, dim3((unsigned)grid), dim3(256), G_SMEM, (cudaStream_t)stream, maps, g));
  b200_count_launch();
  return 0;
}

extern "C" int b200_gemm_set_option(int option, int value) {
  B200_CHECK_ARG(option >= 0 && option < OPT_N && value >= -1 && value <= 1, "gemm_set_option: option %d / value %d out of range", option, value);
  gemm_option(option);          // make sure the environment defaults are in before overriding one
  g_opt[option] = value;
  return 0;
}

extern "C" int b200_gemm_path_counts(uint64_t* out, int n) {
  B200_CHECK_ARG(out && n >= 0, "gemm_path_counts: null pointer");
  for (int i = 0; i < n; ++i) out[i] = i < PATH_N ? g_path[i].load() : 0;
  return PATH_N;
}

extern "C" size_t b200_gemm_workspace_bytes(void) {
  // one 256 x 256 fp32 slot per (tail tile, K split): up to two rounds of 74 CTA pairs' worth, + the counters
  return (size_t)2 * 74 * 256 * P_BN * sizeof(float) + 4096;
}

extern "C" int b200_gemm_set_workspace(void* workspace, size_t bytes) {
  int dev = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < 64, "gemm_set_workspace: device index %d out of range", dev);
  B200_CHECK_ARG(workspace == nullptr || (((uintptr_t)workspace & 255) == 0 && bytes >= 4096),
                 "gemm_set_workspace: workspace must be 256-B aligned and at least 4096 bytes");
  std::lock_guard<std::mutex> lk(g_ws_mu);
  g_tail_ws[dev] = TailWs{workspace, workspace ? bytes : 0};
  return 0;
}

extern "C" int b200_gemm_set_stream_workspace(void* stream, void* workspace, size_t bytes) {
  int dev = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(workspace == nullptr || (((uintptr_t)workspace & 255) == 0 && bytes >= 4096),
                 "gemm_set_stream_workspace: workspace must be 256-B aligned and at least 4096 bytes");
  std::lock_guard<std::mutex> lk(g_ws_mu);
  for (size_t i = 0; i < g_stream_ws.size(); ++i)
    if (g_stream_ws[i].dev == dev && g_stream_ws[i].stream == stream) {
      if (workspace) g_stream_ws[i].ws = TailWs{workspace, bytes};
      else g_stream_ws.erase(g_stream_ws.begin() + (long)i);
      return 0;
    }
  if (workspace) g_stream_ws.push_back(StreamWs{dev, stream, TailWs{workspace, bytes}});
  return 0;
}

extern "C" int b200_gemm_nt_bf16(const void* A1, int64_t lda1, const void* B1, int64_t ldb1,
                                 int64_t K1, const void* A2, int64_t lda2, const void* B2,
                                 int64_t ldb2, int64_t K2, const void* bias, const void* R,
                                 int64_t ldr, void* C, int64_t ldc, int64_t M, int64_t N, int flags,
                                 void* stream) {
  return b200_gemm_ex(A1, lda1, 0, B1, ldb1, 0, K1, A2, lda2, 0, B2, ldb2, 0, K2, bias, R, ldr, C,
                      ldc, M, N, flags, 0, 0, 0, 1.0f, 0, stream);
}

extern "C" int b200_gemm_bf16(const void* A, int64_t lda, int transa, const void* B, int64_t ldb,
                              int transb, void* C, int64_t ldc, int out_f32, int accumulate,
                              float alpha, int64_t M, int64_t N, int64_t K, void* stream) {
  return b200_gemm_ex(A, lda, transa, B, ldb, transb, K, nullptr, 0, 0, nullptr, 0, 0, 0, nullptr,
                      nullptr, 0, C, ldc, M, N, 0, out_f32, accumulate, 0, alpha, 0, stream);
}
