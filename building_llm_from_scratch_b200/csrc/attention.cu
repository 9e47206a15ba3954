// K2 — fused causal MHA / GQA attention: C-ABI entry points and the head_dim-128 kernels.
//
// Flash-style: the T x T score matrix of Models/Llama/Llama3.py:148-153 / Models/GPT2/GPT2.py:38-46
// is never materialised, repeat_interleave of K/V (Llama3.py:144-145) is replaced by indexing the kv
// head, and the causal mask buffer by a predicate.
//
// head_dim 64 (GPT-2, Llama-3.2-1B) dispatches to the tcgen05 / TMEM kernels in attention_tc.cu.
// head_dim 128 (Llama-2-7B, Llama-3-8B) still runs the kernels in this file on the legacy warp-level
// tensor path (mma.sync m16n8k16, bf16 in / fp32 accumulate) with cp.async double-buffered,
// XOR-swizzled shared-memory tiles; one warp owns 16 query rows (forward, dQ) or 16 key rows
// (dK/dV), so no cross-warp reductions or atomics are needed and the result is deterministic. GQA:
// the dK/dV CTA loops over the query heads of its group. B200_ATTN_LEGACY=1 forces these kernels for
// head_dim 64 as well (A/B comparisons).
//
// Roofline: tensor pipe. Algorithmic FLOPs (causal): fwd 2*T*T*hd per (batch, head),
// bwd 2.5x that (recompute included: 3.5x executed).
#include <cstdlib>

#include "common.cuh"

// tcgen05 paths (attention_tc.cu), head_dim 64 and 128
int b200_attention_fwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, void* o, int64_t ldo, float* lse, int64_t B, int64_t T,
                          int64_t h, int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed,
                          cudaStream_t st);

size_t b200_attention_bwd_tc_workspace(int64_t B, int64_t T, int64_t h);
int b200_attention_bwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, const void* o, int64_t ldo, const void* dout, int64_t lddo,
                          const float* lse, void* workspace, void* dq, int64_t lddq, void* dk,
                          int64_t lddk, void* dv, int64_t lddv, int64_t B, int64_t T, int64_t h,
                          int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed, cudaStream_t st);

static bool attn_legacy() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_ATTN_LEGACY");
    v = (e && e[0] == '1') ? 1 : 0;
  }
  return v == 1;
}

namespace {

constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Shared-memory tile of ROWS x HD bf16, 16-byte chunks XOR-swizzled by (row & 7).
template <int HD>
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
  return base + (uint32_t)(row * (HD * 2) + (((chunk & ~7) | ((chunk ^ row) & 7)) << 4));
}

// Cooperative load of ROWS rows (row0.. of a [*, ld] matrix, HD columns from g) into a tile.
template <int ROWS, int HD, int NTHREADS>
__device__ __forceinline__ void load_tile(uint32_t tile, const bf16* g, long long ld, int row0,
                                          int row_limit) {
  constexpr int CPR = HD / 8;
  constexpr int TOTAL = ROWS * CPR;
  for (int c = threadIdx.x; c < TOTAL; c += NTHREADS) {
    const int r = c / CPR, ch = c % CPR;
    const bool ok = (row0 + r) < row_limit;
    const bf16* src = g + (long long)(ok ? row0 + r : 0) * ld + ch * 8;
    const uint32_t dst = tile_addr<HD>(tile, r, ch);
    const int sz = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
  }
}

// A-operand fragments (16 rows x HD) of rows [row0, row0+16) of a tile
template <int HD>
__device__ __forceinline__ void load_a_frags(uint32_t (*f)[4], uint32_t tile, int row0, int lane) {
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks)
    ldsm_x4(f[ks], tile_addr<HD>(tile, row0 + (lane & 15), ks * 2 + (lane >> 4)));
}

// acc[8][4] (16 x 64) = A(16 x HD, frags) * Tile(64 x HD)^T       (tile rows are the N index)
template <int HD>
__device__ __forceinline__ void mma_a_tile_nt(float (*acc)[4], const uint32_t (*af)[4], uint32_t tile,
                                              int lane) {
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b[4];
      ldsm_x4(b, tile_addr<HD>(tile, np * 16 + (lane & 7) + ((lane >> 4) << 3), ks * 2 + ((lane >> 3) & 1)));
      mma16816(acc[2 * np], af[ks], b[0], b[1]);
      mma16816(acc[2 * np + 1], af[ks], b[2], b[3]);
    }
  }
}

// acc[HD/8][4] (16 x HD) += P(16 x 64, bf16 A-frags pf[4][4]) * Tile(64 x HD)   (tile rows = K index)
template <int HD>
__device__ __forceinline__ void mma_p_tile_nn(float (*acc)[4], const uint32_t (*pf)[4], uint32_t tile,
                                              int lane) {
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
    for (int dp = 0; dp < HD / 16; ++dp) {
      uint32_t b[4];
      ldsm_x4_t(b, tile_addr<HD>(tile, kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), dp * 2 + (lane >> 4)));
      mma16816(acc[2 * dp], pf[kk], b[0], b[1]);
      mma16816(acc[2 * dp + 1], pf[kk], b[2], b[3]);
    }
  }
}

// fp32 C-fragments of a 16 x 64 tile -> bf16 A-fragments (4 k-steps of 16)
__device__ __forceinline__ void c_to_a_frags(uint32_t (*pf)[4], const float (*s)[4]) {
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    pf[kk][0] = f2_to_bf2(s[2 * kk][0], s[2 * kk][1]);
    pf[kk][1] = f2_to_bf2(s[2 * kk][2], s[2 * kk][3]);
    pf[kk][2] = f2_to_bf2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
    pf[kk][3] = f2_to_bf2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
  }
}

// =============================== forward ===============================
template <int HD>
__global__ void __launch_bounds__(256)
attn_fwd_kernel(const bf16* __restrict__ q, long long ldq, const bf16* __restrict__ k, long long ldk,
                const bf16* __restrict__ v, long long ldv, bf16* __restrict__ o, long long ldo,
                float* __restrict__ lse, int T, int H, int HKV, float scale) {
  constexpr int BQ = 128, BKV = 64;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sK = sQ + BQ * HD * 2;
  const uint32_t sV = sK + 2 * BKV * HD * 2;

  const int qt = gridDim.x - 1 - blockIdx.x;  // heavy tiles first
  const int head = blockIdx.y, b = blockIdx.z;
  const int kvh = head / (H / HKV);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = qt * BQ;
  const long long tok0 = (long long)b * T;
  const bf16* qg = q + tok0 * ldq + (long long)head * HD;
  const bf16* kg = k + tok0 * ldk + (long long)kvh * HD;
  const bf16* vg = v + tok0 * ldv + (long long)kvh * HD;

  const int q_hi = min(q0 + BQ, T);                // exclusive
  const int ntiles = (q_hi + BKV - 1) / BKV;       // kv tiles 0 .. ntiles-1 cover keys < q_hi

  load_tile<BQ, HD, 256>(sQ, qg, ldq, q0, T);
  load_tile<BKV, HD, 256>(sK, kg, ldk, 0, T);
  load_tile<BKV, HD, 256>(sV, vg, ldv, 0, T);
  cp_async_commit();

  uint32_t qf[HD / 16][4];
  float oacc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const float sl2 = scale * LOG2E;
  const int wrow0 = q0 + warp * 16;  // first query row of this warp
  const int r0 = wrow0 + (lane >> 2), r1 = r0 + 8;

  for (int j = 0; j < ntiles; ++j) {
    const int buf = j & 1;
    if (j + 1 < ntiles) {
      load_tile<BKV, HD, 256>(sK + (buf ^ 1) * BKV * HD * 2, kg, ldk, (j + 1) * BKV, T);
      load_tile<BKV, HD, 256>(sV + (buf ^ 1) * BKV * HD * 2, vg, ldv, (j + 1) * BKV, T);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (j == 0) load_a_frags<HD>(qf, sQ, warp * 16, lane);
    const int kv0 = j * BKV;
    if (kv0 <= wrow0 + 15) {  // warp-uniform: tile not entirely above the diagonal
      float s[8][4];
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
      mma_a_tile_nt<HD>(s, qf, sK + buf * BKV * HD * 2, lane);
      const bool diag = kv0 + BKV - 1 > wrow0;
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float x = s[nt][e] * sl2;
          if (diag) {
            const int col = kv0 + nt * 8 + 2 * (lane & 3) + (e & 1);
            const int row = (e < 2) ? r0 : r1;
            if (col > row) x = -INFINITY;
          }
          s[nt][e] = x;
        }
        mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
        mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      const float a0 = exp2f(m0 - mn0), a1 = exp2f(m1 - mn1);  // m = -inf first time -> 0
      m0 = mn0;
      m1 = mn1;
      float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][0] = exp2f(s[nt][0] - mn0);
        s[nt][1] = exp2f(s[nt][1] - mn0);
        s[nt][2] = exp2f(s[nt][2] - mn1);
        s[nt][3] = exp2f(s[nt][3] - mn1);
        ps0 += s[nt][0] + s[nt][1];
        ps1 += s[nt][2] + s[nt][3];
      }
      l0 = l0 * a0 + ps0;
      l1 = l1 * a1 + ps1;
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        oacc[i][0] *= a0;
        oacc[i][1] *= a0;
        oacc[i][2] *= a1;
        oacc[i][3] *= a1;
      }
      uint32_t pf[4][4];
      c_to_a_frags(pf, s);
      mma_p_tile_nn<HD>(oacc, pf, sV + buf * BKV * HD * 2, lane);
    }
    __syncthreads();  // everyone done with buf before it is refilled two iterations later
  }

  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;
  bf16* og = o + tok0 * ldo + (long long)head * HD;
#pragma unroll
  for (int nt = 0; nt < HD / 8; ++nt) {
    const int col = nt * 8 + 2 * (lane & 3);
    if (r0 < T) *reinterpret_cast<uint32_t*>(og + (long long)r0 * ldo + col) = f2_to_bf2(oacc[nt][0] * inv0, oacc[nt][1] * inv0);
    if (r1 < T) *reinterpret_cast<uint32_t*>(og + (long long)r1 * ldo + col) = f2_to_bf2(oacc[nt][2] * inv1, oacc[nt][3] * inv1);
  }
  if ((lane & 3) == 0) {
    float* lg = lse + ((long long)b * H + head) * T;
    if (r0 < T) lg[r0] = (m0 + log2f(l0)) / LOG2E;
    if (r1 < T) lg[r1] = (m1 + log2f(l1)) / LOG2E;
  }
}

// =============================== backward: delta ===============================
// delta[b, h, t] = sum_c dO[t, h, c] * O[t, h, c]; one warp per (token, head)
__global__ void __launch_bounds__(256)
attn_delta_kernel(const bf16* __restrict__ o, long long ldo, const bf16* __restrict__ dout,
                  long long lddo, float* __restrict__ delta, long long B, int T, int H, int HD) {
  const long long total = B * T * H;
  const int lane = threadIdx.x & 31;
  for (long long wi = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); wi < total;
       wi += (long long)gridDim.x * 8) {
    const int head = (int)(wi % H);
    const long long tok = wi / H;
    const bf16* op = o + tok * ldo + (long long)head * HD;
    const bf16* dp = dout + tok * lddo + (long long)head * HD;
    float acc = 0.f;
    for (int c = lane * 8; c < HD; c += 256) {
      float a[8], g[8];
      unpack8(ld_nc_v4(op + c), a);
      unpack8(ld_nc_v4(dp + c), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += a[j] * g[j];
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      const long long bb = tok / T;
      const int t = (int)(tok % T);
      delta[(bb * H + head) * T + t] = acc;
    }
  }
}

// =============================== backward: dQ ===============================
template <int HD>
__global__ void __launch_bounds__(256)
attn_bwd_dq_kernel(const bf16* __restrict__ q, long long ldq, const bf16* __restrict__ k,
                   long long ldk, const bf16* __restrict__ v, long long ldv,
                   const bf16* __restrict__ dout, long long lddo, const float* __restrict__ lse,
                   const float* __restrict__ delta, bf16* __restrict__ dq, long long lddq, int T,
                   int H, int HKV, float scale) {
  constexpr int BQ = 128, BKV = 64;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sdO = sQ + BQ * HD * 2;
  const uint32_t sK = sdO + BQ * HD * 2;
  const uint32_t sV = sK + 2 * BKV * HD * 2;

  const int qt = gridDim.x - 1 - blockIdx.x;
  const int head = blockIdx.y, b = blockIdx.z;
  const int kvh = head / (H / HKV);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = qt * BQ;
  const long long tok0 = (long long)b * T;
  const bf16* qg = q + tok0 * ldq + (long long)head * HD;
  const bf16* dog = dout + tok0 * lddo + (long long)head * HD;
  const bf16* kg = k + tok0 * ldk + (long long)kvh * HD;
  const bf16* vg = v + tok0 * ldv + (long long)kvh * HD;
  const int q_hi = min(q0 + BQ, T);
  const int ntiles = (q_hi + BKV - 1) / BKV;

  load_tile<BQ, HD, 256>(sQ, qg, ldq, q0, T);
  load_tile<BQ, HD, 256>(sdO, dog, lddo, q0, T);
  load_tile<BKV, HD, 256>(sK, kg, ldk, 0, T);
  load_tile<BKV, HD, 256>(sV, vg, ldv, 0, T);
  cp_async_commit();

  uint32_t qf[HD / 16][4], dof[HD / 16][4];
  float acc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
  const float sl2 = scale * LOG2E;
  const int wrow0 = q0 + warp * 16;
  const int r0 = wrow0 + (lane >> 2), r1 = r0 + 8;
  const float* lg = lse + ((long long)b * H + head) * T;
  const float* dg = delta + ((long long)b * H + head) * T;
  const float lse0 = r0 < T ? lg[r0] * LOG2E : INFINITY, lse1 = r1 < T ? lg[r1] * LOG2E : INFINITY;
  const float dl0 = r0 < T ? dg[r0] : 0.f, dl1 = r1 < T ? dg[r1] : 0.f;

  for (int j = 0; j < ntiles; ++j) {
    const int buf = j & 1;
    if (j + 1 < ntiles) {
      load_tile<BKV, HD, 256>(sK + (buf ^ 1) * BKV * HD * 2, kg, ldk, (j + 1) * BKV, T);
      load_tile<BKV, HD, 256>(sV + (buf ^ 1) * BKV * HD * 2, vg, ldv, (j + 1) * BKV, T);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (j == 0) {
      load_a_frags<HD>(qf, sQ, warp * 16, lane);
      load_a_frags<HD>(dof, sdO, warp * 16, lane);
    }
    const int kv0 = j * BKV;
    if (kv0 <= wrow0 + 15) {
      float s[8][4], dp[8][4];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
        dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f;
      }
      mma_a_tile_nt<HD>(s, qf, sK + buf * BKV * HD * 2, lane);
      mma_a_tile_nt<HD>(dp, dof, sV + buf * BKV * HD * 2, lane);
      const bool diag = kv0 + BKV - 1 > wrow0;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int row = (e < 2) ? r0 : r1;
          float p = exp2f(s[nt][e] * sl2 - ((e < 2) ? lse0 : lse1));
          if (diag) {
            const int col = kv0 + nt * 8 + 2 * (lane & 3) + (e & 1);
            if (col > row) p = 0.f;
          }
          s[nt][e] = p * (dp[nt][e] - ((e < 2) ? dl0 : dl1)) * scale;  // dS
        }
      }
      uint32_t pf[4][4];
      c_to_a_frags(pf, s);
      mma_p_tile_nn<HD>(acc, pf, sK + buf * BKV * HD * 2, lane);
    }
    __syncthreads();
  }
  bf16* dqg = dq + tok0 * lddq + (long long)head * HD;
#pragma unroll
  for (int nt = 0; nt < HD / 8; ++nt) {
    const int col = nt * 8 + 2 * (lane & 3);
    if (r0 < T) *reinterpret_cast<uint32_t*>(dqg + (long long)r0 * lddq + col) = f2_to_bf2(acc[nt][0], acc[nt][1]);
    if (r1 < T) *reinterpret_cast<uint32_t*>(dqg + (long long)r1 * lddq + col) = f2_to_bf2(acc[nt][2], acc[nt][3]);
  }
}

// =============================== backward: dK, dV ===============================
// One CTA owns 64 keys of one kv head (4 warps x 16 keys) and streams every (query head of the
// group) x (query tile at or below the diagonal) through shared memory.
template <int HD>
__global__ void __launch_bounds__(128)
attn_bwd_dkv_kernel(const bf16* __restrict__ q, long long ldq, const bf16* __restrict__ k,
                    long long ldk, const bf16* __restrict__ v, long long ldv,
                    const bf16* __restrict__ dout, long long lddo, const float* __restrict__ lse,
                    const float* __restrict__ delta, bf16* __restrict__ dk, long long lddk,
                    bf16* __restrict__ dv, long long lddv, int T, int H, int HKV, float scale) {
  constexpr int BQ = 64, BKV = 64;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sK = smem_u32(smem);
  const uint32_t sV = sK + BKV * HD * 2;
  const uint32_t sQ = sV + BKV * HD * 2;          // [2][BQ][HD]
  const uint32_t sdO = sQ + 2 * BQ * HD * 2;      // [2][BQ][HD]
  float* sLse = (float*)(smem + 2 * BKV * HD * 2 + 4 * BQ * HD * 2);  // [2][BQ]
  float* sDel = sLse + 2 * BQ;                                        // [2][BQ]

  const int kt = blockIdx.x;  // light tiles (late keys) have fewer q tiles; launch order is fine
  const int kvh = blockIdx.y, b = blockIdx.z;
  const int G = H / HKV;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kv0 = kt * BKV;
  const long long tok0 = (long long)b * T;
  const bf16* kg = k + tok0 * ldk + (long long)kvh * HD;
  const bf16* vg = v + tok0 * ldv + (long long)kvh * HD;
  const int nq_tiles = (T + BQ - 1) / BQ;
  const int nq = nq_tiles - kt;  // q tiles kt .. nq_tiles-1
  const int iters = G * nq;

  auto issue = [&](int it, int buf) {
    const int hh = kvh * G + it / nq;
    const int qi = kt + it % nq;
    const bf16* qg = q + tok0 * ldq + (long long)hh * HD;
    const bf16* dog = dout + tok0 * lddo + (long long)hh * HD;
    load_tile<BQ, HD, 128>(sQ + buf * BQ * HD * 2, qg, ldq, qi * BQ, T);
    load_tile<BQ, HD, 128>(sdO + buf * BQ * HD * 2, dog, lddo, qi * BQ, T);
    if (threadIdx.x < BQ) {
      const int t = qi * BQ + threadIdx.x;
      const long long base = ((long long)b * H + hh) * T;
      sLse[buf * BQ + threadIdx.x] = t < T ? lse[base + t] * LOG2E : INFINITY;
      sDel[buf * BQ + threadIdx.x] = t < T ? delta[base + t] : 0.f;
    }
  };

  load_tile<BKV, HD, 128>(sK, kg, ldk, kv0, T);
  load_tile<BKV, HD, 128>(sV, vg, ldv, kv0, T);
  issue(0, 0);
  cp_async_commit();

  uint32_t kf[HD / 16][4], vf[HD / 16][4];
  float dkacc[HD / 8][4], dvacc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) {
    dkacc[i][0] = dkacc[i][1] = dkacc[i][2] = dkacc[i][3] = 0.f;
    dvacc[i][0] = dvacc[i][1] = dvacc[i][2] = dvacc[i][3] = 0.f;
  }
  const float sl2 = scale * LOG2E;
  const int key0 = kv0 + warp * 16 + (lane >> 2), key1 = key0 + 8;

  for (int it = 0; it < iters; ++it) {
    const int buf = it & 1;
    if (it + 1 < iters) {
      issue(it + 1, buf ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (it == 0) {
      load_a_frags<HD>(kf, sK, warp * 16, lane);
      load_a_frags<HD>(vf, sV, warp * 16, lane);
    }
    const int qi = kt + it % nq;
    const int qrow0 = qi * BQ;
    const uint32_t tQ = sQ + buf * BQ * HD * 2, tdO = sdO + buf * BQ * HD * 2;
    float s[8][4], dp[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
      dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f;
    }
    mma_a_tile_nt<HD>(s, kf, tQ, lane);    // S^T[key, q]
    mma_a_tile_nt<HD>(dp, vf, tdO, lane);  // dP^T[key, q]
    const bool diag = (qi == kt);
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int qc = nt * 8 + 2 * (lane & 3) + (e & 1);
        float p = exp2f(s[nt][e] * sl2 - sLse[buf * BQ + qc]);
        if (diag) {
          const int key = (e < 2) ? key0 : key1;
          if (key > qrow0 + qc) p = 0.f;
        }
        dp[nt][e] = p * (dp[nt][e] - sDel[buf * BQ + qc]) * scale;  // dS^T
        s[nt][e] = p;                                                // P^T
      }
    }
    uint32_t pf[4][4];
    c_to_a_frags(pf, s);
    mma_p_tile_nn<HD>(dvacc, pf, tdO, lane);  // dV += P^T dO
    c_to_a_frags(pf, dp);
    mma_p_tile_nn<HD>(dkacc, pf, tQ, lane);   // dK += dS^T Q
    __syncthreads();
  }

  bf16* dkg = dk + tok0 * lddk + (long long)kvh * HD;
  bf16* dvg = dv + tok0 * lddv + (long long)kvh * HD;
#pragma unroll
  for (int nt = 0; nt < HD / 8; ++nt) {
    const int col = nt * 8 + 2 * (lane & 3);
    if (key0 < T) {
      *reinterpret_cast<uint32_t*>(dkg + (long long)key0 * lddk + col) = f2_to_bf2(dkacc[nt][0], dkacc[nt][1]);
      *reinterpret_cast<uint32_t*>(dvg + (long long)key0 * lddv + col) = f2_to_bf2(dvacc[nt][0], dvacc[nt][1]);
    }
    if (key1 < T) {
      *reinterpret_cast<uint32_t*>(dkg + (long long)key1 * lddk + col) = f2_to_bf2(dkacc[nt][2], dkacc[nt][3]);
      *reinterpret_cast<uint32_t*>(dvg + (long long)key1 * lddv + col) = f2_to_bf2(dvacc[nt][2], dvacc[nt][3]);
    }
  }
}

template <typename K>
int set_smem(K kernel, int bytes, bool* flag) {
  if (!*flag) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    *flag = true;
  }
  return 0;
}

int check_common(const void* q, const void* k, const void* v, int64_t B, int64_t T, int64_t h,
                 int64_t hkv, int64_t hd, const int64_t* lds, int nld) {
  B200_CHECK_ARG(q && k && v && B > 0 && T > 0 && h > 0 && hkv > 0, "attention: null pointer or empty");
  B200_CHECK_ARG(hd == 64 || hd == 128, "attention: head_dim %lld not supported (64 or 128)", (long long)hd);
  B200_CHECK_ARG(h % hkv == 0, "attention: heads %lld not divisible by kv heads %lld", (long long)h, (long long)hkv);
  B200_CHECK_ARG(B <= 65535 && h <= 65535, "attention: batch / heads exceed grid limits");
  for (int i = 0; i < nld; ++i) B200_CHECK_ARG(lds[i] % 8 == 0, "attention: row pitches must be multiples of 8");
  B200_CHECK_ARG(b200_aligned16(q) && b200_aligned16(k) && b200_aligned16(v), "attention: pointers must be 16-B aligned");
  return 0;
}

}  // namespace

extern "C" int b200_attention_fwd_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk,
                                       const void* v, int64_t ldv, void* o, int64_t ldo, float* lse,
                                       int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd,
                                       float scale, float dropout_p, uint64_t seed, void* stream) {
  const int64_t lds[4] = {ldq, ldk, ldv, ldo};
  int rc = check_common(q, k, v, B, T, h, hkv, hd, lds, 4);
  if (rc) return rc;
  B200_CHECK_ARG(o && lse, "attention_fwd: null output");
  cudaStream_t st = (cudaStream_t)stream;
  B200_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "attention: dropout_p=%f must be in [0, 1)", (double)dropout_p);
  if (!attn_legacy())
    return b200_attention_fwd_tc(q, ldq, k, ldk, v, ldv, o, ldo, lse, B, T, h, hkv, hd, scale, dropout_p, seed, st);
  B200_CHECK_ARG(dropout_p == 0.f, "attention: dropout is only implemented on the tcgen05 path");
  dim3 grid((unsigned)((T + 127) / 128), (unsigned)h, (unsigned)B);
  static bool f64[64] = {false}, f128[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (hd == 64) {
    const int smem = (128 + 4 * 64) * 64 * 2;
    if ((rc = set_smem(attn_fwd_kernel<64>, smem, &f64[dev & 63]))) return rc;
    attn_fwd_kernel<64><<<grid, 256, smem, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (bf16*)o, ldo, lse, (int)T, (int)h, (int)hkv, scale);
  } else {
    const int smem = (128 + 4 * 64) * 128 * 2;
    if ((rc = set_smem(attn_fwd_kernel<128>, smem, &f128[dev & 63]))) return rc;
    attn_fwd_kernel<128><<<grid, 256, smem, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (bf16*)o, ldo, lse, (int)T, (int)h, (int)hkv, scale);
  }
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" size_t b200_attention_bwd_workspace_bytes(int64_t B, int64_t T, int64_t h, int64_t hd) {
  (void)hd;   // the tcgen05 layout (lse and delta per 64-query step) also covers the legacy kernels' [B, h, T] floats
  return b200_attention_bwd_tc_workspace(B, T, h);
}

extern "C" int b200_attention_bwd_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk,
                                       const void* v, int64_t ldv, const void* o, int64_t ldo,
                                       const void* dout, int64_t lddo, const float* lse,
                                       void* workspace, size_t workspace_bytes, void* dq,
                                       int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                                       int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd,
                                       float scale, float dropout_p, uint64_t seed, void* stream) {
  float* delta = (float*)workspace;
  const int64_t lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
  int rc = check_common(q, k, v, B, T, h, hkv, hd, lds, 8);
  if (rc) return rc;
  B200_CHECK_ARG(o && dout && lse && delta && dq && dk && dv, "attention_bwd: null pointer");
  B200_CHECK_ARG(workspace_bytes >= b200_attention_bwd_workspace_bytes(B, T, h, hd), "attention_bwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  B200_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "attention: dropout_p=%f must be in [0, 1)", (double)dropout_p);
  if (!attn_legacy())
    return b200_attention_bwd_tc(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, dq, lddq, dk,
                                 lddk, dv, lddv, B, T, h, hkv, hd, scale, dropout_p, seed, st);
  B200_CHECK_ARG(dropout_p == 0.f, "attention: dropout is only implemented on the tcgen05 path");
  {
    const long long total = B * T * h;
    long long blocks = (total + 7) / 8;
    const long long cap = (long long)b200_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    attn_delta_kernel<<<(int)blocks, 256, 0, st>>>((const bf16*)o, ldo, (const bf16*)dout, lddo, delta, B, (int)T, (int)h, (int)hd);
    B200_CHECK_LAUNCH();
  }
  static bool fq64[64] = {false}, fq128[64] = {false}, fk64[64] = {false}, fk128[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  dim3 gq((unsigned)((T + 127) / 128), (unsigned)h, (unsigned)B);
  dim3 gk((unsigned)((T + 63) / 64), (unsigned)hkv, (unsigned)B);
  if (hd == 64) {
    const int smq = (2 * 128 + 4 * 64) * 64 * 2;
    const int smk = (2 * 64 + 4 * 64) * 64 * 2 + 4 * 64 * 4;
    if ((rc = set_smem(attn_bwd_dq_kernel<64>, smq, &fq64[dev & 63]))) return rc;
    if ((rc = set_smem(attn_bwd_dkv_kernel<64>, smk, &fk64[dev & 63]))) return rc;
    attn_bwd_dq_kernel<64><<<gq, 256, smq, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (const bf16*)dout, lddo, lse, delta, (bf16*)dq, lddq, (int)T, (int)h, (int)hkv, scale);
    B200_CHECK_LAUNCH();
    attn_bwd_dkv_kernel<64><<<gk, 128, smk, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (const bf16*)dout, lddo, lse, delta, (bf16*)dk, lddk, (bf16*)dv, lddv, (int)T, (int)h, (int)hkv, scale);
    B200_CHECK_LAUNCH();
  } else {
    const int smq = (2 * 128 + 4 * 64) * 128 * 2;
    const int smk = (2 * 64 + 4 * 64) * 128 * 2 + 4 * 64 * 4;
    if ((rc = set_smem(attn_bwd_dq_kernel<128>, smq, &fq128[dev & 63]))) return rc;
    if ((rc = set_smem(attn_bwd_dkv_kernel<128>, smk, &fk128[dev & 63]))) return rc;
    attn_bwd_dq_kernel<128><<<gq, 256, smq, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (const bf16*)dout, lddo, lse, delta, (bf16*)dq, lddq, (int)T, (int)h, (int)hkv, scale);
    B200_CHECK_LAUNCH();
    attn_bwd_dkv_kernel<128><<<gk, 128, smk, st>>>((const bf16*)q, ldq, (const bf16*)k, ldk, (const bf16*)v, ldv, (const bf16*)dout, lddo, lse, delta, (bf16*)dk, lddk, (bf16*)dv, lddv, (int)T, (int)h, (int)hkv, scale);
    B200_CHECK_LAUNCH();
  }
  b200_count_launch(3);
  return 0;
}

// ======================================================================================
// Decode step against a KV cache (generation; the reference's generate.py:4-75 has
// no cache and re-runs the whole window per token). One new query token per sequence:
//   o[b, head] = softmax(q[b, head] . K[b, 0..T) * scale) V[b, 0..T)
// CTA = one (batch, kv head): the G = h / hkv query heads of the group share every K / V row read.
// Phase 1: thread per cached position, dot products against the G queries (fp32) -> scores in smem;
// softmax per head (one warp each); phase 2: (head, 8-column vector, position slice) per thread,
// coalesced 16-byte V reads, slices summed through smem. HBM-bound: the cache is read exactly once.
// ======================================================================================
namespace {

template <int HD>
__global__ void __launch_bounds__(256)
attn_decode_kernel(const bf16* __restrict__ q, long long ldq, const bf16* __restrict__ kc,
                   const bf16* __restrict__ vc, long long ldc, long long tmax, bf16* __restrict__ o,
                   long long ldo, int T, int H, int HKV, float scale) {
  extern __shared__ float dsm[];
  const int G = H / HKV;
  const int kvh = blockIdx.x, b = blockIdx.y;
  float* sq = dsm;                       // [G][HD]
  float* ssum = sq + G * HD;             // [G] 1 / sum
  float* red = ssum + 8;                 // [NS][G * HD]
  const int items = G * HD / 8;          // (head, 8-column vector) pairs
  const int NS = 256 / items;            // position slices of phase 2
  float* sc = red + NS * G * HD;         // [G][T] scores, then probabilities
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < G * HD; i += 256)
    sq[i] = __bfloat162float(q[(long long)b * ldq + (long long)(kvh * G + i / HD) * HD + i % HD]) * scale;
  __syncthreads();
  const bf16* kb = kc + (long long)b * tmax * ldc + (long long)kvh * HD;
  const bf16* vb = vc + (long long)b * tmax * ldc + (long long)kvh * HD;
  for (int t = tid; t < T; t += 256) {
    float acc[8];
#pragma unroll
    for (int g = 0; g < 8; ++g) acc[g] = 0.f;
    const bf16* kr = kb + (long long)t * ldc;
#pragma unroll 4
    for (int c = 0; c < HD / 8; ++c) {
      float k8[8];
      unpack8(ld_nc_v4(kr + c * 8), k8);
#pragma unroll
      for (int g = 0; g < 8; ++g)
        if (g < G) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[g] += sq[g * HD + c * 8 + j] * k8[j];
        }
    }
#pragma unroll
    for (int g = 0; g < 8; ++g)
      if (g < G) sc[(long long)g * T + t] = acc[g];
  }
  __syncthreads();
  for (int g = warp; g < G; g += 8) {
    float m = -INFINITY;
    for (int t = lane; t < T; t += 32) m = fmaxf(m, sc[(long long)g * T + t]);
    m = warp_max(m);
    float sum = 0.f;
    for (int t = lane; t < T; t += 32) {
      const float p = __expf(sc[(long long)g * T + t] - m);
      sc[(long long)g * T + t] = p;
      sum += p;
    }
    sum = warp_sum(sum);
    if (lane == 0) ssum[g] = 1.f / sum;
  }
  __syncthreads();
  {
    const int item = tid % items, slice = tid / items;
    const int g = item / (HD / 8), c = item % (HD / 8);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (slice < NS) {
      for (int t = slice; t < T; t += NS) {
        float v8[8];
        unpack8(ld_nc_v4(vb + (long long)t * ldc + c * 8), v8);
        const float p = sc[(long long)g * T + t];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += p * v8[j];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) red[(slice * G + g) * HD + c * 8 + j] = acc[j];
    }
  }
  __syncthreads();
  for (int i = tid; i < G * HD; i += 256) {
    float v = 0.f;
    for (int s2 = 0; s2 < NS; ++s2) v += red[s2 * G * HD + i];
    o[(long long)b * ldo + (long long)(kvh * G + i / HD) * HD + i % HD] = __float2bfloat16_rn(v * ssum[i / HD]);
  }
}

}  // namespace

extern "C" int b200_attention_decode_bf16(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                                          int64_t ldc, int64_t t_max, void* o, int64_t ldo, int64_t B, int64_t T,
                                          int64_t h, int64_t hkv, int64_t hd, float scale, void* stream) {
  B200_CHECK_ARG(q && k_cache && v_cache && o && B > 0 && T > 0 && T <= t_max, "attention_decode: null pointer, empty, or T > t_max");
  B200_CHECK_ARG(hd == 64 || hd == 128, "attention_decode: head_dim %lld not supported (64 or 128)", (long long)hd);
  B200_CHECK_ARG(hkv > 0 && h % hkv == 0 && h / hkv <= 8, "attention_decode: heads / kv heads must divide with a group of at most 8");
  B200_CHECK_ARG(ldc % 8 == 0 && ldc >= hkv * hd && b200_aligned16(k_cache) && b200_aligned16(v_cache),
                 "attention_decode: cache rows must be 16-B aligned with pitch >= hkv * hd");
  B200_CHECK_ARG(B <= 65535 && hkv <= 65535, "attention_decode: batch / kv heads exceed grid limits");
  const int G = (int)(h / hkv);
  const int items = G * (int)hd / 8;
  const int NS = 256 / items;
  const size_t smem = ((size_t)G * hd + 8 + (size_t)NS * G * hd + (size_t)G * T) * sizeof(float);
  B200_CHECK_ARG(smem <= 200 * 1024, "attention_decode: %lld cached positions x %d heads exceed the score buffer (200 KB)", (long long)T, G);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)hkv, (unsigned)B);
  static bool f64[64] = {false}, f128[64] = {false};
  int dev = 0, rc;
  cudaGetDevice(&dev);
  if (hd == 64) {
    if ((rc = set_smem(attn_decode_kernel<64>, 200 * 1024, &f64[dev & 63]))) return rc;
    attn_decode_kernel<64><<<grid, 256, smem, st>>>((const bf16*)q, ldq, (const bf16*)k_cache, (const bf16*)v_cache, ldc, t_max, (bf16*)o, ldo, (int)T, (int)h, (int)hkv, scale);
  } else {
    if ((rc = set_smem(attn_decode_kernel<128>, 200 * 1024, &f128[dev & 63]))) return rc;
    attn_decode_kernel<128><<<grid, 256, smem, st>>>((const bf16*)q, ldq, (const bf16*)k_cache, (const bf16*)v_cache, ldc, t_max, (bf16*)o, ldo, (int)T, (int)h, (int)hkv, scale);
  }
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
