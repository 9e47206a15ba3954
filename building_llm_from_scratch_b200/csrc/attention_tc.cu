// K2 on tcgen05 — fused causal MHA/GQA attention forward for head_dim 64.
//
// CTA = 128 query rows of one (batch, head); KV streamed in 64-key tiles by TMA (3-stage ring).
//   warp 0      : TMA producer (Q once, then K/V tiles)
//   warp 1      : TMEM allocator + single-thread MMA issuer
//                   S_j = Q K_j^T      (128 x 64, K-major operands)      -> TMEM, double buffered
//                   O  += P_j V_j      (A = P from smem, B = V as MN-major) -> TMEM
//   warps 2..5  : softmax — one thread per query row (TMEM lane), so row max / row sum need no
//                 shuffles: tcgen05.ld S, exp2 online softmax, P -> bf16 into 128B-swizzled smem,
//                 rescale O in TMEM (tcgen05.ld/st) only when a row maximum moved, final O / l.
// 96 KB smem and 256 TMEM columns per CTA -> two CTAs per SM, so one CTA's softmax overlaps the
// other's MMAs. Causality is a predicate on the diagonal tiles; fully masked tiles are never loaded.
//
// Roofline: tensor pipe / MUFU (head_dim 64: 8192 exp2 per 128x64 tile vs 256 MMA cycles).
#include "common.cuh"

namespace {

constexpr int BQ = 128, BKV = 64, HD = 64, NS = 3;
constexpr int Q_BYTES = BQ * HD * 2;     // 16 KB
constexpr int KV_BYTES = BKV * HD * 2;   // 8 KB
constexpr int P_BYTES = BQ * BKV * 2;    // 16 KB
constexpr int FWD_SMEM = Q_BYTES + NS * 2 * KV_BYTES + 2 * P_BYTES + 1024 + 256;
constexpr uint32_t TMEM_COLS = 256;      // S0 [0,64) S1 [64,128) O [128,192)
constexpr float LOG2E = 1.4426950408889634f;

struct FwdParams {
  int B, T, H, HKV, nqt;
  float scale_log2;
  bf16* o;
  long long ldo;
  float* lse;
};

__global__ void __launch_bounds__(192, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const FwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Q_BYTES;
  uint8_t* sV = sK + NS * KV_BYTES;
  uint8_t* sP = sV + NS * KV_BYTES;
  uint64_t* bars = (uint64_t*)(sP + 2 * P_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // NS
  uint64_t* v_full = k_full + NS;     // NS
  uint64_t* kv_empty = v_full + NS;   // NS
  uint64_t* s_full = kv_empty + NS;   // 2
  uint64_t* s_empty = s_full + 2;     // 2
  uint64_t* p_full = s_empty + 2;     // 2
  uint64_t* o_done = p_full + 2;      // 1
  uint32_t* tmem_slot = (uint32_t*)(o_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_qt = p.B * p.H;
  const int qt = p.nqt - 1 - (int)(blockIdx.x / per_qt);  // heaviest query tiles first
  const int rem = (int)(blockIdx.x % per_qt);
  const int head = rem % p.H, b = rem / p.H;
  const int kvh = head / (p.H / p.HKV);
  const int q0 = qt * BQ;
  const int row_base = b * p.T;
  const int q_hi = min(q0 + BQ, p.T);
  const int n_kv = (q_hi + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < NS; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 4);
      mbar_init(&p_full[i], 4);
    }
    mbar_init(o_done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, Q_BYTES);
      tma_load_2d(sQ, &tmQ, q_full, head * HD, row_base + q0);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % NS;
        mbar_wait(&kv_empty[s], ((j / NS) & 1) ^ 1);
        mbar_expect_tx(&k_full[s], KV_BYTES);
        tma_load_2d(sK + s * KV_BYTES, &tmK, &k_full[s], kvh * HD, row_base + j * BKV);
        mbar_expect_tx(&v_full[s], KV_BYTES);
        tma_load_2d(sV + s * KV_BYTES, &tmV, &v_full[s], kvh * HD, row_base + j * BKV);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, 0, 0);
      const uint32_t idesc_pv = umma_idesc_bf16(BQ, HD, 0, 1);
      const uint32_t aQ = smem_u32(sQ);
      mbar_wait(q_full, 0);
      auto issue_s = [&](int j) {
        const int s = j % NS, buf = j & 1, use = j >> 1;
        mbar_wait(&k_full[s], (j / NS) & 1);
        if (use > 0) mbar_wait(&s_empty[buf], (use - 1) & 1);
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + buf * BKV, umma_desc_k128(aQ + k * 32), umma_desc_k128(aK + k * 32),
                       idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&s_full[buf]);
      };
      issue_s(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) issue_s(j + 1);
        const int s = j % NS, buf = j & 1;
        mbar_wait(&v_full[s], (j / NS) & 1);
        mbar_wait(&p_full[buf], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t aP = smem_u32(sP + buf * P_BYTES);
        const uint32_t aV = smem_u32(sV + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k)
          umma_bf16_ss(tmem_base + 128, umma_desc_k128(aP + k * 32), umma_desc_mn128(aV + k * 2048, 8192),
                       idesc_pv, (j > 0 || k > 0) ? 1u : 0u);
        umma_commit(&kv_empty[s]);
        umma_commit(o_done);
      }
    }
  } else {
    // ---------------- softmax / correction / epilogue: thread <-> query row ----------------
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < n_kv; ++j) {
      const int buf = j & 1;
      mbar_wait(&s_full[buf], (j >> 1) & 1);
      tc_fence_after();
      uint32_t r[64];
      tmem_ld_32x32(tlane + buf * BKV, r);
      tmem_ld_32x32(tlane + buf * BKV + 32, r + 32);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_empty[buf]);

      const int kv0 = j * BKV;
      const bool diag = kv0 + BKV - 1 > q0 + quad * 32;  // warp-uniform
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        float x = __uint_as_float(r[i]) * p.scale_log2;
        if (diag && kv0 + i > q0 + row) x = -INFINITY;
        r[i] = __float_as_uint(x);
        mx = fmaxf(mx, x);
      }
      const float m_new = fmaxf(m, mx);
      const float alpha = exp2f(m - m_new);
      float sum = 0.f;
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float p0 = exp2f(__uint_as_float(r[2 * i]) - m_new);
        const float p1 = exp2f(__uint_as_float(r[2 * i + 1]) - m_new);
        sum += p0 + p1;
        pk[i] = f2_to_bf2(p0, p1);
      }
      l = l * alpha + sum;
      m = m_new;
      // P row -> smem, K-major 128B-swizzled (16-byte chunk index XOR row % 8)
      uint8_t* prow = sP + buf * P_BYTES + row * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(prow + ((c ^ (row & 7)) << 4)) =
            make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      fence_proxy_async_smem();
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);
        tc_fence_after();
        if (!__all_sync(0xffffffffu, alpha == 1.f)) {
          uint32_t o[32];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            tmem_ld_32x32(tlane + 128 + h2 * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(tlane + 128 + h2 * 32, o);
          }
          tmem_st_wait();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[buf]);
    }
    mbar_wait(o_done, (n_kv - 1) & 1);
    tc_fence_after();
    const float inv = 1.f / l;
    const int t = q0 + row;
    bf16* orow = p.o + (long long)(row_base + t) * p.ldo + (long long)head * HD;
#pragma unroll
    for (int h2 = 0; h2 < 2; ++h2) {
      uint32_t o[32];
      tmem_ld_32x32(tlane + 128 + h2 * 32, o);
      tmem_ld_wait();
      if (t < p.T) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(o[c * 8 + i]) * inv;
          st_v4(orow + h2 * 32 + c * 8, pack8(f));
        }
      }
    }
    if (t < p.T) p.lse[((long long)b * p.H + head) * p.T + t] = (m + log2f(l)) / LOG2E;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace

// Returns 0 on success; called by b200_attention_fwd_bf16 for head_dim 64.
int b200_attention_fwd_tc64(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                            int64_t ldv, void* o, int64_t ldo, float* lse, int64_t B, int64_t T,
                            int64_t h, int64_t hkv, float scale, cudaStream_t st) {
  CUtensorMap mq, mk, mv;
  int rc;
  if ((rc = b200_make_tmap_2d_bf16(&mq, q, (uint64_t)(B * T), (uint64_t)(h * HD), (uint64_t)ldq, BQ, HD))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mk, k, (uint64_t)(B * T), (uint64_t)(hkv * HD), (uint64_t)ldk, BKV, HD))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mv, v, (uint64_t)(B * T), (uint64_t)(hkv * HD), (uint64_t)ldv, BKV, HD))) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr_set[dev & 63] = true;
  }
  FwdParams p;
  p.B = (int)B; p.T = (int)T; p.H = (int)h; p.HKV = (int)hkv;
  p.nqt = (int)((T + BQ - 1) / BQ);
  p.scale_log2 = scale * LOG2E;
  p.o = (bf16*)o; p.ldo = ldo; p.lse = lse;
  const long long grid = (long long)p.nqt * B * h;
  B200_CHECK_ARG(grid < (1ll << 31), "attention: grid too large");
  attn_fwd_tc_kernel<<<(unsigned)grid, 192, FWD_SMEM, st>>>(mq, mk, mv, p);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

// =====================================================================================
// backward (head_dim 64): dK, dV and dQ in one kernel.
//
// CTA = 128 keys of one (batch, kv head); it streams every (query head of the GQA group) x
// (128-row query tile at or below the diagonal). Working transposed (keys on the M / TMEM-lane
// axis) lets one thread own one key row:
//   S^T  = K Q^T            dP^T = V dO^T                      (128 x 128 each, TMEM)
//   P^T  = exp2(S^T c - lse[q]),  dS^T = P^T (dP^T - delta[q]) scale   -> bf16, swizzled smem
//   dV  += P^T dO           dK  += dS^T Q                      (accumulate in TMEM over the loop)
//   dQ   = dS K             (fresh per query tile; reduced across key tiles with fp32 vector
//                            atomics into a [tokens, h*64] workspace)
// The same swizzled Q / dO tiles serve as K-major operands (first two products) and as MN-major
// operands (dV, dK); dS^T in smem is read back as an MN-major A operand for dQ. TMEM: 448 of 512
// columns, one CTA per SM.
// =====================================================================================
namespace {

constexpr int BK2 = 128;                  // keys per CTA and query rows per step
constexpr int T2_BYTES = BK2 * HD * 2;    // 16 KB  [128 x 64] bf16 tile
constexpr int PT_BYTES = BK2 * BK2 * 2;   // 32 KB  [128 x 128] bf16 (two 64-wide panels)
constexpr int BWD_SMEM = 2 * T2_BYTES + 4 * T2_BYTES + 2 * PT_BYTES + 4 * BK2 * 4 + 1024 + 256;

struct BwdParams {
  int B, T, H, HKV, nqt;
  float scale, scale_log2;
  const float* lse;
  const float* delta;
  float* dq_acc;      // [B*T, H*64] fp32
  bf16* dk;
  long long lddk;
  bf16* dv;
  long long lddv;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

__global__ void __launch_bounds__(192, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sV = sK + T2_BYTES;
  uint8_t* sQ = sV + T2_BYTES;          // [2]
  uint8_t* sdO = sQ + 2 * T2_BYTES;     // [2]
  uint8_t* sPt = sdO + 2 * T2_BYTES;
  uint8_t* sdSt = sPt + PT_BYTES;
  float* sLse = (float*)(sdSt + PT_BYTES);  // [2][128]
  float* sDel = sLse + 2 * BK2;             // [2][128]
  uint64_t* bars = (uint64_t*)(sDel + 2 * BK2);
  uint64_t* kv_full = bars;            // 1
  uint64_t* qdo_full = bars + 1;       // 2
  uint64_t* qdo_empty = bars + 3;      // 2
  uint64_t* s_full = bars + 5;         // 1
  uint64_t* s_empty = bars + 6;        // 1 (4 arrivals)
  uint64_t* ps_full = bars + 7;        // 1 (4 arrivals)
  uint64_t* dq_full = bars + 8;        // 1
  uint64_t* dq_empty = bars + 9;       // 1 (4 arrivals)
  uint64_t* acc_done = bars + 10;      // 1
  uint32_t* tmem_slot = (uint32_t*)(bars + 11);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_kt = p.B * p.HKV;
  const int kt = (int)(blockIdx.x / per_kt);   // early key tiles (most query tiles) first
  const int rem = (int)(blockIdx.x % per_kt);
  const int kvh = rem % p.HKV, b = rem / p.HKV;
  const int G = p.H / p.HKV;
  const int kv0 = kt * BK2;
  const int row_base = b * p.T;
  const int nq = p.nqt - kt;
  const int iters = G * nq;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_empty, 4);
    mbar_init(ps_full, 4);
    mbar_init(dq_full, 1);
    mbar_init(dq_empty, 4);
    mbar_init(acc_done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmdO);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t C_ST = 0, C_DPT = 128, C_DV = 256, C_DK = 320, C_DQ = 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * T2_BYTES);
      tma_load_2d(sK, &tmK, kv_full, kvh * HD, row_base + kv0);
      tma_load_2d(sV, &tmV, kv_full, kvh * HD, row_base + kv0);
      for (int it = 0; it < iters; ++it) {
        const int st = it & 1;
        const int head = kvh * G + it / nq, qi = kt + it % nq;
        mbar_wait(&qdo_empty[st], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[st], 2 * T2_BYTES);
        tma_load_2d(sQ + st * T2_BYTES, &tmQ, &qdo_full[st], head * HD, row_base + qi * BK2);
        tma_load_2d(sdO + st * T2_BYTES, &tmdO, &qdo_full[st], head * HD, row_base + qi * BK2);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t id_s = umma_idesc_bf16(128, 128, 0, 0);    // S^T, dP^T
      const uint32_t id_acc = umma_idesc_bf16(128, 64, 0, 1);   // dV, dK: A K-major, B MN-major
      const uint32_t id_dq = umma_idesc_bf16(128, 64, 1, 1);    // dQ: A (dS^T read transposed) and B MN-major
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aPt = smem_u32(sPt), adSt = smem_u32(sdSt);
      mbar_wait(kv_full, 0);
      for (int it = 0; it < iters; ++it) {
        const int st = it & 1;
        const uint32_t aQ = smem_u32(sQ + st * T2_BYTES), adO = smem_u32(sdO + st * T2_BYTES);
        mbar_wait(&qdo_full[st], (it >> 1) & 1);
        if (it > 0) mbar_wait(s_empty, (it - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + C_ST, umma_desc_k128(aK + k * 32), umma_desc_k128(aQ + k * 32), id_s, k > 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + C_DPT, umma_desc_k128(aV + k * 32), umma_desc_k128(adO + k * 32), id_s, k > 0);
        umma_commit(s_full);
        mbar_wait(ps_full, it & 1);
        if (it > 0) mbar_wait(dq_empty, (it - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BK2 / 16; ++k)
          umma_bf16_ss(tmem_base + C_DV, umma_desc_k128(aPt + (k >> 2) * T2_BYTES + (k & 3) * 32),
                       umma_desc_mn128(adO + k * 2048, 8192), id_acc, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < BK2 / 16; ++k)
          umma_bf16_ss(tmem_base + C_DK, umma_desc_k128(adSt + (k >> 2) * T2_BYTES + (k & 3) * 32),
                       umma_desc_mn128(aQ + k * 2048, 8192), id_acc, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < BK2 / 16; ++k)
          umma_bf16_ss(tmem_base + C_DQ, umma_desc_mn128(adSt + k * 2048, T2_BYTES),
                       umma_desc_mn128(aK + k * 2048, 8192), id_dq, k > 0 ? 1u : 0u);
        umma_commit(dq_full);
        umma_commit(&qdo_empty[st]);
      }
      umma_commit(acc_done);
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;     // key row (S^T, dV, dK) and query row (dQ)
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int key = kv0 + row;
    for (int it = 0; it < iters; ++it) {
      const int st = it & 1;
      const int head = kvh * G + it / nq, qi = kt + it % nq;
      const int qrow0 = qi * BK2;
      {
        const int t = qrow0 + row;
        const long long base = ((long long)b * p.H + head) * p.T;
        sLse[st * BK2 + row] = t < p.T ? p.lse[base + t] * LOG2E : INFINITY;
        sDel[st * BK2 + row] = t < p.T ? p.delta[base + t] : 0.f;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      const bool diag = (qi == kt);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t s[32], d[32];
        tmem_ld_32x32(tlane + C_ST + c * 32, s);
        tmem_ld_32x32(tlane + C_DPT + c * 32, d);
        tmem_ld_wait();
        uint32_t pp[16], dd[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float pv[2], dv2[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int q = c * 32 + 2 * i + e;
            float pr = exp2f(__uint_as_float(s[2 * i + e]) * p.scale_log2 - sLse[st * BK2 + q]);
            if (diag && key > qrow0 + q) pr = 0.f;
            pv[e] = pr;
            dv2[e] = pr * (__uint_as_float(d[2 * i + e]) - sDel[st * BK2 + q]) * p.scale;
          }
          pp[i] = f2_to_bf2(pv[0], pv[1]);
          dd[i] = f2_to_bf2(dv2[0], dv2[1]);
        }
        // columns [c*32, c*32+32) of row `row`: panel c/2, 16-byte chunks (c&1)*4 .. +3
        uint8_t* prow = sPt + (c >> 1) * T2_BYTES + row * 128;
        uint8_t* drow = sdSt + (c >> 1) * T2_BYTES + row * 128;
#pragma unroll
        for (int t4 = 0; t4 < 4; ++t4) {
          const int ch = ((c & 1) * 4 + t4) ^ (row & 7);
          *reinterpret_cast<uint4*>(prow + (ch << 4)) = make_uint4(pp[4 * t4], pp[4 * t4 + 1], pp[4 * t4 + 2], pp[4 * t4 + 3]);
          *reinterpret_cast<uint4*>(drow + (ch << 4)) = make_uint4(dd[4 * t4], dd[4 * t4 + 1], dd[4 * t4 + 2], dd[4 * t4 + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(s_empty);
        mbar_arrive(ps_full);
      }
      // dQ tile of this step: thread <-> query row
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      {
        const int t = qrow0 + row;
        float* dst = p.dq_acc + (long long)(row_base + t) * ((long long)p.H * HD) + (long long)head * HD;
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          uint32_t o[32];
          tmem_ld_32x32(tlane + C_DQ + h2 * 32, o);
          tmem_ld_wait();
          if (t < p.T) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              red_add_v4(dst + h2 * 32 + i * 4, __uint_as_float(o[4 * i]), __uint_as_float(o[4 * i + 1]),
                         __uint_as_float(o[4 * i + 2]), __uint_as_float(o[4 * i + 3]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_empty);
    }
    mbar_wait(acc_done, 0);
    tc_fence_after();
    bf16* dkrow = p.dk + (long long)(row_base + key) * p.lddk + (long long)kvh * HD;
    bf16* dvrow = p.dv + (long long)(row_base + key) * p.lddv + (long long)kvh * HD;
#pragma unroll
    for (int h2 = 0; h2 < 2; ++h2) {
      uint32_t a[32], c2[32];
      tmem_ld_32x32(tlane + C_DV + h2 * 32, a);
      tmem_ld_32x32(tlane + C_DK + h2 * 32, c2);
      tmem_ld_wait();
      if (key < p.T) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float f[8], g[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            f[i] = __uint_as_float(a[c * 8 + i]);
            g[i] = __uint_as_float(c2[c * 8 + i]);
          }
          st_v4(dvrow + h2 * 32 + c * 8, pack8(f));
          st_v4(dkrow + h2 * 32 + c * 8, pack8(g));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// delta[b, h, t] = sum_c dO * O over one head: 8 lanes per (token, head), 4 pairs per warp
__global__ void __launch_bounds__(256)
attn_delta64_kernel(const bf16* __restrict__ o, long long ldo, const bf16* __restrict__ dout,
                    long long lddo, float* __restrict__ delta, long long B, int T, int H) {
  const long long total = B * T * H;
  const int lane = threadIdx.x & 31, sub = lane >> 3, l8 = lane & 7;
  for (long long w = ((long long)blockIdx.x * 8 + (threadIdx.x >> 5)) * 4 + sub; w < total + 3;
       w += (long long)gridDim.x * 8 * 4) {
    float acc = 0.f;
    const bool ok = w < total;
    const int head = ok ? (int)(w % H) : 0;
    const long long tok = ok ? w / H : 0;
    if (ok) {
      float a[8], g[8];
      unpack8(ld_nc_v4(o + tok * ldo + (long long)head * HD + l8 * 8), a);
      unpack8(ld_nc_v4(dout + tok * lddo + (long long)head * HD + l8 * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += a[j] * g[j];
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (ok && l8 == 0) delta[((tok / T) * H + head) * T + (tok % T)] = acc;
  }
}

// dq (bf16, pitch lddq) = dq_acc (fp32, contiguous [rows, cols])
__global__ void __launch_bounds__(256)
dq_finish_kernel(const float* __restrict__ acc, bf16* __restrict__ dq, long long lddq, long long rows,
                 int cols) {
  const int cv = cols >> 3;
  const long long total = rows * cv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cv;
    const int c = (int)(i % cv) * 8;
    const uint4 u0 = ld_nc_v4(acc + r * cols + c), u1 = ld_nc_v4(acc + r * cols + c + 4);
    const float f[8] = {__uint_as_float(u0.x), __uint_as_float(u0.y), __uint_as_float(u0.z), __uint_as_float(u0.w),
                        __uint_as_float(u1.x), __uint_as_float(u1.y), __uint_as_float(u1.z), __uint_as_float(u1.w)};
    st_v4(dq + r * lddq + c, pack8(f));
  }
}

}  // namespace

size_t b200_attention_bwd_tc64_workspace(int64_t B, int64_t T, int64_t h) {
  return (size_t)(B * h * T) * 4 + (size_t)(B * T * h * HD) * 4 + 256;
}

int b200_attention_bwd_tc64(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                            int64_t ldv, const void* o, int64_t ldo, const void* dout, int64_t lddo,
                            const float* lse, void* workspace, void* dq, int64_t lddq, void* dk,
                            int64_t lddk, void* dv, int64_t lddv, int64_t B, int64_t T, int64_t h,
                            int64_t hkv, float scale, cudaStream_t st) {
  float* delta = (float*)workspace;
  float* dq_acc = (float*)(((uintptr_t)(delta + B * h * T) + 255) & ~(uintptr_t)255);
  const long long rows = B * T;
  {
    const long long total = rows * h;
    long long blocks = (total + 31) / 32;
    const long long cap = (long long)b200_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    attn_delta64_kernel<<<(int)blocks, 256, 0, st>>>((const bf16*)o, ldo, (const bf16*)dout, lddo, delta, B, (int)T, (int)h);
    B200_CHECK_LAUNCH();
  }
  B200_CHECK_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)rows * h * HD * 4, st));
  CUtensorMap mq, mdo, mk, mv;
  int rc;
  if ((rc = b200_make_tmap_2d_bf16(&mq, q, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)ldq, BK2, HD))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mdo, dout, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)lddo, BK2, HD))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mk, k, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldk, BK2, HD))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mv, v, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldv, BK2, HD))) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));
    attr_set[dev & 63] = true;
  }
  BwdParams p;
  p.B = (int)B; p.T = (int)T; p.H = (int)h; p.HKV = (int)hkv;
  p.nqt = (int)((T + BK2 - 1) / BK2);
  p.scale = scale; p.scale_log2 = scale * LOG2E;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.dk = (bf16*)dk; p.lddk = lddk; p.dv = (bf16*)dv; p.lddv = lddv;
  const long long grid = (long long)p.nqt * B * hkv;
  attn_bwd_tc_kernel<<<(unsigned)grid, 192, BWD_SMEM, st>>>(mq, mdo, mk, mv, p);
  B200_CHECK_LAUNCH();
  {
    const long long total = rows * (h * HD / 8);
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)b200_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    dq_finish_kernel<<<(int)blocks, 256, 0, st>>>(dq_acc, (bf16*)dq, lddq, rows, (int)(h * HD));
    B200_CHECK_LAUNCH();
  }
  b200_count_launch(4);
  return 0;
}
