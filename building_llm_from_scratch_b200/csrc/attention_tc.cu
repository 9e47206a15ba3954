// K2 on tcgen05 — fused causal MHA/GQA attention, forward and backward, head_dim 64 and 128
// (kernels are templates on HDIM; a [rows x HDIM] tile is HDIM/64 swizzled 64-column panels).
// Forward:
//
// CTA = 128 query rows of one (batch, head); KV streamed in 64-key tiles by TMA (3-stage ring).
//   warp 0      : TMA producer (Q once, then K/V tiles)
//   warp 1      : TMEM allocator + issuer of  S_j = Q K_j^T  (128 x 64, K-major operands) -> TMEM
//   warp 2      : issuer of  O += P_j V_j  (A = P from smem, B = V as MN-major) -> TMEM
//   warps 3..6  : softmax — one thread per query row (TMEM lane), so row max / row sum need no
//                 shuffles: tcgen05.ld S, exp2 online softmax, P -> bf16 into 128B-swizzled smem,
//                 rescale O in TMEM (tcgen05.ld/st) only when a row maximum moved, final O / l.
// head_dim 64: 65 KB smem and 128 TMEM columns per CTA -> three CTAs per SM (head_dim 128: 113 KB, 256
// columns, two CTAs), so one CTA's softmax overlaps the others' MMAs and barrier / TMEM-load latencies. Causality is a predicate on the diagonal tiles; fully masked tiles are never loaded.
//
// Roofline: tensor pipe / MUFU (head_dim 64: 8192 exp2 per 128x64 tile vs 256 MMA cycles).
#include <cstdlib>
#include <type_traits>

#include "common.cuh"

namespace {

constexpr int BQ = 128, BKV = 64;
constexpr int P_BYTES = BQ * BKV * 2;    // 16 KB
constexpr float LOG2E = 1.4426950408889634f;
constexpr float RESCALE_SLACK = 8.f;     // forward softmax: tolerated growth of the row maximum before O is rescaled

template <int HDIM>
struct FwdCfg {
  static constexpr int Q_BYTES = BQ * HDIM * 2;     // 16 / 32 KB
  static constexpr int KV_BYTES = BKV * HDIM * 2;   // 8 / 16 KB
  // K tiles are released as soon as their score product has retired (they are not needed for P V), V tiles after
  // P V: separate rings, K one stage deeper where shared memory allows (head_dim 64: 73 KB, still three CTAs per SM).
  // With one shared two-stage ring the score issuer spun on k_full: a K tile could only be requested once the P V
  // product two tiles back had retired, one softmax (~500 clk) ahead of its use against ~1 us of load latency.
  static constexpr int NSK = HDIM == 64 ? 3 : 2;
  static constexpr int NSV = 2;
  static constexpr int SMEM = Q_BYTES + (NSK + NSV) * KV_BYTES + P_BYTES + 1024 + 256;
  static constexpr uint32_t TMEM_COLS = HDIM == 64 ? 128 : 256;   // S [0,64)  O [64, 64 + HDIM)
  static constexpr int CTAS = HDIM == 64 ? 3 : 2;
};

// [rows x HDIM] bf16 tile = HDIM/64 panels of [rows x 64] (128 B per row, 128B-swizzled), panel after panel
template <int HDIM>
__device__ __forceinline__ void tma_load_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int col0,
                                              int row0, int rows) {
#pragma unroll
  for (int pn = 0; pn < HDIM / 64; ++pn) tma_load_2d(dst + pn * rows * 128, tm, bar, col0 + pn * 64, row0);
}
// byte offset of the k-th 16-element K step of a K-major [ROWS x HDIM] tile
template <int ROWS>
__device__ __forceinline__ uint32_t kstep(int k) {
  return (uint32_t)((k >> 2) * (ROWS * 128) + (k & 3) * 32);
}

// attention-weight dropout (Models/GPT2/GPT2.py:43): keep flag of score (b, h, q, k) =
// dropout_keep2(((b*H + h)*T + q) * ceil(T/2) + k/2) half k&1; drop_thr16 == 0 disables it.
struct DropParams {
  uint32_t thr16;
  float inv_keep;
  unsigned long long seed;
  const unsigned long long* seed_dev;   // optional per-step offset in device memory, added to seed (graph replay)
};
__device__ __forceinline__ unsigned long long drop_seed(const DropParams& d) {
  return d.seed + (d.seed_dev != nullptr ? *d.seed_dev : 0ull);
}

struct FwdParams {
  int B, T, H, HKV, nqt;
  DropParams drop;
  float scale_log2;
  bf16* o;
  long long ldo;
  float* lse;
};

// PT: P (the bf16 softmax weights) goes to tensor memory and feeds O += P V as the MMA's A operand from there, instead
// of a swizzled shared-memory tile: 16 KB of stores and 16 KB of operand reads less per tile on a kernel whose MMAs
// are bound by shared-memory bandwidth. Costs 32 more TMEM columns: 256 per CTA, two CTAs per SM.
template <int HDIM, bool DROP, bool PT>
// (the dropout variant needs ~30 more registers: two CTAs per SM instead of three rather than 1.3 KB of spills)
__global__ void __launch_bounds__(224, (DROP || PT) ? 2 : FwdCfg<HDIM>::CTAS)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const FwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int HD = HDIM;
  constexpr int Q_BYTES = FwdCfg<HDIM>::Q_BYTES, KV_BYTES = FwdCfg<HDIM>::KV_BYTES;
  constexpr uint32_t TMEM_COLS = PT ? 256 : FwdCfg<HDIM>::TMEM_COLS;
  constexpr uint32_t P_COL = 64 + HDIM;     // PT: P [128 lanes x 64 keys] bf16 = 32 columns behind S (64) and O (HDIM)
  constexpr int NSK = FwdCfg<HDIM>::NSK, NSV = FwdCfg<HDIM>::NSV;
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Q_BYTES;
  uint8_t* sV = sK + NSK * KV_BYTES;
  uint8_t* sP = sV + NSV * KV_BYTES;
  uint64_t* bars = (uint64_t*)(sP + P_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // NSK
  uint64_t* v_full = k_full + NSK;    // NSV
  uint64_t* k_empty = v_full + NSV;   // NSK
  uint64_t* v_empty = k_empty + NSK;  // NSV
  uint64_t* s_full = v_empty + NSV;   // 1
  uint64_t* s_empty = s_full + 1;     // 1 (4 arrivals)
  uint64_t* p_full = s_empty + 1;     // 1 (4 arrivals)
  uint64_t* o_done = p_full + 1;      // 1
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
    for (int i = 0; i < NSK; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    for (int i = 0; i < NSV; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_empty, 4);
    mbar_init(p_full, 4);
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
  pdl_launch_dependents();   // barriers, TMEM and descriptors are set up: the rest waits for the previous kernel's results
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, Q_BYTES);
      tma_load_tile<HD>(sQ, &tmQ, q_full, head * HD, row_base + q0, BQ);
      for (int j = 0; j < n_kv; ++j) {
        const int sk = j % NSK, sv = j % NSV;
        mbar_wait(&k_empty[sk], ((j / NSK) & 1) ^ 1);
        mbar_expect_tx(&k_full[sk], KV_BYTES);
        tma_load_tile<HD>(sK + sk * KV_BYTES, &tmK, &k_full[sk], kvh * HD, row_base + j * BKV, BKV);
        mbar_wait(&v_empty[sv], ((j / NSV) & 1) ^ 1);
        mbar_expect_tx(&v_full[sv], KV_BYTES);
        tma_load_tile<HD>(sV + sv * KV_BYTES, &tmV, &v_full[sv], kvh * HD, row_base + j * BKV, BKV);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, 0, 0);
      const uint32_t aQ = smem_u32(sQ);
      mbar_wait(q_full, 0);
      auto issue_s = [&](int j) {
        const int s = j % NSK;
        mbar_wait(&k_full[s], (j / NSK) & 1);
        if (j > 0) mbar_wait(s_empty, (j - 1) & 1);     // softmax threads hold S_{j-1} in registers
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base, umma_desc_k128(aQ + kstep<BQ>(k)), umma_desc_k128(aK + kstep<BKV>(k)), idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&k_empty[s]);                       // the K tile is free once S_j has been formed
        umma_commit(s_full);
      };
      // score products only; P V is issued by warp 2 (one thread doing both was the slowest link of the
      // dK/dV kernel, see there). s_empty throttles this loop: S is a single TMEM buffer.
      for (int j = 0; j < n_kv; ++j) issue_s(j);
    }
  } else if (warp == 2) {
    if (lane == 0) {
      const uint32_t idesc_pv = umma_idesc_bf16(BQ, HD, 0, 1);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % NSV;
        mbar_wait(&v_full[s], (j / NSV) & 1);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint32_t aP = smem_u32(sP);
        const uint32_t aV = smem_u32(sV + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          if (PT)      // A = 16 keys of P from tensor memory: 8 columns per k step
            umma_bf16_ts(tmem_base + 64, tmem_base + P_COL + k * 8, umma_desc_mn128(aV + k * 2048, 8192), idesc_pv,
                         (j > 0 || k > 0) ? 1u : 0u);
          else
            umma_bf16_ss(tmem_base + 64, umma_desc_k128(aP + k * 32), umma_desc_mn128(aV + k * 2048, 8192),
                         idesc_pv, (j > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&v_empty[s]);
        umma_commit(o_done);
      }
    }
  } else {
    // ---------------- softmax / correction / epilogue: thread <-> query row ----------------
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    float m = -INFINITY, l = 0.f;
    const unsigned long long dseed = DROP ? drop_seed(p.drop) : 0ull;
    const unsigned long long drop_row =
        (((unsigned long long)b * p.H + head) * p.T + (unsigned long long)(q0 + row)) * (unsigned long long)((p.T + 1) >> 1);
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      uint32_t r[64];
      tmem_ld_32x32(tlane, r);
      tmem_ld_32x32(tlane + 32, r + 32);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty);

      const int kv0 = j * BKV;
      const bool diag = kv0 + BKV - 1 > q0 + quad * 32;  // warp-uniform
      if (diag) {
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (kv0 + i > q0 + row) r[i] = 0xff800000u;     // -inf
      }
      // row maximum on the raw scores (the scale is positive), two columns per FMNMX3
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) mx = fmax3(mx, __uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
      // The running maximum only has to bound the exponents, not equal the true maximum: while the new tile raises it
      // by less than RESCALE_SLACK (log2 units) the old reference is kept — p = exp2(s - m) then stays below
      // 2^RESCALE_SLACK, well inside fp32 / bf16 range — and O and l need no rescaling (alpha = 1). With the exact
      // maximum 82 % of the iterations paid a TMEM load / multiply / store round trip over O on the critical path.
      const float m_cand = fmaxf(m, mx * p.scale_log2);
      const float m_new = (m_cand - m > RESCALE_SLACK) ? m_cand : m;     // first tile: m = -inf -> takes m_cand
      const float alpha = fast_exp2(m - m_new);
      // p = exp2(s * scale - m_new): one packed FFMA and one packed FADD per column pair
      const unsigned long long sc2 = pack_f2(p.scale_log2, p.scale_log2), neg2 = pack_f2(-m_new, -m_new);
      unsigned long long sum2 = pack_f2(0.f, 0.f);
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float a0, a1;
        unpack_f2(ffma2(pack_f2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), sc2, neg2), a0, a1);
        float p0, p1;
        if (((HDIM == 128 ? B200_ATTN_POLY_FWD_HD128 : B200_ATTN_POLY_FWD) >> (i & 7)) & 1) {
          exp2_poly_pair(a0, a1, p0, p1);
        } else {
          p0 = fast_exp2(a0);
          p1 = fast_exp2(a1);
        }
        sum2 = fadd2(sum2, pack_f2(p0, p1));          // the softmax denominator is taken before dropout
        if (DROP) {
          bool k0, k1;
          dropout_keep2(drop_row + (unsigned long long)((kv0 >> 1) + i), dseed, p.drop.thr16, k0, k1);
          p0 = k0 ? p0 * p.drop.inv_keep : 0.f;
          p1 = k1 ? p1 * p.drop.inv_keep : 0.f;
        }
        pk[i] = f2_to_bf2(p0, p1);
      }
      float sum, sum_hi;
      unpack_f2(sum2, sum, sum_hi);
      sum += sum_hi;
      l = l * alpha + sum;
      m = m_new;
      // the single P buffer and the O accumulator are free once PV_{j-1} has retired
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);
        tc_fence_after();
      }
      if (PT) {
        // P row -> tensor memory: this thread's lane, 32 columns of two bf16 each (keys 2c, 2c + 1 in column c)
        tmem_st_32x32(tlane + P_COL, pk);
        tmem_st_wait();
      } else {
        // P row -> smem, K-major 128B-swizzled (16-byte chunk index XOR row % 8)
        const uint32_t prow = smem_u32(sP) + row * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          sts_v4(prow + ((c ^ (row & 7)) << 4), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        fence_proxy_async_smem();
      }
      if (j > 0 && !__all_sync(0xffffffffu, alpha == 1.f)) {
        uint32_t o[32];
#pragma unroll
        for (int h2 = 0; h2 < HD / 32; ++h2) {
          tmem_ld_32x32(tlane + 64 + h2 * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x32(tlane + 64 + h2 * 32, o);
        }
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    mbar_wait(o_done, (n_kv - 1) & 1);
    tc_fence_after();
    const float inv = 1.f / l;
    const int t = q0 + row;
    bf16* orow = p.o + (long long)(row_base + t) * p.ldo + (long long)head * HD;
#pragma unroll
    for (int h2 = 0; h2 < HD / 32; ++h2) {
      uint32_t o[32];
      tmem_ld_32x32(tlane + 64 + h2 * 32, o);
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

static DropParams make_drop(float p, uint64_t seed, const uint64_t* seed_dev = nullptr) {
  DropParams d;
  d.seed_dev = (const unsigned long long*)seed_dev;
  const float t = p * 65536.f + 0.5f;
  d.thr16 = p <= 0.f ? 0u : (t > 65535.f ? 65535u : (uint32_t)t);
  d.inv_keep = 1.f / (1.f - p);
  d.seed = (unsigned long long)seed;
  return d;
}

namespace {
template <int HD>
int attention_fwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                     void* o, int64_t ldo, float* lse, int64_t B, int64_t T, int64_t h, int64_t hkv,
                     float scale, float dropout_p, uint64_t seed, const uint64_t* seed_dev, cudaStream_t st) {
  constexpr int SMEM = FwdCfg<HD>::SMEM;
  CUtensorMap mq, mk, mv;
  int rc;
  if ((rc = b200_make_tmap_2d_bf16(&mq, q, (uint64_t)(B * T), (uint64_t)(h * HD), (uint64_t)ldq, BQ, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mk, k, (uint64_t)(B * T), (uint64_t)(hkv * HD), (uint64_t)ldk, BKV, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mv, v, (uint64_t)(B * T), (uint64_t)(hkv * HD), (uint64_t)ldv, BKV, 64))) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<HD, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<HD, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<HD, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_set[dev & 63] = true;
  }
  // measured (B200): P through TMEM pays at head_dim 128 (218 -> 202 us at B1 T4096; both variants run two CTAs per SM)
  // and costs at head_dim 64 (130 -> 136 us: it gives up the third CTA per SM); B200_ATTN_P_TMEM=0/1 overrides
  static const int p_tmem_env = [] { const char* e = getenv("B200_ATTN_P_TMEM"); return e ? (e[0] == '1' ? 1 : 0) : -1; }();
  const bool p_tmem = p_tmem_env >= 0 ? p_tmem_env == 1 : HD == 128;
  FwdParams p;
  p.B = (int)B; p.T = (int)T; p.H = (int)h; p.HKV = (int)hkv;
  p.nqt = (int)((T + BQ - 1) / BQ);
  p.scale_log2 = scale * LOG2E;
  p.drop = make_drop(dropout_p, seed, seed_dev);
  p.o = (bf16*)o; p.ldo = ldo; p.lse = lse;
  const long long grid = (long long)p.nqt * B * h;
  B200_CHECK_ARG(grid < (1ll << 31), "attention: grid too large");
  if (p.drop.thr16) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_fwd_tc_kernel<HD, true, false>, dim3((unsigned)grid), dim3(224), SMEM, st, mq, mk, mv, p));
  else if (p_tmem) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_fwd_tc_kernel<HD, false, true>, dim3((unsigned)grid), dim3(224), SMEM, st, mq, mk, mv, p));
  else B200_CHECK_CUDA(b200_launch_pdl_wide(attn_fwd_tc_kernel<HD, false, false>, dim3((unsigned)grid), dim3(224), SMEM, st, mq, mk, mv, p));
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
}  // namespace

// Returns 0 on success; called by b200_attention_fwd_bf16 (head_dim 64 or 128).
int b200_attention_fwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, void* o, int64_t ldo, float* lse, int64_t B, int64_t T,
                          int64_t h, int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed,
                          const uint64_t* seed_dev, cudaStream_t st) {
  if (hd == 64) return attention_fwd_tc<64>(q, ldq, k, ldk, v, ldv, o, ldo, lse, B, T, h, hkv, scale, dropout_p, seed, seed_dev, st);
  return attention_fwd_tc<128>(q, ldq, k, ldk, v, ldv, o, ldo, lse, B, T, h, hkv, scale, dropout_p, seed, seed_dev, st);
}

// =====================================================================================
// backward, dK / dV kernel.
//
// CTA = 128 keys of one (batch, kv head); it streams every (query head of the GQA group) x
// (64-row query step at or below the diagonal). Working transposed (keys on the M / TMEM-lane
// axis) lets a thread own one key row:
//   S^T  = K Q^T            dP^T = V dO^T                      (128 x 64 each, TMEM, double buffered)
//   P^T  = exp2(S^T c - lse[q]),  dS^T = P^T (dP^T - delta[q]) scale   -> bf16, swizzled smem panels
//   dV  += P^T dO           dK  += dS^T Q                      (accumulate in TMEM over the whole loop)
// The same swizzled Q / dO tiles serve as K-major operands (first two products) and as MN-major
// operands (the accumulating two). The MMA issuer runs one step ahead (S/dP of step s+1 and the
// accumulation of step s-1 execute while the 8 softmax warps work on step s).
// TMEM: all 512 columns (head_dim 64: 3 S^T/dP^T buffer pairs + 2 x 64 accumulator columns; head_dim 128:
// 2 buffer pairs + 2 x 128), one CTA per SM.
// =====================================================================================
namespace {

constexpr int BK2 = 128;                  // keys per CTA (also the query tile of the dQ kernel's maps)
constexpr int QS = 64;                    // query rows per step
constexpr int PANEL_BYTES = BK2 * QS * 2; // 16 KB  [128 keys x 64 q] bf16
constexpr int LD_BYTES = 2 * QS * 4;       // 64 x (lse * log2e) then 64 x delta for one query step

template <int HDIM>
struct DkvCfg {
  static constexpr int NG = HDIM == 64 ? 3 : 2;      // softmax groups = S^T/dP^T TMEM buffers = smem panel pairs
  static constexpr int NSTG = HDIM == 64 ? 5 : 3;    // Q / dO stages
  static constexpr int T2_BYTES = BK2 * HDIM * 2;    // 16 / 32 KB  [128 x HDIM] bf16 tile
  static constexpr int QS_BYTES = QS * HDIM * 2;     // 8 / 16 KB
  static constexpr int THREADS = 96 + NG * 128;      // TMA warp, score-MMA warp, accumulate-MMA warp, NG x 4 softmax warps
  static constexpr int SMEM = 2 * T2_BYTES + NSTG * 2 * QS_BYTES + 2 * NG * PANEL_BYTES + NSTG * LD_BYTES + 1024 + 256;
  static_assert(SMEM <= 232448, "dK/dV kernel exceeds the 227 KB shared memory of an SM");
  static_assert(2 * NG * QS + 2 * HDIM <= 512, "dK/dV kernel exceeds TMEM");
};

struct BwdParams {
  int B, T, H, HKV, nkt, nq64;
  DropParams drop;
  float scale, scale_log2;
  const float* ld2;   // [B, H, nq64, 2, 64]: -lse * log2e (-inf beyond T) and -delta (0 beyond T)
  bf16* dk;
  long long lddk;
  bf16* dv;
  long long lddv;
  const bf16* rope_cos;   // non-null: K was rotated before the kernel (RoPE fused into the projection's epilogue), so dK
  const bf16* rope_sin;   // leaves through the transposed rotation; tables [>= T, HD] bf16 with equal halves
};

// Gradient of rotate-half RoPE (the transpose of the rotation): for the chunk pair (first half a, second half b) of a
// head and this row's table entries, dx1 = a cos + b sin, dx2 = b cos - a sin (compute_rope under autograd,
// Models/Llama/common_components.py:24-35; the two table halves are equal).
__device__ __forceinline__ void rope_inverse_pair(uint32_t* a, uint32_t* b, const bf16* __restrict__ cosr,
                                                  const bf16* __restrict__ sinr) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float c[8], s[8];
    unpack8(ld_v4(cosr + q * 8), c);
    unpack8(ld_v4(sinr + q * 8), s);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x1 = __uint_as_float(a[q * 8 + j]), x2 = __uint_as_float(b[q * 8 + j]);
      a[q * 8 + j] = __float_as_uint(x1 * c[j] + x2 * s[j]);
      b[q * 8 + j] = __float_as_uint(x2 * c[j] - x1 * s[j]);
    }
  }
}

// DT: P^T and dS^T go to tensor memory — over the S^T / dP^T columns their softmax group has just read — and feed the
// accumulating products as A operands from there, instead of swizzled shared-memory panels (32 KB of stores and 32 KB
// of operand reads less per step; the kernel is bound by shared-memory bandwidth). The score issuer then has to wait
// for the accumulating products of the buffer's previous step (panel_free) before it overwrites the columns.
template <int HDIM, bool DROP, bool DT>
__global__ void __launch_bounds__(DkvCfg<HDIM>::THREADS, 1)
attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                       const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                       const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int HD = HDIM, NG = DkvCfg<HDIM>::NG, NSTG = DkvCfg<HDIM>::NSTG;
  constexpr int T2_BYTES = DkvCfg<HDIM>::T2_BYTES, QS_BYTES = DkvCfg<HDIM>::QS_BYTES;
  uint8_t* sK = smem;
  uint8_t* sV = sK + T2_BYTES;
  uint8_t* sQ = sV + T2_BYTES;              // [NSTG] x 8 KB
  uint8_t* sdO = sQ + NSTG * QS_BYTES;      // [NSTG] x 8 KB
  uint8_t* sPt = sdO + NSTG * QS_BYTES;     // [NG] x 16 KB
  uint8_t* sdSt = sPt + NG * PANEL_BYTES;   // [NG] x 16 KB
  uint8_t* sLd = sdSt + NG * PANEL_BYTES;          // [NSTG] x 512 B
  uint64_t* bars = (uint64_t*)(sLd + NSTG * LD_BYTES);
  uint64_t* kv_full = bars;                 // 1
  uint64_t* qdo_full = bars + 1;            // NSTG
  uint64_t* qdo_empty = qdo_full + NSTG;    // NSTG
  uint64_t* s_full = qdo_empty + NSTG;      // NG
  uint64_t* st_empty = s_full + NG;         // NG (4 arrivals: one softmax group each)
  uint64_t* ps_full = st_empty + NG;        // NG (4 arrivals)
  uint64_t* panel_free = ps_full + NG;      // NG
  uint64_t* acc_done = panel_free + NG;     // 1
  uint32_t* tmem_slot = (uint32_t*)(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_kt = p.B * p.HKV;
  const int kt = (int)(blockIdx.x / per_kt);   // early key tiles (most query steps) first
  const int rem = (int)(blockIdx.x % per_kt);
  const int kvh = rem % p.HKV, b = rem / p.HKV;
  const int G = p.H / p.HKV;
  const int kv0 = kt * BK2;
  const int row_base = b * p.T;
  const int spp = p.nq64 - 2 * kt;             // query steps per head
  const int steps = G * spp;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < NSTG; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    for (int i = 0; i < NG; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&st_empty[i], 4);
      mbar_init(&ps_full[i], 4);
      mbar_init(&panel_free[i], 1);
    }
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
  pdl_launch_dependents();   // barriers, TMEM and descriptors are set up: the rest waits for the previous kernel's results
  pdl_wait();
  constexpr uint32_t C_ST = 0, C_DPT = NG * QS, C_DV = 2 * NG * QS, C_DK = 2 * NG * QS + HD;   // 512 columns for NG = 3

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * T2_BYTES);
      tma_load_tile<HD>(sK, &tmK, kv_full, kvh * HD, row_base + kv0, BK2);
      tma_load_tile<HD>(sV, &tmV, kv_full, kvh * HD, row_base + kv0, BK2);
      int head = kvh * G, q64 = 2 * kt;
      for (int s = 0; s < steps; ++s) {
        const int stg = s % NSTG;
        mbar_wait(&qdo_empty[stg], ((s / NSTG) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[stg], 2 * QS_BYTES + LD_BYTES);
        tma_load_tile<HD>(sQ + stg * QS_BYTES, &tmQ, &qdo_full[stg], head * HD, row_base + q64 * QS, QS);
        tma_load_tile<HD>(sdO + stg * QS_BYTES, &tmdO, &qdo_full[stg], head * HD, row_base + q64 * QS, QS);
        bulk_load_1d(sLd + stg * LD_BYTES, p.ld2 + (((long long)b * p.H + head) * p.nq64 + q64) * (2 * QS),
                     LD_BYTES, &qdo_full[stg]);
        if (++q64 == p.nq64) {
          q64 = 2 * kt;
          ++head;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t id_s = umma_idesc_bf16(128, QS, 0, 0);     // S^T, dP^T
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
      mbar_wait(kv_full, 0);
      auto issue_s = [&](int s) {
        const int stg = s % NSTG, buf = s % NG, use = s / NG;
        const uint32_t aQ = smem_u32(sQ + stg * QS_BYTES), adO = smem_u32(sdO + stg * QS_BYTES);
        mbar_wait(&qdo_full[stg], (s / NSTG) & 1);
        if (use > 0) mbar_wait(DT ? &panel_free[buf] : &st_empty[buf], (use - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + C_ST + buf * QS, umma_desc_k128(aK + kstep<BK2>(k)), umma_desc_k128(aQ + kstep<QS>(k)), id_s, k > 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + C_DPT + buf * QS, umma_desc_k128(aV + kstep<BK2>(k)), umma_desc_k128(adO + kstep<QS>(k)), id_s, k > 0);
        umma_commit(&s_full[buf]);
      };
      // This thread only issues the score products; it runs as far ahead as the S^T/dP^T buffers
      // (st_empty) and the Q/dO ring (qdo_full) allow. The accumulating products are issued by warp 2:
      // one thread doing both spent a barrier round trip per product group and was the slowest link
      // (the softmax warps sat on s_full for a quarter of their samples).
      for (int s = 0; s < steps; ++s) issue_s(s);
    }
  } else if (warp == 2) {
    if (lane == 0) {
      const uint32_t id_acc = umma_idesc_bf16(128, HD, 0, 1);   // dV, dK: A K-major, B MN-major
      for (int s = 0; s < steps; ++s) {
        const int stg = s % NSTG, buf = s % NG;
        const uint32_t aQ = smem_u32(sQ + stg * QS_BYTES), adO = smem_u32(sdO + stg * QS_BYTES);
        const uint32_t aPt = smem_u32(sPt + buf * PANEL_BYTES), adSt = smem_u32(sdSt + buf * PANEL_BYTES);
        mbar_wait(&qdo_full[stg], (s / NSTG) & 1);
        mbar_wait(&ps_full[buf], (s / NG) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < QS / 16; ++k) {
          if (DT)
            umma_bf16_ts(tmem_base + C_DV, tmem_base + C_ST + buf * QS + k * 8, umma_desc_mn128(adO + k * 2048, 8192), id_acc,
                         (s > 0 || k > 0) ? 1u : 0u);
          else
            umma_bf16_ss(tmem_base + C_DV, umma_desc_k128(aPt + k * 32), umma_desc_mn128(adO + k * 2048, 8192), id_acc,
                         (s > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < QS / 16; ++k) {
          if (DT)
            umma_bf16_ts(tmem_base + C_DK, tmem_base + C_DPT + buf * QS + k * 8, umma_desc_mn128(aQ + k * 2048, 8192), id_acc,
                         (s > 0 || k > 0) ? 1u : 0u);
          else
            umma_bf16_ss(tmem_base + C_DK, umma_desc_k128(adSt + k * 32), umma_desc_mn128(aQ + k * 2048, 8192), id_acc,
                         (s > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&panel_free[buf]);
        umma_commit(&qdo_empty[stg]);
      }
      umma_commit(acc_done);
    }
  } else {
    // NG softmax groups of 4 warps take the steps round-robin (group = step % NG): each owns one of
    // the S^T/dP^T TMEM buffers and one P^T/dS^T smem panel pair, so the barrier / TMEM-load /
    // proxy-fence latencies of one group hide behind the other groups' arithmetic.
    const int quad = warp & 3;                        // TMEM lane quadrant of this warp
    const int grp = (warp - 3) >> 2;
    const int row = quad * 32 + lane;                 // key row
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int key = kv0 + row;
    const int buf = grp;
    const unsigned long long dseed = DROP ? drop_seed(p.drop) : 0ull;
    for (int s = grp; s < steps; s += NG) {
      const int stg = s % NSTG, use = s / NG;
      const int hq = s / spp;
      const int qrow0 = (2 * kt + (s - hq * spp)) * QS;
      const unsigned long long drop_bh = (unsigned long long)b * p.H + (unsigned long long)(kvh * G + hq);
      mbar_wait(&qdo_full[stg], (s / NSTG) & 1);      // lse/delta travelled with this stage
      mbar_wait(&s_full[buf], use & 1);
      tc_fence_after();
      const bool diag = qrow0 < kv0 + BK2;
      const uint32_t lse_s = smem_u32(sLd) + stg * LD_BYTES;
      const uint32_t del_s = lse_s + QS * 4;
      // (the wait for this group's P^T / dS^T panels to be free — the accumulate MMAs of step s - NG have retired — sits
      // right before the first store into them, not here: the exponentials and products of this step do not need the
      // panels, and with the wait up front each group idled through its own MMAs' issue-to-retire latency; the spin on
      // panel_free was the most executed instruction of the kernel)
      const uint32_t prow = smem_u32(sPt) + buf * PANEL_BYTES + row * 128;
      const uint32_t drow = smem_u32(sdSt) + buf * PANEL_BYTES + row * 128;
      const unsigned long long sc2 = pack_f2(p.scale_log2, p.scale_log2), scl2 = pack_f2(p.scale, p.scale);
      // Two copies of the step body, chosen per step: only the steps on the diagonal (two per head) carry the causal
      // compare / select per element; a single body with `if (diag && ...)` still computed every index and comparison
      // on every step — a third of its instructions.
      auto half_steps = [&](auto masked) {
        constexpr bool MASK = decltype(masked)::value;
  #pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t sv[32], dv[32];
          tmem_ld_32x32(tlane + C_ST + buf * QS + hf * 32, sv);
          tmem_ld_32x32(tlane + C_DPT + buf * QS + hf * 32, dv);
          tmem_ld_wait();
          if (hf == 1) {   // both halves are in registers: the MMA issuer may overwrite this TMEM buffer
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&st_empty[buf]);
          }
          uint32_t pp[16], dd[16];
  #pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 l4 = lds_f4(lse_s + hf * 128 + i4 * 16);     // -lse * log2e
            const float4 d4 = lds_f4(del_s + hf * 128 + i4 * 16);     // -delta
            const float nls[4] = {l4.x, l4.y, l4.z, l4.w}, nds[4] = {d4.x, d4.y, d4.z, d4.w};
            float pv[4], dsv[4];
  #pragma unroll
            for (int e2 = 0; e2 < 2; ++e2) {
              const int c = i4 * 4 + 2 * e2;
              // packed pairs: P^T = exp2(s c - lse[q]), dS^T = P^T (dP^T - delta[q]) scale
              float a0, a1;
              unpack_f2(ffma2(pack_f2(__uint_as_float(sv[c]), __uint_as_float(sv[c + 1])), sc2,
                              pack_f2(nls[2 * e2], nls[2 * e2 + 1])), a0, a1);
              float pr[2];
              if ((B200_ATTN_POLY_DKV >> ((i4 * 2 + e2) & 7)) & 1) {
                exp2_poly_pair(a0, a1, pr[0], pr[1]);
              } else {
                pr[0] = fast_exp2(a0);
                pr[1] = fast_exp2(a1);
              }
              float dpv[2] = {__uint_as_float(dv[c]), __uint_as_float(dv[c + 1])};
              float pdrop[2];
  #pragma unroll
              for (int e = 0; e < 2; ++e) {
                if (MASK && key > qrow0 + hf * 32 + c + e) pr[e] = 0.f;
                pdrop[e] = pr[e];
                if (DROP) {
                  const unsigned long long qg = (unsigned long long)(qrow0 + hf * 32 + c + e);
                  const bool kp = dropout_keep((((drop_bh * p.T + qg) * (unsigned long long)((p.T + 1) >> 1)) << 1) +
                                                   ((unsigned long long)(key >> 1) << 1) + (unsigned long long)(key & 1),
                                               dseed, p.drop.thr16);
                  dpv[e] = kp ? dpv[e] * p.drop.inv_keep : 0.f;
                  pdrop[e] = kp ? pr[e] * p.drop.inv_keep : 0.f;
                }
                pv[2 * e2 + e] = pdrop[e];                    // dV uses the dropped, rescaled P
              }
              unpack_f2(fmul2(fmul2(pack_f2(pr[0], pr[1]), fadd2(pack_f2(dpv[0], dpv[1]), pack_f2(nds[2 * e2], nds[2 * e2 + 1]))),
                              scl2), dsv[2 * e2], dsv[2 * e2 + 1]);
            }
            pp[i4 * 2] = f2_to_bf2(pv[0], pv[1]);
            pp[i4 * 2 + 1] = f2_to_bf2(pv[2], pv[3]);
            dd[i4 * 2] = f2_to_bf2(dsv[0], dsv[1]);
            dd[i4 * 2 + 1] = f2_to_bf2(dsv[2], dsv[3]);
          }
            if (DT) {
            // this thread's key row: 32 query columns of this half = 16 packed columns, over the S^T / dP^T columns
            // already in registers (half 1 of S^T / dP^T sits in columns 32..63 and is untouched by half 0's write)
            tmem_st_32x16(tlane + C_ST + buf * QS + hf * 16, pp);
            tmem_st_32x16(tlane + C_DPT + buf * QS + hf * 16, dd);
          } else {
            if (hf == 0 && use > 0) mbar_wait(&panel_free[buf], (use - 1) & 1);
#pragma unroll
            for (int t4 = 0; t4 < 4; ++t4) {
              const int ch = (hf * 4 + t4) ^ (row & 7);
              sts_v4(prow + (ch << 4), pp[4 * t4], pp[4 * t4 + 1], pp[4 * t4 + 2], pp[4 * t4 + 3]);
              sts_v4(drow + (ch << 4), dd[4 * t4], dd[4 * t4 + 1], dd[4 * t4 + 2], dd[4 * t4 + 3]);
            }
          }
        }
      };
      if (diag) half_steps(std::true_type{});
      else half_steps(std::false_type{});
      if (DT) tmem_st_wait();
      else fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ps_full[buf]);
    }
    mbar_wait(acc_done, 0);
    tc_fence_after();
    // epilogue: the groups share the work of their key row round-robin. Unit u < HD/64 = the chunk pair
    // (u, u + HD/64) of dV and of dK (first / second half of the head: what the RoPE gradient combines).
    constexpr int HC = HD / 64;
#pragma unroll 1
    for (int e = grp; e < 2 * HC; e += NG) {
      bf16* dkrow = p.dk + (long long)(row_base + key) * p.lddk + (long long)kvh * HD;
      bf16* dvrow = p.dv + (long long)(row_base + key) * p.lddv + (long long)kvh * HD;
      const int which = e / HC, u = e % HC;              // which = 0: dV, 1: dK
      {
        uint32_t a[32], c2[32];
        const uint32_t col = which ? C_DK : C_DV;
        tmem_ld_32x32(tlane + col + u * 32, a);
        tmem_ld_32x32(tlane + col + (u + HC) * 32, c2);
        tmem_ld_wait();
        if (key < p.T) {
          if (which && p.rope_cos != nullptr)
            rope_inverse_pair(a, c2, p.rope_cos + (long long)key * HD + u * 32, p.rope_sin + (long long)key * HD + u * 32);
          bf16* dst = which ? dkrow : dvrow;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float f[8], g[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              f[i] = __uint_as_float(a[c * 8 + i]);
              g[i] = __uint_as_float(c2[c * 8 + i]);
            }
            st_v4(dst + u * 32 + c * 8, pack8(f));
            st_v4(dst + (u + HC) * 32 + c * 8, pack8(g));
          }
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

// ---- dQ: CTA = 128 query rows of one (batch, head); streams 64-key tiles up to the diagonal ----
//   S = Q K_j^T, dP = dO V_j^T (TMEM)  ->  dS = exp2(S c - lse) (dP - delta) scale  (thread = query
//   row, lse / delta live in registers)  ->  bf16 smem  ->  dQ += dS K_j (K as MN-major B), dQ
//   accumulates in TMEM over the key loop and is written once: no atomics, deterministic.
// Fused dQ inside the key-tile kernel was limited by the SM's RED issue rate (8192 fp32 atomics per
// 128x128 step ~ 10.5 k cycles); recomputing S / dP here costs two extra MMAs per tile instead.
constexpr int DQ_NS = 2;
template <int HDIM>
struct DqCfg {
  static constexpr int Q_BYTES = BQ * HDIM * 2, KV_BYTES = BKV * HDIM * 2;
  static constexpr int SMEM = 2 * Q_BYTES + DQ_NS * 2 * KV_BYTES + 2 * P_BYTES + 1024 + 256;   // 97 / 161 KB
  static constexpr int CTAS = HDIM == 64 ? 2 : 1;
};

struct DqParams {
  int B, T, H, HKV, nqt;
  DropParams drop;
  float scale, scale_log2;
  const float* ld2;   // [B, H, nq64, 2, 64]
  int nq64;
  bf16* dq;
  long long lddq;
  const bf16* rope_cos;   // see BwdParams
  const bf16* rope_sin;
  // DT kernels only, lse != nullptr: no attn_delta_tc_kernel ran — this kernel forms -lse * log2e and -delta = -sum dO * O
  // for its own query rows (the O tile comes in next to Q and dO, into the shared memory the dS panels do not use) and
  // writes them to ld2_out for the dK/dV kernel that follows
  const float* lse;
  float* ld2_out;
};

// DT: dS goes to tensor memory (two 32-column buffers) and feeds dQ += dS K as the MMA's A operand from there instead
// of two swizzled shared-memory tiles (see attn_fwd_tc_kernel's PT): 256 TMEM columns at head_dim 64, 512 at 128.
template <int HDIM, bool DROP, bool DT>
__global__ void __launch_bounds__(224, DqCfg<HDIM>::CTAS)
attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                      const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                      const __grid_constant__ CUtensorMap tmO, const DqParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int HD = HDIM, Q_BYTES = DqCfg<HDIM>::Q_BYTES, KV_BYTES = DqCfg<HDIM>::KV_BYTES;
  uint8_t* sQ = smem;
  uint8_t* sdO = sQ + Q_BYTES;
  uint8_t* sK = sdO + Q_BYTES;
  uint8_t* sV = sK + DQ_NS * KV_BYTES;
  uint8_t* sdS = sV + DQ_NS * KV_BYTES;   // [2] x [128 x 64] bf16
  uint64_t* bars = (uint64_t*)(sdS + 2 * P_BYTES);
  uint64_t* q_full = bars;                // 1
  uint64_t* kv_full = bars + 1;           // DQ_NS
  uint64_t* kv_empty = kv_full + DQ_NS;   // DQ_NS
  uint64_t* s_full = kv_empty + DQ_NS;    // 1
  uint64_t* s_empty = s_full + 1;         // 1 (4 arrivals)
  uint64_t* ds_full = s_empty + 1;        // 2 (4 arrivals)
  uint64_t* dq_done = ds_full + 2;        // 1
  uint64_t* ds_free = dq_done + 1;        // 2: the dQ product that read this dS buffer has retired
  uint32_t* tmem_slot = (uint32_t*)(ds_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_qt = p.B * p.H;
  const int qt = p.nqt - 1 - (int)(blockIdx.x / per_qt);
  const int rem = (int)(blockIdx.x % per_qt);
  const int head = rem % p.H, b = rem / p.H;
  const int kvh = head / (p.H / p.HKV);
  const int q0 = qt * BQ;
  const int row_base = b * p.T;
  const int n_kv = (min(q0 + BQ, p.T) + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < DQ_NS; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_empty, 4);
    mbar_init(&ds_full[0], 4);
    mbar_init(&ds_full[1], 4);
    mbar_init(dq_done, 1);
    mbar_init(&ds_free[0], 1);
    mbar_init(&ds_free[1], 1);
    mbar_fence_init();
  }
  constexpr uint32_t DQ_TMEM = (DT && HDIM == 128) ? 512 : 256;
  if (warp == 1) tmem_alloc(tmem_slot, DQ_TMEM);
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
  pdl_launch_dependents();   // barriers, TMEM and descriptors are set up: the rest waits for the previous kernel's results
  pdl_wait();
  constexpr uint32_t D_S = 0, D_DP = 64, D_DQ = 128, D_DS = 128 + HDIM;    // D_DS: DT only, 2 x 32 columns

  if (warp == 0) {
    if (lane == 0) {
      const bool fold = DT && p.lse != nullptr;
      mbar_expect_tx(q_full, (fold ? 3 : 2) * Q_BYTES);
      tma_load_tile<HD>(sQ, &tmQ, q_full, head * HD, row_base + q0, BQ);
      tma_load_tile<HD>(sdO, &tmdO, q_full, head * HD, row_base + q0, BQ);
      if (fold) tma_load_tile<HD>(sdS, &tmO, q_full, head * HD, row_base + q0, BQ);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % DQ_NS;
        mbar_wait(&kv_empty[s], ((j / DQ_NS) & 1) ^ 1);
        mbar_expect_tx(&kv_full[s], 2 * KV_BYTES);
        tma_load_tile<HD>(sK + s * KV_BYTES, &tmK, &kv_full[s], kvh * HD, row_base + j * BKV, BKV);
        tma_load_tile<HD>(sV + s * KV_BYTES, &tmV, &kv_full[s], kvh * HD, row_base + j * BKV, BKV);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t id_s = umma_idesc_bf16(BQ, BKV, 0, 0);
      const uint32_t aQ = smem_u32(sQ), adO = smem_u32(sdO);
      mbar_wait(q_full, 0);
      auto issue_s = [&](int j) {
        const int s = j % DQ_NS;
        mbar_wait(&kv_full[s], (j / DQ_NS) & 1);
        if (j > 0) mbar_wait(s_empty, (j - 1) & 1);
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + s * KV_BYTES), aV = smem_u32(sV + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + D_S, umma_desc_k128(aQ + kstep<BQ>(k)), umma_desc_k128(aK + kstep<BKV>(k)), id_s, k > 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16_ss(tmem_base + D_DP, umma_desc_k128(adO + kstep<BQ>(k)), umma_desc_k128(aV + kstep<BKV>(k)), id_s, k > 0);
        umma_commit(s_full);
      };
      // score products only (throttled by s_empty: S / dP are single TMEM buffers); dQ += dS K is issued by warp 2
      for (int j = 0; j < n_kv; ++j) issue_s(j);
    }
  } else if (warp == 2) {
    if (lane == 0) {
      const uint32_t id_dq = umma_idesc_bf16(BQ, HD, 0, 1);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % DQ_NS, buf = j & 1;
        mbar_wait(&kv_full[s], (j / DQ_NS) & 1);
        mbar_wait(&ds_full[buf], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t adS = smem_u32(sdS + buf * P_BYTES), aK = smem_u32(sK + s * KV_BYTES);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          if (DT)
            umma_bf16_ts(tmem_base + D_DQ, tmem_base + D_DS + buf * 32 + k * 8, umma_desc_mn128(aK + k * 2048, 8192), id_dq,
                         (j > 0 || k > 0) ? 1u : 0u);
          else
            umma_bf16_ss(tmem_base + D_DQ, umma_desc_k128(adS + k * 32), umma_desc_mn128(aK + k * 2048, 8192),
                         id_dq, (j > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&kv_empty[s]);
        umma_commit(&ds_free[buf]);
      }
      umma_commit(dq_done);
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int t = q0 + row;
    const long long ldoff = (((long long)b * p.H + head) * p.nq64 + (t >> 6)) * (2 * QS) + (t & 63);
    float nlse2, ndl;                // -lse * log2e (-inf for padding rows), -delta
    if (DT && p.lse != nullptr) {
      mbar_wait(q_full, 0);
      float acc = 0.f;
#pragma unroll
      for (int pnl = 0; pnl < HD / 64; ++pnl) {
        const uint32_t off = (uint32_t)(pnl * BQ * 128 + row * 128);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint32_t o16 = off + (uint32_t)((c ^ (row & 7)) << 4);
          const float4 fg = lds_f4(smem_u32(sdO) + o16), fo = lds_f4(smem_u32(sdS) + o16);
          float g[8], a[8];
          unpack8(make_uint4(__float_as_uint(fg.x), __float_as_uint(fg.y), __float_as_uint(fg.z), __float_as_uint(fg.w)), g);
          unpack8(make_uint4(__float_as_uint(fo.x), __float_as_uint(fo.y), __float_as_uint(fo.z), __float_as_uint(fo.w)), a);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc += a[i] * g[i];
        }
      }
      const bool real = t < p.T;
      nlse2 = real ? -p.lse[((long long)b * p.H + head) * p.T + t] * LOG2E : -INFINITY;
      ndl = real ? -acc : 0.f;
      if ((t >> 6) < p.nq64) {
        p.ld2_out[ldoff] = nlse2;
        p.ld2_out[ldoff + QS] = ndl;
      }
    } else {
      nlse2 = p.ld2[ldoff];
      ndl = p.ld2[ldoff + QS];
    }
    const unsigned long long sc2 = pack_f2(p.scale_log2, p.scale_log2), nl2 = pack_f2(nlse2, nlse2);
    const unsigned long long nd2 = pack_f2(ndl, ndl), scl2 = pack_f2(p.scale, p.scale);
    const unsigned long long drop_row =
        (((unsigned long long)b * p.H + head) * p.T + (unsigned long long)t) * (unsigned long long)((p.T + 1) >> 1);
    const unsigned long long dseed = DROP ? drop_seed(p.drop) : 0ull;
    for (int j = 0; j < n_kv; ++j) {
      const int buf = j & 1;
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int kv0 = j * BKV;
      const bool diag = kv0 + BKV - 1 > q0 + quad * 32;      // warp-uniform
      uint32_t pk[32];
      // 32 key columns at a time (S and dP halves: 64 live registers instead of 128, no spills at two CTAs per SM).
      // Two copies of the element loop, chosen per tile: only the (few) tiles on the diagonal carry the causal
      // compare / select per element. As one loop with `if (diag && ...)` the compiler predicated the selects but
      // still computed every column index and comparison on every tile — a third of the loop's instructions.
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t s[32], d[32];
        tmem_ld_32x32(tlane + D_S + hf * 32, s);
        tmem_ld_32x32(tlane + D_DP + hf * 32, d);
        tmem_ld_wait();
        if (hf == 1) {       // S and dP are in registers: the score MMAs of the next tile may overwrite them
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(s_empty);
        }
        auto elements = [&](auto masked) {
          constexpr bool MASK = decltype(masked)::value;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int col = kv0 + hf * 32 + 2 * i;
            bool keep[2] = {true, true};
            if (DROP)
              dropout_keep2(drop_row + (unsigned long long)(col >> 1), dseed, p.drop.thr16, keep[0], keep[1]);
            // packed pairs: P = exp2(s c - lse), dS = P (dP - delta) scale
            float a0, a1;
            unpack_f2(ffma2(pack_f2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sc2, nl2), a0, a1);
            float pr0, pr1;
            if ((B200_ATTN_POLY_DQ >> (i & 7)) & 1) {
              exp2_poly_pair(a0, a1, pr0, pr1);
            } else {
              pr0 = fast_exp2(a0);
              pr1 = fast_exp2(a1);
            }
            if (MASK) {
              if (col > t) pr0 = 0.f;
              if (col + 1 > t) pr1 = 0.f;
            }
            float dp0 = __uint_as_float(d[2 * i]), dp1 = __uint_as_float(d[2 * i + 1]);
            if (DROP) {
              dp0 = keep[0] ? dp0 * p.drop.inv_keep : 0.f;
              dp1 = keep[1] ? dp1 * p.drop.inv_keep : 0.f;
            }
            float v0, v1;
            unpack_f2(fmul2(fmul2(pack_f2(pr0, pr1), fadd2(pack_f2(dp0, dp1), nd2)), scl2), v0, v1);
            pk[hf * 16 + i] = f2_to_bf2(v0, v1);
          }
        };
        if (diag) elements(std::true_type{});
        else elements(std::false_type{});
      }
      // buffer `buf` was last read by the dQ product of tile j - 2 (another warp issues those)
      if (j >= 2) mbar_wait(&ds_free[buf], ((j >> 1) - 1) & 1);
      if (DT) {      // dS row -> this thread's TMEM lane: keys 2c, 2c + 1 in column c of the buffer
        tc_fence_after();
        tmem_st_32x32(tlane + D_DS + buf * 32, pk);
        tmem_st_wait();
      } else {
        const uint32_t drow = smem_u32(sdS) + buf * P_BYTES + row * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          sts_v4(drow + ((c ^ (row & 7)) << 4), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        fence_proxy_async_smem();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ds_full[buf]);
    }
    mbar_wait(dq_done, 0);
    tc_fence_after();
    bf16* orow = p.dq + (long long)(row_base + t) * p.lddq + (long long)head * HD;
    constexpr int HC = HD / 64;          // chunk u and chunk u + HC: first / second half of the head
#pragma unroll 1
    for (int u = 0; u < HC; ++u) {
      uint32_t o[32], o2[32];
      tmem_ld_32x32(tlane + D_DQ + u * 32, o);
      tmem_ld_32x32(tlane + D_DQ + (u + HC) * 32, o2);
      tmem_ld_wait();
      if (t < p.T) {
        if (p.rope_cos != nullptr)      // Q was rotated in the projection's epilogue: dQ leaves through the transposed rotation
          rope_inverse_pair(o, o2, p.rope_cos + (long long)t * HD + u * 32, p.rope_sin + (long long)t * HD + u * 32);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float f[8], g[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            f[i] = __uint_as_float(o[c * 8 + i]);
            g[i] = __uint_as_float(o2[c * 8 + i]);
          }
          st_v4(orow + u * 32 + c * 8, pack8(f));
          st_v4(orow + (u + HC) * 32 + c * 8, pack8(g));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, DQ_TMEM);
  }
}

// ld2[b, h, step, 0, j] = -lse[b, h, 64*step + j] * log2e (-inf beyond T),
// ld2[b, h, step, 1, j] = -delta = -sum_c dO * O (0 beyond T); HDIM/8 lanes per (token, head).
template <int HDIM>
__global__ void __launch_bounds__(256)
attn_delta_tc_kernel(const bf16* __restrict__ o, long long ldo, const bf16* __restrict__ dout,
                     long long lddo, const float* __restrict__ lse, float* __restrict__ ld2, long long B,
                     int T, int H, int nq64) {
  constexpr int LPH = HDIM / 8;      // lanes per (token, head)
  constexpr int IPW = 32 / LPH;      // (token, head) items per warp
  const long long Tpad = (long long)nq64 * QS;
  const long long total = B * Tpad * H;
  const int lane = threadIdx.x & 31, sub = lane / LPH, l8 = lane % LPH;
  for (long long w = ((long long)blockIdx.x * 8 + (threadIdx.x >> 5)) * IPW + sub; w < total + IPW - 1;
       w += (long long)gridDim.x * 8 * IPW) {
    float acc = 0.f;
    const bool ok = w < total;
    const int head = ok ? (int)(w % H) : 0;
    const long long tp = ok ? w / H : 0;
    const long long bb = tp / Tpad;
    const int t = (int)(tp % Tpad);
    const bool real = ok && t < T;
    if (real) {
      const long long tok = bb * T + t;
      float a[8], g[8];
      unpack8(ld_nc_v4(o + tok * ldo + (long long)head * HDIM + l8 * 8), a);
      unpack8(ld_nc_v4(dout + tok * lddo + (long long)head * HDIM + l8 * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += a[j] * g[j];
    }
#pragma unroll
    for (int m = 1; m < LPH; m <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
    if (ok && l8 == 0) {
      float* dst = ld2 + ((bb * H + head) * nq64 + (t >> 6)) * (2 * QS) + (t & 63);
      dst[0] = real ? -lse[(bb * H + head) * T + t] * LOG2E : -INFINITY;     // addends of the packed FFMA / FADD
      dst[QS] = real ? -acc : 0.f;
    }
  }
}

}  // namespace

size_t b200_attention_bwd_tc_workspace(int64_t B, int64_t T, int64_t h) {
  const int64_t nq64 = (T + QS - 1) / QS;
  return (size_t)(B * h * nq64 * 2 * QS) * 4 + 256;
}

namespace {
template <int HD>
int attention_bwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                     const void* o, int64_t ldo, const void* dout, int64_t lddo, const float* lse,
                     void* workspace, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv,
                     int64_t lddv, int64_t B, int64_t T, int64_t h, int64_t hkv, float scale,
                     float dropout_p, uint64_t seed, const uint64_t* seed_dev, const void* rope_cos, const void* rope_sin,
                     cudaStream_t st) {
  constexpr int DKV_SMEM = DkvCfg<HD>::SMEM, DKV_THREADS = DkvCfg<HD>::THREADS, DQ_SMEM = DqCfg<HD>::SMEM;
  float* ld2 = (float*)workspace;   // cudaMalloc / caching-allocator memory: at least 256-B aligned
  B200_CHECK_ARG((((uintptr_t)ld2) & 15u) == 0, "attention_bwd: workspace must be 16-B aligned");
  const long long rows = B * T;
  const int nq64 = (int)((T + QS - 1) / QS);
  // dS through TMEM: 10 us (hd 64, B4 T2048) / 19 us (hd 128, B1 T4096) faster; B200_ATTN_DS_TMEM=0 restores the panels
  static const bool ds_tmem = [] { const char* e = getenv("B200_ATTN_DS_TMEM"); return !(e && e[0] == '0'); }();
  // ... and leaves the dS panels' shared memory free for the O tile: the dQ kernel forms lse2 / delta itself
  // (B200_ATTN_DELTA_FOLD=0 keeps the separate pass; the dropout variant always uses it)
  static const bool delta_fold = [] { const char* e = getenv("B200_ATTN_DELTA_FOLD"); return !(e && e[0] == '0'); }();
  const bool fold = ds_tmem && delta_fold && dropout_p <= 0.f;
  if (!fold) {
    const long long total = B * nq64 * QS * h;
    long long blocks = (total + 31) / 32;
    const long long cap = (long long)b200_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    attn_delta_tc_kernel<HD><<<(int)blocks, 256, 0, st>>>((const bf16*)o, ldo, (const bf16*)dout, lddo, lse, ld2, B, (int)T, (int)h, nq64);
    B200_CHECK_LAUNCH();
  }
  // every map has a 64-column box: a [rows x HD] tile is fetched as HD/64 panels
  CUtensorMap mq, mdo, mk, mv, mk64, mv64, mo;
  int rc;
  if ((rc = b200_make_tmap_2d_bf16(&mo, o, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)ldo, BK2, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mq, q, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)ldq, BK2, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mdo, dout, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)lddo, BK2, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mk, k, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldk, BK2, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mv, v, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldv, BK2, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mk64, k, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldk, BKV, 64))) return rc;
  if ((rc = b200_make_tmap_2d_bf16(&mv64, v, (uint64_t)rows, (uint64_t)(hkv * HD), (uint64_t)ldv, BKV, 64))) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<HD, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DKV_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<HD, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DKV_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<HD, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DKV_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<HD, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<HD, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<HD, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM));
    attr_set[dev & 63] = true;
  }
  const int nqt = (int)((T + BK2 - 1) / BK2);
  {
    DqParams p;
    p.B = (int)B; p.T = (int)T; p.H = (int)h; p.HKV = (int)hkv; p.nqt = nqt;
    p.scale = scale; p.scale_log2 = scale * LOG2E;
    p.drop = make_drop(dropout_p, seed, seed_dev);
    p.ld2 = ld2; p.nq64 = nq64; p.dq = (bf16*)dq; p.lddq = lddq;
    p.rope_cos = (const bf16*)rope_cos; p.rope_sin = (const bf16*)rope_sin;
    p.lse = fold ? lse : nullptr; p.ld2_out = ld2;
    const long long grid = (long long)nqt * B * h;
    if (p.drop.thr16) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dq_tc_kernel<HD, true, false>, dim3((unsigned)grid), dim3(224), DQ_SMEM, st, mq, mdo, mk64, mv64, mo, p));
    else if (ds_tmem) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dq_tc_kernel<HD, false, true>, dim3((unsigned)grid), dim3(224), DQ_SMEM, st, mq, mdo, mk64, mv64, mo, p));
    else B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dq_tc_kernel<HD, false, false>, dim3((unsigned)grid), dim3(224), DQ_SMEM, st, mq, mdo, mk64, mv64, mo, p));
    B200_CHECK_LAUNCH();
  }
  {
    CUtensorMap mq64, mdo64;
    if ((rc = b200_make_tmap_2d_bf16(&mq64, q, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)ldq, QS, 64))) return rc;
    if ((rc = b200_make_tmap_2d_bf16(&mdo64, dout, (uint64_t)rows, (uint64_t)(h * HD), (uint64_t)lddo, QS, 64))) return rc;
    BwdParams p;
    p.B = (int)B; p.T = (int)T; p.H = (int)h; p.HKV = (int)hkv;
    p.nkt = nqt; p.nq64 = nq64;
    p.scale = scale; p.scale_log2 = scale * LOG2E;
    p.drop = make_drop(dropout_p, seed, seed_dev);
    p.ld2 = ld2;
    p.dk = (bf16*)dk; p.lddk = lddk; p.dv = (bf16*)dv; p.lddv = lddv;
    p.rope_cos = (const bf16*)rope_cos; p.rope_sin = (const bf16*)rope_sin;
    const long long grid = (long long)p.nkt * B * hkv;
    // P^T / dS^T through TMEM: 14 us faster at hd 64 (B4 T2048); B200_ATTN_DKV_TMEM=0 restores the shared-memory panels
    static const bool pt_tmem = [] { const char* e = getenv("B200_ATTN_DKV_TMEM"); return !(e && e[0] == '0'); }();
    if (p.drop.thr16) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dkv_tc_kernel<HD, true, false>, dim3((unsigned)grid), dim3(DKV_THREADS), DKV_SMEM, st, mq64, mdo64, mk, mv, p));
    else if (pt_tmem) B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dkv_tc_kernel<HD, false, true>, dim3((unsigned)grid), dim3(DKV_THREADS), DKV_SMEM, st, mq64, mdo64, mk, mv, p));
    else B200_CHECK_CUDA(b200_launch_pdl_wide(attn_bwd_dkv_tc_kernel<HD, false, false>, dim3((unsigned)grid), dim3(DKV_THREADS), DKV_SMEM, st, mq64, mdo64, mk, mv, p));
    B200_CHECK_LAUNCH();
  }
  b200_count_launch(fold ? 2 : 3);
  return 0;
}
}  // namespace

int b200_attention_bwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, const void* o, int64_t ldo, const void* dout, int64_t lddo,
                          const float* lse, void* workspace, void* dq, int64_t lddq, void* dk,
                          int64_t lddk, void* dv, int64_t lddv, int64_t B, int64_t T, int64_t h,
                          int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed, const uint64_t* seed_dev,
                          const void* rope_cos, const void* rope_sin, cudaStream_t st) {
  if (hd == 64)
    return attention_bwd_tc<64>(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, dq, lddq, dk, lddk, dv,
                                lddv, B, T, h, hkv, scale, dropout_p, seed, seed_dev, rope_cos, rope_sin, st);
  return attention_bwd_tc<128>(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, dq, lddq, dk, lddk, dv,
                               lddv, B, T, h, hkv, scale, dropout_p, seed, seed_dev, rope_cos, rope_sin, st);
}


// keep mask of the attention-weight dropout for tests / oracles: out[b, h, q, k] in {0, 1}
namespace {
__global__ void __launch_bounds__(256)
attn_dropout_mask_kernel(unsigned char* __restrict__ out, long long B, int H, int T, DropParams d) {
  const long long total = B * H * (long long)T * T;
  const unsigned long long tp2 = (unsigned long long)((T + 1) >> 1);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % T);
    const unsigned long long row = (unsigned long long)(i / T);   // (b*H + h)*T + q
    out[i] = dropout_keep(((row * tp2 + (unsigned long long)(k >> 1)) << 1) + (unsigned long long)(k & 1), d.seed, d.thr16) ? 1 : 0;
  }
}
}  // namespace

extern "C" int b200_attention_dropout_mask_u8(void* out, int64_t B, int64_t H, int64_t T, float dropout_p,
                                              uint64_t seed, void* stream) {
  B200_CHECK_ARG(out && B > 0 && H > 0 && T > 0 && dropout_p >= 0.f && dropout_p < 1.f, "attention_dropout_mask: bad arguments");
  attn_dropout_mask_kernel<<<1024, 256, 0, (cudaStream_t)stream>>>((unsigned char*)out, B, (int)H, (int)T, make_drop(dropout_p, seed));
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
