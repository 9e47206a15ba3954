// K2 — fused causal MHA / GQA attention: C-ABI entry points (argument checks, dispatch) and the KV-cache
// decode kernel.
//
// Flash-style: the T x T score matrix of Models/Llama/Llama3.py:148-153 / Models/GPT2/GPT2.py:38-46
// is never materialised, repeat_interleave of K/V (Llama3.py:144-145) is replaced by indexing the kv
// head, and the causal mask buffer by a predicate. Head dims 64 (GPT-2, Llama-3.2-1B) and 128
// (Llama-2-7B, Llama-3-8B) both run the tcgen05 / TMEM kernels of attention_tc.cu.
//
// Roofline: tensor pipe / MUFU. Algorithmic FLOPs (causal): fwd 2*T*T*hd per (batch, head),
// bwd 2.5x that (recompute included: 3.5x executed).
#include "common.cuh"


// tcgen05 paths (attention_tc.cu), head_dim 64 and 128
int b200_attention_fwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, void* o, int64_t ldo, float* lse, int64_t B, int64_t T,
                          int64_t h, int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed,
                          const uint64_t* seed_dev, cudaStream_t st);

size_t b200_attention_bwd_tc_workspace(int64_t B, int64_t T, int64_t h);
int b200_attention_bwd_tc(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                          int64_t ldv, const void* o, int64_t ldo, const void* dout, int64_t lddo,
                          const float* lse, void* workspace, void* dq, int64_t lddq, void* dk,
                          int64_t lddk, void* dv, int64_t lddv, int64_t B, int64_t T, int64_t h,
                          int64_t hkv, int64_t hd, float scale, float dropout_p, uint64_t seed, const uint64_t* seed_dev,
                          const void* rope_cos, const void* rope_sin, cudaStream_t st);

namespace {

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
                                       float scale, float dropout_p, uint64_t seed, const uint64_t* seed_dev, void* stream) {
  const int64_t lds[4] = {ldq, ldk, ldv, ldo};
  int rc = check_common(q, k, v, B, T, h, hkv, hd, lds, 4);
  if (rc) return rc;
  B200_CHECK_ARG(o && lse, "attention_fwd: null output");
  B200_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "attention: dropout_p=%f must be in [0, 1)", (double)dropout_p);
  return b200_attention_fwd_tc(q, ldq, k, ldk, v, ldv, o, ldo, lse, B, T, h, hkv, hd, scale, dropout_p, seed, seed_dev,
                               (cudaStream_t)stream);
}

extern "C" size_t b200_attention_bwd_workspace_bytes(int64_t B, int64_t T, int64_t h, int64_t hd) {
  (void)hd;
  return b200_attention_bwd_tc_workspace(B, T, h);
}

namespace {
int attention_bwd_entry(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                        const void* o, int64_t ldo, const void* dout, int64_t lddo, const float* lse, void* workspace,
                        size_t workspace_bytes, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                        int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd, float scale, float dropout_p,
                        uint64_t seed, const uint64_t* seed_dev, const void* rope_cos, const void* rope_sin, void* stream) {
  const int64_t lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
  int rc = check_common(q, k, v, B, T, h, hkv, hd, lds, 8);
  if (rc) return rc;
  B200_CHECK_ARG(o && dout && lse && workspace && dq && dk && dv, "attention_bwd: null pointer");
  B200_CHECK_ARG(workspace_bytes >= b200_attention_bwd_workspace_bytes(B, T, h, hd), "attention_bwd: workspace too small");
  B200_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "attention: dropout_p=%f must be in [0, 1)", (double)dropout_p);
  B200_CHECK_ARG((rope_cos == nullptr) == (rope_sin == nullptr) && b200_aligned16(rope_cos) && b200_aligned16(rope_sin),
                 "attention_bwd: the RoPE tables come as a 16-B aligned pair");
  return b200_attention_bwd_tc(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, dq, lddq, dk, lddk, dv,
                               lddv, B, T, h, hkv, hd, scale, dropout_p, seed, seed_dev, rope_cos, rope_sin,
                               (cudaStream_t)stream);
}
}  // namespace

extern "C" int b200_attention_bwd_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk,
                                       const void* v, int64_t ldv, const void* o, int64_t ldo,
                                       const void* dout, int64_t lddo, const float* lse,
                                       void* workspace, size_t workspace_bytes, void* dq,
                                       int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                                       int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd,
                                       float scale, float dropout_p, uint64_t seed, const uint64_t* seed_dev, void* stream) {
  return attention_bwd_entry(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, workspace_bytes, dq, lddq, dk,
                             lddk, dv, lddv, B, T, h, hkv, hd, scale, dropout_p, seed, seed_dev, nullptr, nullptr, stream);
}

// Backward for q / k that were rotated (RoPE) BEFORE the attention kernels saw them — by the q|k|v projection's
// epilogue (b200_gemm_qkv_rope_bf16): dq and dk come out already multiplied by the transposed rotation, i.e. as the
// gradients of the un-rotated projections, so no separate inverse-RoPE pass runs. cos / sin: [>= T, hd] bf16 tables
// whose two column halves are equal (precompute_rope_params builds them as cat([angles, angles])).
extern "C" int b200_attention_bwd_rope_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk,
                                            const void* v, int64_t ldv, const void* o, int64_t ldo,
                                            const void* dout, int64_t lddo, const float* lse,
                                            void* workspace, size_t workspace_bytes, void* dq,
                                            int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                                            int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd,
                                            float scale, const void* cos, const void* sin, void* stream) {
  B200_CHECK_ARG(cos && sin, "attention_bwd_rope: null RoPE tables");
  return attention_bwd_entry(q, ldq, k, ldk, v, ldv, o, ldo, dout, lddo, lse, workspace, workspace_bytes, dq, lddq, dk,
                             lddk, dv, lddv, B, T, h, hkv, hd, scale, 0.f, 0, nullptr, cos, sin, stream);
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
