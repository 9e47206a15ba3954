// fp32 parity mode (north star: per-op outputs and gradients within 1e-4 relative in fp32).
//
// The reference's GPT-2 modules are fp32 as written (Models/GPT2/GPT2.py:20-23,59-61 ignore cfg["dtype"])
// and its Llama pieces can be built with dtype=float32 at sub-module level (RMSNorm, compute_rope, FeedForward,
// GroupedQueryAttention). These kernels evaluate the same operators in plain fp32 — FFMA accumulation, no tensor
// cores (bf16 / tf32 inputs would cost 1e-3) — so the fp32 fixtures recorded from the reference replay at 1e-4 on the
// GPU. They are a CORRECTNESS mode: simple SIMT kernels, not tuned (no BASELINE config trains in fp32 on a GPU).
//
//   b200_gemm_f32            C = alpha * A * B^T (+ bias) (+ C), arbitrary row / column strides, batched
//   b200_softmax_causal_f32  row softmax of the scaled causal scores, and its backward
//   b200_layernorm_*_f32, b200_gelu_*_f32, b200_swiglu_*_f32, b200_rope_f32, b200_cross_entropy_f32,
//   b200_embedding_*_f32, b200_add_f32
#include "common.cuh"

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

// C[b][m, n] = alpha * sum_k A[b][m, k] * B[b][n, k] (+ bias[n]) (+ beta * C[b][m, n]);  X[b][i, k] = X[b * bs + i * rs + k * cs]
__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ A, long long a_bs, long long a_rs, long long a_cs, const float* __restrict__ B,
                long long b_bs, long long b_rs, long long b_cs, const float* __restrict__ bias, float* __restrict__ C,
                long long c_bs, long long c_rs, long long c_cs, int M, int N, int K, float alpha, float beta) {
  __shared__ float sa[TK][TM + 1], sb[TK][TN + 1];
  const int b = blockIdx.z;
  A += (long long)b * a_bs;
  B += (long long)b * b_bs;
  C += (long long)b * c_bs;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;      // 16 x 16 threads, 4 x 4 outputs each
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int i = threadIdx.x; i < TM * TK; i += 256) {
      // consecutive threads walk the dimension that is contiguous in memory
      const int r = a_cs == 1 ? i / TK : i % TM, k = a_cs == 1 ? i % TK : i / TM;
      sa[k][r] = (m0 + r < M && k0 + k < K) ? A[(long long)(m0 + r) * a_rs + (long long)(k0 + k) * a_cs] : 0.f;
    }
    for (int i = threadIdx.x; i < TN * TK; i += 256) {
      const int r = b_cs == 1 ? i / TK : i % TN, k = b_cs == 1 ? i % TK : i / TN;
      sb[k][r] = (n0 + r < N && k0 + k < K) ? B[(long long)(n0 + r) * b_rs + (long long)(k0 + k) * b_cs] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = sa[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = sb[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) {
        float v = alpha * acc[i][j];
        if (bias) v += bias[n];
        float* c = C + (long long)m * c_rs + (long long)n * c_cs;
        *c = beta != 0.f ? v + beta * *c : v;
      }
    }
}

// P[b, h, q, k] = softmax_k(S[b, h, q, k] * scale) over k <= q, zeros beyond (one CTA per row, in place)
__global__ void __launch_bounds__(256)
softmax_causal_f32_kernel(float* __restrict__ S, int T, float scale) {
  __shared__ float red[33];
  const long long row = blockIdx.x;
  const int q = (int)(row % T);
  float* s = S + row * T;
  float m = -INFINITY;
  for (int k = threadIdx.x; k <= q; k += blockDim.x) m = fmaxf(m, s[k] * scale);
  m = block_max(m, red);
  float sum = 0.f;
  for (int k = threadIdx.x; k <= q; k += blockDim.x) sum += expf(s[k] * scale - m);
  sum = block_sum(sum, red);
  const float inv = 1.f / sum;
  for (int k = threadIdx.x; k < T; k += blockDim.x) s[k] = k <= q ? expf(s[k] * scale - m) * inv : 0.f;
}

// dS = P * (dP - sum_k dP * P) * scale, in place over dP (one CTA per row)
__global__ void __launch_bounds__(256)
softmax_bwd_f32_kernel(const float* __restrict__ P, float* __restrict__ dP, int T, float scale) {
  __shared__ float red[33];
  const long long row = blockIdx.x;
  const float* p = P + row * T;
  float* d = dP + row * T;
  float dot = 0.f;
  for (int k = threadIdx.x; k < T; k += blockDim.x) dot += p[k] * d[k];
  dot = block_sum(dot, red);
  for (int k = threadIdx.x; k < T; k += blockDim.x) d[k] = p[k] * (d[k] - dot) * scale;
}

// LayerNorm (nn.LayerNorm: biased variance, eps inside the sqrt); one CTA per row
__global__ void __launch_bounds__(256)
layernorm_fwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                         float* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int d, float eps) {
  __shared__ float red[33];
  const long long row = blockIdx.x;
  const float* xr = x + row * d;
  float s = 0.f;
  for (int c = threadIdx.x; c < d; c += blockDim.x) s += xr[c];
  const float mu = block_sum(s, red) / d;
  float ss = 0.f;
  for (int c = threadIdx.x; c < d; c += blockDim.x) ss += (xr[c] - mu) * (xr[c] - mu);
  const float r = rsqrtf(block_sum(ss, red) / d + eps);
  if (threadIdx.x == 0) {
    mean[row] = mu;
    rstd[row] = r;
  }
  for (int c = threadIdx.x; c < d; c += blockDim.x) y[row * d + c] = (xr[c] - mu) * r * w[c] + b[c];
}

__global__ void __launch_bounds__(256)
layernorm_bwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ dy,
                         const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dx,
                         float* __restrict__ dw, float* __restrict__ db, int d) {
  __shared__ float red[33];
  const long long row = blockIdx.x;
  const float mu = mean[row], r = rstd[row];
  const float* xr = x + row * d;
  const float* gr = dy + row * d;
  float s1 = 0.f, s2 = 0.f;
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const float g = gr[c] * w[c];
    s1 += g;
    s2 += g * (xr[c] - mu) * r;
  }
  s1 = block_sum(s1, red) / d;
  s2 = block_sum(s2, red) / d;
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const float xn = (xr[c] - mu) * r;
    dx[row * d + c] = r * (gr[c] * w[c] - s1 - xn * s2);
    if (dw) {
      atomicAdd(&dw[c], gr[c] * xn);
      atomicAdd(&db[c], gr[c]);
    }
  }
}

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_g(float x) {
  return 0.5f * (1.f + erff(x * 0.70710678118654752440f)) + x * 0.39894228040143267794f * expf(-0.5f * x * x);
}

// op: 0 gelu fwd (y = gelu(a)), 1 gelu bwd (y = b * gelu'(a)), 2 add (y = a + b), 3 swiglu fwd (y = silu(a) * b),
//     4 y = a * b[0] (b: one scalar in device memory)
__global__ void __launch_bounds__(256)
elementwise_f32_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ y, long long n, int op) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float av = a[i];
    float v;
    if (op == 0) v = gelu_f(av);
    else if (op == 1) v = b[i] * gelu_g(av);
    else if (op == 2) v = av + b[i];
    else if (op == 3) v = av / (1.f + expf(-av)) * b[i];
    else v = av * b[0];
    y[i] = v;
  }
}

// SwiGLU backward (common_components.py:87,123 under autograd): da = dh * b * s * (1 + a (1 - s)), db = dh * a * s
__global__ void __launch_bounds__(256)
swiglu_bwd_f32_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ dh,
                      float* __restrict__ da, float* __restrict__ db, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float av = a[i], s = 1.f / (1.f + expf(-av));
    da[i] = dh[i] * b[i] * s * (1.f + av * (1.f - s));
    db[i] = dh[i] * av * s;
  }
}

// rotate-half RoPE on [B*T, nheads*hd] (row = b*T + t), fp32 tables [>= T, hd]; inverse = the transposed map
__global__ void __launch_bounds__(256)
rope_f32_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ cosT,
                const float* __restrict__ sinT, long long rows, int T, int nheads, int hd, int inverse) {
  const int half = hd >> 1;
  const long long total = rows * nheads * half;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % half);
    const long long rh = i / half;
    const int head = (int)(rh % nheads);
    const long long row = rh / nheads;
    const int t = (int)(row % T);
    const long long base = row * nheads * hd + (long long)head * hd;
    const float x1 = x[base + j], x2 = x[base + half + j];
    const float c1 = cosT[(long long)t * hd + j], c2 = cosT[(long long)t * hd + half + j];
    const float s1 = sinT[(long long)t * hd + j], s2 = sinT[(long long)t * hd + half + j];
    if (!inverse) {
      y[base + j] = x1 * c1 - x2 * s1;
      y[base + half + j] = x2 * c2 + x1 * s2;
    } else {
      y[base + j] = x1 * c1 + x2 * s2;
      y[base + half + j] = x2 * c2 - x1 * s1;
    }
  }
}

// cross-entropy over fp32 logits, one CTA per row: row_loss[m] = lse - logit[target] (0 for ignore_index),
// dlogits = (softmax - onehot) * inv_valid in place when write_grad; inv_valid is read from device memory
__global__ void __launch_bounds__(256)
ce_f32_kernel(float* __restrict__ logits, long long ld, const long long* __restrict__ targets,
              float* __restrict__ row_loss, const float* __restrict__ inv_valid, int V, int write_grad) {
  __shared__ float red[33];
  const long long row = blockIdx.x;
  float* l = logits + row * ld;
  const long long tgt = targets[row];
  if (tgt == -100) {
    if (threadIdx.x == 0) row_loss[row] = 0.f;
    if (write_grad)
      for (int c = threadIdx.x; c < V; c += blockDim.x) l[c] = 0.f;
    return;
  }
  float m = -INFINITY;
  for (int c = threadIdx.x; c < V; c += blockDim.x) m = fmaxf(m, l[c]);
  m = block_max(m, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += expf(l[c] - m);
  s = block_sum(s, red);
  const float lse = m + logf(s);
  if (threadIdx.x == 0) row_loss[row] = (tgt >= 0 && tgt < V) ? lse - l[tgt] : __int_as_float(0x7fc00000);
  __syncthreads();
  if (write_grad) {
    const float inv = inv_valid[0];
    for (int c = threadIdx.x; c < V; c += blockDim.x) l[c] = (expf(l[c] - lse) - (c == tgt ? 1.f : 0.f)) * inv;
  }
}

// stats[0] = 1 / #valid targets (inf when none), then loss = sum(row_loss) * stats[0] (NaN when none, like the reference)
__global__ void __launch_bounds__(256)
ce_count_f32_kernel(const long long* __restrict__ targets, long long M, float* __restrict__ stats) {
  __shared__ float red[33];
  float c = 0.f;
  for (long long i = threadIdx.x; i < M; i += blockDim.x) c += targets[i] != -100 ? 1.f : 0.f;
  c = block_sum(c, red);
  if (threadIdx.x == 0) stats[0] = 1.f / c;
}
__global__ void __launch_bounds__(256)
ce_finish_f32_kernel(const float* __restrict__ row_loss, long long M, const float* __restrict__ stats, float* __restrict__ loss) {
  __shared__ float red[33];
  float a = 0.f;
  for (long long i = threadIdx.x; i < M; i += blockDim.x) a += row_loss[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) loss[0] = a * stats[0];
}

// out[m, :] = tok[idx[m], :] (+ pos[m % T, :]);  backward: scatter-add
__global__ void __launch_bounds__(256)
embedding_fwd_f32_kernel(const long long* __restrict__ idx, const float* __restrict__ tok, const float* __restrict__ pos,
                         float* __restrict__ out, long long M, int T, int d) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < M * d; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / d;
    const int c = (int)(i % d);
    float v = tok[idx[m] * d + c];
    if (pos) v += pos[(m % T) * d + c];
    out[i] = v;
  }
}
__global__ void __launch_bounds__(256)
embedding_bwd_f32_kernel(const long long* __restrict__ idx, const float* __restrict__ dout, float* __restrict__ dtok,
                         float* __restrict__ dpos, long long M, int T, int d) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < M * d; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / d;
    const int c = (int)(i % d);
    atomicAdd(&dtok[idx[m] * d + c], dout[i]);
    if (dpos) atomicAdd(&dpos[(m % T) * d + c], dout[i]);
  }
}

inline int grid1d(long long n) {
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)b200_num_sms() * 16;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

}  // namespace

extern "C" int b200_gemm_f32(const float* A, int64_t a_bs, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_bs,
                             int64_t b_rs, int64_t b_cs, const float* bias, float* C, int64_t c_bs, int64_t c_rs,
                             int64_t c_cs, int64_t batch, int64_t M, int64_t N, int64_t K, float alpha, float beta,
                             void* stream) {
  B200_CHECK_ARG(A && B && C && batch > 0 && M > 0 && N > 0 && K > 0, "gemm_f32: null pointer or empty");
  B200_CHECK_ARG(batch <= 65535 && (M + TM - 1) / TM <= 65535, "gemm_f32: batch / M exceed grid limits");
  dim3 grid((unsigned)((N + TN - 1) / TN), (unsigned)((M + TM - 1) / TM), (unsigned)batch);
  gemm_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, a_bs, a_rs, a_cs, B, b_bs, b_rs, b_cs, bias, C, c_bs, c_rs, c_cs,
                                                          (int)M, (int)N, (int)K, alpha, beta);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_softmax_causal_f32(float* S, int64_t rows, int64_t T, float scale, void* stream) {
  B200_CHECK_ARG(S && rows > 0 && T > 0 && rows < (1ll << 31), "softmax_causal_f32: bad arguments");
  softmax_causal_f32_kernel<<<(unsigned)rows, 256, 0, (cudaStream_t)stream>>>(S, (int)T, scale);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_softmax_bwd_f32(const float* P, float* dP, int64_t rows, int64_t T, float scale, void* stream) {
  B200_CHECK_ARG(P && dP && rows > 0 && T > 0 && rows < (1ll << 31), "softmax_bwd_f32: bad arguments");
  softmax_bwd_f32_kernel<<<(unsigned)rows, 256, 0, (cudaStream_t)stream>>>(P, dP, (int)T, scale);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_layernorm_fwd_f32(const float* x, const float* w, const float* b, float* y, float* mean, float* rstd,
                                      int64_t M, int64_t d, float eps, void* stream) {
  B200_CHECK_ARG(x && w && b && y && mean && rstd && M > 0 && d > 0 && M < (1ll << 31), "layernorm_fwd_f32: bad arguments");
  layernorm_fwd_f32_kernel<<<(unsigned)M, 256, 0, (cudaStream_t)stream>>>(x, w, b, y, mean, rstd, (int)d, eps);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_layernorm_bwd_f32(const float* x, const float* w, const float* dy, const float* mean, const float* rstd,
                                      float* dx, float* dw, float* db, int64_t M, int64_t d, void* stream) {
  B200_CHECK_ARG(x && w && dy && mean && rstd && dx && M > 0 && d > 0 && M < (1ll << 31), "layernorm_bwd_f32: bad arguments");
  B200_CHECK_ARG((dw == nullptr) == (db == nullptr), "layernorm_bwd_f32: dw and db come together");
  layernorm_bwd_f32_kernel<<<(unsigned)M, 256, 0, (cudaStream_t)stream>>>(x, w, dy, mean, rstd, dx, dw, db, (int)d);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_elementwise_f32(const float* a, const float* b, float* y, int64_t n, int op, void* stream) {
  B200_CHECK_ARG(a && y && n > 0 && op >= 0 && op <= 4 && (op == 0 || b), "elementwise_f32: bad arguments");
  elementwise_f32_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(a, b, y, n, op);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_swiglu_bwd_f32(const float* a, const float* b, const float* dh, float* da, float* db, int64_t n,
                                   void* stream) {
  B200_CHECK_ARG(a && b && dh && da && db && n > 0, "swiglu_bwd_f32: bad arguments");
  swiglu_bwd_f32_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(a, b, dh, da, db, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rope_f32(const float* x, float* y, const float* cos, const float* sin, int64_t rows, int64_t T,
                             int64_t nheads, int64_t hd, int inverse, void* stream) {
  B200_CHECK_ARG(x && y && cos && sin && rows > 0 && T > 0 && nheads > 0 && hd > 0 && hd % 2 == 0, "rope_f32: bad arguments");
  rope_f32_kernel<<<grid1d(rows * nheads * (hd / 2)), 256, 0, (cudaStream_t)stream>>>(x, y, cos, sin, rows, (int)T, (int)nheads,
                                                                                      (int)hd, inverse);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" size_t b200_cross_entropy_f32_workspace_bytes(int64_t M) { return (size_t)(M + 1) * sizeof(float); }

extern "C" int b200_cross_entropy_f32(float* logits, int64_t ld, const int64_t* targets, float* loss_out, int64_t M, int64_t V,
                                      int write_grad, void* workspace, size_t workspace_bytes, void* stream) {
  B200_CHECK_ARG(logits && targets && loss_out && workspace && M > 0 && V > 0 && ld >= V && M < (1ll << 31), "cross_entropy_f32: bad arguments");
  B200_CHECK_ARG(workspace_bytes >= b200_cross_entropy_f32_workspace_bytes(M), "cross_entropy_f32: workspace too small");
  float* ws = (float*)workspace;       // [0] 1 / valid, [1 .. M] per-row losses
  cudaStream_t st = (cudaStream_t)stream;
  ce_count_f32_kernel<<<1, 256, 0, st>>>((const long long*)targets, M, ws);
  B200_CHECK_LAUNCH();
  ce_f32_kernel<<<(unsigned)M, 256, 0, st>>>(logits, ld, (const long long*)targets, ws + 1, ws, (int)V, write_grad);
  B200_CHECK_LAUNCH();
  ce_finish_f32_kernel<<<1, 256, 0, st>>>(ws + 1, M, ws, loss_out);
  B200_CHECK_LAUNCH();
  b200_count_launch(3);
  return 0;
}

extern "C" int b200_embedding_fwd_f32(const int64_t* idx, const float* tok_w, const float* pos_w, float* out, int64_t M,
                                      int64_t T, int64_t d, void* stream) {
  B200_CHECK_ARG(idx && tok_w && out && M > 0 && T > 0 && d > 0, "embedding_fwd_f32: bad arguments");
  embedding_fwd_f32_kernel<<<grid1d(M * d), 256, 0, (cudaStream_t)stream>>>((const long long*)idx, tok_w, pos_w, out, M, (int)T, (int)d);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_embedding_bwd_f32(const int64_t* idx, const float* dout, float* dtok_w, float* dpos_w, int64_t M,
                                      int64_t T, int64_t d, void* stream) {
  B200_CHECK_ARG(idx && dout && dtok_w && M > 0 && T > 0 && d > 0, "embedding_bwd_f32: bad arguments");
  embedding_bwd_f32_kernel<<<grid1d(M * d), 256, 0, (cudaStream_t)stream>>>((const long long*)idx, dout, dtok_w, dpos_w, M, (int)T, (int)d);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
