// K3 (materialised-logits form), K8a, K8b — cross-entropy fwd+bwd, gradient norm, AdamW.
// All HBM-bound. Algorithmic bytes: CE 6 B per logit for a non-ignored row (read twice from
// L2/HBM, write grad) and 2 B for an ignored row (write zeros); grad-norm 2 B/param;
// AdamW 14 B/param (bf16) or 28 B/param (fp32).
#include <mutex>

#include "common.cuh"

namespace {

constexpr int CE_THREADS = 512;

// ws layout (floats): [0] valid count, [1] 1/valid, [2..2+M) per-row loss
__global__ void __launch_bounds__(256)
ce_count_kernel(const long long* __restrict__ targets, long long M, float* __restrict__ ws) {
  __shared__ float red[33];
  float c = 0.f;
  for (long long i = threadIdx.x; i < M; i += blockDim.x) c += (targets[i] != -100) ? 1.f : 0.f;
  c = block_sum(c, red);
  if (threadIdx.x == 0) {
    ws[0] = c;
    ws[1] = 1.f / c;  // inf when everything is ignored -> loss = 0 * inf = NaN like the reference
  }
}

// Rows in flight per SM (one CTA each). The second pass over a row only hits L2 if the rows in flight fit there:
// V = 128256 is 256 KB per row, 4 per SM = 151 MB > the 126 MB L2 (ncu: 2.9 passes' worth of DRAM traffic), 2 per SM = 76 MB.
// Two 512-thread CTAs per SM only stream at full rate with CE_UNROLL loads in flight per thread (555 -> 466 us per 3584 rows;
// without the batching 2 rows per SM took 671 us).
constexpr int CE_UNROLL = 4;
static int ce_rows_per_sm() {
  static const int v = [] { const char* e = getenv("B200_CE_ROWS_PER_SM"); const int n = e ? atoi(e) : 0; return (n >= 1 && n <= 4) ? n : 2; }();
  return v;
}

// One CTA per row. Pass 1: online max / sum-exp over the row; pass 2 (row now L2 resident):
// grad = (softmax - onehot) / n_valid written in place. Ignored rows just get zeros.
__global__ void __launch_bounds__(CE_THREADS)
ce_row_kernel(bf16* __restrict__ logits, long long ld, const long long* __restrict__ targets,
              float* __restrict__ ws, long long M, int V, int write_grad) {
  __shared__ float red[33];
  const int nvec = (V + 7) >> 3;
  for (long long row = blockIdx.x; row < M; row += gridDim.x) {
    bf16* lrow = logits + row * ld;
    const long long tgt = targets[row];
    if (tgt == -100) {
      if (threadIdx.x == 0) ws[2 + row] = 0.f;
      if (write_grad) {
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int i = threadIdx.x; i < nvec; i += CE_THREADS) {
          if (i * 8 + 8 <= V) st_v4(lrow + i * 8, z);
          else
            for (int j = i * 8; j < V; ++j) lrow[j] = __float2bfloat16_rn(0.f);
        }
      }
      continue;
    }
    float m = -INFINITY, s = 0.f;
    // CE_UNROLL vectors per thread are requested before the first is used: with few rows in flight per SM (so that the
    // second pass finds the row in L2) the loads of one thread must overlap each other
    for (int i0 = threadIdx.x; i0 < nvec; i0 += CE_THREADS * CE_UNROLL) {
      uint4 raw[CE_UNROLL];
#pragma unroll
      for (int u = 0; u < CE_UNROLL; ++u) {
        const int i = i0 + u * CE_THREADS;
        raw[u] = i < nvec ? ld_v4(lrow + i * 8) : make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);   // -inf
      }
#pragma unroll
      for (int u = 0; u < CE_UNROLL; ++u) {
        const int i = i0 + u * CE_THREADS;
        float f[8];
        unpack8(raw[u], f);
        if (i < nvec && i * 8 + 8 > V) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (i * 8 + j >= V) f[j] = -INFINITY;
        }
        float vm = f[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) vm = fmaxf(vm, f[j]);
        if (vm > m) {
          s *= __expf(m - vm);
          m = vm;
        }
        if (vm > -INFINITY) {
#pragma unroll
          for (int j = 0; j < 8; ++j) s += __expf(f[j] - m);
        }
      }
    }
    const float gm = block_max(m, red);
    s *= (m == -INFINITY) ? 0.f : __expf(m - gm);
    const float gs = block_sum(s, red);
    const float lse = gm + logf(gs);
    if (threadIdx.x == 0) {
      // a target outside [0, V) is refused: NaN poisons the mean loss (F.cross_entropy asserts on the device)
      const float lt = (tgt >= 0 && tgt < V) ? __bfloat162float(lrow[tgt]) : __int_as_float(0x7fc00000);
      ws[2 + row] = lse - lt;
    }
    if (write_grad) {
      const float inv = ws[1];
      __syncthreads();  // lrow[tgt] was read above before anyone overwrites it
      for (int i0 = threadIdx.x; i0 < nvec; i0 += CE_THREADS * CE_UNROLL) {
        uint4 raw[CE_UNROLL];
#pragma unroll
        for (int u = 0; u < CE_UNROLL; ++u) {
          const int i = i0 + u * CE_THREADS;
          if (i < nvec) raw[u] = ld_v4(lrow + i * 8);
        }
#pragma unroll
        for (int u = 0; u < CE_UNROLL; ++u) {
          const int i = i0 + u * CE_THREADS;
          if (i >= nvec) break;
          float f[8];
          unpack8(raw[u], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float p = __expf(f[j] - lse);
            if ((long long)(i * 8 + j) == tgt) p -= 1.f;
            f[j] = p * inv;
          }
          if (i * 8 + 8 <= V) st_v4(lrow + i * 8, pack8(f));
          else
            for (int j = i * 8; j < V; ++j) lrow[j] = __float2bfloat16_rn(f[j - i * 8]);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(1024)
ce_finish_kernel(const float* __restrict__ ws, long long M, float* __restrict__ loss) {
  __shared__ float red[33];
  float acc = 0.f;
  for (long long i = threadIdx.x; i < M; i += blockDim.x) acc += ws[2 + i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) loss[0] = acc * ws[1];
}

// ---------------- grad norm ----------------
constexpr int GN_MAX_BLOCKS = 1024;

__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const bf16* __restrict__ g, long long n, float* __restrict__ partial) {
  __shared__ float red[33];
  const long long nv = n >> 3;
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_nc_v4(g + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) {
      const float f = __bfloat162float(g[k]);
      acc += f * f;
    }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(256)
sumsq_partial_f32_kernel(const float* __restrict__ g, long long n, float* __restrict__ partial) {
  __shared__ float red[33];
  const long long nv = n >> 2;
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nv; i0 += 4 * stride) {
    uint4 u[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) u[k] = (i0 + k * stride < nv) ? ld_nc_v4(g + (i0 + k * stride) * 4) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float a = __uint_as_float(u[k].x), b = __uint_as_float(u[k].y), c = __uint_as_float(u[k].z),
                  d = __uint_as_float(u[k].w);
      acc += a * a + b * b + c * c + d * d;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 4; k < n; ++k) acc += g[k] * g[k];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(1024)
norm_finish_kernel(const float* __restrict__ partial, int nblocks, float max_norm,
                   float* __restrict__ out) {
  __shared__ float red[33];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) acc += partial[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    const float nrm = sqrtf(acc);
    out[0] = nrm;
    // torch.nn.utils.clip_grad_norm_: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1
    out[1] = fminf(1.f, max_norm / (nrm + 1e-6f));
  }
}

// ---------------- AdamW ----------------
// torch.optim.AdamW single step (decoupled weight decay, no amsgrad):
//   p *= 1 - lr*wd;  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;
//   p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
struct AdamConsts {
  float lr, b1, b2, eps, wd, step_size, inv_sqrt_bc2;
};

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v,
                                            const AdamConsts& c) {
  p *= 1.f - c.lr * c.wd;
  m = c.b1 * m + (1.f - c.b1) * g;
  v = c.b2 * v + (1.f - c.b2) * g * g;
  const float denom = sqrtf(v) * c.inv_sqrt_bc2 + c.eps;
  p -= c.step_size * (m / denom);
}

// The same update with the MUFU square root and reciprocal (2 ulp each) instead of the IEEE sequences (~30 instructions
// per element together): for parameters and moments that are rounded to bf16 right after, the result is the same bf16
// value except at rounding ties, and the kernel stops spending half of its issue slots on two Newton iterations.
__device__ __forceinline__ void adam_update_fast(float& p, float g, float& m, float& v, const AdamConsts& c) {
  p *= 1.f - c.lr * c.wd;
  m = c.b1 * m + (1.f - c.b1) * g;
  v = c.b2 * v + (1.f - c.b2) * g * g;
  float sq;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
  const float denom = sq * c.inv_sqrt_bc2 + c.eps;
  p -= c.step_size * __fdividef(m, denom);
}

// Hyper-parameters may come from device memory (7 floats: lr, beta1, beta2, eps, weight_decay,
// step_size = lr / bias_correction1, 1/sqrt(bias_correction2)) so that a captured CUDA graph
// can be replayed with a new learning rate / step count.
__device__ __forceinline__ AdamConsts load_consts(const AdamConsts& c, const float* __restrict__ dev) {
  if (dev == nullptr) return c;
  AdamConsts r;
  r.lr = dev[0]; r.b1 = dev[1]; r.b2 = dev[2]; r.eps = dev[3]; r.wd = dev[4];
  r.step_size = dev[5]; r.inv_sqrt_bc2 = dev[6];
  return r;
}

template <bool G32>
__global__ void __launch_bounds__(256)
adamw_bf16_kernel(bf16* __restrict__ p, const void* __restrict__ gv, bf16* __restrict__ m,
                  bf16* __restrict__ v, long long n, AdamConsts c0, const float* __restrict__ dev_hyper,
                  const float* __restrict__ grad_scale) {
  const AdamConsts c = load_consts(c0, dev_hyper);
  const float gs = grad_scale ? grad_scale[0] : 1.f;
  const long long nv = n >> 3;
  // two vectors per thread and pass: ten 16-byte loads in flight before the first is used
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nv; i0 += 2 * stride) {
    uint4 rp[2], rg0[2], rg1[2], rm[2], rv[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i < nv) {
        rp[u] = ld_v4(p + i * 8);
        if (G32) {
          const float* g = (const float*)gv;
          rg0[u] = ld_nc_v4(g + i * 8);
          rg1[u] = ld_nc_v4(g + i * 8 + 4);
        } else {
          rg0[u] = ld_nc_v4((const bf16*)gv + i * 8);
        }
        rm[u] = ld_v4(m + i * 8);
        rv[u] = ld_v4(v + i * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i >= nv) break;
      float fp[8], fg[8], fm[8], fv[8];
      unpack8(rp[u], fp);
      if (G32) {
        fg[0] = __uint_as_float(rg0[u].x); fg[1] = __uint_as_float(rg0[u].y); fg[2] = __uint_as_float(rg0[u].z); fg[3] = __uint_as_float(rg0[u].w);
        fg[4] = __uint_as_float(rg1[u].x); fg[5] = __uint_as_float(rg1[u].y); fg[6] = __uint_as_float(rg1[u].z); fg[7] = __uint_as_float(rg1[u].w);
      } else {
        unpack8(rg0[u], fg);
      }
      unpack8(rm[u], fm);
      unpack8(rv[u], fv);
#pragma unroll
      for (int j = 0; j < 8; ++j) adam_update_fast(fp[j], bf16_round(fg[j] * gs), fm[j], fv[j], c);
      st_v4(p + i * 8, pack8(fp));
      st_v4(m + i * 8, pack8(fm));
      st_v4(v + i * 8, pack8(fv));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) {
      float fp = __bfloat162float(p[k]), fm = __bfloat162float(m[k]), fv = __bfloat162float(v[k]);
      const float g = G32 ? ((const float*)gv)[k] : __bfloat162float(((const bf16*)gv)[k]);
      adam_update(fp, bf16_round(g * gs), fm, fv, c);
      p[k] = __float2bfloat16_rn(fp);
      m[k] = __float2bfloat16_rn(fm);
      v[k] = __float2bfloat16_rn(fv);
    }
}

__global__ void __launch_bounds__(256)
adamw_f32_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, long long n, AdamConsts c0, const float* __restrict__ dev_hyper,
                 const float* __restrict__ grad_scale) {
  const AdamConsts c = load_consts(c0, dev_hyper);
  const float gs = grad_scale ? grad_scale[0] : 1.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float fp = p[i], fm = m[i], fv = v[i];
    adam_update(fp, g[i] * gs, fm, fv, c);
    p[i] = fp;
    m[i] = fm;
    v[i] = fv;
  }
}

// fp32 master weights (the "bf16_hybrid" policy of datautils/mixed_precision.py:24-28: parameters and optimizer
// state in fp32, compute copy in bf16): one pass updates the master and rewrites the bf16 copy the GEMMs read.
// 28 B/param of fp32 state traffic + 2 B/param for the copy.
__global__ void __launch_bounds__(256)
adamw_master_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                    float* __restrict__ v, bf16* __restrict__ p_lo, long long n, AdamConsts c0,
                    const float* __restrict__ dev_hyper, const float* __restrict__ grad_scale) {
  const AdamConsts c = load_consts(c0, dev_hyper);
  const float gs = grad_scale ? grad_scale[0] : 1.f;
  const long long nv = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float4 fp = *reinterpret_cast<const float4*>(p + i * 4);
    const uint4 ug = ld_nc_v4(g + i * 4);
    float4 fm = *reinterpret_cast<const float4*>(m + i * 4);
    float4 fv = *reinterpret_cast<const float4*>(v + i * 4);
    adam_update(fp.x, __uint_as_float(ug.x) * gs, fm.x, fv.x, c);
    adam_update(fp.y, __uint_as_float(ug.y) * gs, fm.y, fv.y, c);
    adam_update(fp.z, __uint_as_float(ug.z) * gs, fm.z, fv.z, c);
    adam_update(fp.w, __uint_as_float(ug.w) * gs, fm.w, fv.w, c);
    *reinterpret_cast<float4*>(p + i * 4) = fp;
    *reinterpret_cast<float4*>(m + i * 4) = fm;
    *reinterpret_cast<float4*>(v + i * 4) = fv;
    *reinterpret_cast<uint2*>(p_lo + i * 4) = make_uint2(f2_to_bf2(fp.x, fp.y), f2_to_bf2(fp.z, fp.w));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 4; k < n; ++k) {
      float fp = p[k], fm = m[k], fv = v[k];
      adam_update(fp, g[k] * gs, fm, fv, c);
      p[k] = fp; m[k] = fm; v[k] = fv;
      p_lo[k] = __float2bfloat16_rn(fp);
    }
}

AdamConsts make_consts(float lr, float b1, float b2, float eps, float wd, long long step) {
  AdamConsts c;
  c.lr = lr; c.b1 = b1; c.b2 = b2; c.eps = eps; c.wd = wd;
  const double bc1 = 1.0 - pow((double)b1, (double)step);
  const double bc2 = 1.0 - pow((double)b2, (double)step);
  c.step_size = (float)((double)lr / bc1);
  c.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  return c;
}

inline int ew_grid(long long nvec, int threads) {
  long long blocks = (nvec + threads - 1) / threads;
  const long long cap = (long long)b200_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

extern "C" size_t b200_cross_entropy_workspace_bytes(int64_t M) {
  return (size_t)(M + 2) * sizeof(float);
}

extern "C" int b200_cross_entropy_bf16(void* logits, int64_t ld, const int64_t* targets,
                                       float* loss_out, int64_t M, int64_t V, int write_grad,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  B200_CHECK_ARG(logits && targets && loss_out && workspace && M > 0 && V > 0, "cross_entropy: null pointer or empty");
  B200_CHECK_ARG(ld % 8 == 0 && ld >= V, "cross_entropy: ld=%lld must be a multiple of 8 and >= V", (long long)ld);
  B200_CHECK_ARG(b200_aligned16(logits), "cross_entropy: logits must be 16-B aligned");
  B200_CHECK_ARG(workspace_bytes >= b200_cross_entropy_workspace_bytes(M), "cross_entropy: workspace too small");
  B200_CHECK_ARG(V < (1ll << 31), "cross_entropy: V too large");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = (float*)workspace;
  ce_count_kernel<<<1, 256, 0, st>>>((const long long*)targets, M, ws);
  B200_CHECK_LAUNCH();
  const long long cap = (long long)b200_num_sms() * ce_rows_per_sm();
  ce_row_kernel<<<(int)(M < cap ? M : cap), CE_THREADS, 0, st>>>((bf16*)logits, ld, (const long long*)targets, ws, M, (int)V, write_grad);
  B200_CHECK_LAUNCH();
  ce_finish_kernel<<<1, 1024, 0, st>>>(ws, M, loss_out);
  B200_CHECK_LAUNCH();
  b200_count_launch(3);
  return 0;
}

extern "C" size_t b200_grad_norm_workspace_bytes(int64_t n) {
  (void)n;
  return (size_t)GN_MAX_BLOCKS * sizeof(float);
}

static int grad_norm_impl(const void* g, int g_is_f32, int64_t n, float max_norm, float* out,
                          void* workspace, size_t workspace_bytes, void* stream) {
  B200_CHECK_ARG(g && out && workspace && n > 0, "grad_norm: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(g), "grad_norm: gradient buffer must be 16-B aligned");
  B200_CHECK_ARG(workspace_bytes >= b200_grad_norm_workspace_bytes(n), "grad_norm: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = ew_grid(n / 8 + 1, 256);
  if (blocks > GN_MAX_BLOCKS) blocks = GN_MAX_BLOCKS;
  if (g_is_f32) sumsq_partial_f32_kernel<<<blocks, 256, 0, st>>>((const float*)g, n, (float*)workspace);
  else sumsq_partial_kernel<<<blocks, 256, 0, st>>>((const bf16*)g, n, (float*)workspace);
  B200_CHECK_LAUNCH();
  norm_finish_kernel<<<1, 1024, 0, st>>>((const float*)workspace, blocks, max_norm, out);
  B200_CHECK_LAUNCH();
  b200_count_launch(2);
  return 0;
}

extern "C" int b200_grad_norm_bf16(const void* g, int64_t n, float max_norm, float* out,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  return grad_norm_impl(g, 0, n, max_norm, out, workspace, workspace_bytes, stream);
}

extern "C" int b200_grad_norm_f32(const float* g, int64_t n, float max_norm, float* out,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  return grad_norm_impl(g, 1, n, max_norm, out, workspace, workspace_bytes, stream);
}

extern "C" int b200_adam_hyper(float* host7, float lr, float beta1, float beta2, float eps,
                               float weight_decay, int64_t step) {
  B200_CHECK_ARG(host7 && step >= 1, "adam_hyper: null pointer or step < 1");
  const AdamConsts c = make_consts(lr, beta1, beta2, eps, weight_decay, step);
  host7[0] = c.lr; host7[1] = c.b1; host7[2] = c.b2; host7[3] = c.eps; host7[4] = c.wd;
  host7[5] = c.step_size; host7[6] = c.inv_sqrt_bc2;
  return 0;
}

extern "C" int b200_adamw_bf16(void* p, const void* g, int g_is_f32, void* m, void* v, int64_t n,
                               float lr, float beta1, float beta2, float eps, float weight_decay,
                               int64_t step, const float* dev_hyper, const float* grad_scale,
                               void* stream) {
  B200_CHECK_ARG(p && g && m && v && n > 0 && (step >= 1 || dev_hyper), "adamw: null pointer, empty, or step < 1");
  B200_CHECK_ARG(b200_aligned16(p) && b200_aligned16(g) && b200_aligned16(m) && b200_aligned16(v), "adamw: buffers must be 16-B aligned");
  const AdamConsts c = make_consts(lr, beta1, beta2, eps, weight_decay, step >= 1 ? step : 1);
  const int grid = ew_grid(n / 8 + 1, 256);
  if (g_is_f32)
    adamw_bf16_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>((bf16*)p, g, (bf16*)m, (bf16*)v, n, c, dev_hyper, grad_scale);
  else
    adamw_bf16_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>((bf16*)p, g, (bf16*)m, (bf16*)v, n, c, dev_hyper, grad_scale);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_adamw_f32(float* p, const float* g, float* m, float* v, int64_t n, float lr,
                              float beta1, float beta2, float eps, float weight_decay, int64_t step,
                              const float* dev_hyper, const float* grad_scale, void* stream) {
  B200_CHECK_ARG(p && g && m && v && n > 0 && (step >= 1 || dev_hyper), "adamw: null pointer, empty, or step < 1");
  adamw_f32_kernel<<<ew_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(
      p, g, m, v, n, make_consts(lr, beta1, beta2, eps, weight_decay, step >= 1 ? step : 1), dev_hyper, grad_scale);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_adamw_master_f32(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n,
                                     float lr, float beta1, float beta2, float eps, float weight_decay,
                                     int64_t step, const float* dev_hyper, const float* grad_scale, void* stream) {
  B200_CHECK_ARG(p && g && m && v && p_bf16 && n > 0 && (step >= 1 || dev_hyper), "adamw_master: null pointer, empty, or step < 1");
  B200_CHECK_ARG(b200_aligned16(p) && b200_aligned16(g) && b200_aligned16(m) && b200_aligned16(v) && (((uintptr_t)p_bf16) & 7u) == 0,
                 "adamw_master: fp32 buffers must be 16-B aligned, the bf16 copy 8-B aligned");
  adamw_master_kernel<<<ew_grid(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      p, g, m, v, (bf16*)p_bf16, n, make_consts(lr, beta1, beta2, eps, weight_decay, step >= 1 ? step : 1), dev_hyper, grad_scale);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

// ======================================================================================
// Allocation-free cross-entropy.
//
// For hosts that may not obtain device memory while they run: a trainer that has carved all of
// HBM into a fixed plan up front (nothing left for a workspace, and cudaMalloc would stall every
// stream or simply fail), or an embedding application that hands this library only its input and
// output buffers and forbids it to ask the driver for more. b200_cross_entropy_bf16 needs
// (M + 2) floats of caller-provided workspace; this variant keeps that workspace inside the
// library image instead, in the file-scope device array below. Every GPU that loads the library
// carries those 256 KiB until the process exits.
// ======================================================================================
static __device__ __align__(256) float g_ce_noalloc_scratch[B200_CE_NOALLOC_MAX_ROWS + 2];

// In host code the array's name is only a handle; the device address must be asked for, once
// per device, while that device is current.
static float* g_ce_noalloc_ptr[64] = {nullptr};
static std::mutex g_ce_noalloc_mu;

static int ce_noalloc_scratch(float** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < 64, "cross_entropy_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_ce_noalloc_mu);
  if (g_ce_noalloc_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_ce_noalloc_scratch));
    B200_CHECK_ARG(ptr != nullptr, "cross_entropy_noalloc: could not resolve the scratch array");
    g_ce_noalloc_ptr[dev] = ptr;
  }
  *out = g_ce_noalloc_ptr[dev];
  return 0;
}

extern "C" int b200_cross_entropy_bf16_noalloc(void* logits, int64_t ld, const int64_t* targets,
                                               float* loss_out, int64_t M, int64_t V,
                                               int write_grad, void* stream) {
  B200_CHECK_ARG(logits && targets && loss_out && M > 0 && V > 0, "cross_entropy_noalloc: null pointer or empty");
  B200_CHECK_ARG(M <= B200_CE_NOALLOC_MAX_ROWS,
                 "cross_entropy_noalloc: M=%lld rows exceed the built-in scratch (max %d rows); use "
                 "b200_cross_entropy_bf16 with a caller workspace, or split the batch",
                 (long long)M, (int)B200_CE_NOALLOC_MAX_ROWS);
  B200_CHECK_ARG(ld % 8 == 0 && ld >= V, "cross_entropy_noalloc: ld=%lld must be a multiple of 8 and >= V", (long long)ld);
  B200_CHECK_ARG(b200_aligned16(logits), "cross_entropy_noalloc: logits must be 16-B aligned");
  B200_CHECK_ARG(V < (1ll << 31), "cross_entropy_noalloc: V too large");
  float* ws = nullptr;
  int rc = ce_noalloc_scratch(&ws);
  if (rc) return rc;
  // Nothing in the scratch has to start at zero: slot 0 / 1 (valid count, its reciprocal) and one
  // slot per row are each rewritten by this call before anything reads them, and slots beyond M
  // left over from an earlier, longer call are never read.
  cudaStream_t st = (cudaStream_t)stream;
  ce_count_kernel<<<1, 256, 0, st>>>((const long long*)targets, M, ws);
  B200_CHECK_LAUNCH();
  const long long cap = (long long)b200_num_sms() * ce_rows_per_sm();
  ce_row_kernel<<<(int)(M < cap ? M : cap), CE_THREADS, 0, st>>>((bf16*)logits, ld, (const long long*)targets, ws, M, (int)V, write_grad);
  B200_CHECK_LAUNCH();
  ce_finish_kernel<<<1, 1024, 0, st>>>(ws, M, loss_out);
  B200_CHECK_LAUNCH();
  b200_count_launch(3);
  return 0;
}
