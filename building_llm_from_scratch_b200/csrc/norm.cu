// K4 / K4' — RMSNorm and LayerNorm, forward and backward (HBM-bound).
//
// One warp owns one row: the row is read once with 128-bit loads, kept in registers, reduced
// with warp shuffles (no block barrier) and written once. Algorithmic bytes per element:
// fwd 4 B (read x, write y; +4 B with the fused residual add), bwd 6 B (read x, dy; write dx).
#include "common.cuh"

namespace {

constexpr int WARPS = 8;

// ---------------- RMSNorm forward (bf16) ----------------
// Reference: Models/Llama/common_components.py:66-70. In bf16 eager mode every op rounds:
// x.pow(2) -> bf16, mean -> bf16 (fp32 accumulate), + eps -> bf16, rsqrt -> bf16,
// x * r -> bf16, * weight(fp32) -> fp32, .to(bf16).
template <int MAXV>
__global__ void __launch_bounds__(WARPS * 32)
rmsnorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res,
                   const float* __restrict__ w, bf16* __restrict__ sum_out, bf16* __restrict__ y,
                   long long M, int d, float eps, int bf16_steps) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS;
  const int nvec = d >> 3;
  for (long long row = warp0; row < M; row += nwarps) {
    uint4 v[MAXV];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        v[i] = ld_nc_v4(x + row * d + idx * 8);
        float f[8];
        unpack8(v[i], f);
        if (res != nullptr) {
          float g[8];
          unpack8(ld_nc_v4(res + row * d + idx * 8), g);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = bf16_round(f[j] + g[j]);
          v[i] = pack8(f);
          st_v4(sum_out + row * d + idx * 8, v[i]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += bf16_steps ? bf16_round(f[j] * f[j]) : f[j] * f[j];
      }
    }
    ss = warp_sum(ss);
    float mean = ss / (float)d;
    float r;
    if (bf16_steps) {
      mean = bf16_round(mean);
      r = bf16_round(rsqrtf(bf16_round(mean + eps)));
    } else {
      r = rsqrtf(mean + eps);
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        float f[8];
        unpack8(v[i], f);
        const float4 w0 = *reinterpret_cast<const float4*>(w + idx * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(w + idx * 8 + 4);
        const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float xn = f[j] * r;
          if (bf16_steps) xn = bf16_round(xn);
          f[j] = xn * ww[j];
        }
        st_v4(y + row * d + idx * 8, pack8(f));
      }
    }
  }
}

// ---------------- RMSNorm backward (bf16) ----------------
// Two sweeps over the row: the first forms sum(x^2) and sum(dy*w*x) together (the second dot product
// only needs 1/rms as a factor afterwards), the second writes dx. Only x stays in registers; dy is
// re-read (an L1/L2 hit: the same warp touched it a few hundred cycles earlier), which halves the
// register footprint and doubles the resident warps of this bandwidth-bound kernel.
template <int MAXV, bool HAS_DW>
__global__ void __launch_bounds__(WARPS * 32, (MAXV <= 8 && !HAS_DW) ? 4 : 1)
rmsnorm_bwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w,
                   const bf16* __restrict__ dy, const bf16* __restrict__ dres, bf16* __restrict__ dx,
                   float* __restrict__ dw, long long M, int d, float eps) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS;
  const int nvec = d >> 3;
  float dwacc[HAS_DW ? MAXV * 8 : 1];
  if (HAS_DW) {
#pragma unroll
    for (int i = 0; i < (HAS_DW ? MAXV * 8 : 1); ++i) dwacc[i] = 0.f;
  }
  for (long long row = warp0; row < M; row += nwarps) {
    uint4 vx[MAXV];
    float ss = 0.f, dot = 0.f;   // sum x^2, sum dy*w*x
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        vx[i] = ld_nc_v4(x + row * d + idx * 8);
        float f[8], g[8];
        unpack8(vx[i], f);
        unpack8(ld_v4(dy + row * d + idx * 8), g);
        const float4 w0 = *reinterpret_cast<const float4*>(w + idx * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(w + idx * 8 + 4);
        const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          ss += f[j] * f[j];
          dot += g[j] * ww[j] * f[j];
        }
      }
    }
    ss = warp_sum(ss);
    dot = warp_sum(dot);
    const float r = rsqrtf(ss / (float)d + eps);
    const float c = dot * r / (float)d;   // mean(g * xn), xn = x * r
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        float f[8], g[8], o[8];
        unpack8(vx[i], f);
        unpack8(ld_v4(dy + row * d + idx * 8), g);
        const float4 w0 = *reinterpret_cast<const float4*>(w + idx * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(w + idx * 8 + 4);
        const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xn = f[j] * r;
          o[j] = r * (g[j] * ww[j] - xn * c);
          if (HAS_DW) dwacc[i * 8 + j] += g[j] * xn;
        }
        if (dres != nullptr) {
          float e[8];
          unpack8(ld_nc_v4(dres + row * d + idx * 8), e);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += e[j];
        }
        st_v4(dx + row * d + idx * 8, pack8(o));
      }
    }
  }
  if (HAS_DW) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(&dw[idx * 8 + j], dwacc[i * 8 + j]);
      }
    }
  }
}

// ---------------- RMSNorm fp32 (parity mode) ----------------
__global__ void __launch_bounds__(256)
rmsnorm_fwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                       float* __restrict__ y, long long M, int d, float eps) {
  __shared__ float red[33];
  for (long long row = blockIdx.x; row < M; row += gridDim.x) {
    float ss = 0.f;
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
      const float f = x[row * d + c];
      ss += f * f;
    }
    ss = block_sum(ss, red);
    const float r = rsqrtf(ss / (float)d + eps);
    for (int c = threadIdx.x; c < d; c += blockDim.x) y[row * d + c] = x[row * d + c] * r * w[c];
  }
}

__global__ void __launch_bounds__(256)
rmsnorm_bwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                       const float* __restrict__ dy, float* __restrict__ dx, float* __restrict__ dw,
                       long long M, int d, float eps) {
  __shared__ float red[33];
  for (long long row = blockIdx.x; row < M; row += gridDim.x) {
    float ss = 0.f;
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
      const float f = x[row * d + c];
      ss += f * f;
    }
    ss = block_sum(ss, red);
    const float r = rsqrtf(ss / (float)d + eps);
    float dot = 0.f;
    for (int c = threadIdx.x; c < d; c += blockDim.x)
      dot += dy[row * d + c] * w[c] * x[row * d + c] * r;
    dot = block_sum(dot, red);
    const float cc = dot / (float)d;
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
      const float xn = x[row * d + c] * r;
      dx[row * d + c] = r * (dy[row * d + c] * w[c] - xn * cc);
      if (dw) atomicAdd(&dw[c], dy[row * d + c] * xn);
    }
  }
}

// ---------------- LayerNorm (bf16 io, fp32 math) ----------------
template <int MAXV>
__global__ void __launch_bounds__(WARPS * 32)
layernorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                     const bf16* __restrict__ b, bf16* __restrict__ y, float* __restrict__ mean_out,
                     float* __restrict__ rstd_out, long long M, int d, float eps) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS;
  const int nvec = d >> 3;
  for (long long row = warp0; row < M; row += nwarps) {
    uint4 v[MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        v[i] = ld_nc_v4(x + row * d + idx * 8);
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += f[j];
      }
    }
    const float mu = warp_sum(s) / (float)d;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += (f[j] - mu) * (f[j] - mu);
      }
    }
    const float r = rsqrtf(warp_sum(ss) / (float)d + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mu;
      if (rstd_out) rstd_out[row] = r;
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        float f[8], ww[8], bb[8];
        unpack8(v[i], f);
        unpack8(ld_v4(w + idx * 8), ww);
        unpack8(ld_v4(b + idx * 8), bb);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = (f[j] - mu) * r * ww[j] + bb[j];
        st_v4(y + row * d + idx * 8, pack8(f));
      }
    }
  }
}

// dx = r * (g - mean(g) - xn * mean(g * xn)), g = dy * w; dw += dy * xn; db += dy.
// dw / db partials live in registers across the rows a warp owns; one atomic per column per warp.
template <int MAXV, bool HAS_DW>
__global__ void __launch_bounds__(WARPS * 32)
layernorm_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                     const bf16* __restrict__ dy, const float* __restrict__ mean,
                     const float* __restrict__ rstd, bf16* __restrict__ dx, float* __restrict__ dw,
                     float* __restrict__ db, long long M, int d) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS;
  const int nvec = d >> 3;
  constexpr int NACC = HAS_DW ? MAXV * 8 : 1;
  float dwacc[NACC], dbacc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) dwacc[i] = dbacc[i] = 0.f;
  for (long long row = warp0; row < M; row += nwarps) {
    const float mu = mean[row], r = rstd[row];
    uint4 vx[MAXV], vg[MAXV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        vx[i] = ld_nc_v4(x + row * d + idx * 8);
        vg[i] = ld_nc_v4(dy + row * d + idx * 8);
        float f[8], g[8], ww[8];
        unpack8(vx[i], f);
        unpack8(vg[i], g);
        unpack8(ld_v4(w + idx * 8), ww);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gg = g[j] * ww[j];
          s1 += gg;
          s2 += gg * (f[j] - mu) * r;
        }
      }
    }
    s1 = warp_sum(s1) / (float)d;
    s2 = warp_sum(s2) / (float)d;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        float f[8], g[8], ww[8], o[8];
        unpack8(vx[i], f);
        unpack8(vg[i], g);
        unpack8(ld_v4(w + idx * 8), ww);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xn = (f[j] - mu) * r;
          o[j] = r * (g[j] * ww[j] - s1 - xn * s2);
          if (HAS_DW) {
            dwacc[i * 8 + j] += g[j] * xn;
            dbacc[i * 8 + j] += g[j];
          }
        }
        st_v4(dx + row * d + idx * 8, pack8(o));
      }
    }
  }
  if (HAS_DW) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          atomicAdd(&dw[idx * 8 + j], dwacc[i * 8 + j]);
          atomicAdd(&db[idx * 8 + j], dbacc[i * 8 + j]);
        }
      }
    }
  }
}

int grid_for_rows(long long M) {
  const long long blocks = (M + WARPS - 1) / WARPS;
  const long long cap = (long long)b200_num_sms() * 8;
  return (int)(blocks < cap ? blocks : cap);
}

}  // namespace

#define DISPATCH_MAXV(d, CALL)                 \
  do {                                         \
    const int _nv = (int)(((d) / 8 + 31) / 32); \
    if (_nv <= 4) { CALL(4); }                 \
    else if (_nv <= 8) { CALL(8); }            \
    else if (_nv <= 16) { CALL(16); }          \
    else if (_nv <= 32) { CALL(32); }          \
    else { CALL(64); }                         \
  } while (0)

extern "C" int b200_rmsnorm_fwd_bf16(const void* x, const void* res, const float* w, void* sum_out,
                                     void* y, int64_t M, int64_t d, float eps, int bf16_steps,
                                     void* stream) {
  B200_CHECK_ARG(x && w && y && M > 0, "rmsnorm_fwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "rmsnorm_fwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  B200_CHECK_ARG(res == nullptr || sum_out != nullptr, "rmsnorm_fwd: res given without sum_out");
  B200_CHECK_ARG(b200_aligned16(x) && b200_aligned16(y) && b200_aligned16(w), "rmsnorm_fwd: pointers must be 16-B aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for_rows(M);
#define CALL(NV) rmsnorm_fwd_kernel<NV><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)res, w, (bf16*)sum_out, (bf16*)y, M, (int)d, eps, bf16_steps)
  DISPATCH_MAXV(d, CALL);
#undef CALL
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rmsnorm_bwd_bf16(const void* x, const float* w, const void* dy,
                                     const void* dres_in, void* dx, float* dw, int64_t M, int64_t d,
                                     float eps, void* stream) {
  B200_CHECK_ARG(x && w && dy && dx && M > 0, "rmsnorm_bwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "rmsnorm_bwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  B200_CHECK_ARG(dw == nullptr || d <= 4096, "rmsnorm_bwd: weight gradient supported for d <= 4096");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for_rows(M);
  if (dw) {
    const int nv = (int)((d / 8 + 31) / 32);
#define CALLW(NV) rmsnorm_bwd_kernel<NV, true><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, w, (const bf16*)dy, (const bf16*)dres_in, (bf16*)dx, dw, M, (int)d, eps)
    if (nv <= 4) { CALLW(4); } else if (nv <= 8) { CALLW(8); } else { CALLW(16); }
#undef CALLW
  } else {
#define CALL(NV) rmsnorm_bwd_kernel<NV, false><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, w, (const bf16*)dy, (const bf16*)dres_in, (bf16*)dx, nullptr, M, (int)d, eps)
    DISPATCH_MAXV(d, CALL);
#undef CALL
  }
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rmsnorm_fwd_f32(const float* x, const float* w, float* y, int64_t M, int64_t d,
                                    float eps, void* stream) {
  B200_CHECK_ARG(x && w && y && M > 0 && d > 0, "rmsnorm_fwd_f32: bad arguments");
  const long long cap = (long long)b200_num_sms() * 8;
  rmsnorm_fwd_f32_kernel<<<(int)(M < cap ? M : cap), 256, 0, (cudaStream_t)stream>>>(x, w, y, M, (int)d, eps);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rmsnorm_bwd_f32(const float* x, const float* w, const float* dy, float* dx,
                                    float* dw, int64_t M, int64_t d, float eps, void* stream) {
  B200_CHECK_ARG(x && w && dy && dx && M > 0 && d > 0, "rmsnorm_bwd_f32: bad arguments");
  const long long cap = (long long)b200_num_sms() * 8;
  rmsnorm_bwd_f32_kernel<<<(int)(M < cap ? M : cap), 256, 0, (cudaStream_t)stream>>>(x, w, dy, dx, dw, M, (int)d, eps);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_layernorm_fwd_bf16(const void* x, const void* w, const void* b, void* y,
                                       float* mean, float* rstd, int64_t M, int64_t d, float eps,
                                       void* stream) {
  B200_CHECK_ARG(x && w && b && y && M > 0, "layernorm_fwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "layernorm_fwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for_rows(M);
#define CALL(NV) layernorm_fwd_kernel<NV><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)b, (bf16*)y, mean, rstd, M, (int)d, eps)
  DISPATCH_MAXV(d, CALL);
#undef CALL
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_layernorm_bwd_bf16(const void* x, const void* w, const void* dy,
                                       const float* mean, const float* rstd, void* dx, float* dw,
                                       float* db, int64_t M, int64_t d, void* stream) {
  B200_CHECK_ARG(x && w && dy && mean && rstd && dx && M > 0, "layernorm_bwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "layernorm_bwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for_rows(M);
B200_CHECK_ARG((dw == nullptr) == (db == nullptr), "layernorm_bwd: dw and db must be given together");
  if (dw) {
    B200_CHECK_ARG(d <= 4096, "layernorm_bwd: weight gradient supported for d <= 4096");
    const int nv = (int)((d / 8 + 31) / 32);
#define CALLW(NV) layernorm_bwd_kernel<NV, true><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, dw, db, M, (int)d)
    if (nv <= 4) { CALLW(4); } else if (nv <= 8) { CALLW(8); } else { CALLW(16); }
#undef CALLW
  } else {
#define CALL(NV) layernorm_bwd_kernel<NV, false><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, nullptr, nullptr, M, (int)d)
    DISPATCH_MAXV(d, CALL);
#undef CALL
  }
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
