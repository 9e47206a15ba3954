// K4 / K4' — RMSNorm and LayerNorm, forward and backward (HBM-bound).
//
// One warp owns one row (RMSNorm: 1-8 cooperating warps, by row width): the row is read once with
// 128-bit loads, kept in registers, reduced with warp shuffles and written once. Algorithmic bytes per element:
// fwd 4 B (read x, write y; +4 B with the fused residual add), bwd 6 B (read x, dy; write dx).
#include <cstdlib>

#include "common.cuh"

namespace {

constexpr int WARPS = 8;
constexpr int DW_MAX_D = 4096;   // widest row whose weight-gradient partials are reduced in shared memory (2 x 16 KB)

// All loads of a row are issued before anything depends on them (plain streaming loads the compiler
// may batch; lanes past the row end read nothing and hold zeros), so a warp has its whole 4-16 KB in
// flight at once instead of one 512-byte round trip per 256 columns.
__device__ __forceinline__ uint4 ld_row_v4(const bf16* p, bool ok) {
  return ok ? __ldcs(reinterpret_cast<const uint4*>(p)) : make_uint4(0u, 0u, 0u, 0u);
}
__device__ __forceinline__ void ld_w8(const float* __restrict__ w, int idx, bool ok, float* ww) {
  const int c = ok ? idx * 8 : 0;
  const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + c));
  const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + c + 4));
  ww[0] = w0.x; ww[1] = w0.y; ww[2] = w0.z; ww[3] = w0.w;
  ww[4] = w1.x; ww[5] = w1.y; ww[6] = w1.z; ww[7] = w1.w;
}

// Row partition shared by the RMSNorm kernels: a row is covered by WPR cooperating warps of the
// 8-warp CTA (WPR = 1 up to 2048 columns, 2 up to 4096, ...), each lane holding NV <= 8 vectors, so
// the register footprint does not grow with d. Partial sums meet in shared memory.
template <int WPR>
__device__ __forceinline__ void row_sum2(float& a, float& b, float (*red)[2], int warp) {
  a = warp_sum(a);
  b = warp_sum(b);
  if (WPR > 1) {
    const int lane = threadIdx.x & 31;
    __syncthreads();                       // previous use of red[] is over
    if (lane == 0) {
      red[warp][0] = a;
      red[warp][1] = b;
    }
    __syncthreads();
    const int w0 = warp - warp % WPR;
    a = 0.f;
    b = 0.f;
#pragma unroll
    for (int k = 0; k < WPR; ++k) {
      a += red[w0 + k][0];
      b += red[w0 + k][1];
    }
  }
}

// ---------------- RMSNorm forward (bf16) ----------------
// Reference: Models/Llama/common_components.py:66-70. In bf16 eager mode every op rounds:
// x.pow(2) -> bf16, mean -> bf16 (fp32 accumulate), + eps -> bf16, rsqrt -> bf16,
// x * r -> bf16, * weight(fp32) -> fp32, .to(bf16).
template <int NV, int WPR, bool RES, bool STEPS>
__global__ void __launch_bounds__(WARPS * 32)
rmsnorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res,
                   const float* __restrict__ w, bf16* __restrict__ sum_out, bf16* __restrict__ y,
                   long long M, int d, float eps) {
  __shared__ float red[WARPS][2];
  pdl_launch_dependents();                  // see gemm.cu: the next kernel's prologue may overlap this grid's tail
  pdl_wait();
  constexpr int RPB = WARPS / WPR;          // rows per CTA pass
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col0 = (warp % WPR) * (NV * 32) + lane;   // first vector of this lane
  const int nvec = d >> 3;
  for (long long row0 = (long long)blockIdx.x * RPB; row0 < M; row0 += (long long)gridDim.x * RPB) {
    const long long row = row0 + warp / WPR;
    const bool rok = row < M;
    uint4 v[NV], g4[RES ? NV : 1];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      v[i] = ld_row_v4(x + row * d + idx * 8, rok && idx < nvec);
      if (RES) g4[i] = ld_row_v4(res + row * d + idx * 8, rok && idx < nvec);
    }
    float ss = 0.f, unused = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      float f[8];
      unpack8(v[i], f);
      if (RES) {
        float g[8];
        unpack8(g4[i], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = bf16_round(f[j] + g[j]);
        v[i] = pack8(f);
        if (rok && idx < nvec) st_v4(sum_out + row * d + idx * 8, v[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += STEPS ? bf16_round(f[j] * f[j]) : f[j] * f[j];
    }
    row_sum2<WPR>(ss, unused, red, warp);
    float mean = ss / (float)d;
    float r;
    if (STEPS) {
      mean = bf16_round(mean);
      r = bf16_round(rsqrtf(bf16_round(mean + eps)));
    } else {
      r = rsqrtf(mean + eps);
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      const bool ok = rok && idx < nvec;
      float f[8], ww[8];
      unpack8(v[i], f);
      ld_w8(w, idx, ok, ww);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float xn = f[j] * r;
        if (STEPS) xn = bf16_round(xn);
        f[j] = xn * ww[j];
      }
      if (ok) st_v4(y + row * d + idx * 8, pack8(f));
    }
  }
}

// ---------------- RMSNorm backward (bf16) ----------------
// x and dy of the row are fetched up front and stay in registers (with a weight gradient to
// accumulate, dy is re-read in the second sweep instead — an L1/L2 hit — to stay within the register
// file); sum(x^2) and sum(dy*w*x) are formed together (the second only needs 1/rms as a factor
// afterwards); the residual-stream gradient is fetched at the start of the second sweep.
template <int NV, int WPR, bool HAS_DW, bool DRES>
__global__ void __launch_bounds__(WARPS * 32, 2)   // two CTAs per SM: 16 rows in flight
rmsnorm_bwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w,
                   const bf16* __restrict__ dy, const bf16* __restrict__ dres, bf16* __restrict__ dx,
                   float* __restrict__ dw, long long M, int d, float eps) {
  __shared__ float red[WARPS][2];
  pdl_launch_dependents();
  pdl_wait();
  constexpr int RPB = WARPS / WPR;
  constexpr bool KEEP_G = !HAS_DW;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col0 = (warp % WPR) * (NV * 32) + lane;
  const int nvec = d >> 3;
  float dwacc[HAS_DW ? NV * 8 : 1];
  if (HAS_DW) {
#pragma unroll
    for (int i = 0; i < (HAS_DW ? NV * 8 : 1); ++i) dwacc[i] = 0.f;
  }
  for (long long row0 = (long long)blockIdx.x * RPB; row0 < M; row0 += (long long)gridDim.x * RPB) {
    const long long row = row0 + warp / WPR;
    const bool rok = row < M;
    uint4 vx[NV], vg[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      vx[i] = ld_row_v4(x + row * d + idx * 8, rok && idx < nvec);
      vg[i] = ld_row_v4(dy + row * d + idx * 8, rok && idx < nvec);
      // the residual-stream gradient is only needed after the row reduction and there are no registers left to hold it
      // until then: ask for it now, into L2, so that the loads after the barrier do not start a second DRAM round trip
      if (DRES && rok && idx < nvec) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres + row * d + idx * 8));
    }
    float ss = 0.f, dot = 0.f;   // sum x^2, sum dy*w*x
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      float f[8], g[8], ww[8];
      unpack8(vx[i], f);
      unpack8(vg[i], g);
      ld_w8(w, idx, idx < nvec, ww);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ss += f[j] * f[j];
        dot += g[j] * ww[j] * f[j];
      }
    }
    row_sum2<WPR>(ss, dot, red, warp);
    const float r = rsqrtf(ss / (float)d + eps);
    const float c = dot * r / (float)d;   // mean(g * xn), xn = x * r
    uint4 ve[DRES ? NV : 1];
    if (DRES || !KEEP_G) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int idx = col0 + 32 * i;
        if (DRES) ve[i] = ld_row_v4(dres + row * d + idx * 8, rok && idx < nvec);
        if (!KEEP_G) vg[i] = ld_row_v4(dy + row * d + idx * 8, rok && idx < nvec);
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      const bool ok = rok && idx < nvec;
      float f[8], g[8], o[8], ww[8];
      unpack8(vx[i], f);
      unpack8(vg[i], g);
      ld_w8(w, idx, idx < nvec, ww);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xn = f[j] * r;
        o[j] = r * (g[j] * ww[j] - xn * c);
        if (HAS_DW) dwacc[i * 8 + j] += g[j] * xn;
      }
      if (DRES) {
        float e[8];
        unpack8(ve[i], e);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += e[j];
      }
      if (ok) st_v4(dx + row * d + idx * 8, pack8(o));
    }
  }
  if (HAS_DW) {
    // rows of a CTA meet in shared memory, then one global atomic per column per CTA (see layernorm_bwd_kernel)
    __shared__ float sdw[HAS_DW ? 2 * DW_MAX_D : 1];
    const bool in_smem = d <= 2 * DW_MAX_D;
    if (in_smem) {
      for (int c = threadIdx.x; c < d; c += blockDim.x) sdw[c] = 0.f;
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = col0 + 32 * i;
      if (idx < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(in_smem ? &sdw[idx * 8 + j] : &dw[idx * 8 + j], dwacc[i * 8 + j]);
      }
    }
    if (in_smem) {
      __syncthreads();
      for (int c = threadIdx.x; c < d; c += blockDim.x) atomicAdd(&dw[c], sdw[c]);
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
                     float* __restrict__ db, long long M, int d, const bf16* __restrict__ dres) {
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
        // gradient of the residual stream (x also feeds the block's residual add): added while dx is written, asked
        // for now so that it is in L2 by then (see rmsnorm_bwd_kernel)
        if (dres != nullptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres + row * d + idx * 8));
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
    // The warps of the CTA meet in shared memory first and the CTA issues ONE global atomic per column: with one
    // atomic per column per WARP (and a warp per row) 4096 x 1600 rows x columns meant 13 M atomics on 3200
    // addresses — 438 us for a kernel that streams 39 MB (GPT2-1.5B, 24 % of its training step).
    __shared__ float sdw[DW_MAX_D], sdb[DW_MAX_D];
    for (int c = threadIdx.x; c < d; c += blockDim.x) sdw[c] = sdb[c] = 0.f;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          atomicAdd(&sdw[idx * 8 + j], dwacc[i * 8 + j]);
          atomicAdd(&sdb[idx * 8 + j], dbacc[i * 8 + j]);
        }
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
      atomicAdd(&dw[c], sdw[c]);
      atomicAdd(&db[c], sdb[c]);
    }
  }
}

// dw[c] += sum_rows dy * xn, db[c] += sum_rows dy as a pass of their own (tiles of 256 columns x rows_per_block rows, a warp
// reads 512 contiguous bytes of a row, the eight warps take every eighth row, ONE atomic per column and CTA). Keeping
// these partial sums in registers inside the dx kernel costs it 128 registers per thread — one CTA of eight rows per SM,
// 38 us for GPT2-1.5B's [4096 x 1600] against a 6 us byte floor; re-reading x and dy here (26 MB) is cheaper.
__global__ void __launch_bounds__(256)
layernorm_dwdb_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, const float* __restrict__ mean,
                      const float* __restrict__ rstd, float* __restrict__ dw, float* __restrict__ db, long long M, int d,
                      int rows_per_block) {
  __shared__ float pw[8][264], pb[8][264];
  const int cv = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + cv) * 8;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long r1 = r0 + rows_per_block < M ? r0 + rows_per_block : M;
  float aw[8], ab[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) aw[j] = ab[j] = 0.f;
  if (c < d) {
    for (long long r = r0 + rl; r < r1; r += 16) {
      const bool two = r + 8 < r1;
      const uint4 vx0 = ld_nc_v4(x + r * d + c), vg0 = ld_nc_v4(dy + r * d + c);
      const uint4 vx1 = two ? ld_nc_v4(x + (r + 8) * d + c) : make_uint4(0, 0, 0, 0);
      const uint4 vg1 = two ? ld_nc_v4(dy + (r + 8) * d + c) : make_uint4(0, 0, 0, 0);
      const float mu0 = mean[r], rs0 = rstd[r];
      const float mu1 = two ? mean[r + 8] : 0.f, rs1 = two ? rstd[r + 8] : 0.f;
      float f[8], g[8];
      unpack8(vx0, f);
      unpack8(vg0, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        aw[j] += g[j] * ((f[j] - mu0) * rs0);
        ab[j] += g[j];
      }
      unpack8(vx1, f);
      unpack8(vg1, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        aw[j] += g[j] * ((f[j] - mu1) * rs1);
        ab[j] += g[j];
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    pw[rl][cv * 8 + j] = aw[j];
    pb[rl][cv * 8 + j] = ab[j];
  }
  __syncthreads();
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (col < d) {
    float tw = 0.f, tb = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      tw += pw[k][threadIdx.x];
      tb += pb[k][threadIdx.x];
    }
    atomicAdd(&dw[col], tw);
    atomicAdd(&db[col], tb);
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

// (vectors per lane, warps per row) for a row of d columns. Many light warps beat few heavy ones:
// at d = 2048, 2 warps x 4 vectors ran the backward in 20.9 us against 27.2 us for 1 warp x 8 vectors
// (more CTAs resident, loads of one row overlap the arithmetic of another).
#define DISPATCH_ROW(d, CALL)                      \
  do {                                             \
    if ((d) <= 1024) { CALL(4, 1); }               \
    else if ((d) <= 2048) { CALL(4, 2); }          \
    else if ((d) <= 4096) { CALL(4, 4); }          \
    else if ((d) <= 8192) { CALL(8, 4); }          \
    else { CALL(8, 8); }                           \
  } while (0)

namespace {
int grid_for_rows_wpr(long long M, long long d) {
  const int wpr = d <= 1024 ? 1 : (d <= 2048 ? 2 : (d <= 8192 ? 4 : 8));
  const long long rpb = WARPS / wpr;
  const long long blocks = (M + rpb - 1) / rpb;
  const long long cap = (long long)b200_num_sms() * 8;
  return (int)(blocks < cap ? blocks : cap);
}
}  // namespace

extern "C" int b200_rmsnorm_fwd_bf16(const void* x, const void* res, const float* w, void* sum_out,
                                     void* y, int64_t M, int64_t d, float eps, int bf16_steps,
                                     void* stream) {
  B200_CHECK_ARG(x && w && y && M > 0, "rmsnorm_fwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "rmsnorm_fwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  B200_CHECK_ARG(res == nullptr || sum_out != nullptr, "rmsnorm_fwd: res given without sum_out");
  B200_CHECK_ARG(b200_aligned16(x) && b200_aligned16(y) && b200_aligned16(w), "rmsnorm_fwd: pointers must be 16-B aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for_rows_wpr(M, d);
#define FWD_ARGS (const bf16*)x, (const bf16*)res, w, (bf16*)sum_out, (bf16*)y, M, (int)d, eps
#define CALL(NV, WPR)                                                                                    \
  do {                                                                                                   \
    if (res != nullptr) {                                                                                \
      if (bf16_steps) b200_launch_pdl_wide(rmsnorm_fwd_kernel<NV, WPR, true, true>, dim3(grid), dim3(WARPS * 32), 0, st, FWD_ARGS);   \
      else b200_launch_pdl_wide(rmsnorm_fwd_kernel<NV, WPR, true, false>, dim3(grid), dim3(WARPS * 32), 0, st, FWD_ARGS);             \
    } else {                                                                                             \
      if (bf16_steps) b200_launch_pdl_wide(rmsnorm_fwd_kernel<NV, WPR, false, true>, dim3(grid), dim3(WARPS * 32), 0, st, FWD_ARGS);  \
      else b200_launch_pdl_wide(rmsnorm_fwd_kernel<NV, WPR, false, false>, dim3(grid), dim3(WARPS * 32), 0, st, FWD_ARGS);            \
    }                                                                                                    \
  } while (0)
  DISPATCH_ROW(d, CALL);
#undef CALL
#undef FWD_ARGS
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rmsnorm_bwd_bf16(const void* x, const float* w, const void* dy,
                                     const void* dres_in, void* dx, float* dw, int64_t M, int64_t d,
                                     float eps, void* stream) {
  B200_CHECK_ARG(x && w && dy && dx && M > 0, "rmsnorm_bwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "rmsnorm_bwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for_rows_wpr(M, d);
  if (dw && grid > b200_num_sms()) grid = b200_num_sms();   // fat CTAs: column partials span many rows
#define BWD_ARGS (const bf16*)x, w, (const bf16*)dy, (const bf16*)dres_in, (bf16*)dx, dw, M, (int)d, eps
#define CALL(NV, WPR)                                                                                   \
  do {                                                                                                  \
    if (dw) {                                                                                           \
      if (dres_in) b200_launch_pdl_wide(rmsnorm_bwd_kernel<NV, WPR, true, true>, dim3(grid), dim3(WARPS * 32), 0, st, BWD_ARGS);     \
      else b200_launch_pdl_wide(rmsnorm_bwd_kernel<NV, WPR, true, false>, dim3(grid), dim3(WARPS * 32), 0, st, BWD_ARGS);            \
    } else {                                                                                            \
      if (dres_in) b200_launch_pdl_wide(rmsnorm_bwd_kernel<NV, WPR, false, true>, dim3(grid), dim3(WARPS * 32), 0, st, BWD_ARGS);    \
      else b200_launch_pdl_wide(rmsnorm_bwd_kernel<NV, WPR, false, false>, dim3(grid), dim3(WARPS * 32), 0, st, BWD_ARGS);           \
    }                                                                                                   \
  } while (0)
  DISPATCH_ROW(d, CALL);
#undef CALL
#undef BWD_ARGS
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

extern "C" int b200_layernorm_bwd_res_bf16(const void* x, const void* w, const void* dy, const void* dres,
                                           const float* mean, const float* rstd, void* dx, float* dw,
                                           float* db, int64_t M, int64_t d, void* stream);
extern "C" int b200_layernorm_bwd_bf16(const void* x, const void* w, const void* dy,
                                       const float* mean, const float* rstd, void* dx, float* dw,
                                       float* db, int64_t M, int64_t d, void* stream) {
  return b200_layernorm_bwd_res_bf16(x, w, dy, nullptr, mean, rstd, dx, dw, db, M, d, stream);
}

extern "C" int b200_layernorm_bwd_res_bf16(const void* x, const void* w, const void* dy, const void* dres,
                                           const float* mean, const float* rstd, void* dx, float* dw,
                                           float* db, int64_t M, int64_t d, void* stream) {
  B200_CHECK_ARG(x && w && dy && mean && rstd && dx && M > 0, "layernorm_bwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0 && d <= 16384, "layernorm_bwd: d=%lld must be a multiple of 8 and <= 16384", (long long)d);
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for_rows(M);
  B200_CHECK_ARG((dw == nullptr) == (db == nullptr), "layernorm_bwd: dw and db must be given together");
  static const bool split_dw = [] { const char* e = getenv("B200_LN_SPLIT_DW"); return !(e && e[0] == '0'); }();
  if (dw && split_dw && b200_aligned16(x) && b200_aligned16(dy)) {
    // dx by the light kernel (no column partials in registers), dw / db by a tiled pass of their own
#define CALL(NV) layernorm_bwd_kernel<NV, false><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, nullptr, nullptr, M, (int)d, (const bf16*)dres)
    DISPATCH_MAXV(d, CALL);
#undef CALL
    B200_CHECK_LAUNCH();
    const long long col_tiles = (d + 255) / 256;
    long long rpb = (M * col_tiles + 4LL * b200_num_sms() - 1) / (4LL * b200_num_sms());
    rpb = (rpb + 31) / 32 * 32;
    if (rpb < 32) rpb = 32;
    dim3 g2((unsigned)col_tiles, (unsigned)((M + rpb - 1) / rpb));
    B200_CHECK_ARG(g2.y <= 65535, "layernorm_bwd: too many rows for one launch");
    layernorm_dwdb_kernel<<<g2, 256, 0, st>>>((const bf16*)x, (const bf16*)dy, mean, rstd, dw, db, M, (int)d, (int)rpb);
    B200_CHECK_LAUNCH();
    b200_count_launch(2);
    return 0;
  }
  if (dw) {
    // weight gradients: every warp keeps its column partials over MANY rows (one CTA per SM), see the kernel
    const int fat = b200_num_sms();
    if (grid > fat) grid = fat;
    B200_CHECK_ARG(d <= 4096, "layernorm_bwd: weight gradient supported for d <= 4096");
    const int nv = (int)((d / 8 + 31) / 32);
#define CALLW(NV) layernorm_bwd_kernel<NV, true><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, dw, db, M, (int)d, (const bf16*)dres)
    if (nv <= 4) { CALLW(4); } else if (nv <= 8) { CALLW(8); } else { CALLW(16); }
#undef CALLW
  } else {
#define CALL(NV) layernorm_bwd_kernel<NV, false><<<grid, WARPS * 32, 0, st>>>((const bf16*)x, (const bf16*)w, (const bf16*)dy, mean, rstd, (bf16*)dx, nullptr, nullptr, M, (int)d, (const bf16*)dres)
    DISPATCH_MAXV(d, CALL);
#undef CALL
  }
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
