// K5/K5'/K6/K7/K9 — SwiGLU, GELU, RoPE, embedding gather/scatter, casts, transpose, add.
// All HBM-bound: 128-bit loads/stores, grid sized as a multiple of the SM count, grid-stride.
#include "common.cuh"

namespace {

inline int ew_grid(long long nvec, int threads) {
  long long blocks = (nvec + threads - 1) / threads;
  const long long cap = (long long)b200_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// MUFU.EX2 + MUFU.RCP: the IEEE division this replaces cost more issue slots than the rest of SwiGLU and
// made the forward kernel instruction-bound (4.5 TB/s) instead of HBM-bound
__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// (rows, 16-byte column vectors) launch shape for row-pitched elementwise kernels: blockIdx.x covers
// the vectors of a row, blockIdx.y strides over rows — no 64-bit division per element
inline dim3 row_grid(long long M, int vec_per_row, int threads) {
  long long gx = (vec_per_row + threads - 1) / threads;
  long long gy = (long long)b200_num_sms() * 16 / gx;
  if (gy > M) gy = M;
  if (gy > 65535) gy = 65535;
  if (gy < 1) gy = 1;
  return dim3((unsigned)gx, (unsigned)gy);
}

// ---------------- SwiGLU ----------------
// Reference (common_components.py:87,123): silu(a) = a * sigmoid(a) -> bf16 each op, then * b.
// Algorithmic bytes: fwd 6 B per output element (read a, b; write h), bwd 10 B.
__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const bf16* __restrict__ pa, long long lda, const bf16* __restrict__ pb,
                  long long ldb, bf16* __restrict__ h, long long M, int F, int bf16_steps) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= F) return;
  for (long long row = blockIdx.y; row < M; row += gridDim.y) {
    float a[8], b[8], o[8];
    unpack8(ld_nc_v4(pa + row * lda + c), a);
    unpack8(ld_nc_v4(pb + row * ldb + c), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = sigmoidf_(a[j]);
      if (bf16_steps) {
        s = bf16_round(s);
        o[j] = bf16_round(a[j] * s) * b[j];
      } else {
        o[j] = a[j] * s * b[j];
      }
    }
    st_v4(h + row * F + c, pack8(o));
  }
}

__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const bf16* __restrict__ pa, long long lda, const bf16* __restrict__ pb,
                  long long ldb, const bf16* __restrict__ dh, bf16* __restrict__ pda, long long ldda,
                  bf16* __restrict__ pdb, long long lddb, long long M, int F) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= F) return;
  for (long long row = blockIdx.y; row < M; row += gridDim.y) {
    float a[8], b[8], g[8], da[8], db[8];
    unpack8(ld_nc_v4(pa + row * lda + c), a);
    unpack8(ld_nc_v4(pb + row * ldb + c), b);
    unpack8(ld_nc_v4(dh + row * F + c), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = sigmoidf_(a[j]);
      const float silu = a[j] * s;
      da[j] = g[j] * b[j] * (s * (1.f + a[j] * (1.f - s)));
      db[j] = g[j] * silu;
    }
    st_v4(pda + row * ldda + c, pack8(da));
    st_v4(pdb + row * lddb + c, pack8(db));
  }
}

// ---------------- GELU (erf) ----------------
__global__ void __launch_bounds__(256)
gelu_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long n) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_nc_v4(x + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = 0.5f * f[j] * (1.f + erff(f[j] * 0.70710678118654752440f));
    st_v4(y + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) {
      const float f = __bfloat162float(x[k]);
      y[k] = __float2bfloat16_rn(0.5f * f * (1.f + erff(f * 0.70710678118654752440f)));
    }
}

__device__ __forceinline__ float gelu_grad(float x) {
  const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

__global__ void __launch_bounds__(256)
gelu_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ dx,
                long long n) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8], g[8];
    unpack8(ld_nc_v4(x + i * 8), f);
    unpack8(ld_nc_v4(dy + i * 8), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = g[j] * gelu_grad(f[j]);
    st_v4(dx + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k)
      dx[k] = __float2bfloat16_rn(__bfloat162float(dy[k]) * gelu_grad(__bfloat162float(x[k])));
}

// ---------------- RoPE (rotate-half, in place) ----------------
// Reference (common_components.py:24-35): x*cos + cat(-x2, x1)*sin, bf16 tables, every op
// rounds to bf16. Each thread owns 8 columns of the first half and the matching 8 of the second.
__global__ void __launch_bounds__(256)
rope_kernel(const bf16* x, long long ld, bf16* y, long long ldy, const bf16* __restrict__ cosT,
            const bf16* __restrict__ sinT, long long B, int T, int nheads, int hd, int inverse,
            int bf16_steps) {
  // block = (vectors of one token, tokens): threadIdx.x walks the nheads * vh vectors of a token,
  // threadIdx.y + blockIdx.x * blockDim.y strides over tokens; 32-bit index arithmetic only
  const int half = hd >> 1;
  const int vh = half >> 3;  // vectors per half head
  const int per_tok = nheads * vh;
  const long long ntok = B * T;
  for (long long tok = (long long)blockIdx.x * blockDim.y + threadIdx.y; tok < ntok; tok += (long long)gridDim.x * blockDim.y)
  for (int j = threadIdx.x; j < per_tok; j += blockDim.x) {
    const int head = j / vh;
    const int v = j - head * vh;
    const int t = (int)(tok % T);
    const bf16* p1 = x + tok * ld + (long long)head * hd + v * 8;
    const bf16* p2 = p1 + half;
    bf16* q1 = y + tok * ldy + (long long)head * hd + v * 8;
    bf16* q2 = q1 + half;
    float x1[8], x2[8], c1[8], c2[8], s1[8], s2[8], o1[8], o2[8];
    unpack8(ld_v4(p1), x1);
    unpack8(ld_v4(p2), x2);
    unpack8(ld_nc_v4(cosT + (long long)t * hd + v * 8), c1);
    unpack8(ld_nc_v4(cosT + (long long)t * hd + half + v * 8), c2);
    unpack8(ld_nc_v4(sinT + (long long)t * hd + v * 8), s1);
    unpack8(ld_nc_v4(sinT + (long long)t * hd + half + v * 8), s2);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (!inverse) {
        if (bf16_steps) {
          o1[j] = bf16_round(x1[j] * c1[j]) + bf16_round(-x2[j] * s1[j]);
          o2[j] = bf16_round(x2[j] * c2[j]) + bf16_round(x1[j] * s2[j]);
        } else {
          o1[j] = x1[j] * c1[j] - x2[j] * s1[j];
          o2[j] = x2[j] * c2[j] + x1[j] * s2[j];
        }
      } else {
        // transpose of the map above: dx1 = dy1*c1 + dy2*s2, dx2 = dy2*c2 - dy1*s1
        o1[j] = x1[j] * c1[j] + x2[j] * s2[j];
        o2[j] = x2[j] * c2[j] - x1[j] * s1[j];
      }
    }
    st_v4(q1, pack8(o1));
    st_v4(q2, pack8(o2));
  }
}

// ---------------- embedding ----------------
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const long long* __restrict__ idx, const bf16* __restrict__ tok_w,
                     const bf16* __restrict__ pos_w, bf16* __restrict__ out, long long M, int T,
                     int d, long long vocab) {
  const int dv = d >> 3;
  const long long total = M * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / dv;
    const int c = (int)(i % dv) * 8;
    long long id = idx[row];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    uint4 v = ld_nc_v4(tok_w + id * d + c);
    if (pos_w != nullptr) {
      float a[8], b[8];
      unpack8(v, a);
      unpack8(ld_nc_v4(pos_w + (row % T) * d + c), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += b[j];
      v = pack8(a);
    }
    st_v4(out + row * d + c, v);
  }
}

__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const long long* __restrict__ idx, const bf16* __restrict__ dout,
                     float* __restrict__ dtok, float* __restrict__ dpos, long long M, int T, int d,
                     long long vocab) {
  const int dv = d >> 3;
  const long long total = M * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / dv;
    const int c = (int)(i % dv) * 8;
    const long long id = idx[row];
    if (id < 0 || id >= vocab) continue;
    float g[8];
    unpack8(ld_nc_v4(dout + row * d + c), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&dtok[id * d + c + j], g[j]);
      if (dpos != nullptr) atomicAdd(&dpos[(row % T) * d + c + j], g[j]);
    }
  }
}

// ---------------- casts ----------------
__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 a = *reinterpret_cast<const float4*>(in + i * 8);
    const float4 b = *reinterpret_cast<const float4*>(in + i * 8 + 4);
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    st_v4(out + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) out[k] = __float2bfloat16_rn(in[k]);
}

__global__ void __launch_bounds__(256)
cast_bf16_f32_kernel(const bf16* __restrict__ in, float* __restrict__ out, long long n) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_nc_v4(in + i * 8), f);
    *reinterpret_cast<float4*>(out + i * 8) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(f[4], f[5], f[6], f[7]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) out[k] = __bfloat162float(in[k]);
}

// ---------------- add ----------------
__global__ void __launch_bounds__(256)
add_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, bf16* __restrict__ out,
           long long n) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8], g[8];
    unpack8(ld_nc_v4(a + i * 8), f);
    unpack8(ld_nc_v4(b + i * 8), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] += g[j];
    st_v4(out + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k)
      out[k] = __float2bfloat16_rn(__bfloat162float(a[k]) + __bfloat162float(b[k]));
}

// ---------------- transpose (out[c][r] = alpha * in[r][c]) ----------------
__global__ void __launch_bounds__(256)
transpose_kernel(const bf16* __restrict__ in, long long ldi, bf16* __restrict__ out, long long ldo,
                 int rows, int cols, float alpha) {
  __shared__ bf16 tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = r0 + ty + k * 8, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + k * 8][tx] = in[(long long)r * ldi + c];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = c0 + ty + k * 8, r = r0 + tx;
    if (r < rows && c < cols)
      out[(long long)c * ldo + r] = __float2bfloat16_rn(alpha * __bfloat162float(tile[tx][ty + k * 8]));
  }
}

// ---------------- column sum (bias gradient): out[n] += sum_m x[m, n] ----------------
__global__ void __launch_bounds__(256)
colsum_kernel(const bf16* __restrict__ x, long long ld, float* __restrict__ out, long long M, int N,
              int rows_per_block) {
  const int c = (blockIdx.x * 256 + threadIdx.x) * 8;
  if (c >= N) return;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long r1 = r0 + rows_per_block < M ? r0 + rows_per_block : M;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c + 8 <= N && (ld & 7) == 0) {
    // eight rows' loads are issued before the first one is used: one 16-byte load in flight per thread at a time made
    // this a latency chain (36 us for a 52 MB matrix)
    long long r = r0;
    for (; r + 8 <= r1; r += 8) {
      uint4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = ld_nc_v4(x + (r + i) * ld + c);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < r1; ++r) {
      float f[8];
      unpack8(ld_nc_v4(x + r * ld + c), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  } else {
    for (long long r = r0; r < r1; ++r)
      for (int j = 0; j < 8 && c + j < N; ++j) acc[j] += __bfloat162float(x[r * ld + c + j]);
  }
  for (int j = 0; j < 8 && c + j < N; ++j) atomicAdd(&out[c + j], acc[j]);
}

// Aligned form (N % 8 == 0, 16-byte rows): a CTA takes 256 columns x rows_per_block rows; a warp reads 512 contiguous
// bytes of one row per load, the eight warps take every eighth row, four rows' loads are in flight per thread; the warps
// meet in shared memory and the CTA issues ONE atomic per column. The first form gave every thread its own 8 columns
// and 32 rows and let it add them to the result itself: for GPT2-1.5B's [4096 x 1600] bias gradients that is 128
// atomics per address from 128 CTAs (27 us for 13-52 MB; 3 launches per block, 5 % of that config's step).
__global__ void __launch_bounds__(256)
colsum_tile_kernel(const bf16* __restrict__ x, long long ld, float* __restrict__ out, long long M, int N,
                   int rows_per_block) {
  __shared__ float part[8][264];
  const int cv = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + cv) * 8;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long r1 = r0 + rows_per_block < M ? r0 + rows_per_block : M;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < N) {
    long long r = r0 + rl;
    for (; r + 24 < r1; r += 32) {
      uint4 v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = ld_nc_v4(x + (r + 8 * i) * ld + c);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < r1; r += 8) {
      float f[8];
      unpack8(ld_nc_v4(x + r * ld + c), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[rl][cv * 8 + j] = acc[j];
  __syncthreads();
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (col < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
    atomicAdd(&out[col], t);
  }
}

// ---------------- row compaction for the vocab projection ----------------
// Rows whose target is ignore_index (-100) contribute nothing to the loss or to any gradient
// (train.py:91: mean over non-ignored rows), so the fused vocab-projection + CE op only needs the
// other rows. idx[j] (j < capacity) = j-th non-ignored row in ascending order, -1 for padding;
// meta[0] = number of non-ignored rows, meta[1] = 1 if that number exceeds `capacity`.
__global__ void __launch_bounds__(1024)
compact_rows_kernel(const long long* __restrict__ targets, int M, int capacity, int* __restrict__ idx,
                    int* __restrict__ meta) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  for (int j = threadIdx.x; j < capacity; j += 1024) idx[j] = -1;
  __syncthreads();
  for (int base = 0; base < M; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = (i < M && targets[i] != -100) ? 1 : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (w == 0) {
      int t = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += n;
      }
      warp_tot[lane] = t;   // inclusive totals per warp
    }
    __syncthreads();
    const int pos = carry + (w > 0 ? warp_tot[w - 1] : 0) + incl - v;
    if (v && pos < capacity) idx[pos] = i;
    __syncthreads();
    if (threadIdx.x == 0) carry += warp_tot[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    meta[0] = carry;
    meta[1] = carry > capacity ? 1 : 0;
  }
}

// out[j, :] = x[idx[j], :] (zeros for padding); tout[j] = targets[idx[j]] (-100 for padding).
// If the compaction overflowed (meta[1]) the gathered activations are poisoned with NaNs so the
// loss turns into NaN instead of silently dropping rows.
__global__ void __launch_bounds__(256)
gather_rows_kernel(const bf16* __restrict__ x, long long ldx, const long long* __restrict__ targets,
                   const int* __restrict__ idx, const int* __restrict__ meta, bf16* __restrict__ out,
                   long long* __restrict__ tout, int capacity, int d) {
  const int dv = d >> 3;
  const long long total = (long long)capacity * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i / dv), c = (int)(i % dv) * 8;
    const int r = idx[j];
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r >= 0) v = ld_nc_v4(x + (long long)r * ldx + c);
    if (meta[1]) v = make_uint4(0x7FC07FC0u, 0x7FC07FC0u, 0x7FC07FC0u, 0x7FC07FC0u);  // bf16 NaNs
    st_v4(out + (long long)j * d + c, v);
    if (c == 0) tout[j] = r >= 0 ? targets[r] : -100;
  }
}

// dst[idx[j], :] = src[j, :]; dst rows that are not listed must have been zeroed by the caller
__global__ void __launch_bounds__(256)
scatter_rows_kernel(const bf16* __restrict__ src, const int* __restrict__ idx, bf16* __restrict__ dst,
                    long long ldd, int capacity, int d) {
  const int dv = d >> 3;
  const long long total = (long long)capacity * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i / dv), c = (int)(i % dv) * 8;
    const int r = idx[j];
    if (r >= 0) st_v4(dst + (long long)r * ldd + c, ld_nc_v4(src + (long long)j * d + c));
  }
}

// ---------------- dropout (nn.Dropout in Models/GPT2/GPT2.py:24,81,102; K10) ----------------
// y = (res) + keep(i) * x / (1 - p); the mask is a pure function of (seed, element index), so the
// same call on the upstream gradient is the backward. 6 B/elt (+2 with the residual).
__global__ void __launch_bounds__(256)
dropout_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res, bf16* __restrict__ y, long long n,
               uint32_t thr16, float inv_keep, unsigned long long seed, const unsigned long long* __restrict__ seed_dev) {
  if (seed_dev != nullptr) seed += *seed_dev;   // per-step offset in device memory: a replayed graph draws new masks
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_nc_v4(x + i * 8), f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      bool k0, k1;
      dropout_keep2((unsigned long long)(i * 4 + q), seed, thr16, k0, k1);
      f[2 * q] = k0 ? f[2 * q] * inv_keep : 0.f;
      f[2 * q + 1] = k1 ? f[2 * q + 1] * inv_keep : 0.f;
    }
    if (res != nullptr) {
      float g[8];
      unpack8(ld_nc_v4(res + i * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = bf16_round(f[j]) + g[j];
    }
    st_v4(y + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) {
      float v = dropout_keep((unsigned long long)k, seed, thr16) ? __bfloat162float(x[k]) * inv_keep : 0.f;
      if (res != nullptr) v = bf16_round(v) + __bfloat162float(res[k]);
      y[k] = __float2bfloat16_rn(v);
    }
}

__global__ void __launch_bounds__(256)
dropout_mask_kernel(unsigned char* __restrict__ out, long long n, uint32_t thr16, unsigned long long seed) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = dropout_keep((unsigned long long)i, seed, thr16) ? 1 : 0;
}

// x *= *scale, skipped entirely when the device scalar is exactly 1 (the gradient of a loss that .backward() was
// called on directly): lets the fused vocab-projection backward honour a scaled loss without a host read.
__global__ void __launch_bounds__(256)
scale_dev_kernel(bf16* __restrict__ x, long long n, const float* __restrict__ scale) {
  const float s = *scale;
  if (s == 1.f) return;
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld_v4(x + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= s;
    st_v4(x + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = nv * 8; k < n; ++k) x[k] = __float2bfloat16_rn(__bfloat162float(x[k]) * s);
}

}  // namespace

extern "C" int b200_scale_bf16_dev(void* x, int64_t n, const float* scale_dev, void* stream) {
  B200_CHECK_ARG(x && scale_dev && n > 0, "scale_dev: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(x), "scale_dev: buffer must be 16-B aligned");
  scale_dev_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>((bf16*)x, n, scale_dev);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

static inline uint32_t drop_thr16(float p) {
  const float t = p * 65536.f + 0.5f;
  return t < 0.f ? 0u : (t > 65535.f ? 65535u : (uint32_t)t);
}

extern "C" int b200_dropout_bf16(const void* x, const void* res, void* y, int64_t n, float p, uint64_t seed,
                                 const uint64_t* seed_dev, void* stream) {
  B200_CHECK_ARG(x && y && n > 0, "dropout: null pointer or empty");
  B200_CHECK_ARG(p >= 0.f && p < 1.f, "dropout: p=%f must be in [0, 1)", (double)p);
  B200_CHECK_ARG(b200_aligned16(x) && b200_aligned16(y) && (res == nullptr || b200_aligned16(res)), "dropout: pointers must be 16-B aligned");
  dropout_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)x, (const bf16*)res, (bf16*)y, n, drop_thr16(p), 1.f / (1.f - p), (unsigned long long)seed,
      (const unsigned long long*)seed_dev);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_dropout_mask_u8(void* out, int64_t n, float p, uint64_t seed, void* stream) {
  B200_CHECK_ARG(out && n > 0 && p >= 0.f && p < 1.f, "dropout_mask: bad arguments");
  dropout_mask_kernel<<<ew_grid(n, 256), 256, 0, (cudaStream_t)stream>>>((unsigned char*)out, n, drop_thr16(p), (unsigned long long)seed);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_compact_rows(const int64_t* targets, int64_t M, int64_t capacity, int32_t* idx,
                                 int32_t* meta, void* stream) {
  B200_CHECK_ARG(targets && idx && meta && M > 0 && capacity > 0 && M < (1ll << 31), "compact_rows: bad arguments");
  compact_rows_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const long long*)targets, (int)M, (int)capacity, idx, meta);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_gather_rows_bf16(const void* x, int64_t ldx, const int64_t* targets, const int32_t* idx,
                                     const int32_t* meta, void* out, int64_t* targets_out,
                                     int64_t capacity, int64_t d, void* stream) {
  B200_CHECK_ARG(x && targets && idx && meta && out && targets_out && capacity > 0, "gather_rows: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && ldx % 8 == 0, "gather_rows: d and ldx must be multiples of 8");
  gather_rows_kernel<<<ew_grid(capacity * (d / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)x, ldx, (const long long*)targets, idx, meta, (bf16*)out, (long long*)targets_out, (int)capacity, (int)d);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_scatter_rows_bf16(const void* src, const int32_t* idx, void* dst, int64_t ldd,
                                      int64_t capacity, int64_t d, void* stream) {
  B200_CHECK_ARG(src && idx && dst && capacity > 0, "scatter_rows: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && ldd % 8 == 0, "scatter_rows: d and ldd must be multiples of 8");
  scatter_rows_kernel<<<ew_grid(capacity * (d / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)src, idx, (bf16*)dst, ldd, (int)capacity, (int)d);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_colsum_bf16(const void* x, int64_t ld, float* out, int64_t M, int64_t N,
                                void* stream) {
  B200_CHECK_ARG(x && out && M > 0 && N > 0 && ld >= N, "colsum: null pointer, empty or bad pitch");
  if (N % 8 == 0 && ld % 8 == 0 && b200_aligned16(x)) {
    const long long col_tiles = (N + 255) / 256;
    // ~4 CTAs per SM worth of tiles, at least 32 rows each
    long long rpb = (M * col_tiles + 4LL * b200_num_sms() - 1) / (4LL * b200_num_sms());
    rpb = (rpb + 31) / 32 * 32;
    if (rpb < 32) rpb = 32;
    dim3 grid((unsigned)col_tiles, (unsigned)((M + rpb - 1) / rpb));
    B200_CHECK_ARG(grid.y <= 65535, "colsum: too many rows for one launch");
    colsum_tile_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, ld, out, M, (int)N, (int)rpb);
    B200_CHECK_LAUNCH();
    b200_count_launch();
    return 0;
  }
  const int rows_per_block = M >= 2048 ? 32 : 128;     // enough CTAs to fill the GPU for the usual token counts
  dim3 grid((unsigned)((N + 2047) / 2048), (unsigned)((M + rows_per_block - 1) / rows_per_block));
  B200_CHECK_ARG(grid.y <= 65535, "colsum: too many rows for one launch");
  colsum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, ld, out, M, (int)N, rows_per_block);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_swiglu_fwd_bf16(const void* a, int64_t lda, const void* b, int64_t ldb, void* h,
                                    int64_t M, int64_t F, int bf16_steps, void* stream) {
  B200_CHECK_ARG(a && b && h && M > 0 && F > 0, "swiglu_fwd: null pointer or empty");
  B200_CHECK_ARG(F % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && lda >= F && ldb >= F, "swiglu_fwd: F, lda, ldb must be multiples of 8 and pitches >= F");
  B200_CHECK_ARG(b200_aligned16(a) && b200_aligned16(b) && b200_aligned16(h), "swiglu_fwd: pointers must be 16-B aligned");
  swiglu_fwd_kernel<<<row_grid(M, (int)(F / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)a, lda, (const bf16*)b, ldb, (bf16*)h, M, (int)F, bf16_steps);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_swiglu_bwd_bf16(const void* a, int64_t lda, const void* b, int64_t ldb,
                                    const void* dh, void* da, int64_t ldda, void* db, int64_t lddb,
                                    int64_t M, int64_t F, void* stream) {
  B200_CHECK_ARG(a && b && dh && da && db && M > 0 && F > 0, "swiglu_bwd: null pointer or empty");
  B200_CHECK_ARG(F % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldda % 8 == 0 && lddb % 8 == 0,
                 "swiglu_bwd: F and all pitches must be multiples of 8");
  B200_CHECK_ARG(b200_aligned16(a) && b200_aligned16(b) && b200_aligned16(dh) && b200_aligned16(da) && b200_aligned16(db), "swiglu_bwd: pointers must be 16-B aligned");
  swiglu_bwd_kernel<<<row_grid(M, (int)(F / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)a, lda, (const bf16*)b, ldb, (const bf16*)dh, (bf16*)da, ldda, (bf16*)db, lddb, M, (int)F);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_gelu_fwd_bf16(const void* x, void* y, int64_t n, void* stream) {
  B200_CHECK_ARG(x && y && n > 0, "gelu_fwd: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(x) && b200_aligned16(y), "gelu_fwd: pointers must be 16-B aligned");
  gelu_fwd_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)y, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_gelu_bwd_bf16(const void* x, const void* dy, void* dx, int64_t n, void* stream) {
  B200_CHECK_ARG(x && dy && dx && n > 0, "gelu_bwd: null pointer or empty");
  gelu_bwd_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)x, (const bf16*)dy, (bf16*)dx, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_rope_bf16(const void* x, int64_t ldx, void* y, int64_t ldy, const void* cos,
                              const void* sin, int64_t B, int64_t T, int64_t nheads, int64_t hd,
                              int inverse, int bf16_steps, void* stream) {
  B200_CHECK_ARG(x && y && cos && sin && B > 0 && T > 0 && nheads > 0, "rope: null pointer or empty");
  B200_CHECK_ARG(hd % 16 == 0 && ldx % 8 == 0 && ldy % 8 == 0 && ldx >= nheads * hd && ldy >= nheads * hd,
                 "rope: hd must be a multiple of 16, pitches multiples of 8 and >= nheads*hd");
  B200_CHECK_ARG(b200_aligned16(x) && b200_aligned16(y) && b200_aligned16(cos) && b200_aligned16(sin), "rope: pointers must be 16-B aligned");
  const int per_tok = (int)(nheads * (hd / 16));
  int bx = 1;
  while (bx < per_tok && bx < 256) bx <<= 1;           // power of two <= 256 covering a token's vectors
  const dim3 block((unsigned)bx, (unsigned)(256 / bx));
  long long blocks = (B * T + block.y - 1) / block.y;
  const long long cap = (long long)b200_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  rope_kernel<<<(unsigned)blocks, block, 0, (cudaStream_t)stream>>>(
      (const bf16*)x, ldx, (bf16*)y, ldy, (const bf16*)cos, (const bf16*)sin, B, (int)T, (int)nheads, (int)hd, inverse, bf16_steps);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_embedding_fwd_bf16(const int64_t* idx, const void* tok_w, const void* pos_w,
                                       void* out, int64_t M, int64_t T, int64_t d, int64_t vocab,
                                       void* stream) {
  B200_CHECK_ARG(idx && tok_w && out && M > 0 && T > 0 && vocab > 0, "embedding_fwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0, "embedding_fwd: d must be a multiple of 8");
  embedding_fwd_kernel<<<ew_grid(M * (d / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)idx, (const bf16*)tok_w, (const bf16*)pos_w, (bf16*)out, M, (int)T, (int)d, vocab);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_embedding_bwd_f32acc(const int64_t* idx, const void* dout, float* dtok_w,
                                         float* dpos_w, int64_t M, int64_t T, int64_t d,
                                         int64_t vocab, void* stream) {
  B200_CHECK_ARG(idx && dout && dtok_w && M > 0 && T > 0 && vocab > 0, "embedding_bwd: null pointer or empty");
  B200_CHECK_ARG(d % 8 == 0 && d > 0, "embedding_bwd: d must be a multiple of 8");
  embedding_bwd_kernel<<<ew_grid(M * (d / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)idx, (const bf16*)dout, dtok_w, dpos_w, M, (int)T, (int)d, vocab);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream) {
  B200_CHECK_ARG(in && out && n > 0, "cast: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(in) && b200_aligned16(out), "cast: pointers must be 16-B aligned");
  cast_f32_bf16_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(in, (bf16*)out, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_cast_bf16_to_f32(const void* in, float* out, int64_t n, void* stream) {
  B200_CHECK_ARG(in && out && n > 0, "cast: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(in) && b200_aligned16(out), "cast: pointers must be 16-B aligned");
  cast_bf16_f32_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)in, out, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_add_bf16(const void* a, const void* b, void* out, int64_t n, void* stream) {
  B200_CHECK_ARG(a && b && out && n > 0, "add: null pointer or empty");
  B200_CHECK_ARG(b200_aligned16(a) && b200_aligned16(b) && b200_aligned16(out), "add: pointers must be 16-B aligned");
  add_kernel<<<ew_grid(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)a, (const bf16*)b, (bf16*)out, n);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_transpose_bf16(const void* in, int64_t ldi, void* out, int64_t ldo, int64_t rows,
                                   int64_t cols, float alpha, void* stream) {
  B200_CHECK_ARG(in && out && rows > 0 && cols > 0, "transpose: null pointer or empty");
  B200_CHECK_ARG(ldi >= cols && ldo >= rows, "transpose: pitches too small");
  dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
  B200_CHECK_ARG(grid.y <= 65535, "transpose: too many rows for one launch");
  transpose_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)in, ldi, (bf16*)out, ldo, (int)rows, (int)cols, alpha);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

// ======================================================================================
// Data side of the step on the device, integer work, bit-exact:
//   collate: datautils/dataloader_instruction_finetune.py:10-50 on a packed token buffer;
//   windows: datautils/dataset.py:21-34 (input = tokens[i : i+L], target = tokens[i+1 : i+L+1]).
// ======================================================================================
namespace {

// One CTA per sample. `width` = longest item of the batch + 1 (the appended end-of-text token) - 1 after
// the shift, i.e. the untruncated row length of inputs/targets; out rows hold min(width, max_len) columns.
__global__ void __launch_bounds__(256)
collate_kernel(const long long* __restrict__ tokens, const long long* __restrict__ offsets,
               const long long* __restrict__ instr_len, long long pad_id, long long ignore_index,
               int width, int out_cols, long long* __restrict__ inputs, long long* __restrict__ targets) {
  __shared__ int s_first;
  const int r = blockIdx.x;
  const long long* item = tokens + offsets[r];
  const int len = (int)(offsets[r + 1] - offsets[r]);
  if (threadIdx.x == 0) s_first = 0x7fffffff;
  __syncthreads();
  // padded[j] = item[j] for j < len, pad_id otherwise; target column j is padded[j + 1]
  int first = 0x7fffffff;
  for (int j = threadIdx.x; j < width; j += blockDim.x) {
    const long long t = (j + 1 < len) ? item[j + 1] : pad_id;
    if (t == pad_id) {
      first = j;
      break;                       // columns visited by a thread are increasing: its first hit is its minimum
    }
  }
  if (first != 0x7fffffff) atomicMin(&s_first, first);
  __syncthreads();
  const int first_pad = s_first;
  long long il = instr_len[r] - 1;          // targets[:instruction_length - 1] with Python slice semantics
  if (il < 0) il = (long long)width + il;
  if (il < 0) il = 0;
  if (il > width) il = width;
  for (int j = threadIdx.x; j < out_cols; j += blockDim.x) {
    const long long in = (j < len) ? item[j] : pad_id;
    long long t = (j + 1 < len) ? item[j + 1] : pad_id;
    if (t == pad_id && j != first_pad) t = ignore_index;
    if (j < il) t = ignore_index;
    inputs[(long long)r * out_cols + j] = in;
    targets[(long long)r * out_cols + j] = t;
  }
}

__global__ void __launch_bounds__(256)
window_batch_kernel(const long long* __restrict__ tokens, const long long* __restrict__ starts, int L,
                    long long* __restrict__ inputs, long long* __restrict__ targets) {
  const long long s = starts[blockIdx.x];
  for (int j = threadIdx.x; j < L; j += blockDim.x) {
    inputs[(long long)blockIdx.x * L + j] = tokens[s + j];
    targets[(long long)blockIdx.x * L + j] = tokens[s + j + 1];
  }
}

}  // namespace

extern "C" int b200_collate_i64(const int64_t* tokens, const int64_t* offsets, const int64_t* instr_len,
                                int64_t batch, int64_t width, int64_t out_cols, int64_t pad_id,
                                int64_t ignore_index, int64_t* inputs, int64_t* targets, void* stream) {
  B200_CHECK_ARG(tokens && offsets && instr_len && inputs && targets, "collate: null pointer");
  B200_CHECK_ARG(batch > 0 && batch <= 65535 * 32 && width > 0 && out_cols > 0 && out_cols <= width && width < (1ll << 30),
                 "collate: bad sizes batch=%lld width=%lld out_cols=%lld", (long long)batch, (long long)width, (long long)out_cols);
  collate_kernel<<<(unsigned)batch, 256, 0, (cudaStream_t)stream>>>(
      (const long long*)tokens, (const long long*)offsets, (const long long*)instr_len, (long long)pad_id,
      (long long)ignore_index, (int)width, (int)out_cols, (long long*)inputs, (long long*)targets);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}

extern "C" int b200_window_batch_i64(const int64_t* tokens, int64_t n_tokens, const int64_t* starts, int64_t batch,
                                     int64_t L, int64_t* inputs, int64_t* targets, void* stream) {
  B200_CHECK_ARG(tokens && starts && inputs && targets && batch > 0 && L > 0 && n_tokens > L && L < (1ll << 30),
                 "window_batch: null pointer or bad sizes (every start must satisfy start + L < n_tokens)");
  window_batch_kernel<<<(unsigned)batch, 256, 0, (cudaStream_t)stream>>>(
      (const long long*)tokens, (const long long*)starts, (int)L, (long long*)inputs, (long long*)targets);
  B200_CHECK_LAUNCH();
  b200_count_launch();
  return 0;
}
