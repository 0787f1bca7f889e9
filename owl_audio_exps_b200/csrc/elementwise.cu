// Memory-bound kernels of the DiT block: every one is a single pass over its tensors with 128-bit
// loads/stores, fp32 math in registers and warp-shuffle row reductions.  Roofline = HBM.
//
// Row kernels use one warp per row; lane `l` owns the 8-element (16-byte) column groups
// c = (i*32 + l)*8, i = 0..NV-1, so every warp-wide access is a contiguous 512-byte run.
#include <cuda_bf16.h>

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

static constexpr int kRowThreads = 128;  // 4 warps = 4 rows per CTA pass (register-heavy row kernels: more CTAs fit an SM)

__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void stg16(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------
// cast / silu
// ------------------------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int64_t n) {
  pdl_prologue();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t n8 = n >> 3;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
    uint4 o;
    o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
    stg16(dst + 8 * i, o);
  }
  // tail (n % 8 elements)
  for (int64_t i = (n8 << 3) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

__global__ void silu_bf16_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int64_t n) {
  pdl_prologue();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t n8 = n >> 3;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float f[8];
    unpack8(ldg16(x + 8 * i), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = f[j] * sigmoidf_fast(f[j]);
    stg16(y + 8 * i, pack8(f));
  }
  for (int64_t i = (n8 << 3) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = __bfloat162float(x[i]);
    y[i] = __float2bfloat16_rn(v * sigmoidf_fast(v));
  }
}

// dz = dy * silu'(z)   (backward of F.silu when it is not fused into a GEMM epilogue)
__global__ void dsilu_mul_bf16_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z, bf16* __restrict__ dz,
                                      int64_t n) {
  pdl_prologue();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t n8 = n >> 3;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float g[8], f[8];
    unpack8(ldg16(dy + 8 * i), g);
    unpack8(ldg16(z + 8 * i), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = sigmoidf_fast(f[j]);
      g[j] *= s * (1.f + f[j] * (1.f - s));
    }
    stg16(dz + 8 * i, pack8(g));
  }
  for (int64_t i = (n8 << 3) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = __bfloat162float(z[i]);
    const float s = sigmoidf_fast(v);
    dz[i] = __float2bfloat16_rn(__bfloat162float(dy[i]) * s * (1.f + v * (1.f - s)));
  }
}

// Several fp32 -> bf16 casts in ONE launch (a block's weight matrices are separate nn.Parameters; eight launches of
// 1-4 us each per layer were mostly launch gaps).  Segment i owns blocks [blk0[i], blk0[i+1]).
struct CastSegs {
  const float* src[8];
  bf16* dst[8];
  long long n8[8];   // elements / 8
  int blk0[9];
  int count;
};
__global__ void __launch_bounds__(256) cast_multi_kernel(const CastSegs s) {
  pdl_prologue();
  int i = 0;
  while (i + 1 < s.count && static_cast<int>(blockIdx.x) >= s.blk0[i + 1]) ++i;
  const int local = static_cast<int>(blockIdx.x) - s.blk0[i];
  const long long stride = static_cast<long long>(s.blk0[i + 1] - s.blk0[i]) * blockDim.x;
  const float4* src = reinterpret_cast<const float4*>(s.src[i]);
  bf16* dst = s.dst[i];
  long long j = static_cast<long long>(local) * blockDim.x + threadIdx.x;
  // four independent 32-byte reads in flight per thread (one resident wave of CTAs covers the whole launch: the weights of
  // a block are 75 MB at audio.yml's width and this kernel sits on the step's critical path 17 times)
  for (; j + 3 * stride < s.n8[i]; j += 4 * stride) {
    float4 a[4], b[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      a[u] = __ldg(src + 2 * (j + u * stride));
      b[u] = __ldg(src + 2 * (j + u * stride) + 1);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint4 o;
      o.x = pack_bf16(a[u].x, a[u].y); o.y = pack_bf16(a[u].z, a[u].w);
      o.z = pack_bf16(b[u].x, b[u].y); o.w = pack_bf16(b[u].z, b[u].w);
      stg16(dst + 8 * (j + u * stride), o);
    }
  }
  for (; j < s.n8[i]; j += stride) {
    const float4 a = __ldg(src + 2 * j);
    const float4 b = __ldg(src + 2 * j + 1);
    uint4 o;
    o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
    stg16(dst + 8 * j, o);
  }
}

// one-thread kernel that stores the GPU's global nanosecond timer: in-graph timeline probe (tools/step_trace.py)
__global__ void timestamp_kernel(unsigned long long* dst) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  *dst = t;
}

static inline int grid_for(int64_t work_items, int threads) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

// ------------------------------------------------------------------------------------------------
// adaLN modulate forward: h = rms_norm(x) * (1 + a[frame]) + b[frame]      (a, b may be NULL)
// ------------------------------------------------------------------------------------------------
// RES: the residual variant (x <- x + gate * y first).  A template parameter, not a null test: the plain variant must not
// carry the y / gate register arrays (ncu r02: 12 % warps active with them, registers bound the occupancy).
template <int NV, bool RES>
__global__ void __launch_bounds__(kRowThreads)
adaln_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ a, const bf16* __restrict__ b, int64_t ld_ab,
                 bf16* __restrict__ h, float* __restrict__ rstd_out, int64_t M, int d, int T, float eps,
                 const bf16* __restrict__ y, const bf16* __restrict__ gate, int64_t ld_g, bf16* __restrict__ xsum) {
  pdl_prologue();
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < M; row += nwarps) {
    const bf16* xr = x + row * d;
    const int64_t frow = row / T;
    // every load of the row is issued before the first use: NV (x) to 5 NV (x, y, gate, a, b) independent 16-byte
    // requests per lane in flight (one row per warp at a time otherwise leaves HBM idle: ncu r02, 34 % warps active)
    uint4 xv[NV], yv[RES ? NV : 1], gtv[RES ? NV : 1], av[NV], bv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        xv[i] = ldg16(xr + c);
        if (RES) {
          yv[RES ? i : 0] = ldg16(y + row * d + c);
          gtv[RES ? i : 0] = ldg16(gate + frow * ld_g + c);
        }
        if (a != nullptr) {
          av[i] = ldg16(a + frow * ld_ab + c);
          bv[i] = ldg16(b + frow * ld_ab + c);
        }
      }
    }
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8];
        unpack8(xv[i], f);
        if (RES) {  // x <- x + gate[frame] * y, the residual of the previous branch (attn.py:135,141)
          float fy[8], fg[8];
          unpack8(yv[RES ? i : 0], fy);
          unpack8(gtv[RES ? i : 0], fg);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = fmaf(fg[j], fy[j], f[j]);
          xv[i] = pack8(f);  // the residual stream is bf16: normalise exactly what is stored
          stg16(xsum + row * d + c, xv[i]);
          unpack8(xv[i], f);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += f[j] * f[j];
      }
    }
    ss = warp_sum(ss);
    const float rstd = rsqrtf(ss / static_cast<float>(d) + eps);
    if (rstd_out != nullptr && lane == 0) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8];
        unpack8(xv[i], f);
        if (a != nullptr) {
          float fa[8], fb[8];
          unpack8(av[i], fa);
          unpack8(bv[i], fb);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = f[j] * rstd * (1.f + fa[j]) + fb[j];
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = f[j] * rstd;
        }
        stg16(h + row * d + c, pack8(f));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// adaLN modulate backward, row part:
//   g = dh * (1 + a);  xh = x * rstd;  dx = rstd * (g - xh * mean(g * xh)) (+ dres)
//   if T == 1 also da = dh * xh, db = dh (a frame is a row; no reduction needed)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
adaln_bwd_row_kernel(const bf16* __restrict__ dh, const bf16* __restrict__ x, const bf16* __restrict__ a,
                     int64_t ld_ab, const float* __restrict__ rstd_in, const bf16* __restrict__ dres,
                     bf16* __restrict__ dx, bf16* __restrict__ da, bf16* __restrict__ db, int64_t ld_dab, int64_t M,
                     int d, int T) {
  pdl_prologue();
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const bool write_ab = (T == 1) && (da != nullptr);
  for (int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < M; row += nwarps) {
    const float rstd = rstd_in[row];
    const int64_t frow = row / T;
    uint4 xv[NV], gv[NV], av[NV], rv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {   // all loads of the row first (see adaln_fwd_kernel)
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        xv[i] = ldg16(x + row * d + c);
        gv[i] = ldg16(dh + row * d + c);
        if (a != nullptr) av[i] = ldg16(a + frow * ld_ab + c);
        if (dres != nullptr) rv[i] = ldg16(dres + row * d + c);
      }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float fx[8], fg[8];
        unpack8(xv[i], fx);
        unpack8(gv[i], fg);
        if (write_ab) {
          float t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = fg[j] * fx[j] * rstd;
          stg16(da + frow * ld_dab + c, pack8(t));
          stg16(db + frow * ld_dab + c, gv[i]);
        }
        if (a != nullptr) {
          float fa[8];
          unpack8(av[i], fa);
#pragma unroll
          for (int j = 0; j < 8; ++j) fg[j] *= (1.f + fa[j]);
          gv[i] = pack8(fg);  // keep g (bf16-rounded) for the second pass
          unpack8(gv[i], fg);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += fg[j] * fx[j] * rstd;
      }
    }
    s = warp_sum(s) / static_cast<float>(d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float fx[8], fg[8], o[8];
        unpack8(xv[i], fx);
        unpack8(gv[i], fg);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (fg[j] - fx[j] * rstd * s);
        if (dres != nullptr) {
          float fr[8];
          unpack8(rv[i], fr);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += fr[j];
        }
        stg16(dx + row * d + c, pack8(o));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// adaLN backward fused with the gate backward of the branch that produced this norm's input, one pass over the
// residual-stream gradient (block: x_in = x_prev + c * y;  h = rms_norm(x_in) * (1 + a) + b):
//   dx   = rstd * (g - xh * mean(g * xh)) + dres        g = dh * (1 + a), xh = x_in * rstd     (as adaln_bwd_row_kernel)
//   dy   = dx * c[frame]                                the gradient of the branch output y
//   dc   = dx * y   (T == 1; frames of T > 1 rows are reduced by frame_reduce_kernel afterwards)
//   dbias[n] += sum_m dy[m, n]                          bias gradient of the Linear that produced y (COLSUM only)
// Replaces two launches (adaLN backward, gate backward + column sum) and one re-read of dx per branch.
// ------------------------------------------------------------------------------------------------
template <int NV, bool COLSUM>
__global__ void __launch_bounds__(kRowThreads)
adaln_gate_bwd_kernel(const bf16* __restrict__ dh, const bf16* __restrict__ x, const bf16* __restrict__ a,
                      int64_t ld_ab, const float* __restrict__ rstd_in, const bf16* __restrict__ dres,
                      bf16* __restrict__ dx, bf16* __restrict__ da, bf16* __restrict__ db, int64_t ld_dab,
                      const bf16* __restrict__ y, const bf16* __restrict__ c, int64_t ld_c, bf16* __restrict__ dy,
                      bf16* __restrict__ dc, int64_t ld_dc, float* __restrict__ dbias, int64_t M, int d, int T) {
  pdl_prologue();
  __shared__ float s_col[COLSUM ? NV * 256 : 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const bool write_ab = (T == 1) && (da != nullptr);
  const bool write_dc = (T == 1) && (dc != nullptr);
  float acc[COLSUM ? NV : 1][8];
  if (COLSUM) {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  }
  for (int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < M; row += nwarps) {
    const float rstd = rstd_in[row];
    const int64_t frow = row / T;
    uint4 xv[NV], gv[NV], av[NV], rv[NV], cv[NV], yv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {   // all loads of the row first (see adaln_fwd_kernel): up to 6 NV requests per lane in flight
      const int col = (i * 32 + lane) * 8;
      if (col < d) {
        xv[i] = ldg16(x + row * d + col);
        gv[i] = ldg16(dh + row * d + col);
        if (a != nullptr) av[i] = ldg16(a + frow * ld_ab + col);
        if (dres != nullptr) rv[i] = ldg16(dres + row * d + col);
        cv[i] = ldg16(c + frow * ld_c + col);
        if (write_dc) yv[i] = ldg16(y + row * d + col);
      }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (i * 32 + lane) * 8;
      if (col < d) {
        float fx[8], fg[8];
        unpack8(xv[i], fx);
        unpack8(gv[i], fg);
        if (write_ab) {
          float t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = fg[j] * fx[j] * rstd;
          stg16(da + frow * ld_dab + col, pack8(t));
          stg16(db + frow * ld_dab + col, gv[i]);
        }
        if (a != nullptr) {
          float fa[8];
          unpack8(av[i], fa);
#pragma unroll
          for (int j = 0; j < 8; ++j) fg[j] *= (1.f + fa[j]);
          gv[i] = pack8(fg);
          unpack8(gv[i], fg);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += fg[j] * fx[j] * rstd;
      }
    }
    s = warp_sum(s) / static_cast<float>(d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (i * 32 + lane) * 8;
      if (col < d) {
        float fx[8], fg[8], o[8];
        unpack8(xv[i], fx);
        unpack8(gv[i], fg);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (fg[j] - fx[j] * rstd * s);
        if (dres != nullptr) {
          float fr[8];
          unpack8(rv[i], fr);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += fr[j];
        }
        const uint4 dxp = pack8(o);
        stg16(dx + row * d + col, dxp);
        unpack8(dxp, o);  // the gate backward sees the bf16 residual gradient, exactly as the two-kernel path does
        float fc[8], g[8];
        unpack8(cv[i], fc);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = o[j] * fc[j];
        const uint4 dyp = pack8(g);
        stg16(dy + row * d + col, dyp);
        if (COLSUM) {
          unpack8(dyp, g);  // column sum over the bf16 values the next GEMM reads
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] += g[j];
        }
        if (write_dc) {
          float fy[8];
          unpack8(yv[i], fy);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = o[j] * fy[j];
          stg16(dc + frow * ld_dc + col, pack8(g));
        }
      }
    }
  }
  if (COLSUM) {
    // the 8 warps of the CTA add their partial column sums into one shared row, then one fp32 atomic per column per CTA
    for (int w = 0; w < kRowThreads / 32; ++w) {
      if (warp == w) {
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int idx = (i * 32 + lane) * 8 + j;
            s_col[idx] = (w == 0 ? 0.f : s_col[idx]) + acc[i][j];
          }
      }
      __syncthreads();
    }
    for (int idx = threadIdx.x; idx < d; idx += kRowThreads) atomicAdd(dbias + idx, s_col[idx]);
  }
}

// ------------------------------------------------------------------------------------------------
// Per-frame column reductions (T > 1):
//   out_pq[f, c] = sum_{r in frame f} p[r, c] * q[r, c] * (scale ? scale[r] : 1)
//   out_p [f, c] = sum_{r in frame f} p[r, c]                       (optional)
// grid = (frames, ceil(d / 512)); 64 threads x 8 columns.
// ------------------------------------------------------------------------------------------------
static constexpr int kFrLanes = 8;  // row lanes per CTA: the T rows of a frame are split 8 ways, then summed in smem

__global__ void __launch_bounds__(64 * kFrLanes)
frame_reduce_kernel(const bf16* __restrict__ p, const bf16* __restrict__ q, const float* __restrict__ scale,
                    bf16* __restrict__ out_pq, bf16* __restrict__ out_p, int64_t ld_out, int d, int T) {
  pdl_prologue();
  __shared__ float red[kFrLanes - 1][64][17];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.y * 64 + tx) * 8;
  const bool act = c < d;
  const int64_t f = blockIdx.x;
  float apq[8], ap[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) apq[j] = ap[j] = 0.f;
  const int64_t r0 = f * T;
  if (act) {
#pragma unroll 4
    for (int r = ty; r < T; r += kFrLanes) {
      float fp[8], fq[8];
      unpack8(ldg16(p + (r0 + r) * d + c), fp);
      unpack8(ldg16(q + (r0 + r) * d + c), fq);
      const float s = scale != nullptr ? __ldg(scale + r0 + r) : 1.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        apq[j] += fp[j] * fq[j] * s;
        ap[j] += fp[j];
      }
    }
  }
  if (ty > 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      red[ty - 1][tx][j] = apq[j];
      red[ty - 1][tx][8 + j] = ap[j];
    }
  }
  __syncthreads();
  if (ty == 0 && act) {
#pragma unroll
    for (int l = 0; l < kFrLanes - 1; ++l) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        apq[j] += red[l][tx][j];
        ap[j] += red[l][tx][8 + j];
      }
    }
    stg16(out_pq + f * ld_out + c, pack8(apq));
    if (out_p != nullptr) stg16(out_p + f * ld_out + c, pack8(ap));
  }
}

// ------------------------------------------------------------------------------------------------
// gate forward (standalone Gate module / cond_gate): out = c[frame] * x (+ res)
// gate backward, elementwise part: dy = dx * c[frame]; if T == 1 also dc = dx * y
// Flat grid-stride over 16-byte groups.
// ------------------------------------------------------------------------------------------------
__global__ void gate_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ c, int64_t ld_c,
                                const bf16* __restrict__ res, bf16* __restrict__ out, int64_t M, int d, int T) {
  pdl_prologue();
  const int dv = d >> 3;
  const int64_t total = M * dv;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t row = i / dv;
    const int col = static_cast<int>(i - row * dv) * 8;
    float fx[8], fc[8];
    unpack8(ldg16(x + row * d + col), fx);
    unpack8(ldg16(c + (row / T) * ld_c + col), fc);
#pragma unroll
    for (int j = 0; j < 8; ++j) fx[j] *= fc[j];
    if (res != nullptr) {
      float fr[8];
      unpack8(ldg16(res + row * d + col), fr);
#pragma unroll
      for (int j = 0; j < 8; ++j) fx[j] += fr[j];
    }
    stg16(out + row * d + col, pack8(fx));
  }
}

__global__ void gate_bwd_kernel(const bf16* __restrict__ dx, const bf16* __restrict__ y, const bf16* __restrict__ c,
                                int64_t ld_c, bf16* __restrict__ dy, bf16* __restrict__ dc, int64_t ld_dc, int64_t M,
                                int d, int T) {
  pdl_prologue();
  const int dv = d >> 3;
  const int64_t total = M * dv;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const bool write_dc = (T == 1) && (dc != nullptr);
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t row = i / dv;
    const int col = static_cast<int>(i - row * dv) * 8;
    float fd[8], fc[8], o[8];
    unpack8(ldg16(dx + row * d + col), fd);
    unpack8(ldg16(c + (row / T) * ld_c + col), fc);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fd[j] * fc[j];
    stg16(dy + row * d + col, pack8(o));
    if (write_dc) {
      float fy[8];
      unpack8(ldg16(y + row * d + col), fy);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fd[j] * fy[j];
      stg16(dc + row * ld_dc + col, pack8(o));
    }
  }
}

// gate backward with the bias gradient of the producing Linear folded in (one pass over dx instead of two kernels):
//   dy = dx * c[frame];  dbias[n] += sum_m dy[m, n];  if T == 1 also dc = dx * y.
// Same block shape as colsum_kernel: 32 column groups (256 columns) x 8 row lanes, grid = (ceil(d/256), row_splits).
__global__ void __launch_bounds__(256)
gate_bwd_colsum_kernel(const bf16* __restrict__ dx, const bf16* __restrict__ y, const bf16* __restrict__ c,
                       int64_t ld_c, bf16* __restrict__ dy, bf16* __restrict__ dc, int64_t ld_dc,
                       float* __restrict__ dbias, int64_t M, int d, int T) {
  pdl_prologue();
  __shared__ float part[8][256 + 8];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cg) * 8;
  const bool write_dc = (T == 1) && (dc != nullptr);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (col < d) {
#pragma unroll 4
    for (int64_t r = static_cast<int64_t>(blockIdx.y) * 8 + rl; r < M; r += static_cast<int64_t>(gridDim.y) * 8) {
      float fd[8], fc[8], o[8];
      unpack8(ldg16(dx + r * d + col), fd);
      unpack8(ldg16(c + (r / T) * ld_c + col), fc);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fd[j] * fc[j];
      const uint4 packed = pack8(o);
      stg16(dy + r * d + col, packed);
      unpack8(packed, o);  // the column sum is taken over the bf16 values the next GEMM will read
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += o[j];
      if (write_dc) {
        float fy[8];
        unpack8(ldg16(y + r * d + col), fy);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fd[j] * fy[j];
        stg16(dc + r * ld_dc + col, pack8(o));
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[rl][cg * 8 + j] = acc[j];
  __syncthreads();
  const int cc = threadIdx.x;
  if (blockIdx.x * 256 + cc < d) {
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) sum += part[i][cc];
    atomicAdd(dbias + blockIdx.x * 256 + cc, sum);
  }
}

// ------------------------------------------------------------------------------------------------
// column sum: out[n] (+)= sum_m x[m, n].  Block = 32 column groups (256 columns) x 8 row lanes;
// grid = (ceil(N/256), row_splits).  Partial sums meet in shared memory, then one fp32 atomic per
// column per CTA (the output is zeroed by the launcher unless accumulating).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_kernel(const bf16* __restrict__ x, int64_t ld, float* __restrict__ out, int64_t M, int N) {
  pdl_prologue();
  __shared__ float part[8][256 + 8];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + cg) * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < N) {
    for (int64_t r = static_cast<int64_t>(blockIdx.y) * 8 + rl; r < M; r += static_cast<int64_t>(gridDim.y) * 8) {
      float f[8];
      unpack8(ldg16(x + r * ld + c), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[rl][cg * 8 + j] = acc[j];
  __syncthreads();
  const int col = threadIdx.x;  // 256 threads -> 256 columns
  if (blockIdx.x * 256 + col < N) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) s += part[r][col];
    atomicAdd(out + blockIdx.x * 256 + col, s);
  }
}

// Column sums of up to 4 matrices in ONE launch (the four bias gradients of a DiT block: fc2, fc1, out, qkv).  Segment i owns
// blocks [blk0[i], blk0[i+1]) laid out as gx[i] column groups (256 columns) x row splits; outputs are ADDED to (the caller
// zeroes them: engine.GradBuckets zeroes every bias sink once per step).
struct ColsumSegs {
  const bf16* x[4];
  float* out[4];
  long long ld[4];
  long long M[4];
  int N[4];
  int gx[4];
  int blk0[5];
  int count;
};
__global__ void __launch_bounds__(256) colsum_multi_kernel(const ColsumSegs s) {
  pdl_prologue();
  __shared__ float part[8][256 + 8];
  int i = 0;
  while (i + 1 < s.count && static_cast<int>(blockIdx.x) >= s.blk0[i + 1]) ++i;
  const int local = static_cast<int>(blockIdx.x) - s.blk0[i];
  const int gx = s.gx[i];
  const int bx = local % gx, by = local / gx;
  const int splits = (s.blk0[i + 1] - s.blk0[i]) / gx;
  const bf16* x = s.x[i];
  const long long ld = s.ld[i], M = s.M[i];
  const int N = s.N[i];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = (bx * 32 + cg) * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < N) {
#pragma unroll 4
    for (long long r = static_cast<long long>(by) * 8 + rl; r < M; r += static_cast<long long>(splits) * 8) {
      float f[8];
      unpack8(ldg16(x + r * ld + c), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[rl][cg * 8 + j] = acc[j];
  __syncthreads();
  const int col = threadIdx.x;
  if (bx * 256 + col < N) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += part[r][col];
    atomicAdd(s.out[i] + bx * 256 + col, t);
  }
}

// ------------------------------------------------------------------------------------------------
// KV-cache append: copy [B,H,Ln,dh] rows into cache[B,H,cap,dh] at row `pos` (16-byte groups).
// ------------------------------------------------------------------------------------------------
__global__ void kv_append_kernel(const bf16* __restrict__ src, int64_t src_sb, int64_t src_sh,
                                 bf16* __restrict__ cache, int64_t c_sb, int64_t c_sh, int64_t pos, int64_t B, int H,
                                 int64_t Ln, int dh) {
  pdl_prologue();
  const int64_t per_head = Ln * dh / 8;
  const int64_t total = B * H * per_head;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t bh = i / per_head, e = (i - bh * per_head) * 8;
    const int64_t b = bh / H, h = bh - b * H;
    stg16(cache + b * c_sb + h * c_sh + pos * dh + e, ldg16(src + b * src_sb + h * src_sh + e));
  }
}

// ------------------------------------------------------------------------------------------------
// weight-less LayerNorm forward (normalization.py:6-7)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
layer_norm_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int64_t M, int d, float eps) {
  pdl_prologue();
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < M; row += nwarps) {
    uint4 xv[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        xv[i] = ldg16(x + row * d + c);
        float f[8];
        unpack8(xv[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += f[j];
      }
    }
    const float mean = warp_sum(s) / static_cast<float>(d);
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8];
        unpack8(xv[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) v += (f[j] - mean) * (f[j] - mean);
      }
    }
    const float rstd = rsqrtf(warp_sum(v) / static_cast<float>(d) + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8];
        unpack8(xv[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = (f[j] - mean) * rstd;
        stg16(y + row * d + c, pack8(f));
      }
    }
  }
}

// NV dispatch: NV*256 >= d.  d up to 4096.
#define OWL_DISPATCH_NV(d, ...)                         \
  do {                                                  \
    const int nv_ = ((d) + 255) / 256;                  \
    if (nv_ <= 1) { constexpr int NV = 1; __VA_ARGS__; }        \
    else if (nv_ <= 2) { constexpr int NV = 2; __VA_ARGS__; }   \
    else if (nv_ <= 4) { constexpr int NV = 4; __VA_ARGS__; }   \
    else if (nv_ <= 6) { constexpr int NV = 6; __VA_ARGS__; }   \
    else if (nv_ <= 8) { constexpr int NV = 8; __VA_ARGS__; }   \
    else if (nv_ <= 10) { constexpr int NV = 10; __VA_ARGS__; } \
    else { constexpr int NV = 16; __VA_ARGS__; }                \
  } while (0)

// Sin/cos feature row of one scalar per row (embeddings.py:30-72 SinCosEmbed): out[r] = [sin(t_r*mult*f_i) | cos(t_r*mult*f_i)],
// i < half.  The argument is formed in fp32 with the reference's two roundings (t*mult, then *f_i); the file is built with fast
// math, so sin/cos go through the double-precision routines (arguments reach ~1e3 rad, far outside __sinf's accurate range).
__global__ void __launch_bounds__(256) sincos_embed_kernel(const float* __restrict__ t, const float* __restrict__ freqs,
                                                          bf16* __restrict__ out, int64_t n_rows, int half, float mult) {
  pdl_prologue();
  const int64_t total = n_rows * half;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / half;
    const int c = static_cast<int>(i - r * half);
    const float arg = __fmul_rn(__fmul_rn(t[r], mult), freqs[c]);
    double sn, cs;
    sincos(static_cast<double>(arg), &sn, &cs);
    out[r * (2 * half) + c] = __float2bfloat16(static_cast<float>(sn));
    out[r * (2 * half) + half + c] = __float2bfloat16(static_cast<float>(cs));
  }
}

// weight-less LayerNorm backward: dx = rstd * (dy - mean(dy) - xhat * mean(dy * xhat)); mean / rstd are recomputed from x
// (one more pass over registers, no saved statistics).
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
layer_norm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, bf16* __restrict__ dx, int64_t M, int d, float eps) {
  pdl_prologue();
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const float inv_d = 1.f / static_cast<float>(d);
  for (int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < M; row += nwarps) {
    uint4 xv[NV], gv[NV];
    float s = 0.f, sg = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        xv[i] = ldg16(x + row * d + c);
        gv[i] = ldg16(dy + row * d + c);
        float f[8], g[8];
        unpack8(xv[i], f);
        unpack8(gv[i], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s += f[j]; sg += g[j]; }
      }
    }
    const float mean = warp_sum(s) * inv_d;
    const float gmean = warp_sum(sg) * inv_d;
    float v = 0.f, gx = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8], g[8];
        unpack8(xv[i], f);
        unpack8(gv[i], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) { v += (f[j] - mean) * (f[j] - mean); gx += g[j] * (f[j] - mean); }
      }
    }
    const float rstd = rsqrtf(warp_sum(v) * inv_d + eps);
    const float proj = warp_sum(gx) * inv_d * rstd * rstd;   // mean(dy * xhat) * rstd, applied to (x - mean)
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 8;
      if (c < d) {
        float f[8], g[8], o[8];
        unpack8(xv[i], f);
        unpack8(gv[i], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[j] - gmean - (f[j] - mean) * proj);
        stg16(dx + row * d + c, pack8(o));
      }
    }
  }
}

static inline int row_grid(int64_t M) {
  int64_t blocks = (M + (kRowThreads / 32) - 1) / (kRowThreads / 32);
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  return static_cast<int>(blocks < 1 ? 1 : blocks);
}

}  // namespace owl

using namespace owl;

#define OWL_REQUIRE(cond, code) \
  do {                          \
    if (!(cond)) return (code); \
  } while (0)

extern "C" {

int owl_cast_f32_bf16(const float* src, void* dst, int64_t n, owl_stream_t stream) {
  OWL_REQUIRE(src && dst && n > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(aligned16(src) && aligned16(dst), OWL_ERR_ALIGN);
  (void)launch_k(cast_f32_bf16_kernel, dim3(grid_for((n + 7) / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), src, static_cast<bf16*>(dst), n);
  return launch_status();
}

int owl_silu_bf16(const void* x, void* y, int64_t n, owl_stream_t stream) {
  OWL_REQUIRE(x && y && n > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(aligned16(x) && aligned16(y), OWL_ERR_ALIGN);
  (void)launch_k(silu_bf16_kernel, dim3(grid_for((n + 7) / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), static_cast<bf16*>(y), n);
  return launch_status();
}

int owl_sincos_embed(const float* t, const float* freqs, void* out, int64_t n_rows, int half, float mult,
                     owl_stream_t stream) {
  OWL_REQUIRE(t && freqs && out && n_rows > 0 && half > 0, OWL_ERR_BAD_ARG);
  (void)launch_k(sincos_embed_kernel, dim3(grid_for(n_rows * half, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), t, freqs,
                 static_cast<bf16*>(out), n_rows, half, mult);
  return launch_status();
}

int owl_dsilu_mul_bf16(const void* dy, const void* z, void* dz, int64_t n, owl_stream_t stream) {
  OWL_REQUIRE(dy && z && dz && n > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(aligned16(dy) && aligned16(z) && aligned16(dz), OWL_ERR_ALIGN);
  (void)launch_k(dsilu_mul_bf16_kernel, dim3(grid_for((n + 7) / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(dy), static_cast<const bf16*>(z), static_cast<bf16*>(dz), n);
  return launch_status();
}

int owl_adaln_modulate_fwd(const void* x, const void* a, const void* b, int64_t ld_ab, void* h, float* rstd,
                           int64_t M, int64_t d, int64_t T, float eps, owl_stream_t stream) {
  OWL_REQUIRE(x && h && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE((a == nullptr) == (b == nullptr), OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(x) && aligned16(h) && aligned16(a) && aligned16(b) && (a == nullptr || ld_ab % 8 == 0),
              OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  OWL_DISPATCH_NV(d, ((void)launch_k(adaln_fwd_kernel<NV, false>, dim3(row_grid(M)), dim3(kRowThreads), 0, s, static_cast<const bf16*>(x), static_cast<const bf16*>(a), static_cast<const bf16*>(b), ld_ab,
                         static_cast<bf16*>(h), rstd, M, static_cast<int>(d), static_cast<int>(T), eps, nullptr, nullptr,
                         0, nullptr)));
  return launch_status();
}

int owl_residual_adaln_fwd(const void* x, const void* y, const void* gate, int64_t ld_g, const void* a, const void* b,
                           int64_t ld_ab, void* xsum, void* h, float* rstd, int64_t M, int64_t d, int64_t T, float eps,
                           owl_stream_t stream) {
  OWL_REQUIRE(x && y && gate && xsum && h && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE((a == nullptr) == (b == nullptr), OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(x) && aligned16(y) && aligned16(gate) && aligned16(xsum) && aligned16(h) && aligned16(a) &&
                  aligned16(b) && ld_g % 8 == 0 && (a == nullptr || ld_ab % 8 == 0),
              OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  OWL_DISPATCH_NV(d, ((void)launch_k(adaln_fwd_kernel<NV, true>, dim3(row_grid(M)), dim3(kRowThreads), 0, s, static_cast<const bf16*>(x), static_cast<const bf16*>(a), static_cast<const bf16*>(b), ld_ab,
                         static_cast<bf16*>(h), rstd, M, static_cast<int>(d), static_cast<int>(T), eps,
                         static_cast<const bf16*>(y), static_cast<const bf16*>(gate), ld_g, static_cast<bf16*>(xsum))));
  return launch_status();
}

int owl_adaln_modulate_bwd(const void* dh, const void* x, const void* a, int64_t ld_ab, const float* rstd,
                           const void* dres, void* dx, void* da, void* db, int64_t ld_dab, int64_t M, int64_t d,
                           int64_t T, owl_stream_t stream) {
  OWL_REQUIRE(dh && x && rstd && dx && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE((da == nullptr) == (db == nullptr), OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(dh) && aligned16(x) && aligned16(a) && aligned16(dres) && aligned16(dx) && aligned16(da) &&
                  aligned16(db) && (a == nullptr || ld_ab % 8 == 0) && (da == nullptr || ld_dab % 8 == 0),
              OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  OWL_DISPATCH_NV(d, ((void)launch_k(adaln_bwd_row_kernel<NV>, dim3(row_grid(M)), dim3(kRowThreads), 0, s, static_cast<const bf16*>(dh), static_cast<const bf16*>(x), static_cast<const bf16*>(a), ld_ab,
                         rstd, static_cast<const bf16*>(dres), static_cast<bf16*>(dx), static_cast<bf16*>(da),
                         static_cast<bf16*>(db), ld_dab, M, static_cast<int>(d), static_cast<int>(T))));
  int rc = launch_status();
  if (rc != OWL_OK) return rc;
  if (T > 1 && da != nullptr) {
    dim3 grid(static_cast<unsigned>(M / T), static_cast<unsigned>((d + 511) / 512));
    (void)launch_k(frame_reduce_kernel, dim3(grid), dim3(64 * kFrLanes), 0, s, static_cast<const bf16*>(dh), static_cast<const bf16*>(x), rstd,
                                            static_cast<bf16*>(da), static_cast<bf16*>(db), ld_dab,
                                            static_cast<int>(d), static_cast<int>(T));
    rc = launch_status();
  }
  return rc;
}

int owl_gate_fwd(const void* x, const void* c, int64_t ld_c, const void* res, void* out, int64_t M, int64_t d,
                 int64_t T, owl_stream_t stream) {
  OWL_REQUIRE(x && c && out && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(x) && aligned16(c) && aligned16(res) && aligned16(out) && ld_c % 8 == 0, OWL_ERR_ALIGN);
  (void)launch_k(gate_fwd_kernel, dim3(grid_for(M * (d / 8), 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), static_cast<const bf16*>(c), ld_c, static_cast<const bf16*>(res),
      static_cast<bf16*>(out), M, static_cast<int>(d), static_cast<int>(T));
  return launch_status();
}

int owl_colsum_bf16(const void* x, int64_t ld, float* out, int accumulate, int64_t M, int64_t N,
                    owl_stream_t stream) {
  OWL_REQUIRE(x && out && M > 0 && N > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(N % 8 == 0, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(x) && ld % 8 == 0, OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!accumulate) {
    cudaError_t e = cudaMemsetAsync(out, 0, static_cast<size_t>(N) * sizeof(float), s);
    if (e != cudaSuccess) return static_cast<int>(e);
  }
  const unsigned gx = static_cast<unsigned>((N + 255) / 256);
  int64_t splits = (M + 63) / 64;  // >= 8 rows per row-lane
  const int64_t cap = (static_cast<int64_t>(sm_count()) * 4 + gx - 1) / gx;
  if (splits > cap) splits = cap;
  if (splits < 1) splits = 1;
  LaunchOpts o;
  o.pdl = accumulate != 0;  // otherwise the kernel follows the memset node above, not a kernel
  (void)launch_kx(o, colsum_kernel, dim3(gx, static_cast<unsigned>(splits)), dim3(256), 0, s,
                  static_cast<const bf16*>(x), ld, out, M, static_cast<int>(N));
  return launch_status();
}

int owl_gate_bwd(const void* dx, const void* y, const void* c, int64_t ld_c, void* dy, void* dc, int64_t ld_dc,
                 float* dbias, int dbias_accumulate, int64_t M, int64_t d, int64_t T, owl_stream_t stream) {
  OWL_REQUIRE(dx && c && dy && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(dc == nullptr || y != nullptr, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(dx) && aligned16(y) && aligned16(c) && aligned16(dy) && aligned16(dc) && ld_c % 8 == 0 &&
                  (dc == nullptr || ld_dc % 8 == 0),
              OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  if (dbias != nullptr) {
    LaunchOpts o;
    if (!dbias_accumulate) {
      cudaError_t e = cudaMemsetAsync(dbias, 0, static_cast<size_t>(d) * sizeof(float), s);
      if (e != cudaSuccess) return static_cast<int>(e);
      o.pdl = false;  // follows a memset node
    }
    const unsigned gx = static_cast<unsigned>((d + 255) / 256);
    int64_t splits = (M + 31) / 32;
    const int64_t cap = (static_cast<int64_t>(sm_count()) * 8 + gx - 1) / gx;
    if (splits > cap) splits = cap;
    (void)launch_kx(o, gate_bwd_colsum_kernel, dim3(gx, static_cast<unsigned>(splits)), dim3(256), 0, s,
                    static_cast<const bf16*>(dx), static_cast<const bf16*>(y), static_cast<const bf16*>(c), ld_c,
                    static_cast<bf16*>(dy), static_cast<bf16*>(dc), ld_dc, dbias, M, static_cast<int>(d),
                    static_cast<int>(T));
  } else {
    (void)launch_k(gate_bwd_kernel, dim3(grid_for(M * (d / 8), 256)), dim3(256), 0, s, static_cast<const bf16*>(dx),
                   static_cast<const bf16*>(y), static_cast<const bf16*>(c), ld_c, static_cast<bf16*>(dy),
                   static_cast<bf16*>(dc), ld_dc, M, static_cast<int>(d), static_cast<int>(T));
  }
  rc = launch_status();
  if (rc != OWL_OK) return rc;
  if (T > 1 && dc != nullptr) {
    dim3 grid(static_cast<unsigned>(M / T), static_cast<unsigned>((d + 511) / 512));
    (void)launch_k(frame_reduce_kernel, dim3(grid), dim3(64 * kFrLanes), 0, s, static_cast<const bf16*>(dx), static_cast<const bf16*>(y), nullptr,
                                            static_cast<bf16*>(dc), nullptr, ld_dc, static_cast<int>(d),
                                            static_cast<int>(T));
    rc = launch_status();
    if (rc != OWL_OK) return rc;
  }
  return rc;
}

int owl_adaln_gate_bwd(const void* dh, const void* x, const void* a, int64_t ld_ab, const float* rstd,
                       const void* dres, void* dx, void* da, void* db, int64_t ld_dab, const void* y, const void* c,
                       int64_t ld_c, void* dy, void* dc, int64_t ld_dc, float* dbias, int dbias_accumulate, int64_t M,
                       int64_t d, int64_t T, owl_stream_t stream) {
  OWL_REQUIRE(dh && x && rstd && dx && y && c && dy && M > 0 && d > 0 && T > 0 && M % T == 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE((da == nullptr) == (db == nullptr), OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(dh) && aligned16(x) && aligned16(a) && aligned16(dres) && aligned16(dx) && aligned16(da) &&
                  aligned16(db) && aligned16(y) && aligned16(c) && aligned16(dy) && aligned16(dc) &&
                  (a == nullptr || ld_ab % 8 == 0) && (da == nullptr || ld_dab % 8 == 0) && ld_c % 8 == 0 &&
                  (dc == nullptr || ld_dc % 8 == 0),
              OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // per-lane column-sum accumulators cost 8*NV registers: fused up to d = 2048, a separate column sum beyond
  const bool colsum = dbias != nullptr && d <= 2048;
  LaunchOpts o;
  if (colsum && !dbias_accumulate) {
    cudaError_t e = cudaMemsetAsync(dbias, 0, static_cast<size_t>(d) * sizeof(float), s);
    if (e != cudaSuccess) return static_cast<int>(e);
    o.pdl = false;  // follows a memset node
  }
  // with the column sum every CTA ends in d atomics: two CTAs per SM keep that below the row traffic
  int grid = row_grid(M);
  if (colsum && grid > 2 * sm_count()) grid = 2 * sm_count();
#define OWL_AGB_ARGS                                                                                                  \
  static_cast<const bf16*>(dh), static_cast<const bf16*>(x), static_cast<const bf16*>(a), ld_ab, rstd,                \
      static_cast<const bf16*>(dres), static_cast<bf16*>(dx), static_cast<bf16*>(da), static_cast<bf16*>(db), ld_dab, \
      static_cast<const bf16*>(y), static_cast<const bf16*>(c), ld_c, static_cast<bf16*>(dy), static_cast<bf16*>(dc), \
      ld_dc, dbias, M, static_cast<int>(d), static_cast<int>(T)
  if (colsum) {
    const int nv = static_cast<int>((d + 255) / 256);
    if (nv <= 1) (void)launch_kx(o, adaln_gate_bwd_kernel<1, true>, dim3(grid), dim3(kRowThreads), 0, s, OWL_AGB_ARGS);
    else if (nv <= 2) (void)launch_kx(o, adaln_gate_bwd_kernel<2, true>, dim3(grid), dim3(kRowThreads), 0, s, OWL_AGB_ARGS);
    else if (nv <= 4) (void)launch_kx(o, adaln_gate_bwd_kernel<4, true>, dim3(grid), dim3(kRowThreads), 0, s, OWL_AGB_ARGS);
    else if (nv <= 6) (void)launch_kx(o, adaln_gate_bwd_kernel<6, true>, dim3(grid), dim3(kRowThreads), 0, s, OWL_AGB_ARGS);
    else (void)launch_kx(o, adaln_gate_bwd_kernel<8, true>, dim3(grid), dim3(kRowThreads), 0, s, OWL_AGB_ARGS);
  } else {
    OWL_DISPATCH_NV(d, ((void)launch_kx(o, adaln_gate_bwd_kernel<NV, false>, dim3(grid), dim3(kRowThreads), 0, s,
                                        OWL_AGB_ARGS)));
  }
#undef OWL_AGB_ARGS
  int rc = launch_status();
  if (rc != OWL_OK) return rc;
  if (dbias != nullptr && !colsum) {
    rc = owl_colsum_bf16(dy, d, dbias, dbias_accumulate, M, d, stream);
    if (rc != OWL_OK) return rc;
  }
  if (T > 1) {
    dim3 grid2(static_cast<unsigned>(M / T), static_cast<unsigned>((d + 511) / 512));
    if (da != nullptr) {
      (void)launch_k(frame_reduce_kernel, dim3(grid2), dim3(64 * kFrLanes), 0, s, static_cast<const bf16*>(dh),
                     static_cast<const bf16*>(x), rstd, static_cast<bf16*>(da), static_cast<bf16*>(db), ld_dab,
                     static_cast<int>(d), static_cast<int>(T));
      if ((rc = launch_status()) != OWL_OK) return rc;
    }
    if (dc != nullptr) {
      (void)launch_k(frame_reduce_kernel, dim3(grid2), dim3(64 * kFrLanes), 0, s, static_cast<const bf16*>(dx),
                     static_cast<const bf16*>(y), static_cast<const float*>(nullptr), static_cast<bf16*>(dc),
                     static_cast<bf16*>(nullptr), ld_dc, static_cast<int>(d), static_cast<int>(T));
      rc = launch_status();
    }
  }
  return rc;
}

int owl_colsum_bf16_multi(const void* const* xs, const int64_t* lds, const int64_t* Ms, const int64_t* Ns,
                          float* const* outs, int count, owl_stream_t stream) {
  OWL_REQUIRE(xs && lds && Ms && Ns && outs && count > 0 && count <= 4, OWL_ERR_BAD_ARG);
  ColsumSegs s{};
  const long long budget = static_cast<long long>(sm_count()) * 6;  // CTAs over all segments
  long long cells = 0;
  for (int i = 0; i < count; ++i) {
    OWL_REQUIRE(xs[i] && outs[i] && Ms[i] > 0 && Ns[i] > 0, OWL_ERR_BAD_ARG);
    OWL_REQUIRE(Ns[i] % 8 == 0 && Ns[i] <= INT32_MAX, OWL_ERR_UNSUPPORTED);
    OWL_REQUIRE(aligned16(xs[i]) && lds[i] % 8 == 0, OWL_ERR_ALIGN);
    cells += Ms[i] * ((Ns[i] + 255) / 256);
  }
  int blk = 0;
  for (int i = 0; i < count; ++i) {
    const int gx = static_cast<int>((Ns[i] + 255) / 256);
    long long splits = (budget * Ms[i] + cells - 1) / cells;       // CTAs in proportion to the bytes of the segment
    const long long max_splits = (Ms[i] + 31) / 32;                 // >= 4 rows per row-lane
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    s.x[i] = static_cast<const bf16*>(xs[i]);
    s.out[i] = outs[i];
    s.ld[i] = lds[i];
    s.M[i] = Ms[i];
    s.N[i] = static_cast<int>(Ns[i]);
    s.gx[i] = gx;
    s.blk0[i] = blk;
    blk += gx * static_cast<int>(splits);
  }
  s.blk0[count] = blk;
  s.count = count;
  (void)launch_k(colsum_multi_kernel, dim3(static_cast<unsigned>(blk)), dim3(256), 0, static_cast<cudaStream_t>(stream), s);
  return launch_status();
}

int owl_kv_append(const void* src, int64_t src_sb, int64_t src_sh, void* cache, int64_t cache_sb, int64_t cache_sh,
                  int64_t pos, int64_t B, int64_t H, int64_t Ln, int64_t dh, owl_stream_t stream) {
  OWL_REQUIRE(src && cache && B > 0 && H > 0 && Ln > 0 && dh > 0 && pos >= 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(dh % 8 == 0, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(src) && aligned16(cache) && src_sb % 8 == 0 && src_sh % 8 == 0 && cache_sb % 8 == 0 &&
                  cache_sh % 8 == 0,
              OWL_ERR_ALIGN);
  (void)launch_k(kv_append_kernel, dim3(grid_for(B * H * Ln * dh / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(src), src_sb, src_sh, static_cast<bf16*>(cache), cache_sb, cache_sh, pos, B,
      static_cast<int>(H), Ln, static_cast<int>(dh));
  return launch_status();
}

int owl_layer_norm_fwd(const void* x, void* y, int64_t M, int64_t d, float eps, owl_stream_t stream) {
  OWL_REQUIRE(x && y && M > 0 && d > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(x) && aligned16(y), OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  OWL_DISPATCH_NV(d, ((void)launch_k(layer_norm_fwd_kernel<NV>, dim3(row_grid(M)), dim3(kRowThreads), 0, s, static_cast<const bf16*>(x), static_cast<bf16*>(y), M, static_cast<int>(d), eps)));
  return launch_status();
}

int owl_layer_norm_bwd(const void* dy, const void* x, void* dx, int64_t M, int64_t d, float eps, owl_stream_t stream) {
  OWL_REQUIRE(dy && x && dx && M > 0 && d > 0, OWL_ERR_BAD_ARG);
  OWL_REQUIRE(d % 8 == 0 && d <= 4096, OWL_ERR_UNSUPPORTED);
  OWL_REQUIRE(aligned16(dy) && aligned16(x) && aligned16(dx), OWL_ERR_ALIGN);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  OWL_DISPATCH_NV(d, ((void)launch_k(layer_norm_bwd_kernel<NV>, dim3(row_grid(M)), dim3(kRowThreads), 0, s, static_cast<const bf16*>(dy), static_cast<const bf16*>(x), static_cast<bf16*>(dx), M, static_cast<int>(d), eps)));
  return launch_status();
}

int owl_debug_timestamp(uint64_t* dst, owl_stream_t stream) {
  OWL_REQUIRE(dst != nullptr, OWL_ERR_BAD_ARG);
  timestamp_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(dst));
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? OWL_OK : static_cast<int>(e);
}

int owl_cast_f32_bf16_multi(const float* const* srcs, void* const* dsts, const int64_t* ns, int count,
                            owl_stream_t stream) {
  OWL_REQUIRE(srcs && dsts && ns && count > 0, OWL_ERR_BAD_ARG);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long cap_total = static_cast<long long>(sm_count()) * 8;   // 8 x 256 threads = one resident wave per SM
  for (int base = 0; base < count; base += 8) {
    const int c = count - base < 8 ? count - base : 8;
    CastSegs s{};
    long long total8 = 0;
    for (int i = 0; i < c; ++i) {
      OWL_REQUIRE(srcs[base + i] && dsts[base + i] && ns[base + i] > 0, OWL_ERR_BAD_ARG);
      OWL_REQUIRE(ns[base + i] % 8 == 0, OWL_ERR_UNSUPPORTED);
      OWL_REQUIRE(aligned16(srcs[base + i]) && aligned16(dsts[base + i]), OWL_ERR_ALIGN);
      total8 += ns[base + i] / 8;
    }
    long long want = (total8 + 256 * 4 - 1) / (256 * 4);  // ~4 chunks of 16 bytes out per thread
    if (want > cap_total) want = cap_total;
    if (want < c) want = c;
    int blk = 0;
    for (int i = 0; i < c; ++i) {
      s.src[i] = srcs[base + i];
      s.dst[i] = static_cast<bf16*>(dsts[base + i]);
      s.n8[i] = ns[base + i] / 8;
      s.blk0[i] = blk;
      long long b = (want * s.n8[i] + total8 - 1) / total8;
      if (b < 1) b = 1;
      blk += static_cast<int>(b);
    }
    s.blk0[c] = blk;
    s.count = c;
    (void)launch_k(cast_multi_kernel, dim3(static_cast<unsigned>(blk)), dim3(256), 0, st, s);
    int rc = launch_status();
    if (rc != OWL_OK) return rc;
  }
  return OWL_OK;
}

}  // extern "C"
