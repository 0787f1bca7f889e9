// Rectified-flow training glue either side of the model call: the noising of the clean latents and the
// MSE against the velocity target, each as one pass.
//
//   owl_rf_noise      lerpd = x * (1 - t) + z * t ;  target = z - x        (models/gamerft.py:93-96, audiorft.py:75-79,
//                     gamerft_audio.py:139-146: three elementwise torch kernels + two temporaries per stream)
//   owl_mse_fwd / bwd loss = mean((pred - target)^2) in fp32 without the `.float()` copies of models/*.py
//                     (F.mse_loss(pred.float(), target.float())), and its gradient 2 (pred - target) dloss / n as bf16
//
// bf16 tensors, 16-byte vectors, grid-stride; the reduction is warp shuffles -> shared memory -> ONE fp32 atomic per CTA.
#include <cuda_bf16.h>

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

__device__ __forceinline__ void fl_unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 fl_pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// x, z: [rows, inner] bf16 (inner % 8 == 0), t: [rows] (one timestep per (batch, frame) row), bf16
__global__ void __launch_bounds__(256)
rf_noise_kernel(const bf16* __restrict__ x, const bf16* __restrict__ z, const bf16* __restrict__ t, bf16* __restrict__ lerpd,
                bf16* __restrict__ target, int64_t rows, int inner8) {
  pdl_prologue();
  const int64_t total = rows * inner8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / inner8;
    const float tt = __bfloat162float(t[r]);
    float fx[8], fz[8], l[8], tg[8];
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(x) + i), fx);
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(z) + i), fz);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // the reference evaluates x * (1 - t) + z * t in the tensors' dtype: each product and the sum round to bf16
      l[j] = round_bf16(round_bf16(fx[j] * round_bf16(1.f - tt)) + round_bf16(fz[j] * tt));
      tg[j] = fz[j] - fx[j];
    }
    reinterpret_cast<uint4*>(lerpd)[i] = fl_pack8(l);
    reinterpret_cast<uint4*>(target)[i] = fl_pack8(tg);
  }
}

__global__ void __launch_bounds__(256)
mse_fwd_kernel(const bf16* __restrict__ pred, const bf16* __restrict__ target, float* __restrict__ loss, int64_t n8,
               float inv_n) {
  pdl_prologue();
  __shared__ float s_part[8];
  float acc = 0.f;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float p[8], t[8];
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(pred) + i), p);
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(target) + i), t);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = p[j] - t[j];
      acc = fmaf(d, d, acc);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += s_part[w];
    atomicAdd(loss, s * inv_n);
  }
}

__global__ void __launch_bounds__(256)
mse_bwd_kernel(const bf16* __restrict__ pred, const bf16* __restrict__ target, const float* __restrict__ dloss,
               bf16* __restrict__ dpred, int64_t n8, float two_over_n) {
  pdl_prologue();
  const float g = __ldg(dloss) * two_over_n;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float p[8], t[8];
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(pred) + i), p);
    fl_unpack8(__ldg(reinterpret_cast<const uint4*>(target) + i), t);
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] = (p[j] - t[j]) * g;
    reinterpret_cast<uint4*>(dpred)[i] = fl_pack8(p);
  }
}

static inline int fl_grid(int64_t items) {
  int64_t b = (items + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace owl

using namespace owl;

extern "C" {

int owl_rf_noise(const void* x, const void* z, const void* t, void* lerpd, void* target, int64_t rows, int64_t inner,
                 owl_stream_t stream) {
  if (!x || !z || !t || !lerpd || !target || rows <= 0 || inner <= 0) return OWL_ERR_BAD_ARG;
  if (inner % 8 != 0 || inner / 8 > INT32_MAX) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(x) || !aligned16(z) || !aligned16(lerpd) || !aligned16(target)) return OWL_ERR_ALIGN;
  (void)launch_k(rf_noise_kernel, dim3(fl_grid(rows * (inner / 8))), dim3(256), 0, static_cast<cudaStream_t>(stream),
                 static_cast<const bf16*>(x), static_cast<const bf16*>(z), static_cast<const bf16*>(t),
                 static_cast<bf16*>(lerpd), static_cast<bf16*>(target), rows, static_cast<int>(inner / 8));
  return launch_status();
}

int owl_mse_fwd(const void* pred, const void* target, float* loss, int64_t n, owl_stream_t stream) {
  if (!pred || !target || !loss || n <= 0) return OWL_ERR_BAD_ARG;
  if (n % 8 != 0) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(pred) || !aligned16(target)) return OWL_ERR_ALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(loss, 0, sizeof(float), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  LaunchOpts lo;
  lo.pdl = false;   // follows a memset node
  e = launch_kx(lo, mse_fwd_kernel, dim3(fl_grid(n / 8)), dim3(256), 0, s, static_cast<const bf16*>(pred),
                static_cast<const bf16*>(target), loss, n / 8, 1.0f / static_cast<float>(n));
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return static_cast<int>(e);
  }
  return launch_status();
}

int owl_mse_bwd(const void* pred, const void* target, const float* dloss, void* dpred, int64_t n, owl_stream_t stream) {
  if (!pred || !target || !dloss || !dpred || n <= 0) return OWL_ERR_BAD_ARG;
  if (n % 8 != 0) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(pred) || !aligned16(target) || !aligned16(dpred)) return OWL_ERR_ALIGN;
  (void)launch_k(mse_bwd_kernel, dim3(fl_grid(n / 8)), dim3(256), 0, static_cast<cudaStream_t>(stream),
                 static_cast<const bf16*>(pred), static_cast<const bf16*>(target), dloss, static_cast<bf16*>(dpred), n / 8,
                 2.0f / static_cast<float>(n));
  return launch_status();
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Shard reduction of the copy-engine gradient exchange (engine.PeerExchange): dst[i] = (dst[i] + sum_j src_j[i]) * scale
// with src_j = srcs + j * src_stride, fp32, 16-byte vectors.  The peers' shards arrive by copy-engine peer copies; this
// kernel is the only SM work of a gradient all-reduce (memory-bound, a few CTAs, runs next to the backward GEMMs).
// ------------------------------------------------------------------------------------------------
namespace owl {

// 128 threads x <= 32 registers: one such CTA fits NEXT TO a resident 384-thread x 160-register GEMM CTA (4096 registers of the
// SM's 65536 are left), so the reduction of a bucket shard runs on the SMs the backward GEMMs occupy instead of waiting for
// a kernel boundary; it is bandwidth-bound and short (a 37 MB shard at N = 2).
__global__ void __launch_bounds__(128, 16)
shard_reduce_kernel(float* __restrict__ dst, const float* __restrict__ srcs, long long src_stride, int n_src, long long n4,
                    long long n, float scale) {
  pdl_prologue();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + stride < n4; i += 2 * stride) {   // two independent 16-byte streams per thread
    float4 a = reinterpret_cast<const float4*>(dst)[i];
    float4 c = reinterpret_cast<const float4*>(dst)[i + stride];
    for (int j = 0; j < n_src; ++j) {
      const float4* sp = reinterpret_cast<const float4*>(srcs + j * src_stride);
      const float4 b = __ldg(sp + i), d = __ldg(sp + i + stride);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      c.x += d.x; c.y += d.y; c.z += d.z; c.w += d.w;
    }
    a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
    c.x *= scale; c.y *= scale; c.z *= scale; c.w *= scale;
    reinterpret_cast<float4*>(dst)[i] = a;
    reinterpret_cast<float4*>(dst)[i + stride] = c;
  }
  for (; i < n4; i += stride) {
    float4 a = reinterpret_cast<const float4*>(dst)[i];
    for (int j = 0; j < n_src; ++j) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(srcs + j * src_stride) + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
    reinterpret_cast<float4*>(dst)[i] = a;
  }
  // tail (n not a multiple of 4)
  for (long long t = n4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += stride) {
    float a = dst[t];
    for (int j = 0; j < n_src; ++j) a += srcs[j * src_stride + t];
    dst[t] = a * scale;
  }
}

}  // namespace owl

extern "C" int owl_shard_reduce(float* dst, const float* srcs, int64_t src_stride, int n_src, int64_t n, float scale,
                                int max_ctas, owl_stream_t stream) {
  if (!dst || n <= 0 || n_src < 0 || (n_src > 0 && !srcs)) return OWL_ERR_BAD_ARG;
  if (!owl::aligned16(dst) || (n_src > 0 && (!owl::aligned16(srcs) || src_stride % 4))) return OWL_ERR_ALIGN;
  long long blocks = (n / 4 + 255) / 256;     // two 16-byte groups per thread
  const long long cap = max_ctas > 0 ? max_ctas : owl::sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  (void)owl::launch_k(owl::shard_reduce_kernel, dim3(static_cast<unsigned>(blocks)), dim3(128), 0,
                      static_cast<cudaStream_t>(stream), dst, srcs, static_cast<long long>(src_stride), n_src,
                      static_cast<long long>(n / 4), static_cast<long long>(n), scale);
  return owl::launch_status();
}

// Same reduction with the sources read IN PLACE from the peers' buckets over NVLink (peer-mapped pointers of the symmetric
// allocation) instead of from a local staging copy: no copy-engine pulls and no staging traffic for the reduce-scatter half
// of the exchange; 2 x 16 bytes x (n_src) loads in flight per thread cover the NVLink latency.
namespace owl {

struct ShardSrcs {
  const float* p[15];
};

__global__ void __launch_bounds__(128, 16)
shard_reduce_peers_kernel(float* __restrict__ dst, const ShardSrcs srcs, int n_src, long long n4, long long n, float scale) {
  pdl_prologue();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 a = reinterpret_cast<const float4*>(dst)[i];
#pragma unroll 5
    for (int j = 0; j < n_src; ++j) {
      const float4 b = __ldcg(reinterpret_cast<const float4*>(srcs.p[j]) + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
    reinterpret_cast<float4*>(dst)[i] = a;
  }
  for (long long t = n4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += stride) {
    float a = dst[t];
    for (int j = 0; j < n_src; ++j) a += __ldcg(srcs.p[j] + t);
    dst[t] = a * scale;
  }
}

}  // namespace owl

extern "C" int owl_shard_reduce_peers(float* dst, const float* const* srcs, int n_src, int64_t n, float scale, int max_ctas,
                                      owl_stream_t stream) {
  if (!dst || n <= 0 || n_src < 0 || n_src > 15 || (n_src > 0 && !srcs)) return OWL_ERR_BAD_ARG;
  if (!owl::aligned16(dst)) return OWL_ERR_ALIGN;
  owl::ShardSrcs s{};
  for (int j = 0; j < n_src; ++j) {
    if (!srcs[j] || !owl::aligned16(srcs[j])) return OWL_ERR_ALIGN;
    s.p[j] = srcs[j];
  }
  long long blocks = (n / 4 + 127) / 128;
  const long long cap = max_ctas > 0 ? max_ctas : owl::sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  (void)owl::launch_k(owl::shard_reduce_peers_kernel, dim3(static_cast<unsigned>(blocks)), dim3(128), 0,
                      static_cast<cudaStream_t>(stream), dst, s, n_src, static_cast<long long>(n / 4),
                      static_cast<long long>(n), scale);
  return owl::launch_status();
}

// ------------------------------------------------------------------------------------------------
// In-switch all-reduce of a bucket shard (NVLink SHARP / multimem): `mc` is the MULTICAST address of this rank's shard inside
// the symmetric bucket allocation.  multimem.ld_reduce makes NVSwitch fetch the 16 bytes from every rank and add them on
// the way (one fp32 sum arrives), multimem.st writes the scaled sum back to every rank's copy.  Per rank and exchange the
// links carry the bucket once per direction (reduce-scatter + all-gather by copies: 2 (N-1)/N times), nothing is staged
// in HBM and no copy is enqueued.  Same co-residency recipe as shard_reduce_kernel (128 threads, <= 32 registers); four
// independent 16-byte requests per thread cover the ~3.5 us round trip through the switch.
// ------------------------------------------------------------------------------------------------
namespace owl {

__device__ __forceinline__ float4 mc_ld_reduce(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float* p, float4 v, float s) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(p), "f"(v.x * s), "f"(v.y * s), "f"(v.z * s), "f"(v.w * s) : "memory");
}

__global__ void __launch_bounds__(128, 16) mc_allreduce_kernel(float* __restrict__ mc, long long n4, long long n, float scale) {
  pdl_prologue();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    const float4 a = mc_ld_reduce(mc + 4 * i);
    const float4 b = mc_ld_reduce(mc + 4 * (i + stride));
    const float4 c = mc_ld_reduce(mc + 4 * (i + 2 * stride));
    const float4 d = mc_ld_reduce(mc + 4 * (i + 3 * stride));
    mc_st(mc + 4 * i, a, scale);
    mc_st(mc + 4 * (i + stride), b, scale);
    mc_st(mc + 4 * (i + 2 * stride), c, scale);
    mc_st(mc + 4 * (i + 3 * stride), d, scale);
  }
  for (; i < n4; i += stride) mc_st(mc + 4 * i, mc_ld_reduce(mc + 4 * i), scale);
  for (long long t = n4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += stride) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc + t) : "memory");
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(mc + t), "f"(v * scale) : "memory");
  }
}

}  // namespace owl

extern "C" int owl_mc_allreduce(float* mc_shard, int64_t n, float scale, int max_ctas, owl_stream_t stream) {
  if (!mc_shard || n <= 0) return OWL_ERR_BAD_ARG;
  if (!owl::aligned16(mc_shard)) return OWL_ERR_ALIGN;
  long long blocks = (n / 4 + 511) / 512;     // four 16-byte requests per thread
  const long long cap = max_ctas > 0 ? max_ctas : owl::sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  (void)owl::launch_k(owl::mc_allreduce_kernel, dim3(static_cast<unsigned>(blocks)), dim3(128), 0,
                      static_cast<cudaStream_t>(stream), mc_shard, static_cast<long long>(n / 4), static_cast<long long>(n),
                      scale);
  return owl::launch_status();
}
