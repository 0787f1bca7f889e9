// CUDA-core attention forward/backward with the contract of owl_attn_fwd / owl_attn_bwd
// (include/owl_dit.h).  VERIFICATION AID ONLY: an independent on-device answer for the tests at
// sizes where the CPU oracle is too slow.  The product path never calls these.
//
// One warp per query row (forward, dq) or per key row (dk, dv).  Keys/queries are visited 32 at a
// time: lane t scores element (j0 + t) with a full dot product, then the 32 probabilities are
// broadcast with shuffles while every lane accumulates its dh/32 output features.
#include <cuda_bf16.h>
#include <math.h>

#include "attn_mask.cuh"
#include "host_common.h"
#include "sm100.cuh"

namespace owl {

struct AttnCheckParams {
  const bf16 *q, *k, *v, *o, *dout;
  bf16 *out, *dq, *dk, *dv;
  float* lse;
  const float* lse_in;
  float* delta;
  long long q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, o_sb, o_sl;
  long long dq_sb, dq_sh, dk_sb, dk_sh, dv_sb, dv_sh;
  int B, H, Lq, Lk, dh;
  float scale;
  MaskParams m;
};

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float dot_row(const float* a_smem, const bf16* __restrict__ b, int dh) {
  float s = 0.f;
  for (int d = 0; d < dh; d += 8) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(b + d));
    s += a_smem[d] * bf16_lo(v.x) + a_smem[d + 1] * bf16_hi(v.x) + a_smem[d + 2] * bf16_lo(v.y) +
         a_smem[d + 3] * bf16_hi(v.y) + a_smem[d + 4] * bf16_lo(v.z) + a_smem[d + 5] * bf16_hi(v.z) +
         a_smem[d + 6] * bf16_lo(v.w) + a_smem[d + 7] * bf16_hi(v.w);
  }
  return s;
}

static constexpr int kWarps = 4;

// FPL = features per lane (dh / 32)
template <int FPL>
__global__ void __launch_bounds__(kWarps * 32) attn_fwd_check_kernel(const AttnCheckParams p) {
  __shared__ float qs[kWarps][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kWarps + warp;
  const long long total = static_cast<long long>(p.B) * p.H * p.Lq;
  if (row >= total) return;
  const int i = static_cast<int>(row % p.Lq);
  const int h = static_cast<int>((row / p.Lq) % p.H);
  const int b = static_cast<int>(row / (static_cast<long long>(p.Lq) * p.H));
  const bf16* qr = p.q + b * p.q_sb + h * p.q_sh + static_cast<long long>(i) * p.dh;
  for (int d = lane; d < p.dh; d += 32) qs[warp][d] = __bfloat162float(qr[d]);
  __syncwarp();
  const bf16* kb = p.k + b * p.k_sb + h * p.k_sh;
  const bf16* vb = p.v + b * p.v_sb + h * p.v_sh;
  int lo, hi;
  mask_row_range(p.m, i, p.Lk, lo, hi);
  float mrun = -INFINITY, lrun = 0.f, acc[FPL];
#pragma unroll
  for (int f = 0; f < FPL; ++f) acc[f] = 0.f;
  for (int j0 = lo; j0 < hi; j0 += 32) {
    const int j = j0 + lane;
    float s = -INFINITY;
    if (j < hi && mask_allowed(p.m, b, i, j)) s = dot_row(qs[warp], kb + static_cast<long long>(j) * p.dh, p.dh) * p.scale;
    const float mnew = fmaxf(mrun, wmax(s));
    if (mnew == -INFINITY) continue;
    const float corr = __expf(mrun - mnew);
    const float pr = s == -INFINITY ? 0.f : __expf(s - mnew);
    lrun = lrun * corr + wsum(pr);
#pragma unroll
    for (int f = 0; f < FPL; ++f) acc[f] *= corr;
    const int n = min(32, hi - j0);
    for (int t = 0; t < n; ++t) {
      const float pt = __shfl_sync(0xffffffffu, pr, t);
      if (pt != 0.f) {
        const bf16* vr = vb + static_cast<long long>(j0 + t) * p.dh;
#pragma unroll
        for (int f = 0; f < FPL; ++f) acc[f] += pt * __bfloat162float(vr[lane + 32 * f]);
      }
    }
    mrun = mnew;
  }
  bf16* orow = p.out + b * p.o_sb + static_cast<long long>(i) * p.o_sl + static_cast<long long>(h) * p.dh;
  const float inv = lrun > 0.f ? 1.f / lrun : 0.f;
#pragma unroll
  for (int f = 0; f < FPL; ++f) orow[lane + 32 * f] = __float2bfloat16_rn(acc[f] * inv);
  if (p.lse != nullptr && lane == 0) p.lse[row] = lrun > 0.f ? mrun + __logf(lrun) : -INFINITY;
}

// delta + dq: one warp per query row
template <int FPL>
__global__ void __launch_bounds__(kWarps * 32) attn_bwd_dq_check_kernel(const AttnCheckParams p) {
  __shared__ float qs[kWarps][128];
  __shared__ float gs[kWarps][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kWarps + warp;
  const long long total = static_cast<long long>(p.B) * p.H * p.Lq;
  if (row >= total) return;
  const int i = static_cast<int>(row % p.Lq);
  const int h = static_cast<int>((row / p.Lq) % p.H);
  const int b = static_cast<int>(row / (static_cast<long long>(p.Lq) * p.H));
  const bf16* qr = p.q + b * p.q_sb + h * p.q_sh + static_cast<long long>(i) * p.dh;
  const long long ooff = b * p.o_sb + static_cast<long long>(i) * p.o_sl + static_cast<long long>(h) * p.dh;
  float dl = 0.f;
  for (int d = lane; d < p.dh; d += 32) {
    qs[warp][d] = __bfloat162float(qr[d]);
    const float g = __bfloat162float(p.dout[ooff + d]);
    gs[warp][d] = g;
    dl += g * __bfloat162float(p.o[ooff + d]);
  }
  dl = wsum(dl);
  if (lane == 0) p.delta[row] = dl;
  __syncwarp();
  const float lse = p.lse_in[row];
  const bf16* kb = p.k + b * p.k_sb + h * p.k_sh;
  const bf16* vb = p.v + b * p.v_sb + h * p.v_sh;
  int lo, hi;
  mask_row_range(p.m, i, p.Lk, lo, hi);
  float acc[FPL];
#pragma unroll
  for (int f = 0; f < FPL; ++f) acc[f] = 0.f;
  for (int j0 = lo; j0 < hi; j0 += 32) {
    const int j = j0 + lane;
    float ds = 0.f;
    if (j < hi && lse != -INFINITY && mask_allowed(p.m, b, i, j)) {
      const float s = dot_row(qs[warp], kb + static_cast<long long>(j) * p.dh, p.dh) * p.scale;
      const float pr = __expf(s - lse);
      const float dp = dot_row(gs[warp], vb + static_cast<long long>(j) * p.dh, p.dh);
      ds = pr * (dp - dl) * p.scale;
    }
    const int n = min(32, hi - j0);
    for (int t = 0; t < n; ++t) {
      const float dt = __shfl_sync(0xffffffffu, ds, t);
      if (dt != 0.f) {
        const bf16* kr = kb + static_cast<long long>(j0 + t) * p.dh;
#pragma unroll
        for (int f = 0; f < FPL; ++f) acc[f] += dt * __bfloat162float(kr[lane + 32 * f]);
      }
    }
  }
  bf16* dqr = p.dq + b * p.dq_sb + h * p.dq_sh + static_cast<long long>(i) * p.dh;
#pragma unroll
  for (int f = 0; f < FPL; ++f) dqr[lane + 32 * f] = __float2bfloat16_rn(acc[f]);
}

// dk, dv: one warp per key row
template <int FPL>
__global__ void __launch_bounds__(kWarps * 32) attn_bwd_dkv_check_kernel(const AttnCheckParams p) {
  __shared__ float ks[kWarps][128];
  __shared__ float vs[kWarps][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = static_cast<long long>(blockIdx.x) * kWarps + warp;
  const long long total = static_cast<long long>(p.B) * p.H * p.Lk;
  if (row >= total) return;
  const int j = static_cast<int>(row % p.Lk);
  const int h = static_cast<int>((row / p.Lk) % p.H);
  const int b = static_cast<int>(row / (static_cast<long long>(p.Lk) * p.H));
  const bf16* kr = p.k + b * p.k_sb + h * p.k_sh + static_cast<long long>(j) * p.dh;
  const bf16* vr = p.v + b * p.v_sb + h * p.v_sh + static_cast<long long>(j) * p.dh;
  for (int d = lane; d < p.dh; d += 32) {
    ks[warp][d] = __bfloat162float(kr[d]);
    vs[warp][d] = __bfloat162float(vr[d]);
  }
  __syncwarp();
  // query rows that may see key j
  const long long fk = j / p.m.T;
  long long ilo = 0, ihi = p.Lq;
  if (p.m.causal) ilo = max(ilo, fk * p.m.T - p.m.q_offset);
  if (p.m.W > 0) {
    ihi = min(ihi, (fk + p.m.W) * p.m.T - p.m.q_offset);
    if (!p.m.causal) ilo = max(ilo, (fk - p.m.W + 1) * p.m.T - p.m.q_offset);
  }
  const bf16* qb = p.q + b * p.q_sb + h * p.q_sh;
  const float* lse = p.lse_in + (static_cast<long long>(b) * p.H + h) * p.Lq;
  const float* delta = p.delta + (static_cast<long long>(b) * p.H + h) * p.Lq;
  float ak[FPL], av[FPL];
#pragma unroll
  for (int f = 0; f < FPL; ++f) ak[f] = av[f] = 0.f;
  for (long long i0 = ilo; i0 < ihi; i0 += 32) {
    const long long i = i0 + lane;
    float pr = 0.f, ds = 0.f;
    if (i < ihi && mask_allowed(p.m, b, static_cast<int>(i), j) && lse[i] != -INFINITY) {
      const float s = dot_row(ks[warp], qb + i * p.dh, p.dh) * p.scale;
      pr = __expf(s - lse[i]);
      const float dp = dot_row(vs[warp], p.dout + b * p.o_sb + i * p.o_sl + static_cast<long long>(h) * p.dh, p.dh);
      ds = pr * (dp - delta[i]) * p.scale;
    }
    const int n = static_cast<int>(min(32LL, ihi - i0));
    for (int t = 0; t < n; ++t) {
      const float pt = __shfl_sync(0xffffffffu, pr, t);
      const float dt = __shfl_sync(0xffffffffu, ds, t);
      if (pt != 0.f || dt != 0.f) {
        const bf16* qr = qb + (i0 + t) * p.dh;
        const bf16* gr = p.dout + b * p.o_sb + (i0 + t) * p.o_sl + static_cast<long long>(h) * p.dh;
#pragma unroll
        for (int f = 0; f < FPL; ++f) {
          ak[f] += dt * __bfloat162float(qr[lane + 32 * f]);
          av[f] += pt * __bfloat162float(gr[lane + 32 * f]);
        }
      }
    }
  }
  bf16* dkr = p.dk + b * p.dk_sb + h * p.dk_sh + static_cast<long long>(j) * p.dh;
  bf16* dvr = p.dv + b * p.dv_sb + h * p.dv_sh + static_cast<long long>(j) * p.dh;
#pragma unroll
  for (int f = 0; f < FPL; ++f) {
    dkr[lane + 32 * f] = __float2bfloat16_rn(ak[f]);
    dvr[lane + 32 * f] = __float2bfloat16_rn(av[f]);
  }
}

}  // namespace owl

using namespace owl;

static int fill_mask(const owl_attn_mask* mask, int64_t B, int64_t Lq, int64_t Lk, MaskParams* m) {
  if (!mask || mask->tokens_per_frame <= 0 || mask->q_offset < 0) return OWL_ERR_BAD_ARG;
  m->T = mask->tokens_per_frame;
  m->W = mask->window_frames;
  m->q_offset = mask->q_offset;
  m->causal = mask->causal;
  m->doc = mask->doc_id;
  m->doc_ld = mask->doc_ld;
  (void)B; (void)Lq; (void)Lk;
  return OWL_OK;
}

extern "C" {

int owl_attn_fwd_check(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                       const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       owl_stream_t stream) {
  if (!q || !k || !v || !o || B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  AttnCheckParams p{};
  int rc = fill_mask(mask, B, Lq, Lk, &p.m);
  if (rc != OWL_OK) return rc;
  p.q = static_cast<const bf16*>(q); p.k = static_cast<const bf16*>(k); p.v = static_cast<const bf16*>(v);
  p.out = static_cast<bf16*>(o); p.lse = lse;
  p.q_sb = q_sb; p.q_sh = q_sh; p.k_sb = k_sb; p.k_sh = k_sh; p.v_sb = v_sb; p.v_sh = v_sh; p.o_sb = o_sb; p.o_sl = o_sl;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.dh = static_cast<int>(dh);
  p.scale = 1.0f / sqrtf(static_cast<float>(dh));
  const long long rows = B * H * Lq;
  const unsigned grid = static_cast<unsigned>((rows + kWarps - 1) / kWarps);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dh == 64) attn_fwd_check_kernel<2><<<grid, kWarps * 32, 0, s>>>(p);
  else attn_fwd_check_kernel<4><<<grid, kWarps * 32, 0, s>>>(p);
  return launch_status();
}

int owl_attn_bwd_check(const void* dout, const void* o, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb,
                       int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb,
                       int64_t v_sh, const float* lse, float* delta, float* dq_acc, void* dq, int64_t dq_sb,
                       int64_t dq_sh, void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       owl_stream_t stream) {
  (void)dq_acc;
  if (!dout || !o || !q || !k || !v || !lse || !delta || !dq || !dk || !dv || B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  AttnCheckParams p{};
  int rc = fill_mask(mask, B, Lq, Lk, &p.m);
  if (rc != OWL_OK) return rc;
  p.q = static_cast<const bf16*>(q); p.k = static_cast<const bf16*>(k); p.v = static_cast<const bf16*>(v);
  p.o = static_cast<const bf16*>(o); p.dout = static_cast<const bf16*>(dout);
  p.dq = static_cast<bf16*>(dq); p.dk = static_cast<bf16*>(dk); p.dv = static_cast<bf16*>(dv);
  p.lse_in = lse; p.delta = delta;
  p.q_sb = q_sb; p.q_sh = q_sh; p.k_sb = k_sb; p.k_sh = k_sh; p.v_sb = v_sb; p.v_sh = v_sh; p.o_sb = o_sb; p.o_sl = o_sl;
  p.dq_sb = dq_sb; p.dq_sh = dq_sh; p.dk_sb = dk_sb; p.dk_sh = dk_sh; p.dv_sb = dv_sb; p.dv_sh = dv_sh;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.dh = static_cast<int>(dh);
  p.scale = 1.0f / sqrtf(static_cast<float>(dh));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const unsigned gq = static_cast<unsigned>((B * H * Lq + kWarps - 1) / kWarps);
  const unsigned gk = static_cast<unsigned>((B * H * Lk + kWarps - 1) / kWarps);
  if (dh == 64) {
    attn_bwd_dq_check_kernel<2><<<gq, kWarps * 32, 0, s>>>(p);
    attn_bwd_dkv_check_kernel<2><<<gk, kWarps * 32, 0, s>>>(p);
  } else {
    attn_bwd_dq_check_kernel<4><<<gq, kWarps * 32, 0, s>>>(p);
    attn_bwd_dkv_check_kernel<4><<<gk, kWarps * 32, 0, s>>>(p);
  }
  return launch_status(2);
}

}  // extern "C"
