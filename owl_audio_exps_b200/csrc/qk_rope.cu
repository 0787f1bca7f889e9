// QK-RMSNorm + RoPE + head split (attn.py:83-89, rope.py:43-51), forward and backward.
//
// One thread owns 8 contiguous features (16 bytes) of one (token, q|k|v, head) row, so the
// [B, L, 3, H, dh] side is read/written as one flat contiguous stream; G = dh/8 neighbouring lanes
// share a head row and reduce its sum of squares with xor-shuffles.  The head-major side
// ([B, H, L, dh] with caller strides, e.g. the tail of a KV cache) is touched in 128/256-byte rows.
// RoPE rotates pairs (2i, 2i+1) and emits [y_even | y_odd] halves exactly as rope.py:48-51.
#include <cuda_bf16.h>

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

__device__ __forceinline__ float group_sum(float v, int G) {
  for (int o = G >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Thread mapping (both kernels): one CTA of kQkThreads threads per token (b, l); its 3*H*G 16-byte chunks are walked
// with a block-wide stride, so all index math is 32-bit with G a power of two (one 64-bit division per thread
// in total), and the unrolled loop keeps several independent 16-byte loads in flight per thread.
static constexpr int kQkThreads = 128;

__global__ void __launch_bounds__(kQkThreads)
qk_norm_rope_fwd_kernel(const bf16* __restrict__ qkv, const float* __restrict__ cos_tab,
                        const float* __restrict__ sin_tab, int64_t pos_offset, bf16* __restrict__ q, int64_t q_sb,
                        int64_t q_sh, bf16* __restrict__ k, int64_t k_sb, int64_t k_sh, bf16* __restrict__ v,
                        int64_t v_sb, int64_t v_sh, float* __restrict__ rstd_q, float* __restrict__ rstd_k, int64_t B,
                        int64_t L, int H, int dh, float eps) {
  pdl_prologue();
  const int lgG = dh == 64 ? 3 : 4, G = 1 << lgG;
  const int HG = H << lgG, per_tok = 3 * HG;
  const int64_t token = blockIdx.x;
  const int64_t b = token / L, l = token - b * L;
  const bf16* src = qkv + token * per_tok * 8;
  const int half = dh >> 1;
  const int64_t pos = pos_offset + l;
  const float inv_dh = 1.f / static_cast<float>(dh);
#pragma unroll 3
  for (int c0 = 0; c0 < per_tok; c0 += kQkThreads) {
    const int c = c0 + threadIdx.x;
    const bool active = c < per_tok;
    const int cc = active ? c : 0;
    const int which = cc / HG;
    const int rem = cc - which * HG;
    const int h = rem >> lgG, chunk = rem & (G - 1);
    uint4 raw = make_uint4(0, 0, 0, 0);
    if (active) raw = __ldg(reinterpret_cast<const uint4*>(src + cc * 8));
    float f[8];
    f[0] = bf16_lo(raw.x); f[1] = bf16_hi(raw.x); f[2] = bf16_lo(raw.y); f[3] = bf16_hi(raw.y);
    f[4] = bf16_lo(raw.z); f[5] = bf16_hi(raw.z); f[6] = bf16_lo(raw.w); f[7] = bf16_hi(raw.w);
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) ss += f[j] * f[j];
    ss = group_sum(ss, G);  // all 32 lanes participate
    if (!active) continue;
    if (which == 2) {
      *reinterpret_cast<uint4*>(v + b * v_sb + h * v_sh + l * dh + chunk * 8) = raw;
      continue;
    }
    const float rstd = rsqrtf(ss * inv_dh + eps);
    float* rs = which == 0 ? rstd_q : rstd_k;
    if (rs != nullptr && chunk == 0) rs[(b * H + h) * L + l] = rstd;
    const float4 c4 = __ldg(reinterpret_cast<const float4*>(cos_tab + pos * half + chunk * 4));
    const float4 s4 = __ldg(reinterpret_cast<const float4*>(sin_tab + pos * half + chunk * 4));
    const float cs[4] = {c4.x, c4.y, c4.z, c4.w};
    const float sn[4] = {s4.x, s4.y, s4.z, s4.w};
    float y0[4], y1[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      // the reference rounds rms_norm's output to bf16 before the fp32 rotation (attn.py:84, rope.py:47)
      const float x0 = round_bf16(f[2 * p] * rstd), x1 = round_bf16(f[2 * p + 1] * rstd);
      y0[p] = x0 * cs[p] - x1 * sn[p];
      y1[p] = x1 * cs[p] + x0 * sn[p];
    }
    bf16* dst = (which == 0 ? q + b * q_sb + h * q_sh : k + b * k_sb + h * k_sh) + l * dh;
    uint2 o0, o1;
    o0.x = pack_bf16(y0[0], y0[1]); o0.y = pack_bf16(y0[2], y0[3]);
    o1.x = pack_bf16(y1[0], y1[1]); o1.y = pack_bf16(y1[2], y1[3]);
    *reinterpret_cast<uint2*>(dst + chunk * 4) = o0;
    *reinterpret_cast<uint2*>(dst + half + chunk * 4) = o1;
  }
}

// Backward: un-rotate the saved (rotated) q/k to recover the normalised row xh, un-rotate the
// incoming gradient (rotation is orthogonal: transpose = inverse), then the RMSNorm backward
//   dx = rstd * (dxh - xh * mean(dxh * xh)).
__global__ void __launch_bounds__(kQkThreads)
qk_norm_rope_bwd_kernel(const bf16* __restrict__ dq, const bf16* __restrict__ dk, const bf16* __restrict__ dv,
                        int64_t g_sb, int64_t g_sh, const bf16* __restrict__ q, int64_t q_sb, int64_t q_sh,
                        const bf16* __restrict__ k, int64_t k_sb, int64_t k_sh, const float* __restrict__ rstd_q,
                        const float* __restrict__ rstd_k, const float* __restrict__ cos_tab,
                        const float* __restrict__ sin_tab, int64_t pos_offset, bf16* __restrict__ dqkv, int64_t B,
                        int64_t L, int H, int dh) {
  pdl_prologue();
  const int lgG = dh == 64 ? 3 : 4, G = 1 << lgG;
  const int HG = H << lgG, per_tok = 3 * HG;
  const int64_t token = blockIdx.x;
  const int64_t b = token / L, l = token - b * L;
  bf16* dst = dqkv + token * per_tok * 8;
  const int half = dh >> 1;
  const int64_t pos = pos_offset + l;
  const float inv_dh = 1.f / static_cast<float>(dh);
#pragma unroll 3
  for (int c0 = 0; c0 < per_tok; c0 += kQkThreads) {
    const int c = c0 + threadIdx.x;
    const bool active = c < per_tok;
    const int cc = active ? c : 0;
    const int which = cc / HG;
    const int rem = cc - which * HG;
    const int h = rem >> lgG, chunk = rem & (G - 1);

    float xh[8], dxh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) xh[j] = dxh[j] = 0.f;
    float rstd = 0.f;
    uint4 out = make_uint4(0, 0, 0, 0);
    if (active && which < 2) {
      const bf16* gy = (which == 0 ? dq : dk) + b * g_sb + h * g_sh + l * dh;
      const bf16* yy = which == 0 ? q + b * q_sb + h * q_sh + l * dh : k + b * k_sb + h * k_sh + l * dh;
      const uint2 g0 = __ldg(reinterpret_cast<const uint2*>(gy + chunk * 4));
      const uint2 g1 = __ldg(reinterpret_cast<const uint2*>(gy + half + chunk * 4));
      const uint2 y0 = __ldg(reinterpret_cast<const uint2*>(yy + chunk * 4));
      const uint2 y1 = __ldg(reinterpret_cast<const uint2*>(yy + half + chunk * 4));
      const float4 c4 = __ldg(reinterpret_cast<const float4*>(cos_tab + pos * half + chunk * 4));
      const float4 s4 = __ldg(reinterpret_cast<const float4*>(sin_tab + pos * half + chunk * 4));
      rstd = __ldg((which == 0 ? rstd_q : rstd_k) + (b * H + h) * L + l);
      const float cs[4] = {c4.x, c4.y, c4.z, c4.w};
      const float sn[4] = {s4.x, s4.y, s4.z, s4.w};
      const float a0[4] = {bf16_lo(y0.x), bf16_hi(y0.x), bf16_lo(y0.y), bf16_hi(y0.y)};
      const float a1[4] = {bf16_lo(y1.x), bf16_hi(y1.x), bf16_lo(y1.y), bf16_hi(y1.y)};
      const float d0[4] = {bf16_lo(g0.x), bf16_hi(g0.x), bf16_lo(g0.y), bf16_hi(g0.y)};
      const float d1[4] = {bf16_lo(g1.x), bf16_hi(g1.x), bf16_lo(g1.y), bf16_hi(g1.y)};
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        xh[2 * p] = a0[p] * cs[p] + a1[p] * sn[p];
        xh[2 * p + 1] = a1[p] * cs[p] - a0[p] * sn[p];
        dxh[2 * p] = d0[p] * cs[p] + d1[p] * sn[p];
        dxh[2 * p + 1] = d1[p] * cs[p] - d0[p] * sn[p];
      }
    } else if (active) {
      out = __ldg(reinterpret_cast<const uint4*>(dv + b * g_sb + h * g_sh + l * dh + chunk * 8));
    }
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) dot += xh[j] * dxh[j];
    dot = group_sum(dot, G);
    if (!active) continue;
    if (which < 2) {
      const float m = dot * inv_dh;
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rstd * (dxh[j] - xh[j] * m);
      out.x = pack_bf16(o[0], o[1]); out.y = pack_bf16(o[2], o[3]);
      out.z = pack_bf16(o[4], o[5]); out.w = pack_bf16(o[6], o[7]);
    }
    *reinterpret_cast<uint4*>(dst + cc * 8) = out;
  }
}

// Bare RoPE on contiguous [R, L, dh] rows (RoPE.forward called on its own, rope.py:43-51).
// forward: pairs (2i, 2i+1) -> halves [y0 | y1];  inverse (its autograd): halves -> pairs with the
// transposed rotation.  One thread = 4 pairs.
__global__ void __launch_bounds__(256)
rope_apply_kernel(const bf16* __restrict__ x, const float* __restrict__ cos_tab, const float* __restrict__ sin_tab,
                  int64_t pos_offset, bf16* __restrict__ y, int64_t R, int64_t L, int dh, int inverse) {
  pdl_prologue();
  const int half = dh >> 1, Q = half >> 2;
  const int64_t total = R * L * Q;
  const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int chunk = static_cast<int>(t % Q);
  const int64_t row = t / Q;  // r * L + l
  const int64_t l = row % L;
  const int64_t pos = pos_offset + l;
  const float4 c4 = __ldg(reinterpret_cast<const float4*>(cos_tab + pos * half + chunk * 4));
  const float4 s4 = __ldg(reinterpret_cast<const float4*>(sin_tab + pos * half + chunk * 4));
  const float cs[4] = {c4.x, c4.y, c4.z, c4.w};
  const float sn[4] = {s4.x, s4.y, s4.z, s4.w};
  const bf16* xr = x + row * dh;
  bf16* yr = y + row * dh;
  if (!inverse) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(xr + chunk * 8));
    const float f[8] = {bf16_lo(raw.x), bf16_hi(raw.x), bf16_lo(raw.y), bf16_hi(raw.y),
                        bf16_lo(raw.z), bf16_hi(raw.z), bf16_lo(raw.w), bf16_hi(raw.w)};
    float y0[4], y1[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      y0[p] = f[2 * p] * cs[p] - f[2 * p + 1] * sn[p];
      y1[p] = f[2 * p + 1] * cs[p] + f[2 * p] * sn[p];
    }
    uint2 o0, o1;
    o0.x = pack_bf16(y0[0], y0[1]); o0.y = pack_bf16(y0[2], y0[3]);
    o1.x = pack_bf16(y1[0], y1[1]); o1.y = pack_bf16(y1[2], y1[3]);
    *reinterpret_cast<uint2*>(yr + chunk * 4) = o0;
    *reinterpret_cast<uint2*>(yr + half + chunk * 4) = o1;
  } else {
    const uint2 g0 = __ldg(reinterpret_cast<const uint2*>(xr + chunk * 4));
    const uint2 g1 = __ldg(reinterpret_cast<const uint2*>(xr + half + chunk * 4));
    const float d0[4] = {bf16_lo(g0.x), bf16_hi(g0.x), bf16_lo(g0.y), bf16_hi(g0.y)};
    const float d1[4] = {bf16_lo(g1.x), bf16_hi(g1.x), bf16_lo(g1.y), bf16_hi(g1.y)};
    float o[8];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      o[2 * p] = d0[p] * cs[p] + d1[p] * sn[p];
      o[2 * p + 1] = d1[p] * cs[p] - d0[p] * sn[p];
    }
    uint4 out;
    out.x = pack_bf16(o[0], o[1]); out.y = pack_bf16(o[2], o[3]);
    out.z = pack_bf16(o[4], o[5]); out.w = pack_bf16(o[6], o[7]);
    *reinterpret_cast<uint4*>(yr + chunk * 8) = out;
  }
}

}  // namespace owl

using namespace owl;

extern "C" {

int owl_qk_norm_rope_fwd(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t pos_offset, void* q,
                         int64_t q_sb, int64_t q_sh, void* k, int64_t k_sb, int64_t k_sh, void* v, int64_t v_sb,
                         int64_t v_sh, float* rstd_q, float* rstd_k, int64_t B, int64_t L, int64_t H, int64_t dh,
                         float eps, owl_stream_t stream) {
  if (!qkv || !cos_tab || !sin_tab || !q || !k || !v || B <= 0 || L <= 0 || H <= 0 || pos_offset < 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(qkv) || !aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(cos_tab) ||
      !aligned16(sin_tab) || (q_sb | q_sh | k_sb | k_sh | v_sb | v_sh) % 8)
    return OWL_ERR_ALIGN;
  if (B * L > INT32_MAX || 3 * H * dh > (1 << 24)) return OWL_ERR_UNSUPPORTED;
  const int64_t blocks = B * L;  // one CTA per token
  (void)launch_k(qk_norm_rope_fwd_kernel, dim3(static_cast<unsigned>(blocks)), dim3(kQkThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(qkv), cos_tab, sin_tab, pos_offset, static_cast<bf16*>(q), q_sb, q_sh,
      static_cast<bf16*>(k), k_sb, k_sh, static_cast<bf16*>(v), v_sb, v_sh, rstd_q, rstd_k, B, L, static_cast<int>(H),
      static_cast<int>(dh), eps);
  return launch_status();
}

int owl_qk_norm_rope_bwd(const void* dq, const void* dk, const void* dv, int64_t g_sb, int64_t g_sh, const void* q,
                         int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                         const float* rstd_q, const float* rstd_k, const float* cos_tab, const float* sin_tab,
                         int64_t pos_offset, void* dqkv, int64_t B, int64_t L, int64_t H, int64_t dh,
                         owl_stream_t stream) {
  if (!dq || !dk || !dv || !q || !k || !rstd_q || !rstd_k || !cos_tab || !sin_tab || !dqkv || B <= 0 || L <= 0 ||
      H <= 0 || pos_offset < 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(dq) || !aligned16(dk) || !aligned16(dv) || !aligned16(q) || !aligned16(k) || !aligned16(dqkv) ||
      !aligned16(cos_tab) || !aligned16(sin_tab) || (g_sb | g_sh | q_sb | q_sh | k_sb | k_sh) % 8)
    return OWL_ERR_ALIGN;
  if (B * L > INT32_MAX || 3 * H * dh > (1 << 24)) return OWL_ERR_UNSUPPORTED;
  const int64_t blocks = B * L;  // one CTA per token
  (void)launch_k(qk_norm_rope_bwd_kernel, dim3(static_cast<unsigned>(blocks)), dim3(kQkThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(dq), static_cast<const bf16*>(dk), static_cast<const bf16*>(dv), g_sb, g_sh,
      static_cast<const bf16*>(q), q_sb, q_sh, static_cast<const bf16*>(k), k_sb, k_sh, rstd_q, rstd_k, cos_tab,
      sin_tab, pos_offset, static_cast<bf16*>(dqkv), B, L, static_cast<int>(H), static_cast<int>(dh));
  return launch_status();
}

int owl_rope_apply(const void* x, const float* cos_tab, const float* sin_tab, int64_t pos_offset, void* y, int64_t R,
                   int64_t L, int64_t dh, int inverse, owl_stream_t stream) {
  if (!x || !y || !cos_tab || !sin_tab || R <= 0 || L <= 0 || dh <= 0 || pos_offset < 0) return OWL_ERR_BAD_ARG;
  if (dh % 8 != 0) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(x) || !aligned16(y) || !aligned16(cos_tab) || !aligned16(sin_tab)) return OWL_ERR_ALIGN;
  const int64_t total = R * L * (dh / 8);
  (void)launch_k(rope_apply_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), cos_tab, sin_tab, pos_offset, static_cast<bf16*>(y), R, L, static_cast<int>(dh),
      inverse);
  return launch_status();
}

}  // extern "C"
