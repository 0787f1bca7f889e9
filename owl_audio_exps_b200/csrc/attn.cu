// Frame-block-causal attention on tcgen05 / TMEM, fed by TMA (owl_attn_fwd, owl_attn_bwd).
//
// Forward: one CTA per (batch, head, 128-query tile); 6 warps:
//   warp 0     TMA producer: Q once, then a 2-deep ring of {K tile, V tile}
//   warp 1     MMA issuer + TMEM owner: S = Q K^T (128x128xDH) and O += P V (128xDHx128)
//   warps 2-5  softmax, thread = query row: tcgen05.ld S, frame-block mask as a 32-bit validity
//              word per 32 keys, online softmax in the exp2 domain, P (bf16) into swizzled smem,
//              O rescale in TMEM only when some row's running max moved
// KV tiles outside the union of the rows' causal/window ranges are never loaded (tile-granular
// skipping of the reference's BlockMask, attn.py:43-62), and tiles strictly inside every row's
// range skip the per-element test.  Two CTAs are co-resident per SM at DH=64 (TMEM 2 x 256
// columns, smem 2 x 112 KB) so one CTA's softmax overlaps the other's MMAs.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "attn_common.cuh"
#include "host_common.h"

namespace owl {

struct AttnFwdParams {
  int B, H, Lq, Lk, n_qtiles;
  float scale_log2;  // log2(e) / sqrt(dh)
  bf16* o;
  long long o_sb, o_sl;
  float* lse;
  MaskParams m;
};

template <int DH>
struct FwdCfg {
  static constexpr int kStages = 2;
  static constexpr int kTileBytes = kTile * DH * 2;  // Q, K or V tile
  static constexpr int kPBytes = kTile * kTile * 2;
  static constexpr int kOffK = kTileBytes;
  static constexpr int kOffV = kOffK + kStages * kTileBytes;
  static constexpr int kOffP = kOffV + kStages * kTileBytes;
  static constexpr int kOffBar = kOffP + kPBytes;
  static constexpr int kSmemBytes = kOffBar + 128;
  static constexpr int kTmemCols = 256;  // S: [0,128), O: [128, 128+DH)
};

static constexpr int kFwdThreads = 192;

template <int DH>
__global__ void __launch_bounds__(kFwdThreads, DH == 64 ? 2 : 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const AttnFwdParams p) {
  using Cfg = FwdCfg<DH>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBar);
  uint64_t* q_full = bars;          // 1
  uint64_t* kv_full = bars + 1;     // 2
  uint64_t* kv_empty = bars + 3;    // 2
  uint64_t* s_full = bars + 5;      // MMA -> softmax
  uint64_t* s_free = bars + 6;      // softmax -> MMA (128)
  uint64_t* p_full = bars + 7;      // softmax -> MMA (128)
  uint64_t* pv_done = bars + 8;     // MMA -> softmax
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late, causal) query tiles first
  const int qt = p.n_qtiles - 1 - static_cast<int>(blockIdx.x % p.n_qtiles);
  const int bh = static_cast<int>(blockIdx.x / p.n_qtiles);
  const int h = bh % p.H, b = bh / p.H;
  const int i0 = qt * kTile;
  const int i_last = min(i0 + kTile, p.Lq) - 1;
  int t0, t1;
  kv_tile_range_doc(p.m, b, i0, i_last, p.Lk, t0, t1);
  const int n_kv = t1 - t0;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) {
      printf("owl: attention smem base not 1024-aligned\n");
      __trap();
    }
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 128);
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // setup above overlapped the previous kernel; global memory (TMA loads, lse/delta, stores) from here on
  pdl_launch();
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  // Producer / issuer: the whole warp walks the loop, one elected lane issues each batch (a divergent `lane == 0` region
  // makes ptxas wrap every UTMALDG / UTCHMMA in an ELECT / BRA.U.ANY loop, ~100 clk per instruction).
  if (warp == 0) {
    if (n_kv > 0) {
      if (elect_one()) {
        mbar_arrive_expect_tx(q_full, Cfg::kTileBytes);
#pragma unroll
        for (int pn = 0; pn < DH / 64; ++pn) tma_load_4d(smem + pn * kPanelBytes, &tmQ, q_full, pn * 64, i0, h, b);
      }
      __syncwarp();
      for (int it = 0; it < n_kv; ++it) {
        const int s = it & 1;
        mbar_wait(&kv_empty[s], ((it >> 1) & 1) ^ 1);
        const int j0 = (t0 + it) * kTile;
        if (elect_one()) {
          mbar_arrive_expect_tx(&kv_full[s], 2 * Cfg::kTileBytes);
#pragma unroll
          for (int pn = 0; pn < DH / 64; ++pn) {
            tma_load_4d(smem + Cfg::kOffK + s * Cfg::kTileBytes + pn * kPanelBytes, &tmK, &kv_full[s], pn * 64, j0, h, b);
            tma_load_4d(smem + Cfg::kOffV + s * Cfg::kTileBytes + pn * kPanelBytes, &tmV, &kv_full[s], pn * 64, j0, h, b);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    if (n_kv > 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(kTile, DH, false, true);
      const uint32_t sQ = smem_u32(smem), sP = smem_u32(smem + Cfg::kOffP);
      auto issue_s = [&](int it) {
        const int s = it & 1;
        mbar_wait(&kv_full[s], (it >> 1) & 1);
        if (it > 0) mbar_wait(s_free, (it - 1) & 1);
        tc_fence_after();
        const uint32_t sK = smem_u32(smem + Cfg::kOffK + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tS, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc_s, k > 0);
          umma_commit(s_full);
        }
        __syncwarp();
      };
      mbar_wait(q_full, 0);
      issue_s(0);
      for (int it = 0; it < n_kv; ++it) {
        const int s = it & 1;
        if (it + 1 < n_kv) issue_s(it + 1);
        mbar_wait(p_full, it & 1);
        tc_fence_after();
        const uint32_t sV = smem_u32(smem + Cfg::kOffV + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kTile / 16; ++k)
            umma_bf16_ss(tO, desc_kmajor(sP, k), desc_mnmajor(sV, k), idesc_o, (it > 0 || k > 0) ? 1u : 0u);
          umma_commit(&kv_empty[s]);
          umma_commit(pv_done);
        }
        __syncwarp();
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int i = i0 + row;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    int lo, hi;
    mask_row_range(p.m, min(i, p.Lq - 1), p.Lk, lo, hi);
    const int* doc_row = p.m.doc != nullptr ? p.m.doc + b * p.m.doc_ld : nullptr;
    const int dq = doc_row != nullptr ? __ldg(doc_row + (min(i, p.Lq - 1) + p.m.q_offset) / p.m.T) : 0;
    float m_run = -INFINITY, l_run = 0.f;
    uint8_t* sP = smem + Cfg::kOffP;
    for (int it = 0; it < n_kv; ++it) {
      const int j0 = (t0 + it) * kTile;
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      const bool full = doc_row == nullptr && lo <= j0 && j0 + kTile <= hi;
      // One pass over TMEM: the whole score row (128 fp32) is pulled into registers with a single wait, and the S
      // buffer is handed back to the MMA warp at once, so QK^T of the next KV tile runs under this tile's softmax.
      uint32_t valid[4];
      uint32_t r[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_x32(tS + lane_addr + c * 32, r[c]);
#pragma unroll
      for (int c = 0; c < 4; ++c)
        valid[c] = full ? 0xffffffffu : chunk_valid_bits(j0 + c * 32, lo, hi, doc_row, dq, p.m.T);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if ((valid[c] >> e) & 1u) mx = fmaxf(mx, __uint_as_float(r[c][e]));
      }
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = m_new == -INFINITY ? 0.f : m_new;
      const float alpha = exp2f(m_run - m_use);  // m_run = -inf -> 0
      if (it > 0) {
        mbar_wait(pv_done, (it - 1) & 1);  // P buffer free, O stable
        tc_fence_after();
      }
      float rs = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float p0 = ((valid[c] >> e) & 1u) ? exp2f(__uint_as_float(r[c][e]) * p.scale_log2 - m_use) : 0.f;
          const float p1 = ((valid[c] >> (e + 1)) & 1u) ? exp2f(__uint_as_float(r[c][e + 1]) * p.scale_log2 - m_use) : 0.f;
          rs += p0 + p1;
          pk[e >> 1] = pack_bf16(p0, p1);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
          *reinterpret_cast<uint4*>(sP + tile_chunk_offset(row, c * 4 + g)) =
              make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
      }
      if (it > 0 && !__all_sync(0xffffffffu, alpha == 1.f)) {
#pragma unroll
        for (int c = 0; c < DH / 16; ++c) {
          uint32_t r[16];
          tmem_ld_x16(tO + lane_addr + c * 16, r);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 16; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
          tmem_st_x16(tO + lane_addr + c * 16, r);
        }
        tmem_st_wait();
      }
      l_run = l_run * alpha + rs;
      m_run = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    // epilogue: O / l -> global (token-major [B, Lq, H*DH]), lse
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    if (n_kv > 0) {
      mbar_wait(pv_done, (n_kv - 1) & 1);
      tc_fence_after();
    }
    bf16* orow = p.o + b * p.o_sb + static_cast<long long>(i) * p.o_sl + static_cast<long long>(h) * DH;
#pragma unroll
    for (int c = 0; c < DH / 16; ++c) {
      uint32_t r[16];
      if (n_kv > 0) {
        tmem_ld_x16(tO + lane_addr + c * 16, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int e = 0; e < 16; ++e) r[e] = 0u;
      }
      if (i < p.Lq) {
        uint4 v0, v1;
        v0.x = pack_bf16(__uint_as_float(r[0]) * inv, __uint_as_float(r[1]) * inv);
        v0.y = pack_bf16(__uint_as_float(r[2]) * inv, __uint_as_float(r[3]) * inv);
        v0.z = pack_bf16(__uint_as_float(r[4]) * inv, __uint_as_float(r[5]) * inv);
        v0.w = pack_bf16(__uint_as_float(r[6]) * inv, __uint_as_float(r[7]) * inv);
        v1.x = pack_bf16(__uint_as_float(r[8]) * inv, __uint_as_float(r[9]) * inv);
        v1.y = pack_bf16(__uint_as_float(r[10]) * inv, __uint_as_float(r[11]) * inv);
        v1.z = pack_bf16(__uint_as_float(r[12]) * inv, __uint_as_float(r[13]) * inv);
        v1.w = pack_bf16(__uint_as_float(r[14]) * inv, __uint_as_float(r[15]) * inv);
        *reinterpret_cast<uint4*>(orow + c * 16) = v0;
        *reinterpret_cast<uint4*>(orow + c * 16 + 8) = v1;
      }
    }
    if (p.lse != nullptr && i < p.Lq)
      p.lse[(static_cast<long long>(b) * p.H + h) * p.Lq + i] =
          l_run > 0.f ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int DH>
static int launch_attn_fwd(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const AttnFwdParams& p, cudaStream_t s) {
  using Cfg = FwdCfg<DH>;
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    return static_cast<int>(
        cudaFuncSetAttribute(attn_fwd_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  });
  if (rc0 != OWL_OK) return rc0;
  const long long grid = static_cast<long long>(p.B) * p.H * p.n_qtiles;
  (void)launch_k(attn_fwd_kernel<DH>, dim3(static_cast<unsigned>(grid)), dim3(kFwdThreads), Cfg::kSmemBytes, s, tmQ, tmK, tmV, p);
  return launch_status();
}

int fill_mask_params(const owl_attn_mask* mask, MaskParams* m) {
  if (!mask || mask->tokens_per_frame <= 0 || mask->q_offset < 0) return OWL_ERR_BAD_ARG;
  m->T = mask->tokens_per_frame;
  m->W = mask->window_frames;
  m->q_offset = mask->q_offset;
  m->causal = mask->causal;
  m->doc = mask->doc_id;
  m->doc_ld = mask->doc_ld;
  const bool ext = mask->doc_id != nullptr && mask->doc_first != nullptr && mask->doc_last != nullptr &&
                   mask->doc_breaks != nullptr;
  m->doc_first = ext ? mask->doc_first : nullptr;
  m->doc_last = ext ? mask->doc_last : nullptr;
  m->doc_breaks = ext ? mask->doc_breaks : nullptr;
  return OWL_OK;
}

int attn_fwd2_dispatch(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                       const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       cudaStream_t s);  // attn_fwd2.cu

// Per-frame document extents for tile skipping (see MaskParams): one thread per frame scans its batch row, which
// sits in shared memory when it fits (F <= 8192).
__global__ void __launch_bounds__(256)
attn_doc_extents_kernel(const int* __restrict__ doc, long long ld, int F, int* __restrict__ first,
                        int* __restrict__ last, int* __restrict__ breaks) {
  extern __shared__ int s_doc[];
  pdl_prologue();
  const int b = blockIdx.y;
  const int* row = doc + b * ld;
  const bool in_smem = F <= 8192;
  if (in_smem) {
    for (int g = threadIdx.x; g < F; g += blockDim.x) s_doc[g] = row[g];
    __syncthreads();
    row = s_doc;
  }
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const int id = row[f];
  int lo = f, hi = f;
  for (int g = 0; g < f; ++g)
    if (row[g] == id) { lo = g; break; }
  for (int g = F - 1; g > f; --g)
    if (row[g] == id) { hi = g; break; }
  first[b * ld + f] = lo;
  last[b * ld + f] = hi;
  if (f > 0 && row[f - 1] != id && lo != f) atomicAdd(breaks + b, 1);  // this id was open before and re-opens here
}

}  // namespace owl

using namespace owl;

extern "C" {

int owl_attn_fwd(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                 const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                 int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                 owl_stream_t stream) {
  if (!q || !k || !v || !o || B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (Lq > (1 << 30) || Lk > (1 << 30) || B * H * ((Lq + kTile - 1) / kTile) > INT32_MAX) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(o) || o_sb % 8 || o_sl % 8) return OWL_ERR_ALIGN;
  {
    // Lq > 128: the two-tile persistent kernel (attn_fwd2.cu).  OWL_ATTN_FWD=1 forces the single-tile kernel (A/B aid).
    static const int force = [] { const char* e = getenv("OWL_ATTN_FWD"); return e ? atoi(e) : 0; }();
    if ((Lq > kTile && force != 1) || force == 2)
      return attn_fwd2_dispatch(q, q_sb, q_sh, k, k_sb, k_sh, v, v_sb, v_sh, o, o_sb, o_sl, lse, B, H, Lq, Lk, dh, mask,
                                static_cast<cudaStream_t>(stream));
  }
  AttnFwdParams p{};
  int rc = fill_mask_params(mask, &p.m);
  if (rc != OWL_OK) return rc;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.n_qtiles = static_cast<int>((Lq + kTile - 1) / kTile);
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(dh));
  p.o = static_cast<bf16*>(o); p.o_sb = o_sb; p.o_sl = o_sl; p.lse = lse;
  CUtensorMap tmQ, tmK, tmV;
  if ((rc = make_tmap_4d_bf16(&tmQ, q, dh, Lq, H, B, dh, q_sh, q_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k, dh, Lk, H, B, dh, k_sh, k_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v, dh, Lk, H, B, dh, v_sh, v_sb, 64, kTile)) != OWL_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return dh == 64 ? launch_attn_fwd<64>(tmQ, tmK, tmV, p, s) : launch_attn_fwd<128>(tmQ, tmK, tmV, p, s);
}

int owl_attn_doc_extents(const int32_t* doc_id, int64_t doc_ld, int64_t B, int64_t F, int32_t* first, int32_t* last,
                         int32_t* breaks, owl_stream_t stream) {
  if (!doc_id || !first || !last || !breaks || B <= 0 || F <= 0 || doc_ld < F) return OWL_ERR_BAD_ARG;
  if (B > 65535 || F > (1 << 30)) return OWL_ERR_UNSUPPORTED;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(breaks, 0, static_cast<size_t>(B) * sizeof(int32_t), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  LaunchOpts lo;
  lo.pdl = false;  // follows a memset node, not a kernel
  const size_t smem = F <= 8192 ? static_cast<size_t>(F) * sizeof(int) : 0;
  e = launch_kx(lo, attn_doc_extents_kernel, dim3(static_cast<unsigned>((F + 255) / 256), static_cast<unsigned>(B)),
                dim3(256), smem, s, doc_id, static_cast<long long>(doc_ld), static_cast<int>(F), first, last, breaks);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return static_cast<int>(e);
  }
  return launch_status();
}

}  // extern "C"
