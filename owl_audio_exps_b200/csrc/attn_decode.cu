// Decode-regime attention of the KV-cached causal rollout (owl_wms/nn/attn.py:92-109 with block_mask None;
// owl_wms/sampling/av_caching_v2.py:89-140, audio_caching.py:120-150, inference/causvid_pipeline.py:112-163).
//
// One model evaluation attends <= 128 new query rows (one frame: 65 tokens, or one audio token) to the whole cache plus
// the new rows, with no mask.  Three things make this its own kernel instead of a call of attn.cu:
//
//   * split-KV over the whole machine.  attn.cu gives a B = 1 decode call H CTAs that walk all KV tiles serially
//     (24 CTAs x up to 480 tiles at the dit_v4 sampler's Lk = 61 440).  Here the physical KV tiles of a head are cut
//     into n_splits ranges, one CTA each (B*H*n_splits ~ #SMs); every CTA keeps the attn.cu pipeline (TMA ring ->
//     tcgen05 S = Q K^T -> exp2-domain online softmax, thread = row -> P -> tcgen05 O += P V in TMEM) and writes an
//     unnormalised partial (O, m, l); the LAST CTA of a head to arrive (atomic counter) merges the partials with the
//     log-sum-exp weights and writes bf16 O.  One launch, no second "combine" kernel, the counter resets itself.
//   * the cache is a RING of `cap` physical rows per (batch, head) and its state lives in DEVICE memory
//     (state[0] = rows cached, state[1] = ring start slot, state[2] = RoPE position of the next new token), so the
//     same captured CUDA graph serves every frame of a rollout: appending a frame is state[0] += T, dropping the
//     oldest one (KVCache.truncate) is state[1] += T -- no pointer in the graph changes.  Without a mask the order of
//     the keys is irrelevant to softmax(QK^T)V (RoPE is already applied to the cached K), so the kernel walks the
//     PHYSICAL tiles and masks rows by ring membership.
//   * local layers (attn.py:102-105: only the last local_window frames of [cache, new]) are a membership test too.
//
// owl_qk_norm_rope_decode is the matching producer: QK-RMSNorm + RoPE + head split of the new rows (attn.py:83-89,
// rope.py:43-51) written straight into the ring slots behind the cached rows, position and slot read from `state`.
#include <cuda_bf16.h>
#include <math.h>

#include "attn_common.cuh"
#include "host_common.h"

namespace owl {

struct RingView {
  int pa, len, cap;  // visible physical rows: [pa, pa + len) modulo cap
  __device__ __forceinline__ bool has(int p) const {
    if (p >= cap) return false;
    int d = p - pa;
    if (d < 0) d += cap;
    return d < len;
  }
  __device__ __forceinline__ bool tile_any(int t0) const {  // any visible row in [t0, t0 + 128)?
    if (len <= 0) return false;
    const int t1 = min(t0 + kTile, cap);
    if (pa + len <= cap) return pa < t1 && pa + len > t0;
    return pa < t1 || pa + len - cap > t0;  // wrapped: [pa, cap) or [0, pa + len - cap)
  }
};

__device__ __forceinline__ RingView ring_view(const int* state, int n_new, int vis_rows, int cap) {
  const int n_valid = state[0], start = state[1];
  const int n_tot = n_valid + n_new;             // the new rows are already in their slots
  const int vis_lo = (vis_rows > 0 && n_tot > vis_rows) ? n_tot - vis_rows : 0;
  RingView r;
  r.cap = cap;
  r.len = n_tot - vis_lo;
  r.pa = (start + vis_lo) % cap;
  return r;
}

struct AttnDecodeParams {
  int B, H, Lq, cap, n_splits, tiles_total, tiles_per_split, vis_rows;
  float scale_log2;
  bf16* o;
  long long o_sb, o_sl;
  const int* state;
  float* part_o;       // [B*H][n_splits][128][DH]
  float* part_ml;      // [B*H][n_splits][128][2]
  unsigned* counters;  // [B*H], zero between launches
  // FUSED variant (owl_attn_decode_fused): Q / K / V of the new rows come straight from the qkv GEMM output
  const bf16* qkv;     // [B, Lq, 3, H, DH]
  const float* cos_tab;
  const float* sin_tab;
  long long table_rows;
  bf16* k_cache;       // ring buffers [B, H, cap, DH] (strides c_sb, c_sh): the new rows are written here too
  bf16* v_cache;
  long long c_sb, c_sh;
  float eps;
};

// One head row of qkv (DH bf16 at `src`) -> QK-RMSNorm + RoPE (attn.py:83-89, rope.py:43-51; the arithmetic of
// qk_norm_rope_fwd_kernel) as DH/8 16-byte chunks in the output order [y_even | y_odd]; `rotate` false: raw copy (V).
template <int DH>
__device__ __forceinline__ void norm_rope_row(const bf16* __restrict__ src, const float* __restrict__ cos_row,
                                              const float* __restrict__ sin_row, float eps, bool rotate, uint4 (&out)[DH / 8]) {
  uint4 raw[DH / 8];
#pragma unroll
  for (int c = 0; c < DH / 8; ++c) raw[c] = __ldg(reinterpret_cast<const uint4*>(src) + c);
  if (!rotate) {
#pragma unroll
    for (int c = 0; c < DH / 8; ++c) out[c] = raw[c];
    return;
  }
  float ss = 0.f;
#pragma unroll
  for (int c = 0; c < DH / 8; ++c) {
    const uint32_t w[4] = {raw[c].x, raw[c].y, raw[c].z, raw[c].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float lo = bf16_lo(w[j]), hi = bf16_hi(w[j]);
      ss += lo * lo;
      ss += hi * hi;
    }
  }
  const float rstd = rsqrtf(ss * (1.f / static_cast<float>(DH)) + eps);
  // chunk c of the raw row holds pairs 4c .. 4c+3; their rotated values go to positions 4c.. of each half
  uint32_t y0[DH / 4], y1[DH / 4];   // packed pairs of the even / odd half: index p/2 holds pairs (p, p+1)
#pragma unroll
  for (int c = 0; c < DH / 8; ++c) {
    const float4 c4 = __ldg(reinterpret_cast<const float4*>(cos_row) + c);
    const float4 s4 = __ldg(reinterpret_cast<const float4*>(sin_row) + c);
    const float cs[4] = {c4.x, c4.y, c4.z, c4.w}, sn[4] = {s4.x, s4.y, s4.z, s4.w};
    const uint32_t w[4] = {raw[c].x, raw[c].y, raw[c].z, raw[c].w};
    float e[4], o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float x0 = round_bf16(bf16_lo(w[j]) * rstd), x1 = round_bf16(bf16_hi(w[j]) * rstd);
      e[j] = x0 * cs[j] - x1 * sn[j];
      o[j] = x1 * cs[j] + x0 * sn[j];
    }
    y0[2 * c] = pack_bf16(e[0], e[1]); y0[2 * c + 1] = pack_bf16(e[2], e[3]);
    y1[2 * c] = pack_bf16(o[0], o[1]); y1[2 * c + 1] = pack_bf16(o[2], o[3]);
  }
  // halves: [y0 (DH/2 values = DH/16 chunks) | y1]
#pragma unroll
  for (int c = 0; c < DH / 16; ++c) {
    out[c] = make_uint4(y0[4 * c], y0[4 * c + 1], y0[4 * c + 2], y0[4 * c + 3]);
    out[DH / 16 + c] = make_uint4(y1[4 * c], y1[4 * c + 1], y1[4 * c + 2], y1[4 * c + 3]);
  }
}

template <int DH>
struct DecCfg {
  static constexpr int kStages = 2;
  static constexpr int kTileBytes = kTile * DH * 2;
  static constexpr int kPBytes = kTile * kTile * 2;
  static constexpr int kOffK = kTileBytes;
  static constexpr int kOffV = kOffK + kStages * kTileBytes;
  static constexpr int kOffP = kOffV + kStages * kTileBytes;
  static constexpr int kOffBar = kOffP + kPBytes;
  static constexpr int kSmemBytes = kOffBar + 128;
  static constexpr int kTmemCols = 256;  // S: [0,128), O: [128, 128+DH)
};

static constexpr int kDecThreads = 192;

template <int DH, bool FUSED>
__global__ void __launch_bounds__(kDecThreads, DH == 64 ? 2 : 1)
attn_decode_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const AttnDecodeParams p) {
  using Cfg = DecCfg<DH>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBar);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* s_free = bars + 6;
  uint64_t* p_full = bars + 7;
  uint64_t* pv_done = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
  uint32_t* last_flag = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = static_cast<int>(blockIdx.x) % p.n_splits;
  const int bh = static_cast<int>(blockIdx.x) / p.n_splits;
  const int h = bh % p.H, b = bh / p.H;

  if (threadIdx.x == 0) {
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
  pdl_wait();
  pdl_launch();
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  RingView ring = ring_view(p.state, p.Lq, p.vis_rows, p.cap);
  // FUSED: the new rows are not read back from their ring slots; the LAST split of a head builds them in shared memory
  // as one extra KV tile (it = 0, ring stage 0) and writes them to the ring for the calls that follow
  const bool has_new = FUSED && split == p.n_splits - 1;
  if (FUSED) ring.len -= p.Lq;   // cache tiles: cached rows only (vis_lo was taken over [cache, new]: len >= Lq here or < 0)
  const int tile_lo = split * p.tiles_per_split;
  const int tile_hi = min(tile_lo + p.tiles_per_split, p.tiles_total);
  int n_kv = has_new ? 1 : 0;  // tiles of this split that hold at least one visible row (every role walks the same list)
  for (int t = tile_lo; t < tile_hi; ++t) n_kv += ring.tile_any(t * kTile) ? 1 : 0;

  if (warp == 0) {
    // (whole warp walks the loop, one elected lane issues each batch: no per-instruction ELECT / BRA.U.ANY loop)
    if (n_kv > 0) {
      if (!FUSED) {
        if (elect_one()) {
          mbar_arrive_expect_tx(q_full, Cfg::kTileBytes);
#pragma unroll
          for (int pn = 0; pn < DH / 64; ++pn) tma_load_4d(smem + pn * kPanelBytes, &tmQ, q_full, pn * 64, 0, h, b);
        }
        __syncwarp();
      }
      int it = has_new ? 1 : 0;    // stage 0 of the first round belongs to the tile of new rows
      for (int t = tile_lo; t < tile_hi; ++t) {
        if (!ring.tile_any(t * kTile)) continue;
        const int s = it & 1;
        mbar_wait(&kv_empty[s], ((it >> 1) & 1) ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&kv_full[s], 2 * Cfg::kTileBytes);
#pragma unroll
          for (int pn = 0; pn < DH / 64; ++pn) {
            tma_load_4d(smem + Cfg::kOffK + s * Cfg::kTileBytes + pn * kPanelBytes, &tmK, &kv_full[s], pn * 64, t * kTile, h,
                        b);
            tma_load_4d(smem + Cfg::kOffV + s * Cfg::kTileBytes + pn * kPanelBytes, &tmV, &kv_full[s], pn * 64, t * kTile, h,
                        b);
          }
        }
        __syncwarp();
        ++it;
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
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const bool row_ok = row < p.Lq;
    float m_run = -INFINITY, l_run = 0.f;
    uint8_t* sP = smem + Cfg::kOffP;
    if (FUSED && n_kv > 0) {
      // ---- Q tile (and, in the last split, the K / V tile of the new rows) straight from the qkv GEMM output:
      // QK-RMSNorm + RoPE applied on the way into shared memory (thread = row), swizzled as TMA would have written it
      const long long pos = min(static_cast<long long>(p.state[2]) + row, p.table_rows - 1);
      const float* cr = p.cos_tab + pos * (DH / 2);
      const float* sr = p.sin_tab + pos * (DH / 2);
      const bf16* tok = p.qkv + (static_cast<long long>(b) * p.Lq + row) * (3LL * p.H * DH) + static_cast<long long>(h) * DH;
      uint4 chunk[DH / 8];
      if (row_ok) {
        norm_rope_row<DH>(tok, cr, sr, p.eps, true, chunk);
      } else {
#pragma unroll
        for (int c = 0; c < DH / 8; ++c) chunk[c] = make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int c = 0; c < DH / 8; ++c) *reinterpret_cast<uint4*>(smem + tile_chunk_offset(row, c)) = chunk[c];
      if (has_new) {
        const int slot = static_cast<int>((static_cast<long long>(p.state[1]) + p.state[0] + row) % p.cap);
        bf16* kdst = p.k_cache + b * p.c_sb + h * p.c_sh + static_cast<long long>(slot) * DH;
        bf16* vdst = p.v_cache + b * p.c_sb + h * p.c_sh + static_cast<long long>(slot) * DH;
        if (row_ok) norm_rope_row<DH>(tok + static_cast<long long>(p.H) * DH, cr, sr, p.eps, true, chunk);
#pragma unroll
        for (int c = 0; c < DH / 8; ++c) {
          *reinterpret_cast<uint4*>(smem + Cfg::kOffK + tile_chunk_offset(row, c)) = chunk[c];   // rows >= Lq: zeros
          if (row_ok) reinterpret_cast<uint4*>(kdst)[c] = chunk[c];
        }
        if (row_ok) norm_rope_row<DH>(tok + 2LL * p.H * DH, cr, sr, p.eps, false, chunk);
#pragma unroll
        for (int c = 0; c < DH / 8; ++c) {
          *reinterpret_cast<uint4*>(smem + Cfg::kOffV + tile_chunk_offset(row, c)) = chunk[c];
          if (row_ok) reinterpret_cast<uint4*>(vdst)[c] = chunk[c];
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
      if (threadIdx.x == 64) {
        mbar_arrive(q_full);
        if (has_new) mbar_arrive(&kv_full[0]);
      }
    }
    int it = 0;
    for (int t = has_new ? tile_lo - 1 : tile_lo; t < tile_hi; ++t) {
      const bool new_tile = has_new && t == tile_lo - 1;
      const int j0 = t * kTile;
      if (!new_tile && !ring.tile_any(j0)) continue;
      // every query row sees the same keys: one ballot per 32 keys gives the whole warp the validity word
      uint32_t valid[4];
#pragma unroll
      for (int c = 0; c < 4; ++c)
        valid[c] = __ballot_sync(0xffffffffu, new_tile ? (c * 32 + lane < p.Lq) : ring.has(j0 + c * 32 + lane));
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      uint32_t r[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_x32(tS + lane_addr + c * 32, r[c]);
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
      const float alpha = exp2f(m_run - m_use);
      if (it > 0) {
        mbar_wait(pv_done, (it - 1) & 1);
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
          uint32_t o[16];
          tmem_ld_x16(tO + lane_addr + c * 16, o);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 16; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
          tmem_st_x16(tO + lane_addr + c * 16, o);
        }
        tmem_st_wait();
      }
      l_run = l_run * alpha + rs;
      m_run = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);
      ++it;
    }
    if (n_kv > 0) {
      mbar_wait(pv_done, (n_kv - 1) & 1);
      tc_fence_after();
    }
    bf16* orow = p.o + b * p.o_sb + static_cast<long long>(row) * p.o_sl + static_cast<long long>(h) * DH;
    if (p.n_splits == 1) {
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll
      for (int c = 0; c < DH / 16; ++c) {
        uint32_t o[16];
        if (n_kv > 0) {
          tmem_ld_x16(tO + lane_addr + c * 16, o);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) o[e] = 0u;
        }
        if (row_ok) {
          uint4 v0, v1;
          v0.x = pack_bf16(__uint_as_float(o[0]) * inv, __uint_as_float(o[1]) * inv);
          v0.y = pack_bf16(__uint_as_float(o[2]) * inv, __uint_as_float(o[3]) * inv);
          v0.z = pack_bf16(__uint_as_float(o[4]) * inv, __uint_as_float(o[5]) * inv);
          v0.w = pack_bf16(__uint_as_float(o[6]) * inv, __uint_as_float(o[7]) * inv);
          v1.x = pack_bf16(__uint_as_float(o[8]) * inv, __uint_as_float(o[9]) * inv);
          v1.y = pack_bf16(__uint_as_float(o[10]) * inv, __uint_as_float(o[11]) * inv);
          v1.z = pack_bf16(__uint_as_float(o[12]) * inv, __uint_as_float(o[13]) * inv);
          v1.w = pack_bf16(__uint_as_float(o[14]) * inv, __uint_as_float(o[15]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 16) = v0;
          *reinterpret_cast<uint4*>(orow + c * 16 + 8) = v1;
        }
      }
    } else {
      // partial (O unnormalised, m, l) of this split -> workspace
      const long long slot = (static_cast<long long>(bh) * p.n_splits + split) * kTile + row;
      float* po = p.part_o + slot * DH;
      if (n_kv > 0) {
#pragma unroll
        for (int c = 0; c < DH / 16; ++c) {
          uint32_t o[16];
          tmem_ld_x16(tO + lane_addr + c * 16, o);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int g = 0; g < 4; ++g)
              *reinterpret_cast<uint4*>(po + c * 16 + g * 4) = make_uint4(o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
          }
        }
      }
      if (row_ok) *reinterpret_cast<float2*>(p.part_ml + slot * 2) = make_float2(m_run, l_run);
      __threadfence();
      named_bar_sync(1, 128);
      if (threadIdx.x == 64) {
        const unsigned prev = atomicAdd(p.counters + bh, 1u);
        const unsigned last = prev == static_cast<unsigned>(p.n_splits - 1) ? 1u : 0u;
        if (last) {
          __threadfence();
          p.counters[bh] = 0u;   // every split of this head has arrived: nobody reads the counter again
        }
        *last_flag = last;
      }
      named_bar_sync(1, 128);
      if (*last_flag != 0 && row_ok) {
        const long long base = static_cast<long long>(bh) * p.n_splits * kTile + row;
        float M = -INFINITY;
        for (int s = 0; s < p.n_splits; ++s) {
          const float2 ml = __ldcg(reinterpret_cast<const float2*>(p.part_ml + (base + static_cast<long long>(s) * kTile) * 2));
          if (ml.y > 0.f) M = fmaxf(M, ml.x);
        }
        float Lsum = 0.f;
#pragma unroll 1
        for (int half = 0; half < DH / 64; ++half) {
          float acc[64];
#pragma unroll
          for (int e = 0; e < 64; ++e) acc[e] = 0.f;
          float Lh = 0.f;
          for (int s = 0; s < p.n_splits; ++s) {
            const long long sl = base + static_cast<long long>(s) * kTile;
            const float2 ml = __ldcg(reinterpret_cast<const float2*>(p.part_ml + sl * 2));
            if (!(ml.y > 0.f)) continue;
            const float w = exp2f(ml.x - M);
            Lh += ml.y * w;
            const float4* src = reinterpret_cast<const float4*>(p.part_o + sl * DH + half * 64);
#pragma unroll
            for (int g = 0; g < 16; ++g) {
              const float4 v = __ldcg(src + g);
              acc[4 * g] = fmaf(w, v.x, acc[4 * g]);
              acc[4 * g + 1] = fmaf(w, v.y, acc[4 * g + 1]);
              acc[4 * g + 2] = fmaf(w, v.z, acc[4 * g + 2]);
              acc[4 * g + 3] = fmaf(w, v.w, acc[4 * g + 3]);
            }
          }
          Lsum = Lh;
          const float inv = Lsum > 0.f ? 1.f / Lsum : 0.f;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            uint4 v;
            v.x = pack_bf16(acc[8 * g] * inv, acc[8 * g + 1] * inv);
            v.y = pack_bf16(acc[8 * g + 2] * inv, acc[8 * g + 3] * inv);
            v.z = pack_bf16(acc[8 * g + 4] * inv, acc[8 * g + 5] * inv);
            v.w = pack_bf16(acc[8 * g + 6] * inv, acc[8 * g + 7] * inv);
            *reinterpret_cast<uint4*>(orow + half * 64 + g * 8) = v;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// QK-RMSNorm + RoPE + head split of the new rows into ring slots (see the header comment).  Same arithmetic and thread
// mapping as qk_norm_rope_fwd_kernel (qk_rope.cu): one CTA per token, one thread = 8 features, dh/8-lane shuffle groups.
__global__ void __launch_bounds__(128)
qk_norm_rope_decode_kernel(const bf16* __restrict__ qkv, const float* __restrict__ cos_tab, const float* __restrict__ sin_tab,
                           int64_t table_rows, const int* __restrict__ state, bf16* __restrict__ q, int64_t q_sb,
                           int64_t q_sh, bf16* __restrict__ k, bf16* __restrict__ v, int64_t c_sb, int64_t c_sh, int cap,
                           int64_t L, int H, int dh, float eps) {
  pdl_prologue();
  const int lgG = dh == 64 ? 3 : 4, G = 1 << lgG;
  const int HG = H << lgG, per_tok = 3 * HG;
  const int64_t token = blockIdx.x;
  const int64_t b = token / L, l = token - b * L;
  const bf16* src = qkv + token * per_tok * 8;
  const int half = dh >> 1;
  const int n_valid = state[0], start = state[1];
  int64_t pos = static_cast<int64_t>(state[2]) + l;
  if (pos >= table_rows) pos = table_rows - 1;   // host checks the bound before launch; never read past the table
  const int slot = static_cast<int>((static_cast<int64_t>(start) + n_valid + l) % cap);
  const float inv_dh = 1.f / static_cast<float>(dh);
#pragma unroll 3
  for (int c0 = 0; c0 < per_tok; c0 += 128) {
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
    for (int o = G >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (!active) continue;
    if (which == 2) {
      *reinterpret_cast<uint4*>(v + b * c_sb + h * c_sh + static_cast<int64_t>(slot) * dh + chunk * 8) = raw;
      continue;
    }
    const float rstd = rsqrtf(ss * inv_dh + eps);
    const float4 c4 = __ldg(reinterpret_cast<const float4*>(cos_tab + pos * half + chunk * 4));
    const float4 s4 = __ldg(reinterpret_cast<const float4*>(sin_tab + pos * half + chunk * 4));
    const float cs[4] = {c4.x, c4.y, c4.z, c4.w};
    const float sn[4] = {s4.x, s4.y, s4.z, s4.w};
    float y0[4], y1[4];
#pragma unroll
    for (int pp = 0; pp < 4; ++pp) {
      const float x0 = round_bf16(f[2 * pp] * rstd), x1 = round_bf16(f[2 * pp + 1] * rstd);
      y0[pp] = x0 * cs[pp] - x1 * sn[pp];
      y1[pp] = x1 * cs[pp] + x0 * sn[pp];
    }
    bf16* dst = which == 0 ? q + b * q_sb + h * q_sh + l * dh : k + b * c_sb + h * c_sh + static_cast<int64_t>(slot) * dh;
    uint2 o0, o1;
    o0.x = pack_bf16(y0[0], y0[1]); o0.y = pack_bf16(y0[2], y0[3]);
    o1.x = pack_bf16(y1[0], y1[1]); o1.y = pack_bf16(y1[2], y1[3]);
    *reinterpret_cast<uint2*>(dst + chunk * 4) = o0;
    *reinterpret_cast<uint2*>(dst + half + chunk * 4) = o1;
  }
}

// state <- state after committing `commit_rows` new rows and ejecting the `drop_rows` oldest ones
// (KVCache.update + KVCache.truncate, kv_cache.py:50-75); one thread.
__global__ void decode_state_advance_kernel(int* state, int commit_rows, int drop_rows, int cap) {
  pdl_prologue();
  int n_valid = state[0] + commit_rows, start = state[1];
  const int pos = state[2] + commit_rows;
  if (drop_rows > n_valid) drop_rows = n_valid;
  start = (start + drop_rows) % cap;
  n_valid -= drop_rows;
  state[0] = n_valid;
  state[1] = start;
  state[2] = pos;
}

template <int DH, bool FUSED>
static int launch_attn_decode(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                              const AttnDecodeParams& p, cudaStream_t s) {
  using Cfg = DecCfg<DH>;
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    return static_cast<int>(cudaFuncSetAttribute(attn_decode_kernel<DH, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 Cfg::kSmemBytes));
  });
  if (rc0 != OWL_OK) return rc0;
  const long long grid = static_cast<long long>(p.B) * p.H * p.n_splits;
  (void)launch_k(attn_decode_kernel<DH, FUSED>, dim3(static_cast<unsigned>(grid)), dim3(kDecThreads), Cfg::kSmemBytes, s, tmQ,
                 tmK, tmV, p);
  return launch_status();
}

static constexpr int64_t kDecCounterBytes = 4096;

static int decode_splits(int64_t B, int64_t H, int tiles_total) {
  int s = static_cast<int>((sm_count() + B * H - 1) / (B * H));
  if (s > tiles_total) s = tiles_total;
  if (s > 32) s = 32;
  return s < 1 ? 1 : s;
}

}  // namespace owl

using namespace owl;

extern "C" {

int64_t owl_attn_decode_workspace_bytes(int64_t B, int64_t H, int64_t cap, int64_t dh) {
  const int tiles = static_cast<int>((cap + kTile - 1) / kTile);
  const int64_t s = decode_splits(B, H, tiles);
  return kDecCounterBytes + B * H * s * kTile * (dh + 2) * 4;
}

int owl_attn_decode(const void* q, int64_t q_sb, int64_t q_sh, const void* k_cache, const void* v_cache, int64_t c_sb,
                    int64_t c_sh, int64_t cap, const int32_t* state, int64_t vis_rows, void* o, int64_t o_sb, int64_t o_sl,
                    int64_t B, int64_t H, int64_t Lq, int64_t dh, void* workspace, int64_t workspace_bytes,
                    owl_stream_t stream) {
  if (!q || !k_cache || !v_cache || !state || !o || B <= 0 || H <= 0 || Lq <= 0 || cap <= 0 || vis_rows < 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (Lq > kTile || cap > (1 << 30) || B * H > (1 << 20)) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(o) || o_sb % 8 || o_sl % 8) return OWL_ERR_ALIGN;
  AttnDecodeParams p{};
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.cap = static_cast<int>(cap);
  p.tiles_total = static_cast<int>((cap + kTile - 1) / kTile);
  p.n_splits = decode_splits(B, H, p.tiles_total);
  p.tiles_per_split = (p.tiles_total + p.n_splits - 1) / p.n_splits;
  p.n_splits = (p.tiles_total + p.tiles_per_split - 1) / p.tiles_per_split;
  p.vis_rows = static_cast<int>(vis_rows);
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(dh));
  p.o = static_cast<bf16*>(o); p.o_sb = o_sb; p.o_sl = o_sl;
  p.state = state;
  if (p.n_splits > 1) {
    if (!workspace || !aligned16(workspace) || workspace_bytes < owl_attn_decode_workspace_bytes(B, H, cap, dh))
      return OWL_ERR_BAD_ARG;
    // the counters sit at a FIXED place: the partials of an earlier launch with another geometry must never be read
    // as arrival counts (more than one split only when B*H < #SMs, so B*H counters always fit)
    if (B * H * 4 > kDecCounterBytes) return OWL_ERR_UNSUPPORTED;
    const int64_t n_part = B * H * p.n_splits * kTile;
    p.counters = static_cast<unsigned*>(workspace);
    p.part_o = reinterpret_cast<float*>(static_cast<char*>(workspace) + kDecCounterBytes);
    p.part_ml = p.part_o + n_part * dh;
  }
  CUtensorMap tmQ, tmK, tmV;
  int rc;
  if ((rc = make_tmap_4d_bf16(&tmQ, q, dh, Lq, H, B, dh, q_sh, q_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k_cache, dh, cap, H, B, dh, c_sh, c_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v_cache, dh, cap, H, B, dh, c_sh, c_sb, 64, kTile)) != OWL_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return dh == 64 ? launch_attn_decode<64, false>(tmQ, tmK, tmV, p, s) : launch_attn_decode<128, false>(tmQ, tmK, tmV, p, s);
}

int owl_attn_decode_fused(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t table_rows, void* k_cache,
                          void* v_cache, int64_t c_sb, int64_t c_sh, int64_t cap, const int32_t* state, int64_t vis_rows,
                          void* o, int64_t o_sb, int64_t o_sl, int64_t B, int64_t H, int64_t Lq, int64_t dh, float eps,
                          void* workspace, int64_t workspace_bytes, owl_stream_t stream) {
  if (!qkv || !cos_tab || !sin_tab || !k_cache || !v_cache || !state || !o || B <= 0 || H <= 0 || Lq <= 0 || cap < Lq ||
      vis_rows < 0 || table_rows <= 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (Lq > kTile || cap > (1 << 30) || B * H > (1 << 20)) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(qkv) || !aligned16(o) || !aligned16(k_cache) || !aligned16(v_cache) || !aligned16(cos_tab) ||
      !aligned16(sin_tab) || o_sb % 8 || o_sl % 8 || c_sb % 8 || c_sh % 8)
    return OWL_ERR_ALIGN;
  AttnDecodeParams p{};
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.cap = static_cast<int>(cap);
  p.tiles_total = static_cast<int>((cap + kTile - 1) / kTile);
  p.n_splits = decode_splits(B, H, p.tiles_total);
  p.tiles_per_split = (p.tiles_total + p.n_splits - 1) / p.n_splits;
  p.n_splits = (p.tiles_total + p.tiles_per_split - 1) / p.tiles_per_split;
  p.vis_rows = static_cast<int>(vis_rows);
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(dh));
  p.o = static_cast<bf16*>(o); p.o_sb = o_sb; p.o_sl = o_sl;
  p.state = state;
  p.qkv = static_cast<const bf16*>(qkv); p.cos_tab = cos_tab; p.sin_tab = sin_tab; p.table_rows = table_rows;
  p.k_cache = static_cast<bf16*>(k_cache); p.v_cache = static_cast<bf16*>(v_cache); p.c_sb = c_sb; p.c_sh = c_sh;
  p.eps = eps;
  if (p.n_splits > 1) {
    if (!workspace || !aligned16(workspace) || workspace_bytes < owl_attn_decode_workspace_bytes(B, H, cap, dh))
      return OWL_ERR_BAD_ARG;
    if (B * H * 4 > kDecCounterBytes) return OWL_ERR_UNSUPPORTED;
    const int64_t n_part = B * H * p.n_splits * kTile;
    p.counters = static_cast<unsigned*>(workspace);
    p.part_o = reinterpret_cast<float*>(static_cast<char*>(workspace) + kDecCounterBytes);
    p.part_ml = p.part_o + n_part * dh;
  }
  CUtensorMap tmK, tmV;
  int rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k_cache, dh, cap, H, B, dh, c_sh, c_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v_cache, dh, cap, H, B, dh, c_sh, c_sb, 64, kTile)) != OWL_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return dh == 64 ? launch_attn_decode<64, true>(tmK, tmK, tmV, p, s) : launch_attn_decode<128, true>(tmK, tmK, tmV, p, s);
}

int owl_qk_norm_rope_decode(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t table_rows,
                            const int32_t* state, void* q, int64_t q_sb, int64_t q_sh, void* k_cache, void* v_cache,
                            int64_t c_sb, int64_t c_sh, int64_t cap, int64_t B, int64_t L, int64_t H, int64_t dh, float eps,
                            owl_stream_t stream) {
  if (!qkv || !cos_tab || !sin_tab || !state || !q || !k_cache || !v_cache || B <= 0 || L <= 0 || H <= 0 || cap < L ||
      table_rows <= 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(qkv) || !aligned16(q) || !aligned16(k_cache) || !aligned16(v_cache) || !aligned16(cos_tab) ||
      !aligned16(sin_tab) || (q_sb | q_sh | c_sb | c_sh) % 8)
    return OWL_ERR_ALIGN;
  if (B * L > INT32_MAX || 3 * H * dh > (1 << 24) || cap > (1 << 30)) return OWL_ERR_UNSUPPORTED;
  (void)launch_k(qk_norm_rope_decode_kernel, dim3(static_cast<unsigned>(B * L)), dim3(128), 0,
                 static_cast<cudaStream_t>(stream), static_cast<const bf16*>(qkv), cos_tab, sin_tab, table_rows, state,
                 static_cast<bf16*>(q), q_sb, q_sh, static_cast<bf16*>(k_cache), static_cast<bf16*>(v_cache), c_sb, c_sh,
                 static_cast<int>(cap), L, static_cast<int>(H), static_cast<int>(dh), eps);
  return launch_status();
}

int owl_decode_state_advance(int32_t* state, int64_t commit_rows, int64_t drop_rows, int64_t cap, owl_stream_t stream) {
  if (!state || commit_rows < 0 || drop_rows < 0 || cap <= 0) return OWL_ERR_BAD_ARG;
  (void)launch_k(decode_state_advance_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state,
                 static_cast<int>(commit_rows), static_cast<int>(drop_rows), static_cast<int>(cap));
  return launch_status();
}

}  // extern "C"
