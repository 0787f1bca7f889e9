// Attention backward on tcgen05 / TMEM (owl_attn_bwd): three launches, no atomics, deterministic.
//   1. delta[b,h,i] = sum_d dO * O                       (memory-bound pre-pass)
//   2. dq kernel : CTA per 128-query tile, loops over its KV tiles:
//        S = Q K^T, dP = dO V^T  -> threads: P = exp2(S c - lse), dS = P (dP - delta) / sqrt(dh) -> smem
//        dQ += dS K
//   3. dkv kernel: CTA per 128-key tile, loops over the query tiles that can see it:
//        S, dP as above -> threads write P and dS ([q][k] panels)
//        dV += P^T dO,  dK += dS^T Q       (A operands read MN-major from the same panels)
// Same warp roles, barriers and mask logic as the forward kernel (attn.cu); tiles outside the
// causal/window range are skipped in both loops.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "attn_common.cuh"
#include "host_common.h"

namespace owl {

int fill_mask_params(const owl_attn_mask* mask, MaskParams* m);  // attn.cu
int attn_bwd2_dispatch(const void* dout, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb, int64_t q_sh,
                       const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb, int64_t v_sh,
                       const float* lse, const float* delta, float* dq_acc, void* dq, int64_t dq_sb, int64_t dq_sh,
                       void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh, int64_t B,
                       int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask, cudaStream_t s);  // attn_bwd2.cu

struct AttnBwdParams {
  int B, H, Lq, Lk, n_qtiles, n_ktiles;
  float scale_log2, scale;
  const float* lse;
  const float* delta;
  bf16 *dq, *dk, *dv;
  long long dq_sb, dq_sh, dk_sb, dk_sh, dv_sb, dv_sh;
  MaskParams m;
};

// ------------------------------------------------------------------------------------------------
// delta = rowsum(dO * O); dO, O token-major (b, i, h, :)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
attn_delta_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ o, long long o_sb, long long o_sl,
                  float* __restrict__ delta, int B, int H, int Lq, int dh) {
  pdl_prologue();
  const int G = dh >> 3;
  const long long total = static_cast<long long>(B) * Lq * H * G;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool active = t < total;
  const long long tt = active ? t : 0;
  const int chunk = static_cast<int>(tt % G);
  const int h = static_cast<int>((tt / G) % H);
  const long long bi = tt / (static_cast<long long>(G) * H);
  const int i = static_cast<int>(bi % Lq);
  const int b = static_cast<int>(bi / Lq);
  float s = 0.f;
  if (active) {
    const long long off = b * o_sb + static_cast<long long>(i) * o_sl + static_cast<long long>(h) * dh + chunk * 8;
    const uint4 g = __ldg(reinterpret_cast<const uint4*>(dout + off));
    const uint4 y = __ldg(reinterpret_cast<const uint4*>(o + off));
    s = bf16_lo(g.x) * bf16_lo(y.x) + bf16_hi(g.x) * bf16_hi(y.x) + bf16_lo(g.y) * bf16_lo(y.y) +
        bf16_hi(g.y) * bf16_hi(y.y) + bf16_lo(g.z) * bf16_lo(y.z) + bf16_hi(g.z) * bf16_hi(y.z) +
        bf16_lo(g.w) * bf16_lo(y.w) + bf16_hi(g.w) * bf16_hi(y.w);
  }
  for (int ofs = G >> 1; ofs > 0; ofs >>= 1) s += __shfl_xor_sync(0xffffffffu, s, ofs);
  if (active && chunk == 0) delta[(static_cast<long long>(b) * H + h) * Lq + i] = s;
}

static constexpr int kBwdThreads = 192;

// One 32-key chunk of a score tile: P and dS (both bf16-packed) from S and dP accumulators.
__device__ __forceinline__ void pds_chunk(const uint32_t (&rs)[32], const uint32_t (&rp)[32], uint32_t valid,
                                          float scale_log2, float lse2, float delta, float scale, uint32_t (&pk)[16],
                                          uint32_t (&dk)[16]) {
#pragma unroll
  for (int e = 0; e < 32; e += 2) {
    const float p0 = ((valid >> e) & 1u) ? exp2f(__uint_as_float(rs[e]) * scale_log2 - lse2) : 0.f;
    const float p1 = ((valid >> (e + 1)) & 1u) ? exp2f(__uint_as_float(rs[e + 1]) * scale_log2 - lse2) : 0.f;
    const float d0 = p0 * (__uint_as_float(rp[e]) - delta) * scale;
    const float d1 = p1 * (__uint_as_float(rp[e + 1]) - delta) * scale;
    pk[e >> 1] = pack_bf16(p0, p1);
    dk[e >> 1] = pack_bf16(d0, d1);
  }
}

__device__ __forceinline__ void store_chunk(uint8_t* tile, int row, int c, const uint32_t (&pk)[16]) {
#pragma unroll
  for (int g = 0; g < 4; ++g)
    *reinterpret_cast<uint4*>(tile + tile_chunk_offset(row, c * 4 + g)) =
        make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
}

// TMEM accumulator row (DH fp32 columns at taddr) -> bf16 row in global
template <int DH>
__device__ __forceinline__ void store_acc_row(uint32_t taddr, bf16* dst, bool write) {
#pragma unroll
  for (int c = 0; c < DH / 16; ++c) {
    uint32_t r[16];
    tmem_ld_x16(taddr + c * 16, r);
    tmem_ld_wait();
    if (write) {
      uint4 v0, v1;
      v0.x = pack_bf16(__uint_as_float(r[0]), __uint_as_float(r[1]));
      v0.y = pack_bf16(__uint_as_float(r[2]), __uint_as_float(r[3]));
      v0.z = pack_bf16(__uint_as_float(r[4]), __uint_as_float(r[5]));
      v0.w = pack_bf16(__uint_as_float(r[6]), __uint_as_float(r[7]));
      v1.x = pack_bf16(__uint_as_float(r[8]), __uint_as_float(r[9]));
      v1.y = pack_bf16(__uint_as_float(r[10]), __uint_as_float(r[11]));
      v1.z = pack_bf16(__uint_as_float(r[12]), __uint_as_float(r[13]));
      v1.w = pack_bf16(__uint_as_float(r[14]), __uint_as_float(r[15]));
      *reinterpret_cast<uint4*>(dst + c * 16) = v0;
      *reinterpret_cast<uint4*>(dst + c * 16 + 8) = v1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// dq kernel
// ------------------------------------------------------------------------------------------------
template <int DH>
struct DqCfg {
  static constexpr int kStages = 2;
  static constexpr int kTileBytes = kTile * DH * 2;
  static constexpr int kOffDO = kTileBytes;
  static constexpr int kOffK = 2 * kTileBytes;
  static constexpr int kOffV = kOffK + kStages * kTileBytes;
  static constexpr int kOffDS = kOffV + kStages * kTileBytes;
  static constexpr int kOffBar = kOffDS + kTile * kTile * 2;
  static constexpr int kSmemBytes = kOffBar + 128;
  static constexpr int kTmemCols = 512;  // S [0,128), dP [128,256), dQ [256,256+DH)
};

template <int DH>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const AttnBwdParams p) {
  using Cfg = DqCfg<DH>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBar);
  uint64_t* qdo_full = bars;       // 1
  uint64_t* kv_full = bars + 1;    // 2
  uint64_t* kv_empty = bars + 3;   // 2
  uint64_t* sdp_full = bars + 5;   // MMA -> threads
  uint64_t* sdp_free = bars + 6;   // threads -> MMA (128)
  uint64_t* ds_full = bars + 7;    // threads -> MMA (128)
  uint64_t* dq_done = bars + 8;    // MMA -> threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = p.n_qtiles - 1 - static_cast<int>(blockIdx.x % p.n_qtiles);
  const int bh = static_cast<int>(blockIdx.x / p.n_qtiles);
  const int h = bh % p.H, b = bh / p.H;
  const int i0 = qt * kTile;
  const int i_last = min(i0 + kTile, p.Lq) - 1;
  int t0, t1;
  kv_tile_range_doc(p.m, b, i0, i_last, p.Lk, t0, t1);
  const int n_kv = t1 - t0;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) __trap();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    mbar_init(qdo_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(sdp_full, 1); mbar_init(sdp_free, 128); mbar_init(ds_full, 128); mbar_init(dq_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, Cfg::kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // setup above overlapped the previous kernel; global memory (TMA loads, lse/delta, stores) from here on
  pdl_launch();
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDQ = tmem_base + 256;

  if (warp == 0) {
    // (whole warp walks the loop, one elected lane issues each batch: no per-instruction ELECT / BRA.U.ANY loop)
    if (n_kv > 0) {
      if (elect_one()) {
        mbar_arrive_expect_tx(qdo_full, 2 * Cfg::kTileBytes);
#pragma unroll
        for (int pn = 0; pn < DH / 64; ++pn) {
          tma_load_4d(smem + pn * kPanelBytes, &tmQ, qdo_full, pn * 64, i0, h, b);
          tma_load_4d(smem + Cfg::kOffDO + pn * kPanelBytes, &tmDO, qdo_full, pn * 64, i0, h, b);
        }
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
      constexpr uint32_t idesc_dq = make_idesc_bf16(kTile, DH, false, true);
      const uint32_t sQ = smem_u32(smem), sDO = smem_u32(smem + Cfg::kOffDO), sDS = smem_u32(smem + Cfg::kOffDS);
      auto issue_sdp = [&](int it) {
        const int s = it & 1;
        mbar_wait(&kv_full[s], (it >> 1) & 1);
        if (it > 0) mbar_wait(sdp_free, (it - 1) & 1);
        tc_fence_after();
        const uint32_t sK = smem_u32(smem + Cfg::kOffK + s * Cfg::kTileBytes);
        const uint32_t sV = smem_u32(smem + Cfg::kOffV + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tS, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc_s, k > 0);
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tDP, desc_kmajor(sDO, k), desc_kmajor(sV, k), idesc_s, k > 0);
          umma_commit(sdp_full);
        }
        __syncwarp();
      };
      mbar_wait(qdo_full, 0);
      issue_sdp(0);
      for (int it = 0; it < n_kv; ++it) {
        const int s = it & 1;
        if (it + 1 < n_kv) issue_sdp(it + 1);
        mbar_wait(ds_full, it & 1);
        tc_fence_after();
        const uint32_t sK = smem_u32(smem + Cfg::kOffK + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kTile / 16; ++k)
            umma_bf16_ss(tDQ, desc_kmajor(sDS, k), desc_mnmajor(sK, k), idesc_dq, (it > 0 || k > 0) ? 1u : 0u);
          umma_commit(&kv_empty[s]);
          umma_commit(dq_done);
        }
        __syncwarp();
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int i = i0 + row;
    const int ic = min(i, p.Lq - 1);
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    int lo, hi;
    mask_row_range(p.m, ic, p.Lk, lo, hi);
    if (i >= p.Lq) hi = lo;  // padding rows contribute nothing
    const int* doc_row = p.m.doc != nullptr ? p.m.doc + b * p.m.doc_ld : nullptr;
    const int dq_doc = doc_row != nullptr ? __ldg(doc_row + (ic + p.m.q_offset) / p.m.T) : 0;
    const long long ridx = (static_cast<long long>(b) * p.H + h) * p.Lq + ic;
    const float lse = p.lse[ridx];
    const float lse2 = lse == -INFINITY ? INFINITY : lse * 1.4426950408889634f;
    const float delta = p.delta[ridx];
    uint8_t* sDS = smem + Cfg::kOffDS;
    for (int it = 0; it < n_kv; ++it) {
      const int j0 = (t0 + it) * kTile;
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
      const bool full = doc_row == nullptr && lo <= j0 && j0 + kTile <= hi;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t rs[32], rp[32], pk[16], dk[16];
        tmem_ld_x32(tS + lane_addr + c * 32, rs);
        tmem_ld_x32(tDP + lane_addr + c * 32, rp);
        const uint32_t valid = full ? 0xffffffffu : chunk_valid_bits(j0 + c * 32, lo, hi, doc_row, dq_doc, p.m.T);
        tmem_ld_wait();
        pds_chunk(rs, rp, valid, p.scale_log2, lse2, delta, p.scale, pk, dk);
        if (c == 0 && it > 0) mbar_wait(dq_done, (it - 1) & 1);  // dS buffer free
        store_chunk(sDS, row, c, dk);
      }
      tc_fence_before();
      mbar_arrive(sdp_free);
      fence_proxy_async_smem();
      mbar_arrive(ds_full);
    }
    if (n_kv > 0) {
      mbar_wait(dq_done, (n_kv - 1) & 1);
      tc_fence_after();
      store_acc_row<DH>(tDQ + lane_addr, p.dq + b * p.dq_sb + h * p.dq_sh + static_cast<long long>(i) * DH, i < p.Lq);
    } else if (i < p.Lq) {
      bf16* dst = p.dq + b * p.dq_sb + h * p.dq_sh + static_cast<long long>(i) * DH;
      for (int c = 0; c < DH / 8; ++c) *reinterpret_cast<uint4*>(dst + c * 8) = make_uint4(0, 0, 0, 0);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, Cfg::kTmemCols); }
}

// ------------------------------------------------------------------------------------------------
// dkv kernel
// ------------------------------------------------------------------------------------------------
template <int DH, bool SINGLE = false>
struct DkvCfg {
  static constexpr int kStages = DH == 64 ? 2 : 1;
  static constexpr int kTileBytes = kTile * DH * 2;
  static constexpr int kOffV = kTileBytes;
  static constexpr int kOffQ = 2 * kTileBytes;
  static constexpr int kOffDO = kOffQ + kStages * kTileBytes;
  static constexpr int kOffP = kOffDO + kStages * kTileBytes;
  static constexpr int kOffDS = kOffP + kTile * kTile * 2;
  static constexpr int kOffBar = kOffDS + kTile * kTile * 2;
  static constexpr int kSmemBytes = kOffBar + 128;
  static constexpr int kTmemCols = 512;  // S [0,128), dP [128,256), dV [256,256+DH), dK [256+DH, 256+2DH)
  static constexpr uint32_t kColDV = 256, kColDK = 256 + DH, kColDQ = 256 + 2 * DH;
};
// Single-tile problem (one query tile x one key tile, DH = 64): sized for TWO CTAs per SM, because 512 (b, h) CTAs of a
// few microseconds each are latency-bound, not throughput-bound.  V is dead once dP = dO V^T has completed, so the
// P tile starts in V's 16 KB; the dV / dK / dQ accumulators reuse the S / dP columns, which every thread has finished
// reading (P and dS are in shared memory) before the second group of MMAs is issued.
template <>
struct DkvCfg<64, true> {
  static constexpr int kStages = 1;
  static constexpr int kTileBytes = kTile * 64 * 2;      // 16 KB
  static constexpr int kOffV = kTileBytes;               // K at 0
  static constexpr int kOffP = kOffV;                    // 32 KB: V's panel + 16 KB
  static constexpr int kOffDS = kOffP + kTile * kTile * 2;
  static constexpr int kOffQ = kOffDS + kTile * kTile * 2;
  static constexpr int kOffDO = kOffQ + kTileBytes;
  static constexpr int kOffBar = kOffDO + kTileBytes;    // 112 KB
  static constexpr int kSmemBytes = kOffBar + 128;
  static constexpr int kTmemCols = 256;                  // S [0,128), dP [128,256); then dV [0,64), dK [64,128), dQ [128,192)
  static constexpr uint32_t kColDV = 0, kColDK = 64, kColDQ = 128;
};

// FUSE_DQ (DH = 64, the whole problem is ONE query tile x ONE key tile, e.g. audio.yml's L = 120): this CTA also
// holds the only dS tile any query row will ever see, so dQ = dS K is one more MMA into TMEM columns [384, 448)
// and the separate dq kernel (which recomputes S and dP) is not launched.
template <int DH, bool FUSE_DQ>
__global__ void __launch_bounds__(kBwdThreads, FUSE_DQ ? 2 : 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                    const AttnBwdParams p) {
  using Cfg = DkvCfg<DH, FUSE_DQ>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBar);
  uint64_t* kv_full = bars;         // 1
  uint64_t* qdo_full = bars + 1;    // 2
  uint64_t* qdo_empty = bars + 3;   // 2
  uint64_t* sdp_full = bars + 5;
  uint64_t* sdp_free = bars + 6;    // 128
  uint64_t* pds_full = bars + 7;    // 128
  uint64_t* acc_done = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kt = static_cast<int>(blockIdx.x % p.n_ktiles);  // early key tiles are seen by the most queries
  const int bh = static_cast<int>(blockIdx.x / p.n_ktiles);
  const int h = bh % p.H, b = bh / p.H;
  const int j0 = kt * kTile;
  const int j_last = min(j0 + kTile, p.Lk) - 1;
  int t0, t1;
  q_tile_range(p.m, j0, j_last, p.Lq, t0, t1, b);
  const int n_q = t1 - t0;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) __trap();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
    mbar_init(sdp_full, 1); mbar_init(sdp_free, 128); mbar_init(pds_full, 128); mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, Cfg::kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // setup above overlapped the previous kernel; global memory (TMA loads, lse/delta, stores) from here on
  pdl_launch();
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + Cfg::kColDV, tDK = tmem_base + Cfg::kColDK;
  const uint32_t tDQ = tmem_base + Cfg::kColDQ;  // FUSE_DQ only

  if (warp == 0) {
    // (whole warp walks the loop, one elected lane issues each batch: no per-instruction ELECT / BRA.U.ANY loop)
    if (n_q > 0) {
      if (elect_one()) {
        mbar_arrive_expect_tx(kv_full, 2 * Cfg::kTileBytes);
#pragma unroll
        for (int pn = 0; pn < DH / 64; ++pn) {
          tma_load_4d(smem + pn * kPanelBytes, &tmK, kv_full, pn * 64, j0, h, b);
          tma_load_4d(smem + Cfg::kOffV + pn * kPanelBytes, &tmV, kv_full, pn * 64, j0, h, b);
        }
      }
      __syncwarp();
      for (int it = 0; it < n_q; ++it) {
        const int s = it % kStages;
        mbar_wait(&qdo_empty[s], ((it / kStages) & 1) ^ 1);
        const int i0 = (t0 + it) * kTile;
        if (elect_one()) {
          mbar_arrive_expect_tx(&qdo_full[s], 2 * Cfg::kTileBytes);
#pragma unroll
          for (int pn = 0; pn < DH / 64; ++pn) {
            tma_load_4d(smem + Cfg::kOffQ + s * Cfg::kTileBytes + pn * kPanelBytes, &tmQ, &qdo_full[s], pn * 64, i0, h, b);
            tma_load_4d(smem + Cfg::kOffDO + s * Cfg::kTileBytes + pn * kPanelBytes, &tmDO, &qdo_full[s], pn * 64, i0, h,
                        b);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    if (n_q > 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_acc = make_idesc_bf16(kTile, DH, true, true);
      const uint32_t sK = smem_u32(smem), sV = smem_u32(smem + Cfg::kOffV);
      const uint32_t sP = smem_u32(smem + Cfg::kOffP), sDS = smem_u32(smem + Cfg::kOffDS);
      auto issue_sdp = [&](int it) {
        const int s = it % kStages;
        mbar_wait(&qdo_full[s], (it / kStages) & 1);
        if (it > 0) mbar_wait(sdp_free, (it - 1) & 1);
        tc_fence_after();
        const uint32_t sQ = smem_u32(smem + Cfg::kOffQ + s * Cfg::kTileBytes);
        const uint32_t sDO = smem_u32(smem + Cfg::kOffDO + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tS, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc_s, k > 0);
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tDP, desc_kmajor(sDO, k), desc_kmajor(sV, k), idesc_s, k > 0);
          umma_commit(sdp_full);
        }
        __syncwarp();
      };
      mbar_wait(kv_full, 0);
      issue_sdp(0);
      for (int it = 0; it < n_q; ++it) {
        const int s = it % kStages;
        if (kStages > 1 && it + 1 < n_q) issue_sdp(it + 1);
        mbar_wait(pds_full, it & 1);
        tc_fence_after();
        const uint32_t sQ = smem_u32(smem + Cfg::kOffQ + s * Cfg::kTileBytes);
        const uint32_t sDO = smem_u32(smem + Cfg::kOffDO + s * Cfg::kTileBytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kTile / 16; ++k)
            umma_bf16_ss(tDV, desc_mnmajor(sP, k), desc_mnmajor(sDO, k), idesc_acc, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
          for (int k = 0; k < kTile / 16; ++k)
            umma_bf16_ss(tDK, desc_mnmajor(sDS, k), desc_mnmajor(sQ, k), idesc_acc, (it > 0 || k > 0) ? 1u : 0u);
          if (FUSE_DQ) {  // dQ[i, :] = sum_j dS[i, j] K[j, :]: dS as the K-major A operand, K as the MN-major B operand
            constexpr uint32_t idesc_dq = make_idesc_bf16(kTile, DH, false, true);
#pragma unroll
            for (int k = 0; k < kTile / 16; ++k)
              umma_bf16_ss(tDQ, desc_kmajor(sDS, k), desc_mnmajor(sK, k), idesc_dq, k > 0 ? 1u : 0u);
          }
          umma_commit(&qdo_empty[s]);
          umma_commit(acc_done);
        }
        __syncwarp();
        if (kStages == 1 && it + 1 < n_q) issue_sdp(it + 1);
      }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const int* doc_row = p.m.doc != nullptr ? p.m.doc + b * p.m.doc_ld : nullptr;
    uint8_t* sP = smem + Cfg::kOffP;
    uint8_t* sDS = smem + Cfg::kOffDS;
    for (int it = 0; it < n_q; ++it) {
      const int i = (t0 + it) * kTile + row;
      const int ic = min(i, p.Lq - 1);
      int lo, hi;
      mask_row_range(p.m, ic, p.Lk, lo, hi);
      if (i >= p.Lq) hi = lo;
      const int dq_doc = doc_row != nullptr ? __ldg(doc_row + (ic + p.m.q_offset) / p.m.T) : 0;
      const long long ridx = (static_cast<long long>(b) * p.H + h) * p.Lq + ic;
      const float lse = __ldg(p.lse + ridx);
      const float lse2 = lse == -INFINITY ? INFINITY : lse * 1.4426950408889634f;
      const float delta = __ldg(p.delta + ridx);
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
      const bool full = doc_row == nullptr && lo <= j0 && j0 + kTile <= hi;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t rs[32], rp[32], pk[16], dk[16];
        tmem_ld_x32(tS + lane_addr + c * 32, rs);
        tmem_ld_x32(tDP + lane_addr + c * 32, rp);
        const uint32_t valid = full ? 0xffffffffu : chunk_valid_bits(j0 + c * 32, lo, hi, doc_row, dq_doc, p.m.T);
        tmem_ld_wait();
        pds_chunk(rs, rp, valid, p.scale_log2, lse2, delta, p.scale, pk, dk);
        if (c == 0 && it > 0) mbar_wait(acc_done, (it - 1) & 1);  // P / dS buffers free
        store_chunk(sP, row, c, pk);
        store_chunk(sDS, row, c, dk);
      }
      tc_fence_before();
      mbar_arrive(sdp_free);
      fence_proxy_async_smem();
      mbar_arrive(pds_full);
    }
    const int j = j0 + row;
    bf16* dkr = p.dk + b * p.dk_sb + h * p.dk_sh + static_cast<long long>(j) * DH;
    bf16* dvr = p.dv + b * p.dv_sb + h * p.dv_sh + static_cast<long long>(j) * DH;
    if (n_q > 0) {
      mbar_wait(acc_done, (n_q - 1) & 1);
      tc_fence_after();
      store_acc_row<DH>(tDV + lane_addr, dvr, j < p.Lk);
      store_acc_row<DH>(tDK + lane_addr, dkr, j < p.Lk);
      if (FUSE_DQ)  // query row `row` of the single query tile
        store_acc_row<DH>(tDQ + lane_addr, p.dq + b * p.dq_sb + h * p.dq_sh + static_cast<long long>(row) * DH,
                          row < p.Lq);
    } else {
      if (j < p.Lk) {
        for (int c = 0; c < DH / 8; ++c) {
          *reinterpret_cast<uint4*>(dkr + c * 8) = make_uint4(0, 0, 0, 0);
          *reinterpret_cast<uint4*>(dvr + c * 8) = make_uint4(0, 0, 0, 0);
        }
      }
      if (FUSE_DQ && row < p.Lq) {
        bf16* dqr = p.dq + b * p.dq_sb + h * p.dq_sh + static_cast<long long>(row) * DH;
        for (int c = 0; c < DH / 8; ++c) *reinterpret_cast<uint4*>(dqr + c * 8) = make_uint4(0, 0, 0, 0);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, Cfg::kTmemCols); }
}

template <int DH>
static int launch_attn_bwd(const CUtensorMap& tmQ, const CUtensorMap& tmDO, const CUtensorMap& tmK,
                           const CUtensorMap& tmV, const AttnBwdParams& p, cudaStream_t s) {
  constexpr bool kCanFuse = DH == 64;
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_dq_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         DqCfg<DH>::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    e = cudaFuncSetAttribute(attn_bwd_dkv_kernel<DH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             DkvCfg<DH>::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (kCanFuse) {
      e = cudaFuncSetAttribute(attn_bwd_dkv_kernel<DH, kCanFuse>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               DkvCfg<DH, kCanFuse>::kSmemBytes);
      if (e != cudaSuccess) return static_cast<int>(e);
    }
    return static_cast<int>(OWL_OK);
  });
  if (rc0 != OWL_OK) return rc0;
  const long long gq = static_cast<long long>(p.B) * p.H * p.n_qtiles;
  const long long gk = static_cast<long long>(p.B) * p.H * p.n_ktiles;
  if (kCanFuse && p.n_qtiles == 1 && p.n_ktiles == 1) {  // single tile: dq rides on the dkv kernel
    (void)launch_k(attn_bwd_dkv_kernel<DH, kCanFuse>, dim3(static_cast<unsigned>(gk)), dim3(kBwdThreads),
                   DkvCfg<DH, kCanFuse>::kSmemBytes, s, tmQ, tmDO, tmK, tmV, p);
    return launch_status();
  }
  (void)launch_k(attn_bwd_dq_kernel<DH>, dim3(static_cast<unsigned>(gq)), dim3(kBwdThreads), DqCfg<DH>::kSmemBytes, s, tmQ, tmDO, tmK, tmV, p);
  int rc = launch_status();
  if (rc != OWL_OK) return rc;
  (void)launch_k(attn_bwd_dkv_kernel<DH, false>, dim3(static_cast<unsigned>(gk)), dim3(kBwdThreads), DkvCfg<DH>::kSmemBytes, s, tmQ, tmDO, tmK, tmV, p);
  return launch_status();
}

}  // namespace owl

using namespace owl;

extern "C" {

int owl_attn_bwd(const void* dout, const void* o, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb,
                 int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb, int64_t v_sh,
                 const float* lse, float* delta, float* dq_acc, void* dq, int64_t dq_sb, int64_t dq_sh, void* dk,
                 int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh, int64_t B, int64_t H,
                 int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask, owl_stream_t stream) {
  if (!dout || !o || !q || !k || !v || !lse || !delta || !dq || !dk || !dv || B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0)
    return OWL_ERR_BAD_ARG;
  if (dh != 64 && dh != 128) return OWL_ERR_UNSUPPORTED;
  if (Lq > (1 << 30) || Lk > (1 << 30)) return OWL_ERR_UNSUPPORTED;
  if (!aligned16(dout) || !aligned16(o) || !aligned16(dq) || !aligned16(dk) || !aligned16(dv) ||
      (o_sb | o_sl | dq_sb | dq_sh | dk_sb | dk_sh | dv_sb | dv_sh) % 8)
    return OWL_ERR_ALIGN;
  AttnBwdParams p{};
  int rc = fill_mask_params(mask, &p.m);
  if (rc != OWL_OK) return rc;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.n_qtiles = static_cast<int>((Lq + kTile - 1) / kTile);
  p.n_ktiles = static_cast<int>((Lk + kTile - 1) / kTile);
  p.scale = 1.0f / sqrtf(static_cast<float>(dh));
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta;
  p.dq = static_cast<bf16*>(dq); p.dk = static_cast<bf16*>(dk); p.dv = static_cast<bf16*>(dv);
  p.dq_sb = dq_sb; p.dq_sh = dq_sh; p.dk_sb = dk_sb; p.dk_sh = dk_sh; p.dv_sb = dv_sb; p.dv_sh = dv_sh;
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  const long long total = B * Lq * H * (dh / 8);
  (void)launch_k(attn_delta_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, s,
      static_cast<const bf16*>(dout), static_cast<const bf16*>(o), o_sb, o_sl, delta, p.B, p.H, p.Lq,
      static_cast<int>(dh));
  if ((rc = launch_status()) != OWL_OK) return rc;

  {
    // more than one tile: the single-pass kernel (attn_bwd2.cu; S and dP once per tile pair, dq accumulated in fp32 in
    // dq_acc: TMA reduce-add at dh = 64, red.global.add at dh = 128).  OWL_ATTN_BWD=1 forces the two-pass kernels below
    // (A/B aid), OWL_ATTN_BWD=3 keeps dh = 128 on them.
    static const int force = [] { const char* e = getenv("OWL_ATTN_BWD"); return e ? atoi(e) : 0; }();
    if (dq_acc != nullptr && force != 1 && !(force == 3 && dh == 128) && (p.n_qtiles > 1 || p.n_ktiles > 1))
      return attn_bwd2_dispatch(dout, o_sb, o_sl, q, q_sb, q_sh, k, k_sb, k_sh, v, v_sb, v_sh, lse, delta, dq_acc, dq,
                                dq_sb, dq_sh, dk, dk_sb, dk_sh, dv, dv_sb, dv_sh, B, H, Lq, Lk, dh, mask, s);
  }
  CUtensorMap tmQ, tmDO, tmK, tmV;
  if ((rc = make_tmap_4d_bf16(&tmQ, q, dh, Lq, H, B, dh, q_sh, q_sb, 64, kTile)) != OWL_OK) return rc;
  // dO is token-major (b, i, h, :): row stride o_sl, head stride dh
  if ((rc = make_tmap_4d_bf16(&tmDO, dout, dh, Lq, H, B, o_sl, dh, o_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k, dh, Lk, H, B, dh, k_sh, k_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v, dh, Lk, H, B, dh, v_sh, v_sb, 64, kTile)) != OWL_OK) return rc;
  return dh == 64 ? launch_attn_bwd<64>(tmQ, tmDO, tmK, tmV, p, s) : launch_attn_bwd<128>(tmQ, tmDO, tmK, tmV, p, s);
}

}  // extern "C"
