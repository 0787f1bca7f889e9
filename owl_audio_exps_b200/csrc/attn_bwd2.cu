// Attention backward, second generation (more than one tile): ONE persistent kernel per call instead of a dq pass
// plus a dk/dv pass, so S and dP are computed once per (query tile, key tile) pair - 5 MMAs per pair instead of 7 - and
// neither P nor dS makes a round trip through shared memory on its way back to the tensor core.
//
//   A CTA owns one 128-key tile (K_j, V_j resident in shared memory, dK_j / dV_j accumulating in TMEM) and walks the query
//   tiles i that can see it (causal / window / document extents, heaviest key tiles first).  Everything is computed
//   TRANSPOSED - thread = key row - so that P^T and dS^T land in tensor memory in the layout the A operand of the dV / dK
//   MMAs wants:
//     MMA 1   S^T  = K_j Q_i^T           dP^T = V_j dO_i^T                     (SS, 128 x 128 x 64 each)
//     softmax P^T  = exp2(S^T c - lse_i) dS^T = P^T (dP^T - delta_i) / sqrt(dh)   -> bf16 over S^T / dP^T in TMEM,
//                                                                              dS^T also into shared memory
//     MMA 2   dV_j += P^T dO_i           dK_j += dS^T Q_i                      (A operand from TMEM)
//             dQ_i(partial) = dS K_j                                           (A = the shared-memory dS^T read MN-major)
//     drain   dQ partial: TMEM -> fp32 staging -> TMA reduce-add into the fp32 dq accumulator (owl_attn_bwd's dq_acc)
//
//   512 threads = 4 warpgroups, every warp touching TMEM owns lanes 32*(warp % 4):
//     WG0 / WG1  softmax over query columns [0,64) / [64,128) of the pair (thread = key row), epilogue dV / dK
//     WG2        warp 8 TMA producer, warp 9 MMA issuer, warps 10 / 11 load lse / delta of the query tile into shared memory
//     WG3        dQ drain
//   TMEM, dh = 64:  S^T [0,128)  dP^T [128,256)  dV [256,320)  dK [320,384)  dQ partial, double-buffered [384,448) [448,512).
//   TMEM, dh = 128: S^T [0,128)  dP^T [128,256)  dV [256,384)  dK [384,512) fill all 512 columns.  The dQ partial lives in the
//     four 32-column gaps the softmax leaves inside S^T / dP^T (P^T and dS^T are bf16: each warpgroup's 64 fp32 columns
//     shrink to 32), as four N = 32 MMAs; it is issued FIRST in MMA 2 and drained (red.global.add.v4.f32 straight from
//     registers: no shared memory is left for a staging tile) while dV / dK run, so the next pair's MMA 1 - which overwrites
//     those columns - finds them free.  Shared memory: K, V, 2 stages of {Q, dO}, dS^T = 226 KB.
//
// Replaces the autograd of flex_attention (owl_wms/nn/attn.py:106-109 under rft_trainer.py:189) for training / prefill
// shapes at dh = 64 and dh = 128; single-tile problems stay on attn_bwd.cu.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "attn_common.cuh"
#include "host_common.h"

namespace owl {

int fill_mask_params(const owl_attn_mask* mask, MaskParams* m);  // attn.cu

struct AttnBwd2Params {
  int B, H, Lq, Lk, n_qtiles, n_ktiles, total_items, Lq_pad;
  float scale_log2, scale;
  const float* lse;
  const float* delta;
  float* dq_acc;   // fp32 [B*H, Lq_pad, dh]
  bf16 *dk, *dv;
  long long dk_sb, dk_sh, dv_sb, dv_sh;
  MaskParams m;
};

template <int DH_>
struct Bwd2Cfg {
  static constexpr int DH = DH_;
  static constexpr int kThreads = 512;
  static constexpr int kStages = 2;                        // {Q_i, dO_i, lse_i, delta_i} in flight
  static constexpr int kDSBufs = DH == 64 ? 2 : 1;         // dh 64: dS^T of pair n + 1 is written while dQ of pair n reads its own
  static constexpr int kTileBytes = kTile * DH * 2;        // 16 / 32 KB
  static constexpr int kOffK = 0;
  static constexpr int kOffV = kTileBytes;
  static constexpr int kOffRing = 2 * kTileBytes;          // per stage: Q then dO
  static constexpr int kStageBytes = 2 * kTileBytes;
  static constexpr int kOffDS = kOffRing + kStages * kStageBytes;   // dS^T [128 keys][128 q] bf16, two swizzled panels
  static constexpr int kOffDQ = kOffDS + kDSBufs * kTile * kTile * 2;         // dh = 64: fp32 staging [128 q][64], two 128B-swizzled panels
  static constexpr int kOffStat = kOffDQ + (DH == 64 ? kTile * DH * 4 : 0);   // per stage: lse2[128], delta[128] fp32
  static constexpr int kOffBar = kOffStat + kStages * 1024;
  static constexpr int kSmemBytes = kOffBar + 512;
  static constexpr uint32_t kColST = 0, kColDPT = 128, kColDV = 256, kColDK = 256 + DH, kColDQ = 384;
};

struct Bwd2Item {
  int b, h, j0, t0, n_q;
};

__device__ __forceinline__ Bwd2Item bwd2_item(const AttnBwd2Params& p, int w) {
  Bwd2Item it;
  const int bh = w / p.n_ktiles;
  const int kt = w % p.n_ktiles;  // early key tiles are seen by the most queries under the causal mask: heaviest first
  it.b = bh / p.H;
  it.h = bh % p.H;
  it.j0 = kt * kTile;
  int t1;
  q_tile_range(p.m, it.j0, min(it.j0 + kTile, p.Lk) - 1, p.Lq, it.t0, t1, it.b);
  it.n_q = t1 - it.t0;
  return it;
}

// 32 query columns of one key row: P^T = exp2(S^T c - lse), dS^T = P^T (dP^T / sqrt(dh) - delta / sqrt(dh)), both packed to bf16.
// FULL = every (key, query) of the chunk is visible: no per-element mask test (the common case away from the band edges:
// 5 instructions per element instead of 8).
template <bool FULL>
__device__ __forceinline__ void bwd2_softmax_chunk(const uint32_t (&rs)[32], const uint32_t (&rp)[32], uint32_t (&pk)[16],
                                                   uint32_t (&dk)[16], const float4* st_lse, const float4* st_del,
                                                   float scale_log2, float scale, uint32_t valid) {
#pragma unroll
  for (int e4 = 0; e4 < 8; ++e4) {
    const float4 l4 = st_lse[e4], d4 = st_del[e4];
    const float ls[4] = {l4.x, l4.y, l4.z, l4.w}, dl[4] = {d4.x, d4.y, d4.z, d4.w};
    float pv[4], dv[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int idx = e4 * 4 + e;
      float pe = exp2f(fmaf(__uint_as_float(rs[idx]), scale_log2, -ls[e]));
      if (!FULL && !((valid >> idx) & 1u)) pe = 0.f;
      pv[e] = pe;
      dv[e] = pe * fmaf(__uint_as_float(rp[idx]), scale, -dl[e]);
    }
    pk[e4 * 2] = pack_bf16(pv[0], pv[1]);
    pk[e4 * 2 + 1] = pack_bf16(pv[2], pv[3]);
    dk[e4 * 2] = pack_bf16(dv[0], dv[1]);
    dk[e4 * 2 + 1] = pack_bf16(dv[2], dv[3]);
  }
}

template <int DH>
__global__ void __launch_bounds__(512, 1)
attn_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                 const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                 const __grid_constant__ CUtensorMap tmDQ, const AttnBwd2Params p) {
  using Cfg = Bwd2Cfg<DH>;
  constexpr int kStages = Cfg::kStages, kTileBytes = Cfg::kTileBytes, kStageBytes = Cfg::kStageBytes;
  constexpr int kOffK = Cfg::kOffK, kOffV = Cfg::kOffV, kOffRing = Cfg::kOffRing, kOffDS = Cfg::kOffDS;
  constexpr int kOffDQ = Cfg::kOffDQ, kOffStat = Cfg::kOffStat, kOffBar = Cfg::kOffBar;
  constexpr uint32_t kColST = Cfg::kColST, kColDPT = Cfg::kColDPT, kColDV = Cfg::kColDV, kColDK = Cfg::kColDK;
  constexpr uint32_t kColDQ = Cfg::kColDQ;
  (void)kOffDQ; (void)kColDQ;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* kv_full = bars;                 // TMA -> MMA
  uint64_t* kv_empty = bars + 1;            // MMA -> TMA (every MMA of the item has read K_j / V_j)
  uint64_t* qdo_full = bars + 2;            // [kStages] TMA -> MMA
  uint64_t* qdo_empty = bars + 2 + kStages; // [kStages] MMA -> TMA and the lse / delta loaders
  uint64_t* stat_full = bars + 2 + 2 * kStages;  // [kStages] lse / delta loaders -> softmax (2 arrivals)
  uint64_t* sdp_full = bars + 2 + 3 * kStages;   // MMA -> softmax
  uint64_t* pds_full = sdp_full + 1;        // softmax -> MMA (256 arrivals)
  uint64_t* dq_full = sdp_full + 2;         // [2] MMA -> drain
  uint64_t* dq_free = sdp_full + 4;         // [2] drain -> MMA (128 arrivals)
  uint64_t* acc_done = sdp_full + 6;        // MMA -> softmax epilogue (dV / dK complete)
  uint64_t* acc_free = sdp_full + 7;        // softmax epilogue -> MMA (256 arrivals)
  uint64_t* pds_half = sdp_full + 8;        // softmax -> MMA: first 32 query columns of both halves are in TMEM (256 arrivals)
  uint64_t* s_half = sdp_full + 9;          // [2] dh 64, MMA -> softmax warpgroup h: S^T / dP^T of ITS 64 query columns
  uint64_t* p_half = sdp_full + 11;         // [2] dh 64, softmax warpgroup h -> MMA (128 arrivals)
  uint64_t* ds_free = sdp_full + 13;        // dh 128, drain -> softmax: the dQ staging (which borrows the dS^T buffer) has been read
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sdp_full + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) __trap();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDQ);
    mbar_init(kv_full, 1);
    mbar_init(kv_empty, 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&qdo_full[s], 1);
      mbar_init(&qdo_empty[s], 1);
      mbar_init(&stat_full[s], 2);
    }
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, 256);
    mbar_init(pds_half, 256);
    mbar_init(ds_free, 1);
    for (int h = 0; h < 2; ++h) {
      mbar_init(&s_half[h], 1);
      mbar_init(&p_half[h], 128);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&dq_full[s], 1);
      mbar_init(&dq_free[s], 128);
    }
    mbar_init(acc_done, 1);
    mbar_init(acc_free, 256);
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_launch();

  // Registers follow the work (setmaxnreg at the head of each warpgroup's role; 512 x 128 = 65536 in total): the TMA / MMA / stat-loader
  // warpgroup needs few.  dh 64: the softmax warpgroups get 176 (the second chunk's TMEM loads are in flight while the
  // first is computed).  dh 128: the drain gets 192 and reads the whole 128-column dQ partial at once, so the S^T / dP^T
  // columns it sits in go back to the next pair's MMA 1 before the ~32 reduce-adds per thread are issued.
  if (warp >= 8 && warp < 12) {
  reg_dealloc<64>();
  if (warp == 8) {
    // ---------------------------------------------------------------- TMA producer (whole warp; one elected lane issues)
    {
      uint32_t item_n = 0, ring_n = 0;
      for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
        const Bwd2Item it = bwd2_item(p, w);
        if (it.n_q <= 0) continue;
        mbar_wait(kv_empty, (item_n & 1) ^ 1);
        ++item_n;
        if (elect_one()) {
          mbar_arrive_expect_tx(kv_full, 2 * kTileBytes);
#pragma unroll
          for (int pn = 0; pn < DH / 64; ++pn) {
            tma_load_4d(smem + kOffK + pn * kPanelBytes, &tmK, kv_full, pn * 64, it.j0, it.h, it.b);
            tma_load_4d(smem + kOffV + pn * kPanelBytes, &tmV, kv_full, pn * 64, it.j0, it.h, it.b);
          }
        }
        __syncwarp();
        for (int q = 0; q < it.n_q; ++q) {
          const uint32_t s = ring_n % kStages;
          mbar_wait(&qdo_empty[s], ((ring_n / kStages) & 1) ^ 1);
          ++ring_n;
          const int i0 = (it.t0 + q) * kTile;
          if (elect_one()) {
            mbar_arrive_expect_tx(&qdo_full[s], kStageBytes);
#pragma unroll
            for (int pn = 0; pn < DH / 64; ++pn) {
              tma_load_4d(smem + kOffRing + s * kStageBytes + pn * kPanelBytes, &tmQ, &qdo_full[s], pn * 64, i0, it.h, it.b);
              tma_load_4d(smem + kOffRing + s * kStageBytes + kTileBytes + pn * kPanelBytes, &tmDO, &qdo_full[s], pn * 64,
                          i0, it.h, it.b);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 10 || warp == 11) {
    // ---------------------------------------------------------------- lse (warp 10) / delta (warp 11) of the query tile
    const bool is_lse = warp == 10;
    const float* src = is_lse ? p.lse : p.delta;
    uint32_t ring_n = 0;
    for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
      const Bwd2Item it = bwd2_item(p, w);
      if (it.n_q <= 0) continue;
      const float* row = src + (static_cast<long long>(it.b) * p.H + it.h) * p.Lq;
      for (int q = 0; q < it.n_q; ++q) {
        const uint32_t s = ring_n % kStages;
        mbar_wait(&qdo_empty[s], ((ring_n / kStages) & 1) ^ 1);
        ++ring_n;
        const int i0 = (it.t0 + q) * kTile;
        float* dst = reinterpret_cast<float*>(smem + kOffStat + s * 1024) + (is_lse ? 0 : 128);
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = i0 + lane * 4 + e;
          float x = i < p.Lq ? __ldg(row + i) : (is_lse ? -INFINITY : 0.f);
          // rows with nothing visible (lse = -inf) and padding rows: exponent offset +inf -> P = 0
          if (is_lse) x = x == -INFINITY ? INFINITY : x * 1.4426950408889634f;
          else x *= p.scale;               // delta / sqrt(dh): dS = P (dP / sqrt(dh) - delta / sqrt(dh)) is one FFMA + one FMUL
          v[e] = x;
        }
        *reinterpret_cast<float4*>(dst + lane * 4) = make_float4(v[0], v[1], v[2], v[3]);
        __syncwarp();
        if (lane == 0) mbar_arrive(&stat_full[s]);
      }
    }
  } else {
    // ---------------------------------------------------------------- MMA issuer (warp 9)
    // The WHOLE warp walks the schedule (waits included) and one elected lane issues each BATCH of tcgen05 instructions.  Under a
    // divergent `if (lane == 0)` ptxas wraps every UTCHMMA in an ELECT / BRA.U.ANY loop and moves its operands to uniform
    // registers one at a time: ~100 clk per MMA, and with 40 MMAs of 32-64 clk per pair this warp, not the tensor pipe or the
    // softmax, set the pace of the kernel (ncu source view: the issuer warp never waited).
    {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);   // S^T, dP^T
      constexpr uint32_t idesc_acc = make_idesc_bf16(kTile, DH, false, true);     // dV, dK: A from TMEM, B MN-major
      constexpr uint32_t idesc_dq = make_idesc_bf16(kTile, DH, true, true);       // dQ: A = dS^T tile read MN-major
      constexpr uint32_t idesc_dq32 = make_idesc_bf16(kTile, 32, true, true);     // dh = 128: one 32-column piece of dQ
      (void)idesc_dq; (void)idesc_dq32;
      const uint32_t sK = smem_u32(smem + kOffK), sV = smem_u32(smem + kOffV), sDS = smem_u32(smem + kOffDS);
      const uint32_t tST = tmem_base + kColST, tDPT = tmem_base + kColDPT;
      const uint32_t tDV = tmem_base + kColDV, tDK = tmem_base + kColDK;
      uint32_t item_n = 0, ring_n = 0, pair_n = 0;
      for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
        const Bwd2Item it = bwd2_item(p, w);
        if (it.n_q <= 0) continue;
        mbar_wait(kv_full, item_n & 1);
        if (DH == 64) {
          // dh 64: the two softmax warpgroups run HALF A PERIOD APART.  Warpgroup h owns query columns [64h, 64h + 64) of every
          // pair: its S^T / dP^T columns, its k-steps of dV / dK and its half of dS^T touch nothing of the other half, so
          //   MMA 2 (half h, pair q) -> MMA 1 (half h, pair q + 1)
          // is issued the moment warpgroup h hands over P^T / dS^T, while the other warpgroup is still in its softmax:
          // each warpgroup waits for 2 x 256 clk of MMA per pair instead of the whole pair's 1280.  dQ (all 128 queries,
          // M = 128) follows half B and reads the dS^T buffer of the pair's parity.
          constexpr uint32_t idesc_h = make_idesc_bf16(kTile, 64, false, false);   // S^T / dP^T of one half: N = 64 queries
          // (plain issue, to be called inside ONE elect_one() region per batch: one ELECT and one set of uniform operands
          // per batch instead of per instruction - the issuer warp's instruction count is what bounds this kernel)
          auto mma1_half = [&](uint32_t st, uint32_t h) {
            const uint32_t sQh = smem_u32(smem + kOffRing + st * kStageBytes) + h * (64 * 128), sDOh = sQh + kTileBytes;
#pragma unroll
            for (int k = 0; k < DH / 16; ++k)
              umma_bf16_ss(tST + h * 64, desc_kmajor(sK, k), desc_kmajor(sQh, k), idesc_h, k > 0);
#pragma unroll
            for (int k = 0; k < DH / 16; ++k)
              umma_bf16_ss(tDPT + h * 64, desc_kmajor(sV, k), desc_kmajor(sDOh, k), idesc_h, k > 0);
            umma_commit(&s_half[h]);
          };
          uint32_t s = ring_n % kStages;
          mbar_wait(&qdo_full[s], (ring_n / kStages) & 1);
          ++ring_n;
          tc_fence_after();
          if (elect_one()) {
            mma1_half(s, 0);
            mma1_half(s, 1);
          }
          __syncwarp();
          for (int q = 0; q < it.n_q; ++q) {
            const uint32_t sQ = smem_u32(smem + kOffRing + s * kStageBytes), sDO = sQ + kTileBytes;
            const bool more = q + 1 < it.n_q;
            uint32_t s_next = s;
            if (more) {   // {Q, dO} of the next pair (prefetched by the TMA warp while this pair was computed)
              s_next = ring_n % kStages;
              mbar_wait(&qdo_full[s_next], (ring_n / kStages) & 1);
              ++ring_n;
            }
            if (q == 0) mbar_wait(acc_free, (item_n & 1) ^ 1);   // the previous item's epilogue has read dV / dK
#pragma unroll
            for (uint32_t h = 0; h < 2; ++h) {
              mbar_wait(&p_half[h], pair_n & 1);
              const uint32_t acc = (q > 0 || h > 0) ? 1u : 0u;
              tc_fence_after();
              if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {  // P^T: query columns 16k.. sit at 32-bit columns (k/4)*64 + (k%4)*8
                  const uint32_t k = h * 4 + kk;
                  umma_bf16_ts(tDV, tST + h * 64 + kk * 8, desc_mnmajor(sDO, k), idesc_acc, (acc | (kk > 0)) ? 1u : 0u);
                }
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                  const uint32_t k = h * 4 + kk;
                  umma_bf16_ts(tDK, tDPT + h * 64 + kk * 8, desc_mnmajor(sQ, k), idesc_acc, (acc | (kk > 0)) ? 1u : 0u);
                }
                if (h == 1) umma_commit(&qdo_empty[s]);   // Q_i / dO_i / lse / delta of this pair are no longer read
                if (more) mma1_half(s_next, h);
              }
              __syncwarp();
            }
            const uint32_t buf = pair_n & 1;
            mbar_wait(&dq_free[buf], ((pair_n >> 1) & 1) ^ 1);     // the drain has read this dQ buffer's previous content
            tc_fence_after();
            const uint32_t tDQ = tmem_base + kColDQ + buf * DH;
            const uint32_t sDSb = sDS + buf * (kTile * kTile * 2);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < kTile / 16; ++k)
                umma_bf16_ss(tDQ, desc_mnmajor(sDSb, k), desc_mnmajor(sK, k), idesc_dq, k > 0 ? 1u : 0u);
              umma_commit(&dq_full[buf]);
            }
            __syncwarp();
            s = s_next;
            ++pair_n;
          }
        } else
        for (int q = 0; q < it.n_q; ++q) {   // dh 128: one S^T / dP^T buffer, three batches per pair
          const uint32_t s = ring_n % kStages;
          mbar_wait(&qdo_full[s], (ring_n / kStages) & 1);
          ++ring_n;
          const uint32_t sQ = smem_u32(smem + kOffRing + s * kStageBytes), sDO = sQ + kTileBytes;
          if (pair_n > 0) {
            // the previous pair's dQ partial sits inside the S^T / dP^T columns MMA 1 is about to overwrite
            mbar_wait(&dq_free[0], (pair_n - 1) & 1);
          }
          tc_fence_after();
          // MMA 1 (in issue order behind MMA 2 of the previous pair, which still reads P^T / dS^T from these columns)
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tST, desc_kmajor(sK, k), desc_kmajor(sQ, k), idesc_s, k > 0);
#pragma unroll
            for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tDPT, desc_kmajor(sV, k), desc_kmajor(sDO, k), idesc_s, k > 0);
            umma_commit(sdp_full);
          }
          __syncwarp();
          // MMA 2, first part: the k-steps of dV / dK over the query columns the softmax finishes first (chunk 0 of both
          // warpgroups = columns [0,32) and [64,96) = k 0,1,4,5) run while the softmax is still on the other 64 columns
          mbar_wait(pds_half, pair_n & 1);
          if (q == 0) mbar_wait(acc_free, (item_n & 1) ^ 1);       // the previous item's epilogue has read dV / dK
          const uint32_t acc = q > 0 ? 1u : 0u;
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {  // P^T: query columns 16k.. sit at 32-bit columns (k/4)*64 + (k%4)*8
              const int k = (kk >> 1) * 4 + (kk & 1);
              const uint32_t a = static_cast<uint32_t>((k >> 2) * 64 + (k & 3) * 8);
              umma_bf16_ts(tDV, tST + a, desc_mnmajor(sDO, k), idesc_acc, (acc | (kk > 0)) ? 1u : 0u);
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const int k = (kk >> 1) * 4 + (kk & 1);
              const uint32_t a = static_cast<uint32_t>((k >> 2) * 64 + (k & 3) * 8);
              umma_bf16_ts(tDK, tDPT + a, desc_mnmajor(sQ, k), idesc_acc, (acc | (kk > 0)) ? 1u : 0u);
            }
          }
          __syncwarp();
          // second part: everything is in TMEM / shared memory
          mbar_wait(pds_full, pair_n & 1);
          tc_fence_after();
          if (elect_one()) {
            // dQ partial first, as four N = 32 pieces into the column gaps [32,64) [96,128) [160,192) [224,256): piece c
            // holds head-dim columns [32c, 32c + 32); B = K_j read MN-major at a 64-byte offset inside its 64-column panel
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const uint32_t tDQc = tmem_base + 32u + 64u * static_cast<uint32_t>(c);
              const uint32_t sKc = sK + (c >> 1) * kPanelBytes + (c & 1) * 64;
#pragma unroll
              for (int k = 0; k < kTile / 16; ++k)
                umma_bf16_ss(tDQc, desc_mnmajor(sDS, k), make_smem_desc(sKc + k * 2048, kPanelBytes, 1024), idesc_dq32,
                             k > 0 ? 1u : 0u);
            }
            umma_commit(&dq_full[0]);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const int k = (kk >> 1) * 4 + (kk & 1) + 2;
              const uint32_t a = static_cast<uint32_t>((k >> 2) * 64 + (k & 3) * 8);
              umma_bf16_ts(tDV, tST + a, desc_mnmajor(sDO, k), idesc_acc, 1u);
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const int k = (kk >> 1) * 4 + (kk & 1) + 2;
              const uint32_t a = static_cast<uint32_t>((k >> 2) * 64 + (k & 3) * 8);
              umma_bf16_ts(tDK, tDPT + a, desc_mnmajor(sQ, k), idesc_acc, 1u);
            }
            umma_commit(&qdo_empty[s]);
          }
          __syncwarp();
          ++pair_n;
        }
        if (elect_one()) {
          umma_commit(acc_done);
          umma_commit(kv_empty);
        }
        __syncwarp();
        ++item_n;
      }
    }
  }
  } else if (warp >= 12) {
    // ---------------------------------------------------------------- WG3: dQ drain (thread = query row)
    if (DH == 64) reg_dealloc<96>();
    else reg_alloc<192>();
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const bool leader = threadIdx.x == 12 * 32;
    uint8_t* stag = smem + kOffDQ;
    uint32_t pair_n = 0;
    for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
      const Bwd2Item it = bwd2_item(p, w);
      if (it.n_q <= 0) continue;
      const long long row_base = (static_cast<long long>(it.b) * p.H + it.h) * p.Lq_pad;
      for (int q = 0; q < it.n_q; ++q) {
        if (DH == 64) {
          const uint32_t buf = pair_n & 1;
          mbar_wait(&dq_full[buf], (pair_n >> 1) & 1);
          ++pair_n;
          tc_fence_after();
          uint32_t v[2][32];
          tmem_ld_x32(tmem_base + kColDQ + buf * DH + lane_addr, v[0]);
          tmem_ld_x32(tmem_base + kColDQ + buf * DH + lane_addr + 32, v[1]);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&dq_free[buf]);
          if (leader) tma_store_wait_read0();   // the previous reduce has finished reading the staging panels
          named_bar_sync(1, 128);
#pragma unroll
          for (int pn = 0; pn < 2; ++pn)
#pragma unroll
            for (int c = 0; c < 8; ++c)
              *reinterpret_cast<uint4*>(stag + pn * 16384 + r * 128 + ((c ^ (r & 7)) << 4)) =
                  make_uint4(v[pn][4 * c], v[pn][4 * c + 1], v[pn][4 * c + 2], v[pn][4 * c + 3]);
          fence_proxy_async_smem();
          named_bar_sync(1, 128);
          if (leader) {
            const int row0 = static_cast<int>(row_base + static_cast<long long>(it.t0 + q) * kTile);
            tma_reduce_add_2d(&tmDQ, stag, 0, row0);
            tma_reduce_add_2d(&tmDQ, stag + 16384, 32, row0);
            tma_store_commit();
          }
        } else {
          // dh = 128: the partial sits in the four 32-column gaps of S^T / dP^T; read it out at once (the next pair's
          // MMA 1 waits for these columns), then add it to the fp32 accumulator straight from registers
          mbar_wait(&dq_full[0], pair_n & 1);
          ++pair_n;
          tc_fence_after();
          uint32_t v[4][32];
#pragma unroll
          for (int c = 0; c < 4; ++c) tmem_ld_x32(tmem_base + 32u + 64u * static_cast<uint32_t>(c) + lane_addr, v[c]);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&dq_free[0]);         // the columns go back to the next pair's MMA 1
          // No shared memory is left for a 64 KB staging tile, and 32 red.global per thread (16 bytes per lane, 32 rows per
          // warp instruction) made the dQ accumulation the bound of this kernel (5B shape: 208 ms with, 132 ms without it).
          // The dS^T buffer is idle from here (dq_full: the dQ MMAs have read it) until the NEXT pair's softmax writes it -
          // after MMA 1 (1024 clk) and a chunk of exponentials: the partial goes through it in two 32 KB rounds of
          // full-row TMA reduce-adds, and the softmax waits for ds_free before its first dS^T store.
          uint8_t* stag2 = smem + kOffDS;
          const int row0 = static_cast<int>(row_base + static_cast<long long>(it.t0 + q) * kTile);
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            if (half == 1) {
              if (leader) tma_store_wait_read0();   // round 0 has been read
              named_bar_sync(1, 128);
            }
#pragma unroll
            for (int pn = 0; pn < 2; ++pn)
#pragma unroll
              for (int c = 0; c < 8; ++c)
                *reinterpret_cast<uint4*>(stag2 + pn * 16384 + r * 128 + ((c ^ (r & 7)) << 4)) =
                    make_uint4(v[half * 2 + pn][4 * c], v[half * 2 + pn][4 * c + 1], v[half * 2 + pn][4 * c + 2],
                               v[half * 2 + pn][4 * c + 3]);
            fence_proxy_async_smem();
            named_bar_sync(1, 128);
            if (leader) {
              tma_reduce_add_2d(&tmDQ, stag2, half * 64, row0);
              tma_reduce_add_2d(&tmDQ, stag2 + 16384, half * 64 + 32, row0);
              tma_store_commit();
            }
          }
          if (leader) {
            tma_store_wait_read0();
            mbar_arrive(ds_free);
          }
        }
      }
    }
    if (leader) tma_store_wait_all0();
  } else {
    // ---------------------------------------------------------------- WG0 / WG1: softmax, thread = key row
    if (DH == 64) reg_alloc<176>();
    const int hf = warp >> 2;                 // query columns [64 hf, 64 hf + 64) of every pair
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tST = tmem_base + kColST + lane_addr + hf * 64;
    const uint32_t tDPT = tmem_base + kColDPT + lane_addr + hf * 64;
    uint8_t* sDS = smem + kOffDS;
    uint32_t item_n = 0, ring_n = 0, pair_n = 0;
    for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
      const Bwd2Item it = bwd2_item(p, w);
      const int j = it.j0 + r;
      bf16* dst = (hf == 0 ? p.dv + it.b * p.dv_sb + it.h * p.dv_sh : p.dk + it.b * p.dk_sb + it.h * p.dk_sh) +
                  static_cast<long long>(j) * DH;
      if (it.n_q <= 0) {  // no query sees these keys: zero gradients
        if (j < p.Lk) {
#pragma unroll
          for (int c = 0; c < DH / 8; ++c) *reinterpret_cast<uint4*>(dst + c * 8) = make_uint4(0, 0, 0, 0);
        }
        continue;
      }
      int ilo, ihi;
      bool doc_test;
      key_q_range(p.m, it.b, min(j, p.Lk - 1), p.Lq, p.Lk, ilo, ihi, doc_test);
      if (j >= p.Lk) ihi = ilo;
      const int* doc_row = doc_test ? p.m.doc + it.b * p.m.doc_ld : nullptr;
      const int dkey = doc_row != nullptr ? __ldg(doc_row + min(j, p.Lk - 1) / p.m.T) : 0;
      for (int q = 0; q < it.n_q; ++q) {
        const uint32_t s = ring_n % kStages;
        const int i0 = (it.t0 + q) * kTile + hf * 64;   // first query column of this warpgroup's half
        mbar_wait(&stat_full[s], (ring_n / kStages) & 1);
        ++ring_n;
        const float4* st_lse = reinterpret_cast<const float4*>(smem + kOffStat + s * 1024) + hf * 16;
        const float4* st_del = st_lse + 32;
        mbar_wait(DH == 64 ? &s_half[hf] : sdp_full, pair_n & 1);
        uint8_t* sDSp = sDS + (DH == 64 ? (pair_n & 1) * (kTile * kTile * 2) : 0);   // dh 64: dS^T buffer of the pair's parity
        ++pair_n;
        tc_fence_after();
        const bool full = doc_row == nullptr && ilo <= i0 && i0 + 64 <= ihi;
        uint32_t rsx[2][32], rpx[2][32];
        if (DH == 64) {   // both chunks requested at once (176 registers): one TMEM round trip per pair instead of two
          tmem_ld_x32(tST, rsx[0]);
          tmem_ld_x32(tDPT, rpx[0]);
          tmem_ld_x32(tST + 32, rsx[1]);
          tmem_ld_x32(tDPT + 32, rpx[1]);
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t(&rs)[32] = rsx[DH == 64 ? c : 0];
          uint32_t(&rp)[32] = rpx[DH == 64 ? c : 0];
          uint32_t pk[16], dk[16];
          if (DH == 128) {
            tmem_ld_x32(tST + c * 32, rs);
            tmem_ld_x32(tDPT + c * 32, rp);
          }
          uint32_t valid = 0xffffffffu;
          if (!full) {
            const int qc = i0 + c * 32;
            int a = ilo - qc, z = ihi - qc;
            a = a < 0 ? 0 : a;
            z = z > 32 ? 32 : z;
            valid = z <= a ? 0u : ((z >= 32 ? 0xffffffffu : ((1u << z) - 1u)) & ~((1u << a) - 1u));
            if (doc_row != nullptr) {
              for (int e = a; e < z; ++e)
                if (__ldg(doc_row + (qc + e + p.m.q_offset) / p.m.T) != dkey) valid &= ~(1u << e);
            }
          }
          if (DH == 128 || c == 0) tmem_ld_wait();
          if (full) bwd2_softmax_chunk<true>(rs, rp, pk, dk, st_lse + c * 8, st_del + c * 8, p.scale_log2, p.scale, valid);
          else bwd2_softmax_chunk<false>(rs, rp, pk, dk, st_lse + c * 8, st_del + c * 8, p.scale_log2, p.scale, valid);
          // P^T / dS^T over the columns this thread has already read (first half of its own 64-column region)
          tmem_st_x16(tST + c * 16, pk);
          tmem_st_x16(tDPT + c * 16, dk);
          if (DH == 128 && c == 0 && pair_n > 1) mbar_wait(ds_free, pair_n & 1);   // the previous pair's dQ staging is out
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *reinterpret_cast<uint4*>(sDSp + tile_chunk_offset(r, hf * 8 + c * 4 + g)) =
                make_uint4(dk[4 * g], dk[4 * g + 1], dk[4 * g + 2], dk[4 * g + 3]);
          if (DH == 128 && c == 0) {    // dV / dK over these 32 query columns may start (they read P^T / dS^T from TMEM only)
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(pds_half);
          }
        }
        tmem_st_wait();
        tc_fence_before();
        fence_proxy_async_smem();
        mbar_arrive(DH == 64 ? &p_half[hf] : pds_full);
      }
      // epilogue: WG0 stores dV_j, WG1 stores dK_j
      mbar_wait(acc_done, item_n & 1);
      ++item_n;
      tc_fence_after();
      const uint32_t tAcc = tmem_base + (hf == 0 ? kColDV : kColDK) + lane_addr;
#pragma unroll
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t o[32];
        tmem_ld_x32(tAcc + c * 32, o);
        tmem_ld_wait();
        if (j < p.Lk) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[8 * g + 0]), __uint_as_float(o[8 * g + 1]));
            v.y = pack_bf16(__uint_as_float(o[8 * g + 2]), __uint_as_float(o[8 * g + 3]));
            v.z = pack_bf16(__uint_as_float(o[8 * g + 4]), __uint_as_float(o[8 * g + 5]));
            v.w = pack_bf16(__uint_as_float(o[8 * g + 6]), __uint_as_float(o[8 * g + 7]));
            *reinterpret_cast<uint4*>(dst + c * 32 + g * 8) = v;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(acc_free);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// fp32 dq accumulator [B*H, Lq_pad, dh] -> bf16 dq [B, H, Lq, dh] (strides in elements); lg8 = log2(dh / 8)
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ acc, bf16* __restrict__ dq, long long dq_sb, long long dq_sh, int H,
                       int Lq, int Lq_pad, long long total8, int lg8) {
  pdl_prologue();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int dh = 8 << lg8;
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total8; t += stride) {
    const int c = static_cast<int>(t & ((1 << lg8) - 1));   // 8-element group inside the row
    const long long row = t >> lg8;                   // (b*H + h) * Lq + i
    const int i = static_cast<int>(row % Lq);
    const long long bh = row / Lq;
    const float4* src = reinterpret_cast<const float4*>(acc + (bh * Lq_pad + i) * dh + c * 8);
    const float4 a = __ldg(src), b = __ldg(src + 1);
    uint4 o;
    o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
    *reinterpret_cast<uint4*>(dq + (bh / H) * dq_sb + (bh % H) * dq_sh + static_cast<long long>(i) * dh + c * 8) = o;
  }
}

// called by owl_attn_bwd (attn_bwd.cu) after the delta pre-pass, for more than one tile
template <int DH>
static int attn_bwd2_launch(const void* dout, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb, int64_t q_sh,
                            const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb, int64_t v_sh,
                            const float* lse, const float* delta, float* dq_acc, void* dq, int64_t dq_sb, int64_t dq_sh,
                            void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh, int64_t B,
                            int64_t H, int64_t Lq, int64_t Lk, const owl_attn_mask* mask, cudaStream_t s) {
  using Cfg = Bwd2Cfg<DH>;
  if (dq_acc == nullptr) return OWL_ERR_BAD_ARG;
  AttnBwd2Params p{};
  int rc = fill_mask_params(mask, &p.m);
  if (rc != OWL_OK) return rc;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.n_qtiles = static_cast<int>((Lq + kTile - 1) / kTile);
  p.n_ktiles = static_cast<int>((Lk + kTile - 1) / kTile);
  p.Lq_pad = p.n_qtiles * kTile;
  const long long total = B * H * static_cast<long long>(p.n_ktiles);
  const long long acc_rows = B * H * static_cast<long long>(p.Lq_pad);
  if (total > INT32_MAX || acc_rows > INT32_MAX) return OWL_ERR_UNSUPPORTED;
  p.total_items = static_cast<int>(total);
  p.scale = 1.0f / sqrtf(static_cast<float>(DH));
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.dk = static_cast<bf16*>(dk); p.dv = static_cast<bf16*>(dv);
  p.dk_sb = dk_sb; p.dk_sh = dk_sh; p.dv_sb = dv_sb; p.dv_sh = dv_sh;

  CUtensorMap tmQ, tmDO, tmK, tmV, tmDQ;
  if ((rc = make_tmap_4d_bf16(&tmQ, q, DH, Lq, H, B, DH, q_sh, q_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmDO, dout, DH, Lq, H, B, o_sl, DH, o_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k, DH, Lk, H, B, DH, k_sh, k_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v, DH, Lk, H, B, DH, v_sh, v_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_2d_f32(&tmDQ, dq_acc, static_cast<uint64_t>(acc_rows), DH, DH, 32, kTile)) != OWL_OK) return rc;

  static PerDeviceOnce once;
  rc = once.run([] {
    return static_cast<int>(
        cudaFuncSetAttribute(attn_bwd2_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  });
  if (rc != OWL_OK) return rc;
  cudaError_t e = cudaMemsetAsync(dq_acc, 0, static_cast<size_t>(acc_rows) * DH * sizeof(float), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  LaunchOpts lo;
  lo.pdl = false;  // follows a memset node
  const int grid = min(p.total_items, sm_count());
  e = launch_kx(lo, attn_bwd2_kernel<DH>, dim3(static_cast<unsigned>(grid)), dim3(Cfg::kThreads), Cfg::kSmemBytes, s, tmQ,
                tmDO, tmK, tmV, tmDQ, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return static_cast<int>(e);
  }
  if ((rc = launch_status()) != OWL_OK) return rc;
  const long long total8 = B * H * Lq * (DH / 8);
  long long blocks = (total8 + 255) / 256;
  const long long cap = static_cast<long long>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  (void)launch_k(attn_dq_convert_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, s,
                 static_cast<const float*>(dq_acc), static_cast<bf16*>(dq), static_cast<long long>(dq_sb),
                 static_cast<long long>(dq_sh), p.H, p.Lq, p.Lq_pad, total8, DH == 64 ? 3 : 4);
  return launch_status();
}

int attn_bwd2_dispatch(const void* dout, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb, int64_t q_sh,
                       const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb, int64_t v_sh,
                       const float* lse, const float* delta, float* dq_acc, void* dq, int64_t dq_sb, int64_t dq_sh,
                       void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh, int64_t B,
                       int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask, cudaStream_t s) {
  if (dh == 64)
    return attn_bwd2_launch<64>(dout, o_sb, o_sl, q, q_sb, q_sh, k, k_sb, k_sh, v, v_sb, v_sh, lse, delta, dq_acc, dq, dq_sb,
                                dq_sh, dk, dk_sb, dk_sh, dv, dv_sb, dv_sh, B, H, Lq, Lk, mask, s);
  return attn_bwd2_launch<128>(dout, o_sb, o_sl, q, q_sb, q_sh, k, k_sb, k_sh, v, v_sb, v_sh, lse, delta, dq_acc, dq, dq_sb,
                               dq_sh, dk, dk_sb, dk_sh, dv, dv_sb, dv_sh, B, H, Lq, Lk, mask, s);
}

}  // namespace owl
