// Frame-block-causal attention forward, second generation (Lq > 128): persistent, two query tiles per CTA ping-ponging
// over ONE K/V stream, probabilities kept in tensor memory.
//
//   grid = #SMs (1 CTA/SM, the whole 512-column TMEM), 384 threads = 3 warpgroups:
//     WG0 (warps 0-3)   softmax of query tile 0   thread = query row; S_0 -> P_0 (bf16, written over S_0's columns)
//     WG1 (warps 4-7)   softmax of query tile 1
//     WG2 warp 8        TMA producer: Q_0, Q_1, then the ring K_j, V_j, K_j+1, V_j+1, ...
//         warp 9        MMA issuer:   S_t = Q_t K_j^T (SS), O_t += P_t V_j (A operand = P_t read from TMEM)
//   TMEM columns: S_0 [0,128)  S_1 [128,256)  O_0 [256,256+DH)  O_1 [256+DH,256+2DH);  P_t = first 64 columns of S_t.
//
// Why this shape (profiles/r01_ncu_attn_full.md: tensor pipe 15 % active, nothing saturated, one softmax chain per CTA):
//   * while WG0 does exp2 on tile 0 the tensor core runs S/PV of tile 1 and vice versa - the MUFU and tensor pipes
//     overlap inside one CTA instead of only across co-resident CTAs;
//   * P never touches shared memory (no st.shared + proxy fence + smem read by the MMA): shared memory holds only Q and
//     a 4-8 tile K/V ring, and each K/V tile is used by two query tiles (half the L2 -> SM traffic per flop);
//   * persistent CTAs over (batch, head, query-pair) work items, heaviest causal rows first, so barrier init / TMEM
//     allocation happen once and the tail is one work item long;
//   * lazy rescaling: the running maximum used for the exponent only moves when the true maximum grew by more than 2^8
//     (exact: the final O / l normalisation cancels it), so O is read-modify-written in TMEM a handful of times per row;
//   * tiles fully inside every row's visible range take an unpredicated path; tiles outside the union of the rows'
//     ranges - causal, window AND document extents (owl_attn_doc_extents) - are never loaded.
// Replaces flex_attention + BlockMask of owl_wms/nn/attn.py:24-62,106-109 for training / prefill shapes; Lq <= 128
// (audio.yml's L = 120, decode) stays on the single-tile kernel in attn.cu.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include <type_traits>

#include "attn_common.cuh"
#include "host_common.h"

namespace owl {

int fill_mask_params(const owl_attn_mask* mask, MaskParams* m);  // attn.cu

struct AttnFwd2Params {
  int B, H, Lq, Lk, n_qpairs, total_items;
  float scale_log2;  // log2(e) / sqrt(dh)
  bf16* o;
  long long o_sb, o_sl;
  float* lse;
  MaskParams m;
};

template <int DH>
struct Fwd2Cfg {
  static constexpr int kTileBytes = kTile * DH * 2;       // one Q / K / V tile
  static constexpr int kRing = DH == 64 ? 8 : 4;          // K/V tiles in flight (K and V alternate)
  static constexpr int kOffRing = 2 * kTileBytes;         // Q_0, Q_1 first
  static constexpr int kOffBar = kOffRing + kRing * kTileBytes;
  static constexpr int kSmemBytes = kOffBar + 512;
  static constexpr uint32_t kColS0 = 0, kColS1 = 128, kColO0 = 256, kColO1 = 256 + DH;
};

static constexpr int kFwd2Threads = 384;
static constexpr float kRescaleThreshold = 8.0f;  // log2 units: P stays below 2^8 between rescales

// Work item -> (b, h, first query row).  Items are ordered head-major with the heaviest (latest, causal) query pairs of
// a head first; CTAs take them round-robin, so at any moment the resident CTAs work on one or two heads and their K/V
// stream stays in L2, and since n_qpairs is rarely a multiple of the grid the heavy items rotate over the CTAs.
struct WorkItem {
  int b, h, i0;
  int a[2], z[2];  // per query tile: KV tile range [a, z); a == z: the tile does not take part
  int u0, u1;      // union
};

__device__ __forceinline__ WorkItem decode_item(const AttnFwd2Params& p, int w) {
  WorkItem it;
  const int bh = w / p.n_qpairs;
  const int qp = p.n_qpairs - 1 - (w % p.n_qpairs);
  it.b = bh / p.H;
  it.h = bh % p.H;
  it.i0 = qp * 2 * kTile;
  it.u0 = INT32_MAX;
  it.u1 = 0;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int r0 = it.i0 + t * kTile;
    if (r0 >= p.Lq) {
      it.a[t] = it.z[t] = 0;
      continue;
    }
    const int r1 = min(r0 + kTile, p.Lq) - 1;
    kv_tile_range_doc(p.m, it.b, r0, r1, p.Lk, it.a[t], it.z[t]);
    if (it.z[t] > it.a[t]) {
      it.u0 = min(it.u0, it.a[t]);
      it.u1 = max(it.u1, it.z[t]);
    }
  }
  if (it.u1 <= it.u0) it.u0 = it.u1 = 0;
  return it;
}

template <int DH>
__global__ void __launch_bounds__(kFwd2Threads, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const AttnFwd2Params p) {
  using Cfg = Fwd2Cfg<DH>;
  constexpr int kRing = Cfg::kRing;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBar);
  uint64_t* q_full = bars;                   // [2] TMA -> MMA
  uint64_t* q_empty = bars + 2;              // [2] MMA -> TMA (all S_t of the item issued and complete)
  uint64_t* ring_full = bars + 4;            // [kRing]
  uint64_t* ring_empty = bars + 4 + kRing;   // [kRing]
  uint64_t* s_full = bars + 4 + 2 * kRing;   // [2] MMA -> softmax t
  uint64_t* p_full = s_full + 2;             // [2] softmax t -> MMA (128 arrivals)
  uint64_t* pv_done = s_full + 4;            // [2] MMA -> softmax t (O_t holds every PV issued so far)
  uint64_t* o_free = s_full + 6;             // [2] softmax t -> MMA (epilogue has read O_t; 128 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) {
      printf("owl: attention smem base not 1024-aligned\n");
      __trap();
    }
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    for (int t = 0; t < 2; ++t) {
      mbar_init(&q_full[t], 1);
      mbar_init(&q_empty[t], 1);
      mbar_init(&s_full[t], 1);
      mbar_init(&p_full[t], 128);
      mbar_init(&pv_done[t], 1);
      mbar_init(&o_free[t], 128);
    }
    for (int s = 0; s < kRing; ++s) {
      mbar_init(&ring_full[s], 1);
      mbar_init(&ring_empty[s], 1);
    }
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
  pdl_wait();  // everything above overlapped the previous kernel's tail
  pdl_launch();

  if (warp >= 8) {
    // ------------------------------------------------------------------------------------------------------------
    // WG2: TMA producer (warp 8) and MMA issuer (warp 9); warps 10, 11 only hand their registers over
    // ------------------------------------------------------------------------------------------------------------
    reg_dealloc<80>();
    // Both roles: the WHOLE warp walks the schedule (waits included), one elected lane issues each batch of TMA / tcgen05
    // instructions.  (Under `lane == 0` ptxas wraps every UTCHMMA / UTMALDG in an ELECT / BRA.U.ANY loop, ~100 clk per
    // instruction: 30 tcgen05 instructions per KV step then take longer than the step's MMAs and exponentials.)
    if (warp == 8) {
      uint32_t ring_n = 0;       // tiles pushed so far (ring position, persistent over work items)
      uint32_t q_n[2] = {0, 0};  // Q_t loads so far (indexed only inside fully unrolled loops)
      for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
        const WorkItem it = decode_item(p, w);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (it.z[t] <= it.a[t]) continue;
          mbar_wait(&q_empty[t], (q_n[t] & 1) ^ 1);
          ++q_n[t];
          if (elect_one()) {
            mbar_arrive_expect_tx(&q_full[t], Cfg::kTileBytes);
#pragma unroll
            for (int pn = 0; pn < DH / 64; ++pn)
              tma_load_4d(smem + t * Cfg::kTileBytes + pn * kPanelBytes, &tmQ, &q_full[t], pn * 64, it.i0 + t * kTile, it.h,
                          it.b);
          }
          __syncwarp();
        }
        for (int j = it.u0; j < it.u1; ++j) {
#pragma unroll
          for (int kv = 0; kv < 2; ++kv) {
            const uint32_t s = ring_n % kRing;
            mbar_wait(&ring_empty[s], ((ring_n / kRing) & 1) ^ 1);
            ++ring_n;
            uint8_t* dst = smem + Cfg::kOffRing + s * Cfg::kTileBytes;
            if (elect_one()) {
              mbar_arrive_expect_tx(&ring_full[s], Cfg::kTileBytes);
#pragma unroll
              for (int pn = 0; pn < DH / 64; ++pn)
                tma_load_4d(dst + pn * kPanelBytes, kv == 0 ? &tmK : &tmV, &ring_full[s], pn * 64, j * kTile, it.h, it.b);
            }
            __syncwarp();
          }
        }
      }
    } else if (warp == 9) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(kTile, DH, false, true);
      uint32_t ring_n = 0;             // tiles consumed so far
      uint32_t q_n0 = 0, q_n1 = 0;     // items in which tile t took part
      uint32_t p_n0 = 0, p_n1 = 0;     // P_t tiles consumed
      for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
        const WorkItem it = decode_item(p, w);
        if (it.u1 <= it.u0) continue;
        // S_t = Q_t K^T into S_t's columns; commit -> softmax t
        auto issue_s = [&](auto tc, uint32_t sK) {
          constexpr int t = decltype(tc)::value;
          const uint32_t sQ = smem_u32(smem + t * Cfg::kTileBytes);
          const uint32_t tS = tmem_base + (t == 0 ? Cfg::kColS0 : Cfg::kColS1);
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) umma_bf16_ss(tS, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc_s, k > 0);
          umma_commit(&s_full[t]);
        };
        // one (tile, KV tile j) step: O_t += P_t V_j, then S_t of KV tile j+1 right behind it in issue order (it
        // overwrites S_t / P_t only after PV_t(j) has read P_t: the tensor pipe executes in order)
        auto step = [&](auto tc, int j, uint32_t sV, bool more, uint32_t sK, uint32_t& q_n, uint32_t& p_n) {
          constexpr int t = decltype(tc)::value;
          const int a = it.a[t], z = it.z[t];
          const bool pv = j >= a && j < z;
          const bool nxt = more && j + 1 >= a && j + 1 < z;
          if (pv) {
            mbar_wait(&p_full[t], p_n & 1);
            ++p_n;
            if (j == a) mbar_wait(&o_free[t], q_n & 1);  // the previous item's epilogue has read O_t
          }
          if (pv || nxt) {
            tc_fence_after();
            if (elect_one()) {
              if (pv) {
                const uint32_t tP = tmem_base + (t == 0 ? Cfg::kColS0 : Cfg::kColS1);
                const uint32_t tO = tmem_base + (t == 0 ? Cfg::kColO0 : Cfg::kColO1);
#pragma unroll
                for (int k = 0; k < kTile / 16; ++k)
                  umma_bf16_ts(tO, tP + k * 8, desc_mnmajor(sV, k), idesc_o, (j > a || k > 0) ? 1u : 0u);
                umma_commit(&pv_done[t]);
              }
              if (nxt) {
                issue_s(tc, sK);
                if (j + 2 == z) umma_commit(&q_empty[t]);  // that was this item's last S_t: Q_t may be overwritten
              }
            }
            __syncwarp();
          }
        };
        using T0 = std::integral_constant<int, 0>;
        using T1 = std::integral_constant<int, 1>;
        if (it.z[0] > it.a[0]) { mbar_wait(&q_full[0], q_n0 & 1); ++q_n0; }
        if (it.z[1] > it.a[1]) { mbar_wait(&q_full[1], q_n1 & 1); ++q_n1; }
        {  // K of the first KV tile
          const uint32_t s = ring_n % kRing;
          mbar_wait(&ring_full[s], (ring_n / kRing) & 1);
          ++ring_n;
          tc_fence_after();
          const uint32_t sK = smem_u32(smem + Cfg::kOffRing + s * Cfg::kTileBytes);
          if (elect_one()) {
            if (it.u0 >= it.a[0] && it.u0 < it.z[0]) {
              issue_s(T0{}, sK);
              if (it.u0 + 1 == it.z[0]) umma_commit(&q_empty[0]);
            }
            if (it.u0 >= it.a[1] && it.u0 < it.z[1]) {
              issue_s(T1{}, sK);
              if (it.u0 + 1 == it.z[1]) umma_commit(&q_empty[1]);
            }
            umma_commit(&ring_empty[s]);
          }
          __syncwarp();
        }
        for (int j = it.u0; j < it.u1; ++j) {
          const uint32_t sv = ring_n % kRing;  // V_j
          mbar_wait(&ring_full[sv], (ring_n / kRing) & 1);
          ++ring_n;
          const uint32_t sV = smem_u32(smem + Cfg::kOffRing + sv * Cfg::kTileBytes);
          const bool more = j + 1 < it.u1;
          uint32_t sk = 0, sK = 0;
          if (more) {  // K_{j+1} follows V_j in the ring
            sk = ring_n % kRing;
            mbar_wait(&ring_full[sk], (ring_n / kRing) & 1);
            ++ring_n;
            sK = smem_u32(smem + Cfg::kOffRing + sk * Cfg::kTileBytes);
          }
          step(T0{}, j, sV, more, sK, q_n0, p_n0);
          step(T1{}, j, sV, more, sK, q_n1, p_n1);
          if (elect_one()) {
            umma_commit(&ring_empty[sv]);
            if (more) umma_commit(&ring_empty[sk]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------------------------------------------------
    // WG0 / WG1: softmax of query tile t, thread = query row
    // ------------------------------------------------------------------------------------------------------------
    reg_alloc<208>();
    const int t = warp >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tS = tmem_base + (t == 0 ? Cfg::kColS0 : Cfg::kColS1) + lane_addr;
    const uint32_t tO = tmem_base + (t == 0 ? Cfg::kColO0 : Cfg::kColO1) + lane_addr;
    uint32_t n_s = 0;  // S tiles consumed so far (== P tiles produced == PV commits expected), persistent
    uint32_t n_items = 0;  // items in which this tile took part (o_free phase)
    for (int w = blockIdx.x; w < p.total_items; w += gridDim.x) {
      const WorkItem it = decode_item(p, w);
      const int i = it.i0 + t * kTile + row;
      const int a_t = t == 0 ? it.a[0] : it.a[1];   // (no dynamic indexing: keeps the item in registers)
      const int n_kv = (t == 0 ? it.z[0] : it.z[1]) - a_t;
      bf16* orow = p.o + it.b * p.o_sb + static_cast<long long>(i) * p.o_sl + static_cast<long long>(it.h) * DH;
      float* lse_ptr = p.lse != nullptr ? p.lse + (static_cast<long long>(it.b) * p.H + it.h) * p.Lq + i : nullptr;
      if (n_kv <= 0) {  // nothing visible (or rows beyond Lq): zeros, lse = -inf
        if (i < p.Lq) {
#pragma unroll
          for (int c = 0; c < DH / 8; ++c) *reinterpret_cast<uint4*>(orow + c * 8) = make_uint4(0, 0, 0, 0);
          if (lse_ptr != nullptr) *lse_ptr = -INFINITY;
        }
        continue;
      }
      int lo, hi;
      bool doc_test;
      row_range_doc(p.m, it.b, min(i, p.Lq - 1), p.Lk, lo, hi, doc_test);
      if (i >= p.Lq) hi = lo;  // padding rows see nothing
      const int* doc_row = doc_test ? p.m.doc + it.b * p.m.doc_ld : nullptr;
      const int dq = doc_row != nullptr ? __ldg(doc_row + (min(i, p.Lq - 1) + p.m.q_offset) / p.m.T) : 0;
      float m_run = -INFINITY, l_run = 0.f;
      for (int jj = 0; jj < n_kv; ++jj) {
        const int j0 = (a_t + jj) * kTile;
        mbar_wait(&s_full[t], n_s & 1);
        tc_fence_after();
        uint32_t r[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_x32(tS + c * 32, r[c]);
        tmem_ld_wait();
        const bool full = doc_row == nullptr && lo <= j0 && j0 + kTile <= hi;
        if (!full) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint32_t valid = chunk_valid_bits(j0 + c * 32, lo, hi, doc_row, dq, p.m.T);
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (!((valid >> e) & 1u)) r[c][e] = 0xff800000u;  // -inf
          }
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          mx0 = fmaxf(mx0, __uint_as_float(r[0][e]));
          mx1 = fmaxf(mx1, __uint_as_float(r[1][e]));
          mx2 = fmaxf(mx2, __uint_as_float(r[2][e]));
          mx3 = fmaxf(mx3, __uint_as_float(r[3][e]));
        }
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;  // -inf stays -inf (scale > 0)
        // lazy running maximum
        float alpha = 1.f;
        if (m_run == -INFINITY) {
          m_run = mx;  // O_t row and l are still exactly zero: nothing to rescale
        } else if (mx > m_run + kRescaleThreshold) {
          alpha = exp2f(m_run - mx);
          m_run = mx;
        }
        if (jj > 0 && !__all_sync(0xffffffffu, alpha == 1.f)) {
          mbar_wait(&pv_done[t], (n_s - 1) & 1);  // O_t holds PV(jj-1)
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < DH / 32; ++c) {
            uint32_t o[32];
            tmem_ld_x32(tO + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
            tmem_st_x32(tO + c * 32, o);
          }
          l_run *= alpha;
        }
        const float m_use = m_run == -INFINITY ? 0.f : m_run;
        const float neg_m = -m_use;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int c = 0; c < 4; c += 2) {
          uint32_t pk[32];
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              const float p0 = exp2f(fmaf(__uint_as_float(r[c + cc][e]), p.scale_log2, neg_m));
              const float p1 = exp2f(fmaf(__uint_as_float(r[c + cc][e + 1]), p.scale_log2, neg_m));
              rs0 += p0;
              rs1 += p1;
              pk[cc * 16 + (e >> 1)] = pack_bf16(p0, p1);
            }
          }
          tmem_st_x32(tS + (c >> 1) * 32, pk);  // P_t: keys (2n, 2n+1) of this row in 32-bit column n of S_t's region
        }
        l_run += rs0 + rs1;
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_full[t]);
        ++n_s;
      }
      // epilogue: O / l -> global (token-major [B, Lq, H*DH]), lse
      mbar_wait(&pv_done[t], (n_s - 1) & 1);
      tc_fence_after();
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t o[32];
        tmem_ld_x32(tO + c * 32, o);
        tmem_ld_wait();
        if (i < p.Lq) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[8 * g + 0]) * inv, __uint_as_float(o[8 * g + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[8 * g + 2]) * inv, __uint_as_float(o[8 * g + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[8 * g + 4]) * inv, __uint_as_float(o[8 * g + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[8 * g + 6]) * inv, __uint_as_float(o[8 * g + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c * 32 + g * 8) = v;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&o_free[t]);
      ++n_items;
      if (lse_ptr != nullptr && i < p.Lq)
        *lse_ptr = l_run > 0.f ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
    }
    (void)n_items;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int DH>
static int launch_attn_fwd2(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                            const AttnFwd2Params& p, cudaStream_t s) {
  using Cfg = Fwd2Cfg<DH>;
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    return static_cast<int>(
        cudaFuncSetAttribute(attn_fwd2_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  });
  if (rc0 != OWL_OK) return rc0;
  const int grid = min(p.total_items, sm_count());
  (void)launch_k(attn_fwd2_kernel<DH>, dim3(static_cast<unsigned>(grid)), dim3(kFwd2Threads), Cfg::kSmemBytes, s, tmQ,
                 tmK, tmV, p);
  return launch_status();
}

// called by owl_attn_fwd (attn.cu) when Lq > 128
int attn_fwd2_dispatch(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                       const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       cudaStream_t s) {
  AttnFwd2Params p{};
  int rc = fill_mask_params(mask, &p.m);
  if (rc != OWL_OK) return rc;
  p.B = static_cast<int>(B); p.H = static_cast<int>(H); p.Lq = static_cast<int>(Lq); p.Lk = static_cast<int>(Lk);
  p.n_qpairs = static_cast<int>((Lq + 2 * kTile - 1) / (2 * kTile));
  const long long total = B * H * static_cast<long long>(p.n_qpairs);
  if (total > INT32_MAX) return OWL_ERR_UNSUPPORTED;
  p.total_items = static_cast<int>(total);
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(dh));
  p.o = static_cast<bf16*>(o); p.o_sb = o_sb; p.o_sl = o_sl; p.lse = lse;
  CUtensorMap tmQ, tmK, tmV;
  if ((rc = make_tmap_4d_bf16(&tmQ, q, dh, Lq, H, B, dh, q_sh, q_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmK, k, dh, Lk, H, B, dh, k_sh, k_sb, 64, kTile)) != OWL_OK) return rc;
  if ((rc = make_tmap_4d_bf16(&tmV, v, dh, Lk, H, B, dh, v_sh, v_sb, 64, kTile)) != OWL_OK) return rc;
  return dh == 64 ? launch_attn_fwd2<64>(tmQ, tmK, tmV, p, s) : launch_attn_fwd2<128>(tmQ, tmK, tmV, p, s);
}

}  // namespace owl
