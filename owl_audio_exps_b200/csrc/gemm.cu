// bf16 GEMM on the 5th-gen tensor cores: TMA -> 128B-swizzled shared memory -> tcgen05.mma with the
// fp32 accumulator in TMEM -> tcgen05.ld -> fused epilogue -> global.
//
// One persistent CTA per SM, 8 warps, warp-specialised:
//   warp 0     TMA producer   (one lane): fills the kStages-deep ring of {A tile, B tile}
//   warp 1     MMA issuer     (one lane): 4 x tcgen05.mma (K=16) per ring slot, commits free the slot
//   warp 2     TMEM allocator (2 x BN fp32 columns: the accumulator is double-buffered so the
//              epilogue of tile i overlaps the main loop of tile i+1)
//   warps 4-7  epilogue: warp w owns TMEM lanes 32*(w%4).., i.e. thread = one output row; reads
//              32 columns at a time and applies bias / SiLU / gate+residual / dSiLU / fp32(+=).
//
// Operand storage is chosen per launch (see owl_dit.h): "K-major" operands (contraction index
// contiguous) are loaded as [rows][64 k] boxes, "MN-major" operands (row/col index contiguous, used by
// dgrad's B = W and both wgrad operands) as 64-wide panels of [64 k][64 mn]; both are what the
// tcgen05 shared-memory descriptors in sm100.cuh describe, so no operand is ever transposed in memory.
//
// Split-K (fp32 output only) turns the store into red.global.add so small weight-gradient GEMMs
// still fill 148 SMs.
#include <cuda_bf16.h>
#include <stdlib.h>

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int kGemmThreads = 384;  // 4 control warps + 2 epilogue groups of 4 warps
static constexpr int kSlotBytes = 16384;   // one staging panel: 128 rows x 128 bytes, 128B-swizzled
static constexpr int kBarRegion = 2048;    // mbarriers + TMEM slot (first 1 KB), bias rows of the two epilogue groups (2 x 512 B);
                                           // 2 KB still leaves room for 4 x 48 KB stages + 2 staging panels in 227 KB
static constexpr int kMaxStages = 8;
static constexpr int kSmemLimit = 232448;  // 227 KB opt-in maximum per CTA

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, splits, kb_per_split, kb_total;
  int mode, rows_per_frame;
  int stages;      // depth of the operand ring (runtime: whatever the staging panels leave free)
  int f32_reduce;  // fp32 output goes out as TMA reduce-add (accumulate / split-K) instead of a store
  int has_out1, aux0_tma, aux1_tma;
  const float* bias;
  const bf16* aux1;  // gate rows for rows_per_frame > 1 (read directly: 32 consecutive rows share <= 2 gate rows)
  long long lda1;
};

// CTAS = 2: a CTA pair (cta_group::2) computes a 256 x BN tile; each CTA owns 128 rows of A / D and loads
// BN/2 rows of B.
template <int BN, int CTAS = 1>
struct GemmCfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = (BN / CTAS) * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
};

// sigmoid through ONE MUFU op (tanh.approx) instead of ex2 + rcp: the epilogue of a K = d GEMM has only
// ~8k cycles per 128x256 tile, and 2 MUFU ops x 32k elements at 16/clk/SM would eat half of that.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float silu_f(float z) {
  const float h = 0.5f * z;
  return fmaf(h, tanh_approx(h), h);
}
__device__ __forceinline__ float dsilu_f(float z) {
  const float s = fmaf(0.5f, tanh_approx(0.5f * z), 0.5f);
  return s * fmaf(z, 1.f - s, 1.f);
}
__device__ __forceinline__ void unpack8u(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8u(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}
// byte offset of 16-byte chunk c (0..7) of row r inside a 128B-swizzled [128][128 B] panel
__device__ __forceinline__ uint32_t panel_off(int r, int c) { return r * 128 + ((c ^ (r & 7)) << 4); }

// Persistent warp-specialised GEMM.  Epilogue: each of the 128 epilogue threads owns one accumulator row
// (the tcgen05.ld 32x32b layout); results are packed into a 128B-swizzled shared-memory panel and leave the
// SM as ONE TMA store per 128x64 (bf16) / 128x32 (fp32) panel, so global writes are full 128-byte lines.
// Residual / pre-activation / gate operands arrive the same way (TMA load, prefetched one panel ahead).
// F32ONLY = the fp32-output epilogue (weight gradients) compiled on its own: without the bf16 epilogues' live ranges
// it fits kF32Regs registers per thread, so 384 threads leave a third of the register file to a co-resident CTA
// of a memory-bound kernel from another stream (weight gradients run on the side stream, see functional._Side).
static constexpr int kF32Regs = 104;

template <int BN, bool A_MN, bool B_MN, int CTAS, bool F32ONLY>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO0,
                                          const CUtensorMap& tmO1, const CUtensorMap& tmX0, const CUtensorMap& tmX1,
                                          const GemmParams& p) {
  using Cfg = GemmCfg<BN, CTAS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* tfull_bar = empty_bar + kMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* aux_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux_bar + 2);
  const int kStages = p.stages;
  uint8_t* smem_a = smem + kBarRegion;
  uint8_t* smem_b = smem_a + kStages * Cfg::kABytes;
  uint8_t* stag = smem_b + kStages * Cfg::kBBytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = CTAS == 2 ? static_cast<int>(cluster_ctarank()) : 0;          // 0 = leader of the pair
  const int worker = CTAS == 2 ? static_cast<int>(cluster_id_x()) : static_cast<int>(blockIdx.x);
  const int n_workers = CTAS == 2 ? static_cast<int>(n_clusters_x()) : static_cast<int>(gridDim.x);
  constexpr int TM = BM * CTAS;  // rows of one scheduled tile

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO0);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 256 * CTAS);  // pair: the peer's epilogue threads arrive on the leader's barrier
      mbar_init(&aux_bar[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if (CTAS == 2) {
      tmem_alloc_2cta(tmem_slot, Cfg::kTmemCols);
      tmem_relinquish_2cta();
    } else {
      tmem_alloc(tmem_slot, Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (CTAS == 2) cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above (barriers, TMEM, descriptor prefetch) overlapped the previous kernel's tail; from here on
  // global memory is touched (TMA loads, bias, stores)
  pdl_wait();
  pdl_launch();

  const int total_tiles = p.tiles_m * p.tiles_n * p.splits;

  // Producer and MMA issuer: the WHOLE warp walks the loop (barrier waits included) and one elected lane issues the TMA /
  // tcgen05 instructions of a k-block as one batch.  Under a divergent `if (lane == 0)` ptxas wraps every UTCHMMA / UTMALDG
  // in an ELECT / BRA.U.ANY loop and moves its operands to uniform registers one at a time (~100 clk per instruction, seen
  // in the ncu source view of the attention backward) - as long as the 128 clk a 128x256x16 MMA takes on the tensor pipe.
  if (warp == 0) {
    {
      int stage = 0;
      uint32_t phase = 0;
      auto load = [&](void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
        if (CTAS == 2) tma_load_2d_2cta(dst, tm, bar, c0, c1);  // bytes are counted on the leader's barrier
        else tma_load_2d(dst, tm, bar, c0, c1);
      };
      for (int t = worker; t < total_tiles; t += n_workers) {
        const int tm = t % p.tiles_m;
        const int tn = (t / p.tiles_m) % p.tiles_n;
        const int ks = t / (p.tiles_m * p.tiles_n);
        const int m0 = tm * TM + rank * BM, n0 = tn * BN + rank * (BN / CTAS);
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * Cfg::kABytes;
          uint8_t* sb = smem_b + stage * Cfg::kBBytes;
          const int k0 = kb * BK;
          if (elect_one()) {
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes * CTAS);
            if (A_MN) {
#pragma unroll
              for (int pn = 0; pn < BM / 64; ++pn) load(sa + pn * 8192, &tmA, &full_bar[stage], m0 + pn * 64, k0);
            } else {
              load(sa, &tmA, &full_bar[stage], k0, m0);
            }
            if (B_MN) {
#pragma unroll
              for (int pn = 0; pn < BN / CTAS / 64; ++pn) load(sb + pn * 8192, &tmB, &full_bar[stage], n0 + pn * 64, k0);
            } else {
              load(sb, &tmB, &full_bar[stage], k0, n0);
            }
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(TM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int t = worker; t < total_tiles; t += n_workers) {
        const int ks = t / (p.tiles_m * p.tiles_n);
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + static_cast<uint32_t>(as * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * Cfg::kABytes);
          const uint32_t sb = smem_u32(smem_b + stage * Cfg::kBBytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, 8192, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
              const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, 8192, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
              if (CTAS == 2) umma_bf16_ss_2cta(tacc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
              else umma_bf16_ss(tacc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            // slot reusable once these MMAs have read it (pair: in both CTAs)
            if (CTAS == 2) umma_commit_2cta(&empty_bar[stage]);
            else umma_commit(&empty_bar[stage]);
            // accumulator complete (pair: each CTA's epilogue reads its own 128 rows from its own TMEM)
            if (kb + 1 == kb1) {
              if (CTAS == 2) umma_commit_2cta(&tfull_bar[as]);
              else umma_commit(&tfull_bar[as]);
            }
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (kb1 <= kb0) {   // (no k-block: cannot happen with the planner's splits; the epilogue must still be released)
          if (elect_one()) {
            if (CTAS == 2) umma_commit_2cta(&tfull_bar[as]);
            else umma_commit(&tfull_bar[as]);
          }
          __syncwarp();
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // Two epilogue groups of 4 warps; group g owns panels g, g+2, ... of every tile, with its own staging
    // slots, named barrier, TMA-store queue and aux operand slot, so the latencies of one group's panel
    // (tcgen05.ld, TMA-store drain, barriers) overlap the other group's.
    const int grp = (warp - 4) >> 2;
    const int q = warp & 3;           // TMEM lane quarter this warp may access
    const int r = q * 32 + lane;      // row inside the tile
    const bool leader = (threadIdx.x & 127) == 0;
    const int bar_id = 1 + grp;
    const bool f32out = F32ONLY || p.mode == OWL_EPI_F32;
    const bool has_aux = !F32ONLY && (p.aux0_tma | p.aux1_tma) != 0;
    const int P = f32out ? BN / 32 : BN / 64;  // panels per tile
    const int pcols = f32out ? 32 : 64;
    const int n_out = 1 + p.has_out1;
    uint8_t* s_out0 = stag + grp * n_out * kSlotBytes;
    uint8_t* s_out1 = s_out0 + kSlotBytes;
    uint8_t* s_aux0 = stag + 2 * n_out * kSlotBytes + grp * kSlotBytes;            // [2] slots: one per group
    uint8_t* s_aux1 = stag + (2 * n_out + 2 * p.aux0_tma) * kSlotBytes + grp * kSlotBytes;
    // this tile's bias row, staged once per tile: a global load per 8-column chunk inside the panel loop put
    // an L2 round trip on the epilogue's critical path 8 times per panel (ncu: 30 % of epilogue samples)
    // (a group only needs the 2 x 64 columns of ITS panels: 512 B, second KB of the barrier region)
    float4* s_bias = reinterpret_cast<float4*>(smem + 1024) + grp * 32;
    const bool has_bias = p.bias != nullptr && !f32out;
    const uint32_t aux_bytes = static_cast<uint32_t>((p.aux0_tma + p.aux1_tma) * kSlotBytes);
    auto issue_aux = [&](int t, int pn) {  // group leader only
      const int m0 = (t % p.tiles_m) * TM + rank * BM, n0 = ((t / p.tiles_m) % p.tiles_n) * BN + pn * 64;
      mbar_arrive_expect_tx(&aux_bar[grp], aux_bytes);
      if (p.aux0_tma) tma_load_2d(s_aux0, &tmX0, &aux_bar[grp], n0, m0);
      if (p.aux1_tma) tma_load_2d(s_aux1, &tmX1, &aux_bar[grp], n0, m0);
    };
    int as = 0;
    uint32_t aphase = 0;
    uint32_t xphase = 0;  // parity of this group's aux barrier
    auto release_tmem = [&](int buf) {  // this thread is done reading accumulator buffer `buf`
      tc_fence_before();
      if (CTAS == 2) mbar_arrive_cluster(&tempty_bar[buf], 0);
      else mbar_arrive(&tempty_bar[buf]);
    };
    if (leader && has_aux && grp < P && worker < total_tiles) issue_aux(worker, grp);
    for (int t = worker; t < total_tiles; t += n_workers) {
      const int tm = t % p.tiles_m;
      const int tn = (t / p.tiles_m) % p.tiles_n;
      const int m0 = tm * TM + rank * BM, n0 = tn * BN;
      const long long row = m0 + r;
      if (has_bias) {
        // safe to overwrite: every thread of the group finished reading the previous tile's row before the
        // last named barrier of that tile; visible to the group after the first named barrier below
        const int i = threadIdx.x & 127;
        if (i < 32 && (grp + 2 * (i >> 4)) * 64 < BN) {  // 16 float4 per panel; local panel i>>4 = panel grp + 2*(i>>4)
          const int col = n0 + (grp + 2 * (i >> 4)) * 64 + (i & 15) * 4;
          s_bias[i] = col < p.N ? __ldg(reinterpret_cast<const float4*>(p.bias + col)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as * BN);
      if (grp >= P) release_tmem(as);  // BN = 64 bf16: a single panel, nothing for the second group in this tile
#pragma unroll 1
      for (int pn = grp; pn < P; pn += 2) {
        const bool last = pn + 2 >= P;
        uint4 x0r[8];
        if (has_aux) {
          mbar_wait(&aux_bar[grp], xphase);
          xphase ^= 1;
          if (p.aux0_tma) {
#pragma unroll
            for (int c = 0; c < 8; ++c) x0r[c] = *reinterpret_cast<const uint4*>(s_aux0 + panel_off(r, c));
          }
        }
        uint32_t acc[64];
        {
          uint32_t (&lo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&acc[0]);
          tmem_ld_x32(taddr + pn * pcols, lo);
          if (!f32out) {
            uint32_t (&hi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&acc[32]);
            tmem_ld_x32(taddr + pn * pcols + 32, hi);
          }
          tmem_ld_wait();
        }
        if (last) release_tmem(as);  // this thread has read all it needs: hand the buffer back to the MMA warp
        if (leader) tma_store_wait_read0();  // this group's previous stores have drained its staging slots
        named_bar_sync(bar_id, 128);         // ... and every thread holds its aux0 chunks in registers
        if (leader && has_aux && !p.aux1_tma) {  // refill the aux slot for this group's next panel right away
          if (!last) issue_aux(t, pn + 2);
          else if (t + n_workers < total_tiles) issue_aux(t + n_workers, grp);
        }
        const int c0 = n0 + pn * pcols;
        if (f32out) {
#pragma unroll
          for (int c = 0; c < 8; ++c)
            *reinterpret_cast<uint4*>(s_out0 + panel_off(r, c)) =
                make_uint4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
        } else {
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int col = c0 + c * 8;
            float a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = __uint_as_float(acc[c * 8 + j]);
            if (has_bias) {
              const float4 b0 = s_bias[(pn >> 1) * 16 + c * 2];
              const float4 b1 = s_bias[(pn >> 1) * 16 + c * 2 + 1];
              a[0] += b0.x; a[1] += b0.y; a[2] += b0.z; a[3] += b0.w;
              a[4] += b1.x; a[5] += b1.y; a[6] += b1.z; a[7] += b1.w;
            }
            if (p.mode == OWL_EPI_STORE) {
              *reinterpret_cast<uint4*>(s_out0 + panel_off(r, c)) = pack8u(a);
            } else if (p.mode == OWL_EPI_BIAS_SILU) {
              // the reference's Linear output is bf16 before F.silu (mlp.py:34-35): pack once (F2FP, full rate) and
              // read the rounded values back with shifts instead of a per-element F2F round trip (quarter rate)
              const uint4 zp = pack8u(a);
              unpack8u(zp, a);
              float sv[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) sv[j] = silu_f(a[j]);
              *reinterpret_cast<uint4*>(s_out0 + panel_off(r, c)) = zp;
              *reinterpret_cast<uint4*>(s_out1 + panel_off(r, c)) = pack8u(sv);
            } else if (p.mode == OWL_EPI_GATE_RESIDUAL) {
              float res[8], gt[8];
              unpack8u(x0r[c], res);
              if (p.aux1_tma) {
                unpack8u(*reinterpret_cast<const uint4*>(s_aux1 + panel_off(r, c)), gt);
              } else if (row < p.M && col < p.N) {
                unpack8u(__ldg(reinterpret_cast<const uint4*>(p.aux1 + (row / p.rows_per_frame) * p.lda1 + col)), gt);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) gt[j] = 0.f;
              }
              const uint4 yp = pack8u(a);
              unpack8u(yp, a);  // bf16-rounded branch output
              if (p.has_out1) *reinterpret_cast<uint4*>(s_out1 + panel_off(r, c)) = yp;
#pragma unroll
              for (int j = 0; j < 8; ++j) res[j] += gt[j] * a[j];
              *reinterpret_cast<uint4*>(s_out0 + panel_off(r, c)) = pack8u(res);
            } else {  // OWL_EPI_DSILU
              float z[8];
              unpack8u(x0r[c], z);
#pragma unroll
              for (int j = 0; j < 8; ++j) a[j] *= dsilu_f(z[j]);
              *reinterpret_cast<uint4*>(s_out0 + panel_off(r, c)) = pack8u(a);
            }
          }
        }
        fence_proxy_async_smem();
        named_bar_sync(bar_id, 128);
        if (leader) {
          if (c0 < p.N) {
            if (p.f32_reduce) tma_reduce_add_2d(&tmO0, s_out0, c0, m0);
            else tma_store_2d(&tmO0, s_out0, c0, m0);
            if (p.has_out1) tma_store_2d(&tmO1, s_out1, c0, m0);
            tma_store_commit();
          }
          if (p.aux1_tma) {  // the gate panel was read from smem during the math: refill only now
            if (!last) issue_aux(t, pn + 2);
            else if (t + n_workers < total_tiles) issue_aux(t + n_workers, grp);
          }
        }
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (leader) tma_store_wait_all0();
  }

  tc_fence_before();
  if (CTAS == 2) cluster_sync_all();  // the peer may still be reading its half of this pair's work
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (CTAS == 2) tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols);
    else tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BN, bool A_MN, bool B_MN, int CTAS>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                 const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                 const GemmParams p) {
  gemm_body<BN, A_MN, B_MN, CTAS, false>(tmA, tmB, tmO0, tmO1, tmX0, tmX1, p);
}

template <int BN, bool A_MN, bool B_MN, int CTAS>
__global__ void __maxnreg__(kF32Regs)
gemm_f32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                const GemmParams p) {
  gemm_body<BN, A_MN, B_MN, CTAS, true>(tmA, tmB, tmO0, tmO1, tmX0, tmX1, p);
}

// ------------------------------------------------------------------------------------------------
// CUDA-core check GEMM (verification aid only)
// ------------------------------------------------------------------------------------------------
__global__ void gemm_check_kernel(int layout, int M, int N, int K, const bf16* __restrict__ A, long long lda,
                                  const bf16* __restrict__ B, long long ldb, float* __restrict__ C, long long ldc) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  for (int m = blockIdx.y; m < M; m += gridDim.y) {
    float acc = 0.f;
    for (int k = 0; k < K; ++k) {
      const float a = __bfloat162float(layout == OWL_GEMM_TN ? A[static_cast<long long>(k) * lda + m]
                                                             : A[static_cast<long long>(m) * lda + k]);
      const float b = __bfloat162float(layout == OWL_GEMM_NT ? B[static_cast<long long>(n) * ldb + k]
                                                             : B[static_cast<long long>(k) * ldb + n]);
      acc = fmaf(a, b, acc);
    }
    C[static_cast<long long>(m) * ldc + n] = acc;
  }
}

struct GemmMaps {
  CUtensorMap A, B, O0, O1, X0, X1;
};

template <int BN, bool A_MN, bool B_MN, int CTAS>
static int launch_gemm(const GemmMaps& tm, const GemmParams& p, int grid, int smem_bytes, bool pdl, cudaStream_t stream) {
  auto kern = p.mode == OWL_EPI_F32 ? gemm_f32_kernel<BN, A_MN, B_MN, CTAS> : gemm_bf16_kernel<BN, A_MN, B_MN, CTAS>;
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_kernel<BN, A_MN, B_MN, CTAS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return static_cast<int>(e);
    e = cudaFuncSetAttribute(gemm_f32_kernel<BN, A_MN, B_MN, CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             kSmemLimit);
    return static_cast<int>(e);
  });
  if (rc0 != OWL_OK) return rc0;
  LaunchOpts o;
  o.pdl = pdl;
  o.cluster_x = CTAS;  // the two CTAs of a pair share a TPC
  cudaError_t e = launch_kx(o, kern, dim3(static_cast<unsigned>(grid)), dim3(kGemmThreads),
                            static_cast<size_t>(smem_bytes), stream, tm.A, tm.B, tm.O0, tm.O1, tm.X0, tm.X1, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return static_cast<int>(e);
  }
  return launch_status();
}

template <int BN, int CTAS>
static int launch_gemm_layout(int layout, const GemmMaps& tm, const GemmParams& p, int grid, int smem_bytes, bool pdl,
                              cudaStream_t stream) {
  switch (layout) {
    case OWL_GEMM_NT: return launch_gemm<BN, false, false, CTAS>(tm, p, grid, smem_bytes, pdl, stream);
    case OWL_GEMM_NN: return launch_gemm<BN, false, true, CTAS>(tm, p, grid, smem_bytes, pdl, stream);
    case OWL_GEMM_TN: return launch_gemm<BN, true, true, CTAS>(tm, p, grid, smem_bytes, pdl, stream);
    default: return OWL_ERR_BAD_ARG;
  }
}

// Tile width and split-K factor by a wave-quantisation model:
//   time ~ waves x (k-blocks per tile x BN x eff(BN) + fixed per-tile cost)
// Narrow tiles pay more shared-memory bandwidth per flop (eff) but quantise better on 148 SMs; split-K
// (fp32 reduce-add outputs only) fills the machine when there are few output tiles.
struct GemmPlan {
  int bn, splits, stages, n_slots, ctas;
};

static int slots_for(const GemmParams& p) { return 2 * (1 + p.has_out1) + 2 * p.aux0_tma + 2 * p.aux1_tma; }

static int stages_for(int bn, int n_slots, int ctas = 1) {
  const int stage_bytes = BM * BK * 2 + (bn / ctas) * BK * 2;
  // coresident_smem_reserve(): shared memory left for a small CTA of ANOTHER kernel on the same SM (the shard reduction of
  // the copy-engine gradient exchange runs next to the backward GEMMs; every CTA costs 1 KB of reserved shared memory)
  int st = (kSmemLimit - coresident_smem_reserve() - 1024 - kBarRegion - n_slots * kSlotBytes) / stage_bytes;
  return st > kMaxStages ? kMaxStages : st;
}

static GemmPlan plan_gemm(long long M, long long N, int kb_total, bool can_split, int n_slots, int sms, int force_ctas) {
  GemmPlan best{64, 1, 2, n_slots, 1};
  double best_cost = 1e300;
  // OWL_GEMM_MODEL (A/B aid): bit 1 = charge the split-K reduce-add traffic (default), bit 0 = prefer single CTAs for
  // one-wave problems. Bit 0 wins every isolated timing (tools/gemm_sweep.py: 0.5-1.2 us per GEMM) and LOSES 0.3 ms on
  // the audio step (9.36 -> 9.68 ms, same box): back-to-back launches of an under-filled grid overlap through PDL in
  // the microbenchmark, dependent kernels of the real step do not. Off.
  static const int model = getenv("OWL_GEMM_MODEL") ? atoi(getenv("OWL_GEMM_MODEL")) : 2;
  // {tile width, CTAs per tile, relative time per k-block of one tile (measured on B200, tools/gemm_one.py and
  // tools/gemm_sweep.py)}: a CTA pair finishes its 256x256 tile in ~0.95x the time one CTA needs for 128x256 (each SM
  // pulls only half of B through L2 and shared memory: -5..9 % on the conditioning-bank GEMMs, equal on the K = 1024
  // block shapes); 128-/64-wide single-CTA tiles are operand-bandwidth bound.
  const int cand_bn[4] = {256, 256, 128, 64};
  const int cand_ctas[4] = {2, 1, 1, 1};
  const double eff[4] = {0.95, 1.0, 1.6, 2.8};
  for (int i = 0; i < 4; ++i) {
    const int bn = cand_bn[i], ctas = cand_ctas[i];
    if (force_ctas && ctas != force_ctas) continue;
    if (bn > 64 && N <= bn / 2) continue;  // mostly padding
    if (ctas == 2 && M <= BM) continue;
    const int stages = stages_for(bn, n_slots, ctas);
    if (stages < 3) continue;
    const int workers = sms / ctas;
    const long long tiles = ((M + BM * ctas - 1) / (BM * ctas)) * ((N + bn - 1) / bn);
    const int max_splits = can_split ? (kb_total / 4 > 0 ? kb_total / 4 : 1) : 1;
    for (int sp = 1; sp <= max_splits && sp <= 32; ++sp) {
      const int kbps = (kb_total + sp - 1) / sp;
      const int real_sp = (kb_total + kbps - 1) / kbps;
      const long long waves = (tiles * real_sp + workers - 1) / workers;
      const bool lone = (model & 1) && ctas == 2 && waves == 1;
      const double e = lone ? 1.0 : eff[i];
      const double fixed = 400.0 + (lone ? 200.0 : 0.0) + (!(model & 2) && real_sp > 1 ? 150.0 : 0.0);
      // Split-K reduce-add: every split pushes its fp32 tile through the L2 atomics (~2.5 TB/s measured; one cost unit
      // is ~1.64 ns, so 4100 bytes per unit); the last wave's share is exposed, earlier waves hide theirs under the
      // next tile's main loop.
      double red = 0.0;
      if ((model & 2) && real_sp > 1) {
        const double bytes = static_cast<double>(tiles) * real_sp * (BM * ctas) * bn * 4.0;
        red = bytes / 4100.0 * (waves > 2 ? 2.0 / static_cast<double>(waves) : 1.0);
      }
      const double cost = static_cast<double>(waves) * (kbps * bn * e + fixed) + red;
      if (cost < best_cost) { best_cost = cost; best = GemmPlan{bn, real_sp, stages, n_slots, ctas}; }
      if (tiles * sp >= 2LL * workers) break;
    }
  }
  return best;
}

}  // namespace owl

using namespace owl;

extern "C" {

int owl_gemm_bf16(int layout, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* B,
                  int64_t ldb, const owl_gemm_epilogue* epi, owl_stream_t stream) {
  if (!A || !B || !epi || M <= 0 || N <= 0 || K <= 0) return OWL_ERR_BAD_ARG;
  if (layout < OWL_GEMM_NT || layout > OWL_GEMM_TN) return OWL_ERR_BAD_ARG;
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) return OWL_ERR_UNSUPPORTED;
  if (N % 8 != 0) return OWL_ERR_UNSUPPORTED;
  if (lda % 8 || ldb % 8 || !aligned16(A) || !aligned16(B)) return OWL_ERR_ALIGN;
  const bool a_mn = layout == OWL_GEMM_TN;
  const bool b_mn = layout != OWL_GEMM_NT;
  if ((a_mn ? lda < M : lda < K) || (b_mn ? ldb < N : ldb < K)) return OWL_ERR_BAD_ARG;

  GemmParams p{};
  p.M = static_cast<int>(M); p.N = static_cast<int>(N); p.K = static_cast<int>(K);
  p.mode = epi->mode;
  p.rows_per_frame = epi->rows_per_frame > 0 ? epi->rows_per_frame : 1;
  p.bias = epi->bias;
  const void* out0 = epi->out0;
  const void* out1 = epi->out1;
  const void* aux0 = epi->aux0;
  const void* aux1 = epi->aux1;

  // per-mode argument validation
  bool f32_out = false;
  switch (p.mode) {
    case OWL_EPI_STORE:
      if (!out0) return OWL_ERR_BAD_ARG;
      out1 = nullptr;
      break;
    case OWL_EPI_BIAS_SILU:
      if (!out0 || !out1) return OWL_ERR_BAD_ARG;
      break;
    case OWL_EPI_GATE_RESIDUAL:
      if (!out0 || !aux0 || !aux1) return OWL_ERR_BAD_ARG;
      if (M % p.rows_per_frame) return OWL_ERR_BAD_ARG;
      break;
    case OWL_EPI_DSILU:
      if (!aux0) return OWL_ERR_BAD_ARG;
      if (!out0) return OWL_ERR_UNSUPPORTED;  // fp32 destination for the silu' epilogue is not implemented
      out1 = nullptr;
      break;
    case OWL_EPI_F32:
      if (!epi->outf) return OWL_ERR_BAD_ARG;
      f32_out = true;
      out1 = nullptr;
      break;
    default: return OWL_ERR_BAD_ARG;
  }
  if ((out0 && !f32_out && (epi->ld0 % 8 || !aligned16(out0))) || (out1 && (epi->ld1 % 8 || !aligned16(out1))) ||
      (f32_out && (epi->ldf % 4 || !aligned16(epi->outf))) || (aux0 && (epi->lda0 % 8 || !aligned16(aux0))) ||
      (aux1 && (epi->lda1 % 8 || !aligned16(aux1))) || (p.bias && !aligned16(p.bias)))
    return OWL_ERR_ALIGN;
  const bool uses_aux = p.mode == OWL_EPI_GATE_RESIDUAL || p.mode == OWL_EPI_DSILU;
  p.has_out1 = out1 != nullptr;
  p.aux0_tma = uses_aux ? 1 : 0;
  p.aux1_tma = (p.mode == OWL_EPI_GATE_RESIDUAL && p.rows_per_frame == 1) ? 1 : 0;
  p.aux1 = static_cast<const bf16*>(aux1);
  p.lda1 = epi->lda1;

  const int sms = gemm_sm_count() & ~1;  // even: CTA pairs
  const int kb_total = static_cast<int>((K + BK - 1) / BK);
  int force_ctas = 0;
  if (const char* e = getenv("OWL_GEMM_CTAS")) force_ctas = atoi(e) == 2 ? 2 : (atoi(e) == 1 ? 1 : 0);
  GemmPlan plan = plan_gemm(M, N, kb_total, f32_out, slots_for(p), sms, force_ctas);
  if (const char* e = getenv("OWL_GEMM_BN")) {  // tuning aid (tools/gemm_sweep.py): force the tile width
    const int v = atoi(e);
    if ((v == 64 || v == 128 || v == 256) && stages_for(v, plan.n_slots) >= 2) {
      plan.bn = v;
      plan.ctas = (v == 256 && force_ctas == 2) ? 2 : 1;
      plan.stages = stages_for(v, plan.n_slots, plan.ctas);
      if (!f32_out) plan.splits = 1;
    }
  }
  if (const char* e = getenv("OWL_GEMM_SPLITS")) {
    const int v = atoi(e);
    if (f32_out && v >= 1 && v <= kb_total) plan.splits = v;
  }
  if (const char* e = getenv("OWL_GEMM_STAGES")) {
    const int v = atoi(e);
    if (v >= 2 && v <= plan.stages) plan.stages = v;
  }
  const int bn = plan.bn, ctas = plan.ctas;
  const int tiles_m = static_cast<int>((M + BM * ctas - 1) / (BM * ctas));
  const int tiles_n = static_cast<int>((N + bn - 1) / bn);
  int splits = plan.splits;
  const int kb_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per_split - 1) / kb_per_split;
  p.tiles_m = tiles_m; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per_split; p.kb_total = kb_total;
  p.f32_reduce = (f32_out && (splits > 1 || epi->accumulate)) ? 1 : 0;

  // shared memory: barrier region + operand ring + staging panels
  const int stage_bytes = BM * BK * 2 + (bn / ctas) * BK * 2;
  int stages = plan.stages;
  if (stages > kb_per_split + 1) stages = kb_per_split + 1;
  if (stages < 2) stages = 2;
  p.stages = stages;
  const int smem_bytes = 1024 + kBarRegion + stages * stage_bytes + plan.n_slots * kSlotBytes;

  cudaStream_t s = static_cast<cudaStream_t>(stream);
  bool pdl = true;
  if (f32_out && splits > 1 && !epi->accumulate) {
    pdl = false;  // the kernel follows a memset node, not a kernel
    cudaError_t e = cudaMemset2DAsync(epi->outf, static_cast<size_t>(epi->ldf) * 4, 0, static_cast<size_t>(N) * 4,
                                      static_cast<size_t>(M), s);
    if (e != cudaSuccess) return static_cast<int>(e);
  }

  GemmMaps tm;
  int rc;
  if (a_mn) rc = make_tmap_2d_bf16(&tm.A, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 64, 64);
  else rc = make_tmap_2d_bf16(&tm.A, A, static_cast<uint64_t>(M), static_cast<uint64_t>(K), lda, 64, BM);
  if (rc != OWL_OK) return rc;
  if (b_mn) rc = make_tmap_2d_bf16(&tm.B, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64, 64);
  else rc = make_tmap_2d_bf16(&tm.B, B, static_cast<uint64_t>(N), static_cast<uint64_t>(K), ldb, 64, bn / ctas);
  if (rc != OWL_OK) return rc;
  if (f32_out) rc = make_tmap_2d_f32(&tm.O0, epi->outf, static_cast<uint64_t>(M), static_cast<uint64_t>(N), epi->ldf, 32, BM);
  else rc = make_tmap_2d_bf16(&tm.O0, out0, static_cast<uint64_t>(M), static_cast<uint64_t>(N), epi->ld0, 64, BM);
  if (rc != OWL_OK) return rc;
  tm.O1 = tm.O0; tm.X0 = tm.O0; tm.X1 = tm.O0;  // placeholders for unused maps
  if (out1 && (rc = make_tmap_2d_bf16(&tm.O1, out1, static_cast<uint64_t>(M), static_cast<uint64_t>(N), epi->ld1, 64, BM)) != OWL_OK)
    return rc;
  if (p.aux0_tma && (rc = make_tmap_2d_bf16(&tm.X0, aux0, static_cast<uint64_t>(M), static_cast<uint64_t>(N), epi->lda0, 64, BM)) != OWL_OK)
    return rc;
  if (p.aux1_tma && (rc = make_tmap_2d_bf16(&tm.X1, aux1, static_cast<uint64_t>(M), static_cast<uint64_t>(N), epi->lda1, 64, BM)) != OWL_OK)
    return rc;

  const long long total = static_cast<long long>(tiles_m) * tiles_n * splits;
  const int workers = sms / ctas;
  const int grid = static_cast<int>(total < workers ? total : workers) * ctas;
  if (ctas == 2) return launch_gemm_layout<256, 2>(layout, tm, p, grid, smem_bytes, pdl, s);
  switch (bn) {
    case 64: return launch_gemm_layout<64, 1>(layout, tm, p, grid, smem_bytes, pdl, s);
    case 128: return launch_gemm_layout<128, 1>(layout, tm, p, grid, smem_bytes, pdl, s);
    default: return launch_gemm_layout<256, 1>(layout, tm, p, grid, smem_bytes, pdl, s);
  }
}

int owl_gemm_bf16_check(int layout, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* B,
                        int64_t ldb, float* C, int64_t ldc, owl_stream_t stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) return OWL_ERR_BAD_ARG;
  if (layout < OWL_GEMM_NT || layout > OWL_GEMM_TN) return OWL_ERR_BAD_ARG;
  dim3 grid(static_cast<unsigned>((N + 127) / 128), static_cast<unsigned>(M < 32768 ? M : 32768));
  gemm_check_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      layout, static_cast<int>(M), static_cast<int>(N), static_cast<int>(K), static_cast<const bf16*>(A), lda,
      static_cast<const bf16*>(B), ldb, C, ldc);
  return launch_status();
}

}  // extern "C"
