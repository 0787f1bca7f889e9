// Weight-streaming bf16 GEMM for the decode regime: y[M <= 128, N] = epi(x[M, K] . W[N, K]^T + bias).
//
// One KV-cached model evaluation of the causal rollout (owl_wms/sampling/av_caching_v2.py:89-140,
// inference/causvid_pipeline.py:112-163) runs every nn.Linear of the stack on 1 frame = 65 tokens (or 1 token for the
// audio model, 1 conditioning row per frame): the weights (2 GB at AV geometry) are read once per evaluation and
// nothing is reused, so the GEMMs are HBM-bound and what matters is that ALL SMs pull weight bytes at the same time.
// The training GEMM (gemm.cu) gives a [<=128, N] problem N/BN CTAs that each walk all of K serially (measured:
// 19.8 us per call where the weights stream in 2 us, profiles/r01_launches_av_decode.md).  Here:
//
//   * the problem is cut into tiles_n x splits (n-tile, K-range) chunks, about one per SM, so every SM streams
//     ~1/148th of the weight matrix through a deep TMA ring (up to 8 x 24 KB in flight per SM);
//   * each chunk is a 128 x BN x (64 * kb) tcgen05 MMA with the fp32 accumulator in TMEM (the token rows are padded
//     to 128 by TMA zero fill: no traffic for them, and the tensor pipe is idle anyway);
//   * split-K partials meet in an fp32 workspace tile with red.global.add.v4.f32 (L2 atomics); a per-tile arrival
//     counter elects the LAST chunk's CTA, which reads the sum back, applies the epilogue (bias, bias+SiLU,
//     bias+gate*y+residual) and writes bf16.  That CTA also zeroes the tile and the counter again, so the workspace
//     is zero between launches and no memset node precedes the kernel (one launch per Linear, graph-friendly).
//     OWL_SMALLM_CLUSTER=1 instead makes the <= 8 CTAs of a tile ONE thread-block cluster that reduces through
//     distributed shared memory (no atomics, no workspace): correct (tests) but measured slower, see owl_gemm_smallm;
//   * persistent over chunks (conditioning bank at decode: M = 1 row, N = 6d * n_layers -> 1728 tiles).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

static constexpr int kSmThreads = 192;   // warp 0 TMA, warp 1 MMA + TMEM, warps 2-5 epilogue (thread = token row)
static constexpr int kSmBK = 64;
static constexpr int kSmABytes = 128 * kSmBK * 2;
static constexpr int kSmMaxStages = 8;
static constexpr int kSmSmemLimit = 232448;
// Measured (tools/smallm_bench.py, B200): a ring of 8 stages and one CTA per SM beats 3-4 stages with two CTAs per SM
// (qkv 13.3 vs 14.0 us, conditioning bank 29.5 vs 40.6 us): all k-blocks of a chunk in flight at once means 1 KB of
// consecutive bytes per weight row at the DRAM, and the next kernel's CTA still starts (and prefetches) as soon as this
// kernel's CTA on the same SM has exited.
static constexpr int kSmSmemPerCta = kSmSmemLimit;

struct SmallMParams {
  int M, N, K;
  int tiles_n, splits, kb_per_split, kb_total, stages;
  int mode, rows_per_frame;
  const float* bias;
  bf16* out;
  long long ldo;
  const bf16* res;
  long long ldres;
  const bf16* gate;
  long long ldgate;
  int cluster;          // 1: the splits of a tile form a thread-block cluster and reduce through distributed shared memory
  float* ws;            // [tiles_n][128][BN] fp32, zero between launches
  unsigned* counters;   // [tiles_n], zero between launches
  const char* next_w;   // weights of the NEXT Linear of the schedule (or nullptr): pulled into L2 while this one reduces
  long long next_bytes;
  // adaLN prologue (modulation.py:20-25 as the A-operand transform): A = rms_norm(x) * (1 + a[frame]) + b[frame]
  const bf16* ln_x;     // raw residual rows [M, K] (nullptr: A arrives by TMA as is)
  long long ln_ldx;
  const bf16* ln_a;
  const bf16* ln_b;
  long long ln_ldab;
  int ln_T;
  float ln_eps;
};

// L2 prefetch of [addr, addr + bytes): one bulk request per 16 KB (cp.async.bulk.prefetch.L2; no destination, no barrier)
__device__ __forceinline__ void l2_prefetch_bulk(const char* addr, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(addr), "r"(bytes) : "memory");
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float sm_tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sm_silu(float z) {
  const float h = 0.5f * z;
  return fmaf(h, sm_tanh_approx(h), h);
}

// 32 consecutive output columns of one row: bias (staged in shared memory once per tile), activation or gate+residual,
// four 16-byte bf16 stores.  All global loads of the chunk (residual, gate) are issued before the first use: the
// epilogue of a decode GEMM is a latency chain, not a bandwidth problem.
__device__ __forceinline__ void finish32(const SmallMParams& p, const float* s_bias, int row, int col_in_tile, int n0,
                                         float (&a)[32]) {
  const int col = n0 + col_in_tile;
  if (col >= p.N) return;
  const int ng = min(4, (p.N - col) >> 3);   // N % 8 == 0: whole 8-column groups
  uint4 rv[4], gv[4];
  if (p.mode == OWL_EPI_GATE_RESIDUAL) {
    const bf16* rp = p.res + static_cast<long long>(row) * p.ldres + col;
    const bf16* gp = p.gate + static_cast<long long>(row / p.rows_per_frame) * p.ldgate + col;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (g < ng) {
        rv[g] = __ldg(reinterpret_cast<const uint4*>(rp + g * 8));
        gv[g] = __ldg(reinterpret_cast<const uint4*>(gp + g * 8));
      }
    }
  }
  bf16* op = p.out + static_cast<long long>(row) * p.ldo + col;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    if (g >= ng) break;
    float v[8];
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) v[jj] = a[g * 8 + jj];
    if (p.bias != nullptr) {
      const float4 b0 = *reinterpret_cast<const float4*>(s_bias + col_in_tile + g * 8);
      const float4 b1 = *reinterpret_cast<const float4*>(s_bias + col_in_tile + g * 8 + 4);
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
      v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    if (p.mode != OWL_EPI_STORE) {
      // the Linear's own output is a bf16 tensor in the reference (mlp.py:34-35, modulation.py:41-43): round once
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) v[jj] = round_bf16(v[jj]);
      if (p.mode == OWL_EPI_BIAS_SILU) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) v[jj] = sm_silu(v[jj]);
      } else {  // OWL_EPI_GATE_RESIDUAL: out = res + gate[frame] * y
        const uint32_t rr[4] = {rv[g].x, rv[g].y, rv[g].z, rv[g].w}, gg[4] = {gv[g].x, gv[g].y, gv[g].z, gv[g].w};
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          v[2 * jj] = fmaf(bf16_lo(gg[jj]), v[2 * jj], bf16_lo(rr[jj]));
          v[2 * jj + 1] = fmaf(bf16_hi(gg[jj]), v[2 * jj + 1], bf16_hi(rr[jj]));
        }
      }
    }
    *reinterpret_cast<uint4*>(op + g * 8) =
        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  }
}

// 8 consecutive output columns (one 16-byte bf16 store): the granule of the cluster reduction
__device__ __forceinline__ void finish8(const SmallMParams& p, const float* s_bias, int row, int col_in_tile, int n0,
                                        float (&v)[8]) {
  const int col = n0 + col_in_tile;
  if (col >= p.N) return;
  uint4 rv = make_uint4(0, 0, 0, 0), gv = rv;
  if (p.mode == OWL_EPI_GATE_RESIDUAL) {
    rv = __ldg(reinterpret_cast<const uint4*>(p.res + static_cast<long long>(row) * p.ldres + col));
    gv = __ldg(reinterpret_cast<const uint4*>(p.gate + static_cast<long long>(row / p.rows_per_frame) * p.ldgate + col));
  }
  if (p.bias != nullptr) {
    const float4 b0 = *reinterpret_cast<const float4*>(s_bias + col_in_tile);
    const float4 b1 = *reinterpret_cast<const float4*>(s_bias + col_in_tile + 4);
    v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
    v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
  }
  if (p.mode != OWL_EPI_STORE) {
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) v[jj] = round_bf16(v[jj]);
    if (p.mode == OWL_EPI_BIAS_SILU) {
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) v[jj] = sm_silu(v[jj]);
    } else {
      const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        v[2 * jj] = fmaf(bf16_lo(gg[jj]), v[2 * jj], bf16_lo(rr[jj]));
        v[2 * jj + 1] = fmaf(bf16_hi(gg[jj]), v[2 * jj + 1], bf16_hi(rr[jj]));
      }
    }
  }
  *reinterpret_cast<uint4*>(p.out + static_cast<long long>(row) * p.ldo + col) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

__device__ __forceinline__ float4 ld_dsmem_v4(uint32_t local_saddr, uint32_t rank) {
  float4 v;
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %4, %5;\n\t"
      "ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [ra];\n\t"
      "}\n"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
      : "r"(local_saddr), "r"(rank)
      : "memory");
  return v;
}

template <int BN>
__global__ void __launch_bounds__(kSmThreads, 1)
gemm_smallm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const SmallMParams p) {
  constexpr int kBBytes = BN * kSmBK * 2;
  constexpr int kStageBytes = kSmABytes + kBBytes;
  constexpr uint32_t kTmemCols = 2 * BN;   // double-buffered accumulator
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty_bar = full_bar + kSmMaxStages;
  uint64_t* tfull_bar = empty_bar + kSmMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint32_t* last_flag = tmem_slot + 1;
  const int kStages = p.stages;
  uint8_t* smem_a = smem + 1024;
  uint8_t* smem_b = smem_a + kStages * kSmABytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], p.ln_x != nullptr ? 129 : 1);   // prologue: + the 128 threads that write the A tile
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total = p.tiles_n * p.splits;
  // WEIGHT PREFETCH AHEAD OF THE GRID DEPENDENCY.  The weights do not depend on the previous kernel, only the activations
  // do: under programmatic dependent launch this CTA becomes resident while the previous kernel is still draining (its
  // CTA on this SM has exited, others are in their reduction round trips), so the producer fills the ring with the first
  // weight tiles BEFORE griddepcontrol.wait and HBM keeps streaming across the kernel boundary.
  int pre = 0;
  // (producer and issuer: the whole warp walks the loops, one elected lane issues each batch of TMA / tcgen05 instructions -
  // under `lane == 0` ptxas wraps each of them in an ELECT / BRA.U.ANY loop, ~100 clk per instruction)
  if (warp == 0 && static_cast<int>(blockIdx.x) < total) {
    const int t = blockIdx.x;
    const int tn = p.cluster ? t / p.splits : t % p.tiles_n, ks = p.cluster ? t % p.splits : t / p.tiles_n;
    const int kb0 = ks * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.kb_total);
    const int n_pre = min(kb1 - kb0, kStages);
    if (elect_one()) {
      for (int i = 0; i < n_pre; ++i) {
        mbar_arrive_expect_tx(&full_bar[i], p.ln_x != nullptr ? kBBytes : kStageBytes);
        tma_load_2d(smem_b + i * kBBytes, &tmB, &full_bar[i], (kb0 + i) * kSmBK, tn * BN);
      }
    }
    __syncwarp();
    pre = n_pre > 0 ? n_pre : 0;
  }
  pdl_wait();     // the activations (and the zeroed workspace) come from the previous kernel
  pdl_launch();

  if (warp == 0) {
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int tn = p.cluster ? t / p.splits : t % p.tiles_n, ks = p.cluster ? t % p.splits : t / p.tiles_n;
        const int kb0 = ks * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          const bool tma_a = p.ln_x == nullptr;   // adaLN prologue: the epilogue warps write the A tile instead
          if (pre > 0) {   // this stage's weights are already on their way: add the activations
            --pre;
            if (tma_a && elect_one()) tma_load_2d(smem_a + stage * kSmABytes, &tmA, &full_bar[stage], kb * kSmBK, 0);
          } else {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&full_bar[stage], tma_a ? kStageBytes : kBBytes);
              tma_load_2d(smem_b + stage * kBBytes, &tmB, &full_bar[stage], kb * kSmBK, tn * BN);   // weights: HBM
              if (tma_a) tma_load_2d(smem_a + stage * kSmABytes, &tmA, &full_bar[stage], kb * kSmBK, 0);   // activations: L2
            }
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN, false, false);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int ks = p.cluster ? t % p.splits : t / p.tiles_n;
        const int kb0 = ks * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + static_cast<uint32_t>(as * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * kSmABytes), sb = smem_u32(smem_b + stage * kBBytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kSmBK / 16; ++k)
              umma_bf16_ss(tacc, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                           (kb > kb0 || k > 0) ? 1u : 0u);
            umma_commit(&empty_bar[stage]);
            if (kb + 1 == kb1) umma_commit(&tfull_bar[as]);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (kb1 <= kb0) {
          if (elect_one()) umma_commit(&tfull_bar[as]);
          __syncwarp();
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;            // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;       // token row
    const int et = static_cast<int>(threadIdx.x) - 64;   // 0..127 among the epilogue threads
    const bool row_ok = r < p.M;
    const bool elected = et == 0;
    float* s_bias = reinterpret_cast<float*>(smem + 512);   // this tile's BN bias values
    int as = 0;
    uint32_t aphase = 0;
    if (elected && p.next_w != nullptr) {
      // this CTA's share of the next Linear's weights -> L2.  The decode step is a chain of short dependent kernels;
      // between two GEMMs HBM would idle through reduction round trips, launches and the small kernels in between.  The
      // requests ride behind this kernel's own (already issued) ring fill; the next GEMM then streams from the 126 MB L2.
      const long long chunk = 16384;
      const long long n_chunks = (p.next_bytes + chunk - 1) / chunk;
      for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const long long off = c * chunk;
        const long long left = p.next_bytes - off;
        l2_prefetch_bulk(p.next_w + off, static_cast<unsigned>((left < chunk ? left : chunk) & ~15LL));
      }
    }
    // ---- adaLN prologue: this thread's row statistic once, then the A tile of every k-block of the chunk -------------
    float ln_rstd = 0.f;
    int ln_stage = 0;
    uint32_t ln_phase = 0;
    if (p.ln_x != nullptr && row_ok) {
      // rstd over the WHOLE row (all of K), whatever K-range this CTA multiplies: 16 independent 16-byte loads in flight
      const bf16* xr = p.ln_x + static_cast<long long>(r) * p.ln_ldx;
      float ss = 0.f;
      for (int c0 = 0; c0 < p.K; c0 += 128) {
        uint4 v[16];
#pragma unroll
        for (int g = 0; g < 16; ++g)
          v[g] = c0 + g * 8 < p.K ? __ldg(reinterpret_cast<const uint4*>(xr + c0 + g * 8)) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int g = 0; g < 16; ++g) {
          const uint32_t w[4] = {v[g].x, v[g].y, v[g].z, v[g].w};
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const float lo = bf16_lo(w[jj]), hi = bf16_hi(w[jj]);
            ss = fmaf(lo, lo, fmaf(hi, hi, ss));
          }
        }
      }
      ln_rstd = rsqrtf(ss / static_cast<float>(p.K) + p.ln_eps);
    }
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int tn = p.cluster ? t / p.splits : t % p.tiles_n;
      const int n0 = tn * BN;
      if (p.ln_x != nullptr) {
        const int ks = p.cluster ? t % p.splits : t / p.tiles_n;
        const int kb0 = ks * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        const long long frow = r / p.ln_T;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[ln_stage], ln_phase ^ 1);
          if (row_ok) {
            const int k0 = kb * kSmBK;
            uint8_t* dstrow = smem_a + ln_stage * kSmABytes + r * 128;
            uint4 xv[8], av[8], bv[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const bool in = k0 + c * 8 < p.K;
              xv[c] = in ? __ldg(reinterpret_cast<const uint4*>(p.ln_x + static_cast<long long>(r) * p.ln_ldx + k0 + c * 8))
                         : make_uint4(0, 0, 0, 0);
              av[c] = in ? __ldg(reinterpret_cast<const uint4*>(p.ln_a + frow * p.ln_ldab + k0 + c * 8)) : make_uint4(0, 0, 0, 0);
              bv[c] = in ? __ldg(reinterpret_cast<const uint4*>(p.ln_b + frow * p.ln_ldab + k0 + c * 8)) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const uint32_t xw[4] = {xv[c].x, xv[c].y, xv[c].z, xv[c].w};
              const uint32_t aw[4] = {av[c].x, av[c].y, av[c].z, av[c].w};
              const uint32_t bw[4] = {bv[c].x, bv[c].y, bv[c].z, bv[c].w};
              uint32_t o[4];
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                // same arithmetic as adaln_fwd_kernel (elementwise.cu): f * rstd * (1 + a) + b, one bf16 rounding
                const float lo = bf16_lo(xw[jj]) * ln_rstd * (1.f + bf16_lo(aw[jj])) + bf16_lo(bw[jj]);
                const float hi = bf16_hi(xw[jj]) * ln_rstd * (1.f + bf16_hi(aw[jj])) + bf16_hi(bw[jj]);
                o[jj] = (k0 + c * 8 < p.K) ? pack_bf16(lo, hi) : 0u;
              }
              *reinterpret_cast<uint4*>(dstrow + ((c ^ (r & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          }
          fence_proxy_async_smem();      // generic-proxy writes of the A tile -> visible to tcgen05.mma
          mbar_arrive(&full_bar[ln_stage]);
          if (++ln_stage == kStages) { ln_stage = 0; ln_phase ^= 1; }
        }
      }
      if (p.bias != nullptr) {
        named_bar_sync(1, 128);        // the previous tile's epilogue has read its bias row
        if (et < BN) s_bias[et] = n0 + et < p.N ? __ldg(p.bias + n0 + et) : 0.f;
        named_bar_sync(1, 128);
      }
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as * BN);
      float* wrow = p.ws + (static_cast<long long>(tn) * 128 + r) * BN;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t acc[32];
        tmem_ld_x32(taddr + c * 32, acc);
        tmem_ld_wait();
        if (p.cluster && p.splits > 1) {
          // partial sums of this K-range -> own shared memory, chunk-major [BN/4][128 rows] x 16 B (conflict-free for
          // thread = row); the ring is free: every MMA that read it has completed (tfull)
          float4* P = reinterpret_cast<float4*>(smem_a);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            P[(c * 8 + g) * 128 + r] = make_float4(__uint_as_float(acc[4 * g]), __uint_as_float(acc[4 * g + 1]),
                                                   __uint_as_float(acc[4 * g + 2]), __uint_as_float(acc[4 * g + 3]));
          continue;
        }
        if (!row_ok) continue;
        if (p.splits == 1) {
          float a[32];
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) a[jj] = __uint_as_float(acc[jj]);
          finish32(p, s_bias, r, c * 32, n0, a);
        } else {
#pragma unroll
          for (int g = 0; g < 8; ++g)
            red_add_v4(wrow + c * 32 + g * 4, __uint_as_float(acc[4 * g]), __uint_as_float(acc[4 * g + 1]),
                       __uint_as_float(acc[4 * g + 2]), __uint_as_float(acc[4 * g + 3]));
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[as]);    // accumulator buffer back to the MMA warp
      if (++as == 2) { as = 0; aphase ^= 1; }
      if (p.splits > 1 && !p.cluster) {
        __threadfence();               // this thread's partial sums are at L2 before the arrival is counted
        named_bar_sync(1, 128);
        if (elected) {
          const unsigned prev = atomicAdd(p.counters + tn, 1u);
          const unsigned last = prev == static_cast<unsigned>(p.splits - 1) ? 1u : 0u;
          if (last) __threadfence();   // order the counter read before the workspace reads below
          *last_flag = last;
        }
        named_bar_sync(1, 128);
        const bool last = *last_flag != 0;
        named_bar_sync(1, 128);        // everyone has read the flag before the next tile's election rewrites it
        if (last) {
          if (row_ok) {
#pragma unroll
            for (int hf = 0; hf < BN / 64; ++hf) {
              // the whole 64-column half in flight at once (16 independent 16-byte L2 reads), then zeroed for the next launch
              float4 v[16];
#pragma unroll
              for (int g = 0; g < 16; ++g) v[g] = __ldcg(reinterpret_cast<const float4*>(wrow + hf * 64 + g * 4));
#pragma unroll
              for (int g = 0; g < 16; ++g)
                *reinterpret_cast<float4*>(wrow + hf * 64 + g * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int cc = 0; cc < 2; ++cc) {
                float a[32];
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                  a[4 * g] = v[cc * 8 + g].x; a[4 * g + 1] = v[cc * 8 + g].y;
                  a[4 * g + 2] = v[cc * 8 + g].z; a[4 * g + 3] = v[cc * 8 + g].w;
                }
                finish32(p, s_bias, r, hf * 64 + cc * 32, n0, a);
              }
            }
          }
          if (elected) p.counters[tn] = 0u;
        }
      }
    }
  }

  if (p.cluster && p.splits > 1) {
    // ---- split-K reduction through DISTRIBUTED SHARED MEMORY: the `splits` CTAs of a tile are one thread-block cluster.
    // Every CTA has parked its fp32 partial in its own shared memory; after the cluster barrier CTA `rank` sums rows
    // rank, rank + S, ... over all S partials (ld.shared::cluster), applies the epilogue and stores bf16.  No global
    // atomics, no fences, no arrival counter, no workspace: ~1 us instead of the ~4 us L2 round-trip chain.
    __syncwarp();
    cluster_sync_all();
    if (warp >= 2) {
      const int et = static_cast<int>(threadIdx.x) - 64;
      const int S = p.splits;
      const int rank = static_cast<int>(cluster_ctarank());
      const int tn = static_cast<int>(blockIdx.x) / S;
      const int n0 = tn * BN;
      const float* s_bias = reinterpret_cast<const float*>(smem + 512);
      const int rows_r = p.M > rank ? (p.M - rank + S - 1) / S : 0;
      const int items = rows_r * (BN / 8);
      const uint32_t p_base = smem_u32(smem_a);
      for (int i = et; i < items; i += 128) {
        const int lr = i % rows_r, g = i / rows_r;
        const int row = rank + lr * S;
        float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int sct = 0; sct < S; ++sct) {
          const float4 a0 = ld_dsmem_v4(p_base + static_cast<uint32_t>(((2 * g) * 128 + row) * 16), static_cast<uint32_t>(sct));
          const float4 a1 = ld_dsmem_v4(p_base + static_cast<uint32_t>(((2 * g + 1) * 128 + row) * 16), static_cast<uint32_t>(sct));
          v[0] += a0.x; v[1] += a0.y; v[2] += a0.z; v[3] += a0.w;
          v[4] += a1.x; v[5] += a1.y; v[6] += a1.z; v[7] += a1.w;
        }
        finish8(p, s_bias, row, g * 8, n0, v);
      }
    }
    __syncwarp();
    cluster_sync_all();   // nobody leaves while a peer still reads its partial
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int BN>
static int launch_smallm(const CUtensorMap& tmA, const CUtensorMap& tmB, const SmallMParams& p, int grid, int smem_bytes,
                         cudaStream_t s) {
  static PerDeviceOnce once;
  const int rc0 = once.run([] {
    return static_cast<int>(
        cudaFuncSetAttribute(gemm_smallm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmSmemLimit));
  });
  if (rc0 != OWL_OK) return rc0;
  LaunchOpts lo;
  lo.cluster_x = p.cluster ? p.splits : 1;
  cudaError_t e = launch_kx(lo, gemm_smallm_kernel<BN>, dim3(static_cast<unsigned>(grid)), dim3(kSmThreads),
                            static_cast<size_t>(smem_bytes), s, tmA, tmB, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return static_cast<int>(e);
  }
  return launch_status();
}

struct SmallMPlan {
  int bn, splits;
};

// One wave of ~equal chunks: time ~ waves x (fixed + bytes one CTA streams / per-SM bandwidth) + the reduce round trip.
static SmallMPlan plan_smallm(long long M, long long N, int kb_total, int sms) {
  SmallMPlan best{64, 1};
  double best_cost = 1e300;
  const int cand[2] = {128, 64};
  for (int i = 0; i < 2; ++i) {
    const int bn = cand[i];
    if (bn == 128 && N <= 64) continue;
    const long long tiles = (N + bn - 1) / bn;
    for (int sp = 1; sp <= 16 && sp <= kb_total; ++sp) {
      const int kbps = (kb_total + sp - 1) / sp;
      const int real_sp = (kb_total + kbps - 1) / kbps;
      if (real_sp != sp) continue;
      const long long ctas = tiles * sp;
      const long long waves = (ctas + sms - 1) / sms;
      const double bytes = static_cast<double>(kbps) * (bn * 128.0 + static_cast<double>(M) * 128.0);
      const double cost = static_cast<double>(waves) * (1500.0 + bytes / 45.0) + (sp > 1 ? 700.0 : 0.0);  // ns
      if (cost < best_cost) { best_cost = cost; best = SmallMPlan{bn, sp}; }
      if (ctas >= 2LL * sms) break;
    }
  }
  return best;
}

}  // namespace owl

using namespace owl;

extern "C" {

int64_t owl_gemm_smallm_workspace_bytes(int64_t N) {
  // fp32 [ceil(N/64)][128][64] partial-sum tiles (covers BN = 128 too) + one arrival counter per 64 columns
  const int64_t tiles = (N + 63) / 64;
  return tiles * 128 * 64 * 4 + tiles * 4;
}

int owl_gemm_smallm(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* W, int64_t ldw,
                    const owl_gemm_epilogue* epi, const owl_adaln_prologue* ln, void* workspace, int64_t workspace_bytes,
                    const void* next_w, int64_t next_bytes, owl_stream_t stream) {
  if (!A || !W || !epi || !epi->out0 || M <= 0 || N <= 0 || K <= 0) return OWL_ERR_BAD_ARG;
  if (M > 128) return OWL_ERR_UNSUPPORTED;
  if (N % 8 != 0 || K % 8 != 0) return OWL_ERR_UNSUPPORTED;
  if (lda % 8 || ldw % 8 || epi->ld0 % 8 || !aligned16(A) || !aligned16(W) || !aligned16(epi->out0)) return OWL_ERR_ALIGN;
  if (lda < K || ldw < K) return OWL_ERR_BAD_ARG;
  if (epi->bias && !aligned16(epi->bias)) return OWL_ERR_ALIGN;
  SmallMParams p{};
  p.M = static_cast<int>(M); p.N = static_cast<int>(N); p.K = static_cast<int>(K);
  p.mode = epi->mode;
  p.rows_per_frame = epi->rows_per_frame > 0 ? epi->rows_per_frame : 1;
  p.bias = epi->bias;
  p.out = static_cast<bf16*>(epi->out0);
  p.ldo = epi->ld0;
  if (ln != nullptr) {
    if (!ln->a || !ln->b || ln->rows_per_frame <= 0 || M % ln->rows_per_frame) return OWL_ERR_BAD_ARG;
    if (ln->ld_ab % 8 || !aligned16(ln->a) || !aligned16(ln->b)) return OWL_ERR_ALIGN;
    p.ln_x = static_cast<const bf16*>(A);
    p.ln_ldx = lda;
    p.ln_a = static_cast<const bf16*>(ln->a);
    p.ln_b = static_cast<const bf16*>(ln->b);
    p.ln_ldab = ln->ld_ab;
    p.ln_T = ln->rows_per_frame;
    p.ln_eps = ln->eps;
  }
  if (next_w != nullptr && next_bytes > 0 && aligned16(next_w)) {
    p.next_w = static_cast<const char*>(next_w);
    p.next_bytes = next_bytes;
  }
  switch (p.mode) {
    case OWL_EPI_STORE:
    case OWL_EPI_BIAS_SILU:   // single output here: silu(bf16(x W^T + b)); the pre-activation is not kept (no backward)
      break;
    case OWL_EPI_GATE_RESIDUAL:
      if (!epi->aux0 || !epi->aux1) return OWL_ERR_BAD_ARG;
      if (epi->lda0 % 8 || epi->lda1 % 8 || !aligned16(epi->aux0) || !aligned16(epi->aux1)) return OWL_ERR_ALIGN;
      if (M % p.rows_per_frame) return OWL_ERR_BAD_ARG;
      p.res = static_cast<const bf16*>(epi->aux0); p.ldres = epi->lda0;
      p.gate = static_cast<const bf16*>(epi->aux1); p.ldgate = epi->lda1;
      break;
    default: return OWL_ERR_UNSUPPORTED;
  }
  const int sms = gemm_sm_count();
  const int kb_total = static_cast<int>((K + kSmBK - 1) / kSmBK);
  SmallMPlan plan = plan_smallm(M, N, kb_total, sms);
  if (const char* e = getenv("OWL_SMALLM_BN")) {
    const int v = atoi(e);
    if (v == 64 || v == 128) plan.bn = v;
  }
  if (const char* e = getenv("OWL_SMALLM_SPLITS")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kb_total) plan.splits = v;
  }
  const int bn = plan.bn;
  p.tiles_n = static_cast<int>((N + bn - 1) / bn);
  p.kb_total = kb_total;
  p.kb_per_split = (kb_total + plan.splits - 1) / plan.splits;
  p.splits = (kb_total + p.kb_per_split - 1) / p.kb_per_split;
  // the splits of a tile as one thread-block cluster (<= 8 CTAs, one tile per CTA): reduction through distributed shared memory
  // (opt-in: OWL_SMALLM_CLUSTER=1.  Measured on B200, tools/smallm_bench.py: qkv 16.3 vs 12.0 us, out 13.4 vs 10.8, fc1 18.3 vs
  // 13.8, fc2 21.6 vs 14.8 against the L2-atomic path -- a cluster is only scheduled once 2..8 SMs of ONE GPC are free at
  // the same time, which gives back the overlap with the previous kernel's tail that programmatic dependent launch buys.)
  static const int no_cluster = [] { const char* e = getenv("OWL_SMALLM_CLUSTER"); return e && e[0] == '1' ? 0 : 1; }();
  p.cluster = (!no_cluster && p.splits > 1 && p.splits <= 8 && static_cast<long long>(p.tiles_n) * p.splits <= 2LL * sms) ? 1 : 0;
  if (p.splits > 1 && !p.cluster) {
    if (!workspace || !aligned16(workspace)) return OWL_ERR_BAD_ARG;
    const int64_t ws_floats = static_cast<int64_t>(p.tiles_n) * 128 * bn;
    if (workspace_bytes < ws_floats * 4 + static_cast<int64_t>(p.tiles_n) * 4) return OWL_ERR_BAD_ARG;
    p.ws = static_cast<float*>(workspace);
    p.counters = reinterpret_cast<unsigned*>(static_cast<float*>(workspace) + ws_floats);
  }
  const int stage_bytes = kSmABytes + bn * kSmBK * 2;
  int stages = (kSmSmemPerCta - 2048) / stage_bytes;
  if (stages > kSmMaxStages) stages = kSmMaxStages;
  if (stages > p.kb_per_split && static_cast<long long>(p.tiles_n) * p.splits <= sms) stages = p.kb_per_split;
  if (stages < 2) stages = 2;
  p.stages = stages;
  const int smem_bytes = 2048 + stages * stage_bytes;

  CUtensorMap tmA, tmB;
  int rc;
  if ((rc = make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(M), static_cast<uint64_t>(K), lda, 64, 128)) != OWL_OK) return rc;
  if ((rc = make_tmap_2d_bf16(&tmB, W, static_cast<uint64_t>(N), static_cast<uint64_t>(K), ldw, 64, bn)) != OWL_OK) return rc;
  const long long total = static_cast<long long>(p.tiles_n) * p.splits;
  const int grid = p.cluster ? static_cast<int>(total) : static_cast<int>(total < sms ? total : sms);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return bn == 128 ? launch_smallm<128>(tmA, tmB, p, grid, smem_bytes, s) : launch_smallm<64>(tmA, tmB, p, grid, smem_bytes, s);
}

// ------------------------------------------------------------------------------------------------
// owl_gemm_smallm_static: the same GEMM for callers that cannot allocate (no cudaMalloc under graph capture, in a tight
// serving loop, in a harness that forbids it).  The split-K scratch (fp32 partial tiles + arrival counters) is a file-scope
// __device__ array instead of a caller-provided workspace.  No counterpart in the
// reference.
//   * limit: N <= OWL_SMALLM_STATIC_MAX_N (8192): (8192 / 64) x (128 x 64 x 4 + 4) = 4 194 816 bytes, held by every GPU that
//     loads the library until the process exits;
//   * in host code `g_smallm_scratch` is only a stand-in for the symbol: its device address comes from
//     cudaGetSymbolAddress, once per device, with that device current;
//   * every call shares the array and it keeps the previous call's contents: the part this call uses is cleared on `stream`
//     first (the partial tiles are accumulated with red.add and the counters count arrivals: both must start at zero);
//   * two calls that overlap on the GPU (different streams, or graph branches) race on the array: serialise them.
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace owl {

constexpr int64_t kStaticMaxN = 8192;
constexpr int64_t kStaticTiles = kStaticMaxN / 64;
constexpr int64_t kStaticScratchBytes = kStaticTiles * 128 * 64 * 4 + kStaticTiles * 4;

static __device__ __align__(256) unsigned char g_smallm_scratch[kStaticScratchBytes];

// clears the part of the scratch a call uses (a kernel, not a memset node: the GEMM behind it is launched with programmatic
// stream serialisation, which wants a kernel in front of it)
__global__ void __launch_bounds__(256) smallm_scratch_clear_kernel(uint4* __restrict__ dst, long long n16) {
  pdl_prologue();   // the previous user of the array (an earlier call on this stream) has finished
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += stride)
    dst[i] = make_uint4(0u, 0u, 0u, 0u);
}

static void* g_smallm_scratch_ptr[64];          // per device: the array's DEVICE address
static std::atomic<uint64_t> g_smallm_scratch_have{0};

static int smallm_scratch_address(void** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= 64) return OWL_ERR_UNSUPPORTED;
  const uint64_t bit = 1ull << dev;
  if (!(g_smallm_scratch_have.load(std::memory_order_acquire) & bit)) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_smallm_scratch);   // the current device's copy of the array
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return static_cast<int>(e);
    }
    if (ptr == nullptr) return OWL_ERR_DRIVER;
    g_smallm_scratch_ptr[dev] = ptr;                    // two threads racing here store the same value
    g_smallm_scratch_have.fetch_or(bit, std::memory_order_release);
  }
  *out = g_smallm_scratch_ptr[dev];
  return OWL_OK;
}

}  // namespace owl

extern "C" {

int64_t owl_gemm_smallm_static_max_n(void) { return owl::kStaticMaxN; }

int owl_gemm_smallm_static(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* W, int64_t ldw,
                           const owl_gemm_epilogue* epi, owl_stream_t stream) {
  if (N > owl::kStaticMaxN) return OWL_ERR_TOO_LARGE;
  if (N <= 0) return OWL_ERR_BAD_ARG;
  void* scratch = nullptr;
  int rc = owl::smallm_scratch_address(&scratch);
  if (rc != OWL_OK) return rc;
  const int64_t need = owl_gemm_smallm_workspace_bytes(N);   // <= kStaticScratchBytes for N <= kStaticMaxN
  if (need > owl::kStaticScratchBytes) return OWL_ERR_TOO_LARGE;
  // the array keeps whatever the previous call (any shape, possibly an aborted one) left: start this call's part at zero
  const long long n16 = (need + 15) / 16;                    // the array is 256-byte aligned and a multiple of 16 bytes long
  long long blocks = (n16 + 255) / 256;
  if (blocks > 4LL * owl::sm_count()) blocks = 4LL * owl::sm_count();
  (void)owl::launch_k(owl::smallm_scratch_clear_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0,
                      static_cast<cudaStream_t>(stream), static_cast<uint4*>(scratch), n16);
  rc = owl::launch_status();
  if (rc != OWL_OK) return rc;
  return owl_gemm_smallm(M, N, K, A, lda, W, ldw, epi, nullptr, scratch, need, nullptr, 0, stream);
}

}  // extern "C"
