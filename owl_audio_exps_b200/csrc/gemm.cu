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

#include "host_common.h"
#include "sm100.cuh"

namespace owl {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int kGemmThreads = 256;

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, splits, kb_per_split, kb_total;
  int mode, accumulate, rows_per_frame, atomic_out;
  const float* bias;
  bf16* out0;
  long long ld0;
  bf16* out1;
  long long ld1;
  float* outf;
  long long ldf;
  const bf16* aux0;
  long long lda0;
  const bf16* aux1;
  long long lda1;
};

template <int BN>
struct GemmCfg {
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
  // ring + barriers (full, empty per stage; tmem full/empty x2) + tmem slot, + 1024 alignment slack
  static constexpr int kSmemBytes = kStages * kStageBytes + (2 * kStages + 4) * 8 + 16 + 1024;
};

__device__ __forceinline__ float silu_f(float z) { return z * sigmoidf_fast(z); }
__device__ __forceinline__ float dsilu_f(float z) {
  const float s = sigmoidf_fast(z);
  return s * (1.f + z * (1.f - s));
}

__device__ __forceinline__ void load8_bf16(const bf16* p, float (&f)[8]) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ void store8_bf16(bf16* p, const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = v;
}

// Epilogue for 8 consecutive columns [col, col+8) of output row `row`; acc holds the fp32 accumulators.
__device__ __forceinline__ void epilogue8(const GemmParams& p, long long row, int col, float (&acc)[8]) {
  if (p.bias != nullptr) {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col + 4));
    acc[0] += b0.x; acc[1] += b0.y; acc[2] += b0.z; acc[3] += b0.w;
    acc[4] += b1.x; acc[5] += b1.y; acc[6] += b1.z; acc[7] += b1.w;
  }
  switch (p.mode) {
    case OWL_EPI_STORE: {
      store8_bf16(p.out0 + row * p.ld0 + col, acc);
      break;
    }
    case OWL_EPI_BIAS_SILU: {
      float s[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[j] = round_bf16(acc[j]);  // the reference's Linear output is bf16 before F.silu (mlp.py:34-35)
        s[j] = silu_f(acc[j]);
      }
      store8_bf16(p.out0 + row * p.ld0 + col, acc);
      store8_bf16(p.out1 + row * p.ld1 + col, s);
      break;
    }
    case OWL_EPI_GATE_RESIDUAL: {
      float r[8], g[8];
      load8_bf16(p.aux0 + row * p.lda0 + col, r);
      load8_bf16(p.aux1 + (row / p.rows_per_frame) * p.lda1 + col, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = round_bf16(acc[j]);
      if (p.out1 != nullptr) store8_bf16(p.out1 + row * p.ld1 + col, acc);
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] += g[j] * acc[j];
      store8_bf16(p.out0 + row * p.ld0 + col, r);
      break;
    }
    case OWL_EPI_DSILU: {
      float z[8];
      load8_bf16(p.aux0 + row * p.lda0 + col, z);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] *= dsilu_f(z[j]);
      if (p.out0 != nullptr) {
        store8_bf16(p.out0 + row * p.ld0 + col, acc);
        break;
      }
    }  // fallthrough: fp32 destination
    case OWL_EPI_F32: {
      float* o = p.outf + row * p.ldf + col;
      if (p.atomic_out) {
        atomicAdd(reinterpret_cast<float4*>(o), make_float4(acc[0], acc[1], acc[2], acc[3]));
        atomicAdd(reinterpret_cast<float4*>(o + 4), make_float4(acc[4], acc[5], acc[6], acc[7]));
      } else {
        float4 v0 = make_float4(acc[0], acc[1], acc[2], acc[3]);
        float4 v1 = make_float4(acc[4], acc[5], acc[6], acc[7]);
        if (p.accumulate) {
          const float4 o0 = *reinterpret_cast<const float4*>(o);
          const float4 o1 = *reinterpret_cast<const float4*>(o + 4);
          v0.x += o0.x; v0.y += o0.y; v0.z += o0.z; v0.w += o0.w;
          v1.x += o1.x; v1.y += o1.y; v1.z += o1.z; v1.w += o1.w;
        }
        *reinterpret_cast<float4*>(o) = v0;
        *reinterpret_cast<float4*>(o + 4) = v1;
      }
      break;
    }
    default: break;
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tfull_bar = empty_bar + kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.tiles_m * p.tiles_n * p.splits;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int tm = t % p.tiles_m;
        const int tn = (t / p.tiles_m) % p.tiles_n;
        const int ks = t / (p.tiles_m * p.tiles_n);
        const int m0 = tm * BM, n0 = tn * BN;
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          uint8_t* sa = smem_a + stage * Cfg::kABytes;
          uint8_t* sb = smem_b + stage * Cfg::kBBytes;
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int pn = 0; pn < BM / 64; ++pn) tma_load_2d(sa + pn * 8192, &tmA, &full_bar[stage], m0 + pn * 64, k0);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if (B_MN) {
#pragma unroll
            for (int pn = 0; pn < BN / 64; ++pn) tma_load_2d(sb + pn * 8192, &tmB, &full_bar[stage], n0 + pn * 64, k0);
          } else {
            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
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
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, 8192, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, 8192, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
            umma_bf16_ss(tacc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // slot reusable once these MMAs have read it
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[as]);  // accumulator complete
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    int as = 0;
    uint32_t aphase = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int tm = t % p.tiles_m;
      const int tn = (t / p.tiles_m) % p.tiles_n;
      const int m0 = tm * BM, n0 = tn * BN;
      const long long row = m0 + q * 32 + lane;
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(as * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (row < p.M) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = n0 + c * 32 + g * 8;
            if (col < p.N) {
              float acc[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[j] = __uint_as_float(r[g * 8 + j]);
              epilogue8(p, row, col, acc);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
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

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool configured = false;  // idempotent attribute; a benign race at worst sets it twice
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_kernel<BN, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  gemm_bf16_kernel<BN, A_MN, B_MN><<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, p);
  return launch_status();
}

template <int BN>
static int launch_gemm_layout(int layout, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int grid,
                              cudaStream_t stream) {
  switch (layout) {
    case OWL_GEMM_NT: return launch_gemm<BN, false, false>(tmA, tmB, p, grid, stream);
    case OWL_GEMM_NN: return launch_gemm<BN, false, true>(tmA, tmB, p, grid, stream);
    case OWL_GEMM_TN: return launch_gemm<BN, true, true>(tmA, tmB, p, grid, stream);
    default: return OWL_ERR_BAD_ARG;
  }
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
  p.mode = epi->mode; p.accumulate = epi->accumulate; p.rows_per_frame = epi->rows_per_frame > 0 ? epi->rows_per_frame : 1;
  p.bias = epi->bias;
  p.out0 = static_cast<bf16*>(epi->out0); p.ld0 = epi->ld0;
  p.out1 = static_cast<bf16*>(epi->out1); p.ld1 = epi->ld1;
  p.outf = epi->outf; p.ldf = epi->ldf;
  p.aux0 = static_cast<const bf16*>(epi->aux0); p.lda0 = epi->lda0;
  p.aux1 = static_cast<const bf16*>(epi->aux1); p.lda1 = epi->lda1;

  // per-mode argument validation
  bool f32_out = false;
  switch (p.mode) {
    case OWL_EPI_STORE:
      if (!p.out0) return OWL_ERR_BAD_ARG;
      break;
    case OWL_EPI_BIAS_SILU:
      if (!p.out0 || !p.out1) return OWL_ERR_BAD_ARG;
      break;
    case OWL_EPI_GATE_RESIDUAL:
      if (!p.out0 || !p.aux0 || !p.aux1) return OWL_ERR_BAD_ARG;
      if (M % p.rows_per_frame) return OWL_ERR_BAD_ARG;
      break;
    case OWL_EPI_DSILU:
      if (!p.aux0 || (!p.out0 && !p.outf)) return OWL_ERR_BAD_ARG;
      f32_out = p.out0 == nullptr;
      break;
    case OWL_EPI_F32:
      if (!p.outf) return OWL_ERR_BAD_ARG;
      f32_out = true;
      break;
    default: return OWL_ERR_BAD_ARG;
  }
  if ((p.out0 && (p.ld0 % 8 || !aligned16(p.out0))) || (p.out1 && (p.ld1 % 8 || !aligned16(p.out1))) ||
      (p.outf && (p.ldf % 4 || !aligned16(p.outf))) || (p.aux0 && (p.lda0 % 8 || !aligned16(p.aux0))) ||
      (p.aux1 && (p.lda1 % 8 || !aligned16(p.aux1))) || (p.bias && !aligned16(p.bias)))
    return OWL_ERR_ALIGN;

  // tile shape: widest BN that still gives every SM a tile
  const int sms = sm_count();
  const int tiles_m = static_cast<int>((M + BM - 1) / BM);
  int bn = 256;
  if (N <= 64) bn = 64;
  else if (N <= 128) bn = 128;
  else {
    const long long t256 = static_cast<long long>(tiles_m) * ((N + 255) / 256);
    if (t256 < sms || N % 256 != 0) bn = 128;
  }
  const int tiles_n = static_cast<int>((N + bn - 1) / bn);
  const int kb_total = static_cast<int>((K + BK - 1) / BK);
  int splits = 1;
  if (p.mode == OWL_EPI_F32) {
    const int mn_tiles = tiles_m * tiles_n;
    if (mn_tiles < sms && kb_total >= 8) {
      splits = (sms + mn_tiles - 1) / mn_tiles;
      const int max_splits = kb_total / 4;  // >= 4 k-blocks (256 contraction steps) per split
      if (splits > max_splits) splits = max_splits;
      if (splits < 1) splits = 1;
    }
  }
  int kb_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per_split - 1) / kb_per_split;  // no empty splits
  p.tiles_m = tiles_m; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per_split; p.kb_total = kb_total;
  p.atomic_out = (f32_out && splits > 1) ? 1 : 0;

  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (p.atomic_out && !p.accumulate) {
    cudaError_t e = cudaMemset2DAsync(p.outf, static_cast<size_t>(p.ldf) * 4, 0, static_cast<size_t>(N) * 4,
                                      static_cast<size_t>(M), s);
    if (e != cudaSuccess) return static_cast<int>(e);
  }

  CUtensorMap tmA, tmB;
  int rc;
  if (a_mn) rc = make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(K), static_cast<uint64_t>(M), lda, 64, 64);
  else rc = make_tmap_2d_bf16(&tmA, A, static_cast<uint64_t>(M), static_cast<uint64_t>(K), lda, 64, BM);
  if (rc != OWL_OK) return rc;
  if (b_mn) rc = make_tmap_2d_bf16(&tmB, B, static_cast<uint64_t>(K), static_cast<uint64_t>(N), ldb, 64, 64);
  else rc = make_tmap_2d_bf16(&tmB, B, static_cast<uint64_t>(N), static_cast<uint64_t>(K), ldb, 64, bn);
  if (rc != OWL_OK) return rc;

  const long long total = static_cast<long long>(tiles_m) * tiles_n * splits;
  const int grid = static_cast<int>(total < sms ? total : sms);
  switch (bn) {
    case 64: return launch_gemm_layout<64>(layout, tmA, tmB, p, grid, s);
    case 128: return launch_gemm_layout<128>(layout, tmA, tmB, p, grid, s);
    default: return launch_gemm_layout<256>(layout, tmA, tmB, p, grid, s);
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
