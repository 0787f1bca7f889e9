// Host-side helpers shared by the C-ABI launchers: error codes, TMA tensor-map encoding
// (through the driver entry point, so the library has no link-time dependency on libcuda),
// device properties cache.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/owl_dit.h"

namespace owl {

// 2-D bf16 row-major tensor [rows, cols] with leading dimension ld (elements); box = {box_cols, box_rows},
// 128-byte swizzle (box_cols * 2 bytes must be <= 128).
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_cols, uint32_t box_rows);

// 2-D fp32 row-major tensor [rows, cols]; box = {box_cols (<= 32), box_rows}; 128-byte swizzle.
int make_tmap_2d_f32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                     uint32_t box_cols, uint32_t box_rows);

// 3-D bf16 tensor [n2, n1, n0] (n0 contiguous) with element strides s1, s2; box {b0, b1, 1}; 128B swizzle.
int make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint64_t s1_elems,
                      uint64_t s2_elems, uint32_t b0, uint32_t b1);

// 4-D bf16 head-major tensor (b, h, row, col): dims {cols, rows, H, B}, row stride = cols (contiguous rows),
// head / batch strides in elements; box {box_cols, box_rows, 1, 1}; 128B swizzle; out-of-bounds rows read as 0.
int make_tmap_4d_bf16(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t H, uint64_t B,
                      uint64_t row_stride, uint64_t sh_elems, uint64_t sb_elems, uint32_t box_cols, uint32_t box_rows);

int sm_count();
int coresident_smem_reserve();  // bytes of shared memory the persistent GEMM leaves free per SM (owl_set_coresident_reserve)
int gemm_sm_count();  // sm_count() minus the SMs reserved for concurrent communication kernels (even for CTA pairs)

extern std::atomic<uint64_t> g_launches;

// Call right after a <<<>>> launch: counts it (owl_launch_count) and maps the launch error to the ABI code.
inline int launch_status(int n = 1) {
  g_launches.fetch_add(static_cast<uint64_t>(n), std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? OWL_OK : static_cast<int>(e);
}

// OWL_PDL=0 turns programmatic dependent launch off (plain stream serialisation).
bool pdl_enabled();

// Kernel launch used by every product kernel: cudaLaunchKernelEx with the programmatic-stream-serialisation
// attribute (see sm100.cuh pdl_wait) and, for CTA pairs, the cluster dimension.
struct LaunchOpts {
  bool pdl = true;     // false: the preceding stream operation is not a kernel (memset) -> plain serialisation
  int cluster_x = 1;   // 2: CTA pair
};

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kx(const LaunchOpts& o, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                             cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (o.pdl && pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (o.cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = static_cast<unsigned>(o.cluster_x);
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<Args&&>(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                            Args&&... args) {
  return launch_kx(LaunchOpts{}, kern, grid, block, smem, stream, static_cast<Args&&>(args)...);
}

// One-time, PER-DEVICE kernel configuration (cudaFuncSetAttribute is per device: a process that drives several GPUs --
// the per-GPU-stream rollout replicas -- must opt every device into > 48 KB of dynamic shared memory).
class PerDeviceOnce {
  std::atomic<uint64_t> done_{0};

 public:
  template <class F>
  int run(F&& f) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    const uint64_t bit = 1ull << (dev & 63);
    if (done_.load(std::memory_order_acquire) & bit) return OWL_OK;
    const int rc = f();   // idempotent attribute writes: a race between two threads at worst sets them twice
    if (rc != OWL_OK) return rc;
    done_.fetch_or(bit, std::memory_order_release);
    return OWL_OK;
  }
};

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace owl
