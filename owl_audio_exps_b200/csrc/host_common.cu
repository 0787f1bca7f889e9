// Host-side helpers shared by the C-ABI launchers (see host_common.h).
#include "host_common.h"
#include <cstdio>

#include <stdlib.h>

#include <atomic>
#include <mutex>

namespace owl {

std::atomic<uint64_t> g_launches{0};

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn g_encode_raw = nullptr;

// The driver call needs a current context.  A thread that has not made a runtime call yet has none -- autograd's backward
// thread when a GEMM is the very first op of the backward pass -- and the encode answers CUDA_ERROR_INVALID_CONTEXT:
// bind the device's primary context (cudaFree(0)) and encode again.
static CUresult encode_bound(CUtensorMap* m, CUtensorMapDataType dt, cuuint32_t rank, void* base, const cuuint64_t* dims,
                             const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* estr, CUtensorMapInterleave il,
                             CUtensorMapSwizzle sw, CUtensorMapL2promotion l2, CUtensorMapFloatOOBfill oob) {
  CUresult r = g_encode_raw(m, dt, rank, base, dims, strides, box, estr, il, sw, l2, oob);
  if (r == CUDA_ERROR_INVALID_CONTEXT || r == CUDA_ERROR_NOT_INITIALIZED) {
    (void)cudaFree(nullptr);
    r = g_encode_raw(m, dt, rank, base, dims, strides, box, estr, il, sw, l2, oob);
  }
  return r;
}

static encode_tiled_fn get_encode() {
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) g_encode_raw = reinterpret_cast<encode_tiled_fn>(p);
    (void)cudaGetLastError();
  });
  return g_encode_raw ? &encode_bound : nullptr;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_cols, uint32_t box_rows) {
  encode_tiled_fn enc = get_encode();
  if (!enc) return OWL_ERR_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld_elems * 2) % 16) return OWL_ERR_ALIGN;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    fprintf(stderr, "owl: cuTensorMapEncodeTiled(bf16 2d) -> %d: base %p dims (%llu,%llu) ld %llu box (%u,%u)\n", static_cast<int>(r),
            base, static_cast<unsigned long long>(cols), static_cast<unsigned long long>(rows),
            static_cast<unsigned long long>(ld_elems), box_cols, box_rows);
  return r == CUDA_SUCCESS ? OWL_OK : OWL_ERR_DRIVER;
}

int make_tmap_2d_f32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                     uint32_t box_cols, uint32_t box_rows) {
  encode_tiled_fn enc = get_encode();
  if (!enc) return OWL_ERR_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld_elems * 4) % 16) return OWL_ERR_ALIGN;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 4};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OWL_OK : OWL_ERR_DRIVER;
}

int make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint64_t s1_elems,
                      uint64_t s2_elems, uint32_t b0, uint32_t b1) {
  encode_tiled_fn enc = get_encode();
  if (!enc) return OWL_ERR_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (s1_elems * 2) % 16 || (s2_elems * 2) % 16) return OWL_ERR_ALIGN;
  cuuint64_t dims[3] = {n0, n1, n2};
  cuuint64_t strides[2] = {s1_elems * 2, s2_elems * 2};
  cuuint32_t box[3] = {b0, b1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OWL_OK : OWL_ERR_DRIVER;
}

int make_tmap_4d_bf16(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t H, uint64_t B,
                      uint64_t row_stride, uint64_t sh_elems, uint64_t sb_elems, uint32_t box_cols, uint32_t box_rows) {
  encode_tiled_fn enc = get_encode();
  if (!enc) return OWL_ERR_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (row_stride * 2) % 16 || (sh_elems * 2) % 16 || (sb_elems * 2) % 16)
    return OWL_ERR_ALIGN;
  // a size-1 dimension may carry any stride; the driver still wants it to be a non-zero multiple of 16 bytes
  if (sh_elems == 0) sh_elems = rows * row_stride;
  if (sb_elems == 0) sb_elems = H * sh_elems;
  cuuint64_t dims[4] = {cols, rows, H, B};
  cuuint64_t strides[3] = {row_stride * 2, sh_elems * 2, sb_elems * 2};
  cuuint32_t box[4] = {box_cols, box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OWL_OK : OWL_ERR_DRIVER;
}

int sm_count() {
  static std::atomic<int> cache[64];   // per device (zero-initialised)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  std::atomic<int>& slot = cache[dev & 63];
  int n = slot.load(std::memory_order_relaxed);
  if (n == 0) {
    int v = 0;
    n = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : 148;
    (void)cudaGetLastError();
    slot.store(n, std::memory_order_relaxed);
  }
  return n;
}

static std::atomic<int> g_smem_reserve{0};
int coresident_smem_reserve() { return g_smem_reserve.load(std::memory_order_relaxed); }

static std::atomic<int> g_reserved_sms{0};
int gemm_sm_count() {
  const int n = sm_count() - g_reserved_sms.load(std::memory_order_relaxed);
  return n < 2 ? 2 : n;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("OWL_PDL");
    on = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

}  // namespace owl

extern "C" {

int owl_abi_version(void) { return 1; }

const char* owl_error_string(int code) {
  switch (code) {
    case OWL_OK: return "ok";
    case OWL_ERR_BAD_ARG: return "bad argument (null pointer, non-positive or inconsistent size)";
    case OWL_ERR_ALIGN: return "pointer or leading dimension not 16-byte aligned";
    case OWL_ERR_UNSUPPORTED: return "shape not supported by the sm_100a kernels";
    case OWL_ERR_DRIVER: return "CUDA driver entry point / tensor-map encoding failed";
    case OWL_ERR_TOO_LARGE: return "size beyond the fixed capacity of a no-allocation entry point (static device scratch)";
    default: break;
  }
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "unknown error";
}

uint64_t owl_launch_count(void) { return owl::g_launches.load(); }

int owl_set_coresident_reserve(int bytes) {
  if (bytes < 0 || bytes > 65536) return OWL_ERR_BAD_ARG;
  owl::g_smem_reserve.store(bytes);
  return OWL_OK;
}

int owl_set_reserved_sms(int n) {
  if (n < 0 || n > 64) return OWL_ERR_BAD_ARG;
  owl::g_reserved_sms.store(n);
  return OWL_OK;
}

}  // extern "C"
