// Shared pieces of the tcgen05 attention kernels (forward, dq, dkv).
//
// Tiles are 128 query rows x 128 keys.  Q/K/V/dO tiles arrive by TMA as [128 rows][DH] bf16 in
// 64-column panels of 16 KB ([128 rows][128 B], 128B-swizzled).  The same panel serves as
//   - a K-major operand (contraction over DH):   start + (k/4)*16384 + (k%4)*32, SBO 1024
//   - an MN-major operand (contraction over rows): start + k*2048, LBO 16384, SBO 1024
// (k = index of the 16-wide UMMA_K step), so no tile is ever transposed in memory.
// Probability / dS tiles are written by the softmax threads (thread = query row) into the same
// panel layout: [128 q rows][64 keys] x 2 panels.
#pragma once
#include "attn_mask.cuh"
#include "sm100.cuh"

namespace owl {

static constexpr int kTile = 128;
static constexpr int kPanelBytes = 128 * 128;  // [128 rows][64 bf16]

// Descriptor of the same tile `bytes` further on: the start address sits in the low 14 bits of the low word (in 16-byte
// units; shared memory ends far below 2^18 bytes, so the sum cannot carry into the next field).  One add on the low word:
// the unrolled k-loops of the MMA issuers derive every k-step's descriptor from ONE base per tile instead of re-encoding
// it (shift / mask / or / or), which is what the issuer warps spend their instructions on.
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) {
  return (d & 0xFFFFFFFF00000000ull) | static_cast<uint64_t>(static_cast<uint32_t>(d) + (bytes >> 4));
}
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_saddr, int k) {
  return desc_advance(make_smem_desc(tile_saddr, 16, 1024), static_cast<uint32_t>((k >> 2) * kPanelBytes + (k & 3) * 32));
}
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_saddr, int k) {
  return desc_advance(make_smem_desc(tile_saddr, kPanelBytes, 1024), static_cast<uint32_t>(k * 2048));
}

// byte offset of the 16-byte chunk holding columns [8*chunk16, 8*chunk16+8) of row `row` in a
// [128][128] bf16 tile stored as two swizzled panels
__device__ __forceinline__ uint32_t tile_chunk_offset(int row, int chunk16) {
  const int panel = chunk16 >> 3, c = chunk16 & 7;
  return panel * kPanelBytes + row * 128 + ((c ^ (row & 7)) << 4);
}

// Validity bits of keys [jc, jc+32) for a query row whose causal/window range is [lo, hi) and whose
// doc id is dq (doc == nullptr: no doc test).  Bit e <=> key jc+e allowed.
__device__ __forceinline__ uint32_t chunk_valid_bits(int jc, int lo, int hi, const int* __restrict__ doc_row, int dq,
                                                     int T) {
  int a = lo - jc, b = hi - jc;
  a = a < 0 ? 0 : a;
  b = b > 32 ? 32 : b;
  if (b <= a) return 0u;
  uint32_t bits = (b >= 32 ? 0xffffffffu : ((1u << b) - 1u)) & ~((1u << a) - 1u);  // a < 32 here
  if (doc_row != nullptr) {
    const int f0 = (jc + a) / T, f1 = (jc + b - 1) / T;
    for (int f = f0; f <= f1; ++f) {
      if (__ldg(doc_row + f) != dq) {
        int s = f * T - jc, e = s + T;
        s = s < 0 ? 0 : s;
        e = e > 32 ? 32 : e;
        const uint32_t fb = (e >= 32 ? 0xffffffffu : ((1u << e) - 1u)) & ~((1u << s) - 1u);
        bits &= ~fb;
      }
    }
  }
  return bits;
}

// KV tile range [t0, t1) needed by query rows [i0, i1] (inclusive), from the causal/window terms.
__device__ __forceinline__ void kv_tile_range(const MaskParams& m, int i0, int i1, int Lk, int& t0, int& t1) {
  int lo0, hi0, lo1, hi1;
  mask_row_range(m, i0, Lk, lo0, hi0);
  mask_row_range(m, i1, Lk, lo1, hi1);
  t0 = lo0 / kTile;
  t1 = (hi1 + kTile - 1) / kTile;
  if (t1 < t0) t1 = t0;
}

// The same with the document term: tiles that hold only other documents' frames are dropped as well (what the
// reference's BlockMask does at 128x128 granularity, attn.py:52-55,62).
__device__ __forceinline__ void kv_tile_range_doc(const MaskParams& m, int b, int i0, int i1, int Lk, int& t0, int& t1) {
  int lo0, hi0, lo1, hi1;
  mask_row_range(m, i0, Lk, lo0, hi0);
  mask_row_range(m, i1, Lk, lo1, hi1);
  int lo = lo0, hi = hi1;
  if (m.doc_first != nullptr) {
    int dlo, dhi;
    doc_token_range(m, b, (i0 + m.q_offset) / m.T, (i1 + m.q_offset) / m.T, dlo, dhi);
    lo = dlo > lo ? dlo : lo;
    hi = dhi < hi ? dhi : hi;
  }
  if (hi < lo) hi = lo;
  t0 = lo / kTile;
  t1 = (hi + kTile - 1) / kTile;
  if (t1 < t0) t1 = t0;
}

// One query row's key range [lo, hi) with its document extent folded in; need_doc_test = an id compare per key is
// still required inside the range (ids present and not known to be contiguous runs).
__device__ __forceinline__ void row_range_doc(const MaskParams& m, int b, int i, int Lk, int& lo, int& hi,
                                              bool& need_doc_test) {
  mask_row_range(m, i, Lk, lo, hi);
  need_doc_test = m.doc != nullptr;
  if (m.doc_first != nullptr) {
    const int fq = (i + m.q_offset) / m.T;
    const int dlo = m.doc_first[b * m.doc_ld + fq] * m.T, dhi = (m.doc_last[b * m.doc_ld + fq] + 1) * m.T;
    lo = dlo > lo ? dlo : lo;
    hi = dhi < hi ? dhi : hi;
    if (hi < lo) hi = lo;
    if (m.doc_breaks[b] == 0) need_doc_test = false;
  }
}

// Query tile range [t0, t1) that can see keys [j0, j1] (inclusive): causal / window terms, and - when the mask carries
// document extents and `b` >= 0 - only queries of the documents those keys belong to.
__device__ __forceinline__ void q_tile_range(const MaskParams& m, int j0, int j1, int Lq, int& t0, int& t1, int b = -1) {
  long long ilo = 0, ihi = Lq;
  const long long fk0 = j0 / m.T, fk1 = j1 / m.T;
  if (m.causal) ilo = fk0 * m.T - m.q_offset;
  if (m.W > 0) {
    const long long h = (fk1 + m.W) * m.T - m.q_offset;
    if (h < ihi) ihi = h;
    if (!m.causal) ilo = (fk0 - m.W + 1) * m.T - m.q_offset;
  }
  if (b >= 0 && m.doc_first != nullptr) {
    int dlo, dhi;  // key-frame extents are query-frame extents too (same frame axis, shifted by q_offset tokens)
    doc_token_range(m, b, static_cast<int>(fk0), static_cast<int>(fk1), dlo, dhi);
    const long long a = static_cast<long long>(dlo) - m.q_offset, z = static_cast<long long>(dhi) - m.q_offset;
    if (a > ilo) ilo = a;
    if (z < ihi) ihi = z;
  }
  if (ilo < 0) ilo = 0;
  if (ihi > Lq) ihi = Lq;
  if (ihi < ilo) ihi = ilo;
  t0 = static_cast<int>(ilo / kTile);
  t1 = static_cast<int>((ihi + kTile - 1) / kTile);
}

// One KEY's visible query range [ilo, ihi) (the transpose of row_range_doc: used by kernels whose threads own key rows),
// document extent folded in; need_doc_test as in row_range_doc.
__device__ __forceinline__ void key_q_range(const MaskParams& m, int b, int j, int Lq, int Lk, int& ilo_out, int& ihi_out,
                                            bool& need_doc_test) {
  long long ilo = 0, ihi = Lq;
  const long long fk = j / m.T;
  if (m.causal) ilo = fk * m.T - m.q_offset;
  if (m.W > 0) {
    const long long h = (fk + m.W) * m.T - m.q_offset;
    if (h < ihi) ihi = h;
    if (!m.causal) ilo = (fk - m.W + 1) * m.T - m.q_offset;
  }
  need_doc_test = m.doc != nullptr;
  if (m.doc_first != nullptr && j < Lk) {
    const long long a = static_cast<long long>(m.doc_first[b * m.doc_ld + fk]) * m.T - m.q_offset;
    const long long z = (static_cast<long long>(m.doc_last[b * m.doc_ld + fk]) + 1) * m.T - m.q_offset;
    if (a > ilo) ilo = a;
    if (z < ihi) ihi = z;
    if (m.doc_breaks[b] == 0) need_doc_test = false;
  }
  if (ilo < 0) ilo = 0;
  if (ihi > Lq) ihi = Lq;
  if (j >= Lk || ihi < ilo) ihi = ilo;
  ilo_out = static_cast<int>(ilo);
  ihi_out = static_cast<int>(ihi);
}

}  // namespace owl
