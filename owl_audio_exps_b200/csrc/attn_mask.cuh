// The reference's frame-block attention mask (owl_wms/nn/attn.py:43-59) as closed-form token ranges.
//
//   frame(t) = t / T;  fq = frame(i + q_offset);  fk = frame(j)
//   allowed(i, j) = (!causal || fk <= fq) && (W <= 0 || |fq - fk| < W) && (doc == NULL || doc[fq] == doc[fk])
//
// For one query row the causal and window terms select a contiguous key range [lo, hi):
//   lo = max(0, (fq - W + 1) * T)        (W > 0)       else 0
//   hi = (fq + 1) * T  (causal)   or   (fq + W) * T  (bidirectional, W > 0)   else Lk
// so a KV tile is skipped when it lies outside the union of its query rows' ranges, and inside a
// tile the per-element test is two compares (plus one doc compare when doc ids are present).
#pragma once
#include <stdint.h>

namespace owl {

struct MaskParams {
  int T;         // tokens per frame
  int W;         // window in frames, <= 0: unlimited
  int q_offset;  // tokens already in the KV prefix
  int causal;
  const int* doc;  // [B, doc_ld] frame-indexed doc ids or nullptr
  long long doc_ld;
  // Optional per-frame document extents (owl_attn_doc_extents; same [B, doc_ld] indexing), all three or none:
  //   doc_first[b][f] / doc_last[b][f] = first / last frame whose id equals doc[b][f]  (a bounding range for ANY id
  //   pattern); doc_breaks[b] = 0 when every id of batch b is one contiguous run (count of frames that re-open an earlier id), in which case the bounding range is
  //   exact and no per-element id compare is needed.  They let the kernels skip tiles that lie in other documents.
  const int* doc_first;
  const int* doc_last;
  const int* doc_breaks;
};

// Key-token range [lo, hi) that the rows of frames [f0, f1] (inclusive, frames of q + q_offset) may see under the
// document term alone: the union of their documents' extents.  Only with doc_first / doc_last present.
__device__ __forceinline__ void doc_token_range(const MaskParams& m, int b, int f0, int f1, int& lo, int& hi) {
  const int* fr = m.doc_first + b * m.doc_ld;
  const int* la = m.doc_last + b * m.doc_ld;
  int a = fr[f0], z = la[f0];
  for (int f = f0 + 1; f <= f1; ++f) {
    const int x = fr[f], y = la[f];
    a = x < a ? x : a;
    z = y > z ? y : z;
  }
  lo = a * m.T;
  hi = (z + 1) * m.T;
}

__host__ __device__ __forceinline__ void mask_row_range(const MaskParams& m, int i, int Lk, int& lo, int& hi) {
  const int fq = (i + m.q_offset) / m.T;
  lo = 0;
  hi = Lk;
  if (m.W > 0) {
    const long long l = (static_cast<long long>(fq) - m.W + 1) * m.T;
    if (l > lo) lo = static_cast<int>(l < Lk ? l : Lk);
  }
  if (m.causal) {
    const long long h = (static_cast<long long>(fq) + 1) * m.T;
    if (h < hi) hi = static_cast<int>(h);
  } else if (m.W > 0) {
    const long long h = (static_cast<long long>(fq) + m.W) * m.T;
    if (h < hi) hi = static_cast<int>(h);
  }
  if (hi < lo) hi = lo;
}

__device__ __forceinline__ bool mask_allowed(const MaskParams& m, int b, int i, int j) {
  const int fq = (i + m.q_offset) / m.T;
  const int fk = j / m.T;
  if (m.causal && fk > fq) return false;
  if (m.W > 0) {
    const int d = fq - fk;
    if (d >= m.W || -d >= m.W) return false;
  }
  if (m.doc != nullptr && m.doc[b * m.doc_ld + fq] != m.doc[b * m.doc_ld + fk]) return false;
  return true;
}

}  // namespace owl
