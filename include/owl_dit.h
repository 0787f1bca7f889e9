/*
 * owl_dit.h -- C ABI of libowl_dit.so: hand-written sm_100a (B200) CUDA kernels for the owl_wms
 * DiT-block hot path (block-causal attention with QK-RMSNorm + RoPE + KV cache, adaLN
 * modulate / gate / residual, MLP), forward and backward.
 *
 * The reference (shahbuland/owl-audio-exps) has NO native / FFI interface for this path: every op
 * below is issued through PyTorch from `owl_wms/nn/*.py`.  Each entry point therefore cites the
 * reference Python lines whose arithmetic it replaces.  Paths are relative to the reference root.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless stated otherwise.
 *   - the caller owns every buffer (inputs, outputs, workspaces).  Nothing is allocated or freed
 *     by the library; pointers are borrowed for the duration of the call.
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*); no call
 *     synchronises or reads device memory on the host => CUDA-graph-capture safe.
 *   - return value: OWL_OK (0), a negative OWL_ERR_* for rejected arguments, or a positive
 *     cudaError_t from the launch.  Never throws, never aborts.
 *   - "bf16" = __nv_bfloat16 storage; all accumulation is fp32.
 *   - stateless and re-entrant (only a read-only cache of the driver entry point / SM count).
 */
#ifndef OWL_DIT_H_
#define OWL_DIT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OWL_OK 0
#define OWL_ERR_BAD_ARG (-1)     /* null pointer / non-positive size / inconsistent sizes      */
#define OWL_ERR_ALIGN (-2)       /* pointer or leading dimension violates 16-byte alignment    */
#define OWL_ERR_UNSUPPORTED (-3) /* shape outside what the kernels implement (e.g. head dim)   */
#define OWL_ERR_DRIVER (-4)      /* CUDA driver entry point / tensor-map encoding failed       */
#define OWL_ERR_TOO_LARGE (-5)   /* size beyond the fixed capacity of a no-allocation entry point */

typedef void* owl_stream_t; /* cudaStream_t */

/* ABI version (bumped on any signature change) and human-readable error text. */
int owl_abi_version(void);
const char* owl_error_string(int code);
/* Number of kernels this library has launched from this process (monotonic; for bench.py's
 * `gpu_launches`).  Host-side counter only. */
uint64_t owl_launch_count(void);
/* SMs the persistent GEMM leaves free (0..64, default 0).  The data-parallel engine sets this to the number of
 * NCCL CTAs when gradients are exchanged concurrently with backward (DDP overlap, rft_trainer.py:96): a persistent
 * grid sized to ALL SMs would queue one CTA behind the communication kernel and stretch that GEMM to two waves. */
int owl_set_reserved_sms(int n);
/* Bytes of shared memory per SM the persistent GEMM leaves unallocated from now on (host-side launch state, like
 * owl_set_reserved_sms), so that a small CTA of a concurrently running kernel - the shard reduction of the copy-engine
 * gradient exchange - can become resident next to a GEMM CTA instead of waiting for a kernel boundary.  0 = none. */
int owl_set_coresident_reserve(int bytes);
/* Debug aid: enqueue a one-thread kernel that stores the GPU's %globaltimer (ns) to *dst.  Used by
 * tools/step_trace.py to time phases INSIDE a replayed CUDA graph; never called on the product path. */
int owl_debug_timestamp(uint64_t* dst, owl_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * GEMM  (tcgen05.mma, TMEM accumulators, TMA-fed, persistent, warp-specialised)
 *
 * Replaces every nn.Linear on the path and its autograd:
 *   owl_wms/nn/attn.py:72-73,82,113 (qkv, out), owl_wms/nn/mlp.py:10-11,34-36 (fc1, fc2),
 *   owl_wms/nn/modulation.py:10,19,31,40 (adaLN / gate projections of silu(cond)).
 *
 * C[m,n] = sum_k opA(m,k) * opB(n,k), bf16 operands, fp32 accumulation, followed by a fused
 * epilogue.  `layout` picks the operand storage:
 *   OWL_GEMM_NT : A is [M,K] (ld = lda), B is [N,K]      forward   y  = x W^T
 *   OWL_GEMM_NN : A is [M,K],            B is [K,N]      dgrad     dx = dy W
 *   OWL_GEMM_TN : A is [K,M],            B is [K,N]      wgrad     dW = dy^T x
 * Leading dimensions are in elements and must be multiples of 8; base pointers 16-byte aligned.
 * ---------------------------------------------------------------------------------------------- */
enum { OWL_GEMM_NT = 0, OWL_GEMM_NN = 1, OWL_GEMM_TN = 2 };

enum {
  /* out0[m,n] (bf16) = acc + bias[n]                                                             */
  OWL_EPI_STORE = 0,
  /* z = bf16(acc + bias); out0 = z; out1 = bf16(silu(z))            (mlp.py:34-35)               */
  OWL_EPI_BIAS_SILU = 1,
  /* y = bf16(acc + bias); out1 = y (optional);                                                   */
  /* out0 = bf16(aux0[m,n] + aux1[m / rows_per_frame, n] * y)        (modulation.py:41-43 +       */
  /*                                                                  attn.py:135,141 residual)   */
  OWL_EPI_GATE_RESIDUAL = 2,
  /* out0 = bf16(acc * silu'(aux0[m,n]))   (bf16 destination only)                                 */
  /*                                   (backward of F.silu, mlp.py:35 / modulation.py:17,38)      */
  OWL_EPI_DSILU = 3,
  /* outf[m,n] (fp32) = acc (+ outf[m,n] if accumulate)              (weight gradients)           */
  OWL_EPI_F32 = 4
};

typedef struct owl_gemm_epilogue {
  int32_t mode;           /* OWL_EPI_*                                                            */
  int32_t accumulate;     /* fp32 outputs: add to the existing contents                           */
  int32_t rows_per_frame; /* OWL_EPI_GATE_RESIDUAL: tokens per frame T (gate row = m / T)         */
  int32_t reserved;
  const float* bias;      /* [N] fp32 or NULL                                                     */
  void* out0;             /* bf16 [M, ld0]                                                        */
  int64_t ld0;
  void* out1;             /* bf16 [M, ld1] or NULL                                                */
  int64_t ld1;
  float* outf;            /* fp32 [M, ldf] or NULL                                                */
  int64_t ldf;
  const void* aux0;       /* bf16 [M, lda0]: residual (GATE_RESIDUAL) or pre-activation (DSILU)   */
  int64_t lda0;
  const void* aux1;       /* bf16 [M / rows_per_frame, lda1]: gate                                */
  int64_t lda1;
} owl_gemm_epilogue;

int owl_gemm_bf16(int layout, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* B,
                  int64_t ldb, const owl_gemm_epilogue* epi, owl_stream_t stream);

/* Weight-streaming GEMM of the decode regime: out0[M <= 128, N] = epi(A[M, K] . W[N, K]^T + bias), bf16.
 * Replaces every nn.Linear of one KV-cached model evaluation of the causal rollout (attn.py:82,113, mlp.py:34-36 at
 * 1 frame = 65 tokens or 1 token; owl_wms/sampling/av_caching_v2.py:89-140, inference/causvid_pipeline.py:112-163):
 * the problem is split over N AND K into about one chunk per SM so all SMs stream weights at once; split-K partial
 * sums meet in `workspace` (fp32 L2 atomics) and the last-arriving chunk applies the epilogue.  Modes: OWL_EPI_STORE,
 * OWL_EPI_BIAS_SILU (ONE output: out0 = silu(bf16(acc + bias)), the pre-activation is not kept), OWL_EPI_GATE_RESIDUAL.
 * `workspace`: owl_gemm_smallm_workspace_bytes(N) bytes, ZERO before the first call; every call leaves it zero again.
 * Calls sharing a workspace must be ordered on one stream.
 * `next_w` / `next_bytes` (optional, NULL / 0): the weight matrix the NEXT Linear of the schedule will read; its bytes are
 * requested into L2 (cp.async.bulk.prefetch.L2) while this call reduces, so HBM keeps streaming across kernel boundaries.
 * `ln` (optional, NULL): adaLN as the A-operand transform (modulation.py:20-25; north_star: adaLN scale/shift fused into
 * the GEMM prologue): A is the RAW residual stream and the kernel multiplies rms_norm(A) * (1 + a[m / T]) + b[m / T], built
 * tile by tile in shared memory (the arithmetic of owl_adaln_modulate_fwd; the row statistic is summed per thread). */
typedef struct owl_adaln_prologue {
  const void* a;          /* bf16 [M / rows_per_frame, ld_ab]: scale                               */
  const void* b;          /* bf16 [M / rows_per_frame, ld_ab]: shift                               */
  int64_t ld_ab;
  int32_t rows_per_frame; /* tokens per frame T                                                    */
  float eps;              /* F.rms_norm epsilon                                                    */
} owl_adaln_prologue;

int64_t owl_gemm_smallm_workspace_bytes(int64_t N);
int owl_gemm_smallm(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* W, int64_t ldw,
                    const owl_gemm_epilogue* epi, const owl_adaln_prologue* ln, void* workspace, int64_t workspace_bytes,
                    const void* next_w, int64_t next_bytes, owl_stream_t stream);

/* Plain CUDA-core GEMM with the same contract (OWL_EPI_STORE / OWL_EPI_F32 only, no bias).
 * Verification aid for tests: an independent on-device answer at sizes where the CPU oracle is
 * too slow.  Never called by the product path. */
int owl_gemm_bf16_check(int layout, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* B,
                        int64_t ldb, float* C, int64_t ldc, owl_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Memory-bound kernels: coalesced, 128-bit vectorised, warp-shuffle reductions.
 * ---------------------------------------------------------------------------------------------- */

/* dst[i] = bf16(src[i])  -- fp32 master weights -> bf16 operands (what autocast does per forward,
 * rft_trainer.py:146).  n elements, n % 8 == 0 not required. */
int owl_cast_f32_bf16(const float* src, void* dst, int64_t n, owl_stream_t stream);
/* `count` casts in one launch per 8 tensors (every ns[i] a multiple of 8): a block's weight matrices staged together. */
int owl_cast_f32_bf16_multi(const float* const* srcs, void* const* dsts, const int64_t* ns, int count,
                            owl_stream_t stream);

/* y = silu(x), bf16, n elements                                    (modulation.py:17,38)         */
int owl_silu_bf16(const void* x, void* y, int64_t n, owl_stream_t stream);

/* SinCosEmbed (embeddings.py:30-72): out[r] = [sin(t[r]*mult*freqs[i]) | cos(...)], i < half; t fp32 [n_rows],
 * freqs fp32 [half] (the caller builds exp(-i*log(theta)/(half-1)) as the reference does), out bf16 [n_rows, 2*half]. */
int owl_sincos_embed(const float* t, const float* freqs, void* out, int64_t n_rows, int half, float mult,
                     owl_stream_t stream);

/* dz = dy * silu'(z), bf16, n elements: autograd of F.silu where it is not fused into a GEMM
 * epilogue (attn.py:274 FinalLayer.act; standalone MLP)                                           */
int owl_dsilu_mul_bf16(const void* dy, const void* z, void* dz, int64_t n, owl_stream_t stream);

/* adaLN modulate forward                                           (modulation.py:20-25,
 *                                                                   normalization.py:10-11)
 *   h[m,:] = rms_norm(x[m,:]) * (1 + a[m / T,:]) + b[m / T,:]
 * x,h: bf16 [M,d]; a,b: bf16 rows of length d with row stride ld_ab (they are column slices of
 * the fused cond projection); rstd: fp32 [M] out (saved for backward; may be NULL).
 * eps: value added to mean(x^2) (see DESIGN.md on F.rms_norm's default).  d % 8 == 0.           */
int owl_adaln_modulate_fwd(const void* x, const void* a, const void* b, int64_t ld_ab, void* h, float* rstd,
                           int64_t M, int64_t d, int64_t T, float eps, owl_stream_t stream);

/* Gated residual + adaLN modulate in one pass                       (attn.py:135-138, modulation.py:20-25,41-43)
 *   xsum[m,:] = x[m,:] + gate[m / T,:] * y[m,:]          (the residual of the attention branch)
 *   h[m,:]    = rms_norm(xsum[m,:]) * (1 + a[m / T,:]) + b[m / T,:]
 * Same traffic as doing the residual in the producing GEMM's epilogue (x, y read once; xsum, h written
 * once) but keeps that GEMM on the wide-tile store-only epilogue.                                   */
int owl_residual_adaln_fwd(const void* x, const void* y, const void* gate, int64_t ld_g, const void* a, const void* b,
                           int64_t ld_ab, void* xsum, void* h, float* rstd, int64_t M, int64_t d, int64_t T, float eps,
                           owl_stream_t stream);

/* adaLN modulate backward.  Given dh [M,d]:
 *   dx[m,:]  = rms_norm_bwd(dh * (1 + a)) (+ dres[m,:] if dres != NULL)   (residual path grad)
 *   da[f,:]  = sum_{m in frame f} dh[m,:] * xhat[m,:]
 *   db[f,:]  = sum_{m in frame f} dh[m,:]
 * da, db: bf16 rows with row stride ld_dab (column slices of the cond-projection gradient).      */
int owl_adaln_modulate_bwd(const void* dh, const void* x, const void* a, int64_t ld_ab, const float* rstd,
                           const void* dres, void* dx, void* da, void* db, int64_t ld_dab, int64_t M, int64_t d,
                           int64_t T, owl_stream_t stream);

/* gate forward, unfused (standalone Gate module / cond_gate; the block path fuses this into the
 * GEMM epilogue OWL_EPI_GATE_RESIDUAL)                               (modulation.py:41-43,57-63)
 *   out[m,:] = c[m / T,:] * x[m,:] (+ res[m,:] if res != NULL)                                   */
int owl_gate_fwd(const void* x, const void* c, int64_t ld_c, const void* res, void* out, int64_t M, int64_t d,
                 int64_t T, owl_stream_t stream);

/* gate + residual backward                                          (modulation.py:41-43)
 *   x_out = x_in + c[m / T] * y    =>   dy[m,:] = dx[m,:] * c[m / T,:]
 *                                       dc[f,:] = sum_{m in frame f} dx[m,:] * y[m,:]
 *   dbias[n] (+)= sum_m dy[m,n]   (fp32, optional: bias gradient of the Linear that produced y)  */
int owl_gate_bwd(const void* dx, const void* y, const void* c, int64_t ld_c, void* dy, void* dc, int64_t ld_dc,
                 float* dbias, int dbias_accumulate, int64_t M, int64_t d, int64_t T, owl_stream_t stream);

/* adaLN backward fused with the gate backward of the branch that produced the normalised tensor, one pass
 * (block structure attn.py:130-143: x_in = x_prev + c * y, h = rms_norm(x_in) * (1 + a) + b):
 *   dx = adaLN-backward(dh, x_in, a, rstd) + dres;  dy = dx * c[frame];  dc[frame] = sum dx * y;
 *   dbias (+)= colsum(dy)  (bias gradient of the Linear that produced y; NULL: not wanted).
 * Same argument meaning as owl_adaln_modulate_bwd followed by owl_gate_bwd(dx, ...); autograd of
 * modulation.py:20-25 and 41-43.                                                                   */
int owl_adaln_gate_bwd(const void* dh, const void* x, const void* a, int64_t ld_ab, const float* rstd,
                       const void* dres, void* dx, void* da, void* db, int64_t ld_dab, const void* y, const void* c,
                       int64_t ld_c, void* dy, void* dc, int64_t ld_dc, float* dbias, int dbias_accumulate, int64_t M,
                       int64_t d, int64_t T, owl_stream_t stream);

/* out[n] (+)= sum_m x[m,n]; x bf16 [M, ld], out fp32 [N]           (bias gradients)              */
int owl_colsum_bf16(const void* x, int64_t ld, float* out, int accumulate, int64_t M, int64_t N,
                    owl_stream_t stream);

/* Up to 4 column sums in one launch: outs[i][n] += sum_m xs[i][m, n] (bf16 [Ms[i], Ns[i]], row stride lds[i]).
 * The four bias gradients of one DiT block (autograd of the nn.Linear biases, attn.py:72-73, mlp.py:10-11);
 * outputs are added to, the caller zeroes them.                                                        */
int owl_colsum_bf16_multi(const void* const* xs, const int64_t* lds, const int64_t* Ms, const int64_t* Ns,
                          float* const* outs, int count, owl_stream_t stream);

/* QK-RMSNorm + RoPE + head split                                    (attn.py:83-89, rope.py:43-51)
 *   qkv: bf16 [B, L, 3, H, dh] (the qkv Linear output)
 *   q   -> bf16, element (b,h,l,:) at q + b*q_sb + h*q_sh + l*dh   (same for k, v with their strides;
 *          k/v strides let the kernel write straight into a preallocated KV cache at its tail)
 *   q,k = rope(rms_norm(.)) with cos/sin fp32 [*, dh/2] rows (pos_offset + l); v copied.
 *   RoPE output layout is [even-rotated | odd-rotated] halves, as rope.py:48-51.
 *   rstd_q, rstd_k: fp32 [B,H,L] out (NULL allowed).  dh in {64, 128}.                           */
int owl_qk_norm_rope_fwd(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t pos_offset, void* q,
                         int64_t q_sb, int64_t q_sh, void* k, int64_t k_sb, int64_t k_sh, void* v, int64_t v_sb,
                         int64_t v_sh, float* rstd_q, float* rstd_k, int64_t B, int64_t L, int64_t H, int64_t dh,
                         float eps, owl_stream_t stream);

/* Backward of the above: (dq, dk, dv) in [B,H,L,dh] (given strides) + saved rotated q,k and rstd
 * -> dqkv bf16 [B, L, 3, H, dh].                                                                  */
int owl_qk_norm_rope_bwd(const void* dq, const void* dk, const void* dv, int64_t g_sb, int64_t g_sh, const void* q,
                         int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                         const float* rstd_q, const float* rstd_k, const float* cos_tab, const float* sin_tab,
                         int64_t pos_offset, void* dqkv, int64_t B, int64_t L, int64_t H, int64_t dh,
                         owl_stream_t stream);

/* Bare RoPE on contiguous rows x [R, L, dh] (R = B*H): RoPE.forward on its own     (rope.py:43-51)
 *   inverse == 0: pairs (2i, 2i+1) -> [y_even | y_odd] halves;  inverse != 0: the transposed
 *   rotation from halves back to pairs (the autograd of the former).  dh % 8 == 0.                */
int owl_rope_apply(const void* x, const float* cos_tab, const float* sin_tab, int64_t pos_offset, void* y, int64_t R,
                   int64_t L, int64_t dh, int inverse, owl_stream_t stream);

/* KV-cache append                                                   (attn.py:92-99, kv_cache.py:50-58)
 * Copies new rows src [B,H,Ln,dh] (strides src_sb, src_sh) into cache [B,H,cap,dh] at row `pos`.
 * The reference re-concatenates the whole cache (O(cache) bytes per call); this moves only the
 * new rows.                                                                                       */
int owl_kv_append(const void* src, int64_t src_sb, int64_t src_sh, void* cache, int64_t cache_sb, int64_t cache_sh,
                  int64_t pos, int64_t B, int64_t H, int64_t Ln, int64_t dh, owl_stream_t stream);

/* Weight-less LayerNorm over the last dim (normalization.py:6-7), bf16 [M,d].                     */
int owl_layer_norm_fwd(const void* x, void* y, int64_t M, int64_t d, float eps, owl_stream_t stream);
/* Its autograd: dx = rstd * (dy - mean(dy) - xhat * mean(dy * xhat)), statistics recomputed from x
 * (gamerft_audio.py:86 puts layer_norm on the training path of the video head).                   */
int owl_layer_norm_bwd(const void* dy, const void* x, void* dx, int64_t M, int64_t d, float eps, owl_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Attention                                                          (attn.py:24-62,102-111)
 *
 * softmax(q k^T / sqrt(dh) + mask) v with the reference's frame-block mask evaluated in-kernel:
 *   frame(i) = i / tokens_per_frame
 *   allowed(i,j) = (!causal || frame(j) <= frame(i + q_offset))
 *                  && |frame(i + q_offset) - frame(j)| < window_frames
 *                  && (doc_id == NULL || doc_id[b, frame(i + q_offset)] == doc_id[b, frame(j)])
 * window_frames <= 0 means "no window".  KV tiles that are fully masked are skipped.
 *   q: bf16 (b,h,i,:) at q + b*q_sb + h*q_sh + i*dh;  k, v likewise with Lk rows
 *   o: bf16 (b,i,h,:) at o + b*o_sb + i*o_sl + h*dh   (i.e. [B,L,H*dh] ready for the out Linear,
 *      replacing attn.py:111's permute+contiguous copy)
 *   lse: fp32 [B,H,Lq] natural-log-sum-exp of the scaled scores (saved for backward; NULL ok)
 *   doc_id: int32 [B, doc_ld] frame-indexed, or NULL.
 * dh in {64, 128}.
 * ---------------------------------------------------------------------------------------------- */
typedef struct owl_attn_mask {
  int32_t tokens_per_frame;
  int32_t window_frames; /* <= 0: unlimited */
  int32_t q_offset;      /* tokens already in the KV prefix (kv_cache.length_at(0))               */
  int32_t causal;        /* 0/1; decode mode (attn.py:106-107, block_mask None) = causal 0, window 0 */
  const int32_t* doc_id; /* [B, doc_ld] or NULL                                                   */
  int64_t doc_ld;
  /* Optional (all three or none; same [B, doc_ld] indexing, filled by owl_attn_doc_extents): per-frame document
   * extents that let the kernels skip KV / query tiles lying entirely in other documents, the way the reference's
   * BlockMask does for packed sequences (attn.py:52-55,62; data/latent_seq_packing.py:94-126).                  */
  const int32_t* doc_first;  /* first frame carrying the same id as frame f                        */
  const int32_t* doc_last;   /* last  frame carrying the same id as frame f                        */
  const int32_t* doc_breaks; /* [B] frames that re-open an id seen earlier (0: ids are contiguous runs, the
                                extents are exact and no per-key id compare is needed)              */
} owl_attn_mask;

/* doc_id int32 [B, doc_ld] (F valid frames per row) -> first/last int32 [B, doc_ld], breaks int32 [B].
 * Replaces the document term of create_block_mask (attn.py:52-55,62) at tile granularity.          */
int owl_attn_doc_extents(const int32_t* doc_id, int64_t doc_ld, int64_t B, int64_t F, int32_t* first, int32_t* last,
                         int32_t* breaks, owl_stream_t stream);

int owl_attn_fwd(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                 const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                 int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                 owl_stream_t stream);

/* Backward.  dout/o: bf16 in the o layout above.  dq/dk/dv: bf16 [B,H,L*,dh] with strides g?_sb, g?_sh.
 * delta: fp32 [B,H,Lq] workspace (rowsum(dout * o)).  dq_acc: fp32 [B,H,Lq,dh] workspace.        */
int owl_attn_bwd(const void* dout, const void* o, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb,
                 int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb,
                 int64_t v_sh, const float* lse, float* delta, float* dq_acc, void* dq, int64_t dq_sb, int64_t dq_sh,
                 void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh, int64_t B,
                 int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask, owl_stream_t stream);

/* CUDA-core versions of the two calls above with identical contracts (one warp per query row /
 * per key row).  Verification aid for tests, as owl_gemm_bf16_check.                              */
int owl_attn_fwd_check(const void* q, int64_t q_sb, int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh,
                       const void* v, int64_t v_sb, int64_t v_sh, void* o, int64_t o_sb, int64_t o_sl, float* lse,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       owl_stream_t stream);
int owl_attn_bwd_check(const void* dout, const void* o, int64_t o_sb, int64_t o_sl, const void* q, int64_t q_sb,
                       int64_t q_sh, const void* k, int64_t k_sb, int64_t k_sh, const void* v, int64_t v_sb,
                       int64_t v_sh, const float* lse, float* delta, float* dq_acc, void* dq, int64_t dq_sb,
                       int64_t dq_sh, void* dk, int64_t dk_sb, int64_t dk_sh, void* dv, int64_t dv_sb, int64_t dv_sh,
                       int64_t B, int64_t H, int64_t Lq, int64_t Lk, int64_t dh, const owl_attn_mask* mask,
                       owl_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Decode regime of the KV-cached causal rollout (attn.py:92-109 with block_mask None; av_caching_v2.py:89-140,
 * audio_caching.py:120-150, causvid_pipeline.py:112-163): <= 128 new query rows against a RING cache whose state
 * lives in device memory, so one captured CUDA graph serves every frame.
 *   state (int32[4], device): [0] rows cached, [1] ring start slot, [2] RoPE position of the next new token.
 *   k_cache / v_cache: bf16 [B, H, cap, dh] (strides c_sb, c_sh in elements; rows contiguous).
 * owl_qk_norm_rope_decode: QK-RMSNorm + RoPE + head split of qkv [B, L, 3*H*dh] (attn.py:83-89, rope.py:43-51);
 *   q -> [B, H, L, dh]; K (post-norm, post-RoPE) and V rows -> ring slots (state[1] + state[0] + l) % cap.  The rows
 *   are NOT counted in state[0] until owl_decode_state_advance commits them (KVCache.update, kv_cache.py:50-58).
 * owl_attn_decode: o[B, Lq, H*dh] = softmax(q K^T / sqrt(dh)) V over the cached rows plus the Lq rows just written
 *   (no mask); vis_rows > 0: only the last vis_rows rows of [cache, new] (local layers, attn.py:102-105).  The KV
 *   tiles of a head are split over n CTAs (B*H*n ~ #SMs); the last CTA of a head merges the partials.
 *   `workspace`: owl_attn_decode_workspace_bytes() bytes, ZERO before the first call; left zero-countered by each call.
 * owl_decode_state_advance: state after committing commit_rows and ejecting the drop_rows oldest rows
 *   (KVCache.truncate, kv_cache.py:60-75). */
int64_t owl_attn_decode_workspace_bytes(int64_t B, int64_t H, int64_t cap, int64_t dh);
int owl_attn_decode(const void* q, int64_t q_sb, int64_t q_sh, const void* k_cache, const void* v_cache, int64_t c_sb,
                    int64_t c_sh, int64_t cap, const int32_t* state, int64_t vis_rows, void* o, int64_t o_sb, int64_t o_sl,
                    int64_t B, int64_t H, int64_t Lq, int64_t dh, void* workspace, int64_t workspace_bytes,
                    owl_stream_t stream);
/* owl_attn_decode_fused = owl_qk_norm_rope_decode + owl_attn_decode in ONE launch (north_star: RoPE fused into the Q/K
 * load): every CTA builds its Q tile from qkv [B, Lq, 3*H*dh] with QK-RMSNorm + RoPE applied on the way into shared
 * memory; the last KV split of each head builds the K / V tile of the new rows the same way, attends to it as one more
 * tile and writes the rows into their ring slots for the calls that follow. */
int owl_attn_decode_fused(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t table_rows, void* k_cache,
                          void* v_cache, int64_t c_sb, int64_t c_sh, int64_t cap, const int32_t* state, int64_t vis_rows,
                          void* o, int64_t o_sb, int64_t o_sl, int64_t B, int64_t H, int64_t Lq, int64_t dh, float eps,
                          void* workspace, int64_t workspace_bytes, owl_stream_t stream);
int owl_qk_norm_rope_decode(const void* qkv, const float* cos_tab, const float* sin_tab, int64_t table_rows,
                            const int32_t* state, void* q, int64_t q_sb, int64_t q_sh, void* k_cache, void* v_cache,
                            int64_t c_sb, int64_t c_sh, int64_t cap, int64_t B, int64_t L, int64_t H, int64_t dh, float eps,
                            owl_stream_t stream);
int owl_decode_state_advance(int32_t* state, int64_t commit_rows, int64_t drop_rows, int64_t cap, owl_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Rectified-flow glue either side of the model call (bf16, 16-byte vectors).
 * owl_rf_noise: lerpd = x * (1 - t[row]) + z * t[row]; target = z - x, x / z [rows, inner], t [rows]
 *   (models/gamerft.py:93-96, audiorft.py:75-79, gamerft_audio.py:139-146).
 * owl_mse_fwd: *loss = mean((pred - target)^2) accumulated in fp32 (F.mse_loss(pred.float(), target.float()) of
 *   models/*.py without the fp32 copies); owl_mse_bwd: dpred = 2 (pred - target) * *dloss / n. */
int owl_rf_noise(const void* x, const void* z, const void* t, void* lerpd, void* target, int64_t rows, int64_t inner,
                 owl_stream_t stream);
int owl_mse_fwd(const void* pred, const void* target, float* loss, int64_t n, owl_stream_t stream);
int owl_mse_bwd(const void* pred, const void* target, const float* dloss, void* dpred, int64_t n, owl_stream_t stream);

/* Shard reduction of the copy-engine gradient exchange (engine.PeerExchange; replaces the reduction half of DDP's
 * ncclAllReduce, rft_trainer.py:96): dst[i] = (dst[i] + sum_{j < n_src} srcs[j * src_stride + i]) * scale, fp32.  The peers'
 * shards arrive by copy-engine peer copies; `max_ctas` bounds the grid (the kernel runs next to the backward GEMMs). */
int owl_shard_reduce(float* dst, const float* srcs, int64_t src_stride, int n_src, int64_t n, float scale, int max_ctas,
                     owl_stream_t stream);
/* The same with the n_src <= 15 sources given as (host array of) device pointers, e.g. peer-mapped pointers into the other
 * ranks' buckets: the shard is then summed straight over NVLink, without a staging copy. */
int owl_shard_reduce_peers(float* dst, const float* const* srcs, int n_src, int64_t n, float scale, int max_ctas,
                           owl_stream_t stream);

/* In-switch variant (NVLink SHARP): `mc_shard` is the MULTICAST address of this rank's shard of a bucket that every rank
 * allocated symmetrically (torch symmetric memory's multicast_ptr + offset).  Every 16 bytes are fetched from all ranks and
 * summed inside NVSwitch (multimem.ld_reduce), scaled, and written back to all ranks (multimem.st): all-reduce of the shard
 * with no staging and no copies.  The caller brackets the call with device-side barriers across the ranks. */
int owl_mc_allreduce(float* mc_shard, int64_t n, float scale, int max_ctas, owl_stream_t stream);

/* owl_gemm_smallm for callers that cannot allocate (no cudaMalloc under CUDA graph capture, in a tight serving loop, in a
 * harness that forbids it).  NO COUNTERPART IN THE REFERENCE.  Same
 * arithmetic and epilogues as owl_gemm_smallm (no adaLN prologue, no next-weight prefetch); the split-K scratch is a
 * file-scope __device__ array of the library (4 194 816 bytes, resident on every GPU that loads the library until the
 * process exits) instead of a caller workspace, so the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine).  Limit: N <= owl_gemm_smallm_static_max_n() = 8192; a larger N returns OWL_ERR_TOO_LARGE.  All calls on
 * a device share the array and it keeps the previous call's contents: the part a call uses is cleared on `stream` first;
 * calls that overlap on the GPU (other streams, parallel graph branches) race on it and must be serialised by the caller. */
int64_t owl_gemm_smallm_static_max_n(void);
int owl_gemm_smallm_static(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, const void* W, int64_t ldw,
                           const owl_gemm_epilogue* epi, owl_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* OWL_DIT_H_ */
