"""Decode regime of the KV-cached causal rollout: a ring KV cache whose state lives on the device and the lean,
gradient-free launch schedule of a DiT stack for <= 128 new tokens.

The reference's rollout (owl_wms/sampling/av_caching_v2.py:89-140, audio_caching.py:120-150; the per-frame loop of
inference/causvid_pipeline.py:112-163) calls the model once per Euler step on ONE new frame with `kv_cache=...` and
`transformer.decoding = True` (attn.py:102-109: no mask, local layers see the last `local_window` frames).  Every call
reads all weights (2 GB at av_v5 / causvid geometry) for 65 tokens, so the step is weight-streaming + launch bound.

`RingKVCache` keeps per layer ONE zero-initialised [B, H, cap, dh] buffer for K and V used as a ring, and three device
integers (rows cached, ring start, RoPE position).  New rows are always written behind the cached rows
(`owl_qk_norm_rope_decode`); a model call never changes the state, `commit()` counts the rows just written in
(KVCache.update, kv_cache.py:50-58) and `truncate()` moves the start (kv_cache.py:60-75).  Nothing a captured CUDA
graph bakes in (pointers, launch dimensions, scalar arguments) depends on the frame index, so ONE graph of the model call
serves every denoise step and every cache commit of a rollout.

`dit_decode_forward` is the schedule: per layer qkv GEMM with the adaLN PROLOGUE (rms_norm * (1 + a) + b is the A-operand
transform) -> split-KV attention with QK-norm + RoPE + ring append in its load path -> out GEMM with gate+residual epilogue
-> fc1 GEMM with adaLN prologue and bias+SiLU epilogue -> fc2 GEMM with gate+residual epilogue; the GEMMs are the
weight-streaming split-K kernel (`owl_gemm_smallm`), 5 launches per layer."""
from __future__ import annotations

import torch

from .. import functional as OF
from .. import ops
from .._lib import OWL_EPI_BIAS_SILU, OWL_EPI_GATE_RESIDUAL, OWL_EPI_STORE

bf16 = torch.bfloat16
MAX_NEW_ROWS = 128
import os as _os

# Both fusions of the north_star contract exist on this path and are switchable; the defaults are what measured fastest
# (tools/decode_eval.py, DESIGN.md "Decode schedule"):
#   FUSED_ATTENTION  QK-norm + RoPE + ring append inside the attention kernel's load path (one launch instead of two)
#   ADALN_PROLOGUE   adaLN as the A-operand transform of the qkv / fc1 GEMMs (no separate adaLN launch)
FUSED_ATTENTION = _os.environ.get("OWL_DECODE_FUSE_ATTN", "1") != "0"
ADALN_PROLOGUE = _os.environ.get("OWL_DECODE_ADALN_PROLOGUE", "0") != "0"


class RingKVCache:
    """Device-state ring cache (see module docstring).  API subset of SingleKVCache that the decode loop uses, plus
    `commit` / `from_cache`.  `window_rows`: the most rows ever cached at the moment a model call starts; the ring holds
    `window_rows + new_rows` rows so the rows of the frame being denoised always have free slots."""

    def __init__(self, config, batch: int, window_rows: int, new_rows: int, device="cuda"):
        self.config = config
        self.n_layers, self.n_heads = config.n_layers, config.n_heads
        self.head_dim = config.d_model // config.n_heads
        self.batch, self.new_rows = batch, new_rows
        self.cap = int(window_rows + new_rows)
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        # zero-initialised: rows outside the live ring are multiplied by P = 0 in the PV product and must stay finite
        shape = (self.n_layers, batch, self.n_heads, self.cap, self.head_dim)
        self.k = torch.zeros(shape, device=self.device, dtype=bf16)
        self.v = torch.zeros(shape, device=self.device, dtype=bf16)
        self.state = torch.zeros(4, device=self.device, dtype=torch.int32)
        self.n_valid = self.start = self.pos = 0      # host mirror of `state`
        self.should_update = False
        self.noise_caches = 0.0

    # ---- reference-shaped API --------------------------------------------------------------------
    def enable_cache_updates(self):
        self.should_update = True

    def disable_cache_updates(self):
        self.should_update = False

    def length_at(self, idx=0):
        return self.n_valid

    def get_offset(self, idx=0):
        return self.pos

    def __len__(self):
        return self.n_valid

    def n_frames(self):
        assert self.n_valid % self.config.tokens_per_frame == 0
        return self.n_valid // self.config.tokens_per_frame

    def truncate(self, truncate_amt, front=False):
        """kv_cache.py:60-75: eject the `truncate_amt` OLDEST frames (front=False)."""
        if front:
            raise NotImplementedError("RingKVCache drops the oldest frames only (the rollout's use of truncate)")
        self._advance(0, min(truncate_amt * self.config.tokens_per_frame, self.n_valid))

    # ---- ring bookkeeping ----------------------------------------------------------------------
    def commit(self, rows: int):
        """The `rows` rows written by the last model call become part of the cache (KVCache.update)."""
        if self.n_valid + rows > self.cap:
            raise RuntimeError(f"ring cache of {self.cap} rows cannot take {rows} more on top of {self.n_valid}")
        self._advance(rows, 0)

    def _advance(self, commit_rows: int, drop_rows: int):
        ops.decode_state_advance(self.state, commit_rows, drop_rows, self.cap)
        self.n_valid += commit_rows
        self.pos += commit_rows
        self.start = (self.start + drop_rows) % self.cap
        self.n_valid -= drop_rows

    def check_room(self, new_rows: int, table_rows: int):
        if new_rows > MAX_NEW_ROWS:
            raise RuntimeError(f"decode path takes at most {MAX_NEW_ROWS} new rows per call, got {new_rows}")
        if self.n_valid + new_rows > self.cap:
            raise RuntimeError(f"ring cache full: {self.n_valid} cached + {new_rows} new > {self.cap} rows "
                               "(truncate before the call or size the ring for the window)")
        if self.pos + new_rows > table_rows:
            raise RuntimeError(f"RoPE table has {table_rows} positions; need {self.pos + new_rows}")

    @classmethod
    def from_cache(cls, cache, window_rows: int, new_rows: int):
        """Ring with the contents, length and RoPE offset of a prefilled SingleKVCache (the masked prefill of the
        context runs on the training-shape kernels; the per-frame loop continues on the ring)."""
        ring = cls(cache.config, cache._batch, max(window_rows, len(cache)), new_rows, cache._k[0].device)
        n = len(cache)
        for i in range(ring.n_layers):
            k, v = cache.get(i)
            if k.shape[2] != n:
                raise RuntimeError("layers of the source cache hold different lengths")
            ring.k[i, :, :, :n].copy_(k)
            ring.v[i, :, :, :n].copy_(v)
        ring.n_valid, ring.start, ring.pos = n, 0, cache.get_offset(0)
        ring.state.copy_(torch.tensor([ring.n_valid, ring.start, ring.pos, 0], dtype=torch.int32))
        return ring


def _stack_operands(dit):
    """bf16 operands of every block + the stack-level conditioning bank (staged once inside frozen_weights())."""
    per_layer = [OF.stage_block_weights(b._stage, b.block_params(), with_cond=False) for b in dit.blocks]
    cps = [b.cond_params() for b in dit.blocks]
    wc = dit._stage.get("wc_all", [p for ps in cps for p in ps[0::2]])
    bc = dit._stage.get("bc_all", [p for ps in cps for p in ps[1::2]])
    return per_layer, wc, bc


# Which kernel multiplies the square-ish Linears (qkv, out, fc1: N >= K) of a decode evaluation.  "smallm": the
# weight-streaming split-K kernel for all of them.  "tiles": the training GEMM's 64- / 128-wide tiles, no split-K (no
# fp32 atomics, no arrival counter); fc2 (K = 4 N) always streams through the split-K kernel.  OWL_DECODE_GEMM selects;
# measured on B200 (tools/decode_eval.py, causvid.yml model, graph replay): tiles 2.52 ms, smallm 2.69 ms per evaluation.
DECODE_GEMM = _os.environ.get("OWL_DECODE_GEMM", "tiles")


def _lin(a, w, M, N, K, *, mode=0, bias=None, out0=None, aux0=None, aux1=None, rows_per_frame=1, next_w=None, scratch=None):
    if DECODE_GEMM == "tiles" and N >= K:
        kw = {}
        if mode == OWL_EPI_BIAS_SILU:     # the training epilogue writes the pre-activation too
            kw["out1"] = out0
            out0 = scratch
        ops.gemm(0, a, w, M, N, K, mode=mode, bias=bias, out0=out0, aux0=aux0, aux1=aux1, rows_per_frame=rows_per_frame, **kw)
    else:
        ops.gemm_smallm(a, w, M, N, K, mode=mode, bias=bias, out0=out0, aux0=aux0, aux1=aux1, rows_per_frame=rows_per_frame,
                        next_w=next_w)


@torch.no_grad()
def dit_decode_forward(dit, x, cond, ring: RingKVCache):
    """DiT.forward (attn.py:175-191) for <= 128 new tokens against a RingKVCache, decode mode (no mask).  x [B, L, d],
    cond [B, F, d] with L = F * tokens-per-frame rows.  Does not change the ring state: the caller commits."""
    B, L, d = x.shape
    cfg = dit.config
    H = cfg.n_heads
    dh = d // H
    M = B * L
    R = cond.shape[0] * cond.shape[1]
    if M > MAX_NEW_ROWS or R > MAX_NEW_ROWS:
        raise RuntimeError(f"decode path: at most {MAX_NEW_ROWS} rows per call (got {M} tokens, {R} conditioning rows)")
    if M % R:
        raise RuntimeError("tokens must be a multiple of conditioning frames")
    if B != ring.batch:
        raise RuntimeError(f"ring cache holds batch {ring.batch}, forward has batch {B}")
    T = M // R
    rope0 = dit.blocks[0].attn.rope
    ring.check_room(L, rope0.cos.shape[0])
    dev = x.device
    x2 = OF._as_bf16(x).view(M, d)
    scond = ops.silu(OF._as_bf16(cond).view(R, d))
    per_layer, wc, bc = _stack_operands(dit)
    n_layers = len(dit.blocks)
    mod_all = torch.empty(R, 6 * d * n_layers, device=dev, dtype=bf16)
    ops.gemm_smallm(scond, wc, R, 6 * d * n_layers, d, bias=bc, out0=mod_all, next_w=per_layer[0]["wqkv"])
    q = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
    qkv = torch.empty(B, L, 3 * d, device=dev, dtype=bf16)
    o = torch.empty(B, L, d, device=dev, dtype=bf16)
    s = torch.empty(M, 4 * d, device=dev, dtype=bf16)
    z = torch.empty(M, 4 * d, device=dev, dtype=bf16) if DECODE_GEMM == "tiles" else None
    for i, blk in enumerate(dit.blocks):
        W = per_layer[i]
        mod = mod_all[:, i * 6 * d:(i + 1) * 6 * d]
        a1, b1, c1 = mod[:, 0:d], mod[:, d:2 * d], mod[:, 2 * d:3 * d]
        a2, b2, c2 = mod[:, 3 * d:4 * d], mod[:, 4 * d:5 * d], mod[:, 5 * d:6 * d]
        attn = blk.attn
        nxt = per_layer[i + 1]["wqkv"] if i + 1 < n_layers else None
        if ADALN_PROLOGUE:   # adaLN-1 as the A-operand transform of the qkv GEMM (no separate pass over the residual stream)
            ops.gemm_smallm(x2, W["wqkv"], M, 3 * d, d, bias=W["bqkv"], out0=qkv.view(M, 3 * d), next_w=W["wout"],
                            adaln=(a1, b1, T, OF.RMS_EPS))
        else:
            h1, _ = ops.adaln_modulate_fwd(x2, a1, b1, T, OF.RMS_EPS, want_rstd=False)
            _lin(h1, W["wqkv"], M, 3 * d, d, bias=W["bqkv"], out0=qkv.view(M, 3 * d), next_w=W["wout"])
        vis = attn.local_offset if attn.local else 0
        if FUSED_ATTENTION:
            # QK-norm + RoPE + ring append applied in the attention kernel's own load path (one launch)
            ops.attn_decode_fused(qkv, attn.rope.cos, attn.rope.sin, ring.k[i], ring.v[i], ring.state, H, OF.RMS_EPS, vis, out=o)
        else:
            ops.qk_norm_rope_decode(qkv, attn.rope.cos, attn.rope.sin, ring.state, q, ring.k[i], ring.v[i], H, OF.RMS_EPS)
            ops.attn_decode(q, ring.k[i], ring.v[i], ring.state, vis, out=o)
        x1 = torch.empty(M, d, device=dev, dtype=bf16)
        _lin(o.view(M, d), W["wout"], M, d, d, mode=OWL_EPI_GATE_RESIDUAL, bias=W["bout"], out0=x1, aux0=x2, aux1=c1,
             rows_per_frame=T, next_w=W["w1"])
        if ADALN_PROLOGUE:
            ops.gemm_smallm(x1, W["w1"], M, 4 * d, d, mode=OWL_EPI_BIAS_SILU, bias=W["b1"], out0=s, next_w=W["w2"],
                            adaln=(a2, b2, T, OF.RMS_EPS))
        else:
            h2, _ = ops.adaln_modulate_fwd(x1, a2, b2, T, OF.RMS_EPS, want_rstd=False)
            _lin(h2, W["w1"], M, 4 * d, d, mode=OWL_EPI_BIAS_SILU, bias=W["b1"], out0=s, next_w=W["w2"], scratch=z)
        x2n = torch.empty(M, d, device=dev, dtype=bf16)
        ops.gemm_smallm(s, W["w2"], M, d, 4 * d, mode=OWL_EPI_GATE_RESIDUAL, bias=W["b2"], out0=x2n, aux0=x1, aux1=c2,
                        rows_per_frame=T, next_w=nxt)
        x2 = x2n
    out = x2.view(B, L, d)
    return out if x.dtype == bf16 else out.to(x.dtype)
