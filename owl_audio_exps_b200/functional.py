"""Autograd layer over the C-ABI kernels: the DiT block (fused schedule) and its standalone pieces.

This is the host side of the hot path.  Every function here only sequences launches of
libowl_dit.so (through `ops`) and keeps the tensors the backward needs; there is no torch math on
the path except allocation, views and (outside the block) gradient accumulation by autograd.

Reference arithmetic being replaced (paths relative to the reference root):
  DiTBlock.forward   owl_wms/nn/attn.py:130-143
  Attn.forward       owl_wms/nn/attn.py:79-113
  AdaLN / Gate       owl_wms/nn/modulation.py:12-26, 33-43
  MLP.forward        owl_wms/nn/mlp.py:33-37
  rms_norm           owl_wms/nn/normalization.py:10-11
  RoPE.forward       owl_wms/nn/rope.py:43-51

Numerics: bf16 storage, fp32 accumulation, one bf16 rounding per fused op (the eager reference rounds
after every op).  Parameters may be fp32 (autocast regime, rft_trainer.py:146) or bf16 (inference
regime, causvid_pipeline.py:36); activations on the path are always bf16.
"""
from __future__ import annotations

from dataclasses import dataclass
import contextlib
from typing import Optional

import torch

from . import ops
from ._lib import (OWL_EPI_BIAS_SILU, OWL_EPI_DSILU, OWL_EPI_F32, OWL_EPI_GATE_RESIDUAL, OWL_EPI_STORE,
                   OWL_GEMM_NN, OWL_GEMM_NT, OWL_GEMM_TN)

bf16 = torch.bfloat16
f32 = torch.float32

# F.rms_norm(x, (D,)) with eps=None: torch 2.11 adds fp32 machine epsilon to mean(x^2) for bf16 inputs
# (tests/test_block_gpu.py::test_rms_eps_matches_torch re-probes this on the GPU box).
RMS_EPS = 1.1920928955078125e-07
LN_EPS = 1e-5


@dataclass(frozen=True)
class FrameMask:
    """The reference's frame-block mask (attn.py:24-62) as the closed-form descriptor the attention
    kernels evaluate in place of a materialised BlockMask.  `window_frames` <= 0: unlimited."""

    tokens_per_frame: int
    window_frames: int = 0
    q_offset: int = 0
    causal: bool = True
    doc_id: Optional[torch.Tensor] = None  # int32 [B, n_frames_total] on the device, or None
    # (first, last, breaks) from ops.attn_doc_extents(doc_id): per-frame document extents that let the kernels skip
    # tiles of other documents (the BlockMask sparsity of attn.py:52-55,62); None = per-element id compare only
    doc_ext: Optional[tuple] = None

    def kw(self):
        return dict(tokens_per_frame=self.tokens_per_frame, window_frames=self.window_frames,
                    q_offset=self.q_offset, causal=self.causal, doc_id=self.doc_id, doc_ext=self.doc_ext)

    def mask_mod(self, b, h, q, kv):
        """Same predicate as the closure in attn.py:43-59 (for tests / dense rendering)."""
        T = self.tokens_per_frame
        fq, fk = (q + self.q_offset) // T, kv // T
        m = torch.ones_like(fq + fk, dtype=torch.bool)
        if self.causal:
            m = m & (fk <= fq)
        if self.window_frames > 0:
            m = m & ((fq - fk).abs() < self.window_frames)
        if self.doc_id is not None:
            m = m & (self.doc_id[b, fq] == self.doc_id[b, fk])
        return m


DECODE_MASK_T = 1  # decode mode (block_mask None, attn.py:106-107): every key visible


def _decode_mask() -> FrameMask:
    return FrameMask(tokens_per_frame=DECODE_MASK_T, window_frames=0, q_offset=0, causal=False, doc_id=None)


# ------------------------------------------------------------------------------------------------
# parameter staging: fp32 master weights -> bf16 operands (what autocast does per forward)
# ------------------------------------------------------------------------------------------------
def _capturing() -> bool:
    return torch.cuda.is_current_stream_capturing()


_FROZEN = False
_EPOCH = 0   # bumped whenever staged operands may have gone stale without a visible `_version` change


def restage() -> None:
    """Invalidate every staged bf16 operand.  Needed only inside a `frozen_weights()` region after parameters were
    mutated through `.data` (ema_pytorch's EMA.update / copy_params_from_model_to_ema do that: in-place writes through
    `.data` do not bump `Tensor._version`); outside such a region nothing is cached across forwards."""
    global _EPOCH
    _EPOCH += 1


@contextlib.contextmanager
def frozen_weights():
    """Declares that parameters do not change inside this context (one sampling call, one CUDA-graph capture for
    inference): staged bf16 operands are reused across forwards and a graph capture may reuse operands staged before
    it instead of recording the fp32 -> bf16 casts.  Entering the outermost context re-stages once, so whatever
    happened to the parameters before it (optimizer steps, EMA updates through `.data`) is picked up.

    OUTSIDE this context every forward re-casts fp32 master weights / re-concatenates operand groups -- exactly what
    autocast does per forward (its weight cache lives only inside one autocast region)."""
    global _FROZEN, _EPOCH
    prev = _FROZEN
    if not prev:
        _EPOCH += 1
    _FROZEN = True
    try:
        yield
    finally:
        _FROZEN = prev


class WeightStage:
    """bf16 operand copies of a module's parameters.

    Cached across forwards ONLY inside `frozen_weights()` (keyed on data_ptr / _version / dtype / restage epoch);
    otherwise re-staged every forward, because `_version` misses in-place updates made through `.data` (EMA).
    A single contiguous bf16 parameter is never copied (the operand aliases it), so the pure-bf16 inference regime
    pays nothing.  Inside CUDA-graph capture of a training step nothing is cached either: the casts are recorded
    into the graph so each replay sees the current master weights."""

    def __init__(self):
        self._slots = {}
        self._bufs = {}             # persistent bf16 staging buffers (shapes are fixed: same memory every step)
        self.after_backward = None  # engine.GradBuckets: called when the owning block's backward is done
        self.cond_sink = None       # engine.CondSink on a DiT stack's stage: bucket views for the CondBank

    def get(self, key: str, params, rows=None) -> torch.Tensor:
        """bf16 [sum(rows), K] concatenation of 2-D `params` (or fp32 concatenation of 1-D biases)."""
        return self.get_many([(key, params)])[0]

    def get_many(self, items):
        """[(key, params), ...] -> staged operands; every fp32 -> bf16 cast that is due goes into ONE multi-tensor
        launch per 8 matrices (ops.cast_f32_bf16_multi) instead of one launch per parameter."""
        outs, pairs = [], []
        for key, params in items:
            vers = (_EPOCH,) + tuple((p.data_ptr(), p._version, p.dtype) for p in params)
            hit = self._slots.get(key)
            if hit is not None and hit[0] == vers and _FROZEN:
                outs.append(hit[1])
                continue
            p0 = params[0]
            if p0.dim() == 1:
                out = params[0].detach().float() if len(params) == 1 else torch.cat([p.detach().float() for p in params])
                out = out.contiguous()
            elif len(params) == 1 and p0.dtype == bf16 and p0.is_contiguous():
                out = p0.detach()
            else:
                n = sum(p.shape[0] for p in params)
                out = self._bufs.get(key)
                if out is None or out.shape != (n, p0.shape[1]) or out.device != p0.device:
                    out = torch.empty(n, p0.shape[1], device=p0.device, dtype=bf16)
                    self._bufs[key] = out
                r = 0
                for p in params:
                    dst = out[r:r + p.shape[0]]
                    if p.dtype == f32:
                        pairs.append((p.detach().contiguous(), dst))
                    else:
                        dst.copy_(p.detach())
                    r += p.shape[0]
            if _FROZEN and not _capturing():
                self._slots[key] = (vers, out)
            outs.append(out)
        if pairs:
            if all(s_.numel() % 8 == 0 for s_, _ in pairs):
                ops.cast_f32_bf16_multi(pairs)
            else:
                for s_, d_ in pairs:
                    ops.cast_f32_bf16(s_, d_)
        return outs

    def clear(self):
        self._slots.clear()
        self._bufs.clear()

    # EMA deep-copies the model and checkpoints pickle it: the staged operands never travel
    def __deepcopy__(self, memo):
        return WeightStage()

    def __reduce__(self):
        return (WeightStage, ())


def _as_bf16(t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError("owl_audio_exps_b200: CUDA tensors required (sm_100a kernels only; there is no CPU path)")
    if t.dtype == bf16:
        return t if t.is_contiguous() else t.contiguous()
    if t.dtype == f32 and not t.requires_grad:
        return ops.cast_f32_bf16(t.contiguous())
    return t.to(bf16).contiguous()


def _grad_like(g32: Optional[torch.Tensor], p: torch.Tensor):
    if g32 is None:
        return None
    return g32 if p.dtype == f32 else g32.to(p.dtype)


def _sink(p: torch.Tensor):
    """engine.GradSink installed by engine.GradBuckets (.view = fp32 slice of the flat all-reduce bucket,
    .accumulate = add or overwrite): wgrad GEMMs write straight into the bucket instead of
    materialising a gradient tensor for autograd to copy."""
    return getattr(p, "_owl_grad_sink", None)


class _Side:
    """Weight-gradient side stream (installed by engine.GradBuckets; None = everything on the current stream).

    Weight gradients and bias column sums have no consumer until the gradient exchange, so with bucket sinks
    they are enqueued on a second stream: their persistent CTAs fill the SMs the main chain's partial waves,
    small memory-bound kernels and kernel boundaries leave idle.  Operands are kept referenced (`keep`) until
    the main stream has waited on an event recorded behind the side work, so the caching allocator cannot hand
    their memory to a later main-stream allocation while the side stream still reads it."""
    stream = None
    keep: list = []
    pending: list = []   # [(event recorded on the side stream, tensors to release once the main stream waited on it)]


def set_side_stream(stream) -> None:
    _Side.stream = stream
    _Side.keep, _Side.pending = [], []


def _on_side(fn, *keep) -> None:
    side = _Side.stream
    if side is None:
        fn()
        return
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    _Side.keep.extend(keep)


def side_mark() -> None:
    """End of a unit of backward (one block): stamp the side stream, retire what older stamps protected."""
    side = _Side.stream
    if side is None:
        return
    if _Side.keep:
        ev = torch.cuda.Event()
        ev.record(side)
        _Side.pending.append((ev, _Side.keep))
        _Side.keep = []
    while len(_Side.pending) > 2:
        ev, _ = _Side.pending.pop(0)
        torch.cuda.current_stream().wait_event(ev)


def side_join() -> None:
    side = _Side.stream
    if side is None:
        return
    torch.cuda.current_stream().wait_stream(side)
    _Side.keep, _Side.pending = [], []


def _wgrad(dy: torch.Tensor, x: torch.Tensor, p: torch.Tensor, M: int) -> Optional[torch.Tensor]:
    """dW[N,K] = dy[M,N]^T x[M,K] in fp32.  Returns the gradient for autograd, or None when it was
    accumulated into the parameter's bucket sink."""
    N, K = p.shape
    sink = _sink(p)
    if sink is not None:
        _on_side(lambda: ops.gemm(OWL_GEMM_TN, dy, x, N, K, M, mode=OWL_EPI_F32, outf=sink.view,
                                  accumulate=sink.accumulate), dy, x)
        return None
    g = torch.empty(N, K, device=dy.device, dtype=f32)
    ops.gemm(OWL_GEMM_TN, dy, x, N, K, M, mode=OWL_EPI_F32, outf=g)
    return _grad_like(g, p)


def _bgrad_multi(items):
    """[(dy [M, N], bias parameter, wanted), ...] (<= 4) -> bias gradients for autograd (None where the gradient went into
    the parameter's bucket sink or is not wanted); all column sums in ONE launch."""
    out = [None] * len(items)
    sunk, fresh = [], []
    for i, (dy, p, want) in enumerate(items):
        if not want or p is None:
            continue
        sink = _sink(p)
        if sink is not None and sink.pre_zeroed:
            _ = sink.accumulate          # marks the sink written this step (engine bookkeeping)
            sunk.append((dy, sink.view))
        elif sink is not None:
            if not sink.accumulate:
                sink.view.zero_()
            sunk.append((dy, sink.view))
        else:
            fresh.append((i, dy, p))
    if sunk:
        _on_side(lambda: ops.colsum_multi(sunk), *[dy for dy, _ in sunk])
    if fresh:
        buf = torch.zeros(sum(p.shape[0] for _, _, p in fresh), device=fresh[0][1].device, dtype=f32)
        views, o = [], 0
        for i, dy, p in fresh:
            v = buf[o:o + p.shape[0]]
            o += p.shape[0]
            views.append((dy, v))
            out[i] = v
        ops.colsum_multi(views)
        for i, dy, p in fresh:
            out[i] = _grad_like(out[i], p)
    return out


def _bgrad(dy: torch.Tensor, p: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if p is None:
        return None
    sink = _sink(p)
    if sink is not None:
        _on_side(lambda: ops.colsum(dy, sink.view, accumulate=sink.accumulate), dy)
        return None
    g = torch.empty(p.shape[0], device=dy.device, dtype=f32)
    ops.colsum(dy, g)
    return _grad_like(g, p)


# ------------------------------------------------------------------------------------------------
# standalone pieces (used by the AdaLN / Gate / MLP / Attn / FinalLayer modules when they are called
# on their own; DiTBlock uses the fused schedule further down)
# ------------------------------------------------------------------------------------------------
class _Linear(torch.autograd.Function):
    """y = act(x W^T + b).  silu_in: apply SiLU to x first (modulation.py:17-19: fc(silu(cond)));
    silu_out: y = silu(.) (mlp.py:34-35).  dgrad / wgrad / dbias are the NN / TN GEMMs and a column sum."""

    @staticmethod
    def forward(ctx, x, weight, bias, w16, b32, silu_in: bool, silu_out: bool):
        shp = x.shape
        x16 = _as_bf16(x).view(-1, shp[-1])
        M, K = x16.shape
        N = w16.shape[0]
        a = ops.silu(x16) if silu_in else x16
        y = torch.empty(M, N, device=x16.device, dtype=bf16)
        z = None
        if silu_out:
            z = torch.empty_like(y)
            ops.gemm(OWL_GEMM_NT, a, w16, M, N, K, mode=OWL_EPI_BIAS_SILU, bias=b32, out0=z, out1=y)
        else:
            ops.gemm(OWL_GEMM_NT, a, w16, M, N, K, mode=OWL_EPI_STORE, bias=b32, out0=y)
        ctx.save_for_backward(x16 if silu_in else None, a, z, w16, weight, bias)
        ctx.flags = (silu_in, silu_out, shp, x.dtype)
        return y.view(*shp[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x16, a, z, w16, weight, bias = ctx.saved_tensors
        silu_in, silu_out, shp, xdtype = ctx.flags
        N, K = w16.shape
        dy = _as_bf16(dy).view(-1, N)
        M = dy.shape[0]
        if silu_out:  # dz = dy * silu'(z): identity-free path = tiny GEMM-less elementwise via torch is avoided:
            dz = torch.empty_like(dy)
            ops.dsilu_mul(dy, z, dz)
            dy = dz
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, device=dy.device, dtype=bf16)
            if silu_in:
                ops.gemm(OWL_GEMM_NN, dy, w16, M, K, N, mode=OWL_EPI_DSILU, aux0=x16, out0=dx)
            else:
                ops.gemm(OWL_GEMM_NN, dy, w16, M, K, N, mode=OWL_EPI_STORE, out0=dx)
            dx = dx.view(shp).to(xdtype)
        dw = _wgrad(dy, a, weight, M) if ctx.needs_input_grad[1] else None
        db = _bgrad(dy, bias) if (bias is not None and ctx.needs_input_grad[2]) else None
        return dx, dw, db, None, None, None, None


def linear(x, weight, bias, stage: WeightStage, key: str, silu_in=False, silu_out=False):
    K = weight.shape[1]
    if K % 8:
        # TMA rows are 16-byte multiples: a narrow input (ButtonEmbeddding's n_buttons, MouseEmbedding's 2 angle features)
        # is zero-padded to the next multiple of 8; the pad's own autograd slices dW back to [N, K].
        pad = 8 - K % 8
        x = torch.nn.functional.pad(x if x.requires_grad else _as_bf16(x), (0, pad))
        weight = torch.nn.functional.pad(weight, (0, pad))
        b32 = stage.get(key + ".b", [bias]) if bias is not None else None
        return _Linear.apply(x, weight, bias, weight.detach().to(bf16), b32, silu_in, silu_out)
    w16 = stage.get(key + ".w", [weight])
    b32 = stage.get(key + ".b", [bias]) if bias is not None else None
    if not torch.is_grad_enabled():
        M = x.numel() // K
        if 0 < M <= 128:
            # inference at decode shapes (one frame / one token per call): the weight-streaming split-K GEMM
            x16 = _as_bf16(x).view(M, K)
            a = ops.silu(x16) if silu_in else x16
            y = torch.empty(M, w16.shape[0], device=x16.device, dtype=bf16)
            ops.gemm_smallm(a, w16, M, w16.shape[0], K, mode=OWL_EPI_BIAS_SILU if silu_out else OWL_EPI_STORE, bias=b32,
                            out0=y)
            return y.view(*x.shape[:-1], w16.shape[0])
    return _Linear.apply(x, weight, bias, w16, b32, silu_in, silu_out)


class _LayerNorm(torch.autograd.Function):
    """Weight-less LayerNorm over the last dim (normalization.py:6-7) with its backward on the kernel path."""

    @staticmethod
    def forward(ctx, x):
        x16 = _as_bf16(x).view(-1, x.shape[-1])
        ctx.save_for_backward(x16)
        ctx.meta = (x.shape, x.dtype)
        return ops.layer_norm(x16, LN_EPS).view(x.shape).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        (x16,) = ctx.saved_tensors
        shp, dt = ctx.meta
        return ops.layer_norm_bwd(_as_bf16(dy).view(-1, shp[-1]), x16, LN_EPS).view(shp).to(dt)


def layer_norm(x):
    return _LayerNorm.apply(x)


class _Silu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x16 = _as_bf16(x)
        ctx.save_for_backward(x16)
        ctx.xdtype = x.dtype
        return ops.silu(x16)

    @staticmethod
    def backward(ctx, dy):
        (x16,) = ctx.saved_tensors
        dx = torch.empty_like(x16)
        ops.dsilu_mul(_as_bf16(dy), x16, dx)
        return dx.to(ctx.xdtype)


def silu(x):
    return _Silu.apply(x)


class _AdaLNModulate(torch.autograd.Function):
    """h = rms_norm(x) * (1 + a[frame]) + b[frame]; ab = [B, F, 2d] (a | b) or None for plain rms_norm."""

    @staticmethod
    def forward(ctx, x, ab):
        shp = x.shape
        d = shp[-1]
        x16 = _as_bf16(x).view(-1, d)
        M = x16.shape[0]
        if ab is not None:
            ab16 = _as_bf16(ab).view(-1, 2 * d)
            T = M // ab16.shape[0]
            a, b = ab16[:, :d], ab16[:, d:]
        else:
            ab16, T, a, b = None, 1, None, None
        h, rstd = ops.adaln_modulate_fwd(x16, a, b, T, RMS_EPS)
        ctx.save_for_backward(x16, ab16, rstd)
        ctx.meta = (shp, T, x.dtype, None if ab is None else (ab.shape, ab.dtype))
        return h.view(shp)

    @staticmethod
    def backward(ctx, dh):
        x16, ab16, rstd = ctx.saved_tensors
        shp, T, xdtype, abmeta = ctx.meta
        d = shp[-1]
        dh = _as_bf16(dh).view(-1, d)
        if ab16 is not None:
            dab = torch.empty_like(ab16)
            dx = ops.adaln_modulate_bwd(dh, x16, ab16[:, :d], rstd, None, dab[:, :d], dab[:, d:], T)
            return dx.view(shp).to(xdtype), dab.view(abmeta[0]).to(abmeta[1])
        dx = ops.adaln_modulate_bwd(dh, x16, None, rstd, None, None, None, 1)
        return dx.view(shp).to(xdtype), None


def adaln_modulate(x, ab):
    return _AdaLNModulate.apply(x, ab)


def rms_norm(x):
    """Weight-less RMSNorm over the last dim (normalization.py:10-11)."""
    if x.shape[-1] % 8 != 0:
        raise RuntimeError("rms_norm: last dim must be a multiple of 8 for the 128-bit vectorised kernel")
    return _AdaLNModulate.apply(x, None)


class _Gate(torch.autograd.Function):
    """out = c[frame] * x (+ res)   (modulation.py:41-43, cond_gate 57-63)."""

    @staticmethod
    def forward(ctx, x, c, res):
        shp = x.shape
        d = shp[-1]
        x16 = _as_bf16(x).view(-1, d)
        c16 = _as_bf16(c).view(-1, d)
        T = x16.shape[0] // c16.shape[0]
        r16 = _as_bf16(res).view(-1, d) if res is not None else None
        out = ops.gate_fwd(x16, c16, r16, T)
        ctx.save_for_backward(x16, c16)
        ctx.meta = (shp, T, x.dtype, c.shape, c.dtype, res is not None)
        return out.view(shp)

    @staticmethod
    def backward(ctx, dout):
        x16, c16 = ctx.saved_tensors
        shp, T, xdtype, cshape, cdtype, has_res = ctx.meta
        d = shp[-1]
        dout = _as_bf16(dout).view(-1, d)
        dc = torch.empty_like(c16)
        dx = ops.gate_bwd(dout, x16, c16, dc, None, T)
        return dx.view(shp).to(xdtype), dc.view(cshape).to(cdtype), (dout.view(shp) if has_res else None)


def gate(x, c, res=None):
    return _Gate.apply(x, c, res)


# ------------------------------------------------------------------------------------------------
# attention core: qkv -> (QK-RMSNorm, RoPE, head split, KV-cache append) -> masked attention
# ------------------------------------------------------------------------------------------------
@dataclass
class AttnPlan:
    """Everything Attn.forward (attn.py:79-113) decides on the host before launching kernels."""

    n_heads: int
    cos: torch.Tensor            # fp32 [L_max, dh/2]
    sin: torch.Tensor
    mask: Optional[FrameMask]    # None = decode mode (block_mask None)
    local: bool = False
    local_offset: int = 0        # local_window * tokens_per_frame (attn.py:77)
    kv_cache: object = None
    layer_idx: int = 0


def _attn_core_fwd(qkv: torch.Tensor, plan: AttnPlan, need_grad: bool):
    """qkv bf16 [B, L, 3*d] -> o bf16 [B, L, d] plus the tensors the backward needs."""
    B, L, three_d = qkv.shape
    H = plan.n_heads
    dh = three_d // (3 * H)
    if dh not in (64, 128):
        raise RuntimeError(f"head dim {dh} unsupported by the tcgen05 attention kernels (64 or 128)")
    dev = qkv.device
    cache = plan.kv_cache
    offset = cache.get_offset(plan.layer_idx) if cache is not None else 0
    if offset + L > plan.cos.shape[0]:
        raise RuntimeError(f"RoPE table has {plan.cos.shape[0]} positions; need {offset + L}")
    q = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
    n_old = 0
    if cache is not None and (offset > 0 or cache.should_update):
        n_old = cache.length_at(plan.layer_idx) if offset > 0 else 0
        # in-place: new K/V rows are written by the kernel straight behind the cached rows.  When the
        # backward will need them later (self-forcing, sf_vid_only.py:180-190) the attention operands
        # must not alias the cache (it is overwritten by later frames), so a private copy is used.
        private = need_grad and not cache.should_update
        k_all, v_all = cache._operands(plan.layer_idx, L, use_old=offset > 0, private=private, batch=B, device=dev)
        k_new, v_new = k_all[:, :, n_old:], v_all[:, :, n_old:]
    else:
        k_all = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
        v_all = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
        k_new, v_new = k_all, v_all
    rq, rk = ops.qk_norm_rope_fwd(qkv, plan.cos, plan.sin, offset, H, RMS_EPS, q, k_new, v_new, want_rstd=need_grad)
    if cache is not None and cache.should_update:
        cache._commit(plan.layer_idx, L, k_all, v_all)
        if need_grad:
            # the committed rows may be overwritten / compacted later: detach the autograd copy
            k_all, v_all = k_all.clone(), v_all.clone()
    mask = plan.mask
    if mask is None:
        mask = _decode_mask()
        if plan.local and plan.local_offset < k_all.shape[2]:
            k_all, v_all = k_all[:, :, -plan.local_offset:], v_all[:, :, -plan.local_offset:]
    o, lse = ops.attn_fwd(q, k_all, v_all, want_lse=need_grad, **mask.kw())
    return o, (q, k_all, v_all, lse, rq, rk, offset, mask)


def _attn_core_bwd(do: torch.Tensor, o: torch.Tensor, saved, plan_cos, plan_sin, L: int):
    """-> dqkv bf16 [B, L, 3*d].  Cached key/value rows carry no gradient (they were produced under
    no_grad, av_caching_v2.py:68-76); only the last L rows of dk/dv flow back."""
    q, k_all, v_all, lse, rq, rk, offset, mask = saved
    dq, dk, dv = ops.attn_bwd(do, o, q, k_all, v_all, lse, **mask.kw())
    Lk = k_all.shape[2]
    if Lk != L:
        dk, dv = dk[:, :, Lk - L:], dv[:, :, Lk - L:]
        k_new = k_all[:, :, Lk - L:]
        if dq.stride() != dk.stride():
            dk, dv = dk.contiguous(), dv.contiguous()
    else:
        k_new = k_all
    if not (k_new.stride(3) == 1 and k_new.stride(2) == k_new.shape[3]):
        k_new = k_new.contiguous()
    return ops.qk_norm_rope_bwd(dq, dk, dv, q, k_new, rq, rk, plan_cos, plan_sin, offset)


class _AttnCore(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, plan: AttnPlan):
        need_grad = ctx.needs_input_grad[0]
        qkv16 = _as_bf16(qkv)
        o, saved = _attn_core_fwd(qkv16, plan, need_grad)
        if need_grad:
            q, k_all, v_all, lse, rq, rk, offset, mask = saved
            ctx.save_for_backward(o, q, k_all, v_all, lse, rq, rk, plan.cos, plan.sin)
            ctx.meta = (offset, mask, qkv.shape[1], qkv.dtype)
        return o

    @staticmethod
    def backward(ctx, do):
        o, q, k_all, v_all, lse, rq, rk, cos, sin = ctx.saved_tensors
        offset, mask, L, dtype = ctx.meta
        dqkv = _attn_core_bwd(_as_bf16(do), o, (q, k_all, v_all, lse, rq, rk, offset, mask), cos, sin, L)
        return dqkv.to(dtype), None


def attention(qkv, plan: AttnPlan):
    return _AttnCore.apply(qkv, plan)


class _Rope(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cos, sin, offset: int):
        ctx.save_for_backward(cos, sin)
        ctx.meta = (offset, x.dtype)
        L = x.shape[-2]
        if offset + L > cos.shape[0]:
            raise RuntimeError(f"RoPE table has {cos.shape[0]} positions; need {offset + L}")
        return ops.rope_apply(_as_bf16(x), cos, sin, offset).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        cos, sin = ctx.saved_tensors
        offset, dtype = ctx.meta
        return ops.rope_apply(_as_bf16(dy), cos, sin, offset, inverse=True).to(dtype), None, None, None


def rope_only(x, cos, sin, offset: int = 0):
    return _Rope.apply(x, cos, sin, offset)


# ------------------------------------------------------------------------------------------------
# conditioning bank: the 4 x n_layers modulation projections of a DiT stack as ONE GEMM
# ------------------------------------------------------------------------------------------------
COND_PARAMS_PER_LAYER = 8   # adaln1.fc.{weight,bias}, gate1.fc_c.{weight,bias}, adaln2.fc.{...}, gate2.fc_c.{...}


class CondBank:
    """Per-forward state of the batched conditioning projections of a DiT stack.

    `cond` is layer-invariant (attn.py:185-189 passes the same tensor to every block), and every block
    applies four Linear(silu(cond)) to it (modulation.py:10,31).  With one token per frame (audio.yml)
    these are a third of all linear FLOPs, issued as 16 x 6 small GEMMs.  The bank computes
        mod_all[R, n_layers*6d] = silu(cond) Wc_all^T + bc_all
    once per DiT.forward, blocks read their [R, 6d] column slice, write their dmod slice into
    `dmod_all` during backward, and the weight gradients are taken `group` layers at a time as those
    layers' backward completes (so the all-reduce of that slice overlaps the remaining backward)."""

    def __init__(self, n_layers: int, d: int, stage: "WeightStage", group: int = 4):
        self.n_layers, self.d, self.stage, self.group = n_layers, d, stage, max(1, group)
        self.mod_all = None      # bf16 [R, n_layers*6d]
        self.dmod_all = None     # bf16 [R, n_layers*6d], allocated by the first block backward
        self.scond16 = None
        self.gw = self.gb = None  # fp32 [n_layers*6d, d] / [n_layers*6d]: bucket sink views or private buffers
        self.accumulate = False
        self.gb_pre_zeroed = False   # engine zeroes the bias-gradient sinks at begin_step: column sums always add
        self.after_group = None  # engine hook: (row_lo, row_hi) of gw finished
        self.want_wgrad = False
        self.sunk = False        # gw/gb are engine bucket views (then the group wgrads may run on the side stream)
        self.sink = None         # engine.CondSink (decides add-vs-overwrite per row group at write time)
        self.token_grad = None
        # Cross-layer fusion inside a linear stack (DiT.forward sets `chain`): the memory-bound passes either side of a
        # layer boundary are one kernel, handed over through these slots (a cache of work done early, no autograd edge):
        #   pre[l]      = (x_in, h1, rstd1)  block l-1 already normalised block l's input (residual+gate+adaLN pass)
        #   tail[l]     = (y2, c2)     what block l+1's backward needs to run block l's last gate backward
        #   pre_bwd[l]  = (dx2, dy2)   block l+1's backward already did it
        self.chain = False
        self.pre, self.tail, self.pre_bwd = {}, {}, {}
        self._done = set()
        self._flushed = 0        # layers [_flushed_lo, n_layers) already went through the group wgrad
        self._lo = n_layers

    def mod(self, layer: int) -> torch.Tensor:
        w = 6 * self.d
        return self.mod_all[:, layer * w:(layer + 1) * w]

    def dmod(self, layer: int) -> torch.Tensor:
        if self.dmod_all is None:
            self.dmod_all = torch.empty_like(self.mod_all)
        w = 6 * self.d
        return self.dmod_all[:, layer * w:(layer + 1) * w]

    def block_done(self, layer: int):
        """Called by DiTBlockFn.backward once the block's dmod slice is complete (blocks finish in
        reverse order): takes the weight gradient of every complete group of layers at the tail."""
        self._done.add(layer)
        while self._lo > 0:
            lo = max(0, (self._lo - 1) // self.group * self.group)
            if not all(l in self._done for l in range(lo, self._lo)):
                break
            self._wgrad_rows(lo, self._lo)
            self._lo = lo

    def flush(self):
        """Take whatever has not been through a group wgrad yet (blocks that never ran backward
        contribute zeros: their dmod slice is cleared first)."""
        if self._lo > 0 and self.want_wgrad:
            if self.dmod_all is None:
                self.dmod_all = torch.zeros_like(self.mod_all)
            for l in range(self._lo):
                if l not in self._done:
                    self.dmod(l).zero_()
            self._wgrad_rows(0, self._lo)
        self._lo = 0

    def _wgrad_rows(self, lo: int, hi: int):
        if not self.want_wgrad:
            return
        w = 6 * self.d
        r0, r1 = lo * w, hi * w
        R = self.mod_all.shape[0]
        dm = self.dmod_all[:, r0:r1]

        acc = self.sink.acc_rows(r0) if self.sink is not None else self.accumulate

        def run():
            ops.gemm(OWL_GEMM_TN, dm, self.scond16, r1 - r0, self.d, R, mode=OWL_EPI_F32, outf=self.gw[r0:r1],
                     accumulate=acc)
            ops.colsum(dm, self.gb[r0:r1], accumulate=acc or self.gb_pre_zeroed)
        if self.sunk:
            _on_side(run, self.dmod_all, self.scond16)
        else:
            run()   # private gradient buffers are handed to autograd on the current stream
        if self.after_group is not None:
            self.after_group(r0, r1)


class CondBankFn(torch.autograd.Function):
    """token = f(scond, P...) with side effect bank.mod_all.  The returned 1-element `token` only
    carries the autograd edge: every block takes it as an input, so this node's backward runs after all
    blocks' backward and finishes the job (remaining group wgrads, then dscond = dmod_all Wc_all)."""

    @staticmethod
    def forward(ctx, scond, bank: CondBank, *P):
        d = bank.d
        scond16 = _as_bf16(scond).view(-1, d)
        R = scond16.shape[0]
        ws, bs = list(P[0::2]), list(P[1::2])
        wc = bank.stage.get("wc_all", ws)
        bc = bank.stage.get("bc_all", bs)
        N = wc.shape[0]
        bank.mod_all = torch.empty(R, N, device=scond16.device, dtype=bf16)
        ops.gemm(OWL_GEMM_NT, scond16, wc, R, N, d, mode=OWL_EPI_STORE, bias=bc, out0=bank.mod_all)
        bank.scond16 = scond16
        bank.want_wgrad = any(ctx.needs_input_grad[2:])
        if any(ctx.needs_input_grad):
            sink = getattr(bank.stage, "cond_sink", None)
            if bank.want_wgrad:
                if sink is not None:
                    bank.gw, bank.gb, bank.accumulate, bank.after_group = sink.gw, sink.gb, sink.accumulate, sink.after_group
                    bank.sunk = True
                    bank.sink = sink
                    bank.gb_pre_zeroed = getattr(sink, "gb_pre_zeroed", False)
                else:
                    bank.gw = torch.empty(N, d, device=scond16.device, dtype=f32)
                    bank.gb = torch.empty(N, device=scond16.device, dtype=f32)
            ctx.sunk = sink is not None
            ctx.save_for_backward(wc, *P)
            ctx.bank = bank
            ctx.meta = (scond.shape, scond.dtype)
            bank.token_grad = torch.zeros(1, device=scond16.device, dtype=f32)
        return torch.zeros(1, device=scond16.device, dtype=f32)

    @staticmethod
    def backward(ctx, dtoken):
        bank = ctx.bank
        wc = ctx.saved_tensors[0]
        P = ctx.saved_tensors[1:]
        d, N = bank.d, wc.shape[0]
        bank.flush()
        dscond = None
        if ctx.needs_input_grad[0]:
            shape, dtype = ctx.meta
            R = bank.mod_all.shape[0]
            if bank.dmod_all is None:
                bank.dmod_all = torch.zeros_like(bank.mod_all)
            # [R, d] output = few tiles, K = 6d*n_layers = many k-blocks: fp32 output so the planner may split K over
            # all SMs (TMA reduce-add), then one small cast
            acc = torch.empty(R, d, device=wc.device, dtype=f32)
            ops.gemm(OWL_GEMM_NN, bank.dmod_all, wc, R, d, N, mode=OWL_EPI_F32, outf=acc)
            dscond = ops.cast_f32_bf16(acc).view(shape).to(dtype)
        grads = [None] * len(P)
        if bank.want_wgrad and not ctx.sunk:
            r = 0
            for i in range(0, len(P), 2):
                n = P[i].shape[0]
                if ctx.needs_input_grad[2 + i]:
                    grads[i] = _grad_like(bank.gw[r:r + n], P[i])
                if ctx.needs_input_grad[3 + i]:
                    grads[i + 1] = _grad_like(bank.gb[r:r + n], P[i + 1])
                r += n
        bank.dmod_all = bank.mod_all = None
        bank.pre.clear(), bank.tail.clear(), bank.pre_bwd.clear()
        side_mark()
        return (dscond, None, *grads)


def cond_bank(scond, blocks_params, d: int, stage: "WeightStage", group: int = 4):
    """-> (bank, token).  blocks_params: per layer (wa1, ba1, wg1, bg1, wa2, ba2, wg2, bg2)."""
    bank = CondBank(len(blocks_params), d, stage, group)
    flat = [p for ps in blocks_params for p in ps]
    token = CondBankFn.apply(scond, bank, *flat)
    return bank, token


# ------------------------------------------------------------------------------------------------
# the fused DiT block (attn.py:130-143): 13 launches forward, 22 backward
# ------------------------------------------------------------------------------------------------
_BLOCK_PARAMS = ("attn.qkv.weight", "attn.qkv.bias", "attn.out.weight", "attn.out.bias",
                 "mlp.fc1.weight", "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias",
                 "adaln1.fc.weight", "adaln1.fc.bias", "gate1.fc_c.weight", "gate1.fc_c.bias",
                 "adaln2.fc.weight", "adaln2.fc.bias", "gate2.fc_c.weight", "gate2.fc_c.bias")


def stage_block_weights(stage: WeightStage, P, with_cond: bool = True):
    """bf16 operands + fp32 biases of one block.  The four conditioning projections (adaln1.fc,
    gate1.fc_c, adaln2.fc, gate2.fc_c; modulation.py:10,31) share their input silu(cond), so they are
    staged as ONE [6d, d] operand: columns [a1 | b1 | c1 | a2 | b2 | c2] of a single GEMM."""
    (wqkv, bqkv, wout, bout, w1, b1, w2, b2, wa1, ba1, wg1, bg1, wa2, ba2, wg2, bg2) = P
    items = [("wqkv", [wqkv]), ("bqkv", [bqkv]), ("wout", [wout]), ("bout", [bout]),
             ("w1", [w1]), ("b1", [b1]), ("w2", [w2]), ("b2", [b2])]
    if with_cond:  # not needed when the stack's CondBank already projected the conditioning
        items += [("wc", [wa1, wg1, wa2, wg2]), ("bc", [ba1, bg1, ba2, bg2])]
    W = dict(zip([k for k, _ in items], stage.get_many(items)))
    return W


class DiTBlockFn(torch.autograd.Function):
    """x_out = DiTBlock(x, cond).  Inputs: x [B,L,d], cond [B,F,d] (L = F*T), the 16 block parameters.

    Forward schedule (M = B*L rows, R = B*F conditioning rows):
      1  scond = silu(cond)                                   [R,d]      (skipped when the DiT stack passes it in)
      2  mod   = scond Wc^T + bc                              [R,6d]     GEMM
      3  h1    = rms_norm(x)*(1+a1)+b1 ; rstd1                [M,d]      adaLN kernel
      4  qkv   = h1 Wqkv^T + bqkv                             [M,3d]     GEMM
      5  q,k,v = rope(rms_norm(.)) + head split (+KV append)  [B,H,L,dh] QK-norm+RoPE kernel
      6  o,lse = attention(q,k,v,mask)                        [M,d]      tcgen05 attention (token-major out)
      7  y1 = o Wout^T + bout                                 [M,d]      GEMM
      8  x1 = x + c1*y1 ; h2 = rms_norm(x1)*(1+a2)+b2 ; rstd2 [M,d]      residual+adaLN kernel (one pass)
      9  z = h2 W1^T + b1 ; s = silu(z)                       [M,4d]     GEMM, bias+SiLU epilogue
      10 y2 = s W2^T + b2                                     [M,d]      GEMM
      11 x2 = x1 + c2*y2                                      [M,d]      gate+residual kernel
    The backward mirrors it with dgrad (NN) and wgrad (TN, fp32) GEMMs; silu', the residual-gradient
    add, the per-frame da/db/dc reductions and the bias column sums are fused into kernel epilogues or
    the single-pass memory-bound kernels.
    """

    @staticmethod
    def forward(ctx, x, cond, scond, token, plan: AttnPlan, stage: WeightStage, bank, light, *P):
        """`light`: activation checkpointing sized for 180 GB (attn.py:187-188, dit_v4_5B.yml `gradient_checkpointing`): keep
        only the block input, the modulation vectors, the attention output and its log-sum-exp; the backward re-runs the
        cheap part of the forward (adaLN, the four GEMMs, QK-norm/RoPE) and NOT the attention kernel, which torch's
        per-block checkpoint would run a second time."""
        B, L, d = x.shape
        R = cond.shape[0] * cond.shape[1]
        M = B * L
        if M % R != 0:
            raise RuntimeError(f"tokens ({L}) must be a multiple of conditioning frames ({cond.shape[1]})")
        T = M // R
        need_grad = any(ctx.needs_input_grad)
        W = stage_block_weights(stage, P, with_cond=bank is None)
        ops.stamp("f0 cast")
        x16 = _as_bf16(x).view(M, d)
        cond16 = _as_bf16(cond).view(R, d)
        if bank is not None:
            # conditioning bank (CondBank): the stack already projected silu(cond) for every layer
            scond16 = bank.scond16
            mod = bank.mod(plan.layer_idx)
        else:
            scond16 = ops.silu(cond16) if scond is None else scond.view(R, d)
            mod = torch.empty(R, 6 * d, device=x16.device, dtype=bf16)
            ops.gemm(OWL_GEMM_NT, scond16, W["wc"], R, 6 * d, d, mode=OWL_EPI_STORE, bias=W["bc"], out0=mod)
        a1, b1, c1 = mod[:, 0:d], mod[:, d:2 * d], mod[:, 2 * d:3 * d]
        a2, b2, c2 = mod[:, 3 * d:4 * d], mod[:, 4 * d:5 * d], mod[:, 5 * d:6 * d]

        pre = bank.pre.pop(plan.layer_idx, None) if bank is not None else None
        if pre is not None and pre[0].data_ptr() == x16.data_ptr() and pre[0].shape == x16.shape:
            h1, rstd1 = pre[1], pre[2]   # the previous block's last pass already produced them
        else:
            h1, rstd1 = ops.adaln_modulate_fwd(x16, a1, b1, T, RMS_EPS, want_rstd=need_grad)
        ops.stamp("f1 adaln1")
        qkv = torch.empty(B, L, 3 * d, device=x16.device, dtype=bf16)
        ops.gemm(OWL_GEMM_NT, h1, W["wqkv"], M, 3 * d, d, mode=OWL_EPI_STORE, bias=W["bqkv"], out0=qkv)
        ops.stamp("f2 qkv")
        o, asaved = _attn_core_fwd(qkv, plan, need_grad)
        ops.stamp("f3 rope+attn")
        del qkv
        o2 = o.view(M, d)

        # out projection (store-only epilogue); its gate and residual ride on the next norm's single pass
        y1 = torch.empty(M, d, device=x16.device, dtype=bf16)
        ops.gemm(OWL_GEMM_NT, o2, W["wout"], M, d, d, mode=OWL_EPI_STORE, bias=W["bout"], out0=y1)
        ops.stamp("f4 out")
        x1, h2, rstd2 = ops.residual_adaln_fwd(x16, y1, c1, a2, b2, T, RMS_EPS, want_rstd=need_grad)
        ops.stamp("f5 res+adaln2")
        z = torch.empty(M, 4 * d, device=x16.device, dtype=bf16)
        s = torch.empty(M, 4 * d, device=x16.device, dtype=bf16)
        ops.gemm(OWL_GEMM_NT, h2, W["w1"], M, 4 * d, d, mode=OWL_EPI_BIAS_SILU, bias=W["b1"], out0=z, out1=s)
        ops.stamp("f6 fc1")
        y2 = torch.empty(M, d, device=x16.device, dtype=bf16)
        ops.gemm(OWL_GEMM_NT, s, W["w2"], M, d, 4 * d, mode=OWL_EPI_STORE, bias=W["b2"], out0=y2)
        ops.stamp("f7 fc2")
        nxt = plan.layer_idx + 1
        if bank is not None and bank.chain and nxt < bank.n_layers:
            # x2 = x1 + c2*y2 and the NEXT block's h1 = rms_norm(x2)*(1+a1')+b1' in one pass (its modulation vectors are
            # already in the bank)
            nm = bank.mod(nxt)
            x2, h1n, r1n = ops.residual_adaln_fwd(x1, y2, c2, nm[:, 0:d], nm[:, d:2 * d], T, RMS_EPS, want_rstd=True)
            bank.pre[nxt] = (x2, h1n, r1n)
        else:
            x2 = ops.gate_fwd(y2, c2, x1, T)
        ops.stamp("f8 gate2")

        light = bool(light) and plan.kv_cache is None
        ctx.light = light
        if need_grad:
            q, k_all, v_all, lse, rq, rk, offset, mask = asaved
            if light:
                ctx.save_for_backward(x16, cond16, scond16, mod, None, None, None, None, None, lse, None, None, o2, None, None,
                                      None, None, None, None, None, W["wqkv"], W["wout"], W["w1"], W["w2"],
                                      W["wc"] if bank is None else None, plan.cos, plan.sin, *P)
                ctx.n_heads = plan.n_heads
                ctx.biases = (W["bqkv"], W["bout"], W["b1"], W["b2"])
            else:
                ctx.save_for_backward(x16, cond16, scond16, mod, rstd1, h1, q, k_all, v_all, lse, rq, rk, o2, y1, x1, rstd2,
                                      h2, z, s, y2, W["wqkv"], W["wout"], W["w1"], W["w2"],
                                      W["wc"] if bank is None else None, plan.cos, plan.sin, *P)
            ctx.meta = (B, L, d, R, T, offset, mask, x.dtype, cond.dtype, cond.shape,
                        None if scond is None else scond.shape)
            ctx.stage = stage
            ctx.bank = bank
            ctx.layer = plan.layer_idx
            if bank is not None and bank.chain and not light:
                bank.tail[plan.layer_idx] = (y2, c2)
        return x2.view(B, L, d) if x.dtype == bf16 else x2.view(B, L, d).to(x.dtype)

    @staticmethod
    def backward(ctx, dx2):
        sv = ctx.saved_tensors
        (x16, cond16, scond16, mod, rstd1, h1, q, k_all, v_all, lse, rq, rk, o2, y1, x1, rstd2, h2, z, s, y2,
         wqkv16, wout16, w116, w216, wc16, cos, sin) = sv[:27]
        P = sv[27:]
        (wqkv, bqkv, wout, bout, w1, b1, w2, b2, wa1, ba1, wg1, bg1, wa2, ba2, wg2, bg2) = P
        B, L, d, R, T, offset, mask, xdtype, cdtype, cond_shape, scond_shape = ctx.meta
        M = B * L
        dev = x16.device
        need = ctx.needs_input_grad
        np0 = 8  # index of the first parameter among forward's inputs
        if ctx.light:
            # re-run the forward up to what the backward reads, except the attention (its output and lse were kept)
            bq, bo_, b1_, b2_ = ctx.biases
            H = ctx.n_heads
            dh = d // H
            a1_, b1m, c1_ = mod[:, 0:d], mod[:, d:2 * d], mod[:, 2 * d:3 * d]
            a2_, b2m = mod[:, 3 * d:4 * d], mod[:, 4 * d:5 * d]
            h1, rstd1 = ops.adaln_modulate_fwd(x16, a1_, b1m, T, RMS_EPS, want_rstd=True)
            qkv = torch.empty(B, L, 3 * d, device=dev, dtype=bf16)
            ops.gemm(OWL_GEMM_NT, h1, wqkv16, M, 3 * d, d, mode=OWL_EPI_STORE, bias=bq, out0=qkv)
            q = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
            k_all, v_all = torch.empty_like(q), torch.empty_like(q)
            rq, rk = ops.qk_norm_rope_fwd(qkv, cos, sin, offset, H, RMS_EPS, q, k_all, v_all, want_rstd=True)
            del qkv
            y1 = torch.empty(M, d, device=dev, dtype=bf16)
            ops.gemm(OWL_GEMM_NT, o2, wout16, M, d, d, mode=OWL_EPI_STORE, bias=bo_, out0=y1)
            x1, h2, rstd2 = ops.residual_adaln_fwd(x16, y1, c1_, a2_, b2m, T, RMS_EPS, want_rstd=True)
            z = torch.empty(M, 4 * d, device=dev, dtype=bf16)
            s = torch.empty(M, 4 * d, device=dev, dtype=bf16)
            ops.gemm(OWL_GEMM_NT, h2, w116, M, 4 * d, d, mode=OWL_EPI_BIAS_SILU, bias=b1_, out0=z, out1=s)
            y2 = torch.empty(M, d, device=dev, dtype=bf16)
            ops.gemm(OWL_GEMM_NT, s, w216, M, d, 4 * d, mode=OWL_EPI_STORE, bias=b2_, out0=y2)
        bank = ctx.bank
        dx2 = _as_bf16(dx2).view(M, d)
        c1, a2, c2, a1 = mod[:, 2 * d:3 * d], mod[:, 3 * d:4 * d], mod[:, 5 * d:6 * d], mod[:, 0:d]
        dmod = bank.dmod(ctx.layer) if bank is not None else torch.empty(R, 6 * d, device=dev, dtype=bf16)
        ops.stamp("b0 start")

        # ---- MLP branch: x2 = x1 + c2 * (s W2^T + b2)
        pb = bank.pre_bwd.pop(ctx.layer, None) if bank is not None else None
        if pb is not None:
            # the next block's backward ran this block's last gate backward together with its own adaLN backward
            if pb[0].data_ptr() != dx2.data_ptr():
                raise RuntimeError("DiT stack: the gradient reaching block %d is not the tensor block %d produced (the "
                                   "stack output of an inner block was used elsewhere); run the blocks without the "
                                   "conditioning bank (DiT.cond_group = 0)" % (ctx.layer, ctx.layer + 1))
            dy2 = pb[1]
        else:
            dy2 = ops.gate_bwd(dx2, y2, c2, dmod[:, 5 * d:6 * d], None, T)
        ops.stamp("b1 gate2")
        dz = torch.empty(M, 4 * d, device=dev, dtype=bf16)
        ops.gemm(OWL_GEMM_NN, dy2, w216, M, 4 * d, d, mode=OWL_EPI_DSILU, aux0=z, out0=dz)
        ops.stamp("b2 dz")
        g_w2 = _wgrad(dy2, s, w2, M) if need[np0 + 6] else None
        dh2 = torch.empty(M, d, device=dev, dtype=bf16)
        ops.gemm(OWL_GEMM_NN, dz, w116, M, d, 4 * d, mode=OWL_EPI_STORE, out0=dh2)
        ops.stamp("b3 dh2")
        g_w1 = _wgrad(dz, h2, w1, M) if need[np0 + 4] else None
        # ---- adaLN-2 backward and the attention branch's gate backward (x1 = x + c1 * (o Wout^T + bout)) in ONE pass
        dx1, dy1 = ops.adaln_gate_bwd(dh2, x1, a2, rstd2, dx2, dmod[:, 3 * d:4 * d], dmod[:, 4 * d:5 * d],
                                      y1, c1, dmod[:, 2 * d:3 * d], None, T)
        ops.stamp("b5 adaln2+gate1")
        del dh2
        do = torch.empty(B, L, d, device=dev, dtype=bf16)
        ops.gemm(OWL_GEMM_NN, dy1, wout16, M, d, d, mode=OWL_EPI_STORE, out0=do)
        ops.stamp("b6 do")
        g_wout = _wgrad(dy1, o2, wout, M) if need[np0 + 2] else None
        dqkv = _attn_core_bwd(do, o2.view(B, L, d), (q, k_all, v_all, lse, rq, rk, offset, mask), cos, sin, L)
        ops.stamp("b7 attn+rope")
        dqkv2 = dqkv.view(M, 3 * d)
        # the block's four bias gradients = column sums of dy2 / dz / dy1 / dqkv: ONE launch (next to the weight gradients
        # on the side stream when the engine's buckets are installed)
        g_b2, g_b1, g_bout, g_bqkv = _bgrad_multi([(dy2, b2, need[np0 + 7]), (dz, b1, need[np0 + 5]),
                                                   (dy1, bout, need[np0 + 3]), (dqkv2, bqkv, need[np0 + 1])])
        del dy2, dz, dy1
        dh1 = torch.empty(M, d, device=dev, dtype=bf16)
        ops.gemm(OWL_GEMM_NN, dqkv2, wqkv16, M, d, 3 * d, mode=OWL_EPI_STORE, out0=dh1)
        ops.stamp("b8 dh1")
        g_wqkv = _wgrad(dqkv2, h1, wqkv, M) if need[np0 + 0] else None
        del dqkv, dqkv2
        tail = bank.tail.pop(ctx.layer - 1, None) if (bank is not None and bank.chain and need[0]) else None
        if tail is not None and xdtype == bf16:
            # adaLN-1 backward of this block + the last gate backward of the PREVIOUS block in one pass
            y2p, c2p = tail
            dx, dy2p = ops.adaln_gate_bwd(dh1, x16, a1, rstd1, dx1, dmod[:, 0:d], dmod[:, d:2 * d], y2p, c2p,
                                          bank.dmod(ctx.layer - 1)[:, 5 * d:6 * d], None, T)
            bank.pre_bwd[ctx.layer - 1] = (dx, dy2p)
        else:
            dx = ops.adaln_modulate_bwd(dh1, x16, a1, rstd1, dx1, dmod[:, 0:d], dmod[:, d:2 * d], T)
        ops.stamp("b9 adaln1")

        # ---- conditioning projections: mod = silu(cond) Wc^T + bc
        g_wa1 = g_wg1 = g_wa2 = g_wg2 = g_ba1 = g_bg1 = g_ba2 = g_bg2 = None
        dtoken = None
        if bank is not None:
            # the bank takes the projections' gradients for whole groups of layers (CondBank.block_done)
            bank.block_done(ctx.layer)
            if need[3] and ctx.layer == 0:
                dtoken = bank.token_grad
        elif any(need[np0 + 8:np0 + 16]):
            sinks = [_sink(p) for p in (wa1, wg1, wa2, wg2)]
            if all(s_ is not None for s_ in sinks):
                for p, lo, hi in ((wa1, 0, 2 * d), (wg1, 2 * d, 3 * d), (wa2, 3 * d, 5 * d), (wg2, 5 * d, 6 * d)):
                    _wgrad(dmod[:, lo:hi], scond16, p, R)
                for p, lo, hi in ((ba1, 0, 2 * d), (bg1, 2 * d, 3 * d), (ba2, 3 * d, 5 * d), (bg2, 5 * d, 6 * d)):
                    _bgrad(dmod[:, lo:hi], p)
            else:
                gwc = torch.empty(6 * d, d, device=dev, dtype=f32)
                ops.gemm(OWL_GEMM_TN, dmod, scond16, 6 * d, d, R, mode=OWL_EPI_F32, outf=gwc)
                gbc = torch.empty(6 * d, device=dev, dtype=f32)
                ops.colsum(dmod, gbc)
                g_wa1, g_wg1 = _grad_like(gwc[0:2 * d], wa1), _grad_like(gwc[2 * d:3 * d], wg1)
                g_wa2, g_wg2 = _grad_like(gwc[3 * d:5 * d], wa2), _grad_like(gwc[5 * d:6 * d], wg2)
                g_ba1, g_bg1 = _grad_like(gbc[0:2 * d], ba1), _grad_like(gbc[2 * d:3 * d], bg1)
                g_ba2, g_bg2 = _grad_like(gbc[3 * d:5 * d], ba2), _grad_like(gbc[5 * d:6 * d], bg2)
        dcond = dscond = None
        if bank is not None:
            pass
        elif scond_shape is not None:
            if need[2]:
                dscond = torch.empty(R, d, device=dev, dtype=bf16)
                ops.gemm(OWL_GEMM_NN, dmod, wc16, R, d, 6 * d, mode=OWL_EPI_STORE, out0=dscond)
                dscond = dscond.view(scond_shape)
        elif need[1]:
            dcond = torch.empty(R, d, device=dev, dtype=bf16)
            ops.gemm(OWL_GEMM_NN, dmod, wc16, R, d, 6 * d, mode=OWL_EPI_DSILU, aux0=cond16, out0=dcond)
            dcond = dcond.view(cond_shape).to(cdtype)

        dx = dx.view(B, L, d)
        if xdtype != bf16:
            dx = dx.to(xdtype)
        side_mark()
        if ctx.stage.after_backward is not None:
            ctx.stage.after_backward()
        return (dx if need[0] else None, dcond, dscond, dtoken, None, None, None, None,
                g_wqkv, g_bqkv, g_wout, g_bout, g_w1, g_b1, g_w2, g_b2,
                g_wa1, g_ba1, g_wg1, g_bg1, g_wa2, g_ba2, g_wg2, g_bg2)


class _MSE(torch.autograd.Function):
    """mean((pred - target)^2) in fp32 from bf16 operands, one pass each way (models/*.py: F.mse_loss(pred.float(),
    target.float()))."""

    @staticmethod
    def forward(ctx, pred, target):
        p16, t16 = _as_bf16(pred), _as_bf16(target)
        ctx.save_for_backward(p16, t16)
        ctx.meta = (pred.dtype, pred.shape)
        return ops.mse_fwd(p16, t16)

    @staticmethod
    def backward(ctx, dloss):
        p16, t16 = ctx.saved_tensors
        dtype, shape = ctx.meta
        d = ops.mse_bwd(p16, t16, dloss.to(f32).contiguous())
        return d.view(shape).to(dtype), None


def mse_loss(pred, target):
    """Fused MSE for bf16 predictions whose element count is a multiple of 8; anything else takes torch's kernel."""
    if pred.is_cuda and pred.dtype == bf16 and target.dtype == bf16 and pred.numel() % 8 == 0 and pred.shape == target.shape:
        return _MSE.apply(pred.contiguous(), target.detach().contiguous())
    return torch.nn.functional.mse_loss(pred.float(), target.float())


def rf_noise(x, z, t):
    """(lerpd, target) = (x (1 - t) + z t, z - x) with t broadcast over the trailing dims of x (one launch for bf16)."""
    inner = x.numel() // max(t.numel(), 1)
    if x.is_cuda and x.dtype == bf16 and z.dtype == bf16 and inner % 8 == 0 and x.shape[:t.dim()] == t.shape:
        return ops.rf_noise(x.contiguous(), z.contiguous(), t.to(bf16).contiguous())
    tt = t.view(*t.shape, *([1] * (x.dim() - t.dim())))
    return x * (1 - tt) + z * tt, z - x


def _fresh_or_sink(p: torch.Tensor, needed: bool):
    """(fp32 destination for a fused bias column-sum, accumulate flag, destination is a bucket sink)."""
    if not needed:
        return None, False, False
    sink = _sink(p)
    if sink is not None:
        return sink.view, sink.accumulate, True
    return torch.empty(p.shape[0], device=p.device, dtype=f32), False, False
