"""Tensor-level wrappers over the C ABI (include/owl_dit.h).

PyTorch is plumbing here: it owns device memory and the stream; every function below only passes
raw device pointers, sizes and the current CUDA stream to libowl_dit.so.  All outputs/workspaces
are allocated by the caller side (here) -- the library never allocates.
"""
from __future__ import annotations

import ctypes as C
import functools
from typing import Optional

import torch

from . import _lib
from ._lib import (OWL_EPI_BIAS_SILU, OWL_EPI_DSILU, OWL_EPI_F32, OWL_EPI_GATE_RESIDUAL, OWL_EPI_STORE,
                   OWL_GEMM_NN, OWL_GEMM_NT, OWL_GEMM_TN, AttnMask, GemmEpilogue)

bf16 = torch.bfloat16


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def _stream():
    """cudaStream_t of torch's current stream.  The raw accessors skip building a torch.cuda.Stream object per launch
    (a quarter of the host time of a launch-bound decode step)."""
    if _raw_stream is not None and _raw_device is not None:
        return _raw_stream(_raw_device())
    return torch.cuda.current_stream().cuda_stream


def _req(t: torch.Tensor, dtype=bf16):
    if not t.is_cuda:
        raise RuntimeError("owl ops need CUDA tensors: there is no CPU fallback")
    if t.dtype != dtype:
        raise RuntimeError(f"expected {dtype}, got {t.dtype}")


def _rows(t: torch.Tensor):
    """2-D view info of a tensor whose last dim is contiguous and leading dims are collapsible."""
    if t.stride(-1) != 1:
        raise RuntimeError("last dim must be contiguous")
    if t.dim() == 1:
        return 1, t.shape[0], t.shape[0]
    if t.dim() > 2:
        t = t.reshape(-1, t.shape[-1]) if t.is_contiguous() else t.flatten(0, -2)
    return t.shape[0], t.shape[1], t.stride(0)


def _ld(t: torch.Tensor) -> int:
    """Leading dimension of a 2-D operand (fast path of _rows)."""
    if t.dim() == 2 and t.stride(1) == 1:
        return t.stride(0)
    return _rows(t)[2]


_GEMM_FN = None


def _gemm_fn():
    global _GEMM_FN
    if _GEMM_FN is None:
        _GEMM_FN = _lib.load().owl_gemm_bf16
    return _GEMM_FN


# in-graph timeline probe (tools/step_trace.py): TRACE = (uint64 device buffer, [labels]) or None
TRACE = None


def stamp(label: str) -> None:
    if TRACE is None:
        return
    buf, labels = TRACE
    i = len(labels)
    if i >= buf.numel():
        return
    labels.append(label)
    _lib.check(_lib.load().owl_debug_timestamp(buf.data_ptr() + 8 * i, _stream()), "owl_debug_timestamp")


def launch_count() -> int:
    return int(_lib.load().owl_launch_count())


# ------------------------------------------------------------------------------------------------
# per-op CUDA-event timing (bench.py roofline / profiles): off unless profile_begin() was called
# ------------------------------------------------------------------------------------------------
_PROFILE = None


def profile_begin():
    global _PROFILE
    _PROFILE = []


def profile_end():
    """-> {kind: {"ms": total, "n": launches, "flops": total algorithmic flops (GEMMs only)}}"""
    global _PROFILE
    rec, _PROFILE = _PROFILE, None
    torch.cuda.synchronize()
    out = {}
    for kind, e0, e1, fl in rec or []:
        d = out.setdefault(kind, {"ms": 0.0, "n": 0, "flops": 0.0})
        d["ms"] += e0.elapsed_time(e1)
        d["n"] += 1
        d["flops"] += fl
    return out


def _prof(kind, flops=None):
    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*a, **k):
            if _PROFILE is None:
                return fn(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            kd = kind(*a, **k) if callable(kind) else kind
            _PROFILE.append((kd, e0, e1, flops(*a, **k) if flops else 0.0))
            return r
        return wrapped
    return deco


# ------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------
@_prof(lambda layout, *a, **k: ("gemm_nt", "gemm_nn", "gemm_tn")[layout],
       lambda layout, a, b, M, N, K, **k: 2.0 * M * N * K)
def gemm(layout: int, a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, *, mode: int = OWL_EPI_STORE,
         bias: Optional[torch.Tensor] = None, out0: Optional[torch.Tensor] = None,
         out1: Optional[torch.Tensor] = None, outf: Optional[torch.Tensor] = None,
         aux0: Optional[torch.Tensor] = None, aux1: Optional[torch.Tensor] = None, rows_per_frame: int = 1,
         accumulate: bool = False) -> None:
    """C[M,N] = opA . opB with a fused epilogue; a/b are 2-D bf16 (row stride = leading dimension)."""
    _req(a), _req(b)
    e = GemmEpilogue()
    e.mode, e.accumulate, e.rows_per_frame = mode, int(accumulate), rows_per_frame
    if bias is not None:
        _req(bias, torch.float32)
        e.bias = bias.data_ptr()
    # spelled out (no loop / setattr): this wrapper runs ~70 times per decode call, where host time is the bound
    if out0 is not None:
        _req(out0)
        e.out0, e.ld0 = out0.data_ptr(), _ld(out0)
    if out1 is not None:
        _req(out1)
        e.out1, e.ld1 = out1.data_ptr(), _ld(out1)
    if outf is not None:
        _req(outf, torch.float32)
        e.outf, e.ldf = outf.data_ptr(), _ld(outf)
    if aux0 is not None:
        _req(aux0)
        e.aux0, e.lda0 = aux0.data_ptr(), _ld(aux0)
    if aux1 is not None:
        _req(aux1)
        e.aux1, e.lda1 = aux1.data_ptr(), _ld(aux1)
    rc = _gemm_fn()(layout, M, N, K, a.data_ptr(), _ld(a), b.data_ptr(), _ld(b), C.byref(e), _stream())
    if rc != 0:
        _lib.check(rc, "owl_gemm_bf16")


_SMALLM_WS = {}   # device index -> zeroed workspace tensor (the kernel leaves it zero)


def _smallm_workspace(dev: torch.device, N: int) -> torch.Tensor:
    """Workspace of the split-K reduce, one per device: the decode-regime launches of a device are ordered on one
    stream (eager) or inside one graph; concurrent decode streams on the SAME device would need one workspace each.
    Grown (never shrunk) to the widest N seen; allocated outside any graph capture so its address is stable."""
    key = dev.index
    need = int(_lib.load().owl_gemm_smallm_workspace_bytes(N))
    ws = _SMALLM_WS.get(key)
    if ws is None or ws.numel() < need:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("owl: small-M GEMM workspace must exist before graph capture (run one eager call first)")
        ws = torch.zeros(max(need, 1 << 22), device=dev, dtype=torch.uint8)
        _SMALLM_WS[key] = ws
    return ws


@_prof("gemm_smallm", lambda a, w, M, N, K, **k: 2.0 * M * N * K)
def gemm_smallm(a: torch.Tensor, w: torch.Tensor, M: int, N: int, K: int, *, mode: int = OWL_EPI_STORE,
                bias: Optional[torch.Tensor] = None, out0: Optional[torch.Tensor] = None,
                aux0: Optional[torch.Tensor] = None, aux1: Optional[torch.Tensor] = None, rows_per_frame: int = 1,
                next_w: Optional[torch.Tensor] = None, adaln=None) -> None:
    """out0[M <= 128, N] = epi(a[M, K] . w[N, K]^T + bias): the decode regime's weight-streaming GEMM (owl_gemm_smallm).
    `next_w`: the (contiguous) weight matrix the next Linear of the schedule reads, prefetched into L2 meanwhile.
    `adaln` = (a, b, T, eps): `a` is the raw residual stream and the GEMM multiplies rms_norm(a)*(1+a_mod[frame])+b_mod[frame]."""
    _req(a), _req(w), _req(out0)
    e = GemmEpilogue()
    e.mode, e.rows_per_frame = mode, rows_per_frame
    if bias is not None:
        _req(bias, torch.float32)
        e.bias = bias.data_ptr()
    e.out0, e.ld0 = out0.data_ptr(), _ld(out0)
    if aux0 is not None:
        e.aux0, e.lda0 = aux0.data_ptr(), _ld(aux0)
    if aux1 is not None:
        e.aux1, e.lda1 = aux1.data_ptr(), _ld(aux1)
    ws = _smallm_workspace(a.device, N)
    nw, nb = (next_w.data_ptr(), next_w.numel() * next_w.element_size()) if next_w is not None else (None, 0)
    ln = None
    if adaln is not None:
        am, bm, T, eps = adaln
        _req(am), _req(bm)
        ln = _lib.AdaLNPrologue()
        ln.a, ln.b, ln.ld_ab, ln.rows_per_frame, ln.eps = am.data_ptr(), bm.data_ptr(), _ld(am), T, eps
        ln = C.byref(ln)
    rc = _lib.load().owl_gemm_smallm(M, N, K, a.data_ptr(), _ld(a), w.data_ptr(), _ld(w), C.byref(e), ln, ws.data_ptr(),
                                     ws.numel(), nw, nb, _stream())
    if rc != 0:
        _lib.check(rc, "owl_gemm_smallm")


def gemm_check(layout: int, a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int) -> torch.Tensor:
    lib = _lib.load()
    c = torch.empty(M, N, device=a.device, dtype=torch.float32)
    rc = lib.owl_gemm_bf16_check(layout, M, N, K, a.data_ptr(), _rows(a)[2], b.data_ptr(), _rows(b)[2], c.data_ptr(), N,
                                 _stream())
    _lib.check(rc, "owl_gemm_bf16_check")
    return c


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = x w^T + bias, x [..., K] bf16, w [N, K] bf16, bias fp32."""
    M, K, _ = _rows(x)
    N = w.shape[0]
    y = torch.empty(*x.shape[:-1], N, device=x.device, dtype=bf16)
    gemm(OWL_GEMM_NT, x, w, M, N, K, mode=OWL_EPI_STORE, bias=bias, out0=y)
    return y


# ------------------------------------------------------------------------------------------------
# memory-bound ops
# ------------------------------------------------------------------------------------------------
@_prof("cast_f32_bf16")
def cast_f32_bf16(src: torch.Tensor, dst: Optional[torch.Tensor] = None) -> torch.Tensor:
    _req(src, torch.float32)
    if dst is None:
        dst = torch.empty(src.shape, device=src.device, dtype=bf16)
    _lib.check(_lib.load().owl_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()), "owl_cast_f32_bf16")
    return dst


@_prof("cast_f32_bf16")
def cast_f32_bf16_multi(pairs) -> None:
    """[(src fp32, dst bf16), ...] contiguous tensors of equal numel per pair, numel % 8 == 0: one launch per 8."""
    n = len(pairs)
    srcs = (C.c_void_p * n)(*[s.data_ptr() for s, _ in pairs])
    dsts = (C.c_void_p * n)(*[d.data_ptr() for _, d in pairs])
    ns = (C.c_int64 * n)(*[s.numel() for s, _ in pairs])
    _lib.check(_lib.load().owl_cast_f32_bf16_multi(srcs, dsts, ns, n, _stream()), "owl_cast_f32_bf16_multi")


@_prof("silu")
def silu(x: torch.Tensor) -> torch.Tensor:
    _req(x)
    y = torch.empty_like(x)
    _lib.check(_lib.load().owl_silu_bf16(x.data_ptr(), y.data_ptr(), x.numel(), _stream()), "owl_silu_bf16")
    return y


@_prof("sincos_embed")
def sincos_embed(t: torch.Tensor, freqs: torch.Tensor, mult: float) -> torch.Tensor:
    """t fp32 [n] contiguous, freqs fp32 [half] -> bf16 [n, 2*half] = [sin | cos] (embeddings.py:30-72)."""
    _req(t, torch.float32)
    _req(freqs, torch.float32)
    n, half = t.numel(), freqs.numel()
    out = torch.empty(n, 2 * half, device=t.device, dtype=torch.bfloat16)
    _lib.check(_lib.load().owl_sincos_embed(t.data_ptr(), freqs.data_ptr(), out.data_ptr(), n, half, float(mult), _stream()),
               "owl_sincos_embed")
    return out


@_prof("dsilu_mul")
def dsilu_mul(dy: torch.Tensor, z: torch.Tensor, dz: torch.Tensor) -> torch.Tensor:
    _req(dy), _req(z), _req(dz)
    _lib.check(_lib.load().owl_dsilu_mul_bf16(dy.data_ptr(), z.data_ptr(), dz.data_ptr(), z.numel(), _stream()),
               "owl_dsilu_mul_bf16")
    return dz


@_prof("adaln_modulate_fwd")
def adaln_modulate_fwd(x, a, b, T: int, eps: float, want_rstd: bool = True):
    """h = rms_norm(x) * (1 + a[frame]) + b[frame]; x [M,d]; a,b [M/T,d] column slices (or None)."""
    _req(x)
    M, d, _ = _rows(x)
    h = torch.empty_like(x)
    rstd = torch.empty(M, device=x.device, dtype=torch.float32) if want_rstd else None
    ld_ab = _rows(a)[2] if a is not None else 0
    rc = _lib.load().owl_adaln_modulate_fwd(x.data_ptr(), _p(a), _p(b), ld_ab, h.data_ptr(), _p(rstd), M, d, T, eps,
                                            _stream())
    _lib.check(rc, "owl_adaln_modulate_fwd")
    return h, rstd


@_prof("residual_adaln_fwd")
def residual_adaln_fwd(x, y, gate, a, b, T: int, eps: float, want_rstd: bool = True):
    """xsum = x + gate[frame]*y ; h = rms_norm(xsum)*(1+a[frame])+b[frame].  Returns (xsum, h, rstd)."""
    _req(x), _req(y)
    M, d, _ = _rows(x)
    xsum, h = torch.empty_like(x), torch.empty_like(x)
    rstd = torch.empty(M, device=x.device, dtype=torch.float32) if want_rstd else None
    rc = _lib.load().owl_residual_adaln_fwd(x.data_ptr(), y.data_ptr(), gate.data_ptr(), _rows(gate)[2], _p(a), _p(b),
                                            _rows(a)[2] if a is not None else 0, xsum.data_ptr(), h.data_ptr(),
                                            _p(rstd), M, d, T, eps, _stream())
    _lib.check(rc, "owl_residual_adaln_fwd")
    return xsum, h, rstd


@_prof("adaln_modulate_bwd")
def adaln_modulate_bwd(dh, x, a, rstd, dres, da, db, T: int):
    """Returns dx; writes da, db (views with a row stride) in place."""
    M, d, _ = _rows(x)
    dx = torch.empty_like(x)
    ld_ab = _rows(a)[2] if a is not None else 0
    ld_dab = _rows(da)[2] if da is not None else 0
    rc = _lib.load().owl_adaln_modulate_bwd(dh.data_ptr(), x.data_ptr(), _p(a), ld_ab, rstd.data_ptr(), _p(dres),
                                            dx.data_ptr(), _p(da), _p(db), ld_dab, M, d, T, _stream())
    _lib.check(rc, "owl_adaln_modulate_bwd")
    return dx


@_prof("adaln_gate_bwd")
def adaln_gate_bwd(dh, x, a, rstd, dres, da, db, y, c, dc, dbias, T: int, dbias_accumulate: bool = False):
    """adaLN backward + gate backward of the producing branch in one pass -> (dx, dy); da/db/dc written in place,
    dbias (+)= colsum(dy).  See owl_adaln_gate_bwd."""
    M, d, _ = _rows(x)
    dx, dy = torch.empty_like(x), torch.empty_like(x)
    rc = _lib.load().owl_adaln_gate_bwd(dh.data_ptr(), x.data_ptr(), _p(a), _rows(a)[2] if a is not None else 0,
                                        rstd.data_ptr(), _p(dres), dx.data_ptr(), _p(da), _p(db),
                                        _rows(da)[2] if da is not None else 0, y.data_ptr(), c.data_ptr(), _rows(c)[2],
                                        dy.data_ptr(), _p(dc), _rows(dc)[2] if dc is not None else 0, _p(dbias),
                                        int(dbias_accumulate), M, d, T, _stream())
    _lib.check(rc, "owl_adaln_gate_bwd")
    return dx, dy


@_prof("gate_fwd")
def gate_fwd(x, c, res, T: int):
    M, d, _ = _rows(x)
    out = torch.empty_like(x)
    rc = _lib.load().owl_gate_fwd(x.data_ptr(), c.data_ptr(), _rows(c)[2], _p(res), out.data_ptr(), M, d, T, _stream())
    _lib.check(rc, "owl_gate_fwd")
    return out


@_prof("gate_bwd")
def gate_bwd(dx, y, c, dc, dbias, T: int, dbias_accumulate: bool = False):
    """dy = dx * c[frame]; dc[frame] = sum dx*y (written in place); dbias (+)= colsum(dy)."""
    M, d, _ = _rows(dx)
    dy = torch.empty_like(dx)
    rc = _lib.load().owl_gate_bwd(dx.data_ptr(), _p(y), c.data_ptr(), _rows(c)[2], dy.data_ptr(), _p(dc),
                                  _rows(dc)[2] if dc is not None else 0, _p(dbias), int(dbias_accumulate), M, d, T,
                                  _stream())
    _lib.check(rc, "owl_gate_bwd")
    return dy


@_prof("colsum")
def colsum(x, out, accumulate: bool = False):
    M, N, ld = _rows(x)
    _lib.check(_lib.load().owl_colsum_bf16(x.data_ptr(), ld, out.data_ptr(), int(accumulate), M, N, _stream()),
               "owl_colsum_bf16")
    return out


@_prof("colsum")
def colsum_multi(items) -> None:
    """[(x bf16 [M, N], out fp32 [N]), ...] (at most 4): out += column sums of x, one launch.  Outputs must be zeroed /
    hold what is to be accumulated onto."""
    n = len(items)
    info = [_rows(x) for x, _ in items]
    xs = (C.c_void_p * n)(*[x.data_ptr() for x, _ in items])
    outs = (C.c_void_p * n)(*[o.data_ptr() for _, o in items])
    lds = (C.c_int64 * n)(*[i[2] for i in info])
    Ms = (C.c_int64 * n)(*[i[0] for i in info])
    Ns = (C.c_int64 * n)(*[i[1] for i in info])
    _lib.check(_lib.load().owl_colsum_bf16_multi(xs, lds, Ms, Ns, outs, n, _stream()), "owl_colsum_bf16_multi")


@_prof("layer_norm")
def layer_norm(x, eps: float = 1e-5):
    M, d, _ = _rows(x)
    y = torch.empty_like(x)
    _lib.check(_lib.load().owl_layer_norm_fwd(x.data_ptr(), y.data_ptr(), M, d, eps, _stream()), "owl_layer_norm_fwd")
    return y


@_prof("layer_norm_bwd")
def layer_norm_bwd(dy, x, eps: float = 1e-5):
    M, d, _ = _rows(x)
    dx = torch.empty_like(x)
    _lib.check(_lib.load().owl_layer_norm_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), M, d, eps, _stream()),
               "owl_layer_norm_bwd")
    return dx


@_prof("qk_norm_rope_fwd")
def qk_norm_rope_fwd(qkv, cos, sin, pos_offset: int, H: int, eps: float, q, k, v, want_rstd: bool = True):
    """qkv [B,L,3*H*dh] -> q,k,v written into caller views [B,H,L,dh] (dh contiguous, L stride dh)."""
    B, L, three_d = qkv.shape
    dh = three_d // (3 * H)
    _req(cos, torch.float32), _req(sin, torch.float32)      # a bf16 table read as fp32 is an out-of-bounds read
    if cos.shape[0] < pos_offset + L:
        raise RuntimeError(f"RoPE table has {cos.shape[0]} positions; need {pos_offset + L}")
    for t in (q, k, v):
        assert t.stride(3) == 1 and t.stride(2) == dh, "head-major views need contiguous [L, dh] rows"
    rq = torch.empty(B, H, L, device=qkv.device, dtype=torch.float32) if want_rstd else None
    rk = torch.empty(B, H, L, device=qkv.device, dtype=torch.float32) if want_rstd else None
    rc = _lib.load().owl_qk_norm_rope_fwd(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), pos_offset, q.data_ptr(),
                                          q.stride(0), q.stride(1), k.data_ptr(), k.stride(0), k.stride(1),
                                          v.data_ptr(), v.stride(0), v.stride(1), _p(rq), _p(rk), B, L, H, dh, eps,
                                          _stream())
    _lib.check(rc, "owl_qk_norm_rope_fwd")
    return rq, rk


@_prof("qk_norm_rope_bwd")
def qk_norm_rope_bwd(dq, dk, dv, q, k, rq, rk, cos, sin, pos_offset: int):
    """dq/dk/dv share strides ([B,H,L,dh]); returns dqkv [B,L,3*H*dh]."""
    B, H, L, dh = dq.shape
    _req(cos, torch.float32), _req(sin, torch.float32)
    assert dq.stride() == dk.stride() == dv.stride()
    dqkv = torch.empty(B, L, 3 * H * dh, device=dq.device, dtype=bf16)
    rc = _lib.load().owl_qk_norm_rope_bwd(dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dq.stride(0), dq.stride(1),
                                          q.data_ptr(), q.stride(0), q.stride(1), k.data_ptr(), k.stride(0),
                                          k.stride(1), rq.data_ptr(), rk.data_ptr(), cos.data_ptr(), sin.data_ptr(),
                                          pos_offset, dqkv.data_ptr(), B, L, H, dh, _stream())
    _lib.check(rc, "owl_qk_norm_rope_bwd")
    return dqkv


@_prof("rope_apply")
def rope_apply(x, cos, sin, pos_offset: int, inverse: bool = False):
    """x contiguous bf16 [..., L, dh] -> rotated copy (see owl_rope_apply)."""
    _req(x)
    _req(cos, torch.float32), _req(sin, torch.float32)
    L, dh = x.shape[-2], x.shape[-1]
    y = torch.empty_like(x)
    rc = _lib.load().owl_rope_apply(x.data_ptr(), cos.data_ptr(), sin.data_ptr(), pos_offset, y.data_ptr(),
                                    x.numel() // (L * dh), L, dh, int(inverse), _stream())
    _lib.check(rc, "owl_rope_apply")
    return y


@_prof("kv_append")
def kv_append(src, cache, pos: int):
    B, H, Ln, dh = src.shape
    rc = _lib.load().owl_kv_append(src.data_ptr(), src.stride(0), src.stride(1), cache.data_ptr(), cache.stride(0),
                                   cache.stride(1), pos, B, H, Ln, dh, _stream())
    _lib.check(rc, "owl_kv_append")


# ------------------------------------------------------------------------------------------------
# attention
# ------------------------------------------------------------------------------------------------
def _mask(tokens_per_frame: int, window_frames: int, q_offset: int, causal: bool, doc_id, doc_ext=None) -> AttnMask:
    m = AttnMask()
    m.tokens_per_frame, m.window_frames, m.q_offset, m.causal = tokens_per_frame, window_frames, q_offset, int(causal)
    if doc_id is not None:
        _req(doc_id, torch.int32)
        m.doc_id, m.doc_ld = doc_id.data_ptr(), doc_id.stride(0)
        if doc_ext is not None:
            first, last, breaks = doc_ext
            if first.stride(0) != doc_id.stride(0) or last.stride(0) != doc_id.stride(0):
                raise RuntimeError("document extents must share doc_id's row stride")
            m.doc_first, m.doc_last, m.doc_breaks = first.data_ptr(), last.data_ptr(), breaks.data_ptr()
    return m


@_prof("attn_doc_extents")
def attn_doc_extents(doc_id: torch.Tensor):
    """doc_id int32 [B, F] -> (first, last int32 [B, F], breaks int32 [B]); see include/owl_dit.h."""
    _req(doc_id, torch.int32)
    if doc_id.dim() != 2 or doc_id.stride(1) != 1:
        raise RuntimeError("doc_id must be [B, F] with contiguous rows")
    B, F = doc_id.shape
    ld = doc_id.stride(0)
    first = torch.empty_strided((B, F), (ld, 1), device=doc_id.device, dtype=torch.int32)
    last = torch.empty_strided((B, F), (ld, 1), device=doc_id.device, dtype=torch.int32)
    breaks = torch.empty(B, device=doc_id.device, dtype=torch.int32)
    rc = _lib.load().owl_attn_doc_extents(doc_id.data_ptr(), ld, B, F, first.data_ptr(), last.data_ptr(),
                                          breaks.data_ptr(), _stream())
    _lib.check(rc, "owl_attn_doc_extents")
    return first, last, breaks


def _hm(t: torch.Tensor):
    """[B,H,L,dh] head-major view: dh contiguous, row stride dh; returns (ptr, batch stride, head stride)."""
    _req(t)
    if t.stride(3) != 1 or (t.shape[2] > 1 and t.stride(2) != t.shape[3]):
        raise RuntimeError("head-major tensors need contiguous [L, dh] rows")
    return t.data_ptr(), t.stride(0), t.stride(1)


@_prof("attn_fwd")
def attn_fwd(q, k, v, *, tokens_per_frame: int, window_frames: int = 0, q_offset: int = 0, causal: bool = True,
             doc_id=None, doc_ext=None, want_lse: bool = True, check: bool = False):
    """q [B,H,Lq,dh], k/v [B,H,Lk,dh] -> o [B,Lq,H*dh] (token-major, ready for the out Linear), lse [B,H,Lq]."""
    lib = _lib.load()
    B, H, Lq, dh = q.shape
    Lk = k.shape[2]
    o = torch.empty(B, Lq, H * dh, device=q.device, dtype=bf16)
    lse = torch.empty(B, H, Lq, device=q.device, dtype=torch.float32) if want_lse else None
    m = _mask(tokens_per_frame, window_frames, q_offset, causal, doc_id, doc_ext)
    fn = lib.owl_attn_fwd_check if check else lib.owl_attn_fwd
    rc = fn(*_hm(q), *_hm(k), *_hm(v), o.data_ptr(), o.stride(0), o.stride(1), _p(lse), B, H, Lq, Lk, dh, C.byref(m),
            _stream())
    _lib.check(rc, "owl_attn_fwd")
    return o, lse


@_prof("attn_bwd")
def attn_bwd(dout, o, q, k, v, lse, *, tokens_per_frame: int, window_frames: int = 0, q_offset: int = 0,
             causal: bool = True, doc_id=None, doc_ext=None, check: bool = False, dq=None, dk=None, dv=None):
    """Returns (dq, dk, dv) in [B,H,L,dh]."""
    lib = _lib.load()
    B, H, Lq, dh = q.shape
    Lk = k.shape[2]
    _req(dout), _req(o)
    if dout.stride() != o.stride():
        dout = dout.contiguous()
    dq = torch.empty(B, H, Lq, dh, device=q.device, dtype=bf16) if dq is None else dq
    dk = torch.empty(B, H, Lk, dh, device=q.device, dtype=bf16) if dk is None else dk
    dv = torch.empty(B, H, Lk, dh, device=q.device, dtype=bf16) if dv is None else dv
    delta = torch.empty(B, H, Lq, device=q.device, dtype=torch.float32)
    dq_acc = None if check else torch.empty(attn_bwd_workspace_elems(B, H, Lq, dh), device=q.device, dtype=torch.float32)
    m = _mask(tokens_per_frame, window_frames, q_offset, causal, doc_id, doc_ext)
    fn = lib.owl_attn_bwd_check if check else lib.owl_attn_bwd
    rc = fn(dout.data_ptr(), o.data_ptr(), o.stride(0), o.stride(1), *_hm(q), *_hm(k), *_hm(v), lse.data_ptr(),
            delta.data_ptr(), _p(dq_acc), *_hm(dq), *_hm(dk), *_hm(dv), B, H, Lq, Lk, dh, C.byref(m), _stream())
    _lib.check(rc, "owl_attn_bwd")
    return dq, dk, dv


def attn_bwd_workspace_elems(B: int, H: int, Lq: int, dh: int) -> int:
    """fp32 elements of the dq accumulator workspace owl_attn_bwd needs (query rows padded to 128)."""
    return B * H * ((Lq + 127) // 128 * 128) * dh


# ------------------------------------------------------------------------------------------------
# decode regime: ring KV cache with device-resident state (owl_qk_norm_rope_decode / owl_attn_decode)
# ------------------------------------------------------------------------------------------------
@_prof("qk_norm_rope_decode")
def qk_norm_rope_decode(qkv, cos, sin, state, q, k_cache, v_cache, H: int, eps: float):
    """qkv [B, L, 3*H*dh] -> q [B, H, L, dh]; new K / V rows -> ring slots of k_cache / v_cache [B, H, cap, dh]
    behind the cached rows; slot and RoPE position come from the device `state` (int32[4])."""
    B, L, three_d = qkv.shape
    dh = three_d // (3 * H)
    _req(qkv), _req(q), _req(k_cache), _req(v_cache), _req(state, torch.int32)
    _req(cos, torch.float32), _req(sin, torch.float32)
    if k_cache.stride() != v_cache.stride() or k_cache.stride(3) != 1 or k_cache.stride(2) != dh:
        raise RuntimeError("ring cache buffers must be [B, H, cap, dh] with contiguous rows and equal strides")
    rc = _lib.load().owl_qk_norm_rope_decode(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), cos.shape[0], state.data_ptr(),
                                             q.data_ptr(), q.stride(0), q.stride(1), k_cache.data_ptr(), v_cache.data_ptr(),
                                             k_cache.stride(0), k_cache.stride(1), k_cache.shape[2], B, L, H, dh, eps,
                                             _stream())
    _lib.check(rc, "owl_qk_norm_rope_decode")


_DECODE_WS = {}


def _decode_workspace(dev, B: int, H: int, cap: int, dh: int) -> torch.Tensor:
    key = dev.index
    need = int(_lib.load().owl_attn_decode_workspace_bytes(B, H, cap, dh))
    ws = _DECODE_WS.get(key)
    if ws is None or ws.numel() < need:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("owl: decode attention workspace must exist before graph capture (run one eager call first)")
        ws = torch.zeros(need, device=dev, dtype=torch.uint8)
        _DECODE_WS[key] = ws
    return ws


@_prof("attn_decode")
def attn_decode(q, k_cache, v_cache, state, vis_rows: int = 0, out=None):
    """q [B, H, Lq <= 128, dh] against the ring cache (cached rows + the Lq rows just written), no mask ->
    o [B, Lq, H*dh].  vis_rows > 0: only the last vis_rows rows of [cache, new] (local layers)."""
    B, H, Lq, dh = q.shape
    o = torch.empty(B, Lq, H * dh, device=q.device, dtype=bf16) if out is None else out
    cap = k_cache.shape[2]
    ws = _decode_workspace(q.device, B, H, cap, dh)
    qp, q_sb, q_sh = _hm(q)
    rc = _lib.load().owl_attn_decode(qp, q_sb, q_sh, k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
                                     k_cache.stride(1), cap, state.data_ptr(), vis_rows, o.data_ptr(), o.stride(0),
                                     o.stride(1), B, H, Lq, dh, ws.data_ptr(), ws.numel(), _stream())
    _lib.check(rc, "owl_attn_decode")
    return o


@_prof("attn_decode_fused")
def attn_decode_fused(qkv, cos, sin, k_cache, v_cache, state, H: int, eps: float, vis_rows: int = 0, out=None):
    """qkv [B, Lq <= 128, 3*H*dh] straight from the qkv GEMM: QK-norm + RoPE + ring append + split-KV attention in one
    launch -> o [B, Lq, H*dh] (owl_attn_decode_fused)."""
    B, Lq, three_d = qkv.shape
    dh = three_d // (3 * H)
    _req(qkv), _req(k_cache), _req(v_cache), _req(state, torch.int32)
    _req(cos, torch.float32), _req(sin, torch.float32)
    if k_cache.stride() != v_cache.stride() or k_cache.stride(3) != 1 or k_cache.stride(2) != dh or not qkv.is_contiguous():
        raise RuntimeError("ring cache buffers must be [B, H, cap, dh] with contiguous rows and equal strides; qkv contiguous")
    o = torch.empty(B, Lq, H * dh, device=qkv.device, dtype=bf16) if out is None else out
    cap = k_cache.shape[2]
    ws = _decode_workspace(qkv.device, B, H, cap, dh)
    rc = _lib.load().owl_attn_decode_fused(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), cos.shape[0], k_cache.data_ptr(),
                                           v_cache.data_ptr(), k_cache.stride(0), k_cache.stride(1), cap, state.data_ptr(),
                                           vis_rows, o.data_ptr(), o.stride(0), o.stride(1), B, H, Lq, dh, eps, ws.data_ptr(),
                                           ws.numel(), _stream())
    _lib.check(rc, "owl_attn_decode_fused")
    return o


def decode_state_advance(state, commit_rows: int, drop_rows: int, cap: int) -> None:
    _lib.check(_lib.load().owl_decode_state_advance(state.data_ptr(), commit_rows, drop_rows, cap, _stream()),
               "owl_decode_state_advance")


# ------------------------------------------------------------------------------------------------
# rectified-flow glue (owl_rf_noise / owl_mse_fwd / owl_mse_bwd)
# ------------------------------------------------------------------------------------------------
@_prof("rf_noise")
def rf_noise(x, z, t):
    """x, z bf16 [rows..., inner...] contiguous, t bf16 with one value per leading row -> (lerpd, target) bf16."""
    _req(x), _req(z), _req(t)
    rows = t.numel()
    inner = x.numel() // rows
    lerpd, target = torch.empty_like(x), torch.empty_like(x)
    rc = _lib.load().owl_rf_noise(x.data_ptr(), z.data_ptr(), t.data_ptr(), lerpd.data_ptr(), target.data_ptr(), rows, inner,
                                  _stream())
    _lib.check(rc, "owl_rf_noise")
    return lerpd, target


@_prof("mse_fwd")
def mse_fwd(pred, target):
    _req(pred), _req(target)
    loss = torch.empty((), device=pred.device, dtype=torch.float32)
    _lib.check(_lib.load().owl_mse_fwd(pred.data_ptr(), target.data_ptr(), loss.data_ptr(), pred.numel(), _stream()),
               "owl_mse_fwd")
    return loss


@_prof("mse_bwd")
def mse_bwd(pred, target, dloss):
    dpred = torch.empty_like(pred)
    _lib.check(_lib.load().owl_mse_bwd(pred.data_ptr(), target.data_ptr(), dloss.data_ptr(), dpred.data_ptr(), pred.numel(),
                                       _stream()), "owl_mse_bwd")
    return dpred


def shard_reduce(dst, srcs, n_src: int, scale: float, max_ctas: int = 32) -> None:
    """dst (fp32 [n]) <- (dst + sum of the first n_src rows of srcs (fp32 [>= n_src, stride])) * scale."""
    _req(dst, torch.float32)
    stride = srcs.stride(0) if n_src > 0 else 0
    rc = _lib.load().owl_shard_reduce(dst.data_ptr(), _p(srcs) if n_src > 0 else None, stride, n_src, dst.numel(), float(scale),
                                      max_ctas, _stream())
    _lib.check(rc, "owl_shard_reduce")


def shard_reduce_peers(dst, srcs, scale: float, max_ctas: int = 0) -> None:
    """dst (fp32 [n]) <- (dst + sum of the tensors in `srcs` (fp32 [n] each, e.g. peer-mapped bucket slices)) * scale."""
    _req(dst, torch.float32)
    n = len(srcs)
    arr = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in srcs])
    rc = _lib.load().owl_shard_reduce_peers(dst.data_ptr(), arr, n, dst.numel(), float(scale), max_ctas, _stream())
    _lib.check(rc, "owl_shard_reduce_peers")


def mc_allreduce(mc_ptr: int, n: int, scale: float, max_ctas: int = 0) -> None:
    """All-reduce (sum * scale) of n fp32 at the MULTICAST address `mc_ptr` (this rank's shard of a symmetric bucket),
    reduced inside NVSwitch.  The caller orders it between two cross-rank barriers."""
    rc = _lib.load().owl_mc_allreduce(C.c_void_p(int(mc_ptr)), int(n), float(scale), max_ctas, _stream())
    _lib.check(rc, "owl_mc_allreduce")
