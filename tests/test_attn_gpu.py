"""Attention kernels against torch SDPA with the dense mask of attn.py:43-59 (fp32 math on the same
bf16 inputs).  Covers T=1/64/65, windows, q_offset (KV prefix), non-causal, doc ids, dh 64/128.
Bar: max|out-ref|/max|ref| <= 1e-2 forward, 2e-2 gradients (north_star tolerance)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
bf16 = torch.bfloat16


def _ops():
    from owl_audio_exps_b200 import ops
    return ops


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def dense_mask(Lq, Lk, T, W, q_offset, causal, doc, dev):
    qi = torch.arange(Lq, device=dev)[:, None] + q_offset
    kj = torch.arange(Lk, device=dev)[None, :]
    fq, fk = qi // T, kj // T
    m = torch.ones(Lq, Lk, dtype=torch.bool, device=dev)
    if causal:
        m &= fk <= fq
    if W > 0:
        m &= (fq - fk).abs() < W
    m = m[None].clone()
    if doc is not None:
        m = m & (doc[:, fq.squeeze(1)][:, :, None] == doc[:, fk.squeeze(0)][:, None, :])
    return m[:, None]  # [B or 1, 1, Lq, Lk]


CASES = [
    # B, H, Lq, Lk, dh, T, W, causal, doc
    (2, 4, 120, 120, 64, 1, 16, True, False),
    (2, 4, 120, 120, 64, 1, 0, True, False),
    (1, 3, 1040, 1040, 64, 65, 0, True, False),
    (1, 3, 1040, 1040, 64, 65, 4, True, False),
    (1, 2, 2048, 2048, 64, 64, 8, True, True),
    (1, 2, 1024, 1024, 128, 64, 0, True, False),
    (1, 2, 65, 65 * 7, 64, 65, 0, False, False),       # decode mode: block_mask None, all keys visible
    (1, 2, 130, 65 * 5, 64, 65, 3, True, False),       # cached, masked path with q_offset
    (2, 2, 520, 520, 64, 65, 3, False, False),         # bidirectional window
    (1, 2, 300, 300, 128, 1, 16, True, False),
    # two-tile persistent forward (attn_fwd2.cu): odd tile counts, ragged tails, dh 128, prefix, many work items per CTA
    (1, 2, 384, 384, 64, 64, 0, True, False),          # 3 query tiles: the last pair has only tile 0
    (1, 1, 129, 129, 64, 1, 0, True, False),           # tile 1 holds a single row
    (2, 3, 1100, 1100, 128, 1, 5, True, False),        # narrow window: tile 0 / tile 1 use different KV ranges
    (1, 2, 700, 1470, 64, 70, 0, True, False),         # q_offset = 770 tokens of KV prefix, T = 70
    (3, 20, 1280, 1280, 64, 64, 0, True, False),       # 300 work items > 148 CTAs: persistence, barrier phases
    (1, 2, 1040, 1040, 128, 65, 0, True, False),
    # document masks with extents (tile skipping): contiguous runs, and ids that re-open (bounding range + id compare)
    (1, 2, 2048, 2048, 64, 64, 8, True, "ext"),
    (1, 2, 4096, 4096, 64, 64, 0, True, "ext"),
    (1, 2, 4096, 4096, 64, 64, 0, True, "ext_nc"),
    (2, 2, 1536, 1536, 128, 64, 0, False, "ext"),      # bidirectional + docs
    (1, 2, 8192, 8192, 64, 64, 0, True, "ext"),        # parity at L >= 8192 with doc ids
]


def _inputs(case, dev):
    B, H, Lq, Lk, dh, T, W, causal, use_doc = case
    g = torch.Generator(device="cpu").manual_seed(Lq + Lk + dh)
    q = torch.randn(B, H, Lq, dh, generator=g).to(bf16).to(dev)
    k = torch.randn(B, H, Lk, dh, generator=g).to(bf16).to(dev)
    v = torch.randn(B, H, Lk, dh, generator=g).to(bf16).to(dev)
    doc = None
    if use_doc:
        nF = (Lk + T - 1) // T
        doc = (torch.arange(nF)[None] // 11).repeat(B, 1)
        if use_doc == "ext_nc":
            doc = doc % 3                      # ids 0,1,2,0,1,2,...: every id re-opens (not contiguous runs)
        if B > 1:
            doc[1] = (torch.arange(nF) // 7)   # different packing per batch row
        doc = doc.to(torch.int32).to(dev)
    return q, k, v, doc


def _reference(q, k, v, mask, dout):
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    s = (qf @ kf.transpose(-1, -2)) / math.sqrt(q.shape[-1])
    s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, -1)
    o = p @ vf
    lse = torch.logsumexp(s, -1)
    o_tm = o.permute(0, 2, 1, 3).reshape(q.shape[0], q.shape[2], -1)
    o_tm.backward(dout.float())
    return o_tm, lse, qf.grad, kf.grad, vf.grad


def _run(case, impl, want_bwd):
    ops = _ops()
    dev = torch.device("cuda:0")
    B, H, Lq, Lk, dh, T, W, causal, use_doc = case
    q_offset = Lk - Lq
    q, k, v, doc = _inputs(case, dev)
    mask = dense_mask(Lq, Lk, T, W, q_offset, causal, doc, dev)
    g = torch.Generator(device="cpu").manual_seed(5)
    dout = torch.randn(B, Lq, H * dh, generator=g).to(bf16).to(dev)
    o_ref, lse_ref, dq_ref, dk_ref, dv_ref = _reference(q, k, v, mask, dout)
    kw = dict(tokens_per_frame=T, window_frames=W, q_offset=q_offset, causal=causal, doc_id=doc, check=impl == "check")
    if isinstance(use_doc, str):
        first, last, breaks = ops.attn_doc_extents(doc)
        kw["doc_ext"] = (first, last, breaks)
        d = doc.cpu()
        for b in range(B):       # the extents kernel against a plain scan
            ids = d[b].tolist()
            assert first[b].cpu().tolist() == [ids.index(x) for x in ids]
            assert last[b].cpu().tolist() == [len(ids) - 1 - ids[::-1].index(x) for x in ids]
            nb = sum(1 for f in range(1, len(ids)) if ids[f] != ids[f - 1] and ids.index(ids[f]) != f)
            assert int(breaks[b]) == nb
        assert (int(breaks[0]) == 0) == (use_doc == "ext")
    o, lse = ops.attn_fwd(q, k, v, **kw)
    torch.cuda.synchronize()
    assert relerr(o, o_ref) <= 1e-2
    assert (lse - lse_ref).abs().max().item() <= 2e-2
    if not want_bwd:
        return
    dq, dk, dv = ops.attn_bwd(dout, o, q, k, v, lse, **kw)
    torch.cuda.synchronize()
    assert relerr(dq, dq_ref) <= 2e-2
    assert relerr(dk, dk_ref) <= 2e-2
    assert relerr(dv, dv_ref) <= 2e-2


@pytest.mark.parametrize("impl", ["check", "tcgen05"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(str(x) for x in c))
def test_attn_fwd(case, impl):
    _run(case, impl, False)


@pytest.mark.parametrize("impl", ["check", "tcgen05"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(str(x) for x in c))
def test_attn_bwd(case, impl):
    _run(case, impl, True)
