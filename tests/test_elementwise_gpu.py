"""Memory-bound kernels (adaLN modulate, gate, colsum, qk-norm+rope, kv append, layer norm, cast, silu)
against plain torch fp32 math on the same bf16 inputs.  Bar: max|out-ref|/max|ref| <= 1e-2 (one bf16
rounding of an fp32 result is 2^-9 relative; the reference itself rounds after every op)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
bf16 = torch.bfloat16
EPS = 1.1920928955078125e-07


def _ops():
    from owl_audio_exps_b200 import ops
    return ops


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def rms(x, eps=EPS):
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps)


@pytest.mark.parametrize("M,d,T", [(3840, 1024, 1), (1040, 1536, 65), (4096, 1536, 64), (128, 2560, 64), (7, 64, 1)])
def test_adaln_fwd_bwd(M, d, T):
    ops = _ops()
    dev = "cuda:0"
    torch.manual_seed(0)
    x = torch.randn(M, d, device=dev).to(bf16)
    ab = (0.5 * torch.randn(M // T, 2 * d, device=dev)).to(bf16)
    a, b = ab[:, :d], ab[:, d:]
    h, rstd = ops.adaln_modulate_fwd(x, a, b, T, EPS)
    xf = x.float().requires_grad_(True)
    af = a.float().requires_grad_(True)
    bf = b.float().requires_grad_(True)
    ref = rms(xf) * (1 + af.repeat_interleave(T, 0)) + bf.repeat_interleave(T, 0)
    assert relerr(h, ref) <= 1e-2
    assert relerr(rstd, torch.rsqrt(x.float().pow(2).mean(-1) + EPS)) <= 1e-3
    dh = torch.randn(M, d, device=dev).to(bf16)
    dres = torch.randn(M, d, device=dev).to(bf16)
    dab = torch.empty(M // T, 2 * d, device=dev, dtype=bf16)
    dx = ops.adaln_modulate_bwd(dh, x, a, rstd, dres, dab[:, :d], dab[:, d:], T)
    ref.backward(dh.float())
    assert relerr(dx, xf.grad + dres.float()) <= 1e-2
    assert relerr(dab[:, :d], af.grad) <= 1.5e-2
    assert relerr(dab[:, d:], bf.grad) <= 1.5e-2
    # no modulation: plain rms_norm
    h2, _ = ops.adaln_modulate_fwd(x, None, None, T, EPS)
    assert relerr(h2, rms(x.float())) <= 1e-2


@pytest.mark.parametrize("M,d,T", [(3840, 1024, 1), (1040, 1536, 65), (2048, 1536, 64)])
def test_gate_fwd_bwd(M, d, T):
    ops = _ops()
    dev = "cuda:0"
    torch.manual_seed(1)
    y = torch.randn(M, d, device=dev).to(bf16)
    res = torch.randn(M, d, device=dev).to(bf16)
    c = torch.randn(M // T, d, device=dev).to(bf16)
    out = ops.gate_fwd(y, c, res, T)
    cf = c.float().repeat_interleave(T, 0)
    assert relerr(out, res.float() + cf * y.float()) <= 1e-2
    out2 = ops.gate_fwd(y, c, None, T)
    assert relerr(out2, cf * y.float()) <= 1e-2
    dx = torch.randn(M, d, device=dev).to(bf16)
    dc = torch.empty(M // T, d, device=dev, dtype=bf16)
    dbias = torch.empty(d, device=dev)
    dy = ops.gate_bwd(dx, y, c, dc, dbias, T)
    assert relerr(dy, dx.float() * cf) <= 1e-2
    assert relerr(dc, (dx.float() * y.float()).view(M // T, T, d).sum(1)) <= 1.5e-2
    assert relerr(dbias, dy.float().sum(0)) <= 1e-3


@pytest.mark.parametrize("M,N", [(3840, 3072), (1000, 1032), (5, 8)])
def test_colsum(M, N):
    ops = _ops()
    x = torch.randn(M, N, device="cuda:0").to(bf16)
    out = torch.full((N,), 3.0, device="cuda:0")
    ops.colsum(x, out)
    assert relerr(out, x.float().sum(0)) <= 1e-4
    ops.colsum(x, out, accumulate=True)
    assert relerr(out, 2 * x.float().sum(0)) <= 1e-4


def test_cast_silu_layernorm():
    ops = _ops()
    dev = "cuda:0"
    w = torch.randn(1000, 1027, device=dev)
    assert torch.equal(ops.cast_f32_bf16(w), w.to(bf16))
    x = torch.randn(333, 1031, device=dev).to(bf16)
    assert relerr(ops.silu(x), F.silu(x.float())) <= 1e-2
    x = torch.randn(1040, 1536, device=dev).to(bf16)
    assert relerr(ops.layer_norm(x, 1e-5), F.layer_norm(x.float(), (1536,), eps=1e-5)) <= 1e-2


def rope_ref(x, cos, sin):
    """rope.py:43-51 restated: pairs (2i, 2i+1), output [even-rotated | odd-rotated]."""
    x0, x1 = x.float().unfold(-1, 2, 2).unbind(-1)
    return torch.cat((x0 * cos - x1 * sin, x1 * cos + x0 * sin), dim=-1)


@pytest.mark.parametrize("B,L,H,dh,off", [(2, 120, 16, 64, 0), (1, 65, 24, 64, 130), (1, 256, 20, 128, 0)])
def test_qk_norm_rope_fwd_bwd(B, L, H, dh, off):
    ops = _ops()
    dev = "cuda:0"
    torch.manual_seed(2)
    qkv = torch.randn(B, L, 3 * H * dh, device=dev).to(bf16)
    ang = torch.rand(off + L + 3, dh // 2, device=dev) * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    q = torch.empty(B, H, L, dh, device=dev, dtype=bf16)
    # k and v land in the tail of a larger (cache-like) buffer to exercise the strides
    kbuf = torch.zeros(B, H, L + 40, dh, device=dev, dtype=bf16)
    vbuf = torch.zeros(B, H, L + 40, dh, device=dev, dtype=bf16)
    k, v = kbuf[:, :, 40:], vbuf[:, :, 40:]
    rq, rk = ops.qk_norm_rope_fwd(qkv, cos, sin, off, H, EPS, q, k, v)
    x = qkv.float().view(B, L, 3, H, dh).permute(2, 0, 3, 1, 4).clone().requires_grad_(True)
    c, s = cos[off:off + L], sin[off:off + L]
    qr, kr = rope_ref(rms(x[0]), c, s), rope_ref(rms(x[1]), c, s)
    assert relerr(q, qr) <= 1e-2 and relerr(k, kr) <= 1e-2
    assert torch.equal(v, qkv.view(B, L, 3, H, dh)[:, :, 2].permute(0, 2, 1, 3))
    assert kbuf[:, :, :40].abs().max().item() == 0
    dq = torch.randn(B, H, L, dh, device=dev).to(bf16)
    dk = torch.randn(B, H, L, dh, device=dev).to(bf16)
    dv = torch.randn(B, H, L, dh, device=dev).to(bf16)
    dqkv = ops.qk_norm_rope_bwd(dq, dk, dv, q, k, rq, rk, cos, sin, off)
    (qr * dq.float()).sum().add((kr * dk.float()).sum()).add((x[2] * dv.float()).sum()).backward()
    ref = x.grad.permute(1, 3, 0, 2, 4).reshape(B, L, 3 * H * dh)
    assert relerr(dqkv, ref) <= 1.5e-2


def test_kv_append():
    ops = _ops()
    dev = "cuda:0"
    B, H, cap, dh = 2, 24, 300, 64
    cache = torch.zeros(B, H, cap, dh, device=dev, dtype=bf16)
    new = torch.randn(B, H, 65, dh, device=dev).to(bf16)
    ops.kv_append(new, cache, 130)
    assert torch.equal(cache[:, :, 130:195], new)
    assert cache[:, :, :130].abs().max().item() == 0 and cache[:, :, 195:].abs().max().item() == 0


@pytest.mark.parametrize("M,d", [(48, 128), (130, 1024), (1040, 1536)])
def test_layer_norm_fwd_bwd_vs_torch(M, d):
    """normalization.py:6-7 weight-less LayerNorm and its backward (gamerft_audio.py:86 trains through it)."""
    import torch.nn.functional as F
    from owl_audio_exps_b200 import functional as OF
    torch.manual_seed(M + d)
    x = (torch.randn(M, d, device="cuda:0") * 2 + 0.5).to(torch.bfloat16)
    dy = torch.randn(M, d, device="cuda:0").to(torch.bfloat16)
    xa = x.clone().requires_grad_(True)
    y = OF.layer_norm(xa)
    y.backward(dy)
    xr = x.float().requires_grad_(True)
    yr = F.layer_norm(xr, (d,))
    yr.backward(dy.float())
    assert y.dtype == torch.bfloat16
    assert (y.float() - yr).abs().max() <= 2 ** -7 * yr.abs().max()
    assert (xa.grad.float() - xr.grad).abs().max() <= 2 ** -7 * xr.grad.abs().max()


def test_rf_noise_and_fused_mse_vs_torch():
    """owl_rf_noise / owl_mse_fwd / owl_mse_bwd against the reference's torch expressions (models/gamerft.py:93-96,
    F.mse_loss(pred.float(), target.float())) on bf16 latents."""
    from owl_audio_exps_b200 import functional as OF
    dev = "cuda:0"
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 5, 16, 8, 8, generator=g).to(torch.bfloat16).to(dev)
    z = torch.randn(2, 5, 16, 8, 8, generator=g).to(torch.bfloat16).to(dev)
    t = torch.rand(2, 5, generator=g).to(torch.bfloat16).to(dev)
    lerpd, target = OF.rf_noise(x, z, t)
    tv = t[:, :, None, None, None]
    assert torch.equal(lerpd, x * (1 - tv) + z * tv) and torch.equal(target, z - x)
    pred = torch.randn(2, 5, 16, 8, 8, generator=g).to(torch.bfloat16).to(dev).requires_grad_(True)
    loss = OF.mse_loss(pred, target)
    ref_in = pred.detach().clone().requires_grad_(True)
    ref = torch.nn.functional.mse_loss(ref_in.float(), target.float())
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    (loss * 3.0).backward()
    (ref * 3.0).backward()
    assert relerr(pred.grad, ref_in.grad) <= 1e-2
    # shapes the fused kernels do not take fall back to the torch expressions
    xo = torch.randn(3, 7, 5, device=dev)
    l2, t2 = OF.rf_noise(xo, torch.ones_like(xo), torch.full((3, 7), 0.25, device=dev))
    assert torch.allclose(l2, xo * 0.75 + 0.25) and torch.allclose(t2, 1 - xo)
