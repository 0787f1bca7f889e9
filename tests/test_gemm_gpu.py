"""tcgen05 GEMM (owl_gemm_bf16) against a torch fp32 reference and the CUDA-core check kernel.

Tolerance: operands are bf16 (exact in fp32), accumulation fp32, so the only error of a bf16
output is its final rounding: |err| <= 2^-8 * |ref| + accumulation-order noise.  Stated bar:
max|out - ref| / max|ref| <= 1e-2 for bf16 outputs, 1e-4 for fp32 outputs.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

bf16 = torch.bfloat16


def _ops():
    from owl_audio_exps_b200 import ops
    return ops


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def _operands(layout, M, N, K, dev):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K + layout)
    A = torch.randn(M, K, generator=g).to(bf16)
    Bm = torch.randn(N, K, generator=g).to(bf16)
    ref = A.float() @ Bm.float().t()
    a = A if layout != 2 else A.t().contiguous()          # TN: A stored [K, M]
    b = Bm if layout == 0 else Bm.t().contiguous()        # NN/TN: B stored [K, N]
    return a.to(dev), b.to(dev), ref.to(dev)


SHAPES = [
    (128, 128, 64), (128, 256, 128), (256, 512, 1024), (120, 1024, 1024), (3840, 3072, 1024),
    (3840, 1024, 4096), (200, 72, 136), (1, 64, 64), (65, 1536, 1536), (1000, 2048, 520),
]


@pytest.mark.parametrize("layout", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_store(layout, M, N, K):
    ops = _ops()
    dev = torch.device("cuda:0")
    if layout == 2 and M % 8:
        pytest.skip("TN needs lda = M multiple of 8")
    a, b, ref = _operands(layout, M, N, K, dev)
    out = torch.empty(M, N, device=dev, dtype=bf16)
    ops.gemm(layout, a, b, M, N, K, mode=ops.OWL_EPI_STORE, out0=out)
    torch.cuda.synchronize()
    assert relerr(out, ref) <= 1e-2
    chk = ops.gemm_check(layout, a, b, M, N, K)
    assert relerr(chk, ref) <= 1e-4


@pytest.mark.parametrize("layout", [0, 2])
@pytest.mark.parametrize("M,N,K", [(1024, 1024, 3840), (3072, 1024, 3840), (256, 128, 98304), (1024, 4096, 120)])
def test_gemm_f32_splitk_and_accumulate(layout, M, N, K):
    ops = _ops()
    dev = torch.device("cuda:0")
    a, b, ref = _operands(layout, M, N, K, dev)
    out = torch.full((M, N), 7.0, device=dev, dtype=torch.float32)
    ops.gemm(layout, a, b, M, N, K, mode=ops.OWL_EPI_F32, outf=out)
    assert relerr(out, ref) <= 2e-4  # fp32 reduce-add order of split-K partials
    ops.gemm(layout, a, b, M, N, K, mode=ops.OWL_EPI_F32, outf=out, accumulate=True)
    assert relerr(out, 2 * ref) <= 1e-4


def test_gemm_bias_silu():
    ops = _ops()
    dev = torch.device("cuda:0")
    M, N, K = 3840, 4096, 1024
    a, b, ref = _operands(0, M, N, K, dev)
    ref = ref / 32
    a = (a.float() / 32).to(bf16)  # exact power-of-two scaling
    bias = torch.randn(N, device=dev)
    z = torch.empty(M, N, device=dev, dtype=bf16)
    s = torch.empty(M, N, device=dev, dtype=bf16)
    ops.gemm(0, a, b, M, N, K, mode=ops.OWL_EPI_BIAS_SILU, bias=bias, out0=z, out1=s)
    zr = (ref + bias).to(bf16)
    assert relerr(z, zr) <= 1e-2
    assert relerr(s, torch.nn.functional.silu(zr.float())) <= 1e-2


@pytest.mark.parametrize("T", [1, 64, 65])
def test_gemm_gate_residual(T):
    ops = _ops()
    dev = torch.device("cuda:0")
    F_, N, K = 24, 1536, 1536
    M = F_ * T
    a, b, ref = _operands(0, M, N, K, dev)
    ref = ref / 32
    a = (a.float() / 32).to(bf16)
    bias = torch.randn(N, device=dev)
    res = torch.randn(M, N, device=dev).to(bf16)
    gate = torch.randn(F_, N, device=dev).to(bf16)
    out = torch.empty(M, N, device=dev, dtype=bf16)
    y = torch.empty(M, N, device=dev, dtype=bf16)
    ops.gemm(0, a, b, M, N, K, mode=ops.OWL_EPI_GATE_RESIDUAL, bias=bias, out0=out, out1=y, aux0=res, aux1=gate,
             rows_per_frame=T)
    yr = (ref + bias).to(bf16)
    expect = res.float() + gate.float().repeat_interleave(T, 0) * yr.float()
    assert relerr(y, yr) <= 1e-2
    assert relerr(out, expect) <= 1e-2


def test_gemm_dsilu():
    ops = _ops()
    dev = torch.device("cuda:0")
    M, N, K = 1024, 4096, 1024
    a, b, ref = _operands(1, M, N, K, dev)
    z = torch.randn(M, N, device=dev).to(bf16)
    out = torch.empty(M, N, device=dev, dtype=bf16)
    ops.gemm(1, a, b, M, N, K, mode=ops.OWL_EPI_DSILU, out0=out, aux0=z)
    zf = z.float()
    sg = torch.sigmoid(zf)
    expect = ref * (sg * (1 + zf * (1 - sg)))
    assert relerr(out, expect) <= 1e-2


def test_gemm_rejects_bad_args():
    ops = _ops()
    dev = torch.device("cuda:0")
    a = torch.zeros(128, 64, device=dev, dtype=bf16)
    b = torch.zeros(128, 64, device=dev, dtype=bf16)
    with pytest.raises(RuntimeError):
        ops.gemm(0, a, b, 128, 128, 64, mode=ops.OWL_EPI_STORE)  # no output
    with pytest.raises(RuntimeError):
        ops.gemm(0, a[:, 1:], b, 128, 128, 56, mode=ops.OWL_EPI_STORE, out0=torch.empty(128, 128, device=dev, dtype=bf16))


@pytest.mark.parametrize("ctas", ["1", "2"])
@pytest.mark.parametrize("layout", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", [(384, 320, 200), (3840, 1024, 1024), (1000, 2048, 520)])
def test_gemm_single_cta_and_cta_pair(monkeypatch, ctas, layout, M, N, K):
    """Both tile schedules (one CTA per 128x256 tile; cta_group::2 pair per 256x256 tile) on the same operands,
    incl. ragged M / N / K: the planner normally picks one per shape, OWL_GEMM_CTAS forces it."""
    ops = _ops()
    dev = torch.device("cuda:0")
    monkeypatch.setenv("OWL_GEMM_CTAS", ctas)
    a, b, ref = _operands(layout, M, N, K, dev)
    out = torch.empty(M, N, device=dev, dtype=bf16)
    bias = torch.randn(N, device=dev)
    ops.gemm(layout, a, b, M, N, K, mode=ops.OWL_EPI_STORE, bias=bias, out0=out)
    outf = torch.zeros(M, N, device=dev)
    ops.gemm(layout, a, b, M, N, K, mode=ops.OWL_EPI_F32, outf=outf)
    torch.cuda.synchronize()
    assert relerr(out, ref + bias) <= 1e-2
    assert relerr(outf, ref) <= 1e-4


# ------------------------------------------------------------------------------------------------
# weight-streaming small-M GEMM of the decode regime (owl_gemm_smallm)
# ------------------------------------------------------------------------------------------------
SMALLM_SHAPES = [
    # M, N, K: the Linears of one AV decode evaluation (65 tokens, d = 1536), the conditioning bank at one frame,
    # the audio model's single token, and ragged / tiny shapes
    (65, 1536, 1536), (65, 4608, 1536), (65, 6144, 1536), (65, 1536, 6144), (1, 6 * 1536 * 4, 1536),
    (1, 1024, 1024), (1, 3072, 1024), (128, 192, 320), (7, 64, 64), (130 // 2, 72, 136),
]


def _smallm_ref(mode, A, W, bias, res, gate, T):
    y = A.float() @ W.float().t() + (bias if bias is not None else 0.0)
    if mode == 0:
        return y
    y = y.to(bf16).float()
    if mode == 1:
        return torch.nn.functional.silu(y)
    return res.float() + gate.float().repeat_interleave(T, dim=0) * y


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", SMALLM_SHAPES)
def test_gemm_smallm(mode, M, N, K):
    import os
    ops = _ops()
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(M + 3 * N + 7 * K + mode)
    A = torch.randn(M, K, generator=g).to(bf16).to(dev)
    W = (torch.randn(N, K, generator=g) * K ** -0.5).to(bf16).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    T = 65 if M % 65 == 0 else 1
    res = torch.randn(M, N, generator=g).to(bf16).to(dev)
    gate = torch.randn(M // T, N, generator=g).to(bf16).to(dev)
    ref = _smallm_ref(mode, A, W, bias, res, gate, T)
    kw = dict(mode=mode, bias=bias)
    if mode == 2:
        kw.update(aux0=res, aux1=gate, rows_per_frame=T)
    outs = []
    # the planner's choice, no split, an odd split count, 8 splits (largest cluster when OWL_SMALLM_CLUSTER=1), 12 splits
    for forced in (None, "1", "3", "8", "12"):
        if forced is None:
            os.environ.pop("OWL_SMALLM_SPLITS", None)
        else:
            os.environ["OWL_SMALLM_SPLITS"] = forced
        try:
            for rep in range(2):               # twice: the kernel must leave its workspace zeroed
                out = torch.full((M, N), float("nan"), device=dev, dtype=bf16)
                ops.gemm_smallm(A, W, M, N, K, out0=out, **kw)
                torch.cuda.synchronize()
                assert relerr(out, ref) <= 1e-2, (forced, rep)
                outs.append(out)
        finally:
            os.environ.pop("OWL_SMALLM_SPLITS", None)
    from owl_audio_exps_b200 import ops as O
    for ws in O._SMALLM_WS.values():
        assert int(ws.count_nonzero()) == 0
    # same problem through the training GEMM (store mode): both tensor-core paths agree to rounding
    if mode == 0:
        big = torch.empty(M, N, device=dev, dtype=bf16)
        ops.gemm(0, A, W, M, N, K, mode=0, bias=bias, out0=big)
        assert relerr(outs[0], big) <= 1e-2


def test_gemm_smallm_rejects_large_m():
    ops = _ops()
    dev = torch.device("cuda:0")
    A = torch.zeros(129, 64, device=dev, dtype=bf16)
    W = torch.zeros(64, 64, device=dev, dtype=bf16)
    with pytest.raises(RuntimeError):
        ops.gemm_smallm(A, W, 129, 64, 64, out0=torch.empty(129, 64, device=dev, dtype=bf16))


@pytest.mark.parametrize("M,N,K,T", [(65, 4608, 1536, 65), (65, 6144, 1536, 65), (1, 3072, 1024, 1), (128, 192, 320, 1),
                                     (130 // 2, 72, 136, 5), (2, 64, 64, 1)])
def test_gemm_smallm_adaln_prologue_matches_the_separate_pass(M, N, K, T):
    """adaLN as the A-operand transform of the decode GEMM (north_star: adaLN scale/shift fused into the GEMM prologue):
    same arithmetic as owl_adaln_modulate_fwd up to the summation order of the row statistic (per-thread here, warp
    shuffles there: the normalised operand can differ by one bf16 ulp), so the product matches the two-kernel result to
    bf16 rounding."""
    import os
    ops = _ops()
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).to(bf16).to(dev)
    mod = torch.randn(M // T, 3 * K, generator=g).to(bf16).to(dev)      # a, b are column slices with a row stride
    a, b = mod[:, :K], mod[:, K:2 * K]
    W = (torch.randn(N, K, generator=g) * K ** -0.5).to(bf16).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    eps = 1.1920928955078125e-07
    h, _ = ops.adaln_modulate_fwd(x, a, b, T, eps, want_rstd=False)
    ref_math = torch.nn.functional.rms_norm(x.float(), (K,), eps=eps) * (1 + a.float().repeat_interleave(T, 0)) \
        + b.float().repeat_interleave(T, 0)
    assert relerr(h, ref_math) <= 1e-2
    for forced in (None, "1", "2"):
        if forced is None:
            os.environ.pop("OWL_SMALLM_SPLITS", None)
        else:
            os.environ["OWL_SMALLM_SPLITS"] = forced
        try:
            two = torch.empty(M, N, device=dev, dtype=bf16)
            ops.gemm_smallm(h, W, M, N, K, mode=1, bias=bias, out0=two)
            one = torch.full((M, N), float("nan"), device=dev, dtype=bf16)
            ops.gemm_smallm(x, W, M, N, K, mode=1, bias=bias, out0=one, adaln=(a, b, T, eps))
            torch.cuda.synchronize()
            assert relerr(one, two) <= 1e-2, forced
            assert relerr(one, torch.nn.functional.silu((ref_math.to(bf16).float() @ W.float().t() + bias).to(bf16).float())) <= 2e-2
        finally:
            os.environ.pop("OWL_SMALLM_SPLITS", None)


def test_gemm_smallm_cluster_reduction_subprocess():
    """OWL_SMALLM_CLUSTER=1 (read once per process): the splits of a tile as one thread-block cluster reducing through
    distributed shared memory, against fp32 math, for clusters of 2, 3 and 8 CTAs and all three epilogues."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r"""
import os, sys, torch
sys.path.insert(0, %r)
from owl_audio_exps_b200 import ops
bf16, dev = torch.bfloat16, "cuda:0"
g = torch.Generator().manual_seed(0)
for M, N, K, T in [(65, 1536, 1536, 65), (1, 3072, 1024, 1), (128, 192, 512, 1)]:
    A = torch.randn(M, K, generator=g).to(bf16).to(dev)
    W = (torch.randn(N, K, generator=g) * K ** -0.5).to(bf16).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    res = torch.randn(M, N, generator=g).to(bf16).to(dev)
    gate = torch.randn(M // T, N, generator=g).to(bf16).to(dev)
    y = A.float() @ W.float().t() + bias
    yb = y.to(bf16).float()
    refs = {0: y, 1: torch.nn.functional.silu(yb), 2: res.float() + gate.float().repeat_interleave(T, 0) * yb}
    for sp in ("2", "3", "8"):
        os.environ["OWL_SMALLM_SPLITS"] = sp
        for mode in (0, 1, 2):
            out = torch.full((M, N), float("nan"), device=dev, dtype=bf16)
            kw = dict(aux0=res, aux1=gate, rows_per_frame=T) if mode == 2 else {}
            ops.gemm_smallm(A, W, M, N, K, mode=mode, bias=bias, out0=out, **kw)
            torch.cuda.synchronize()
            err = ((out.float() - refs[mode]).abs().max() / refs[mode].abs().max()).item()
            assert err <= 1e-2, (M, N, K, sp, mode, err)
print("ok")
""" % root
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300,
                         env=dict(os.environ, OWL_SMALLM_CLUSTER="1"))
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), (out.stdout[-800:], out.stderr[-1500:])
