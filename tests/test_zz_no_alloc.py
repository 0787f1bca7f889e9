"""owl_gemm_smallm_static: the no-allocation variant of the decode-regime GEMM (split-K scratch in a file-scope __device__
array of libowl_dit.so), called through the C ABI and checked against the existing entry point owl_gemm_smallm.

Collected last on purpose: a failure here must not hide another test.  Tolerance: the
bar of tests/test_gemm_gpu.py for bf16 outputs, max|out - ref| / max|ref| <= 1e-2 (split-K partial sums meet through fp32
L2 atomics, so the summation order - not the arithmetic - may differ between two launches).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

bf16 = torch.bfloat16
OWL_ERR_TOO_LARGE = -5


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def _problem(M, N, K, mode, dev):
    g = torch.Generator(device="cpu").manual_seed(11 * M + 3 * N + 7 * K + mode)
    A = torch.randn(M, K, generator=g).to(bf16).to(dev)
    W = (torch.randn(N, K, generator=g) * K ** -0.5).to(bf16).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    T = 65 if M % 65 == 0 else 1
    res = torch.randn(M, N, generator=g).to(bf16).to(dev)
    gate = torch.randn(M // T, N, generator=g).to(bf16).to(dev)
    return A, W, bias, res, gate, T


def _epilogue(_lib, mode, bias, out, res, gate, T):
    e = _lib.GemmEpilogue()
    e.mode, e.rows_per_frame = mode, T
    e.bias = bias.data_ptr()
    e.out0, e.ld0 = out.data_ptr(), out.stride(0)
    if mode == 2:
        e.aux0, e.lda0 = res.data_ptr(), res.stride(0)
        e.aux1, e.lda1 = gate.data_ptr(), gate.stride(0)
    return e


def _static_call(_lib, A, W, M, N, K, e):
    stream = torch.cuda.current_stream().cuda_stream
    return _lib.load().owl_gemm_smallm_static(M, N, K, A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0), C.byref(e),
                                              C.c_void_p(stream))


# shapes of one AV decode evaluation (split-K with different tile counts, so consecutive calls reuse the SAME scratch bytes
# with different geometry), the one-token conditioning bank slice, and a shape the planner does not split
SHAPES = [(65, 4608, 1536, 0), (65, 1536, 6144, 2), (65, 6144, 1536, 1), (1, 8192, 1536, 0), (7, 64, 64, 0)]


def test_static_scratch_gemm_matches_the_workspace_entry_point():
    from owl_audio_exps_b200 import _lib, ops
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    assert lib.owl_gemm_smallm_static_max_n() == 8192
    # everything the test needs exists before the no-allocation claim is checked
    cases = []
    for M, N, K, mode in SHAPES:
        A, W, bias, res, gate, T = _problem(M, N, K, mode, dev)
        ref = torch.empty(M, N, device=dev, dtype=bf16)
        kw = dict(mode=mode, bias=bias)
        if mode == 2:
            kw.update(aux0=res, aux1=gate, rows_per_frame=T)
        ops.gemm_smallm(A, W, M, N, K, out0=ref, **kw)           # the existing entry point (caller-provided workspace)
        outs = [torch.full((M, N), float("nan"), device=dev, dtype=bf16) for _ in range(2)]
        cases.append((M, N, K, mode, A, W, bias, res, gate, T, ref, outs))
    # one warm call (module load, cudaGetSymbolAddress of this device), then measure
    M, N, K, mode, A, W, bias, res, gate, T, ref, outs = cases[0]
    assert _static_call(_lib, A, W, M, N, K, _epilogue(_lib, mode, bias, outs[0], res, gate, T)) == 0
    torch.cuda.synchronize()
    outs[0].fill_(float("nan"))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    alloc0 = torch.cuda.memory_allocated(dev)
    for rep in range(2):                                          # two calls in a row per shape, shapes interleaved
        for M, N, K, mode, A, W, bias, res, gate, T, ref, outs in cases:
            rc = _static_call(_lib, A, W, M, N, K, _epilogue(_lib, mode, bias, outs[rep], res, gate, T))
            assert rc == 0, (rc, M, N, K)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert torch.cuda.memory_allocated(dev) == alloc0             # nothing through torch's allocator
    assert free1 >= free0                                         # nothing through cudaMalloc either
    for M, N, K, mode, A, W, bias, res, gate, T, ref, outs in cases:
        for rep in range(2):
            assert torch.isfinite(outs[rep].float()).all(), (M, N, K, rep)
            assert relerr(outs[rep], ref) <= 1e-2, (M, N, K, mode, rep)


def test_static_scratch_gemm_rejects_sizes_beyond_its_capacity():
    from owl_audio_exps_b200 import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    M, N, K = 4, 8192 + 64, 64
    A = torch.zeros(M, K, device=dev, dtype=bf16)
    W = torch.zeros(N, K, device=dev, dtype=bf16)
    out = torch.zeros(M, N, device=dev, dtype=bf16)
    e = _lib.GemmEpilogue()
    e.mode, e.rows_per_frame = 0, 1
    e.out0, e.ld0 = out.data_ptr(), N
    rc = _static_call(_lib, A, W, M, N, K, e)
    assert rc == OWL_ERR_TOO_LARGE
    msg = _lib.load().owl_error_string(rc).decode()
    assert "capacity" in msg and "no-allocation" in msg
