"""Decode-regime GEMMs: owl_gemm_smallm vs the training GEMM at M = 65 / 1, weights streamed from HBM (every launch of the
timed graph reads a different weight copy; the copies together exceed the 126 MB L2).  GPU box only.
    python tools/smallm_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops

bf16, dev = torch.bfloat16, "cuda:0"
SHAPES = [("qkv", 65, 4608, 1536), ("out", 65, 1536, 1536), ("fc1", 65, 6144, 1536), ("fc2", 65, 1536, 6144),
          ("cond bank 4 layers", 1, 6 * 1536 * 4, 1536), ("audio qkv", 1, 3072, 1024), ("audio fc1", 1, 4096, 1024)]


def timeit(fns, reps=5):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for f in fns[:2]:
            f()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for f in fns:
            f()
    gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (len(fns) * reps) * 1e3


def main():
    print(f"{'shape':22s} {'M':>4s} {'N':>6s} {'K':>5s} {'MB':>7s} {'smallm us':>10s} {'GB/s':>7s} {'train-gemm us':>14s}")
    for name, M, N, K in SHAPES:
        copies = max(2, int(300e6 // (N * K * 2)))
        copies = min(copies, 40)
        Ws = [(torch.randn(N, K, device=dev) * K ** -0.5).to(bf16) for _ in range(copies)]
        A = torch.randn(M, K, device=dev).to(bf16)
        bias = torch.zeros(N, device=dev)
        out = torch.empty(M, N, device=dev, dtype=bf16)
        ops.gemm_smallm(A, Ws[0], M, N, K, bias=bias, out0=out)   # workspace exists before capture
        t_s = timeit([lambda W=W: ops.gemm_smallm(A, W, M, N, K, bias=bias, out0=out) for W in Ws])
        t_g = timeit([lambda W=W: ops.gemm(0, A, W, M, N, K, bias=bias, out0=out) for W in Ws])
        mb = N * K * 2 / 1e6
        print(f"{name:22s} {M:4d} {N:6d} {K:5d} {mb:7.1f} {t_s:10.2f} {mb / t_s * 1e3:7.0f} {t_g:14.2f}", flush=True)


if __name__ == "__main__":
    main()
