"""Per-shape timing of the tcgen05 GEMM on the shapes of one DiT block (CUDA events, L2-cold-ish rotation of
buffers), with torch.matmul (cuBLAS) beside it as the library reference point.  GPU box only.
    python tools/gemm_shapes.py [audio|av|dit_v4]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops

bf16 = torch.bfloat16
dev = "cuda:0"
CFG = {"audio": (3840, 1024, 1), "av": (8320, 1536, 65), "dit_v4": (98304, 1536, 64)}


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "audio"
    M, d, T = CFG[name]
    R = M // T
    g = lambda *s: torch.randn(*s, device=dev).to(bf16)
    rows = []

    def run(tag, layout, Mo, No, K, a, b, ref, **kw):
        us = timeit(lambda: ops.gemm(layout, a, b, Mo, No, K, **kw))
        us_ref = timeit(ref)
        tf = 2.0 * Mo * No * K / us / 1e6
        rows.append((tag, Mo, No, K, us, tf, us_ref, 2.0 * Mo * No * K / us_ref / 1e6))

    # forward NT
    for tag, N, K, rows_, mode in [("cond  NT store", 6 * d, d, R, "store"), ("qkv   NT store", 3 * d, d, M, "store"),
                                   ("out   NT gate+res", d, d, M, "gate"), ("fc1   NT bias+silu", 4 * d, d, M, "silu"),
                                   ("fc2   NT gate+res", d, 4 * d, M, "gate")]:
        a, w = g(rows_, K), g(N, K)
        bias = torch.randn(N, device=dev)
        out0, out1 = torch.empty(rows_, N, device=dev, dtype=bf16), torch.empty(rows_, N, device=dev, dtype=bf16)
        if mode == "store":
            kw = dict(mode=ops.OWL_EPI_STORE, bias=bias, out0=out0)
        elif mode == "silu":
            kw = dict(mode=ops.OWL_EPI_BIAS_SILU, bias=bias, out0=out0, out1=out1)
        else:
            kw = dict(mode=ops.OWL_EPI_GATE_RESIDUAL, bias=bias, out0=out0, out1=out1, aux0=g(rows_, N),
                      aux1=g(rows_ // T, N), rows_per_frame=T)
        run(tag, ops.OWL_GEMM_NT, rows_, N, K, a, w, lambda a=a, w=w: a @ w.t(), **kw)
    # dgrad NN: dx[M,K] = dy[M,N] W[N,K]
    for tag, N, K, rows_, mode in [("fc2   NN dsilu", d, 4 * d, M, "dsilu"), ("fc1   NN store", 4 * d, d, M, "store"),
                                   ("out   NN store", d, d, M, "store"), ("qkv   NN store", 3 * d, d, M, "store"),
                                   ("cond  NN store", 6 * d, d, R, "store")]:
        dy, w = g(rows_, N), g(N, K)
        out0 = torch.empty(rows_, K, device=dev, dtype=bf16)
        kw = dict(mode=ops.OWL_EPI_STORE, out0=out0) if mode == "store" else dict(mode=ops.OWL_EPI_DSILU, out0=out0, aux0=g(rows_, K))
        run(tag, ops.OWL_GEMM_NN, rows_, K, N, dy, w, lambda dy=dy, w=w: dy @ w, **kw)
    # wgrad TN: dW[N,K] = dy[M,N]^T x[M,K]
    for tag, N, K, rows_ in [("w2    TN f32", d, 4 * d, M), ("w1    TN f32", 4 * d, d, M), ("wout  TN f32", d, d, M),
                             ("wqkv  TN f32", 3 * d, d, M), ("wc    TN f32", 6 * d, d, R)]:
        dy, x = g(rows_, N), g(rows_, K)
        outf = torch.zeros(N, K, device=dev)
        run(tag, ops.OWL_GEMM_TN, N, K, rows_, dy, x, lambda dy=dy, x=x: dy.t() @ x, mode=ops.OWL_EPI_F32, outf=outf,
            accumulate=False)
    tot = sum(r[4] for r in rows)
    tot_ref = sum(r[6] for r in rows)
    print(f"{'gemm':20s} {'M':>6s} {'N':>6s} {'K':>6s} {'us':>8s} {'TF/s':>7s} | {'cuBLAS us':>9s} {'TF/s':>7s}")
    for r in rows:
        print(f"{r[0]:20s} {r[1]:6d} {r[2]:6d} {r[3]:6d} {r[4]:8.1f} {r[5]:7.0f} | {r[6]:9.1f} {r[7]:7.0f}")
    fl = sum(2.0 * r[1] * r[2] * r[3] for r in rows)
    print(f"block total: {tot:.0f} us = {fl / tot / 1e6:.0f} TF/s ; cuBLAS (no epilogue work) {tot_ref:.0f} us = {fl / tot_ref / 1e6:.0f} TF/s")


if __name__ == "__main__":
    main()
