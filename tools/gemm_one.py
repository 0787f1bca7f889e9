"""A handful of launches of selected block GEMMs (for `ncu --set full -k regex:gemm_bf16`), audio.yml shapes.
    python tools/gemm_one.py [fc1nt qkvnt fc2nn wqkvtn ...] [--iters N]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops

dev, bf16 = "cuda:0", torch.bfloat16
M, d = 3840, 1024
g = lambda *s: torch.randn(*s, device=dev).to(bf16)


def case(tag):
    if tag == "fc1nt":
        a, w, b = g(M, d), g(4 * d, d), torch.randn(4 * d, device=dev)
        z, s = torch.empty(M, 4 * d, device=dev, dtype=bf16), torch.empty(M, 4 * d, device=dev, dtype=bf16)
        return lambda: ops.gemm(0, a, w, M, 4 * d, d, mode=ops.OWL_EPI_BIAS_SILU, bias=b, out0=z, out1=s)
    if tag == "qkvnt":
        a, w, b = g(M, d), g(3 * d, d), torch.randn(3 * d, device=dev)
        o = torch.empty(M, 3 * d, device=dev, dtype=bf16)
        return lambda: ops.gemm(0, a, w, M, 3 * d, d, mode=ops.OWL_EPI_STORE, bias=b, out0=o)
    if tag == "fc2nt":
        a, w, b = g(M, 4 * d), g(d, 4 * d), torch.randn(d, device=dev)
        o = torch.empty(M, d, device=dev, dtype=bf16)
        return lambda: ops.gemm(0, a, w, M, d, 4 * d, mode=ops.OWL_EPI_STORE, bias=b, out0=o)
    if tag == "fc2nn":
        dy, w, z = g(M, d), g(d, 4 * d), g(M, 4 * d)
        o = torch.empty(M, 4 * d, device=dev, dtype=bf16)
        return lambda: ops.gemm(1, dy, w, M, 4 * d, d, mode=ops.OWL_EPI_DSILU, aux0=z, out0=o)
    if tag == "fc1nn":
        dy, w = g(M, 4 * d), g(4 * d, d)
        o = torch.empty(M, d, device=dev, dtype=bf16)
        return lambda: ops.gemm(1, dy, w, M, d, 4 * d, mode=ops.OWL_EPI_STORE, out0=o)
    if tag == "wqkvtn":
        dy, x = g(M, 3 * d), g(M, d)
        o = torch.zeros(3 * d, d, device=dev)
        return lambda: ops.gemm(2, dy, x, 3 * d, d, M, mode=ops.OWL_EPI_F32, outf=o)
    if tag == "wouttn":
        dy, x = g(M, d), g(M, d)
        o = torch.zeros(d, d, device=dev)
        return lambda: ops.gemm(2, dy, x, d, d, M, mode=ops.OWL_EPI_F32, outf=o)
    if tag == "w1tn":
        dy, x = g(M, 4 * d), g(M, d)
        o = torch.zeros(4 * d, d, device=dev)
        return lambda: ops.gemm(2, dy, x, 4 * d, d, M, mode=ops.OWL_EPI_F32, outf=o)
    if tag == "condnt":   # the conditioning bank's forward: all 16 layers at once
        a, w, b = g(M, d), g(96 * d, d), torch.randn(96 * d, device=dev)
        o = torch.empty(M, 96 * d, device=dev, dtype=bf16)
        return lambda: ops.gemm(0, a, w, M, 96 * d, d, mode=ops.OWL_EPI_STORE, bias=b, out0=o)
    if tag == "condnn":
        dy, w = g(M, 96 * d), g(96 * d, d)
        o = torch.empty(M, d, device=dev, dtype=bf16)
        return lambda: ops.gemm(1, dy, w, M, d, 96 * d, mode=ops.OWL_EPI_STORE, out0=o)
    if tag == "condtn":   # one group of 4 layers
        dy, x = g(M, 24 * d), g(M, d)
        o = torch.zeros(24 * d, d, device=dev)
        return lambda: ops.gemm(2, dy, x, 24 * d, d, M, mode=ops.OWL_EPI_F32, outf=o)
    raise SystemExit(f"unknown case {tag}")


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    iters = 3
    if "--iters" in sys.argv:
        iters = int(sys.argv[sys.argv.index("--iters") + 1])
        args = [a for a in args if a != str(iters)]
    tags = args or ["fc1nt", "qkvnt", "fc2nn", "wqkvtn"]
    fns = [(t, case(t)) for t in tags]
    for t, fn in fns:
        for _ in range(iters):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{t}: {e0.elapsed_time(e1) / iters * 1e3:.1f} us")


if __name__ == "__main__":
    main()
