"""Steady-state throughput of the GEMM kernel on a large square problem, 1-CTA vs CTA-pair, vs cuBLAS."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from owl_audio_exps_b200 import ops
    from tools.gemm_shapes import timeit
    dev, bf16 = "cuda:0", torch.bfloat16
    out = []
    for n in (4096, 8192):
        a = torch.randn(n, n, device=dev).to(bf16); b = torch.randn(n, n, device=dev).to(bf16)
        o = torch.empty(n, n, device=dev, dtype=bf16)
        us = timeit(lambda: ops.gemm(0, a, b, n, n, n, mode=ops.OWL_EPI_STORE, out0=o), n=10)
        usr = timeit(lambda: a @ b.t(), n=10)
        out.append(f"n={n}: {us:.0f}us {2.0*n**3/us/1e6:.0f}TF (cuBLAS {usr:.0f}us {2.0*n**3/usr/1e6:.0f}TF)")
    print(" | ".join(out))
else:
    for env in [{"OWL_GEMM_CTAS": "1"}, {"OWL_GEMM_CTAS": "2"}, {"OWL_GEMM_CTAS": "1", "OWL_GEMM_STAGES": "3"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_STAGES": "4"}, {"OWL_GEMM_CTAS": "2", "OWL_GEMM_STAGES": "3"}]:
        r = subprocess.run([sys.executable, __file__, "child"], env={**os.environ, **env}, capture_output=True, text=True)
        print(env, r.stdout.strip() or r.stderr[-400:])
