"""Sweep forced tile widths / stages for a few GEMM shapes (tuning aid; GPU box only)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from owl_audio_exps_b200 import ops
    from tools.gemm_shapes import timeit
    dev, bf16 = "cuda:0", torch.bfloat16
    g = lambda *s: torch.randn(*s, device=dev).to(bf16)
    out = []
    for tag, layout, M, N, K in [("qkvNT", 0, 3840, 3072, 1024), ("fc1NT", 0, 3840, 4096, 1024), ("outNT", 0, 3840, 1024, 1024),
                                 ("fc2NT", 0, 3840, 1024, 4096), ("qkvNN", 1, 3840, 1024, 3072), ("outNN", 1, 3840, 1024, 1024),
                                 ("woutTN", 2, 1024, 1024, 3840), ("wqkvTN", 2, 3072, 1024, 3840), ("wcTN", 2, 6144, 1024, 3840),
                                 ("wfc1TN", 2, 4096, 1024, 3840), ("wfc2TN", 2, 1024, 4096, 3840), ("dfc2NN", 1, 3840, 4096, 1024)]:
        if layout == 0:
            a, b = g(M, K), g(N, K)
        elif layout == 1:
            a, b = g(M, K), g(K, N)
        else:
            a, b = g(K, M), g(K, N)
        if layout == 2:
            o = torch.zeros(M, N, device=dev)
            kw = dict(mode=ops.OWL_EPI_F32, outf=o)
        elif tag == "fc1NT":
            kw = dict(mode=ops.OWL_EPI_BIAS_SILU, out0=torch.empty(M, N, device=dev, dtype=bf16), out1=torch.empty(M, N, device=dev, dtype=bf16), bias=torch.randn(N, device=dev))
        else:
            kw = dict(mode=ops.OWL_EPI_STORE, out0=torch.empty(M, N, device=dev, dtype=bf16))
        us = timeit(lambda: ops.gemm(layout, a, b, M, N, K, **kw))
        out.append(f"{tag}:{us:.1f}us/{2.0*M*N*K/us/1e6:.0f}TF")
    print(" ".join(out))
else:
    quick = [{}, {"OWL_GEMM_CTAS": "1"}, {"OWL_GEMM_CTAS": "2"}, {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "1"}]
    full = [{}, {"OWL_GEMM_BN": "256"}, {"OWL_GEMM_BN": "128"}, {"OWL_GEMM_BN": "64"}, {"OWL_GEMM_BN": "256", "OWL_GEMM_STAGES": "3"},
                {"OWL_GEMM_BN": "128", "OWL_GEMM_STAGES": "3"}, {"OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "4"},
                {"OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "2"}, {"OWL_GEMM_BN": "128", "OWL_GEMM_SPLITS": "2"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "1"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "2"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "3"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "4"},
                {"OWL_GEMM_CTAS": "2", "OWL_GEMM_BN": "256", "OWL_GEMM_SPLITS": "6"},
                {"OWL_GEMM_BN": "128", "OWL_GEMM_SPLITS": "3"}, {"OWL_GEMM_BN": "128", "OWL_GEMM_SPLITS": "4"}]
    for env in (quick if "--quick" in sys.argv else full):
        r = subprocess.run([sys.executable, __file__, "child"], env={**os.environ, **env}, capture_output=True, text=True)
        print(env, "\n   ", r.stdout.strip() or r.stderr[-300:])
