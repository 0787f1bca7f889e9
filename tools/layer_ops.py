"""Back-to-back timing (CUDA events around CUDA-graph replays of 20 launches) of every non-GEMM kernel of one DiT layer at a config's
shapes, with the algorithmic bytes and the achieved GB/s.  GPU box only.
    python tools/layer_ops.py [audio|av]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops

bf16, f32, dev = torch.bfloat16, torch.float32, "cuda:0"
CFG = {"audio": dict(B=32, L=120, d=1024, H=16, T=1, W=16), "av": dict(B=8, L=1040, d=1536, H=24, T=65, W=16),
       "dit_v4": dict(B=1, L=98304, d=1536, H=24, T=64, W=16)}
ONCE = "--once" in sys.argv      # one eager launch per op (for ncu); attention is left out (tools/attn_one.py)


def timeit(fn, n=20, reps=5):
    """n launches captured into a CUDA graph (so the CPU launch path is out of the picture), replayed reps times."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(n):
            fn()
    gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * reps) * 1e3


def main():
    c = CFG[sys.argv[1] if len(sys.argv) > 1 else "audio"]
    big = c["L"] > 20000
    B, L, d, H, T, W = c["B"], c["L"], c["d"], c["H"], c["T"], c["W"]
    M, R, dh = B * L, B * L // T, d // H
    g = lambda *s: torch.randn(*s, device=dev).to(bf16)
    x, y, dhh = g(M, d), g(M, d), g(M, d)
    mod, dmod = g(R, 6 * d), torch.empty(R, 6 * d, device=dev, dtype=bf16)
    a, b, cc = mod[:, :d], mod[:, d:2 * d], mod[:, 2 * d:3 * d]
    rstd = torch.rand(M, device=dev) + 0.5
    qkv = g(B, L, 3 * d)
    cos, sin = torch.rand(max(4096, L), dh // 2, device=dev), torch.rand(max(4096, L), dh // 2, device=dev)
    q, k, v = (torch.empty(B, H, L, dh, device=dev, dtype=bf16) for _ in range(3))
    rq, rk = ops.qk_norm_rope_fwd(qkv, cos, sin, 0, H, 1e-6, q, k, v)
    mk = dict(tokens_per_frame=T, window_frames=0, q_offset=0, causal=True, doc_id=None)
    do = g(B, L, d)
    if ONCE or big:
        o, lse, dq, dk, dv = None, None, g(B, H, L, dh), g(B, H, L, dh), g(B, H, L, dh)
    else:
        o, lse = ops.attn_fwd(q, k, v, **mk)
        dq, dk, dv = ops.attn_bwd(do, o, q, k, v, lse, **mk)
    kc = torch.empty(B, H, L + 256, dh, device=dev, dtype=bf16)
    z, dy4 = g(M, 4 * d), g(M, 4 * d)
    bias32 = torch.zeros(4 * d, device=dev)
    w32 = torch.randn(4 * d, d, device=dev)
    w16 = torch.empty(4 * d, d, device=dev, dtype=bf16)
    E = 2 * M * d  # bytes of one [M, d] bf16 tensor
    rows = [
        ("adaln_fwd", lambda: ops.adaln_modulate_fwd(x, a, b, T, 1e-6), 2 * E),
        ("residual_adaln_fwd", lambda: ops.residual_adaln_fwd(x, y, cc, a, b, T, 1e-6), 4 * E),
        ("gate_fwd (+res)", lambda: ops.gate_fwd(y, cc, x, T), 3 * E),
        ("qk_norm_rope_fwd", lambda: ops.qk_norm_rope_fwd(qkv, cos, sin, 0, H, 1e-6, q, k, v), 6 * E),
        ("attn_fwd", lambda: ops.attn_fwd(q, k, v, **mk), 4 * E),
        ("attn_bwd (delta+dq+dkv)", lambda: ops.attn_bwd(do, o, q, k, v, lse, **mk), 8 * E),
        ("qk_norm_rope_bwd", lambda: ops.qk_norm_rope_bwd(dq, dk, dv, q, k, rq, rk, cos, sin, 0), 8 * E),
        ("adaln_gate_bwd", lambda: ops.adaln_gate_bwd(dhh, x, a, rstd, y, dmod[:, 3 * d:4 * d], dmod[:, 4 * d:5 * d], y, cc,
                                                      dmod[:, 2 * d:3 * d], None, T), 6 * E),
        ("kv_append", lambda: ops.kv_append(k, kc, 128), 2 * E),
        ("gate_bwd (+dbias)", lambda: ops.gate_bwd(dhh, y, cc, dmod[:, 5 * d:], bias32[:d], T, dbias_accumulate=True), 3 * E),
        ("adaln_bwd (+dres)", lambda: ops.adaln_modulate_bwd(dhh, x, a, rstd, y, dmod[:, :d], dmod[:, d:2 * d], T), 4 * E),
        ("colsum [M,3d]", lambda: ops.colsum(qkv.view(M, 3 * d), bias32[:3 * d], accumulate=True), 3 * E),
        ("colsum [M,4d]", lambda: ops.colsum(dy4, bias32, accumulate=True), 4 * E),
        ("cast_f32_bf16 [4d,d]", lambda: ops.cast_f32_bf16(w32, w16), 6 * 4 * d * d),
    ]
    if ONCE or big:
        rows = [r for r in rows if not r[0].startswith("attn_")]
    if ONCE:
        for name, fn, nbytes in rows:
            fn()
            fn()
        torch.cuda.synchronize()
        print("ok")
        return
    print(f"{'op':28s} {'us':>8s} {'MB':>8s} {'GB/s':>8s}")
    tot = 0.0
    for name, fn, nbytes in rows:
        us = timeit(fn)
        tot += us
        print(f"{name:28s} {us:8.1f} {nbytes / 1e6:8.1f} {nbytes / us / 1e3:8.0f}")
    print(f"sum {tot:.1f} us")


if __name__ == "__main__":
    main()
