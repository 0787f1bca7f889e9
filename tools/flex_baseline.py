"""The bar SURVEY 2a sets for the attention kernels: torch.compile(flex_attention) with the reference's own mask
predicate (owl_wms/nn/attn.py:43-62) on the same B200, next to this library's kernels at the same shapes.  GPU box only.
    python tools/flex_baseline.py [dit_v4|dit_v4_local|5b|av|long8k ...]
Prints one line per (shape, implementation): fwd / bwd ms and TFLOP/s over the UNMASKED pairs (4*dh flops per pair
forward, 2.5x that backward)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CC", "/usr/bin/gcc")
os.environ.setdefault("CXX", "/usr/bin/g++")
import torch

SHAPES = {
    # name: B, H, L, T, dh, window_frames (0 = global)
    "dit_v4": (1, 24, 98304, 64, 64, 0),
    "dit_v4_local": (1, 24, 98304, 64, 64, 16),
    "5b": (1, 20, 98304, 64, 128, 0),
    "5b_local": (1, 20, 98304, 64, 128, 16),
    "av": (8, 24, 1040, 65, 64, 0),
    "long8k": (1, 24, 8192, 64, 64, 0),
    "long16k": (1, 24, 16384, 64, 64, 0),
}


def pairs(L, T, W):
    F = L // T
    if W <= 0 or W >= F:
        return T * T * F * (F + 1) // 2
    return T * T * (W * (W + 1) // 2 + (F - W) * W)


def timeit(fn, n):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def run_flex(name, B, H, L, T, dh, W):
    from torch.nn.attention.flex_attention import flex_attention, create_block_mask
    import torch._functorch.config as fcfg
    fcfg.donated_buffer = False      # the timing loop re-runs backward on one graph (retain_graph=True)
    torch._dynamo.reset()
    flex = torch.compile(flex_attention)
    cbm = torch.compile(create_block_mask)

    def mask_mod(b, h, q, kv):   # attn.py:43-59 without doc ids
        fq, fk = q // T, kv // T
        m = fk <= fq
        if W > 0:
            m = m & ((fq - fk).abs() < W)
        return m

    t0 = time.time()
    bm = cbm(mask_mod, B=None, H=None, Q_LEN=L, KV_LEN=L, device="cuda")
    torch.cuda.synchronize()
    t_mask = time.time() - t0
    g = lambda: torch.randn(B, H, L, dh, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    q, k, v = g(), g(), g()
    n = 3 if L > 20000 else 10
    with torch.no_grad():
        fwd = timeit(lambda: flex(q, k, v, block_mask=bm), n)
    o = flex(q, k, v, block_mask=bm)
    do = torch.randn_like(o)

    def bwd():
        torch.autograd.grad(o, (q, k, v), do, retain_graph=True)
    b = timeit(bwd, n)
    return fwd, b, t_mask


def run_owl(name, B, H, L, T, dh, W):
    from owl_audio_exps_b200 import ops
    g = lambda *s: torch.randn(*s, device="cuda", dtype=torch.bfloat16)
    q, k, v = g(B, H, L, dh), g(B, H, L, dh), g(B, H, L, dh)
    mk = dict(tokens_per_frame=T, window_frames=W, q_offset=0, causal=True, doc_id=None)
    n = 3 if L > 20000 else 10
    fwd = timeit(lambda: ops.attn_fwd(q, k, v, **mk), n)
    o, lse = ops.attn_fwd(q, k, v, **mk)
    do = g(B, L, H * dh)
    b = timeit(lambda: ops.attn_bwd(do, o, q, k, v, lse, **mk), n)
    return fwd, b, 0.0


def main():
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["dit_v4", "dit_v4_local"]
    impls = ["owl"] if "--owl-only" in sys.argv else (["flex"] if "--flex-only" in sys.argv else ["owl", "flex"])
    for name in names:
        B, H, L, T, dh, W = SHAPES[name]
        fl = 4.0 * dh * B * H * pairs(L, T, W)
        for impl in impls:
            try:
                fwd, bwd, tm = (run_owl if impl == "owl" else run_flex)(name, B, H, L, T, dh, W)
                print(f"{name:14s} {impl:5s} fwd {fwd:9.3f} ms {fl / fwd / 1e9:7.1f} TF/s | bwd {bwd:9.3f} ms "
                      f"{2.5 * fl / bwd / 1e9:7.1f} TF/s | mask build {tm:.2f} s", flush=True)
            except Exception as e:  # noqa: BLE001
                print(f"{name:14s} {impl:5s} FAILED: {type(e).__name__}: {str(e)[:300]}", flush=True)
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
