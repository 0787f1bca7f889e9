"""Do two concurrent streams of persistent GEMMs fill each other's partial waves?  Backward GEMMs of one audio
block: dgrad chain on one stream, wgrad chain on another, vs everything on one stream."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops
dev, bf16 = "cuda:0", torch.bfloat16
M, d = 3840, 1024
g = lambda *s: torch.randn(*s, device=dev).to(bf16)
dgr, wgr = [], []
for N, K, mode in [(d, 4 * d, "dsilu"), (4 * d, d, "store"), (d, d, "store"), (3 * d, d, "store"), (6 * d, d, "store")]:
    dy, w, out = g(M, N), g(N, K), torch.empty(M, K, device=dev, dtype=bf16)
    kw = dict(mode=ops.OWL_EPI_STORE, out0=out) if mode == "store" else dict(mode=ops.OWL_EPI_DSILU, out0=out, aux0=g(M, K))
    dgr.append((lambda dy=dy, w=w, N=N, K=K, kw=kw: ops.gemm(1, dy, w, M, K, N, **kw)))
    x, of = g(M, K), torch.zeros(N, K, device=dev)
    wgr.append((lambda dy=dy, x=x, N=N, K=K, of=of: ops.gemm(2, dy, x, N, K, M, mode=ops.OWL_EPI_F32, outf=of)))
side = torch.cuda.Stream()

def seq():
    for a, b in zip(dgr, wgr):
        a(); b()

def par():
    ev = torch.cuda.Event(); ev.record()
    with torch.cuda.stream(side):
        side.wait_event(ev)
        for b in wgr:
            b()
    for a in dgr:
        a()
    torch.cuda.current_stream().wait_stream(side)

def graphed(fn):
    fn(); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(4):
            fn()
    return gr

for name, fn in (("one stream", seq), ("two streams", par)):
    gr = graphed(fn)
    for _ in range(3):
        gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10):
        gr.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 40 * 1e3:.1f} us per block-backward GEMM set")
