"""Timeline INSIDE the replayed CUDA graph of one training step: one-thread %globaltimer kernels between the phases of
every block (ops.stamp), averaged over layers.  Main stream only (weight gradients run on the side stream).  The
stamps cost ~2 us each and break the PDL chain, so totals are a little above bench.py's.  GPU box only.
    python tools/step_trace.py [audio|av] [--no-side]"""
import os, sys, types, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import engine, ops
import owl_audio_exps_b200.nn as N
from bench import WORKLOADS, loss_fn

w = WORKLOADS[[a for a in sys.argv[1:] if not a.startswith("--")][0] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else "audio"]
cfg = types.SimpleNamespace(**w["cfg"])
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = N.DiT(cfg).to(dev)
buckets = engine.GradBuckets(model, wgrad_side_stream="--no-side" not in sys.argv)
B, F_, T, d = w["B"], w["F"], cfg.tokens_per_frame, cfg.d_model
L = F_ * T
host = [torch.randn(B, L, d).to(torch.bfloat16), torch.randn(B, F_, d).to(torch.bfloat16), torch.randn(B, L, d).to(torch.bfloat16)]
buf = torch.zeros(4096, device=dev, dtype=torch.int64)
labels = []
ops.TRACE = (buf, labels)
step = engine.TrainStep(model, loss_fn, host, buckets=buckets, use_graph=False)
for _ in range(2):
    labels.clear()
    step(*step.static_in)
labels.clear()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    step._eager()
n = len(labels)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
t = buf[:n].cpu().tolist()
acc = collections.OrderedDict()
for i in range(1, n):
    key = f"{labels[i - 1]:14s} -> {labels[i]}"
    acc.setdefault(key, []).append((t[i] - t[i - 1]) / 1e3)
print(f"step {e0.elapsed_time(e1):.2f} ms with {n} stamps; first->last stamp {(t[-1] - t[0]) / 1e6:.2f} ms")
tot = 0.0
for k, v in acc.items():
    tot += sum(v)
    print(f"{k:34s} n={len(v):3d} mean {sum(v) / len(v):7.1f} us  total {sum(v) / 1e3:6.2f} ms")
print(f"sum of intervals {tot / 1e3:.2f} ms")
