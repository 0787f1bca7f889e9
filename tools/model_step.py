"""Full-model numbers around the block stack (GPU box only), audio.yml:
  * AudioRFT train step (noise + t_embed + proj_in + 16 blocks + FinalLayer + MSE, fwd+bwd), eager, B=32 L=120;
  * AudioCachingSampler decode: ms per model call and latents/s at B=1 against a ~120-latent cache.
    python tools/model_step.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import audio_cfg
from owl_audio_exps_b200.models import AudioRFT
from owl_audio_exps_b200.sampling import AudioCachingSampler

dev = "cuda:0"
cfg = audio_cfg()
torch.manual_seed(0)
rft = AudioRFT(cfg).to(dev)


def timed(fn, n):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


x = torch.randn(32, 120, cfg.channels, device=dev)


def train():
    for p in rft.parameters():
        p.grad = None
    rft(x).backward()


ms = timed(train, 10)
print(f"AudioRFT train step eager B=32 L=120: {ms:.2f} ms  ({32 * 120 / ms * 1e3:.0f} tokens/s)")

from owl_audio_exps_b200 import engine
buckets = engine.GradBuckets(rft)
step = engine.TrainStep(rft, lambda m, x: m(x), [x], buckets=buckets, use_graph=True, warmup=3)
ms = timed(lambda: step(x), 20)
print(f"AudioRFT train step CUDA graph (engine.TrainStep, noise + loss inside) B=32 L=120: {ms:.2f} ms  "
      f"({32 * 120 / ms * 1e3:.0f} tokens/s), {step.launches_per_step} launches")
buckets.detach()
del step

core = rft.core.eval()
calls = [0]
fwd = core.forward


def counted(*a, **k):
    calls[0] += 1
    return fwd(*a, **k)


core.forward = counted
ctx = torch.randn(1, 104, cfg.channels, device=dev)
for n_steps in (4, 16):
    for graph in (False, True):
        sampler = AudioCachingSampler(n_steps=n_steps, num_tokens=16, max_window=120, graph_decode=graph, graph_min_steps=2)
        sampler(core, ctx)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sampler(core, ctx)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"AudioCachingSampler B=1, 104-latent context, 16 latents x {n_steps} steps, graph_decode={graph}: "
              f"{dt * 1e3:.1f} ms total, {dt * 1e3 / (16 * (n_steps + 1)):.2f} ms per model evaluation, {16 / dt:.1f} latents/s")
