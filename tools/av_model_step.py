"""Full AV model numbers (GPU box only), av_v5_8x8_sf geometry (d=1536, 24 layers, 16 frames x (8x8 video + 1 audio) tokens):
  * GameRFTAudio train step (noise, embeddings incl. controls, proj_in x2, 24 blocks, two heads, two MSEs; fwd+bwd), B=8,
    eager and through engine.TrainStep under the CUDA graph;
  * AVCachingSamplerV2 on GameRFTCore (video only, T=64), B=1, cfg_scale 1.3: frames/s eager vs per-frame graph.
    python tools/av_model_step.py"""
import os, sys, time, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import engine
from owl_audio_exps_b200.models import GameRFTAudio, GameRFTCore
from owl_audio_exps_b200.sampling import AVCachingSamplerV2

dev = "cuda:0"
base = dict(d_model=1536, n_heads=24, n_layers=24, n_frames=16, sample_size=8, causal=True, rope_impl="ortho",
            local_window=16, global_window=None, backbone="dit", channels=128, audio_channels=64, n_buttons=11,
            uncond=False, cfg_prob=0.1, gradient_checkpointing=False)


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


torch.manual_seed(0)
cfg = types.SimpleNamespace(**base, tokens_per_frame=65, has_audio=True)
rft = GameRFTAudio(cfg).to(dev)
B, n = 8, 16
x = torch.randn(B, n, 128, 8, 8, device=dev)
audio = torch.randn(B, n, 64, device=dev)
mouse = torch.randn(B, n, 2, device=dev)
btn = (torch.rand(B, n, 11, device=dev) < 0.3).float()


def train():
    for p in rft.parameters():
        p.grad = None
    rft(x, audio, mouse, btn)[0].backward()


ms = timed(train, 5)
toks = B * n * 65
print(f"GameRFTAudio train step eager B=8: {ms:.1f} ms ({toks / ms * 1e3:.0f} tokens/s)")
buckets = engine.GradBuckets(rft)
step = engine.TrainStep(rft, lambda m, *a: m(*a)[0], [x, audio, mouse, btn], buckets=buckets, use_graph=True, warmup=2)
ms = timed(lambda: step(x, audio, mouse, btn), 10)
print(f"GameRFTAudio train step CUDA graph (engine.TrainStep): {ms:.1f} ms ({toks / ms * 1e3:.0f} tokens/s), "
      f"{step.launches_per_step} launches")
buckets.detach()
del step, rft, buckets
torch.cuda.empty_cache()

cfg = types.SimpleNamespace(**base, tokens_per_frame=64, has_audio=False)
core = GameRFTCore(cfg).to(dev).eval()
ctx = torch.randn(1, 8, 128, 8, 8, device=dev)
mouse = torch.randn(1, 16, 2, device=dev)
btn = (torch.rand(1, 16, 11, device=dev) < 0.3).float()
for graph in (False, True):
    s = AVCachingSamplerV2(n_steps=16, cfg_scale=1.3, num_frames=4, max_window=16, graph_decode=graph)
    s(core, ctx, mouse, btn)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s(core, ctx, mouse, btn)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"AVCachingSamplerV2 B=1, 8-frame context, 4 frames x 16 guided steps, graph_decode={graph}: {dt * 1e3:.0f} ms, "
          f"{4 / dt:.2f} frames/s, {dt * 1e3 / (4 * 33):.2f} ms per model evaluation")
