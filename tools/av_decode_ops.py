"""Per-kernel GPU time of one AV decode evaluation (GameRFTCore, d=1536, 24 layers, 64 tokens against a 512-token cache, B=1).
    python tools/av_decode_ops.py"""
import os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops
from owl_audio_exps_b200.models import GameRFTCore
from owl_audio_exps_b200.nn import KVCache

dev = "cuda:0"
cfg = types.SimpleNamespace(d_model=1536, n_heads=24, n_layers=24, n_frames=16, sample_size=8, causal=True, rope_impl="ortho",
                            local_window=16, global_window=None, backbone="dit", channels=128, n_buttons=11, uncond=False,
                            cfg_prob=0.0, gradient_checkpointing=False, tokens_per_frame=64, has_audio=False)
core = GameRFTCore(cfg).to(dev).eval()
with torch.no_grad():
    cache = KVCache(cfg)
    cache.reset(1)
    cache.enable_cache_updates()
    core(torch.randn(1, 8, 128, 8, 8, device=dev), torch.full((1, 8), 0.2, device=dev), torch.randn(1, 8, 2, device=dev),
         torch.zeros(1, 8, 11, device=dev), kv_cache=cache)
    cache.disable_cache_updates()
    core.transformer.enable_decoding()
    a = (torch.randn(1, 1, 128, 8, 8, device=dev), torch.ones(1, 1, device=dev), torch.randn(1, 1, 2, device=dev),
         torch.zeros(1, 1, 11, device=dev))
    for _ in range(3):
        core(*a, kv_cache=cache)
    torch.cuda.synchronize()
    if os.environ.get("OWL_NCU_ONE"):      # under ncu: mark the one evaluation to list, then stop
        torch.cuda.nvtx.range_push("decode_eval")
        core(*a, kv_cache=cache)
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        sys.exit(0)
    ops.profile_begin()
    core(*a, kv_cache=cache)
    torch.cuda.synchronize()
    rec = ops.profile_end()
tot = 0.0
for k, v in sorted(rec.items(), key=lambda kv: -kv[1]["ms"]):
    tot += v["ms"]
    print(f"{k:24s} n={v['n']:4d}  {v['ms'] * 1e3:9.1f} us  ({v['ms'] * 1e3 / v['n']:7.1f} us each)")
print(f"sum {tot * 1e3:.0f} us")

# the same evaluation as a CUDA graph: GPU time per replay, capture + instantiate cost
import time
from owl_audio_exps_b200 import functional as OF
with torch.no_grad():
    cache.make_room(64)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = torch.cuda.CUDAGraph()
    with OF.frozen_weights(), torch.cuda.graph(g):
        out = core(*a, kv_cache=cache)
    torch.cuda.synchronize()
    t_cap = time.perf_counter() - t0
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"graph: capture+instantiate {t_cap * 1e3:.1f} ms, replay {e0.elapsed_time(e1) / 20 * 1e3:.0f} us per evaluation")
