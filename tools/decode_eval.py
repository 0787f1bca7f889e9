"""One KV-cached decode evaluation of the causvid.yml model (GameRFTAudioCore, 24 layers, d=1536, 65 tokens per frame) on
the ring-cache path: eager and as graph replays, ms per evaluation.  `--once`: two eager evaluations only (for an ncu
launch list).  GPU box only.
    python tools/decode_eval.py [--once] [--layers N]"""
import os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import functional as OF, ops
from owl_audio_exps_b200.models import GameRFTAudioCore
from owl_audio_exps_b200.nn.decode import RingKVCache
import owl_audio_exps_b200.nn as N

dev = "cuda:0"
layers = int(sys.argv[sys.argv.index("--layers") + 1]) if "--layers" in sys.argv else 24
cfg = types.SimpleNamespace(d_model=1536, n_heads=24, n_layers=layers, tokens_per_frame=65, n_frames=40, sample_size=8, channels=64,
                            audio_channels=64, n_buttons=11, n_mouse_axes=2, causal=True, uncond=False, has_audio=True,
                            rope_impl="ortho", local_window=16, global_window=None, backbone="dit", cfg_prob=0.0)
torch.manual_seed(0)
model = GameRFTAudioCore(cfg).to(dev).bfloat16().eval()
T, ctx = 65, 15
with torch.no_grad(), OF.frozen_weights():
    cache = N.KVCache(cfg)
    cache.reset(1)
    cache.enable_cache_updates()
    g = lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16)
    model(g(1, ctx, 64, 8, 8), g(1, ctx, 64), torch.full((1, ctx), 0.2, device=dev, dtype=torch.bfloat16), g(1, ctx, 2),
          g(1, ctx, 11), kv_cache=cache)
    cache.disable_cache_updates()
    model.transformer.enable_decoding()
    ring = RingKVCache.from_cache(cache, window_rows=16 * T, new_rows=T)
    x, a, t, m, b = g(1, 1, 64, 8, 8), g(1, 1, 64), torch.ones(1, 1, device=dev, dtype=torch.bfloat16), g(1, 1, 2), g(1, 1, 11)
    call = lambda: model(x, a, t, m, b, kv_cache=ring)
    call()
    torch.cuda.synchronize()
    if "--once" in sys.argv:
        call()
        torch.cuda.synchronize()
        print("ok")
        sys.exit(0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = ops.launch_count()
    e0.record()
    for _ in range(10):
        call()
    e1.record()
    torch.cuda.synchronize()
    print(f"eager: {e0.elapsed_time(e1) / 10:.3f} ms per evaluation, {(ops.launch_count() - n0) // 10} library launches")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        call()
    torch.cuda.current_stream().wait_stream(side)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        call()
    gr.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"graph: {e0.elapsed_time(e1) / 20:.3f} ms per evaluation")
    for pdl in ("0",):
        pass
