"""cProfile of one eager B=1 decode call (audio.yml, 120-latent cache): where the host time of ~200 launches goes.
    python tools/decode_profile.py"""
import cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import audio_cfg
from owl_audio_exps_b200.models import AudioRFTCore
from owl_audio_exps_b200.nn import KVCache

dev = "cuda:0"
cfg = audio_cfg()
core = AudioRFTCore(cfg).to(dev).eval()
with torch.no_grad():
    cache = KVCache(cfg)
    cache.reset(1)
    cache.enable_cache_updates()
    core(torch.randn(1, 119, cfg.channels, device=dev), torch.full((1, 119), 0.2, device=dev), kv_cache=cache)
    cache.disable_cache_updates()
    core.transformer.enable_decoding()
    x, t = torch.randn(1, 1, cfg.channels, device=dev), torch.ones(1, 1, device=dev)
    for _ in range(3):
        core(x, t, kv_cache=cache)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(10):
        core(x, t, kv_cache=cache)
    torch.cuda.synchronize()
    pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
