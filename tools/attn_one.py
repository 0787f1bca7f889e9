"""A few launches of the attention forward / backward kernels at a config's shapes (for ncu).  GPU box only.
    python tools/attn_one.py [audio|av]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from owl_audio_exps_b200 import ops
CFG = {"audio": dict(B=32, L=120, H=16, T=1), "av": dict(B=8, L=1040, H=24, T=65), "long": dict(B=1, L=8192, H=24, T=64),
       "long16k": dict(B=1, L=16384, H=24, T=64), "long16k_dh128": dict(B=1, L=16384, H=20, T=64, dh=128)}
c = CFG[sys.argv[1] if len(sys.argv) > 1 else "av"]
B, L, H, T, dh = c["B"], c["L"], c["H"], c["T"], c.get("dh", 64)
dev, bf16 = "cuda:0", torch.bfloat16
g = lambda *s: torch.randn(*s, device=dev).to(bf16)
q, k, v = g(B, H, L, dh), g(B, H, L, dh), g(B, H, L, dh)
mk = dict(tokens_per_frame=T, window_frames=0, q_offset=0, causal=True, doc_id=None)
for _ in range(2):
    o, lse = ops.attn_fwd(q, k, v, **mk)
    do = g(B, L, H * dh)
    ops.attn_bwd(do, o, q, k, v, lse, **mk)
torch.cuda.synchronize()
print("ok")
