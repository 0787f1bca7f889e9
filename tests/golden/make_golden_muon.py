"""Golden vectors for Muon's Newton-Schulz from the reference's own file (owl_wms/muon.py:11-38), run on CPU in the build
container: `python tests/golden/make_golden_muon.py` -> tests/golden/muon_ns5.pt.  torch.compile is disabled (same math)."""
import os
import sys

os.environ["TORCHDYNAMO_DISABLE"] = "1"
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as G_  # noqa: E402

G_.install_stubs()          # owl_vaes / diffusers / omegaconf import-time stubs, reference on sys.path
from owl_wms.muon import zeropower_via_newtonschulz5  # noqa: E402

g = torch.Generator().manual_seed(0)
Gs = [torch.randn(64, 128, generator=g), torch.randn(192, 64, generator=g), torch.randn(128, 128, generator=g)]
outs = [zeropower_via_newtonschulz5(G, 5).float() for G in Gs]
torch.save({"G": Gs, "out": outs, "steps": 5}, os.path.join(os.path.dirname(os.path.abspath(__file__)), "muon_ns5.pt"))
print("ok", [tuple(o.shape) for o in outs])
