"""owl_audio_exps_b200.install(): the reference's own model / sampler files, unedited, resolve `..nn.*` to this package.
Build-container only (needs the reference checkout; skipped on the GPU box) and run in a subprocess so the aliased
modules do not leak into the other tests."""
import os
import subprocess
import sys

import pytest

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, types
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests/golden")
import make_golden as G
G.install_stubs()                       # owl_vaes / diffusers / omegaconf import-time stubs, reference on sys.path
del sys.modules["rotary_embedding_torch"]   # not needed once nn.rope / nn.embeddings are ours
import owl_audio_exps_b200
owl_audio_exps_b200.install("owl_wms")
from owl_wms.models.audiorft import AudioRFTCore
from owl_wms.models.gamerft import GameRFTCore
import owl_wms.sampling.audio_caching as S
cfg = G.make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=1, n_frames=64, sample_size=24, has_audio=True,
                 rope_impl="audio1d", local_window=4, channels=64)
m = AudioRFTCore(cfg)
mods = {{type(x).__module__ for x in (m.transformer, m.t_embed, m.proj_out, m.transformer.blocks[0].attn.rope)}}
assert all(x.startswith("owl_audio_exps_b200.nn.") for x in mods), mods
assert S.KVCache.__module__ == "owl_audio_exps_b200.nn.kv_cache"
cfg = G.make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=4, n_frames=8, sample_size=2, has_audio=False,
                 rope_impl="motion", local_window=2, channels=16, n_buttons=11, uncond=False)
g = GameRFTCore(cfg)
assert type(g.control_embed).__module__ == "owl_audio_exps_b200.nn.embeddings"
import owl_audio_exps_b200.models as M
ours = M.GameRFTCore(cfg).state_dict()
assert list(g.state_dict().keys()) == list(ours.keys())
assert all(g.state_dict()[k].shape == v.shape for k, v in ours.items())
print("ok")
"""


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")
def test_reference_models_resolve_to_this_package():
    out = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT)], capture_output=True, text=True, timeout=300,
                         env=dict(os.environ, TORCHDYNAMO_DISABLE="1"))
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
