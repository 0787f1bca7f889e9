"""compat.compile on the host: OptimizedModule's surface without a GPU (construction, attribute pass-through, keys)."""
import types

import torch

from owl_audio_exps_b200 import compat
import owl_audio_exps_b200.nn as N


def _cfg():
    return types.SimpleNamespace(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=4, n_frames=8, sample_size=2,
                                 causal=True, has_audio=False, rope_impl="motion", rope_ats_delta=2.0, local_window=2,
                                 global_window=None, backbone="dit")


def test_compile_wraps_owl_modules_and_passes_everything_else_on():
    dit = N.DiT(_cfg())
    w = compat.compile(dit, mode="max-autotune", dynamic=False, fullgraph=True)
    assert isinstance(w, compat.CompiledModule) and w._orig_mod is dit
    assert w.config is dit.config and w.blocks is dit.blocks and w.decoding is False
    w.enable_decoding()
    assert dit.decoding is True
    assert list(w.state_dict().keys()) == ["_orig_mod." + k for k in dit.state_dict().keys()]
    assert compat.compile(w) is w
    holder = torch.nn.Sequential(torch.nn.Identity())
    holder.add_module("core", dit)
    assert isinstance(compat.compile(holder), compat.CompiledModule)      # e.g. DDP(model) around the blocks
    deco = compat.compile(fullgraph=True)
    assert isinstance(deco(dit), compat.CompiledModule)


def test_patch_is_reversible():
    real = torch.compile
    compat.patch_torch_compile()
    try:
        assert torch.compile is compat.compile
        compat.patch_torch_compile()
    finally:
        compat.unpatch_torch_compile()
    assert torch.compile is real
