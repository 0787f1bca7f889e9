"""owl_audio_exps_b200: B200-native (sm_100a) implementation of the owl_wms DiT-block hot path.

Layout: csrc/ (CUDA kernels + C ABI, built into libowl_dit.so), _lib.py (ctypes binding), ops.py
(tensor-level wrappers), functional.py (autograd schedule of the block), nn/ (drop-in mirror of
owl_wms/nn), models/ + sampling/ (the callers either side of the stack: AudioRFT, GameRFT, AudioCachingSampler), engine.py (data-parallel step: flat gradient buckets, overlapped NCCL all-reduce, CUDA graphs).
"""
import sys

__version__ = "0.1.0"

_NN_MODULES = ("attn", "embeddings", "kv_cache", "mlp", "modulation", "normalization", "rope")


def install(package: str = "owl_wms", patch_compile: bool = True):
    """Make `<package>.nn.{attn,embeddings,kv_cache,mlp,modulation,normalization,rope}` resolve to this
    implementation, so the reference's models / samplers / trainers / inference pipeline (which do
    `from ..nn.attn import DiT, FinalLayer`, `from ..nn.kv_cache import KVCache`, ...) pick up the
    sm_100a path without any edit.  Call before importing `<package>.models`.  See INTEGRATION.md.

    `patch_compile`: also route `torch.compile` through `compat.compile`, so the reference's `torch.compile(model)` /
    `torch.compile(DDP(model))` / `fullgraph=True` lines (rft_trainer.py:96-99, causvid_pipeline.py:63,
    gamerft_audio.py:36) wrap these modules in a pass-through with OptimizedModule's surface instead of handing
    ctypes launches to Dynamo."""
    from . import nn as _nn
    if patch_compile:
        from . import compat
        compat.patch_torch_compile()
    sys.modules[f"{package}.nn"] = sys.modules.get(f"{package}.nn", _nn)
    for name in _NN_MODULES:
        sys.modules[f"{package}.nn.{name}"] = getattr(_nn, name)
    from . import muon as _muon          # `from ..muon import init_muon` in the trainers (av_trainer.py:19)
    sys.modules[f"{package}.muon"] = _muon
    return _nn
