"""Shared helpers for the test-suite: fixture loading, config objects, error metric."""
import os
import types

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["audio_T1", "av_T5_ortho", "video_T4_motion_doc", "av_T5_bidir"]


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), map_location="cpu", weights_only=False)


def cfg_of(fx):
    return types.SimpleNamespace(**fx["cfg"])


def relerr(a, b):
    """max|a-b| / max|b|: the tensor-normalised relative error the parity bar is stated in."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def audio_cfg(**kw):
    """configs/audio.yml model section (reference configs/audio.yml:1-23)."""
    base = dict(model_id="audio_rft", sample_size=120, channels=64, n_layers=16, n_heads=16, d_model=1024,
                tokens_per_frame=1, n_frames=10000, cfg_prob=0.0, causal=True, uncond=True, backbone="dit",
                has_audio=True, rope_impl="audio1d", local_window=16, global_window=None,
                gradient_checkpointing=False)
    base.update(kw)
    return types.SimpleNamespace(**base)
