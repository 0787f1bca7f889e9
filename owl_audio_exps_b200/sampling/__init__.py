"""Samplers that drive the KV-cached decode path, mirroring owl_wms/sampling."""
from .audio_caching import AudioCachingSampler, get_deltas  # noqa: F401
from .av_caching_v2 import AVCachingSamplerV2  # noqa: F401
from .schedulers import get_sd3_euler  # noqa: F401


def get_sampler_cls(sampler_id):
    """owl_wms/sampling/__init__.py:1-40; the KV-cached samplers are built here, the window samplers consume the drop-in
    unchanged."""
    if sampler_id == "audio_caching":
        return AudioCachingSampler
    if sampler_id == "av_caching":
        return AVCachingSamplerV2
    raise NotImplementedError(f"sampler '{sampler_id}': use the reference's owl_wms.sampling on top of owl_audio_exps_b200.install()")
