"""Drop-in for owl_wms/sampling/av_caching_v2.py:24-144: one video frame at a time through the KV-cached decode path of a
GameRFTCore-style model `model(x, t, mouse, btn, kv_cache=...)`, with classifier-free guidance on the controls.

Same rollout as the audio sampler (prefill the noised context, n_steps Euler steps per new frame against the cache, re-noise
and append, drop the oldest frame once the window is full); with cfg_scale != 1 every step evaluates the model on the real
and on zeroed controls and extrapolates.  Both evaluations of a step sit in the per-frame CUDA graph."""
import torch

from .. import functional as OF
from ..nn.kv_cache import KVCache
from .audio_caching import AudioCachingSampler, get_deltas
from .schedulers import get_sd3_euler


class AVCachingSamplerV2(AudioCachingSampler):
    def __init__(self, n_steps: int = 16, cfg_scale: float = 1.3, num_frames: int = 60, noise_prev: float = 0.2,
                 max_window=None, custom_schedule=None, noise_fn=None, graph_decode: bool = True,
                 graph_min_steps: int = 10, use_ring: bool = True) -> None:
        super().__init__(n_steps=n_steps, num_tokens=num_frames, noise_prev=noise_prev, custom_schedule=custom_schedule,
                         max_window=max_window, noise_fn=noise_fn, graph_decode=graph_decode,
                         graph_min_steps=graph_min_steps, use_ring=use_ring)
        self.cfg_scale = cfg_scale
        self.num_frames = num_frames

    @torch.no_grad()
    def __call__(self, model, x, mouse, btn, compile_on_decode=False):
        with OF.frozen_weights():    # see AudioCachingSampler.__call__
            return self._run(model, x, mouse, btn)

    def _run(self, model, x, mouse, btn):
        """x [b, init_len, c, h, w], mouse [b, n, 2], btn [b, n, n_buttons] (n >= init_len: controls of the frames to
        generate follow those of the context) -> [b, init_len + frames, c, h, w]."""
        b, init_len = x.size(0), x.size(1)
        dt = get_sd3_euler(self.n_steps).tolist() if self.custom_schedule is None else get_deltas(self.custom_schedule)
        cache = KVCache(model.config)
        cache.reset(b)
        latents = [x.clone()]
        rows = model.config.tokens_per_frame

        def commit(frames, m, bt):
            t = frames.new_full((b, frames.size(1)), self.noise_prev)
            cache.enable_cache_updates()
            model(self.zlerp(frames, self.noise_prev), t, m, bt, kv_cache=cache)
            cache.disable_cache_updates()

        commit(x, mouse[:, :init_len], btn[:, :init_len])
        n_frames = min(self.num_frames, mouse.size(1) - init_len)
        model.transformer.enable_decoding()
        try:
            if self._ring_ok(model, b, rows, x):
                cur = torch.empty_like(x[:, :1])
                t = torch.empty(b, 1, device=x.device, dtype=x.dtype)
                m = torch.empty_like(mouse[:, :1])
                bt = torch.empty_like(btn[:, :1])
                m0, bt0 = torch.zeros_like(m), torch.zeros_like(bt)

                def controls(idx):
                    at = init_len + idx
                    m.copy_(mouse[:, at:at + 1])
                    bt.copy_(btn[:, at:at + 1])

                def call(ring):
                    v = model(cur, t, m, bt, kv_cache=ring)
                    if self.cfg_scale != 1.0:
                        v0 = model(cur, t, m0, bt0, kv_cache=ring)
                        v = v0 + self.cfg_scale * (v - v0)
                    return v

                commit_call = (lambda ring: model(cur, t, m, bt, kv_cache=ring)) if self.cfg_scale != 1.0 else None
                new = self._ring_rollout(model, cache, [cur, t, m, bt], call, n_frames, rows, dt,
                                         fresh=lambda i: self._randn(x[:, :1].shape, x),
                                         renoise=lambda z: self.zlerp(z, self.noise_prev), extra_inputs=controls,
                                         commit_call=commit_call)
                return torch.cat(latents + new, dim=1)
            for idx in range(n_frames):
                cur = self._randn(x[:, :1].shape, x)
                t = x.new_ones(b, 1)
                at = init_len + idx
                m, bt = mouse[:, at:at + 1].contiguous(), btn[:, at:at + 1].contiguous()
                m0, bt0 = torch.zeros_like(m), torch.zeros_like(bt)

                def velocity(c, tt):
                    v = model(c, tt, m, bt, kv_cache=cache)
                    if self.cfg_scale != 1.0:
                        v0 = model(c, tt, m0, bt0, kv_cache=cache)
                        v = v0 + self.cfg_scale * (v - v0)
                    return v

                cur = self._denoise(velocity, cache, rows, cur, t, dt)
                latents.append(cur)
                commit(cur, m, bt)
                if self.max_window is not None and len(latents) > self.max_window:
                    cache.truncate(1, front=False)
        finally:
            model.transformer.disable_decoding()
        return torch.cat(latents, dim=1)
