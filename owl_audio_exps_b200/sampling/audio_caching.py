"""Drop-in for owl_wms/sampling/audio_caching.py:22-160: one audio latent at a time through the KV-cached decode path.

Prefill the noised context with cache updates on, then for every new latent run `n_steps` Euler steps against the cache
(decode mode: no mask), re-noise the result to `noise_prev`, append it to the cache and drop the oldest latent once the
window is full.  Differences from the reference loop: dt lives on the host (no per-step device read), there is no
per-token `empty_cache()`, the noise source is injectable (`noise_fn`) so a rollout is reproducible across devices, and
the Euler steps of one latent replay a CUDA graph: within a latent the cache, its length and the RoPE offset are fixed, so
step 0 runs eagerly (and stages the weights), the same call is captured once and steps 1..n-1 are replays (B=1 decode is
launch-bound: ~200 kernels per call)."""
import torch

from .. import functional as OF
from ..nn.kv_cache import KVCache
from .schedulers import get_sd3_euler


def get_deltas(custom_schedule):
    """audio_caching.py:9-19: |difference| of consecutive schedule points, a final 0.0 appended when missing."""
    pts = list(custom_schedule)
    if pts[-1] != 0.0:
        pts.append(0.0)
    return [abs(b - a) for a, b in zip(pts[:-1], pts[1:])]


class AudioCachingSampler:
    def __init__(self, n_steps: int = 16, num_tokens: int = 120, noise_prev: float = 0.2, custom_schedule=None,
                 max_window=None, noise_fn=None, graph_decode: bool = True, graph_min_steps: int = 10) -> None:
        self.n_steps = n_steps
        self.num_tokens = num_tokens
        self.noise_prev = noise_prev
        self.custom_schedule = custom_schedule
        self.max_window = max_window
        self.noise_fn = noise_fn        # (shape, like) -> N(0,1) tensor; default torch.randn on like's device / dtype
        self.graph_decode = graph_decode
        # capturing + instantiating the decode call costs ~5x its eager time, a replay saves ~60 % of it (measured, audio.yml
        # and av_v5 geometry: tools/model_step.py, tools/av_decode_ops.py): the graph pays off from ~10 steps per latent
        self.graph_min_steps = graph_min_steps
        self._pool = None

    def _denoise(self, velocity, cache, new_rows, cur, t, dt):
        """n_steps Euler steps of one latent / frame against a fixed cache: cur <- cur - dt*v(cur, t), t <- t - dt.
        `velocity(cur, t)` is the model evaluation (two of them under classifier-free guidance)."""
        n = self.n_steps
        if not (self.graph_decode and cur.is_cuda and n >= max(self.graph_min_steps, 2)):
            for step in range(n):
                cur = cur - dt[step] * velocity(cur, t).to(cur.dtype)
                t = t - dt[step]
            return cur
        cache.make_room(new_rows)                         # buffer management happens here, not inside the capture
        cur, t = cur.clone(), t.clone()                   # the graph's input buffers

        def euler(step, v):                               # same arithmetic as the eager loop, written into the graph's inputs
            cur.copy_(cur - dt[step] * v.to(cur.dtype))
            t.copy_(t - dt[step])

        euler(0, velocity(cur, t))                        # step 0, eager
        graph = torch.cuda.CUDAGraph()
        if self._pool is None:
            self._pool = torch.cuda.graph_pool_handle()
        with OF.frozen_weights(), torch.cuda.graph(graph, pool=self._pool):
            v = velocity(cur, t)
        for step in range(1, n):
            graph.replay()
            euler(step, v)
        out = cur.clone()
        del graph, v
        return out

    def _randn(self, shape, like):
        if self.noise_fn is not None:
            return self.noise_fn(tuple(shape), like)
        return torch.randn(*shape, device=like.device, dtype=like.dtype)

    def zlerp(self, x, alpha):
        return x * (1.0 - alpha) + self._randn(x.shape, x) * alpha

    @torch.no_grad()
    def __call__(self, model, x, decode_fn=None, vae_scale=1.0, compile_on_decode=False):
        # parameters do not change during one sampling call: operands are staged once on entry (picking up any EMA /
        # optimizer update made since the last call) and reused by every model evaluation and graph capture inside
        with OF.frozen_weights():
            return self._run(model, x, decode_fn, vae_scale)

    def _run(self, model, x, decode_fn=None, vae_scale=1.0):
        """model: AudioRFTCore; x [b, init_len, channels] -> [b, init_len + num_tokens, channels] (and waveforms when
        decode_fn is given).  `compile_on_decode` is accepted and ignored: the decode step already is hand-written kernels."""
        b, _, channels = x.shape
        if self.custom_schedule is None:
            dt = get_sd3_euler(self.n_steps).tolist()
        else:
            dt = get_deltas(self.custom_schedule)
        cache = KVCache(model.config)
        cache.reset(b)
        latents = [x.clone()]

        def commit(tokens):     # noise to `noise_prev`, run once with cache updates on
            t = tokens.new_full((b, tokens.size(1)), self.noise_prev)
            cache.enable_cache_updates()
            model(self.zlerp(tokens, self.noise_prev), t, doc_id=None, kv_cache=cache)
            cache.disable_cache_updates()

        commit(x)
        model.transformer.enable_decoding()
        try:
            for _ in range(self.num_tokens):
                cur = self._randn((b, 1, channels), x)
                t = x.new_ones(b, 1)
                cur = self._denoise(lambda c, tt: model(c, tt, doc_id=None, kv_cache=cache), cache, 1, cur, t, dt)
                latents.append(cur)
                commit(cur)
                if self.max_window is not None and len(latents) > self.max_window:
                    cache.truncate(1, front=False)
        finally:
            model.transformer.disable_decoding()
        full = torch.cat(latents, dim=1)
        if decode_fn is not None:
            return full, decode_fn(full * vae_scale)
        return full
