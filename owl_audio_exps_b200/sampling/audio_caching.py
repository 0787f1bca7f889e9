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
                 max_window=None, noise_fn=None, graph_decode: bool = True, graph_min_steps: int = 10,
                 use_ring: bool = True) -> None:
        self.use_ring = use_ring        # device-state ring cache + ONE decode graph per rollout (nn/decode.py)
        self.last_ring = None
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

    # ---- ring path: ONE graph of the model call for the whole rollout (nn/decode.py) ----------------------------
    def _ring_ok(self, model, b, rows, x):
        from ..nn.attn import DiT
        return (self.graph_decode and self.use_ring and x.is_cuda and isinstance(getattr(model, "transformer", None), DiT)
                and b * rows <= 128 and x.dtype in (torch.bfloat16, torch.float32))

    def _ring_window_rows(self, init_frames, n_new, rows):
        """Most rows the cache holds when a model call starts, by replaying the loop's commit / truncate bookkeeping."""
        frames, n_lat, most = init_frames, 1, init_frames
        for _ in range(n_new):
            frames += 1
            n_lat += 1
            if self.max_window is not None and n_lat > self.max_window:
                frames -= 1
            most = max(most, frames)
        return (most + 1) * rows

    def _ring_rollout(self, model, cache, statics, call, n_new, rows, dt, fresh, renoise, extra_inputs, commit_call=None):
        """statics = [cur, t, *controls] (device buffers the captured call reads); call(ring) -> velocity tensor.
        Per new latent / frame: n_steps replays with the Euler update in between, one more replay on the re-noised
        result to write its K/V rows, then the ring state is advanced (commit + optional truncate).  `commit_call`: the
        model call whose K/V rows are kept when it differs from `call` (guidance evaluates the model twice per step; the
        cache commit of av_caching_v2.py:68-76 is one conditional evaluation)."""
        from ..nn.decode import RingKVCache
        cur, t = statics[0], statics[1]
        ring = RingKVCache.from_cache(cache, self._ring_window_rows(len(cache) // rows, n_new, rows), rows)
        extra_inputs(0)
        cur.copy_(fresh(0))
        t.fill_(1.0)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            call(ring)                                   # eager once: stages weights, creates the kernel workspaces
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        if self._pool is None:
            self._pool = torch.cuda.graph_pool_handle()
        with torch.cuda.graph(graph, pool=self._pool):
            v = call(ring)
        commit_graph = graph
        if commit_call is not None:
            commit_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(commit_graph, pool=self._pool):
                commit_call(ring)
        out, n_lat = [], 1
        for idx in range(n_new):
            if idx > 0:
                extra_inputs(idx)
                cur.copy_(fresh(idx))
                t.fill_(1.0)
            for step in range(self.n_steps):
                graph.replay()
                cur.copy_(cur - dt[step] * v.to(cur.dtype))      # the eager loop's arithmetic, into the graph's inputs
                t.copy_(t - dt[step])
            done = cur.clone()
            out.append(done)
            cur.copy_(renoise(done))
            t.fill_(self.noise_prev)
            commit_graph.replay()                        # writes this latent's K/V rows behind the cached ones
            ring.commit(rows)
            n_lat += 1
            if self.max_window is not None and n_lat > self.max_window:
                ring.truncate(1, front=False)
        self.last_ring = ring
        del graph, commit_graph
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
            if self._ring_ok(model, b, 1, x):
                cur = torch.empty(b, 1, channels, device=x.device, dtype=x.dtype)
                t = torch.empty(b, 1, device=x.device, dtype=x.dtype)
                new = self._ring_rollout(
                    model, cache, [cur, t], lambda ring: model(cur, t, doc_id=None, kv_cache=ring), self.num_tokens, 1, dt,
                    fresh=lambda i: self._randn((b, 1, channels), x), renoise=lambda z: self.zlerp(z, self.noise_prev),
                    extra_inputs=lambda i: None)
                latents.extend(new)
                full = torch.cat(latents, dim=1)
                if decode_fn is not None:
                    return full, decode_fn(full * vae_scale)
                return full
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
