"""KV cache for causal rollout: drop-in for owl_wms/nn/kv_cache.py (reference lines cited inline).

The reference keeps, per layer, a tuple of tensors and replaces it with `torch.cat([old, new])` on
every call (attn.py:92-99, kv_cache.py:50-58): O(cache) bytes moved per layer per call.  Here each
layer owns one preallocated [B, H, capacity, dh] buffer per K and V; the QK-norm+RoPE kernel writes
the new (post-norm, post-RoPE) K rows and raw V rows straight behind the cached rows, `update`
becomes a length bump, `truncate` a start-pointer bump, and rows behind `length` double as scratch
for denoise steps that read but do not extend the cache.  The Python API and observable semantics
(`offsets`, `length_at`, `get`, `cache`, `clone/detach`, `noise_caches`, `shape`) are the reference's.
"""
from __future__ import annotations

import torch

from .. import ops


def KVCache(config):
    """kv_cache.py:5-9."""
    if getattr(config, "backbone", "dit") in ("dit", "mmdit"):
        return SingleKVCache(config)
    raise ValueError(f"Invalid backbone: {config.backbone}")


class SingleKVCache:
    MIN_CAPACITY = 256

    def __init__(self, config):
        self.config = config
        self.device = "cuda"
        self.dtype = torch.bfloat16
        self.should_update = False
        self.noise_caches = 0.0
        self.offsets = [0] * config.n_layers
        self._k = None        # per layer [B, H, cap, dh]
        self._v = None
        self._start = None    # per layer first live row
        self._len = None      # per layer live rows
        self._batch = 0
        self._reserve = 0

    # ---- kv_cache.py:25-34 -------------------------------------------------------------------
    def enable_cache_updates(self):
        self.should_update = True

    def disable_cache_updates(self):
        self.should_update = False

    def to(self, device="cuda", dtype=torch.bfloat16):
        self.device = device
        self.dtype = dtype
        return self

    # ---- sizing ------------------------------------------------------------------------------
    def reserve(self, n_tokens: int):
        """Preallocate room for `n_tokens` rows per layer at the next reset/growth (keeps device
        pointers stable, e.g. for CUDA-graph capture of the decode step).  Extension, not in the reference."""
        self._reserve = int(n_tokens)
        return self

    def make_room(self, n_new: int):
        """Do now any compaction / growth the next `n_new`-row call would trigger, so that call finds its rows in place
        (a CUDA-graph capture of the decode step must not record buffer management).  Extension, not in the reference."""
        self._check()
        for i in range(self.config.n_layers):
            self._ensure_room(i, n_new)
        return self

    @property
    def _head_dim(self):
        return self.config.d_model // self.config.n_heads

    def _alloc(self, cap: int):
        return torch.empty(self._batch, self.config.n_heads, cap, self._head_dim, device=self.device, dtype=self.dtype)

    def reset(self, batch_size=1):
        """kv_cache.py:36-40: empty caches, offsets back to 0."""
        n = self.config.n_layers
        self._batch = batch_size
        cap = max(self._reserve, 0)
        self._k = [self._alloc(cap) for _ in range(n)]
        self._v = [self._alloc(cap) for _ in range(n)]
        self._start = [0] * n
        self._len = [0] * n
        self.offsets = [0] * n

    def _check(self):
        assert self._k is not None, "Must reset cache before using"

    def _ensure_room(self, i: int, extra: int):
        """Make rows [start+len, start+len+extra) of layer i addressable, compacting or growing."""
        k, v, st, ln = self._k[i], self._v[i], self._start[i], self._len[i]
        cap = k.shape[2]
        if st + ln + extra <= cap:
            return
        if ln + extra <= cap and st > 0:  # slide live rows to the front (amortised O(1) per appended row)
            if ln > 0:  # buffer management (like realloc), not the per-step append
                k[:, :, :ln] = k[:, :, st:st + ln].clone() if st < ln else k[:, :, st:st + ln]
                v[:, :, :ln] = v[:, :, st:st + ln].clone() if st < ln else v[:, :, st:st + ln]
            self._start[i] = 0
            return
        new_cap = max(2 * (ln + extra), self.MIN_CAPACITY, self._reserve)
        nk, nv = self._alloc(new_cap), self._alloc(new_cap)
        if ln > 0:
            nk[:, :, :ln] = k[:, :, st:st + ln]
            nv[:, :, :ln] = v[:, :, st:st + ln]
        self._k[i], self._v[i], self._start[i] = nk, nv, 0

    # ---- hot-path hooks used by functional._attn_core_fwd --------------------------------------
    def _operands(self, i: int, n_new: int, use_old: bool, private: bool, batch=None, device=None):
        """(k_all, v_all) views of [B, H, n_old + n_new, dh]: cached rows followed by n_new writable
        rows.  `private`: return a copy that does not alias the cache buffers.  `batch` / `device`: those of the
        incoming qkv -- the kernel sizes its writes from qkv, so a mismatch must fail here, on the host."""
        self._check()
        if batch is not None and self.dtype != torch.bfloat16:   # (called from the kernel path: batch / device given)
            raise RuntimeError(f"KV cache dtype {self.dtype}: the sm_100a attention path stores K/V in bf16 "
                               "(kv_cache.py:18 default); call cache.to(dtype=torch.bfloat16)")
        if device is not None and torch.device(self.device).type != device.type:
            raise RuntimeError(f"KV cache lives on {self.device}, activations on {device}")
        if batch is not None and batch != self._batch:
            if use_old and self._len[i] > 0:   # the reference fails in torch.cat (attn.py:94) in this case
                raise RuntimeError(f"KV cache holds batch {self._batch}, forward has batch {batch}")
            # empty cache: the reference accepts any batch on the first update (kv_cache.py:50-58)
            if any(self._len):
                raise RuntimeError(f"KV cache holds batch {self._batch} in other layers, forward has batch {batch}")
            off = list(self.offsets)
            self.reset(batch)
            self.offsets = off
        n_old = self._len[i] if use_old else 0
        if not use_old and self.should_update:
            self._start[i], self._len[i] = 0, 0  # offset 0: cache restarts (attn.py:92 skips the prepend)
        if private or self.noise_caches > 0.0:
            k_all = torch.empty(self._batch, self.config.n_heads, n_old + n_new, self._head_dim,
                                device=self._k[i].device, dtype=self.dtype)
            v_all = torch.empty_like(k_all)
            if n_old > 0:
                k_old, v_old = self.get(i)  # applies noise_caches (kv_cache.py:45-47)
                if self.noise_caches > 0.0:
                    k_all[:, :, :n_old] = k_old
                    v_all[:, :, :n_old] = v_old
                else:
                    ops.kv_append(k_old, k_all, 0)
                    ops.kv_append(v_old, v_all, 0)
            return k_all, v_all
        self._ensure_room(i, n_new)
        st = self._start[i]
        return (self._k[i][:, :, st:st + n_old + n_new], self._v[i][:, :, st:st + n_old + n_new])

    def _commit(self, i: int, n_new: int, k_all, v_all):
        """The rows written behind the cache become part of it (the reference's update(), kv_cache.py:50-58)."""
        st = self._start[i]
        buf = self._k[i]
        if k_all.data_ptr() == buf.data_ptr() + st * buf.stride(2) * buf.element_size() and \
                k_all.stride() == buf.stride():
            self._len[i] = k_all.shape[2]
            self.offsets[i] += n_new
        else:
            self.update(k_all, v_all, i)

    # ---- reference API -----------------------------------------------------------------------
    def get(self, layer_ind):
        """kv_cache.py:42-48."""
        self._check()
        st, ln = self._start[layer_ind], self._len[layer_ind]
        k = self._k[layer_ind][:, :, st:st + ln]
        v = self._v[layer_ind][:, :, st:st + ln]
        if self.noise_caches > 0.0:
            k = k + torch.randn_like(k) * self.noise_caches
            v = v + torch.randn_like(v) * self.noise_caches
        return k, v

    def update(self, new_k, new_v, layer_ind):
        """kv_cache.py:50-58: `new_k/new_v` are the FULL concatenated tensors."""
        self._check()
        old_len = self.length_at(layer_ind)
        new_len = new_k.shape[2]
        self.offsets[layer_ind] += new_len - old_len
        new_k, new_v = new_k.detach(), new_v.detach()
        if new_len > 0 and self._k[layer_ind].numel() > 0 and \
                new_k.untyped_storage().data_ptr() == self._k[layer_ind].untyped_storage().data_ptr():
            new_k, new_v = new_k.clone(), new_v.clone()  # caller handed back a view of this cache
        self._start[layer_ind], self._len[layer_ind] = 0, 0
        if new_k.shape[0] != self._batch:
            self._batch = new_k.shape[0]
            self._k[layer_ind], self._v[layer_ind] = self._alloc(0), self._alloc(0)
        self._ensure_room(layer_ind, new_len)
        if new_len > 0:
            self._k[layer_ind][:, :, :new_len] = new_k
            self._v[layer_ind][:, :, :new_len] = new_v
        self._len[layer_ind] = new_len

    def truncate(self, truncate_amt, front=False):
        """kv_cache.py:60-75: eject frames; front=False drops the OLDEST frames, front=True the newest."""
        self._check()
        amt = truncate_amt * self.config.tokens_per_frame
        for i in range(self.config.n_layers):
            amt_i = min(amt, self._len[i])
            if not front:
                self._start[i] += amt_i
            self._len[i] -= amt_i

    def length_at(self, idx):
        self._check()
        return self._len[idx]

    def get_offset(self, idx=0):
        return self.offsets[idx]

    def __len__(self):
        self._check()
        return self._len[0]

    def n_frames(self):
        assert len(self) % self.config.tokens_per_frame == 0
        return len(self) // self.config.tokens_per_frame

    def clone(self):
        """kv_cache.py:93-97 (fresh storage for every layer)."""
        self._check()
        for i in range(self.config.n_layers):
            st, ln = self._start[i], self._len[i]
            nk, nv = self._alloc(self._k[i].shape[2]), self._alloc(self._k[i].shape[2])
            if ln > 0:
                nk[:, :, :ln] = self._k[i][:, :, st:st + ln]
                nv[:, :, :ln] = self._v[i][:, :, st:st + ln]
            self._k[i], self._v[i], self._start[i] = nk, nv, 0
        return self

    def detach(self):
        """kv_cache.py:99-102: the buffers never carry autograd history here."""
        return self

    @property
    def cache(self):
        """List of per-layer (k, v) views, as the reference attribute (None before reset)."""
        if self._k is None:
            return None
        out = []
        for i in range(self.config.n_layers):
            st, ln = self._start[i], self._len[i]
            out.append((self._k[i][:, :, st:st + ln], self._v[i][:, :, st:st + ln]))
        return out

    @cache.setter
    def cache(self, value):
        if value is None:
            self._k = self._v = self._start = self._len = None
            return
        if self._k is None:
            self.reset(value[0][0].shape[0])
        for i, (k, v) in enumerate(value):
            off = self.offsets[i]
            self.update(k, v, i)
            self.offsets[i] = off

    @property
    def shape(self):
        self._check()
        return torch.Size((self._batch, self.config.n_heads, self._len[0], self._head_dim))
