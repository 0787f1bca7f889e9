"""Host-side bookkeeping of the preallocated KV cache against the reference's list-of-tensors semantics
(kv_cache.py:12-104, restated in oracle.OracleKVCache).  CPU tensors: buffer management only, no kernels."""
import types

import torch

from oracle.dit_oracle import OracleKVCache
from owl_audio_exps_b200.nn.kv_cache import KVCache

CFG = types.SimpleNamespace(n_layers=3, n_heads=2, d_model=16, tokens_per_frame=5, backbone="dit")


def _mk(B=2):
    c = KVCache(CFG).to("cpu", torch.float32)
    c.reset(B)
    return c


def _rows(n, seed, B=2):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, CFG.n_heads, n, 8, generator=g), torch.randn(B, CFG.n_heads, n, 8, generator=g)


def test_reset_state():
    c = KVCache(CFG)
    assert c.cache is None and c.offsets == [0, 0, 0] and c.should_update is False
    c = _mk()
    assert len(c) == 0 and c.length_at(2) == 0 and tuple(c.shape) == (2, 2, 0, 8)
    assert len(c.cache) == 3 and c.cache[0][0].shape == (2, 2, 0, 8)


def test_update_truncate_follow_reference():
    c, o = _mk(), OracleKVCache(CFG.n_layers, CFG.tokens_per_frame)
    for step in range(12):
        for i in range(CFG.n_layers):
            k, v = _rows(5, 100 * step + i)
            if o.cache[i] is not None:
                ok, ov = o.cache[i]
                k, v = torch.cat([ok, k], 2), torch.cat([ov, v], 2)
            o.update(k, v, i)
            c.update(k, v, i)
        if step % 3 == 2:
            o.truncate(1, front=False)
            c.truncate(1, front=False)
        if step == 7:
            o.truncate(1, front=True)
            c.truncate(1, front=True)
        for i in range(CFG.n_layers):
            assert c.length_at(i) == o.length_at(i)
            assert c.get_offset(i) == o.offsets[i]
            assert torch.equal(c.get(i)[0], o.cache[i][0]) and torch.equal(c.get(i)[1], o.cache[i][1])
    assert c.n_frames() == len(c) // 5


def test_in_place_tail_and_commit():
    """The hot-path hooks: rows written behind the live region become cache content on commit, and a
    non-committed write (denoise step) leaves the cache unchanged."""
    c = _mk()
    c.enable_cache_updates()
    k_all, v_all = c._operands(0, 10, use_old=False, private=False)
    nk, nv = _rows(10, 1)
    k_all.copy_(nk), v_all.copy_(nv)
    c._commit(0, 10, k_all, v_all)
    assert c.length_at(0) == 10 and c.get_offset(0) == 10
    c.disable_cache_updates()
    k2, v2 = c._operands(0, 5, use_old=True, private=False)   # scratch rows behind the cache
    assert k2.shape[2] == 15 and torch.equal(k2[:, :, :10], nk)
    k2[:, :, 10:] = 7.0
    assert c.length_at(0) == 10 and torch.equal(c.get(0)[0], nk)
    c.enable_cache_updates()
    k3, v3 = c._operands(0, 5, use_old=True, private=False)
    mk, mv = _rows(5, 2)
    k3[:, :, 10:] = mk
    v3[:, :, 10:] = mv
    c._commit(0, 5, k3, v3)
    assert c.length_at(0) == 15 and c.get_offset(0) == 15
    assert torch.equal(c.get(0)[0], torch.cat([nk, mk], 2))
    # sliding window: drop the oldest frame many times, cache keeps its content through compaction/growth
    want = torch.cat([nk, mk], 2)
    for s in range(40):
        c.truncate(1, front=False)
        want = want[:, :, 5:]
        k4, v4 = c._operands(0, 5, use_old=True, private=False)
        a, b = _rows(5, 10 + s)
        k4[:, :, -5:] = a
        v4[:, :, -5:] = b
        c._commit(0, 5, k4, v4)
        want = torch.cat([want, a], 2)
        assert torch.equal(c.get(0)[0], want)
    assert c.get_offset(0) == 15 + 40 * 5 and c.length_at(0) == 15


def test_clone_detach_noise_and_cache_setter():
    c = _mk()
    k, v = _rows(10, 5)
    for i in range(CFG.n_layers):
        c.update(k, v, i)
    p = c._k[0].data_ptr()
    assert c.clone() is c and c._k[0].data_ptr() != p and torch.equal(c.get(0)[0], k)
    assert c.detach() is c
    c.noise_caches = 0.5
    assert not torch.equal(c.get(0)[0], k)
    c.noise_caches = 0.0
    c.cache = [(k[:, :, :5], v[:, :, :5])] * CFG.n_layers
    assert c.length_at(1) == 5 and torch.equal(c.cache[1][1], v[:, :, :5])
    c.reserve(64).reset(1)
    assert c._k[0].shape[2] == 64 and len(c) == 0
