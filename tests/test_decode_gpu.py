"""Decode regime (ring KV cache with device-resident state): the kernels against dense fp32 math, the ring schedule of a
DiT stack against the reference's golden KV rollout and against the oracle's cached decode, and one CUDA graph replayed
over every frame of a rollout.  Bar: 2e-2 (north_star) on activations; ring path == graph replay bit for bit."""
import math
import types

import pytest
import torch

from helpers import cfg_of, load_golden, relerr
from oracle import dit_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
bf16 = torch.bfloat16


def _ops():
    from owl_audio_exps_b200 import ops
    return ops


def _nn():
    import owl_audio_exps_b200.nn as N
    return N


ATTN_CASES = [
    # B, H, Lq, dh, cap, n_valid, start, vis_rows
    (1, 24, 65, 64, 1105, 975, 0, 0),          # AV geometry, full window: Lk = 1040
    (1, 24, 65, 64, 1105, 1040, 390, 0),       # wrapped ring, full
    (1, 24, 65, 64, 1105, 1040, 1100, 260),    # wrapped, local layer: last 4 frames of [cache, new]
    (2, 3, 1, 64, 300, 120, 250, 16),          # audio token, window 16, wrap
    (1, 2, 65, 128, 700, 130, 0, 0),           # dh = 128
    (1, 24, 64, 64, 61440 + 64, 61440, 0, 0),  # dit_v4 sampler: Lk = 61 504, whole-machine split
    (1, 2, 128, 64, 256, 0, 0, 0),             # empty cache: only the new rows
    (3, 20, 7, 64, 100, 93, 50, 0),            # more heads than SMs / split: n_splits = 1
]


@pytest.mark.parametrize("case", ATTN_CASES, ids=lambda c: "-".join(map(str, c)))
def test_attn_decode_vs_dense(case):
    ops = _ops()
    B, H, Lq, dh, cap, n_valid, start, vis = case
    g = torch.Generator().manual_seed(sum(case))
    q = torch.randn(B, H, Lq, dh, generator=g).to(bf16).to(DEV)
    k = torch.randn(B, H, cap, dh, generator=g).to(bf16).to(DEV)
    v = torch.randn(B, H, cap, dh, generator=g).to(bf16).to(DEV)
    n_tot = n_valid + Lq
    idx = (start + torch.arange(n_tot)) % cap
    live = torch.zeros(cap, dtype=torch.bool)
    live[idx] = True
    dead = (~live).to(DEV)
    k[:, :, dead] = 0        # rows outside the ring are zero (stale rows are finite in the product path)
    v[:, :, dead] = 0
    state = torch.tensor([n_valid, start, 0, 0], dtype=torch.int32, device=DEV)
    o = ops.attn_decode(q, k, v, state, vis)
    if vis > 0 and n_tot > vis:
        idx = idx[n_tot - vis:]
    kk, vv = k[:, :, idx.to(DEV)].float(), v[:, :, idx.to(DEV)].float()
    s = (q.float() @ kk.transpose(-1, -2)) / math.sqrt(dh)
    ref = (torch.softmax(s, -1) @ vv).permute(0, 2, 1, 3).reshape(B, Lq, H * dh)
    torch.cuda.synchronize()
    assert relerr(o, ref) <= 1e-2
    # same answer from the training-shape kernel on the gathered rows (decode mask: all keys visible)
    o2, _ = ops.attn_fwd(q, k[:, :, idx.to(DEV)].contiguous(), v[:, :, idx.to(DEV)].contiguous(), tokens_per_frame=1,
                         window_frames=0, q_offset=0, causal=False, want_lse=False)
    assert relerr(o, o2) <= 1e-2


@pytest.mark.parametrize("dh,H,L,cap,n_valid,start", [(64, 24, 65, 1105, 1040, 1100), (128, 2, 3, 64, 10, 60), (64, 16, 1, 40, 0, 0)])
def test_qk_norm_rope_decode_vs_training_kernel(dh, H, L, cap, n_valid, start):
    ops = _ops()
    B, pos = 2, 37
    g = torch.Generator().manual_seed(dh + L)
    qkv = torch.randn(B, L, 3 * H * dh, generator=g).to(bf16).to(DEV)
    cos = torch.rand(pos + L + 5, dh // 2, generator=g).to(DEV)
    sin = torch.rand(pos + L + 5, dh // 2, generator=g).to(DEV)
    q0, k0, v0 = (torch.empty(B, H, L, dh, device=DEV, dtype=bf16) for _ in range(3))
    ops.qk_norm_rope_fwd(qkv, cos, sin, pos, H, 1e-6, q0, k0, v0, want_rstd=False)
    kc = torch.full((B, H, cap, dh), 7.0, device=DEV, dtype=bf16)
    vc = torch.full((B, H, cap, dh), 7.0, device=DEV, dtype=bf16)
    q1 = torch.empty(B, H, L, dh, device=DEV, dtype=bf16)
    state = torch.tensor([n_valid, start, pos, 0], dtype=torch.int32, device=DEV)
    ops.qk_norm_rope_decode(qkv, cos, sin, state, q1, kc, vc, H, 1e-6)
    slots = ((start + n_valid + torch.arange(L)) % cap).to(DEV)
    assert torch.equal(q1, q0)
    assert torch.equal(kc[:, :, slots], k0) and torch.equal(vc[:, :, slots], v0)
    other = torch.ones(cap, dtype=torch.bool, device=DEV)
    other[slots] = False
    assert bool((kc[:, :, other] == 7.0).all()) and bool((vc[:, :, other] == 7.0).all())


@pytest.mark.parametrize("case", [(1, 24, 65, 64, 1105, 975, 0, 0), (1, 24, 65, 64, 1105, 1040, 1100, 260),
                                  (2, 3, 1, 64, 300, 120, 250, 16), (1, 2, 65, 128, 700, 130, 0, 0),
                                  (1, 2, 128, 64, 256, 0, 0, 0), (3, 20, 7, 64, 100, 93, 50, 0)],
                         ids=lambda c: "-".join(map(str, c)))
def test_attn_decode_fused_equals_two_kernel_path(case):
    """owl_attn_decode_fused (QK-norm + RoPE + ring append in the attention's load path) against owl_qk_norm_rope_decode
    followed by owl_attn_decode: the ring contents must be bit for bit the same, the outputs equal to summation order."""
    ops = _ops()
    B, H, Lq, dh, cap, n_valid, start, vis = case
    g = torch.Generator().manual_seed(sum(case) + 1)
    qkv = torch.randn(B, Lq, 3 * H * dh, generator=g).to(bf16).to(DEV)
    pos = 11
    cos = torch.rand(pos + Lq + 3, dh // 2, generator=g).to(DEV)
    sin = torch.rand(pos + Lq + 3, dh // 2, generator=g).to(DEV)
    k0 = torch.randn(B, H, cap, dh, generator=g).to(bf16).to(DEV)
    v0 = torch.randn(B, H, cap, dh, generator=g).to(bf16).to(DEV)
    state = torch.tensor([n_valid, start, pos, 0], dtype=torch.int32, device=DEV)
    k1, v1 = k0.clone(), v0.clone()
    q = torch.empty(B, H, Lq, dh, device=DEV, dtype=bf16)
    ops.qk_norm_rope_decode(qkv, cos, sin, state, q, k1, v1, H, 1e-6)
    o_two = ops.attn_decode(q, k1, v1, state, vis)
    k2, v2 = k0.clone(), v0.clone()
    o_one = ops.attn_decode_fused(qkv, cos, sin, k2, v2, state, H, 1e-6, vis)
    torch.cuda.synchronize()
    assert torch.equal(k2, k1) and torch.equal(v2, v1)
    assert relerr(o_one, o_two) <= 1e-2


def _build(cfg, state, rope, dtype=torch.bfloat16):
    N = _nn()
    model = N.DiT(cfg)
    model.load_state_dict({k: v.float() for k, v in state.items()}, strict=True)
    model = model.to(DEV).to(dtype)
    for blk, (cos, sin) in zip(model.blocks, rope):
        blk.attn.rope.cos = cos.to(DEV).float().contiguous()
        blk.attn.rope.sin = sin.to(DEV).float().contiguous()
    return model.eval()


def test_ring_rollout_matches_reference_golden():
    """The reference's own recorded sequence (tests/golden/kv_rollout.pt: prefill -> decode -> decode+cache update ->
    truncate -> decode): the masked prefill runs on the training-shape path, everything after it on the ring."""
    from owl_audio_exps_b200.nn.decode import RingKVCache
    N = _nn()
    fx = load_golden("kv_rollout")
    cfg = cfg_of(fx)
    model = _build(cfg, fx["state"], fx["rope"])
    xs, conds, ref = fx["xs"].to(DEV).to(bf16), fx["conds"].to(DEV).to(bf16), fx["out"]
    T = cfg.tokens_per_frame
    with torch.no_grad():
        cache = N.KVCache(cfg)
        cache.reset(fx["B"])
        cache.enable_cache_updates()
        model(xs[:, :3 * T], conds[:, :3], None, cache)
        cache.disable_cache_updates()
        ring = RingKVCache.from_cache(cache, window_rows=4 * T, new_rows=T)
        model.enable_decoding()
        assert relerr(model(xs[:, 3 * T:4 * T], conds[:, 3:4], None, ring), ref["decode3"]) <= 2e-2
        assert relerr(model(xs[:, 3 * T:4 * T], conds[:, 3:4], None, ring), ref["decode3_cache"]) <= 2e-2
        ring.commit(T)                                   # the reference's cache update of that call
        assert ring.length_at(0) == ref["len_after"][0] and ring.get_offset(0) == ref["offsets_after"][0]
        ring.truncate(1, front=False)
        assert relerr(model(xs[:, 4 * T:5 * T], conds[:, 4:5], None, ring), ref["decode4"]) <= 2e-2
        assert ring.state[:3].tolist() == [ring.n_valid, ring.start, ring.pos]


def test_ring_rollout_many_frames_vs_oracle_and_graph_replay():
    """A windowed rollout long enough for the ring to wrap several times (window 3 frames, 9 generated frames), local and
    global layers, against the oracle's cached decode (fp32 CPU) -- and ONE captured graph of the model call replayed for
    every frame must reproduce the eager ring path bit for bit."""
    from owl_audio_exps_b200.nn.decode import RingKVCache
    N = _nn()
    cfg = types.SimpleNamespace(d_model=128, n_heads=2, n_layers=5, tokens_per_frame=5, n_frames=16, sample_size=2,
                                causal=True, has_audio=True, rope_impl="ortho", local_window=2, global_window=None,
                                backbone="dit")
    torch.manual_seed(11)
    model = N.DiT(cfg)
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(p.to(bf16).float())
    W = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    rope = [(b.attn.rope.cos.clone().cpu(), b.attn.rope.sin.clone().cpu()) for b in model.blocks]
    model = model.to(DEV).to(bf16).eval()
    T, d, B, n_ctx, n_gen, win = cfg.tokens_per_frame, cfg.d_model, 2, 3, 9, 3
    g = torch.Generator().manual_seed(12)
    xs = torch.randn(B, (n_ctx + n_gen) * T, d, generator=g).to(bf16)
    conds = torch.randn(B, n_ctx + n_gen, d, generator=g).to(bf16)
    # oracle: prefill, then per frame decode + cache update, truncating to the window
    oc = O.OracleKVCache(cfg.n_layers, T)
    with torch.no_grad():
        oc.should_update = True
        O.dit_forward(xs[:, :n_ctx * T].float(), conds[:, :n_ctx].float(), W, cfg, rope, kv_cache=oc)
        ref = []
        for f in range(n_ctx, n_ctx + n_gen):
            ref.append(O.dit_forward(xs[:, f * T:(f + 1) * T].float(), conds[:, f:f + 1].float(), W, cfg, rope, kv_cache=oc,
                                     decoding=True))
            if oc.length_at(0) // T > win:
                oc.truncate(1)
    from owl_audio_exps_b200 import functional as OF
    xg, cg = xs.to(DEV), conds.to(DEV)
    with torch.no_grad(), OF.frozen_weights():
        cache = N.KVCache(cfg)
        cache.reset(B)
        cache.enable_cache_updates()
        model(xg[:, :n_ctx * T], cg[:, :n_ctx], None, cache)
        cache.disable_cache_updates()
        model.enable_decoding()
        ring = RingKVCache.from_cache(cache, window_rows=(win + 1) * T, new_rows=T)
        ring2 = RingKVCache.from_cache(cache, window_rows=(win + 1) * T, new_rows=T)
        # the graph: one model call on static inputs against ring2
        x_in, c_in = xg[:, :T].clone(), cg[:, :1].clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            model(x_in, c_in, None, ring2)               # warm-up: workspaces exist before the capture
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            y_graph = model(x_in, c_in, None, ring2)
        for i, f in enumerate(range(n_ctx, n_ctx + n_gen)):
            y = model(xg[:, f * T:(f + 1) * T], cg[:, f:f + 1], None, ring)
            assert relerr(y, ref[i]) <= 2e-2, f
            x_in.copy_(xg[:, f * T:(f + 1) * T])
            c_in.copy_(cg[:, f:f + 1])
            graph.replay()
            assert torch.equal(y_graph, y), f
            for r in (ring, ring2):
                r.commit(T)
                if r.n_frames() > win:
                    r.truncate(1)
        assert ring.start != 0 and ring.pos == (n_ctx + n_gen) * T    # the ring did wrap / advance


def test_ring_cache_guards():
    from owl_audio_exps_b200.nn.decode import RingKVCache
    cfg = types.SimpleNamespace(d_model=128, n_heads=2, n_layers=1, tokens_per_frame=4, n_frames=8)
    ring = RingKVCache(cfg, 1, window_rows=8, new_rows=4, device=DEV)
    ring.commit(4)
    ring.commit(4)
    ring.commit(4)
    with pytest.raises(RuntimeError):
        ring.commit(4)
    with pytest.raises(RuntimeError):
        ring.check_room(4, 100)
    ring.truncate(1)
    ring.check_room(4, 100)
    with pytest.raises(RuntimeError):
        ring.check_room(4, 10)      # RoPE table too short for position 12..15
    assert ring.state[:3].tolist() == [8, 4, 12]
