"""Pin the oracle (oracle/dit_oracle.py) to the reference: fixtures in tests/golden/ were produced by the
reference's own owl_wms/nn modules (tests/golden/make_golden.py).  CPU only."""
import pytest
import torch

from helpers import GOLDEN_CASES, cfg_of, load_golden, relerr
from oracle import dit_oracle as O

TOL = 2e-5  # fp32 vs fp32, different op order


def _weights(fx):
    return {k: v.float() for k, v in fx["state"].items()}


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_dit_forward_backward_matches_reference(name):
    fx = load_golden(name)
    cfg = cfg_of(fx)
    assert fx["flex_vs_dense_fwd_maxabs"] < 5e-6  # the SDPA substitution used for the reference's CPU backward
    W = {k: v.requires_grad_(True) for k, v in _weights(fx).items()}
    x = fx["x"].float().requires_grad_(True)
    cond = fx["cond"].float().requires_grad_(True)
    y = O.dit_forward(x, cond, W, cfg, fx["rope"], doc_id=fx["doc_id"])
    assert relerr(y, fx["y"]) < TOL
    (y * fx["dy"].float()).sum().backward()
    assert relerr(x.grad, fx["dx"]) < TOL
    assert relerr(cond.grad, fx["dcond"]) < TOL
    for k, g in fx["grads"].items():
        assert relerr(W[k].grad, g) < 5 * TOL, k


def test_kv_rollout_matches_reference():
    fx = load_golden("kv_rollout")
    cfg = cfg_of(fx)
    W = _weights(fx)
    xs, conds, ref = fx["xs"].float(), fx["conds"].float(), fx["out"]
    T = cfg.tokens_per_frame
    with torch.no_grad():
        cache = O.OracleKVCache(cfg.n_layers, T)
        cache.should_update = True
        y = O.dit_forward(xs[:, :3 * T], conds[:, :3], W, cfg, fx["rope"], kv_cache=cache)
        assert relerr(y, ref["prefill"]) < TOL
        cache.should_update = False
        y = O.dit_forward(xs[:, 3 * T:4 * T], conds[:, 3:4], W, cfg, fx["rope"], kv_cache=cache)
        assert relerr(y, ref["masked_cached"]) < TOL
        y = O.dit_forward(xs[:, 3 * T:4 * T], conds[:, 3:4], W, cfg, fx["rope"], kv_cache=cache, decoding=True)
        assert relerr(y, ref["decode3"]) < TOL
        assert relerr(y, ref["full4"]) < TOL  # cache equivalence (sanity/sampling.py:86-133)
        cache.should_update = True
        y = O.dit_forward(xs[:, 3 * T:4 * T], conds[:, 3:4], W, cfg, fx["rope"], kv_cache=cache, decoding=True)
        assert relerr(y, ref["decode3_cache"]) < TOL
        cache.should_update = False
        assert [cache.length_at(i) for i in range(cfg.n_layers)] == ref["len_after"]
        assert cache.offsets == ref["offsets_after"]
        cache.truncate(1, front=False)
        y = O.dit_forward(xs[:, 4 * T:5 * T], conds[:, 4:5], W, cfg, fx["rope"], kv_cache=cache, decoding=True)
        assert relerr(y, ref["decode4"]) < TOL
        assert relerr(cache.cache[1][0], ref["k_layer1"]) < TOL
        assert relerr(cache.cache[1][1], ref["v_layer1"]) < TOL


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_rope_tables_match_reference_code_over_restated_package(name):
    """nn/rope.py's table builders against the tables the REFERENCE CODE (rope.py:57-179) built on top of the generator's
    restatement of the absent third-party `rotary_embedding_torch` (tests/golden/make_golden.py install_stubs).  This pins
    the in-repo arithmetic (axis layout, audio-token slot, subsampling, MotionRoPE / Audio1DRoPE closed forms); it does
    NOT pin the third-party package itself: PARITY UNPINNED there (nn/rope.py header)."""
    from owl_audio_exps_b200.nn.rope import get_rope_cls
    fx = load_golden(name)
    cfg = cfg_of(fx)
    r = get_rope_cls(cfg.rope_impl)(cfg)
    cos, sin = fx["rope"][0]
    assert r.cos.shape == cos.shape and r.cos.dtype == torch.float32
    assert (r.cos - cos).abs().max().item() < 1e-6
    assert (r.sin - sin).abs().max().item() < 1e-6


def test_ortho_rope_axial_variants_and_real_package_hook():
    """The two release behaviours of get_axial_freqs that nn/rope.py can restate differ only on the leading (time,
    height) axes, and an importable `rotary_embedding_torch` is used verbatim for the table (init-time plumbing)."""
    import sys, types
    from owl_audio_exps_b200.nn import rope as R
    cfg = types.SimpleNamespace(d_model=128, n_heads=2, n_frames=6, sample_size=3, has_audio=True)
    a = R.OrthoRoPE(cfg)
    cfg2 = types.SimpleNamespace(**{**vars(cfg), "rope_axial_variant": "last_two_linspace"})
    b = R.OrthoRoPE(cfg2)
    assert a.cos.shape == b.cos.shape == (6 * 10, 32)
    assert (a.cos - b.cos).abs().max() > 1e-3           # the variants are distinguishable ...
    dh4 = (cfg.d_model // cfg.n_heads) // 4             # ... but agree on the last-two-axes features (width, pad)
    w0 = 2 * (2 * (dh4 // 2)) // 2                      # features [w0:] after the ::2 subsampling: width + pad axes
    assert torch.equal(a.cos[:, w0:], b.cos[:, w0:])

    class Fake:                                         # an "installed" package: whatever it returns is used
        def __init__(self, dim, freqs_for, max_freq):
            self.n = 2 * (dim // 2)
        def get_axial_freqs(self, *dims, offsets=None):
            return torch.full((*dims, 4 * self.n), 0.25)
    mod = types.ModuleType("rotary_embedding_torch")
    mod.RotaryEmbedding = Fake
    old = sys.modules.get("rotary_embedding_torch")
    sys.modules["rotary_embedding_torch"] = mod
    try:
        c = R.OrthoRoPE(cfg)
    finally:
        if old is None:
            del sys.modules["rotary_embedding_torch"]
        else:
            sys.modules["rotary_embedding_torch"] = old
    assert torch.allclose(c.cos, torch.full_like(c.cos, float(torch.cos(torch.tensor(0.25)))))


def test_frame_mask_matches_package_descriptor():
    """oracle.frame_mask == FrameMask.mask_mod (the predicate the kernels evaluate) on ragged cases."""
    from owl_audio_exps_b200.nn.attn import get_block_mask
    for T, nq, off, win, causal in [(1, 24, 0, 4, True), (5, 30, 0, None, True), (5, 10, 15, 3, True),
                                    (4, 32, 0, 2, False), (65, 130, 65, 16, True)]:
        doc = torch.tensor([[0, 0, 1, 1, 1, 2, 2, 2, 2, 2][: (nq + off) // T + 1]]) if T in (4, 5) else None
        if doc is not None and doc.shape[1] < (nq + off) // T:
            doc = None
        fm = get_block_mask(nq + off, T, win, doc, off, causal)
        dense = fm.mask_mod(torch.tensor(0), torch.tensor(0), torch.arange(nq)[:, None], torch.arange(nq + off)[None, :])
        ref = O.frame_mask(nq, nq + off, T, win, doc, off, causal)[0, 0]
        assert torch.equal(dense, ref), (T, nq, off, win, causal)


def test_audiorft_core_matches_reference():
    """Oracle restatement of SinCosEmbed / TimestepEmbedding / AudioRFTCore vs the reference model's own outputs
    (tests/golden/make_golden_audiorft.py): forward, input gradient and every parameter gradient."""
    fx = load_golden("audiorft_core")
    cfg = cfg_of(fx)
    assert fx["flex_vs_dense_fwd_maxabs"] < 5e-6
    t = fx["t"]
    assert relerr(O.sincos_embed(t, 512), fx["sincos"]) < TOL
    W = {k: v.requires_grad_(True) for k, v in _weights(fx).items()}
    assert relerr(O.timestep_embedding(t, W, "t_embed."), fx["cond"]) < TOL
    x = fx["x"].float().requires_grad_(True)
    y = O.audio_rft_core(x, t, W, cfg, fx["rope"])
    assert relerr(y, fx["y"]) < TOL
    (y * fx["dy"].float()).sum().backward()
    assert relerr(x.grad, fx["dx"]) < TOL
    for k, g in fx["grads"].items():
        assert relerr(W[k].grad, g) < 5 * TOL, k


def test_gamerft_core_matches_reference():
    """Mouse/Button/ControlEmbedding, patchify and GameRFTCore of the oracle vs the reference model's own outputs."""
    fx = load_golden("gamerft_core")
    cfg = cfg_of(fx)
    assert fx["flex_vs_dense_fwd_maxabs"] < 5e-6
    W = {k: v.requires_grad_(True) for k, v in _weights(fx).items()}
    assert relerr(O.control_embedding(fx["mouse"], fx["btn"], W, "control_embed."), fx["ctrl"]) < TOL
    x = fx["x"].float().requires_grad_(True)
    y = O.game_rft_core(x, fx["t"], fx["mouse"], fx["btn"], W, cfg, fx["rope"], has_controls=fx["has_controls"])
    assert relerr(y, fx["y"]) < TOL
    (y * fx["dy"].float()).sum().backward()
    assert relerr(x.grad, fx["dx"]) < TOL
    for k, g in fx["grads"].items():
        assert relerr(W[k].grad, g) < (5 * TOL if g.dtype == torch.float32 else 4e-3), k


def test_audio_cached_rollout_matches_reference_sampler():
    """Oracle rollout vs the reference's own AudioCachingSampler run (tests/golden/make_golden_models.py audio_rollout)."""
    fx = load_golden("audio_rollout")
    core = load_golden(fx["weights_from"])
    cfg = cfg_of(core)
    W = _weights(core)
    draws = iter(fx["draws"])

    def noise(shape):
        z = next(draws)
        assert tuple(z.shape) == tuple(shape)
        return z
    sched = list(fx["schedule"]) + [0.0]
    dt = [abs(b - a) for a, b in zip(sched[:-1], sched[1:])]
    with torch.no_grad():
        out = O.audio_cached_rollout(fx["x"], W, cfg, core["rope"], dt, fx["num_tokens"], fx["noise_prev"],
                                     fx["max_window"], noise)
    assert next(draws, None) is None
    assert relerr(out, fx["out"]) < 5 * TOL


def test_gamerft_audio_core_matches_reference():
    """Video + audio token interleave, layer_norm'd video head and the audio head of the oracle vs the reference model."""
    fx = load_golden("gamerft_audio_core")
    cfg = cfg_of(fx)
    assert fx["flex_vs_dense_fwd_maxabs"] < 5e-6
    W = {k: v.requires_grad_(True) for k, v in _weights(fx).items()}
    x = fx["x"].float().requires_grad_(True)
    audio = fx["audio"].float().requires_grad_(True)
    yv, ya = O.game_rft_audio_core(x, audio, fx["t"], W, cfg, fx["rope"])
    assert relerr(yv, fx["yv"]) < TOL and relerr(ya, fx["ya"]) < TOL
    ((yv * fx["dyv"].float()).sum() + (ya * fx["dya"].float()).sum()).backward()
    assert relerr(x.grad, fx["dx"]) < TOL and relerr(audio.grad, fx["daudio"]) < TOL
    for k, g in fx["grads"].items():
        assert relerr(W[k].grad, g) < (5 * TOL if g.dtype == torch.float32 else 4e-3), k


def test_av_cached_rollout_matches_reference_sampler():
    """Oracle rollout vs the reference's own AVCachingSamplerV2 run (cfg_scale 1.3: two model evaluations per step)."""
    fx = load_golden("av_rollout")
    core = load_golden(fx["weights_from"])
    cfg = cfg_of(core)
    W = _weights(core)
    draws = iter(fx["draws"])

    def noise(shape):
        z = next(draws)
        assert tuple(z.shape) == tuple(shape)
        return z
    sched = list(fx["schedule"]) + [0.0]
    dt = [abs(b - a) for a, b in zip(sched[:-1], sched[1:])]
    with torch.no_grad():
        out = O.av_cached_rollout(fx["x"], fx["mouse"], fx["btn"], W, cfg, core["rope"], dt, fx["num_frames"],
                                  fx["cfg_scale"], fx["noise_prev"], fx["max_window"], noise)
    assert next(draws, None) is None
    assert relerr(out, fx["out"]) < 5 * TOL


def test_oracle_newton_schulz_matches_reference_muon():
    """oracle.newton_schulz5 (bf16 iteration) == owl_wms/muon.py:11-38 run by tests/golden/make_golden_muon.py, bit for bit;
    the fp32 iteration of the same recurrence stays within bf16 iteration noise of it."""
    import os
    from helpers import GOLDEN
    fx = torch.load(os.path.join(GOLDEN, "muon_ns5.pt"), map_location="cpu", weights_only=False)
    for G, ref in zip(fx["G"], fx["out"]):
        assert torch.equal(O.newton_schulz5(G, fx["steps"], torch.bfloat16).float(), ref)
        assert relerr(O.newton_schulz5(G, fx["steps"]), ref) <= 6e-2
