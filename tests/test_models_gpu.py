"""Parity of the callers either side of the DiT stack (SinCosEmbed / TimestepEmbedding,
proj_in, FinalLayer, assembled as AudioRFTCore) against the reference model's golden fixture and the CPU oracle.
Bar: tensor-normalised max relative error <= 2e-2 (bf16), the same as the block stack."""
import pytest
import torch

from helpers import audio_cfg, cfg_of, load_golden, relerr
from oracle import dit_oracle as O

pytestmark = pytest.mark.gpu
TOL = 2e-2
DEV = "cuda:0"


def _core(cfg, state, rope, dtype=torch.float32):
    from owl_audio_exps_b200.models import AudioRFTCore
    model = AudioRFTCore(cfg)
    model.load_state_dict({k: v.float() for k, v in state.items()}, strict=True)   # the reference's own key names
    model = model.to(DEV).to(dtype)
    for blk, (cos, sin) in zip(model.transformer.blocks, rope):
        blk.attn.rope.cos = cos.to(DEV).float().contiguous()
        blk.attn.rope.sin = sin.to(DEV).float().contiguous()
    return model


def test_sincos_embed_matches_reference_and_oracle():
    import owl_audio_exps_b200.nn as N
    fx = load_golden("audiorft_core")
    emb = N.SinCosEmbed(512, theta=300, mult=1000)
    got = emb(fx["t"].to(DEV))
    assert got.shape == (*fx["t"].shape, 512) and got.dtype == torch.bfloat16
    # bf16 output of an fp32 argument: one bf16 ulp of a value in [-1, 1] is 2^-8
    assert (got.float().cpu() - fx["sincos"]).abs().max() <= 2 ** -8
    # scalars, [b], large arguments (t*mult up to 1e3 rad: needs real range reduction, not the fast-math intrinsic)
    t = torch.tensor([0.0, 1e-4, 0.5, 0.999, 1.0], device=DEV)
    assert (emb(t).float().cpu() - O.sincos_embed(t.cpu(), 512)).abs().max() <= 2 ** -8
    assert emb(0.25, device=DEV).shape == (1, 512)
    step = N.StepEmbedding(128).to(DEV)
    assert step(torch.tensor([1.0, 4.0, 128.0], device=DEV)).shape == (3, 128)


@pytest.mark.parametrize("regime", ["fp32_params", "bf16_params"])
def test_audiorft_core_golden(regime):
    fx = load_golden("audiorft_core")
    cfg = cfg_of(fx)
    model = _core(cfg, fx["state"], fx["rope"], torch.float32 if regime == "fp32_params" else torch.bfloat16)
    x = fx["x"].to(DEV).float().requires_grad_(True)
    t = fx["t"].to(DEV)
    assert relerr(model.t_embed(t), fx["cond"]) <= TOL
    y = model(x, t)
    assert y.shape == fx["y"].shape
    assert relerr(y, fx["y"]) <= TOL
    y.backward(fx["dy"].to(DEV).to(y.dtype))
    assert relerr(x.grad, fx["dx"]) <= TOL
    worst = max((relerr(p.grad, fx["grads"][k]), k) for k, p in model.named_parameters())
    assert worst[0] <= TOL, worst


def test_audiorft_core_audio_yml_shape_vs_oracle():
    """configs/audio.yml width (d=1024, 16 heads, 64 channels, L=120) on 2 layers, against the fp32 CPU oracle."""
    cfg = audio_cfg(n_layers=2)
    torch.manual_seed(5)
    from owl_audio_exps_b200.models import AudioRFT, AudioRFTCore
    model = AudioRFTCore(cfg).to(DEV)
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
            p.copy_(p.to(torch.bfloat16).float())
    B, L = 2, 120
    x = torch.randn(B, L, cfg.channels, device=DEV).to(torch.bfloat16).float().requires_grad_(True)
    t = torch.rand(B, L, device=DEV)
    dy = torch.randn(B, L, cfg.channels, device=DEV).to(torch.bfloat16)
    y = model(x, t)
    y.backward(dy)
    W = {k: v.detach().float().cpu().requires_grad_(True) for k, v in model.state_dict().items()}
    rope = [(b.attn.rope.cos.cpu(), b.attn.rope.sin.cpu()) for b in model.transformer.blocks]
    xo = x.detach().cpu().requires_grad_(True)
    yo = O.audio_rft_core(xo, t.cpu(), W, cfg, rope)
    (yo * dy.float().cpu()).sum().backward()
    assert relerr(y, yo) <= TOL
    assert relerr(x.grad, xo.grad) <= TOL
    worst = max((relerr(p.grad, W[k].grad), k) for k, p in model.named_parameters())
    assert worst[0] <= TOL, worst
    # the training wrapper (audiorft.py:60-97): scalar loss, gradients reach every parameter
    rft = AudioRFT(cfg).to(DEV)
    loss = rft(torch.randn(B, L, cfg.channels, device=DEV))
    loss.backward()
    assert loss.dim() == 0 and torch.isfinite(loss)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in rft.parameters())


def _game(cfg, state, rope, dtype=torch.float32):
    from owl_audio_exps_b200.models import GameRFTCore
    model = GameRFTCore(cfg)
    model.load_state_dict({k: v.float() for k, v in state.items()}, strict=True)
    model = model.to(DEV).to(dtype)
    for blk, (cos, sin) in zip(model.transformer.blocks, rope):
        blk.attn.rope.cos = cos.to(DEV).float().contiguous()
        blk.attn.rope.sin = sin.to(DEV).float().contiguous()
    return model


@pytest.mark.parametrize("regime", ["fp32_params", "bf16_params"])
def test_gamerft_core_golden(regime):
    """GameRFTCore (gamerft.py:13-60): ControlEmbedding with its narrow inputs (2 angle features, 11 buttons: K padded to 8s),
    has_controls masking, patchify / unpatchify, against the reference model's fixture."""
    fx = load_golden("gamerft_core")
    cfg = cfg_of(fx)
    model = _game(cfg, fx["state"], fx["rope"], torch.float32 if regime == "fp32_params" else torch.bfloat16)
    mouse, btn, t = fx["mouse"].to(DEV), fx["btn"].to(DEV), fx["t"].to(DEV)
    assert relerr(model.control_embed(mouse, btn), fx["ctrl"]) <= TOL
    x = fx["x"].to(DEV).float().requires_grad_(True)
    y = model(x, t, mouse, btn, None, fx["has_controls"].to(DEV))
    assert y.shape == fx["y"].shape
    assert relerr(y, fx["y"]) <= TOL
    y.backward(fx["dy"].to(DEV).to(y.dtype))
    assert relerr(x.grad, fx["dx"]) <= TOL
    worst = max((relerr(p.grad, fx["grads"][k]), k) for k, p in model.named_parameters())
    assert worst[0] <= TOL, worst
    assert all(p.grad.shape == p.shape for p in model.parameters())


def test_gamerft_training_wrapper():
    """gamerft.py:63-125: CFG control dropout + noise + MSE; gradients reach every parameter."""
    from owl_audio_exps_b200.models import GameRFT
    fx = load_golden("gamerft_core")
    cfg = cfg_of(fx)
    cfg.cfg_prob = 0.5
    torch.manual_seed(3)
    rft = GameRFT(cfg).to(DEV)
    B, n = 4, 6
    x = torch.randn(B, n, cfg.channels, cfg.sample_size, cfg.sample_size, device=DEV)
    out = rft(x, torch.randn(B, n, 2, device=DEV), (torch.rand(B, n, cfg.n_buttons, device=DEV) < 0.3).float(),
              return_dict=True)
    out["diffusion_loss"].backward()
    assert torch.isfinite(out["diffusion_loss"]) and out["pred_video"].shape == x.shape
    assert out["cfg_mask"].dtype == torch.bool and out["cfg_mask"].shape == (B,)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in rft.parameters())


def test_audio_caching_sampler_matches_reference_rollout():
    """AudioCachingSampler on the sm_100a decode path vs the reference sampler's recorded rollout (same noise draws):
    prefill with cache updates, decode-mode Euler steps, cache append, window truncation.  4 autoregressive tokens x 3
    steps compound the per-call bf16 error, hence 3e-2 instead of the single-call 2e-2."""
    from owl_audio_exps_b200.sampling import AudioCachingSampler
    fx = load_golden("audio_rollout")
    core = load_golden(fx["weights_from"])
    model = _core(cfg_of(core), core["state"], core["rope"]).eval()
    draws = iter(fx["draws"])

    def noise_fn(shape, like):
        z = next(draws)
        assert tuple(z.shape) == tuple(shape)
        return z.to(device=like.device, dtype=like.dtype)
    for ring in (True, False):      # ONE ring-cache graph per rollout / the per-latent graph on the growing cache
        draws = iter(fx["draws"])
        sampler = AudioCachingSampler(n_steps=len(fx["schedule"]), num_tokens=fx["num_tokens"], noise_prev=fx["noise_prev"],
                                      custom_schedule=list(fx["schedule"]), max_window=fx["max_window"], noise_fn=noise_fn,
                                      use_ring=ring, graph_min_steps=2)
        out = sampler(model, fx["x"].to(DEV))
        assert next(draws, None) is None
        assert out.shape == fx["out"].shape
        assert relerr(out, fx["out"]) <= 3e-2, ring
        assert not model.transformer.decoding
        assert (sampler.last_ring is not None) == ring
    # default schedule + default noise source run end to end
    out = AudioCachingSampler(n_steps=2, num_tokens=2)(model, fx["x"].to(DEV))
    assert out.shape == (2, 8, 64) and torch.isfinite(out).all()


def test_sampler_graph_replay_equals_eager_steps():
    """The per-latent CUDA graph of the decode step launches the same kernels as the eager loop: identical latents.  The
    ring path (ONE graph per rollout, weight-streaming GEMMs, split-KV attention) is a different kernel schedule: same
    latents to the rollout tolerance."""
    from owl_audio_exps_b200.sampling import AudioCachingSampler
    core = load_golden("audiorft_core")
    model = _core(cfg_of(core), core["state"], core["rope"]).eval()
    x = torch.randn(2, 5, 64, device=DEV)
    outs = []
    for graph, ring in ((False, False), (True, False), (True, True)):
        g = torch.Generator(device=DEV).manual_seed(11)
        noise_fn = lambda shape, like: torch.randn(*shape, device=like.device, dtype=like.dtype, generator=g)  # noqa: E731
        outs.append(AudioCachingSampler(n_steps=5, num_tokens=4, max_window=6, noise_fn=noise_fn, graph_decode=graph,
                                        graph_min_steps=2, use_ring=ring)(model, x))
    assert torch.equal(outs[0], outs[1])
    assert relerr(outs[2], outs[0]) <= 3e-2


def test_gamerft_audio_core_golden():
    """GameRFTAudioCore (gamerft_audio.py:19-98) vs the reference model's fixture; state_dict() spells the backbone keys
    `transformer._orig_mod.*` like a compiled reference checkpoint, and both spellings load."""
    from owl_audio_exps_b200.models import GameRFTAudio, GameRFTAudioCore
    fx = load_golden("gamerft_audio_core")
    cfg = cfg_of(fx)
    model = GameRFTAudioCore(cfg)
    model.load_state_dict({k: v.float() for k, v in fx["state"].items()}, strict=True)   # fixture: dynamo off, plain keys
    assert list(model.state_dict().keys()) == [k.replace("transformer.", "transformer._orig_mod.") for k in fx["state"]]
    model.load_state_dict(model.state_dict(), strict=True)                             # compiled-checkpoint spelling
    model = model.to(DEV)
    for blk, (cos, sin) in zip(model.transformer.blocks, fx["rope"]):
        blk.attn.rope.cos = cos.to(DEV).float().contiguous()
        blk.attn.rope.sin = sin.to(DEV).float().contiguous()
    x = fx["x"].to(DEV).float().requires_grad_(True)
    audio = fx["audio"].to(DEV).float().requires_grad_(True)
    yv, ya = model(x, audio, fx["t"].to(DEV), None, None)
    assert relerr(yv, fx["yv"]) <= TOL and relerr(ya, fx["ya"]) <= TOL
    torch.autograd.backward([yv, ya], [fx["dyv"].to(DEV).to(yv.dtype), fx["dya"].to(DEV).to(ya.dtype)])
    assert relerr(x.grad, fx["dx"]) <= TOL and relerr(audio.grad, fx["daudio"]) <= TOL
    named = dict(model.named_parameters())
    worst = max((relerr(named[k].grad, g), k) for k, g in fx["grads"].items())
    assert worst[0] <= TOL, worst
    # training wrapper: three losses, finite gradients everywhere
    cfg.cfg_prob = 0.0
    rft = GameRFTAudio(cfg).to(DEV)
    B, n = 2, 6
    loss, lv, la = rft(torch.randn(B, n, cfg.channels, 2, 2, device=DEV), torch.randn(B, n, cfg.audio_channels, device=DEV),
                       None, None)
    loss.backward()
    assert torch.isfinite(loss) and torch.allclose(loss, lv + la)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in rft.parameters())


def test_step_condition_posenc_embeddings_vs_oracle():
    """StepEmbedding (embeddings.py:86-104), ConditionEmbedding (:106-117), LearnedPosEnc (:14-28) against the oracle's
    sincos / mlp restatements on the same weights."""
    import math
    import owl_audio_exps_b200.nn as N
    torch.manual_seed(9)
    step = N.StepEmbedding(128, d_in=512, max_steps=128).to(DEV)
    steps = torch.tensor([1.0, 2.0, 16.0, 128.0], device=DEV)
    W = {k: v.detach().float().cpu() for k, v in step.state_dict().items()}
    t = math.log2(128) - torch.log2(steps.cpu())
    ref = O.mlp(O.sincos_embed(t, 512, theta=300, mult=1000 / math.log2(128)), W, "mlp")
    assert relerr(step(steps), ref) <= TOL
    assert step(4).shape == (1, 128)                                   # python scalar, as the reference accepts

    cond = N.ConditionEmbedding(10, 128).to(DEV)
    idx = torch.tensor([0, 3, 9], device=DEV)
    W = {k: v.detach().float().cpu() for k, v in cond.state_dict().items()}
    ref = O.mlp(W["embedding.weight"][idx.cpu()], W, "mlp")
    out = cond(idx)
    assert relerr(out, ref) <= TOL
    out.float().sum().backward()
    assert cond.embedding.weight.grad is not None and cond.mlp.fc1.weight.grad is not None

    pe = N.LearnedPosEnc(8, 16).to(DEV)
    x = torch.randn(2, 5, 16, device=DEV)
    assert torch.equal(pe(x), x + pe.p[-5:].unsqueeze(0))              # shorter than n_seq: the LAST n rows
    x = torch.randn(2, 8, 16, device=DEV)
    assert torch.equal(pe(x), x + pe.p.unsqueeze(0))


@pytest.mark.parametrize("graph", [False, True, "ring"])
def test_av_caching_sampler_matches_reference_rollout(graph):
    """AVCachingSamplerV2 on the decode path (T=4 tokens per frame, CFG on the controls, window truncation) vs the reference
    sampler's recorded rollout; eager and with the per-frame CUDA graph (two model evaluations inside it).  3 frames x 3
    guided steps compound the per-call bf16 error (guidance extrapolates by 1.3): 4e-2."""
    from owl_audio_exps_b200.sampling import AVCachingSamplerV2, get_sampler_cls
    assert get_sampler_cls("av_caching") is AVCachingSamplerV2
    fx = load_golden("av_rollout")
    core = load_golden(fx["weights_from"])
    model = _game(cfg_of(core), core["state"], core["rope"]).eval()
    draws = iter(fx["draws"])

    def noise_fn(shape, like):
        z = next(draws)
        assert tuple(z.shape) == tuple(shape)
        return z.to(device=like.device, dtype=like.dtype)
    sampler = AVCachingSamplerV2(n_steps=len(fx["schedule"]), cfg_scale=fx["cfg_scale"], num_frames=fx["num_frames"],
                                 noise_prev=fx["noise_prev"], max_window=fx["max_window"],
                                 custom_schedule=list(fx["schedule"]), noise_fn=noise_fn, graph_decode=bool(graph),
                                 graph_min_steps=2, use_ring=graph == "ring")
    out = sampler(model, fx["x"].to(DEV), fx["mouse"].to(DEV), fx["btn"].to(DEV))
    assert next(draws, None) is None
    assert out.shape == fx["out"].shape
    assert relerr(out, fx["out"]) <= 4e-2
    assert not model.transformer.decoding
