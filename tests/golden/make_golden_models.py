"""Generate tests/golden/audiorft_core.pt and gamerft_core.pt by running the UNMODIFIED reference models
`owl_wms.models.audiorft.AudioRFTCore` (t_embed -> proj_in -> DiT -> FinalLayer) and `owl_wms.models.gamerft.GameRFTCore`
(+ ControlEmbedding, patchify) on CPU in fp32, forward + backward.  Build container only (needs
/root/reference):   python tests/golden/make_golden_models.py
Same recipe and the same SDPA-for-flex backward substitution as make_golden.py (whose helpers are reused)."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402  (sets TORCHDYNAMO_DISABLE before torch is imported)
import torch  # noqa: E402


def audiorft(A, flex_real):
    from owl_wms.models.audiorft import AudioRFTCore
    cfg = G.make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=1, n_frames=64, sample_size=24,
                     has_audio=True, rope_impl="audio1d", local_window=4, channels=64)
    torch.manual_seed(321)
    model = AudioRFTCore(cfg)
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
            p.copy_(G.bf16r(p))
    B, L = 2, 24
    g = torch.Generator().manual_seed(4321)
    x = G.bf16r(torch.randn(B, L, cfg.channels, generator=g)).requires_grad_(True)
    t = torch.randn(B, L, generator=g).sigmoid()
    dy = G.bf16r(torch.randn(B, L, cfg.channels, generator=g))
    A.flex_attention = flex_real
    with torch.no_grad():
        y_flex = model(x, t)
    A.flex_attention = G.dense_flex
    y = model(x, t)
    sub_diff = (y - y_flex).abs().max().item()
    (y * dy).sum().backward()
    with torch.no_grad():
        emb = model.t_embed.sincos(t)
        cond = model.t_embed(t)
    fx = dict(
        name="audiorft_core", cfg=dict(vars(cfg)), B=B, L=L, flex_vs_dense_fwd_maxabs=sub_diff,
        state=dict((k, v.detach().to(torch.bfloat16)) for k, v in model.state_dict().items()),
        rope=[(blk.attn.rope.cos.clone(), blk.attn.rope.sin.clone()) for blk in model.transformer.blocks],
        x=x.detach().to(torch.bfloat16), t=t, dy=dy.to(torch.bfloat16), sincos=emb, cond=cond,
        y=y.detach(), dx=x.grad.clone(), grads=dict((k, p.grad.clone()) for k, p in model.named_parameters()),
    )
    torch.save(fx, os.path.join(HERE, "audiorft_core.pt"))
    print(f"audiorft_core: y absmax {y.abs().max():.3f}  flex-vs-dense fwd diff {sub_diff:.2e}")


def gamerft(A, flex_real):
    from owl_wms.models.gamerft import GameRFTCore
    cfg = G.make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=4, n_frames=8, sample_size=2,
                     has_audio=False, rope_impl="motion", local_window=2, channels=16, n_buttons=11, uncond=False)
    torch.manual_seed(654)
    model = GameRFTCore(cfg)
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
            p.copy_(G.bf16r(p))
    B, n, c, h = 2, 6, cfg.channels, cfg.sample_size
    g = torch.Generator().manual_seed(8765)
    x = G.bf16r(torch.randn(B, n, c, h, h, generator=g)).requires_grad_(True)
    t = torch.randn(B, n, generator=g).sigmoid()
    mouse = 3.0 * torch.randn(B, n, 2, generator=g)
    btn = (torch.rand(B, n, cfg.n_buttons, generator=g) < 0.3).float()
    has_controls = torch.tensor([True, False])
    dy = G.bf16r(torch.randn(B, n, c, h, h, generator=g))
    A.flex_attention = flex_real
    with torch.no_grad():
        y_flex = model(x, t, mouse, btn, None, has_controls)
    A.flex_attention = G.dense_flex
    y = model(x, t, mouse, btn, None, has_controls)
    sub_diff = (y - y_flex).abs().max().item()
    (y * dy).sum().backward()
    with torch.no_grad():
        ctrl = model.control_embed(mouse, btn)
    fx = dict(
        name="gamerft_core", cfg=dict(vars(cfg)), B=B, flex_vs_dense_fwd_maxabs=sub_diff,
        state=dict((k, v.detach().to(torch.bfloat16)) for k, v in model.state_dict().items()),
        rope=[(blk.attn.rope.cos.clone(), blk.attn.rope.sin.clone()) for blk in model.transformer.blocks],
        x=x.detach().to(torch.bfloat16), t=t, mouse=mouse, btn=btn, has_controls=has_controls, dy=dy.to(torch.bfloat16),
        ctrl=ctrl, y=y.detach(), dx=x.grad.clone(),
        # fixture size: gradients of the wide embedding matrices are kept in bf16 (checked to 4e-3), the rest in fp32
        grads=dict((k, p.grad.to(torch.bfloat16) if p.numel() > 100_000 else p.grad.clone())
                   for k, p in model.named_parameters()),
    )
    torch.save(fx, os.path.join(HERE, "gamerft_core.pt"))
    print(f"gamerft_core: y absmax {y.abs().max():.3f}  flex-vs-dense fwd diff {sub_diff:.2e}")


def gamerft_audio(A, flex_real):
    """GameRFTAudioCore (video patches + one audio token per frame, T = h*w + 1).  The reference wraps its backbone in
    torch.compile(dynamic=True); with TORCHDYNAMO_DISABLE=1 that wrapper runs the module eagerly, and its state-dict keys
    keep the `transformer._orig_mod.` spelling."""
    import types
    mm = types.ModuleType("owl_wms.nn.mmattn")   # the reference's mmattn.py:12 imports a name attn.py no longer has, so
    mm.MMDIT = object                            # gamerft_audio.py cannot be imported as shipped; only its `dit` branch
    sys.modules["owl_wms.nn.mmattn"] = mm        # is exercised here and that branch never touches MMDIT
    from owl_wms.models.gamerft_audio import GameRFTAudioCore
    cfg = G.make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=5, n_frames=6, sample_size=2,
                     has_audio=True, rope_impl="ortho", local_window=3, channels=16, audio_channels=8, n_buttons=11,
                     uncond=True)
    torch.manual_seed(987)
    model = GameRFTAudioCore(cfg)
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
            p.copy_(G.bf16r(p))
    B, n, c, h = 2, 6, cfg.channels, cfg.sample_size
    g = torch.Generator().manual_seed(1357)
    x = G.bf16r(torch.randn(B, n, c, h, h, generator=g)).requires_grad_(True)
    audio = G.bf16r(torch.randn(B, n, cfg.audio_channels, generator=g)).requires_grad_(True)
    t = torch.randn(B, n, generator=g).sigmoid()
    dyv = G.bf16r(torch.randn(B, n, c, h, h, generator=g))
    dya = G.bf16r(torch.randn(B, n, cfg.audio_channels, generator=g))
    A.flex_attention = flex_real
    with torch.no_grad():
        yv_flex, _ = model(x, audio, t, None, None)
    A.flex_attention = G.dense_flex
    yv, ya = model(x, audio, t, None, None)
    sub_diff = (yv - yv_flex).abs().max().item()
    ((yv * dyv).sum() + (ya * dya).sum()).backward()
    blocks = model.transformer._orig_mod.blocks if hasattr(model.transformer, "_orig_mod") else model.transformer.blocks
    fx = dict(
        name="gamerft_audio_core", cfg=dict(vars(cfg)), B=B, flex_vs_dense_fwd_maxabs=sub_diff,
        state=dict((k, v.detach().to(torch.bfloat16)) for k, v in model.state_dict().items()),
        rope=[(blk.attn.rope.cos.clone(), blk.attn.rope.sin.clone()) for blk in blocks],
        x=x.detach().to(torch.bfloat16), audio=audio.detach().to(torch.bfloat16), t=t, dyv=dyv.to(torch.bfloat16),
        dya=dya.to(torch.bfloat16), yv=yv.detach(), ya=ya.detach(), dx=x.grad.clone(), daudio=audio.grad.clone(),
        grads=dict((k, p.grad.to(torch.bfloat16) if p.numel() > 100_000 else p.grad.clone())
                   for k, p in model.named_parameters()),
    )
    torch.save(fx, os.path.join(HERE, "gamerft_audio_core.pt"))
    print(f"gamerft_audio_core: yv absmax {yv.abs().max():.3f} ya absmax {ya.abs().max():.3f}  flex-vs-dense {sub_diff:.2e}")


def audio_rollout(A):
    """The reference's own AudioCachingSampler (sampling/audio_caching.py) on the reference AudioRFTCore carrying the
    audiorft_core.pt weights: real CPU flex forward, every torch.randn / randn_like draw served from a recorded queue,
    a custom schedule (the default one needs diffusers), the KV cache moved to CPU fp32."""
    from owl_wms.models.audiorft import AudioRFTCore
    import owl_wms.nn.kv_cache as K
    import owl_wms.sampling.audio_caching as S
    core = torch.load(os.path.join(HERE, "audiorft_core.pt"), weights_only=False)
    cfg = G.make_cfg(**core["cfg"])
    model = AudioRFTCore(cfg).eval()
    model.load_state_dict({k: v.float() for k, v in core["state"].items()})
    S.KVCache = lambda config: K.KVCache(config).to("cpu", torch.float32)
    g = torch.Generator().manual_seed(2468)
    B, init_len, num_tokens, schedule = 2, 6, 4, [1.0, 0.75, 0.4]
    x = G.bf16r(torch.randn(B, init_len, cfg.channels, generator=g))
    draws = []

    def draw(shape):
        z = torch.randn(*shape, generator=g)
        draws.append(z)
        return z
    real_randn, real_like = torch.randn, torch.randn_like
    torch.randn_like = lambda t, **kw: draw(tuple(t.shape))
    torch.randn = lambda *shape, generator=None, **kw: real_randn(*shape, generator=generator) if generator is not None \
        else draw(tuple(shape[0]) if isinstance(shape[0], (tuple, list)) else shape)
    try:
        sampler = S.AudioCachingSampler(n_steps=len(schedule), num_tokens=num_tokens, noise_prev=0.2,
                                        custom_schedule=list(schedule), max_window=4)
        out = sampler(model, x)
    finally:
        torch.randn, torch.randn_like = real_randn, real_like
    fx = dict(name="audio_rollout", weights_from="audiorft_core", x=x, schedule=schedule, num_tokens=num_tokens,
              noise_prev=0.2, max_window=4, draws=draws, out=out)
    torch.save(fx, os.path.join(HERE, "audio_rollout.pt"))
    print(f"audio_rollout: {len(draws)} noise draws, out {tuple(out.shape)} absmax {out.abs().max():.3f}")


def av_rollout(A):
    """The reference's own AVCachingSamplerV2 (sampling/av_caching_v2.py) on the reference GameRFTCore carrying the
    gamerft_core.pt weights, cfg_scale 1.3 (two evaluations per step), same recording harness as audio_rollout."""
    from owl_wms.models.gamerft import GameRFTCore
    import owl_wms.nn.kv_cache as K
    import owl_wms.sampling.av_caching_v2 as S
    core = torch.load(os.path.join(HERE, "gamerft_core.pt"), weights_only=False)
    cfg = G.make_cfg(**core["cfg"])
    model = GameRFTCore(cfg).eval()
    model.load_state_dict({k: v.float() for k, v in core["state"].items()})
    S.KVCache = lambda config: K.KVCache(config).to("cpu", torch.float32)
    g = torch.Generator().manual_seed(97531)
    B, init_len, n_total, schedule = 2, 3, 6, [1.0, 0.7, 0.3]
    x = G.bf16r(torch.randn(B, init_len, cfg.channels, cfg.sample_size, cfg.sample_size, generator=g))
    mouse = 2.0 * torch.randn(B, n_total, 2, generator=g)
    btn = (torch.rand(B, n_total, cfg.n_buttons, generator=g) < 0.3).float()
    draws = []

    def draw(shape):
        z = torch.randn(*shape, generator=g)
        draws.append(z)
        return z
    real_randn, real_like = torch.randn, torch.randn_like
    torch.randn_like = lambda t, **kw: draw(tuple(t.shape))
    try:
        sampler = S.AVCachingSamplerV2(n_steps=len(schedule), cfg_scale=1.3, num_frames=8, noise_prev=0.2, max_window=4,
                                       custom_schedule=list(schedule))
        out = sampler(model, x, mouse, btn)
    finally:
        torch.randn, torch.randn_like = real_randn, real_like
    fx = dict(name="av_rollout", weights_from="gamerft_core", x=x, mouse=mouse, btn=btn, schedule=schedule, cfg_scale=1.3,
              num_frames=8, noise_prev=0.2, max_window=4, draws=draws, out=out)
    torch.save(fx, os.path.join(HERE, "av_rollout.pt"))
    print(f"av_rollout: {len(draws)} noise draws, out {tuple(out.shape)} absmax {out.abs().max():.3f}")


def main():
    G.install_stubs()
    sys.modules["rotary_embedding_torch"].apply_rotary_emb = None   # imported by embeddings.py:9-12, never called on this path
    import owl_wms.nn.attn as A
    flex_real = A.flex_attention
    if "--only-av" in sys.argv:
        return gamerft_audio(A, flex_real)
    if "--only-av-rollout" in sys.argv:
        return av_rollout(A)
    if "--only-rollout" in sys.argv:
        return audio_rollout(A)
    if "--only-game" not in sys.argv:
        audiorft(A, flex_real)
    gamerft(A, flex_real)
    gamerft_audio(A, flex_real)
    A.flex_attention = flex_real
    audio_rollout(A)
    av_rollout(A)


if __name__ == "__main__":
    main()
