"""Generate tests/golden/*.pt by running the UNMODIFIED reference `owl_wms/nn` modules on CPU (fp32).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
Nothing in the test-suite, smoke() or bench.py reads /root/reference -- they read the .pt files.

Recipe:
  * stub the import-time-only dependencies that are absent here (owl_vaes, diffusers, omegaconf) and
    provide `rotary_embedding_torch.RotaryEmbedding` (closed forms of the published package; used only
    at module init to build cos/sin tables -- the tables are stored in the fixture, so neither the
    oracle nor the CUDA path depends on this restatement);
  * TORCHDYNAMO_DISABLE=1 (attn.py:15-16 wraps flex_attention in torch.compile at import);
  * FlexAttention has no CPU backward in torch 2.11, so for fwd+bwd cases the one call
    `flex_attention(q,k,v,block_mask)` (attn.py:106-109) is served by SDPA with the dense mask rendered
    from the reference's OWN block_mask.mask_mod (the idiom of attn.py:286-292).  The substitution is
    checked here against real CPU flex forward and the difference is recorded in every fixture.
All weights / inputs are rounded to bf16-representable values first, so a bf16 implementation sees
exactly the same numbers.
"""
import math
import os
import sys
import types

os.environ["TORCHDYNAMO_DISABLE"] = "1"
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
from torch import nn  # noqa: E402

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def install_stubs():
    ov = types.ModuleType("owl_vaes")
    ov.from_pretrained = lambda *a, **k: None
    sys.modules["owl_vaes"] = ov
    df = types.ModuleType("diffusers")
    df.AutoencoderDC = object
    df.FlowMatchEulerDiscreteScheduler = object
    sys.modules["diffusers"] = df
    oc = types.ModuleType("omegaconf")

    class OmegaConf:
        @staticmethod
        def create(x):
            return x

        @staticmethod
        def structured(x):
            return x

        @staticmethod
        def load(x):
            raise RuntimeError("stub")
    oc.OmegaConf = OmegaConf
    sys.modules["omegaconf"] = oc

    ret = types.ModuleType("rotary_embedding_torch")

    class RotaryEmbedding(nn.Module):
        def __init__(self, dim, freqs_for="lang", theta=10000, max_freq=10, **kw):
            super().__init__()
            self.freqs_for = freqs_for
            if freqs_for == "lang":
                freqs = 1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))
            elif freqs_for == "pixel":
                freqs = torch.linspace(1.0, max_freq / 2, dim // 2) * math.pi
            else:
                raise ValueError(freqs_for)
            self.freqs = nn.Parameter(freqs, requires_grad=False)

        def forward(self, t, seq_len=None, offset=0):
            return torch.einsum("..., f -> ... f", t.type(self.freqs.dtype), self.freqs).repeat_interleave(2, dim=-1)

        def get_axial_freqs(self, *dims, offsets=None):
            out = []
            for i, n in enumerate(dims):
                pos = torch.linspace(-1, 1, steps=n) if self.freqs_for == "pixel" else torch.arange(n)
                pos = pos + (0 if offsets is None else offsets[i])
                idx = [None] * len(dims)
                idx[i] = slice(None)
                out.append(self.forward(pos)[(Ellipsis, *idx, slice(None))])
            return torch.cat(torch.broadcast_tensors(*out), dim=-1)
    ret.RotaryEmbedding = RotaryEmbedding
    sys.modules["rotary_embedding_torch"] = ret
    sys.path.insert(0, REF)


def dense_flex(q, k, v, block_mask=None, **kw):
    """SDPA over the dense mask rendered from the reference's own mask_mod (attn.py:286-292 idiom)."""
    if block_mask is None:
        return F.scaled_dot_product_attention(q, k, v)
    B, Lq, Lk = q.shape[0], q.shape[2], k.shape[2]
    qi = torch.arange(Lq)[:, None]
    ki = torch.arange(Lk)[None, :]
    masks = [block_mask.mask_mod(torch.tensor(b), torch.tensor(0), qi, ki) for b in range(B)]
    return F.scaled_dot_product_attention(q, k, v, attn_mask=torch.stack(masks)[:, None])


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float32)


def make_cfg(**kw):
    base = dict(backbone="dit", causal=True, cfg_prob=0.0, uncond=True, gradient_checkpointing=False,
                global_window=None)
    base.update(kw)
    return types.SimpleNamespace(**base)


CASES = {
    # name: (cfg kwargs, B, n_frames_in_batch, doc_id)
    "audio_T1": (dict(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=1, n_frames=64, sample_size=24,
                      has_audio=True, rope_impl="audio1d", local_window=4), 2, 24, None),
    "av_T5_ortho": (dict(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=5, n_frames=6, sample_size=2,
                         has_audio=True, rope_impl="ortho", local_window=3), 2, 6, None),
    "video_T4_motion_doc": (dict(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=4, n_frames=8, sample_size=2,
                                 has_audio=False, rope_impl="motion", local_window=2), 1, 8,
                            [[0, 0, 0, 1, 1, 1, 1, 2]]),
    "av_T5_bidir": (dict(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=5, n_frames=6, sample_size=2,
                         has_audio=True, rope_impl="ortho", local_window=2, causal=False), 1, 6, None),
}


def build(A, cfg, seed):
    torch.manual_seed(seed)
    model = A.DiT(cfg)
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() == 1:  # reference init leaves MLP biases at 0; make every bias exercise its path
                p.copy_(0.1 * torch.randn_like(p))
            p.copy_(bf16r(p))
    return model


def run_case(A, name, spec, flex_real):
    kw, B, nF, doc = spec
    cfg = make_cfg(**kw)
    model = build(A, cfg, seed=sum(map(ord, name)) % 1000)
    d, T = cfg.d_model, cfg.tokens_per_frame
    L = nF * T
    g = torch.Generator().manual_seed(1234)
    x = bf16r(torch.randn(B, L, d, generator=g)).requires_grad_(True)
    cond = bf16r(torch.randn(B, nF, d, generator=g)).requires_grad_(True)
    dy = bf16r(torch.randn(B, L, d, generator=g))
    doc_id = torch.tensor(doc, dtype=torch.long) if doc is not None else None

    # forward agreement of the SDPA substitution with the real CPU flex kernel
    A.flex_attention = flex_real
    with torch.no_grad():
        y_flex = model(x, cond, doc_id)
    A.flex_attention = dense_flex
    y = model(x, cond, doc_id)
    sub_diff = (y - y_flex).abs().max().item()
    (y * dy).sum().backward()
    fx = dict(
        name=name, cfg=dict(vars(cfg)), B=B, L=L, doc_id=doc_id, flex_vs_dense_fwd_maxabs=sub_diff,
        state=dict((k, v.detach().to(torch.bfloat16)) for k, v in model.state_dict().items()),
        rope=[(blk.attn.rope.cos.clone(), blk.attn.rope.sin.clone()) for blk in model.blocks],
        x=x.detach().to(torch.bfloat16), cond=cond.detach().to(torch.bfloat16), dy=dy.to(torch.bfloat16),
        y=y.detach(), dx=x.grad.clone(), dcond=cond.grad.clone(),
        grads=dict((k, p.grad.clone()) for k, p in model.named_parameters()),
    )
    torch.save(fx, os.path.join(HERE, f"{name}.pt"))
    print(f"{name}: y absmax {y.abs().max():.3f}  flex-vs-dense fwd diff {sub_diff:.2e}")


def run_kv_rollout(A, K, flex_real):
    """Prefill 3 frames (cache on) -> decode frame 3 twice (denoise step, then caching step) ->
    truncate oldest frame -> decode frame 4.  Mirrors av_caching_v2.py:55-140 on the DiT stack."""
    cfg = make_cfg(d_model=128, n_heads=2, n_layers=2, tokens_per_frame=5, n_frames=8, sample_size=2,
                   has_audio=True, rope_impl="ortho", local_window=2)
    model = build(A, cfg, seed=77).eval()
    A.flex_attention = flex_real  # forward only: the real CPU flex path
    d, T, B = cfg.d_model, cfg.tokens_per_frame, 2
    g = torch.Generator().manual_seed(99)
    xs = bf16r(torch.randn(B, 6 * T, d, generator=g))
    conds = bf16r(torch.randn(B, 6, d, generator=g))
    out = {}
    with torch.no_grad():
        cache = K.KVCache(cfg).to("cpu", torch.float32)
        cache.reset(B)
        cache.enable_cache_updates()
        out["prefill"] = model(xs[:, :3 * T], conds[:, :3], None, cache)
        cache.disable_cache_updates()
        out["full4"] = model(xs[:, :4 * T], conds[:, :4])[:, 3 * T:]          # no-cache reference of frame 3
        out["masked_cached"] = model(xs[:, 3 * T:4 * T], conds[:, 3:4], None, cache)  # cached, masked (q_offset) path
        model.enable_decoding()
        out["decode3"] = model(xs[:, 3 * T:4 * T], conds[:, 3:4], None, cache)
        cache.enable_cache_updates()
        out["decode3_cache"] = model(xs[:, 3 * T:4 * T], conds[:, 3:4], None, cache)
        cache.disable_cache_updates()
        out["len_after"] = [cache.length_at(i) for i in range(cfg.n_layers)]
        out["offsets_after"] = list(cache.offsets)
        cache.truncate(1, front=False)
        out["decode4"] = model(xs[:, 4 * T:5 * T], conds[:, 4:5], None, cache)
        out["k_layer1"] = cache.cache[1][0].clone()
        out["v_layer1"] = cache.cache[1][1].clone()
    fx = dict(name="kv_rollout", cfg=dict(vars(cfg)), B=B,
              state=dict((k, v.detach().to(torch.bfloat16)) for k, v in model.state_dict().items()),
              rope=[(blk.attn.rope.cos.clone(), blk.attn.rope.sin.clone()) for blk in model.blocks],
              xs=xs.to(torch.bfloat16), conds=conds.to(torch.bfloat16), out=out)
    torch.save(fx, os.path.join(HERE, "kv_rollout.pt"))
    print("kv_rollout: decode3 vs full4 maxabs", (out["decode3"] - out["full4"]).abs().max().item(),
          " masked_cached vs full4", (out["masked_cached"] - out["full4"]).abs().max().item())


def main():
    install_stubs()
    import owl_wms.nn.attn as A
    import owl_wms.nn.kv_cache as K
    flex_real = A.flex_attention
    for name, spec in CASES.items():
        run_case(A, name, spec, flex_real)
    run_kv_rollout(A, K, flex_real)


if __name__ == "__main__":
    main()
