"""ORACLE -- test infrastructure, not product code.

CPU restatement (plain torch, fp32 or fp64, no custom kernels) of the reference algorithm for the
DiT-block hot path of shahbuland/owl-audio-exps.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this; the product package
(owl_audio_exps_b200) never does.

Every function cites the reference lines it restates (paths relative to the reference root).
Pinned against the real reference: tests/golden/*.pt were produced by importing the reference's own
`owl_wms/nn` modules in the build container (tests/golden/make_golden.py) and
tests/test_oracle_golden.py checks this file against them (forward, all gradients, KV-cache
prefill/decode).  The reference's own tests hold no golden vectors for this path, and
the RoPE frequency tables of OrthoRoPE come from the un-vendored `rotary_embedding_torch`:
PARITY UNPINNED at that table-construction boundary only (tables are passed in, not rebuilt here).

Weights are passed as a flat dict with the reference's state-dict key names.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ---------------------------------------------------------------------------------------------
# normalization.py:10-11  rms_norm = F.rms_norm(x, (D,)) with eps=None
# ---------------------------------------------------------------------------------------------
RMS_EPS = 1.1920928955078125e-07  # what torch 2.11 uses for eps=None


def rms_norm(x: Tensor, eps: float = RMS_EPS) -> Tensor:
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps)


def layer_norm(x: Tensor, eps: float = 1e-5) -> Tensor:
    """normalization.py:6-7."""
    mu = x.mean(-1, keepdim=True)
    var = (x - mu).pow(2).mean(-1, keepdim=True)
    return (x - mu) * torch.rsqrt(var + eps)


# ---------------------------------------------------------------------------------------------
# rope.py:43-51  RoPE.forward
# ---------------------------------------------------------------------------------------------
def rope(x: Tensor, cos: Tensor, sin: Tensor, offset: int = 0) -> Tensor:
    """x [B,H,L,dh]; cos/sin [L_max, dh/2].  Pairs (2i, 2i+1) rotate; output = cat(y_even, y_odd)."""
    L = x.size(2)
    c, s = cos[offset:offset + L].to(x.dtype), sin[offset:offset + L].to(x.dtype)
    x0, x1 = x[..., 0::2], x[..., 1::2]
    return torch.cat((x0 * c - x1 * s, x1 * c + x0 * s), dim=-1)


# ---------------------------------------------------------------------------------------------
# attn.py:24-62  get_block_mask -> dense boolean mask
# ---------------------------------------------------------------------------------------------
def frame_mask(n_q: int, n_kv: int, tokens_per_frame: int, window_len: Optional[int], doc_id: Optional[Tensor],
               q_offset: int, is_causal: bool, batch: int = 1) -> Tensor:
    """-> bool [B or 1, 1, n_q, n_kv]; True = attend.  n_kv = n_q + q_offset (attn.py:61-62)."""
    q = torch.arange(n_q)[:, None] + q_offset
    kv = torch.arange(n_kv)[None, :]
    fq, fk = q // tokens_per_frame, kv // tokens_per_frame
    n_frames = n_kv // tokens_per_frame
    if window_len is None:
        window_len = n_frames  # attn.py:40-41
    m = (fq - fk).abs() < window_len
    if is_causal:
        m = m & (fk <= fq)
    m = m[None, None]
    if doc_id is not None:
        d = doc_id.long()
        same = d[:, fq.squeeze(1)][:, :, None] == d[:, fk.squeeze(0)][:, None, :]
        m = m & same[:, None]
    return m


def sdpa(q: Tensor, k: Tensor, v: Tensor, mask: Optional[Tensor]) -> Tensor:
    """flex_attention(q,k,v[,block_mask]) (attn.py:106-109): softmax(q k^T / sqrt(dh) + mask) v."""
    s = (q @ k.transpose(-1, -2)) / math.sqrt(q.size(-1))
    if mask is not None:
        s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)  # fully-masked rows produce 0 (flex convention)
    return p @ v


# ---------------------------------------------------------------------------------------------
# kv_cache.py:12-104  SingleKVCache (only what Attn.forward touches)
# ---------------------------------------------------------------------------------------------
class OracleKVCache:
    def __init__(self, n_layers: int, tokens_per_frame: int):
        self.n_layers, self.tokens_per_frame = n_layers, tokens_per_frame
        self.cache: List[Optional[Tuple[Tensor, Tensor]]] = [None] * n_layers
        self.offsets = [0] * n_layers
        self.should_update = False

    def length_at(self, i: int) -> int:
        return 0 if self.cache[i] is None else self.cache[i][0].shape[2]

    def update(self, k: Tensor, v: Tensor, i: int):            # kv_cache.py:50-58
        self.offsets[i] += k.shape[2] - self.length_at(i)
        self.cache[i] = (k.detach(), v.detach())

    def truncate(self, n_frames: int, front: bool = False):    # kv_cache.py:60-75
        amt = n_frames * self.tokens_per_frame
        for i in range(self.n_layers):
            k, v = self.cache[i]
            self.cache[i] = (k[:, :, :-amt], v[:, :, :-amt]) if front else (k[:, :, amt:], v[:, :, amt:])


# ---------------------------------------------------------------------------------------------
# modulation.py / mlp.py / attn.py modules as functions of a weight dict
# ---------------------------------------------------------------------------------------------
def _lin(x: Tensor, W: Dict[str, Tensor], name: str) -> Tensor:
    return F.linear(x, W[name + ".weight"], W.get(name + ".bias"))


def _per_frame(t: Tensor, m: int) -> Tensor:
    """modulation.py:20-22: [b, n, c] -> [b, n*m, c] by repeating each frame's row m times."""
    b, n, c = t.shape
    return t.view(b, n, 1, c).expand(-1, -1, m, -1).reshape(b, n * m, c)


def adaln(x: Tensor, cond: Tensor, W: Dict[str, Tensor], name: str) -> Tensor:
    """modulation.py:12-26."""
    m = x.shape[1] // cond.shape[1]
    ab = _per_frame(_lin(F.silu(cond), W, name + ".fc"), m)
    a, b = ab.chunk(2, dim=-1)
    return rms_norm(x) * (1 + a) + b


def gate(x: Tensor, cond: Tensor, W: Dict[str, Tensor], name: str) -> Tensor:
    """modulation.py:33-43."""
    m = x.shape[1] // cond.shape[1]
    return _per_frame(_lin(F.silu(cond), W, name + ".fc_c"), m) * x


def mlp(x: Tensor, W: Dict[str, Tensor], name: str) -> Tensor:
    """mlp.py:33-37."""
    return _lin(F.silu(_lin(x, W, name + ".fc1")), W, name + ".fc2")


def attn(x: Tensor, W: Dict[str, Tensor], name: str, n_heads: int, cos: Tensor, sin: Tensor,
         mask: Optional[Tensor], kv_cache: Optional[OracleKVCache] = None, layer_idx: int = 0,
         local: bool = False, local_offset: int = 0, decoding: bool = False) -> Tensor:
    """attn.py:79-113.  `decoding` = block_mask None."""
    B, L, d = x.shape
    dh = d // n_heads
    qkv = _lin(x, W, name + ".qkv")
    q, k, v = qkv.view(B, L, 3, n_heads, dh).permute(2, 0, 3, 1, 4)        # 'b t (three h d) -> three b h t d'
    q, k = rms_norm(q), rms_norm(k)
    offset = kv_cache.offsets[layer_idx] if kv_cache is not None else 0
    q, k = rope(q, cos, sin, offset), rope(k, cos, sin, offset)
    if offset > 0:
        ok, ov = kv_cache.cache[layer_idx]
        k, v = torch.cat([ok, k], dim=2), torch.cat([ov, v], dim=2)
    if kv_cache is not None and kv_cache.should_update:
        kv_cache.update(k, v, layer_idx)
    if local and decoding:
        k, v = k[:, :, -local_offset:], v[:, :, -local_offset:]
    o = sdpa(q, k, v, None if decoding else mask)
    o = o.permute(0, 2, 1, 3).reshape(B, L, d)
    return _lin(o, W, name + ".out")


def dit_block(x: Tensor, cond: Tensor, W: Dict[str, Tensor], prefix: str, n_heads: int, cos: Tensor, sin: Tensor,
              mask: Optional[Tensor], **attn_kw) -> Tensor:
    """attn.py:130-143."""
    p = prefix
    h = adaln(x, cond, W, p + "adaln1")
    h = attn(h, W, p + "attn", n_heads, cos, sin, mask, **attn_kw)
    x = x + gate(h, cond, W, p + "gate1")
    h = adaln(x, cond, W, p + "adaln2")
    h = mlp(h, W, p + "mlp")
    return x + gate(h, cond, W, p + "gate2")


def dit_forward(x: Tensor, cond: Tensor, W: Dict[str, Tensor], cfg, rope_tables: List[Tuple[Tensor, Tensor]],
                doc_id: Optional[Tensor] = None, kv_cache: Optional[OracleKVCache] = None,
                decoding: bool = False, prefix: str = "") -> Tensor:
    """attn.py:175-191 (DiT.forward): local/global layer pattern attn.py:151-153."""
    L = x.shape[1]
    q_off = kv_cache.length_at(0) if kv_cache is not None else 0
    local_idx = getattr(cfg, "local_idx", 4)
    lm = gm = None
    if not decoding:
        lm = frame_mask(L, L + q_off, cfg.tokens_per_frame, cfg.local_window, doc_id, q_off, cfg.causal)
        gm = frame_mask(L, L + q_off, cfg.tokens_per_frame, cfg.global_window, doc_id, q_off, cfg.causal)
    for i in range(cfg.n_layers):
        local = (i % local_idx) != 0
        cos, sin = rope_tables[i]
        x = dit_block(x, cond, W, f"{prefix}blocks.{i}.", cfg.n_heads, cos, sin, lm if local else gm,
                      kv_cache=kv_cache, layer_idx=i, local=local,
                      local_offset=cfg.local_window * cfg.tokens_per_frame, decoding=decoding)
    return x


def final_layer(x: Tensor, cond: Tensor, W: Dict[str, Tensor], prefix: str) -> Tensor:
    """attn.py:264-277."""
    return _lin(F.silu(adaln(x, cond, W, prefix + "norm")), W, prefix + "proj")


def sincos_embed(t: Tensor, dim: int, theta: float = 300, mult: float = 1000) -> Tensor:
    """embeddings.py:30-72 (SinCosEmbed): [..] -> [.., dim] = cat(sin, cos) of t*mult*exp(-i*log(theta)/(dim/2-1))."""
    half = dim // 2
    step = torch.log(torch.tensor(float(theta))) / (half - 1)
    freqs = torch.exp(torch.arange(half) * -step).to(t.dtype)
    arg = (t * mult).unsqueeze(-1) * freqs
    return torch.cat((torch.sin(arg), torch.cos(arg)), dim=-1)


def timestep_embedding(t: Tensor, W: Dict[str, Tensor], prefix: str) -> Tensor:
    """embeddings.py:74-84 (TimestepEmbedding): SinCosEmbed(512, 300, 1000) -> MLPCustom."""
    return mlp(sincos_embed(t, 512), W, prefix + "mlp")


def audio_rft_core(x: Tensor, t: Tensor, W: Dict[str, Tensor], cfg, rope_tables: List[Tuple[Tensor, Tensor]],
                   doc_id: Optional[Tensor] = None, kv_cache: Optional[OracleKVCache] = None,
                   decoding: bool = False) -> Tensor:
    """models/audiorft.py:34-57 (AudioRFTCore.forward): t_embed -> proj_in -> DiT -> FinalLayer."""
    cond = timestep_embedding(t, W, "t_embed.")
    h = _lin(x, W, "proj_in")
    h = dit_forward(h, cond, W, cfg, rope_tables, doc_id=doc_id, kv_cache=kv_cache, decoding=decoding,
                    prefix="transformer.")
    return final_layer(h, cond, W, "proj_out.")


def audio_cached_rollout(x: Tensor, W: Dict[str, Tensor], cfg, rope_tables: List[Tuple[Tensor, Tensor]], dt: List[float],
                         num_tokens: int, noise_prev: float, max_window: Optional[int], noise) -> Tensor:
    """sampling/audio_caching.py:55-151 (AudioCachingSampler.__call__): `noise(shape)` supplies every N(0,1) draw in
    the reference's order (context re-noise, then per token: start latent, committed-latent re-noise)."""
    b, _, ch = x.shape
    cache = OracleKVCache(cfg.n_layers, cfg.tokens_per_frame)
    latents = [x.clone()]

    def commit(tok):
        noisy = tok * (1.0 - noise_prev) + noise(tuple(tok.shape)) * noise_prev
        cache.should_update = True
        audio_rft_core(noisy, torch.full((b, tok.shape[1]), noise_prev), W, cfg, rope_tables, kv_cache=cache,
                       decoding=len(latents) > 1)
        cache.should_update = False

    commit(x)
    for _ in range(num_tokens):
        cur, t = noise((b, 1, ch)), torch.ones(b, 1)
        for d in dt:
            cur = cur - d * audio_rft_core(cur, t, W, cfg, rope_tables, kv_cache=cache, decoding=True)
            t = t - d
        latents.append(cur)
        commit(cur)
        if max_window is not None and len(latents) > max_window:
            cache.truncate(1, front=False)
    return torch.cat(latents, dim=1)


def mouse_embedding(m: Tensor, W: Dict[str, Tensor], prefix: str, dim: int = 512) -> Tensor:
    """embeddings.py:119-156 (MouseEmbedding): symlog, polar split, angle Linear(2, dim/2) | SinCosEmbed(dim/2) -> MLP."""
    m = torch.sign(m) * torch.log1p(torch.abs(m))
    ang = torch.atan2(m[..., 1], m[..., 0])
    mag = torch.norm(m, dim=-1)
    ang_emb = _lin(torch.stack([torch.cos(ang), torch.sin(ang)], dim=-1), W, prefix + "angle_proj")
    return mlp(torch.cat([ang_emb, sincos_embed(mag, dim // 2)], dim=-1), W, prefix + "mlp")


def control_embedding(mouse: Tensor, btn: Tensor, W: Dict[str, Tensor], prefix: str) -> Tensor:
    """embeddings.py:158-184 (ButtonEmbeddding + ControlEmbedding)."""
    return mouse_embedding(mouse, W, prefix + "mouse.") + mlp(btn * 2 - 1, W, prefix + "button.proj")


def game_rft_core(x: Tensor, t: Tensor, mouse: Tensor, btn: Tensor, W: Dict[str, Tensor], cfg,
                  rope_tables: List[Tuple[Tensor, Tensor]], doc_id: Optional[Tensor] = None,
                  has_controls: Optional[Tensor] = None, kv_cache: Optional[OracleKVCache] = None,
                  decoding: bool = False) -> Tensor:
    """models/gamerft.py:33-60 (GameRFTCore.forward)."""
    b, n, c, h, w = x.shape
    cond = timestep_embedding(t, W, "t_embed.")
    if not cfg.uncond:
        ctrl = control_embedding(mouse, btn, W, "control_embed.")
        if has_controls is not None:
            ctrl = torch.where(has_controls[:, None, None], ctrl, torch.zeros_like(ctrl))
        cond = cond + ctrl
    tok = x.permute(0, 1, 3, 4, 2).reshape(b, n * h * w, c)
    tok = dit_forward(_lin(tok, W, "proj_in"), cond, W, cfg, rope_tables, doc_id=doc_id, kv_cache=kv_cache,
                      decoding=decoding, prefix="transformer.")
    tok = final_layer(tok, cond, W, "proj_out.")
    return tok.view(b, n, h, w, c).permute(0, 1, 4, 2, 3)


def av_cached_rollout(x: Tensor, mouse: Tensor, btn: Tensor, W: Dict[str, Tensor], cfg,
                      rope_tables: List[Tuple[Tensor, Tensor]], dt: List[float], num_frames: int, cfg_scale: float,
                      noise_prev: float, max_window: Optional[int], noise) -> Tensor:
    """sampling/av_caching_v2.py:46-144 (AVCachingSamplerV2.__call__): per-frame rollout with classifier-free guidance
    on the controls; `noise(shape)` supplies the N(0,1) draws in the reference's order."""
    b, init_len = x.shape[0], x.shape[1]
    cache = OracleKVCache(cfg.n_layers, cfg.tokens_per_frame)
    latents = [x.clone()]

    def run(frames, t, m, bt, decoding):
        return game_rft_core(frames, t, m, bt, W, cfg, rope_tables, kv_cache=cache, decoding=decoding)

    def commit(frames, m, bt):
        noisy = frames * (1.0 - noise_prev) + noise(tuple(frames.shape)) * noise_prev
        cache.should_update = True
        run(noisy, torch.full((b, frames.shape[1]), noise_prev), m, bt, len(latents) > 1)
        cache.should_update = False

    commit(x, mouse[:, :init_len], btn[:, :init_len])
    for idx in range(min(num_frames, mouse.shape[1] - init_len)):
        cur, t = noise(tuple(x[:, :1].shape)), torch.ones(b, 1)
        at = init_len + idx
        m, bt = mouse[:, at:at + 1], btn[:, at:at + 1]
        for d in dt:
            v = run(cur, t, m, bt, True)
            if cfg_scale != 1.0:
                v0 = run(cur, t, torch.zeros_like(m), torch.zeros_like(bt), True)
                v = v0 + cfg_scale * (v - v0)
            cur, t = cur - d * v, t - d
        latents.append(cur)
        commit(cur, m, bt)
        if max_window is not None and len(latents) > max_window:
            cache.truncate(1, front=False)
    return torch.cat(latents, dim=1)


def game_rft_audio_core(x: Tensor, audio: Tensor, t: Tensor, W: Dict[str, Tensor], cfg,
                        rope_tables: List[Tuple[Tensor, Tensor]]) -> Tuple[Tensor, Tensor]:
    """models/gamerft_audio.py:47-98 (GameRFTAudioCore.forward, `dit` backbone, uncond): per frame h*w video tokens then
    the audio token; video head sees layer_norm(video), layer_norm(cond).  Keys are spelled `transformer.*` as in the
    fixture (made with dynamo disabled; a compiled reference checkpoint spells them `transformer._orig_mod.*`)."""
    b, n, c, h, w = x.shape
    cond = timestep_embedding(t, W, "t_embed.")
    vid = _lin(x.permute(0, 1, 3, 4, 2).reshape(b, n * h * w, c), W, "proj_in")
    aud = _lin(audio, W, "audio_proj_in")
    d = vid.shape[-1]
    tok = torch.cat([vid.view(b, n, h * w, d), aud.view(b, n, 1, d)], dim=2).reshape(b, n * (h * w + 1), d)
    tok = dit_forward(tok, cond, W, cfg, rope_tables, prefix="transformer.").view(b, n, h * w + 1, d)
    vid = final_layer(layer_norm(tok[:, :, :-1].reshape(b, n * h * w, d)), layer_norm(cond), W, "proj_out.")
    return vid.reshape(b, n, h, w, c).permute(0, 1, 4, 2, 3), final_layer(tok[:, :, -1], cond, W, "audio_proj_out.")


# ---------------------------------------------------------------------------------------------
# algorithmic work (SURVEY.md section 8d / BASELINE.md section 3) -- used by bench.py for the roofline basis
# ---------------------------------------------------------------------------------------------
def unmasked_pairs(F_: int, T: int, W: Optional[int]) -> int:
    if W is None or W >= F_:
        return T * T * F_ * (F_ + 1) // 2
    return T * T * (W * (W + 1) // 2 + (F_ - W) * W)


def step_flops(B: int, L: int, d: int, T: int, n_layers: int, local_window: Optional[int],
               global_window: Optional[int], local_idx: int = 4) -> Dict[str, float]:
    """fwd+bwd algorithmic FLOPs: 3x linear + 3.5x attention, masked pairs excluded."""
    F_ = L // T
    lin_fwd = B * L * (24 * d * d + 12 * d * d / T) * n_layers
    att_fwd = 0.0
    for i in range(n_layers):
        w = local_window if (i % local_idx) != 0 else global_window
        att_fwd += 4.0 * d * B * unmasked_pairs(F_, T, w)
    return {"linear_fwd": lin_fwd, "attn_fwd": att_fwd, "step": 3 * lin_fwd + 3.5 * att_fwd}


# ---------------------------------------------------------------------------------------------
# muon.py (optimizer neighbour of the path)
# ---------------------------------------------------------------------------------------------
def newton_schulz5(G: Tensor, steps: int, dtype=torch.float32) -> Tensor:
    """owl_wms/muon.py:11-38 (zeropower_via_newtonschulz5) for a 2-D G, iterated in `dtype` (the reference runs bf16)."""
    a, b, c = (3.4445, -4.7750, 2.0315)
    X = G.to(dtype)
    if G.size(-2) > G.size(-1):
        X = X.mT
    X = X / (X.norm(dim=(-2, -1), keepdim=True) + 1e-7)
    for _ in range(steps):
        A = X @ X.mT
        B = b * A + c * A @ A
        X = a * X + B @ X
    if G.size(-2) > G.size(-1):
        X = X.mT
    return X
