"""Drop-in for owl_wms/nn/rope.py: RoPE base + OrthoRoPE / MotionRoPE / Audio1DRoPE frequency layouts.

Hot path: inside Attn / DiTBlock the rotation is fused with the QK-RMSNorm and the head split in one
kernel (csrc/qk_rope.cu) that reads the fp32 cos/sin tables held here; `RoPE.forward` on a bare
[B,H,L,dh] tensor (the reference signature, rope.py:43-51) is served by the same kernel family.

Tables are init-time work.  The reference builds them with the un-vendored third-party
`rotary_embedding_torch` (requirements.txt:5, `RotaryEmbedding(...).freqs / get_axial_freqs`); that
package's published closed forms are restated in `_lang_freqs`, `_pixel_freqs`, `_axial_freqs`:
  'lang'  freqs = theta ** -(arange(0, dim, 2) / dim)
  'pixel' freqs = linspace(1, max_freq / 2, dim // 2) * pi
  forward(t)      = outer(t, freqs) with every frequency repeated twice
  get_axial_freqs = per axis: pos = linspace(-1, 1, n) ('pixel') (+ offset), forward(pos), broadcast, concat
PARITY UNPINNED for the 'pixel' / axial form (OrthoRoPE): the reference has no golden vector for it
(its only checksum, sanity/rope_test.py:91, targets a deleted class), the package is not installed in the build
image, and its releases differ on which axes of `get_axial_freqs` use linspace(-1, 1) (all of them, or only the
last two with arange on the leading ones).  So:
  * when `rotary_embedding_torch` IS importable (any real deployment of the reference has it: requirements.txt:5),
    OrthoRoPE builds its table by calling the package exactly as the reference does (rope.py:65-73) -- init-time
    plumbing, identical tables by construction, whatever the installed release does;
  * otherwise the restatement below is used; `config.rope_axial_variant` ("all_linspace", the default, or
    "last_two_linspace") selects the release behaviour being restated.
The cos/sin buffers are non-persistent (rope.py:40-41), so a checkpoint never carries them: a reference AV
checkpoint loaded here runs with whatever table this file builds.  The hot-path tests feed identical tables to
both sides, so block parity does not depend on this restatement.
"""
import math

import torch
from torch import nn


def get_rope_cls(cls_name):
    cls_name = cls_name.lower()
    if cls_name == "ortho":
        return OrthoRoPE
    elif cls_name == "motion":
        return MotionRoPE
    elif cls_name == "audio1d":
        return Audio1DRoPE
    raise ValueError(f"Invalid RoPE class: {cls_name}")


def cast_rope_buffers_to_fp32(module):
    """rope.py:22-28."""
    for submodule in module.modules():
        if isinstance(submodule, RoPE):
            if hasattr(submodule, "cos"):
                submodule.cos = submodule.cos.float()
            if hasattr(submodule, "sin"):
                submodule.sin = submodule.sin.float()


def _lang_freqs(dim: int, theta: float = 10000.0) -> torch.Tensor:
    return 1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim))


def _pixel_freqs(dim: int, max_freq: float) -> torch.Tensor:
    return torch.linspace(1.0, max_freq / 2, dim // 2) * math.pi


def _axial_freqs(freqs: torch.Tensor, dims, offsets, pixel: bool, variant: str = "all_linspace") -> torch.Tensor:
    parts = []
    for i, n in enumerate(dims):
        lin = pixel and (variant == "all_linspace" or i >= len(dims) - 2)
        pos = torch.linspace(-1, 1, steps=n) if lin else torch.arange(n).float()
        pos = pos + (0 if offsets is None else offsets[i])
        ang = (pos[:, None] * freqs[None, :]).repeat_interleave(2, dim=-1)  # [n, 2*len(freqs)]
        shape = [1] * len(dims) + [ang.shape[-1]]
        shape[i] = n
        parts.append(ang.view(shape))
    parts = torch.broadcast_tensors(*[p.expand(*dims, p.shape[-1]) for p in parts])
    return torch.cat(parts, dim=-1)


class RoPE(nn.Module):
    def __init__(self, config):
        super().__init__()
        freqs = self.get_freqs(config)
        if not config.has_audio:
            # subclasses' freqs include the audio token: drop the last item of each frame (rope.py:36-39)
            freqs = freqs.view(config.n_frames, -1, freqs.size(-1))[:, :-1].flatten(0, 1)
        self.cos = nn.Buffer(freqs.cos().contiguous(), persistent=False)
        self.sin = nn.Buffer(freqs.sin().contiguous(), persistent=False)

    def _apply(self, fn, recurse=True):
        """`.to(device)` moves the tables, dtype casts (`model.bfloat16()`, causvid_pipeline.py:36) never touch them:
        the kernels read fp32 tables (rope.py:45 asserts fp32; the reference's trainers call cast_rope_buffers_to_fp32
        after the cast, which would rebuild them from bf16-rounded values)."""
        cos, sin = self.cos, self.sin
        super()._apply(fn, recurse) if recurse is not True else super()._apply(fn)
        for name, old in (("cos", cos), ("sin", sin)):
            new = getattr(self, name)
            if new.dtype != torch.float32:
                setattr(self, name, nn.Buffer(old.to(device=new.device), persistent=False))
        return self

    def forward(self, x, offset: int = 0):
        """x [B,H,L,dh]: rotate pairs (2i, 2i+1); output halves are [y_even | y_odd] (rope.py:43-51)."""
        from .. import functional as OF
        assert self.cos.dtype == torch.float32
        return OF.rope_only(x, self.cos, self.sin, offset)

    def get_freqs(self, config):
        raise NotImplementedError


class OrthoRoPE(RoPE):
    """Axial time / height / width rotation (rope.py:57-79)."""

    def get_freqs(self, config):
        p = config.sample_size
        head_dim = config.d_model // config.n_heads
        try:   # the reference's own table builder, when its third-party dependency is installed (rope.py:65-73)
            from rotary_embedding_torch import RotaryEmbedding
            pos_emb = RotaryEmbedding(dim=head_dim // 4, freqs_for="pixel", max_freq=256)
            freqs = pos_emb.get_axial_freqs(config.n_frames, p + 1, p + 1, 1, offsets=(0, 0, 0, 1)).detach().float()
        except ImportError:
            base = _pixel_freqs(head_dim // 4, max_freq=256)
            freqs = _axial_freqs(base, (config.n_frames, p + 1, p + 1, 1), (0, 0, 0, 1), pixel=True,
                                 variant=getattr(config, "rope_axial_variant", "all_linspace"))
        freqs = freqs.view(config.n_frames, p + 1, p + 1, -1)
        vid = freqs[:, :p, :p].reshape(config.n_frames, p ** 2, -1)
        aud = freqs[:, -1, -1].unsqueeze(1)
        freqs = torch.cat([vid, aud], dim=1).flatten(0, 1)
        return freqs[..., ::2]


class MotionRoPE(RoPE):
    """VideoRoPE-style diagonal layout (rope.py:82-152): spatial coordinates drift linearly with time."""

    def get_freqs(self, config):
        H = W = config.sample_size
        F = config.n_frames
        d_head = config.d_model // config.n_heads
        dt = getattr(config, "rope_dim_t", d_head * 2 // 8)
        dx = getattr(config, "rope_dim_x", d_head * 3 // 8)
        dy = getattr(config, "rope_dim_y", d_head * 3 // 8)
        theta = getattr(config, "rope_base", 10000.0)
        delta = getattr(config, "rope_ats_delta", 2.0)

        base = _lang_freqs(dt + dx + dy, theta).float()
        f_sp, f_t = torch.split(base, [(dx + dy) // 2, dt // 2])
        f_x, f_y = f_sp[::2], f_sp[1::2]

        t = torch.arange(F, dtype=torch.float32) * delta
        hg = torch.arange(H, dtype=torch.float32) - (H - 1) / 2.0
        wg = torch.arange(W, dtype=torch.float32) - (W - 1) / 2.0
        # video tokens ordered (f, h, w); one audio token closes each frame
        t_v = t[:, None, None].expand(F, H, W)
        x_v = t_v + wg[None, None, :]
        y_v = t_v + hg[None, :, None]
        x_pos = torch.cat([x_v.reshape(F, -1), t[:, None]], dim=1).flatten()
        y_pos = torch.cat([y_v.reshape(F, -1), (t + (H - 1) / 2.0 + 1.0)[:, None]], dim=1).flatten()
        t_pos = torch.cat([t_v.reshape(F, -1), t[:, None]], dim=1).flatten()

        ang_x = x_pos[:, None] * f_x[None, :]
        ang_y = y_pos[:, None] * f_y[None, :]
        ang_t = t_pos[:, None] * f_t[None, :]
        spatial = torch.stack([ang_x, ang_y], dim=-1).flatten(1)  # interleave x, y
        return torch.cat([spatial, ang_t], dim=-1)


class Audio1DRoPE(RoPE):
    """1-D temporal RoPE for audio latents (rope.py:155-179); n_frames is the table length."""

    def get_freqs(self, config):
        head_dim = config.d_model // config.n_heads
        pos = torch.arange(config.n_frames, dtype=torch.float32)
        return pos[:, None] * _lang_freqs(head_dim)[None, :]
