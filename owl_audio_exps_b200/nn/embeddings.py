"""Drop-in for the conditioning embeddings of owl_wms/nn/embeddings.py that feed the DiT stack's `cond`
(SURVEY.md section 8f rank 2): SinCosEmbed (:30-72), TimestepEmbedding (:74-84), StepEmbedding (:86-104),
ConditionEmbedding (:106-117), LearnedPosEnc (:14-28).  The sin/cos features are one kernel writing bf16
(`owl_sincos_embed`), the MLPs are the tcgen05 GEMMs of nn/mlp.py.  Parameter names and shapes are the
reference's; Mouse/Button/ControlEmbedding (:119-184) build the game models' control half of `cond`."""
import math

import torch
from torch import nn

from .. import functional as OF
from .. import ops
from .mlp import MLPCustom


class LearnedPosEnc(nn.Module):
    def __init__(self, n_seq, dim):
        super().__init__()
        self.n_seq = n_seq
        self.p = nn.Parameter(torch.randn(n_seq, dim) * 0.02)

    def forward(self, x):
        n = x.shape[1]
        p = self.p[-n:] if n < self.n_seq else self.p
        return x + p.unsqueeze(0)


class SinCosEmbed(nn.Module):
    """[..] scalars -> [.., dim] = cat(sin(x*mult*f), cos(x*mult*f)), f_i = exp(-i*log(theta)/(dim/2-1)).
    Output is bf16 (what the consuming Linear sees under the trainers' autocast); the argument is formed in fp32."""

    def __init__(self, dim, theta=300, mult=1000):
        super().__init__()
        self.dim = dim
        self.theta = theta
        self.mult = mult
        self._freqs = {}

    def freqs(self, device):
        f = self._freqs.get(device)
        if f is None:
            half = self.dim // 2
            step = torch.log(torch.tensor(self.theta)) / (half - 1)    # the reference's op order, fp32 on the host
            f = torch.exp(torch.arange(half) * -step).to(device=device, dtype=torch.float32)
            self._freqs[device] = f
        return f

    def forward(self, x, device=None):
        if not isinstance(x, torch.Tensor):
            x = torch.tensor([x] if isinstance(x, float) else x, device=device)
        if x.dim() == 0:
            x = x.unsqueeze(0)
        if x.dim() > 2:
            raise ValueError("SinCosEmbed takes a scalar, [b] or [b, n]")
        if not x.is_cuda:
            raise RuntimeError("owl_audio_exps_b200: CUDA tensors required (sm_100a kernels only; there is no CPU path)")
        t = x.detach().to(torch.float32).contiguous()
        out = ops.sincos_embed(t.view(-1), self.freqs(t.device), float(self.mult))
        return out.view(*x.shape, 2 * (self.dim // 2))


class TimestepEmbedding(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.sincos = SinCosEmbed(512, theta=300, mult=1000)
        self.mlp = MLPCustom(512, dim * 4, dim)

    def forward(self, x):
        if not isinstance(x, torch.Tensor):
            return self.mlp(self.sincos(x, device=self.mlp.fc1.weight.device))
        return self.mlp(self.sincos(x))


class StepEmbedding(nn.Module):
    def __init__(self, d_out, d_in=512, max_steps=128):
        super().__init__()
        self.mlp = MLPCustom(d_in, dim_middle=4 * d_out, dim_out=d_out)
        self.max_steps = max_steps
        self.sincos = SinCosEmbed(d_in, theta=300, mult=1000 / math.log2(max_steps))

    def forward(self, steps):
        if not isinstance(steps, torch.Tensor):
            steps = torch.tensor(steps, device=self.mlp.fc1.weight.device, dtype=self.mlp.fc1.weight.dtype)
        if steps.ndim == 0:
            steps = steps.unsqueeze(0)
        t = (math.log2(self.max_steps) - torch.log2(steps.float())).to(steps.dtype)   # steps -> [0, log2(max_steps)]
        return self.mlp(self.sincos(t))


class ConditionEmbedding(nn.Module):
    def __init__(self, n_classes, dim):
        super().__init__()
        self.embedding = nn.Embedding(n_classes, dim)
        self.mlp = MLPCustom(dim, dim * 4, dim)

    def forward(self, x):
        return self.mlp(self.embedding(x))


class MouseEmbedding(nn.Module):
    """embeddings.py:119-156: symlog -> polar; angle (cos, sin) through Linear(2, dim/2), magnitude through
    SinCosEmbed(dim/2); concatenated into the MLP."""

    def __init__(self, dim_out, dim=512):
        super().__init__()
        self.angle_proj = nn.Linear(2, dim // 2, bias=False)
        self.magnitude_embed = SinCosEmbed(dim // 2)
        self.mlp = MLPCustom(dim, dim * 4, dim_out)
        self._stage = OF.WeightStage()

    def forward(self, x):
        with torch.no_grad():
            x = torch.sign(x) * torch.log1p(torch.abs(x))
            angles = torch.atan2(x[..., 1], x[..., 0])
            magnitudes = torch.norm(x, dim=-1)
            angle_emb = torch.stack([torch.cos(angles), torch.sin(angles)], dim=-1).to(x.dtype)
            magnitude_emb = self.magnitude_embed(magnitudes)
        angle_emb = OF.linear(angle_emb, self.angle_proj.weight, None, self._stage, "angle_proj")
        return self.mlp(torch.cat([angle_emb, magnitude_emb], dim=-1))


class ButtonEmbeddding(nn.Module):
    """embeddings.py:158-168 (the reference's spelling): {0,1} -> {-1,+1} -> MLP."""

    def __init__(self, n_buttons, dim_out, dim=512):
        super().__init__()
        self.proj = MLPCustom(n_buttons, dim * 4, dim_out)

    def forward(self, x):
        return self.proj((x * 2) - 1)


class ControlEmbedding(nn.Module):
    """embeddings.py:170-184: mouse [b,n,2] + button [b,n,n_buttons] -> [b,n,d]."""

    def __init__(self, n_buttons, dim_out, dim=512):
        super().__init__()
        self.mouse = MouseEmbedding(dim_out, dim)
        self.button = ButtonEmbeddding(n_buttons, dim_out, dim)

    def forward(self, mouse, button, has_controls=None):
        return self.mouse(mouse) + self.button(button)
