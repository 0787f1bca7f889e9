"""Drop-in for owl_wms/nn/normalization.py.

`rms_norm` (normalization.py:10-11) is on the hot path (adaLN, QK-norm) and runs the vectorised
warp-shuffle kernel; the fused block never calls it on its own (the norm is folded into the adaLN
modulate and the QK-norm+RoPE kernels).  `layer_norm` (normalization.py:6-7) is used by
gamerft_audio.py:89.  RMSNorm / L2Norm / QKNorm are kept for surface compatibility; no DiT config
instantiates RMSNorm or L2Norm.
"""
import torch
from torch import nn

from .. import functional as OF
from .. import ops


def layer_norm(x: torch.Tensor) -> torch.Tensor:
    """Weight-less LayerNorm over the last dim (normalization.py:6-7), forward and backward on the kernel path."""
    return OF.layer_norm(x)


def rms_norm(x: torch.Tensor) -> torch.Tensor:
    return OF.rms_norm(x)


class RMSNorm(nn.Module):
    """normalization.py:14-27 (learned gain; off the DiT path)."""

    def __init__(self, dim):
        super().__init__()
        self.gain = nn.Parameter(torch.randn(dim) * 0.02)
        self.eps = 1e-6

    def forward(self, x):
        rms = x.float().pow(2).mean(dim=-1, keepdim=True)
        norm = rms.add(self.eps).rsqrt().to(x.dtype)
        return x * norm * (1. + self.gain[None, None, None, :])


class L2Norm(nn.Module):
    def forward(self, x):
        return torch.nn.functional.normalize(x, dim=-1)


class QKNorm(nn.Module):
    def __init__(self, dim):
        super().__init__()

    def forward(self, q, k):
        return rms_norm(q), rms_norm(k)
