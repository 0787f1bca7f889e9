"""Drop-in for owl_wms/nn/modulation.py (AdaLN, Gate, cond_adaln, cond_gate).

Standalone calls run: SiLU folded into the projection GEMM's operand staging, then ONE pass of the
adaLN-modulate kernel (rms_norm, (1+a), +b with the per-frame broadcast done by index, never
materialised -- modulation.py:20-25 expands [b,n,2d] to [b,n*m,2d]).  Inside DiTBlock these modules
only hold parameters; the block runs the fused schedule in functional.DiTBlockFn."""
from torch import nn

from .. import functional as OF


class AdaLN(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.fc = nn.Linear(dim, 2 * dim)
        self._stage = OF.WeightStage()

    def forward(self, x, cond):
        ab = OF.linear(cond, self.fc.weight, self.fc.bias, self._stage, "fc", silu_in=True)  # [b, n, 2d]
        return OF.adaln_modulate(x, ab)



class Gate(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.fc_c = nn.Linear(dim, dim)
        self._stage = OF.WeightStage()

    def forward(self, x, cond):
        c = OF.linear(cond, self.fc_c.weight, self.fc_c.bias, self._stage, "fc_c", silu_in=True)  # [b, n, d]
        return OF.gate(x, c)



def cond_adaln(x, scale, bias):
    """modulation.py:46-55: scale, bias [b, n, d]; x [b, n*m, d]."""
    import torch
    return OF.adaln_modulate(x, torch.cat([scale, bias], dim=-1))


def cond_gate(x, gate):
    """modulation.py:57-63."""
    return OF.gate(x, gate)
