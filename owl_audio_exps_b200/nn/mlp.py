"""Drop-in for owl_wms/nn/mlp.py: fc1 -> SiLU -> fc2 as two tcgen05 GEMMs with bias+SiLU fused into
the first epilogue.  Parameter names, shapes and the custom init (mlp.py:13-20) are the reference's."""
import torch
from torch import nn

from .. import functional as OF


class MLPCustom(nn.Module):
    def __init__(self, dim_in, dim_middle, dim_out):
        super().__init__()
        self.fc1 = nn.Linear(dim_in, dim_middle)
        self.fc2 = nn.Linear(dim_middle, dim_out)

        nn.init.kaiming_normal_(self.fc1.weight)
        nn.init.kaiming_normal_(self.fc2.weight)
        self.fc1.weight.data *= dim_in ** -0.5
        self.fc2.weight.data *= dim_middle ** -0.5
        nn.init.zeros_(self.fc1.bias)
        nn.init.zeros_(self.fc2.bias)
        self._stage = OF.WeightStage()

    def forward(self, x):
        h = OF.linear(x, self.fc1.weight, self.fc1.bias, self._stage, "fc1", silu_out=True)
        return OF.linear(h, self.fc2.weight, self.fc2.bias, self._stage, "fc2")



class MLP(MLPCustom):
    def __init__(self, config):
        super().__init__(config.d_model, config.d_model * 4, config.d_model)
