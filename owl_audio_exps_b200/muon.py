"""Drop-in for owl_wms/muon.py: Muon with the Newton-Schulz orthogonalisation on the tcgen05 GEMMs.

`zeropower_via_newtonschulz5` (muon.py:11-38) is five iterations of
    A = X X^T ;  B = b A + c A A ;  X <- a X + B X            (X bf16 [m, n], m <= n, ||X||_F <= 1)
i.e. 15 bf16 GEMMs per weight matrix per optimizer step (the 5B config trains every 2-D block weight with it:
configs/dit_v4_5B.yml opt: Muon).  Here every product is one launch of the library GEMM with the affine part folded in:
    A  = X X^T                   NT GEMM, store epilogue
    B  = b A + c (A A^T)         NT GEMM (A is symmetric), gate+residual epilogue: residual = b A, gate row = c
    X' = (B + a I) X             NN GEMM, store epilogue (a X + B X = (a I + B) X: `a` goes on B's diagonal, an [m, m] op)
so nothing of size [m, n] is touched outside the GEMMs.  The optimizer around it (momentum, Nesterov, the all-gather of the
per-rank updates over NCCL, the AdamW half of CombinedOptimizer, `init_muon`) keeps the reference's interface and update
rule (muon.py:40-176); parameters are dealt round-robin to ranks per size group as there."""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.optim import AdamW
from torch.optim.optimizer import Optimizer

from . import ops
from ._lib import OWL_EPI_GATE_RESIDUAL, OWL_EPI_STORE, OWL_GEMM_NN, OWL_GEMM_NT

NS_COEFFS = (3.4445, -4.7750, 2.0315)
bf16 = torch.bfloat16


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def zeropower_via_newtonschulz5(G: torch.Tensor, steps: int) -> torch.Tensor:
    """muon.py:11-38 for one matrix (`G.ndim == 2`; Muon.step flattens conv-shaped parameters first).  Returns bf16."""
    if G.ndim != 2:
        raise ValueError("zeropower_via_newtonschulz5: 2-D gradient expected (Muon.step flattens 4-D ones)")
    if not G.is_cuda:
        raise RuntimeError("owl_audio_exps_b200: CUDA tensors required (sm_100a kernels only; there is no CPU path)")
    a, b, c = NS_COEFFS
    X = G.to(bf16)
    transposed = G.size(-2) > G.size(-1)
    if transposed:
        X = X.mT
    X = X / (X.norm(dim=(-2, -1), keepdim=True) + 1e-7)
    m, n = X.shape
    mp, np_ = _pad8(m), _pad8(n)           # GEMM leading dimensions are multiples of 8: zero rows / columns change nothing
    Xp = torch.zeros(mp, np_, device=X.device, dtype=bf16)
    Xp[:m, :n] = X
    A = torch.empty(mp, mp, device=X.device, dtype=bf16)
    B = torch.empty(mp, mp, device=X.device, dtype=bf16)
    Xn = torch.empty_like(Xp)
    c_row = torch.full((1, mp), c, device=X.device, dtype=bf16)
    eye_a = torch.zeros(mp, mp, device=X.device, dtype=bf16)
    eye_a.diagonal().fill_(a)
    for _ in range(steps):
        ops.gemm(OWL_GEMM_NT, Xp, Xp, mp, mp, np_, mode=OWL_EPI_STORE, out0=A)
        bA = A * b
        ops.gemm(OWL_GEMM_NT, A, A, mp, mp, mp, mode=OWL_EPI_GATE_RESIDUAL, out0=B, aux0=bA, aux1=c_row, rows_per_frame=mp)
        B.add_(eye_a)
        ops.gemm(OWL_GEMM_NN, B, Xp, mp, np_, mp, mode=OWL_EPI_STORE, out0=Xn)
        Xp, Xn = Xn, Xp
    X = Xp[:m, :n]
    return X.mT if transposed else X


class Muon(Optimizer):
    """muon.py:40-113.  `rank` / `world_size` are required as in the reference (pass 0 / 1 on a single GPU)."""

    def __init__(self, params, lr=0.02, weight_decay=0.01, momentum=0.95, nesterov=True, ns_steps=5, rank=None,
                 world_size=None):
        if rank is None or world_size is None:
            raise Exception("world_size and rank params required, if you want to use this optimizer on a single GPU, "
                            "pass rank=0 and world_size=1.")
        self.rank, self.world_size = rank, world_size
        defaults = dict(lr=lr, weight_decay=weight_decay, momentum=momentum, nesterov=nesterov, ns_steps=ns_steps)
        params = list(params)
        groups = []
        for size in sorted({p.numel() for p in params}):
            same = [p for p in params if p.numel() == size]
            buf = torch.empty(world_size, size, dtype=bf16, device=same[0].device)
            groups.append(dict(params=same, update_buffer=buf, update_buffer_views=[buf[i] for i in range(world_size)]))
        super().__init__(groups, defaults)

    def _direction(self, p, group):
        """Momentum (+ Nesterov) then the orthogonalised update of one parameter, flattened bf16."""
        g = p.grad
        state = self.state[p]
        if "momentum_buffer" not in state:
            state["momentum_buffer"] = torch.zeros_like(g)
        buf = state["momentum_buffer"]
        buf.lerp_(g, 1 - group["momentum"])
        g = g.lerp_(buf, group["momentum"]) if group["nesterov"] else buf
        if g.ndim == 4:
            g = g.view(len(g), -1)
        return zeropower_via_newtonschulz5(g, steps=group["ns_steps"])

    @staticmethod
    def _apply(p, update, group):
        p.mul_(1 - group["lr"] * group["weight_decay"])
        p.add_(update.view_as(p), alpha=-group["lr"] * max(1, p.size(-2) / p.size(-1)) ** 0.5)

    @torch.no_grad()
    def step(self):
        for group in self.param_groups:
            params = group["params"]
            if self.world_size == 1:
                for p in params:
                    if p.grad is not None:
                        self._apply(p, self._direction(p, group), group)
                continue
            # rank r orthogonalises params[base + r]; the updates of one round travel in ONE all-gather, applied while
            # the next round's Newton-Schulz runs (muon.py:88-113)
            gathered, views = group["update_buffer"], group["update_buffer_views"]
            pending = None        # (handle, parameters of the round in flight)
            for base in range(0, len(params), self.world_size):
                mine = base + self.rank
                if mine < len(params):
                    assert params[mine].grad is not None
                    send = self._direction(params[mine], group).reshape(-1)
                else:
                    send = views[self.rank]
                if pending is not None:
                    pending[0].wait()
                    for p_w, u in zip(pending[1], views):
                        self._apply(p_w, u, group)
                handle = dist.all_gather_into_tensor(gathered, send.contiguous(), async_op=True)
                pending = (handle, params[base:base + self.world_size])
            if pending is not None:
                pending[0].wait()
                for p_w, u in zip(pending[1], views):
                    self._apply(p_w, u, group)


class CombinedOptimizer(Optimizer):
    """muon.py:115-173: AdamW for `adamw_keys` / 1-D parameters, Muon for every other matrix."""

    def __init__(self, model, rank=0, world_size=1, **kwargs):
        self.defaults = {}
        adamw_keys = kwargs.pop("adamw_keys", [])
        named = {n.replace("._orig_mod", ""): p for n, p in model.named_parameters()}
        for key in adamw_keys:
            assert any(key in n for n in named), f"AdamW key '{key}' not found in model parameters" + str(list(named))
        is_adamw = {n: (any(k in n for k in adamw_keys) or p.ndim < 2) for n, p in named.items()}
        self.adamw = AdamW([p for n, p in named.items() if is_adamw[n]], lr=kwargs.get("adamw_lr"),
                           betas=kwargs.get("adamw_betas", (0.9, 0.999)), weight_decay=kwargs.get("adamw_wd", 0.01),
                           eps=kwargs.get("adamw_eps", 1.0e-15))
        self.muon = Muon([p for n, p in named.items() if not is_adamw[n]], lr=kwargs.get("lr"),
                         momentum=kwargs.get("momentum"), rank=rank, world_size=world_size)
        self.param_groups = self.adamw.param_groups + self.muon.param_groups   # LR schedulers walk this

    def zero_grad(self, set_to_none: bool = False):
        self.adamw.zero_grad(set_to_none)
        self.muon.zero_grad(set_to_none)

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.adamw.step()
        self.muon.step()
        return loss

    def state_dict(self):
        return {"adamw": self.adamw.state_dict(), "muon": self.muon.state_dict()}

    def load_state_dict(self, state_dict):
        self.adamw.load_state_dict(state_dict["adamw"])
        self.muon.load_state_dict(state_dict["muon"])


def init_muon(model, rank=0, world_size=1, **kwargs):
    return CombinedOptimizer(model, rank, world_size, **kwargs)
