"""Wrappers the reference puts around the DiT modules and how this package survives them.

`torch.compile(model)` / `torch.compile(DDP(model))` / `torch.compile(core, fullgraph=True)` (rft_trainer.py:96-99,
causvid_pipeline.py:63, gamerft_audio.py:36): the block here already IS hand-written kernels sequenced by one autograd
node per block, launched through ctypes -- there is nothing for a tracing compiler to fuse and Dynamo cannot trace a
ctypes call.  So for modules that contain this package's DiT modules `compile` returns a thin wrapper with the same
observable surface as torch's OptimizedModule (`._orig_mod`, attribute pass-through, `_orig_mod.`-prefixed state-dict
keys, which the reference's checkpoints rely on: `transformer._orig_mod.blocks...`) and runs the module as is; anything
else (plain functions, modules without owl blocks) goes to the real `torch.compile`.  `install()` points `torch.compile`
at this function so the reference's unedited trainers / pipelines keep their `torch.compile(...)` lines.

Launch-overhead removal, which is what `mode="max-autotune"` buys the reference at decode shapes, is done by CUDA graphs
here (engine.TrainStep, the samplers' decode graphs)."""
from __future__ import annotations

import torch
from torch import nn

_REAL_COMPILE = torch.compile


def _has_owl_modules(mod: nn.Module) -> bool:
    return any(type(m).__module__.startswith("owl_audio_exps_b200.") for m in mod.modules())


class CompiledModule(nn.Module):
    """Stand-in for torch._dynamo.eval_frame.OptimizedModule around modules of this package."""

    def __init__(self, mod: nn.Module, **compile_kwargs):
        super().__init__()
        self._orig_mod = mod
        self._compile_kwargs = compile_kwargs

    def forward(self, *args, **kwargs):
        return self._orig_mod(*args, **kwargs)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self._orig_mod, name)

    def __call__(self, *args, **kwargs):   # skip nn.Module's hook machinery twice per call at launch-bound shapes
        return self._orig_mod(*args, **kwargs)


def compile(model=None, **kwargs):  # noqa: A001  (drop-in for torch.compile)
    """torch.compile with the same call forms (direct and decorator)."""
    if model is None:
        return lambda m: compile(m, **kwargs)
    if isinstance(model, nn.Module) and _has_owl_modules(model):
        return model if isinstance(model, CompiledModule) else CompiledModule(model, **kwargs)
    return _REAL_COMPILE(model, **kwargs)


def patch_torch_compile() -> None:
    """Route `torch.compile` through `compile` above (idempotent)."""
    if torch.compile is not compile:
        torch.compile = compile


def unpatch_torch_compile() -> None:
    torch.compile = _REAL_COMPILE
