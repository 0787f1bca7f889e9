"""The wrappers the reference's trainers and pipelines put around the DiT modules: torch.compile (through
compat.compile, what install() points torch.compile at), DistributedDataParallel, EMA deep copies updated through
`.data`, GradScaler + clip_grad_norm_, and a second device in the same process.  Each must give the bare module's
outputs and gradients (rft_trainer.py:96-99,105,186-204; causvid_pipeline.py:36,63)."""
import copy
import os
import socket

import pytest
import torch

from helpers import cfg_of, load_golden, relerr

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(dtype=torch.float32, dev=DEV):
    import owl_audio_exps_b200.nn as N
    fx = load_golden("av_T5_ortho")
    cfg = cfg_of(fx)
    model = N.DiT(cfg)
    model.load_state_dict({k: v.float() for k, v in fx["state"].items()})
    model = model.to(dev).to(dtype)
    for blk, (cos, sin) in zip(model.blocks, fx["rope"]):
        blk.attn.rope.cos, blk.attn.rope.sin = cos.to(dev).float().contiguous(), sin.to(dev).float().contiguous()
    return model, fx


def _step(callable_model, fx, params, dev=DEV, scale=1.0):
    for p in params:
        p.grad = None
    x = fx["x"].to(dev).requires_grad_(True)
    cond = fx["cond"].to(dev).requires_grad_(True)
    y = callable_model(x, cond)
    (y.float() * fx["dy"].to(dev).float()).sum().mul(scale).backward()
    return y.detach(), x.grad.detach(), [p.grad.detach().clone() for p in params]


def _same(a, b, tol=1e-5):
    ya, dxa, ga = a
    yb, dxb, gb = b
    assert torch.equal(ya, yb) and torch.equal(dxa, dxb)
    for u, v in zip(ga, gb):
        assert relerr(u, v) <= tol      # bias column sums meet in fp32 atomics: order is not fixed


def test_torch_compile_wrapper_same_outputs_and_grads():
    from owl_audio_exps_b200 import compat
    model, fx = _model()
    params = list(model.parameters())
    bare = _step(model, fx, params)
    for kw in (dict(), dict(mode="max-autotune", dynamic=False, fullgraph=True)):
        wrapped = compat.compile(model, **kw)
        assert wrapped._orig_mod is model and wrapped.blocks is model.blocks          # OptimizedModule's surface
        assert all(k.startswith("_orig_mod.") for k in wrapped.state_dict())
        _same(_step(wrapped, fx, params), bare)
    # install() routes torch.compile itself (the reference's unedited `torch.compile(model)` lines)
    compat.patch_torch_compile()
    try:
        wrapped = torch.compile(model, mode="max-autotune", fullgraph=True)
        assert isinstance(wrapped, compat.CompiledModule)
        _same(_step(wrapped, fx, params), bare)
        plain = torch.nn.Linear(4, 4)
        assert not isinstance(torch.compile(plain), compat.CompiledModule)           # everything else: real torch.compile
    finally:
        compat.unpatch_torch_compile()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_distributed_data_parallel_wrapper_same_grads():
    """rft_trainer.py:96-99: torch.compile(DDP(model)).  World size 1 is enough for DDP's reducer hooks to run."""
    import torch.distributed as dist
    from owl_audio_exps_b200 import compat
    model, fx = _model()
    params = list(model.parameters())
    bare = _step(model, fx, params)
    own = not dist.is_initialized()
    if own:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(_free_port())
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device(DEV))
    try:
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[0])
        wrapped = compat.compile(ddp)
        got = _step(wrapped, fx, params)
        _same(got, bare)
        # a second step: the reducer re-arms, gradients are rebuilt (zero_grad(set_to_none=True) in between)
        _same(_step(wrapped, fx, params), bare)
    finally:
        if own:
            dist.destroy_process_group()


def test_ema_deepcopy_and_data_updates_are_seen():
    """ema_pytorch deep-copies the model and updates the copy through `.data` (rft_trainer.py:105,204), which does not
    bump Tensor._version: the staged bf16 operands must still follow."""
    model, fx = _model()
    ema = copy.deepcopy(model)
    x, cond = fx["x"].to(DEV), fx["cond"].to(DEV)
    with torch.no_grad():
        y0 = ema(x, cond).clone()
        assert torch.equal(y0, model(x, cond))
        for p_ema, p in zip(ema.parameters(), model.parameters()):
            p_ema.data.lerp_(torch.randn_like(p) * p.std().clamp_min(1e-3), 0.3)      # EMA.update style
        y1 = ema(x, cond).clone()
        fresh, _ = _model()
        fresh.load_state_dict(ema.state_dict())
        y2 = fresh(x, cond)
    assert not torch.equal(y1, y0)
    assert torch.equal(y1, y2)


def test_grad_scaler_and_clip_grad_norm():
    """GradScaler.scale(loss).backward() -> unscale_ -> clip_grad_norm_ -> step (trainers' AMP recipe)."""
    model, fx = _model()
    params = list(model.parameters())
    bare = _step(model, fx, params)
    opt = torch.optim.SGD(params, lr=1e-3)
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
    opt.zero_grad(set_to_none=True)
    x, cond = fx["x"].to(DEV).requires_grad_(True), fx["cond"].to(DEV).requires_grad_(True)
    loss = (model(x, cond).float() * fx["dy"].to(DEV).float()).sum()
    scaler.scale(loss).backward()
    scaler.unscale_(opt)
    for p, g in zip(params, bare[2]):
        assert relerr(p.grad, g) <= 2e-2      # dy * 1024 is rounded to bf16 at other points than dy
    total = torch.nn.utils.clip_grad_norm_(params, 1.0)
    assert torch.isfinite(total) and total > 0
    norm_after = torch.sqrt(sum((p.grad.float() ** 2).sum() for p in params))
    assert norm_after <= 1.0 + 1e-3
    before = [p.detach().clone() for p in params]
    scaler.step(opt)
    scaler.update()
    assert any(not torch.equal(a, p) for a, p in zip(before, params))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two devices in one process")
def test_second_device_in_the_same_process():
    """Per-device kernel configuration (>48 KB shared-memory opt-in is per device): rollout replicas on independent
    streams in one process (north_star: 'inference rollout runs per-GPU on independent streams')."""
    m0, fx = _model()
    m1, _ = _model(dev="cuda:1")
    a = _step(m0, fx, list(m0.parameters()))
    with torch.cuda.device(1):
        b = _step(m1, fx, list(m1.parameters()), dev="cuda:1")
    assert relerr(b[0], a[0]) <= 1e-6 and relerr(b[1], a[1]) <= 1e-6
