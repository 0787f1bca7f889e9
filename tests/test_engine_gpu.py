"""GPU tests of the data-parallel engine: gradients written straight into the flat buckets (block buckets
+ the stack's conditioning bucket) == plain autograd gradients, eagerly and through the CUDA graph."""
import pytest
import torch

from helpers import cfg_of, load_golden, relerr

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(fx):
    import owl_audio_exps_b200.nn as N
    cfg = cfg_of(fx)
    model = N.DiT(cfg)
    model.load_state_dict({k: v.float() for k, v in fx["state"].items()})
    model = model.to(DEV)
    for blk, (cos, sin) in zip(model.blocks, fx["rope"]):
        blk.attn.rope.cos, blk.attn.rope.sin = cos.to(DEV).float().contiguous(), sin.to(DEV).float().contiguous()
    return model


def _loss(model, x, cond, dy):
    x = x.detach().requires_grad_(True)
    cond = cond.detach().requires_grad_(True)
    return (model(x, cond).float() * dy.float()).sum()


@pytest.mark.parametrize("use_graph", [False, True])
def test_bucket_grads_match_autograd(use_graph):
    from owl_audio_exps_b200 import engine
    fx = load_golden("av_T5_ortho")
    model = _model(fx)
    inputs = [fx["x"].to(DEV), fx["cond"].to(DEV), fx["dy"].to(DEV)]
    _loss(model, *inputs).backward()
    ref = {k: p.grad.clone() for k, p in model.named_parameters()}
    model.zero_grad()
    buckets = engine.GradBuckets(model)
    assert len(buckets.cond_stacks) == 1
    step = engine.TrainStep(model, _loss, inputs, buckets=buckets, use_graph=use_graph, warmup=2)
    for _ in range(2):   # sinks are overwritten on the first micro-batch of every step, not accumulated
        step(*inputs)
    torch.cuda.synchronize()
    for k, p in model.named_parameters():
        assert relerr(p.grad, ref[k]) <= 1e-3, k
    # gradient accumulation (DDP no_sync equivalent): two micro-batches add up
    if not use_graph:
        buckets.begin_step()
        _loss(model, *inputs).backward()
        buckets.next_micro_batch()
        _loss(model, *inputs).backward()
        buckets.finish_step()
        torch.cuda.synchronize()
        for k, p in model.named_parameters():
            assert relerr(p.grad, 2 * ref[k]) <= 1e-3, k
    buckets.detach()


def test_prefetch_matches_direct_call():
    """TrainStep.prefetch (H2D of the next batch on a copy stream into staging buffers) + TrainStep() == TrainStep(*batch)."""
    from owl_audio_exps_b200 import engine
    fx = load_golden("audio_T1")
    model = _model(fx)
    dev_in = [fx["x"].to(DEV), fx["cond"].to(DEV), fx["dy"].to(DEV)]
    host_a = [t.cpu().pin_memory() for t in dev_in]
    host_b = [(t * 0.5).cpu().pin_memory() for t in dev_in]
    buckets = engine.GradBuckets(model)
    step = engine.TrainStep(model, _loss, dev_in, buckets=buckets, use_graph=True, warmup=2)
    ref = {}
    for name, host in (("a", host_a), ("b", host_b)):
        loss = step(*host).clone()
        torch.cuda.synchronize()
        ref[name] = (loss, {k: p.grad.clone() for k, p in model.named_parameters()})
    step.prefetch(*host_a)
    la = step().clone()
    step.prefetch(*host_b)            # overlaps the step above
    ga = {k: p.grad.clone() for k, p in model.named_parameters()}
    lb = step().clone()
    torch.cuda.synchronize()
    gb = {k: p.grad.clone() for k, p in model.named_parameters()}
    assert torch.equal(la, ref["a"][0]) and torch.equal(lb, ref["b"][0])
    for k in ga:
        assert relerr(ga[k], ref["a"][1][k]) <= 1e-3, k
        assert relerr(gb[k], ref["b"][1][k]) <= 1e-3, k
    buckets.detach()


def test_cast_multi_matches_single():
    from owl_audio_exps_b200 import ops
    g = torch.Generator(device=DEV).manual_seed(0)
    srcs = [torch.randn(n, device=DEV, generator=g) for n in (1024 * 1024, 8, 3072 * 1024, 64, 1024, 4096 * 1024, 512, 2048,
                                                              1024 * 16, 40)]
    dsts = [torch.empty(s.numel(), device=DEV, dtype=torch.bfloat16) for s in srcs]
    ops.cast_f32_bf16_multi(list(zip(srcs, dsts)))
    for s, d in zip(srcs, dsts):
        assert torch.equal(d, s.to(torch.bfloat16))


@pytest.mark.parametrize("use_graph", [False, True])
def test_full_model_step_buckets_match_autograd(use_graph):
    """The engine around the whole AudioRFTCore (embeddings, proj_in, stack, FinalLayer): stack gradients land in the block /
    conditioning buckets, everything else in the `rest` bucket; graph-captured step == plain autograd."""
    from owl_audio_exps_b200 import engine
    from owl_audio_exps_b200.models import AudioRFTCore
    fx = load_golden("audiorft_core")
    model = AudioRFTCore(cfg_of(fx))
    model.load_state_dict({k: v.float() for k, v in fx["state"].items()})
    model = model.to(DEV)
    for blk, (cos, sin) in zip(model.transformer.blocks, fx["rope"]):
        blk.attn.rope.cos, blk.attn.rope.sin = cos.to(DEV).float().contiguous(), sin.to(DEV).float().contiguous()

    def loss_fn(m, x, t, dy):
        return (m(x, t).float() * dy.float()).sum()
    inputs = [fx["x"].to(DEV).float(), fx["t"].to(DEV), fx["dy"].to(DEV)]
    loss_fn(model, *inputs).backward()
    ref = {k: p.grad.clone() for k, p in model.named_parameters()}
    model.zero_grad()
    buckets = engine.GradBuckets(model)
    step = engine.TrainStep(model, loss_fn, inputs, buckets=buckets, use_graph=use_graph, warmup=2)
    for _ in range(2):
        step(*inputs)
    torch.cuda.synchronize()
    for k, p in model.named_parameters():
        assert relerr(p.grad, ref[k]) <= 1e-3, k
        assert relerr(p.grad, fx["grads"][k]) <= 2e-2, k     # and the reference model's own gradients
    buckets.detach()


def test_multi_step_optimizer_with_zero_grad_set_to_none():
    """optimizer.step() + zero_grad(set_to_none=True) with buckets, several steps: every parameter keeps
    receiving its gradient (p.grad is re-attached to the bucket view) and the trajectory matches plain autograd."""
    import copy
    from owl_audio_exps_b200 import engine
    fx = load_golden("audio_T1")
    model = _model(fx)
    ref_model = copy.deepcopy(model)
    inputs = [fx["x"].to(DEV), fx["cond"].to(DEV), fx["dy"].to(DEV)]
    opt_ref = torch.optim.SGD(ref_model.parameters(), lr=1e-2)
    for _ in range(3):
        opt_ref.zero_grad(set_to_none=True)
        _loss(ref_model, *inputs).backward()
        opt_ref.step()
    buckets = engine.GradBuckets(model)
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    p0 = {k: p.detach().clone() for k, p in model.named_parameters()}
    for _ in range(3):
        buckets.begin_step()
        _loss(model, *inputs).backward()
        buckets.finish_step()
        assert all(p.grad is not None for p in model.parameters())
        opt.step()
        opt.zero_grad(set_to_none=True)      # what the reference trainers do (rft_trainer.py:200)
    torch.cuda.synchronize()
    ref = dict(ref_model.named_parameters())
    for k, p in model.named_parameters():
        assert not torch.equal(p, p0[k]), k                     # nothing was silently skipped by the optimizer
        assert relerr(p - p0[k], ref[k] - p0[k]) <= 2e-2, k     # same update as plain autograd
    buckets.detach()


def test_two_model_calls_in_one_loss_accumulate_into_the_buckets():
    """a loss that runs the stack twice (self-forcing / distillation trainers) with backwards_per_step=2:
    the second backward through a block adds to the first instead of overwriting it."""
    from owl_audio_exps_b200 import engine
    fx = load_golden("av_T5_ortho")
    model = _model(fx)
    inputs = [fx["x"].to(DEV), fx["cond"].to(DEV), fx["dy"].to(DEV)]

    def loss2(m, x, cond, dy):
        return _loss(m, x, cond, dy) + 0.5 * _loss(m, x * 0.5, cond, dy)
    loss2(model, *inputs).backward()
    ref = {k: p.grad.clone() for k, p in model.named_parameters()}
    model.zero_grad()
    buckets = engine.GradBuckets(model, backwards_per_step=2)
    for _ in range(2):
        buckets.begin_step()
        loss2(model, *inputs).backward()
        buckets.finish_step()
    torch.cuda.synchronize()
    for k, p in model.named_parameters():
        assert relerr(p.grad, ref[k]) <= 2e-3, k
    buckets.detach()
    buckets = engine.GradBuckets(model)   # default: one backward per block per micro-batch -> the second is refused
    buckets.begin_step()
    with pytest.raises(RuntimeError, match="backwards_per_step"):
        loss2(model, *inputs).backward()
    buckets.detach()


def test_weight_stage_sees_data_mutations():
    """`p.data.<op>_()` (ema_pytorch's update path) does not bump `_version`; outside frozen_weights() every
    forward re-stages, inside it operands are cached until restage() / the next outermost frozen_weights()."""
    from owl_audio_exps_b200 import functional as OF
    fx = load_golden("audio_T1")
    model = _model(fx)
    x, cond = fx["x"].to(DEV), fx["cond"].to(DEV)
    with torch.no_grad():
        y0 = model(x, cond).clone()
        w = model.blocks[0].mlp.fc2.weight
        v0 = w._version
        w.data.mul_(0.0)                       # EMA-style in-place update through .data
        assert w._version == v0
        y1 = model(x, cond).clone()
        assert not torch.equal(y0, y1)         # picked up without any version bump
        with OF.frozen_weights():
            ya = model(x, cond).clone()
            assert torch.equal(ya, y1)
            model.blocks[0].attn.out.weight.data.mul_(0.0)
            yb = model(x, cond).clone()
            assert torch.equal(yb, ya)         # declared frozen: cached operands are reused ...
            OF.restage()
            yc = model(x, cond).clone()
            assert not torch.equal(yc, ya)     # ... until restage()
        with OF.frozen_weights():              # entering the outermost context re-stages as well
            model.blocks[1].mlp.fc2.weight.data.mul_(0.0)
        with OF.frozen_weights():
            yd = model(x, cond).clone()
        assert not torch.equal(yd, yc)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_copy_engine_exchange_matches_nccl_two_ranks():
    """engine.PeerExchange (copy-engine peer copies over symmetric memory + owl_shard_reduce) == ncclAllReduce on the same
    model and per-rank batches, eager and captured in the step's graph, bitwise identical on every rank
    (tools/exchange_check.py under torchrun, 2 ranks)."""
    import os
    import socket
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port),
                          os.path.join(root, "tools", "exchange_check.py")], capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0 and "OK" in out.stdout, (out.stdout[-1500:], out.stderr[-1500:])
