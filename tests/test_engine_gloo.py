"""Host logic of the data-parallel engine on CPU: bucket layout, sinks, step protocol, and a
world_size-2 gloo all-reduce of the flat buckets (no kernels are launched: gradients are written by hand)."""
import os
import types

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from owl_audio_exps_b200 import engine
import owl_audio_exps_b200.nn as N

CFG = dict(d_model=64, n_heads=1, n_layers=3, tokens_per_frame=1, n_frames=32, sample_size=8, causal=True,
           has_audio=True, rope_impl="audio1d", local_window=4, global_window=None, backbone="dit")


class Wrapped(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.transformer = N.DiT(types.SimpleNamespace(**CFG))
        self.proj_in = torch.nn.Linear(8, 64, bias=False)


def test_bucket_layout_and_sinks():
    m = Wrapped()
    b = engine.GradBuckets(m)
    nl = CFG["n_layers"]
    # one bucket per block (its 8 non-conditioning parameters), one for the whole stack's conditioning
    # projections (functional.CondBank layout), one for everything else
    assert len(b.flats) == nl + 2 and b.rest_index == nl + 1
    d = CFG["d_model"]
    assert sum(p.numel() for p in m.transformer.blocks[0].parameters()) == 18 * d * d + 15 * d
    cond_flat = b.flats[nl]
    assert cond_flat.numel() == nl * (6 * d * d + 6 * d)
    sink = m.transformer._stage.cond_sink
    assert sink.gw.shape == (nl * 6 * d, d) and sink.gb.shape == (nl * 6 * d,)
    assert sink.gw.data_ptr() == cond_flat.data_ptr()
    for i, blk in enumerate(m.transformer.blocks):
        cond_ids = {id(p) for p in blk.cond_params()}
        for p in blk.parameters():
            s = p._owl_grad_sink
            assert s.view.data_ptr() == p.grad.data_ptr() and p.grad.dtype == torch.float32
            assert s.view.data_ptr() % 16 == 0           # TMA / float4 alignment of every sink
            flat = cond_flat if id(p) in cond_ids else b.flats[i]
            lo, hi = flat.data_ptr(), flat.data_ptr() + flat.numel() * 4
            assert lo <= s.view.data_ptr() < hi
        assert blk._stage.after_backward is not None
        # rows of the stacked conditioning gradient are in CondBank order: a1|b1, c1, a2|b2, c2 per layer
        wa1, _, wg1, _, wa2, _, wg2, _ = blk.cond_params()
        base = sink.gw.data_ptr() + i * 6 * d * d * 4
        assert wa1.grad.data_ptr() == base and wg1.grad.data_ptr() == base + 2 * d * d * 4
        assert wa2.grad.data_ptr() == base + 3 * d * d * 4 and wg2.grad.data_ptr() == base + 5 * d * d * 4
    assert not hasattr(m.proj_in.weight, "_owl_grad_sink") and m.proj_in.weight.grad is not None
    assert b.accumulate is False
    b.begin_step()
    assert b.accumulate is False
    b.next_micro_batch()
    assert m.transformer.blocks[0].attn.qkv.weight._owl_grad_sink.accumulate is True
    b.detach()
    assert not hasattr(m.transformer.blocks[0].attn.qkv.weight, "_owl_grad_sink")


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    m = Wrapped()
    b = engine.GradBuckets(m)
    b.begin_step()
    for p in m.parameters():                      # stand-in for the wgrad kernels: rank-dependent gradients
        p.grad.fill_(float(rank + 1))
    d6 = 6 * CFG["d_model"]
    for i in reversed(range(CFG["n_layers"])):    # blocks finish backward in reverse order
        m.transformer.blocks[i]._stage.after_backward()
        if i == 2:                                # CondBank: the tail group of layers is reduced early ...
            m.transformer._stage.cond_sink.after_group(2 * d6, 3 * d6)
    b.finish_step()                               # ... the remaining rows and the biases here
    ok = all(torch.allclose(p.grad, torch.full_like(p.grad, 1.5)) for p in m.parameters())
    # no_sync equivalent: nothing is exchanged
    b.sync_grads = False
    b.begin_step()
    for p in m.parameters():
        p.grad.fill_(float(rank + 1))
    for i in reversed(range(CFG["n_layers"])):
        m.transformer.blocks[i]._stage.after_backward()
    b.finish_step()
    ok = ok and all(torch.allclose(p.grad, torch.full_like(p.grad, float(rank + 1))) for p in m.parameters())
    out[rank] = ok
    dist.destroy_process_group()


def test_gloo_world2_bucket_allreduce():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, 29533, out), nprocs=2, join=True)
    assert out[0] and out[1]


def test_zero_grad_set_to_none_is_reattached_and_sinks_overwrite_then_add():
    """the reference trainers call opt.zero_grad(set_to_none=True) every step (rft_trainer.py:200); the
    bucket views must be p.grad again at the next begin_step, and a sink's first write of a step overwrites while any
    later write in the same step (second micro-batch or second backward through the block) adds."""
    m = Wrapped()
    b = engine.GradBuckets(m)
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    b.begin_step()
    views = {id(p): p.grad for p in m.parameters()}
    opt.zero_grad(set_to_none=True)
    assert all(p.grad is None for p in m.parameters())
    b.begin_step()
    assert all(p.grad is views[id(p)] for p in m.parameters())
    w = m.transformer.blocks[0].attn.qkv.weight._owl_grad_sink
    assert w.accumulate is False        # first write of the step
    assert w.accumulate is True         # the same sink written again in the same step
    b.begin_step()
    assert w.accumulate is False
    bias = m.transformer.blocks[0].attn.qkv.bias._owl_grad_sink
    assert bias.accumulate is True      # column sums always add into buffers zeroed by begin_step
    cs = m.transformer._stage.cond_sink
    assert cs.acc_rows(0) is False and cs.acc_rows(0) is True and cs.acc_rows(384) is False
    b.detach()


def test_backwards_per_step_defers_the_reduce_and_rejects_extra_passes():
    m = Wrapped()
    b = engine.GradBuckets(m, backwards_per_step=2)
    sent = []
    b._reduce = lambda flat: sent.append(flat.data_ptr())
    b.begin_step()
    blk = m.transformer.blocks[1]
    blk._stage.after_backward()
    assert sent == []                                   # first of two expected backward passes: nothing handed over
    blk._stage.after_backward()
    assert sent == [b.flats[1].data_ptr()]
    try:
        blk._stage.after_backward()
        raise AssertionError("third backward pass must be rejected")
    except RuntimeError as e:
        assert "backwards_per_step" in str(e)
    b.begin_step()                                      # counters reset per step
    blk._stage.after_backward()
    assert len(sent) == 1
    b.detach()
