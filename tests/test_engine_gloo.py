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
    assert len(b.flats) == CFG["n_layers"] + 1 and b.rest_index == CFG["n_layers"]
    d = CFG["d_model"]
    assert sum(p.numel() for p in m.transformer.blocks[0].parameters()) == 18 * d * d + 15 * d
    for i, blk in enumerate(m.transformer.blocks):
        for p in blk.parameters():
            s = p._owl_grad_sink
            assert s.view.data_ptr() == p.grad.data_ptr() and p.grad.dtype == torch.float32
            assert s.view.data_ptr() % 16 == 0           # TMA / float4 alignment of every sink
            lo, hi = b.flats[i].data_ptr(), b.flats[i].data_ptr() + b.flats[i].numel() * 4
            assert lo <= s.view.data_ptr() < hi
        assert blk._stage.after_backward is not None
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
    for i in reversed(range(CFG["n_layers"])):    # blocks finish backward in reverse order
        m.transformer.blocks[i]._stage.after_backward()
    b.finish_step()
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
