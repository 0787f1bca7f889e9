"""Copy-engine gradient exchange (engine.PeerExchange) against ncclAllReduce on the same model and per-rank batches:
eager and captured in the step's CUDA graph.  torchrun --nproc-per-node N tools/exchange_check.py"""
import os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from owl_audio_exps_b200 import engine
import owl_audio_exps_b200.nn as N

cfg = types.SimpleNamespace(d_model=256, n_heads=4, n_layers=5, tokens_per_frame=1, n_frames=64, sample_size=24,
                            causal=True, has_audio=True, rope_impl="audio1d", local_window=4, global_window=None,
                            backbone="dit")
torch.manual_seed(0)
model = N.DiT(cfg).to(dev)
g = torch.Generator().manual_seed(100 + rank)
x = torch.randn(4, 24, 256, generator=g).to(torch.bfloat16).to(dev)
cond = torch.randn(4, 24, 256, generator=g).to(torch.bfloat16).to(dev)
tgt = torch.randn(4, 24, 256, generator=g).to(torch.bfloat16).to(dev)


def loss_fn(m, x, c, t):
    x = x.detach().requires_grad_(True)
    c = c.detach().requires_grad_(True)
    return torch.nn.functional.mse_loss(m(x, c).float(), t.float())


def grads(mode, graph):
    os.environ["OWL_EXCHANGE"] = mode
    b = engine.GradBuckets(model)
    assert (b.exchange is not None) == (mode != "nccl"), mode
    if b.exchange is not None:
        assert b.exchange.mode == mode, (b.exchange.mode, mode)
    step = engine.TrainStep(model, loss_fn, [x, cond, tgt], buckets=b, use_graph=graph, warmup=2)
    for _ in range(2):
        step(x, cond, tgt)
    torch.cuda.synchronize()
    out = torch.cat([f.detach().clone().flatten() for f in b.flats])
    b.detach()
    return out


ref = grads("nccl", False)
ok = True
import torch.distributed._symmetric_memory as _sm
_probe = _sm.empty(1024, dtype=torch.float32, device=dev)
has_mc = bool(int(_sm.rendezvous(_probe, dist.group.WORLD).multicast_ptr))
if rank == 0:
    print("multicast mapping:", has_mc, flush=True)
for mode, graph in [(m, g_) for m in (["ce", "mc"] if has_mc else ["ce"]) for g_ in (False, True)]:
    got = grads(mode, graph)
    err = ((got - ref).abs().max() / ref.abs().max()).item()
    same_across = got.clone()
    dist.broadcast(same_across, 0)
    bit_equal = bool(torch.equal(same_across, got))
    if rank == 0:
        print(f"{mode} exchange ({'graph' if graph else 'eager'}) vs nccl: max rel err {err:.2e}; identical on every rank: {bit_equal}",
              flush=True)
    ok = ok and err <= 1e-5 and bit_equal
dist.barrier()
torch.cuda.synchronize()
if rank == 0:
    print("OK" if ok else "MISMATCH", flush=True)
os._exit(0 if ok else 1)
