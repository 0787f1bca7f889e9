"""Probe of torch symmetric memory on this box (N ranks, one node): peer copies through the copy engines, device-side
barriers, multimem all-reduce, inside and outside a CUDA graph, next to ncclAllReduce; and how much a concurrent GEMM
slows down under each.  torchrun --nproc-per-node N tools/symm_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
P = lambda *a: print(*a, flush=True) if rank == 0 else None

n = 75 * 1024 * 1024 // 4          # one block bucket: 75 MB of fp32
buf = symm_mem.empty(n, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
P("multicast support:", hdl.has_multicast_support, "world", hdl.world_size)
buf.fill_(rank + 1.0)
stage = torch.empty(n, device=dev)
peer = hdl.get_buffer((rank + 1) % world, (n,), torch.float32)
hdl.barrier()
torch.cuda.synchronize()


def timeit(fn, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


t = timeit(lambda: stage.copy_(peer))
P(f"peer pull copy_ 75 MB: {t:.3f} ms = {n * 4 / t / 1e6:.0f} GB/s;  value {stage[0].item()} (expect {(rank + 1) % world + 1})")
nc = torch.empty(n, device=dev).fill_(1.0)
t = timeit(lambda: dist.all_reduce(nc))
P(f"nccl all_reduce 75 MB fp32: {t:.3f} ms = busbw {2 * (world - 1) / world * n * 4 / t / 1e6:.0f} GB/s")
for name in ("two_shot_all_reduce_", "multimem_all_reduce_"):
    try:
        op = getattr(torch.ops.symm_mem, name)
        t = timeit(lambda: op(buf, "sum", dist.group.WORLD.group_name))
        P(f"symm_mem.{name} 75 MB fp32: {t:.3f} ms = busbw {2 * (world - 1) / world * n * 4 / t / 1e6:.0f} GB/s")
    except Exception as e:  # noqa: BLE001
        P(f"symm_mem.{name}: {type(e).__name__}: {str(e)[:200]}")
# a GEMM alone, then with each transfer running on a side stream
a = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
b = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
t0 = timeit(lambda: a @ b)
P(f"gemm 8192^3 alone: {t0:.3f} ms")
side = torch.cuda.Stream()


def with_side(fn):
    def run():
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(4):
                fn()
        for _ in range(6):
            a @ b
        torch.cuda.current_stream().wait_stream(side)
    return timeit(run, 5)


P(f"6 gemms + 4 peer copies on a side stream: {with_side(lambda: stage.copy_(peer)):.3f} ms (6 gemms alone {6 * t0:.3f})")
P(f"6 gemms + 4 nccl all_reduce on a side stream: {with_side(lambda: dist.all_reduce(nc)):.3f} ms")
# graph capture of barrier + peer copy
try:
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        hdl.barrier()
        stage.copy_(peer)
        hdl.barrier()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        hdl.barrier()
        stage.copy_(peer)
        hdl.barrier()
    buf.fill_(10.0 * (rank + 1))
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    P("graph replay of barrier + peer copy + barrier ok; value", stage[0].item(), "expect", 10.0 * ((rank + 1) % world + 1))
    t = timeit(lambda: g.replay())
    P(f"graph (barrier, 75 MB pull, barrier): {t:.3f} ms")
except Exception as e:  # noqa: BLE001
    P(f"graph capture: {type(e).__name__}: {str(e)[:300]}")
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
