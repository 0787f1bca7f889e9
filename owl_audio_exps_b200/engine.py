"""Data-parallel training step for the DiT hot path: flat fp32 gradient buckets, all-reduce over
NCCL/NVLink overlapped with the rest of backward, and the whole step replayed as one CUDA graph.

Replaces what the reference gets from `DistributedDataParallel(model)` + `torch.compile`
(owl_wms/trainers/rft_trainer.py:96-99, audio_rft_trainer.py:93) for this path:

  * one bucket per DiT block (its 16 parameters, 18 d^2 + 15 d fp32 values) plus one for everything
    else.  `param.grad` IS a view of the bucket, and the block's weight-gradient GEMMs / bias column
    sums write straight into it (functional._wgrad sink), so there is no autograd copy and no
    bucket-packing pass;
  * a block's bucket is handed to NCCL on a side stream the moment that block's backward has
    finished (blocks finish in reverse order), so the exchange of block i overlaps the backward of
    blocks i-1..0.  NVSwitch gives every GPU full bandwidth to every peer, so bucket size is chosen for
    overlap granularity (one block), not link count;
  * launch-bound shapes (audio.yml: ~600 kernels of 5-50 us per step) are captured into a CUDA
    graph -- the C ABI only enqueues on the current stream and never syncs, so capture is legal.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

f32 = torch.float32


class GradSink:
    """What functional._sink returns: the fp32 bucket view of one parameter + whether to add to it."""

    __slots__ = ("view", "owner", "pre_zeroed", "written")

    def __init__(self, view: torch.Tensor, owner: "GradBuckets", pre_zeroed: bool = False):
        self.view, self.owner, self.pre_zeroed = view, owner, pre_zeroed
        self.written = -1    # owner.step_id of the last step this sink was written in

    @property
    def accumulate(self) -> bool:
        """Read ONCE per write by the kernel launcher (functional._wgrad / _bgrad / _fresh_or_sink): the first write of a
        step overwrites, every later one (second micro-batch, or a second backward through the same block inside one
        loss: self-forcing / distillation trainers call the model twice) adds."""
        # bias gradients are column sums built with atomics: the engine zeroes all of them with ONE multi-tensor
        # kernel at begin_step, so the kernels always add (no memset node in front of every column sum)
        again = self.written == self.owner.step_id
        self.written = self.owner.step_id
        return self.pre_zeroed or self.owner.accumulate or again


class CondSink:
    """What functional.CondBank looks for on the DiT stack's WeightStage: the fp32 gradient of ALL layers'
    conditioning projections as one [n_layers*6d, d] matrix (+ [n_layers*6d] biases) inside one flat bucket,
    so a group of layers' weight gradients is ONE GEMM written in place."""

    __slots__ = ("gw", "gb", "owner", "index", "gb_pre_zeroed", "written")

    def __init__(self, gw, gb, owner: "GradBuckets", index: int):
        self.gw, self.gb, self.owner, self.index = gw, gb, owner, index
        self.gb_pre_zeroed = True   # GradBuckets.begin_step zeroes every bias sink (incl. gb) in one kernel
        self.written = {}           # first weight row of a group -> owner.step_id of its last write

    @property
    def accumulate(self) -> bool:
        return self.owner.accumulate

    def acc_rows(self, r0: int) -> bool:
        """Asked by CondBank at the moment it writes the weight gradient of the row group starting at r0: the first
        write of a step overwrites, later ones (next micro-batch, second backward of the stack in one loss) add."""
        again = self.written.get(r0) == self.owner.step_id
        self.written[r0] = self.owner.step_id
        return self.owner.accumulate or again

    def after_group(self, r0: int, r1: int):
        self.owner._cond_rows_done(self.index, r0, r1)

    # the sink belongs to a live engine: copies of the model (EMA) and pickles do not carry it
    def __deepcopy__(self, memo):
        return None

    def __reduce__(self):
        return (type(None), ())


def _align(n: int, a: int = 64) -> int:
    return (n + a - 1) // a * a


import os


class PeerExchange:
    """Gradient all-reduce whose data movement runs on the COPY ENGINES, not on SMs.

    `ncclAllReduce` next to the backward GEMMs costs twice: its CTAs take SMs (the persistent GEMM grid has to leave
    16 of 148 free, or a GEMM CTA queues behind a communication CTA) and measured at N=2 (profiles/r02_symm_probe_n2.txt)
    four 75 MB all-reduces on a side stream stretch six 8192^3 GEMMs from 4.02 to 4.97 ms, while four 75 MB peer copies
    leave them at 4.02 ms.  So the buckets live in ONE symmetric-memory allocation (torch symmetric memory: every rank can
    address every peer's buckets over NVLink), and an all-reduce of a bucket slice is
        barrier                                   every rank's slice is complete
        pull   my shard of the slice from each peer into local staging      (N-1 copy-engine peer copies)
        reduce my shard = (mine + staged) / N                               (the only kernel: `owl_shard_reduce`, <= 32 CTAs)
        barrier                                   every rank's shard is reduced
        pull   every peer's reduced shard into my slice                     (N-1 copy-engine peer copies)
    i.e. reduce-scatter + all-gather with NVSwitch carrying each byte once per direction, fp32 end to end (bitwise the
    same on every rank).  Barriers are torch symmetric memory's device-side signal barriers; everything is enqueued on the
    communication stream and is captured into the step's CUDA graph like the NCCL calls it replaces."""

    def __init__(self, total_elems: int, device, group, average: bool):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.average = average
        self.sym = symm_mem.empty(total_elems, dtype=f32, device=device)
        self.hdl = symm_mem.rendezvous(self.sym, self.group)
        self.sym.zero_()
        self.peers = [self.sym if p == self.rank else self.hdl.get_buffer(p, (total_elems,), f32)
                      for p in range(self.world)]
        self._staging = {}          # lane -> [world - 1, shard] fp32
        self.max_ctas = 32
        # measurement aid: OWL_CE_SKIP=reduce|copy|all leaves out the reduction kernel / the peer copies / both (results are
        # then wrong; tools/n2_sweep.sh uses it to attribute the exchange's cost)
        self._skip = os.environ.get("OWL_CE_SKIP", "")
        # OWL_CE_DIRECT=1: reduce-scatter half without staging copies - the reduction kernel sums the peers' shards in place
        # over NVLink.  Correct (tools/exchange_check.py) but measured slower at N=2 (10.59 vs 10.13 ms/step): the 128-thread
        # co-resident CTAs cannot keep enough remote loads in flight, and the longer kernel sits next to the GEMMs longer.
        self.direct = os.environ.get("OWL_CE_DIRECT", "0") == "1" and self.world <= 16
        # In-switch reduction (NVLink SHARP), OWL_EXCHANGE=mc: when the symmetric allocation has a multicast mapping, a
        # bucket's all-reduce is
        #   barrier -> owl_mc_allreduce on my shard (multimem.ld_reduce: NVSwitch sums the ranks' copies; multimem.st: the
        #   result lands in every rank's copy) -> barrier
        # with no staging traffic and no copies.  Correct (tools/exchange_check.py: 4.9e-8 of ncclAllReduce, bitwise the same
        # on every rank, eager and in the graph) but NOT the default: measured at N=2 on audio.yml 10.45 ms (2 lanes, 64 CTAs)
        # and 10.88 ms (4 lanes, 148 CTAs) against 9.90 ms for the copy-engine exchange, and the 4-lane runs with 32 / 64
        # CTAs did not finish inside 300 s (profiles/r02_mc_sweep_n2.txt) - the multimem CTAs sit next to the GEMM CTAs for
        # the whole ~3.5 us switch round trip of every request, where a copy engine costs the SMs nothing.
        mode = os.environ.get("OWL_EXCHANGE", "ce").lower()
        try:
            self.mc_ptr = int(self.hdl.multicast_ptr) if mode == "mc" else 0
        except Exception:  # noqa: BLE001
            self.mc_ptr = 0
        if mode == "mc" and not self.mc_ptr:
            raise RuntimeError("OWL_EXCHANGE=mc: the symmetric allocation has no multicast mapping on this system")
        self.mode = "mc" if self.mc_ptr else "ce"
        self.mc_ctas = int(os.environ.get("OWL_MC_CTAS", "64"))

    def _shard(self, n: int) -> int:
        s = (n + self.world - 1) // self.world
        return (s + 3) // 4 * 4          # 16-byte shards

    def _stage(self, shard: int, lane: int):
        st = self._staging.get(lane)
        if st is None or st.shape[1] < shard:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("PeerExchange staging must exist before graph capture (run one eager step first)")
            st = self._staging[lane] = torch.empty(self.world - 1, shard, device=self.sym.device, dtype=f32)
        return st

    def barrier(self, lane: int = 0):
        self.hdl.barrier(channel=lane)

    def all_reduce(self, flat: torch.Tensor, lane: int = 0):
        """In place on `flat` (a contiguous fp32 slice of the symmetric allocation), on the CURRENT stream.  `lane`: exchanges
        in flight at the same time (on different streams) use different lanes = their own staging buffer and barrier channel;
        every rank must give the same exchange the same lane."""
        from . import ops
        n = flat.numel()
        off = (flat.data_ptr() - self.sym.data_ptr()) // 4
        if off < 0 or off + n > self.sym.numel():
            raise RuntimeError("PeerExchange.all_reduce: the tensor is not a slice of the symmetric bucket memory")
        S = self._shard(n)
        lo, hi = min(self.rank * S, n), min((self.rank + 1) * S, n)
        stage = None if (self.direct or self.mc_ptr) else self._stage(S, lane)
        # pull order rotated by rank: at any moment the ranks read from different peers
        others = [(self.rank + k) % self.world for k in range(1, self.world)]
        skip_copy, skip_red = self._skip in ("copy", "all"), self._skip in ("reduce", "all")
        self.barrier(lane)
        if self.mc_ptr:
            if hi > lo and not skip_red:
                ops.mc_allreduce(self.mc_ptr + 4 * (off + lo), hi - lo, 1.0 / self.world if self.average else 1.0,
                                 self.mc_ctas)
            self.barrier(lane)
            return
        if hi > lo and self.direct:
            if not skip_red:
                ops.shard_reduce_peers(flat[lo:hi], [self.peers[p][off + lo:off + hi] for p in others],
                                       1.0 / self.world if self.average else 1.0, self.max_ctas)
        elif hi > lo:
            if not skip_copy:
                for j, p in enumerate(others):
                    stage[j, :hi - lo].copy_(self.peers[p][off + lo:off + hi], non_blocking=True)
            if not skip_red:
                ops.shard_reduce(flat[lo:hi], stage, len(others), 1.0 / self.world if self.average else 1.0, self.max_ctas)
        self.barrier(lane)
        if not skip_copy:
            for p in others:
                plo, phi = min(p * S, n), min((p + 1) * S, n)
                if phi > plo:
                    flat[plo:phi].copy_(self.peers[p][off + plo:off + phi], non_blocking=True)


class GradBuckets:
    def __init__(self, model: torch.nn.Module, process_group=None, comm_stream: Optional[torch.cuda.Stream] = None,
                 average: bool = True, wgrad_side_stream: bool = True, backwards_per_step: int = 1,
                 comm_dtype: Optional[torch.dtype] = None):
        """backwards_per_step: how many times every block runs backward in one micro-batch (2 when the loss calls the
        model twice, as the reference's self-forcing / causvid / distillation trainers do).  A block's bucket is handed
        to NCCL after its LAST expected backward; one more than declared raises (the bucket would be reduced twice)."""
        from .nn.attn import DiT, DiTBlock
        import os
        # Wire format of the gradient exchange.  fp32 (default) is what DistributedDataParallel sends (rft_trainer.py:96).
        # bf16 (comm_dtype=torch.bfloat16 or OWL_COMM_DTYPE=bf16) halves the bytes NCCL moves and the time its CTAs share
        # the SMs / HBM with backward: the fp32 bucket is cast on the communication stream, all-reduced and written back
        # as fp32 (each rank's contribution is rounded to bf16 once: relative error 2^-9 per element, inside the 2e-2
        # gradient tolerance the path is held to; the optimizer still sees fp32 gradients).
        if comm_dtype is None:
            comm_dtype = torch.bfloat16 if os.environ.get("OWL_COMM_DTYPE", "fp32").lower() in ("bf16", "bfloat16") else f32
        self.comm_dtype = comm_dtype
        self._wire: dict = {}        # bucket slice (data_ptr, numel) -> persistent bf16 wire buffer
        self.backwards_per_step = max(1, int(backwards_per_step))
        self.step_id = 0
        self._seen: dict = {}
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.average = average
        self.accumulate = False      # False on the first micro-batch of a step: sinks are overwritten, not added to
        self._pending: List[int] = []
        self.blocks: List[torch.nn.Module] = [m for m in model.modules() if isinstance(m, DiTBlock)]
        groups: List[List[torch.nn.Parameter]] = []
        seen = set()
        # conditioning projections of a whole DiT stack live in ONE bucket, weights of all layers first
        # (rows in CondBank order: per layer adaln1.fc | gate1.fc_c | adaln2.fc | gate2.fc_c), then biases
        self.cond_stacks = []        # (dit, bucket index, n weight elements, d)
        cond_groups = []
        for dit in (m for m in model.modules() if isinstance(m, DiT)):
            cps = [b.cond_params() for b in dit.blocks]
            ws = [p for ps in cps for p in ps[0::2]]
            bs = [p for ps in cps for p in ps[1::2]]
            d = ws[0].shape[1] if ws else 0
            ok = ws and all(p.requires_grad and p.dtype == f32 and id(p) not in seen for p in ws + bs) and d % 64 == 0
            if ok:
                cond_groups.append((dit, ws + bs, sum(p.numel() for p in ws), d))
                seen.update(id(p) for p in ws + bs)
        for blk in self.blocks:
            ps = [p for p in blk.parameters() if p.requires_grad and id(p) not in seen]
            groups.append(ps)
            seen.update(id(p) for p in ps)
        for dit, ps, nw, d in cond_groups:
            self.cond_stacks.append((dit, len(groups), nw, d))
            groups.append(ps)
        rest = [p for p in model.parameters() if p.requires_grad and id(p) not in seen]
        self.rest_index = len(groups)
        groups.append(rest)
        self.flats: List[torch.Tensor] = []
        self._views: List = []       # (param, its bucket view): what begin_step re-attaches as p.grad
        self._bias_views: List[torch.Tensor] = []
        dev = next(model.parameters()).device
        sizes = [max(sum(_align(p.numel()) for p in ps), 64) for ps in groups]
        # all buckets in ONE allocation: symmetric memory (peer-addressable, PeerExchange) when several GPUs take part
        self.exchange: Optional[PeerExchange] = None
        base = None
        if dev.type == "cuda" and self.world > 1 and os.environ.get("OWL_EXCHANGE", "ce").lower() != "nccl":
            try:
                self.exchange = PeerExchange(sum(sizes), dev, process_group, average)
                base = self.exchange.sym
            except Exception as e:  # noqa: BLE001  (no symmetric memory on this system: NCCL carries the exchange)
                if dist.get_rank(process_group) == 0:
                    print(f"owl engine: symmetric-memory exchange unavailable ({type(e).__name__}: {str(e)[:120]}); "
                          "using ncclAllReduce", flush=True)
                self.exchange = None
        if base is None:
            base = torch.zeros(sum(sizes), device=dev, dtype=f32)
        goff = 0
        for gi, ps in enumerate(groups):
            flat = base[goff:goff + sizes[gi]]
            goff += sizes[gi]
            off = 0
            for p in ps:
                if p.dtype != f32:
                    raise RuntimeError("GradBuckets needs fp32 master parameters (the autocast training regime)")
                view = flat[off:off + p.numel()].view_as(p)
                p.grad = view
                self._views.append((p, view))
                if gi != self.rest_index:
                    p._owl_grad_sink = GradSink(view, self, pre_zeroed=p.dim() == 1)
                    if p.dim() == 1:
                        self._bias_views.append(view)
                off += _align(p.numel())
            self.flats.append(flat)
        self.is_cuda = dev.type == "cuda"
        self.comm_stream = comm_stream if comm_stream is not None else (torch.cuda.Stream(dev) if self.is_cuda else None)
        # copy-engine exchange: one exchange is a dependent chain (barrier, pulls, reduce, barrier, pulls: ~0.3 ms for a 75 MB
        # bucket); several chains in flight on their own streams keep the copy engines and the reduction busy
        # measured (audio.yml): N=2 10.16 / 10.09 / 10.00 ms with 3 / 2 / 4 lanes; N=8 11.12 / 10.95 ms with 4 / 8 lanes
        default_lanes = 4 if self.world <= 4 else 8
        n_lanes = max(1, int(os.environ.get("OWL_CE_LANES", str(default_lanes)))) if self.exchange is not None else 1
        self.comm_streams = [self.comm_stream] + [torch.cuda.Stream(dev) for _ in range(n_lanes - 1)]
        self._lane = 0
        for i, blk in enumerate(self.blocks):
            blk._stage.after_backward = (lambda i=i: self._block_done(i))
        self._cond_low: List[int] = []   # per cond bucket: weight rows [low, n_rows) already handed to NCCL this step
        for ci, (dit, gi, nw, d) in enumerate(self.cond_stacks):
            flat = self.flats[gi]
            n_rows = nw // d
            dit._stage.cond_sink = CondSink(flat[:nw].view(n_rows, d), flat[nw:nw + n_rows], self, ci)
            self._cond_low.append(n_rows)
        self.sync_grads = True       # set False on non-final accumulation micro-steps (DDP no_sync equivalent)
        # weight-gradient GEMMs / bias column sums go to a second stream (functional._Side)
        self.side_stream = torch.cuda.Stream(dev) if (self.is_cuda and wgrad_side_stream) else None
        from . import functional as OF
        OF.set_side_stream(self.side_stream)
        # gradient exchange concurrent with backward: keep NCCL on a few CTAs and leave them their SMs
        self.comm_ctas = 0
        if self.is_cuda and self.world > 1:
            import os
            from . import _lib
            self.comm_ctas = min(int(os.environ.get("OWL_COMM_CTAS", "16")), 64)  # measured at N=2: 0 -> 11.90, 16 -> 11.54, 24 -> 11.91 ms
            if self.exchange is not None:
                self.comm_ctas = 0   # copy engines move the data: backward keeps every SM
                # ... and the shard reduction (128 threads x 32 registers per CTA) rides next to the GEMM CTAs: they leave
                # 4 KB of shared memory per SM for it (every resident CTA costs 1 KB of reserved shared memory)
                _lib.check(_lib.load().owl_set_coresident_reserve(4096), "owl_set_coresident_reserve")
                self.exchange.max_ctas = 0      # one reduction CTA per SM
        self._reserved = 0

    # ---- step protocol -------------------------------------------------------------------------
    def begin_step(self):
        """Call before the first micro-batch's forward."""
        self.accumulate = False
        self.step_id += 1
        self._lane = 0                        # the same exchange takes the same lane on every rank, every step
        self._seen.clear()
        self._reattach()
        self._reserve_sms(0)                  # no exchange in flight during forward: GEMMs use every SM
        self.flats[self.rest_index].zero_()   # these grads arrive through autograd's accumulate-in-place
        if self._bias_views:
            torch._foreach_zero_(self._bias_views)
        self._pending.clear()
        self._reset_cond()

    def _reattach(self):
        """`optimizer.zero_grad(set_to_none=True)` (the reference trainers: rft_trainer.py:200, audio_rft_trainer.py:206)
        drops `p.grad`; the sinks keep writing into the buckets, so the views are put back before every step.  With the
        buckets in place zero_grad() is unnecessary: begin_step zeroes what accumulates and the rest is overwritten."""
        for p, view in self._views:
            if p.grad is not view:
                p.grad = view

    def _reset_cond(self):
        for ci, (dit, gi, nw, d) in enumerate(self.cond_stacks):
            self._cond_low[ci] = nw // d

    def next_micro_batch(self):
        self.accumulate = True
        self._seen.clear()
        self._reset_cond()

    def _cond_rows_done(self, ci: int, r0: int, r1: int):
        """CondBank finished the weight gradient of rows [r0, r1) (groups complete from the tail)."""
        dit, gi, nw, d = self.cond_stacks[ci]
        if not self._last_pass(("cond", ci, r0)):
            return
        if self.sync_grads and r1 == self._cond_low[ci]:
            self._reduce(self.flats[gi][r0 * d:r1 * d])
            self._cond_low[ci] = r0

    def _reserve_sms(self, n: int):
        """SMs the persistent GEMM grids leave free from now on (host-side launch state, so it is also what a graph
        capture bakes into the kernels launched after this point): 0 during forward, comm_ctas once the first
        bucket of the step is in flight."""
        if n != self._reserved and self.is_cuda and self.world > 1:
            from . import _lib
            _lib.check(_lib.load().owl_set_reserved_sms(n), "owl_set_reserved_sms")
            self._reserved = n

    def _reduce(self, flat: torch.Tensor):
        if self.world == 1:
            return
        self._reserve_sms(self.comm_ctas)
        if self.is_cuda:
            lane = self._lane % len(self.comm_streams)
            self._lane += 1
            stream = self.comm_streams[lane]
            stream.wait_stream(torch.cuda.current_stream())
            if self.side_stream is not None:
                stream.wait_stream(self.side_stream)
            op = dist.ReduceOp.AVG if self.average else dist.ReduceOp.SUM
            with torch.cuda.stream(stream):
                if self.exchange is not None:
                    self.exchange.all_reduce(flat, lane)
                elif self.comm_dtype == torch.bfloat16 and flat.numel() >= 4096 and flat.numel() % 8 == 0:
                    from . import ops
                    key = (flat.data_ptr(), flat.numel())
                    wire = self._wire.get(key)
                    if wire is None:
                        wire = self._wire[key] = torch.empty(flat.numel(), device=flat.device, dtype=torch.bfloat16)
                    ops.cast_f32_bf16(flat, wire)
                    dist.all_reduce(wire, op=op, group=self.group)
                    flat.copy_(wire)
                else:
                    dist.all_reduce(flat, op=op, group=self.group)
        else:  # gloo (CPU tests of the host logic): no AVG op
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
            if self.average:
                flat.div_(self.world)

    def _last_pass(self, key) -> bool:
        n = self._seen.get(key, 0) + 1
        self._seen[key] = n
        if n > self.backwards_per_step:
            raise RuntimeError(
                f"GradBuckets: {key} ran backward {n} times in one micro-batch but backwards_per_step="
                f"{self.backwards_per_step}; its bucket has already been handed to the all-reduce. Construct "
                f"GradBuckets(model, backwards_per_step=<model calls per loss>).")
        return n == self.backwards_per_step

    def _block_done(self, i: int):
        if self._last_pass(("block", i)) and self.sync_grads:
            self._reduce(self.flats[i])

    def finish_step(self):
        """Call after backward: reduce the non-block bucket and join the communication stream."""
        from . import functional as OF
        OF.side_join()
        if self.sync_grads:
            for ci, (dit, gi, nw, d) in enumerate(self.cond_stacks):
                low = self._cond_low[ci]           # weight rows not reduced yet (all of them on the per-block path)
                if low > 0:
                    self._reduce(self.flats[gi][:low * d])
                self._reduce(self.flats[gi][nw:])  # biases
                self._cond_low[ci] = 0
            self._reduce(self.flats[self.rest_index])
        if self.is_cuda and self.world > 1:
            for st in self.comm_streams[1:]:
                self.comm_stream.wait_stream(st)
            if self.exchange is not None and self.sync_grads:
                with torch.cuda.stream(self.comm_stream):
                    self.exchange.barrier()   # no peer still pulls from these buckets when the next step writes them
            torch.cuda.current_stream().wait_stream(self.comm_stream)
            self._reserve_sms(0)

    def grad_bytes(self) -> int:
        return sum(f.numel() * 4 for f in self.flats)

    def detach(self):
        from . import functional as OF
        if OF._Side.stream is self.side_stream:   # another engine may have installed its own since
            OF.set_side_stream(None)
        for dit, gi, nw, d in self.cond_stacks:
            dit._stage.cond_sink = None
        for blk in self.blocks:
            blk._stage.after_backward = None
            for p in blk.parameters():
                if hasattr(p, "_owl_grad_sink"):
                    del p._owl_grad_sink


class TrainStep:
    """fwd + bwd (+ bucketed all-reduce) of `loss_fn(model, *inputs)` with static input buffers; after
    `warmup` eager runs the step is captured into a CUDA graph and replayed.

    `inputs` given to __call__ may live on the host (pinned): they are copied into the static device
    buffers on the step's stream, inside what the caller times."""

    def __init__(self, model: torch.nn.Module, loss_fn: Callable, example_inputs: Sequence[torch.Tensor],
                 buckets: Optional[GradBuckets] = None, use_graph: bool = True, warmup: int = 3):
        self.model, self.loss_fn, self.buckets = model, loss_fn, buckets
        self.static_in = [torch.empty_like(t, device=next(model.parameters()).device) for t in example_inputs]
        for s, t in zip(self.static_in, example_inputs):
            s.copy_(t)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.loss: Optional[torch.Tensor] = None
        self.launches_per_step = 0
        self._copy_stream = None
        self._have_staged = False
        if use_graph:
            self._capture(warmup)

    def _eager(self):
        b = self.buckets
        if b is not None:
            b.begin_step()
        else:
            for p in self.model.parameters():
                p.grad = None
        loss = self.loss_fn(self.model, *self.static_in)
        loss.backward()
        if b is not None:
            b.finish_step()
        return loss.detach()

    def _capture(self, warmup: int):
        from . import ops
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):
                n0 = ops.launch_count()
                self._eager()
                self.launches_per_step = ops.launch_count() - n0
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._eager()

    def prefetch(self, *inputs: torch.Tensor) -> None:
        """Start the host->device copy of the NEXT step's inputs (pinned host tensors) on a copy stream, into staging
        buffers, while the current step is still running; the next __call__ without arguments consumes them with a
        device-to-device copy.  This is the data-loader side of the step (DataLoader(pin_memory=True) +
        `.to(device, non_blocking=True)` in rft_trainer.py:176-180), double-buffered."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream()
            self._staging = [torch.empty_like(s) for s in self.static_in]
            self._staged = torch.cuda.Event()
            self._consumed = torch.cuda.Event()
            self._consumed.record()
        self._copy_stream.wait_event(self._consumed)        # the previous staging contents have been read
        with torch.cuda.stream(self._copy_stream):
            for s, t in zip(self._staging, inputs):
                s.copy_(t, non_blocking=True)
            self._staged.record()
        self._have_staged = True

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        if not inputs and self._have_staged:
            cur = torch.cuda.current_stream()
            cur.wait_event(self._staged)
            for s, t in zip(self.static_in, self._staging):
                s.copy_(t, non_blocking=True)
            self._consumed.record()
            self._have_staged = False
        for s, t in zip(self.static_in, inputs):
            if t is not s:
                s.copy_(t, non_blocking=True)
        if self.graph is not None:
            self.graph.replay()
            return self.loss
        from . import ops
        n0 = ops.launch_count()
        self.loss = self._eager()
        self.launches_per_step = ops.launch_count() - n0
        return self.loss
