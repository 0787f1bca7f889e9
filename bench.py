#!/usr/bin/env python
"""Benchmark of the DiT hot path: train tokens/s (fwd + bwd + gradient all-reduce of the DiT stack).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload audio|dit_v4|av] [--impl owl|reference]

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for what every field means.
  value      whole-job tokens/s, inputs resident in HBM when the timed region starts (CUDA-graph replay)
  e2e        same metric through the public call (DiT forward/backward via engine.TrainStep) with the
             step's inputs in PINNED HOST memory: H2D copy of x/cond/target (double-buffered on a copy stream:
             step i+1's copy overlaps step i's compute, TrainStep.prefetch) and D2H read of the loss are inside
             the timed region, every step
  roofline   the dominant kernel (tcgen05 GEMM for audio/av, tcgen05 attention for dit_v4) against the
             measured bf16 peak: algorithmic FLOPs / CUDA-event time of its launches inside a step
  cpu_baseline   the oracle port (oracle/dit_oracle.py, plain torch CPU) on the host cores, bounded sample
--impl reference times that same CPU port as the reference arm (the reference is pure Python/PyTorch;
its CPU path for this block IS torch CPU ops -- /root/reference does not exist on the GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOADS = {
    # configs/audio.yml: model section + train.batch_size 32, window 88200 samples / 735 = 120 latents
    "audio": dict(cfg=dict(d_model=1024, n_heads=16, n_layers=16, tokens_per_frame=1, n_frames=10000, sample_size=120,
                           causal=True, has_audio=True, rope_impl="audio1d", local_window=16, global_window=None,
                           backbone="dit"), B=32, F=120, name="configs/audio.yml"),
    # configs/dit_v4.yml: B=1, 1536 frames x 64 tokens
    "dit_v4": dict(cfg=dict(d_model=1536, n_heads=24, n_layers=16, tokens_per_frame=64, n_frames=1536, sample_size=8,
                            causal=True, has_audio=False, rope_impl="motion", rope_ats_delta=2.0, local_window=16,
                            global_window=None, backbone="dit"), B=1, F=1536, name="configs/dit_v4.yml"),
    # configs/av_v5_8x8_sf.yml block geometry (shape only): B=8, 16 frames x 65 tokens, 24 layers
    "av": dict(cfg=dict(d_model=1536, n_heads=24, n_layers=24, tokens_per_frame=65, n_frames=16, sample_size=8,
                        causal=True, has_audio=True, rope_impl="ortho", local_window=16, global_window=None,
                        backbone="dit"), B=8, F=16, name="configs/av_v5_8x8_sf.yml (block geometry)"),
}
# configs/dit_v4_5B.yml: d=2560, H=20 (dh=128), 36 layers, per-block gradient checkpointing, B=1, 1536 frames x 64 tokens
WORKLOADS["dit_v4_5B"] = dict(cfg=dict(d_model=2560, n_heads=20, n_layers=36, tokens_per_frame=64, n_frames=1536,
                                       sample_size=8, causal=True, has_audio=False, rope_impl="motion", rope_ats_delta=2.0,
                                       local_window=16, global_window=None, backbone="dit", gradient_checkpointing=True),
                              B=1, F=1536, name="configs/dit_v4_5B.yml")
# configs/causvid.yml: game_rft_audio, 24 layers, d=1536, H=24, T=65 (8x8 video tokens + 1 audio token), window 16 frames,
# B=1 causal rollout (inference/causvid_pipeline.py; KV-cached per-frame loop of av_caching_v2.py:89-140)
WORKLOADS["causvid"] = dict(cfg=dict(d_model=1536, n_heads=24, n_layers=24, tokens_per_frame=65, n_frames=16, sample_size=8,
                                     channels=64, audio_channels=64, n_buttons=11, n_mouse_axes=2, causal=True,
                                     uncond=False, has_audio=True, rope_impl="ortho", local_window=16, global_window=None,
                                     backbone="dit", cfg_prob=0.0), B=1, F=16, name="configs/causvid.yml")
METRIC_ROLLOUT = "rollout frames/sec (causvid.yml KV-cached causal rollout, 1 flow step + cache commit per frame), whole job"
METRIC = "train tokens/sec (audio DiT stack fwd+bwd+grad all-reduce), whole job"


def unmasked_pairs(F_, T, W):
    """(query, key) token pairs the frame-block causal mask leaves visible (attn.py:43-59, no doc ids)."""
    if W is None or W >= F_:
        return T * T * F_ * (F_ + 1) // 2
    return T * T * (W * (W + 1) // 2 + (F_ - W) * W)


def step_flops(B, L, d, T, n_layers, local_window, global_window, local_idx=4):
    """fwd+bwd algorithmic FLOPs (SURVEY 8d closed forms): 3x linear + 3.5x attention, masked pairs excluded, activation
    recomputation under gradient checkpointing NOT counted."""
    F_ = L // T
    lin_fwd = B * L * (24 * d * d + 12 * d * d / T) * n_layers
    att_fwd = 0.0
    for i in range(n_layers):
        w = local_window if (i % local_idx) != 0 else global_window
        att_fwd += 4.0 * d * B * unmasked_pairs(F_, T, w)
    return {"linear_fwd": lin_fwd, "attn_fwd": att_fwd, "step": 3 * lin_fwd + 3.5 * att_fwd}


def measured_traffic(workload, dom_gemm):
    """dram__bytes_read+write per launch of the dominant kernel, from the ncu --set full capture summarised in
    profiles/traffic.json (one entry per workload, each naming the .md file it was read from); None when there is no
    capture for this workload/kernel."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None
    ent = json.load(open(p)).get(workload, {}).get("gemm" if dom_gemm else "attn")
    if not ent:
        return None, None
    return float(ent["dram_bytes_per_launch"]), ent.get("source")


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(tf=float(d["bf16_tflops"]), tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    hbm=float(d["hbm_gbs"]), src="measured (MEASURED_PEAKS.json, burst)")
    return dict(tf=1590.0, tf_sustained=1400.0, hbm=6650.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit())
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        mx = max(float(r[1]) for r in self.rows if len(r) >= 7)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port timed on the host cores (cpu_baseline, and --impl reference)
# ------------------------------------------------------------------------------------------------
def host_threads():
    """All the host cores this process may use.  torchrun exports OMP_NUM_THREADS=1 to every rank, which would silently
    make the CPU arm single-threaded at N > 1: set the count explicitly (physical cores, capped by the affinity mask)."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    try:
        import psutil
        phys = psutil.cpu_count(logical=False)
        if phys:
            n = min(n, phys)
    except Exception:
        pass
    return max(n, 1)


def cpu_port_step_time(workload, B, n_layers, steps, warmup, threads=None):
    """Seconds per fwd+bwd of the DiT stack on CPU (fp32) via oracle.dit_forward; returns (s/step, tokens/step, sample)."""
    from oracle import dit_oracle as O
    import owl_audio_exps_b200.nn.rope as R
    w = WORKLOADS[workload]
    cfg = types.SimpleNamespace(**{**w["cfg"], "n_layers": n_layers})
    F_ = w["F"] if not workload.startswith("dit_v4") else 32   # dense-mask SDPA is O(L^2): 98k tokens does not fit; 32 frames = 2048 tokens
    L = F_ * cfg.tokens_per_frame
    if threads:
        torch.set_num_threads(threads)
    torch.manual_seed(0)
    d = cfg.d_model
    W = {}
    for i in range(n_layers):
        pre = f"blocks.{i}."
        for name, (n, k) in {"attn.qkv": (3 * d, d), "attn.out": (d, d), "mlp.fc1": (4 * d, d), "mlp.fc2": (d, 4 * d),
                             "adaln1.fc": (2 * d, d), "gate1.fc_c": (d, d), "adaln2.fc": (2 * d, d),
                             "gate2.fc_c": (d, d)}.items():
            W[pre + name + ".weight"] = (torch.randn(n, k) * k ** -0.5).requires_grad_(True)
            W[pre + name + ".bias"] = torch.zeros(n, requires_grad=True)
    rope = R.get_rope_cls(cfg.rope_impl)(cfg)
    tables = [(rope.cos, rope.sin)] * n_layers
    x = torch.randn(B, L, d)
    cond = torch.randn(B, F_, d)
    tgt = torch.randn(B, L, d)

    def step():
        for v in W.values():
            v.grad = None
        y = O.dit_forward(x, cond, W, cfg, tables)
        torch.nn.functional.mse_loss(y, tgt).backward()

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    sample = f"{w['name']} DiT stack fwd+bwd, fp32 torch CPU, B={B}, L={L}, {n_layers} of {w['cfg']['n_layers']} layers"
    return dt, B * L, sample, torch.get_num_threads()


def run_reference(args):
    """--impl reference: the reference's CPU path for the block (torch CPU ops via the oracle port) on the GPU arm's
    config: same per-GPU batch, same sequence, 2 of the stack's layers timed and scaled to full depth (layers are
    identical in cost).  BASELINE.json's own CPU case (batch 1) is timed beside it and reported under `batch_1`."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    n_layers = 2
    B = args.batch or w["B"]
    steps, warmup = max(min(args.steps, 10), 1), max(min(args.warmup, 2), 1)   # bounded: seconds per step on the host
    dt, toks, sample, cores = cpu_port_step_time(args.workload, B, n_layers, steps, warmup, threads=host_threads())
    # per-layer cost is uniform: scale the measured layers to the full stack
    dt_full = dt * w["cfg"]["n_layers"] / n_layers
    val = toks / dt_full
    dt1, toks1, sample1, _ = cpu_port_step_time(args.workload, 1, n_layers, 3, 1, threads=host_threads())
    val1 = toks1 / (dt1 * w["cfg"]["n_layers"] / n_layers)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": dt_full * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "per_gpu_batch": B,
                   "note": "CPU reference arm; time of the measured layers scaled to the full depth"},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
        "batch_1": {"value": val1, "unit": "tokens/s", "sample": sample1},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def loss_fn(model, x, cond, target):
    # in the trainers x = proj_in(latents) and cond = t_embed(ts) carry gradients (audiorft.py:46-49), so the
    # stack's input gradients (dx of block 0, dcond through every conditioning projection) are part of the step
    x = x.detach().requires_grad_(True)
    cond = cond.detach().requires_grad_(True)
    y = model(x, cond)
    from owl_audio_exps_b200 import functional as OF
    return OF.mse_loss(y, target)      # mean((y - target)^2) in fp32 from the bf16 tensors, as models/*.py compute it


def run_owl(args):
    import torch.distributed as dist
    from owl_audio_exps_b200 import engine, ops
    import owl_audio_exps_b200.nn as N

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the sm_100a kernels are the only implementation)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NVSwitch carries the 1.24 GB/step exchange on a handful of CTAs; the rest of the SMs stay with backward
        if os.environ.get("OWL_NCCL_CTAS"):
            os.environ["NCCL_MAX_CTAS"] = os.environ["OWL_NCCL_CTAS"]
        dist.init_process_group("nccl", device_id=dev)
    w = WORKLOADS[args.workload]
    cfg = types.SimpleNamespace(**w["cfg"])
    if args.layers:
        cfg.n_layers = args.layers
    B, F_ = (args.batch or w["B"]), w["F"]
    T, d = cfg.tokens_per_frame, cfg.d_model
    L = F_ * T
    torch.manual_seed(0)
    model = N.DiT(cfg).to(dev)                       # fp32 master weights, reference init
    model.train()                                    # (gradient_checkpointing of dit_v4_5B is a training-mode switch)
    n_params = sum(p.numel() for p in model.parameters())
    buckets = engine.GradBuckets(model, wgrad_side_stream=os.environ.get("OWL_WGRAD_SIDE", "1") != "0")
    g = torch.Generator().manual_seed(1000 + rank)   # each rank draws its own batch (DistributedSampler analogue)
    host = [torch.randn(B, L, d, generator=g).to(torch.bfloat16).pin_memory(),
            torch.randn(B, F_, d, generator=g).to(torch.bfloat16).pin_memory(),
            torch.randn(B, L, d, generator=g).to(torch.bfloat16).pin_memory()]
    if args.workload.startswith("dit_v4"):
        # 98 304 tokens: ~90 GB of saved activations per step; a graph-private pool next to the eager warm-up's pool
        # does not fit 180 GB, and with 50-3000 us kernels there is no launch overhead for a graph to remove
        args.no_graph = True
    step = engine.TrainStep(model, loss_fn, host, buckets=buckets, use_graph=not args.no_graph, warmup=3)
    h2d = sum(t.numel() * t.element_size() for t in host)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, from_host):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        if from_host:
            step.prefetch(*host)                     # H2D of the first step's inputs (inside the timed region)
        for i in range(n):
            if from_host:
                loss = step()                        # consumes the staged inputs, runs the step
                if i + 1 < n:
                    step.prefetch(*host)             # H2D of the next step's inputs overlaps this step's compute
                _ = loss.item()                      # D2H read of the step's result, every step
            else:
                step(*step.static_in)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    for _ in range(max(args.warmup, 3)):
        step(*step.static_in)
    with ClockSampler(local) as clk:
        ms = timed(args.steps, False)
    for _ in range(2):
        step.prefetch(*host)
        step()
    ms_e2e = timed(args.steps, True)
    tokens = B * L * world
    value = tokens * args.steps / (ms * 1e-3)
    e2e = tokens * args.steps / (ms_e2e * 1e-3)

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel: CUDA events around every launch of one eager step
        fl = step_flops(B, L, d, T, cfg.n_layers, cfg.local_window, cfg.global_window)
        buckets.sync_grads = False   # rank 0 only from here on: no collectives (the other ranks are at the barrier)
        from owl_audio_exps_b200 import functional as OF
        OF.set_side_stream(None)     # per-launch event timing needs every kernel serialised on one stream
        eager = engine.TrainStep(model, loss_fn, host, buckets=buckets, use_graph=False)
        eager(*step.static_in)
        prof = None
        for _ in range(3):
            ops.profile_begin()
            eager(*step.static_in)
            prof = ops.profile_end()
        gemm_ms = sum(v["ms"] for k, v in prof.items() if k.startswith("gemm"))
        gemm_n = sum(v["n"] for k, v in prof.items() if k.startswith("gemm"))
        att_ms = sum(v["ms"] for k, v in prof.items() if k.startswith("attn"))
        att_n = sum(v["n"] for k, v in prof.items() if k.startswith("attn"))
        total_ms = sum(v["ms"] for v in prof.values())
        pk = _peaks()
        gemm_tf = 3 * fl["linear_fwd"] / (gemm_ms * 1e-3) / 1e12
        att_tf = 3.5 * fl["attn_fwd"] / (att_ms * 1e-3) / 1e12 if att_ms > 0 else 0.0
        dom_gemm = gemm_ms >= att_ms
        ach = gemm_tf if dom_gemm else att_tf
        traffic, traffic_src = measured_traffic(args.workload, dom_gemm)
        roofline = {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05)" if dom_gemm else "attn fwd/bwd kernels (tcgen05)",
                    "achieved": ach, "peak": pk["tf"], "unit": "TFLOP/s", "frac": ach / pk["tf"],
                    "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": pk["src"], "launches_per_step": gemm_n if dom_gemm else att_n,
                    "avg_launch_us": 1e3 * (gemm_ms / gemm_n if dom_gemm else att_ms / max(att_n, 1)),
                    "share_of_step_kernel_time": (gemm_ms if dom_gemm else att_ms) / total_ms,
                    "other": {"gemm_tflops": gemm_tf, "attn_tflops": att_tf,
                              "kernel_ms_per_step": {k: round(v["ms"], 4) for k, v in sorted(prof.items())}}}
        step_tf = fl["step"] * world * args.steps / (ms * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "per_gpu_batch": B, "tokens_per_sample": L, "d_model": d,
                       "n_layers": cfg.n_layers, "n_heads": cfg.n_heads, "params": n_params, "parallelism": f"dp{world}",
                       "step": "DiT stack fwd + bwd + grad all-reduce (bucket per block, overlapped; "
                               + (("in-switch reduction over the symmetric buckets (multimem.ld_reduce / multimem.st kernel), fp32"
                                   if getattr(buckets.exchange, "mode", "ce") == "mc" else
                                   "copy-engine peer copies over symmetric memory + shard-reduce kernel, fp32")
                                  if getattr(buckets, "exchange", None) is not None else
                                  "ncclAllReduce, wire dtype " + ("bf16" if buckets.comm_dtype == torch.bfloat16 else "fp32"))
                               + "); optimizer excluded",
                       "cuda_graph": not args.no_graph,
                       "l2": "per-step working set (weights+grads+activations, >4 GB) exceeds the 126 MB L2; no flush"},
            "tokens_per_s_per_gpu": value / world,
            "step_tflops_per_gpu": step_tf / world, "frac_of_bf16_peak_per_gpu": step_tf / world / pk["tf"],
            "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": step.launches_per_step * args.steps,
            "roofline": roofline, "clocks": clk.summary(),
        }
        if world == 1 and not args.no_cpu:
            # same per-GPU batch as the GPU arm (bounded: 2 layers, 3 steps), and BASELINE.json's batch-1 CPU case beside it
            dt, toks, sample, cores = cpu_port_step_time(args.workload, B, 2, 3, 1, threads=host_threads())
            dt_full = dt * cfg.n_layers / 2
            line["cpu_baseline"] = {"value": toks / dt_full, "unit": "tokens/s", "cores": cores, "kind": "port",
                                    "sample": sample + "; scaled to full depth"}
            dt1, toks1, sample1, _ = cpu_port_step_time(args.workload, 1, 2, 3, 1, threads=host_threads())
            line["cpu_baseline"]["batch_1"] = {"value": toks1 / (dt1 * cfg.n_layers / 2), "unit": "tokens/s",
                                               "sample": sample1 + "; scaled to full depth"}
        print(json.dumps(line), flush=True)
    if world > 1:
        # End of job without NCCL: the other ranks wait on the rendezvous store (not on a collective) until rank 0
        # has finished its rank-0-only profiling passes, then every rank leaves without tearing the communicator
        # down (destroy_process_group after captured NCCL graphs can block).
        import datetime
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            store.set("owl_bench_done", "1")
        else:
            store.wait(["owl_bench_done"], datetime.timedelta(seconds=1800))
        sys.stdout.flush()
        torch.cuda.synchronize()
        os._exit(0)



# ------------------------------------------------------------------------------------------------
# causvid.yml: causal rollout inference (frames/s)
# ------------------------------------------------------------------------------------------------
def cpu_rollout_frame_time(n_layers_sample=2, window=15):
    """Seconds per generated frame on the host cores: the oracle port's KV-cached decode of one 65-token frame against a
    `window`-frame cache (one denoise evaluation + one cache-commit evaluation), `n_layers_sample` layers, scaled by the
    caller.  fp32 torch CPU."""
    from oracle import dit_oracle as O
    import owl_audio_exps_b200.nn.rope as R
    w = WORKLOADS["causvid"]
    cfg = types.SimpleNamespace(**{**w["cfg"], "n_layers": n_layers_sample, "n_frames": window + 4})
    torch.manual_seed(0)
    d, T = cfg.d_model, cfg.tokens_per_frame
    W = {}
    for i in range(n_layers_sample):
        pre = f"blocks.{i}."
        for name, (n, k) in {"attn.qkv": (3 * d, d), "attn.out": (d, d), "mlp.fc1": (4 * d, d), "mlp.fc2": (d, 4 * d),
                             "adaln1.fc": (2 * d, d), "gate1.fc_c": (d, d), "adaln2.fc": (2 * d, d),
                             "gate2.fc_c": (d, d)}.items():
            W[pre + name + ".weight"] = torch.randn(n, k) * k ** -0.5
            W[pre + name + ".bias"] = torch.zeros(n)
    rope = R.get_rope_cls(cfg.rope_impl)(cfg)
    tables = [(rope.cos, rope.sin)] * n_layers_sample
    with torch.no_grad():
        cache = O.OracleKVCache(n_layers_sample, T)
        cache.should_update = True
        O.dit_forward(torch.randn(1, window * T, d), torch.randn(1, window, d), W, cfg, tables, kv_cache=cache)
        x, c = torch.randn(1, T, d), torch.randn(1, 1, d)

        def frame():
            cache.should_update = False
            O.dit_forward(x, c, W, cfg, tables, kv_cache=cache, decoding=True)          # denoise evaluation
            cache.should_update = True
            O.dit_forward(x, c, W, cfg, tables, kv_cache=cache, decoding=True)          # cache-commit evaluation
            cache.truncate(1)
        frame()
        t0 = time.perf_counter()
        n = 3
        for _ in range(n):
            frame()
        return (time.perf_counter() - t0) / n


def run_causvid(args):
    """configs/causvid.yml rollout.  Per GPU one independent replica on its own stream (north_star: 'inference rollout runs
    per-GPU on independent streams'); no collective on the data path, frames/s adds up over ranks."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    w = WORKLOADS["causvid"]
    if args.impl == "reference":
        if rank != 0:
            return
        torch.set_num_threads(host_threads())
        dt = cpu_rollout_frame_time(2) * w["cfg"]["n_layers"] / 2
        val = 1.0 / dt
        print(json.dumps({
            "impl": "reference", "metric": METRIC_ROLLOUT, "value": val, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": w["name"], "note": "CPU reference arm: oracle port of the DiT stack's KV-cached decode "
                       "(2 evaluations of one 65-token frame against a 15-frame cache), 2 of 24 layers scaled to full depth"},
            "cpu_baseline": {"value": val, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
                             "sample": "3 frames x 2 evaluations, 2 of 24 layers, fp32 torch CPU"},
            "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}),
            flush=True)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the sm_100a kernels are the only implementation)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from owl_audio_exps_b200 import functional as OF, ops
    from owl_audio_exps_b200.models import GameRFTAudioCore
    from owl_audio_exps_b200.nn.decode import RingKVCache
    import owl_audio_exps_b200.nn as N
    window, frames_per_step = w["F"], 16            # a "step" = 16 generated frames
    n_warm = max(args.warmup, 3)
    total_frames = (n_warm + 2 * args.steps + 4) * frames_per_step
    cfg = types.SimpleNamespace(**{**w["cfg"], "n_frames": window + total_frames + 2})   # RoPE table spans the rollout
    torch.manual_seed(0)
    model = GameRFTAudioCore(cfg).to(dev).bfloat16().eval()          # causvid_pipeline.py:36 (.cuda().bfloat16().eval())
    n_params = sum(p.numel() for p in model.parameters())
    T, C = cfg.tokens_per_frame, cfg.channels
    g = torch.Generator().manual_seed(7 + rank)
    ctx = window - 1
    hist = torch.randn(1, ctx, C, 8, 8, generator=g).to(dev, torch.bfloat16)
    aud = torch.randn(1, ctx, cfg.audio_channels, generator=g).to(dev, torch.bfloat16)
    mouse_h = torch.randn(total_frames + ctx, 2, generator=g).to(torch.bfloat16).pin_memory()
    btn_h = (torch.rand(total_frames + ctx, 11, generator=g) > 0.5).to(torch.bfloat16).pin_memory()
    out_h = torch.empty(C * 64 + cfg.audio_channels, dtype=torch.bfloat16).pin_memory()
    alpha = 0.2
    with torch.no_grad(), OF.frozen_weights():
        # prefill of the noised context with cache updates on (masked, training-shape kernels), then the ring
        cache = N.KVCache(cfg)
        cache.reset(1)
        cache.enable_cache_updates()
        t_ctx = torch.full((1, ctx), alpha, device=dev, dtype=torch.bfloat16)
        model(hist, aud, t_ctx, mouse_h[:ctx].to(dev)[None], btn_h[:ctx].to(dev)[None], kv_cache=cache)
        cache.disable_cache_updates()
        model.transformer.enable_decoding()
        ring = RingKVCache.from_cache(cache, window_rows=window * T, new_rows=T)
        x_in = torch.empty(1, 1, C, 8, 8, device=dev, dtype=torch.bfloat16)
        a_in = torch.empty(1, 1, cfg.audio_channels, device=dev, dtype=torch.bfloat16)
        t_in = torch.empty(1, 1, device=dev, dtype=torch.bfloat16)
        m_in = torch.empty(1, 1, 2, device=dev, dtype=torch.bfloat16)
        b_in = torch.empty(1, 1, 11, device=dev, dtype=torch.bfloat16)
        x_in.normal_(), a_in.normal_(), t_in.fill_(1.0), m_in.zero_(), b_in.zero_()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            n0 = ops.launch_count()
            model(x_in, a_in, t_in, m_in, b_in, kv_cache=ring)
            launches_per_eval = ops.launch_count() - n0
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            v_x, v_a = model(x_in, a_in, t_in, m_in, b_in, kv_cache=ring)
        out_d = torch.empty(C * 64 + cfg.audio_channels, device=dev, dtype=torch.bfloat16)
        frame_no = [0]

        def one_frame(from_host):
            i = frame_no[0]
            frame_no[0] += 1
            if from_host:                                   # this frame's controls arrive from pinned host memory
                m_in.copy_(mouse_h[ctx + i].view(1, 1, 2), non_blocking=True)
                b_in.copy_(btn_h[ctx + i].view(1, 1, 11), non_blocking=True)
            x_in.normal_(), a_in.normal_(), t_in.fill_(1.0)
            graph.replay()                                  # ONE flow step (causvid.yml rollout_steps: 1): x <- x - 1.0 * v
            x_in.sub_(v_x), a_in.sub_(v_a)
            out_d[:C * 64].copy_(x_in.view(-1)), out_d[C * 64:].copy_(a_in.view(-1))
            # re-noise the new frame to alpha (zlerp, causvid_pipeline.py:15-16) and run it once more to write its K/V rows
            x_in.mul_(1.0 - alpha).add_(torch.randn_like(x_in), alpha=alpha)
            a_in.mul_(1.0 - alpha).add_(torch.randn_like(a_in), alpha=alpha)
            t_in.fill_(alpha)
            graph.replay()
            ring.commit(T)
            if ring.n_frames() >= window:
                ring.truncate(1)
            if from_host:
                out_h.copy_(out_d, non_blocking=True)       # the generated latents go back to the host every frame
                torch.cuda.current_stream().synchronize()   # the caller consumes the frame before producing the next controls

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        def timed(n_frames, from_host):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record()
            for _ in range(n_frames):
                one_frame(from_host)
            e1.record()
            barrier()
            ms = e0.elapsed_time(e1)
            if world > 1:
                tt = torch.tensor([ms], device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                ms = tt.item()
            return ms

        for _ in range(n_warm * frames_per_step):
            one_frame(False)
        with ClockSampler(local) as clk:
            ms = timed(args.steps * frames_per_step, False)
        for _ in range(4):
            one_frame(True)
        ms_e2e = timed(args.steps * frames_per_step, True)
        # one decode evaluation alone (graph replays back to back), for the roofline of the weight-streaming schedule
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        eval_ms = e0.elapsed_time(e1) / 20
        # the pipeline's own call shape (inference/causvid_pipeline.py:112-163): the whole 16-frame window through the model,
        # sampling_steps = 1, no cache
        model.transformer.disable_decoding()
        xw = torch.randn(1, window, C, 8, 8, device=dev, dtype=torch.bfloat16)
        aw = torch.randn(1, window, cfg.audio_channels, device=dev, dtype=torch.bfloat16)
        tw = torch.full((1, window), alpha, device=dev, dtype=torch.bfloat16)
        mw, bw = mouse_h[:window].to(dev)[None], btn_h[:window].to(dev)[None]
        for _ in range(3):
            model(xw, aw, tw, mw, bw)
        gw = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gw):
            model(xw, aw, tw, mw, bw)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            gw.replay()
        e1.record()
        torch.cuda.synchronize()
        window_ms = e0.elapsed_time(e1) / 10
    n_frames = args.steps * frames_per_step
    value = world * n_frames / (ms * 1e-3)
    e2e = world * n_frames / (ms_e2e * 1e-3)
    if rank == 0:
        pk = _peaks()
        weight_bytes = 2.0 * n_params
        ach = weight_bytes / (eval_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC_ROLLOUT, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": n_warm, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "model": "GameRFTAudioCore (24 layers, d=1536, H=24, 65 tokens per frame)",
                       "params": n_params, "batch": 1, "window_frames": window, "frames_per_step": frames_per_step,
                       "flow_steps_per_frame": 1, "evaluations_per_frame": 2, "parallelism": f"{world} independent replicas",
                       "cuda_graph": "one graph of the model call, replayed for every denoise step and cache commit",
                       "l2": "2 GB of weights per evaluation exceed the 126 MB L2; no flush"},
            "ms_per_frame": ms / n_frames, "ms_per_evaluation": eval_ms, "launches_per_evaluation": launches_per_eval,
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": frames_per_step * (2 + 11) * 2,
                    "d2h_bytes_per_step": frames_per_step * out_h.numel() * 2, "ms_per_frame": ms_e2e / n_frames},
            "gpu_launches": launches_per_eval * 2 * n_frames,
            "roofline": {"bound": "hbm", "kernel": "gemm_smallm_kernel (weight-streaming split-K, tcgen05) over one model evaluation",
                         "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"],
                         "traffic": None, "peak_source": pk["src"],
                         "algorithmic_bytes_per_evaluation": weight_bytes,
                         "note": "achieved = bf16 weight bytes of the model (every weight is read once per evaluation) / "
                                 "time of one evaluation (graph replays back to back)"},
            "other": {"pipeline_full_window_ms": window_ms, "pipeline_full_window_frames_per_s": 1e3 / window_ms,
                      "note": "inference/causvid_pipeline.py's own call: all 16 frames through the model per new frame"},
            "clocks": clk.summary(),
        }
        if world == 1 and not args.no_cpu:
            torch.set_num_threads(host_threads())
            dtc = cpu_rollout_frame_time(2) * cfg.n_layers / 2
            line["cpu_baseline"] = {"value": 1.0 / dtc, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": "oracle port, KV-cached decode of one 65-token frame vs a 15-frame cache, 2 "
                                              "evaluations per frame, 2 of 24 layers scaled to full depth, fp32 torch CPU"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="owl", choices=["owl", "reference"])
    ap.add_argument("--workload", default="audio", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.workload == "causvid":
        run_causvid(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_owl(args)


if __name__ == "__main__":
    main()
