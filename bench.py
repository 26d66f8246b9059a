#!/usr/bin/env python
"""bench.py — train samples/sec of the B200 data-parallel training step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path (task.py:303-312) over one synthetic batch: K-pre -> forward GEMMs ->
loss -> backward GEMMs -> gradient all-reduce (N>1) -> optimizer.  Rank 0 prints ONE JSON line.

Workloads (BASELINE.json configs):
  t0_bf16_b65536  (default) T0 tabular MLP, bf16, batch 65536 per GPU — the per-GPU workload of config 3, the one
                  the 1/2/4/8-GPU metric is quoted on (weak scaling: per-GPU batch fixed)
  t0_fp32_b4096   config 2: T0, fp32 storage / TF32 tensor cores, batch 4096, single GPU
  t1_bf16_b65536  config 4: widened MLP (4x1024, 26 x 65536 x 64 embeddings), bf16, batch 65536 per GPU
  t0_fp32_b4096_cpu_gloo2  config 1 (`--impl reference` only): the oracle as real DDP over gloo, 2 CPU processes
Timing: CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks.  The per-step
working set (activations of a 65536-row batch, > 500 MB) is larger than the 126 MB L2, so no explicit flush.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

WORKLOADS = {
    "t0_bf16_b65536": dict(arch="tabular_mlp", precision="bf16", batch=65536, optimizer="sgd"),
    "t0_fp32_b4096": dict(arch="tabular_mlp", precision="fp32", batch=4096, optimizer="sgd"),   # 3xTF32 = fp32-grade
    "t0_tf32_b4096": dict(arch="tabular_mlp", precision="tf32", batch=4096, optimizer="sgd"),   # single TF32 pass
    "t1_bf16_b65536": dict(arch="tabular_mlp_wide", precision="bf16", batch=65536, optimizer="sgd"),
    "t1_bf16_b65536_adam": dict(arch="tabular_mlp_wide", precision="bf16", batch=65536, optimizer="adam"),
    # the reference-literal model (task.py:63,171: --arch resnet18, batch 1024 per process) — bench_conv.py
    "resnet18_fp32_b1024": dict(arch="resnet18", precision="fp32", batch=1024, optimizer="sgd", conv=True),
    "resnet18_tf32_b1024": dict(arch="resnet18", precision="tf32", batch=1024, optimizer="sgd", conv=True),
    "resnet18_bf16_b1024": dict(arch="resnet18", precision="bf16", batch=1024, optimizer="sgd", conv=True),
    # BASELINE config 1 (CPU only, `--impl reference`): the oracle as real DistributedDataParallel over gloo,
    # 2 processes x (host cores / 2) threads, 4096 rows per rank
    "t0_fp32_b4096_cpu_gloo2": dict(arch="tabular_mlp", precision="fp32", batch=4096, optimizer="sgd", cpu_gloo=2),
}
BATCHES_PER_RANK = 8          # distinct batches per rank, cycled (N = W*B*8 rows)
REF_SAMPLE_CAP = 65536        # rows per step of the CPU arms (bounded sample)


def _measure_tf32_peak(dev):
    """TF32 tensor-core peak measured the way MEASURED_PEAKS.json measures bf16: torch.matmul 8192^3 on fp32 operands
    with TF32 allowed, best of 10 (CUDA events).  A library GEMM, used as the yardstick only."""
    import torch

    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        best = float("inf")
        for i in range(13):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize(dev)
            if i >= 3:
                best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best / 1e3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                toks = [t for t in vis.split(",") if t.strip() != ""]
                if index < len(toks) and toks[index].strip().isdigit():
                    phys = int(toks[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------- CPU arms
def _cpu_oracle_throughput(spec, ds_rows, batch, world, steps, warmup, threads=None, time_budget_s=None):
    """Times the oracle port (oracle/reference_step.py: the reference's step with .cuda() removed, true fp32)
    on the host cores.  `batch` rows per step (a bounded sample of the GPU arm's per-step work)."""
    import torch

    from b200step.data import make_synthetic
    from oracle.reference_step import OracleTrainer

    if threads:
        torch.set_num_threads(threads)
    sd = dataclasses.asdict(spec)
    sd.pop("name")
    ds = make_synthetic(ds_rows, spec)
    tr = OracleTrainer(sd, ds, world=1, batch_size=batch, precision="fp32")
    for _ in range(warmup):
        tr.step()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        tr.step()
        times.append(time.perf_counter() - t0)
        if time_budget_s is not None and sum(times) > time_budget_s:   # bounded: the arm must end within minutes
            break
    steps = len(times)
    return {"samples_per_s": batch * steps / sum(times), "ms_per_step": 1e3 * sum(times) / steps,
            "cores": torch.get_num_threads(), "min_ms": 1e3 * min(times), "steps": steps}


def run_reference(args, rank):
    """`--impl reference`: the reference's own CPU path for this step.  task.py itself cannot run without CUDA
    and the network, so this is the oracle PORT on all host threads; rank 0 only."""
    if rank != 0:
        return
    import torch

    from b200step.model import ARCHS

    wl = WORKLOADS[args.workload]
    if wl.get("conv"):
        import bench_conv

        bench_conv.run_reference(args, wl)
        return
    spec = ARCHS[wl["arch"]]
    # --steps / --warmup are honoured up to a time budget (a 65536-row CPU step takes ~0.3 s on 32 threads)
    steps = max(1, args.steps)
    warm = max(1, min(args.warmup, 3))
    if wl.get("cpu_gloo"):
        _run_reference_gloo(args, wl, spec, steps, warm)
        return
    sample = min(wl["batch"] * args.gpus, REF_SAMPLE_CAP)
    r = _cpu_oracle_throughput(spec, sample * 2, sample, 1, steps, warm, threads=os.cpu_count(), time_budget_s=60.0)
    steps = r["steps"]
    out = {
        "impl": "reference", "metric": "train samples/sec", "value": r["samples_per_s"], "unit": "samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "arch": wl["arch"], "per_gpu_batch": wl["batch"],
                   "optimizer": wl["optimizer"],
                   "arm": "ONE process, all host threads, oracle port, at every --gpus value (task.py has no CPU path; "
                          "BASELINE config 1, DDP over gloo with 2 CPU processes, is --workload "
                          "t0_fp32_b4096_cpu_gloo2)"},
        "cpu_baseline": {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{steps} steps of {sample} rows (global batch capped at {REF_SAMPLE_CAP}), "
                                   f"fp32 torch CPU, host threads={r['cores']} of {os.cpu_count()} logical cpus"},
        "e2e": {"value": r["samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def _run_reference_gloo(args, wl, spec, steps, warm):
    """Config 1: the oracle under torch's own DistributedDataParallel, gloo backend, one process per rank."""
    import torch

    from b200step.data import make_synthetic
    from oracle.reference_step import run_ddp_gloo

    world, B = wl["cpu_gloo"], wl["batch"]
    threads = max(1, (os.cpu_count() or 2) // world)
    sd = dataclasses.asdict(spec)
    sd.pop("name")
    ds = make_synthetic(world * B * 2, spec)
    res = run_ddp_gloo(sd, ds, world=world, batch_size=B, steps=warm + steps, threads=threads,
                       port=29500 + (os.getpid() % 400), time_it=True)[0]
    times = res["times"][warm:]
    sps = world * B * len(times) / sum(times)
    out = {
        "impl": "reference", "metric": "train samples/sec", "value": sps, "unit": "samples/s", "n_gpus": 0,
        "steps": len(times), "warmup": warm, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "arch": wl["arch"], "per_rank_batch": B, "ranks": world,
                   "backend": "gloo", "optimizer": wl["optimizer"]},
        "cpu_baseline": {"value": sps, "unit": "samples/s", "cores": threads * world, "kind": "port",
                         "sample": f"{len(times)} DDP steps of {world} x {B} rows, {world} processes x {threads} "
                                   f"threads, torch {torch.__version__} CPU fp32 over gloo"},
        "e2e": {"value": sps, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "final_loss": res["losses"][-1],
    }
    print(json.dumps(out), flush=True)


# --------------------------------------------------------------------------------------------- parity leg
PARITY_LOSS_REL = 2e-3      # same bars as tests/test_step_parity_gpu.py (bf16 path against the bf16-storage oracle)
PARITY_PARAM_REL_L2 = 1e-2


def parity_check(rank, world, dev, comm, dist):
    """Correctness evidence printed with every bench line, OUTSIDE the timed region: 4 steps (eager, then graph
    capture + replay) of a T0-shaped model through the same TrainStep / Communicator / bucket schedule as the
    measured run, against OracleTrainer(world=N) on rank 0 (the oracle here is the checker, never the thing
    measured).  The tables are cut to 100 rows and the batch to 512 rows per rank so that the CPU side takes
    seconds; layer widths (448/512/512/1000) and hence the gradient buckets are T0's."""
    import torch

    from b200step.data import make_synthetic
    from b200step.engine import StepConfig, TrainStep
    from b200step.model import ARCHS, init_parameters
    from b200step.sampler import rank_indices

    spec = dataclasses.replace(ARCHS["tabular_mlp"], cardinality=100)
    B, steps = 512, 4
    rows = world * B * 2
    ds = make_synthetic(rows, spec)
    eng = TrainStep(spec, init_parameters(spec, 0), ds, rank_indices(rows, world, rank),
                    StepConfig(batch_size=B, world=world, rank=rank, precision="bf16"), comm, device=dev)
    losses = []
    for _ in range(steps):
        eng.step()
        losses.append(eng.loss_value())
    eng.synchronize()
    flat = eng.p.detach().clone()
    mine = torch.tensor(losses, dtype=torch.float64, device=dev)
    layout, n_buckets = eng.layout, len(eng.buckets)
    eng.close()
    identical, all_losses = True, [mine]
    if world > 1:
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        identical = all(torch.equal(g, gathered[0]) for g in gathered)
        all_losses = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(all_losses, mine)
    if rank != 0:
        return None
    from oracle.reference_step import OracleTrainer

    sd = dataclasses.asdict(spec)
    sd.pop("name")
    oracle = OracleTrainer(sd, ds, world=world, batch_size=B, precision="bf16")
    loss_rel = 0.0
    for s in range(steps):
        ref = oracle.step()                     # one local loss per rank (task.py:310 on that rank)
        for r in range(world):
            loss_rel = max(loss_rel, abs(ref[r] - float(all_losses[r][s])) / max(1.0, abs(ref[r])))
    so, sg = oracle.state_dict(), layout.unflatten(flat)
    param_rel = max(((so[k] - sg[k]).norm() / so[k].norm().clamp_min(1e-12)).item() for k in so)
    ok = bool(identical and loss_rel <= PARITY_LOSS_REL and param_rel <= PARITY_PARAM_REL_L2)
    return {"ok": ok, "loss_rel_max": loss_rel, "param_rel_l2_max": param_rel, "ranks_bit_identical": bool(identical),
            "steps": steps, "world": world, "gradient_buckets": n_buckets,
            "model": "T0 layer widths, 26 x 100 x 16 tables, 512 rows per rank, bf16, SGD",
            "against": "OracleTrainer(world=N) on rank 0 (oracle/reference_step.py)",
            "tolerance": {"loss_rel": PARITY_LOSS_REL, "param_rel_l2": PARITY_PARAM_REL_L2}}


# --------------------------------------------------------------------------------------------- GPU arm
def run_b200(args, rank, local_rank, world):
    import torch

    from b200step.data import make_synthetic
    from b200step.engine import Communicator, StepConfig, TrainStep
    from b200step.model import ARCHS, FlatLayout, init_parameters
    from b200step.sampler import rank_indices

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: the training step has no CPU path")
    wl = WORKLOADS[args.workload]
    if wl.get("conv"):
        import bench_conv

        bench_conv.run_b200(args, rank, local_rank, world, wl, _peaks(), ClockSampler, _measure_tf32_peak)
        return
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if wl.get("cpu_gloo"):
        raise SystemExit(f"--workload {args.workload} is the CPU/gloo baseline (config 1): run it with --impl reference")
    spec = ARCHS[wl["arch"]]
    B = wl["batch"]
    comm = None
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group(backend="nccl", init_method="env://", world_size=world, rank=rank,
                                device_id=dev)
        uid = [Communicator.make_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        comm = Communicator(rank, world, uid[0])

    parity = None if args.no_parity else parity_check(rank, world, dev, comm, dist)

    rows = world * B * BATCHES_PER_RANK
    ds = make_synthetic(rows, spec, zipf_alpha=args.zipf_alpha)
    cfg = StepConfig(batch_size=B, world=world, rank=rank, precision=wl["precision"], optimizer=wl["optimizer"])
    if os.environ.get("B200STEP_BUCKET_CAP_BYTES"):      # experiments with the DDP-style bucket cap
        cfg.bucket_cap_bytes = int(os.environ["B200STEP_BUCKET_CAP_BYTES"])
        cfg.first_bucket_bytes = min(cfg.first_bucket_bytes, cfg.bucket_cap_bytes)
    if os.environ.get("B200STEP_SM_RESERVE"):             # experiments: SMs the wgrads leave to the all-reduce
        cfg.comm_sm_reserve = int(os.environ["B200STEP_SM_RESERVE"])
    if os.environ.get("B200STEP_PREFETCH_RESERVE"):
        cfg.prefetch_sm_reserve = int(os.environ["B200STEP_PREFETCH_RESERVE"])
    if os.environ.get("B200STEP_L2_ORDER"):               # A/B: alternating row-tile direction of consecutive GEMMs
        cfg.l2_tile_order = os.environ["B200STEP_L2_ORDER"] != "0"
    if os.environ.get("B200STEP_NCCL_REG"):               # A/B: gradients in NCCL-registered memory (NVLS-capable)
        cfg.nccl_registered_grads = os.environ["B200STEP_NCCL_REG"] != "0"
    if os.environ.get("B200STEP_GROUPED_WGRAD"):          # A/B at world > 1: one wgrad launch, buckets reduced after it
        cfg.grouped_wgrad = os.environ["B200STEP_GROUPED_WGRAD"] != "0"
    if os.environ.get("B200STEP_PREFETCH_CTAS"):          # A/B: CTA slots per SM of the overlapped K-pre
        cfg.prefetch_ctas_per_sm = int(os.environ["B200STEP_PREFETCH_CTAS"])
    if os.environ.get("B200STEP_EMB_SIDE"):               # A/B: K-embbwd on a side stream under the grouped wgrad
        cfg.emb_bwd_side = os.environ["B200STEP_EMB_SIDE"] != "0"
    if os.environ.get("B200STEP_PREFETCH_PRE"):           # A/B: K-pre of the next batch under wgrad / the all-reduces
        cfg.prefetch_pre = os.environ["B200STEP_PREFETCH_PRE"] != "0"
    idx = rank_indices(rows, world, rank)
    # keep only this rank's rows resident (same values as the global dataset, re-indexed 0..n-1)
    local = dataclasses.replace(ds, dense=ds.dense[idx], cat=ds.cat[idx], label=ds.label[idx])
    host_local = local
    eng = TrainStep(spec, init_parameters(spec, 0), local, torch.arange(idx.numel()), cfg, comm, device=dev)
    del ds

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident measurement ("value") -------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        eng.step()
    eng.synchronize()
    sampler = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record(eng.main)
    for _ in range(args.steps):
        eng.step()
    e1.record(eng.main)
    eng.synchronize()
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    loss_after = eng.loss_value()
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * B * args.steps / (ms / 1e3)

    # ---- sustained leg: the same graph-replayed step for >= --sustain-seconds, with its own clock record ---------
    # (`value` above is a burst of --steps steps, a few ms at boost clocks; a training run settles under the power cap)
    sustained = None
    if args.sustain_seconds > 0:
        n_sus = max(args.steps, int(args.sustain_seconds * 1e3 / (ms / args.steps)) + 1)
        sampler2 = ClockSampler(local_rank)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sampler2.start()
        s0.record(eng.main)
        for _ in range(n_sus):
            eng.step()
        s1.record(eng.main)
        eng.synchronize()
        barrier()
        clocks2 = sampler2.stop()
        sus_ms = s0.elapsed_time(s1)
        if dist is not None:
            t = torch.tensor([sus_ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sus_ms = float(t.item())
        sustained = {"seconds": sus_ms / 1e3, "steps": n_sus, "ms_per_step": sus_ms / n_sus,
                     "value": world * B * n_sus / (sus_ms / 1e3), "unit": "samples/s", "clocks": clocks2}

    # ---- end-to-end measurement: host batches, H2D inside the timed region, loss read back every step --------
    nb = min(BATCHES_PER_RANK, 4)
    host = [eng.pack_host_batch(host_local.dense[i * B:(i + 1) * B], host_local.cat[i * B:(i + 1) * B],
                                host_local.label[i * B:(i + 1) * B]) for i in range(nb)]
    e2e_steps = max(3, min(args.steps, 50))
    for i in range(5):                      # warm-up: eager step, then graph capture for every staging buffer
        eng.host_loss(eng.step_from_host(host[i % nb]))
    eng.synchronize()
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(eng.main)
    pending = []
    for i in range(e2e_steps):
        pending.append(eng.step_from_host(host[i % nb]))   # H2D of this step's batch, then the step, then loss D2H
        if len(pending) > 2:
            _ = eng.host_loss(pending.pop(0))              # every step's loss is read on the host, two steps behind
    while pending:
        _ = eng.host_loss(pending.pop(0))
    g1.record(eng.main)
    eng.synchronize()
    barrier()
    e2e_ms = g0.elapsed_time(g1)   # device time from the first H2D wait to the last step, host gaps included
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e = {"value": world * B * e2e_steps / (e2e_ms / 1e3), "unit": "samples/s",
           "h2d_bytes_per_step": int(host[0].numel()), "d2h_bytes_per_step": 4, "steps": e2e_steps}

    # ---- per-kernel timing of one eager step (CUDA events around each launch) -> roofline --------------------
    prof = profile_kernels(eng, spec, B, world)
    barrier()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peaks = _peaks()
    # Roofline of the dominant kernel.  The nine GEMM launches of a step are ONE kernel (gemm_kernel / its grouped
    # wgrad twin) and together the largest share of the step: achieved = their algorithmic FLOPs / their summed
    # device time (CUDA events around each launch, measured in this run).  The largest memory-bound launch is
    # reported next to it as `roofline_hbm`.
    # `ms` of a kernel = CUDA events around its launch in EAGER steps; their sum exceeds the graph-replayed step
    # (`ms_per_step`) because in the graph each kernel's prologue overlaps its predecessor's tail (programmatic
    # dependent launch) and no event records sit between launches.  `share` = ms / sum of the kernels' ms;
    # `share_of_step_ms` = ms / ms_per_step (these add up to kernel_ms_sum / ms_per_step > 1).
    total_ms = sum(r["ms"] * r["per_step"] for r in prof)
    step_ms = ms / args.steps
    for r in prof:
        r["share"] = r["ms"] * r["per_step"] / total_ms
        r["share_of_step_ms"] = r["ms"] * r["per_step"] / step_ms
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")  # dram bytes per launch from `ncu --set full` captures
    traffic_tab = {}
    if os.path.exists(tpath):
        with open(tpath) as fh:
            blob = json.load(fh)
            # newest capture of this workload: keys "<workload>@<round>" -> {"<launch index>:<kernel>": DRAM bytes}
            tags = sorted(k for k in blob if k.startswith(args.workload + "@"))
            traffic_tab = blob[tags[-1]] if tags else {}
    gem = [r for r in prof if r["bound"] == "tensor"]
    mem = [r for r in prof if r["bound"] == "hbm"]
    gem_ms = sum(r["ms"] * r["per_step"] for r in gem)
    gem_tflops = sum(r["tflops"] * r["ms"] * r["per_step"] for r in gem) / gem_ms      # = sum(flops) / sum(time)
    if wl["precision"] == "bf16":
        tensor_peak, peak_note = peaks["bf16_tflops"], peaks["source"] + " bf16 burst"
    else:
        # fp32 storage: the yardstick is the TF32 tensor-core rate, measured in this run.  "fp32" spends three TF32
        # products per algorithmic multiply-add (error compensation), so its ceiling is a third of it.
        tf32_peak = _measure_tf32_peak(dev)
        passes = 3.0 if wl["precision"] == "fp32" else 1.0
        tensor_peak = tf32_peak / passes
        peak_note = (f"TF32 peak measured in this run (torch.matmul 8192^3, best of 10): {tf32_peak:.1f} TFLOP/s"
                     + (" / 3 passes of the error-compensated product" if passes > 1 else ""))
    # DRAM bytes (read + write, ncu --set full, scripts/ncu_step_summary.py) of the step's GEMM-type launches, per launch
    gem_traffic = [v for k, v in traffic_tab.items()
                   if k.split(":")[1] in ("gemm_kernel", "linear_ce_kernel", "wgrad_grouped_kernel")]
    roof_tensor = {"bound": "tensor", "kernel": "gemm_kernel (fwd + dgrad + wgrad launches of the step)",
                   "achieved": gem_tflops, "peak": tensor_peak, "unit": "TFLOP/s", "frac": gem_tflops / tensor_peak,
                   "traffic": (sum(gem_traffic) / len(gem_traffic)) if gem_traffic else None,
                   "traffic_note": "mean DRAM bytes per GEMM-type launch of one step (profiles/ncu_traffic.json)",
                   "share_of_step": gem_ms / total_ms, "share_of_step_ms": gem_ms / step_ms,
                   "kernel_ms": gem_ms, "flops_basis": "2*M*N*K on unpadded dims (SURVEY.md 8d)",
                   "launches_per_step": sum(r["per_step"] for r in gem),
                   "peak_source": peak_note}
    topm = max(mem, key=lambda r: r["ms"])
    roof_hbm = {"bound": "hbm", "kernel": topm["name"], "achieved": topm["gbs"], "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": topm["gbs"] / peaks["hbm_gbs"],
                "traffic": next((v for k, v in traffic_tab.items() if topm["name"].split("[")[0] in k), None),
                "share_of_step": topm["share"], "peak_source": peaks["source"] + " copy bandwidth"}
    roof = roof_tensor if gem_ms >= topm["ms"] * topm["per_step"] else roof_hbm
    flops = spec.flops_per_sample()["total"]
    step_tflops = flops * B / (ms / args.steps / 1e3) / 1e12
    out = {
        "metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": {"bf16": "bf16", "tf32": "tf32", "fp32": "f32"}[wl["precision"]], "data": "synthetic",
        "config": {"workload": args.workload, "arch": wl["arch"], "per_gpu_batch": B, "global_batch": B * world,
                   "optimizer": wl["optimizer"], "params": spec.num_params, "parallelism": f"dp{world}",
                   "l2": "per-step working set > 126 MB L2 (no flush needed)" if B >= 32768 else
                         "working set fits L2: dataset cycles 8 batches",
                   "cuda_graph": True, "prefetch_pre": eng.prefetch,
                   "categorical_ids": "uniform" if args.zipf_alpha is None else f"zipf(alpha={args.zipf_alpha})"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": eng.kernels_per_step * args.steps,
        "roofline": roof, "roofline_hbm": roof_hbm, "kernels": prof, "kernel_ms_sum": total_ms,
        "step_tflops_per_gpu": step_tflops,
        "final_loss": loss_after,
    }
    if sustained is not None:
        out["sustained"] = sustained
    if parity is not None:
        out["parity"] = parity
    if world == 1 and not args.no_cpu_baseline:
        sample = min(B, REF_SAMPLE_CAP)
        r = _cpu_oracle_throughput(spec, sample * 2, sample, 1, 5, 2)
        out["cpu_baseline"] = {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["cores"], "kind": "port",
                               "sample": f"5 steps of {sample} rows after 2 warm-up, oracle port (torch CPU fp32), "
                                         f"{r['cores']} threads"}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"parity check failed: {json.dumps(parity)}")


def profile_kernels(eng, spec, B, world):
    """Average device time of each kernel class over a few eager steps (events on the launching stream), with the
    algorithmic FLOPs / bytes per launch -> achieved TFLOP/s or GB/s."""
    import torch

    from b200step import ops

    names = ["preprocess", "linear_fwd", "linear_ce", "ce_loss_grad", "linear_dgrad", "linear_dgrad_emb", "linear_wgrad",
             "linear_wgrad_grouped",
             "colsum", "emb_bwd", "zero", "sgd_step", "adam_step", "split_tf32"]
    records = []
    orig = {n: getattr(ops, n) for n in names}
    es = 2 if eng.dt == torch.bfloat16 else 4

    def wrap(name, fn):
        def inner(*a, **k):
            st = k.get("stream") or eng.main
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(st)
            r = fn(*a, **k)
            a1.record(st)
            label, work = name, {}
            # FLOPs follow SURVEY.md §8(d): 2*M*N*K on UNPADDED dims (the first layer's K is stored padded to a
            # whole number of 128-byte rows), and the layer-0 dgrad only for the embedding columns
            unpad = lambda kk: spec.in_features if (kk == eng.k0 and eng.k0 != spec.in_features) else kk
            x3 = 3 if getattr(eng, "x3", False) else 1   # "fp32": contraction dimension 3x as long (hi/lo copies);
            #                                               FLOPs stay the algorithmic ones
            if name == "linear_wgrad_grouped":
                label = f"linear_wgrad_grouped[{k['M'] // x3}x{len(k['Ns'])} layers]"
                work = {"flops": sum(2.0 * (k["M"] // x3) * n_ * unpad(k_) for n_, k_ in zip(k["Ns"], k["Ks"]))}
            elif name == "linear_dgrad_emb":
                kk = k["num_cat"] * k["embed_dim"]
                label = f"linear_dgrad_emb[{k['M']}x{k['N']}x{kk}]"
                work = {"flops": 2.0 * k["M"] * k["N"] * kk}   # embedding columns only; the scatter rides along
            elif name.startswith("linear_"):
                M, N, K = k["M"], k["N"], k["K"]
                if name in ("linear_fwd", "linear_ce"):
                    K //= x3
                elif name == "linear_dgrad":
                    N //= x3
                elif name == "linear_wgrad":
                    M //= x3
                label = f"{name}[{M}x{N}x{K}]" + ("+dbias" if k.get("dbias") is not None else "")
                keff = unpad(K)
                if name == "linear_dgrad" and K == eng.k0 and spec.num_cat:
                    keff = spec.num_cat * spec.embed_dim          # input gradient of the embedding columns only
                work = {"flops": 2.0 * M * N * keff}  # linear_ce: the output layer's GEMM; the loss rides along
            elif name == "preprocess":
                raw = 4 * spec.num_dense + 8 * spec.num_cat + 8 + 8
                # `bytes` = SURVEY.md §8(d)'s per-sample figure (it counts the gathered table rows, which on T0 come
                # from the L2-resident 1.7 MB tables); `dram_bytes` = what has to cross HBM: raw row in, x0/y/idx out
                work = {"bytes": B * (raw + 4 * spec.num_cat * spec.embed_dim + es * eng.k0 + 8 + 4 * spec.num_cat),
                        "dram_bytes": B * (raw + es * eng.k0 + 8 + 4 * spec.num_cat)}
            elif name == "ce_loss_grad":
                work = {"bytes": B * (2 * es * spec.num_classes + 8)}
            elif name == "split_tf32":
                outs = (a[1] is not None) + (a[3] is not None)
                work = {"bytes": 4.0 * k["rows"] * k["cols"] * (1 + 3 * outs)}
            elif name == "colsum":
                work = {"bytes": k["M"] * k["N"] * es}
            elif name == "emb_bwd":
                # dX columns + ids read once; the table gradients are L2-resident atomics (counted once: zero + final)
                work = {"bytes": B * spec.num_cat * (spec.embed_dim * es + 4) + 8 * spec.num_cat * spec.cardinality * spec.embed_dim}
            elif name == "zero":
                work = {"bytes": a[0].numel() * a[0].element_size()}
            elif name == "sgd_step":
                work = {"bytes": a[0].numel() * (24 if k.get("zero_grad") else 20) + 2 * k.get("n_shadow", 0)}
            elif name == "adam_step":
                work = {"bytes": a[0].numel() * (32 if k.get("zero_grad") else 28) + 2 * k.get("n_shadow", 0)}
            records.append((label, a0, a1, work))
            return r
        return inner

    reps = 3
    try:
        for n in names:
            setattr(ops, n, wrap(n, orig[n]))
        eng._pre_ready = False      # the eager steps below run K-pre themselves and overwrite the prefetched batch
        for _ in range(reps):
            eng._enqueue(B, None, eng.row_idx, eng.cursor, True)
        eng.synchronize()
    finally:
        for n in names:
            setattr(ops, n, orig[n])
    agg = {}
    for label, a0, a1, work in records:
        d = agg.setdefault(label, {"ms": 0.0, "launches": 0, "work": work})
        d["ms"] += a0.elapsed_time(a1)
        d["launches"] += 1
    out = []
    for label, d in agg.items():
        per = d["ms"] / d["launches"]
        rec = {"name": label, "ms": per, "per_step": d["launches"] // reps}
        if "flops" in d["work"]:
            rec.update(bound="tensor", tflops=d["work"]["flops"] / (per / 1e3) / 1e12)
        else:
            rec.update(bound="hbm", gbs=d["work"]["bytes"] / (per / 1e3) / 1e9)
            if "dram_bytes" in d["work"]:
                rec["dram_gbs"] = d["work"]["dram_bytes"] / (per / 1e3) / 1e9
        out.append(rec)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="t0_bf16_b65536")
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    ap.add_argument("--sustain-seconds", dest="sustain_seconds", type=float, default=2.0,
                    help="after the timed --steps burst, run the same step back to back for this long and report it as "
                         "`sustained` with its own clock / throttle record (0 = skip)")
    ap.add_argument("--no-parity", dest="no_parity", action="store_true",
                    help="skip the parity leg (4 small steps against the CPU oracle, outside the timed region)")
    ap.add_argument("--zipf-alpha", dest="zipf_alpha", type=float, default=None,
                    help="categorical ids ~ Zipf(alpha) instead of uniform (SURVEY.md §8d secondary dataset row)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # launched without torchrun: re-exec under torch.distributed.run
            import subprocess

            port = 29500 + (os.getpid() % 2000)
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
            sys.exit(subprocess.call(cmd))
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
