"""bench.py's legs for the reference-literal workloads (`--workload resnet18_*`): ResNet-18 on 3x32x32 images, the
model `task.py` trains by default (task.py:63,171), batch 1024 per process (task.py:58).  Same JSON contract as the
tabular workloads; imported by bench.py only."""
from __future__ import annotations

import json
import os
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
BATCHES_PER_RANK = 8


def resnet_flops_per_sample(lay) -> dict:
    """Algorithmic FLOPs of one sample: 2*Ho*Wo*Cout*K per convolution (unpadded K) + the fc layer; backward = dgrad
    + wgrad for every layer except conv1's dgrad (the input needs no gradient)."""
    convs = [lay.stem] + [c for b in lay.blocks for c in (b.conv1, b.conv2, b.down) if c is not None]
    fwd = sum(2.0 * c.Ho * c.Wo * c.cout * c.K for c in convs) + 2.0 * 512 * lay.num_classes
    bwd = 2 * fwd - 2.0 * lay.stem.Ho * lay.stem.Wo * lay.stem.cout * lay.stem.K
    return {"fwd": fwd, "bwd": bwd, "total": fwd + bwd}


def cpu_oracle_throughput(batch, steps, warmup, threads=None, time_budget_s=None):
    """oracle/resnet_step.py (torchvision resnet18 + CrossEntropyLoss + SGD, fp32, the reference's own objects) timed on
    the host cores; `batch` images per step, augmentation included."""
    import torch

    from b200step.convnet import draw_augmentation, make_synthetic_images
    from oracle.resnet_step import ResNetOracle

    if threads:
        torch.set_num_threads(threads)
    ds = make_synthetic_images(batch * 2, seed=0)
    rows = torch.arange(batch * 2).repeat(steps + warmup + 1)
    torch.manual_seed(0)
    dr = draw_augmentation(batch * 2).repeat(steps + warmup + 1, 1)
    tr = ResNetOracle("resnet18", ds.images, ds.labels, [rows], batch_size=batch, rank_draws=[dr])
    for _ in range(warmup):
        tr.step()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        tr.step()
        times.append(time.perf_counter() - t0)
        if time_budget_s is not None and sum(times) > time_budget_s:
            break
    n = len(times)
    return {"samples_per_s": batch * n / sum(times), "ms_per_step": 1e3 * sum(times) / n,
            "cores": torch.get_num_threads(), "steps": n}


def run_reference(args, wl):
    """`--impl reference` for the ResNet workloads: the oracle IS the reference's model / criterion / optimizer objects
    on the CPU (kind "reference-objects port": task.py itself needs CUDA and the network)."""
    steps, warm = max(1, args.steps), max(1, min(args.warmup, 2))
    sample = min(wl["batch"], 256)
    r = cpu_oracle_throughput(sample, steps, warm, threads=os.cpu_count(), time_budget_s=60.0)
    out = {
        "impl": "reference", "metric": "train samples/sec", "value": r["samples_per_s"], "unit": "samples/s",
        "n_gpus": args.gpus, "steps": r["steps"], "warmup": warm, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "arch": "resnet18", "per_gpu_batch": wl["batch"], "optimizer": "sgd",
                   "arm": "ONE process, all host threads: torchvision resnet18 + nn.CrossEntropyLoss + torch.optim.SGD "
                          "fed by the restated transform (oracle/resnet_step.py)"},
        "cpu_baseline": {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{r['steps']} steps of {sample} images (a bounded sample of the {wl['batch']}-image "
                                   f"step), fp32 torch CPU, {r['cores']} threads"},
        "e2e": {"value": r["samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def parity_check(rank, world, dev, comm, dist):
    """Two steps of ResNet-18 (32 images per rank, fp32 mode) through the same ConvTrainStep / Communicator / bucket
    schedule against ResNetOracle(world ranks) on rank 0; ranks must hold bit-identical parameters."""
    import torch

    from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout
    from b200step.engine import StepConfig
    from b200step.sampler import rank_indices

    B, rows = 32, 256 * world
    ds = make_synthetic_images(rows, seed=1)
    rank_rows = [rank_indices(rows, world, r) for r in range(world)]
    torch.manual_seed(5)
    rank_draws = [draw_augmentation(len(rr)) for rr in rank_rows]
    cfg = StepConfig(batch_size=B, world=world, rank=rank, precision="fp32", lr=0.01)
    eng = ConvTrainStep(resnet_layout(), init_state(), ds, rank_rows[rank], cfg, comm, device=dev)
    eng.set_draws(rank_draws[rank])
    lg = []
    for _ in range(2):
        eng.step()
        lg.append(eng.loss_value())
    eng.synchronize()
    flat = eng.p.clone()
    same = True
    if dist is not None:
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        ok = torch.tensor([int(torch.equal(ref, flat))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        same = bool(ok.item())
    sd = eng.state_dict()
    eng.close()
    if rank != 0:
        return None
    from oracle.resnet_step import ResNetOracle

    oracle = ResNetOracle("resnet18", ds.images, ds.labels, rank_rows, batch_size=B, lr=0.01, rank_draws=rank_draws)
    lo = [oracle.step()[0] for _ in range(2)]
    so = oracle.state_dict()
    rel = [abs(a - b) / max(1.0, abs(a)) for a, b in zip(lo, lg)]
    init = {"module." + k: v for k, v in init_state().items()}
    keys = [k for k in so if so[k].dtype.is_floating_point and "running" not in k]
    uo = torch.cat([(so[k] - init[k]).double().flatten() for k in keys])
    ug = torch.cat([(sd[k] - init[k]).double().flatten() for k in keys])
    cos = torch.nn.functional.cosine_similarity(uo, ug, dim=0).item()
    upd_rel = ((ug - uo).norm() / uo.norm()).item()
    # bars: step-1 loss = the forward pass (1e-5); step-2 loss = forward of the once-updated weights, and the update
    # itself, with room for single ReLU sign flips (each is ~1 % of a layer4 gradient: profiles/
    # r02_resnet_parity_conditioning.md)
    tol = {"loss_rel_step1": 1e-5, "loss_rel_step2": 1e-3, "update_cosine_min": 0.97}
    ok = rel[0] <= tol["loss_rel_step1"] and rel[1] <= tol["loss_rel_step2"] and cos >= tol["update_cosine_min"] and same
    return {"ok": bool(ok), "loss_rel": rel, "update_rel_l2": upd_rel, "update_cosine": cos,
            "ranks_bit_identical": same, "steps": 2, "world": world, "gradient_buckets": len(eng.buckets),
            "model": "resnet18, 32 images per rank, fp32 (3xTF32), SGD", "against": "ResNetOracle(world ranks) on rank "
            "0 (oracle/resnet_step.py: torchvision model, criterion, optimizer)", "tolerance": tol}


def profile_kernels(eng, B):
    """CUDA events around every launch of a few eager steps -> per kernel class: mean time, launches per step and the
    algorithmic FLOPs (GEMMs: 2*M*N*K, unpadded K) or bytes (everything else: every operand read or written once)."""
    import torch

    from b200step import ops

    es = 2 if eng.dt == torch.bfloat16 else 4
    x3 = 3 if eng.x3 else 1
    names = ["image_augment", "im2col", "col2im", "bn_fwd_stats", "bn_fwd_apply", "bn_bwd_reduce", "bn_bwd_apply", "bn_fwd", "bn_bwd",
             "maxpool_fwd", "maxpool_bwd", "avgpool_fwd", "avgpool_bwd", "linear_fwd", "linear_ce", "linear_dgrad",
             "linear_wgrad", "conv_fwd", "conv_wgrad", "conv_dgrad", "sgd_step", "adam_step", "split_tf32"]
    orig = {n: getattr(ops, n) for n in names}
    records = []

    def work_of(name, a, k):
        if name in ("linear_fwd", "linear_ce"):
            return "gemm_fwd", {"flops": 2.0 * k["M"] * k["N"] * (k["K"] // x3)}
        if name == "linear_dgrad":
            return "gemm_dgrad", {"flops": 2.0 * k["M"] * (k["N"] // x3) * k["K"]}
        if name == "linear_wgrad":
            return "gemm_wgrad", {"flops": 2.0 * (k["M"] // x3) * k["N"] * k["K"]}
        if name in ("conv_fwd", "conv_wgrad"):
            Ho = ops.conv_out_size(k["H"], k["kh"], k["stride"], k["pad"])
            Wo = ops.conv_out_size(k["W"], k["kw"], k["stride"], k["pad"])
            return ("gemm_fwd" if name == "conv_fwd" else "gemm_wgrad"), \
                {"flops": 2.0 * k["B"] * Ho * Wo * k["Cout"] * k["kh"] * k["kw"] * k["C"]}
        if name == "conv_dgrad":
            return "gemm_dgrad", {"flops": 2.0 * k["B"] * k["H"] * k["W"] * k["C"] * k["kh"] * k["kw"] * k["Cout"]}
        if name == "image_augment":
            return name, {"bytes": k["batch"] * k["H"] * k["W"] * 3 * (1 + es)}
        if name == "im2col":
            Ho, Wo = ops.conv_out_size(k["H"], k["kh"], k["stride"], k["pad"]), ops.conv_out_size(k["W"], k["kw"], k["stride"], k["pad"])
            return name, {"bytes": es * (k["B"] * k["H"] * k["W"] * k["C"] + k["B"] * Ho * Wo * a[1].shape[1])}
        if name == "col2im":
            return name, {"bytes": es * (a[0].shape[0] * a[0].shape[1] + k["B"] * k["H"] * k["W"] * k["C"] * (2 if a[1] is not None else 1))}
        if name == "bn_fwd":       # y read (twice, the second time from cache), out written, residual read
            return name, {"bytes": es * k["M"] * k["C"] * (3 if a[1] is not None else 2)}
        if name == "bn_bwd":       # dout (+dout2), out, y read; dy (+dz) written
            ins = 2 + (a[1] is not None) + (k.get("dout2") is not None)
            return name, {"bytes": es * k["M"] * k["C"] * (ins + (2 if a[4] is not None else 1))}
        if name == "bn_fwd_stats":
            return name, {"bytes": es * k["M"] * k["C"]}
        if name == "bn_fwd_apply":
            return name, {"bytes": es * k["M"] * k["C"] * (3 if a[1] is not None else 2)}
        if name == "bn_bwd_reduce":
            return name, {"bytes": es * k["M"] * k["C"] * (3 if a[1] is not None else 2)}
        if name == "bn_bwd_apply":
            return name, {"bytes": es * k["M"] * k["C"] * ((3 if a[1] is not None else 2) + (2 if a[4] is not None else 1))}
        if name.startswith("maxpool") or name.startswith("avgpool"):
            return name, {"bytes": sum(t.numel() * t.element_size() for t in a if t is not None)}
        if name == "split_tf32":
            outs = (a[1] is not None) + (a[3] is not None)
            return name, {"bytes": 4.0 * k["rows"] * k["cols"] * (1 + 3 * outs)}
        if name == "sgd_step":
            return name, {"bytes": a[0].numel() * 24 + 2 * k.get("n_shadow", 0)}
        return name, {"bytes": a[0].numel() * 32 + 2 * k.get("n_shadow", 0)}

    def wrap(name, fn):
        def inner(*a, **k):
            st = k.get("stream") or eng.main
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            r = fn(*a, **k)
            e1.record(st)
            label, work = work_of(name, a, k)
            records.append((label, e0, e1, work))
            return r
        return inner

    reps = 3
    try:
        for n in names:
            setattr(ops, n, wrap(n, orig[n]))
        for _ in range(reps):
            eng._enqueue(B, eng.row_idx, eng.cursor, eng._draws(), True)
        eng.synchronize()
    finally:
        for n in names:
            setattr(ops, n, orig[n])
    agg = {}
    for label, e0, e1, work in records:
        d = agg.setdefault(label, {"ms": 0.0, "launches": 0, "flops": 0.0, "bytes": 0.0})
        d["ms"] += e0.elapsed_time(e1)
        d["launches"] += 1
        d["flops"] += work.get("flops", 0.0)
        d["bytes"] += work.get("bytes", 0.0)
    out = []
    for label, d in agg.items():
        rec = {"name": label, "ms_per_step": d["ms"] / reps, "per_step": d["launches"] // reps,
               "ms": d["ms"] / d["launches"]}
        if d["flops"]:
            rec.update(bound="tensor", tflops=d["flops"] / (d["ms"] / 1e3) / 1e12, flops_per_step=d["flops"] / reps)
        else:
            rec.update(bound="hbm", gbs=d["bytes"] / (d["ms"] / 1e3) / 1e9, bytes_per_step=d["bytes"] / reps)
        out.append(rec)
    return out


def run_b200(args, rank, local_rank, world, wl, peaks, ClockSampler, measure_tf32_peak):
    import torch

    from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout
    from b200step.engine import Communicator, StepConfig

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: the training step has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    B = wl["batch"]
    comm = dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group(backend="nccl", init_method="env://", world_size=world, rank=rank, device_id=dev)
        uid = [Communicator.make_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        comm = Communicator(rank, world, uid[0])
    parity = None if args.no_parity else parity_check(rank, world, dev, comm, dist)

    lay = resnet_layout("resnet18")
    rows = B * BATCHES_PER_RANK                                   # this rank's images (weak scaling), cycled
    ds = make_synthetic_images(rows, seed=100 + rank)
    cfg = StepConfig(batch_size=B, world=world, rank=rank, precision=wl["precision"], optimizer="sgd")
    eng = ConvTrainStep(lay, init_state("resnet18", 0), ds, torch.arange(rows), cfg, comm, device=dev)
    torch.manual_seed(rank)
    draws = draw_augmentation(rows)
    eng.set_draws(draws)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(n, fn):
        sampler = ClockSampler(local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sampler.start()
        e0.record(eng.main)
        fn(n)
        e1.record(eng.main)
        eng.synchronize()
        barrier()
        clocks = sampler.stop()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks

    def resident(n):
        for _ in range(n):
            eng.step()

    for _ in range(max(args.warmup, 3)):
        eng.step()
    eng.synchronize()
    ms, clocks = timed(args.steps, resident)
    loss_after = eng.loss_value()
    value = world * B * args.steps / (ms / 1e3)
    sustained = None
    if args.sustain_seconds > 0:
        n_sus = max(args.steps, int(args.sustain_seconds * 1e3 / (ms / args.steps)) + 1)
        sus_ms, clocks2 = timed(n_sus, resident)
        sustained = {"seconds": sus_ms / 1e3, "steps": n_sus, "ms_per_step": sus_ms / n_sus,
                     "value": world * B * n_sus / (sus_ms / 1e3), "unit": "samples/s", "clocks": clocks2}

    # ---- end to end: uint8 images + labels + draws in pinned host memory, H2D and loss read-back every step ---------
    nb = 4
    host = [(ds.images[i * B:(i + 1) * B].pin_memory(), ds.labels[i * B:(i + 1) * B].pin_memory(),
             draws[i * B:(i + 1) * B].pin_memory()) for i in range(nb)]
    for i in range(4):
        eng.host_loss(eng.step_from_host(*host[i % nb]))
    eng.synchronize()
    e2e_steps = max(3, min(args.steps, 50))

    def hostfed(n):
        pending = []
        for i in range(n):
            pending.append(eng.step_from_host(*host[i % nb]))
            if len(pending) > 2:
                eng.host_loss(pending.pop(0))
        while pending:
            eng.host_loss(pending.pop(0))

    e2e_ms, _ = timed(e2e_steps, hostfed)
    e2e = {"value": world * B * e2e_steps / (e2e_ms / 1e3), "unit": "samples/s",
           "h2d_bytes_per_step": eng.host_batch_bytes(), "d2h_bytes_per_step": 4, "steps": e2e_steps}

    prof = profile_kernels(eng, B)
    barrier()
    launches = eng.kernels_per_step
    buckets = len(eng.buckets)
    eng.close()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    step_ms = ms / args.steps
    total_ms = sum(r["ms_per_step"] for r in prof)
    for r in prof:
        r["share"] = r["ms_per_step"] / total_ms
        r["share_of_step_ms"] = r["ms_per_step"] / step_ms
    gem = [r for r in prof if r["bound"] == "tensor"]
    mem = [r for r in prof if r["bound"] == "hbm"]
    gem_ms = sum(r["ms_per_step"] for r in gem)
    mem_ms = sum(r["ms_per_step"] for r in mem)
    flops = resnet_flops_per_sample(lay)
    if wl["precision"] == "bf16":
        tensor_peak, peak_note = peaks["bf16_tflops"], peaks["source"] + " bf16 burst"
    else:
        tf32 = measure_tf32_peak(dev)
        passes = 3.0 if wl["precision"] == "fp32" else 1.0
        tensor_peak = tf32 / passes
        peak_note = f"TF32 peak measured in this run: {tf32:.1f} TFLOP/s" + (" / 3 passes" if passes > 1 else "")
    gem_tflops = flops["total"] * B / (gem_ms / 1e3) / 1e12
    # DRAM bytes per launch from the ncu pass of scripts/gpu_ncu_resnet_traffic.sh (profiles/ncu_traffic.json)
    gem_traffic = mem_traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as fh:
            blob = json.load(fh)
        tags = sorted(k for k in blob if k.startswith(args.workload + "@"))
        if tags:
            tab = {k: v for k, v in blob[tags[-1]].items() if not k.startswith("_")}
            g = [v for k, v in tab.items() if k in ("gemm_kernel", "linear_ce_kernel")]
            m = [v for k, v in tab.items() if k not in ("gemm_kernel", "linear_ce_kernel", "reduce_partials_kernel")]
            gem_traffic = sum(v["dram_bytes"] for v in g) / max(1, sum(v["launches"] for v in g))
            mem_traffic = sum(v["dram_bytes"] for v in m) / max(1, sum(v["launches"] for v in m))
    roof_tensor = {"bound": "tensor", "kernel": "gemm_kernel / linear_ce_kernel (all conv + fc GEMM launches of the step)",
                   "achieved": gem_tflops, "peak": tensor_peak, "unit": "TFLOP/s", "frac": gem_tflops / tensor_peak,
                   "traffic": gem_traffic, "traffic_note": "mean DRAM bytes per GEMM launch of one step (ncu)",
                   "share_of_step": gem_ms / total_ms, "share_of_step_ms": gem_ms / step_ms,
                   "kernel_ms": gem_ms, "launches_per_step": sum(r["per_step"] for r in gem),
                   "flops_basis": "2*Ho*Wo*Cout*kh*kw*Cin per conv, fwd + dgrad + wgrad (conv1: no dgrad)",
                   "peak_source": peak_note}
    mem_bytes = sum(r["bytes_per_step"] for r in mem)
    mem_gbs = mem_bytes / (mem_ms / 1e3) / 1e9
    roof_hbm = {"bound": "hbm", "kernel": "conv.cu kernels (im2col, col2im, BatchNorm, pooling) + K-opt, all launches of "
                "the step", "achieved": mem_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": mem_gbs / peaks["hbm_gbs"], "traffic": mem_traffic,
                "traffic_note": "mean DRAM bytes per launch of these kernels (ncu); below the algorithmic bytes because "
                                "im2col matrices and activations are re-read from the 126 MB L2",
                "share_of_step": mem_ms / total_ms,
                "share_of_step_ms": mem_ms / step_ms, "kernel_ms": mem_ms,
                "launches_per_step": sum(r["per_step"] for r in mem),
                "bytes_basis": "every operand of every launch read or written once", "peak_source": peaks["source"] +
                " copy bandwidth"}
    out = {
        "metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": {"bf16": "bf16", "tf32": "tf32", "fp32": "f32"}[wl["precision"]],
        "data": "synthetic",
        "config": {"workload": args.workload, "arch": "resnet18", "image": "3x32x32 uint8 (CIFAR-10 shape)",
                   "per_gpu_batch": B, "global_batch": B * world, "optimizer": "sgd", "params": 11689512,
                   "parallelism": f"dp{world}", "gradient_buckets": buckets, "cuda_graph": True,
                   "l2": "activations + im2col matrices of one step exceed the 126 MB L2 (no flush needed)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches * args.steps,
        "roofline": roof_tensor if gem_ms >= mem_ms else roof_hbm,
        "roofline_tensor": roof_tensor, "roofline_hbm": roof_hbm, "kernels": prof, "kernel_ms_sum": total_ms,
        "step_tflops_per_gpu": flops["total"] * B / (step_ms / 1e3) / 1e12, "final_loss": loss_after,
    }
    if sustained is not None:
        out["sustained"] = sustained
    if parity is not None:
        out["parity"] = parity
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_oracle_throughput(256, 4, 1)
        out["cpu_baseline"] = {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["cores"], "kind": "port",
                               "sample": f"{r['steps']} steps of 256 images after 1 warm-up: torchvision resnet18 + "
                                         f"CrossEntropyLoss + SGD on the CPU (fp32), {r['cores']} threads"}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"parity check failed: {json.dumps(parity)}")
