#!/usr/bin/env python
"""Context row of SURVEY.md §8(d): the SAME training step written with stock PyTorch-eager CUDA ops
(cuBLAS GEMMs, ATen embedding / cross_entropy / SGD kernels, bf16 autocast over fp32 master weights) and, for
N > 1, torch's own DistributedDataParallel over NCCL — "the existing Blackwell library kernels" bar.

Not part of the product and not the oracle: none of this repo's kernels run here.  Data is resident in HBM
and batches are picked by the sampler's index list, exactly as in bench.py's `value` leg; timing is CUDA
events over K steps after W warm-up steps, max over ranks.

    python scripts/gpu_torch_eager_context.py [--arch tabular_mlp] [--precision bf16] [--batch 65536]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 ... scripts/gpu_torch_eager_context.py
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200"))

from b200step.data import make_synthetic  # noqa: E402  (dataset generator only; no kernels)
from b200step.model import ARCHS  # noqa: E402
from b200step.sampler import rank_indices  # noqa: E402


class EagerTabularMLP(nn.Module):
    def __init__(self, spec):
        super().__init__()
        self.emb = nn.ModuleList(nn.Embedding(spec.cardinality, spec.embed_dim) for _ in range(spec.num_cat))
        self.layers = nn.ModuleList(nn.Linear(k, o) for k, o in spec.layer_dims)

    def forward(self, dense_std, cat):
        x = torch.cat([dense_std] + [e(cat[:, t]) for t, e in enumerate(self.emb)], dim=1)
        for li, lin in enumerate(self.layers):
            x = lin(x)
            if li + 1 < len(self.layers):
                x = F.relu(x)
        return x


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="tabular_mlp")
    ap.add_argument("--precision", choices=["bf16", "fp32", "tf32"], default="bf16")
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", init_method="env://", world_size=world, rank=rank, device_id=dev)
    if args.precision == "tf32":
        torch.backends.cuda.matmul.allow_tf32 = True
    spec = ARCHS[args.arch]
    B = args.batch
    rows = world * B * 8
    ds = make_synthetic(rows, spec)
    idx = rank_indices(rows, world, rank)
    dense, cat, label = ds.dense[idx].to(dev), ds.cat[idx].to(dev), ds.label[idx].to(dev)
    mean, std = ds.mean.to(dev), ds.std.to(dev)
    order = torch.arange(idx.numel(), device=dev)
    torch.manual_seed(0)
    model = EagerTabularMLP(spec).to(dev)
    net = model
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel

        net = DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    crit = nn.CrossEntropyLoss()
    nb = idx.numel() // B

    def step(i):
        sel = order[(i % nb) * B:(i % nb + 1) * B]
        x = (dense.index_select(0, sel) - mean) / std
        c = cat.index_select(0, sel)
        y = label.index_select(0, sel)
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=args.precision == "bf16"):
            out = net(x, c)
        loss = crit(out.float(), y)
        loss.backward()
        opt.step()
        return loss

    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        print(json.dumps({
            "impl": "torch-eager-context", "metric": "train samples/sec", "value": world * B * args.steps / (ms / 1e3),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "ms_per_step": ms / args.steps,
            "dtype": args.precision, "config": {"arch": args.arch, "per_gpu_batch": B, "optimizer": "sgd",
                                                "torch": torch.__version__,
                                                "ddp": world > 1, "note": "stock ATen/cuBLAS kernels, no CUDA graph"},
            "final_loss": float(loss.item())}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
