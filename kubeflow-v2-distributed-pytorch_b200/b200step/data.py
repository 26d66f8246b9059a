"""Synthetic tabular dataset ("criteo-shape") standing in for the reference's CIFAR-10 download.

The reference trains on `torchvision.datasets.CIFAR10(..., download=True)` (task.py:256), which needs the network;
BASELINE.json's configs ask for "synthetic tabular data" instead.  Everything here is generated on the CPU from a
fixed `torch.Generator` so that the GPU step and the CPU oracle see byte-identical inputs:

  dense  fp32 [N, num_dense]   column j ~ N(mu_j = j, sigma_j = 1 + j/4)
  mean / std                   the exact fp32 column statistics of `dense` (the tabular twin of the constants
                               in transforms.Normalize at task.py:251)
  cat    int64 [N, num_cat]    uniform in [0, cardinality)  (or Zipf(alpha) when `zipf_alpha` is given)
  label  int64 [N]             argmax(standardised dense @ T), T fixed [num_dense, 10] from seed 4321
                               -> 10 classes like CIFAR-10 under a 1000-way head (task.py:171), and learnable.
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch

from .model import ModelSpec

DATA_SEED = 1234
TEACHER_SEED = 4321
NUM_LABEL_CLASSES = 10


@dataclasses.dataclass
class TabularDataset:
    dense: torch.Tensor   # fp32 [N, num_dense]
    cat: torch.Tensor     # int64 [N, num_cat]
    label: torch.Tensor   # int64 [N]
    mean: torch.Tensor    # fp32 [num_dense]
    std: torch.Tensor     # fp32 [num_dense]

    def __len__(self) -> int:
        return self.dense.shape[0]

    def to(self, device) -> "TabularDataset":
        return TabularDataset(*(getattr(self, f.name).to(device) for f in dataclasses.fields(self)))

    def pin(self) -> "TabularDataset":
        return TabularDataset(*(getattr(self, f.name).pin_memory() for f in dataclasses.fields(self)))


def make_synthetic(num_rows: int, spec: ModelSpec, seed: int = DATA_SEED,
                   zipf_alpha: Optional[float] = None) -> TabularDataset:
    g = torch.Generator(device="cpu").manual_seed(seed)
    nd, nc = spec.num_dense, spec.num_cat
    j = torch.arange(nd, dtype=torch.float32)
    dense = torch.randn(num_rows, nd, generator=g, dtype=torch.float32) * (1.0 + j / 4.0) + j
    if zipf_alpha is None:
        cat = torch.randint(0, spec.cardinality, (num_rows, nc), generator=g, dtype=torch.int64)
    else:
        ranks = torch.arange(1, spec.cardinality + 1, dtype=torch.float64)
        probs = ranks.pow(-zipf_alpha)
        cat = torch.multinomial(probs / probs.sum(), num_rows * nc, replacement=True, generator=g)
        cat = cat.view(num_rows, nc).to(torch.int64)
    mean = dense.mean(dim=0)
    std = dense.std(dim=0)
    gt = torch.Generator(device="cpu").manual_seed(TEACHER_SEED)
    teacher = torch.randn(nd, NUM_LABEL_CLASSES, generator=gt, dtype=torch.float32)
    label = ((dense - mean) / std).matmul(teacher).argmax(dim=1).to(torch.int64)
    return TabularDataset(dense.contiguous(), cat.contiguous(), label.contiguous(), mean.contiguous(),
                          std.contiguous())
