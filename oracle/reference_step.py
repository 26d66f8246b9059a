"""CPU oracle: PyTorch restatement of the reference's data-parallel training step.

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs, never by the product path (see oracle/__init__.py).

Why a restatement: the reference's task.py cannot run without a GPU and the network (every branch calls
.cuda(): task.py:179,191,198,205,207,211,305,307; CIFAR10(download=True): task.py:256) and it ships no tests or
golden vectors, so the reference leaves parity UNPINNED.  The pin used instead: this file makes
the *same torch calls in the same order* with `.cuda()` removed, on the installed torch (2.11.0) — the library
the reference's arithmetic lives in — and tests/golden/ holds vectors generated from it
(tests/golden/make_golden.py).  The tabular model, Adam, dropout and bf16 have no counterpart in the reference
at all; for those this file *is* the definition, anchored on torch's own modules/optimizers.

What follows the reference line by line:
  set_random_seeds            task.py:22-28, called at task.py:162
  model built under the seed  task.py:166-171   (here: TabularMLP instead of torchvision resnet18)
  DDP mean-gradient semantics task.py:189/194   (emulated exactly in one process, or run for real over gloo)
  nn.CrossEntropyLoss()       task.py:211, used at task.py:310
  torch.optim.SGD(lr, momentum, weight_decay)   task.py:212-214
  Normalize -> sub(mean).div(std)               task.py:250-251 (torchvision _functional_tensor.py:928)
  DistributedSampler + DataLoader batching      task.py:260, 263
  step order zero_grad -> forward -> loss -> backward -> step   task.py:308-312
"""
from __future__ import annotations

import os
import random
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.data import DataLoader, Dataset
from torch.utils.data.distributed import DistributedSampler

from .philox import dropout_keep_mask


# ---------------------------------------------------------------------------------------------- seeding
def set_random_seeds(random_seed: int = 0) -> None:
    """task.py:22-28."""
    torch.manual_seed(random_seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    np.random.seed(random_seed)
    random.seed(random_seed)


# ------------------------------------------------------------------------------------------------ model
class TabularMLP(nn.Module):
    """Embeddings for the categorical columns, concatenated behind the (already standardised) dense columns,
    then Linear/ReLU(/dropout) blocks and a Linear head.  Module construction order = parameter order =
    RNG consumption order: embeddings 0..num_cat-1, then layers 0..L-1."""

    def __init__(self, num_dense: int, num_cat: int, cardinality: int, embed_dim: int,
                 hidden: Sequence[int], num_classes: int, dropout: float = 0.0, **_unused):
        super().__init__()
        self.num_dense, self.num_cat, self.embed_dim = num_dense, num_cat, embed_dim
        self.dropout = float(dropout)
        self.emb = nn.ModuleList([nn.Embedding(cardinality, embed_dim) for _ in range(num_cat)])
        dims = [num_dense + num_cat * embed_dim, *hidden, num_classes]
        self.layers = nn.ModuleList([nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1)])

    def features(self, dense_std: torch.Tensor, cat: torch.Tensor) -> torch.Tensor:
        cols = [dense_std] + [self.emb[t](cat[:, t]) for t in range(self.num_cat)]
        return torch.cat(cols, dim=1)

    def forward(self, dense_std: torch.Tensor, cat: torch.Tensor,
                keep_masks: Optional[List[torch.Tensor]] = None) -> torch.Tensor:
        x = self.features(dense_std, cat)
        last = len(self.layers) - 1
        for i, lin in enumerate(self.layers):
            x = lin(x)
            if i < last:
                x = F.relu(x)
                if keep_masks is not None:
                    inv_keep = torch.tensor(1.0, dtype=torch.float32) / (1.0 - torch.tensor(self.dropout, dtype=torch.float32))
                    x = torch.where(keep_masks[i], x * inv_keep, torch.zeros((), dtype=x.dtype))
        return x


def build_model(spec: Dict, seed: int) -> TabularMLP:
    """task.py:162 then task.py:171: seed, then construct."""
    set_random_seeds(seed)
    return TabularMLP(**spec)


def standardise(dense: torch.Tensor, mean: torch.Tensor, std: torch.Tensor) -> torch.Tensor:
    """transforms.Normalize (task.py:251) == tensor.sub_(mean).div_(std): subtract, then true divide, fp32."""
    return dense.clone().sub_(mean).div_(std)


# ------------------------------------------------------------------------------------- sampler / loader
class _IndexDataset(Dataset):
    def __init__(self, n: int):
        self.n = n

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int) -> int:
        return i


def rank_batches(num_rows: int, world: int, rank: int, batch_size: int) -> List[torch.Tensor]:
    """Row indices of every batch rank `rank` sees in one epoch: the reference's own classes
    (DistributedSampler defaults, task.py:260; DataLoader(batch_size, sampler), task.py:263; no set_epoch)."""
    ds = _IndexDataset(num_rows)
    sampler = DistributedSampler(dataset=ds, num_replicas=world, rank=rank)
    loader = DataLoader(dataset=ds, batch_size=batch_size, sampler=sampler, num_workers=0)
    return [b.clone() for b in loader]


# --------------------------------------------------------------------------------- manual fwd/bwd (bf16)
def _q(t: torch.Tensor, precision: str) -> torch.Tensor:
    """Storage rounding of an activation/gradient tensor: identity for fp32, RNE to bf16 and back for bf16."""
    return t if precision == "fp32" else t.to(torch.bfloat16).to(torch.float32)


def manual_forward_backward(model: TabularMLP, dense_std: torch.Tensor, cat: torch.Tensor, labels: torch.Tensor,
                            grad_scale: float, precision: str,
                            keep_masks: Optional[List[torch.Tensor]] = None, logits_rounded: bool = False,
                            dx0_rounded: bool = False) -> Tuple[torch.Tensor, Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """One forward + backward written out op by op, with the storage rounding points of the bf16 GPU path:
    hidden activations and activation-gradients are stored in bf16, accumulation is fp32, weights enter the
    GEMMs as bf16 copies of the fp32 masters, parameter gradients are fp32.  The logits are rounded to bf16 only
    when `logits_rounded` (the GPU path that stores them: output layer and loss as two kernels); by default the
    output layer is fused with the loss, the logits stay in fp32 on chip and are never rounded.  Likewise the
    gradient with respect to the model input is rounded only when `dx0_rounded` (the path that stores it for a
    separate embedding-backward kernel); by default the first layer's dgrad adds it onto the table gradients in
    fp32.  With
    precision='fp32' it must agree with autograd (checked in tests/test_oracle.py), which is what ties this
    restatement to torch's own backward.

    Returns (loss, grads keyed like state_dict, intermediates)."""
    L = len(model.layers)
    x0 = _q(model.features(dense_std, cat).detach(), precision)
    acts = [x0]
    inv_keep = float(torch.tensor(1.0, dtype=torch.float32) / (1.0 - torch.tensor(model.dropout, dtype=torch.float32)))
    wq = [_q(l.weight.detach(), precision) for l in model.layers]
    for i, lin in enumerate(model.layers):
        z = acts[-1] @ wq[i].t() + lin.bias.detach()
        if i < L - 1:
            z = z.clamp_min(0)
            if keep_masks is not None:
                z = torch.where(keep_masks[i], z * inv_keep, torch.zeros(()))
        acts.append(_q(z, precision) if (i < L - 1 or logits_rounded) else z)
    logits = acts[-1]
    lse = torch.logsumexp(logits, dim=1)
    loss = (lse - logits.gather(1, labels[:, None])[:, 0]).mean()
    p = torch.softmax(logits, dim=1)
    onehot = F.one_hot(labels, logits.shape[1]).to(torch.float32)
    dy = _q((p - onehot) * grad_scale, precision)
    grads: Dict[str, torch.Tensor] = {}
    inter = {"x0": x0, "logits": logits, "dlogits": dy}
    for i in range(L - 1, -1, -1):
        grads[f"layers.{i}.weight"] = dy.t() @ acts[i]
        grads[f"layers.{i}.bias"] = dy.sum(0)
        dx = dy @ wq[i]
        if i > 0:
            if keep_masks is not None:
                dx = dx * inv_keep
            dx = torch.where(acts[i] > 0, dx, torch.zeros(()))
        dy = _q(dx, precision) if (i > 0 or dx0_rounded) else dx
    inter["dx0"] = dy
    nd, d = model.num_dense, model.embed_dim
    for t in range(model.num_cat):
        g = torch.zeros_like(model.emb[t].weight)
        g.index_add_(0, cat[:, t], dy[:, nd + t * d: nd + (t + 1) * d])
        grads[f"emb.{t}.weight"] = g
    return loss, grads, inter


# ------------------------------------------------------------------------------------------ the trainer
class OracleTrainer:
    """The reference's epoch loop body (task.py:303-312) for `world` data-parallel ranks, in ONE process.

    Each step: every rank runs zero_grad/forward/loss/backward on its own batch; DistributedDataParallel
    (task.py:189) then leaves the MEAN of the ranks' gradients in every .grad; optimizer.step().  Averaging the
    per-rank gradients here is that same arithmetic (tests/test_oracle.py checks it against real DDP over gloo).
    """

    def __init__(self, spec: Dict, dataset, *, world: int = 1, batch_size: int, seed: int = 0,
                 lr: float = 0.1, momentum: float = 0.9, weight_decay: float = 1e-4, optimizer: str = "sgd",
                 precision: str = "fp32", dropout_seed: int = 0, use_autograd: Optional[bool] = None,
                 logits_rounded: bool = False, dx0_rounded: bool = False):
        assert precision in ("fp32", "bf16")
        self.logits_rounded = logits_rounded   # bf16 only: True = logits stored (and rounded) between layer and loss
        self.dx0_rounded = dx0_rounded         # bf16 only: True = input gradient stored (and rounded) before the scatter
        self.spec, self.ds, self.world, self.batch_size = dict(spec), dataset, world, batch_size
        self.precision = precision
        self.model = build_model(self.spec, seed)                                   # task.py:162,171
        self.criterion = nn.CrossEntropyLoss()                                      # task.py:211
        if optimizer == "sgd":                                                      # task.py:212-214
            self.optimizer = torch.optim.SGD(self.model.parameters(), lr, momentum=momentum,
                                             weight_decay=weight_decay)
        elif optimizer == "adam":
            self.optimizer = torch.optim.Adam(self.model.parameters(), lr, weight_decay=weight_decay)
        else:
            raise ValueError(optimizer)
        self.use_autograd = (precision == "fp32") if use_autograd is None else use_autograd
        self.dropout_seed = dropout_seed
        self.batches = [rank_batches(len(dataset), world, r, batch_size) for r in range(world)]  # task.py:260-263
        self.steps_per_epoch = len(self.batches[0])
        self.step_count = 0
        self.model.train()                                                          # task.py:301

    def _keep_masks(self, rows: int) -> Optional[List[torch.Tensor]]:
        p = float(self.spec.get("dropout", 0.0))
        if p <= 0.0:
            return None
        hidden = list(self.spec["hidden"])
        return [torch.from_numpy(dropout_keep_mask(rows, h, p, self.dropout_seed, self.step_count, li))
                for li, h in enumerate(hidden)]

    def rank_loss_and_grads(self, rows: torch.Tensor):
        dense_std = standardise(self.ds.dense[rows], self.ds.mean, self.ds.std)     # task.py:250-251
        cat, labels = self.ds.cat[rows], self.ds.label[rows]
        masks = self._keep_masks(rows.numel())
        if self.use_autograd:
            self.optimizer.zero_grad()                                              # task.py:308
            outputs = self.model(dense_std, cat, masks)                             # task.py:309
            loss = self.criterion(outputs, labels)                                  # task.py:310
            loss.backward()                                                         # task.py:311
            grads = {k: p.grad.detach().clone() for k, p in self.model.named_parameters()}
            return loss.detach(), grads
        loss, grads, _ = manual_forward_backward(self.model, dense_std, cat, labels, 1.0 / rows.numel(),
                                                 self.precision, masks, self.logits_rounded, self.dx0_rounded)
        return loss, grads

    def step(self) -> List[float]:
        """One optimizer step; returns each rank's local loss (what `loss` holds on that rank at task.py:310)."""
        b = self.step_count % self.steps_per_epoch
        losses, acc = [], None
        for r in range(self.world):
            loss, grads = self.rank_loss_and_grads(self.batches[r][b])
            losses.append(float(loss))
            if acc is None:
                acc = {k: g.clone() for k, g in grads.items()}
            else:
                for k, g in grads.items():
                    acc[k] += g
        for k, p in self.model.named_parameters():
            p.grad = acc[k] / self.world if self.world > 1 else acc[k]              # DDP mean, task.py:189
        self.optimizer.step()                                                       # task.py:312
        self.step_count += 1
        return losses

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Keys as task.py:283/287 would save them: the DDP wrapper adds 'module.'."""
        return {"module." + k: v.detach().clone() for k, v in self.model.state_dict().items()}


# --------------------------------------------------------------------- real DDP over gloo (config 1)
def _ddp_worker(rank: int, world: int, port: int, spec: Dict, ds, batch_size: int, steps: int, seed: int,
                hyper: Dict, threads: int, out_path: str, time_it: bool) -> None:
    import time

    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if threads > 0:
        torch.set_num_threads(threads)
    dist.init_process_group(backend="gloo", init_method="env://", world_size=world, rank=rank)  # task.py:148
    model = build_model(spec, seed)                                                  # task.py:162,171
    ddp = torch.nn.parallel.DistributedDataParallel(model)                           # task.py:194
    criterion = nn.CrossEntropyLoss()                                                # task.py:211
    optimizer = torch.optim.SGD(ddp.parameters(), hyper["lr"], momentum=hyper["momentum"],
                                weight_decay=hyper["weight_decay"])                  # task.py:212-214
    batches = rank_batches(len(ds), world, rank, batch_size)
    ddp.train()
    losses, times = [], []
    for s in range(steps):
        rows = batches[s % len(batches)]
        t0 = time.perf_counter()
        dense_std = standardise(ds.dense[rows], ds.mean, ds.std)
        optimizer.zero_grad()
        outputs = ddp(dense_std, ds.cat[rows])
        loss = criterion(outputs, ds.label[rows])
        loss.backward()
        optimizer.step()
        times.append(time.perf_counter() - t0)
        losses.append(float(loss))
    if rank == 0 or not time_it:
        torch.save({"losses": losses, "times": times,
                    "state": {"module." + k: v.detach().clone() for k, v in model.state_dict().items()}},
                   f"{out_path}.rank{rank}")
    dist.barrier()
    dist.destroy_process_group()


def run_ddp_gloo(spec: Dict, ds, *, world: int, batch_size: int, steps: int, seed: int = 0,
                 lr: float = 0.1, momentum: float = 0.9, weight_decay: float = 1e-4, threads: int = 0,
                 port: int = 29533, out_path: str = "/tmp/oracle_ddp", time_it: bool = False) -> List[Dict]:
    """BASELINE config 1: the reference's DDP path on CPU with the gloo backend, one process per rank
    (task.py:124 mp.spawn -> task.py:148 init_process_group -> task.py:194 DDP)."""
    import torch.multiprocessing as mp

    hyper = {"lr": lr, "momentum": momentum, "weight_decay": weight_decay}
    mp.spawn(_ddp_worker, nprocs=world,
             args=(world, port, spec, ds, batch_size, steps, seed, hyper, threads, out_path, time_it))
    ranks = [0] if time_it else range(world)
    out = [torch.load(f"{out_path}.rank{r}") for r in ranks]
    for r in ranks:
        os.remove(f"{out_path}.rank{r}")
    return out
