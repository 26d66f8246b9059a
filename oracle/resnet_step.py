"""CPU oracle for the reference-literal path: torchvision ResNet-18 on 3x32x32 images (task.py:171,247-257,303-312).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl
reference` legs, never by the product path (see oracle/__init__.py).

Unlike the tabular oracle this one is PINNED to the reference's own objects: the model is
`torchvision.models.__dict__[arch]()` (task.py:171), the input pipeline is torchvision's own
`transforms.Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(), ToTensor(), Normalize(...)])` object built
exactly as at task.py:250-255 and applied to PIL images the way `CIFAR10.__getitem__` does, the criterion is
`nn.CrossEntropyLoss()` (task.py:211), the optimizer `torch.optim.SGD(params, lr, momentum, weight_decay)`
(task.py:212-214), the step order zero_grad -> forward -> loss -> backward -> step (task.py:308-312), all on the
installed torch 2.11 / torchvision 0.26 with `.cuda()` removed.  What is defined here, because CIFAR-10 cannot be
downloaded (task.py:256 needs the network): the images are seeded synthetic uint8 arrays of CIFAR's shape.

DistributedDataParallel (task.py:189) is emulated in one process: every rank's batch goes through the same weights
with ITS OWN BatchNorm batch statistics (DDP does not synchronise them), the rank losses are scaled by 1/world before
backward (DDP's gradient mean), and the BatchNorm buffers that survive are rank 0's (DDP broadcasts rank 0's buffers
before each forward, so every rank's forward starts from them and rank 0 is the process that saves: task.py:283).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)        # task.py:254
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def reference_transform():
    """The transform object of task.py:250-255."""
    from torchvision import transforms

    return transforms.Compose([
        transforms.RandomCrop(32, padding=4),
        transforms.RandomHorizontalFlip(),
        transforms.ToTensor(),
        transforms.Normalize(CIFAR_MEAN, CIFAR_STD),
    ])


def augment_with_torchvision(images_u8: torch.Tensor, seed: int) -> torch.Tensor:
    """Run the reference's transform over uint8 HWC images in order, drawing from torch's global generator seeded with
    `seed` (what a DataLoader with num_workers=0 does after torch.manual_seed).  Returns fp32 [n, 3, 32, 32]."""
    from PIL import Image

    tf = reference_transform()
    torch.manual_seed(seed)
    out = []
    for i in range(images_u8.shape[0]):
        img = Image.fromarray(images_u8[i].numpy())          # CIFAR10.__getitem__: Image.fromarray(img)
        out.append(tf(img))
    return torch.stack(out)


def augment_with_draws(images_u8: torch.Tensor, draws: Optional[torch.Tensor]) -> torch.Tensor:
    """The same pipeline restated on whole tensors with the random numbers given: draws[i] = (top, left, flip, -).
    Pinned against `augment_with_torchvision` in tests/test_conv_oracle.py.  draws None = centre crop, no flip."""
    n, H, W, _ = images_u8.shape
    pad = 4
    padded = np.zeros((n, H + 2 * pad, W + 2 * pad, 3), dtype=np.uint8)        # RandomCrop(padding=4): fill 0
    padded[:, pad:pad + H, pad:pad + W] = images_u8.numpy()
    out = np.empty((n, H, W, 3), dtype=np.uint8)
    for i in range(n):
        top, left, flip = (int(draws[i, 0]), int(draws[i, 1]), int(draws[i, 2])) if draws is not None else (pad, pad, 0)
        crop = padded[i, top:top + H, left:left + W]
        out[i] = crop[:, ::-1] if flip else crop
    x = torch.from_numpy(out).permute(0, 3, 1, 2).to(torch.float32).div(255)  # ToTensor
    mean = torch.tensor(CIFAR_MEAN, dtype=torch.float32).view(1, 3, 1, 1)
    std = torch.tensor(CIFAR_STD, dtype=torch.float32).view(1, 3, 1, 1)
    return x.sub_(mean).div_(std)                                              # Normalize: sub_(mean).div_(std)


class ResNetOracle:
    """`world` ranks of the reference's training loop in one process."""

    def __init__(self, arch: str, images_u8: torch.Tensor, labels: torch.Tensor, rank_rows: Sequence[torch.Tensor], *,
                 batch_size: int, lr: float = 0.1, momentum: float = 0.9, weight_decay: float = 1e-4, seed: int = 0,
                 num_classes: int = 1000, rank_draws: Optional[Sequence[Optional[torch.Tensor]]] = None,
                 optimizer: str = "sgd", dtype: torch.dtype = torch.float32):
        """dtype=torch.float64 runs the same objects in double precision: the yardstick for how far the fp32 CPU
        path itself is from the exact result (tests/test_conv_step_gpu.py, profiles/r02_resnet_parity_conditioning.md)."""
        import torchvision

        torch.manual_seed(seed)                                                # set_random_seeds, task.py:162
        self.model = torchvision.models.__dict__[arch](num_classes=num_classes).to(dtype)   # task.py:171
        self.dtype = dtype
        self.model.train()
        self.criterion = nn.CrossEntropyLoss()                                 # task.py:211
        if optimizer == "sgd":
            self.opt = torch.optim.SGD(self.model.parameters(), lr, momentum=momentum, weight_decay=weight_decay)
        else:
            self.opt = torch.optim.Adam(self.model.parameters(), lr, weight_decay=weight_decay)
        self.images, self.labels = images_u8, labels
        self.rank_rows = [r.to(torch.int64) for r in rank_rows]
        self.rank_draws = list(rank_draws) if rank_draws is not None else [None] * len(rank_rows)
        self.world, self.B, self.k = len(rank_rows), batch_size, 0

    def _batch(self, rank: int, k: int):
        rows = self.rank_rows[rank][k * self.B: (k + 1) * self.B]
        d = self.rank_draws[rank]
        x = augment_with_draws(self.images[rows], d[k * self.B: (k + 1) * self.B] if d is not None else None)
        return x.to(self.dtype), self.labels[rows]

    def step(self) -> List[float]:
        """One optimizer step; returns every rank's local loss (what each process prints at task.py:310)."""
        self.opt.zero_grad()                                                   # task.py:308
        start = {k: v.clone() for k, v in self.model.state_dict().items() if "running" in k or "num_batches" in k}
        keep: Dict[str, torch.Tensor] = {}
        losses = []
        for r in range(self.world):
            if r > 0:                                                          # every rank starts from rank 0's buffers
                self.model.load_state_dict(start, strict=False)
            x, y = self._batch(r, self.k)
            loss = self.criterion(self.model(x), y)                            # task.py:309-310
            (loss / self.world).backward()                                     # task.py:311 + DDP's mean
            losses.append(float(loss.detach()))
            if r == 0:
                keep = {k: v.clone() for k, v in self.model.state_dict().items() if k in start}
        self.model.load_state_dict(keep, strict=False)
        self.opt.step()                                                        # task.py:312
        self.k += 1
        return losses

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {"module." + k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def evaluate(self, images_u8: torch.Tensor, labels: torch.Tensor, draws: Optional[torch.Tensor] = None,
                 batch: int = 256):
        """task.py:30-46: model.eval(), argmax, accuracy."""
        self.model.eval()
        preds = []
        with torch.no_grad():
            for s in range(0, images_u8.shape[0], batch):
                x = augment_with_draws(images_u8[s: s + batch], draws[s: s + batch] if draws is not None else None)
                preds.append(self.model(x.to(self.dtype)).argmax(dim=1))
        self.model.train()
        p = torch.cat(preds)
        return float((p == labels).sum().item()) / labels.numel(), p
