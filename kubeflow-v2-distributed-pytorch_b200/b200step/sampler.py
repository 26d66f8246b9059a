"""Index lists of the reference's DistributedSampler + DataLoader batching, built once on the host.

Reference behaviour (task.py:260 `DistributedSampler(dataset=train_set)` with defaults shuffle=True, seed=0,
drop_last=False; task.py:263 `DataLoader(batch_size=batch_size, sampler=...)`; `set_epoch` is never called, so
every epoch sees the same order — SURVEY.md App. B-4):
  perm    = randperm(N) from torch.Generator().manual_seed(seed + epoch)
  padded  = perm + perm[: total_size - N]   (wrap-around so that total_size = ceil(N/W)*W; repeated if needed)
  rank r  = padded[r : total_size : W]
  batches = consecutive chunks of `batch_size`; the last partial chunk is kept (drop_last=False).
"""
from __future__ import annotations

import math
from typing import List

import torch


def rank_indices(num_rows: int, world: int, rank: int, seed: int = 0, epoch: int = 0,
                 shuffle: bool = True) -> torch.Tensor:
    if not 0 <= rank < world:
        raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {world - 1}]")
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(num_rows, generator=g)
    else:
        idx = torch.arange(num_rows)
    num_samples = math.ceil(num_rows / world)
    total = num_samples * world
    pad = total - num_rows
    if pad > 0:
        if pad <= num_rows:
            idx = torch.cat([idx, idx[:pad]])
        else:
            idx = torch.cat([idx, idx.repeat(math.ceil(pad / num_rows))[:pad]])
    out = idx[rank:total:world].contiguous()
    assert out.numel() == num_samples
    return out


def batch_bounds(num_samples: int, batch_size: int) -> List[range]:
    """[start, end) of each DataLoader batch over the rank's index list (last one may be partial)."""
    return [range(s, min(s + batch_size, num_samples)) for s in range(0, num_samples, batch_size)]
