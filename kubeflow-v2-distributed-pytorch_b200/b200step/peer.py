"""Peer-mapped buffers for the K-opt-peer path (csrc/peer.cu; EXPERIMENTAL, `StepConfig.peer_optimizer`).

Every rank keeps ONE device allocation `[parameters fp32 | gradients fp32 | bf16 operand copy | flags]` and maps
the allocation of every other rank of the node (CUDA IPC over NVLink), so that the per-bucket kernel can read all
ranks' gradient slices and write all ranks' parameters.  Replaces DDP's bucket all-reduce + optimizer.step()
(task.py:189, 312) for ranks that share a node; the reference's multi-node case stays on the NCCL path.

The allocation comes from torch's caching allocator and is shared with torch's own CUDA-IPC reduction
(`torch.multiprocessing.reductions.reduce_tensor`, the mechanism behind sending CUDA tensors through
multiprocessing queues); the handles travel through `torch.distributed.all_gather_object`.  A host without torch
uses `b200_peer_alloc` / `b200_peer_open` of the C ABI for the same purpose.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from .lib import B200Error, get_lib

ALIGN_BYTES = 256


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def slice_bounds(start: int, end: int, world: int, rank: int):
    """[lo, hi) of the bucket [start, end) that `rank` reduces and updates — the same arithmetic as peer_opt_kernel
    (csrc/peer.cu): `world` equal slices, each a multiple of 4 elements, the last one possibly shorter or empty."""
    n = end - start
    per = (((n + world - 1) // world) + 3) & ~3
    lo = start + per * rank
    hi = min(lo + per, end)
    return (lo, max(lo, hi)) if lo < end else (end, end)


class PeerBuffers:
    def __init__(self, total: int, n_shadow: int, rank: int, world: int, device: torch.device):
        lib = get_lib()
        self.rank, self.world, self.dev = rank, world, device
        self.total, self.n_shadow = total, n_shadow
        self.flag_words = int(lib.peer_flag_words())
        self.off_p = 0
        self.off_g = _round_up(self.off_p + 4 * total, ALIGN_BYTES)
        self.off_shadow = _round_up(self.off_g + 4 * total, ALIGN_BYTES)
        self.off_flags = _round_up(self.off_shadow + 2 * n_shadow, ALIGN_BYTES)
        self.nbytes = _round_up(self.off_flags + 4 * self.flag_words, ALIGN_BYTES)
        # (torch shares the cudaMalloc block this tensor lives in plus its offset; peers see the same bytes)
        self.arena = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        torch.cuda.synchronize(device)
        self.arenas: List[Optional[torch.Tensor]] = [None] * world
        self.arenas[rank] = self.arena
        if world > 1:
            import torch.distributed as dist
            from torch.multiprocessing.reductions import reduce_tensor

            if not dist.is_initialized():
                raise B200Error("peer_optimizer with world > 1 needs an initialised torch.distributed process group "
                                "to exchange the CUDA IPC handles")
            rebuild, args = reduce_tensor(self.arena)
            shared = [None] * world
            dist.all_gather_object(shared, (rebuild, args))
            for q in range(world):
                if q != rank:
                    fn, a = shared[q]
                    self.arenas[q] = fn(*a)             # maps rank q's allocation into this process
                    if self.arenas[q].numel() != self.nbytes:
                        raise B200Error(f"rank {q} shared {self.arenas[q].numel()} bytes, expected {self.nbytes}")
                    # torch maps the allocation in the context of the device that owns it; kernels of THIS device
                    # dereference it, which needs peer access switched on explicitly
                    with torch.cuda.device(device):
                        lib.peer_enable_access(int(self.arenas[q].device.index))
            dist.barrier()                               # every rank has mapped every arena before anyone goes on
        base = [int(a.data_ptr()) for a in self.arenas]
        if any(b % 16 for b in base):
            raise B200Error("peer arenas must be 16-byte aligned")
        tab = lambda off: (C.c_void_p * world)(*[b + off for b in base])
        self.p_tab, self.g_tab = tab(self.off_p), tab(self.off_g)
        self.shadow_tab = tab(self.off_shadow) if n_shadow else None
        self.flags_tab = tab(self.off_flags)
        a = self.arena
        self.p = a[self.off_p: self.off_p + 4 * total].view(torch.float32)
        self.g = a[self.off_g: self.off_g + 4 * total].view(torch.float32)
        self.shadow = a[self.off_shadow: self.off_shadow + 2 * n_shadow].view(torch.bfloat16) if n_shadow else None
        self.flags = a[self.off_flags: self.off_flags + 4 * self.flag_words].view(torch.int32)
        # local-only words: [0] ticket of the per-bucket kernels, [1] ticket of the finishing kernel, [2] error
        self.words = torch.zeros(4, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)

    def error_code(self) -> int:
        return int(self.words[2].item())

    def close(self) -> None:
        self.arenas = []
        self.arena = None
        self.p = self.g = self.shadow = self.flags = None
