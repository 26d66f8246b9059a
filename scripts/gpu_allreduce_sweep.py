#!/usr/bin/env python
"""BASELINE config 5: all-reduce sweep of one flat fp32 gradient buffer of S MB, cut into
DDP-style buckets (first 1 MiB, then 25 MiB), reduced through the library's own K-comm entry point
(`b200_comm_allreduce_sum`, NCCL over NVLink) on a high-priority side stream WHILE a backward-shaped GEMM chain
(K-gemm dgrad + wgrad launches at the T0 layer shape, 65 536 x 512 x 512 bf16) runs on the main stream.

Per S it reports (max over ranks, CUDA events):
  comm_alone_ms   all buckets back to back, nothing else running
  gemm_alone_ms   the GEMM chain alone (its length is chosen so that it lasts about as long as comm_alone)
  both_ms         both started together, until both are done
  exposed_ms      both_ms - gemm_alone_ms  (the part of the all-reduce the backward could not hide)
  algbw / busbw   S / t  and  S * 2(W-1)/W / t  for t = comm_alone, against 900 GB/s per direction.
Every leg is captured into a CUDA graph and replayed (like the training step), so host enqueue order plays no part.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 --master-port P \
        scripts/gpu_allreduce_sweep.py [--max-mb 1024] [--reps 5]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200"))

from b200step import ops  # noqa: E402
from b200step.engine import Communicator  # noqa: E402
from b200step.lib import get_lib  # noqa: E402

NVLINK_GBS = 900.0
M, N, K = 65536, 512, 512


def buckets(numel, first=1 << 20, cap=25 << 20):
    out, s, c = [], 0, first // 4
    while s < numel:
        e = min(numel, s + c)
        out.append((s, e))
        s, c = e, cap // 4
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--min-mb", type=int, default=1)
    ap.add_argument("--sm-reserve", type=int, default=0, help="SMs the GEMM grids leave free (b200_set_sm_budget)")
    args = ap.parse_args()
    import torch.distributed as dist

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", init_method="env://", world_size=world, rank=rank, device_id=dev)
    uid = [Communicator.make_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    comm = Communicator(rank, world, uid[0])
    main_s = torch.cuda.Stream(dev)
    side = torch.cuda.Stream(dev, priority=-1)

    # operands of the GEMM chain
    g = torch.Generator(device=dev).manual_seed(1)
    dy = (torch.randn(M, N, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    x = (torch.randn(M, K, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    dx = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    dw = torch.zeros(N, K, device=dev, dtype=torch.float32)

    def gemm_pair():
        ops.linear_dgrad(dy, w, None, dx, M=M, N=N, K=K, relu=False, stream=main_s)
        ops.linear_wgrad(dy, x, dw, M=M, N=N, K=K, accumulate=True, stream=main_s)

    if args.sm_reserve:
        ops.set_sm_budget(ops.sm_count() - args.sm_reserve)   # every GEMM grid below leaves these SMs free

    def maxms(v):
        t = torch.tensor([v], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lib = get_lib()
    fork, join = torch.cuda.Event(), torch.cuda.Event()

    def capture(fn):
        """`fn` enqueues on main_s (and, forked from it, on side); returns a replayable graph like the step's."""
        import ctypes as C

        torch.cuda.synchronize(dev)
        lib.graph_begin(main_s.cuda_stream)
        try:
            fn()
        finally:
            out = C.c_void_p()
            lib.graph_end(main_s.cuda_stream, C.byref(out))
        return out

    def timed(graph):
        """median over reps of one replay, max over ranks"""
        vals = []
        for _ in range(args.reps + 1):
            dist.barrier()
            torch.cuda.synchronize(dev)
            a, b = ev(), ev()
            a.record(main_s)
            lib.graph_launch(graph, main_s.cuda_stream)
            b.record(main_s)
            torch.cuda.synchronize(dev)
            vals.append(a.elapsed_time(b))
        vals = sorted(vals[1:])
        return maxms(vals[len(vals) // 2])

    ev = lambda: torch.cuda.Event(enable_timing=True)

    # one GEMM pair alone (for sizing the chain)
    for _ in range(3):
        gemm_pair()
    torch.cuda.synchronize(dev)
    a, b = ev(), ev()
    a.record(main_s)
    for _ in range(8):
        gemm_pair()
    b.record(main_s)
    torch.cuda.synchronize(dev)
    pair_ms = maxms(a.elapsed_time(b) / 8)

    rows = []
    S = args.min_mb
    while S <= args.max_mb:
        numel = S * (1 << 20) // 4
        flat = torch.zeros(numel, device=dev, dtype=torch.float32)
        bks = buckets(numel)

        def comm_part():
            fork.record(main_s)
            side.wait_event(fork)
            for s, e in bks:
                comm.allreduce_sum(flat[s:e], side)
            join.record(side)

        def comm_only():
            comm_part()
            main_s.wait_event(join)

        for s, e in bks:                         # eager once: NCCL sets up its channels for this size
            comm.allreduce_sum(flat[s:e], side)
        torch.cuda.synchronize(dev)
        g_comm = capture(comm_only)
        comm_alone = timed(g_comm)
        pairs = max(4, int(comm_alone / pair_ms) + 1)

        def gemm_only():
            for _ in range(pairs):
                gemm_pair()

        def both():
            comm_part()
            for _ in range(pairs):
                gemm_pair()
            main_s.wait_event(join)

        g_gemm = capture(gemm_only)
        g_both = capture(both)
        gemm_alone = timed(g_gemm)
        both_ms = timed(g_both)
        bus = S * (1 << 20) * 2 * (world - 1) / world
        rows.append({
            "S_MB": S, "buckets": len(bks), "gemm_pairs": pairs,
            "comm_alone_ms": comm_alone, "gemm_alone_ms": gemm_alone, "both_ms": both_ms,
            "exposed_ms": max(0.0, both_ms - gemm_alone),
            "slowdown_vs_max": both_ms / max(comm_alone, gemm_alone),
            "algbw_alone_gbs": S * (1 << 20) / comm_alone / 1e6, "busbw_alone_gbs": bus / comm_alone / 1e6,
            "busbw_alone_frac_of_900": bus / comm_alone / 1e6 / NVLINK_GBS,
        })
        for gph in (g_comm, g_gemm, g_both):
            lib.graph_destroy(gph)
        del flat
        S *= 2
    if rank == 0:
        print(json.dumps({"config": "allreduce_sweep", "world": world, "gemm_shape": [M, N, K],
                          "gemm_pair_ms": pair_ms, "nvlink_gbs_per_dir": NVLINK_GBS, "sm_reserve": args.sm_reserve,
                          "nccl_max_ctas": os.environ.get("NCCL_MAX_CTAS"), "rows": rows}), flush=True)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
