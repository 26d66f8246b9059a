"""Device time of the output layer + loss: b200_linear_ce (one kernel) against b200_linear_fwd + b200_ce_loss_grad.
Run on the GPU box: python scripts/gpu_linear_ce_bench.py [M N K]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
from b200step import ops  # noqa: E402


def timeit(fn, reps=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def main():
    M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (65536, 1000, 512)
    for dtype in (torch.bfloat16, torch.float32):
        g = torch.Generator().manual_seed(0)
        x = torch.randn(M, K, generator=g).to(dtype).cuda()
        w = (torch.randn(N, K, generator=g) * 0.05).to(dtype).cuda()
        b = torch.randn(N, generator=g).cuda()
        y = torch.randint(0, 10, (M,), generator=g).cuda()
        logits = torch.empty(M, N, dtype=dtype, device="cuda")
        d = torch.empty(M, N, dtype=dtype, device="cuda")
        loss, db = torch.zeros(1, device="cuda"), torch.zeros(N, device="cuda")
        s1 = torch.empty(ops.ce_scratch_floats(N), device="cuda")
        s2 = torch.empty(ops.linear_ce_scratch_floats(N), device="cuda")
        # a buffer larger than L2 written between repetitions is not needed: 2 x 131 MB of traffic per repetition
        t_fwd = timeit(lambda: ops.linear_fwd(x, w, b, logits, M=M, N=N, K=K, relu=False))
        t_ce = timeit(lambda: ops.ce_loss_grad(logits, y, d, loss, db, s1, batch=M, num_classes=N, grad_scale=1.0 / M))
        t_both = timeit(lambda: (ops.linear_fwd(x, w, b, logits, M=M, N=N, K=K, relu=False),
                                 ops.ce_loss_grad(logits, y, d, loss, db, s1, batch=M, num_classes=N,
                                                  grad_scale=1.0 / M)))
        t_fused = timeit(lambda: ops.linear_ce(x, w, b, y, d, loss, db, s2, M=M, N=N, K=K, grad_scale=1.0 / M))
        fl = 2.0 * M * N * K
        print(f"{str(dtype):16s} M={M} N={N} K={K}: linear_fwd {t_fwd:7.1f} us  ce_loss_grad {t_ce:7.1f} us  "
              f"both {t_both:7.1f} us | linear_ce {t_fused:7.1f} us = {fl / t_fused / 1e6:7.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
