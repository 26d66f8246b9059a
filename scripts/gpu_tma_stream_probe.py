"""Operand-streaming rate of the GEMM mainloop (B200STEP_GEMM_DEBUG=10: MMAs and epilogue skipped) for the same
bytes of A laid out with a 1 KiB row pitch (M=65536, K=512) vs contiguous 16 KiB tiles (M=524288, K=64)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "kubeflow-v2-distributed-pytorch_b200"))
import torch
from b200step import ops
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for (M, K) in [(65536, 512), (131072, 256), (262144, 128), (524288, 64)]:
    N = 256
    x = torch.randn(M, K, device="cuda").bfloat16(); w = torch.randn(N, K, device="cuda").bfloat16()
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    def run():
        flush.zero_()                     # evict x from L2 between repetitions
        ops.linear_fwd(x, w, None, y, M=M, N=N, K=K, relu=False)
    def run_flush_only():
        flush.zero_()
    us = t(run) - t(run_flush_only)
    print(f"M={M} K={K} (row pitch {K*2} B): {us:.1f} us for {M*K*2/1e6:.0f} MB of A -> {M*K*2/us/1e6:.2f} TB/s")
