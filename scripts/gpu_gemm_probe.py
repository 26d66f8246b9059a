"""Timing probe of single GEMM launches (CUDA events, 20 reps) — used with the B200STEP_GEMM_DEBUG knobs."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "kubeflow-v2-distributed-pytorch_b200"))
import torch
from b200step import ops
M = 65536
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
for (N, K) in [(512, 448), (512, 512), (1000, 512)]:
    x = torch.randn(M, K, device="cuda").bfloat16(); w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    b = torch.randn(N, device="cuda"); y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    mask = torch.zeros(M, ((N + 31) // 32 + 3) // 4 * 4, dtype=torch.int32, device="cuda")
    us = t(lambda: ops.linear_fwd(x, w, b, y, M=M, N=N, K=K, relu=True, relu_mask=mask))
    us2 = t(lambda: ops.linear_fwd(x, w, b, y, M=M, N=N, K=K, relu=True))
    dy = torch.randn(M, N, device="cuda").bfloat16(); dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    us3 = t(lambda: ops.linear_dgrad(dy, w, None, dx, M=M, N=N, K=K, relu=False))
    dw = torch.empty(N, K, device="cuda")
    us4 = t(lambda: ops.linear_wgrad(dy, x, dw, M=M, N=N, K=K))
    fl = 2.0 * M * N * K
    print(f"N={N} K={K}: fwd+mask {us:.1f} us ({fl/us/1e6:.0f} TF) | fwd {us2:.1f} us ({fl/us2/1e6:.0f} TF) | dgrad {us3:.1f} us ({fl/us3/1e6:.0f} TF) | wgrad {us4:.1f} us ({fl/us4/1e6:.0f} TF)")
xx = torch.randn(8192, 8192, device="cuda").bfloat16(); ww = torch.randn(8192, 8192, device="cuda").bfloat16(); yy = torch.empty(8192, 8192, device="cuda", dtype=torch.bfloat16)
us = t(lambda: ops.linear_fwd(xx, ww, None, yy, M=8192, N=8192, K=8192, relu=False), 5)
print(f"8192^3 fwd: {us:.1f} us  {2*8192**3/us/1e6:.0f} TF;  cuBLAS:", end=" ")
us = t(lambda: torch.matmul(xx, ww.t(), out=yy), 5); print(f"{us:.1f} us {2*8192**3/us/1e6:.0f} TF")
x = torch.randn(M, 512, device="cuda").bfloat16(); w = torch.randn(512, 512, device="cuda").bfloat16()
us = t(lambda: torch.nn.functional.linear(x, w)); print(f"cuBLAS 65536x512x512: {us:.1f} us {2*M*512*512/us/1e6:.0f} TF")
