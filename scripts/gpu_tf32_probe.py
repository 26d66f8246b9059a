"""Probe: error of the TF32 GEMM path against fp64, with tensor-map rounding vs tensor-core truncation."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "kubeflow-v2-distributed-pytorch_b200"))
import torch
from b200step import ops
torch.manual_seed(0)
M, N, K = 4096, 512, 512
x = torch.randn(M, K, device="cuda"); w = torch.randn(N, K, device="cuda") * 0.05
y = torch.empty(M, N, device="cuda")
ops.linear_fwd(x, w, None, y, M=M, N=N, K=K, relu=False); torch.cuda.synchronize()
ref = x.double() @ w.double().t()
err = (y.double() - ref)
print("mode", "truncate" if os.environ.get("B200STEP_TF32_TRUNCATE") else "round", "max_abs", err.abs().max().item(), "rms", err.pow(2).mean().sqrt().item(), "mean_signed_rel", (err * ref.sign()).mean().item() / ref.abs().mean().item())
torch.backends.cuda.matmul.allow_tf32 = True
yt = x @ w.t(); e2 = yt.double() - ref
print("torch tf32 cublas: max_abs", e2.abs().max().item(), "rms", e2.pow(2).mean().sqrt().item())
torch.backends.cuda.matmul.allow_tf32 = False
yt = x @ w.t(); e2 = yt.double() - ref
print("torch fp32 cublas: max_abs", e2.abs().max().item(), "rms", e2.pow(2).mean().sqrt().item())
