"""Device time of K-embbwd: privatised (shared-memory) kernel against the plain global scatter, uniform and Zipf ids."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]


def main():
    from b200step import ops

    B, nc, rows, dim, ldx = 65536, 26, 1000, 16, 448
    g = torch.Generator().manual_seed(0)
    dx = (torch.randn(B, ldx, generator=g) * 0.01).to(torch.bfloat16).cuda()
    probs = torch.arange(1, rows + 1, dtype=torch.float64).pow(-1.05)
    ids = {"uniform": torch.randint(0, rows, (B, nc), generator=g).int().cuda(),
           "zipf1.05": torch.multinomial(probs / probs.sum(), B * nc, replacement=True, generator=g).view(B, nc).int().cuda()}
    dt = torch.zeros(nc * rows * dim, device="cuda")
    big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for name, idx in ids.items():
        ts = []
        for _ in range(10):
            big.zero_()      # flush L2
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.emb_bwd(dx, idx, dt, batch=B, num_cat=nc, table_rows=rows, embed_dim=dim)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        print(f"B200STEP_EMB_BWD={os.environ.get('B200STEP_EMB_BWD', 'auto'):7s} {name:9s} min {min(ts):6.1f} us  median {sorted(ts)[5]:6.1f} us", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        main()
    else:
        for mode in ("smem", "global"):
            subprocess.run([sys.executable, os.path.abspath(__file__), "child"], env=dict(os.environ, B200STEP_EMB_BWD=mode))
