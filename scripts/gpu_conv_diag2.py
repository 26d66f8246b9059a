"""Which tensors differ from the oracle after 1 and 2 steps, for several dataset sizes (prints only)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout  # noqa
from b200step.engine import StepConfig  # noqa
from b200step.sampler import rank_indices  # noqa
from oracle.resnet_step import ResNetOracle  # noqa

for rows, batch, seed in ((64, 32, 1), (256, 32, 1), (128, 64, 0), (64, 32, 7)):
    ds = make_synthetic_images(rows, seed=seed)
    rr = rank_indices(rows, 1, 0)
    torch.manual_seed(5)
    dr = draw_augmentation(rows)
    o = ResNetOracle("resnet18", ds.images, ds.labels, [rr], batch_size=batch, lr=0.01, rank_draws=[dr])
    o64 = ResNetOracle("resnet18", ds.images, ds.labels, [rr], batch_size=batch, lr=0.01, rank_draws=[dr],
                       dtype=torch.float64)
    e = ConvTrainStep(resnet_layout(), init_state(), ds, rr, StepConfig(batch_size=batch, precision="fp32", lr=0.01,
                                                                        use_graph=False))
    e.set_draws(dr)
    init = {"module." + k: v for k, v in init_state().items()}
    for s in range(2):
        lo = o.step()[0]
        l64 = o64.step()[0]
        e.step()
        lg = e.loss_value()
        so, sg, s64 = o.state_dict(), e.state_dict(), o64.state_dict()
        errs = []
        def worst(a, b):
            out = []
            for k in b:
                if b[k].dtype.is_floating_point and "running" not in k:
                    du = (b[k].double() - init[k].double())
                    out.append((((a[k].double() - b[k].double()).norm() / du.norm().clamp_min(1e-30)).item(), k))
            out.sort(reverse=True)
            return f"worst {out[0][0]:.1e} ({out[0][1]}) median {out[len(out) // 2][0]:.1e}"
        print(f"   loss fp64 {l64:.9f}: engine rel {abs(lg - l64) / l64:.1e}, cpu fp32 rel {abs(lo - l64) / l64:.1e}" if False else "", end="")
        for k in so:
            if not so[k].dtype.is_floating_point:
                continue
            du = (so[k] - init[k]).float()
            err_u = ((sg[k] - so[k]).float().norm() / du.norm().clamp_min(1e-20)).item()
            errs.append((err_u, k, du.norm().item()))
        errs.sort(reverse=True)
        print(f"rows={rows} B={batch} seed={seed} step {s + 1}: loss oracle {lo:.7f} engine {lg:.7f} rel {abs(lo - lg) / lo:.1e}")
        print(f"     loss vs fp64 oracle ({l64:.9f}): engine {abs(lg - l64) / l64:.1e}, cpu fp32 {abs(lo - l64) / l64:.1e}")
        print(f"     update vs fp64 oracle: engine   {worst(sg, s64)}")
        print(f"     update vs fp64 oracle: cpu fp32 {worst(so, s64)}")
        print(f"     update engine vs cpu fp32:      {worst(sg, so)}")
    e.close()
