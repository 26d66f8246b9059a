"""Per-tensor error of the first step's update against the fp64 oracle, in parameter order (prints only)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout  # noqa
from b200step.engine import StepConfig  # noqa
from b200step.sampler import rank_indices  # noqa
from oracle.resnet_step import ResNetOracle  # noqa

rows, batch, seed = 64, 32, 1
ds = make_synthetic_images(rows, seed=seed)
rr = rank_indices(rows, 1, 0)
torch.manual_seed(5)
dr = draw_augmentation(rows)
init = {"module." + k: v for k, v in init_state().items()}
o64 = ResNetOracle("resnet18", ds.images, ds.labels, [rr], batch_size=batch, lr=0.01, rank_draws=[dr], dtype=torch.float64)
o64.step()
s64 = o64.state_dict()
for prec, det in (("fp32", False), ("fp32", True), ("tf32", False)):
    e = ConvTrainStep(resnet_layout(), init_state(), ds, rr, StepConfig(batch_size=batch, precision=prec, lr=0.01,
                                                                        use_graph=False, deterministic_wgrad=det))
    e.set_draws(dr)
    e.step()
    sg = e.state_dict()
    print(f"--- {prec} deterministic_wgrad={det}: loss {e.loss_value():.7f}")
    for k in s64:
        if s64[k].dtype.is_floating_point and "running" not in k:
            du = s64[k].double() - init[k].double()
            err = ((sg[k].double() - s64[k].double()).norm() / du.norm().clamp_min(1e-30)).item()
            print(f"   {err:.1e}  {k}")
    e.close()
