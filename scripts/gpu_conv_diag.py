"""Diagnostics for the ResNet path: run-to-run determinism, graph vs eager, sensitivity to lr (prints only)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout  # noqa
from b200step.engine import StepConfig  # noqa
from b200step.sampler import rank_indices  # noqa
from oracle.resnet_step import ResNetOracle  # noqa


def run(precision, lr, use_graph, steps=5, oracle=False, batch=32, rows=256):
    ds = make_synthetic_images(rows, seed=1)
    rr = rank_indices(rows, 1, 0)
    torch.manual_seed(5)
    dr = draw_augmentation(rows)
    if oracle:
        o = ResNetOracle("resnet18", ds.images, ds.labels, [rr], batch_size=batch, lr=lr, rank_draws=[dr])
        return [o.step()[0] for _ in range(steps)]
    eng = ConvTrainStep(resnet_layout(), init_state(), ds, rr,
                        StepConfig(batch_size=batch, precision=precision, lr=lr, use_graph=use_graph))
    eng.set_draws(dr)
    out = []
    for _ in range(steps):
        eng.step()
        out.append(eng.loss_value())
    eng.close()
    return out


for lr in (0.05, 0.005):
    ref = run("fp32", lr, False, oracle=True)
    print(f"lr={lr} oracle      ", ref)
    for prec in ("fp32", "tf32", "bf16"):
        e1 = run(prec, lr, False)
        e2 = run(prec, lr, False)
        g1 = run(prec, lr, True)
        g2 = run(prec, lr, True)
        rel = lambda a: [f"{abs(x - y) / max(1, abs(y)):.1e}" for x, y in zip(a, ref)]
        print(f"lr={lr} {prec} eager  ", e1, rel(e1))
        print(f"lr={lr} {prec} eager2 ", "same" if e1 == e2 else e2)
        print(f"lr={lr} {prec} graph  ", "same as eager" if g1 == e1 else g1, rel(g1))
        print(f"lr={lr} {prec} graph2 ", "same as graph" if g1 == g2 else g2)
