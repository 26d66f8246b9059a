"""ReLU masks of the engine's forward pass against the fp64 oracle's, per block (failing batch): counts sign flips."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
import torchvision  # noqa
from b200step.convnet import ConvTrainStep, draw_augmentation, init_state, make_synthetic_images, resnet_layout  # noqa
from b200step.engine import StepConfig  # noqa
from b200step.sampler import rank_indices  # noqa
from oracle.resnet_step import augment_with_draws  # noqa

for rows, batch, seed in ((64, 32, 1), (256, 32, 1)):
    ds = make_synthetic_images(rows, seed=seed)
    rr = rank_indices(rows, 1, 0)
    torch.manual_seed(5)
    dr = draw_augmentation(rows)
    x = augment_with_draws(ds.images[rr[:batch]], dr[:batch]).double()
    torch.manual_seed(0)
    m = torchvision.models.resnet18().double()
    m.train()
    pre = {}
    for l in (1, 2, 3, 4):
        for b in (0, 1):
            name, blk = f"layer{l}.{b}", getattr(m, f"layer{l}")[b]

            def fh(md, i, o, name=name):
                xin = i[0]
                idt = md.downsample(xin) if md.downsample is not None else xin
                mid = md.bn1(md.conv1(xin))
                pre[name + ".conv1"] = mid.detach()
                pre[name + ".conv2"] = (md.bn2(md.conv2(md.relu(mid))) + idt).detach()
            blk.register_forward_hook(fh)
    m(x)
    e = ConvTrainStep(resnet_layout(), init_state(), ds, rr, StepConfig(batch_size=batch, precision="fp32", lr=0.01,
                                                                        use_graph=False))
    e.set_draws(dr)
    e.step()
    e.synchronize()
    print(f"rows={rows} seed={seed}")
    for k, p in pre.items():
        B, C, H, W = p.shape
        mine = e.a[k][: B * H * W].view(B, H, W, C).permute(0, 3, 1, 2).cpu().double()
        flips = ((mine > 0) != (p > 0))
        err = (mine - p.clamp_min(0)).abs().max().item()
        msg = f"  {k}: {int(flips.sum())} ReLU sign flips of {p.numel()}, max |activation error| {err:.1e}"
        if flips.any():
            msg += f", |pre-activation| at the flips {p[flips].abs().tolist()} -> one flip is {1 / p.numel() ** 0.5:.1e} of the layer's gradient norm"
        print(msg)
    e.close()
