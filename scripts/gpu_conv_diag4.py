"""layer4.1's backward, piece by piece, on the real tensors of the failing batch: GPU ops vs fp64 autograd."""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")]
import torchvision  # noqa
from b200step import ops  # noqa
from b200step.convnet import draw_augmentation, filter_to_gemm, make_synthetic_images  # noqa
from b200step.sampler import rank_indices  # noqa
from oracle.resnet_step import augment_with_draws  # noqa

rows, batch, seed = 64, 32, 1
ds = make_synthetic_images(rows, seed=seed)
rr = rank_indices(rows, 1, 0)
torch.manual_seed(5)
dr = draw_augmentation(rows)
xin = augment_with_draws(ds.images[rr[:batch]], dr[:batch]).double()
yl = ds.labels[rr[:batch]]
torch.manual_seed(0)
m = torchvision.models.resnet18().double()
m.train()
cap = {}
blk = m.layer4[1]
blk.register_forward_pre_hook(lambda md, i: cap.__setitem__("x", i[0].detach().clone()))
def _fh(md, i, o):
    o.register_hook(lambda g: cap.__setitem__("g", g.detach().clone()))
blk.register_forward_hook(_fh)
F.cross_entropy(m(xin), yl).backward()
x, g = cap["x"], cap["g"]                      # [32, 512, 1, 1] fp64
W1, W2 = blk.conv1.weight.detach(), blk.conv2.weight.detach()
# fp64 replay of the block with every intermediate kept
xr = x.clone().requires_grad_(True)
y1 = F.conv2d(xr, W1, padding=1); y1.retain_grad()
a1 = F.relu(F.batch_norm(y1, None, None, torch.ones(512, dtype=torch.float64), torch.zeros(512, dtype=torch.float64), True)); a1.retain_grad()
y2 = F.conv2d(a1, W2, padding=1); y2.retain_grad()
out = F.relu(F.batch_norm(y2, None, None, torch.ones(512, dtype=torch.float64), torch.zeros(512, dtype=torch.float64), True) + xr)
out.backward(g)
print("fp64 replay: |dx| =", xr.grad.norm().item(), " x sparsity", (x == 0).double().mean().item(),
      " min channel var of y1 / y2:", y1.var(dim=0, unbiased=False).min().item(), y2.var(dim=0, unbiased=False).min().item())

dev = "cuda"
f32 = dict(dtype=torch.float32, device=dev)
def mat(t):   # [B, C, 1, 1] -> [B, C] fp32 cuda
    return t.reshape(t.shape[0], -1).float().to(dev).contiguous()
def rel(a, b):
    return ((a.double().cpu() - b.double().reshape(a.shape)).norm() / b.double().norm()).item()
B, C, Kp = 32, 512, 9 * 512
scr = torch.zeros(ops.bn_scratch_floats(C), **f32)
ones, zeros = torch.ones(C, **f32), torch.zeros(C, **f32)
def stats(y):
    mu, iv = torch.empty(C, **f32), torch.empty(C, **f32)
    ops.bn_fwd_stats(y, mu, iv, None, None, None, scr, M=B, C=C)
    return mu, iv
def gemm3(kind, a, w, out, **kw):
    """3xTF32 product through the split kernels, like ConvTrainStep in fp32 mode."""
    rows_, cols_ = a.shape
    ak = torch.empty(rows_, 3 * cols_, **f32); ar = torch.empty(3 * rows_, cols_, **f32)
    return ak, ar
wg1, wg2 = filter_to_gemm(W1.float(), Kp).to(dev), filter_to_gemm(W2.float(), Kp).to(dev)
def split_w(w):
    wk, wr = torch.empty(w.shape[0], 3 * w.shape[1], **f32), torch.empty(3 * w.shape[0], w.shape[1], **f32)
    ops.split_tf32(w, wk, ops.SPLIT_RIGHT, wr, ops.SPLIT_RIGHT, rows=w.shape[0], cols=w.shape[1])
    return wk, wr
wk1, wr1 = split_w(wg1); wk2, wr2 = split_w(wg2)
def dgrad3(dy, wr):
    dyk, dyr = torch.empty(B, 3 * C, **f32), torch.empty(3 * B, C, **f32)
    ops.split_tf32(dy, dyk, ops.SPLIT_LEFT, dyr, ops.SPLIT_LEFT, rows=B, cols=C)
    dcol = torch.empty(B, Kp, **f32)
    ops.linear_dgrad(dyk, wr, None, dcol, M=B, N=3 * C, K=Kp, relu=False)
    return dcol
y2g, outg, y1g, a1g, gg, xg = mat(y2), mat(out), mat(y1), mat(a1), mat(g), mat(x)
mu2, iv2 = stats(y2g)
dg, db = torch.empty(C, **f32), torch.empty(C, **f32)
dy2, dz = torch.empty(B, C, **f32), torch.empty(B, C, **f32)
ops.bn_bwd_reduce(gg, outg, y2g, mu2, iv2, dg, db, scr, M=B, C=C)
ops.bn_bwd_apply(gg, outg, y2g, dy2, dz, mu2, iv2, ones, dg, db, M=B, C=C)
print("bn2 bwd: dy2 rel", rel(dy2, y2.grad), " dz rel", rel(dz, g * (out > 0)))
dcol2 = dgrad3(dy2, wr2)
ref_dcol2 = (y2.grad.reshape(B, C) @ filter_to_gemm(W2, Kp))
print("conv2 dgrad (3xTF32): dcol rel", rel(dcol2, ref_dcol2), " centre-tap columns rel",
      rel(dcol2[:, 4 * C:5 * C], ref_dcol2[:, 4 * C:5 * C]))
dcol2_exact_in = dgrad3(mat(y2.grad), wr2)
print("conv2 dgrad with the fp64 dy2 rounded to fp32 as input: centre-tap rel", rel(dcol2_exact_in[:, 4 * C:5 * C], ref_dcol2[:, 4 * C:5 * C]))
da1 = torch.empty(B, C, **f32)
ops.col2im(dcol2, None, da1, B=B, H=1, W=1, C=C, kh=3, kw=3, stride=1, pad=1)
print("col2im: d_a1 rel", rel(da1, a1.grad))
mu1, iv1 = stats(y1g)
dy1 = torch.empty(B, C, **f32)
ops.bn_bwd_reduce(da1, a1g, y1g, mu1, iv1, dg, db, scr, M=B, C=C)
ops.bn_bwd_apply(da1, a1g, y1g, dy1, None, mu1, iv1, ones, dg, db, M=B, C=C)
print("bn1 bwd: dy1 rel", rel(dy1, y1.grad))
dcol1 = dgrad3(dy1, wr1)
dx = torch.empty(B, C, **f32)
ops.col2im(dcol1, dz, dx, B=B, H=1, W=1, C=C, kh=3, kw=3, stride=1, pad=1)
print("conv1 dgrad + col2im + residual: dx rel", rel(dx, xr.grad))
ref_main = (y1.grad.reshape(B, C) @ filter_to_gemm(W1, Kp))[:, 4 * C:5 * C]
print("   main branch alone rel", rel(dcol1[:, 4 * C:5 * C], ref_main), " |main| / |dz| =", ref_main.norm().item() / (g * (out > 0)).norm().item())
