"""The reference-literal model path: torchvision ResNet (BasicBlock nets) on 3x32x32 images (task.py:171,247-257).

`task.py` trains `models.__dict__[args.arch]()` — ResNet-18 by default — on CIFAR-10 with RandomCrop / flip /
normalize, CrossEntropyLoss and SGD under DistributedDataParallel.  This module is that training step on the B200
kernels: NHWC activations, every convolution an im2col gather + the tcgen05 GEMMs of K-gemm, BatchNorm as two
memory-bound passes, the output layer fused with the loss (K-gemm-ce), one fused SGD launch over the flat parameter
buffer, DDP's gradient buckets all-reduced on a side stream.  No torch operator runs inside the step.

Parameter layout: one flat fp32 buffer in `model.parameters()` order; a conv filter is stored as the GEMM operand
Wg[Cout][ky][kx][c] (row padded to a multiple of 8 columns), i.e. torch's OIHW permuted to OHWI; `state_dict()` permutes
back, so checkpoints are interchangeable with the reference's (`module.`-prefixed torchvision keys, task.py:189).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

from . import ops
from .lib import B200Error, get_lib

ARCH_BLOCKS = {"resnet18": (2, 2, 2, 2), "resnet34": (3, 4, 6, 3)}     # torchvision/models/resnet.py: BasicBlock nets
WIDTHS = (64, 128, 256, 512)
CIFAR_MEAN = (0.4914, 0.4822, 0.4465)        # task.py:254
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


@dataclass
class ConvDesc:
    name: str                 # torchvision module path, e.g. "layer2.0.downsample.0"
    cin: int
    cout: int
    k: int
    stride: int
    pad: int
    H: int                    # input height / width
    W: int
    bn: str                   # name of the BatchNorm that follows
    Ho: int = 0
    Wo: int = 0
    K: int = 0                # kh*kw*cin
    Kp: int = 0               # K padded to the GEMM's row alignment

    def __post_init__(self):
        self.Ho = ops.conv_out_size(self.H, self.k, self.stride, self.pad)
        self.Wo = ops.conv_out_size(self.W, self.k, self.stride, self.pad)
        self.K = self.k * self.k * self.cin
        self.Kp = _round_up(self.K, 8)


@dataclass
class BlockDesc:
    name: str                 # "layer1.0"
    conv1: ConvDesc
    conv2: ConvDesc
    down: Optional[ConvDesc]


@dataclass
class Segment:
    key: str                  # state_dict key without the DDP "module." prefix
    shape: Tuple[int, ...]    # torch shape
    offset: int               # into the flat buffers
    numel: int                # flat elements (conv filters: cout * Kp)
    conv: Optional[ConvDesc] = None


@dataclass
class ResNetLayout:
    """Where every parameter and buffer of the torchvision model lives in the flat buffers."""
    arch: str
    image: int
    num_classes: int
    stem: ConvDesc = None
    blocks: List[BlockDesc] = field(default_factory=list)
    params: List[Segment] = field(default_factory=list)       # model.parameters() order
    bns: List[Tuple[str, int]] = field(default_factory=list)  # (name, channels) in module order
    n_params: int = 0
    pool_hw: int = 0          # spatial positions entering the average pool
    feat: int = 512

    def seg(self, key: str) -> Segment:
        return self._by_key[key]

    def finish(self):
        self._by_key = {s.key: s for s in self.params}
        self.n_params = self.params[-1].offset + self.params[-1].numel


def resnet_layout(arch: str = "resnet18", image: int = 32, num_classes: int = 1000) -> ResNetLayout:
    if arch not in ARCH_BLOCKS:
        raise B200Error(f"--arch {arch}: only the BasicBlock ResNets {sorted(ARCH_BLOCKS)} have a B200 path")
    lay = ResNetLayout(arch, image, num_classes)
    off = 0

    def add_param(key, shape, numel=None, conv=None):
        nonlocal off
        n = int(numel if numel is not None else torch.Size(shape).numel())
        lay.params.append(Segment(key, tuple(shape), off, n, conv))
        off += _round_up(n, 4)                 # every segment 16-byte aligned

    def add_conv(name, cin, cout, k, stride, pad, H, W, bn):
        c = ConvDesc(name, cin, cout, k, stride, pad, H, W, bn)
        add_param(name + ".weight", (cout, cin, k, k), cout * c.Kp, c)
        return c

    def add_bn(name, ch):
        add_param(name + ".weight", (ch,))
        add_param(name + ".bias", (ch,))
        lay.bns.append((name, ch))

    lay.stem = add_conv("conv1", 3, 64, 7, 2, 3, image, image, "bn1")
    add_bn("bn1", 64)
    H = ops.conv_out_size(lay.stem.Ho, 3, 2, 1)           # maxpool(3, 2, 1)
    cin = 64
    for li, (nblk, width) in enumerate(zip(ARCH_BLOCKS[arch], WIDTHS), start=1):
        for bi in range(nblk):
            stride = 2 if (bi == 0 and li > 1) else 1
            name = f"layer{li}.{bi}"
            c1 = add_conv(name + ".conv1", cin, width, 3, stride, 1, H, H, name + ".bn1")
            add_bn(name + ".bn1", width)
            c2 = add_conv(name + ".conv2", width, width, 3, 1, 1, c1.Ho, c1.Wo, name + ".bn2")
            add_bn(name + ".bn2", width)
            down = None
            if stride != 1 or cin != width:
                down = add_conv(name + ".downsample.0", cin, width, 1, stride, 0, H, H, name + ".downsample.1")
                add_bn(name + ".downsample.1", width)
            lay.blocks.append(BlockDesc(name, c1, c2, down))
            cin, H = width, c1.Ho
    lay.pool_hw = H * H
    add_param("fc.weight", (num_classes, 512))
    add_param("fc.bias", (num_classes,))
    lay.finish()
    return lay


def ddp_bucket_plan(lay: ResNetLayout, first_bucket_bytes: int = 1 << 20, bucket_cap_bytes: int = 25 << 20) -> List[List[str]]:
    """DistributedDataParallel's gradient buckets for this model (task.py:189; torch/nn/parallel/distributed.py builds
    them with `_compute_bucket_assignment_by_size(reversed(parameters), [1 MiB, bucket_cap_mb])`): parameters are taken in
    reverse order — the order backward produces their gradients — and a bucket is closed as soon as its byte size
    reaches the limit (1 MiB for the first, 25 MiB after).  Sizes are the TORCH sizes of the parameters.  Because the flat
    buffer is in parameter order, every bucket is one contiguous range.  Checked against torch's own function in
    tests/test_conv_oracle.py (ResNet-18: 3 buckets of 2 / 12 / 48 tensors = 2.05 / 28.9 / 15.9 MB)."""
    plan: List[List[str]] = []
    cur: List[str] = []
    size = 0
    for s in reversed(lay.params):
        cur.append(s.key)
        size += 4 * int(torch.Size(s.shape).numel())
        if size >= (first_bucket_bytes if not plan else bucket_cap_bytes):
            plan.append(cur)
            cur, size = [], 0
    if cur:
        plan.append(cur)
    return plan


def filter_to_gemm(w: torch.Tensor, Kp: int) -> torch.Tensor:
    """torch OIHW -> Wg[Cout, Kp] with columns ordered (ky, kx, c) and zero padding."""
    o = w.shape[0]
    g = w.permute(0, 2, 3, 1).reshape(o, -1)
    out = torch.zeros(o, Kp, dtype=w.dtype, device=w.device)
    out[:, : g.shape[1]] = g
    return out


def gemm_to_filter(g: torch.Tensor, shape: Tuple[int, ...]) -> torch.Tensor:
    o, i, kh, kw = shape
    return g[:, : i * kh * kw].reshape(o, kh, kw, i).permute(0, 3, 1, 2).contiguous()


def flatten_state(lay: ResNetLayout, sd: Dict[str, torch.Tensor]) -> torch.Tensor:
    flat = torch.zeros(lay.n_params, dtype=torch.float32)
    for s in lay.params:
        t = sd[s.key].detach().to(torch.float32).cpu()
        if tuple(t.shape) != s.shape:
            raise ValueError(f"{s.key}: expected shape {s.shape}, got {tuple(t.shape)}")
        flat[s.offset: s.offset + s.numel] = (filter_to_gemm(t, s.conv.Kp) if s.conv is not None else t).reshape(-1)
    return flat


def init_state(arch: str = "resnet18", seed: int = 0, num_classes: int = 1000) -> Dict[str, torch.Tensor]:
    """The reference's initialisation: set_random_seeds(seed) then `models.__dict__[arch]()` (task.py:22-28,171)."""
    import torchvision

    torch.manual_seed(seed)
    model = torchvision.models.__dict__[arch](num_classes=num_classes)
    return {k: v.detach().clone() for k, v in model.state_dict().items()}


@dataclass
class ImageDataset:
    """A CIFAR-10-shaped dataset resident in memory: uint8 HWC images and int64 labels (torchvision's CIFAR10.data)."""
    images: torch.Tensor      # [n, H, W, 3] uint8
    labels: torch.Tensor      # [n] int64

    def __len__(self):
        return int(self.images.shape[0])

    def to(self, dev) -> "ImageDataset":
        return ImageDataset(self.images.to(dev).contiguous(), self.labels.to(dev).contiguous())


def make_synthetic_images(n: int, seed: int = 0, image: int = 32, num_classes: int = 10) -> ImageDataset:
    """CIFAR-10 is not reachable (no network): same shapes and dtypes, seeded noise with a class-dependent tint so
    that the loss can fall."""
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, num_classes, (n,), generator=g, dtype=torch.int64)
    base = torch.randint(0, 256, (n, image, image, 3), generator=g, dtype=torch.int64)
    tint = torch.randint(0, 96, (num_classes, 3), generator=g, dtype=torch.int64)
    images = ((base * 5) // 8 + tint[labels][:, None, None, :]).clamp_(0, 255).to(torch.uint8)
    return ImageDataset(images, labels)


def load_cifar10_test_split(root: str = "data") -> Optional[ImageDataset]:
    """The dataset the reference trains AND tests on: CIFAR10(root="data", train=False) (task.py:256-257), when its
    files are already on disk (the download the reference asks for needs the network).  None when absent."""
    import os

    if not os.path.isdir(os.path.join(root, "cifar-10-batches-py")):
        return None
    import torchvision

    ds = torchvision.datasets.CIFAR10(root=root, train=False, download=False)
    return ImageDataset(torch.from_numpy(ds.data).contiguous(), torch.tensor(ds.targets, dtype=torch.int64))


def draw_augmentation_loop(n: int, image: int = 32, pad: int = 4) -> torch.Tensor:
    """`draw_augmentation` by the very calls torchvision makes, one sample at a time (the definition; 40 us/sample)."""
    out = torch.zeros(n, 4, dtype=torch.int8)
    hi = 2 * pad + 1
    for s in range(n):
        out[s, 0] = int(torch.randint(0, hi, size=(1,)).item())
        out[s, 1] = int(torch.randint(0, hi, size=(1,)).item())
        out[s, 2] = int(torch.rand(1).item() < 0.5)
    return out


_MT_WORDS, _MT_STATE_BYTES = 624, 5056


def draw_augmentation(n: int, image: int = 32, pad: int = 4) -> torch.Tensor:
    """The random numbers torchvision's transform pipeline (task.py:250-253) consumes for n consecutive samples, from
    torch's global generator in the same order: RandomCrop.get_params -> randint(0, 2*pad+1) for the top, then for the
    left; RandomHorizontalFlip.forward -> rand(1) < 0.5.  Returns int8 [n, 4] = (top, left, flip, 0).

    Each of those calls consumes exactly one 32-bit output of torch's CPU mt19937 (randint: value % range; rand: the low
    24 bits * 2^-24), so the 3n outputs are generated in one vectorised pass by numpy's MT19937 started from torch's
    own generator state, and the advanced state is written back: bit-identical draws and generator state to the
    per-sample loop (tests/test_conv_oracle.py) at 1/300 of its cost — the loop took longer per epoch than the GPU."""
    import numpy as np

    st = torch.get_rng_state()
    if n <= 0 or st.numel() != _MT_STATE_BYTES:
        return draw_augmentation_loop(n, image, pad)
    raw_state = bytearray(st.numpy().tobytes())
    left = int.from_bytes(raw_state[8:12], "little", signed=True)
    nxt = int.from_bytes(raw_state[16:24], "little")
    key = np.frombuffer(bytes(raw_state[24:24 + _MT_WORDS * 8]), dtype=np.uint64)
    if int(raw_state[12]) != 1 or not (0 <= nxt <= _MT_WORDS) or key.max() >= (1 << 32):
        return draw_augmentation_loop(n, image, pad)
    bg = np.random.MT19937()
    state = bg.state
    state["state"]["key"] = key.astype(np.uint32)
    state["state"]["pos"] = _MT_WORDS if left == 1 else nxt          # left == 1: the next output twists first
    bg.state = state
    raw = bg.random_raw(3 * n).reshape(n, 3)
    out = torch.zeros(n, 4, dtype=torch.int8)
    hi = 2 * pad + 1
    out[:, 0] = torch.from_numpy((raw[:, 0] % hi).astype(np.int8))
    out[:, 1] = torch.from_numpy((raw[:, 1] % hi).astype(np.int8))
    out[:, 2] = torch.from_numpy((((raw[:, 2] >> 23) & 1) == 0).astype(np.int8))     # (x & (2^24-1)) * 2^-24 < 0.5
    new = bg.state["state"]
    pos = int(new["pos"])
    raw_state[8:12] = int(_MT_WORDS + 1 - pos).to_bytes(4, "little", signed=True)
    raw_state[16:24] = pos.to_bytes(8, "little")
    raw_state[24:24 + _MT_WORDS * 8] = new["key"].astype(np.uint64).tobytes()
    torch.set_rng_state(torch.frombuffer(raw_state, dtype=torch.uint8).clone())
    return out


class ConvTrainStep:
    """One process's training step of a BasicBlock ResNet (the loop body of task.py:303-312) on one B200."""

    def __init__(self, layout: ResNetLayout, state: Dict[str, torch.Tensor], dataset: ImageDataset,
                 row_idx: torch.Tensor, cfg, comm=None, device: Optional[torch.device] = None, augment: bool = True):
        if not torch.cuda.is_available():
            raise B200Error("no CUDA device: the B200 training step has no CPU fallback")
        if cfg.precision not in ("bf16", "fp32", "tf32"):
            raise ValueError(f"precision must be bf16, fp32 or tf32, got {cfg.precision}")
        if cfg.optimizer not in ("sgd", "adam"):
            raise ValueError(f"optimizer must be sgd or adam, got {cfg.optimizer}")
        if cfg.world > 1 and comm is None:
            raise ValueError("world > 1 needs a Communicator")
        if dataset.images.dtype != torch.uint8 or dataset.images.shape[1:] != (layout.image, layout.image, 3):
            raise ValueError(f"images must be uint8 [n, {layout.image}, {layout.image}, 3]")
        self.lib = get_lib()
        self.lay, self.cfg, self.comm = layout, cfg, comm
        self.dev = dev = device or torch.device("cuda", torch.cuda.current_device())
        self.dt = torch.bfloat16 if cfg.precision == "bf16" else torch.float32
        self.x3 = cfg.precision == "fp32"
        self.augment = augment
        self.det = bool(getattr(cfg, "deterministic_wgrad", False))
        B = self.B = cfg.batch_size
        f32 = dict(dtype=torch.float32, device=dev)
        act = dict(dtype=self.dt, device=dev)

        self.ds = dataset.to(dev)
        self.row_idx = row_idx.to(dev, dtype=torch.int64).contiguous()
        self.num_samples = int(row_idx.numel())
        self.steps_per_epoch_full = self.num_samples // B
        self.tail = self.num_samples - self.steps_per_epoch_full * B
        self.draws = torch.zeros(self.num_samples, 4, dtype=torch.int8, device=dev)
        self.draws[:, 0] = 4
        self.draws[:, 1] = 4                                    # until set_draws(): centre crop, no flip
        self.mean3 = torch.tensor(CIFAR_MEAN, dtype=torch.float32)
        self.std3 = torch.tensor(CIFAR_STD, dtype=torch.float32)

        # ---- flat parameter / gradient / optimizer state
        n = layout.n_params
        self.p = flatten_state(layout, state).to(dev)
        self.g = None
        if cfg.world > 1 and getattr(cfg, "nccl_registered_grads", True):
            try:
                self.g = comm.alloc_registered(n, dev)
            except B200Error:
                self.g = None
        if self.g is None:
            self.g = torch.zeros(n, **f32)
        self.s1 = torch.zeros(n, **f32)
        self.s2 = torch.zeros(n, **f32) if cfg.optimizer == "adam" else None
        self.shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev) if self.dt == torch.bfloat16 else None
        if self.shadow is not None:
            ops.cast_bf16(self.p, self.shadow)

        # ---- BatchNorm buffers (state_dict order) and saved batch statistics
        self.bn_off, o = {}, 0
        for name, ch in layout.bns:
            self.bn_off[name] = (o, ch)
            o += ch
        self.running_mean = torch.zeros(o, **f32)
        self.running_var = torch.ones(o, **f32)
        self.num_batches = torch.zeros(len(layout.bns), dtype=torch.int64, device=dev)
        self.bn_index = {name: i for i, (name, _) in enumerate(layout.bns)}
        for name, (off, ch) in self.bn_off.items():
            self.running_mean[off: off + ch] = state[name + ".running_mean"].to(dev)
            self.running_var[off: off + ch] = state[name + ".running_var"].to(dev)
            self.num_batches[self.bn_index[name]] = int(state[name + ".num_batches_tracked"])
        self.bn_mean = torch.zeros(o, **f32)
        self.bn_invstd = torch.zeros(o, **f32)
        self.bn_scratch = torch.zeros(ops.bn_scratch_floats(512), **f32)
        # statistics + apply (and reduce + apply) as one launch each (blocks wait in-kernel for the per-channel result).
        # Off: measured 2.29 against 2.01 ms/step — the waiting blocks keep their shared memory, so the side stream's
        # wgrad GEMM CTAs cannot start under them, and the apply phase runs with the reduction's narrow grid
        # (profiles/r02_bn_fused_ab.md)
        self.bn_fused = os.environ.get("B200STEP_BN_FUSED", "0") == "1"

        # ---- activations (NHWC matrices) ------------------------------------------------------------------------
        convs = [layout.stem] + [c for b in layout.blocks for c in (b.conv1, b.conv2, b.down) if c is not None]
        self.convs = {c.name: c for c in convs}
        self.x0 = torch.zeros(B * layout.image * layout.image, 3, **act)
        self.y = torch.zeros(B, dtype=torch.int64, device=dev)
        # Implicit GEMM (b200_conv_fwd / b200_conv_wgrad: A / B tiles by im2col-mode TMA straight from the NHWC
        # activation) for every conv whose channel count fills whole 128-byte k-blocks; conv1 (C = 3) and the 3xTF32
        # mode (which needs hi/lo copies of the unfolded operand) keep the explicit col matrix.
        vec = 64 if self.dt == torch.bfloat16 else 32
        self.implicit = {c.name: (not self.x3) and c.cin % vec == 0 and c.K == c.Kp and
                         os.environ.get("B200STEP_IMPLICIT_CONV", "1") != "0" for c in convs}
        # 3xTF32 mode, implicit too: the hi/lo copies are taken of the ACTIVATION (not of its 9x larger unfolding) and
        # laid out so that the unchanged implicit-GEMM entry points see them as a longer contraction:
        #   forward  x -> [hi | hi | lo] along the channels (3C input channels), filter rows [tap][hi | lo | hi]
        #   wgrad    x -> [hi ; lo ; hi] stacked as 3B images against dy [hi ; hi ; lo] (contraction over 3M pixels)
        #   dgrad    dy -> [hi | hi | lo] along Cout (3 Cout channels) against the filter rows stacked [hi ; lo ; hi]
        self.implicit3 = {c.name: self.x3 and c.cin % 32 == 0 and c.K == c.Kp and
                          os.environ.get("B200STEP_IMPLICIT_CONV", "1") != "0" for c in convs}
        self.col = {c.name: torch.zeros(B * c.Ho * c.Wo, c.Kp, **act) for c in convs
                    if not (self.implicit[c.name] or self.implicit3[c.name])}
        self.xin: Dict[str, torch.Tensor] = {}          # implicit convs: the input activation seen in forward
        # fp32 workspace for the split contraction of the small-M convolutions (main stream only); the layers that
        # split have <= 37 output tiles of 256 x 256
        self.conv_ws = torch.zeros(37 * 256 * 256, **f32) if (self.dt == torch.bfloat16 and
                                                              os.environ.get("B200STEP_CONV_SPLIT", "1") != "0") else None
        self.yc = {c.name: torch.zeros(B * c.Ho * c.Wo, c.cout, **act) for c in convs}      # conv output (pre-BN)
        self.a = {c.name: torch.zeros(B * c.Ho * c.Wo, c.cout, **act) for c in convs if ".downsample" not in c.name}
        self.sc = {b.name: torch.zeros_like(self.yc[b.down.name]) for b in layout.blocks if b.down is not None}
        st = layout.stem
        self.pool_hw = ops.conv_out_size(st.Ho, 3, 2, 1)
        self.p0 = torch.zeros(B * self.pool_hw * self.pool_hw, 64, **act)
        self.taps = torch.zeros(B * self.pool_hw * self.pool_hw, 64, dtype=torch.uint8, device=dev)
        self.feat = torch.zeros(B, 512, **act) if layout.pool_hw > 1 else None
        # ---- gradients of activations: scratch shared by all layers (sizes: the largest user)
        big = max(B * c.Ho * c.Wo * c.Kp for c in convs if c is not layout.stem)
        self.dcol = torch.zeros(big, **act)
        amax = max(B * c.Ho * c.Wo * c.cout for c in convs)
        xin = max([B * c.H * c.W * c.cin for c in convs if c is not layout.stem])
        # gradient wrt each conv output: one buffer per conv, because its wgrad GEMM runs on a side stream while the
        # main stream already produces the next layer's gradient
        self.dyc = {c.name: torch.zeros(B * c.Ho * c.Wo, c.cout, **act) for c in convs}
        self.dz2 = [torch.zeros(amax, **act) for _ in range(2)]   # gradient wrt a block output after its ReLU mask
        self.dact = [torch.zeros(max(amax, xin), **act) for _ in range(3)]   # rotating: block-output / mid grads
        self.dlogits = torch.zeros(B, _round_up(layout.num_classes, 8), **act)
        self.dfeat = torch.zeros(B, 512, **act)
        self.loss = torch.zeros(1, **f32)
        self.ce_scratch = torch.zeros(ops.linear_ce_scratch_floats(layout.num_classes), **f32)

        # ---- 3xTF32 operand copies (precision fp32)
        if self.x3:
            self.wk = {c.name: torch.zeros(c.cout, 3 * c.Kp, **f32) for c in convs}
            self.wr = {c.name: torch.zeros(3 * c.cout, c.Kp, **f32) for c in convs}
            self.wk["fc"] = torch.zeros(layout.num_classes, 3 * 512, **f32)
            self.wr["fc"] = torch.zeros(3 * layout.num_classes, 512, **f32)
            ex = [c for c in convs if not self.implicit3[c.name]]
            self.colr = {c.name: torch.zeros(3 * B * c.Ho * c.Wo, c.Kp, **f32) for c in ex}
            self.colk = torch.zeros(max([B * c.Ho * c.Wo * 3 * c.Kp for c in ex] + [4]), **f32)
            im = [c for c in convs if self.implicit3[c.name]]
            self.wk3 = {c.name: torch.zeros(c.cout, 3 * c.K, **f32) for c in im}        # [Cout][tap][hi | lo | hi]
            self.xr3 = {c.name: torch.zeros(3 * B * c.H * c.W, c.cin, **f32) for c in im}   # kept for wgrad
            self.xk3 = torch.zeros(max([B * c.H * c.W * 3 * c.cin for c in im] + [4]), **f32)
            self.dyk = torch.zeros(3 * max(amax, B * self.dlogits.shape[1]), **f32)
            self.dyr = torch.zeros(3 * max(amax, B * self.dlogits.shape[1]), **f32)
            self.featk = torch.zeros(B, 3 * 512, **f32)
            self.featr = torch.zeros(3 * B, 512, **f32)

        # ---- streams, step bookkeeping, DDP buckets
        self.main = torch.cuda.Stream(device=dev)
        self.comm_stream = torch.cuda.Stream(device=dev)
        self.side = torch.cuda.Stream(device=dev)          # filter-gradient GEMMs, off the backward chain's critical path
        self.side_wgrad = (not self.x3) and os.environ.get("B200STEP_SIDE_WGRAD", "1") != "0"
        self.side_wgrad3 = self.x3 and os.environ.get("B200STEP_SIDE_WGRAD", "1") != "0"
        if self.side_wgrad3:     # the side stream reads dy's [hi ; hi ; lo] copy after the main stream has moved on
            self.dyr3 = {c.name: torch.zeros(3 * B * c.Ho * c.Wo, c.cout, **f32) for c in convs}
        self._ev_dy = {c.name: torch.cuda.Event() for c in convs}
        self._ev_wg = {c.name: torch.cuda.Event() for c in convs}
        self.cursor = torch.zeros(1, dtype=torch.int32, device=dev)
        self.step_ctr = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ticket = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status = ops.status_word(dev)
        self._make_buckets()
        self.graph = None
        self.host_steps = 0
        self.kernels_per_step = 0
        if self.x3:
            with torch.cuda.stream(self.main):
                self._split_weights()
        torch.cuda.synchronize(dev)

    # ------------------------------------------------------------------------------------------ views
    def _pv(self, key: str, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        s = self.lay.seg(key)
        t = (self.p if buf is None else buf)[s.offset: s.offset + s.numel]
        if s.conv is not None:
            return t.view(s.conv.cout, s.conv.Kp)
        return t.view(s.shape) if len(s.shape) == 2 else t

    def _w(self, name: str) -> torch.Tensor:
        """GEMM operand of conv `name` / "fc" in the compute dtype."""
        return self._pv(name + ".weight", self.shadow if self.shadow is not None else None)

    def _bn(self, name: str):
        off, ch = self.bn_off[name]
        return (self.bn_mean[off: off + ch], self.bn_invstd[off: off + ch], self.running_mean[off: off + ch],
                self.running_var[off: off + ch], self.num_batches[self.bn_index[name]:])

    def _split_weights(self) -> int:
        names = list(self.convs) + ["fc"]
        for nm in names:
            w = self._pv(nm + ".weight")
            if self.implicit3.get(nm, False):
                c = self.convs[nm]          # one launch: rows of the filter seen as [Cout*taps, C]
                rows = c.cout * c.k * c.k
                ops.split_tf32(w.view(rows, c.cin), self.wk3[nm].view(rows, 3 * c.cin), ops.SPLIT_RIGHT,
                               self.wr[nm].view(3 * rows, c.cin), ops.SPLIT_RIGHT, rows=rows, cols=c.cin,
                               stream=self.main)
            else:
                ops.split_tf32(w, self.wk[nm], ops.SPLIT_RIGHT, self.wr[nm], ops.SPLIT_RIGHT, rows=w.shape[0],
                               cols=w.shape[1], stream=self.main)
        return len(names)

    def _make_buckets(self) -> None:
        """DDP's gradient buckets (task.py:189; reducer defaults: first bucket 1 MiB, then 25 MiB), filled in reverse
        parameter order = the order backward produces gradients.  A bucket is a contiguous range of the flat buffer."""
        cfg = self.cfg
        plan = ddp_bucket_plan(self.lay, getattr(cfg, "first_bucket_bytes", 1 << 20),
                               getattr(cfg, "bucket_cap_bytes", 25 << 20))
        self.buckets: List[List[int]] = []          # [start, end, params outstanding]
        self.bucket_of: Dict[str, int] = {}
        end = self.lay.n_params
        for keys in plan:
            start = self.lay.seg(keys[-1]).offset       # keys run from the last parameter of the bucket to the first
            for k in keys:
                self.bucket_of[k] = len(self.buckets)
            self.buckets.append([start, end, len(keys)])
            end = start
        self._ev_ready = [torch.cuda.Event() for _ in self.buckets]
        self._ev_done = [torch.cuda.Event() for _ in self.buckets]

    def _grad_done(self, key: str) -> int:
        """Called once the gradient of parameter `key` is complete: all-reduce its bucket when it was the last one."""
        if self.cfg.world <= 1:
            return 0
        bi = self.bucket_of[key]
        self._pending[bi] -= 1
        if self._pending[bi] > 0:
            return 0
        s, e, _ = self.buckets[bi]
        self._ev_ready[bi].record(self.main)
        self.comm_stream.wait_event(self._ev_ready[bi])
        for k, ev in self._side_pending:                # filter gradients of this bucket still running on the side stream
            if self.bucket_of[k] == bi:
                self.comm_stream.wait_event(ev)
        self.comm.allreduce_sum(self.g[s:e], self.comm_stream)
        self._ev_done[bi].record(self.comm_stream)
        return 1

    # ------------------------------------------------------------------------------------------ GEMM helpers
    def _conv_fwd(self, c: ConvDesc, x: torch.Tensor, rows_b: int) -> int:
        """yc[c] = conv(x) for `rows_b` images: im2col + GEMM (+ operand split in fp32 mode)."""
        M = rows_b * c.Ho * c.Wo
        if self.implicit[c.name]:
            self.xin[c.name] = x
            ops.conv_fwd(x, self._w(c.name), self.yc[c.name][:M], B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k,
                         stride=c.stride, pad=c.pad, Cout=c.cout, deterministic=self.det, ws=self.conv_ws,
                         stream=self.main)
            return 1
        if self.implicit3[c.name]:
            Min = rows_b * c.H * c.W
            xk = self.xk3[: Min * 3 * c.cin].view(Min, 3 * c.cin)
            ops.split_tf32(x, xk, ops.SPLIT_LEFT, self.xr3[c.name][: 3 * Min], ops.SPLIT_RIGHT, rows=Min, cols=c.cin,
                           stream=self.main)
            ops.conv_fwd(xk, self.wk3[c.name], self.yc[c.name][:M], B=rows_b, H=c.H, W=c.W, C=3 * c.cin, kh=c.k, kw=c.k,
                         stride=c.stride, pad=c.pad, Cout=c.cout, deterministic=self.det, stream=self.main)
            return 2
        col, y = self.col[c.name][:M], self.yc[c.name][:M]
        ops.im2col(x, col, B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k, stride=c.stride, pad=c.pad, stream=self.main)
        n = 2
        if self.x3:
            colk = self.colk[: M * 3 * c.Kp].view(M, 3 * c.Kp)
            ops.split_tf32(col, colk, ops.SPLIT_LEFT, self.colr[c.name][: 3 * M], ops.SPLIT_RIGHT, rows=M, cols=c.Kp,
                           stream=self.main)
            ops.linear_fwd(colk, self.wk[c.name], None, y, M=M, N=c.cout, K=3 * c.Kp, relu=False, stream=self.main)
            n += 1
        else:
            ops.linear_fwd(col, self._w(c.name), None, y, M=M, N=c.cout, K=c.Kp, relu=False, stream=self.main)
        return n

    def _implicit_dgrad(self, c: ConvDesc) -> bool:
        """b200_conv_dgrad applies: implicit conv, stride 1, Cout in whole 128-byte k-blocks."""
        vec = 64 if self.dt == torch.bfloat16 else 32
        return (self.implicit[c.name] or self.implicit3[c.name]) and c.stride == 1 and c.cout % vec == 0 and \
            os.environ.get("B200STEP_IMPLICIT_DGRAD", "1") != "0"

    def _conv_bwd_input(self, c: ConvDesc, dy: torch.Tensor, rows_b: int, dx: torch.Tensor, add, dyk=None) -> int:
        """dx = gradient of conv c wrt its input (+ add).  Implicit GEMM over dy when it applies, else dgrad GEMM into
        the dcol scratch + col2im."""
        M = rows_b * c.Ho * c.Wo
        if add is None and self._implicit_dgrad(c) and self.x3:
            n = 1
            if dyk is None:
                dyk = self.dyk[: M * 3 * c.cout].view(M, 3 * c.cout)
                ops.split_tf32(dy, dyk, ops.SPLIT_LEFT, None, 0, rows=M, cols=c.cout, stream=self.main)
                n += 1
            ops.conv_dgrad(dyk, self.wr[c.name], dx, B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k, pad=c.pad,
                           Cout=3 * c.cout, deterministic=self.det, stream=self.main)
            return n
        if add is None and self._implicit_dgrad(c):
            ops.conv_dgrad(dy, self._w(c.name), dx, B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k, pad=c.pad,
                           Cout=c.cout, deterministic=self.det, ws=self.conv_ws, stream=self.main)
            return 1
        n = 2
        dcol = self.dcol[: M * c.Kp].view(M, c.Kp)
        if self.x3:
            if dyk is None:
                dyk = self.dyk[: M * 3 * c.cout].view(M, 3 * c.cout)
                ops.split_tf32(dy, dyk, ops.SPLIT_LEFT, None, 0, rows=M, cols=c.cout, stream=self.main)
                n += 1
            ops.linear_dgrad(dyk, self.wr[c.name], None, dcol, M=M, N=3 * c.cout, K=c.Kp, relu=False, stream=self.main)
        else:
            ops.linear_dgrad(dy, self._w(c.name), None, dcol, M=M, N=c.cout, K=c.Kp, relu=False, stream=self.main)
        ops.col2im(dcol, add, dx, B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k, stride=c.stride, pad=c.pad,
                   stream=self.main)
        return n

    def _conv_bwd(self, c: ConvDesc, dy: torch.Tensor, rows_b: int, dx: Optional[torch.Tensor], add,
                  defer_input: bool = False) -> int:
        """Backward of conv c given dy: the filter gradient into the flat gradient buffer and, when `dx` is given, the
        input gradient (+ add) into dx.  `defer_input`: only the filter gradient now (the caller issues
        `_conv_bwd_input` later, e.g. after the shortcut branch produced `add`)."""
        M = rows_b * c.Ho * c.Wo
        gw = self._pv(c.name + ".weight", self.g)
        n = 0
        if self.x3:
            dyk = self.dyk[: M * 3 * c.cout].view(M, 3 * c.cout)
            side = self.side_wgrad3
            dyr = (self.dyr3[c.name][: 3 * M] if side else self.dyr[: 3 * M * c.cout].view(3 * M, c.cout))
            ops.split_tf32(dy, dyk, ops.SPLIT_LEFT, dyr, ops.SPLIT_LEFT, rows=M, cols=c.cout, stream=self.main)
            n += 1
            wstream = self.main
            if side:             # fork: the filter gradient (its own hi/lo copy of dy) runs beside the backward chain
                self._ev_dy[c.name].record(self.main)
                self.side.wait_event(self._ev_dy[c.name])
                wstream = self.side
            if self.implicit3[c.name]:
                ops.conv_wgrad(dyr, self.xr3[c.name][: 3 * rows_b * c.H * c.W], gw, B=3 * rows_b, H=c.H, W=c.W, C=c.cin,
                               kh=c.k, kw=c.k, stride=c.stride, pad=c.pad, Cout=c.cout, deterministic=self.det,
                               stream=wstream)
            else:
                ops.linear_wgrad(dyr, self.colr[c.name][: 3 * M], gw, M=3 * M, N=c.cout, K=c.Kp, deterministic=self.det,
                                 stream=wstream)
            if side:
                self._ev_wg[c.name].record(self.side)
                self._side_pending.append((c.name + ".weight", self._ev_wg[c.name]))
            if dx is not None and not defer_input:
                n += self._conv_bwd_input(c, dy, rows_b, dx, add, dyk)
        else:
            wstream = self.main
            if self.side_wgrad:
                # dy is complete here: fork.  The wgrad GEMM only feeds the optimizer (and its DDP bucket), so it runs
                # on the side stream under the dgrad / col2im / BatchNorm-backward chain, whose small kernels leave
                # most SMs idle; the join before K-opt / the bucket's all-reduce wait for its event.
                self._ev_dy[c.name].record(self.main)
                self.side.wait_event(self._ev_dy[c.name])
                wstream = self.side
            if dx is not None and not defer_input:
                n += self._conv_bwd_input(c, dy, rows_b, dx, add)
            if self.implicit[c.name]:
                ops.conv_wgrad(dy, self.xin[c.name], gw, B=rows_b, H=c.H, W=c.W, C=c.cin, kh=c.k, kw=c.k,
                               stride=c.stride, pad=c.pad, Cout=c.cout, deterministic=self.det, stream=wstream)
            else:
                ops.linear_wgrad(dy, self.col[c.name][:M], gw, M=M, N=c.cout, K=c.Kp, deterministic=self.det,
                                 stream=wstream)
            if self.side_wgrad:
                self._ev_wg[c.name].record(self.side)
                self._side_pending.append((c.name + ".weight", self._ev_wg[c.name]))
        n += 1 + self._grad_done(c.name + ".weight")
        return n

    def _bn_fwd(self, bn: str, y: torch.Tensor, res, out: torch.Tensor, relu: bool, training: bool) -> int:
        M, C = y.shape
        mean, invstd, rm, rv, nb = self._bn(bn)
        n = 1
        if training and self.bn_fused:
            ops.bn_fwd(y, res, out, mean, invstd, self._pv(bn + ".weight"), self._pv(bn + ".bias"), rm, rv, nb,
                       self.bn_scratch, M=M, C=C, relu=relu, stream=self.main)
            return 1
        if training:
            ops.bn_fwd_stats(y, mean, invstd, rm, rv, nb, self.bn_scratch, M=M, C=C, stream=self.main)
            n += 1
        else:
            mean, invstd = rm, self._eval_invstd[self.bn_off[bn][0]: self.bn_off[bn][0] + C]
        ops.bn_fwd_apply(y, res, out, mean, invstd, self._pv(bn + ".weight"), self._pv(bn + ".bias"), M=M, C=C,
                         relu=relu, stream=self.main)
        return n

    def _bn_bwd(self, bn: str, dout: torch.Tensor, out, y: torch.Tensor, dy: torch.Tensor, dz, dout2=None) -> int:
        M, C = y.shape
        mean, invstd, _, _, _ = self._bn(bn)
        dg, db = self._pv(bn + ".weight", self.g), self._pv(bn + ".bias", self.g)
        if self.bn_fused:
            ops.bn_bwd(dout, out, y, dy, dz, mean, invstd, self._pv(bn + ".weight"), dg, db, self.bn_scratch, M=M, C=C,
                       dout2=dout2, stream=self.main)
            return 1 + self._grad_done(bn + ".bias") + self._grad_done(bn + ".weight")
        ops.bn_bwd_reduce(dout, out, y, mean, invstd, dg, db, self.bn_scratch, M=M, C=C, dout2=dout2, stream=self.main)
        ops.bn_bwd_apply(dout, out, y, dy, dz, mean, invstd, self._pv(bn + ".weight"), dg, db, M=M, C=C, dout2=dout2,
                         stream=self.main)
        return 2 + self._grad_done(bn + ".bias") + self._grad_done(bn + ".weight")

    # ------------------------------------------------------------------------------------------ the schedule
    def _load(self, b: int, row_idx, cursor, draws, ds: Optional[ImageDataset] = None) -> int:
        ds = ds or self.ds
        img = self.lay.image
        ops.image_augment(ds.images, ds.labels, row_idx, cursor, draws, self.mean3, self.std3, self.x0, self.y, batch=b,
                          H=img, W=img, pad=4, cursor_stride=self.B if cursor is not None else 0, stream=self.main)
        return 1

    def _forward(self, b: int, training: bool) -> int:
        """model(inputs) (task.py:309) for the b images in x0; returns the number of launches."""
        lay, st = self.lay, self.lay.stem
        n = self._conv_fwd(st, self.x0, b)
        M0 = b * st.Ho * st.Wo
        n += self._bn_fwd("bn1", self.yc[st.name][:M0], None, self.a[st.name][:M0], True, training)
        ph = self.pool_hw
        ops.maxpool_fwd(self.a[st.name], self.p0, self.taps, B=b, H=st.Ho, W=st.Wo, C=64, stream=self.main)
        n += 1
        x = self.p0[: b * ph * ph]
        for blk in lay.blocks:
            c1, c2, dn = blk.conv1, blk.conv2, blk.down
            M = b * c1.Ho * c1.Wo
            n += self._conv_fwd(c1, x, b)
            n += self._bn_fwd(c1.bn, self.yc[c1.name][:M], None, self.a[c1.name][:M], True, training)
            n += self._conv_fwd(c2, self.a[c1.name], b)
            if dn is not None:
                n += self._conv_fwd(dn, x, b)
                n += self._bn_fwd(dn.bn, self.yc[dn.name][:M], None, self.sc[blk.name][:M], False, training)
                res = self.sc[blk.name][:M]
            else:
                res = x
            n += self._bn_fwd(c2.bn, self.yc[c2.name][:M], res, self.a[c2.name][:M], True, training)
            x = self.a[c2.name][:M]
        if lay.pool_hw > 1:
            ops.avgpool_fwd(x, self.feat, B=b, HW=lay.pool_hw, C=512, stream=self.main)
            n += 1
            x = self.feat[:b]
        self._features = x
        return n

    def _enqueue(self, b: int, row_idx, cursor, draws, advance: bool, ds: Optional[ImageDataset] = None) -> int:
        """One whole step for b images: load + augment, forward, loss, backward (+ bucket all-reduces), optimizer."""
        lay, cfg = self.lay, self.cfg
        C = lay.num_classes
        self._pending = [bk[2] for bk in self.buckets]
        self._side_pending = []
        n = self._load(b, row_idx, cursor, draws, ds)
        n += self._forward(b, True)
        feat = self._features
        # ---- output layer + CrossEntropyLoss (task.py:309-310) and the start of backward (task.py:311)
        scale = 1.0 / (b * cfg.world)
        dl = self.dlogits[:b]
        gb = self._pv("fc.bias", self.g)
        if self.x3:
            ops.split_tf32(feat, self.featk[:b], ops.SPLIT_LEFT, self.featr[: 3 * b], ops.SPLIT_RIGHT, rows=b, cols=512,
                           stream=self.main)
            ops.linear_ce(self.featk[:b], self.wk["fc"], self._pv("fc.bias"), self.y, dl, self.loss, gb,
                          self.ce_scratch, M=b, N=C, K=3 * 512, grad_scale=scale, stream=self.main)
            n += 1
        else:
            ops.linear_ce(feat, self._w("fc"), self._pv("fc.bias"), self.y, dl, self.loss, gb, self.ce_scratch, M=b, N=C,
                          K=512, grad_scale=scale, stream=self.main)
        n += 2 + self._grad_done("fc.bias")              # K-gemm-ce + its partial reduction
        gw = self._pv("fc.weight", self.g)
        dfeat = self.dfeat[:b]
        if self.x3:
            dyk = self.dyk[: b * 3 * C].view(b, 3 * C)
            dyr = self.dyr[: 3 * b * C].view(3 * b, C)
            ops.split_tf32(dl, dyk, ops.SPLIT_LEFT, dyr, ops.SPLIT_LEFT, rows=b, cols=C, stream=self.main)
            ops.linear_dgrad(dyk, self.wr["fc"], None, dfeat, M=b, N=3 * C, K=512, relu=False, stream=self.main)
            ops.linear_wgrad(dyr, self.featr[: 3 * b], gw, M=3 * b, N=C, K=512, deterministic=self.det, stream=self.main)
            n += 1
        else:
            ops.linear_dgrad(dl, self._w("fc"), None, dfeat, M=b, N=C, K=512, relu=False, stream=self.main)
            ops.linear_wgrad(dl, feat, gw, M=b, N=C, K=512, deterministic=self.det, stream=self.main)
        n += 2 + self._grad_done("fc.weight")
        # ---- backward through the blocks
        last = lay.blocks[-1].conv2
        Ml = b * last.Ho * last.Wo
        if lay.pool_hw > 1:
            dout = self.dact[0][: Ml * 512].view(Ml, 512)
            ops.avgpool_bwd(dfeat, dout, B=b, HW=lay.pool_hw, C=512, stream=self.main)
            n += 1
        else:
            dout = dfeat
        dout2 = None                       # second branch of the incoming gradient (summed by its consumer)
        for bi, blk in enumerate(reversed(lay.blocks)):
            c1, c2, dn = blk.conv1, blk.conv2, blk.down
            M, Cc = b * c2.Ho * c2.Wo, c2.cout
            free = [t for t in self.dact if t.data_ptr() != dout.data_ptr()]     # dout sits in one of the three
            mid, nxt = free[0], free[1]
            dy = self.dyc[c2.name][:M]
            dz = self.dz2[bi & 1][: M * Cc].view(M, Cc)     # alternating: the previous block's dz may be `dout2` here
            n += self._bn_bwd(c2.bn, dout, self.a[c2.name][:M], self.yc[c2.name][:M], dy, dz, dout2)
            d_a1 = mid[: M * Cc].view(M, Cc)
            n += self._conv_bwd(c2, dy, b, d_a1, None)
            dy = self.dyc[c1.name][:M]
            n += self._bn_bwd(c1.bn, d_a1, self.a[c1.name][:M], self.yc[c1.name][:M], dy, None)
            Min = b * c1.H * c1.W
            dx = nxt[: Min * c1.cin].view(Min, c1.cin)
            if dn is not None:
                # projection shortcut: its own BatchNorm + 1x1 stride-2 conv backward, then both branches into dx
                n += self._conv_bwd(c1, dy, b, None, None, defer_input=True)
                dyd = self.dyc[dn.name][:M]
                n += self._bn_bwd(dn.bn, dz, None, self.yc[dn.name][:M], dyd, None)
                add = mid[: Min * c1.cin].view(Min, c1.cin)        # d_a1 is consumed: reuse its buffer
                n += self._conv_bwd(dn, dyd, b, add, None)
                n += self._conv_bwd_input(c1, dy, b, dx, add)
                dout, dout2 = dx, None
            elif self._implicit_dgrad(c1):
                n += self._conv_bwd(c1, dy, b, dx, None)             # dx = main branch only ...
                dout, dout2 = dx, dz                                 # ... the identity branch (dz) is added by the consumer
            else:
                n += self._conv_bwd(c1, dy, b, dx, dz)
                dout, dout2 = dx, None
        # ---- stem: maxpool, bn1 (+ReLU), conv1's filter gradient (the input needs no gradient)
        st = lay.stem
        M0 = b * st.Ho * st.Wo
        free = [t for t in self.dact if t.data_ptr() != dout.data_ptr()]
        d_a0 = free[0][: M0 * 64].view(M0, 64)
        ops.maxpool_bwd(dout, self.taps, d_a0, B=b, H=st.Ho, W=st.Wo, C=64, dout2=dout2, stream=self.main)
        dy0 = self.dyc[st.name][:M0]
        n += 1 + self._bn_bwd("bn1", d_a0, self.a[st.name][:M0], self.yc[st.name][:M0], dy0, None)
        n += self._conv_bwd(st, dy0, b, None, None)
        # ---- optimizer.step() (task.py:312) once every bucket is reduced; the gradients are left zeroed (task.py:308)
        for _, ev in self._side_pending:                 # join: every filter gradient is complete
            self.main.wait_event(ev)
        if cfg.world > 1:
            for ev in self._ev_done:
                self.main.wait_event(ev)
        adv = (self.cursor if advance else None, self.step_ctr, max(1, self.steps_per_epoch_full), self.ticket)
        ns = self.lay.n_params if self.shadow is not None else 0
        if cfg.optimizer == "sgd":
            ops.sgd_step(self.p, self.g, self.s1, self.shadow, n_shadow=ns, lr=cfg.lr, momentum=cfg.momentum,
                         weight_decay=cfg.weight_decay, zero_grad=True, advance=adv, stream=self.main)
        else:
            ops.adam_step(self.p, self.g, self.s1, self.s2, self.shadow, self.step_ctr, n_shadow=ns, lr=cfg.lr,
                          beta1=cfg.beta1, beta2=cfg.beta2, eps=cfg.eps, weight_decay=cfg.weight_decay, zero_grad=True,
                          advance=adv, stream=self.main)
        n += 1
        if self.x3:
            n += self._split_weights()
        return n

    # ------------------------------------------------------------------------------------------ public API
    def set_draws(self, draws: torch.Tensor) -> None:
        """This epoch's augmentation draws for this rank's samples in sampler order (`draw_augmentation`)."""
        if tuple(draws.shape) != (self.num_samples, 4):
            raise ValueError(f"draws must be int8 [{self.num_samples}, 4]")
        with torch.cuda.stream(self.main):
            self.draws.copy_(draws.to(torch.int8), non_blocking=False)

    def set_row_order(self, row_idx: torch.Tensor) -> None:
        """DistributedSampler.set_epoch for the resident path (same buffer, so the captured graph stays valid)."""
        if row_idx.numel() != self.num_samples:
            raise ValueError(f"sample order has {row_idx.numel()} entries, this rank holds {self.num_samples}")
        with torch.cuda.stream(self.main):
            self.row_idx.copy_(row_idx.to(torch.int64))

    def _draws(self):
        return self.draws if self.augment else None

    def capture(self) -> None:
        if self.graph is not None:
            return
        import ctypes as C

        self.main.synchronize()
        self.comm_stream.synchronize()
        self.side.synchronize()
        self.lib.graph_begin(self.main.cuda_stream)
        try:
            self.kernels_per_step = self._enqueue(self.B, self.row_idx, self.cursor, self._draws(), True)
        finally:
            out = C.c_void_p()
            self.lib.graph_end(self.main.cuda_stream, C.byref(out))
        self.graph = out

    def step(self) -> None:
        """One optimizer step on this rank's next full batch (asynchronous)."""
        if self.steps_per_epoch_full == 0:
            raise B200Error("batch_size larger than this rank's sample count: use step_tail()")
        if self.cfg.use_graph and self.host_steps > 0:
            if self.graph is None:
                self.capture()
            self.lib.graph_launch(self.graph, self.main.cuda_stream)
        else:
            self.kernels_per_step = self._enqueue(self.B, self.row_idx, self.cursor, self._draws(), True)
        self.host_steps += 1

    def step_tail(self) -> None:
        """The last, partial batch of an epoch (DataLoader drop_last=False, task.py:263)."""
        if self.tail == 0:
            return
        off = self.steps_per_epoch_full * self.B
        d = self._draws()
        self._enqueue(self.tail, self.row_idx[off:], None, d[off:] if d is not None else None, False)
        self.host_steps += 1

    # ---- host-fed path: what the reference's DataLoader + .cuda() copies do (task.py:303-307) ---------------------
    HOST_SLOTS = 4

    def host_batch_bytes(self) -> int:
        img = self.lay.image
        return self.B * (img * img * 3 + 8 + 4)

    def step_from_host(self, images_u8: torch.Tensor, labels: torch.Tensor, draws: torch.Tensor) -> int:
        """One step on a batch that lives in (pinned) host memory: raw uint8 images, int64 labels and the int8
        augmentation draws cross PCIe (3 084 bytes per sample), everything else happens on the device.  Returns a
        ticket for `host_loss`."""
        B = self.B
        if images_u8.shape[0] != B:
            raise ValueError(f"host batches must have batch_size={B} rows")
        if not hasattr(self, "_stage"):
            dev = self.dev
            self._stage = ImageDataset(torch.empty(B, self.lay.image, self.lay.image, 3, dtype=torch.uint8, device=dev),
                                       torch.empty(B, dtype=torch.int64, device=dev))
            self._stage_draws = torch.empty(B, 4, dtype=torch.int8, device=dev)
            self._loss_host = torch.zeros(self.HOST_SLOTS, dtype=torch.float32).pin_memory()
            self._loss_ev = [torch.cuda.Event() for _ in range(self.HOST_SLOTS)]
            self._host_graph, self._host_ticket = None, 0
        with torch.cuda.stream(self.main):
            self._stage.images.copy_(images_u8, non_blocking=True)
            self._stage.labels.copy_(labels, non_blocking=True)
            self._stage_draws.copy_(draws, non_blocking=True)
        if self.cfg.use_graph and self.host_steps > 0:
            if self._host_graph is None:
                import ctypes as C

                self.main.synchronize()
                self.comm_stream.synchronize()
                self.side.synchronize()
                self.lib.graph_begin(self.main.cuda_stream)
                try:
                    self._enqueue(B, None, None, self._stage_draws, False, self._stage)
                finally:
                    out = C.c_void_p()
                    self.lib.graph_end(self.main.cuda_stream, C.byref(out))
                self._host_graph = out
            self.lib.graph_launch(self._host_graph, self.main.cuda_stream)
        else:
            self.kernels_per_step = self._enqueue(B, None, None, self._stage_draws, False, self._stage)
        k = self._host_ticket % self.HOST_SLOTS
        with torch.cuda.stream(self.main):
            self._loss_host[k: k + 1].copy_(self.loss, non_blocking=True)
        self._loss_ev[k].record(self.main)
        self.host_steps += 1
        self._host_ticket += 1
        return self._host_ticket - 1

    def host_loss(self, ticket: int) -> float:
        k = ticket % self.HOST_SLOTS
        self._loss_ev[k].synchronize()
        return float(self._loss_host[k])

    def loss_value(self) -> float:
        self.main.synchronize()
        return float(self.loss.item())

    def synchronize(self) -> None:
        self.main.synchronize()
        self.comm_stream.synchronize()
        self.side.synchronize()
        bits = int(self.status.item())
        if bits:
            self.status.zero_()
            what = []
            if bits & ops.STATUS_BAD_ROW_INDEX:
                what.append(f"a sample index outside [0, {len(self.ds)}) (Dataset.__getitem__ raises IndexError)")
            if bits & ops.STATUS_BAD_LABEL:
                what.append(f"a label outside [0, {self.lay.num_classes}) (nn.CrossEntropyLoss raises)")
            raise B200Error("the step met " + " and ".join(what or [f"status bits {bits}"]) + "; the offending rows "
                            "were neutralised on the device but the input is invalid")

    def evaluate(self, dataset: Optional[ImageDataset] = None, draws: Optional[torch.Tensor] = None,
                 return_predictions: bool = False):
        """Top-1 accuracy in eval mode (task.py:30-46: model.eval() -> BatchNorm uses its running statistics) over
        `dataset` (default: the resident one) in stored order.  `draws`: the augmentation the reference's test loader
        applies too (task.py:258 passes the training transform); None = centre crop, no flip."""
        ds = (dataset or self.ds).to(self.dev)
        n_rows, B = len(ds), self.B
        correct = torch.zeros((), dtype=torch.int64, device=self.dev)
        preds = torch.empty(n_rows, dtype=torch.int64, device=self.dev) if return_predictions else None
        logits = torch.empty(B, self.lay.num_classes, dtype=self.dt, device=self.dev)
        dr = draws.to(self.dev, dtype=torch.int8).contiguous() if draws is not None else None
        torch.cuda.synchronize(self.dev)
        with torch.cuda.stream(self.main):
            self._eval_invstd = torch.rsqrt(self.running_var + 1e-5)
            for s in range(0, n_rows, B):
                b = min(B, n_rows - s)
                order = torch.arange(s, s + b, dtype=torch.int64, device=self.dev)
                self._load(b, order, None, dr[s:] if dr is not None else None, ds)
                self._forward(b, False)
                f = self._features
                if self.x3:
                    ops.split_tf32(f, self.featk[:b], ops.SPLIT_LEFT, None, 0, rows=b, cols=512, stream=self.main)
                    ops.linear_fwd(self.featk[:b], self.wk["fc"], self._pv("fc.bias"), logits[:b], M=b,
                                   N=self.lay.num_classes, K=3 * 512, relu=False, stream=self.main)
                else:
                    ops.linear_fwd(f, self._w("fc"), self._pv("fc.bias"), logits[:b], M=b, N=self.lay.num_classes,
                                   K=512, relu=False, stream=self.main)
                pred = logits[:b].argmax(dim=1)                   # torch.max(outputs, 1), task.py:40
                correct += (pred == self.y[:b]).sum()
                if preds is not None:
                    preds[s: s + b] = pred
        self.main.synchronize()
        acc = correct.item() / n_rows
        return (acc, preds.cpu()) if return_predictions else acc

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Keys, shapes and order of the reference's `model.state_dict()` under DDP (task.py:283/287)."""
        self.synchronize()
        flat = self.p.detach().cpu()
        params = {}
        for s in self.lay.params:
            t = flat[s.offset: s.offset + s.numel]
            params[s.key] = gemm_to_filter(t.view(s.conv.cout, s.conv.Kp), s.shape) if s.conv is not None \
                else t.view(s.shape).clone()
        rm, rv, nb = self.running_mean.cpu(), self.running_var.cpu(), self.num_batches.cpu()
        out: Dict[str, torch.Tensor] = {}
        bn_names = {name for name, _ in self.lay.bns}
        for s in self.lay.params:                       # torch's module order: each BN's buffers follow its bias
            out["module." + s.key] = params[s.key]
            mod = s.key.rsplit(".", 1)[0]
            if s.key.endswith(".bias") and mod in bn_names:
                off, ch = self.bn_off[mod]
                out[f"module.{mod}.running_mean"] = rm[off: off + ch].clone()
                out[f"module.{mod}.running_var"] = rv[off: off + ch].clone()
                out[f"module.{mod}.num_batches_tracked"] = nb[self.bn_index[mod]].clone()
        return out

    def close(self) -> None:
        self.main.synchronize()
        self.comm_stream.synchronize()
        self.side.synchronize()
        if self.graph is not None:
            self.lib.graph_destroy(self.graph)
            self.graph = None
        if getattr(self, "_host_graph", None) is not None:
            self.lib.graph_destroy(self._host_graph)
            self._host_graph = None
