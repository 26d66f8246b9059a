"""Model description and flat parameter layout of the tabular MLP the B200 step trains.

The reference builds its model with `models.__dict__[args.arch]()` under `set_random_seeds` (task.py:22-28,
162, 166-171) and lets DDP/SGD see `model.parameters()` as a list of tensors.  Here the same parameters live
in ONE flat fp32 buffer (one gradient buffer, one optimizer-state buffer of the same layout) so that the
all-reduce and the optimizer are single streaming passes; `state_dict()` re-creates the per-tensor view with
the `module.`-prefixed keys the reference's `torch.save(model.state_dict())` writes (task.py:283/287).

Architectures (SURVEY.md §8 / BASELINE.json configs; additive values of task.py's --arch):
  tabular_mlp       T0: 13 dense + 26 x Embedding(1000, 16)  -> 429 -> 512 -> 512 -> 1000
  tabular_mlp_wide  T1: 13 dense + 26 x Embedding(65536, 64) -> 1677 -> 1024 x4 -> 1000
The 1000-way head over 10-class labels mirrors `models.resnet18()`'s default head on CIFAR-10 (task.py:171).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Tuple

import torch

ALIGN = 64  # elements: every segment of the flat buffer starts on a 256-byte boundary


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclasses.dataclass(frozen=True)
class ModelSpec:
    name: str = "tabular_mlp"
    num_dense: int = 13
    num_cat: int = 26
    cardinality: int = 1000
    embed_dim: int = 16
    hidden: Tuple[int, ...] = (512, 512)
    num_classes: int = 1000
    dropout: float = 0.0

    @property
    def in_features(self) -> int:
        return self.num_dense + self.num_cat * self.embed_dim

    @property
    def in_features_padded(self) -> int:
        # one 128-byte swizzle row of bf16 (64) / two of fp32: keeps every GEMM K a whole number of k-blocks
        return _round_up(self.in_features, 64)

    @property
    def layer_dims(self) -> List[Tuple[int, int]]:
        """(in_features, out_features) per Linear, unpadded."""
        dims = [self.in_features, *self.hidden, self.num_classes]
        return [(dims[i], dims[i + 1]) for i in range(len(dims) - 1)]

    @property
    def num_params(self) -> int:
        n = self.num_cat * self.cardinality * self.embed_dim
        for k, o in self.layer_dims:
            n += k * o + o
        return n

    def flops_per_sample(self) -> Dict[str, int]:
        """Algorithmic FLOPs (2*M*N*K on unpadded dims); dgrad of layer 0 only for the embedding columns."""
        fwd = sum(2 * k * o for k, o in self.layer_dims)
        wgrad = fwd
        dims = self.layer_dims
        dgrad = sum(2 * k * o for k, o in dims[1:]) + 2 * (self.num_cat * self.embed_dim) * dims[0][1]
        return {"fwd": fwd, "wgrad": wgrad, "dgrad": dgrad, "total": fwd + wgrad + dgrad}


ARCHS: Dict[str, ModelSpec] = {
    "tabular_mlp": ModelSpec(),
    "tabular_mlp_wide": ModelSpec(name="tabular_mlp_wide", cardinality=65536, embed_dim=64,
                                  hidden=(1024, 1024, 1024, 1024)),
}


@dataclasses.dataclass(frozen=True)
class Segment:
    key: str          # state_dict key without the DDP "module." prefix
    offset: int       # element offset in the flat buffer
    rows: int
    cols: int         # logical columns
    ld: int           # stored row stride (>= cols; K padding of the first Linear)

    @property
    def numel_stored(self) -> int:
        return self.rows * self.ld


class FlatLayout:
    """Where each parameter tensor lives in the flat buffer.

    Order: Linear weights/biases first (they also get a bf16 operand copy: `n_shadow` elements), embedding
    tables last.  The first Linear's weight is stored with its K dimension padded to `in_features_padded`
    (pad columns are zero and stay zero: their gradient is X_pad^T dY = 0 and weight decay of 0 is 0) and with
    its input columns in the order the preprocess kernel writes them (`input_perm`).
    """

    def __init__(self, spec: ModelSpec):
        self.spec = spec
        segs: List[Segment] = []
        off = 0
        for li, (k, o) in enumerate(spec.layer_dims):
            ld = spec.in_features_padded if li == 0 else k
            segs.append(Segment(f"layers.{li}.weight", off, o, k, ld))
            off = _round_up(off + o * ld, ALIGN)
            segs.append(Segment(f"layers.{li}.bias", off, 1, o, o))
            off = _round_up(off + o, ALIGN)
        self.n_shadow = off
        self.tables_offset = off
        for t in range(spec.num_cat):
            segs.append(Segment(f"emb.{t}.weight", off, spec.cardinality, spec.embed_dim, spec.embed_dim))
            off += spec.cardinality * spec.embed_dim
        self.tables_numel = off - self.tables_offset
        self.total = _round_up(off, ALIGN)
        self.segments = segs
        self.by_key = {s.key: s for s in segs}

    def weight(self, li: int) -> Segment:
        return self.by_key[f"layers.{li}.weight"]

    def bias(self, li: int) -> Segment:
        return self.by_key[f"layers.{li}.bias"]

    def view(self, flat: torch.Tensor, seg: Segment) -> torch.Tensor:
        """[rows, cols] strided view of a segment in STORED column order (biases: [cols])."""
        v = flat[seg.offset: seg.offset + seg.numel_stored].view(seg.rows, seg.ld)[:, : seg.cols]
        return v[0] if seg.key.endswith(".bias") else v

    def input_perm(self) -> torch.Tensor:
        """perm[j] = model (oracle / state_dict) input column held at stored column j of the preprocessed row x0 and
        of the first Linear's weight.  The model's order is [dense | embeddings] (torch.cat in TabularMLP.features);
        the step stores [embeddings | dense | pad] so that every gathered float4 is an aligned chunk of the row."""
        nd, ne = self.spec.num_dense, self.spec.num_cat * self.spec.embed_dim
        return torch.cat([torch.arange(nd, nd + ne), torch.arange(0, nd)])

    def flatten(self, tensors: Dict[str, torch.Tensor]) -> torch.Tensor:
        flat = torch.zeros(self.total, dtype=torch.float32)
        for seg in self.segments:
            t = tensors[seg.key]
            if seg.key == "layers.0.weight":
                t = t[:, self.input_perm()]
            self.view(flat, seg).copy_(t)
        return flat

    def unflatten(self, flat: torch.Tensor, prefix: str = "module.") -> Dict[str, torch.Tensor]:
        """state_dict in nn.Module order (embeddings, then layers), contiguous CPU copies."""
        out: Dict[str, torch.Tensor] = {}
        order = [s for s in self.segments if s.key.startswith("emb.")] + \
                [s for s in self.segments if s.key.startswith("layers.")]
        inv = torch.argsort(self.input_perm())
        for seg in order:
            t = self.view(flat, seg).detach().to("cpu")
            if seg.key == "layers.0.weight":
                t = t[:, inv]                      # back to the model's [dense | embeddings] column order
            out[prefix + seg.key] = t.contiguous().clone()
        return out

    def buckets(self, first_cap_bytes: int = 1 << 20, cap_bytes: int = 25 << 20) -> List[Tuple[int, int, int]]:
        """Gradient buckets in backward-readiness order, DDP-style caps (first 1 MiB, then 25 MiB;
        torch.distributed._DEFAULT_FIRST_BUCKET_BYTES / bucket_cap_mb).

        Returns (start, end, ready_after_layer): the bucket [start, end) of the flat buffer is complete once
        the backward of Linear `ready_after_layer` has run (-1 = after the embedding backward).
        Buckets are contiguous ranges, so one ncclAllReduce per bucket, in place.
        """
        L = len(self.spec.layer_dims)
        units: List[Tuple[int, int, int]] = []  # per layer, last layer first
        for li in range(L - 1, -1, -1):
            w, b = self.weight(li), self.bias(li)
            end = _round_up(b.offset + b.numel_stored, ALIGN)
            units.append((w.offset, end, li))
        if self.tables_numel:
            units.append((self.tables_offset, self.total, -1))
        out: List[Tuple[int, int, int]] = []
        cur_start = cur_end = None
        cap = first_cap_bytes
        ready = None
        for (s, e, li) in units:
            if cur_start is None:
                cur_start, cur_end, ready = s, e, li
            elif e == cur_start and (cur_end - s) * 4 <= cap:
                cur_start, ready = s, li       # contiguous downwards: merge
            else:
                out.append((cur_start, cur_end, ready))
                cap = cap_bytes
                cur_start, cur_end, ready = s, e, li
        if cur_start is not None:
            out.append((cur_start, cur_end, ready))
        # A single range above the cap (the wide model's 436 MB of embedding tables) is deliberately NOT cut into
        # cap-sized pieces: measured on 8 GPUs, 17 all-reduces of 25 MiB made the T1 step 4.77 ms against 3.73 ms
        # with one 436 MB all-reduce.
        return out


def init_parameters(spec: ModelSpec, seed: int) -> Dict[str, torch.Tensor]:
    """Initial parameters, created on the CPU by torch's own initialisers under `seed`.

    Mirrors task.py:162 (set_random_seeds) followed by task.py:171 (model construction): embeddings first
    (nn.Embedding -> N(0,1)), then the Linear layers (kaiming_uniform(a=sqrt(5)) weights, U(+-1/sqrt(fan_in))
    biases), each drawn in module-construction order from torch's global CPU generator.
    """
    import torch.nn as nn

    torch.manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    for t in range(spec.num_cat):
        out[f"emb.{t}.weight"] = nn.Embedding(spec.cardinality, spec.embed_dim).weight.detach()
    for li, (k, o) in enumerate(spec.layer_dims):
        lin = nn.Linear(k, o)
        out[f"layers.{li}.weight"] = lin.weight.detach()
        out[f"layers.{li}.bias"] = lin.bias.detach()
    return out
