"""TrainStep — the schedule of one data-parallel training step on one B200 (one process per GPU).

This is the body of the reference's hot loop (task.py:303-312) re-laid for the GPU:

    task.py:305,307  H2D copies of the batch          -> none: the dataset is resident in HBM; K-pre gathers rows
                                                         (`step_from_host` keeps the copies for host-fed batches)
    task.py:308      optimizer.zero_grad()            -> folded into the optimizer pass (it leaves the gradients zeroed)
    task.py:309      outputs = model(inputs)          -> K-pre, then one tcgen05 GEMM per Linear (bias/ReLU/dropout and
                                                         the ReLU bit-mask for backward fused into the epilogue)
    task.py:310      loss = criterion(outputs, y)     -> K-loss (loss + dlogits + output-layer bias gradient)
    task.py:311      loss.backward()  [+ DDP hooks]   -> dgrad GEMM chain (ReLU mask and bias-gradient column sums fused),
                                                         embedding scatter-add, then the wgrad GEMMs (split-K, fp32
                                                         red.add into the flat gradient buffer); each finished bucket is
                                                         all-reduced by NCCL on a side stream under the remaining wgrads
    task.py:312      optimizer.step()                 -> fused SGD/Adam pass over the flat buffers, one launch per bucket

All kernels are enqueued through the C ABI (libb200step.so) with programmatic dependent launch.  The steady-state
step (full batches) is captured once into a CUDA graph and replayed; the per-step variables (batch cursor, step
count) live in device memory.  PyTorch owns the memory and the streams and does nothing else here.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional

import torch

from . import ops
from .data import TabularDataset
from .lib import B200Error, get_lib
from .model import FlatLayout, ModelSpec


@dataclasses.dataclass
class StepConfig:
    batch_size: int                      # per process, like the reference's --batch_size
    world: int = 1
    rank: int = 0
    precision: str = "bf16"              # "bf16" (bf16 storage, fp32 accumulate/master) | "tf32" (fp32 storage, one
                                         # TF32 tensor-core pass, ~1e-3) | "fp32" (fp32 storage, three error-
                                         # compensated TF32 passes per product = fp32-grade, ~1e-6: task.py:309's
                                         # arithmetic)
    optimizer: str = "sgd"               # task.py:212-214 | additive "adam"
    lr: float = 0.1                      # task.py:59
    momentum: float = 0.9                # task.py:68
    weight_decay: float = 1e-4           # task.py:70
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    dropout_seed: int = 0
    use_graph: bool = True
    overlap_comm: bool = True            # all-reduce buckets on a side stream under the rest of backward
    first_bucket_bytes: int = 1 << 20    # DDP defaults
    bucket_cap_bytes: int = 25 << 20
    keep_logits: bool = False            # keep logits and dlogits in separate buffers (tests)
    deterministic_bias_grad: bool = False  # hidden-layer bias grads through a fixed-order reduction instead of atomics
    micro_rows: int = 0                  # rows per chunk of a step (0 = whole batch; must be a multiple of 8)
    comm_sm_reserve: Optional[int] = None  # None = DEFAULT_COMM_SM_RESERVE when world > 1 and comm overlaps
    prefetch_sm_reserve: int = 32        # SMs the prefetched K-pre leaves to the all-reduces it runs under
    l2_tile_order: bool = True           # alternate the row-tile direction of consecutive GEMMs (B200_FLAG_REVERSE_M) so
                                         # that each one starts with the rows its producer wrote last (still in L2)
    nccl_registered_grads: bool = True   # world > 1: flat gradient buffer from ncclMemAlloc + ncclCommRegister (in-place,
                                         # NVLS-capable all-reduce; measured on 8 GPUs: T1 3.71 -> 3.37 ms/step, T0 equal)
    prefetch_pre: Optional[bool] = None  # run K-pre of the NEXT batch inside this step (device-resident `step()`
                                         # path only).  world == 1: the table gradients are complete before the wgrad
                                         # GEMMs start, so a side stream updates the tables and gathers the next batch
                                         # UNDER the wgrad launch (x0 is double-buffered: wgrad still reads this
                                         # batch's).  world > 1: K-pre follows the embedding tables' optimizer launch,
                                         # under the all-reduces still in flight (+0.6 % at 2 GPUs, -2 % at 8 GPUs on
                                         # T0, profiles/r01_comm_overlap_experiments.md).  None / False = off
    prefetch_ctas_per_sm: int = 4        # CTA slots per SM the overlapped K-pre takes (the GEMM grid needs the rest)
    fused_ce: Optional[bool] = None      # output layer + cross-entropy as one kernel (K-gemm-ce): the logits never go to
                                         # HBM.  None = whenever it applies (<= 1024 classes, logits not kept)
    fused_emb_bwd: bool = False          # layer-0 dgrad adds its output straight onto the table gradients (no dx0
                                         # matrix, no separate scatter kernel).  Off by default: measured on T0 the
                                         # 6.8 M L2 atomics per step take as long from the GEMM epilogue as from the
                                         # scatter kernel (67 us fused vs 35 + 31 us) and 3x longer on Zipf ids;
                                         # K-embbwd's shared-memory privatisation is what cuts them
    deterministic_wgrad: bool = False    # convolutional path: filter gradients summed in a fixed order (no split of the
                                         # batch dimension, B200_FLAG_DETERMINISTIC): the whole step is then
                                         # bit-reproducible run to run, at the cost of wgrad occupancy
    emb_bwd_side: Optional[bool] = None  # world == 1: K-embbwd on a side stream under the grouped wgrad GEMM (None = on)
    grouped_wgrad: Optional[bool] = None # all layers' wgrad GEMMs in one launch.  None = only when world == 1:
                                         # equal at B = 65536, 3 % faster at B = 4096; per-layer launches let each
                                         # bucket's all-reduce start earlier when world > 1


# SMs the GEMMs that run while an all-reduce is in flight leave free (b200_set_sm_budget around those launches).  The
# GEMM grids are persistent and sized for every SM: when NCCL's CTAs take some, the displaced GEMM CTAs become a
# second wave.  Measured with scripts/gpu_allreduce_sweep.py (profiles/r01_comm_overlap_experiments.md): leaving SMs
# free cost more than it recovered on T0, hence 0.
DEFAULT_COMM_SM_RESERVE = 0


class Communicator:
    """NCCL communicator living inside libb200step.so (one per process)."""

    def __init__(self, rank: int, world: int, unique_id: bytes):
        self.rank, self.world = rank, world
        buf = (torch.frombuffer(bytearray(unique_id), dtype=torch.uint8)).clone()
        get_lib().comm_init(buf.data_ptr(), rank, world)
        self._alive = True

    @staticmethod
    def make_unique_id() -> bytes:
        buf = torch.zeros(128, dtype=torch.uint8)
        get_lib().comm_unique_id(buf.data_ptr())
        return bytes(buf.numpy().tobytes())

    def allreduce_sum(self, t: torch.Tensor, stream: torch.cuda.Stream) -> None:
        get_lib().comm_allreduce_sum(t.data_ptr(), t.numel(), stream.cuda_stream)

    def broadcast(self, t: torch.Tensor, root: int, stream: torch.cuda.Stream) -> None:
        get_lib().comm_broadcast(t.data_ptr(), t.numel(), root, stream.cuda_stream)

    def alloc_registered(self, numel: int, device: torch.device) -> torch.Tensor:
        """A zeroed fp32 buffer from ncclMemAlloc, registered with the communicator: all-reduces on it run in place
        (no staging copies) and may use NVSwitch in-network reduction."""
        import ctypes as C

        out = C.c_void_p()
        get_lib().comm_mem_alloc(4 * numel, C.byref(out))
        get_lib().comm_register(out.value, 4 * numel)
        t = torch.as_tensor(_DevicePtr(out.value, numel), device=device)
        t.zero_()
        self._owned = getattr(self, "_owned", []) + [out.value]
        return t

    def close(self) -> None:
        if self._alive:
            get_lib().comm_destroy()
            for ptr_ in getattr(self, "_owned", []):
                get_lib().comm_mem_free(ptr_)
            self._owned = []
            self._alive = False


class _DevicePtr:
    """Lets torch wrap raw device memory (here: an ncclMemAlloc allocation) as a tensor."""

    def __init__(self, ptr: int, numel: int):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                         "strides": None}


class TrainStep:
    def __init__(self, spec: ModelSpec, init_params: Dict[str, torch.Tensor], dataset: TabularDataset,
                 row_idx: torch.Tensor, cfg: StepConfig, comm: Optional[Communicator] = None,
                 device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise B200Error("no CUDA device: the B200 training step has no CPU fallback")
        if cfg.precision not in ("bf16", "fp32", "tf32"):
            raise ValueError(f"precision must be bf16, fp32 or tf32, got {cfg.precision}")
        if cfg.optimizer not in ("sgd", "adam"):
            raise ValueError(f"optimizer must be sgd or adam, got {cfg.optimizer}")
        if cfg.world > 1 and comm is None:
            raise ValueError("world > 1 needs a Communicator")
        self.lib = get_lib()  # raises when libb200step.so is missing
        self.sms = ops.sm_count()
        self.spec, self.cfg, self.comm = spec, cfg, comm
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.layout = FlatLayout(spec)
        self.dt = torch.bfloat16 if cfg.precision == "bf16" else torch.float32
        B, dev, lay = cfg.batch_size, self.dev, self.layout

        # ---- dataset (resident in HBM) and this rank's sample order (task.py:260-263)
        self.ds = dataset.to(dev)
        self.row_idx = row_idx.to(dev, dtype=torch.int64).contiguous()
        self.num_samples = int(row_idx.numel())
        self.steps_per_epoch_full = self.num_samples // B
        self.tail = self.num_samples - self.steps_per_epoch_full * B
        if self.steps_per_epoch_full == 0 and self.tail == 0:
            raise ValueError("empty sample list")

        # ---- flat parameter / gradient / optimizer-state buffers
        self.p = lay.flatten(init_params).to(dev)
        self.g = None
        if cfg.world > 1 and cfg.nccl_registered_grads:
            try:
                self.g = comm.alloc_registered(lay.total, dev)     # the all-reduce works on the gradients in place
            except B200Error as e:    # an older libnccl without ncclMemAlloc / ncclCommRegister: plain device memory
                print(f"[b200step] NCCL buffer registration unavailable ({e}); using unregistered gradient memory")
        if self.g is None:
            self.g = torch.zeros_like(self.p)
        self.shadow = None
        if self.dt == torch.bfloat16:
            self.shadow = torch.zeros(lay.n_shadow, dtype=torch.bfloat16, device=dev)
        self.s1 = torch.zeros_like(self.p)                    # SGD momentum buffer | Adam exp_avg
        self.s2 = torch.zeros_like(self.p) if cfg.optimizer == "adam" else None
        self.wsrc = self.shadow if self.shadow is not None else self.p   # GEMM weight operands
        self.tables = self.p[lay.tables_offset: lay.tables_offset + max(lay.tables_numel, 1)]
        self.gtables = self.g[lay.tables_offset: lay.total]

        # ---- activations / gradients (row stride = feature count, K of layer 0 padded)
        dims = spec.layer_dims
        self.L = len(dims)
        self.k0 = spec.in_features_padded
        # the preprocessed batch; two buffers so that the next batch can be gathered while wgrad still reads this one
        self.x0s = [torch.zeros(B, self.k0, dtype=self.dt, device=dev) for _ in range(2)]
        self.cur = 0
        self.acts = [torch.zeros(B, o, dtype=self.dt, device=dev) for (_, o) in dims]  # acts[-1] = logits
        self.dlogits = torch.zeros_like(self.acts[-1]) if cfg.keep_logits else self.acts[-1]
        wmax = (max([o for (_, o) in dims[:-1]] + [8]) + 7) // 8 * 8
        # gradient wrt each hidden activation; all stay alive until the wgrad phase
        self.dh = [torch.zeros(B, wmax, dtype=self.dt, device=dev) for _ in range(max(self.L - 1, 1))]
        ed = spec.embed_dim
        pow2 = ed >= 4 and (ed & (ed - 1)) == 0
        self.fused_emb = bool(cfg.fused_emb_bwd)
        if self.fused_emb and cfg.precision == "fp32":
            raise ValueError("fused_emb_bwd runs single-pass TF32 products: use precision tf32 or bf16 with it")
        if self.fused_emb and not (spec.num_cat > 0 and pow2):
            raise ValueError("fused_emb_bwd needs categorical features with a power-of-two embedding width >= 4")
        self.dx0 = None if self.fused_emb else torch.zeros(B, self.k0, dtype=self.dt, device=dev)
        self.y = torch.zeros(B, dtype=torch.int64, device=dev)
        self.idx = torch.zeros(B, max(spec.num_cat, 1), dtype=torch.int32, device=dev)
        self.loss = torch.zeros(1, dtype=torch.float32, device=dev)
        self.cursor = torch.zeros(1, dtype=torch.int32, device=dev)
        self.step_ctr = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ticket = torch.zeros(1, dtype=torch.int32, device=dev)   # last-block detector of the optimizer launch
        # the table optimizer of the overlapped schedule runs beside the dense one: own step counter, own ticket
        self.step_ctr2 = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ticket2 = torch.zeros(1, dtype=torch.int32, device=dev)
        # [0] = rows of the current batch whose label != ignore_index (K-pre counts, K-loss divides by it)
        self.label_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        # bits the kernels set when they meet an id / label torch would have raised on (checked in synchronize())
        self.status = ops.status_word(dev)
        self.fused_ce = (spec.num_classes <= 1024 and not cfg.keep_logits) if cfg.fused_ce is None else cfg.fused_ce
        if self.fused_ce and (spec.num_classes > 1024 or cfg.keep_logits):
            raise ValueError("fused_ce needs <= 1024 classes and keep_logits off (the logits are never stored)")
        self.ce_scratch = torch.empty(max(ops.ce_scratch_floats(spec.num_classes),
                                          ops.linear_ce_scratch_floats(spec.num_classes) if self.fused_ce else 0),
                                      dtype=torch.float32, device=dev)
        # 1 bit per hidden activation (ReLU/dropout mask for backward), row stride padded to 16 bytes
        self.masks = [torch.zeros(B, ((o + 31) // 32 + 3) // 4 * 4, dtype=torch.int32, device=dev)
                      for (_, o) in dims[:-1]]
        self.dg_scratch = torch.empty(max(ops.dgrad_scratch_floats(B, wmax), 4), dtype=torch.float32, device=dev)
        # "fp32" = 3xTF32: every GEMM operand also exists as hi/lo copies laid out along the contraction dimension
        # (ops.split_tf32), so that the TF32 kernels compute a_hi.b_hi + a_hi.b_lo + a_lo.b_hi in one pass over a 3x
        # longer contraction.  k-concatenated copies feed fwd / dgrad, row-stacked copies feed wgrad.
        self.x3 = cfg.precision == "fp32"
        if self.x3:
            if cfg.micro_rows:
                raise ValueError("precision fp32 (3xTF32) does not support micro_rows")
            f32 = dict(dtype=torch.float32, device=dev)
            self.kin = [self.k0 if li == 0 else k for li, (k, _) in enumerate(dims)]
            self.xk = [torch.zeros(B, 3 * kin, **f32) for kin in self.kin]          # layer input  [hi | hi | lo]
            self.xr = [torch.zeros(3 * B * kin, **f32) for kin in self.kin]         # layer input  [hi ; lo ; hi]
            self.dyk = [torch.zeros(B, 3 * o, **f32) for (_, o) in dims]            # output grad  [hi | hi | lo]
            self.dyr = [torch.zeros(3 * B * o, **f32) for (_, o) in dims]           # output grad  [hi ; hi ; lo]
            self.wk = [torch.zeros(o, 3 * kin, **f32) for kin, (_, o) in zip(self.kin, dims)]   # W [hi | lo | hi]
            self.wr = [torch.zeros(3 * o, kin, **f32) for kin, (_, o) in zip(self.kin, dims)]   # W [hi ; lo ; hi]
        self.identity_idx = None  # lazily built for host-fed batches

        # ---- streams / events
        self.main = torch.cuda.Stream(device=dev)
        self.comm_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.buckets = lay.buckets(cfg.first_bucket_bytes, cfg.bucket_cap_bytes)
        self._ev_ready = [torch.cuda.Event() for _ in self.buckets]
        self._ev_done = [torch.cuda.Event() for _ in self.buckets]
        self.graphs = [None, None]     # one captured step per x0 buffer (the prefetch schedule alternates them)
        self.kernels_per_step = 0
        self.host_steps = 0
        # opt-in: measured on T0 (profiles/r02_prefetch_overlap.md) the overlapped K-pre gains 0-2 %: its CTAs and the
        # persistent wgrad grid compete for registers, so little of it actually runs under the GEMMs
        want_prefetch = bool(cfg.prefetch_pre)
        self.prefetch = want_prefetch and len(self._chunks(B)) == 1 and self.steps_per_epoch_full > 0
        # world == 1: the prefetch runs on a side stream UNDER the wgrad GEMMs (see _enqueue)
        self.overlap_pre = self.prefetch and cfg.world == 1
        self.pre_stream = torch.cuda.Stream(device=dev)
        self._ev_fork, self._ev_join = torch.cuda.Event(), torch.cuda.Event()
        self._pre_ready = False   # x0 / y / idx already hold the batch at the current cursor (prefetch mode)

        torch.cuda.synchronize(dev)  # allocations/fills above ran on torch's default stream
        with torch.cuda.stream(self.main):
            if self.shadow is not None:
                ops.cast_bf16(self.p[: lay.n_shadow], self.shadow, stream=self.main)
            if comm is not None and cfg.world > 1:
                # DDP's constructor broadcast of parameters from rank 0 (task.py:189)
                comm.broadcast(self.p, 0, self.main)
                if self.shadow is not None:
                    ops.cast_bf16(self.p[: lay.n_shadow], self.shadow, stream=self.main)
            if self.x3:
                self._split_weights()
        self.main.synchronize()

    # ------------------------------------------------------------------------------------------ helpers
    def _split(self, src: torch.Tensor, kcat, kmask: int, rflat, rmask: int, rows: int, cols: int) -> int:
        """hi/lo operand copies of src[rows, cols] for the 3xTF32 GEMMs (one launch)."""
        r = rflat[: 3 * rows * cols].view(3 * rows, cols) if rflat is not None else None
        ops.split_tf32(src, kcat[:rows] if kcat is not None else None, kmask, r, rmask, rows=rows, cols=cols,
                       stream=self.main)
        return 1

    def _rstack(self, flat: torch.Tensor, rows: int, cols: int) -> torch.Tensor:
        return flat[: 3 * rows * cols].view(3 * rows, cols)

    def _split_weights(self) -> int:
        """Refresh the hi/lo copies of every Linear weight (after the optimizer touched them)."""
        for li, (_, o) in enumerate(self.spec.layer_dims):
            ops.split_tf32(self._w(li), self.wk[li], ops.SPLIT_RIGHT, self.wr[li], ops.SPLIT_RIGHT, rows=o,
                           cols=self.kin[li], stream=self.main)
        return self.L
    def _w(self, li: int) -> torch.Tensor:
        seg = self.layout.weight(li)
        return self.wsrc[seg.offset: seg.offset + seg.numel_stored].view(seg.rows, seg.ld)

    def _gw(self, li: int) -> torch.Tensor:
        seg = self.layout.weight(li)
        return self.g[seg.offset: seg.offset + seg.numel_stored].view(seg.rows, seg.ld)

    def _b(self, li: int, grad: bool = False) -> torch.Tensor:
        seg = self.layout.bias(li)
        return (self.g if grad else self.p)[seg.offset: seg.offset + seg.cols]

    @property
    def x0(self) -> torch.Tensor:
        """The preprocessed batch the step computes on."""
        return self.x0s[self.cur]

    def _chunks(self, batch: int) -> List[tuple]:
        """Row chunks a step is processed in (`StepConfig.micro_rows`; 0 = the whole batch at once, the default).
        Chunking keeps a chunk's activations L2-resident between producer and consumer kernels, but every extra
        kernel launch costs ~6 us of prologue/ramp/tail on this part: measured on T0 (B = 65536) 0.49 ms unchunked
        vs 0.58 / 0.65 / 0.74 ms with 2 / 3 / 4 chunks, so it is only worth it when memory is the constraint."""
        rows = 0 if self.cfg.deterministic_bias_grad else self.cfg.micro_rows  # that reduction overwrites
        if rows <= 0 or rows >= batch:
            return [(0, batch)]
        out, r0 = [], 0
        while r0 < batch:
            out.append((r0, min(rows, batch - r0)))
            r0 += rows
        return out

    def _pre(self, batch: int, ds: TabularDataset, row_idx: torch.Tensor, cursor: Optional[torch.Tensor],
             r0: int = 0, rows: Optional[int] = None, *, into_next: bool = False, stream=None,
             max_ctas: int = 0) -> int:
        """K-pre for rows [r0, r0 + rows) of the batch the cursor points at: x0, y, idx.  `into_next`: fill the OTHER
        x0 buffer (the prefetched next batch)."""
        spec = self.spec
        rows = batch if rows is None else rows
        sl = slice(r0, r0 + rows)
        x0 = self.x0s[self.cur ^ 1] if into_next else self.x0s[self.cur]
        ops.preprocess(ds.dense, ds.cat, ds.label, row_idx[r0:], cursor, ds.mean, ds.std,
                       self.tables if spec.num_cat else None, x0[sl], self.y[sl], self.idx[sl], batch=rows,
                       num_dense=spec.num_dense, num_cat=spec.num_cat, table_rows=spec.cardinality,
                       embed_dim=spec.embed_dim, cursor_stride=batch,
                       label_stats=self.label_stats if rows == batch else None, max_ctas=max_ctas,
                       stream=stream or self.main)
        return 1

    def _forward(self, batch: int, src: Optional[TabularDataset], row_idx: torch.Tensor,
                 cursor: Optional[torch.Tensor], training: bool, r0: int = 0, rows: Optional[int] = None,
                 skip_pre: bool = False, skip_last: bool = False) -> int:
        """K-pre + the Linear stack (task.py:309) for rows [r0, r0 + rows) of the batch; dropout only when training
        (model.train(), task.py:301)."""
        spec, cfg, st = self.spec, self.cfg, self.main
        ds = src or self.ds
        dims = spec.layer_dims
        rows = batch if rows is None else rows
        sl = slice(r0, r0 + rows)
        n = 0
        if not skip_pre:
            n += self._pre(batch, ds, row_idx, cursor, r0, rows)
        inp = self.x0[sl]
        if self.x3:
            n += self._split(self.x0[sl], self.xk[0], ops.SPLIT_LEFT, self.xr[0], ops.SPLIT_RIGHT, rows, self.k0)
            inp = self.xk[0][:rows]
        for li, (k, o) in enumerate(dims):
            kk = self.k0 if li == 0 else k
            last = li == self.L - 1
            if last and skip_last:      # the caller fuses the output layer with the loss (K-gemm-ce)
                break
            ops.linear_fwd(inp, self.wk[li] if self.x3 else self._w(li), self._b(li), self.acts[li][sl], M=rows, N=o,
                           K=3 * kk if self.x3 else kk, relu=not last,
                           dropout_p=spec.dropout if (training and not last) else 0.0, seed=cfg.dropout_seed,
                           step=self.step_ctr, layer_id=li,
                           relu_mask=self.masks[li][sl] if (training and not last) else None, row0=r0,
                           # the loss kernel walks the rows upwards: the layer feeding it runs downwards, the one
                           # before upwards, ...
                           reverse_m=cfg.l2_tile_order and (self.L - 2 - li) % 2 == 0, stream=st)
            n += 1
            inp = self.acts[li][sl]
            if self.x3 and not last:
                n += self._split(inp, self.xk[li + 1], ops.SPLIT_LEFT, self.xr[li + 1], ops.SPLIT_RIGHT, rows, o)
                inp = self.xk[li + 1][:rows]
        return n

    def evaluate(self, return_predictions: bool = False):
        """Top-1 accuracy over the whole resident dataset in its stored order (the reference's evaluate(),
        task.py:30-46, over a shuffle=False loader, task.py:267): forward only, eval mode (no dropout).
        `return_predictions` also hands back each row's predicted class (`predicted` at task.py:40)."""
        n_rows, B = len(self.ds), self.cfg.batch_size
        order = torch.arange(n_rows, dtype=torch.int64, device=self.dev)
        correct = torch.zeros((), dtype=torch.int64, device=self.dev)
        preds = torch.empty(n_rows, dtype=torch.int64, device=self.dev) if return_predictions else None
        torch.cuda.synchronize(self.dev)
        self._pre_ready = False                                  # x0 / y / idx are overwritten below
        with torch.cuda.stream(self.main):
            for s in range(0, n_rows, B):
                b = min(B, n_rows - s)
                self._forward(b, None, order[s:], None, training=False)
                pred = self.acts[-1][:b].argmax(dim=1)          # torch.max(outputs, 1), task.py:40
                correct += (pred == self.y[:b]).sum()           # task.py:42
                if preds is not None:
                    preds[s: s + b] = pred
        self.main.synchronize()
        acc = correct.item() / n_rows
        return (acc, preds.cpu()) if return_predictions else acc

    # --------------------------------------------------------------------------------------- the schedule
    def _enqueue(self, batch: int, src: Optional[TabularDataset], row_idx: torch.Tensor,
                 cursor: Optional[torch.Tensor], advance: bool, prefetch: bool = False) -> int:
        """Enqueue one full step on self.main for `batch` rows; returns the number of kernels launched.
        prefetch: x0 / y / idx already hold this batch; K-pre of the NEXT batch is launched after the first optimizer
        range instead (needs `cursor` and `advance`)."""
        spec, cfg, lay, st = self.spec, self.cfg, self.layout, self.main
        dims = spec.layer_dims
        L = self.L
        n = 0
        inv_keep = 1.0 / (1.0 - spec.dropout) if spec.dropout > 0 else 1.0
        bucket_of: Dict[int, List[int]] = {}
        for bi, (_, _, ready) in enumerate(self.buckets):
            bucket_of.setdefault(ready, []).append(bi)
        chunks = self._chunks(batch)
        reserve = 0
        if cfg.world > 1 and cfg.overlap_comm:
            reserve = DEFAULT_COMM_SM_RESERVE if cfg.comm_sm_reserve is None else cfg.comm_sm_reserve
        comm_live = False
        join_emb = False
        for ci, (r0, rows) in enumerate(chunks):
            sl = slice(r0, r0 + rows)
            last_chunk = ci == len(chunks) - 1
            n += self._forward(batch, src, row_idx, cursor, training=True, r0=r0, rows=rows, skip_pre=prefetch,
                               skip_last=self.fused_ce)
            # loss + dlogits + db_L (summed over the chunks) ---------------------------------------------------
            n_valid = self.label_stats if len(chunks) == 1 else None
            if self.fused_ce:
                # output layer and loss in one kernel: logits live in tensor memory only, dlogits is what is stored
                k_last = dims[L - 1][0] if L > 1 else self.k0
                if self.x3:
                    x_last, w_last, k_last = self.xk[L - 1][:rows], self.wk[L - 1], 3 * k_last
                else:
                    x_last, w_last = (self.x0[sl] if L == 1 else self.acts[L - 2][sl]), self._w(L - 1)
                ops.linear_ce(x_last, w_last, self._b(L - 1),
                              self.y[sl], self.dlogits[sl], self.loss, self._b(L - 1, grad=True), self.ce_scratch,
                              M=rows, N=spec.num_classes, K=k_last, grad_scale=1.0 / (batch * cfg.world),
                              loss_scale=1.0 / batch, accumulate=ci > 0, n_valid=n_valid, stream=st)
            else:
                ops.ce_loss_grad(self.acts[-1][sl], self.y[sl], self.dlogits[sl], self.loss,
                                 self._b(L - 1, grad=True), self.ce_scratch, batch=rows,
                                 num_classes=spec.num_classes, grad_scale=1.0 / (batch * cfg.world),
                                 loss_scale=1.0 / batch, accumulate=ci > 0, n_valid=n_valid, stream=st)
            n += 2
            # backward ------------------------------------------------------------------------------------------
            # Phase 1 walks the dgrad chain down to the embedding gradient, phase 2 runs the wgrad GEMMs.  The table
            # gradients (the largest bucket, and in torch's autograd order the LAST one to become ready) are
            # therefore finished first and their all-reduce hides under the wgrad GEMMs of the last chunk.
            dys = [None] * L
            dys[L - 1] = self.dlogits[sl]
            x3 = self.x3
            if x3:   # hi/lo copies of the output-layer gradient for its dgrad and wgrad
                n += self._split(dys[L - 1], self.dyk[L - 1], ops.SPLIT_LEFT, self.dyr[L - 1], ops.SPLIT_LEFT, rows,
                                 spec.num_classes)
            for li in range(L - 1, 0, -1):
                k, o = dims[li]
                dx = self.dh[li - 1][sl, :k]
                # dY_{l-1} = (dY_l W_l) * mask_{l-1} / keep, and its column sums += bias gradient of layer l-1
                ops.linear_dgrad(self.dyk[li][:rows] if x3 else dys[li], self.wr[li] if x3 else self._w(li), None, dx,
                                 M=rows, N=3 * o if x3 else o, K=k, relu=True, dscale=inv_keep,
                                 relu_mask=self.masks[li - 1][sl], dbias=self._b(li - 1, grad=True),
                                 scratch=self.dg_scratch if cfg.deterministic_bias_grad else None,
                                 reverse_m=cfg.l2_tile_order and (L - 1 - li) % 2 == 0, stream=st)
                n += 2 if cfg.deterministic_bias_grad else 1
                dys[li - 1] = dx
                if x3:
                    n += self._split(dx, self.dyk[li - 1], ops.SPLIT_LEFT, self.dyr[li - 1], ops.SPLIT_LEFT, rows, k)
            if spec.num_cat:
                # (the table gradients were zeroed by the previous optimizer pass / at allocation)
                if self.fused_emb:
                    ops.linear_dgrad_emb(dys[0], self._w(0), self.idx[sl], self.gtables, M=rows, N=dims[0][1],
                                         num_cat=spec.num_cat, table_rows=spec.cardinality,
                                         embed_dim=spec.embed_dim, stream=st)
                    n += 1
                else:
                    ops.linear_dgrad(self.dyk[0][:rows] if x3 else dys[0], self.wr[0] if x3 else self._w(0), None,
                                     self.dx0[sl], M=rows, N=3 * dims[0][1] if x3 else dims[0][1], K=self.k0,
                                     relu=False, reverse_m=cfg.l2_tile_order and (L - 1) % 2 == 0, stream=st)
                    # K-embbwd only feeds the optimizer.  On one GPU (single chunk, no overlapped prefetch) it runs on
                    # the side stream UNDER the grouped wgrad GEMM: its 28-register blocks fit next to the persistent
                    # GEMM CTA (66 registers x 384 threads, all the shared memory) and it is bound by L2 atomics, not
                    # by the tensor pipe.  Joined before K-opt.
                    emb_side = (cfg.emb_bwd_side if cfg.emb_bwd_side is not None else True) and cfg.world == 1 and \
                        len(chunks) == 1 and not (prefetch and self.overlap_pre)
                    es = st
                    if emb_side:
                        self._ev_fork.record(st)
                        self.pre_stream.wait_event(self._ev_fork)
                        es = self.pre_stream
                    ops.emb_bwd(self.dx0[sl], self.idx[sl], self.gtables, batch=rows, emb_col0=0,
                                num_cat=spec.num_cat, table_rows=spec.cardinality, embed_dim=spec.embed_dim,
                                stream=es)
                    if emb_side:
                        self._ev_join.record(self.pre_stream)
                        join_emb = True
                    n += 2
                if last_chunk:
                    for bi in bucket_of.get(-1, []):
                        n += self._reduce_bucket(bi)
                        comm_live = True
            overlap = prefetch and self.overlap_pre
            if overlap:
                # Fork: the table gradients are final, the wgrad GEMMs below neither read nor write the tables.  The
                # side stream updates the tables (its own step counter / ticket: it runs beside the dense optimizer),
                # moves the batch cursor and gathers the NEXT batch into the other x0 buffer, all under the wgrads.
                ps, spe_ = self.pre_stream, max(self.steps_per_epoch_full, 1)
                self._ev_fork.record(st)
                ps.wait_event(self._ev_fork)
                if lay.tables_numel:
                    self._opt_range(lay.tables_offset, lay.total, (cursor, self.step_ctr2, spe_, self.ticket2),
                                    self.step_ctr2, ps)
                else:
                    ops.advance_step(cursor, self.step_ctr2, spe_, stream=ps)
                n += 1 + self._pre(batch, self.ds, row_idx, cursor, into_next=True, stream=ps,
                                   max_ctas=cfg.prefetch_ctas_per_sm * self.sms)
                self._ev_join.record(ps)
            grouped = cfg.grouped_wgrad if cfg.grouped_wgrad is not None else cfg.world == 1
            xs = [self.x0[sl] if li == 0 else self.acts[li - 1][sl] for li in range(L)]
            m_w = rows
            if x3:   # contraction over 3 x rows: [dY_hi ; dY_hi ; dY_lo]^T . [X_hi ; X_lo ; X_hi]
                xs = [self._rstack(self.xr[li], rows, self.kin[li]) for li in range(L)]
                dys = [self._rstack(self.dyr[li], rows, dims[li][1]) for li in range(L)]
                m_w = 3 * rows
            # split-K partials are red.add-ed onto the gradient buffer the optimizer left zeroed: no memset node
            if grouped:
                # one launch for every layer's weight gradient (largest layer first)
                order = list(range(L - 1, -1, -1))
                ops.linear_wgrad_grouped([dys[li] for li in order], [xs[li] for li in order],
                                         [self._gw(li) for li in order], M=m_w,
                                         Ns=[dims[li][1] for li in order],
                                         Ks=[self.k0 if li == 0 else dims[li][0] for li in order], stream=st)
                n += 1
                if last_chunk:
                    for li in order:
                        for bi in bucket_of.get(li, []):
                            n += self._reduce_bucket(bi)
            else:
                for li in range(L - 1, -1, -1):
                    k, o = dims[li]
                    kk = self.k0 if li == 0 else k
                    # while an all-reduce is in flight the persistent GEMM grid leaves `reserve` SMs to NCCL's CTAs
                    if reserve and comm_live:
                        ops.set_sm_budget(self.sms - reserve)
                    ops.linear_wgrad(dys[li], xs[li], self._gw(li), M=m_w, N=o, K=kk, accumulate=True, stream=st)
                    if reserve and comm_live:
                        ops.set_sm_budget(0)
                    n += 1
                    if last_chunk:
                        for bi in bucket_of.get(li, []):
                            n += self._reduce_bucket(bi)
                            comm_live = True
        # optimizer ----------------------------------------------------------------------------------------
        # world > 1: one launch per bucket, each gated only by that bucket's all-reduce, so the update of the early
        # buckets overlaps the all-reduce of the last one; world == 1: a single pass over the whole flat buffer (or,
        # with the overlapped prefetch, over the Linear layers only: the tables were updated on the side stream).
        spe = max(self.steps_per_epoch_full, 1)
        if prefetch and self.overlap_pre:
            self._opt_range(0, lay.tables_offset, (None, self.step_ctr, spe, self.ticket), self.step_ctr, st)
            st.wait_event(self._ev_join)   # join: tables updated, cursor moved, next batch gathered
            return n + 1 + (self._split_weights() if self.x3 else 0)
        if cfg.world > 1 and cfg.overlap_comm:
            ranges = [(self._ev_done[bi], s_, e_) for bi, (s_, e_, _) in enumerate(self.buckets)]
            order = sorted(range(len(ranges)),
                           key=lambda bi: (-1 if self.buckets[bi][2] == -1 else L - self.buckets[bi][2], bi))
            ranges = [ranges[bi] for bi in order]   # in the order their all-reduces were issued
        else:
            ranges = [(None, 0, lay.total)]
        # the last optimizer launch also advances the device-side batch cursor / step count (its last block does)
        # With `prefetch` the batch cursor moves in the FIRST launch (the range that holds the embedding tables, whose
        # all-reduce was issued first), K-pre of the next batch follows it and runs under the all-reduces that are
        # still in flight; the step counter (dropout stream, Adam's t) still moves in the last launch.
        cur = self.cursor if advance else None
        if join_emb:
            st.wait_event(self._ev_join)       # the table gradients from the side stream
        for ri, (ev, s_, e_) in enumerate(ranges):
            first_range, last_range = ri == 0, ri == len(ranges) - 1
            adv_cursor = cur if (first_range if prefetch else last_range) else None
            adv_step = self.step_ctr if last_range else None
            adv = (adv_cursor, adv_step, spe, self.ticket) if (adv_cursor is not None or adv_step is not None) else None
            if ev is not None:
                st.wait_event(ev)
            self._opt_range(s_, e_, adv, self.step_ctr, st)
            n += 1
            if prefetch and first_range:
                # K-pre is a persistent grid too: leave SMs to the all-reduces it is meant to run under
                pre_reserve = cfg.prefetch_sm_reserve if not last_range else 0
                if pre_reserve:
                    ops.set_sm_budget(self.sms - pre_reserve)
                n += self._pre(batch, src or self.ds, row_idx, cursor, into_next=True)
                if pre_reserve:
                    ops.set_sm_budget(0)
        if self.overlap_pre:
            # a step outside the overlapped schedule (tail batch, host-fed batch, profiling): the table optimizer's own
            # step counter has to keep pace
            ops.advance_step(None, self.step_ctr2, spe, stream=st)
            n += 1
        if self.x3:
            n += self._split_weights()     # hi/lo copies of the updated weights for the next step's GEMMs
        return n

    def _opt_range(self, s_: int, e_: int, adv, step_ptr: torch.Tensor, stream) -> None:
        """K-opt over [s_, e_) of the flat buffers (task.py:312; the consumed gradients are left zeroed = the next
        step's optimizer.zero_grad(), task.py:308)."""
        cfg, lay = self.cfg, self.layout
        n_shadow = lay.n_shadow if self.shadow is not None else 0
        ns = max(0, min(n_shadow, e_) - s_)
        sh = self.shadow[s_:] if (self.shadow is not None and ns > 0) else None
        if cfg.optimizer == "sgd":
            ops.sgd_step(self.p[s_:e_], self.g[s_:e_], self.s1[s_:e_], sh, n_shadow=ns, lr=cfg.lr,
                         momentum=cfg.momentum, weight_decay=cfg.weight_decay, zero_grad=True, advance=adv,
                         stream=stream)
        else:
            ops.adam_step(self.p[s_:e_], self.g[s_:e_], self.s1[s_:e_], self.s2[s_:e_], sh, step_ptr, n_shadow=ns,
                          lr=cfg.lr, beta1=cfg.beta1, beta2=cfg.beta2, eps=cfg.eps, weight_decay=cfg.weight_decay,
                          zero_grad=True, advance=adv, stream=stream)

    def _reduce_bucket(self, bi: int) -> int:
        """DDP's bucket all-reduce (task.py:189 hooks): SUM over ranks; 1/world is already in grad_scale."""
        if self.cfg.world <= 1:
            return 0
        s, e, _ = self.buckets[bi]
        if self.cfg.overlap_comm:
            self._ev_ready[bi].record(self.main)
            self.comm_stream.wait_event(self._ev_ready[bi])
            self.comm.allreduce_sum(self.g[s:e], self.comm_stream)
            self._ev_done[bi].record(self.comm_stream)
        else:
            self.comm.allreduce_sum(self.g[s:e], self.main)
        return 1

    # ---------------------------------------------------------------------------------------- public API
    def capture(self) -> None:
        """Capture the steady-state (full-batch) step into a CUDA graph, one per x0 buffer.  One eager step must have
        run before (it configures the kernels and warms NCCL), which `step()` guarantees."""
        if self.graphs[self.cur] is not None:
            return
        import ctypes as C

        self.main.synchronize()
        self.comm_stream.synchronize()
        self.pre_stream.synchronize()
        self.lib.graph_begin(self.main.cuda_stream)
        try:
            self.kernels_per_step = self._enqueue(self.cfg.batch_size, None, self.row_idx, self.cursor, True,
                                                  prefetch=self.prefetch)
        finally:
            out = C.c_void_p()
            self.lib.graph_end(self.main.cuda_stream, C.byref(out))
        self.graphs[self.cur] = out

    def step(self) -> None:
        """One optimizer step on the next full batch of this rank (asynchronous; see `loss_value`)."""
        if self.steps_per_epoch_full == 0:
            raise B200Error("batch_size larger than this rank's sample count: use step_tail()")
        bs = self.cfg.batch_size
        if self.prefetch and not self._pre_ready:
            # first step, or x0 / y / idx were overwritten by a tail batch, an evaluation or a host-fed step
            self._pre(bs, self.ds, self.row_idx, self.cursor)
        if self.cfg.use_graph and self.host_steps > 0:
            if self.graphs[self.cur] is None:
                self.capture()
            self.lib.graph_launch(self.graphs[self.cur], self.main.cuda_stream)
        else:
            self.kernels_per_step = self._enqueue(bs, None, self.row_idx, self.cursor, True, prefetch=self.prefetch)
        if self.prefetch:
            self.cur ^= 1          # the batch gathered during this step sits in the other buffer
        self._pre_ready = self.prefetch
        self.host_steps += 1

    def set_row_order(self, row_idx: torch.Tensor) -> None:
        """`DistributedSampler.set_epoch` for the resident path: replace this rank's sample order (same length) in
        place, so the captured graph keeps reading the same buffer.  The reference never calls set_epoch (SURVEY.md
        App. B-4); `task.py --set_epoch` does, once per epoch, with the cursor back at batch 0."""
        if row_idx.numel() != self.num_samples:
            raise ValueError(f"sample order has {row_idx.numel()} entries, this rank holds {self.num_samples}")
        with torch.cuda.stream(self.main):
            self.row_idx.copy_(row_idx.to(torch.int64))
        self._pre_ready = False

    def step_tail(self) -> None:
        """The last, partial batch of an epoch (DataLoader drop_last=False, task.py:263): eager launch with the
        smaller row count; does not move the full-batch cursor."""
        if self.tail == 0:
            return
        off = self.steps_per_epoch_full * self.cfg.batch_size
        self._pre_ready = False
        self._enqueue(self.tail, None, self.row_idx[off:], None, False)
        self.host_steps += 1

    HOST_STAGES = 3  # staging buffers of the host-fed path: the host can run two steps ahead of the GPU

    HOST_ID_BYTES = 4   # categorical ids cross PCIe as int32 (lossless: cardinality < 2^31), 39 % fewer bytes per batch

    def host_batch_bytes(self, rows: int) -> int:
        spec = self.spec
        return rows * (4 * spec.num_dense + self.HOST_ID_BYTES * spec.num_cat + 8)

    def pack_host_batch(self, dense: torch.Tensor, cat: torch.Tensor, label: torch.Tensor) -> torch.Tensor:
        """Lay one raw batch out as ONE pinned host buffer [label int64 | dense fp32 | cat int32] so that it crosses
        PCIe in a single cudaMemcpyAsync (what a data-loader thread would hand to the training loop).  The ids are
        narrowed from torch's int64 to int32 on the way (lossless)."""
        B = dense.shape[0]
        buf = torch.empty(self.host_batch_bytes(B), dtype=torch.uint8).pin_memory()
        o1 = 8 * B
        o2 = o1 + 4 * self.spec.num_dense * B
        buf[:o1].view(torch.int64).copy_(label)
        buf[o1:o2].view(torch.float32).view(B, -1).copy_(dense)
        if self.spec.num_cat:
            buf[o2:].view(torch.int32).view(B, -1).copy_(cat)
        return buf

    def _stage_views(self, raw: torch.Tensor, B: int) -> TabularDataset:
        o1 = 8 * B
        o2 = o1 + 4 * self.spec.num_dense * B
        o3 = o2 + self.HOST_ID_BYTES * self.spec.num_cat * B
        return TabularDataset(raw[o1:o2].view(torch.float32).view(B, self.spec.num_dense),
                              raw[o2:o3].view(torch.int32).view(B, self.spec.num_cat),
                              raw[:o1].view(torch.int64), self.ds.mean, self.ds.std)

    def _init_host_path(self) -> None:
        bs, dev, n = self.cfg.batch_size, self.dev, self.HOST_STAGES
        self.identity_idx = torch.arange(bs, dtype=torch.int64, device=dev)
        self._stage_raw = [torch.empty(self.host_batch_bytes(bs), dtype=torch.uint8, device=dev) for _ in range(n)]
        self._stages = [self._stage_views(r, bs) for r in self._stage_raw]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._stage_ready = [torch.cuda.Event() for _ in range(n)]
        self._stage_free = [torch.cuda.Event() for _ in range(n)]
        self._loss_ev = [torch.cuda.Event() for _ in range(n)]
        self._loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(n)]
        self._host_graphs = [None] * n
        self._host_k = 0
        torch.cuda.synchronize(dev)

    def step_from_host(self, packed: torch.Tensor, rows: Optional[int] = None) -> int:
        """One step on a batch that lives in pinned HOST memory (`pack_host_batch`): the reference's per-step H2D
        copies (task.py:305,307) followed by the same device step, then an asynchronous device->host copy of the
        step's loss.  Three staging buffers rotate, so the copies of the next batches overlap the compute of this one
        (the reference gets the same overlap from its DataLoader workers + non_blocking copies).
        Returns a ticket for `host_loss(ticket)`; a ticket is valid until HOST_STAGES further steps were issued."""
        bs = self.cfg.batch_size
        B = bs if rows is None else rows
        if B > bs or packed.numel() != self.host_batch_bytes(B):
            raise ValueError("host batch does not match the configured batch size / layout")
        if self.identity_idx is None:
            self._init_host_path()
        k = self._host_k
        self._host_k = (k + 1) % self.HOST_STAGES
        self._pre_ready = False
        self.copy_stream.wait_event(self._stage_free[k])      # the step that last read this stage is done
        with torch.cuda.stream(self.copy_stream):
            self._stage_raw[k][: packed.numel()].copy_(packed, non_blocking=True)
        self._stage_ready[k].record(self.copy_stream)
        self.main.wait_event(self._stage_ready[k])
        if B == bs and self.cfg.use_graph and self.host_steps > 0:
            if self._host_graphs[k] is None:
                import ctypes as C

                self.main.synchronize()
                self.comm_stream.synchronize()
                self.lib.graph_begin(self.main.cuda_stream)
                try:
                    self._enqueue(B, self._stages[k], self.identity_idx, None, False)
                finally:
                    out = C.c_void_p()
                    self.lib.graph_end(self.main.cuda_stream, C.byref(out))
                self._host_graphs[k] = out
            self.lib.graph_launch(self._host_graphs[k], self.main.cuda_stream)
        else:
            stage = self._stages[k] if B == bs else self._stage_views(self._stage_raw[k], B)
            self._enqueue(B, stage, self.identity_idx, None, False)
        self._stage_free[k].record(self.main)
        with torch.cuda.stream(self.main):
            self._loss_host[k].copy_(self.loss, non_blocking=True)
        self._loss_ev[k].record(self.main)
        self.host_steps += 1
        return k

    def host_loss(self, ticket: int) -> float:
        """Loss of the host-fed step that returned `ticket` (blocks until that step's D2H copy has landed)."""
        self._loss_ev[ticket].synchronize()
        return float(self._loss_host[ticket].item())

    def loss_value(self) -> float:
        """Blocking read of the last step's local mean loss (what `loss` holds at task.py:310)."""
        self.main.synchronize()
        return float(self.loss.item())

    def synchronize(self) -> None:
        self.main.synchronize()
        self.comm_stream.synchronize()
        self.pre_stream.synchronize()
        bits = int(self.status.item())
        if bits:
            self.status.zero_()
            what = []
            if bits & ops.STATUS_BAD_CATEGORICAL_ID:
                what.append(f"a categorical id outside [0, {self.spec.cardinality}) (nn.Embedding raises IndexError)")
            if bits & ops.STATUS_BAD_LABEL:
                what.append(f"a label outside [0, {self.spec.num_classes}) that is not ignore_index "
                            "(nn.CrossEntropyLoss raises)")
            raise B200Error("the step met " + " and ".join(what) + "; the offending rows were neutralised on the "
                            "device (id -> row 0, label -> ignored) but the input is invalid")

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Same keys/shapes as the reference's `model.state_dict()` under DDP (task.py:283/287)."""
        self.synchronize()
        return self.layout.unflatten(self.p)


    def close(self) -> None:
        self.synchronize()
        for i, g in enumerate(self.graphs):
            if g is not None:
                self.lib.graph_destroy(g)
                self.graphs[i] = None
        for g in getattr(self, "_host_graphs", []):
            if g is not None:
                self.lib.graph_destroy(g)
        self._host_graphs = [None] * self.HOST_STAGES
