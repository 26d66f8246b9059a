"""Tensor-level wrappers over the C ABI: torch tensors in, raw device pointers out.

PyTorch is only the owner of device memory and streams here; every function below enqueues one (or two)
hand-written sm_100a kernels from libb200step.so on the given stream (default: torch's current stream).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import lib as _l
from .lib import BF16, F32, FLAG_ACCUMULATE, FLAG_DETERMINISTIC, FLAG_RELU, FLAG_REVERSE_M, get_lib, ptr

__all__ = [
    "dtype_code", "preprocess", "linear_fwd", "linear_dgrad", "linear_wgrad", "ce_loss_grad", "colsum",
    "emb_bwd", "sgd_step", "adam_step", "cast_bf16", "advance_step", "F32", "BF16", "FLAG_RELU",
    "FLAG_ACCUMULATE",
]


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {dt}: the B200 step computes in float32 (TF32) or bfloat16")


def _stream(stream) -> int:
    if stream is None:
        return torch.cuda.current_stream().cuda_stream
    if isinstance(stream, torch.cuda.Stream):
        return stream.cuda_stream
    return int(stream)


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _l.B200Error("the B200 step has no CPU path: all tensors must be CUDA tensors")


def preprocess(ds_dense, ds_cat, ds_label, row_idx, cursor, mean, std, tables, x_out, y_out, idx_out, *,
               batch: int, num_dense: int, num_cat: int, table_rows: int, embed_dim: int, cursor_stride: int = 0,
               label_stats=None, ignore_index: int = -100, max_ctas: int = 0, stream=None):
    _cuda(ds_dense, ds_cat, ds_label, row_idx, mean, std, x_out, y_out, idx_out, label_stats)
    get_lib().preprocess(ptr(ds_dense), ptr(ds_cat), ds_cat.element_size() if ds_cat is not None else 8,
                         ptr(ds_label), ptr(row_idx), ptr(cursor), ptr(mean),
                         ptr(std), ptr(tables), ptr(x_out), ptr(y_out), ptr(idx_out), batch, num_dense,
                         num_cat, table_rows, embed_dim, x_out.stride(0), dtype_code(x_out.dtype),
                         cursor_stride, ptr(label_stats), int(ignore_index), int(max_ctas), _stream(stream))


def linear_fwd(x, w, bias, y, *, M: int, N: int, K: int, relu: bool, dropout_p: float = 0.0, seed: int = 0,
               step=None, layer_id: int = 0, row0: int = 0, relu_mask=None, reverse_m: bool = False, stream=None):
    _cuda(x, w, bias, y, relu_mask)
    get_lib().linear_fwd(ptr(x), x.stride(0), ptr(w), w.stride(0), ptr(bias), ptr(y), y.stride(0), M, N, K,
                         dtype_code(x.dtype), (FLAG_RELU if relu else 0) | (FLAG_REVERSE_M if reverse_m else 0),
                         float(dropout_p), int(seed),
                         ptr(step), layer_id, int(row0), ptr(relu_mask),
                         relu_mask.stride(0) if relu_mask is not None else 0, _stream(stream))


def dgrad_scratch_floats(M: int, K: int) -> int:
    return int(get_lib().dgrad_scratch_floats(M, K))


def linear_dgrad(dy, w, act, dx, *, M: int, N: int, K: int, relu: bool, dscale: float = 1.0, relu_mask=None,
                 dbias=None, scratch=None, reverse_m: bool = False, stream=None):
    _cuda(dy, w, act, dx, relu_mask, dbias, scratch)
    get_lib().linear_dgrad(ptr(dy), dy.stride(0), ptr(w), w.stride(0), ptr(act),
                           act.stride(0) if act is not None else 0, ptr(dx), dx.stride(0), M, N, K,
                           dtype_code(dy.dtype), (FLAG_RELU if relu else 0) | (FLAG_REVERSE_M if reverse_m else 0),
                           float(dscale), ptr(relu_mask),
                           relu_mask.stride(0) if relu_mask is not None else 0, ptr(dbias), ptr(scratch),
                           scratch.numel() if scratch is not None else 0, _stream(stream))


def linear_dgrad_emb(dy, w, idx, dtables, *, M: int, N: int, num_cat: int, table_rows: int, embed_dim: int,
                     stream=None):
    """Layer-0 dgrad whose output goes straight onto the embedding-table gradients (no dx0, no emb_bwd launch)."""
    _cuda(dy, w, idx, dtables)
    get_lib().linear_dgrad_emb(ptr(dy), dy.stride(0), ptr(w), w.stride(0), ptr(idx), idx.stride(0), ptr(dtables), M,
                               N, num_cat, table_rows, embed_dim, dtype_code(dy.dtype), _stream(stream))


def linear_wgrad(dy, x, dw, *, M: int, N: int, K: int, accumulate: bool = False, deterministic: bool = False,
                 stream=None):
    _cuda(dy, x, dw)
    get_lib().linear_wgrad(ptr(dy), dy.stride(0), ptr(x), x.stride(0), ptr(dw), dw.stride(0), M, N, K,
                           dtype_code(dy.dtype),
                           (FLAG_ACCUMULATE if accumulate else 0) | (FLAG_DETERMINISTIC if deterministic else 0),
                           _stream(stream))


def linear_wgrad_grouped(dys, xs, dws, *, M: int, Ns, Ks, stream=None):
    """dws[j] += dys[j]^T @ xs[j] for all j in one launch (outputs must be zeroed / hold the value to add onto)."""
    import ctypes as C

    n = len(dys)
    _cuda(*dys, *xs, *dws)
    vp = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])
    i64 = lambda vs: (C.c_int64 * n)(*[int(v) for v in vs])
    get_lib().linear_wgrad_grouped(n, vp(dys), i64([t.stride(0) for t in dys]), vp(xs),
                                   i64([t.stride(0) for t in xs]), vp(dws), i64([t.stride(0) for t in dws]), M,
                                   i64(Ns), i64(Ks), dtype_code(dys[0].dtype), _stream(stream))


def ce_scratch_floats(num_classes: int) -> int:
    return int(get_lib().ce_scratch_floats(num_classes))


def ce_loss_grad(logits, labels, dlogits, loss_out, dbias, scratch, *, batch: int, num_classes: int,
                 grad_scale: float, loss_scale: float = 0.0, accumulate: bool = False, ignore_index: int = -100,
                 n_valid=None, stream=None):
    _cuda(logits, labels, dlogits, loss_out, dbias, scratch, n_valid)
    get_lib().ce_loss_grad(ptr(logits), logits.stride(0), ptr(labels), ptr(dlogits), dlogits.stride(0),
                           ptr(loss_out), ptr(dbias), ptr(scratch), scratch.numel(), batch, num_classes,
                           float(grad_scale), float(loss_scale), int(accumulate), int(ignore_index), ptr(n_valid),
                           dtype_code(logits.dtype), _stream(stream))


def linear_ce_scratch_floats(num_classes: int) -> int:
    return int(get_lib().linear_ce_scratch_floats(num_classes))


def linear_ce(x, w, bias, labels, dlogits, loss_out, dbias, scratch, *, M: int, N: int, K: int, grad_scale: float,
              loss_scale: float = 0.0, accumulate: bool = False, ignore_index: int = -100, n_valid=None, stream=None):
    """Output layer + cross-entropy in one kernel: the logits stay in tensor memory, only dlogits is written."""
    _cuda(x, w, bias, labels, dlogits, loss_out, dbias, scratch, n_valid)
    get_lib().linear_ce(ptr(x), x.stride(0), ptr(w), w.stride(0), ptr(bias), ptr(labels), ptr(dlogits),
                        dlogits.stride(0), ptr(loss_out), ptr(dbias), ptr(scratch), scratch.numel(), M, N, K,
                        dtype_code(x.dtype), float(grad_scale), float(loss_scale), int(accumulate), int(ignore_index),
                        ptr(n_valid), _stream(stream))


def colsum_scratch_floats(n: int) -> int:
    return int(get_lib().colsum_scratch_floats(n))


def colsum(dy, out, scratch, *, M: int, N: int, stream=None):
    _cuda(dy, out, scratch)
    get_lib().colsum(ptr(dy), dy.stride(0), ptr(out), ptr(scratch), scratch.numel(), M, N,
                     dtype_code(dy.dtype), _stream(stream))


def emb_bwd(dx, idx, dtables, *, batch: int, emb_col0: int = 0, num_cat: int, table_rows: int, embed_dim: int,
            stream=None):
    _cuda(dx, idx, dtables)
    get_lib().emb_bwd(ptr(dx), dx.stride(0), ptr(idx), ptr(dtables), batch, emb_col0, num_cat, table_rows,
                      embed_dim, dtype_code(dx.dtype), _stream(stream))


def sgd_step(p, g, buf, shadow, *, n_shadow: int, lr: float, momentum: float, weight_decay: float,
             zero_grad: bool = False, advance=None, stream=None):
    """advance = (cursor, step, steps_per_epoch, ticket) folds the step bookkeeping into this launch."""
    _cuda(p, g, buf, shadow)
    ac, as_, spe, tk = advance if advance is not None else (None, None, 1, None)
    get_lib().sgd_step(ptr(p), ptr(g), ptr(buf), ptr(shadow), p.numel(), n_shadow, lr, momentum, weight_decay,
                       int(zero_grad), ptr(ac), ptr(as_), spe, ptr(tk), _stream(stream))


def adam_step(p, g, m, v, shadow, step, *, n_shadow: int, lr: float, beta1: float, beta2: float, eps: float,
              weight_decay: float, zero_grad: bool = False, advance=None, stream=None):
    _cuda(p, g, m, v, shadow, step)
    ac, as_, spe, tk = advance if advance is not None else (None, None, 1, None)
    get_lib().adam_step(ptr(p), ptr(g), ptr(m), ptr(v), ptr(shadow), p.numel(), n_shadow, lr, beta1, beta2,
                        eps, weight_decay, ptr(step), int(zero_grad), ptr(ac), ptr(as_), spe, ptr(tk),
                        _stream(stream))


SPLIT_LEFT = 0b100    # (hi, hi, lo): the left operand of a 3xTF32 product
SPLIT_RIGHT = 0b010   # (hi, lo, hi): the right operand


def split_tf32(src, kcat, kmask: int, rstack, rmask: int, *, rows: int, cols: int, stream=None):
    """hi/lo copies of an fp32 matrix laid out along the contraction dimension (include/b200step.h)."""
    _cuda(src, kcat, rstack)
    get_lib().split_tf32(ptr(src), src.stride(0), rows, cols, ptr(kcat), int(kmask), ptr(rstack), int(rmask),
                         _stream(stream))


def cast_bf16(src, dst, stream=None):
    _cuda(src, dst)
    get_lib().cast_bf16(ptr(src), ptr(dst), src.numel(), _stream(stream))


def advance_step(cursor, step, steps_per_epoch: int, stream=None):
    _cuda(cursor, step)
    get_lib().advance_step(ptr(cursor), ptr(step), steps_per_epoch, _stream(stream))


def sm_count() -> int:
    import ctypes as C

    n = C.c_int(0)
    get_lib().device_sm_count(C.byref(n))
    return int(n.value)


STATUS_BAD_CATEGORICAL_ID = 1
STATUS_BAD_LABEL = 2


_status_words = {}


def status_word(device=None) -> torch.Tensor:
    """The process-wide status word of `device` (one int32, allocated once and registered with the library): kernels
    OR the B200_STATUS_* bits onto it (include/b200step.h).  One device per process is the rule here (task.py:124
    spawns one process per GPU); a second device gets its own word and re-registers."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    w = _status_words.get(dev)
    if w is None:
        w = torch.zeros(1, dtype=torch.int32, device=dev)
        _status_words[dev] = w
    get_lib().set_status_word(ptr(w))
    return w


def set_sm_budget(sms: int) -> None:
    """Size the persistent grids of the launches that follow for `sms` SMs (0 = all of them)."""
    get_lib().set_sm_budget(int(sms))


def zero(t, stream=None):
    _cuda(t)
    get_lib().zero(ptr(t), t.numel() * t.element_size(), _stream(stream))


# ---- convolutional path (include/b200step.h: "convolutional path") ---------------------------------------------
STATUS_BAD_ROW_INDEX = 4


def conv_out_size(n: int, k: int, stride: int, pad: int) -> int:
    return (n + 2 * pad - k) // stride + 1


def image_augment(images, labels, row_idx, cursor, draws, mean3, std3, x_out, y_out, *, batch: int, H: int, W: int,
                  pad: int = 4, cursor_stride: int = 0, stream=None):
    """mean3 / std3: CPU float32 tensors of 3 elements (read during the call)."""
    _cuda(images, labels, row_idx, draws, x_out, y_out)
    get_lib().image_augment(ptr(images), ptr(labels), ptr(row_idx), ptr(cursor), int(cursor_stride), ptr(draws),
                            images.shape[0], ptr(mean3), ptr(std3), int(pad), ptr(x_out), ptr(y_out), batch, H, W,
                            dtype_code(x_out.dtype), _stream(stream))


def im2col(x, col, *, B: int, H: int, W: int, C: int, kh: int, kw: int, stride: int, pad: int, stream=None):
    _cuda(x, col)
    get_lib().im2col(ptr(x), ptr(col), col.stride(0), B, H, W, C, kh, kw, stride, pad, dtype_code(x.dtype),
                     _stream(stream))


def conv_fwd(x, w, y, *, B: int, H: int, W: int, C: int, kh: int, kw: int, stride: int, pad: int, Cout: int,
             relu: bool = False, deterministic: bool = False, ws=None, stream=None):
    """y = conv2d(x NHWC, w[Cout, kh*kw*C]) as an implicit GEMM (im2col-mode TMA, no col matrix).  ws: optional fp32
    workspace (>= y.numel() floats) that allows the contraction of a small-M convolution to be split over idle SMs."""
    _cuda(x, w, y, ws)
    get_lib().conv_fwd(ptr(x), ptr(w), w.stride(0), ptr(y), y.stride(0), B, H, W, C, kh, kw, stride, pad, Cout,
                       dtype_code(x.dtype), (FLAG_RELU if relu else 0) | (FLAG_DETERMINISTIC if deterministic else 0),
                       ptr(ws), ws.numel() if ws is not None else 0, _stream(stream))


def conv_dgrad(dy, w, dx, *, B: int, H: int, W: int, C: int, kh: int, kw: int, pad: int, Cout: int,
               deterministic: bool = False, ws=None, stream=None):
    """dx[B*H*W, C] of a stride-1 convolution as an implicit GEMM over dy (no dcol matrix, no col2im)."""
    _cuda(dy, w, dx, ws)
    get_lib().conv_dgrad(ptr(dy), ptr(w), w.stride(0), ptr(dx), dx.stride(0), B, H, W, C, kh, kw, pad, Cout,
                         dtype_code(dx.dtype), FLAG_DETERMINISTIC if deterministic else 0, ptr(ws),
                         ws.numel() if ws is not None else 0, _stream(stream))


def conv_wgrad(dy, x, dw, *, B: int, H: int, W: int, C: int, kh: int, kw: int, stride: int, pad: int, Cout: int,
               accumulate: bool = False, deterministic: bool = False, stream=None):
    """dw[Cout, kh*kw*C] = dy^T . unfold(x NHWC) as an implicit GEMM (no col matrix)."""
    _cuda(dy, x, dw)
    get_lib().conv_wgrad(ptr(dy), dy.stride(0), ptr(x), ptr(dw), dw.stride(0), B, H, W, C, kh, kw, stride, pad, Cout,
                         dtype_code(x.dtype),
                         (FLAG_ACCUMULATE if accumulate else 0) | (FLAG_DETERMINISTIC if deterministic else 0),
                         _stream(stream))


def col2im(dcol, add, dx, *, B: int, H: int, W: int, C: int, kh: int, kw: int, stride: int, pad: int, stream=None):
    _cuda(dcol, add, dx)
    get_lib().col2im(ptr(dcol), dcol.stride(0), ptr(add), ptr(dx), B, H, W, C, kh, kw, stride, pad,
                     dtype_code(dx.dtype), _stream(stream))


def bn_scratch_floats(C: int) -> int:
    return int(get_lib().bn_scratch_floats(C))


def bn_fwd_stats(y, mean, invstd, running_mean, running_var, num_batches, scratch, *, M: int, C: int,
                 eps: float = 1e-5, momentum: float = 0.1, stream=None):
    _cuda(y, mean, invstd, running_mean, running_var, num_batches, scratch)
    get_lib().bn_fwd_stats(ptr(y), M, C, float(eps), float(momentum), ptr(mean), ptr(invstd), ptr(running_mean),
                           ptr(running_var), ptr(num_batches), ptr(scratch), scratch.numel(), dtype_code(y.dtype),
                           _stream(stream))


def bn_fwd_apply(y, res, out, mean, invstd, gamma, beta, *, M: int, C: int, relu: bool, stream=None):
    _cuda(y, res, out, mean, invstd, gamma, beta)
    get_lib().bn_fwd_apply(ptr(y), ptr(res), ptr(out), M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta),
                           int(relu), dtype_code(y.dtype), _stream(stream))


def bn_bwd_reduce(dout, out, y, mean, invstd, dgamma, dbeta, scratch, *, M: int, C: int, dout2=None, stream=None):
    _cuda(dout, dout2, out, y, mean, invstd, dgamma, dbeta, scratch)
    get_lib().bn_bwd_reduce(ptr(dout), ptr(dout2), ptr(out), ptr(y), M, C, ptr(mean), ptr(invstd), ptr(dgamma), ptr(dbeta),
                            ptr(scratch), scratch.numel(), dtype_code(y.dtype), _stream(stream))


def bn_bwd_apply(dout, out, y, dy, dz_out, mean, invstd, gamma, dgamma, dbeta, *, M: int, C: int, dout2=None,
                 stream=None):
    _cuda(dout, dout2, out, y, dy, dz_out, mean, invstd, gamma, dgamma, dbeta)
    get_lib().bn_bwd_apply(ptr(dout), ptr(dout2), ptr(out), ptr(y), ptr(dy), ptr(dz_out), M, C, ptr(mean), ptr(invstd),
                           ptr(gamma), ptr(dgamma), ptr(dbeta), dtype_code(y.dtype), _stream(stream))


def bn_fwd(y, res, out, mean, invstd, gamma, beta, running_mean, running_var, num_batches, scratch, *, M: int, C: int,
           relu: bool, eps: float = 1e-5, momentum: float = 0.1, stream=None):
    """bn_fwd_stats + bn_fwd_apply as one launch (include/b200step.h: b200_bn_fwd)."""
    _cuda(y, res, out, mean, invstd, gamma, beta, running_mean, running_var, num_batches, scratch)
    get_lib().bn_fwd(ptr(y), ptr(res), ptr(out), M, C, float(eps), float(momentum), ptr(mean), ptr(invstd), ptr(gamma),
                     ptr(beta), int(relu), ptr(running_mean), ptr(running_var), ptr(num_batches), ptr(scratch),
                     scratch.numel(), dtype_code(y.dtype), _stream(stream))


def bn_bwd(dout, out, y, dy, dz_out, mean, invstd, gamma, dgamma, dbeta, scratch, *, M: int, C: int, dout2=None,
           stream=None):
    """bn_bwd_reduce + bn_bwd_apply as one launch (include/b200step.h: b200_bn_bwd)."""
    _cuda(dout, dout2, out, y, dy, dz_out, mean, invstd, gamma, dgamma, dbeta, scratch)
    get_lib().bn_bwd(ptr(dout), ptr(dout2), ptr(out), ptr(y), ptr(dy), ptr(dz_out), M, C, ptr(mean), ptr(invstd),
                     ptr(gamma), ptr(dgamma), ptr(dbeta), ptr(scratch), scratch.numel(), dtype_code(y.dtype),
                     _stream(stream))


def maxpool_fwd(x, out, taps, *, B: int, H: int, W: int, C: int, k: int = 3, stride: int = 2, pad: int = 1,
                stream=None):
    _cuda(x, out, taps)
    get_lib().maxpool_fwd(ptr(x), ptr(out), ptr(taps), B, H, W, C, k, stride, pad, dtype_code(x.dtype),
                          _stream(stream))


def maxpool_bwd(dout, taps, dx, *, B: int, H: int, W: int, C: int, k: int = 3, stride: int = 2, pad: int = 1,
                dout2=None, stream=None):
    _cuda(dout, dout2, taps, dx)
    get_lib().maxpool_bwd(ptr(dout), ptr(dout2), ptr(taps), ptr(dx), B, H, W, C, k, stride, pad, dtype_code(dx.dtype),
                          _stream(stream))


def avgpool_fwd(x, out, *, B: int, HW: int, C: int, stream=None):
    _cuda(x, out)
    get_lib().avgpool_fwd(ptr(x), ptr(out), B, HW, C, dtype_code(x.dtype), _stream(stream))


def avgpool_bwd(dout, dx, *, B: int, HW: int, C: int, stream=None):
    _cuda(dout, dx)
    get_lib().avgpool_bwd(ptr(dout), ptr(dx), B, HW, C, dtype_code(dx.dtype), _stream(stream))
