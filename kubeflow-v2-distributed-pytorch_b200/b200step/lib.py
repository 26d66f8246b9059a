"""ctypes binding of libb200step.so (the C ABI declared in include/b200step.h).

There is no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

F32 = 0
BF16 = 1
FLAG_RELU = 1
FLAG_ACCUMULATE = 2
FLAG_REVERSE_M = 4
FLAG_DETERMINISTIC = 8

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_LIB = os.path.join(_PKG_DIR, "lib", "libb200step.so")

_p = C.c_void_p
_i = C.c_int
_i64 = C.c_int64
_u64 = C.c_uint64
_f = C.c_float
_d = C.c_double

# name -> (restype, argtypes); mirrors include/b200step.h one to one
SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_abi_version": (_i, []),
    "b200_device_sm_count": (_i, [C.POINTER(_i)]),
    "b200_set_sm_budget": (_i, [_i]),
    "b200_set_status_word": (_i, [_p]),
    "b200_preprocess": (_i, [_p, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _i, _i, _i64, _i, _i64, _i, _i64, _p,
                             _i64, _i, _p]),
    "b200_linear_fwd": (_i, [_p, _i64, _p, _i64, _p, _p, _i64, _i64, _i64, _i64, _i, _i, _f, _u64, _p, _i, _i64, _p, _i64,
                             _p]),
    "b200_linear_dgrad": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i64, _i64, _i64, _i, _i, _f, _p, _i64, _p,
                               _p, _i64, _p]),
    "b200_linear_dgrad_emb": (_i, [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i64, _i, _i64, _i, _i, _p]),
    "b200_linear_wgrad_grouped": (_i, [_i, _p, _p, _p, _p, _p, _p, _i64, _p, _p, _i, _p]),
    "b200_dgrad_scratch_floats": (_i64, [_i64, _i64]),
    "b200_linear_wgrad": (_i, [_p, _i64, _p, _i64, _p, _i64, _i64, _i64, _i64, _i, _i, _p]),
    "b200_ce_scratch_floats": (_i64, [_i]),
    "b200_ce_loss_grad": (_i, [_p, _i64, _p, _p, _i64, _p, _p, _p, _i64, _i64, _i, _f, _f, _i, _i64, _p, _i, _p]),
    "b200_linear_ce_scratch_floats": (_i64, [_i64]),
    "b200_linear_ce": (_i, [_p, _i64, _p, _i64, _p, _p, _p, _i64, _p, _p, _p, _i64, _i64, _i64, _i64, _i, _f, _f, _i,
                            _i64, _p, _p]),
    "b200_colsum_scratch_floats": (_i64, [_i64]),
    "b200_colsum": (_i, [_p, _i64, _p, _p, _i64, _i64, _i64, _i, _p]),
    "b200_colsum_static": (_i, [_p, _i64, _p, _i64, _i64, _i, _p]),
    "b200_emb_bwd": (_i, [_p, _i64, _p, _p, _i64, _i, _i, _i64, _i, _i, _p]),
    "b200_sgd_step": (_i, [_p, _p, _p, _p, _i64, _i64, _d, _d, _d, _i, _p, _p, C.c_int32, _p, _p]),
    "b200_adam_step": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _d, _d, _d, _d, _d, _p, _i, _p, _p, C.c_int32, _p, _p]),
    "b200_split_tf32": (_i, [_p, _i64, _i64, _i64, _p, _i, _p, _i, _p]),
    "b200_cast_bf16": (_i, [_p, _p, _i64, _p]),
    "b200_image_augment": (_i, [_p, _p, _p, _p, _i64, _p, _i64, _p, _p, _i, _p, _p, _i64, _i, _i, _i, _p]),
    "b200_conv_fwd": (_i, [_p, _p, _i64, _p, _i64, _i64, _i, _i, _i, _i, _i, _i, _i, _i64, _i, _i, _p, _i64, _p]),
    "b200_conv_dgrad": (_i, [_p, _p, _i64, _p, _i64, _i64, _i, _i, _i, _i, _i, _i, _i64, _i, _i, _p, _i64, _p]),
    "b200_conv_wgrad": (_i, [_p, _i64, _p, _p, _i64, _i64, _i, _i, _i, _i, _i, _i, _i, _i64, _i, _i, _p]),
    "b200_im2col": (_i, [_p, _p, _i64, _i64, _i, _i, _i, _i, _i, _i, _i, _i, _p]),
    "b200_col2im": (_i, [_p, _i64, _p, _p, _i64, _i, _i, _i, _i, _i, _i, _i, _i, _p]),
    "b200_bn_scratch_floats": (_i64, [_i64]),
    "b200_bn_fwd_stats": (_i, [_p, _i64, _i64, _f, _f, _p, _p, _p, _p, _p, _p, _i64, _i, _p]),
    "b200_bn_fwd_apply": (_i, [_p, _p, _p, _i64, _i64, _p, _p, _p, _p, _i, _i, _p]),
    "b200_bn_bwd_reduce": (_i, [_p, _p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p, _i64, _i, _p]),
    "b200_bn_bwd_apply": (_i, [_p, _p, _p, _p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p, _i, _p]),
    "b200_bn_fwd": (_i, [_p, _p, _p, _i64, _i64, _f, _f, _p, _p, _p, _p, _i, _p, _p, _p, _p, _i64, _i, _p]),
    "b200_bn_bwd": (_i, [_p, _p, _p, _p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p, _p, _i64, _i, _p]),
    "b200_maxpool_fwd": (_i, [_p, _p, _p, _i64, _i, _i, _i, _i, _i, _i, _i, _p]),
    "b200_maxpool_bwd": (_i, [_p, _p, _p, _p, _i64, _i, _i, _i, _i, _i, _i, _i, _p]),
    "b200_avgpool_fwd": (_i, [_p, _p, _i64, _i, _i, _i, _p]),
    "b200_avgpool_bwd": (_i, [_p, _p, _i64, _i, _i, _i, _p]),
    "b200_zero": (_i, [_p, _i64, _p]),
    "b200_advance_step": (_i, [_p, _p, C.c_int32, _p]),
    "b200_comm_unique_id": (_i, [_p]),
    "b200_comm_init": (_i, [_p, _i, _i]),
    "b200_comm_allreduce_sum": (_i, [_p, _i64, _p]),
    "b200_comm_broadcast": (_i, [_p, _i64, _i, _p]),
    "b200_comm_destroy": (_i, []),
    "b200_comm_mem_alloc": (_i, [_i64, C.POINTER(_p)]),
    "b200_comm_mem_free": (_i, [_p]),
    "b200_comm_register": (_i, [_p, _i64]),
    "b200_stream_create": (_i, [C.POINTER(_p), _i]),
    "b200_stream_destroy": (_i, [_p]),
    "b200_stream_sync": (_i, [_p]),
    "b200_event_create": (_i, [C.POINTER(_p), _i]),
    "b200_event_destroy": (_i, [_p]),
    "b200_event_record": (_i, [_p, _p]),
    "b200_stream_wait_event": (_i, [_p, _p]),
    "b200_event_elapsed_ms": (_i, [_p, _p, C.POINTER(_f)]),
    "b200_graph_begin": (_i, [_p]),
    "b200_graph_end": (_i, [_p, C.POINTER(_p)]),
    "b200_graph_launch": (_i, [_p, _p]),
    "b200_graph_destroy": (_i, [_p]),
}

# functions whose int return value is NOT an error code
_NO_CHECK = {"b200_last_error", "b200_abi_version", "b200_ce_scratch_floats", "b200_colsum_scratch_floats",
             "b200_linear_ce_scratch_floats",
             "b200_dgrad_scratch_floats", "b200_bn_scratch_floats"}


class B200Error(RuntimeError):
    pass


class _Lib:
    def __init__(self, path: str):
        if not os.path.exists(path):
            raise B200Error(
                f"{path} not found: build it with `python kubeflow-v2-distributed-pytorch_b200/build.py` "
                "(there is no CPU or PyTorch fallback for the training step)"
            )
        self.path = path
        self.cdll = C.CDLL(path, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(self.cdll, name)  # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
            setattr(self, "_raw_" + name, fn)

    def last_error(self) -> str:
        return self._raw_b200_last_error().decode("utf-8", "replace")

    def __getattr__(self, name: str):
        # lib.preprocess(...) -> b200_preprocess(...) with error checking
        full = "b200_" + name
        if full not in SIGNATURES:
            raise AttributeError(name)
        raw = getattr(self, "_raw_" + full)
        if full in _NO_CHECK:
            return raw

        def checked(*args):
            rc = raw(*args)
            if rc != 0:
                raise B200Error(f"{full} failed (rc={rc}): {self.last_error()}")
            return 0

        checked.__name__ = full
        setattr(self, name, checked)
        return checked


_lib: Optional[_Lib] = None


def lib_path() -> str:
    return os.environ.get("B200STEP_LIB", DEFAULT_LIB)


def get_lib() -> _Lib:
    """Load (once) and return the C-ABI library; raises B200Error when it is absent."""
    global _lib
    if _lib is None:
        _lib = _Lib(lib_path())
    return _lib


def ptr(t) -> Optional[int]:
    """Device/host address of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()
