"""ctypes binding of the C ABI declared in include/pdt.h (lib/libpdt.so, sm_100a only).

There is no fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PDT_LIB") or os.path.normpath(os.path.join(_HERE, "..", "lib", "libpdt.so"))     # PDT_LIB: an experimental build

c_p = ctypes.c_void_p
c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_f = ctypes.c_float

# name -> (restype, argtypes); mirrors include/pdt.h one to one
SIGNATURES = {
    "pdt_last_error": (ctypes.c_char_p, []),
    "pdt_version": (c_int, []),
    "pdt_num_tiles": (c_i64, [c_i64]),
    "pdt_mnist_scratch_bytes": (c_i64, [c_i64]),
    "pdt_nll_scratch_bytes": (c_i64, [c_i64]),
    "pdt_conv1_pool_relu_fwd": (c_int, [c_p, c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_conv1_pool_relu_fwd_u8": (c_int, [c_p, c_f, c_f, c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_conv1_fwd_masks_u8": (c_int, [c_p, c_f, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_i64, c_p]),
    "pdt_conv2_bwd_data_conv1_bwd_weight_u8": (c_int, [c_p, c_p, c_p, c_p, c_p, c_f, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_conv1_bwd_weight_u8": (c_int, [c_p, c_p, c_p, c_f, c_f, c_p, c_i64, c_p]),
    "pdt_conv2_drop_pool_relu_fwd": (c_int, [c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_conv2_drop_pool_relu_fwd_tc": (c_int, [c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_head_fwd": (c_int, [c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_head_nll_scratch_bytes": (c_i64, [c_i64]),
    "pdt_head_fwd_nll": (c_int, [c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_i64, c_p, c_int, c_i64, c_p, c_p, c_p,
                                 c_p, c_int, c_p]),
    "pdt_conv2_head_train": (c_int, [c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_p,
                                     c_int, c_i64, c_p, c_p, c_p, c_p, c_int, c_p, c_p, c_i64, c_p]),
    "pdt_conv2_head_train_tc": (c_int, [c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_p,
                                        c_int, c_i64, c_p, c_p, c_p, c_p, c_int, c_p, c_p, c_i64, c_p]),
    "pdt_dropout_masks": (c_int, [c_p, c_p, c_i64, c_f, c_p, c_p]),
    "pdt_conv1_fwd_masks": (c_int, [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_p, c_i64, c_p]),
    "pdt_nll_metrics_fwd": (c_int, [c_p, c_p, c_i64, c_int, c_int, c_i64, c_p, c_p, c_p]),
    "pdt_nll_metrics_fwd_log": (c_int, [c_p, c_p, c_i64, c_int, c_int, c_i64, c_p, c_p, c_p, c_p, c_int, c_p]),
    "pdt_nll_metrics_fwd_static": (c_int, [c_p, c_p, c_i64, c_int, c_int, c_i64, c_p, c_p]),
    "pdt_nll_bwd": (c_int, [c_p, c_p, c_p, c_p, c_i64, c_int, c_i64, c_p]),
    "pdt_head_bwd": (c_int, [c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_head_bwd_nll": (c_int, [c_p, c_p, c_p, c_i64, c_p, c_p, c_p, c_p, c_p, c_f, c_f, c_p, c_p, c_i64, c_p]),
    "pdt_conv2_bwd_data": (c_int, [c_p, c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_conv2_bwd_data_conv1_bwd_weight": (c_int, [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_conv2_bwd_weight": (c_int, [c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_conv1_bwd_weight": (c_int, [c_p, c_p, c_p, c_p, c_i64, c_p]),
    "pdt_grad_reduce": (c_int, [c_p, c_i64, c_f, c_f, c_p, c_p, c_p, c_p]),
    "pdt_flat_scale_cast": (c_int, [c_p, c_p, c_i64, c_f, c_int, c_p]),
    "pdt_comm_unique_id": (c_int, [c_p]),
    "pdt_comm_init": (c_int, [c_p, c_int, c_int, ctypes.POINTER(c_p)]),
    "pdt_comm_destroy": (c_int, [c_p]),
    "pdt_allreduce_flat": (c_int, [c_p, c_p, c_i64, c_p]),
    "pdt_comm_mem_alloc": (c_int, [c_i64, ctypes.POINTER(c_p)]),
    "pdt_comm_mem_free": (c_int, [c_p]),
    "pdt_comm_register": (c_int, [c_p, c_p, c_i64, ctypes.POINTER(c_p)]),
    "pdt_comm_deregister": (c_int, [c_p, c_p]),
    "pdt_p2p_bytes": (c_i64, [c_i64]),
    "pdt_p2p_alloc": (c_int, [c_i64, ctypes.POINTER(c_p), c_p]),
    "pdt_p2p_open": (c_int, [c_p, ctypes.POINTER(c_p)]),
    "pdt_p2p_close": (c_int, [c_p]),
    "pdt_p2p_free": (c_int, [c_p]),
    "pdt_p2p_allreduce": (c_int, [c_p, c_int, c_int, c_p, c_i64, c_i64, c_p]),
    "pdt_p2p_set_timeout_ms": (c_int, [c_i64]),
    "pdt_p2p_reduce_adam": (c_int, [c_p, c_int, c_int, c_p, c_i64, c_f, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i64,
                                    c_p]),
    "pdt_p2p_error": (c_int, [c_p, ctypes.POINTER(c_int)]),
    "pdt_p2p_error_async": (c_int, [c_p, c_p, c_p]),
    "pdt_sgd_step_flat": (c_int, [c_p, c_p, c_p, c_i64, c_f, c_p, c_p]),
    "pdt_adam_step_flat": (c_int, [c_p, c_p, c_p, c_p, c_p, c_i64, c_f, c_p, c_p]),
    "pdt_adam_step_reduce": (c_int, [c_p, c_i64, c_f, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "pdt_memcpy_async": (c_int, [c_p, c_p, c_i64, c_p]),
    "pdt_ring_push": (c_int, [c_p, c_p, c_p, c_int, c_p]),
    "pdt_fma_peak": (c_int, [c_p, c_int, c_int, c_int, c_int, c_p]),
}

_lib = None


class PdtError(RuntimeError):
    pass


def load():
    """Load lib/libpdt.so; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PdtError(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C pytorch-distributed-template_b200/csrc`" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def call(name, *args):
    """Call an int-returning entry point and raise PdtError on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise PdtError("%s failed (%d): %s" % (name, rc, lib.pdt_last_error().decode()))


def ptr(t):
    return None if t is None else t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise PdtError("this path runs on a CUDA device only (got a %s tensor); there is no CPU fallback"
                           % t.device)


def num_tiles(n):
    return (n + 31) // 32
