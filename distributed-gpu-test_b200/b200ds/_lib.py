"""ctypes binding of libb200ds.so (the C ABI declared in include/b200ds.h).

There is no fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200DS_LIB") or os.path.join(_HERE, "libb200ds.so")   # override: tuning builds only

NCCL_ID_BYTES = 128

c_i64 = C.c_int64
c_u64 = C.c_uint64
c_ptr = C.c_void_p
c_stream = C.c_size_t      # uintptr_t


class StepBufs(C.Structure):
    """struct b200ds_step_bufs (include/b200ds.h)."""
    _fields_ = [
        ("params", c_ptr), ("grads", c_ptr), ("exp_avg", c_ptr), ("exp_avg_sq", c_ptr),
        ("x_stage", c_ptr), ("y_stage", c_ptr),
        ("logits", c_ptr), ("dlogits", c_ptr),
        ("loss", c_ptr),
        ("scratch", c_ptr), ("scratch_bytes", C.c_size_t),
        ("xq", c_ptr), ("wq", c_ptr),
    ]


# name -> (restype, argtypes); must list every symbol include/b200ds.h declares
# (tests/test_abi.py parses the header and checks this table and the .so against it).
PROTOTYPES = {
    "b200ds_last_error": (C.c_char_p, []),
    "b200ds_version": (C.c_int, []),
    "b200ds_device_sm_count": (C.c_int, [C.c_int, C.POINTER(C.c_int)]),
    "b200ds_nccl_unique_id": (C.c_int, [c_ptr]),
    "b200ds_ctx_create": (C.c_int, [C.c_int, C.c_int, C.c_int, c_ptr, C.POINTER(c_ptr)]),
    "b200ds_ctx_destroy": (C.c_int, [c_ptr]),
    "b200ds_ctx_info": (C.c_int, [c_ptr, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b200ds_philox_normal": (C.c_int, [c_ptr, c_i64, c_u64, c_u64, C.POINTER(c_u64), c_stream]),
    "b200ds_philox_normal_counted": (C.c_int, [c_ptr, c_i64, c_u64, c_u64, c_ptr, c_stream]),
    "b200ds_label_rule": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, c_stream]),
    "b200ds_gather_batch": (C.c_int, [c_ptr, c_ptr, c_ptr, c_i64, C.c_int, c_ptr, c_ptr, c_stream]),
    "b200ds_mlp_forward_f32": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_stream]),
    "b200ds_bce_scratch_bytes": (C.c_size_t, [c_i64]),
    "b200ds_bce_loss_grad": (C.c_int, [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_bce_loss_grad_noalloc": (C.c_int, [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_stream]),
    "b200ds_mlp_backward_f32_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_mlp_backward_f32": (C.c_int, [c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_allreduce_bucket": (C.c_int, [c_ptr, c_ptr, c_i64, c_stream]),
    "b200ds_allreduce_bucket_begin": (C.c_int, [c_ptr, c_ptr, c_i64, c_stream]),
    "b200ds_allreduce_bucket_end": (C.c_int, [c_ptr, c_stream]),
    "b200ds_zero1_step": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_double, C.c_double, C.c_double, C.c_double,
                                    c_i64, c_stream]),
    "b200ds_collective": (C.c_int, [c_ptr, C.c_int, C.c_int, c_ptr, c_i64, c_stream]),
    "b200ds_broadcast_params": (C.c_int, [c_ptr, c_ptr, c_i64, c_stream]),
    "b200ds_barrier": (C.c_int, [c_ptr]),
    "b200ds_adam_step": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_double, C.c_double, C.c_double,
                                   C.c_double, c_i64, C.c_float, C.c_int, c_stream]),
    "b200ds_adam_step_counted": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_double, C.c_double, C.c_double,
                                           C.c_double, c_ptr, C.c_float, C.c_int, c_stream]),
    "b200ds_tiny_train_steps_f32": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr,
                                              c_i64, c_i64, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                              C.c_double, C.c_double, c_i64, c_ptr, c_stream]),
    "b200ds_tc_supported": (C.c_int, [c_i64, C.c_int, C.c_int]),
    "b200ds_tc_xq_bytes": (C.c_size_t, [c_i64]),
    "b200ds_tc_wq_bytes": (C.c_size_t, [C.c_int]),
    "b200ds_pack_x_bf16": (C.c_int, [c_ptr, c_i64, C.c_int, c_ptr, c_stream]),
    "b200ds_gen_batch_bf16": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, c_u64, c_u64, C.POINTER(c_u64), c_stream]),
    "b200ds_gen_batch_bf16_counted": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, c_u64, c_u64, c_ptr, c_stream]),
    "b200ds_pack_params_bf16": (C.c_int, [c_ptr, C.c_int, C.c_int, c_ptr, c_stream]),
    "b200ds_pack_x_bf16_rows": (C.c_int, [c_ptr, c_i64, c_i64, C.c_int, c_ptr, c_stream]),
    "b200ds_mlp_forward_tc_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_mlp_forward_tc": (C.c_int, [c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_mlp_backward_tc_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_mlp_backward_tc": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_train_step_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_train_step_host": (C.c_int, [c_ptr, C.POINTER(StepBufs), c_ptr, c_ptr, c_i64, C.c_int, C.c_int,
                                         C.c_double, C.c_double, C.c_double, C.c_double, c_i64, c_ptr, c_stream]),
    "b200ds_mlp_loss_backward_tc": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr,
                                              c_ptr, C.c_size_t, c_stream]),
    "b200ds_mlp_loss_backward_partials_tc": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr,
                                                       C.c_size_t, c_stream]),
    "b200ds_update_tc_exchange_elems": (c_i64, [C.c_int]),
    "b200ds_update_tc": (C.c_int, [c_ptr, c_ptr, C.c_size_t, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                   C.c_double, C.c_double, C.c_double, C.c_double, c_i64, c_ptr, c_ptr, c_stream]),
    "b200ds_update_tc_begin": (C.c_int, [c_ptr, c_ptr, C.c_size_t, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                         C.c_double, C.c_double, C.c_double, C.c_double, c_i64, c_ptr, c_ptr, c_stream]),
    "b200ds_update_tc_end": (C.c_int, [c_ptr, c_ptr, C.c_size_t, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                       C.c_double, C.c_double, C.c_double, C.c_double, c_i64, c_ptr, c_ptr, c_stream]),
    "b200ds_fh_supported": (C.c_int, [c_i64, C.c_int, C.c_int]),
    "b200ds_fh_train_supported": (C.c_int, [c_i64, C.c_int, C.c_int]),
    "b200ds_fh_operand_bytes": (C.c_size_t, [c_i64, c_i64]),
    "b200ds_fh_pack_bf16": (C.c_int, [c_ptr, c_i64, c_i64, c_ptr, c_stream]),
    "b200ds_fh_forward_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_fh_forward": (C.c_int, [c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_fh_gen_supported": (C.c_int, [c_i64, C.c_int]),
    "b200ds_fh_gen_scratch_bytes": (C.c_size_t, [c_i64, C.c_int]),
    "b200ds_fh_gen_batch": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, c_u64, c_u64, C.POINTER(c_u64), c_ptr, C.c_size_t, c_stream]),
    "b200ds_fh_gen_batch_counted": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, c_u64, c_u64, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_fh_logits": (C.c_int, [c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_stream]),
    "b200ds_fh_backward_scratch_bytes": (C.c_size_t, [c_i64, C.c_int, C.c_int]),
    "b200ds_fh_backward": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_int, C.c_int, c_ptr, c_ptr, C.c_size_t, c_stream]),
    "b200ds_ctx_enable_p2p": (C.c_int, [c_ptr, c_i64]),
    "b200ds_allreduce_adam_p2p": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_double, C.c_double, C.c_double,
                                            C.c_double, c_i64, c_ptr, c_stream]),
    "b200ds_train_step_host_prefetch": (C.c_int, [c_ptr, C.POINTER(StepBufs), c_ptr, c_ptr, c_ptr, c_ptr, c_i64, C.c_int,
                                                  C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, c_i64, c_ptr,
                                                  c_stream]),
}


class B200dsError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """Load the in-tree shared library once.  Missing library is a hard error: build it
    with `python -c 'import __graft_entry__ as g; g.build()'` or `make -C .../csrc`."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200dsError(
            f"{LIB_PATH} not found: the b200ds CUDA extension is not built and there is no "
            "CPU or PyTorch fallback. Run __graft_entry__.build().")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)            # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200ds_last_error().decode("utf-8", "replace")
        raise B200dsError(f"{what or 'b200ds call'} failed (code {rc}): {msg}")


def ptr(t) -> int:
    """Raw device/host pointer of a torch tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
