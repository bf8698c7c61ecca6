"""ctypes binding of ``libb200shard.so`` (``include/b200shard.h``).

The library is the product: nothing here falls back to torch or to the CPU.  ``load()``
raises if the shared object is missing; every compute entry point raises if CUDA is not
available.  Status codes map to the exception classes the reference raises at the same
places (``ValueError`` / ``RuntimeError`` with "[ERROR] ..." texts, utils/shard_loader.py:40-41,
utils/node_worker.py:134-135,196-197,234-235).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libb200shard.so")

OK, EINVAL, ESTATE, ECUDA, ENOMEM, ETIMEOUT = 0, -1, -2, -3, -4, -5
DT_BF16, DT_F32, DT_F16 = 0, 1, 2
HOP_HANDLE_BYTES = 80
HOP_HEADER_BYTES = 256

TORCH_TO_DT = {torch.bfloat16: DT_BF16, torch.float32: DT_F32, torch.float16: DT_F16}


class ShardConfig(C.Structure):
    _fields_ = [
        ("hidden_size", C.c_int32), ("intermediate_size", C.c_int32), ("n_heads", C.c_int32),
        ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32), ("vocab_size", C.c_int32),
        ("rms_eps", C.c_float), ("rope_theta", C.c_float), ("n_layers", C.c_int32),
        ("has_final_norm", C.c_int32), ("has_lm_head", C.c_int32), ("has_embedding", C.c_int32),
        ("max_tokens", C.c_int32), ("max_batch", C.c_int32),
    ]


# name -> (restype, argtypes); exactly the entry points include/b200shard.h declares
_P, _I, _I64, _U64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
PROTOTYPES = {
    "b200_shard_create": (_I, [C.POINTER(ShardConfig), _I, C.POINTER(_P)]),
    "b200_shard_set_layer": (_I, [_P, _I] + [_P] * 9 + [_I, _P]),
    "b200_shard_set_final_norm": (_I, [_P, _P, _I, _P]),
    "b200_shard_set_lm_head": (_I, [_P, _P, _I, _P]),
    "b200_shard_set_embedding": (_I, [_P, _P, _I, _P]),
    "b200_shard_set_rope_inv_freq": (_I, [_P, C.POINTER(C.c_float), _I]),
    "b200_shard_destroy": (_I, [_P]),
    "b200_kv_create": (_I, [_P, _I, _I, _I, C.POINTER(_P)]),
    "b200_kv_reset": (_I, [_P, _I, _P]),
    "b200_kv_seq_len": (_I, [_P, _I]),
    "b200_kv_error": (_I, [_P]),
    "b200_kv_destroy": (_I, [_P]),
    "b200_shard_forward": (_I, [_P, _P, C.POINTER(C.c_int32), _P, _I, _I, _I, _P, _P]),
    "b200_shard_lm_head_argmax": (_I, [_P, _P, _I, _I, _P, _P, _P]),
    "b200_shard_forward_argmax": (_I, [_P, _P, C.POINTER(C.c_int32), _P, _I, _I, _I, _P, _P, _P, _P]),
    "b200_shard_embed": (_I, [_P, _P, _I, _P, _P]),
    "b200_shard_step_tokens": (_I, [_P, _P, C.POINTER(C.c_int32), _P, _I, _I, _I, _P, _P, _P, _P]),
    "b200_shard_forward_partial": (_I, [_P, _P, C.POINTER(C.c_int32), _P, _I, _I, _I, _I, _P, _P]),
    "b200_shard_set_vocab_tail": (_I, [_P, _P, _P, _I, _I, _P]),
    "b200_shard_token_from_partial": (_I, [_P, _P, _I, _P, _P, _P]),
    "b200_shard_run_kernel": (_I, [_P, _P, _I, _I, _I, _P]),
    "b200_hop_create": (_I, [_I, _I, _I64, C.POINTER(_P)]),
    "b200_hop_export": (_I, [_P, _P]),
    "b200_hop_connect": (_I, [_P, _P, _I, _I]),
    "b200_hop_send": (_I, [_P, _P, _I64, _U64, _P]),
    "b200_hop_recv": (_I, [_P, _P, _I64, _U64, _I64, _P]),
    "b200_hop_send_msg": (_I, [_P, _P, _I, _P, _I64, _U64, _P]),
    "b200_hop_poll": (_I, [_P, _U64, _P]),
    "b200_hop_peer_slot": (_P, [_P, _U64]),
    "b200_hop_local_slot": (_P, [_P, _U64]),
    "b200_hop_error": (_I, [_P]),
    "b200_hop_destroy": (_I, [_P]),
    "b200_last_error": (C.c_char_p, []),
    "b200_version": (_I, []),
    "b200_shard_counters": (_I, [_P, C.POINTER(_U64)]),
    "b200_shard_reset_counters": (_I, [_P]),
    "b200_shard_set_graphs": (_I, [_P, _I]),
    "b200_shard_set_io_dtype": (_I, [_P, _I]),
    "b200_rmsnorm": (_I, [_P, _P, C.c_float, _I, _I, _P, _P]),
    "b200_attn_decode": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "b200_attn_decode_tc": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
    "b200_attn_decode_static": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    "b200_attn_prefill": (_I, [_P] * 6 + [_I] * 9 + [_P, _P]),
    "b200_gemm_bf16": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "b200_gemm_f16": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "b200_gemv_f16": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "b200_gemm_stream_f16": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the CUDA library (works without a GPU: cudart is linked statically and only
    touches the driver on the first CUDA call)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"[ERROR] {LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU or torch fallback for the shard path)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def require_cuda() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("[ERROR] no CUDA device: the B200 shard path has no CPU fallback")


def last_error() -> str:
    return load().b200_last_error().decode(errors="replace")


def check(status: int) -> int:
    if status >= 0:
        return status
    msg = last_error()
    if status == EINVAL:
        raise ValueError(msg)
    if status == ENOMEM:
        raise torch.cuda.OutOfMemoryError(msg)
    raise RuntimeError(msg)


def stream_ptr(device: torch.device | None = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()
