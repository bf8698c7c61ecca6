"""CPU-only: the C-ABI library loads without a GPU and exports every symbol include/b200shard.h
declares; the ctypes table matches the header."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200shard.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"B200_API\s+(?:const\s+)?\w+\s*\*?\s*(b200_\w+)\s*\(", src)))


def test_header_declares_entry_points():
    names = declared_functions()
    assert "b200_shard_forward" in names and "b200_kv_create" in names and "b200_hop_send" in names
    assert len(names) >= 25


def test_library_exports_every_declared_symbol():
    from llm_sharding_b200 import _lib
    assert os.path.isfile(_lib.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing


def test_ctypes_table_matches_header():
    from llm_sharding_b200 import _lib
    assert sorted(_lib.PROTOTYPES) == declared_functions()
    lib = _lib.load()
    assert lib.b200_version() >= 100
    assert ctypes.sizeof(_lib.ShardConfig) == 14 * 4


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    from llm_sharding_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.require_cuda()
    from llm_sharding_b200.engine import Shard
    with pytest.raises(RuntimeError):
        Shard({"hidden_size": 256, "intermediate_size": 640, "num_attention_heads": 2, "num_key_value_heads": 2,
               "head_dim": 128, "vocab_size": 1000, "rms_norm_eps": 1e-5, "rope_theta": 1e4}, 1, "cuda:0")


def test_invalid_arguments_are_rejected_without_touching_a_gpu():
    """Argument validation happens before any CUDA call, so it is checkable on CPU."""
    from llm_sharding_b200 import _lib
    lib = _lib.load()
    cfg = _lib.ShardConfig(256, 640, 2, 2, 64, 1000, 1e-5, 1e4, 1, 0, 0, 0, 64, 1)  # head_dim 64: not built
    h = ctypes.c_void_p()
    assert lib.b200_shard_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == _lib.EINVAL
    assert b"head_dim" in lib.b200_last_error()
    cfg = _lib.ShardConfig(256, 640, 2, 2, 128, 1000, 1e-5, 1e4, 0, 0, 0, 0, 64, 1)  # empty layer range
    assert lib.b200_shard_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == _lib.EINVAL
    assert b"start or end is invalid" in lib.b200_last_error()
    with pytest.raises(ValueError):
        _lib.check(_lib.EINVAL)
