"""-m gpu, collected last on purpose: b200_attn_decode_static, the allocation-free twin of b200_attn_decode.

It exists for callers that cannot allocate device memory: a process whose HBM already belongs to another pool,
a per-token loop that must not run into cudaMalloc's device-wide synchronisation, a tool probing K/V pages it
owns without a shard handle.  Its split-K partials live in an array compiled into the library.  The test calls
it through the C ABI twice in a row (second call with a different, smaller problem, so stale partials of the first
call are still sitting in the shared array and some splits are empty) and checks each result against
b200_attn_decode run with ordinary caller-provided scratch, at the tolerance the attention unit test uses; it also
checks the size limit is enforced and that free device memory does not move across a call.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5  # same bound as tests/test_gpu_kernels.py::test_attn_decode_split_k


def _st():
    return torch.cuda.current_stream().cuda_stream


def _rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def _problem(rows, n_heads, n_kv, T_list, seed):
    from test_gpu_kernels import _paged_kv
    g = torch.Generator(device="cuda").manual_seed(seed)
    k_pool, v_pool, bt, dense, max_pages = _paged_kv(rows, n_kv, T_list, 64, g)
    q = torch.randn(rows, n_heads * 128, device="cuda", generator=g)
    t_end = torch.tensor(T_list, dtype=torch.int32, device="cuda")
    return q, k_pool, v_pool, bt, t_end, dense, max_pages


def _both(lib, rows, n_heads, n_kv, T_list, n_splits, seed):
    from llm_sharding_b200 import _lib
    q, k_pool, v_pool, bt, t_end, dense, max_pages = _problem(rows, n_heads, n_kv, T_list, seed)
    scratch = torch.empty(rows * n_heads * (n_splits * 130 + 1), device="cuda")
    ref = torch.full_like(q, float("nan"))
    _lib.check(lib.b200_attn_decode(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows,
                                    n_heads, n_kv, 64, max_pages, n_splits, scratch.data_ptr(), ref.data_ptr(), _st()))
    out = torch.full_like(q, float("nan"))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(lib.b200_attn_decode_static(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(),
                                           rows, n_heads, n_kv, 64, max_pages, n_splits, out.data_ptr(), _st()))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return out, ref, q, dense, free_before, free_after


def test_static_scratch_attention_matches_the_allocating_entry_point_twice_in_a_row():
    from llm_sharding_b200 import _lib
    from test_gpu_kernels import attn_decode_reference
    _lib.require_cuda()
    lib = _lib.load()
    # call 1: Llama-2-7B decode shape, every split busy
    out1, ref1, q1, dense1, _, _ = _both(lib, 1, 32, 32, [700], 10, seed=11)
    assert not torch.isnan(out1).any()
    assert _rel_l2(out1, ref1) <= TOL
    assert _rel_l2(out1, attn_decode_reference(q1, dense1, 32, 32)) <= TOL  # and both agree with plain torch
    # call 2, straight after: fewer rows of work, short sequences (empty splits), the array still holds call 1's partials
    out2, ref2, _, _, free_before, free_after = _both(lib, 3, 8, 4, [5, 130, 64], 16, seed=12)
    assert not torch.isnan(out2).any()
    assert _rel_l2(out2, ref2) <= TOL
    # by the second call the per-device set-up is done: the call itself must not change free device memory
    assert free_before == free_after


def test_static_scratch_attention_rejects_problems_beyond_its_built_in_capacity():
    from llm_sharding_b200 import _lib
    _lib.require_cuda()
    lib = _lib.load()
    rows, n_heads, n_kv, n_splits = 64, 8, 8, 16   # 64*8*16 = 8192 partials > 4096
    q, k_pool, v_pool, bt, t_end, _, max_pages = _problem(rows, n_heads, n_kv, [64] * rows, seed=13)
    out = torch.zeros_like(q)
    rc = lib.b200_attn_decode_static(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows,
                                     n_heads, n_kv, 64, max_pages, n_splits, out.data_ptr(), _st())
    assert rc == _lib.EINVAL
    assert "exceeds the built-in scratch" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(rc)
    # the same problem is fine through the allocating entry point
    scratch = torch.empty(rows * n_heads * (n_splits * 130 + 1), device="cuda")
    _lib.check(lib.b200_attn_decode(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows,
                                    n_heads, n_kv, 64, max_pages, n_splits, scratch.data_ptr(), out.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
