"""-m gpu: single kernels through the C ABI against a plain torch fp32 reference of the same op
(floating-point kernels; tolerances stated per test)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from llm_sharding_b200 import _lib
    _lib.require_cuda()
    return _lib.load()


def _st():
    return torch.cuda.current_stream().cuda_stream


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.mark.parametrize("rows,n", [(1, 256), (3, 4096), (130, 4096), (7, 11008)])
def test_rmsnorm_matches_llama_rmsnorm(lib, rows, n):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(rows * 131 + n)
    x = torch.randn(rows, n, device="cuda", generator=g) * 3.0
    w = (1 + 0.1 * torch.randn(n, device="cuda", generator=g)).bfloat16()
    out = torch.empty(rows, n, device="cuda", dtype=torch.bfloat16)
    _lib.check(lib.b200_rmsnorm(x.data_ptr(), w.data_ptr(), 1e-5, rows, n, out.data_ptr(), _st()))
    ref = w.float() * (x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5))
    # output is bf16: one rounding (2^-8 relative) on top of fp32 arithmetic
    assert torch.allclose(out.float(), ref, rtol=2 ** -7, atol=1e-6)
    assert rel_l2(out.float(), ref) < 3e-3


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("N,K", [(256, 256), (1000, 256), (4096, 4096), (4096, 11008), (1536, 640), (32000, 4096), (24, 72)])
def test_gemv_weight_streaming(lib, M, N, K):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.02).bfloat16().half()  # bf16 model weights held as fp16
    out = torch.full((M, N), float("nan"), device="cuda")
    _lib.check(lib.b200_gemv_f16(x.data_ptr(), w.data_ptr(), M, N, K, out.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
    if M == 1:   # CUDA-core kernel: fp32 activations and accumulation, only the summation order differs
        ref = (x.double() @ w.double().t()).float()
        assert rel_l2(out, ref) < 2e-6
    else:        # tensor-core kernel: activations rounded to fp16 (2^-11 relative), exact products, fp32 accumulation
        ref = (x.half().double() @ w.double().t()).float()
        assert rel_l2(out, ref) < 2e-6
        assert rel_l2(out, (x.double() @ w.double().t()).float()) < 1e-3


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 128, 512), (1, 128, 64), (200, 384, 512), (33, 1000, 256),
                                   (512, 4096, 4096), (260, 12288, 4096), (77, 4096, 11008)])
def test_gemm_tcgen05(lib, M, N, K):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
    out = torch.full((M, N), float("nan"), device="cuda")
    _lib.check(lib.b200_gemm_bf16(a.data_ptr(), w.data_ptr(), M, N, K, out.data_ptr(), _st()))
    torch.cuda.synchronize()
    ref = (a.double() @ w.double().t()).float()
    assert not torch.isnan(out).any()
    assert rel_l2(out, ref) < 3e-5  # exact bf16 products; fp32 accumulation order differs (K up to 11008)


def test_gemv_rejects_bad_shapes(lib):
    from llm_sharding_b200 import _lib
    x = torch.zeros(16, 64, device="cuda")
    w = torch.zeros(8, 64, device="cuda", dtype=torch.bfloat16)
    o = torch.zeros(8, 8, device="cuda")
    with pytest.raises(ValueError):
        _lib.check(lib.b200_gemv_f16(x.data_ptr(), w.data_ptr(), 9, 8, 64, o.data_ptr(), _st()))
    with pytest.raises(ValueError):
        _lib.check(lib.b200_gemm_bf16(w.data_ptr(), w.data_ptr(), 8, 8, 60, o.data_ptr(), _st()))


def _paged_kv(rows, n_kv, T_list, page_size, g, fill=float("nan")):
    """random K/V written into a shuffled page pool; returns pools, block table, dense copies"""
    max_pages = max((t + page_size - 1) // page_size for t in T_list)
    n_pages = rows * max_pages + 3
    # rows the sequences never wrote hold NaN: a kernel that lets them leak into a sum (0 x NaN inside an MMA) is caught
    k_pool = torch.full((n_pages, n_kv, page_size, 128), fill, device="cuda", dtype=torch.bfloat16)
    v_pool = torch.full_like(k_pool, fill)
    perm = torch.randperm(n_pages, generator=torch.Generator().manual_seed(7)).tolist()
    bt = torch.full((rows, max_pages), -1, dtype=torch.int32)
    dense = []
    for r, T in enumerate(T_list):
        k = torch.randn(n_kv, T, 128, device="cuda", generator=g).bfloat16()
        v = torch.randn(n_kv, T, 128, device="cuda", generator=g).bfloat16()
        dense.append((k, v))
        for pi in range((T + page_size - 1) // page_size):
            page = perm.pop()
            bt[r, pi] = page
            lo, hi = pi * page_size, min(T, (pi + 1) * page_size)
            k_pool[page, :, : hi - lo] = k[:, lo:hi]
            v_pool[page, :, : hi - lo] = v[:, lo:hi]
    return k_pool, v_pool, bt.cuda(), dense, max_pages


def attn_decode_reference(q, dense, n_heads, n_kv):
    out = torch.empty_like(q)
    rep = n_heads // n_kv
    for r, (k, v) in enumerate(dense):
        qr = q[r].view(n_heads, 128).double()
        kk = k.double().repeat_interleave(rep, dim=0)
        vv = v.double().repeat_interleave(rep, dim=0)
        p = torch.softmax(torch.einsum("hd,htd->ht", qr, kk) * 128 ** -0.5, dim=-1)
        out[r] = torch.einsum("ht,htd->hd", p, vv).reshape(-1).float()
    return out


@pytest.mark.parametrize("rows,n_heads,n_kv,T_list,n_splits", [
    (1, 32, 32, [640], 10), (1, 4, 2, [1], 3), (3, 8, 8, [65, 128, 7], 4), (2, 32, 8, [1000, 513], 16), (4, 2, 2, [64, 64, 64, 64], 1),
    # n_splits == 8 is served by the cluster variant (partials over distributed shared memory): one row = 256-thread CTAs,
    # several rows = 128-thread CTAs; a single key leaves 7 of the 8 CTAs of a cluster with nothing but the barrier
    (1, 32, 32, [640], 8), (1, 4, 2, [1], 8), (3, 8, 8, [65, 128, 7], 8), (2, 32, 8, [3000, 1537], 8), (1, 8, 2, [2500], 8),
    (8, 4, 4, [5, 9, 1, 130, 64, 65, 2, 300], 8)])
def test_attn_decode_split_k(lib, rows, n_heads, n_kv, T_list, n_splits):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(rows * 100 + n_heads)
    k_pool, v_pool, bt, dense, max_pages = _paged_kv(rows, n_kv, T_list, 64, g)
    q = torch.randn(rows, n_heads * 128, device="cuda", generator=g)
    t_end = torch.tensor(T_list, dtype=torch.int32, device="cuda")
    scratch = torch.empty(rows * n_heads * (n_splits * 130 + 1), device="cuda")
    out = torch.full_like(q, float("nan"))
    _lib.check(lib.b200_attn_decode(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows,
                                    n_heads, n_kv, 64, max_pages, n_splits, scratch.data_ptr(), out.data_ptr(), _st()))
    ref = attn_decode_reference(q, dense, n_heads, n_kv)
    assert rel_l2(out, ref) < 1e-5  # fp32 softmax and accumulation over bf16 K/V; only summation order differs


@pytest.mark.parametrize("rows,n_heads,n_kv,T_list", [
    (1, 2, 2, [1]), (2, 4, 4, [128, 129]), (3, 8, 2, [65, 640, 7]), (4, 32, 32, [704, 704, 333, 1]), (2, 32, 8, [1500, 2048]),
    (32, 4, 4, list(range(40, 40 + 32 * 37, 37))), (1, 8, 1, [300]), (2, 2, 2, [4096, 3000]), (1, 2, 2, [9000])])
def test_attn_decode_tensor_core(lib, rows, n_heads, n_kv, T_list):
    """Batched decode attention on tcgen05 + TMA (one CTA per row and K/V head, grouped query heads as the narrow operand,
    two passes over the keys).  fp32 q travels as three bf16 parts, probabilities as hi + lo: the same 1e-5 bound as the
    CUDA-core kernel up to 4096 keys.  Beyond that the output is one fp32 tensor-memory accumulator fed by hundreds of
    MMAs whose adder truncates rather than rounds (measured 2.2e-5 at 9000 keys): bound 4e-5 there.  Rows past a
    sequence's end hold large finite values (stale page contents), not NaN."""
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(rows * 100 + n_heads + T_list[0])
    k_pool, v_pool, bt, dense, max_pages = _paged_kv(rows, n_kv, T_list, 64, g, fill=1.0e4)
    q = torch.randn(rows, n_heads * 128, device="cuda", generator=g)
    t_end = torch.tensor(T_list, dtype=torch.int32, device="cuda")
    out = torch.full_like(q, float("nan"))
    for _ in range(2):   # the second launch meets whatever the first left in shared / tensor memory
        _lib.check(lib.b200_attn_decode_tc(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows,
                                           n_heads, n_kv, 64, max_pages, k_pool.shape[0], max(T_list), out.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    ref = attn_decode_reference(q, dense, n_heads, n_kv)
    assert rel_l2(out, ref) < (1e-5 if max(T_list) <= 4096 else 4e-5)


def test_attn_decode_tensor_core_rejects_what_it_does_not_serve(lib):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(1)
    k_pool, v_pool, bt, dense, max_pages = _paged_kv(1, 1, [70], 64, g, fill=0.0)
    q = torch.randn(1, 16 * 128, device="cuda", generator=g)
    t_end = torch.tensor([70], dtype=torch.int32, device="cuda")
    out = torch.zeros_like(q)
    with pytest.raises(ValueError):   # 16 query heads on one K/V head: more than the narrow operand holds
        _lib.check(lib.b200_attn_decode_tc(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), 1,
                                           16, 1, 64, max_pages, k_pool.shape[0], 70, out.data_ptr(), _st()))
    with pytest.raises(ValueError):   # page size must be a multiple of 64
        _lib.check(lib.b200_attn_decode_tc(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), 1,
                                           16, 16, 48, max_pages, k_pool.shape[0], 70, out.data_ptr(), _st()))


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (200, 384, 512), (512, 4096, 4096), (77, 4096, 11008)])
def test_gemm_tcgen05_fp16_operands(lib, M, N, K):
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).half()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16().half()  # bf16 model weights held as fp16
    out = torch.full((M, N), float("nan"), device="cuda")
    _lib.check(lib.b200_gemm_f16(a.data_ptr(), w.data_ptr(), M, N, K, out.data_ptr(), _st()))
    torch.cuda.synchronize()
    ref = (a.double() @ w.double().t()).float()
    assert not torch.isnan(out).any()
    assert rel_l2(out, ref) < 3e-5  # exact fp16 x fp16 products, fp32 accumulation


@pytest.mark.parametrize("M,N,K,split", [
    (9, 128, 64, 1), (16, 256, 512, 2), (17, 384, 1024, 4), (32, 4096, 4096, 8), (32, 4096, 11008, 8), (24, 12288, 4096, 0),
    (32, 22016, 4096, 0), (12, 200, 320, 1), (32, 32000, 4096, 0), (1, 128, 4096, 8), (31, 4096, 11008, 0)])
def test_gemm_stream_tensor_core(lib, M, N, K, split):
    """The weight-streaming tcgen05 GEMM of batched decode (9..32 rows): weight tile as the 128-row operand, split-K over
    a cluster reduced through distributed shared memory; ragged N (200) and K (320 = 5 k-steps) included."""
    from llm_sharding_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N + K + split)
    x = torch.randn(M, K, device="cuda", generator=g).half()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16().half()
    out = torch.full((M, N), float("nan"), device="cuda")
    for _ in range(2):  # twice: the second launch meets whatever the first left in shared memory / TMEM
        _lib.check(lib.b200_gemm_stream_f16(x.data_ptr(), w.data_ptr(), M, N, K, split, out.data_ptr(), _st()))
    torch.cuda.synchronize()
    ref = (x.double() @ w.double().t()).float()
    assert not torch.isnan(out).any()
    assert rel_l2(out, ref) < 3e-5  # exact fp16 x fp16 products, fp32 accumulation
    out2 = torch.empty_like(out)
    _lib.check(lib.b200_gemm_stream_f16(x.data_ptr(), w.data_ptr(), M, N, K, split, out2.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert torch.equal(out, out2)   # the split-K sum runs in rank order: bit-reproducible


def test_gemm_stream_rejects_bad_arguments(lib):
    from llm_sharding_b200 import _lib
    x = torch.zeros(40, 64, device="cuda").half()
    w = torch.zeros(128, 64, device="cuda").half()
    out = torch.zeros(40, 128, device="cuda")
    for args in ((33, 128, 64, 1), (4, 128, 60, 1), (4, 128, 32, 1), (4, 128, 64, 3), (4, 128, 64, 2)):
        with pytest.raises(ValueError):
            _lib.check(lib.b200_gemm_stream_f16(x.data_ptr(), w.data_ptr(), *args, out.data_ptr(), _st()))
