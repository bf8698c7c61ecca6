#!/usr/bin/env python
"""Times batched decode attention alone: the CUDA-core split-K kernel (b200_attn_decode) against the tcgen05 + TMA kernel
(b200_attn_decode_tc) on Llama-2-7B head geometry (32 heads, MHA), `rows` sequences of T cached keys each in a paged pool.
K/V bytes per call = rows * 2 * T * 4096 * 2; every call reads a pool far larger than L2 at 16+ rows x 704 keys."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from llm_sharding_b200 import _lib  # noqa: E402


def main():
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    n_heads = n_kv = 32
    for rows, T in [(8, 192), (8, 704), (16, 192), (16, 704), (32, 192), (32, 704), (32, 1900)]:
        per = (T + 63) // 64
        n_pages = rows * per
        k_pool = (torch.randn(n_pages, n_kv, 64, 128, device="cuda") * 0.5).bfloat16()
        v_pool = (torch.randn(n_pages, n_kv, 64, 128, device="cuda") * 0.5).bfloat16()
        bt = torch.randperm(n_pages).view(rows, per).int().cuda()
        t_end = torch.full((rows,), T, dtype=torch.int32, device="cuda")
        q = torch.randn(rows, n_heads * 128, device="cuda")
        out = torch.empty_like(q)
        ns = 2 if rows <= 16 else 1
        scratch = torch.zeros(rows * n_heads * (ns * 130 + 1), device="cuda")
        nbytes = rows * 2 * T * n_kv * 128 * 2

        def cuda_core():
            lib.b200_attn_decode(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows, n_heads,
                                 n_kv, 64, per, ns, scratch.data_ptr(), out.data_ptr(), st)

        def tensor_core():
            lib.b200_attn_decode_tc(q.data_ptr(), k_pool.data_ptr(), v_pool.data_ptr(), bt.data_ptr(), t_end.data_ptr(), rows, n_heads,
                                    n_kv, 64, per, n_pages, T, out.data_ptr(), st)
        res = []
        for fn in (cuda_core, tensor_core):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50):
                fn()
            e1.record()
            torch.cuda.synchronize()
            us = 1000 * e0.elapsed_time(e1) / 50
            res.append((us, nbytes / us / 1e3))
        print(f"rows={rows:2d} T={T:4d} {nbytes / 1e6:7.1f} MB  cuda-core {res[0][0]:7.1f} us {res[0][1]:6.0f} GB/s   "
              f"tensor-core {res[1][0]:7.1f} us {res[1][1]:6.0f} GB/s ({100 * res[1][1] / 6551.4:4.1f} % of HBM)", flush=True)


if __name__ == "__main__":
    main()
