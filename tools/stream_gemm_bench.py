#!/usr/bin/env python
"""Times the 4..32-row weight-streaming tensor-core GEMM alone (b200_gemm_stream_f16) on the five Llama-2-7B shapes, cycling
over enough distinct weight buffers that nothing is served from L2.  CUDA events on the launching stream, back-to-back
launches (so programmatic dependent launch overlaps prologues as in the shard path).  B200SHARD_STREAMK=0 selects the
first-generation kernel (tile per CTA, cluster split-K) for the A/B."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from llm_sharding_b200 import _lib  # noqa: E402

SHAPES = [("qkv", 12288, 4096), ("o_proj", 4096, 4096), ("gate_up", 22016, 4096), ("down", 4096, 11008), ("lm_head", 32000, 4096)]


def main():
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    rows = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "4,16,32").split(",")]
    print(f"B200SHARD_STREAMK={os.environ.get('B200SHARD_STREAMK', '1')}")
    for name, N, K in SHAPES:
        n_buf = max(2, int(400e6 // (N * K * 2)) + 1)
        ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(n_buf)]
        for M in rows:
            x = torch.randn(M, K, device="cuda").half()
            out = torch.empty(M, N, device="cuda")
            for w in ws:
                _lib.check(lib.b200_gemm_stream_f16(x.data_ptr(), w.data_ptr(), M, N, K, 0, out.data_ptr(), st))
            torch.cuda.synchronize()
            reps = max(4, 256 // n_buf)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                for w in ws:
                    lib.b200_gemm_stream_f16(x.data_ptr(), w.data_ptr(), M, N, K, 0, out.data_ptr(), st)
            e1.record()
            torch.cuda.synchronize()
            us = 1000 * e0.elapsed_time(e1) / (reps * n_buf)
            gbs = N * K * 2 / us / 1e3
            print(f"{name:8s} M={M:2d}  {us:7.2f} us  {gbs:7.0f} GB/s  {100 * gbs / 6551.4:5.1f} % of HBM", flush=True)
        del ws


if __name__ == "__main__":
    main()
