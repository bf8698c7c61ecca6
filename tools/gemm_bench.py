#!/usr/bin/env python
"""Times the tcgen05 GEMM (C = A W^T, fp16 operands, fp32 out) through the C ABI for prefill-sized problems and prints
TFLOP/s against the measured cuBLAS bf16 peak.  B200SHARD_GEMM_PAIR=0 selects the single-CTA 128x128 kernel, default is
the CTA-pair 256x256 kernel for M > 128.  python tools/gemm_bench.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from llm_sharding_b200 import _lib  # noqa: E402


def main():
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    peak = 1643.1
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        peak = json.load(open(p))["bf16_tflops"]
    res = {}
    for M in (512, 2048, 8192):
        for name, N, K in (("qkv", 12288, 4096), ("o", 4096, 4096), ("gate_up", 22016, 4096), ("down", 4096, 11008)):
            a = torch.randn(M, K, device="cuda").half()
            w = (torch.randn(N, K, device="cuda") * 0.02).half()
            out = torch.empty(M, N, device="cuda")
            for _ in range(3):
                _lib.check(lib.b200_gemm_f16(a.data_ptr(), w.data_ptr(), M, N, K, out.data_ptr(), st))
            torch.cuda.synchronize()
            reps = 10
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                _lib.check(lib.b200_gemm_f16(a.data_ptr(), w.data_ptr(), M, N, K, out.data_ptr(), st))
            e1.record()
            torch.cuda.synchronize()
            us = 1000 * e0.elapsed_time(e1) / reps
            tf = 2.0 * M * N * K / us / 1e6
            res[f"M={M} {name}"] = {"us": round(us, 1), "tflops": round(tf, 1), "frac_of_measured_peak": round(tf / peak, 3)}
            print(f"M={M:5d} {name:8s} {us:9.1f} us  {tf:7.1f} TFLOP/s  {100 * tf / peak:5.1f}% of measured cuBLAS peak", flush=True)
            del a, w, out
    print(json.dumps({"pair_mode": os.environ.get("B200SHARD_GEMM_PAIR", "1"), "results": res}))


if __name__ == "__main__":
    main()
