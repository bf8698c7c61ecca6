#!/usr/bin/env python
"""Times one prefill call (B x S tokens through `--layers` Llama-2-7B-sized layers) with CUDA events.
B200SHARD_ATTN_TC=0 selects the mma.sync attention kernel, B200SHARD_GEMM_PAIR=0 the single-CTA GEMM."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--causal", type=int, default=0)
    ap.add_argument("--cases", default="1x512,1x2048,4x2048,1x8192", help="comma-separated BxS list")
    ap.add_argument("--iters", type=int, default=4, help="timed calls per case (>= 3: the median of all but the first is reported)")
    a = ap.parse_args()
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.engine import Shard, read_model_config
    cfg = dict(synthetic.LLAMA2_7B, num_hidden_layers=a.layers)
    d = tempfile.mkdtemp()
    synthetic.write_model_dir(d, cfg)
    prov = synthetic.RandomStateProvider(cfg, "cuda:0")
    sh = Shard(read_model_config(d), a.layers, torch.device("cuda:0"), max_tokens=16384, max_batch=16)
    for l in range(a.layers):
        sh.set_layer(l, prov(f"block_{l}.pth"))
    for B, S in [tuple(int(v) for v in c.split("x")) for c in a.cases.split(",")]:
        x = (torch.randn(B, S, 4096, device="cuda") * 0.1).bfloat16()
        kv = sh.new_kv(max_slots=B, max_seq_len=S + 64)
        times = []
        for it in range(a.iters):
            kv.reset(-1)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sh.forward(x, kv, list(range(B)), causal=bool(a.causal))
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        ms = sorted(times[1:])[(len(times) - 1) // 2] if len(times) > 1 else times[0]
        flops = a.layers * (2 * 202375168 * B * S + 4 * B * S * S * 4096 * (0.5 if a.causal else 1.0))
        print(f"B={B} S={S}: {ms / a.layers * 1000:8.1f} us/layer  {B * S / ms * 1000 / 1e3 * (32 / a.layers) / 32:8.1f} k tok/s per 32-layer-equivalent x(32/{a.layers})"
              f"  {flops / ms / 1e9:7.1f} TFLOP/s", flush=True)
        kv.close()


if __name__ == "__main__":
    main()
