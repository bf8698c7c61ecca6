#!/usr/bin/env python
"""Batched single-token decode through `--layers` Llama-2-7B-sized layers: B rows (one KV slot each) with `--context`
cached tokens, CUDA-graph replay, CUDA events.  Prints us per layer per step and the HBM fraction of the algorithmic bytes
(weights once + every row's K/V).  B200SHARD_STREAM_GEMM=0 selects the 128 x 128 tile GEMMs for 9..32 rows (A/B)."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--context", type=int, default=128)
    ap.add_argument("--rows", default="1,4,8,9,16,32")
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--page-size", type=int, default=64)
    a = ap.parse_args()
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.engine import Shard, read_model_config
    cfg = dict(synthetic.LLAMA2_7B, num_hidden_layers=a.layers)
    d = tempfile.mkdtemp()
    synthetic.write_model_dir(d, cfg)
    prov = synthetic.RandomStateProvider(cfg, "cuda:0")
    rows_list = [int(v) for v in a.rows.split(",")]
    bmax = max(rows_list)
    sh = Shard(read_model_config(d), a.layers, torch.device("cuda:0"), max_tokens=bmax * a.context, max_batch=bmax)
    for l in range(a.layers):
        sh.set_layer(l, prov(f"block_{l}.pth"))
    H = cfg["hidden_size"]
    w_bytes = 202383360 * 2
    for B in rows_list:
        kv = sh.new_kv(max_slots=B, page_size=a.page_size, max_seq_len=a.context + a.steps + 80)
        x = (torch.randn(B, a.context, H, device="cuda") * 0.1).bfloat16()
        sh.forward(x, kv, list(range(B)))
        step = (torch.randn(B, 1, H, device="cuda") * 0.1).bfloat16()
        for _ in range(8):
            sh.forward(step, kv, list(range(B)))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            sh.forward(step, kv, list(range(B)))
        e1.record()
        torch.cuda.synchronize()
        us_layer = 1000 * e0.elapsed_time(e1) / a.steps / a.layers
        T = a.context + 8 + a.steps // 2
        alg = w_bytes + B * 2 * T * 4096 * 2
        print(f"B={B:3d}: {us_layer:7.1f} us/layer/step  {B / (us_layer * 32 / 1e6):9.0f} tok/s at 32 layers  "
              f"{alg / us_layer / 1e3:7.0f} GB/s = {100 * alg / us_layer / 1e3 / 6551.4:5.1f} % of HBM", flush=True)
        kv.close()


if __name__ == "__main__":
    main()
