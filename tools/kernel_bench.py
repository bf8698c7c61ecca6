#!/usr/bin/env python
"""Times each decode-path kernel of one Llama-2-7B layer stack alone (warm, back to back over the layers, CUDA
events) for 1..4 token rows and prints achieved HBM bandwidth.  python tools/kernel_bench.py [--layers 8]"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--context", type=int, default=576)
    a = ap.parse_args()
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.engine import Shard, read_model_config
    cfg = dict(synthetic.LLAMA2_7B, num_hidden_layers=a.layers)
    d = tempfile.mkdtemp()
    synthetic.write_model_dir(d, cfg)
    prov = synthetic.RandomStateProvider(cfg, "cuda:0")
    sh = Shard(read_model_config(d), a.layers, torch.device("cuda:0"), has_final_norm=True, has_lm_head=True, max_tokens=4 * a.context, max_batch=4)
    for l in range(a.layers):
        sh.set_layer(l, prov(f"block_{l}.pth"))
    sh.set_final_norm(prov("final_norm.pth")["weight"])
    sh.set_lm_head(prov("lm_head.pth")["weight"])
    H, I, V = cfg["hidden_size"], cfg["intermediate_size"], cfg["vocab_size"]
    bytes_ = {"qkv": 3 * H * H * 2, "attn": 2 * a.context * H * 2, "o": H * H * 2, "gate_up": 2 * I * H * 2, "down": H * I * 2, "lm_head": V * H * 2}
    out = {}
    for rows in (1, 2, 3):
        kv = sh.new_kv(max_slots=4, max_seq_len=a.context + 64)
        x = torch.randn(rows, a.context, H, device="cuda").bfloat16() * 0.1
        sh.forward(x, kv, list(range(rows)))                       # context in the cache
        sh.forward(x[:, :1].contiguous(), kv, list(range(rows)))   # one decode step: sets the per-call metadata for `rows` rows
        res = {}
        for name, kind in (("qkv", 0), ("attn", 1), ("o", 2), ("gate_up", 3), ("down", 4), ("lm_head", 5)):
            n_l = 1 if kind == 5 else a.layers
            for l in range(n_l):
                sh.run_kernel(kv, kind, l, rows)
            torch.cuda.synchronize()
            reps = 8 if kind != 5 else 32
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                for l in range(n_l):
                    sh.run_kernel(kv, kind, l, rows)
            e1.record()
            torch.cuda.synchronize()
            us = 1000 * e0.elapsed_time(e1) / (reps * n_l)
            b = bytes_[name] * (rows if name == "attn" else 1)
            res[name] = {"us": round(us, 2), "GBps": round(b / us / 1e3, 1)}
        out[f"rows={rows}"] = res
        kv.close()
        print(f"rows={rows}: " + "  ".join(f"{k} {v['us']}us {v['GBps']}GB/s" for k, v in res.items()), flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
