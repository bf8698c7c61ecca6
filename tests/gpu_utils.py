"""Helpers for the -m gpu parity tests: build an in-process chain of this repo's NodeWorkers on
cuda:0 from a golden fixture's description."""
import itertools
import os

import torch

from oracle import synth

_port = itertools.count(46000)
_shard_dirs: dict = {}


def shard_dir(tmp_root: str, preset: str, seed: int) -> str:
    key = (preset, seed)
    if key not in _shard_dirs:
        d = os.path.join(tmp_root, f"{preset}_{seed}_bfloat16")
        synth.write_shards(d, synth.PRESETS[preset], seed, torch.bfloat16)
        _shard_dirs[key] = d
    return _shard_dirs[key]


def build_chain(tmp_root, preset, seed, splits, max_batch=1, max_tokens=512, causal=False, use_graphs=True,
                max_seq_len=None, device="cuda:0", dtype=torch.bfloat16):
    from llm_sharding_b200.chain import LoopbackCommunicator
    from llm_sharding_b200.node_worker import NodeWorker
    cfg = synth.PRESETS[preset]
    provider = None
    if cfg.hidden_size >= 4096:   # 7B-dimension presets: weights straight from the generators, no gigabyte .pth files
        d = synth.write_model_dir(os.path.join(tmp_root, f"{preset}_cfg"), cfg)
        provider = synth.state_provider(cfg, seed)
    else:
        d = shard_dir(tmp_root, preset, seed)
    n = len(splits)
    ports = [next(_port) for _ in range(n)]
    workers = []
    for i, (s, e) in enumerate(splits):
        src, dst = f"tcp://*:{ports[i]}", f"tcp://127.0.0.1:{ports[(i + 1) % n]}"
        w = NodeWorker(src, dst, i == 0, d, device=device, dtype=dtype,
                       communicator=LoopbackCommunicator(src, dst), max_tokens=max_tokens, max_batch=max_batch,
                       causal=causal, use_graphs=use_graphs, max_seq_len=max_seq_len, state_provider=provider)
        w.load_shards(s, e)
        workers.append(w)
    return workers


def close_chain(workers):
    for w in workers:
        w.past_key_value = None
        if w.shard is not None:
            w.shard.engine.close()
        w.communicator.close()
