"""Helpers shared by the parity tests: load a golden fixture, replay it through the oracle."""
import json
import os

import numpy as np
import torch

from oracle import synth
from oracle.llama_oracle import OracleWorker, run_chain

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names(big=False):
    names = sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))
    return [n for n in names if big or not n.startswith("llama2_7b")]


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    return z, meta


def golden_states(z, n_pass, n_shards):
    return [[torch.from_numpy(z[f"h_p{p}_s{i}"]) for i in range(n_shards - 1)] for p in range(n_pass)]


def oracle_replay(meta, dtype=torch.float32):
    cfg = synth.PRESETS[meta["preset"]]
    workers = [OracleWorker(cfg, meta["seed"], s, e, i == 0, dtype) for i, (s, e) in enumerate(meta["splits"])]
    ids = synth.prompt_ids(cfg, meta.get("prompt_seed", meta["seed"]), meta["batch"], meta["prompt_len"])
    return run_chain(workers, ids, meta["new_tokens"])


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.double().flatten()
    b = b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
