"""TEST / MEASUREMENT INFRASTRUCTURE — times the REAL reference chain on the host cores.

Used only by ``bench.py``'s ``cpu_baseline`` and ``--impl reference`` legs (never by the product path).  The
reference's own ``NodeWorker`` / ``LlamaShardPart`` / ``Communicator`` (``utils/node_worker.py``,
``utils/shard_loader.py``) run unmodified — imported from ``/root/reference`` in the build container, or from the
git-ignored copy ``baseline/_ref/`` that ``__graft_entry__.build()`` makes there so that the files travel to the GPU
box — behind the two-point transformers-5.5 compat shim of ``ref_harness.py``.

Workload: Llama-2-7B architecture, random init, CPU fp32, N ``NodeWorker``s in one process forming a ring over loopback
ZeroMQ (the in-process pattern of ``utils/node_profiler.py:1174-1236``), one request at a time (the reference has a
single ``DynamicCache`` per worker), greedy decode.  Timing protocol of the reference's profiler
(``utils/node_profiler.py:865-901,946-960``): one warm-up request, then ``time.perf_counter`` around every pass.

Bounded sample: the 32 layer files are 4 distinct random-init ``block_i.pth`` (fp32, 809 MB each) plus symlinks, so the
set-up writes 3.2 GB instead of 26 GB; every layer is still its own 809 MB of weights in RAM after
``LlamaShardPart`` loaded it (26 GB in all, streamed from DRAM on every token exactly like distinct weights would be).
"""
from __future__ import annotations

import json
import os
import shutil
import tempfile
import time

import torch

from . import ref_harness, synth

LLAMA2_7B = dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                 num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096,
                 rms_norm_eps=1e-5, rope_theta=10000.0)
DISTINCT_LAYERS = 4


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def available() -> bool:
    return ref_harness.reference_available()


def write_shard_dir(path: str, n_layers: int, seed: int = 0) -> str:
    """ModelSharder layout (utils/model_sharder.py:64-94) for a random-init Llama-2-7B-architecture model."""
    os.makedirs(path, exist_ok=True)
    cfg = dict(LLAMA2_7B, num_hidden_layers=n_layers)
    full = dict(cfg, architectures=["LlamaForCausalLM"], model_type="llama", hidden_act="silu", attention_bias=False,
                mlp_bias=False, tie_word_embeddings=False, bos_token_id=1, eos_token_id=2, pad_token_id=0)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(full, f)
    synth.write_tokenizer(path, cfg["vocab_size"])
    g = torch.Generator().manual_seed(seed)
    H, I, V = cfg["hidden_size"], cfg["intermediate_size"], cfg["vocab_size"]
    mat = lambda r, c: torch.randn(r, c, generator=g) * 0.02
    torch.save({"weight": mat(V, H)}, os.path.join(path, "embedding.pth"))
    torch.save({"weight": mat(V, H)}, os.path.join(path, "lm_head.pth"))
    torch.save({"weight": torch.ones(H)}, os.path.join(path, "final_norm.pth"))
    for i in range(min(DISTINCT_LAYERS, n_layers)):
        torch.save({"self_attn.q_proj.weight": mat(H, H), "self_attn.k_proj.weight": mat(H, H),
                    "self_attn.v_proj.weight": mat(H, H), "self_attn.o_proj.weight": mat(H, H),
                    "mlp.gate_proj.weight": mat(I, H), "mlp.up_proj.weight": mat(I, H), "mlp.down_proj.weight": mat(H, I),
                    "input_layernorm.weight": torch.ones(H), "post_attention_layernorm.weight": torch.ones(H)},
                   os.path.join(path, f"block_{i}.pth"))
    for i in range(DISTINCT_LAYERS, n_layers):
        os.symlink(f"block_{i % DISTINCT_LAYERS}.pth", os.path.join(path, f"block_{i}.pth"))
    return path


def even_splits(n_stages: int, n_layers: int):
    base, extra = divmod(n_layers, n_stages)
    out, s = [], 0
    for r in range(n_stages):
        e = s + base + (1 if r < extra else 0)
        out.append((s, e))
        s = e
    return out


class TimedReferenceChain:
    """The reference chain plus a stopwatch: per pass, the time inside pass_through_shard of each worker and the time in
    Communicator.transfer_data + receive_data (the hop, utils/node_worker.py:44-67)."""

    def __init__(self, n_stages: int = 2, n_layers: int = 32, base_port: int = 41900):
        torch.set_num_threads(host_threads())
        self.threads = torch.get_num_threads()
        self.root = tempfile.mkdtemp(prefix="ref_chain_")
        self.n_layers, self.n_stages = n_layers, n_stages
        t0 = time.perf_counter()
        shards = write_shard_dir(os.path.join(self.root, "llama2_7b_float32"), n_layers)
        self.setup_write_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        # Set-up shortcut, outside the reference's files and outside every timed region: nn.Linear's random init is skipped
        # while the reference builds its layers — load_state_dict (utils/shard_loader.py:42-45) overwrites every weight.
        orig_reset = torch.nn.Linear.reset_parameters
        torch.nn.Linear.reset_parameters = lambda self: None
        try:
            self.chain = ref_harness.ReferenceChain(shards, even_splits(n_stages, n_layers), torch.float32, base_port=base_port,
                                                    workdir=os.path.join(self.root, "work"))
        finally:
            torch.nn.Linear.reset_parameters = orig_reset
        self.setup_load_s = time.perf_counter() - t0
        self.vocab = LLAMA2_7B["vocab_size"]

    def close(self):
        try:
            self.chain.close()
        finally:
            shutil.rmtree(self.root, ignore_errors=True)

    def _pass(self, msg):
        """One message around the ring, exactly the loop body of go_through_every_shards (node_profiler.py:1217-1230)."""
        ws = self.chain.workers
        compute_s = hop_s = 0.0
        data = msg
        with self.chain._mute():
            for i, w in enumerate(ws):
                t0 = time.perf_counter()
                out = w.pass_through_shard(data)
                t1 = time.perf_counter()
                w.communicator.transfer_data(out)
                data = ws[(i + 1) % len(ws)].communicator.receive_data()
                t2 = time.perf_counter()
                compute_s += t1 - t0
                hop_s += t2 - t1
        return data, compute_s, hop_s

    def request(self, prompt_len: int, new_tokens: int, time_budget_s: float = 1e9, seed: int = 1234):
        """One request: prompt_len-token prompt -> up to new_tokens greedy tokens.  Returns timings per pass."""
        head = self.chain.workers[0]
        g = torch.Generator().manual_seed(seed)
        ids = torch.randint(3, self.vocab, (1, prompt_len), generator=g, dtype=torch.int64)
        self.chain.clear()
        with self.chain._mute():
            msg = head.receive_user_request(input_ids=ids)
        t_start = time.perf_counter()
        tok, c, h = self._pass(msg)
        prefill = {"s": time.perf_counter() - t_start, "compute_s": c, "hop_s": h}
        steps = []
        for _ in range(new_tokens - 1):
            t0 = time.perf_counter()
            with self.chain._mute():
                reached_end, msg = head.receive_next_token(tok, 1 << 30)
            tok, c, h = self._pass(msg)
            steps.append({"s": time.perf_counter() - t0, "compute_s": c, "hop_s": h})
            if time.perf_counter() - t_start > time_budget_s:
                break
        return prefill, steps


def run_workload(n_stages: int, prompt_len: int, steps: int, warmup: int, time_budget_s: float, with_c1: bool,
                 n_layers: int = 32) -> dict:
    """What bench.py reports: decode tok/s of one request (prompt_len-token prompt, `warmup` untimed + `steps` timed decode
    passes) through an n_stages reference chain, optionally followed by BASELINE configs[0] exactly (2 workers would be
    n_stages == 2: 32-token prompt -> 16 greedy tokens)."""
    chain = TimedReferenceChain(n_stages, n_layers)
    try:
        chain.request(8, 2)  # warm-up request (the profiler warms up before timing, node_profiler.py:865-878)
        prefill, st = chain.request(prompt_len, 1 + warmup + steps, time_budget_s)
        main = summarize(prefill, st[warmup:], prompt_len, n_stages)
        out = {"main": main, "threads": chain.threads, "n_stages": n_stages, "n_layers": n_layers,
               "setup_write_s": chain.setup_write_s, "setup_load_s": chain.setup_load_s,
               "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")}
        if with_c1:
            pf, st = chain.request(32, 16, time_budget_s)
            c1 = summarize(pf, st, 32, n_stages)
            c1["total_s_32_prompt_16_tokens"] = pf["s"] + sum(x["s"] for x in st)
            out["c1"] = c1
        return out
    finally:
        chain.close()


def summarize(prefill, steps, prompt_len, n_stages):
    n = len(steps)
    tot = sum(s["s"] for s in steps)
    srt = sorted(s["s"] for s in steps)
    hops = sum(s["hop_s"] for s in steps) / max(1, n * n_stages)
    return {"decode_tok_s": n / tot if n else None, "ms_per_token": 1000 * tot / n if n else None,
            "p50_token_ms": 1000 * srt[n // 2] if n else None, "steps_timed": n,
            "prefill_s": prefill["s"], "prefill_tok_s": prompt_len / prefill["s"],
            "hop_ms": 1000 * hops, "prefill_hop_ms": 1000 * prefill["hop_s"] / n_stages}
