#!/usr/bin/env python
"""Benchmark of the shard-execution hot path (NodeWorker.pass_through_shard) on B200.

Contract:  python bench.py --gpus N --steps K --warmup W [--impl reference]
(for N > 1 the driver launches it under torch.distributed.run, one rank per GPU).

N = 1 workload (BASELINE.json configs[1]): Llama-2-7B, bf16, all 32 layers on one B200, batch 1,
512-token prompt, greedy decode with the paged KV cache.  A "step" is ONE decode token through all 32
layers + final norm + lm_head + arg-max.  `value` = decode tok/s with everything resident in HBM
(CUDA-graph replays, token fed back on the device); `e2e` = the same metric through the
reference-shaped NodeWorker API with host token buffers (H2D of the token id, D2H of the next one,
tokenizer decode, every step).  Extra keys: prefill tok/s (512-token prompt, host ids in), p50
per-token ms, the roofline of the dominant kernel, the CPU baseline (oracle port on the host cores).

N > 1: layer-sharded chain, one shard (32/N layers) per GPU, N independent batch-1 request streams
in flight (weak scaling: every GPU streams 32/N layers' weights once per stream per round), hidden
states handed over by in-kernel NVLink peer stores.
"""
import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

PROMPT_LEN = 512
NEW_TOKENS = 256
METRIC = "llama2_7b_decode_tok_s"


# ----------------------------------------------------------------------------- helpers
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu_index, self.rows, self.proc = gpu_index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def quiet():
    return contextlib.redirect_stdout(io.StringIO())


# ----------------------------------------------------------------------------- CPU baseline (oracle port)
def cpu_decode_baseline(steps, warmup, time_budget_s, prompt_len=PROMPT_LEN):
    """The oracle (CPU restatement of the reference's HF-layer shard path, fp32) on the host cores:
    Llama-2-7B dimensions, 32 layers, `prompt_len` cached tokens, batch-1 greedy decode.  Bounded
    sample: 4 distinct random-init layers' weights are cycled over the 32 layer positions (each
    layer is 809 MB fp32, far beyond the LLC, so the DRAM-streaming behaviour is the real one)."""
    from oracle import llama_oracle as lo  # allowed here: the cpu_baseline / --impl reference legs

    torch.set_num_threads(host_threads())
    torch.manual_seed(0)
    H, I, V, L, nh = 4096, 11008, 32000, 32, 32
    distinct = []
    for _ in range(4):
        distinct.append({
            "self_attn.q_proj.weight": torch.randn(H, H) * 0.02, "self_attn.k_proj.weight": torch.randn(H, H) * 0.02,
            "self_attn.v_proj.weight": torch.randn(H, H) * 0.02, "self_attn.o_proj.weight": torch.randn(H, H) * 0.02,
            "mlp.gate_proj.weight": torch.randn(I, H) * 0.02, "mlp.up_proj.weight": torch.randn(I, H) * 0.02,
            "mlp.down_proj.weight": torch.randn(H, I) * 0.02,
            "input_layernorm.weight": torch.ones(H), "post_attention_layernorm.weight": torch.ones(H)})
    lm_head = torch.randn(V, H) * 0.02
    final_norm = torch.ones(H)
    cache = lo.KVCache()

    def run(hidden):
        S = hidden.shape[1]
        pos = lo.build_position_ids(cache.get_seq_length(), S, 1)
        cos, sin = lo.rope_cos_sin(pos, 128, 10000.0, hidden.dtype)
        for li in range(L):
            hidden = lo.decoder_layer(hidden, distinct[li % 4], cos, sin, cache, li, nh, nh, 1e-5)
        hidden = lo.rms_norm(hidden, final_norm, 1e-5)
        logits = torch.nn.functional.linear(hidden, lm_head)  # the reference computes every position (node_worker.py:262)
        return torch.argmax(logits[:, -1, :], dim=-1)

    with torch.inference_mode():
        ids = torch.randint(3, V, (1, prompt_len))
        t0 = time.perf_counter()
        tok = run(torch.nn.functional.embedding(ids, lm_head))
        prefill_s = time.perf_counter() - t0
        for _ in range(warmup):
            tok = run(torch.nn.functional.embedding(tok.unsqueeze(0), lm_head))
        times = []
        t_start = time.perf_counter()
        for _ in range(steps):
            t0 = time.perf_counter()
            tok = run(torch.nn.functional.embedding(tok.unsqueeze(0), lm_head))
            times.append(time.perf_counter() - t0)
            if time.perf_counter() - t_start > time_budget_s:
                break
    n = len(times)
    total = sum(times)
    times.sort()
    return {"value": n / total, "unit": "tok/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": (f"oracle port (fp32 torch-CPU restatement of the reference's HF-layer path), Llama-2-7B dims, 32 layers "
                       f"(4 distinct random-init weight sets cycled), batch 1, {prompt_len}-token context, {n} decode steps after "
                       f"{warmup} warm-up; prefill of the {prompt_len}-token prompt took {prefill_s:.2f}s "
                       f"({prompt_len / prefill_s:.1f} tok/s)"),
            "steps_timed": n, "ms_per_step": 1000 * total / n, "p50_token_ms": 1000 * times[n // 2],
            "prefill_tok_s": prompt_len / prefill_s}


def reference_chain_baseline(n_gpus, steps, warmup, time_budget_s, with_c1):
    """The reference's OWN NodeWorker chain (its files, unmodified, behind the transformers-5.5 shim) on the host cores:
    `kind: "reference"`.  Returns None when the reference files are not on this machine (then the port is used)."""
    from oracle import ref_chain_bench as rb  # allowed here: the cpu_baseline / --impl reference legs
    if not rb.available():
        return None
    try:
        import psutil
        if psutil.virtual_memory().available < 40e9:   # 32 fp32 layers = 26 GB resident
            return None
    except ImportError:
        pass
    n_stages = max(2, n_gpus)
    r = rb.run_workload(n_stages, PROMPT_LEN, steps, warmup, time_budget_s, with_c1)
    m = r["main"]
    if not m["steps_timed"]:
        return None
    sample = (f"the reference's own NodeWorker chain (utils/node_worker.py + utils/shard_loader.py, unmodified, transformers 5.5 compat shim), "
              f"CPU fp32, Llama-2-7B architecture random init (32 layers = 4 distinct weight files + symlinks, 26 GB resident), "
              f"{n_stages} NodeWorkers in one process over loopback ZeroMQ + torch.save/load files (its Communicator), batch 1, "
              f"{PROMPT_LEN}-token prompt, {m['steps_timed']} decode passes timed after {warmup} warm-up; prefill {m['prefill_s']:.2f}s "
              f"({m['prefill_tok_s']:.1f} tok/s); hop {m['hop_ms']:.2f} ms per message; set-up {r['setup_write_s'] + r['setup_load_s']:.0f}s untimed")
    out = {"value": m["decode_tok_s"], "unit": "tok/s", "cores": r["threads"], "kind": "reference", "sample": sample,
           "steps_timed": m["steps_timed"], "ms_per_step": m["ms_per_token"], "p50_token_ms": m["p50_token_ms"],
           "prefill_tok_s": m["prefill_tok_s"], "hop_ms": m["hop_ms"], "omp_num_threads_env": r["omp_num_threads_env"],
           "n_stages": n_stages}
    if "c1" in r:
        out["c1"] = dict(r["c1"], workload="BASELINE.json configs[0] exactly: 2 NodeWorkers (layers 0-16 / 16-32) on localhost, CPU fp32, "
                                           "batch 1, 32-token prompt -> 16 greedy tokens" if n_stages == 2 else
                                           f"configs[0]'s request (32-token prompt -> 16 greedy tokens) on the {n_stages}-worker chain")
    return out


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    torch.set_num_threads(host_threads())   # torchrun exports OMP_NUM_THREADS=1: the CPU arm uses every host core it can
    W = max(1, min(args.warmup, 3))
    base = None
    try:
        base = reference_chain_baseline(args.gpus, min(args.steps, 64), W, time_budget_s=90.0, with_c1=True)
    except Exception as e:  # fall back to the port, say why
        print(f"[bench] reference chain unavailable ({e!r}); timing the oracle port instead", file=sys.stderr, flush=True)
    if base is None:
        base = cpu_decode_baseline(args.steps, W, time_budget_s=150.0, prompt_len=PROMPT_LEN)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": "tok/s", "n_gpus": args.gpus,
        "steps": base["steps_timed"], "warmup": W, "ms_per_step": base["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "prefill_tok_s": base["prefill_tok_s"], "p50_token_ms": base["p50_token_ms"], "gpu_launches": 0,
    }
    for k in ("hop_ms", "c1", "omp_num_threads_env", "n_stages"):
        if k in base:
            line[k] = base[k]
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    if n_gpus == 1:
        return {"workload": "BASELINE.json configs[1]: Llama-2-7B bf16, all 32 layers on 1 B200, batch 1, 512-token prompt, "
                            "greedy decode with KV cache; one step = one decode token",
                "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "batch": 1, "layers_per_gpu": 32,
                "attention": "non-causal prefill (reference parity, SURVEY.md F4); decode unaffected",
                "arithmetic": "model dtype bf16: weights held as fp16 (exact for bf16 values in fp16's normal range), tensor-core operands "
                              "fp16 x fp16, accumulation / residual stream / softmax fp32, K/V cache and hop messages bf16",
                "l2": "inputs larger than L2: 13.2 GB of weights streamed per step vs 126 MB L2"}
    return {"workload": f"Llama-2-7B bf16 layer-sharded {n_gpus}-way (BASELINE.json configs[2] family): {32 // n_gpus} layers per B200, "
                        f"{n_gpus} independent batch-1 request streams in flight, 512-token prompts, greedy decode; "
                        "one step = one decode token for every stream",
            "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "batch": 1, "streams": n_gpus, "layers_per_gpu": 32 // n_gpus,
            "hop": "in-kernel NVLink peer stores + flag",
            "arithmetic": "model dtype bf16: weights held as fp16 (exact for bf16 values in fp16's normal range), tensor-core operands "
                          "fp16 x fp16, accumulation / residual stream / softmax fp32, K/V cache and hop messages bf16",
            "l2": "inputs larger than L2: every GPU streams its layers' weights once per stream per step"}


# ----------------------------------------------------------------------------- N = 1
def run_single(args):
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.chain import LoopbackCommunicator
    from llm_sharding_b200.node_worker import NodeWorker

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    K, W = args.steps, max(args.warmup, 3)
    parity = None
    if not args.no_parity:   # the same public API on a model small enough for the CPU oracle, before anything is timed
        import __graft_entry__ as entry
        with quiet():
            parity = entry.smoke_check()
    cfg = synthetic.LLAMA2_7B
    tmp = tempfile.mkdtemp(prefix="b200shard_bench_")
    synthetic.write_model_dir(tmp, cfg)
    with quiet():
        worker = NodeWorker("tcp://*:48800", "tcp://127.0.0.1:48800", True, tmp, device="cuda:0", dtype=torch.bfloat16,
                            communicator=LoopbackCommunicator("tcp://*:48800", "tcp://127.0.0.1:48800"),
                            max_tokens=max(PROMPT_LEN, 0 if args.no_extras else 4 * 2048), max_batch=1 if args.no_extras else 4,
                            max_seq_len=max(PROMPT_LEN + 2 * (K + W) + 64, 0 if args.no_extras else 2048 + 64),
                            state_provider=synthetic.RandomStateProvider(cfg, "cuda:0", seed=0))
        worker.load_shards(0, cfg["num_hidden_layers"])
    eng, kv = worker.shard.engine, worker.past_key_value
    g = torch.Generator().manual_seed(1234)
    ids_host = torch.randint(3, cfg["vocab_size"], (1, PROMPT_LEN), generator=g, dtype=torch.int64).pin_memory()
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # ---- prefill: host ids in -> first token (through the NodeWorker API)
    prefill_ms = []
    for it in range(2 + 3):
        with quiet():
            worker.clear_KV_cache()
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        with quiet():
            msg = worker.receive_user_request(input_ids=ids_host)
            tok = worker.pass_through_shard(msg)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            prefill_ms.append(e0.elapsed_time(e1))
    prefill_ms.sort()
    prefill_med_ms = prefill_ms[len(prefill_ms) // 2]
    prefill_tok_s = PROMPT_LEN / (prefill_med_ms / 1000)

    def prefill_flops(B, S):
        # 2 x params per token per layer + 4 S T H per sequence per layer (non-causal), lm_head on the last row only
        H, I, L, V = cfg["hidden_size"], cfg["intermediate_size"], cfg["num_hidden_layers"], cfg["vocab_size"]
        return B * (S * L * 2 * (4 * H * H + 3 * H * I) + L * 4 * S * S * H + 2 * H * V)

    pk0 = peaks()
    roofline_prefill = {"bound": "tensor", "workload": f"1 x {PROMPT_LEN}-token prompt through NodeWorker (host ids in, first token out)",
                        "flops": prefill_flops(1, PROMPT_LEN), "ms": prefill_med_ms,
                        "achieved": prefill_flops(1, PROMPT_LEN) / (prefill_med_ms / 1000) / 1e12, "peak": pk0["bf16_tflops"],
                        "unit": "TFLOP/s", "frac": prefill_flops(1, PROMPT_LEN) / (prefill_med_ms / 1000) / 1e12 / pk0["bf16_tflops"],
                        "peak_source": pk0["source"], "peak_note": "burst cuBLAS bf16 figure: an 8 ms call runs at boost clocks"}
    prefill_4x2048 = None
    if not args.no_extras:   # the tensor-bound regime proper: 4 x 2048 tokens in one call
        ids4 = torch.randint(3, cfg["vocab_size"], (4, 2048), generator=g, dtype=torch.int64).pin_memory()
        ms4 = []
        for it in range(1 + 3):
            with quiet():
                worker.clear_KV_cache()
            torch.cuda.synchronize()
            e0, e1 = ev(), ev()
            e0.record()
            with quiet():
                tok4 = worker.pass_through_shard(worker.receive_user_request(input_ids=ids4))
            e1.record()
            torch.cuda.synchronize()
            if it >= 1:
                ms4.append(e0.elapsed_time(e1))
        ms4.sort()
        f4 = prefill_flops(4, 2048)
        prefill_4x2048 = {"bound": "tensor", "workload": "4 x 2048-token prompts in one call through NodeWorker", "flops": f4,
                          "ms": ms4[1], "tok_s": 4 * 2048 * 1000.0 / ms4[1], "achieved": f4 / (ms4[1] / 1000) / 1e12,
                          "peak": pk0["bf16_tflops"], "unit": "TFLOP/s", "frac": f4 / (ms4[1] / 1000) / 1e12 / pk0["bf16_tflops"],
                          "peak_sustained": pk0["bf16_tflops_sustained"],
                          "frac_of_sustained": f4 / (ms4[1] / 1000) / 1e12 / pk0["bf16_tflops_sustained"],
                          "peak_note": "a 0.1 s tensor-bound call runs power-capped: the sustained cuBLAS figure is the fair denominator"}
        with quiet():
            worker.clear_KV_cache()
            msg = worker.receive_user_request(input_ids=ids_host)
            tok = worker.pass_through_shard(msg)

    # ---- decode, device resident: token buffer fed back on the GPU, one CUDA-graph replay per step
    tokbuf = tok.clone()
    for _ in range(W):
        eng.step_tokens(tokbuf, kv, [0], tok_out=tokbuf)
    torch.cuda.synchronize()
    eng.reset_counters()
    ctx0 = kv.get_seq_length(0)
    with ClockSampler(0) as clk:
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(K):
            eng.step_tokens(tokbuf, kv, [0], tok_out=tokbuf)
        e1.record()
        torch.cuda.synchronize()
    total_ms = e0.elapsed_time(e1)
    cnt = eng.counters()
    ms_per_step = total_ms / K
    value = 1000.0 / ms_per_step
    bytes_per_step = cnt["bytes"] / max(1, cnt["calls"])

    # ---- e2e: the reference's loop through the NodeWorker API with host token buffers
    with quiet():
        worker.clear_KV_cache()
        msg = worker.receive_user_request(input_ids=ids_host)
        tok = worker.pass_through_shard(msg)
    tok_host = tok.cpu().pin_memory()
    for _ in range(W):
        with quiet():
            _, info = worker.receive_next_token(tok_host, 1 << 30)
            tok_host = worker.pass_through_shard(info).cpu()
    per_step = []
    torch.cuda.synchronize()
    t_all0 = time.perf_counter()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(K):
        t0 = time.perf_counter()
        with quiet():
            _, info = worker.receive_next_token(tok_host, 1 << 30)  # H2D of the id, embed on the GPU, tokenizer decode
            tok_host = worker.pass_through_shard(info).cpu()          # 32 layers + lm_head + argmax, D2H of the id
        per_step.append(time.perf_counter() - t0)
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1)
    wall_ms = (time.perf_counter() - t_all0) * 1000
    per_step.sort()
    e2e = {"value": 1000.0 * K / max(e2e_ms, wall_ms), "unit": "tok/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 8,
           "api": "NodeWorker.receive_next_token + NodeWorker.pass_through_shard, host int64 token in / out per step"}

    # ---- roofline of the dominant kernel (fused RMSNorm + gate/up GEMV + SwiGLU), timed alone on its stream,
    #      cycling over the 32 layers' weights (5.8 GB >> L2)
    H, I = cfg["hidden_size"], cfg["intermediate_size"]
    L = cfg["num_hidden_layers"]
    for l in range(L):
        eng.run_kernel(kv, 3, l, 1)
    torch.cuda.synchronize()
    reps = 4
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(reps):
        for l in range(L):
            eng.run_kernel(kv, 3, l, 1)
    e1.record()
    torch.cuda.synchronize()
    k_ms = e0.elapsed_time(e1) / (reps * L)
    # per-kernel breakdown of a decode step: each product kernel timed the same way (warm, back to back over the layers)
    kernel_us = {}
    for name, kind in (("qkv_rope_append", 0), ("attention", 1), ("o_proj", 2), ("gate_up_swiglu", 3), ("down_proj", 4), ("lm_head", 5)):
        n_l = 1 if kind == 5 else L
        for l in range(n_l):
            eng.run_kernel(kv, kind, l, 1)
        torch.cuda.synchronize()
        a0, a1 = ev(), ev()
        a0.record()
        for _ in range(reps * (8 if kind == 5 else 1)):
            for l in range(n_l):
                eng.run_kernel(kv, kind, l, 1)
        a1.record()
        torch.cuda.synchronize()
        kernel_us[name] = 1000 * a0.elapsed_time(a1) / (reps * (8 if kind == 5 else 1) * n_l)
    k_bytes = 2 * I * H * 2 + H * 2 + H * 4 + I * 4
    pk = peaks()
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(tpath):
        traffic = json.load(open(tpath)).get("gemv_gateup_dram_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "gemv_kernel<1,2,EPI_SWIGLU,NORM> (RMSNorm + gate/up + SwiGLU, 180 MB of weights)",
                "achieved": k_bytes / (k_ms / 1000) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": k_bytes / (k_ms / 1000) / 1e9 / pk["hbm_gbs"], "traffic": traffic, "peak_source": pk["source"],
                "bytes_per_launch": k_bytes, "us_per_launch": k_ms * 1000}
    step_roofline = {"bound": "hbm", "bytes_per_step": bytes_per_step, "achieved": bytes_per_step / (ms_per_step / 1000) / 1e9,
                     "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": bytes_per_step / (ms_per_step / 1000) / 1e9 / pk["hbm_gbs"],
                     "note": "whole decode step (all kernels, launch gaps included) vs algorithmic bytes of SURVEY.md 8(d)"}

    cpu = None
    if not args.no_cpu_baseline:
        try:
            cpu = None
            try:
                cpu = reference_chain_baseline(1, steps=16, warmup=1, time_budget_s=25.0, with_c1=False)
            except Exception as e:
                print(f"[bench] reference chain unavailable ({e!r}); timing the oracle port instead", file=sys.stderr, flush=True)
            if cpu is None:
                cpu = cpu_decode_baseline(steps=16, warmup=1, time_budget_s=25.0, prompt_len=PROMPT_LEN)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:  # the baseline is reported, never required
            cpu = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {e!r}"}

    line = {
        "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": dict(workload_config(1), context_at_start=ctx0),
        "clocks": clk.summary(), "e2e": e2e, "gpu_launches": cnt["launches"],
        "roofline": roofline, "step_roofline": step_roofline, "cpu_baseline": cpu,
        "prefill_tok_s": prefill_tok_s, "p50_token_ms": 1000 * per_step[len(per_step) // 2],
        "p50_token_ms_device": ms_per_step, "kernel_us": kernel_us,
        "roofline_prefill": roofline_prefill, "prefill_4x2048": prefill_4x2048, "parity": parity,
    }
    print(json.dumps(line), flush=True)


def main():
    global PROMPT_LEN
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=NEW_TOKENS)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--prompt-len", type=int, default=PROMPT_LEN, help="prompt tokens per request (default: BASELINE configs[1]'s 512)")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"], help="hop transport for --gpus > 1")
    ap.add_argument("--streams", type=int, default=0, help="--gpus > 1: request streams in flight (default: one per GPU)")
    ap.add_argument("--rows", type=int, default=1, help="--gpus > 1: batch rows per stream; 8 streams x 4 rows = BASELINE configs[2]'s batch 32")
    ap.add_argument("--splits", default="", help="--gpus > 1: layers per stage, e.g. 4,12,16 (BASELINE configs[4]); default: even")
    ap.add_argument("--mixed", action="store_true", help="--gpus > 1: BASELINE configs[4]'s mixed stream: prompts of 32/512/2048 tokens, "
                                                         "streams are retired and re-prefilled while the others decode")
    ap.add_argument("--retire-every", type=int, default=8, help="--mixed: rounds between two retirements")
    ap.add_argument("--no-parity", action="store_true", help="--gpus > 1: skip the oracle parity check of the ring that precedes the timed region")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra blocks (c3_batch32 / c4_prefill at --gpus > 1, prefill_4x2048 at 1)")
    args = ap.parse_args()
    PROMPT_LEN = args.prompt_len
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the shard path has no CPU fallback (use --impl reference for the CPU baseline)")
    if args.gpus == 1 and not args.mixed:
        return run_single(args)
    from bench_chain import run_chain  # multi-GPU arm
    return run_chain(args, workload_config(args.gpus), METRIC, ClockSampler, peaks)


if __name__ == "__main__":
    main()
