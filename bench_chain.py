"""Multi-GPU arm of bench.py: Llama-2-7B layer-sharded over N B200s of one node, one process per GPU
(torch.distributed.run), one shard per process, N independent batch-1 request streams in flight so that every
stage stays busy (weak scaling: each GPU streams its 32/N layers' weights once per stream per round).  Hidden
states and token ids travel by in-kernel NVLink peer stores + flags (llm-sharding_b200/csrc/hop.cu); `--transport
nccl` switches to ncclSend/ncclRecv for the A/B comparison."""
import contextlib
import io
import json
import os
import tempfile
import time

import torch
import torch.distributed as dist

PROMPT_LEN = 512


def _quiet():
    return contextlib.redirect_stdout(io.StringIO())


def hop_pingpong(rank, world, device, ctrl, n=200, payload=8192):
    """One-way latency of the hop between GPU 0 and GPU 1: kernel store over NVLink + flag release on one side,
    flag acquire on the other, `n` round trips enqueued back to back; rank 0 times them with CUDA events."""
    from llm_sharding_b200.chain import P2PHop
    hop = P2PHop(device, 4, payload)
    gathered = [None] * world
    dist.all_gather_object(gathered, (hop.export(), hop.dev_index), group=ctrl)
    result = None
    if rank in (0, 1):
        peer = 1 - rank
        hop.connect(gathered[peer][0], gathered[peer][1], same_process=False)
    dist.barrier(group=ctrl)
    if rank in (0, 1):
        buf = torch.zeros(payload, dtype=torch.uint8, device=device)
        out = torch.empty_like(buf)
        for phase in range(2):  # warm-up, then timed
            torch.cuda.synchronize()
            base = phase * n
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(n):
                seq = base + i
                if rank == 0:
                    hop.send(buf, seq)
                    hop.recv(out, seq, max_wait_us=5_000_000)
                else:
                    hop.recv(out, seq, max_wait_us=5_000_000)
                    hop.send(buf, seq)
            e1.record()
            torch.cuda.synchronize()
            if phase == 1 and rank == 0:
                result = {"one_way_us": e0.elapsed_time(e1) * 1000 / (2 * n), "payload_bytes": payload, "round_trips": n,
                          "error_flag": hop.error()}
    dist.barrier(group=ctrl)
    hop.close()
    return result


def run_chain(args, config, metric, ClockSampler, peaks):
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.chain import ChainStage, LoopbackCommunicator, layer_range
    from llm_sharding_b200.node_worker import NodeWorker

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        raise SystemExit(f"bench.py --gpus {args.gpus} must run under torch.distributed.run with {args.gpus} ranks (got {world})")
    torch.cuda.set_device(local)
    device = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=device)
    ctrl = dist.new_group(backend="gloo")
    global PROMPT_LEN
    PROMPT_LEN = getattr(args, "prompt_len", PROMPT_LEN)
    K, W = args.steps, max(args.warmup, 3)
    transport = getattr(args, "transport", "p2p")
    cfg = synthetic.LLAMA2_7B
    L = cfg["num_hidden_layers"]
    rows = max(1, getattr(args, "rows", 1))          # batch rows per request stream (micro-batch)
    n_streams = getattr(args, "streams", 0) or world
    start, end = layer_range(rank, world, L)
    splits = [int(v) for v in getattr(args, "splits", "").split(",") if v]
    if splits:
        if len(splits) != world or sum(splits) != L or min(splits) < 1:
            raise SystemExit(f"--splits needs {world} positive layer counts that add up to {L}")
        start = sum(splits[:rank])
        end = start + splits[rank]
    mixed = bool(getattr(args, "mixed", False))
    if mixed:
        return run_mixed(args, config, rank, world, local, device, ctrl, start, end, ClockSampler, peaks)
    tmp = tempfile.mkdtemp(prefix=f"b200shard_bench_r{rank}_")
    synthetic.write_model_dir(tmp, cfg)
    with _quiet():
        worker = NodeWorker(f"tcp://*:{48900 + rank}", f"tcp://127.0.0.1:{48900 + (rank + 1) % world}", rank == 0, tmp,
                            device=str(device), dtype=torch.bfloat16,
                            communicator=LoopbackCommunicator(f"tcp://*:{48900 + rank}", f"tcp://127.0.0.1:{48900 + (rank + 1) % world}"),
                            max_tokens=rows * PROMPT_LEN, max_batch=rows, max_slots=n_streams * rows,
                            max_seq_len=PROMPT_LEN + 2 * (K + W) + 64,
                            state_provider=synthetic.RandomStateProvider(cfg, str(device), seed=0))
        worker.load_shards(start, end)
    hop_lat = None
    try:
        hop_lat = hop_pingpong(rank, world, device, ctrl)
    except Exception as e:  # reported, not required
        hop_lat = {"error": repr(e)} if rank == 0 else None
    try:
        stage = ChainStage(worker, rank, world, n_streams, PROMPT_LEN, ctrl_group=ctrl, transport=transport, rows=rows)
    except Exception as e:
        if transport == "p2p":
            if rank == 0:
                print(f"[bench] P2P mailbox set-up failed ({e!r}); falling back to NCCL send/recv", flush=True)
            transport = "nccl"
            stage = ChainStage(worker, rank, world, n_streams, PROMPT_LEN, ctrl_group=ctrl, transport="nccl", rows=rows)
        else:
            raise
    eng = worker.shard.engine
    g = torch.Generator().manual_seed(1234)
    prompts = [torch.randint(3, cfg["vocab_size"], (rows * PROMPT_LEN,), generator=g, dtype=torch.int64) for _ in range(n_streams)]

    def prefill_all():
        for i in range(n_streams):
            stage.prefill(i, PROMPT_LEN, prompts[i].to(device) if rank == 0 else None)

    def clear_all():
        torch.cuda.synchronize()
        dist.barrier(group=ctrl)
        worker.past_key_value.reset(-1)
        stage.sched.clear()
        torch.cuda.synchronize()
        dist.barrier(group=ctrl)

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def timed(fn):
        torch.cuda.synchronize()
        dist.barrier(group=ctrl)
        e0, e1 = ev(), ev()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ctrl)
        return float(t[0])

    # ---- prefill throughput (all streams' prompts through the whole ring)
    prefill_all()
    clear_all()
    prefill_ms = timed(prefill_all)
    prefill_tok_s = n_streams * rows * PROMPT_LEN / (prefill_ms / 1000)

    # ---- decode, device resident
    for _ in range(W):
        for i in range(n_streams):
            stage.decode(i)
    torch.cuda.synchronize()
    eng.reset_counters()
    with ClockSampler(local) as clk:
        total_ms = timed(lambda: [stage.decode(i) for _ in range(K) for i in range(n_streams)])
    cnt = eng.counters()
    ms_per_step = total_ms / K
    value = n_streams * rows * 1000.0 / ms_per_step
    launches = torch.tensor([cnt["launches"]], dtype=torch.float64)
    bytes_all = torch.tensor([cnt["bytes"]], dtype=torch.float64)
    dist.all_reduce(launches, group=ctrl)
    bmax = bytes_all.clone()
    dist.all_reduce(bmax, op=dist.ReduceOp.MAX, group=ctrl)

    # ---- end to end: every token id makes a round trip through the head's host memory
    clear_all()
    prefill_all()
    for _ in range(W):
        for i in range(n_streams):
            stage.decode(i, via_host=True)
    t0 = time.perf_counter()
    e2e_ms = timed(lambda: [stage.decode(i, via_host=True) for _ in range(K) for i in range(n_streams)])
    e2e_val = n_streams * rows * K * 1000.0 / e2e_ms
    err = torch.tensor([stage.error()], dtype=torch.float64)
    dist.all_reduce(err, op=dist.ReduceOp.MAX, group=ctrl)

    if rank == 0:
        pk = peaks()
        per_gpu_bytes = float(bmax[0]) / max(1, cnt["calls"])  # algorithmic bytes of one stage call (largest stage)
        stage_ms = ms_per_step / n_streams                      # steady state: one stage call per stream per step
        line = {
            "metric": metric, "value": value, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": dict(config, transport=transport, layers_this_rank=[start, end], streams=n_streams, rows_per_stream=rows,
                           batch_in_flight=n_streams * rows),
            "clocks": clk.summary(),
            "e2e": {"value": e2e_val, "unit": "tok/s", "h2d_bytes_per_step": 8 * n_streams * rows, "d2h_bytes_per_step": 8 * n_streams * rows,
                    "api": "ChainStage.decode(via_host=True): every token id D2H to the head's pinned host memory and H2D back"},
            "gpu_launches": int(launches[0]),
            "roofline": {"bound": "hbm", "kernel": "one shard call (all kernels of 32/N layers, CUDA-graph replay) on the busiest stage",
                         "achieved": per_gpu_bytes / (stage_ms / 1000) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": per_gpu_bytes / (stage_ms / 1000) / 1e9 / pk["hbm_gbs"], "traffic": None, "peak_source": pk["source"]},
            "hop": hop_lat, "hop_error_flag": int(err[0]),
            "prefill_tok_s": prefill_tok_s, "p50_token_ms": ms_per_step,
        }
        print(json.dumps(line), flush=True)
    dist.barrier(group=ctrl)
    stage.close()
    dist.destroy_process_group()


MIX_LENGTHS = (32, 512, 2048)


def run_mixed(args, config, rank, world, local, device, ctrl, start, end, ClockSampler, peaks):
    """BASELINE configs[4] / SURVEY.md 8(d) C5: uneven layer split, prompts of mixed length, prefill of new requests
    interleaved with the ongoing decodes.  Streams are served round-robin; every `--retire-every` rounds the next stream
    in rotation is retired at its turn (ChainStage.finish) and takes a fresh prompt of its length straight away."""
    from llm_sharding_b200 import synthetic
    from llm_sharding_b200.chain import ChainStage, LoopbackCommunicator
    from llm_sharding_b200.node_worker import NodeWorker

    K, W = args.steps, max(args.warmup, 3)
    cfg = synthetic.LLAMA2_7B
    rows = max(1, args.rows)
    n_streams = args.streams or 2 * world
    every = max(1, args.retire_every)
    lens = [MIX_LENGTHS[i % len(MIX_LENGTHS)] for i in range(n_streams)]
    max_prompt = max(lens)
    tmp = tempfile.mkdtemp(prefix=f"b200shard_bench_r{rank}_")
    synthetic.write_model_dir(tmp, cfg)
    with _quiet():
        worker = NodeWorker(f"tcp://*:{48900 + rank}", f"tcp://127.0.0.1:{48900 + (rank + 1) % world}", rank == 0, tmp,
                            device=str(device), dtype=torch.bfloat16,
                            communicator=LoopbackCommunicator(f"tcp://*:{48900 + rank}", f"tcp://127.0.0.1:{48900 + (rank + 1) % world}"),
                            max_tokens=rows * max_prompt, max_batch=rows, max_slots=n_streams * rows,
                            max_seq_len=max_prompt + K + W + 64,
                            state_provider=synthetic.RandomStateProvider(cfg, str(device), seed=0))
        worker.load_shards(start, end)
    stage = ChainStage(worker, rank, world, n_streams, max_prompt, ctrl_group=ctrl, transport=args.transport, rows=rows)
    g = torch.Generator().manual_seed(1234)
    n_retire = K // every
    def new_prompt(i):
        t = torch.randint(3, cfg["vocab_size"], (rows * lens[i],), generator=g, dtype=torch.int64)
        return t.to(device) if rank == 0 else None
    first = [new_prompt(i) for i in range(n_streams)]
    later = [new_prompt((j % n_streams)) for j in range(n_retire)]   # staged on the head's GPU before the clock starts

    for i in range(n_streams):
        stage.prefill(i, lens[i], first[i])
    for _ in range(W):
        for i in range(n_streams):
            stage.decode(i)
    torch.cuda.synchronize()
    dist.barrier(group=ctrl)
    eng = worker.shard.engine
    eng.reset_counters()
    counts = {"gen": 0, "prompt": 0, "retired": 0}

    def body():
        for r in range(K):
            victim = (r // every) % n_streams if (r % every == every - 1 and r // every < n_retire) else -1
            for i in range(n_streams):
                if i == victim:
                    stage.finish(i)
                    stage.prefill(i, lens[i], later[r // every])
                    counts["prompt"] += rows * lens[i]
                    counts["retired"] += 1
                else:
                    stage.decode(i)
                counts["gen"] += rows   # a decode yields a token, and so does the prefill that replaces it

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        body()
        e1.record()
        torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ctrl)
    total_ms = float(t[0])
    cnt = eng.counters()
    launches = torch.tensor([cnt["launches"]], dtype=torch.float64)
    dist.all_reduce(launches, group=ctrl)
    err = torch.tensor([stage.error()], dtype=torch.float64)
    dist.all_reduce(err, op=dist.ReduceOp.MAX, group=ctrl)
    layers = [None] * world
    dist.all_gather_object(layers, (start, end), group=ctrl)
    if rank == 0:
        line = {
            "metric": "llama2_7b_mixed_stream_tok_s", "value": counts["gen"] * 1000.0 / total_ms, "unit": "tok/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[4]: Llama-2-7B bf16, uneven layer split, embedding holder = entry stage, "
                                   "request streams with 32/512/2048-token prompts; streams are retired and re-prefilled in "
                                   "rotation while the others decode; one step = one round over all streams",
                       "layers_per_stage": layers, "streams": n_streams, "rows_per_stream": rows, "prompt_lengths": lens,
                       "retire_every_rounds": every, "transport": args.transport,
                       "l2": "inputs larger than L2: every stage streams its layers' weights once per stream per round"},
            "clocks": clk.summary(), "gpu_launches": int(launches[0]),
            "generated_tokens": counts["gen"], "prompt_tokens_in_timed_region": counts["prompt"], "requests_retired": counts["retired"],
            "prompt_tok_s": counts["prompt"] * 1000.0 / total_ms, "hop_error_flag": int(err[0]),
        }
        print(json.dumps(line), flush=True)
    dist.barrier(group=ctrl)
    stage.close()
    dist.destroy_process_group()
