"""``NodeProfiler`` — the reference's acceptance / capacity harness (utils/node_profiler.py:13-1331) on the B200
worker: same entry points and timing protocol (warm-up with the longest and the shortest prompt, ``perf_counter``
bracketed by device synchronisation, ``PROFILE_REPEAT_NUM`` repeats per prompt length, cumulative per-token decode
latency, least-squares linear and quadratic latency models with RMSE / R^2).  What it leaves out: the matplotlib plots
(:154-194) and the two-device assisted mode over ZeroMQ (:478-574, :981-1136) — on one NVSwitch box the in-process ring
(``go_through_every_shards``) and ``bench.py --gpus N`` cover that ground.

Results are returned as dictionaries (the reference prints them and writes PNGs under results/profiling/)."""
from __future__ import annotations

import time

import torch

from .chain import LoopbackCommunicator
from .engine import read_model_config
from .node_worker import NodeWorker


class NodeProfiler:
    PROFILE_INTERVAL_SLEEP_TIME = 0          # the reference sleeps 1 s between measurements (thermal settling on edge devices)
    PROFILE_REPEAT_NUM = 3
    PROFILE_PREFILL_INPUT_TOKEN_LENGTHS = [8, 16, 32, 64, 128, 256, 512]
    PROFILE_DECODE_OUTPUT_TOKEN_LENGTHS = [8, 16, 32, 64, 128, 256, 512]

    def __init__(self, shards_path: str, device="cuda:0", dtype=torch.bfloat16, state_provider=None):
        self.shards_path = shards_path
        self.device = torch.device(device)
        self.dtype = dtype
        self.config = read_model_config(shards_path)
        self.layer_num = self.config["num_hidden_layers"]
        self.state_provider = state_provider
        self._port = 40800

    # ------------------------------------------------------------------ helpers
    def _new_worker(self, can_receive_user_request=True, max_tokens=None, **kw) -> NodeWorker:
        src, dst = f"tcp://*:{self._port}", f"tcp://127.0.0.1:{self._port}"
        self._port += 1
        longest = max(self.PROFILE_PREFILL_INPUT_TOKEN_LENGTHS)
        return NodeWorker(src_addr=src, dst_addr=dst, can_receive_user_request=can_receive_user_request,
                          shards_path=self.shards_path, device=self.device, dtype=self.dtype,
                          communicator=LoopbackCommunicator(src, dst), max_tokens=max_tokens or longest,
                          max_seq_len=longest + max(self.PROFILE_DECODE_OUTPUT_TOKEN_LENGTHS) + 64,
                          state_provider=self.state_provider, **kw)

    def _synchronize_device(self) -> None:
        torch.cuda.synchronize(self.device)

    def _build_profile_input_ids(self, tokenizer=None) -> tuple[list[int], list[torch.Tensor]]:
        """One [1, S] id tensor per profiled prompt length (the reference repeats a text fragment and tokenizes it,
        :346-366; ids are what reaches the worker either way, through receive_user_request(input_ids=...))."""
        g = torch.Generator().manual_seed(1234)
        lengths = [s for s in self.PROFILE_PREFILL_INPUT_TOKEN_LENGTHS if s <= self.config["max_position_embeddings"]]
        return lengths, [torch.randint(3, self.config["vocab_size"], (1, s), generator=g, dtype=torch.int64) for s in lengths]

    @staticmethod
    def _fit_latency_models(token_lengths: list[int], latencies: list[float]) -> dict:
        """T(S) = a S + b and T(S) = a S^2 + b S + c by least squares, with RMSE and R^2 (reference :64-150)."""
        if len(token_lengths) != len(latencies):
            raise ValueError("[ERROR] token_lengths length must be equal to latencies length.")
        if len(token_lengths) < 3:
            raise ValueError("[ERROR] at least 3 points are needed for quadratic fitting.")
        s = torch.tensor(token_lengths, dtype=torch.float64)
        t = torch.tensor(latencies, dtype=torch.float64)
        out = {}
        for name, design in (("linear", torch.stack((s, torch.ones_like(s)), 1)),
                             ("quadratic", torch.stack((s * s, s, torch.ones_like(s)), 1))):
            coef = torch.linalg.lstsq(design, t.unsqueeze(1)).solution.squeeze(1)
            resid = t - design @ coef
            sst = torch.sum((t - t.mean()) ** 2)
            out[name] = {"coefficients": coef.tolist(), "rmse": float(torch.sqrt(torch.mean(resid ** 2))),
                         "r_squared": float("nan") if float(sst) == 0 else float(1 - torch.sum(resid ** 2) / sst)}
        return out

    # ------------------------------------------------------------------ public routines
    def profile_max_layer_num(self) -> int:
        """How many layers (plus the embedding) fit on this device: load 1, 2, ... until CUDA runs out (:46-62)."""
        node = self._new_worker(max_tokens=64)
        max_layer_num = 0
        try:
            for i in range(self.layer_num):
                node.load_shards(0, i + 1)
                max_layer_num = i + 1
        except torch.cuda.OutOfMemoryError:
            print(f"[INFO] max layer num: {max_layer_num}")
        self._release(node)
        return max_layer_num

    def profile_compute_capability(self, max_layer_num: int | None = None, decode_tokens: int | None = None) -> dict:
        """Prefill latency per prompt length and cumulative decode latency per generated token on the layers this node
        can hold, normalised to the whole model (x layer_num / loaded layers, :376-377), plus the fitted models."""
        if max_layer_num is None or max_layer_num < 0:
            max_layer_num = self.layer_num
        loaded = min(max_layer_num, self.layer_num)
        node = self._new_worker()
        node.load_shards(0, loaded)
        lengths, ids = self._build_profile_input_ids(node.tokenizer)

        def one_prefill(x):
            d0 = node.receive_user_request(input_ids=x)
            return node.pass_through_shard(d0)

        for x in (ids[-1], ids[0]):  # warm-up: longest, then shortest prompt (:865-878)
            one_prefill(x)
            self._synchronize_device()
            node.clear_KV_cache()
        repeated = []
        for x in ids:
            cur = []
            for _ in range(self.PROFILE_REPEAT_NUM):
                self._synchronize_device()
                t0 = time.perf_counter()
                one_prefill(x)
                self._synchronize_device()
                cur.append(time.perf_counter() - t0)
                node.clear_KV_cache()
                if self.PROFILE_INTERVAL_SLEEP_TIME:
                    time.sleep(self.PROFILE_INTERVAL_SLEEP_TIME)
            repeated.append(cur)
        scale = self.layer_num / loaded
        prefill_lat = [scale * sum(c) / len(c) for c in repeated]

        # decode: cumulative latency after each generated token (:946-960).  The node holds layer 0 but usually not
        # the lm_head, so the token fed back is a fixed id: the cost of a step does not depend on which token it is.
        n_dec = decode_tokens or max(self.PROFILE_DECODE_OUTPUT_TOKEN_LENGTHS)
        step_ids = torch.full((1, 1), 7, dtype=torch.int64)
        one_prefill(ids[0])
        for _ in range(3):
            node.pass_through_shard({"hidden_states": node.embed_tokens(step_ids), "batch_size": 1, "seq_len": 1})
        self._synchronize_device()
        cumulative, t0 = [], time.perf_counter()
        for _ in range(n_dec):
            node.pass_through_shard({"hidden_states": node.embed_tokens(step_ids), "batch_size": 1, "seq_len": 1})
            self._synchronize_device()
            cumulative.append(scale * (time.perf_counter() - t0))
        node.clear_KV_cache()
        self._release(node)
        result = {
            "loaded_layer_num": loaded, "layer_num": self.layer_num,
            "prefill": {"token_lengths": lengths, "latencies_s": prefill_lat, "repeated_s": repeated,
                        "fit": self._fit_latency_models(lengths, prefill_lat) if len(lengths) >= 3 else None},
            "decode": {"token_counts": list(range(1, n_dec + 1)), "cumulative_latencies_s": cumulative,
                       "fit": self._fit_latency_models(list(range(1, n_dec + 1)), cumulative) if n_dec >= 3 else None},
        }
        result["prefill"]["tokens_per_s"] = [s / t for s, t in zip(lengths, prefill_lat)]
        result["decode"]["tokens_per_s"] = n_dec / cumulative[-1]
        return result

    def profile_cold_start_latency(self, max_layer_num: int | None = None) -> dict:
        """Seconds from nothing to a loaded, ready shard (:1138-1172)."""
        loaded = min(max_layer_num or self.layer_num, self.layer_num)
        self._synchronize_device()
        t0 = time.perf_counter()
        node = self._new_worker(max_tokens=64)
        node.load_shards(0, loaded)
        self._synchronize_device()
        dt = time.perf_counter() - t0
        self._release(node)
        return {"loaded_layer_num": loaded, "cold_start_s": dt}

    def go_through_every_shards(self, out_token_num: int = 50, splits: list[tuple[int, int]] | None = None,
                                input_ids: torch.Tensor | None = None) -> torch.Tensor:
        """The reference's in-process ring (:1174-1236): one NodeWorker per layer range (default: four equal ranges),
        request -> every shard -> next token -> back to the head, `out_token_num` tokens.  Returns the token ids."""
        from .chain import generate, layer_range
        if splits is None:
            n = min(4, self.layer_num)
            splits = [layer_range(r, n, self.layer_num) for r in range(n)]
        workers = []
        for i, (s, e) in enumerate(splits):
            w = self._new_worker(can_receive_user_request=(i == 0))
            w.load_shards(s, e)
            workers.append(w)
        if input_ids is None:
            input_ids = self._build_profile_input_ids()[1][0]
        tokens, _ = generate(workers, input_ids, out_token_num)
        for w in workers:
            self._release(w)
        return tokens

    @staticmethod
    def _release(node: NodeWorker) -> None:
        node.past_key_value = None
        if node.shard is not None:
            node.shard.engine.close()
        node.communicator.close()
