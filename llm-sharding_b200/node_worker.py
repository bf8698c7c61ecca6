"""``Communicator`` / ``NodeWorker`` / ``NodeController`` with the reference's API
(utils/node_worker.py:13-559): same constructor arguments, the same three message kinds
(``input_token_info`` = {hidden_states, batch_size, seq_len}; ``next_state_info`` = {hidden_states,
cos, sin}; next-token ``Tensor``), the same placement rules (embedding + tokenizer on nodes that
``can_receive_user_request``, RoPE on ``start == 0``, final norm + lm_head on ``end ==
layer_num``) and the same exceptions.  The layer stack underneath is ``libb200shard.so``.

Differences, all additive:
* ``next_state_info`` keeps the ``cos`` / ``sin`` keys (so ``is_next_state_info`` is unchanged)
  but their values are ``None``: every stage recomputes RoPE on its GPU from ``pos_offset``,
  which travels instead (8 bytes, not 2 x [B,S,128]).
* messages may carry ``slot`` (KV slot ids, one per row) so that several micro-batches can be in
  flight; absent = slots 0..B-1, the reference's single-request behaviour.
* ``receive_next_token`` also works for B > 1 (the reference's is batch-1 only).
* ``Communicator`` serialises in memory instead of through ``results/*.pt`` files.
"""
from __future__ import annotations

import gc
import io
import os

import torch
import zmq

from .engine import PagedKVCache
from .forwarding_utils import build_position_ids  # noqa: F401  (re-exported like the reference module)
from .shard_loader import LlamaShardPart


class Communicator:
    """ZeroMQ PULL (bind ``src_addr``) / PUSH (connect ``dst_addr``) pair, utils/node_worker.py:13-67.
    Used for the control messages and for hops that leave the NVSwitch box; GPU-to-GPU hops inside
    the box go through ``P2PHop`` (chain.py)."""

    def __init__(self, src_addr: str, dst_addr: str):
        self.src_addr = src_addr
        self._rctx = zmq.Context()
        self.recv_socket = self._rctx.socket(zmq.PULL)
        self.recv_socket.bind(self.src_addr)
        self.actual_src_addr = self.recv_socket.getsockopt_string(zmq.LAST_ENDPOINT)
        self.dst_addr = dst_addr
        self._sctx = zmq.Context()
        self.send_socket = self._sctx.socket(zmq.PUSH)
        self.send_socket.connect(self.dst_addr)

    def change_src_addr(self, new_src_addr: str) -> str:
        self.recv_socket.unbind(self.actual_src_addr)
        self.recv_socket.bind(new_src_addr)
        self.src_addr = new_src_addr
        self.actual_src_addr = self.recv_socket.getsockopt_string(zmq.LAST_ENDPOINT)
        return self.src_addr

    def change_dst_addr(self, new_dst_addr: str) -> str:
        self.send_socket.disconnect(self.dst_addr)
        self.send_socket.connect(new_dst_addr)
        self.dst_addr = new_dst_addr
        return self.dst_addr

    @staticmethod
    def _to_wire(data):
        if isinstance(data, torch.Tensor):
            return data.detach().cpu()
        if isinstance(data, dict):
            return {k: Communicator._to_wire(v) for k, v in data.items()}
        return data

    def transfer_data(self, data, data_path: str = "results/send_data.pt", keep_data: bool = False):
        buf = io.BytesIO()
        torch.save(self._to_wire(data), buf)
        payload = buf.getvalue()
        self.send_socket.send(payload)
        if keep_data:
            os.makedirs(os.path.dirname(data_path), exist_ok=True)
            with open(data_path, "wb") as f:
                f.write(payload)
            return data_path
        return None

    def receive_data(self, no_block: bool = False, data_path: str = "results/recv_data.pt", keep_data: bool = False):
        raw = self.recv_socket.recv(flags=zmq.NOBLOCK) if no_block else self.recv_socket.recv()
        if keep_data:
            os.makedirs(os.path.dirname(data_path), exist_ok=True)
            with open(data_path, "wb") as f:
                f.write(raw)
        # weights_only: the socket is bound to a network address and the messages are dicts of tensors / ints / None / str
        # (or a bare tensor) — nothing that needs the unpickler to run code (the reference's torch.load(path) default)
        return torch.load(io.BytesIO(raw), weights_only=True)

    def close(self) -> None:
        self.recv_socket.close(0)
        self.send_socket.close(0)


class NodeWorker:
    CLEAR_KV_CACHE_COMMAND = "clear_KV_cache"
    CLEAR_KV_CACHE_ORIGIN_KEY = "origin_node"

    def __init__(self, src_addr: str, dst_addr: str, can_receive_user_request: bool, shards_path: str,
                 device="cuda:0", dtype=torch.bfloat16, *, communicator=None, transport: str | None = None, max_tokens: int = 4096,
                 max_batch: int = 1, max_slots: int | None = None, max_seq_len: int | None = None,
                 causal: bool = False, use_graphs: bool = True, state_provider=None):
        """Positional parameters as in the reference (utils/node_worker.py:75-76).  ``transport="p2p"``
        replaces the ZeroMQ data path by the NVLink mailbox (``P2PCommunicator``, same interface; the nodes must share
        one NVSwitch box), ``"zmq"`` is the reference's; ``communicator`` injects a ready-made one.  ``causal=False`` is
        the reference's (non-causal prefill) behaviour."""
        from .engine import read_model_config

        if communicator is None:
            # B200SHARD_TRANSPORT lets the reference's unchanged launch scripts (start_node.py, profiling.py) pick the NVLink path
            transport = transport or os.environ.get("B200SHARD_TRANSPORT", "zmq")
            if transport == "p2p":
                from .p2p_communicator import P2PCommunicator
                communicator = P2PCommunicator(src_addr, dst_addr, device=device)
            elif transport == "zmq":
                communicator = Communicator(src_addr, dst_addr)
            else:
                raise ValueError(f"[ERROR] unknown transport {transport!r} (zmq or p2p)")
        self.communicator = communicator
        self.can_receive_user_request = can_receive_user_request
        self.shards_path = shards_path
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("[ERROR] the B200 shard path runs on CUDA devices only (no CPU fallback)")
        self.dtype = dtype
        self.config = read_model_config(shards_path)
        self.layer_num = self.config["num_hidden_layers"]
        self.max_tokens, self.max_batch = max_tokens, max_batch
        self.max_slots = max_slots if max_slots is not None else max_batch
        self.max_seq_len = max_seq_len or self.config["max_position_embeddings"]
        self.causal = causal
        self.use_graphs = use_graphs
        self.state_provider = state_provider

        self.tokenizer = None
        self.embed_tokens = None   # callable ids -> hidden once the shard is loaded
        self.rope = None           # RoPE is recomputed inside the kernels; kept as a placement marker
        self.shard = None
        self.past_key_value = None
        self.lm_head = None        # placement marker: True on the tail node

        self.start = 0
        self.end = 0
        self.batch_size = 0
        if can_receive_user_request:
            self.generated_ids = []
            self._load_embedding()
            self.input_token_length = None

    def _load_embedding(self) -> None:
        """Tokenizer now, embedding table when the shard handle exists (utils/node_worker.py:110-125)."""
        print("[INFO] loading tokenizer...")
        from transformers import AutoTokenizer  # host-side text handling only

        self.tokenizer = AutoTokenizer.from_pretrained(self.shards_path)
        print("[INFO] tokenizer loaded.")
        if not os.path.isfile(os.path.join(self.shards_path, "embedding.pth")):
            raise FileNotFoundError(os.path.join(self.shards_path, "embedding.pth"))

    def load_shards(self, start: int, end: int) -> None:
        """(Re)load layers [start, end) and drop the previous ones (utils/node_worker.py:127-185)."""
        if start < 0 or start >= end or end > self.layer_num:
            raise ValueError("[ERROR] start or end is invalid")
        self.start = start
        self.end = end
        self.past_key_value = None
        if self.shard is not None:
            self.shard.engine.close()
        self.shard = None
        self.rope = None
        self.lm_head = None
        gc.collect()
        torch.cuda.empty_cache()

        is_tail = self.end == self.layer_num
        print(f"[INFO] loading hidden layer {start}~{end}(end excluded)...")
        self.shard = LlamaShardPart(
            self.shards_path, ["block_" + str(i) + ".pth" for i in range(start, end)], start, end,
            device=self.device, dtype=self.dtype,
            add_final_norm=is_tail, final_norm_weight="final_norm.pth" if is_tail else None,
            lm_head_weight="lm_head.pth" if is_tail else None,
            embedding_weight="embedding.pth" if self.can_receive_user_request else None,
            max_tokens=self.max_tokens, max_batch=self.max_batch, state_provider=self.state_provider)
        self.shard.eval()
        self.shard.engine.set_graphs(self.use_graphs)
        if self.start == 0:
            self.rope = True
        if is_tail:
            self.lm_head = True
        if self.can_receive_user_request:
            self.embed_tokens = self.shard.engine.embed
        print(f"[INFO] hidden layer {start}~{end}(end excluded) loaded.")
        self.past_key_value = self.shard.new_cache(self.max_slots, self.max_seq_len)
        print("[INFO] KV cache loaded.")

    # ------------------------------------------------------------------ message producers / consumers
    @torch.inference_mode()
    def receive_user_request(self, request: str = "Write a poem about the blue sky.",
                             input_ids: torch.Tensor | None = None) -> dict:
        """utils/node_worker.py:187-224."""
        if not self.can_receive_user_request:
            raise RuntimeError("[ERROR] this node does not have embedding layer while receiving user request.")
        if input_ids is None:
            print("[INFO] input: " + request)
            input_ids = self.tokenizer(request, return_tensors="pt")["input_ids"].to(self.device)
        else:
            if input_ids.ndim != 2:
                raise ValueError("[ERROR] input_ids must be a 2D tensor with shape [batch_size, seq_len].")
            input_ids = input_ids.to(device=self.device, dtype=torch.long)
        self.input_token_length = input_ids.shape[1]
        self.generated_ids = [input_ids]
        if self.embed_tokens is None:
            raise RuntimeError("[ERROR] load_shards must run before a user request can be embedded.")
        hidden_states = self.embed_tokens(input_ids)  # [B, S, H] bf16 on the GPU
        batch_size, seq_len, _ = hidden_states.shape
        self.batch_size = batch_size
        return {"hidden_states": hidden_states, "batch_size": batch_size, "seq_len": seq_len}

    @torch.inference_mode()
    def pass_through_shard(self, state_info: dict):
        """utils/node_worker.py:226-272: the hot path."""
        if self.is_input_token_info(state_info):
            if self.start != 0:
                raise RuntimeError("[ERROR] after embedding layer, the states should first passing hidden layer 0!")
            self.batch_size = state_info["batch_size"]
        elif self.is_next_state_info(state_info):
            pass
        else:
            attrs = [a for a in dir(state_info) if not a.startswith("__")]
            raise RuntimeError(f"[ERROR] Received unknown state_info type: {type(state_info)}; "
                               f"Attributes: {attrs if attrs else 'No attributes found'}")
        hidden_states = state_info["hidden_states"]
        slots = state_info.get("slot")
        B = hidden_states.shape[0]
        if slots is None:
            slots = list(range(B))
        pos_offset = self.past_key_value.get_seq_length(slots[0])
        if "pos_offset" in state_info and state_info["pos_offset"] != pos_offset:
            raise RuntimeError(f"[ERROR] stage cache length {pos_offset} disagrees with the message's "
                               f"pos_offset {state_info['pos_offset']}")
        hidden_states = hidden_states.to(device=self.device, dtype=self.dtype)
        engine = self.shard.engine
        if self.end == self.layer_num:
            next_token_id, _, _ = engine.forward_argmax(hidden_states, self.past_key_value, slots, causal=self.causal)
            return next_token_id  # int64 [B]
        next_hidden_states = engine.forward(hidden_states, self.past_key_value, slots, causal=self.causal)
        out = {"hidden_states": next_hidden_states, "cos": None, "sin": None, "pos_offset": pos_offset}
        if "slot" in state_info:
            out["slot"] = state_info["slot"]
        return out

    @torch.inference_mode()
    def receive_next_token(self, next_token_id: torch.Tensor, max_new_tokens: int = 1024) -> tuple:
        """utils/node_worker.py:274-309 (stop rule ``len(generated_ids) > max_new_tokens`` kept)."""
        if not self.can_receive_user_request:
            raise RuntimeError('[ERROR] this node does not store the "generated token id" list, '
                               "but received a next_token_id.")
        single = next_token_id.item() if next_token_id.numel() == 1 else None  # read where the id lives (no extra sync for host ids)
        next_token_id = next_token_id.to(self.device)
        self.generated_ids.append(next_token_id.unsqueeze(-1))
        reached_eos = False
        if single is not None:
            next_token = self.tokenizer.decode(single)
            print(repr(next_token), end=" ", flush=True)
            reached_eos = next_token == self.tokenizer.eos_token
        reached_end = reached_eos or len(self.generated_ids) > max_new_tokens
        if reached_end:
            print()
            final_ids = torch.cat(self.generated_ids, dim=-1)
            print("output: ", self.tokenizer.decode(final_ids[0]))
            print("\n[INFO] output token number: " + str(len(self.generated_ids) - 1))
            return reached_end, None
        hidden_states = self.embed_tokens(next_token_id.reshape(-1, 1))  # [B, 1, H]
        return reached_end, {"hidden_states": hidden_states, "batch_size": self.batch_size, "seq_len": 1}

    @staticmethod
    def is_input_token_info(data: object) -> bool:
        return isinstance(data, dict) and "batch_size" in data and "seq_len" in data

    @staticmethod
    def is_next_state_info(data: object) -> bool:
        return isinstance(data, dict) and "cos" in data and "sin" in data

    def clear_KV_cache(self) -> None:
        """Reset per-request state, keep the loaded weights (utils/node_worker.py:319-355).  The
        pages go back to the pool; nothing is reallocated."""
        self.batch_size = 0
        if self.can_receive_user_request:
            self.generated_ids = []
            self.input_token_length = None
        if isinstance(self.past_key_value, PagedKVCache):
            self.past_key_value.reset(-1)
        elif self.shard is not None:
            self.past_key_value = self.shard.new_cache(self.max_slots, self.max_seq_len)
        print("\n[INFO] KV cache and all states caused by prev user are cleared.")

    def _get_clear_KV_cache_origin(self) -> dict:
        return {"src_addr": self.communicator.src_addr, "dst_addr": self.communicator.dst_addr,
                "shards_start": self.start, "shards_end": self.end}

    def build_clear_KV_cache_command(self) -> dict:
        return {"command": self.CLEAR_KV_CACHE_COMMAND, self.CLEAR_KV_CACHE_ORIGIN_KEY: self._get_clear_KV_cache_origin()}

    @classmethod
    def is_clear_KV_cache_command(cls, data: object) -> bool:
        return isinstance(data, dict) and data.get("command") == cls.CLEAR_KV_CACHE_COMMAND

    def is_clear_KV_cache_command_origin(self, data: object) -> bool:
        return (self.is_clear_KV_cache_command(data)
                and data.get(self.CLEAR_KV_CACHE_ORIGIN_KEY) == self._get_clear_KV_cache_origin())


class NodeController:
    """Receives a JSON config from the master node, builds the ``NodeWorker`` and serves messages
    (utils/node_worker.py:385-559).  Same config keys, ports and dispatch rules."""

    def __init__(self, shards_path: str, device: str, dtype: torch.dtype, listen_port: int = 40700, **worker_kwargs):
        self.shards_path = shards_path
        self.device = device
        self.dtype = dtype
        self.worker_kwargs = worker_kwargs
        self.listen_addr = "tcp://*:" + str(listen_port)
        self._cfg_ctx = zmq.Context()
        self.recv_config_socket = self._cfg_ctx.socket(zmq.PULL)
        self.recv_config_socket.bind(self.listen_addr)
        self.send_request_socket = None
        self.first_node_addr = ""
        self._seen_tokens = 0     # next-token messages handled (head node)
        self._token_log = []      # their ids (batch-1 requests)
        self.received_config = self._receive_config()
        if self.received_config["can_receive_user_request"]:
            self._open_request_socket(self.received_config["first_node_addr"])
        self.node_worker = self._build_worker(self.received_config)
        print("[INFO] Node is ready.")

    def _build_worker(self, cfg: dict) -> NodeWorker:
        w = NodeWorker(src_addr=cfg["src_addr"], dst_addr=cfg["dst_addr"],
                       can_receive_user_request=cfg["can_receive_user_request"], shards_path=self.shards_path,
                       device=self.device, dtype=self.dtype, **self.worker_kwargs)
        w.load_shards(cfg["shards_start"], cfg["shards_end"])
        return w

    def _open_request_socket(self, addr: str) -> None:
        self.first_node_addr = addr
        self._req_ctx = zmq.Context()
        self.send_request_socket = self._req_ctx.socket(zmq.PUSH)
        self.send_request_socket.connect(addr)

    def _receive_config(self, no_block: bool = False):
        if no_block:
            try:
                cfg = self.recv_config_socket.recv_json(flags=zmq.NOBLOCK)
            except zmq.Again:
                return None
        else:
            print("[CONFIG] Waiting for configuration file from master node...")
            cfg = self.recv_config_socket.recv_json()
        print("[CONFIG] Received configuration file from master node:")
        for k, v in cfg.items():
            print(f"  - {k}: {v}")
        return cfg

    def _change_first_node_addr(self, new_first_node_addr: str) -> str:
        if self.send_request_socket is None:  # the reference raises AttributeError here on non-entry nodes
            if new_first_node_addr:
                self._open_request_socket(new_first_node_addr)
            return new_first_node_addr
        self.send_request_socket.disconnect(self.first_node_addr)
        self.send_request_socket.connect(new_first_node_addr)
        self.first_node_addr = new_first_node_addr
        return self.first_node_addr

    def check_new_config(self) -> None:
        new_config = self._receive_config(no_block=True)
        if not new_config:
            return
        if new_config["can_receive_user_request"] != self.received_config["can_receive_user_request"]:
            self.node_worker.communicator.close()
            del self.node_worker
            if new_config["can_receive_user_request"]:
                self._change_first_node_addr(new_config["first_node_addr"])
            self.node_worker = self._build_worker(new_config)
        else:
            self.node_worker.communicator.change_src_addr(new_config["src_addr"])
            self.node_worker.communicator.change_dst_addr(new_config["dst_addr"])
            if new_config["can_receive_user_request"]:
                self._change_first_node_addr(new_config["first_node_addr"])
            self.node_worker.load_shards(new_config["shards_start"], new_config["shards_end"])
        self.received_config = new_config
        print("[INFO] The new configuration node is ready.")

    def _forward_request(self, data: dict, data_path: str = "results/send_request.pt", keep_data: bool = False):
        comm = self.node_worker.communicator
        if hasattr(comm, "forward_request"):   # P2P transport: no pickle on any socket
            comm.forward_request(data, self.first_node_addr, self.send_request_socket)
            return None
        buf = io.BytesIO()
        torch.save(Communicator._to_wire(data), buf)
        self.send_request_socket.send(buf.getvalue())
        return None

    def receive_request(self, request: str = "Write a poem about the blue sky.", input_ids=None) -> None:
        if not self.node_worker.can_receive_user_request:
            raise RuntimeError("[ERROR] this node cannot receive user request.")
        token_info = self.node_worker.receive_user_request(request, input_ids=input_ids)
        self._forward_request(token_info)

    def step(self, max_new_tokens: int = 1024) -> bool:
        """One iteration of the reference's ``while True`` body (utils/node_worker.py:502-559).
        Returns True when a message was handled."""
        w = self.node_worker
        handled = False
        try:
            data = w.communicator.receive_data(no_block=True)
            handled = True
            if w.is_clear_KV_cache_command(data):
                if not w.is_clear_KV_cache_command_origin(data):
                    w.communicator.transfer_data(data)
                    w.clear_KV_cache()
                self.check_new_config()
                return True
            run = True
            if isinstance(data, torch.Tensor):
                if w.start != 0:
                    raise RuntimeError('[ERROR] I\'m not the first node in the model chain, but received "next_token_id".')
                self._seen_tokens += 1
                if data.numel() == 1:
                    self._token_log.append(int(data.item()))
                reached_end, data = w.receive_next_token(data, max_new_tokens)
                if reached_end:
                    w.communicator.transfer_data(w.build_clear_KV_cache_command())
                    w.clear_KV_cache()
                    run = False
            elif w.is_input_token_info(data):
                if w.start != 0:
                    raise RuntimeError('[ERROR] I\'m not the first node in the model chain, but received "input_token_info".')
            elif w.is_next_state_info(data):
                if w.start == 0:
                    raise RuntimeError('[ERROR] I\'m the first node in the model chain, but received "next_state_info".')
            else:
                attrs = [a for a in dir(data) if not a.startswith("__")]
                raise RuntimeError(f"[ERROR] Received unknown data type: {type(data)}; "
                                   f"Attributes: {attrs if attrs else 'No attributes found'}")
            if run:
                out = w.pass_through_shard(data)
                w.communicator.transfer_data(out)
                if w.start != 0:
                    print("*", end=" ", flush=True)
        except zmq.Again:
            pass
        self.check_new_config()
        return handled

    def run_worker_loop(self, max_new_tokens: int = 1024) -> None:
        while True:
            self.step(max_new_tokens)
