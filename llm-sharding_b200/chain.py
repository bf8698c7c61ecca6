"""Chain runtime: the reference's in-process N-worker ring (utils/node_profiler.py:1174-1236) and
the transports that replace its ZeroMQ hop on one NVSwitch box.

* ``LoopbackCommunicator`` — ``Communicator``-shaped (transfer_data / receive_data / change_*),
  in-process queues; device tensors are handed over by reference (same GPU) — used when several
  shards share one GPU and by the single-GPU benchmark.
* ``P2PHop`` — the NVLink hop: ``b200_hop_*`` mailbox + in-kernel 128-bit peer stores + flag.
* ``generate`` — drives a list of ``NodeWorker``s exactly like ``go_through_every_shards``.
"""
from __future__ import annotations

import collections
import ctypes as C
import os

import torch
import zmq

from . import _lib

_QUEUES: dict[str, collections.deque] = {}


def _norm(addr: str) -> str:
    return addr.replace("tcp://*:", "tcp://127.0.0.1:")


class LoopbackCommunicator:
    def __init__(self, src_addr: str, dst_addr: str):
        self.src_addr, self.dst_addr = src_addr, dst_addr
        self.actual_src_addr = _norm(src_addr)
        _QUEUES.setdefault(self.actual_src_addr, collections.deque())

    def change_src_addr(self, new_src_addr: str) -> str:
        self.src_addr, self.actual_src_addr = new_src_addr, _norm(new_src_addr)
        _QUEUES.setdefault(self.actual_src_addr, collections.deque())
        return self.src_addr

    def change_dst_addr(self, new_dst_addr: str) -> str:
        self.dst_addr = new_dst_addr
        return self.dst_addr

    def transfer_data(self, data, data_path: str = "", keep_data: bool = False):
        _QUEUES.setdefault(_norm(self.dst_addr), collections.deque()).append(data)
        return None

    def receive_data(self, no_block: bool = False, data_path: str = "", keep_data: bool = False):
        q = _QUEUES[self.actual_src_addr]
        if not q:
            raise zmq.Again()
        return q.popleft()

    def close(self) -> None:
        _QUEUES.pop(self.actual_src_addr, None)


class P2PHop:
    """One receive mailbox on this GPU and a mapping of the next stage's mailbox."""

    def __init__(self, device: torch.device, n_ring_slots: int, slot_bytes: int):
        _lib.require_cuda()
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        h = C.c_void_p()
        _lib.check(self.lib.b200_hop_create(self.dev_index, n_ring_slots, slot_bytes, C.byref(h)))
        self.handle = h
        self.n_ring_slots, self.slot_bytes = n_ring_slots, slot_bytes

    def export(self) -> bytes:
        buf = C.create_string_buffer(_lib.HOP_HANDLE_BYTES)
        _lib.check(self.lib.b200_hop_export(self.handle, buf))
        return buf.raw

    def connect(self, peer_handle: bytes, peer_device: int, same_process: bool) -> None:
        _lib.check(self.lib.b200_hop_connect(self.handle, peer_handle, peer_device, int(same_process)))

    def send(self, payload: torch.Tensor, seq: int) -> None:
        t = payload.contiguous()
        _lib.check(self.lib.b200_hop_send(self.handle, t.data_ptr(), t.numel() * t.element_size(), seq,
                                          _lib.stream_ptr(self.device)))

    def recv(self, out: torch.Tensor, seq: int, max_wait_us: int = 2_000_000) -> torch.Tensor:
        _lib.check(self.lib.b200_hop_recv(self.handle, out.data_ptr(), out.numel() * out.element_size(), seq,
                                          max_wait_us, _lib.stream_ptr(self.device)))
        return out

    def error(self) -> int:
        return self.lib.b200_hop_error(self.handle)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200_hop_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def generate(workers, input_ids: torch.Tensor, max_new_tokens: int, record_states: bool = False,
             teacher_tokens: torch.Tensor | None = None):
    """The reference's in-process loop: request -> every shard in order -> next token -> re-embed.
    Returns ``(tokens [n, B] int64 (cpu), states per pass (list of lists, fp32 cpu))``.
    ``teacher_tokens`` [n, B] forces the token fed back at each step (used by parity tests to keep
    comparing after a near-tie flips an arg-max)."""
    head = workers[0]
    msg = head.receive_user_request(input_ids=input_ids)
    tokens, states = [], []
    n = 0
    while True:
        per_pass = []
        data = msg
        for w in workers:
            data = w.pass_through_shard(data)
            if isinstance(data, dict) and record_states:
                per_pass.append(data["hidden_states"].float().cpu())
        tokens.append(data.cpu())
        states.append(per_pass)
        fed = data if teacher_tokens is None else teacher_tokens[n].to(data.device)
        n += 1
        head.generated_ids.append(fed.unsqueeze(-1))
        if len(head.generated_ids) > max_new_tokens:
            break
        msg = {"hidden_states": head.embed_tokens(fed.reshape(-1, 1)), "batch_size": head.batch_size, "seq_len": 1}
    return torch.stack(tokens), states


# ----------------------------------------------------------------------------------------------
# Multi-GPU ring: one process per GPU, one shard per process (the reference's NodeController-per-node
# model, utils/node_worker.py:385-559 + run_this.sh), messages handed over inside the NVSwitch box.
# ----------------------------------------------------------------------------------------------
def layer_range(rank: int, world: int, n_layers: int) -> tuple[int, int]:
    """Contiguous, as even as possible split [start, end) of the layers over `world` stages (the kind of
    split ConfigSender.build_config(shards_start, shards_end) describes, utils/config_sender.py:38-39)."""
    if world < 1 or world > n_layers:
        raise ValueError("[ERROR] start or end is invalid")
    base, extra = divmod(n_layers, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class RingSchedule:
    """Sequence numbers and mailbox slots of one stage of the ring.

    Every stage consumes the messages of its predecessor in the order they were produced, so both sides
    can count: message number `seq` of a hop lives in mailbox slot `seq % n_slots`.  With `n_streams`
    request streams served round-robin and n_slots == n_streams, stream i always lands in slot i — which
    keeps the captured decode graphs (keyed by their pointers) to one per stream."""

    def __init__(self, rank: int, world: int, n_streams: int):
        if n_streams < 1:
            raise ValueError("[ERROR] at least one request stream is required")
        self.rank, self.world, self.n_streams = rank, world, n_streams
        self.is_head, self.is_tail = rank == 0, rank == world - 1
        self.seq_in = 0    # next message expected from the predecessor
        self.seq_out = 0   # next message sent to the successor
        self.prefilled = [False] * n_streams
        self._owed = []    # streams whose token is on its way back to the head, in the order the tail produces them

    def _slot(self, seq: int) -> int:
        return seq % self.n_streams

    def _consume_token_of(self, stream: int) -> None:
        """The head's mailbox is a FIFO: a stream may take its token only when it is the oldest one outstanding, i.e.
        streams must be continued (or finished) in the order their previous items were issued — round-robin does."""
        if not self._owed or self._owed[0] != stream:
            raise RuntimeError(f"[ERROR] stream {stream} continued out of order: the next token in the ring belongs to "
                               f"stream {self._owed[0] if self._owed else None}")
        self._owed.pop(0)

    def prefill(self, stream: int) -> dict:
        """Work item for the prompt of `stream`.  The head consumes nothing (the prompt comes from its own host)."""
        if self.prefilled[stream]:
            raise RuntimeError(f"[ERROR] stream {stream} already holds a request; clear it first")
        self.prefilled[stream] = True
        self._owed.append(stream)
        item = {"kind": "prefill", "stream": stream, "kv_slot": stream, "wait": None, "send": None}
        if not self.is_head:
            item["wait"] = (self.seq_in, self._slot(self.seq_in))
            self.seq_in += 1
        item["send"] = (self.seq_out, self._slot(self.seq_out))
        self.seq_out += 1
        return item

    def decode(self, stream: int) -> dict:
        """Work item for one more token of `stream`: every stage (the head too: it waits for the token id the
        tail sent back, utils/node_worker.py:515-520) consumes one message and produces one."""
        if not self.prefilled[stream]:
            raise RuntimeError(f"[ERROR] stream {stream} has no request to continue")
        self._consume_token_of(stream)
        self._owed.append(stream)
        item = {"kind": "decode", "stream": stream, "kv_slot": stream,
                "wait": (self.seq_in, self._slot(self.seq_in)), "send": (self.seq_out, self._slot(self.seq_out))}
        self.seq_in += 1
        self.seq_out += 1
        return item

    def finish(self, stream: int) -> dict:
        """Work item that retires the request of `stream` (its stop rule fired, utils/node_worker.py:291-299): the head
        consumes the last token the tail sent back — every message produced must be consumed or the two sides' counts
        drift apart — and nothing is sent; the other stages only release the stream's KV slots.  The stream can then be
        given a new prompt while the remaining streams keep decoding."""
        if not self.prefilled[stream]:
            raise RuntimeError(f"[ERROR] stream {stream} has no request to finish")
        self._consume_token_of(stream)
        self.prefilled[stream] = False
        item = {"kind": "finish", "stream": stream, "kv_slot": stream, "wait": None, "send": None}
        if self.is_head:
            item["wait"] = (self.seq_in, self._slot(self.seq_in))
            self.seq_in += 1
        return item

    def clear(self) -> None:
        """Drop every request (all stages call this at the same point, after draining their streams).  Tokens the tail
        already sent back and the head never consumed are skipped, so that both ends of every hop keep the same count:
        otherwise the head's next wait would be satisfied by a stale flag of a dropped request."""
        if self.is_head:
            self.seq_in += len(self._owed)
        self.prefilled = [False] * self.n_streams
        self._owed = []


class _DevPtr:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_view(ptr: int, nbytes: int, device) -> torch.Tensor:
    """uint8 view of LOCAL device memory owned by the library (mailbox slots)."""
    return torch.as_tensor(_DevPtr(ptr, nbytes), device=device)


class ChainStage:
    """One stage of the layer-sharded ring on its own GPU.  `worker` is this repo's NodeWorker (already
    `load_shards`-ed with max_slots >= n_streams).  Transport "p2p": hidden states / token ids are written by the
    producing kernel straight into the next GPU's mailbox over NVLink and published with a flag; "nccl":
    ncclSend / ncclRecv through torch.distributed (the A/B alternative)."""

    def __init__(self, worker, rank: int, world: int, n_streams: int, max_prompt: int, ctrl_group=None,
                 transport: str = "p2p", max_wait_us: int = 20_000_000, rows: int = 1, split_tail: bool | None = None):
        import torch.distributed as dist

        self.w, self.rank, self.world, self.n_streams = worker, rank, world, n_streams
        self.sched = RingSchedule(rank, world, n_streams)
        self.eng, self.kv = worker.shard.engine, worker.past_key_value
        if self.eng.io_dtype != torch.bfloat16:
            raise ValueError("[ERROR] the ring's mailbox slots are laid out for bfloat16 hidden states; start the worker with dtype=torch.bfloat16")
        self.lib = self.eng.lib
        self.device = worker.device
        self.H = self.eng.hidden_size
        self.causal = int(worker.causal)
        self.transport = transport
        self.max_wait_us = max_wait_us
        self.dist = dist
        self.nxt, self.prv = (rank + 1) % world, (rank - 1) % world
        self.rows = rows  # batch rows of one stream (micro-batch); stream i owns KV slots [i*rows, (i+1)*rows)
        # Vocabulary-split tail: the reference's last worker carries final norm + lm_head + argmax for every stream
        # (utils/node_worker.py:155-164,260-264) and so bounds the ring; here the tail scores the lower half of lm_head and the
        # head — which receives the token anyway — the upper half, from the fp32 residual row the tail sends along
        # (include/b200shard.h: bit-identical tokens).  Every stage derives the same decision from the same inputs.
        if split_tail is None:
            split_tail = os.environ.get("B200SHARD_SPLIT_TAIL", "1") != "0"
        self.split = 0
        if split_tail and world >= 2 and transport == "p2p" and rows <= 32:
            self.split = int(worker.config["vocab_size"]) // 2
        self.msg_bytes = rows * (self.H * 4 + 16)
        self.slot_bytes = max(rows * max_prompt * self.H * 2, 4096, self.msg_bytes if self.split else 0)
        if self.split and self.sched.is_head:
            load = worker.state_provider or (lambda f: torch.load(os.path.join(worker.shards_path, f), map_location="cpu"))
            lm = load("lm_head.pth")["weight"]
            try:
                fn = load("final_norm.pth")["weight"]
            except FileNotFoundError:
                fn = None
            self.eng.set_vocab_tail(lm[self.split:], fn, self.split)
            del lm
            self._tok_buf = [torch.zeros(rows, dtype=torch.int64, device=self.device) for _ in range(n_streams)]
        if transport == "p2p":
            self.hop = P2PHop(self.device, n_streams, self.slot_bytes)
            mine = (self.hop.export(), self.hop.dev_index)
            gathered = [None] * world
            dist.all_gather_object(gathered, mine, group=ctrl_group)
            self.hop.connect(gathered[self.nxt][0], gathered[self.nxt][1], same_process=(world == 1))
            dist.barrier(group=ctrl_group)
        else:
            # NCCL A/B transport.  Two communicators: hidden states travel forward on one, token ids return to the head
            # on the other, so that a token the head has not asked for yet cannot block the forward traffic.  Sends are
            # non-blocking (isend): like a mailbox store they must not stall the producer's stream.
            self.hop = None
            self.pg_fwd = dist.new_group(backend="nccl")
            self.pg_tok = dist.new_group(backend="nccl")
            self.inbuf = [torch.empty(self.slot_bytes, dtype=torch.uint8, device=self.device) for _ in range(n_streams)]
            self.outbuf = [torch.empty(self.slot_bytes, dtype=torch.uint8, device=self.device) for _ in range(n_streams)]
            self._pending = []
            # NCCL builds the pair-wise communicators lazily and that set-up blocks the host until the peer joins:
            # do it now, in ring order, so that it can never interleave with the asynchronous pipeline
            probe = torch.zeros(8, dtype=torch.uint8, device=self.device)
            if not self.sched.is_head:
                dist.recv(probe, src=self.prv, group=self.pg_fwd)
            if not self.sched.is_tail:
                dist.send(probe, dst=self.nxt, group=self.pg_fwd)
            if world > 1:
                if self.sched.is_tail:
                    dist.send(probe, dst=self.nxt, group=self.pg_tok)
                if self.sched.is_head:
                    dist.recv(probe, src=self.prv, group=self.pg_tok)
            torch.cuda.synchronize()
            dist.barrier(group=ctrl_group)
        self.tokens = []       # head: device int64[1] tensors, one per recorded decode
        self.host_tokens = []  # head, via_host mode: python ints
        self.record_inbound = False  # non-head stages: keep a copy of every hidden state that arrives (parity checks)
        self.inbound = []            # (stream, S, bf16 tensor [rows, S, H])

    # -- raw-pointer plumbing
    def _st(self):
        return _lib.stream_ptr(self.device)

    def _in_ptr(self, item) -> int:
        if self.hop is not None:
            return self.lib.b200_hop_local_slot(self.hop.handle, item["wait"][0])
        return self.inbuf[item["wait"][1]].data_ptr()

    def _out_ptr(self, item) -> int:
        if self.hop is not None:
            return self.lib.b200_hop_peer_slot(self.hop.handle, item["send"][0])
        return self.outbuf[item["send"][1]].data_ptr()

    def _wait(self, item, nbytes: int) -> None:
        if item["wait"] is None:
            return
        if self.hop is not None:
            _lib.check(self.lib.b200_hop_recv(self.hop.handle, None, 0, item["wait"][0], self.max_wait_us, self._st()))
        else:
            self.dist.recv(self.inbuf[item["wait"][1]][:nbytes], src=self.prv,
                           group=self.pg_tok if self.sched.is_head else self.pg_fwd)

    def _publish(self, item, nbytes: int) -> None:
        if self.hop is not None:
            _lib.check(self.lib.b200_hop_send(self.hop.handle, None, 0, item["send"][0], self._st()))
        else:
            self._pending.append(self.dist.isend(self.outbuf[item["send"][1]][:nbytes], dst=self.nxt,
                                                 group=self.pg_tok if self.sched.is_tail else self.pg_fwd))
            if len(self._pending) > 4 * self.n_streams:
                self._pending = [w for w in self._pending if not w.is_completed()]

    def _token_ptr(self, item) -> int:
        """Head: device address of the token ids that the message just waited for carries — the mailbox slot itself, or, with
        the vocabulary-split tail, this stream's token buffer after the head scored its half of lm_head."""
        if not self.split:
            return self._in_ptr(item)
        buf = self._tok_buf[item["stream"]]
        _lib.check(self.lib.b200_shard_token_from_partial(self.eng.handle, self._in_ptr(item), self.rows, buf.data_ptr(), None, self._st()))
        return buf.data_ptr()

    def _token_view(self, item, tok_ptr: int) -> torch.Tensor:
        if self.split:
            return self._tok_buf[item["stream"]]
        if self.hop is not None:
            return tensor_view(tok_ptr, 8 * self.rows, self.device)
        return self.inbuf[item["wait"][1]][:8 * self.rows]

    def _run(self, item, S: int, tok_in_ptr: int | None, record: bool, via_host: bool = False) -> None:
        import ctypes as C
        B = self.rows
        slots = (C.c_int32 * B)(*range(item["kv_slot"] * B, item["kv_slot"] * B + B))
        sch = self.sched
        hid_bytes, tok_bytes = B * S * self.H * 2, 8 * B
        in_bytes = tok_bytes if (sch.is_head and item["kind"] == "decode") else hid_bytes
        if self.split:
            tok_bytes = self.msg_bytes
        out_bytes = tok_bytes if sch.is_tail else hid_bytes
        self._wait(item, in_bytes)
        out_ptr = self._out_ptr(item)
        if self.record_inbound and not sch.is_head:
            src = (tensor_view(self._in_ptr(item), hid_bytes, self.device) if self.hop is not None
                   else self.inbuf[item["wait"][1]][:hid_bytes])
            self.inbound.append((item["stream"], S, src.view(torch.bfloat16).view(B, S, self.H).clone()))
        if sch.is_head:
            if tok_in_ptr is None:
                tok_in_ptr = self._token_ptr(item)
                if via_host:
                    # end-to-end mode: the id leaves the GPU (D2H, the host blocks until the ring delivered it), is seen
                    # by the host loop (the reference decodes / prints it there) and comes back from pinned memory (H2D)
                    src = self._token_view(item, tok_in_ptr)
                    if not hasattr(self, "_host_tok"):
                        self._host_tok = torch.empty(B, dtype=torch.int64).pin_memory()
                        self._dev_tok = [torch.empty(B, dtype=torch.int64, device=self.device) for _ in range(self.n_streams)]
                    self._host_tok.copy_(src.view(torch.int64))  # synchronous D2H
                    self.host_tokens.append(self._host_tok.tolist())
                    self._dev_tok[item["stream"]].copy_(self._host_tok, non_blocking=True)
                    tok_in_ptr = self._dev_tok[item["stream"]].data_ptr()
                elif record:
                    self.tokens.append(self._token_view(item, tok_in_ptr).view(torch.int64).clone())
            if sch.is_tail:  # single-stage ring: token in, token out
                _lib.check(self.lib.b200_shard_step_tokens(self.eng.handle, self.kv.handle, slots, tok_in_ptr, B, S, self.causal,
                                                           None, out_ptr, None, self._st()))
            else:
                _lib.check(self.lib.b200_shard_step_tokens(self.eng.handle, self.kv.handle, slots, tok_in_ptr, B, S, self.causal,
                                                           out_ptr, None, None, self._st()))
        elif sch.is_tail and self.split:
            _lib.check(self.lib.b200_shard_forward_partial(self.eng.handle, self.kv.handle, slots, self._in_ptr(item), B, S,
                                                           self.causal, self.split, out_ptr, self._st()))
        elif sch.is_tail:
            _lib.check(self.lib.b200_shard_forward_argmax(self.eng.handle, self.kv.handle, slots, self._in_ptr(item), B, S,
                                                          self.causal, None, out_ptr, None, self._st()))
        else:
            _lib.check(self.lib.b200_shard_forward(self.eng.handle, self.kv.handle, slots, self._in_ptr(item), B, S,
                                                   self.causal, out_ptr, self._st()))
        self._publish(item, out_bytes)

    # -- public
    def prefill(self, stream: int, prompt_len: int, ids: torch.Tensor | None = None) -> None:
        """ids: int64 [rows * prompt_len] (row-major [rows, prompt_len]) on this GPU (head only)."""
        item = self.sched.prefill(stream)
        if self.sched.is_head and ids is None:
            raise ValueError("[ERROR] the head stage needs the prompt ids")
        self._run(item, prompt_len, ids.data_ptr() if self.sched.is_head else None, False)
        if self.sched.is_head:
            self._keep = getattr(self, "_keep", []) + [ids]  # keep the prompt alive until the kernels ran

    def decode(self, stream: int, record: bool = False, via_host: bool = False) -> None:
        """One more token for `stream`.  `via_host` (head only): the token id that came back through the ring makes a
        round trip through pinned host memory before it is embedded (end-to-end mode)."""
        item = self.sched.decode(stream)
        self._run(item, 1, None, record, via_host and self.sched.is_head)

    def finish(self, stream: int, record: bool = False) -> None:
        """Retire the request of `stream` on this stage (same call on every stage, at the same point of the schedule):
        the head takes the last token out of its mailbox, every stage returns the stream's KV pages to its pool
        (stream-ordered, behind the kernels that still read them)."""
        item = self.sched.finish(stream)
        B = self.rows
        if item["wait"] is not None:
            self._wait(item, self.msg_bytes if self.split else 8 * B)
            if record:
                self.tokens.append(self._token_view(item, self._token_ptr(item)).view(torch.int64).clone())
        for slot in range(stream * B, stream * B + B):
            self.kv.reset(slot)

    def error(self) -> int:
        return self.hop.error() if self.hop is not None else 0

    def close(self) -> None:
        if self.hop is not None:
            self.hop.close()


# ----------------------------------------------------------------------------------------------
# Stop rules inside the ring: the reference's run_worker_loop / receive_next_token decide on the HEAD node when a request
# is over (EOS, or len(generated_ids) > max_new_tokens, utils/node_worker.py:290-299) and then send a clear-cache
# command around the ring (:357-382, :507-524).  With several requests in flight the same decisions are taken per
# stream: RingPlanner (pure host logic, head only) turns "what the head has seen so far" into one action per stream and
# round, RingServer broadcasts that plan over the control group and every stage enqueues the same work items.
# ----------------------------------------------------------------------------------------------
ACT_IDLE, ACT_DECODE, ACT_FINISH, ACT_STOP = 0, -1, -2, -9   # a positive action is a prefill of that many prompt tokens


class RingPlanner:
    """Per-stream request state on the head.  `max_new_tokens` is the reference's rule: exactly that many tokens are
    produced and the last one is not fed back; it is applied without looking at any token, so no work
    is wasted.  The EOS rule needs token values, which reach the head's host a little after the ring has moved on: a
    request that hit EOS is retired at the first round that knows about it and its output is cut at the EOS token."""

    def __init__(self, n_streams: int, rows: int = 1, eos_token_id: int | None = None):
        self.n_streams, self.rows, self.eos = n_streams, rows, eos_token_id
        self.pending = []                       # (request id, prompt_len, max_new_tokens)
        self.active = [None] * n_streams        # per stream: id of the request it serves
        self.reqs = {}                          # request id -> dict(req, issued, max_new, eos_seen, row_eos, tokens)
        self.done = {}                          # request id -> list of per-step token lists
        self._next_id = 0

    def submit(self, prompt_len: int, max_new_tokens: int) -> int:
        if prompt_len < 1 or max_new_tokens < 1:
            raise ValueError("[ERROR] a request needs at least one prompt token and one new token")
        rid = self._next_id
        self._next_id += 1
        self.pending.append((rid, prompt_len, max_new_tokens))
        return rid

    def observe(self, stream: int, request: int, tokens: list) -> None:
        """A token (one id per row) of `request` reached the head's host."""
        st = self.reqs.get(request)
        if st is None or st["eos_seen"]:
            return                                # produced after the EOS: dropped
        st["tokens"].append(list(tokens))
        if self.eos is not None:
            for r, t in enumerate(tokens):
                st["row_eos"][r] = st["row_eos"][r] or t == self.eos
            if all(st["row_eos"]):
                st["eos_seen"] = True

    def plan(self) -> list:
        """Actions of the next round, one per stream: [finish?, action] pairs flattened to 2 * n_streams ints."""
        out = []
        for i in range(self.n_streams):
            a = self.reqs[self.active[i]] if self.active[i] is not None else None
            fin, act = 0, ACT_IDLE
            if a is not None:
                if a["eos_seen"] or a["issued"] >= a["max_new"]:
                    fin = 1
                    self.done[a["req"]] = a["tokens"]
                    self.active[i] = a = None
                else:
                    act = ACT_DECODE
                    a["issued"] += 1
            if a is None and self.pending:
                rid, plen, mnt = self.pending.pop(0)
                self.reqs[rid] = {"req": rid, "issued": 1, "max_new": mnt, "eos_seen": False, "tokens": [],
                                  "row_eos": [False] * self.rows}
                self.active[i] = rid
                act = plen
            out += [fin, act]
        return out

    def idle(self) -> bool:
        return not self.pending and all(a is None for a in self.active)


class RingServer:
    """The generalised run_worker_loop of one stage: rounds of per-stream actions, planned on the head (rank 0) and
    broadcast over the control group (`dist` gloo group; `None` for a one-stage ring).  The head needs the prompts
    (`submit`), the other stages only their turn in `serve()`."""

    def __init__(self, stage: ChainStage, ctrl_group=None, eos_token_id: int | None = None):
        self.stage, self.ctrl = stage, ctrl_group
        self.is_head = stage.sched.is_head
        self.planner = RingPlanner(stage.n_streams, stage.rows, eos_token_id) if self.is_head else None
        self.prompts = {}        # head: request id -> ids [rows, S] on this GPU
        self._stream_req = [None] * stage.n_streams
        self._records = []       # head: (stream, request, index into stage.tokens)
        self._seen = 0

    def submit(self, ids: torch.Tensor, max_new_tokens: int) -> int:
        if not self.is_head:
            raise RuntimeError("[ERROR] this node cannot receive user request.")
        if ids.ndim != 2 or ids.shape[0] != self.stage.rows:
            raise ValueError("[ERROR] input_ids must be a 2D tensor with shape [batch_size, seq_len].")
        rid = self.planner.submit(ids.shape[1], max_new_tokens)
        self.prompts[rid] = ids.to(device=self.stage.device, dtype=torch.long).contiguous()
        return rid

    def _drain_tokens(self, block: bool) -> None:
        """Move the tokens whose copies have completed from the device to the planner (never blocks unless asked)."""
        while self._seen < len(self._records):
            stream, req, idx, ev = self._records[self._seen]
            if not block and not ev.query():
                return
            ev.synchronize()
            self.planner.observe(stream, req, self.stage.tokens[idx].tolist())
            self._seen += 1

    def _exchange(self, plan: list) -> list:
        if self.ctrl is None or self.stage.world == 1:
            return plan
        t = torch.tensor(plan if self.is_head else [0] * (2 * self.stage.n_streams), dtype=torch.int64)
        self.stage.dist.broadcast(t, src=0, group=self.ctrl)
        return t.tolist()

    def _record(self, stream: int) -> None:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.stage.device))
        self._records.append((stream, self._stream_req[stream], len(self.stage.tokens) - 1, ev))

    def serve(self, max_rounds: int = 1 << 30) -> dict:
        """Run rounds until every submitted request is finished (head's decision).  Returns, on the head,
        {request id: int64 tensor [n_tokens, rows]}; {} elsewhere."""
        st = self.stage
        for _ in range(max_rounds):
            if self.is_head:
                self._drain_tokens(block=False)
                plan = [ACT_STOP] * (2 * st.n_streams) if self.planner.idle() else self.planner.plan()
            else:
                plan = None
            plan = self._exchange(plan)
            if plan[1] == ACT_STOP:
                break
            for i in range(st.n_streams):
                fin, act = plan[2 * i], plan[2 * i + 1]
                if fin:
                    st.finish(i, record=self.is_head)
                    if self.is_head:
                        self._record(i)
                if act == ACT_DECODE:
                    st.decode(i, record=self.is_head)
                    if self.is_head:
                        self._record(i)
                elif act > 0:
                    ids = None
                    if self.is_head:
                        rid = self.planner.active[i]
                        self._stream_req[i] = rid
                        ids = self.prompts.pop(rid).reshape(-1)
                    st.prefill(i, act, ids)
        if not self.is_head:
            return {}
        self._drain_tokens(block=True)
        return {rid: torch.tensor(toks, dtype=torch.int64) for rid, toks in self.planner.done.items()}
